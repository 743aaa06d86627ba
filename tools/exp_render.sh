#!/bin/bash
# build the render kernel with an experiment macro and time it (development only)
cd libear_b200
for e in "$@"; do
  nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -w -DEARB_RENDER_EXPERIMENT=$e -c csrc/render_tc_kernels.cu -o build/render_tc_kernels.o && nvcc -gencode arch=compute_100a,code=sm_100a -shared -o libearb200.so build/*.o -lcudart && echo "EXP $e" && (cd .. && timeout 300 python tools/prof_render.py 5625 2>&1 | head -1)
done
