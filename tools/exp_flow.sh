#!/bin/bash
# build the render kernel with an experiment macro and time the Objects signal flow with it (development only;
# the outputs of the experiment builds are wrong by design)
cd libear_b200
for e in "$@"; do
  nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -w -DEARB_RENDER_EXPERIMENT=$e -c csrc/render_tc_kernels.cu -o build/render_tc_kernels.o && nvcc -gencode arch=compute_100a,code=sm_100a -shared -o libearb200.so build/*.o -lcudart && echo "EXP $e" && (cd .. && timeout 300 python tools/flow_time.py 2>&1 | tail -1; timeout 100 python tools/prof_render.py 5625 128 24 | head -1)
done
