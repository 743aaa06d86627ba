#!/bin/bash
# wide convolver kernel: parity tests, then cfg-4 time of the pair and the wide kernel for the warp counts given (development)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_convolver.py tests/test_gpu_flow.py -x -q 2>&1 | tail -4
echo "pair:"; EARB200_CONV_PATH=pair timeout 200 python tools/bench_conv.py 56250
for w in "$@"; do
  (cd libear_b200 && nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -DEARB_WIDE_WARPS=$w -c csrc/convolver_fast_kernels.cu -o build/convolver_fast_kernels.o && nvcc -gencode arch=compute_100a,code=sm_100a -shared -o libearb200.so build/*.o -lcudart) || exit 1
  echo "wide, $w warps:"; timeout 200 python tools/bench_conv.py 56250
done
