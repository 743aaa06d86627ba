#!/bin/bash
# wide convolver kernel variants "warps-per-CTA:units-per-run" (development): parity tests and cfg-4 time for each
mkdir -p gpurun_out
for v in "$@"; do
  w=${v%%:*}; u=${v##*:}
  (cd libear_b200 && nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -DEARB_WIDE_WARPS=$w -DEARB_WIDE_UNITS=$u -c csrc/convolver_fast_kernels.cu -o build/convolver_fast_kernels.o && nvcc -gencode arch=compute_100a,code=sm_100a -shared -o libearb200.so build/*.o -lcudart) || exit 1
  echo "== $w warps per CTA, $u units per run:"
  timeout 600 python -m pytest tests/test_gpu_convolver.py tests/test_gpu_flow.py -x -q 2>&1 | tail -1
  timeout 200 python tools/bench_conv.py 56250
  timeout 200 python tools/flow_time.py 2>&1 | tail -1
done
