#!/bin/bash
# ncu --set full capture of one objects kernel (default: the epilogue); usage: bash tools/gpu_ncu_epi1.sh <tag> [regex]
tag=${1:-e}; rx=${2:-objects_epilogue}
mkdir -p gpurun_out
PCMD="python bench.py --steps 2 --warmup 3 --no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-parity --sustained 0"
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:$rx -s 2 -c 1 -f -o gpurun_out/prof_${tag} $PCMD > gpurun_out/ncu_${tag}.log 2>&1
echo "ncu rc=$?"
