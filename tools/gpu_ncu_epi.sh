#!/bin/bash
# ncu --set full capture of the objects prelude / epilogue kernels (run under gpurun)
tag=${1:-e}
mkdir -p gpurun_out
PCMD="python bench.py --steps 2 --warmup 3 --no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e"
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:objects_epilogue -s 10 -c 1 -f -o gpurun_out/prof_epi_$tag $PCMD > gpurun_out/ncu_epi_$tag.log 2>&1
echo "ncu epi rc=$?"
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:objects_prelude -s 10 -c 1 -f -o gpurun_out/prof_pre_$tag $PCMD > gpurun_out/ncu_pre_$tag.log 2>&1
echo "ncu pre rc=$?"
