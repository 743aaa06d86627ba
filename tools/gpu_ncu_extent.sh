#!/bin/bash
tag=${1:-n}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --objects 262144 --no-render --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/extent_plain_$tag.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:extent_core -s 2 -c 1 -f -o gpurun_out/prof_extent_$tag $CMD > gpurun_out/ncu_extent_$tag.log 2>&1
echo "ncu rc=$?"
