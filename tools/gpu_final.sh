#!/bin/bash
# Final visit of a round: parity tests, smoke, bench (both arms), launch lists, ncu of the objects prelude / epilogue.
# Usage (under gpurun): bash tools/gpu_final.sh <tag>
tag=${1:-f}
bash tools/gpu_verify.sh $tag
PCMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-render-e2e --no-e2e --no-mix"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$tag.csv $PCMD > gpurun_out/ncu_launches_$tag.log 2>&1
echo "ncu launches rc=$?"
bash tools/gpu_launches_cfg2.sh $tag
OBJ="python bench.py --steps 2 --warmup 3 --no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-parity --sustained 0"
for k in epilogue prelude; do
  timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:objects_$k -s 2 -c 1 -f -o gpurun_out/prof_${k}_$tag $OBJ > gpurun_out/ncu_${k}_$tag.log 2>&1
  echo "ncu $k rc=$?"
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:extent_cell -s 3 -c 1 -f -o gpurun_out/prof_cell_$tag $OBJ > gpurun_out/ncu_cell_$tag.log 2>&1
echo "ncu cell rc=$?"
