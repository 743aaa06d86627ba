#!/bin/bash
# launch list of the cfg-2 step alone (objects block only)
tag=${1:-l}
mkdir -p gpurun_out
PCMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-render --no-convolver --no-flow --no-mix --no-points --no-e2e --no-parity --sustained 0"
timeout 600 $PCMD > /dev/null 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_${tag}cfg2.csv $PCMD > /dev/null 2>&1
echo "rc=$?"
