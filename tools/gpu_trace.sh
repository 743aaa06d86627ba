#!/bin/bash
mkdir -p gpurun_out
# needs a build with -DEARB_CELL_DEBUG (make -C libear_b200 clean all EXTRA=-DEARB_CELL_DEBUG)
OBJ="--no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-parity --sustained 0"
EARB200_CELL_DBG=${1:-64} timeout 300 python bench.py --steps 1 --warmup 3 $OBJ 2>&1 | grep trace | tail -${2:-24}
