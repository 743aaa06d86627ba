#!/bin/bash
# debug-flag sweep of the cell core: usage bash tools/gpu_dbg.sh "0 256 512 32 4 36"
mkdir -p gpurun_out
# needs a build with -DEARB_CELL_DEBUG (make -C libear_b200 clean all EXTRA=-DEARB_CELL_DEBUG)
OBJ="--no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-parity --sustained 0"
for d in $1; do
  EARB200_CELL_DBG=$d timeout 300 python bench.py --steps 5 --warmup 3 $OBJ > gpurun_out/dbg_$d.json 2> gpurun_out/dbg_$d.err
  python - <<PY
import json
try:
    b = json.load(open("gpurun_out/dbg_$d.json"))
    print("dbg $d: value %.1f M  core ms/launch %.3f" % (b["value"] / 1e6, b["roofline"]["kernel_ms"]))
except Exception as e:
    print("dbg $d failed", e)
PY
done
