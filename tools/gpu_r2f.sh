#!/bin/bash
tag=${1:-r2f}
mkdir -p gpurun_out
OBJ="--no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-parity --sustained 0"
for d in 0 1 2 3 4 8 12 16 32 33 36 44 47; do
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
