#!/bin/bash
# round 2: where does the cell-culled extent core spend its time?  debug-flag sweep + ncu capture
tag=${1:-r2b}
mkdir -p gpurun_out
OBJ="--no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-parity --sustained 0"
for d in 0 1 2 4 8 12 16 18 31; do
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
CMD="python bench.py --steps 1 --warmup 3 --objects 262144 $OBJ"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:extent_cell -s 2 -c 1 -f -o gpurun_out/prof_cell_$tag $CMD > gpurun_out/ncu_cell_$tag.log 2>&1
echo "ncu rc=$?"
