#!/bin/bash
# round 2: cell-culled extent core v2 (indicator MMA, pooled lists, cleaner): parity, bench, ncu
tag=${1:-r2c}
mkdir -p gpurun_out
OBJ="--no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline"
timeout 900 python -m pytest tests/test_gpu_objects.py tests/test_gpu_golden.py -x -q > gpurun_out/pytest_obj_$tag.log 2>&1; rc=$?; echo "pytest objects rc=$rc"
tail -25 gpurun_out/pytest_obj_$tag.log
timeout 600 python bench.py --steps 10 --warmup 3 $OBJ > gpurun_out/bench_cells_$tag.json 2> gpurun_out/bench_cells_$tag.err; echo "bench cells rc=$?"
python - <<PY
import json
try:
    b = json.load(open("gpurun_out/bench_cells_$tag.json"))
    print("value %.1f M" % (b["value"] / 1e6), "e2e %.1f M" % ((b["e2e"]["value"] or 0) / 1e6), "ms/step %.3f" % b["ms_per_step"],
          "core ms/launch %.3f" % b["roofline"]["kernel_ms"], "share %.3f" % b["roofline"]["kernel_share_of_step"])
except Exception as e:
    print("failed", e)
PY
tail -3 gpurun_out/bench_cells_$tag.err
if [ $rc -eq 0 ]; then
timeout 1500 python -m pytest tests/test_gpu_timed_paths.py -x -q -s -k objects > gpurun_out/pytest_timed_$tag.log 2>&1; echo "pytest timed rc=$?"
tail -5 gpurun_out/pytest_timed_$tag.log
CMD="python bench.py --steps 1 --warmup 3 --objects 262144 $OBJ --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:extent_cell -s 2 -c 1 -f -o gpurun_out/prof_cell_$tag $CMD > gpurun_out/ncu_cell_$tag.log 2>&1
echo "ncu rc=$?"
fi
