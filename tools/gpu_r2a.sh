#!/bin/bash
# round 2, first visit: the cell-culled extent core (parity + A/B against the first-generation core),
# then the parity tests on the timed entry points
tag=${1:-r2a}
mkdir -p gpurun_out
OBJ="--no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline"
timeout 900 python -m pytest tests/test_gpu_objects.py tests/test_gpu_golden.py -x -q > gpurun_out/pytest_obj_$tag.log 2>&1; echo "pytest objects rc=$?"
tail -15 gpurun_out/pytest_obj_$tag.log
timeout 600 python bench.py --steps 10 --warmup 3 $OBJ > gpurun_out/bench_cells_$tag.json 2> gpurun_out/bench_cells_$tag.err; echo "bench cells rc=$?"
EARB200_EXTENT_PATH=tf32 timeout 600 python bench.py --steps 10 --warmup 3 $OBJ > gpurun_out/bench_tf32_$tag.json 2> gpurun_out/bench_tf32_$tag.err; echo "bench tf32 rc=$?"
python - <<PY
import json
for k in ("cells", "tf32"):
    try:
        b = json.load(open("gpurun_out/bench_%s_$tag.json" % k))
        print(k, "value %.1f M" % (b["value"] / 1e6), "e2e %.1f M" % ((b["e2e"]["value"] or 0) / 1e6), "ms/step %.3f" % b["ms_per_step"],
              "core ms/launch %.3f" % b["roofline"]["kernel_ms"], "share %.3f" % b["roofline"]["kernel_share_of_step"])
    except Exception as e:
        print(k, "failed", e)
PY
tail -3 gpurun_out/bench_cells_$tag.err
timeout 1500 python -m pytest tests/test_gpu_timed_paths.py -x -q -s > gpurun_out/pytest_timed_$tag.log 2>&1; echo "pytest timed rc=$?"
tail -25 gpurun_out/pytest_timed_$tag.log
