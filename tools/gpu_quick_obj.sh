#!/bin/bash
# Quick GPU check of the objects path: parity tests, then a short bench without the render block.
tag=${1:-o}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_objects.py tests/test_gpu_golden.py tests/test_gpu_panner.py -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_$tag.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-render --no-cpu-baseline > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"
python - <<PY
import json
b=json.load(open("gpurun_out/bench_$tag.json"))
print("value", b["value"], "e2e", b["e2e"]["value"], "ms/step", b["ms_per_step"], "core ms/launch", b["roofline"]["kernel_ms"], "share", b["roofline"]["kernel_share_of_step"])
PY
tail -3 gpurun_out/bench_$tag.err
