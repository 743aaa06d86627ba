#!/bin/bash
# Quick GPU check: render/interp tests, then the render profile driver.  Usage: bash tools/gpu_quick.sh <tag>
tag=${1:-q}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_render.py tests/test_gpu_golden.py -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_$tag.log
timeout 300 python tools/prof_render.py 5625 > gpurun_out/render_$tag.log 2>&1; echo "prof_render rc=$?"
cat gpurun_out/render_$tag.log
