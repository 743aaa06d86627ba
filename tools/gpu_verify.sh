#!/bin/bash
# One GPU visit without profiler passes: parity tests, smoke, bench (both arms).
# Usage (under gpurun): bash tools/gpu_verify.sh <tag>
tag=${1:-v}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi_$tag.txt 2>&1
SECONDS=0
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$? after ${SECONDS}s"
tail -5 gpurun_out/pytest_$tag.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$tag.log 2>&1; echo "smoke rc=$? after ${SECONDS}s"
tail -2 gpurun_out/smoke_$tag.log
timeout 900 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$? after ${SECONDS}s"
cat gpurun_out/bench_$tag.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err; echo "ref rc=$? after ${SECONDS}s"
cat gpurun_out/bench_ref_$tag.json
