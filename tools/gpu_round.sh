#!/bin/bash
# One GPU visit: parity tests, bench (both arms), ncu launch list and full captures.
# Usage (under gpurun): bash tools/gpu_round.sh <tag>
tag=${1:-r1}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi_$tag.txt 2>&1
nproc > gpurun_out/nproc_$tag.txt; lscpu | head -20 >> gpurun_out/nproc_$tag.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_$tag.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$tag.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"
cat gpurun_out/bench_$tag.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err; echo "ref rc=$?"
cat gpurun_out/bench_ref_$tag.json
PCMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-render-e2e --no-e2e --no-mix"
timeout 600 $PCMD > gpurun_out/plain_$tag.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$tag.csv $PCMD > gpurun_out/ncu_launches_$tag.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:extent_core -s 3 -c 2 -f -o gpurun_out/prof_extent_$tag $PCMD > gpurun_out/ncu_extent_$tag.log 2>&1
echo "ncu extent rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render -s 3 -c 2 -f -o gpurun_out/prof_render_$tag $PCMD > gpurun_out/ncu_render_$tag.log 2>&1
echo "ncu render rc=$?"
ls -la gpurun_out
