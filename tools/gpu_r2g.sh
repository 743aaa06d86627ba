#!/bin/bash
# round 2 state check: all GPU tests, smoke, bench (both arms), launch list, ncu of the cell core
tag=${1:-r2g}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi_$tag.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/pytest_$tag.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$tag.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"
cat gpurun_out/bench_$tag.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err; echo "ref rc=$?"
cat gpurun_out/bench_ref_$tag.json
PCMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-render-e2e --no-e2e --no-mix --no-parity --sustained 0"
timeout 600 $PCMD > gpurun_out/plain_$tag.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$tag.csv $PCMD > gpurun_out/ncu_launches_$tag.log 2>&1
echo "ncu launches rc=$?"
CMD="python bench.py --steps 1 --warmup 3 --objects 262144 --no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-parity --sustained 0"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:extent_cell -s 2 -c 1 -f -o gpurun_out/prof_cell_$tag $CMD > gpurun_out/ncu_cell_$tag.log 2>&1
echo "ncu cell rc=$?"
ls -la gpurun_out
RCMD="python bench.py --steps 2 --warmup 3 --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-render-e2e --no-parity --sustained 0 --objects 8192"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_tc_persistent -s 3 -c 1 -f -o gpurun_out/prof_render_$tag $RCMD > gpurun_out/ncu_render_$tag.log 2>&1
echo "ncu render rc=$?"
CCMD="python bench.py --steps 1 --warmup 3 --no-render --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-render-e2e --no-parity --sustained 0 --objects 8192"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:convolver_fast -s 3 -c 1 -f -o gpurun_out/prof_conv_$tag $CCMD > gpurun_out/ncu_conv_$tag.log 2>&1
echo "ncu conv rc=$?"
