#!/bin/bash
tag=${1:-n}
mkdir -p gpurun_out
timeout 300 python tools/prof_render.py 2960 > gpurun_out/render_plain_$tag.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_tc -s 3 -c 1 -f -o gpurun_out/prof_render_$tag python tools/prof_render.py 2960 > gpurun_out/ncu_render_$tag.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/render_plain_$tag.log | head -3
