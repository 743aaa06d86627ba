#!/bin/bash
# ncu --set full capture of the fast convolver kernel at cfg 4 scale (run under gpurun)
tag=${1:-c}
mkdir -p gpurun_out
timeout 300 python tools/bench_conv.py 56250 > gpurun_out/conv_plain_$tag.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:convolver_(fast|wide)" -s 3 -c 1 -f -o gpurun_out/prof_conv_$tag python tools/bench_conv.py 56250 > gpurun_out/ncu_conv_$tag.log 2>&1
echo "ncu conv rc=$?"; grep blocks gpurun_out/conv_plain_$tag.log
