#!/bin/bash
# per-chunk timeline of the host pipeline (EARB200_HOST_TRACE=1) on the cfg-2 step, last step only
mkdir -p gpurun_out
OBJ="--no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-parity --sustained 0"
EARB200_HOST_TRACE=1 timeout 300 python bench.py --steps 5 --warmup 3 $OBJ 2> gpurun_out/host_trace.err | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('device %.1f M  e2e %.1f M  floor %.1f M' % (b['value']/1e6, b['e2e']['value']/1e6, b['e2e']['floor']['value']/1e6))"
grep host-trace gpurun_out/host_trace.err | tail -${1:-14}
