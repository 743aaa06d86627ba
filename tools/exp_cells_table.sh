#!/bin/bash
# offline-searched cell table of the 1,652-point grid against the run-time bisection: parity tests, cfg-2 bench
OBJ="--no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --sustained 0"
timeout 600 python -m pytest tests/test_gpu_objects.py tests/test_gpu_golden.py tests/test_gpu_timed_paths.py tests/test_gpu_round2.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -1
for m in table bisect table bisect; do
  EARB200_EXTENT_CELLS=$m timeout 300 python bench.py --steps 10 --warmup 3 $OBJ 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$m: %.1f M vectors/s  core %.4f ms  parity %s %.2e' % (b['value']/1e6, b['roofline']['kernel_ms'], b['parity']['ok'], b['parity']['gains_max_abs']))"
done
