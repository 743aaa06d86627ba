#!/bin/bash
# e2e (host buffers) against the chunk size of the host pipeline
OBJ="--no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-parity --sustained 0"
for c in 0 24576 32768 49152 65536 81920 98304 131072 196608; do
  EARB200_OBJECTS_CHUNK=$c python bench.py --steps 5 --warmup 3 $OBJ 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunk $c: device %.1f M  e2e %.1f M' % (b['value']/1e6, b['e2e']['value']/1e6))"
done
python tools/pcie_bw.py 2>&1 | tail -6
