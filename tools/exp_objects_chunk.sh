#!/bin/bash
# chunk-size sweep of the batched objects path (EARB200_OBJECTS_CHUNK fixes the chunk; unset = adaptive)
for c in 49152 98304 147456 196608 adaptive; do
  echo "chunk $c"
  if [ "$c" = adaptive ]; then unset EARB200_OBJECTS_CHUNK; else export EARB200_OBJECTS_CHUNK=$c; fi
  timeout 200 python bench.py --steps 10 --warmup 3 --no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['kernel_ms'], d['e2e']['value'])"
done
