#!/bin/bash
# objects epilogue / prelude timing: parity tests, headline bench, ncu launch times of the two kernels
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_objects.py tests/test_gpu_golden.py tests/test_gpu_custom_layout.py tests/test_gpu_panner.py -m gpu -x -q 2>&1 | tail -3
PCMD="python bench.py --steps 3 --warmup 3 --no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e"
timeout 200 python bench.py --steps 10 --warmup 3 --no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['kernel_ms'], d['e2e']['value'])"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:"objects_(epilogue|prelude)|extent_core" -c 60 --csv --log-file gpurun_out/epi.csv $PCMD > /dev/null 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/epi.csv')) if len(r) > 5]
h = rows[0]; ki = h.index('Kernel Name'); vi = h.index('Metric Value'); ui = h.index('Metric Unit')
acc = collections.defaultdict(list)
for r in rows[1:]:
    try:
        acc[r[ki][:60]].append((float(r[vi].replace(',', '')), r[ui]))
    except ValueError:
        pass
for k, v in acc.items():
    print(k, len(v), sum(x for x, _ in v) / len(v), v[0][1])
PY
