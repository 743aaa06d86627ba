#!/bin/bash
# objects prelude / epilogue variants "min-blocks-per-SM:objects-per-epilogue-block" (development): tests + kernel times under ncu
mkdir -p gpurun_out
PCMD="python bench.py --steps 3 --warmup 3 --no-render --no-convolver --no-mix --no-flow --no-points --no-cpu-baseline --no-e2e --no-parity --sustained 0"
for v in "$@"; do
  mb=${v%%:*}; eo=${v##*:}
  (cd libear_b200 && nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC --fmad=false -DEARB_PRELUDE_MIN_BLOCKS=$mb -DEARB_EPI_OBJECTS=$eo -c csrc/objects_kernels.cu -o build/objects_kernels.o && nvcc -gencode arch=compute_100a,code=sm_100a -shared -o libearb200.so build/*.o -lcudart) || exit 1
  echo "== prelude min blocks $mb, epilogue objects per block $eo"
  timeout 300 python -m pytest tests/test_gpu_objects.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -1
  timeout 200 $PCMD 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('value', d['value'], 'core ms', d['roofline']['kernel_ms'])"
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:"objects_(epilogue|prelude)" -c 12 --csv --log-file gpurun_out/pre_$mb.csv $PCMD > /dev/null 2>&1
  python - <<PY
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/pre_$mb.csv')) if len(r) > 5]
h = rows[0]; ki = h.index('Kernel Name'); vi = h.index('Metric Value')
acc = collections.defaultdict(list)
for r in rows[1:]:
    try: acc[r[ki][:50]].append(float(r[vi].replace(',', '')))
    except ValueError: pass
for k, v in acc.items(): print(k, len(v), sum(v) / len(v) / 1000, 'us')
PY
done
