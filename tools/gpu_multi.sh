#!/bin/bash
# multi-GPU visit: bus tests over NVLink, the bench as the driver runs it, then the exchange arms
# usage: bash tools/gpu_multi.sh <n_gpus> <tag>
N=${1:-2}; tag=${2:-m}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$tag.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_bus.py tests/test_gpu_round2.py -x -q -k "bus or devices" > gpurun_out/pytest_multi_$tag.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_multi_$tag.log
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N"
show() {
  python - <<PY
import json
try:
    b = json.load(open("$1"))
    print("$2: objects %.1f M e2e %.1f M | render %.3f ms | mix %.3f ms tail %s ms | parity %s" % (
        b["value"] / 1e6, (b["e2e"]["value"] or 0) / 1e6, b["render"]["ms_per_step"], b["mix"]["ms_per_step"],
        b["mix"]["bus_reduce_ms"], {k: v for k, v in b["parity"].items() if "sharded" in k or k == "ok"}))
except Exception as e:
    print("$2 failed", e)
PY
}
timeout 1200 $RUN --steps 20 --warmup 3 > gpurun_out/bench_${N}gpu_$tag.json 2> gpurun_out/bench_${N}gpu_$tag.err
echo "bench default rc=$?"; tail -2 gpurun_out/bench_${N}gpu_$tag.err; show gpurun_out/bench_${N}gpu_$tag.json default
OPT="--steps 10 --warmup 3 --no-convolver --no-flow --no-points --no-cpu-baseline --no-e2e --sustained 0"
timeout 900 $RUN $OPT --bus-gather 0 > gpurun_out/bench_${N}gpu_rs_$tag.json 2> gpurun_out/bench_${N}gpu_rs_$tag.err
echo "bench reduce-scatter rc=$?"; show gpurun_out/bench_${N}gpu_rs_$tag.json reduce-scatter
timeout 900 $RUN $OPT --exchange nccl > gpurun_out/bench_${N}gpu_nccl_$tag.json 2> gpurun_out/bench_${N}gpu_nccl_$tag.err
echo "bench nccl rc=$?"; show gpurun_out/bench_${N}gpu_nccl_$tag.json nccl
