"""Objects signal flow of bench.py's flow block (128 objects x 60 s into 9+10+3): time per call,
EARB200_FLOW_PAIRED=0 (two render launches) against the default (both renders on one launch)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import libear_b200 as eb

N, T = 128, 2_880_000
dev = torch.device("cuda", 0)
fr = eb.ObjectsRenderer("9+10+3", N)
C = fr.num_outputs
rng = np.random.default_rng(7)
samples = np.arange(0, T + 512, 512, dtype=np.int64)
gd = (rng.uniform(0, 1, size=(len(samples), N, C)) / np.sqrt(N)).astype(np.float32)
gf = (rng.uniform(0, 1, size=(len(samples), N, C)) / np.sqrt(N)).astype(np.float32)
fr.set_gains(samples, gd, gf)
x = (torch.randn((N, T), device=dev) * 0.25).clamp_(-1, 1)
y = torch.empty((C, T), device=dev)
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    fr.process_device(0, T // 512, x.data_ptr(), T, y.data_ptr(), T, stream=st)
ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fr.process_device(0, T // 512, x.data_ptr(), T, y.data_ptr(), T, stream=st)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print("flow: median %.4f ms  min %.4f ms  checksum %.6f" % (float(np.median(ts)), min(ts), float(y.double().abs().sum())))
