#!/usr/bin/env python
"""Development probe: groups of 256 objects mixed into one bus (cfg5 shape), timing per variant."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import libear_b200 as eb
G = int(sys.argv[1]) if len(sys.argv) > 1 else 16
RAMP = int(sys.argv[2]) if len(sys.argv) > 2 else 48000
ACC = int(sys.argv[3]) if len(sys.argv) > 3 else 1
M, T = 12, 2_880_000
dev = torch.device("cuda", 0)
x = torch.empty((G * 256, T), device=dev).normal_(0, 0.25)
bus = torch.empty((M, T), device=dev)
samples = np.arange(0, T + RAMP, RAMP, dtype=np.int64)
rng = np.random.default_rng(0)
gis = []
for g in range(G):
    gi = eb.GainInterpolator(256, M)
    gi.set_points(samples, (rng.uniform(0, 1, (len(samples), 256, M)) / 16).astype(np.float32))
    gi.set_accumulate(bool(ACC) and g > 0)
    gis.append(gi)
st = torch.cuda.current_stream().cuda_stream
def step():
    for g, gi in enumerate(gis):
        gi.process_device(0, T, x[g * 256].data_ptr(), T, bus.data_ptr(), T, stream=st)
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
for gi in gis: gi.set_profiling(True)
step(); 
ks = [gi.kernel_time()[0] for gi in gis]
print(f"G={G} ramp={RAMP} acc={ACC}: {ms:.3f} ms/step, {4.0*T*G*256/ms/1e6:.0f} GB/s input; per-group kernel ms: " + " ".join(f"{k:.3f}" for k in ks))
