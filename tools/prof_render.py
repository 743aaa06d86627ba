#!/usr/bin/env python
"""Small driver for profiling the K3 render kernel (ncu): BASELINE config 2 shape, shorter T."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import libear_b200 as eb  # noqa: E402

# usage: prof_render.py [tiles-of-512 [N [M [ramp-spacing]]]]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 128
M = int(sys.argv[3]) if len(sys.argv) > 3 else 14
RAMP = int(sys.argv[4]) if len(sys.argv) > 4 else 512
T = 512 * int(sys.argv[1]) if len(sys.argv) > 1 else 512 * 1184
dev = torch.device("cuda", 0)
x = (torch.randn((N, T), device=dev) * 0.25).clamp_(-1, 1)
y = torch.empty((M, T), device=dev)
samples = np.arange(0, T + RAMP, RAMP, dtype=np.int64)
gains = (np.random.default_rng(3).uniform(0, 1, size=(len(samples), N, M)) / np.sqrt(N)).astype(np.float32)
gi = eb.GainInterpolator(N, M)
gi.set_points(samples, gains)
stream = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    gi.process_device(0, T, x.data_ptr(), T, y.data_ptr(), T, stream=stream)
torch.cuda.synchronize()
gi.set_profiling(True)
for _ in range(5):
    gi.process_device(0, T, x.data_ptr(), T, y.data_ptr(), T, stream=stream)
ms, n = gi.kernel_time()
alg = 4.0 * T * (N + M) + 4.0 * N * M * len(samples)
print(f"N={N} M={M} ramp={RAMP} T={T} kernel {ms / n:.4f} ms  {alg / (ms / n * 1e-3) / 1e9:.1f} GB/s")
