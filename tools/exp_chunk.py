#!/usr/bin/env python
"""Cost of rendering a bus in time chunks (1 GPU, no collective): N objects -> M feeds, T samples."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import libear_b200 as eb  # noqa: E402
from libear_b200.sharding import chunked_bus, mix_pipelined  # noqa: E402

N, M, T = int(sys.argv[1]), int(sys.argv[2]), 2_880_000
dev = torch.device("cuda", 0)
x = torch.randn((N, T), device=dev) * 0.25
samples = np.arange(0, T + 512, 512, dtype=np.int64)
gains = (np.random.default_rng(3).uniform(0, 1, size=(len(samples), N, M)) / np.sqrt(N)).astype(np.float32)
gi = eb.GainInterpolator(N, M)
gi.set_points(samples, gains)
for chunks in (1, 2, 4, 8, 16):
    bus, ranges = chunked_bus(M, T, chunks, device=dev)
    per = bus.shape[2]

    def render_chunk(k, t0, t1):
        gi.process_device(t0, t1 - t0, x.data_ptr() + 4 * t0, T, bus[k].data_ptr(), per, stream=0)
    for _ in range(3):
        mix_pipelined(bus, render_chunk, ranges=ranges)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(10):
        mix_pipelined(bus, render_chunk, ranges=ranges)
    e1.record()
    host = (time.perf_counter() - t0) / 10 * 1e3
    torch.cuda.synchronize()
    print(f"N={N} M={M} chunks={len(ranges)}: {e0.elapsed_time(e1) / 10:.3f} ms/step (host issue {host:.3f} ms)")
