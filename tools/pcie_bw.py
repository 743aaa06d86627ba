#!/usr/bin/env python
"""Host<->device copy rates with pinned memory at the sizes of one cfg-2 step (64 MB in, 192 MB out),
each direction alone and both at once: the floor of the end-to-end number."""
import torch

dev = torch.device("cuda", 0)
h_in = torch.empty(64_000_000, dtype=torch.uint8).pin_memory()
h_out = torch.empty(192_000_000, dtype=torch.uint8).pin_memory()
d_in = torch.empty_like(h_in, device=dev)
d_out = torch.empty_like(h_out, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    for s in (s1, s2):
        torch.cuda.current_stream().wait_stream(s)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def h2d():
    with torch.cuda.stream(s1):
        s1.wait_stream(torch.cuda.current_stream())
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        s2.wait_stream(torch.cuda.current_stream())
        h_out.copy_(d_out, non_blocking=True)


def both():
    h2d()
    d2h()


t = timed(h2d)
print(f"H2D  64 MB alone: {t:.3f} ms  {64e6 / t / 1e6:.1f} GB/s")
t = timed(d2h)
print(f"D2H 192 MB alone: {t:.3f} ms  {192e6 / t / 1e6:.1f} GB/s")
t = timed(both)
print(f"both at once:     {t:.3f} ms per step (64 MB in + 192 MB out)")
