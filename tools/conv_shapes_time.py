"""Convolver bank of cfg 4's size (24 channels x 10 min at 48 kHz, 512-tap decorrelators) at every
partitioning of the 512 taps: time of one steady-state call on the device."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import libear_b200 as eb

C, T = 24, 56250 * 512
dev = torch.device("cuda", 0)
x = torch.randn((C, T), device=dev) * 0.25
y = torch.empty_like(x)
filt = eb.design_decorrelators("9+10+3", np.float32)
st = torch.cuda.current_stream().cuda_stream
for B, nb in ((512, 1), (256, 2), (128, 4), (64, 8), (32, 16), (16, 32)):
    conv = eb.BlockConvolver(B, C, nb)
    for c in range(C):
        conv.set_filter(conv.add_filter(filt[c])[0], c)
    K = T // B
    conv.process_device(K, x.data_ptr(), T, y.data_ptr(), T, stream=st)  # first call: warm-up, tail of set_filter
    times = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        conv.process_device(K, x.data_ptr(), T, y.data_ptr(), T, stream=st)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = min(times)
    print("B=%4d x %2d partitions: %.3f ms  (%.2f TB/s of in + out)" % (B, nb, ms, 2 * 4 * C * T / ms / 1e9))
