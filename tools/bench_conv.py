#!/usr/bin/env python
"""Device-time of the decorrelation convolver bank at BASELINE config 4 (24 channels of 9+10+3,
512-tap decorrelators, B = 512): CUDA events around earb200_convolver_process_device."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
import libear_b200 as eb

C, B = 24, 512
K = int(sys.argv[1]) if len(sys.argv) > 1 else 56250 // 4
T = K * B
dev = torch.device("cuda", 0)
x = (torch.randn((C, T), device=dev) * 0.25).clamp_(-1, 1)
y = torch.empty_like(x)
conv = eb.BlockConvolver(B, C, 1)
filt = eb.design_decorrelators("9+10+3", np.float32)
for c in range(C):
    fid, _ = conv.add_filter(filt[c])
    conv.set_filter(fid, c)
stream = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    conv.process_device(K, x.data_ptr(), T, y.data_ptr(), T, stream)
torch.cuda.synchronize()
conv.set_profiling(True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
n = 10
for _ in range(n):
    conv.process_device(K, x.data_ptr(), T, y.data_ptr(), T, stream)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
kms, kl = conv.kernel_time()
print(f"K={K} blocks/channel: {ms:.3f} ms/step, kernel {kms / kl:.3f} ms, "
      f"{8.0 * C * T / (kms / kl * 1e-3) / 1e9:.1f} GB/s algorithmic, {C * T / (ms * 1e-3) / 1e6:.1f} Msamples/s")
