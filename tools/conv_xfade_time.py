"""cfg-4-sized convolver call with a filter crossfade pending on every channel: time of the call."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import libear_b200 as eb

C, B, K = 24, 512, 56250
dev = torch.device("cuda", 0)
x = torch.randn((C, K * B), device=dev) * 0.25
y = torch.empty_like(x)
conv = eb.BlockConvolver(B, C, 1)
filt = eb.design_decorrelators("9+10+3", np.float32)
ids = [conv.add_filter(filt[c])[0] for c in range(C)]
for c in range(C):
    conv.set_filter(ids[c], c)
st = torch.cuda.current_stream().cuda_stream
for label, xf in (("steady", False), ("crossfade pending", True), ("steady", False), ("crossfade pending", True)):
    if xf:
        for c in range(C):
            conv.crossfade_filter(ids[(c + 1) % C], c)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    conv.process_device(K, x.data_ptr(), K * B, y.data_ptr(), K * B, stream=st)
    e1.record()
    torch.cuda.synchronize()
    print("%-18s %.3f ms" % (label, e0.elapsed_time(e1)))
