"""Kernel table of one encoder call (C2 shapes) per precision: what runs besides the repo's own kernels."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from quber_b200.pixel_decoder import MSDeformAttnTransformerEncoderOnly
dev = "cuda"
torch.manual_seed(0)
enc = MSDeformAttnTransformerEncoderOnly(256, 8, 6, 1024, 0.0, "relu", 3, 4).to(dev).eval()
shapes = [(15, 20), (30, 40), (60, 80)]
B = 32
srcs = [torch.randn(B, 256, h, w, device=dev).contiguous(memory_format=torch.channels_last) for h, w in shapes]
pos = [(torch.randn(B, 256, h, w, device=dev) * 0.1).contiguous(memory_format=torch.channels_last) for h, w in shapes]
with torch.no_grad():
    for prec in (sys.argv[1:] or ["bf16", "fp32"]):
        enc.set_precision(prec)
        for _ in range(3):
            enc(srcs, pos)
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            enc(srcs, pos)
            torch.cuda.synchronize()
        print("=====", prec)
        rows = sorted(prof.key_averages(), key=lambda e: -getattr(e, "device_time_total", 0.0))
        tot = sum(getattr(e, "device_time_total", 0.0) for e in rows)
        for e in rows[:18]:
            print(f"{e.device_time_total * 1e-3:8.3f} ms  x{e.count:3d}  {e.key[:110]}")
        print(f"{tot * 1e-3:8.3f} ms total device time")
