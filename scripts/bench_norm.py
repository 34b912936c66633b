"""Micro-benchmark of quber_group_norm_relu and quber_upsample_concat at the C2 network shapes (32 frames)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from quber_b200 import ops
def t(f, n=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
B = 32
for (C, H, W) in ((256, 120, 160), (128, 120, 160), (512, 60, 80), (2048, 30, 40), (32, 120, 160)):
    for dt in (torch.bfloat16, torch.float32):
        for cl in (True, False):
            x = torch.randn(B, C, H, W, device="cuda").to(dt)
            x2 = torch.randn(B, C, H, W, device="cuda").to(dt)      # alternate inputs: 2 x 315 MB > L2
            if cl:
                x, x2 = x.contiguous(memory_format=torch.channels_last), x2.contiguous(memory_format=torch.channels_last)
            w, b = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")
            if not ops.group_norm_supported(x, 32):
                continue
            it = [0]
            def f():
                it[0] ^= 1
                ops.group_norm_relu(x if it[0] else x2, w, b, 32, 1e-5, inplace=True)
            ms = t(f)
            byts = x.numel() * x.element_size() * 3          # 2 reads + 1 write
            ms_t = t(lambda: torch.relu_(torch.nn.functional.group_norm(x, 32, w.to(dt), b.to(dt), 1e-5)))
            print(f"gn_relu C={C} {H}x{W} {str(dt)[6:]} {'nhwc' if cl else 'nchw'}: {ms*1e3:.0f} us, {byts/ms/1e6:.0f} GB/s (2R+1W)   torch group_norm+relu_: {ms_t*1e3:.0f} us", flush=True)
for dt in (torch.bfloat16, torch.float32):
    a = torch.randn(B, 32, 120, 160, device="cuda").to(dt).contiguous(memory_format=torch.channels_last)
    y = torch.randn(B, 128, 60, 80, device="cuda").to(dt).contiguous(memory_format=torch.channels_last)
    ms = t(lambda: ops.upsample_concat(a, y))
    byts = (a.numel() * 2 + y.numel() + B * 128 * 120 * 160) * a.element_size()
    ms_t = t(lambda: torch.cat([a, torch.nn.functional.interpolate(y, size=(120, 160), mode="bilinear", align_corners=False)], 1))
    print(f"upsample_concat 32+128 -> 120x160 {str(dt)[6:]}: {ms*1e3:.0f} us, {byts/ms/1e6:.0f} GB/s   torch interpolate+cat: {ms_t*1e3:.0f} us", flush=True)
