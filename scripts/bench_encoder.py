"""The MSDeformAttn transformer encoder (6 layers, msdeformattn.py:61-161) at the C2 shapes (32 frames, S = 6300):
time per precision and the top kernels of one forward.  Usage: python scripts/bench_encoder.py [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from quber_b200.pixel_decoder import MSDeformAttnTransformerEncoderOnly
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
torch.manual_seed(0)
enc = MSDeformAttnTransformerEncoderOnly(256, 8, 6, 1024, 0.0, "relu", 3, 4).cuda().eval()
shapes = [(15, 20), (30, 40), (60, 80)]                   # res5 -> res3 order (msdeformattn.py:318-324)
srcs = [torch.randn(B, 256, h, w, device="cuda") for h, w in shapes]
pos = [torch.randn(B, 256, h, w, device="cuda") * 0.1 for h, w in shapes]
def t(f, n=5):
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
with torch.no_grad():
    for prec in ("fp32", "tf32", "bf16"):
        enc.set_precision(prec)
        ms = t(lambda: enc(srcs, pos))
        print(f"encoder x6 {prec}: {ms:.2f} ms per 32 frames ({ms / 6:.2f} ms / layer)", flush=True)
    for prec in ("tf32", "bf16"):
        enc.set_precision(prec)
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            enc(srcs, pos); torch.cuda.synchronize()
        print(prec)
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=22, max_name_column_width=60))
