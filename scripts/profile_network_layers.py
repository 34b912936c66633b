"""Per-module GPU time of the network stage (forward hooks + CUDA events): top modules with their shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from quber_b200 import config as qcfg
from quber_b200.meta_arch import build_model
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
cfg = qcfg.get_cfg()
cfg.merge_from_file(os.path.join(ROOT, "configs", "mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8.yaml"))
cfg.MODEL.DEVICE = "cuda"
torch.manual_seed(0)
model = build_model(cfg, prec).eval()
x = torch.randn(B, 9, 480, 640, device="cuda")
recs = []
def pre(m, inp):
    m._e0 = torch.cuda.Event(enable_timing=True); m._e0.record()
def post(name):
    def f(m, inp, out):
        e1 = torch.cuda.Event(enable_timing=True); e1.record()
        t = inp[0]
        recs.append((name, type(m).__name__, tuple(t.shape), str(t.dtype), t.is_contiguous(memory_format=torch.channels_last), m._e0, e1,
                     getattr(m, "kernel_size", None), getattr(m, "dilation", None), getattr(m, "out_channels", None)))
    return f
for n, m in model.named_modules():
    if isinstance(m, (torch.nn.Conv2d, torch.nn.GroupNorm)):
        m.register_forward_pre_hook(pre); m.register_forward_hook(post(n))
with torch.no_grad():
    for _ in range(2):
        model.network_logits(x)
    recs.clear()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); model.network_logits(x); e1.record()
    torch.cuda.synchronize()
print("total ms", e0.elapsed_time(e1))
rows = sorted(((a.elapsed_time(b), n, t, s, d, cl, k, dil, oc) for n, t, s, d, cl, a, b, k, dil, oc in recs), reverse=True)
print("sum of hooked modules", sum(r[0] for r in rows))
for r in rows[:25]:
    print("%8.3f ms  %-60s %s in=%s %s cl=%s k=%s dil=%s out=%s" % r)
