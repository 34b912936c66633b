"""Kernel table of one network forward (C2 batch) per precision."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
from quber_b200 import config as qcfg, ops
from quber_b200.meta_arch import build_model
dev = torch.device("cuda", 0)
cfg = qcfg.get_cfg()
cfg.merge_from_file(os.path.join(ROOT, "configs", "mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8.yaml"))
cfg.MODEL.DEVICE = str(dev)
torch.manual_seed(0)
model = build_model(cfg, "fp32").eval()
B, H, W = 32, 480, 640
img = torch.randint(0, 256, (B, 6, H, W), dtype=torch.uint8, device=dev)
enc = torch.rand(B, 3, H, W, device=dev)
x = ops.pack_input(img, enc, model.pixel_mean, model.pixel_std)
model.use_cuda_graph = False
for prec in (sys.argv[1:] or ["bf16"]):
    model.set_precision(prec)
    with torch.no_grad():
        for _ in range(3):
            model.network_logits(x)
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            model.network_logits(x)
            torch.cuda.synchronize()
    rows = sorted(prof.key_averages(), key=lambda e: -getattr(e, "device_time_total", 0.0))
    tot = sum(getattr(e, "device_time_total", 0.0) for e in rows)
    print("=====", prec, f"{tot * 1e-3:.3f} ms total device time")
    for e in rows[:28]:
        print(f"{e.device_time_total * 1e-3:8.3f} ms  x{e.count:3d}  {e.key[:120]}")
