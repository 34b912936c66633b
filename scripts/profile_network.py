"""Per-kernel breakdown of the network stage with torch.profiler: which cuDNN / elementwise
kernels the 32-frame forward spends its time in.  Usage: python scripts/profile_network.py [fp32|bf16] [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from quber_b200 import config as qcfg
from quber_b200.meta_arch import build_model
prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
cfg = qcfg.get_cfg()
cfg.merge_from_file(os.path.join(ROOT, "configs", "mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8.yaml"))
cfg.MODEL.DEVICE = "cuda"
torch.manual_seed(0)
model = build_model(cfg, prec).eval()
model.set_precision(prec)
x = torch.randn(B, 9, 480, 640, device="cuda")
with torch.no_grad():
    for _ in range(2):
        model.network_logits(x)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        model.network_logits(x)
        torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=70))
