"""Which call sites of the network forward launch torch's copy / cat kernels (bf16 plan)."""
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
model.set_precision("bf16")
with torch.no_grad():
    for _ in range(3):
        model.network_logits(x)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True, record_shapes=True) as prof:
        model.network_logits(x)
        torch.cuda.synchronize()
rows = [e for e in prof.key_averages(group_by_stack_n=6, group_by_input_shape=True)
        if e.key in ("aten::copy_", "aten::cat", "aten::_to_copy", "aten::contiguous", "aten::clone", "aten::add", "aten::mul")]
rows.sort(key=lambda e: -e.device_time_total)
for e in rows[:16]:
    print(f"{e.device_time_total * 1e-3:7.3f} ms x{e.count:3d} {e.key} {str(e.input_shapes)[:90]}")
    for fr in e.stack[:6]:
        if "quber_b200" in fr or "scripts" in fr:
            print("        ", fr[-110:])
