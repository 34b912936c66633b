"""Per-kernel device times of the post-processing sequence (and encode / boundary) at a bench.py config, via the torch
profiler (no ncu).  Usage: python scripts/bench_post.py [c2|c3|c4]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from quber_b200 import ops
from quber_b200.pipeline import HotPath, RefinerConfig
name = sys.argv[1] if len(sys.argv) > 1 else "c2"
c, g = bench.CONFIGS[name], bench.geometry(bench.CONFIGS[name])
masks, head, msda = bench.make_inputs(name, 0)
dev = torch.device("cuda")
hp = HotPath(RefinerConfig(compact_label_map=True, crop=g["crop"], valid_w=c["W"] if g["Wp"] != c["W"] else 0), dev)
dm = torch.from_numpy(masks).to(dev)
dh = {k: torch.from_numpy(v).to(dev) for k, v in head.items()}
d_bnd = ops.boundary_dilation(c["H"], c["W"], 0.01)
def step():
    ops.encode_initial_masks(dm, 10, hp.cfg.valid_w)
    ops.mask_boundary(dm, d_bnd)
    return hp.post(dh)
import ctypes
from quber_b200 import _lib
variants = [None]
if hasattr(_lib.lib(), "quber_debug_post_variant"):
    variants = [0, 1, 2, 3]
for var in variants:
    if var is not None:
        ctypes.CDLL(_lib.SO_PATH).quber_debug_post_variant(var)
        print(f"--- variant {var}: split_eee={var & 1} cta_prune={(var >> 1) & 1}")
    for _ in range(3):
        r = step()
    torch.cuda.synchronize()
    N = 10
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(N):
            step()
        torch.cuda.synchronize()
    tot = 0.0
    for ev in sorted(prof.key_averages(), key=lambda e: -getattr(e, "device_time_total", 0)):
        t = getattr(ev, "device_time_total", 0) / N
        if t < 0.5:
            continue
        tot += t
        print(f"{ev.key[:70]:70s} {t:8.1f} us / step  ({ev.count // N} launches)")
    print(f"sum {tot:.1f} us; instances per step {int(r['inst_offsets'][-1])}")
