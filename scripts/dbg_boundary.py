"""Debug helper: one small quber_mask_boundary call (run under compute-sanitizer on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oracle
from quber_b200 import ops, synth
H, W, K = 120, 160, 3
d = int(sys.argv[1]) if len(sys.argv) > 1 else 2
m = synth.ellipse_masks(K, H, W, seed=1, axes=(10, 30), margin=20)
per, uni, fg = ops.mask_boundary(torch.from_numpy(m).cuda()[None], d)
torch.cuda.synchronize()
u, p = None, np.stack([np.zeros_like(m[0])] * K)
import ctypes
for k in range(K):
    oracle.lib().oracle_mask_to_boundary(m[k].ctypes.data_as(ctypes.c_void_p), H, W, d, p[k].ctypes.data_as(ctypes.c_void_p))
print("per equal:", np.array_equal(per[0].cpu().numpy(), p), "fg sum", int(fg.sum()))
