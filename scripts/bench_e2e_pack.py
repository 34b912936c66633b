"""End-to-end host call with and without the packed mask transfer, with the host-side pack time broken out."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from quber_b200 import ops, synth
from quber_b200.pipeline import HotPath, RefinerConfig

B, K, H, W = 32, 30, 480, 640
NS = 3
sets = []
for s in range(NS):
    masks = np.stack([synth.ellipse_masks(K, H, W, seed=10 * s + b % 4) for b in range(B)])
    heads = [synth.head_outputs(masks[b], seed=s + b % 4) for b in range(B)]
    head = {k: np.stack([h[k] for h in heads]) for k in heads[0]}
    sets.append((torch.from_numpy(masks).pin_memory(), {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}))
orig = ops.host_pack_wait
acc = [0.0, 0]
def timed(*a, **k):                        # time this thread spends blocked on the packing
    t = time.perf_counter(); r = orig(*a, **k); acc[0] += time.perf_counter() - t; acc[1] += 1; return r
ops.host_pack_wait = timed
for graphs in (False, True):
  print("use_graphs =", graphs)
  for pack, th, cf in [(False, 0, 8), (True, 12, 8), (None, 12, 8), (True, 14, 8)]:
    hp = HotPath(RefinerConfig(), "cuda")
    hp.use_graphs = graphs
    hp.pack_masks, hp.pack_threads = pack, th
    for i in range(3):
        hp.refine_batch_host(*sets[i % NS], chunk_frames=cf); torch.cuda.synchronize()
    acc[0], acc[1] = 0.0, 0
    t = time.perf_counter()
    enq = 0.0
    for i in range(12):
        t0 = time.perf_counter()
        hp.refine_batch_host(*sets[i % NS], chunk_frames=cf)
        enq += time.perf_counter() - t0
        torch.cuda.current_stream().synchronize()
    dt = (time.perf_counter() - t) / 12
    print(f"pack={pack} threads={th} chunk={cf}: {dt * 1e3:.2f} ms per 32 frames (call returns after {enq / 12 * 1e3:.2f} ms); "
          f"blocked on pack {acc[0] / 12 * 1e3:.2f} ms in {acc[1] // 12} calls; decided={hp.pack_masks} {getattr(hp, 'pack_calibration', None)}")
