"""Timeline of one host-buffer step as bench.py runs it (MSDA first, then the chunk pipeline): GPU timestamps from events
around every chunk's kernels, host timestamps around every wait for the packing."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from quber_b200 import ops, synth
from quber_b200.pipeline import HotPath, RefinerConfig

B, K, H, W = 32, 30, 480, 640
CH = int(sys.argv[1]) if len(sys.argv) > 1 else 16
TAIL = int(sys.argv[2]) if len(sys.argv) > 2 else 2
sets = []
for s in range(2):
    masks = np.stack([synth.ellipse_masks(K, H, W, seed=10 * s + b % 4) for b in range(B)])
    heads = [synth.head_outputs(masks[b], seed=s + b % 4) for b in range(B)]
    head = {k: np.stack([h[k] for h in heads]) for k in heads[0]}
    inp = synth.msda_inputs(B=B, mode="module", seed=s)
    msda = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in inp.items() if k in ("value", "shapes", "starts", "loc", "attn")}
    sets.append((torch.from_numpy(masks).pin_memory(), {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}, msda))
hp = HotPath(RefinerConfig(), "cuda")
def step(i):
    hm, hh, m = sets[i % 2]
    for _ in range(6):
        ops.msda_forward(m["value"], m["shapes"], m["starts"], m["loc"], m["attn"])
    return hp.refine_batch_host(hm, hh, chunk_frames=CH, min_tail=TAIL)
for i in range(4):
    step(i); torch.cuda.synchronize()
ev, host = [], []
orig_run, orig_wait = hp._run_graphed, ops.host_pack_wait
def run(st, key, body):
    a = torch.cuda.Event(enable_timing=True); a.record()
    r = orig_run(st, key, body)
    b = torch.cuda.Event(enable_timing=True); b.record()
    ev.append((key, a, b)); return r
def wait(c):
    t0 = time.perf_counter(); r = orig_wait(c); host.append((c, t0, time.perf_counter())); return r
hp._run_graphed, ops.host_pack_wait = run, wait
for rep in range(3):
    ev.clear(); host.clear()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e0.record(); t0 = time.perf_counter()
    hm, hh, m = sets[rep % 2]
    for _ in range(6):
        ops.msda_forward(m["value"], m["shapes"], m["starts"], m["loc"], m["attn"])
    em = torch.cuda.Event(enable_timing=True); em.record()
    t_msda_enq = time.perf_counter() - t0
    hp.refine_batch_host(hm, hh, chunk_frames=CH, min_tail=TAIL)
    t_ret = time.perf_counter() - t0
    e1 = torch.cuda.Event(enable_timing=True); e1.record()
    torch.cuda.current_stream().synchronize()
    t_sync = time.perf_counter() - t0
    print(f"rep {rep}: msda enqueued {t_msda_enq * 1e3:.2f} ms, call returned {t_ret * 1e3:.2f} ms, synchronised {t_sync * 1e3:.2f} ms; "
          f"GPU: msda done {e0.elapsed_time(em):.2f} ms, all done {e0.elapsed_time(e1):.2f} ms")
    for key, a, b in ev:
        print(f"   graph {str(key):10s}: GPU {e0.elapsed_time(a):.2f} -> {e0.elapsed_time(b):.2f} ms")
    for c, w0, w1 in host:
        print(f"   pack chunk {c}: host waited {(w0 - t0) * 1e3:.2f} -> {(w1 - t0) * 1e3:.2f} ms")
