"""Host bit-packing rate (quber_host_pack_masks) against the plain pinned H2D copy it would replace.
Buffers are cycled (8 x 74 MB) so that every pass reads DRAM, not the last-level cache."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from quber_b200 import ops

NB = 8
bufs = [(torch.rand(8, 30, 480, 640) < 0.3).to(torch.uint8).pin_memory() for _ in range(NB)]
m = bufs[0]
out = torch.empty(m.numel() // 32, dtype=torch.int32).pin_memory()
print("affinity cores:", len(os.sched_getaffinity(0)))
for th in (1, 4, 8, 12, 16, 24, 32):
    ops.host_pack_masks(m, out=out, threads=th)
    t = time.perf_counter()
    for i in range(2 * NB):
        ops.host_pack_masks(bufs[i % NB], out=out, threads=th)
    dt = (time.perf_counter() - t) / (2 * NB)
    print(f"pack threads={th:2d} (cold, {NB} buffers cycled): {m.numel() / dt / 1e9:6.1f} GB/s  ({dt * 1e3:.2f} ms per 8 frames)")
if torch.cuda.is_available():
    d = torch.empty_like(m, device="cuda")
    d.copy_(m, non_blocking=True); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(2 * NB):
        d.copy_(bufs[i % NB], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(f"pinned H2D: {2 * NB * m.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9:.1f} GB/s")
    # both at once: the copy engine reading host memory while the cores pack the next chunk
    t = time.perf_counter()
    e0.record()
    for i in range(2 * NB):
        d.copy_(bufs[i % NB], non_blocking=True)
        ops.host_pack_masks(bufs[(i + 1) % NB], out=out, threads=16)
    e1.record(); torch.cuda.synchronize()
    dt = time.perf_counter() - t
    print(f"H2D of chunk i while packing chunk i+1 (16 threads): {dt / (2 * NB) * 1e3:.2f} ms per pair of 8 frames")
    dp = out.cuda()
    u = ops.unpack_masks(dp, m.shape)
    _, val = ops.host_pack_masks(m, out=out, threads=8)
    assert torch.equal(ops.unpack_masks(out.cuda(), m.shape, value=val).cpu(), m)
    e0.record()
    for _ in range(10):
        u = ops.unpack_masks(dp, m.shape)
    e1.record(); torch.cuda.synchronize()
    print(f"unpack kernel: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us per 8 frames")
