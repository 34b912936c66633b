"""Micro-benchmark of quber_msda_forward at the C2 shapes (B=32, S=Q=6300, M=8, D=32, L=3, P=4)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from quber_b200 import ops, synth
import oracle
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dt = sys.argv[2] if len(sys.argv) > 2 else "f32"
sets = []
for s in range(2):
    inp = synth.msda_inputs(B=B, mode="module", seed=s)
    d = {k: torch.from_numpy(v).cuda() for k, v in inp.items()}
    if dt == "bf16":
        d["value"] = d["value"].bfloat16()
    sets.append((inp, d))

def run(d):
    return ops.msda_forward(d["value"], d["shapes"], d["starts"], d["loc"], d["attn"])
for i in range(4):
    out = run(sets[i % 2][1])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
N = 20
e0.record()
for i in range(N):
    out = run(sets[i % 2][1])
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / N * 1e3
S = 6300
bytes_ = B * (S * 256 * (4 if dt == "f32" else 2) * 2 + S * 8 * 12 * 2 * 4 + S * 8 * 12 * 4)
print(f"msda {dt} B={B}: {us:.1f} us/launch, {bytes_ / us / 1e3:.0f} GB/s algorithmic")
# fused variant: softmax + ref + off / (W,H) inside the set-up phase, reading the qk GEMM's rows (ms_deform_attn.py:102-117)
M_, L_, P_ = 8, 3, 4
g = torch.Generator(device="cuda").manual_seed(0)
fsets = []
for s in range(2):
    d = sets[s][1]
    oa = torch.randn(B, S, M_ * L_ * P_ * 3, device="cuda", generator=g)
    rp = torch.rand(B, S, L_, 2, device="cuda", generator=g)
    fsets.append((d, oa, rp))
def runf(t):
    d, oa, rp = t
    return ops.msda_forward_fused(d["value"], d["shapes"], d["starts"], oa, rp, L_, P_)
def runp(t):
    d, oa, rp = t
    loc, aw = ops.msda_prepare(oa, rp, d["shapes"], M_, L_, P_)
    return ops.msda_forward(d["value"], d["shapes"], d["starts"], loc, aw)
for name, fn in (("fused (one kernel)", runf), ("prepare + forward (two kernels)", runp)):
    for i in range(4):
        fn(fsets[i % 2])
    torch.cuda.synchronize()
    e0.record()
    for i in range(N):
        fn(fsets[i % 2])
    e1.record(); torch.cuda.synchronize()
    print(f"msda {dt} B={B} {name}: {e0.elapsed_time(e1) / N * 1e3:.1f} us/launch")
inp, d = sets[(N - 1) % 2]
sel = np.r_[0:150, 4700:4900, 6200:6300]
ref = oracle.msda_forward(inp["value"][:1], inp["shapes"], inp["starts"], inp["loc"][:1, sel], inp["attn"][:1, sel])
err = np.abs(out[0, sel].float().cpu().numpy() - ref).max()
print("max abs err vs oracle (450 queries over 3 levels):", err)
