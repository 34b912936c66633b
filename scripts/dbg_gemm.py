"""Debug helper: quber_gemm_bias_act against torch on a few shapes (run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from quber_b200 import ops, _lib
torch.manual_seed(0)
def tf32(x):
    return (x.view(torch.int32) & ~0x1fff).view(torch.float32)
cases = [(128, 128, 64), (128, 64, 64), (256, 128, 256), (300, 256, 256), (1000, 192, 256), (1000, 96, 256), (515, 1024, 256),
         (515, 256, 1024), (201600, 256, 256), (77, 40, 168), (1000, 4, 32), (333, 100, 72)]
for dt in ("bf16", "tf32"):
    for (M, N, K) in cases:
        a = torch.randn(M, K, device="cuda"); w = torch.randn(N, K, device="cuda") / K ** 0.5; b = torch.randn(N, device="cuda")
        if dt == "bf16":
            a, w = a.bfloat16(), w.bfloat16()
        else:
            a, w = tf32(a), tf32(w)
        ref = torch.relu(a.double() @ w.double().T + b.double()).float()
        out = ops.gemm_bias_act(a, w, b, act="relu", out_dtype=torch.float32)
        ob = ops.gemm_bias_act(a, w, b, act="relu", out_dtype=torch.bfloat16) if N % 8 == 0 else None
        torch.cuda.synchronize()
        err = (out - ref).abs().max().item()
        errb = (ob.float() - ref).abs().max().item() if ob is not None else -1
        print(dt, (M, N, K), "max abs err", f"{err:.3e}", f"bf16 out {errb:.3e}", "timeouts", _lib.lib().quber_gemm_timeouts(), flush=True)
# timing at the MSDA value_proj shape
for dt in ("bf16", "tf32"):
    M, N, K = 201600, 256, 256
    a = torch.randn(M, K, device="cuda"); w = torch.randn(N, K, device="cuda"); b = torch.randn(N, device="cuda")
    if dt == "bf16":
        a, w = a.bfloat16(), w.bfloat16()
    for od in (torch.float32, torch.bfloat16):
        for _ in range(3):
            ops.gemm_bias_act(a, w, b, out_dtype=od)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            ops.gemm_bias_act(a, w, b, out_dtype=od)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 20 * 1e3
        byt = M * K * a.element_size() + N * K * a.element_size() + M * N * (4 if od == torch.float32 else 2)
        print(f"{dt}->{od}: {us:.1f} us, {2*M*N*K/us/1e6:.1f} TFLOP/s, {byt/us/1e3:.0f} GB/s")
        af, wf = a.float(), w.float()
        for _ in range(3):
            torch.nn.functional.linear(a, w)
        e0.record()
        for _ in range(20):
            torch.nn.functional.linear(a, w)
        e1.record(); torch.cuda.synchronize()
        print(f"   torch linear same dtype (no bias): {e0.elapsed_time(e1)/20*1e3:.1f} us")
