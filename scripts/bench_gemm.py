"""quber_gemm_bias_act (tcgen05 / TMEM) at the path's GEMM shapes for 32 frames, next to cuBLAS (torch.nn.functional.linear).
rows = 32 * 6300 (encoder) or 32 * 19200 (FusionLayers 1x1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from quber_b200 import ops
def t(f, n=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
cases = [("value_proj / output_proj", 201600, 256, 256), ("offsets|weights", 201600, 288, 256), ("ffn linear1 (+ReLU)", 201600, 1024, 256),
         ("ffn linear2", 201600, 256, 1024), ("FusionLayers 1x1 (+BN+ReLU)", 614400, 128, 168)]
for name, M, N, K in cases:
    for dt, od in ((torch.bfloat16, torch.bfloat16), (torch.bfloat16, torch.float32), (torch.float32, torch.float32)):
        a = torch.randn(M, K, device="cuda").to(dt); a2 = torch.randn(M, K, device="cuda").to(dt)
        w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dt); b = torch.randn(N, device="cuda")
        it = [0]
        def f():
            it[0] ^= 1
            ops.gemm_bias_act(a if it[0] else a2, w, b, act="relu", out_dtype=od)
        us = t(f)
        byt = M * K * a.element_size() + N * K * a.element_size() + M * N * (4 if od == torch.float32 else 2)
        us_t = t(lambda: torch.relu_(torch.nn.functional.linear(a, w, b.to(dt))))
        kind = ("bf16" if dt == torch.bfloat16 else "tf32") + "->" + ("bf16" if od == torch.bfloat16 else "f32")
        print(f"{name:30s} {M}x{N}x{K} {kind}: {us:7.1f} us  {2*M*N*K/us/1e6:6.1f} TFLOP/s  {byt/us/1e3:5.0f} GB/s   cuBLAS linear+relu_: {us_t:7.1f} us", flush=True)

# fp32-accurate split-operand mode (3xTF32) against cuBLAS sgemm (TF32 disabled)
torch.backends.cuda.matmul.allow_tf32 = False
for name, M, N, K in cases[:4]:
    a = torch.randn(M, K, device="cuda"); a2 = torch.randn(M, K, device="cuda")
    w = torch.randn(N, K, device="cuda") / K ** 0.5; b = torch.randn(N, device="cuda")
    hi, lo = ops.split_tf32(w)
    it = [0]
    def f():
        it[0] ^= 1
        ops.gemm_bias_act_x3(a if it[0] else a2, hi, lo, b, act="relu")
    us = t(f)
    us_t = t(lambda: torch.relu_(torch.nn.functional.linear(a, w, b)))
    byt = M * K * 4 + 2 * N * K * 4 + M * N * 4
    print(f"{name:30s} {M}x{N}x{K} 3xTF32->f32: {us:7.1f} us  {2*M*N*K/us/1e6:6.1f} TFLOP/s (x3 issued)  {byt/us/1e3:5.0f} GB/s   cuBLAS sgemm+relu_: {us_t:7.1f} us", flush=True)
