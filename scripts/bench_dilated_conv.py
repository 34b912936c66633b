"""cuDNN 3x3 dilated convolution, bf16 / fp32, channels-last vs NCHW operands, at the path's res5 / ASPP shapes."""
import torch, torch.nn.functional as F
def t(f, n=5):
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for dt in (torch.bfloat16, torch.float32):
    for (cin, cout) in ((512, 512), (2048, 256)):
        for d in (1, 2, 4, 6, 8, 12, 18):
            x = torch.randn(32, cin, 30, 40, device="cuda", dtype=dt)
            w = torch.randn(cout, cin, 3, 3, device="cuda", dtype=dt) * 0.01
            b = torch.zeros(cout, device="cuda", dtype=dt)
            xc, wc = x.contiguous(memory_format=torch.channels_last), w.contiguous(memory_format=torch.channels_last)
            r = {}
            for name, (xx, ww) in {"nchw": (x, w), "cl": (xc, wc)}.items():
                r[name] = t(lambda: F.conv2d(xx, ww, None, 1, d, d))
                r[name + "_relu"] = t(lambda: torch.cudnn_convolution_relu(xx, ww, b, (1, 1), (d, d), (d, d), 1))
            print(str(dt)[6:], cin, cout, "dil", d, " ".join(f"{k}={v:.3f}ms" for k, v in r.items()), flush=True)
