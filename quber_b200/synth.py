"""Seeded synthetic inputs for the maskrefiner hot path.

Pure numpy; used by tests/, bench.py and oracle/gen_golden.py so that every
leg sees the same bytes.  Nothing here is on the product's compute path.
"""
import numpy as np


def ellipse_masks(K, H=480, W=640, seed=0, axes=(20, 60), margin=60, value=255, empty_tail=0):
    """K filled ellipses as [K,H,W] uint8 (0 / value); `empty_tail` all-zero masks appended."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    out = np.zeros((K + empty_tail, H, W), np.uint8)
    my, mx = min(margin, H // 4), min(margin, W // 4)
    for k in range(K):
        cy, cx = rng.uniform(my, H - my), rng.uniform(mx, W - mx)
        ay, ax = rng.uniform(*axes), rng.uniform(*axes)
        th = rng.uniform(0, np.pi)
        dy, dx = yy - cy, xx - cx
        u = dx * np.cos(th) + dy * np.sin(th)
        v = -dx * np.sin(th) + dy * np.cos(th)
        out[k][(u / ax) ** 2 + (v / ay) ** 2 <= 1.0] = value
    return out


def perturb_masks(masks, seed=1, max_shift=4):
    """Jittered copy of `masks` (integer shifts), the "initial" masks of a quadruple-map test."""
    rng = np.random.default_rng(seed)
    out = np.zeros_like(masks)
    for k, m in enumerate(masks):
        dy, dx = rng.integers(-max_shift, max_shift + 1, 2)
        out[k] = np.roll(np.roll(m, dy, 0), dx, 1)
    return out


def rgbd_frame(H=480, W=640, seed=0):
    """RGB uint8 U[0,255] and a 3-channel uint8 depth image (normalised U[250,1500] mm, 5 % holes)."""
    rng = np.random.default_rng(seed)
    rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    d = rng.uniform(250, 1500, (H, W)).astype(np.float32)
    d[rng.random((H, W)) < 0.05] = 0
    dn = np.clip((d - 250.0) / (1500.0 - 250.0) * 255.0, 0, 255).astype(np.uint8)
    dn[d == 0] = 0
    return rgb, np.repeat(dn[:, :, None], 3, axis=2)


def head_outputs(masks, stride=4, seed=0, eee_channels=4, sigma=10):
    """Synthetic 1/stride-resolution head outputs whose post-processing recovers `masks`.

    Returns dict of float32 arrays: foreground [1,h,w], center [1,h,w], offset [2,h,w] (in
    low-res pixel units, i.e. true pixel offset / stride, model.py:698 multiplies it back),
    eee_boundary [C,h,w].
    """
    rng = np.random.default_rng(seed)
    K, H, W = masks.shape
    h, w = H // stride, W // stride
    ys = (np.arange(h) * stride + (stride - 1) / 2.0).astype(np.float32)
    xs = (np.arange(w) * stride + (stride - 1) / 2.0).astype(np.float32)
    center = np.zeros((h, w), np.float32)
    off = np.zeros((2, h, w), np.float32)
    union = np.zeros((h, w), bool)
    sub = masks[:, stride // 2::stride, stride // 2::stride][:, :h, :w] != 0
    for k in range(K):
        idx = np.nonzero(masks[k])
        if len(idx[0]) == 0:
            continue
        cy, cx = idx[0].mean(), idx[1].mean()
        g = np.exp(-((ys[:, None] - cy) ** 2 + (xs[None, :] - cx) ** 2) / (2.0 * sigma ** 2)).astype(np.float32)
        center = np.maximum(center, g)
        m = sub[k]
        off[0][m] = ((cy - ys[:, None]) / stride).astype(np.float32).repeat(w, 1)[m]
        off[1][m] = ((cx - xs[None, :]) / stride).astype(np.float32).repeat(h, 0)[m]
        union |= m
    center = center + rng.normal(0, 0.02, (h, w)).astype(np.float32)
    off = off + rng.normal(0, 0.5 / stride, (2, h, w)).astype(np.float32)
    fg = np.where(union, 6.0, -6.0).astype(np.float32) + rng.normal(0, 1.0, (h, w)).astype(np.float32)
    eee = rng.normal(0, 1.0, (eee_channels, h, w)).astype(np.float32)
    # crude class map: boundary ring of the union = channel 0 (TP), interior/background = channel 1 (TN)
    ring = union & ~(np.roll(union, 1, 0) & np.roll(union, -1, 0) & np.roll(union, 1, 1) & np.roll(union, -1, 1))
    cls = np.where(ring, 0, 1)
    cls[rng.random((h, w)) < 0.05] = min(2, eee_channels - 1)
    cls[rng.random((h, w)) < 0.05] = eee_channels - 1
    for c in range(eee_channels):
        eee[c][cls == c] += 4.0
    return dict(foreground=fg[None], center=center[None], offset=off, eee_boundary=eee)


def msda_inputs(B=1, shapes=((60, 80), (30, 40), (15, 20)), M=8, D=32, P=4, Q=None, seed=0,
                mode="rand", dtype=np.float32, operands=False):
    """MSDeformAttn inputs.  mode "rand": OPS/test.py:36-39 (value=rand*0.01, loc=rand, attn=rand+1e-5
    normalised over L*P); "border": locs in [-0.1, 1.1]; "module": reference points + the module's
    initial sampling-offset grid (OPS/modules/ms_deform_attn.py:66-74), i.e. what a random-init
    encoder layer actually feeds the op (Q = S).  operands=True (mode "module" only) also returns what the MODULE
    computes them from: `oa` [B,Q,M*L*P*3] = the sampling_offsets | attention_weights projection rows (offsets in
    pixels, logits = log(attn) so that their softmax is attn) and `ref` [B,Q,L,2] reference points — the inputs of
    quber_msda_forward_fused; loc / attn are then what ms_deform_attn.py:104-109 makes of them (up to rounding)."""
    rng = np.random.default_rng(seed)
    shapes = np.asarray(shapes, np.int64)
    L = len(shapes)
    S = int((shapes[:, 0] * shapes[:, 1]).sum())
    starts = np.concatenate([[0], np.cumsum(shapes[:, 0] * shapes[:, 1])[:-1]]).astype(np.int64)
    if mode == "module":
        Q = S
    elif Q is None:
        Q = S
    value = (rng.random((B, S, M, D)) * 0.01).astype(dtype)
    attn = rng.random((B, Q, M, L, P)).astype(dtype) + dtype(1e-5)
    attn /= attn.sum((-1, -2), keepdims=True)
    if mode == "rand":
        loc = rng.random((B, Q, M, L, P, 2)).astype(dtype)
    elif mode == "border":
        loc = (rng.random((B, Q, M, L, P, 2)) * 1.2 - 0.1).astype(dtype)
    elif mode == "module":
        ref = []
        for (h, w) in shapes:
            ry, rx = np.meshgrid((np.arange(h) + 0.5) / h, (np.arange(w) + 0.5) / w, indexing="ij")
            ref.append(np.stack([rx.ravel(), ry.ravel()], -1))
        ref = np.concatenate(ref, 0).astype(dtype)                     # [S,2] (x,y)
        th = np.arange(M, dtype=np.float32) * (2.0 * np.pi / M)
        grid = np.stack([np.cos(th), np.sin(th)], -1)
        grid = grid / np.abs(grid).max(-1, keepdims=True)              # [M,2]
        offs = grid[:, None, None, :] * (np.arange(P, dtype=np.float32) + 1)[None, None, :, None]
        offs = np.broadcast_to(offs, (M, L, P, 2)) + rng.normal(0, 0.25, (B, Q, M, L, P, 2))
        norm = np.stack([shapes[:, 1], shapes[:, 0]], -1).astype(np.float32)   # (W_l, H_l)
        loc = (ref[None, :, None, None, None, :] + offs / norm[None, None, None, :, None, :]).astype(dtype)
        if operands:
            oa = np.concatenate([offs.astype(np.float32).reshape(B, Q, M * L * P * 2),
                                 np.log(attn.astype(np.float64)).astype(np.float32).reshape(B, Q, M * L * P)], -1)
            refp = np.ascontiguousarray(np.broadcast_to(ref[None, :, None, :], (B, Q, L, 2)).astype(np.float32))
            return dict(value=value, shapes=shapes, starts=starts, loc=np.ascontiguousarray(loc), attn=attn,
                        oa=np.ascontiguousarray(oa), ref=refp)
    else:
        raise ValueError(mode)
    return dict(value=value, shapes=shapes, starts=starts, loc=np.ascontiguousarray(loc), attn=attn)
