"""Device-side operators: torch CUDA tensors in, torch CUDA tensors out, computed by the hand-written
sm_100a kernels of libquber_b200.so through its C ABI (include/quber_b200.h).

PyTorch is plumbing only here (device memory, the current stream); there is no PyTorch or CPU
implementation of any of these ops in the product.  They are also registered as `torch.ops.quber.*`
custom ops at the bottom of the file.
"""
import ctypes

import numpy as np
import os

import torch

from . import _lib

_GAUSS = {}


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _need_cuda(t, name, dtype=None):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        # same condition the reference asserts with AT_ASSERTM("... must be a CUDA tensor"), ms_deform_attn_cuda.cu:39-43
        raise ValueError(f"{name} must be a CUDA tensor")
    if dtype is not None and t.dtype != dtype:
        raise ValueError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} tensor has to be contiguous")   # ms_deform_attn_cuda.cu:33-37
    return t


def gaussian_table(sigma, device):
    key = (sigma, str(device))
    if key not in _GAUSS:
        n = 6 * sigma + 3
        host = np.empty((n, n), np.float32)
        _lib.lib().quber_gaussian_table(sigma, host.ctypes.data_as(ctypes.c_void_p))
        _GAUSS[key] = torch.from_numpy(host).to(device)
    return _GAUSS[key]


def boundary_dilation(H, W, ratio):
    """explicit_error_estimation/util.py:80-84."""
    return int(_lib.lib().quber_boundary_dilation(int(H), int(W), float(ratio)))


OCC_TILE = (16, 128)     # QUBER_OCC_TILE_H, QUBER_OCC_TILE_W


def _occupancy(B, K, H, W, device):
    th, tw = OCC_TILE
    return torch.empty((B, K, (H + th - 1) // th, (W + tw - 1) // tw), dtype=torch.uint8, device=device)


def encode_initial_masks(masks, sigma=10, valid_w=0, with_occupancy=False):
    """maskrefiner/predictor.py:304-357 for a batch: masks [B,K,H,W] uint8 -> [B,3,H,W] float32.
    valid_w (0 = W): the frames' own width when W is a zero-padded row pitch (see encode_initial_masks_any_width).
    with_occupancy: also return the tile occupancy [B,K,ceil(H/16),ceil(W/128)] uint8 (1 = the tile holds a set pixel of
    the mask), a by-product of the pass over the mask bytes that mask_boundary(occupancy=) uses to skip empty tiles."""
    _need_cuda(masks, "masks", torch.uint8)
    B, K, H, W = masks.shape
    L = _lib.lib()
    out = torch.empty((B, 3, H, W), dtype=torch.float32, device=masks.device)
    occ = _occupancy(B, K, H, W, masks.device) if with_occupancy else None
    nbytes = L.quber_encode_workspace_bytes(B, K, H, W)
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=masks.device)
    g = gaussian_table(sigma, masks.device)
    with torch.cuda.device(masks.device):
        _lib.check(L.quber_encode_initial_masks(_ptr(masks), B, K, H, W, int(valid_w), sigma, _ptr(g), _ptr(out), _ptr(occ), _ptr(ws),
                                                nbytes, _stream()), "quber_encode_initial_masks")
    return (out, occ) if with_occupancy else out


def encode_initial_masks_packed(packed, value, masks_out, sigma=10, valid_w=0, with_occupancy=False, shape=None):
    """encode_initial_masks from bit-packed masks (host_pack_masks / rle_decode): packed int32 CUDA [B*K*H*W / 32], set
    pixels carry `value`; masks_out uint8 CUDA [B,K,H,W] receives the byte planes (0 / value) on the way, or None with
    shape=(B,K,H,W) when nobody needs them (mask_boundary_bits works from the bits) -> [B,3,H,W] float32
    (and the tile occupancy with with_occupancy, see encode_initial_masks)."""
    _need_cuda(packed, "packed", torch.int32)
    if masks_out is not None:
        _need_cuda(masks_out, "masks_out", torch.uint8)
        shape = tuple(masks_out.shape)
    elif shape is None:
        raise ValueError("encode_initial_masks_packed: give masks_out or shape")
    B, K, H, W = shape
    if packed.numel() * 32 != B * K * H * W:
        raise ValueError("packed words do not match the mask shape")
    L = _lib.lib()
    out = torch.empty((B, 3, H, W), dtype=torch.float32, device=packed.device)
    occ = _occupancy(B, K, H, W, packed.device) if with_occupancy else None
    nbytes = L.quber_encode_workspace_bytes(B, K, H, W)
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=packed.device)
    g = gaussian_table(sigma, packed.device)
    with torch.cuda.device(packed.device):
        _lib.check(L.quber_encode_initial_masks_packed(_ptr(packed), int(value), _ptr(masks_out), B, K, H, W, int(valid_w), sigma, _ptr(g), _ptr(out),
                                                       _ptr(occ), _ptr(ws), nbytes, _stream()), "quber_encode_initial_masks_packed")
    return (out, occ) if with_occupancy else out


def encode_initial_masks_any_width(masks, sigma=10):
    """encode_initial_masks for frames whose width is not a multiple of 16 (armbench 750x1333): zero columns are
    appended, which changes neither a centroid nor a paste inside the frame, and the kernel normalises the x-offset
    plane by the frame's own width (valid_w) — bit-exact with the reference on the unpadded frame."""
    B, K, H, W = masks.shape
    pad = (-W) % 16
    if not pad:
        return encode_initial_masks(masks, sigma)
    return encode_initial_masks(torch.nn.functional.pad(masks, (0, pad)).contiguous(), sigma, valid_w=W)[..., :W].contiguous()


_BOUNDARY_STATUS = {}


def boundary_status(device=None, check=False):
    """The per-call status word of the LAST mask_boundary call on `device` (int32 CUDA [1]; bit 0 = a TMA tile load never
    arrived and that call's outputs are invalid).  check=True synchronises and raises instead of returning."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    st = _BOUNDARY_STATUS.get(dev)
    if check and st is not None and int(st[0]) & 1:
        raise RuntimeError("quber_mask_boundary: a TMA tile load timed out; the boundary outputs of that call are invalid")
    return st


def mask_boundary(masks, d, per_instance=True, union=True, fg=True, status=None, occupancy=None):
    """explicit_error_estimation/util.py:62-99 for a batch: masks [B,K,H,W] uint8 ->
    (per-instance boundary [B,K,H,W] | None, union boundary [B,H,W] | None, fg [B,H,W] | None).
    occupancy: the tile occupancy encode_initial_masks(with_occupancy=True) returned for the SAME masks — planes with no
    pixel near a tile are then skipped there (same results, less to read).
    The call's device status word (see boundary_status) is written to `status` (int32 CUDA [1]) or to a per-device
    word that boundary_status() returns; nothing is synchronised here."""
    _need_cuda(masks, "masks", torch.uint8)
    B, K, H, W = masks.shape
    per = torch.empty_like(masks) if per_instance else None
    uni = torch.empty((B, H, W), dtype=torch.uint8, device=masks.device) if union else None
    fgm = torch.empty((B, H, W), dtype=torch.uint8, device=masks.device) if fg else None
    if status is None:
        status = _BOUNDARY_STATUS.get(masks.device)
        if status is None:
            status = _BOUNDARY_STATUS[masks.device] = torch.zeros(1, dtype=torch.int32, device=masks.device)
    else:
        _need_cuda(status, "status", torch.int32)
    if occupancy is not None:
        _need_cuda(occupancy, "occupancy", torch.uint8)
        if tuple(occupancy.shape) != tuple(_occupancy(B, K, H, W, "meta").shape) or not occupancy.is_contiguous():
            raise ValueError("occupancy does not belong to masks of this shape")
    with torch.cuda.device(masks.device):
        _lib.check(_lib.lib().quber_mask_boundary(_ptr(masks), B, K, H, W, int(d), _ptr(per), _ptr(uni), _ptr(fgm),
                                                  _ptr(occupancy), _ptr(status), _stream()), "quber_mask_boundary")
    return per, uni, fgm


def mask_boundary_bits(packed, value, shape, d, per_instance=True, union=True, fg=True, occupancy=None):
    """mask_boundary from bit planes (packed int32 CUDA [B*K*H*W / 32], set pixels carry `value` in the outputs; W % 32
    == 0): the same three outputs without byte planes ever existing on the device."""
    _need_cuda(packed, "packed", torch.int32)
    B, K, H, W = shape
    if packed.numel() * 32 != B * K * H * W:
        raise ValueError("packed words do not match the mask shape")
    dev = packed.device
    per = torch.empty((B, K, H, W), dtype=torch.uint8, device=dev) if per_instance else None
    uni = torch.empty((B, H, W), dtype=torch.uint8, device=dev) if union else None
    fgm = torch.empty((B, H, W), dtype=torch.uint8, device=dev) if fg else None
    if occupancy is not None:
        _need_cuda(occupancy, "occupancy", torch.uint8)
        if tuple(occupancy.shape) != tuple(_occupancy(B, K, H, W, "meta").shape) or not occupancy.is_contiguous():
            raise ValueError("occupancy does not belong to masks of this shape")
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().quber_mask_boundary_bits(_ptr(packed), int(value), B, K, H, W, int(d), _ptr(per), _ptr(uni), _ptr(fgm),
                                                       _ptr(occupancy), _stream()), "quber_mask_boundary_bits")
    return per, uni, fgm


def quadruple_maps(gt, pred):
    """tools/ours/panoptic2eee.py:110-123: gt, pred [B,...] uint8 -> [B,4,...] uint8 (TP,TN,FP,FN)."""
    _need_cuda(gt, "gt", torch.uint8)
    _need_cuda(pred, "pred", torch.uint8)
    if gt.shape != pred.shape:
        raise ValueError("gt and pred must have the same shape")
    B = gt.shape[0]
    n = gt[0].numel()
    out = torch.empty((B, 4) + tuple(gt.shape[1:]), dtype=torch.uint8, device=gt.device)
    with torch.cuda.device(gt.device):
        _lib.check(_lib.lib().quber_quadruple_maps(_ptr(gt), _ptr(pred), B, n, _ptr(out), _stream()), "quber_quadruple_maps")
    return out


def post_process(foreground, center, offset, eee=None, *, scale=4, offset_mul=None, threshold=0.3, nms_kernel=7,
                 top_k=200, min_area=512, stuff_area=2048, label_divisor=1000, void_label=-1.0, max_instances=None,
                 want=("sem_seg", "eee_full", "eee_onehot", "panoptic", "label_map", "masks"), centers=None,
                 n_centers=None, no_topk=False, label_dtype=torch.int32, crop=None, label_map_out=None):
    """One fused call for model.py:689-708 + :260-356 + post_processing.py:9-221 + eval_utils.py:308-328.

    foreground [B,1,h,w], center [B,1,h,w], offset [B,2,h,w], eee [B,C,h,w] float32 CUDA tensors at
    1/scale resolution (scale=1: full-resolution maps).  Returns a dict of device tensors; nothing is
    synchronised or copied to the host.  label_dtype=torch.uint8 (QUBER_POST_LABEL_U8, needs top_k <= 255) makes
    `label_map` a byte map: a quarter of the bytes to write, to send to the host and to all-gather.
    crop=(Hc, Wc): the frames are the top-left Hc x Wc corner of the up-sampled maps (sem_seg_postprocess' crop,
    model.py:266-289; armbench 188 x 334 -> 750 x 1333).  Every map output then has Hc rows and a row PITCH of Wc rounded
    up to 8 (out["valid_w"] = Wc; columns past it are void) — slice `[..., :Wc]` for the reference's shape.
    """
    _need_cuda(foreground, "foreground", torch.float32)
    if label_dtype not in (torch.int32, torch.uint8):
        raise ValueError("label_dtype must be torch.int32 or torch.uint8")
    if label_dtype == torch.uint8 and top_k > 255:
        raise ValueError("a uint8 label map needs top_k <= 255")
    _need_cuda(center, "center", torch.float32)
    _need_cuda(offset, "offset", torch.float32)
    if center.dim() != 4 or foreground.dim() != 4:
        raise ValueError("Center prediction with un-supported dimension: {}.".format(center.dim()))   # post_processing.py:199
    if offset.dim() != 4 or offset.shape[1] != 2:
        raise ValueError("Offset prediction with un-supported dimension: {}.".format(offset.dim()))   # :203
    dev = foreground.device
    B, _, h, w = foreground.shape
    H, W = h * scale, w * scale
    in_h = in_w = valid_w = 0
    if crop is not None:
        Hc, Wc = int(crop[0]), int(crop[1])
        if not (0 < Hc <= h * scale and 0 < Wc <= w * scale):
            raise ValueError("crop must lie inside the up-sampled maps")
        H, W, in_h, in_w, valid_w = Hc, (Wc + 7) // 8 * 8, h, w, Wc
    C = 0
    if eee is not None:
        _need_cuda(eee, "eee", torch.float32)
        C = eee.shape[1]
    if offset_mul is None:
        offset_mul = float(scale)          # model.py:698
    if max_instances is None:
        # worst case by construction: an image never has more instances than centres (<= top_k; one stuff paste when
        # there is no centre), so the default pool can not overflow.  The reference returns up to top_k - 1 instances
        # per image (post_processing.py:40-41).  A caller that passes a smaller pool must check out["status"] & 1.
        max_instances = B * max(top_k, 1)
    flags = 0
    p = _lib.PostParams(B, H, W, scale, offset_mul, C, threshold, nms_kernel, top_k, min_area, stuff_area, label_divisor,
                        void_label, max_instances, 0, in_h, in_w, valid_w)
    out = {}
    i32, f32, u8 = torch.int32, torch.float32, torch.uint8
    mk = lambda name, shape, dt: out.setdefault(name, torch.empty(shape, dtype=dt, device=dev)) if name in want else None
    sem = mk("sem_seg", (B, 1, H, W), f32)
    eee_full = mk("eee_full", (B, C, H, W), f32) if C else None
    onehot = mk("eee_onehot", (B, C, H, W), u8) if C else None
    pan = mk("panoptic", (B, H, W), f32)
    if label_map_out is not None and "label_map" in want:
        # caller-owned payload buffer (the all-gather's send buffer): the map is written straight into it
        _need_cuda(label_map_out, "label_map_out", label_dtype)
        if label_map_out.numel() != B * H * W:
            raise ValueError("label_map_out must hold B*H*W elements")
        lmap = out["label_map"] = label_map_out.view(B, H, W)
    else:
        lmap = mk("label_map", (B, H, W), label_dtype)
    if lmap is not None and label_dtype == u8:
        flags |= _lib.POST_LABEL_U8
    iseg = mk("ins_seg", (B, H, W), i32)
    masks = mk("masks", (max(max_instances, 1), H, W), u8)
    if centers is not None:
        flags |= _lib.POST_GIVEN_CENTERS
        out["centers_yx"] = _need_cuda(centers, "centers", i32)
        out["n_centers"] = _need_cuda(n_centers, "n_centers", i32)
        if tuple(centers.shape) != (B, top_k, 2):
            raise ValueError("centers must be [B, top_k, 2] int32")
    else:
        out["centers_yx"] = torch.empty((B, top_k, 2), dtype=i32, device=dev)
        out["n_centers"] = torch.empty((B,), dtype=i32, device=dev)
    if no_topk:
        flags |= _lib.POST_NO_TOPK
    p.flags = flags
    out["inst_offsets"] = torch.empty((B + 1,), dtype=i32, device=dev)
    out["labels"] = torch.empty((max(max_instances, 1),), dtype=i32, device=dev)
    out["scores"] = torch.empty((max(max_instances, 1),), dtype=f32, device=dev)
    out["boxes"] = torch.empty((max(max_instances, 1), 4), dtype=f32, device=dev)
    out["status"] = torch.empty((1,), dtype=i32, device=dev)
    io = _lib.PostIO(*[None if t is None else t.data_ptr() for t in (
        foreground, center, offset, eee, sem, eee_full, onehot, pan, lmap, iseg, masks, out["centers_yx"],
        out["n_centers"], out["inst_offsets"], out["labels"], out["scores"], out["boxes"], out["status"])])
    L = _lib.lib()
    nbytes = L.quber_post_workspace_bytes(ctypes.byref(p))
    if nbytes == 0:
        _lib.check(L.quber_post_process(ctypes.byref(p), ctypes.byref(io), None, 0, _stream()), "quber_post_process")
    ws = torch.empty(nbytes, dtype=u8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.quber_post_process(ctypes.byref(p), ctypes.byref(io), _ptr(ws), nbytes, _stream()), "quber_post_process")
    out["max_instances"] = max_instances
    out["valid_w"] = valid_w if valid_w else W
    return out


def msda_forward(value, spatial_shapes, level_start_index, sampling_locations, attention_weights, im2col_step=None):
    """MultiScaleDeformableAttention.ms_deform_attn_forward (OPS/src/vision.cpp:18-21).

    value [B,S,M,D] float32 | bfloat16 | float64, spatial_shapes [L,2] int64, level_start_index [L] int64,
    sampling_locations [B,Q,M,L,P,2] and attention_weights [B,Q,M,L,P] float32 (float64 with a float64 value)
    -> [B,Q,M*D] (value dtype).
    `im2col_step` is accepted for signature compatibility and ignored (the batch is one launch).
    """
    _need_cuda(value, "value")
    _need_cuda(spatial_shapes, "spatial_shapes", torch.int64)
    _need_cuda(level_start_index, "level_start_index", torch.int64)
    if value.dtype == torch.float32:
        dt = 0
    elif value.dtype == torch.bfloat16:
        dt = 1
    elif value.dtype == torch.float64:        # real fp64, as the reference's AT_DISPATCH_FLOATING_TYPES (ms_deform_attn_cuda.cu:69)
        dt = 2
    else:
        raise ValueError(f"value must be float32, float64 or bfloat16, got {value.dtype}")
    aux = torch.float64 if dt == 2 else torch.float32
    _need_cuda(sampling_locations, "sampling_loc", aux)
    _need_cuda(attention_weights, "attn_weight", aux)
    B, S, M, D = value.shape
    _, Q, _, Lv, P, _ = sampling_locations.shape
    out = torch.empty((B, Q, M * D), dtype=value.dtype, device=value.device)
    with torch.cuda.device(value.device):
        _lib.check(_lib.lib().quber_msda_forward(_ptr(value), _ptr(spatial_shapes), _ptr(level_start_index),
                                                 _ptr(sampling_locations), _ptr(attention_weights), B, S, M, D, Lv, Q, P, dt,
                                                 _ptr(out), _stream()), "quber_msda_forward")
    return out


def msda_fused_supported(value, n_levels, n_points):
    """True when quber_msda_forward_fused takes this shape (P == 4, L <= 4, a head of a multiple of 64 bytes)."""
    D = value.shape[-1]
    return (n_points == 4 and n_levels <= 4 and value.dtype in (torch.float32, torch.bfloat16)
            and (D * value.element_size()) % 64 == 0 and D in (16, 32, 64))


def msda_forward_fused(value, spatial_shapes, level_start_index, oa, reference_points, n_levels, n_points, levels=None):
    """MSDeformAttn.forward between its Linear layers and the output projection in one kernel
    (OPS/modules/ms_deform_attn.py:102-117 + ms_deform_attn_forward): value [N,S,M,D] float32 | bfloat16, oa [N,Q,ld]
    float32 = the fused sampling_offsets | attention_weights GEMM output, reference_points [N,Q,L,2] float32
    -> [N,Q,M*D] in value's dtype.  The softmax and `ref + off / (W_l, H_l)` happen in the kernel's set-up phase.
    levels: the HOST copy of the level table, ((H_0, W_0), (H_1, W_1), ...) as the caller built spatial_shapes from (level
    starts are the cumulative H*W) — optional; with it the kernel reads the table from its parameter bank (faster, same
    results).  It must describe the same levels as spatial_shapes / level_start_index."""
    _need_cuda(value, "value")
    _need_cuda(spatial_shapes, "spatial_shapes", torch.int64)
    _need_cuda(level_start_index, "level_start_index", torch.int64)
    _need_cuda(oa, "oa", torch.float32)
    _need_cuda(reference_points, "reference_points", torch.float32)
    if not msda_fused_supported(value, n_levels, n_points):
        raise ValueError("msda_forward_fused: unsupported shape (needs P == 4, L <= 4, D in {16, 32, 64})")
    N, S, M, D = value.shape
    Q, ld = oa.shape[1], oa.shape[2]
    if oa.shape[0] != N or ld < M * n_levels * n_points * 3 or tuple(reference_points.shape) != (N, Q, n_levels, 2):
        raise ValueError("msda_forward_fused: oa must be [N,Q,>=M*L*P*3] and reference_points [N,Q,L,2]")
    out = torch.empty((N, Q, M * D), dtype=value.dtype, device=value.device)
    tab = None
    if levels is not None:
        tab = _level_table(levels)
        if len(levels) != n_levels or tab[1] != S:
            raise ValueError("msda_forward_fused: `levels` does not describe value's S positions in n_levels levels")
    with torch.cuda.device(value.device):
        _lib.check(_lib.lib().quber_msda_forward_fused(_ptr(value), _ptr(spatial_shapes), _ptr(level_start_index), _ptr(oa), ld,
                                                       _ptr(reference_points), N, S, M, D, n_levels, Q, n_points,
                                                       0 if value.dtype == torch.float32 else 1, _ptr(out),
                                                       None if tab is None else tab[0], _stream()),
                   "quber_msda_forward_fused")
    return out


_LEVEL_TABLES = {}


def _level_table(levels):
    """((H, W), ...) -> (ctypes int32 [L][3] = (H, W, start), total positions); cached per level tuple."""
    key = tuple((int(h), int(w)) for h, w in levels)
    ent = _LEVEL_TABLES.get(key)
    if ent is None:
        flat, start = [], 0
        for h, w in key:
            flat += [h, w, start]
            start += h * w
        ent = _LEVEL_TABLES[key] = ((ctypes.c_int32 * len(flat))(*flat), start)
    return ent


def gemm_bias_act(a, weight, bias=None, act="none", out_dtype=None):
    """C = act(a @ weight.T + bias) on tcgen05 / TMEM (quber_gemm_bias_act) — the nn.Linear / 1x1-conv shape.

    a [..., K] and weight [N, K] are both bfloat16 (tcgen05 kind::f16) or both float32 (multiplied as TF32);
    bias [N] float32 or None; act "none" | "relu"; out_dtype float32 (default for float32 inputs) or bfloat16.
    The last dimension of `a` must be contiguous; leading dimensions are flattened into rows.
    """
    _need_cuda(weight, "weight")
    if not isinstance(a, torch.Tensor) or not a.is_cuda:
        raise ValueError("a must be a CUDA tensor")
    if a.dtype != weight.dtype or a.dtype not in (torch.bfloat16, torch.float32):
        raise ValueError("a and weight must both be bfloat16 or both float32")
    K = a.shape[-1]
    N = weight.shape[0]
    if weight.dim() != 2 or weight.shape[1] != K:
        raise ValueError(f"weight must be [N, {K}], got {tuple(weight.shape)}")
    a2 = a.reshape(-1, K)
    if a2.stride(1) != 1:
        a2 = a2.contiguous()
    M = a2.shape[0]
    if out_dtype is None:
        out_dtype = a.dtype
    if out_dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("out_dtype must be float32 or bfloat16")
    if bias is not None:
        _need_cuda(bias, "bias", torch.float32)
        if bias.numel() != N:
            raise ValueError("bias must have N elements")
    if act not in ("none", "relu"):
        raise ValueError("act must be 'none' or 'relu'")
    out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    if M > 0:
        with torch.cuda.device(a.device):
            _lib.check(_lib.lib().quber_gemm_bias_act(_ptr(a2), _ptr(weight), _ptr(bias), _ptr(out), M, N, K, a2.stride(0),
                                                      weight.stride(0), N, 1 if a.dtype == torch.bfloat16 else 2,
                                                      0 if out_dtype == torch.float32 else 1, 1 if act == "relu" else 0,
                                                      _stream()), "quber_gemm_bias_act")
    return out.reshape(tuple(a.shape[:-1]) + (N,))


def gemm_bias_act_cat2(a1, a2, weight, bias=None, act="none", out_dtype=None):
    """act(cat([a1, a2], -1) @ weight.T + bias) without the concatenation (quber_gemm_bias_act_cat2): a1 [..., K1],
    a2 [..., K2] with equal leading shapes, weight [N, K1 + K2]; dtypes / act / out_dtype as gemm_bias_act.  K1 must be a
    multiple of 64 (bfloat16) or 32 (float32).  The 1x1 convolution behind the RGB-D concat fusion on channels-last rows."""
    _need_cuda(weight, "weight")
    for t, nm in ((a1, "a1"), (a2, "a2")):
        if not isinstance(t, torch.Tensor) or not t.is_cuda:
            raise ValueError(f"{nm} must be a CUDA tensor")
    if not (a1.dtype == a2.dtype == weight.dtype) or a1.dtype not in (torch.bfloat16, torch.float32):
        raise ValueError("a1, a2 and weight must all be bfloat16 or all float32")
    K1, K2, N = a1.shape[-1], a2.shape[-1], weight.shape[0]
    if weight.dim() != 2 or weight.shape[1] != K1 + K2 or tuple(a1.shape[:-1]) != tuple(a2.shape[:-1]):
        raise ValueError("weight must be [N, K1 + K2] and a1 / a2 must share their leading shape")
    if K1 % (64 if a1.dtype == torch.bfloat16 else 32):
        raise ValueError("K1 must fall on a k-block boundary (64 bfloat16 / 32 float32 elements)")
    r1, r2 = a1.reshape(-1, K1), a2.reshape(-1, K2)
    r1 = r1 if r1.stride(1) == 1 else r1.contiguous()
    r2 = r2 if r2.stride(1) == 1 else r2.contiguous()
    M = r1.shape[0]
    out_dtype = a1.dtype if out_dtype is None else out_dtype
    if out_dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("out_dtype must be float32 or bfloat16")
    if bias is not None:
        _need_cuda(bias, "bias", torch.float32)
        if bias.numel() != N:
            raise ValueError("bias must have N elements")
    if act not in ("none", "relu"):
        raise ValueError("act must be 'none' or 'relu'")
    out = torch.empty((M, N), dtype=out_dtype, device=a1.device)
    if M > 0:
        with torch.cuda.device(a1.device):
            _lib.check(_lib.lib().quber_gemm_bias_act_cat2(_ptr(r1), _ptr(r2), _ptr(weight), _ptr(bias), _ptr(out), M, N, K1, K2,
                                                           r1.stride(0), r2.stride(0), weight.stride(0), N,
                                                           1 if a1.dtype == torch.bfloat16 else 2, 0 if out_dtype == torch.float32 else 1,
                                                           1 if act == "relu" else 0, _stream()), "quber_gemm_bias_act_cat2")
    return out.reshape(tuple(a1.shape[:-1]) + (N,))


def split_rgbd_input(x, pad8=False):
    """x [B,9,H,W] float32 (rgb | depth | encoded offsets, ops.pack_input) -> (rgb6, depth6): NCHW-shaped [B,6,H,W]
    bfloat16 tensors in channels-last memory, cat([rgb, offsets], 1) and cat([depth, offsets], 1) of
    RGBDFusionBackbone.forward (resnet.py:489-493) in one pass.  pad8: [B,8,H,W] with two zero channels behind them."""
    _need_cuda(x, "x", torch.float32)
    if x.dim() != 4 or x.shape[1] != 9 or not x.is_contiguous() or (x.shape[2] * x.shape[3]) % 4:
        raise ValueError("split_rgbd_input: x must be a contiguous [B,9,H,W] tensor with H*W a multiple of 4")
    B, _, H, W = x.shape
    C = 8 if pad8 else 6
    a = torch.empty((B, H, W, C), dtype=torch.bfloat16, device=x.device)
    b = torch.empty((B, H, W, C), dtype=torch.bfloat16, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().quber_split_rgbd_input(_ptr(x), B, H, W, C, _ptr(a), _ptr(b), _stream()), "quber_split_rgbd_input")
    return a.permute(0, 3, 1, 2), b.permute(0, 3, 1, 2)


def split_tf32(w):
    """fp32 CUDA tensor -> (hi, lo): hi = w with the 13 low mantissa bits cleared (what a TF32 tensor-core product
    reads), lo = w - hi (exact).  The pre-split weight of gemm_bias_act_x3."""
    _need_cuda(w, "w", torch.float32)
    hi, lo = torch.empty_like(w), torch.empty_like(w)
    with torch.cuda.device(w.device):
        _lib.check(_lib.lib().quber_split_tf32(_ptr(w), w.numel(), _ptr(hi), _ptr(lo), _stream()), "quber_split_tf32")
    return hi, lo


def gemm_bias_act_x3(a, w_hi, w_lo, bias=None, act="none", out_dtype=torch.float32):
    """C = act(a @ W.T + bias) at fp32 accuracy on the TF32 tensor cores (three split-operand products,
    quber_gemm_bias_act_x3): a [..., K] float32, (w_hi, w_lo) = split_tf32(W [N,K])."""
    _need_cuda(w_hi, "w_hi", torch.float32)
    _need_cuda(w_lo, "w_lo", torch.float32)
    if not isinstance(a, torch.Tensor) or not a.is_cuda or a.dtype != torch.float32:
        raise ValueError("a must be a float32 CUDA tensor")
    K, N = a.shape[-1], w_hi.shape[0]
    if w_hi.dim() != 2 or w_hi.shape[1] != K or w_lo.shape != w_hi.shape:
        raise ValueError(f"w_hi / w_lo must be [N, {K}]")
    a2 = a.reshape(-1, K)
    if a2.stride(1) != 1:
        a2 = a2.contiguous()
    M = a2.shape[0]
    if out_dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("out_dtype must be float32 or bfloat16")
    if bias is not None:
        _need_cuda(bias, "bias", torch.float32)
        if bias.numel() != N:
            raise ValueError("bias must have N elements")
    if act not in ("none", "relu"):
        raise ValueError("act must be 'none' or 'relu'")
    out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    if M > 0:
        with torch.cuda.device(a.device):
            _lib.check(_lib.lib().quber_gemm_bias_act_x3(_ptr(a2), _ptr(w_hi), _ptr(w_lo), _ptr(bias), _ptr(out), M, N, K, a2.stride(0),
                                                         w_hi.stride(0), N, 0 if out_dtype == torch.float32 else 1,
                                                         1 if act == "relu" else 0, _stream()), "quber_gemm_bias_act_x3")
    return out.reshape(tuple(a.shape[:-1]) + (N,))


def pack_input(image, offsets, mean, std):
    """model.py:137-153 for a batch: image [B,C,H,W] uint8, offsets [B,CO,H,W] float32 | None, mean/std [C] float32
    -> [B,C+CO,H,W] float32 = cat((image - mean) / std, offsets)."""
    _need_cuda(image, "image", torch.uint8)
    B, C, H, W = image.shape
    CO = 0
    if offsets is not None:
        _need_cuda(offsets, "offsets", torch.float32)
        CO = offsets.shape[1]
    mean = mean.reshape(-1).to(device=image.device, dtype=torch.float32).contiguous()
    std = std.reshape(-1).to(device=image.device, dtype=torch.float32).contiguous()
    if mean.numel() != C or std.numel() != C:
        raise ValueError("mean / std must have one entry per image channel")
    out = torch.empty((B, C + CO, H, W), dtype=torch.float32, device=image.device)
    with torch.cuda.device(image.device):
        _lib.check(_lib.lib().quber_pack_input(_ptr(image), _ptr(offsets), _ptr(mean), _ptr(std), B, C, CO, H, W, _ptr(out),
                                               _stream()), "quber_pack_input")
    return out


def filter_instances(label_map, counts, fg=None, zero_depth=None, thr=0.3, max_per=64):
    """eval/refiner_model.py:271-288 on the label-map form of the refined instances: keep instance j iff
    sum(mask_j & fg) / sum(mask_j) > thr, then clear the zero-depth pixels of the kept masks (OCID).
    label_map [B,H,W] int32, counts [B] int32, fg / zero_depth [B,H,W] uint8 | None ->
    (filtered label map [B,H,W] int32 with kept instances renumbered, keep [B,max_per] uint8, new counts [B] int32)."""
    _need_cuda(label_map, "label_map", torch.int32)
    _need_cuda(counts, "counts", torch.int32)
    for t, n in ((fg, "fg"), (zero_depth, "zero_depth")):
        if t is not None:
            _need_cuda(t, n, torch.uint8)
            if t.shape != label_map.shape:
                raise ValueError(f"{n} must have the label map's shape")
    B, H, W = label_map.shape
    dev = label_map.device
    out = torch.empty_like(label_map)
    keep = torch.empty((B, max_per), dtype=torch.uint8, device=dev)
    new_counts = torch.empty((B,), dtype=torch.int32, device=dev)
    L = _lib.lib()
    nbytes = L.quber_filter_workspace_bytes(B, max_per)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.quber_filter_instances(_ptr(label_map), _ptr(counts), _ptr(fg), _ptr(zero_depth), B, H, W, max_per, float(thr),
                                            _ptr(out), _ptr(keep), _ptr(new_counts), _ptr(ws), nbytes, _stream()),
                   "quber_filter_instances")
    return out, keep, new_counts


def maxpool3x3s2_supported(x):
    return (isinstance(x, torch.Tensor) and x.is_cuda and x.dim() == 4 and x.dtype in (torch.float32, torch.bfloat16)
            and x.shape[1] % (16 // x.element_size()) == 0 and x.is_contiguous(memory_format=torch.channels_last))


def maxpool3x3s2(x):
    """F.max_pool2d(x, kernel_size=3, stride=2, padding=1) (resnet.py:76, :125) on a channels-last CUDA tensor."""
    if not maxpool3x3s2_supported(x):
        raise ValueError("maxpool3x3s2: needs a channels-last CUDA fp32/bf16 tensor whose channels fill 16-byte vectors")
    N, C, H, W = x.shape
    y = torch.empty((N, C, (H - 1) // 2 + 1, (W - 1) // 2 + 1), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().quber_maxpool3x3s2(_ptr(x), _ptr(y), N, C, H, W, 0 if x.dtype == torch.float32 else 1, _stream()),
                   "quber_maxpool3x3s2")
    return y


def host_pack_masks(masks_host, out=None, threads=None):
    """HOST: bit-pack a contiguous uint8 CPU tensor (numel % 32 == 0) -> (int32 CPU tensor [numel / 32], value).
    `value` = the one value every set pixel carries (1, or 255 as eval/refiner_model.py:230-231 makes the masks), 0 for
    an all-zero buffer, -1 when there is more than one non-zero value (send the bytes then).  The reference moves the
    bytes (predictor.py:304-310)."""
    if masks_host.is_cuda or masks_host.dtype != torch.uint8 or not masks_host.is_contiguous():
        raise ValueError("masks_host must be a contiguous uint8 CPU tensor")
    n = masks_host.numel()
    if n == 0 or n % 32:
        raise ValueError("number of mask bytes must be a positive multiple of 32")
    if out is None:
        out = torch.empty(n // 32, dtype=torch.int32)
    elif out.is_cuda or out.dtype != torch.int32 or out.numel() < n // 32 or not out.is_contiguous():
        raise ValueError("out must be a contiguous int32 CPU tensor of numel / 32 words")
    if threads is None:
        threads = min(len(os.sched_getaffinity(0)), 32)
    value = ctypes.c_int(-1)
    _lib.check(_lib.lib().quber_host_pack_masks(masks_host.data_ptr(), n, out.data_ptr(), int(threads), ctypes.byref(value)),
               "quber_host_pack_masks")
    return out, int(value.value)


def host_pack_begin(masks_host, out, threads, chunk_end_bytes):
    """HOST, asynchronous: start packing `masks_host` (contiguous uint8 CPU tensor) into `out` (int32 CPU tensor of
    numel / 32 words) on `threads` pool threads, chunk after chunk (`chunk_end_bytes`: ascending multiples of 32 ending at
    numel).  Returns at once; host_pack_wait(c) blocks until chunk c is packed.  The caller keeps both tensors alive."""
    if masks_host.is_cuda or masks_host.dtype != torch.uint8 or not masks_host.is_contiguous():
        raise ValueError("masks_host must be a contiguous uint8 CPU tensor")
    n = masks_host.numel()
    if out.is_cuda or out.dtype != torch.int32 or out.numel() < n // 32 or not out.is_contiguous():
        raise ValueError("out must be a contiguous int32 CPU tensor of numel / 32 words")
    ends = (ctypes.c_size_t * len(chunk_end_bytes))(*[int(e) for e in chunk_end_bytes])
    _lib.check(_lib.lib().quber_host_pack_begin(masks_host.data_ptr(), n, out.data_ptr(), int(threads), ends, len(chunk_end_bytes)),
               "quber_host_pack_begin")


def host_pack_wait(chunk):
    """Blocks until `chunk` of the job started by host_pack_begin is packed -> its value (see host_pack_masks)."""
    value = ctypes.c_int(-1)
    _lib.check(_lib.lib().quber_host_pack_wait(int(chunk), ctypes.byref(value)), "quber_host_pack_wait")
    return int(value.value)


def unpack_masks(packed, shape, value=1, out=None):
    """int32 CUDA tensor of packed words -> uint8 CUDA tensor `shape` (prod(shape) == 32 * words), set pixels = value."""
    _need_cuda(packed, "packed", torch.int32)
    n = 1
    for d in shape:
        n *= int(d)
    if n != 32 * packed.numel():
        raise ValueError("shape does not match the number of packed words")
    if not 0 <= int(value) <= 255:
        raise ValueError("value must be a byte")
    if out is None:
        out = torch.empty(tuple(shape), dtype=torch.uint8, device=packed.device)
    else:
        _need_cuda(out, "out", torch.uint8)
        if out.numel() != n or out.device != packed.device:
            raise ValueError("out does not match shape")
    with torch.cuda.device(packed.device):
        _lib.check(_lib.lib().quber_unpack_masks(_ptr(packed), packed.numel(), int(value), _ptr(out), _stream()), "quber_unpack_masks")
    return out


def add_layer_norm(a, b, weight, bias, eps=1e-5, want_bf16=False, pos=None, q_dtype=None):
    """LayerNorm(a + b) over the last dimension (msdeformattn.py:119-120, :126-127) in one pass.  a, b [..., C] float32
    CUDA (b may be None), C a multiple of 128 (<= 1024) -> y float32 [, y_bf16] [, q = y + pos in `q_dtype`, the next
    layer's query (with_pos_embed, msdeformattn.py:124)]."""
    _need_cuda(a, "a", torch.float32)
    if b is not None:
        _need_cuda(b, "b", torch.float32)
        if b.shape != a.shape:
            raise ValueError("a and b must have the same shape")
    C = a.shape[-1]
    if C % 128 or C > 1024 or C // 128 in (5, 7):
        raise ValueError("add_layer_norm: the normalised dimension must be 128 * {1,2,3,4,6,8}")
    rows = a.numel() // C
    y = torch.empty_like(a)
    y16 = torch.empty(a.shape, dtype=torch.bfloat16, device=a.device) if want_bf16 else None
    q = None
    if pos is not None:
        _need_cuda(pos, "pos", torch.float32)
        if pos.shape != a.shape or q_dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("pos must have a's shape and q_dtype must be float32 or bfloat16")
        q = torch.empty(a.shape, dtype=q_dtype, device=a.device)
    g = weight.detach().float().contiguous()
    bt = bias.detach().float().contiguous()
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().quber_add_layer_norm(_ptr(a), _ptr(b), _ptr(g), _ptr(bt), rows, C, float(eps), _ptr(y), _ptr(y16),
                                                   _ptr(pos), _ptr(q), int(q_dtype == torch.bfloat16), _stream()),
                   "quber_add_layer_norm")
    out = (y,) + ((y16,) if want_bf16 else ()) + ((q,) if pos is not None else ())
    return out if len(out) > 1 else y


def msda_prepare(oa, reference_points, spatial_shapes, n_heads, n_levels, n_points):
    """OPS/modules/ms_deform_attn.py:102-109: the fused sampling_offsets | attention_weights projection
    oa [N,Q,M*L*P*3] float32 and reference_points [N,Q,L,2] -> (sampling_locations [N,Q,M,L,P,2], attention weights
    [N,Q,M,L,P] after the softmax over L*P)."""
    _need_cuda(oa, "oa", torch.float32)
    _need_cuda(reference_points, "reference_points", torch.float32)
    _need_cuda(spatial_shapes, "spatial_shapes", torch.int64)
    N, Q, ld = oa.shape
    M, L, P = n_heads, n_levels, n_points
    if ld != M * L * P * 3 or tuple(reference_points.shape) != (N, Q, L, 2):
        raise ValueError("msda_prepare: oa must be [N,Q,M*L*P*3] and reference_points [N,Q,L,2]")
    loc = torch.empty((N, Q, M, L, P, 2), dtype=torch.float32, device=oa.device)
    aw = torch.empty((N, Q, M, L, P), dtype=torch.float32, device=oa.device)
    with torch.cuda.device(oa.device):
        _lib.check(_lib.lib().quber_msda_prepare(_ptr(oa), ld, _ptr(reference_points), _ptr(spatial_shapes), N * Q, M, L, P,
                                                 _ptr(loc), _ptr(aw), _stream()), "quber_msda_prepare")
    return loc, aw


def group_norm_supported(x, groups):
    """True when quber_group_norm_relu takes this tensor as it lies (CUDA, fp32 / bf16, dense NCHW or channels-last,
    16-byte vectors along the contiguous dimension)."""
    if not (isinstance(x, torch.Tensor) and x.is_cuda and x.dim() == 4 and x.dtype in (torch.float32, torch.bfloat16)):
        return False
    N, C, H, W = x.shape
    v = 16 // x.element_size()
    if C % groups or N == 0:
        return False
    if x.is_contiguous():
        return (H * W) % v == 0 and N * groups <= 65535
    if x.is_contiguous(memory_format=torch.channels_last):
        tp = C // v
        return C % v == 0 and (tp <= 256 or tp % 256 == 0)
    return False


def group_norm_relu(x, weight, bias, groups=32, eps=1e-5, relu=True, inplace=False, pre_bias=None):
    """nn.GroupNorm(groups, C)(x) followed by ReLU (detectron2 Conv2d norm="GN" + F.relu, model.py:386-403;
    resnet.py:466-479) in two passes over x (statistics; normalise + affine + ReLU).  x [N,C,H,W] float32 | bfloat16,
    dense NCHW or channels-last; the result has x's dtype and memory format.  weight / bias [C] (any float dtype).
    pre_bias [C] | None: the preceding convolution's bias, added to x before the statistics."""
    if not group_norm_supported(x, groups):
        raise ValueError("group_norm_relu: unsupported tensor (needs a dense CUDA fp32/bf16 NCHW or channels-last tensor)")
    N, C, H, W = x.shape
    layout = 0 if x.is_contiguous() else 1
    y = x if inplace else torch.empty_like(x)           # empty_like preserves the memory format
    g = weight.detach().to(device=x.device, dtype=torch.float32).contiguous()
    b = bias.detach().to(device=x.device, dtype=torch.float32).contiguous()
    pb = None if pre_bias is None else pre_bias.detach().to(device=x.device, dtype=torch.float32).contiguous()
    L = _lib.lib()
    nbytes = L.quber_group_norm_workspace_bytes(N, C, H * W, groups, layout)
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(L.quber_group_norm_relu(_ptr(x), _ptr(y), _ptr(pb), _ptr(g), _ptr(b), N, C, H * W, groups, float(eps), int(bool(relu)),
                                           layout, 0 if x.dtype == torch.float32 else 1, _ptr(ws), nbytes, _stream()),
                   "quber_group_norm_relu")
    return y


def upsample_concat_supported(a, b):
    if not (isinstance(b, torch.Tensor) and b.is_cuda and b.dim() == 4 and b.dtype in (torch.float32, torch.bfloat16)):
        return False
    v = 16 // b.element_size()
    cl = lambda t: t.is_contiguous(memory_format=torch.channels_last)
    if b.shape[1] % v or not cl(b):
        return False
    if a is not None and not (a.is_cuda and a.dtype == b.dtype and a.shape[0] == b.shape[0] and a.shape[1] % v == 0 and cl(a)):
        return False
    return True


def upsample_concat(a, b, size=None):
    """torch.cat([a, F.interpolate(b, size=a.shape[-2:], mode="bilinear", align_corners=False)], dim=1) in one pass
    (DeepLabV3PlusHead.layers, model.py:713).  a [N,Ca,H,W] | None (then `size` = (H, W)), b [N,Cb,h,w]; channels-last
    float32 | bfloat16 CUDA tensors; the result is channels-last [N,Ca+Cb,H,W]."""
    if not upsample_concat_supported(a, b):
        raise ValueError("upsample_concat: needs channels-last CUDA fp32/bf16 tensors with channel counts that fill 16-byte vectors")
    N, Cb, h, w = b.shape
    Ca = 0 if a is None else a.shape[1]
    H, W = (a.shape[2], a.shape[3]) if a is not None else (int(size[0]), int(size[1]))
    out = torch.empty((N, Ca + Cb, H, W), dtype=b.dtype, device=b.device, memory_format=torch.channels_last)
    with torch.cuda.device(b.device):
        _lib.check(_lib.lib().quber_upsample_concat(_ptr(a), _ptr(b), _ptr(out), N, Ca, Cb, H, W, h, w,
                                                    0 if b.dtype == torch.float32 else 1, _stream()), "quber_upsample_concat")
    return out


_FUSE_OPS = {"copy": 0, "sigmoid": 1, "softmax": 2}


def fuse_concat(sources, ops_, layout="nchw", pitch=None):
    """model.py:745-759: channel concat of [B,C_i,h,w] float32 sources, each passed through "copy" | "sigmoid" |
    "softmax" (over its own channels).  layout "nchw" -> [B,sum C,h,w] float32; "nhwc_f32" / "nhwc_bf16" ->
    channels-last rows [B*h*w, pitch] (pitch >= sum C, zero padded) for gemm_bias_act."""
    if len(sources) != len(ops_) or not sources:
        raise ValueError("sources and ops must have the same non-zero length")
    for i, t in enumerate(sources):
        _need_cuda(t, f"sources[{i}]", torch.float32)
    B, _, h, w = sources[0].shape
    if any(t.shape[0] != B or t.shape[2:] != sources[0].shape[2:] for t in sources):
        raise ValueError("Sizes of tensors must match except in dimension 1")      # what torch.cat raises
    ch = [int(t.shape[1]) for t in sources]
    ctot = sum(ch)
    lay = {"nchw": 0, "nhwc_f32": 1, "nhwc_bf16": 2}[layout]
    if lay == 0:
        pitch = ctot
        out = torch.empty((B, ctot, h, w), dtype=torch.float32, device=sources[0].device)
    else:
        pitch = ctot if pitch is None else int(pitch)
        out = torch.empty((B * h * w, pitch), dtype=torch.float32 if lay == 1 else torch.bfloat16, device=sources[0].device)
    n = len(sources)
    ptrs = (ctypes.c_void_p * n)(*[t.data_ptr() for t in sources])
    chs = (ctypes.c_int * n)(*ch)
    opc = (ctypes.c_int * n)(*[_FUSE_OPS[o] for o in ops_])
    with torch.cuda.device(sources[0].device):
        _lib.check(_lib.lib().quber_fuse_concat(ptrs, chs, opc, n, B, h, w, lay, pitch, _ptr(out), _stream()), "quber_fuse_concat")
    return out


def fuse_concat_rows(sources, ops_, pitch=None, out_dtype=torch.bfloat16):
    """fuse_concat for channels-last sources (model.py:745-759): each source [B,C_i,h,w] float32 | bfloat16 dense in
    channels-last memory -> rows [B*h*w, pitch] (zero padded), no layout copies."""
    if len(sources) != len(ops_) or not sources:
        raise ValueError("sources and ops must have the same non-zero length")
    B, _, h, w = sources[0].shape
    for i, t in enumerate(sources):
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype in (torch.float32, torch.bfloat16)
                and t.is_contiguous(memory_format=torch.channels_last)):
            raise ValueError(f"sources[{i}] must be a channels-last CUDA float32 / bfloat16 tensor")
        if t.shape[0] != B or t.shape[2:] != sources[0].shape[2:]:
            raise ValueError("Sizes of tensors must match except in dimension 1")
    ch = [int(t.shape[1]) for t in sources]
    ctot = sum(ch)
    pitch = ctot if pitch is None else int(pitch)
    rows = B * h * w
    out = torch.empty((rows, pitch), dtype=out_dtype, device=sources[0].device)
    n = len(sources)
    ptrs = (ctypes.c_void_p * n)(*[t.data_ptr() for t in sources])
    chs = (ctypes.c_int * n)(*ch)
    opc = (ctypes.c_int * n)(*[_FUSE_OPS[o] for o in ops_])
    bf = (ctypes.c_int * n)(*[int(t.dtype == torch.bfloat16) for t in sources])
    with torch.cuda.device(sources[0].device):
        _lib.check(_lib.lib().quber_fuse_concat_rows(ptrs, chs, opc, bf, n, rows, pitch, int(out_dtype == torch.bfloat16), _ptr(out),
                                                     _stream()), "quber_fuse_concat_rows")
    return out



def upsample_crop_resize(x, scale, crop_hw, out_hw, mul=1.0):
    """model.py:689-708 + detectron2 sem_seg_postprocess (model.py:266-289) in one pass: x [B,C,h,w] float32 head
    outputs -> [B,C,out_h,out_w] = resize(crop(upsample(x, scale) * mul, crop_hw), out_hw), bilinear,
    align_corners=False at both steps (the second is the identity when out_hw == crop_hw)."""
    _need_cuda(x, "x", torch.float32)
    B, C, h, w = x.shape
    ch, cw, oh, ow = int(crop_hw[0]), int(crop_hw[1]), int(out_hw[0]), int(out_hw[1])
    out = torch.empty((B, C, oh, ow), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().quber_upsample_crop_resize(_ptr(x), B, C, h, w, int(scale), float(mul), ch, cw, oh, ow, _ptr(out),
                                                         _stream()), "quber_upsample_crop_resize")
    return out


def resize_nearest(masks, out_hw):
    """cv2.resize(m, (out_w, out_h), interpolation=cv2.INTER_NEAREST) for every plane of masks uint8 [...,h,w] (CUDA)."""
    _need_cuda(masks, "masks", torch.uint8)
    h, w = masks.shape[-2:]
    n = masks.numel() // (h * w)
    oh, ow = int(out_hw[0]), int(out_hw[1])
    out = torch.empty(tuple(masks.shape[:-2]) + (oh, ow), dtype=torch.uint8, device=masks.device)
    if n:
        with torch.cuda.device(masks.device):
            _lib.check(_lib.lib().quber_resize_nearest_u8(_ptr(masks), n, h, w, oh, ow, _ptr(out), _stream()), "quber_resize_nearest_u8")
    return out


def eval_counts(prediction, gt, radius, max_labels=64, boundary=True, want_bmaps=False):
    """The integer counts of multilabel_metrics (eval/evaluation.py:57-274) for a batch of label-map pairs:
    prediction, gt int32 CUDA [B,H,W] (labels in [0, 65535], 0 = background) -> dict of device tensors:
    pred_labels / gt_labels [B,ML], n_pred / n_gt [B], inter [B,ML,ML] (gt index first), area_pred / area_gt [B,ML],
    bcount_pred / bcount_gt [B,ML], btp [B,ML,ML,2], status [1] (+ bmaps [2,B,H,W] uint8 with want_bmaps)."""
    _need_cuda(prediction, "prediction", torch.int32)
    _need_cuda(gt, "gt", torch.int32)
    if prediction.shape != gt.shape or prediction.dim() != 3:
        raise ValueError("prediction and gt must both be [B,H,W]")
    B, H, W = prediction.shape
    ML = int(max_labels)
    if not 0 < ML <= 255:
        raise ValueError("max_labels must be in 1..255")
    dev = prediction.device
    i32 = torch.int32
    e = lambda *shape: torch.empty(shape, dtype=i32, device=dev)
    out = dict(pred_labels=e(B, ML), n_pred=e(B), gt_labels=e(B, ML), n_gt=e(B), bcount_pred=e(B, ML), bcount_gt=e(B, ML),
               btp=e(B, ML, ML, 2), status=e(1))
    full = e(B, ML + 1, ML + 1)
    bm = torch.empty((2, B, H, W), dtype=torch.uint8, device=dev) if want_bmaps else None
    L = _lib.lib()
    nbytes = L.quber_eval_workspace_bytes(B, H, W, ML)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.quber_eval_counts(_ptr(prediction), _ptr(gt), B, H, W, ML, int(radius), int(bool(boundary)), _ptr(out["pred_labels"]),
                                       _ptr(out["n_pred"]), _ptr(out["gt_labels"]), _ptr(out["n_gt"]), _ptr(full), _ptr(out["bcount_pred"]),
                                       _ptr(out["bcount_gt"]), _ptr(out["btp"]), _ptr(bm), _ptr(ws), nbytes, _ptr(out["status"]), _stream()),
                   "quber_eval_counts")
    out["inter"] = full[:, 1:, 1:]
    out["area_gt"] = full[:, 1:, :].sum(-1)                     # row sums incl. the background column
    out["area_pred"] = full[:, :, 1:].sum(-2)
    if not boundary:
        out["bcount_pred"].zero_(); out["bcount_gt"].zero_(); out["btp"].zero_()
    if want_bmaps:
        out["bmaps"] = bm
    return out


def seg2bmap(masks):
    """eval/utilities.py:672-696 for a batch of binary maps: masks [B,H,W] (non-zero = inside, any integer dtype) ->
    uint8 [B,H,W] 0/1, the one-pixel outer boundary (cv2.findContours RETR_EXTERNAL + drawContours)."""
    if not isinstance(masks, torch.Tensor) or not masks.is_cuda:
        raise ValueError("masks must be a CUDA tensor")
    m = (masks != 0).to(torch.int32).contiguous()
    r = eval_counts(m, m, 0, max_labels=1, boundary=True, want_bmaps=True)
    return r["bmaps"][0]


def rle_decode(counts, offsets, n, H, W, value=1, packed=False, out=None, check=False):
    """pycocotools.mask.decode on the device (perturbed_panoptic_dataset_mapper.py:195-214): `counts` uint32-as-int32
    CUDA tensor holding the run lengths of `n` masks back to back (column-major runs, starting with zeros), `offsets`
    int64 CUDA [n+1] -> ROW-major masks: uint8 [n,H,W] (0 / value), or with packed=True int32 [n*H*W/32] words
    (bit i of word w = pixel 32 w + i; W % 32 == 0), the input of encode_initial_masks_packed.
    Returns (masks, status int32 CUDA [1]); check=True synchronises and raises if some counts do not sum to H*W."""
    _need_cuda(counts, "counts", torch.int32)
    _need_cuda(offsets, "offsets", torch.int64)
    if offsets.numel() != n + 1:
        raise ValueError("offsets must have n + 1 entries")
    dev = counts.device
    if out is None:
        out = (torch.empty(n * H * W // 32, dtype=torch.int32, device=dev) if packed
               else torch.empty((n, H, W), dtype=torch.uint8, device=dev))
    else:
        _need_cuda(out, "out", torch.int32 if packed else torch.uint8)
        if out.numel() * (32 if packed else 1) != n * H * W:
            raise ValueError("out does not match n*H*W")
    if packed and W % 32:
        raise ValueError("packed RLE decode needs W % 32 == 0")
    L = _lib.lib()
    nbytes = L.quber_rle_decode_workspace_bytes(counts.numel())
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    status = torch.empty(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.quber_rle_decode(_ptr(counts), _ptr(offsets), counts.numel(), int(n), int(H), int(W), int(value),
                                      _ptr(out) if packed else None, None if packed else _ptr(out), _ptr(ws), nbytes,
                                      _ptr(status), _stream()), "quber_rle_decode")
    if check and int(status[0]) & 1:
        raise ValueError("RLE counts do not sum to H*W")
    return out, status


def rle_encode(masks, cap=None):
    """pycocotools.mask.encode on the device (tools/ours/panoptic2eee.py:61-64,125-132): masks uint8 CUDA [n,H,W]
    (non-zero = set) -> (counts int32 CUDA [cap] holding uint32 run lengths, offsets int64 CUDA [n+1], status int32 [1]).
    `cap` = capacity of the counts buffer (default: 4 runs per column and mask + 2, which every blob-like mask fits;
    status bit 1 is set when it was too small and offsets[n] then holds the capacity needed)."""
    _need_cuda(masks, "masks", torch.uint8)
    n, H, W = masks.shape
    dev = masks.device
    if cap is None:
        cap = n * (4 * W + 2)
    counts = torch.empty(int(cap), dtype=torch.int32, device=dev)
    offsets = torch.empty(n + 1, dtype=torch.int64, device=dev)
    status = torch.empty(1, dtype=torch.int32, device=dev)
    L = _lib.lib()
    nbytes = L.quber_rle_encode_workspace_bytes(n, H, W, int(cap))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.quber_rle_encode(_ptr(masks), n, H, W, _ptr(counts), int(cap), _ptr(offsets), _ptr(ws), nbytes,
                                      _ptr(status), _stream()), "quber_rle_encode")
    return counts, offsets, status


def rle_from_string(s):
    """HOST: compressed COCO RLE string (bytes / str) -> uint32 numpy counts (maskApi.c rleFrString)."""
    if isinstance(s, str):
        s = s.encode("ascii")
    out = np.empty(max(len(s), 1), np.uint32)
    m = _lib.lib().quber_rle_from_string(s, len(s), out.ctypes.data_as(ctypes.c_void_p), out.size)
    if m < 0:
        raise ValueError("malformed RLE string")
    return out[:m].copy()


def rle_to_string(counts):
    """HOST: uint32 counts -> compressed COCO RLE bytes (maskApi.c rleToString)."""
    counts = np.ascontiguousarray(counts, np.uint32)
    buf = ctypes.create_string_buffer(7 * counts.size + 1)
    n = _lib.lib().quber_rle_to_string(counts.ctypes.data_as(ctypes.c_void_p), counts.size, buf, len(buf))
    if n < 0 or n >= len(buf):
        raise ValueError("rle_to_string failed")
    return buf.raw[:n]


# ---- torch.library registration: torch.ops.quber.* ----------------------------------------------
# Every device entry point of the C ABI is a PyTorch custom op with a fake (meta) kernel, so the path is visible to
# torch.compile / torch.export as opaque CUDA ops (no graph break, fullgraph=True works) and has NO decomposition into
# aten ops — there is nothing for a compiler to "fall back" to.  The functions above are the implementations.
POST_OUTPUTS = ("sem_seg", "eee_full", "eee_onehot", "panoptic", "label_map", "masks", "centers_yx", "n_centers",
                "inst_offsets", "labels", "scores", "boxes", "status")


def _register():
    try:
        from torch.library import custom_op
    except Exception:        # pragma: no cover - very old torch
        return
    from typing import Optional, Sequence
    T = torch.Tensor

    @custom_op("quber::msda_forward", mutates_args=(), device_types="cuda")
    def _msda(value: T, spatial_shapes: T, level_start_index: T, sampling_locations: T, attention_weights: T) -> T:
        return msda_forward(value, spatial_shapes, level_start_index, sampling_locations, attention_weights)

    @_msda.register_fake
    def _(value, spatial_shapes, level_start_index, sampling_locations, attention_weights):
        B, S, M, D = value.shape
        return value.new_empty((B, sampling_locations.shape[1], M * D))

    @custom_op("quber::msda_forward_fused", mutates_args=(), device_types="cuda")
    def _msdaf(value: T, spatial_shapes: T, level_start_index: T, oa: T, reference_points: T, n_levels: int, n_points: int) -> T:
        return msda_forward_fused(value, spatial_shapes, level_start_index, oa, reference_points, n_levels, n_points)

    @_msdaf.register_fake
    def _(value, spatial_shapes, level_start_index, oa, reference_points, n_levels, n_points):
        B, S, M, D = value.shape
        return value.new_empty((B, oa.shape[1], M * D))

    @custom_op("quber::msda_prepare", mutates_args=(), device_types="cuda")
    def _prep(oa: T, reference_points: T, spatial_shapes: T, n_heads: int, n_levels: int, n_points: int) -> tuple[T, T]:
        return msda_prepare(oa, reference_points, spatial_shapes, n_heads, n_levels, n_points)

    @_prep.register_fake
    def _(oa, reference_points, spatial_shapes, n_heads, n_levels, n_points):
        N, Q, _ = oa.shape
        return (oa.new_empty((N, Q, n_heads, n_levels, n_points, 2)), oa.new_empty((N, Q, n_heads, n_levels, n_points)))

    @custom_op("quber::encode_initial_masks", mutates_args=(), device_types="cuda")
    def _enc(masks: T, sigma: int) -> T:
        return encode_initial_masks(masks, sigma)

    @_enc.register_fake
    def _(masks, sigma):
        B, K, H, W = masks.shape
        return masks.new_empty((B, 3, H, W), dtype=torch.float32)

    @custom_op("quber::encode_initial_masks_packed", mutates_args=("masks_out",), device_types="cuda")
    def _encp(packed: T, value: int, masks_out: T, sigma: int) -> T:
        return encode_initial_masks_packed(packed, value, masks_out, sigma)

    @_encp.register_fake
    def _(packed, value, masks_out, sigma):
        B, K, H, W = masks_out.shape
        return packed.new_empty((B, 3, H, W), dtype=torch.float32)

    @custom_op("quber::unpack_masks", mutates_args=(), device_types="cuda")
    def _unp(packed: T, shape: Sequence[int], value: int) -> T:
        return unpack_masks(packed, tuple(shape), value)

    @_unp.register_fake
    def _(packed, shape, value):
        return packed.new_empty(tuple(shape), dtype=torch.uint8)

    @custom_op("quber::mask_boundary", mutates_args=(), device_types="cuda")
    def _bnd(masks: T, d: int) -> tuple[T, T, T]:
        return mask_boundary(masks, d)

    @_bnd.register_fake
    def _(masks, d):
        B, K, H, W = masks.shape
        return torch.empty_like(masks), masks.new_empty((B, H, W)), masks.new_empty((B, H, W))

    @custom_op("quber::quadruple_maps", mutates_args=(), device_types="cuda")
    def _quad(gt: T, pred: T) -> T:
        return quadruple_maps(gt, pred)

    @_quad.register_fake
    def _(gt, pred):
        return gt.new_empty((gt.shape[0], 4) + tuple(gt.shape[1:]))

    @custom_op("quber::post_process", mutates_args=(), device_types="cuda")
    def _post(foreground: T, center: T, offset: T, eee: Optional[T], scale: int, offset_mul: float, threshold: float, nms_kernel: int,
              top_k: int, min_area: int, stuff_area: int, label_divisor: int, void_label: float, max_instances: int,
              label_u8: bool) -> tuple[T, T, T, T, T, T, T, T, T, T, T, T, T]:
        r = post_process(foreground, center, offset, eee, scale=scale, offset_mul=offset_mul, threshold=threshold,
                         nms_kernel=nms_kernel, top_k=top_k, min_area=min_area, stuff_area=stuff_area, label_divisor=label_divisor,
                         void_label=void_label, max_instances=max_instances,
                         label_dtype=torch.uint8 if label_u8 else torch.int32)
        B, _, h, w = foreground.shape
        empty = foreground.new_empty((B, 0, h * scale, w * scale))
        r.setdefault("eee_full", empty)
        r.setdefault("eee_onehot", empty.to(torch.uint8))
        return tuple(r[k] for k in POST_OUTPUTS)

    @_post.register_fake
    def _(foreground, center, offset, eee, scale, offset_mul, threshold, nms_kernel, top_k, min_area, stuff_area, label_divisor,
          void_label, max_instances, label_u8):
        B, _, h, w = foreground.shape
        H, W = h * scale, w * scale
        C = 0 if eee is None else eee.shape[1]
        n = max(max_instances, 1)
        f, i32, u8 = torch.float32, torch.int32, torch.uint8
        e = lambda shape, dt: foreground.new_empty(shape, dtype=dt)
        return (e((B, 1, H, W), f), e((B, C, H, W), f), e((B, C, H, W), u8), e((B, H, W), f),
                e((B, H, W), u8 if label_u8 else i32), e((n, H, W), u8), e((B, top_k, 2), i32), e((B,), i32), e((B + 1,), i32),
                e((n,), i32), e((n,), f), e((n, 4), f), e((1,), i32))

    @custom_op("quber::gemm_bias_act", mutates_args=(), device_types="cuda")
    def _gemm(a: T, weight: T, bias: Optional[T], relu: bool, out_bf16: bool) -> T:
        return gemm_bias_act(a, weight, bias, act="relu" if relu else "none", out_dtype=torch.bfloat16 if out_bf16 else torch.float32)

    @_gemm.register_fake
    def _(a, weight, bias, relu, out_bf16):
        return a.new_empty(tuple(a.shape[:-1]) + (weight.shape[0],), dtype=torch.bfloat16 if out_bf16 else torch.float32)

    @custom_op("quber::split_rgbd_input", mutates_args=(), device_types="cuda")
    def _split_rgbd(x: T, pad8: bool) -> tuple[T, T]:
        return split_rgbd_input(x, pad8)

    @_split_rgbd.register_fake
    def _(x, pad8):
        B, _, H, W = x.shape
        e = lambda: torch.empty((B, H, W, 8 if pad8 else 6), dtype=torch.bfloat16, device=x.device).permute(0, 3, 1, 2)
        return e(), e()

    @custom_op("quber::gemm_bias_act_cat2", mutates_args=(), device_types="cuda")
    def _gemm_cat2(a1: T, a2: T, weight: T, bias: Optional[T], relu: bool, out_bf16: bool) -> T:
        return gemm_bias_act_cat2(a1, a2, weight, bias, act="relu" if relu else "none",
                                  out_dtype=torch.bfloat16 if out_bf16 else torch.float32)

    @_gemm_cat2.register_fake
    def _(a1, a2, weight, bias, relu, out_bf16):
        return a1.new_empty(tuple(a1.shape[:-1]) + (weight.shape[0],), dtype=torch.bfloat16 if out_bf16 else torch.float32)

    @custom_op("quber::gemm_bias_act_x3", mutates_args=(), device_types="cuda")
    def _gemm3(a: T, w_hi: T, w_lo: T, bias: Optional[T], relu: bool, out_bf16: bool) -> T:
        return gemm_bias_act_x3(a, w_hi, w_lo, bias, act="relu" if relu else "none", out_dtype=torch.bfloat16 if out_bf16 else torch.float32)

    @_gemm3.register_fake
    def _(a, w_hi, w_lo, bias, relu, out_bf16):
        return a.new_empty(tuple(a.shape[:-1]) + (w_hi.shape[0],), dtype=torch.bfloat16 if out_bf16 else torch.float32)

    @custom_op("quber::split_tf32", mutates_args=(), device_types="cuda")
    def _split(w: T) -> tuple[T, T]:
        return split_tf32(w)

    @_split.register_fake
    def _(w):
        return torch.empty_like(w), torch.empty_like(w)

    @custom_op("quber::pack_input", mutates_args=(), device_types="cuda")
    def _pack(image: T, offsets: Optional[T], mean: T, std: T) -> T:
        return pack_input(image, offsets, mean, std)

    @_pack.register_fake
    def _(image, offsets, mean, std):
        B, C, H, W = image.shape
        return image.new_empty((B, C + (0 if offsets is None else offsets.shape[1]), H, W), dtype=torch.float32)

    @custom_op("quber::fuse_concat", mutates_args=(), device_types="cuda")
    def _fuse(sources: Sequence[T], op_codes: Sequence[int], layout: int, pitch: int) -> T:
        names = {v: k for k, v in _FUSE_OPS.items()}
        return fuse_concat(list(sources), [names[c] for c in op_codes], layout=("nchw", "nhwc_f32", "nhwc_bf16")[layout],
                           pitch=pitch if layout else None)

    @_fuse.register_fake
    def _(sources, op_codes, layout, pitch):
        B, _, h, w = sources[0].shape
        ctot = sum(t.shape[1] for t in sources)
        if layout == 0:
            return sources[0].new_empty((B, ctot, h, w), dtype=torch.float32)
        return sources[0].new_empty((B * h * w, pitch), dtype=torch.float32 if layout == 1 else torch.bfloat16)

    @custom_op("quber::group_norm_relu", mutates_args=(), device_types="cuda")
    def _gn(x: T, weight: T, bias: T, groups: int, eps: float, relu: bool, pre_bias: Optional[T]) -> T:
        return group_norm_relu(x, weight, bias, groups, eps, relu, False, pre_bias)

    @_gn.register_fake
    def _(x, weight, bias, groups, eps, relu, pre_bias):
        return torch.empty_like(x)

    @custom_op("quber::upsample_concat", mutates_args=(), device_types="cuda")
    def _upc(a: T, b: T) -> T:
        return upsample_concat(a, b)

    @_upc.register_fake
    def _(a, b):
        return torch.empty((b.shape[0], a.shape[1] + b.shape[1], a.shape[2], a.shape[3]), dtype=b.dtype, device=b.device,
                           memory_format=torch.channels_last)

    @custom_op("quber::maxpool3x3s2", mutates_args=(), device_types="cuda")
    def _mp(x: T) -> T:
        return maxpool3x3s2(x)

    @_mp.register_fake
    def _(x):
        N, C, H, W = x.shape
        return torch.empty((N, C, (H - 1) // 2 + 1, (W - 1) // 2 + 1), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)

    @custom_op("quber::add_layer_norm", mutates_args=(), device_types="cuda")
    def _aln(a: T, b: Optional[T], weight: T, bias: T, eps: float) -> T:
        return add_layer_norm(a, b, weight, bias, eps)

    @_aln.register_fake
    def _(a, b, weight, bias, eps):
        return torch.empty_like(a)

    @custom_op("quber::filter_instances", mutates_args=(), device_types="cuda")
    def _filt(label_map: T, counts: T, fg: Optional[T], zero_depth: Optional[T], thr: float, max_per: int) -> tuple[T, T, T]:
        return filter_instances(label_map, counts, fg, zero_depth, thr, max_per)

    @_filt.register_fake
    def _(label_map, counts, fg, zero_depth, thr, max_per):
        B = label_map.shape[0]
        return torch.empty_like(label_map), label_map.new_empty((B, max_per), dtype=torch.uint8), label_map.new_empty((B,), dtype=torch.int32)

    @custom_op("quber::upsample_crop_resize", mutates_args=(), device_types="cuda")
    def _ucr(x: T, scale: int, crop_h: int, crop_w: int, out_h: int, out_w: int, mul: float) -> T:
        return upsample_crop_resize(x, scale, (crop_h, crop_w), (out_h, out_w), mul)

    @_ucr.register_fake
    def _(x, scale, crop_h, crop_w, out_h, out_w, mul):
        return x.new_empty((x.shape[0], x.shape[1], out_h, out_w))

    @custom_op("quber::resize_nearest", mutates_args=(), device_types="cuda")
    def _rn(masks: T, out_h: int, out_w: int) -> T:
        return resize_nearest(masks, (out_h, out_w))

    @_rn.register_fake
    def _(masks, out_h, out_w):
        return masks.new_empty(tuple(masks.shape[:-2]) + (out_h, out_w))

    @custom_op("quber::rle_decode", mutates_args=(), device_types="cuda")
    def _rled(counts: T, offsets: T, n: int, H: int, W: int, value: int, packed: bool) -> tuple[T, T]:
        return rle_decode(counts, offsets, n, H, W, value=value, packed=packed)

    @_rled.register_fake
    def _(counts, offsets, n, H, W, value, packed):
        out = counts.new_empty((n * H * W // 32,), dtype=torch.int32) if packed else counts.new_empty((n, H, W), dtype=torch.uint8)
        return out, counts.new_empty((1,), dtype=torch.int32)

    @custom_op("quber::rle_encode", mutates_args=(), device_types="cuda")
    def _rlee(masks: T, cap: int) -> tuple[T, T, T]:
        return rle_encode(masks, cap)

    @_rlee.register_fake
    def _(masks, cap):
        return (masks.new_empty((cap,), dtype=torch.int32), masks.new_empty((masks.shape[0] + 1,), dtype=torch.int64),
                masks.new_empty((1,), dtype=torch.int32))


_register()
