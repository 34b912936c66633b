"""CPU oracle for the maskrefiner hot path — TEST INFRASTRUCTURE, NOT PRODUCT.

ctypes front-end of ``oracle/quber_oracle.c`` (a C restatement of the reference's
arithmetic, each function citing the reference file:line it follows).  Only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; ``quber_b200`` never does.

Parity pinning: ``tests/golden/*.npz`` are outputs of the reference's own Python
(``oracle/gen_golden.py`` imports it from /root/reference, which exists only in
the build container) and ``tests/test_oracle_golden.py`` checks this oracle
against them.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libquber_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "quber_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.oracle_boundary_dilation.restype = ctypes.c_int
        _lib.oracle_boundary_dilation.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_double]
        _lib.oracle_find_instance_center.restype = ctypes.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def gaussian_table(sigma=10):
    n = 6 * sigma + 3
    g = np.empty((n, n), np.float32)
    lib().oracle_gaussian_table(ctypes.c_int(sigma), _p(g))
    return g


def encode_initial_masks(masks, sigma=10, return_centres=False):
    """maskrefiner/predictor.py:304-357 → [3,H,W] float32."""
    masks = _c(masks, np.uint8)
    K, H, W = masks.shape
    out = np.empty((3, H, W), np.float32)
    ctr = np.empty((K, 2), np.float64)
    lib().oracle_encode_initial_masks(_p(masks), K, H, W, sigma, _p(out), _p(ctr))
    return (out, ctr) if return_centres else out


def boundary_dilation(H, W, ratio):
    return int(lib().oracle_boundary_dilation(H, W, ctypes.c_double(ratio)))


def mask_to_boundary(mask, dilation_ratio=0.02):
    """explicit_error_estimation/util.py:72-90."""
    mask = _c(mask, np.uint8)
    H, W = mask.shape
    out = np.empty_like(mask)
    lib().oracle_mask_to_boundary(_p(mask), H, W, boundary_dilation(H, W, dilation_ratio), _p(out))
    return out


def masks_to_fg_mask(masks):
    """explicit_error_estimation/util.py:62-68."""
    masks = _c(masks, np.uint8)
    K, H, W = masks.shape
    out = np.empty((H, W), np.uint8)
    lib().oracle_masks_to_fg_mask(_p(masks), K, H, W, _p(out))
    return out


def masks_to_boundary(masks, dilation_ratio=0.01, return_per_instance=False):
    """explicit_error_estimation/util.py:92-99."""
    masks = _c(masks, np.uint8)
    K, H, W = masks.shape
    out = np.empty((H, W), np.uint8)
    per = np.empty((K, H, W), np.uint8) if return_per_instance else None
    lib().oracle_masks_to_boundary(_p(masks), K, H, W, boundary_dilation(H, W, dilation_ratio), _p(per), _p(out))
    return (out, per) if return_per_instance else out


def quadruple_maps(gt, pred):
    """tools/ours/panoptic2eee.py:110-123 → [4,H,W] uint8 (TP, TN, FP, FN)."""
    gt, pred = _c(gt, np.uint8), _c(pred, np.uint8)
    out = np.empty((4,) + gt.shape, np.uint8)
    lib().oracle_quadruple_maps(_p(gt), _p(pred), ctypes.c_size_t(gt.size), _p(out))
    return out


def upsample_bilinear(src, scale=4, mul=1.0):
    """model.py:689-708: F.interpolate(scale_factor=scale, bilinear, align_corners=False) on [C,h,w]."""
    src = _c(src, np.float32)
    C, h, w = src.shape
    out = np.empty((C, h * scale, w * scale), np.float32)
    lib().oracle_upsample_bilinear(_p(src), C, h, w, scale, ctypes.c_float(mul), _p(out))
    return out


def find_instance_center(heat, threshold=0.1, nms_kernel=3, top_k=None):
    """post_processing.py:9-41 → [K,2] int32 (y,x)."""
    heat = _c(heat, np.float32).reshape(heat.shape[-2:])
    H, W = heat.shape
    out = np.empty((H * W, 2), np.int32)
    n = lib().oracle_find_instance_center(_p(heat), H, W, ctypes.c_float(threshold), nms_kernel,
                                          top_k if top_k else 0, _p(out), H * W)
    return out[:n].copy()


def group_pixels(centres_yx, offsets):
    """post_processing.py:44-76 → [H,W] int32 in 1..K."""
    centres_yx = _c(centres_yx, np.int32)
    offsets = _c(offsets, np.float32)
    _, H, W = offsets.shape
    out = np.empty((H, W), np.int32)
    lib().oracle_group_pixels(_p(centres_yx), len(centres_yx), _p(offsets), H, W, _p(out))
    return out


class _Counts(ctypes.Structure):
    _fields_ = [("n_centres", ctypes.c_int32), ("n_instances", ctypes.c_int32)]


def post_process(fg_logit, centre, offset, eee=None, threshold=0.3, nms_kernel=7, top_k=200,
                 min_area=512, stuff_area=2048, label_divisor=1000, void_label=-1.0, cap=256):
    """get_panoptic_segmentation (post_processing.py:165-221) + instance extraction
    (model.py:313-356) + error-map decode (eval/eval_utils.py:308-328) on full-res maps."""
    fg_logit = _c(fg_logit, np.float32).reshape(fg_logit.shape[-2:])
    H, W = fg_logit.shape
    centre = _c(centre, np.float32).reshape(H, W)
    offset = _c(offset, np.float32).reshape(2, H, W)
    C = 0
    onehot = None
    if eee is not None:
        eee = _c(eee, np.float32)
        C = eee.shape[0]
        onehot = np.empty((C, H, W), np.uint8)
    pan = np.empty((H, W), np.float32)
    labels = np.zeros(cap, np.int32)
    masks = np.zeros((cap, H, W), np.uint8)
    scores = np.zeros(cap, np.float32)
    boxes = np.zeros((cap, 4), np.float32)
    ctr = np.zeros((max(top_k, 1) if top_k else H * W, 2), np.int32)
    cnt = _Counts()
    lib().oracle_post_process(_p(fg_logit), _p(centre), _p(offset), _p(eee), C, H, W,
                              ctypes.c_float(threshold), nms_kernel, top_k if top_k else 0, min_area,
                              stuff_area, label_divisor, ctypes.c_float(void_label), cap, _p(pan),
                              _p(labels), _p(masks), _p(scores), _p(boxes), _p(onehot), _p(ctr),
                              ctypes.byref(cnt))
    n = min(cnt.n_instances, cap)
    return dict(panoptic=pan, labels=labels[:n].copy(), masks=masks[:n].copy(), scores=scores[:n].copy(),
                boxes=boxes[:n].copy(), onehot=onehot, centres=ctr[:cnt.n_centres].copy(),
                n_centres=cnt.n_centres, n_instances=cnt.n_instances)


def msda_forward(value, shapes, starts, loc, attn):
    """OPS/src/cuda/ms_deform_im2col_cuda.cuh:242-304 on CPU; float32 or float64 by value.dtype."""
    dt = np.float64 if value.dtype == np.float64 else np.float32
    value, loc, attn = _c(value, dt), _c(loc, dt), _c(attn, dt)
    shapes, starts = _c(shapes, np.int64), _c(starts, np.int64)
    B, S, M, D = value.shape
    _, Q, _, L, P, _ = loc.shape
    out = np.empty((B, Q, M * D), dt)
    fn = lib().oracle_msda_forward_f64 if dt == np.float64 else lib().oracle_msda_forward_f32
    fn(_p(value), _p(shapes), _p(starts), _p(loc), _p(attn), B, S, M, D, L, Q, P, _p(out))
    return out


def filter_instances(masks, fg_mask=None, zero_depth=None, thr=0.3):
    """eval/refiner_model.py:271-288, restated with numpy: masks [N,H,W] bool -> (kept masks [N',H,W] bool, keep [N] bool).
    keep j iff np.sum(mask & fg) / np.sum(mask) > thr (int64 / int64 true division); the zero-depth pixels of the
    kept masks are cleared afterwards (:279-288)."""
    masks = np.asarray(masks).astype(bool)
    keep = np.ones(len(masks), bool)
    if fg_mask is not None:
        fgb = np.asarray(fg_mask).astype(bool)
        with np.errstate(invalid="ignore", divide="ignore"):
            keep = np.array([np.sum(np.bitwise_and(m, fgb)) / np.sum(m) > thr for m in masks], bool)
    out = masks[keep].copy()
    if zero_depth is not None:
        out[:, np.asarray(zero_depth).astype(bool)] = False
    return out, keep


# ---- (f)4 COCO RLE (pycocotools.mask restated; see the header of that section in quber_oracle.c) ----------------
def rle_encode(mask):
    """pycocotools.mask.encode for one [H,W] mask -> uncompressed counts (uint32), column-major runs starting with zeros."""
    mask = _c(mask, np.uint8)
    H, W = mask.shape
    cnts = np.empty(H * W + 1, np.uint32)
    lib().oracle_rle_encode.restype = ctypes.c_size_t
    m = lib().oracle_rle_encode(_p(mask), H, W, _p(cnts), ctypes.c_size_t(cnts.size))
    return cnts[:m].copy()


def rle_decode(counts, H, W):
    """pycocotools.mask.decode for one RLE -> [H,W] uint8 0/1."""
    counts = _c(counts, np.uint32)
    out = np.empty((H, W), np.uint8)
    if lib().oracle_rle_decode(_p(counts), ctypes.c_size_t(counts.size), H, W, _p(out)) != 0:
        raise ValueError("RLE counts exceed the mask size")
    return out


def rle_to_string(counts):
    counts = _c(counts, np.uint32)
    buf = ctypes.create_string_buffer(6 * counts.size + 1)
    lib().oracle_rle_to_string.restype = ctypes.c_size_t
    n = lib().oracle_rle_to_string(_p(counts), ctypes.c_size_t(counts.size), buf)
    return buf.raw[:n]


def rle_from_string(s):
    if isinstance(s, str):
        s = s.encode("ascii")
    cnts = np.empty(max(len(s), 1), np.uint32)
    lib().oracle_rle_from_string.restype = ctypes.c_size_t
    m = lib().oracle_rle_from_string(ctypes.c_char_p(s), _p(cnts))
    return cnts[:m].copy()


def upsample_crop_resize(src, scale, crop_hw, out_hw, mul=1.0):
    """model.py:689-708 + detectron2 sem_seg_postprocess (model.py:266-289): [C,h,w] -> [C,out_h,out_w]."""
    src = _c(src, np.float32)
    C, h, w = src.shape
    out = np.empty((C, int(out_hw[0]), int(out_hw[1])), np.float32)
    lib().oracle_upsample_crop_resize(_p(src), C, h, w, int(scale), ctypes.c_float(mul), int(crop_hw[0]), int(crop_hw[1]),
                                      int(out_hw[0]), int(out_hw[1]), _p(out))
    return out


def resize_nearest(mask, out_hw):
    """cv2.resize(mask, (out_w, out_h), interpolation=cv2.INTER_NEAREST) for one [h,w] uint8 plane."""
    mask = _c(mask, np.uint8)
    out = np.empty((int(out_hw[0]), int(out_hw[1])), np.uint8)
    lib().oracle_resize_nearest_u8(_p(mask), mask.shape[0], mask.shape[1], out.shape[0], out.shape[1], _p(out))
    return out


# ---- (f)3 evaluation metrics (eval/evaluation.py, eval/utilities.py:672) --------------------------------------------
def seg2bmap(mask):
    """eval/utilities.py:672-696 for one [H,W] mask -> uint8 0/1 boundary map."""
    mask = _c(np.asarray(mask) != 0, np.uint8)
    out = np.empty_like(mask)
    lib().oracle_seg2bmap(_p(mask), mask.shape[0], mask.shape[1], _p(out))
    return out


def eval_counts(prediction, gt, radius, cap=256):
    """The integer counts multilabel_metrics (eval/evaluation.py:57-274) is made of, for one pair of label maps."""
    prediction, gt = _c(prediction, np.int32), _c(gt, np.int32)
    H, W = gt.shape
    pl, gl = np.zeros(cap, np.int32), np.zeros(cap, np.int32)
    n_p, n_g = ctypes.c_int32(), ctypes.c_int32()
    inter, btp = np.zeros((cap, cap), np.int64), np.zeros((cap, cap, 2), np.int64)
    ap, ag, bp, bg = (np.zeros(cap, np.int64) for _ in range(4))
    rc = lib().oracle_eval_counts(_p(prediction), _p(gt), H, W, cap, int(radius), _p(pl), ctypes.byref(n_p), _p(gl), ctypes.byref(n_g),
                                  _p(inter), _p(ap), _p(ag), _p(bp), _p(bg), _p(btp))
    if rc != 0:
        raise ValueError("more labels than cap")
    np_, ng = n_p.value, n_g.value
    return dict(pred_labels=pl[:np_].copy(), gt_labels=gl[:ng].copy(), inter=inter[:ng, :np_].copy(), area_pred=ap[:np_].copy(),
                area_gt=ag[:ng].copy(), bcount_pred=bp[:np_].copy(), bcount_gt=bg[:ng].copy(), btp=btp[:ng, :np_].copy())
