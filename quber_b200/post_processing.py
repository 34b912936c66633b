"""Drop-in for maskrefiner/modeling/mask_refiner/post_processing.py (same names, arguments, error
behaviour), computed by the sm_100a kernels behind quber_post_process.  Inputs are CUDA tensors
shaped as the reference's ([1,H,W] / [2,H,W]); outputs have the reference's shapes and dtypes.

For a whole batch in one launch sequence use quber_b200.ops.post_process directly.
"""
import torch

from . import ops

_CAP = 1024     # capacity for top_k=None (find_instance_center then returns every NMS survivor)


def _as4(t, c):
    if t.dim() != 3:
        raise ValueError("un-supported dimension: {}.".format(t.dim()))
    return t.reshape(1, c, t.shape[-2], t.shape[-1]).float().contiguous()


def find_instance_center(center_heatmap, threshold=0.1, nms_kernel=3, top_k=None):
    """post_processing.py:9-41 -> [K,2] int64 (y,x), raster order.  Unlike the reference it does not
    modify `center_heatmap` in place (the reference's F.threshold / masked assign at :22,:29 do)."""
    c = _as4(center_heatmap, 1)
    z = torch.zeros_like(c)
    r = ops.post_process(z, c, torch.zeros((1, 2) + tuple(c.shape[2:]), device=c.device), None, scale=1,
                         threshold=threshold, nms_kernel=nms_kernel, top_k=top_k if top_k is not None else _CAP,
                         max_instances=1, want=(), no_topk=top_k is None)
    n = int(r["n_centers"][0])
    if top_k is None and int(r["status"][0]) & 2:
        raise ValueError(f"more than {_CAP} centres with top_k=None")
    return r["centers_yx"][0, :n].long()


def group_pixels(center_points, offsets):
    """post_processing.py:44-76 -> [1,H,W] int64 in 1..K."""
    o = _as4(offsets, 2)
    K = center_points.shape[0]
    if K == 0 or K > 1024:
        raise ValueError("group_pixels needs 1..1024 centre points")   # torch.argmin on an empty dim raises too
    H, W = o.shape[2:]
    ones = torch.ones((1, 1, H, W), device=o.device)
    ctr = torch.zeros((1, K, 2), dtype=torch.int32, device=o.device)
    ctr[0] = center_points.to(torch.int32)
    r = ops.post_process(ones, ones, o, None, scale=1, top_k=K, max_instances=1, want=("ins_seg",), centers=ctr,
                         n_centers=torch.tensor([K], dtype=torch.int32, device=o.device))
    return r["ins_seg"].long()


def get_instance_segmentation(center_heatmap, offsets, foreground_seg, threshold=0.1, nms_kernel=3, top_k=None):
    """post_processing.py:79-107 -> (foreground_seg * ins_seg [1,H,W] (dtype of foreground_seg), centres [1,K,2])."""
    c, o, f = _as4(center_heatmap, 1), _as4(offsets, 2), _as4(foreground_seg, 1)
    r = ops.post_process(f, c, o, None, scale=1, threshold=threshold, nms_kernel=nms_kernel,
                         top_k=top_k if top_k is not None else _CAP, max_instances=1, want=("ins_seg",),
                         no_topk=top_k is None)
    n = int(r["n_centers"][0])
    return r["ins_seg"].to(foreground_seg.dtype), r["centers_yx"][0, :n].long().unsqueeze(0)


def get_panoptic_segmentation(foreground_seg, center_heatmap, offsets, thing_ids, label_divisor, stuff_area, void_label,
                              threshold=0.1, nms_kernel=7, top_k=200):
    """post_processing.py:165-221 -> (panoptic [1,H,W] in foreground_seg's dtype, centres [1,K,2]).

    `foreground_seg` is the reference's `r.sigmoid().round()` (0/1).  `thing_ids` must be the
    uoais-sim metadata's dict_values([0]) (register_uoais_sim_panoptic.py:189-193): class 0 is skipped
    as "thing" and class 1 outside every instance is the only stuff candidate (:152-160) — the
    kernels implement exactly that case.
    """
    if center_heatmap.dim() != 3:
        raise ValueError("Center prediction with un-supported dimension: {}.".format(center_heatmap.dim()))
    if offsets.dim() != 3:
        raise ValueError("Offset prediction with un-supported dimension: {}.".format(offsets.dim()))
    if sorted(int(t) for t in thing_ids) != [0]:
        raise ValueError("quber_b200 implements the uoais-sim metadata only (thing_ids == [0])")
    c, o, f = _as4(center_heatmap, 1), _as4(offsets, 2), _as4(foreground_seg, 1)
    r = ops.post_process(f, c, o, None, scale=1, threshold=threshold, nms_kernel=nms_kernel, top_k=top_k,
                         stuff_area=stuff_area, label_divisor=label_divisor, void_label=float(void_label),
                         max_instances=1, want=("panoptic",))
    n = int(r["n_centers"][0])
    return r["panoptic"].to(foreground_seg.dtype), r["centers_yx"][0, :n].long().unsqueeze(0)
