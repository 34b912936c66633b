"""Drop-in for the boundary utilities of explicit_error_estimation/util.py:62-99 (duplicated at
tools/ours/panoptic2eee.py:23-59): same names and arguments.  numpy in -> numpy out (as the
reference), torch CUDA tensor in -> torch CUDA tensor out; the arithmetic runs in quber_mask_boundary.
"""
import numpy as np
import torch

from . import ops


def _to_dev(a):
    if isinstance(a, torch.Tensor):
        return a.to(device="cuda", dtype=torch.uint8).contiguous(), False
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint8)).cuda(), True


def _back(t, as_numpy):
    if not as_numpy:
        return t                                   # device callers read ops.boundary_status() with their results
    out = t.cpu().numpy()                          # the copy synchronises: the call's status word is final
    ops.boundary_status(t.device, check=True)      # never hand back the buffers of a call whose TMA load was lost
    return out


def _pad16(m):
    """The kernels need W % 16 == 0: zero columns on the right behave exactly like the image border
    (erosion treats outside as 0), so padding then cropping does not change any in-image pixel."""
    W = m.shape[-1]
    pad = (-W) % 16
    return (torch.nn.functional.pad(m, (0, pad)) if pad else m), W


def masks_to_fg_mask(masks):
    m, np_ = _to_dev(masks)
    m, W = _pad16(m)
    _, _, fg = ops.mask_boundary(m[None], 1, per_instance=False, union=False, fg=True)
    return _back(fg[0, :, :W].contiguous(), np_)


def mask_to_boundary(mask, dilation_ratio=0.02):
    m, np_ = _to_dev(mask)
    H, W0 = m.shape
    d = ops.boundary_dilation(H, W0, dilation_ratio)
    m, W = _pad16(m)
    per, _, _ = ops.mask_boundary(m[None, None], d, per_instance=True, union=False, fg=False)
    return _back(per[0, 0, :, :W].contiguous(), np_)


def masks_to_boundary(masks, dilation_ratio=0.01):
    m, np_ = _to_dev(masks)
    _, H, W0 = m.shape
    d = ops.boundary_dilation(H, W0, dilation_ratio)
    m, W = _pad16(m)
    _, uni, _ = ops.mask_boundary(m[None], d, per_instance=False, union=True, fg=False)
    return _back(uni[0, :, :W].contiguous(), np_)


def quadruple_boundary_maps(gt_masks, perturbed_masks, dilation_ratio=0.01):
    """tools/ours/panoptic2eee.py:110-123: dict of the eight tp/tn/fp/fn_{mask,boundary} maps."""
    g, np_ = _to_dev(gt_masks)
    p, _ = _to_dev(perturbed_masks)
    _, H, W0 = g.shape
    d = ops.boundary_dilation(H, W0, dilation_ratio)
    g, W = _pad16(g)
    p, _ = _pad16(p)
    _, gb, gf = ops.mask_boundary(g[None], d, per_instance=False)
    _, pb, pf = ops.mask_boundary(p[None], d, per_instance=False)
    qm = ops.quadruple_maps(gf, pf)[0]
    qb = ops.quadruple_maps(gb, pb)[0]
    out = {}
    for i, n in enumerate(("tp", "tn", "fp", "fn")):
        out[f"{n}_mask"] = _back(qm[i, :, :W].contiguous(), np_)
        out[f"{n}_boundary"] = _back(qb[i, :, :W].contiguous(), np_)
    return out
