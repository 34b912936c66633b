"""Evaluation metrics of the refined masks — drop-in for eval/evaluation.py:22-274 (`boundary_overlap`,
`multilabel_metrics`), eval/utilities.py:672-696 (`seg2bmap`) and the assignment step of eval/munkres.py
(SURVEY §8-f rank 3).

What runs where:
  per-pixel work — label census, every (gt, prediction) intersection count, the one-pixel outer boundary of every
  mask (the reference: cv2.findContours per mask), the boundary matches inside a disk of `bound_pix` pixels (the
  reference: two cv2.dilate + two logical_and + two sums PER PAIR of masks)            -> quber_eval_counts (CUDA)
  the K x K' float64 P / R / F / IoU matrices, the Hungarian assignment and the final ratios — a few thousand flops on
  small integer matrices                                                              -> host, numpy float64, written
  with the reference's own operation order so that every returned number is bit-identical to the reference's.

`Munkres.compute` is this repo's own implementation of the published Kuhn-Munkres steps; it reproduces the reference
module's scan order (including its habit of taking the LAST uncovered zero of the first row that has one, munkres.py:536-560),
because the metric `obj_mIOU` depends on which of several equally cheap zero-overlap pairs gets assigned.
"""
import numpy as np
import torch

from . import ops

BACKGROUND_LABEL = 0
OBJECTS_LABEL = 1


# ---------------------------------------------------------------------------------------------- Hungarian assignment
class Munkres:
    """Kuhn-Munkres on a (possibly rectangular) cost matrix; `compute` returns [(row, col), ...] like eval/munkres.py."""

    def compute(self, cost_matrix):
        cost = np.array(cost_matrix, dtype=np.float64)
        rows, cols = cost.shape
        n = max(rows, cols)
        C = np.zeros((n, n), np.float64)                      # padded with zeros (munkres.py:271-317)
        C[:rows, :cols] = cost
        C -= C.min(axis=1, keepdims=True)                     # step 1: subtract the row minima
        star = np.zeros((n, n), bool)
        prime = np.zeros((n, n), bool)
        row_cov, col_cov = np.zeros(n, bool), np.zeros(n, bool)
        for i in range(n):                                    # step 2: greedy starring in row-major order
            for j in range(n):
                if C[i, j] == 0 and not col_cov[j] and not row_cov[i]:
                    star[i, j] = True
                    row_cov[i] = col_cov[j] = True
        row_cov[:] = False
        col_cov[:] = False
        while True:
            col_cov |= star.any(axis=0)                       # step 3: cover the columns of the starred zeros
            if int(star.sum()) >= n:
                break
            while True:                                       # step 4
                free = (C == 0) & ~row_cov[:, None] & ~col_cov[None, :]
                has = free.any(axis=1)
                if not has.any():
                    m = C[~row_cov][:, ~col_cov].min()        # step 6: smallest uncovered value
                    C[row_cov, :] += m                        # added to covered rows first, then subtracted from uncovered
                    C[:, ~col_cov] -= m                       # columns: the reference's order on doubly affected cells
                    continue
                i = int(np.argmax(has))                       # first row with an uncovered zero ...
                j = int(np.flatnonzero(free[i])[-1])          # ... and the LAST such zero in it (munkres.py:536-560)
                prime[i, j] = True
                sc = np.flatnonzero(star[i])
                if sc.size:
                    row_cov[i] = True
                    col_cov[sc[0]] = False
                    continue
                path = [(i, j)]                               # step 5: alternating path of starred / primed zeros
                while True:
                    r = np.flatnonzero(star[:, path[-1][1]])
                    if not r.size:
                        break
                    path.append((int(r[0]), path[-1][1]))
                    c = np.flatnonzero(prime[path[-1][0]])
                    path.append((path[-1][0], int(c[0])))
                for (r_, c_) in path:
                    star[r_, c_] = not star[r_, c_]
                row_cov[:] = False
                col_cov[:] = False
                prime[:] = False
                break
        return [(i, j) for i in range(rows) for j in range(cols) if star[i, j]]


# ---------------------------------------------------------------------------------------------- device counts
def bound_pixels(shape, bound_th=0.003):
    """evaluation.py:34-35."""
    return bound_th if bound_th >= 1 else np.ceil(bound_th * np.linalg.norm(shape))


def eval_counts(prediction, gt, bound_th=0.003, compute_boundary_stuff=True, max_labels=64):
    """prediction, gt: [H,W] (or [B,H,W]) integer label maps, numpy or CUDA tensors -> list of per-image dicts of numpy
    integer arrays (labels, inter [n_gt,n_pred], areas, boundary counts, boundary tps [n_gt,n_pred,2])."""
    def dev(a):
        t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
        return t.to(device="cuda", dtype=torch.int32).contiguous()
    p, g = dev(prediction), dev(gt)
    if p.dim() == 2:
        p, g = p[None], g[None]
    if p.shape != g.shape:
        raise ValueError("prediction and gt must have the same shape")
    radius = int(bound_pixels(p.shape[-2:], bound_th))
    while True:
        r = ops.eval_counts(p, g, radius, max_labels, boundary=compute_boundary_stuff)
        st = int(r["status"][0])
        if st & 1 and max_labels < 255:                       # more labels than the matrices hold: grow and redo
            max_labels = min(255, max_labels * 2)
            continue
        if st:
            raise ValueError("quber_eval_counts: %s" % ("more than 255 labels in a map" if st & 1 else "labels must lie in [0, 65535]"))
        break
    out = []
    host = {k: v.cpu().numpy() for k, v in r.items()}
    for b in range(p.shape[0]):
        n_p, n_g = int(host["n_pred"][b]), int(host["n_gt"][b])
        out.append(dict(pred_labels=host["pred_labels"][b, :n_p], gt_labels=host["gt_labels"][b, :n_g],
                        inter=host["inter"][b, :n_g, :n_p].astype(np.int64), area_pred=host["area_pred"][b, :n_p].astype(np.int64),
                        area_gt=host["area_gt"][b, :n_g].astype(np.int64), bcount_pred=host["bcount_pred"][b, :n_p].astype(np.int64),
                        bcount_gt=host["bcount_gt"][b, :n_g].astype(np.int64), btp=host["btp"][b, :n_g, :n_p].astype(np.int64)))
    return out


# ---------------------------------------------------------------------------------------------- metric assembly (host)
def _edge_case(num_labels_pred, num_labels_gt):
    """evaluation.py:104-160."""
    if num_labels_pred == 0 and num_labels_gt > 0:
        v, p, r, pct = 0., 1., 0., 0.
    elif num_labels_pred > 0 and num_labels_gt == 0:
        v, p, r, pct = 0., 0., 1., 0.
    else:
        v, p, r, pct = 1., 1., 1., 1.
    out = {}
    for grp in ("Objects", "Boundary", "Objects OSN", "Boundary OSN"):
        out[f"{grp} F-measure"], out[f"{grp} Precision"], out[f"{grp} Recall"] = v, p, r
    out.update({"obj_detected": num_labels_pred, "obj_detected_075": 0., "obj_gt": num_labels_gt,
                "obj_detected_075_percentage": pct, "obj_detected_075_percentage_normalized": pct})
    return out


def metrics_from_counts(c, obj_detect_threshold=0.75, compute_boundary_stuff=True):
    """evaluation.py:83-274 from the integer counts of one image: the same float64 expressions in the same order."""
    n_gt, n_pred = len(c["gt_labels"]), len(c["pred_labels"])
    if n_pred == 0 or n_gt == 0:
        return _edge_case(n_pred, n_gt)
    with np.errstate(divide="ignore", invalid="ignore"):
        inter = c["inter"].astype(np.int64)
        obj_tps = inter.astype(np.float64)
        union = c["area_gt"][:, None] + c["area_pred"][None, :] - inter
        obj_iou = inter / union                                # :184 int64 / int64
        obj_union = union.astype(np.float64)
        obj_P = inter / c["area_pred"][None, :]                 # :194
        obj_R = inter / c["area_gt"][:, None]                   # :195
        obj_F = (2 * obj_P * obj_R) / (obj_P + obj_R)           # :196
        if compute_boundary_stuff:
            btp = c["btp"].astype(np.float64)
            bound_P = btp[..., 0] / c["bcount_pred"].astype(np.float64)[None, :]     # :203
            bound_R = btp[..., 1] / c["bcount_gt"].astype(np.float64)[:, None]       # :204
            bound_F = (2 * bound_P * bound_R) / (bound_P + bound_R)
        obj_F[np.isnan(obj_F)] = 0
        assignments = Munkres().compute(obj_F.max() - obj_F.copy())                  # :210-212
        idx = tuple(np.array(assignments).T)
        num_obj_detected = sum(1 for a in assignments if obj_F[a] > obj_detect_threshold)
        # np.sum(prediction.clip(0,1) == OBJECTS_LABEL): every pixel with a positive label (:220-221)
        n_pred_px = int(c["area_pred"][c["pred_labels"] > 0].sum())
        n_gt_px = int(c["area_gt"][c["gt_labels"] > 0].sum())
        precision = np.sum(obj_tps[idx]) / n_pred_px
        recall = np.sum(obj_tps[idx]) / n_gt_px
        F_measure = (2 * precision * recall) / (precision + recall)
        if np.isnan(F_measure):
            F_measure = 0
        obj_F_osn = np.sum(obj_F[idx]) / max(n_pred, n_gt)
        obj_P_osn = np.sum(obj_P[idx]) / n_pred
        obj_R_osn = np.sum(obj_R[idx]) / n_gt
        obj_iou_osn = np.mean(obj_iou[idx])
        i_over_u = np.sum(obj_tps[idx]) / np.sum(obj_union[idx])
        if compute_boundary_stuff:
            bound_F[np.isnan(bound_F)] = 0
            boundary_precision = np.sum(btp[idx][:, 0]) / np.sum(c["bcount_pred"].astype(np.float64))
            boundary_recall = np.sum(btp[idx][:, 1]) / np.sum(c["bcount_gt"].astype(np.float64))
            boundary_F_measure = (2 * boundary_precision * boundary_recall) / (boundary_precision + boundary_recall)
            if np.isnan(boundary_F_measure):
                boundary_F_measure = 0
            bound_F_osn = np.sum(bound_F[idx]) / max(n_pred, n_gt)
            bound_P_osn = np.sum(bound_P[idx]) / n_pred
            bound_R_osn = np.sum(bound_R[idx]) / n_gt
        else:
            boundary_F_measure = boundary_precision = boundary_recall = bound_F_osn = bound_P_osn = bound_R_osn = None
    return {"Objects F-measure": F_measure, "Objects Precision": precision, "Objects Recall": recall,
            "Boundary F-measure": boundary_F_measure, "Boundary Precision": boundary_precision, "Boundary Recall": boundary_recall,
            "Objects OSN F-measure": obj_F_osn, "Objects OSN Precision": obj_P_osn, "Objects OSN Recall": obj_R_osn,
            "Boundary OSN F-measure": bound_F_osn, "Boundary OSN Precision": bound_P_osn, "Boundary OSN Recall": bound_R_osn,
            "obj_detected": n_pred, "obj_detected_075": num_obj_detected, "obj_gt": n_gt,
            "obj_detected_075_percentage": num_obj_detected / n_gt,
            "obj_detected_075_percentage_normalized": num_obj_detected / max(n_gt, n_pred),
            "obj_mIOU_osn": obj_iou_osn, "obj_mIOU": i_over_u}


# ---------------------------------------------------------------------------------------------- the reference's names
def multilabel_metrics(prediction, gt, i=0, N=1, obj_detect_threshold=0.75, compute_boundary_stuff=True, verbose=False):
    """eval/evaluation.py:57-274: same arguments, same dictionary."""
    if verbose:
        print(f"Running example {i}/{N}...")
    c = eval_counts(prediction, gt, compute_boundary_stuff=compute_boundary_stuff)[0]
    return metrics_from_counts(c, obj_detect_threshold, compute_boundary_stuff)


def multilabel_metrics_batch(predictions, gts, obj_detect_threshold=0.75, compute_boundary_stuff=True):
    """[B,H,W] label maps (numpy or CUDA) -> list of B metric dictionaries; one kernel sequence for the batch."""
    return [metrics_from_counts(c, obj_detect_threshold, compute_boundary_stuff)
            for c in eval_counts(predictions, gts, compute_boundary_stuff=compute_boundary_stuff)]


def seg2bmap(seg, return_contour=False):
    """eval/utilities.py:672-696 (the boundary map; the ordered contour of `return_contour` is a CPU-side cv2 notion
    and is not produced here)."""
    if return_contour:
        raise NotImplementedError("seg2bmap(return_contour=True) is not part of the evaluation path")
    as_np = not isinstance(seg, torch.Tensor)
    t = (torch.from_numpy(np.ascontiguousarray(np.asarray(seg) != 0)) if as_np else (seg != 0)).to(device="cuda", dtype=torch.int32)
    b = ops.seg2bmap(t[None])[0]
    return b.cpu().numpy() if as_np else b


def boundary_overlap(predicted_mask, gt_mask, bound_th=0.003):
    """eval/evaluation.py:22-55 -> (precision true positives, recall true positives)."""
    p = np.asarray(predicted_mask) if not isinstance(predicted_mask, torch.Tensor) else predicted_mask
    assert np.atleast_3d(np.empty(tuple(p.shape))).shape[2] == 1
    as_i32 = lambda a: (torch.from_numpy(np.ascontiguousarray(np.asarray(a) != 0)) if not isinstance(a, torch.Tensor) else (a != 0)).to(torch.int32)
    c = eval_counts(as_i32(predicted_mask), as_i32(gt_mask), bound_th=bound_th)[0]
    if len(c["pred_labels"]) == 0 or len(c["gt_labels"]) == 0:
        return 0, 0
    return int(c["btp"][0, 0, 0]), int(c["btp"][0, 0, 1])
