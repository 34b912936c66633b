"""Evaluation metrics on the device against the reference's own eval/evaluation.py outputs
(tests/golden/eval.npz) and the oracle's integer counts, through the C ABI.  B200 only."""
import numpy as np
import pytest
import torch

import oracle
from quber_b200 import evaluation as qe
from quber_b200 import ops, synth
from conftest import golden
from test_oracle_golden import EVAL_CASES, check_eval_metrics, _bits

pytestmark = pytest.mark.gpu


def _maps(case):
    g = golden("eval.npz")
    return g[f"{case}_pred"].astype(np.int32), g[f"{case}_gt"].astype(np.int32)


@pytest.mark.parametrize("case", EVAL_CASES)
def test_multilabel_metrics_bit_identical_to_reference(case):
    pr, gt = _maps(case)
    check_eval_metrics(case, qe.multilabel_metrics(pr, gt, 0, 1))
    # CUDA tensors in: same dictionary
    check_eval_metrics(case, qe.multilabel_metrics(torch.from_numpy(pr).cuda(), torch.from_numpy(gt).cuda(), 0, 1))


@pytest.mark.parametrize("case", ["normal", "more_pred", "holes", "c1"])
def test_eval_counts_match_oracle(case):
    pr, gt = _maps(case)
    radius = int(qe.bound_pixels(gt.shape))
    c = qe.eval_counts(pr, gt)[0]
    ref = oracle.eval_counts(pr, gt, radius)
    for k in ("pred_labels", "gt_labels", "inter", "area_pred", "area_gt", "bcount_pred", "bcount_gt", "btp"):
        np.testing.assert_array_equal(c[k], ref[k], err_msg=k)


def test_seg2bmap_and_boundary_overlap_golden():
    g = golden("eval.npz")
    shape = tuple(int(x) for x in g["bmap_shape"])
    masks, want = _bits(g, "bmap_masks", shape), _bits(g, "bmap_out", shape)
    got = ops.seg2bmap(torch.from_numpy(masks).cuda())
    np.testing.assert_array_equal(got.cpu().numpy().astype(bool), want)          # cv2.findContours + drawContours
    np.testing.assert_array_equal(qe.seg2bmap(masks[5]).astype(bool), want[5])  # the reference-named wrapper, numpy in / out
    pr, gt = _maps("holes")
    assert qe.boundary_overlap(pr == 3, gt == 3) == tuple(int(v) for v in g["overlap_holes_3"])
    assert qe.boundary_overlap(pr == 4, gt == 3) == tuple(int(v) for v in g["overlap_holes_4_vs_3"])


def test_batch_of_frames_equals_frame_by_frame():
    cases = ["normal", "more_pred", "holes", "pred_empty", "gt_empty", "both_empty"]
    H = max(_maps(c)[0].shape[0] for c in cases)
    W = max(_maps(c)[0].shape[1] for c in cases)
    same = [c for c in cases if _maps(c)[0].shape == (120, 160)]
    pr = np.stack([_maps(c)[0] for c in same])
    gt = np.stack([_maps(c)[1] for c in same])
    for c, m in zip(same, qe.multilabel_metrics_batch(pr, gt)):
        check_eval_metrics(c, m)
    assert H >= 120 and W >= 160


def test_many_labels_grow_the_matrices_and_bad_labels_raise():
    rng = np.random.default_rng(3)
    H, W = 96, 160
    gt = np.zeros((H, W), np.int32)
    for k in range(100):                                     # 100 small boxes: more than the default 64 labels
        y, x = (k // 10) * 9 + 1, (k % 10) * 15 + 1
        gt[y:y + 6, x:x + 10] = 1000 + 7 * k
    pr = np.roll(gt, 2, 1)
    pr[pr > 0] -= 500
    c = qe.eval_counts(pr, gt)[0]
    ref = oracle.eval_counts(pr, gt, int(qe.bound_pixels(gt.shape)))
    assert len(c["gt_labels"]) == 100
    for k in ("pred_labels", "gt_labels", "inter", "bcount_pred", "bcount_gt", "btp"):
        np.testing.assert_array_equal(c[k], ref[k], err_msg=k)
    bad = gt.copy(); bad[0, 0] = 70000
    with pytest.raises(ValueError):
        qe.eval_counts(pr, bad)
    neg = gt.copy(); neg[0, 0] = -1
    with pytest.raises(ValueError):
        qe.eval_counts(pr, neg)


def test_armbench_sized_frame_single_plane_path():
    """750 x 1333: the two bit planes of a full-frame bounding box do not fit one SM, so the flood fill keeps only `reach`
    in shared memory for labels with a large box (and both planes for the small ones)."""
    H, W = 750, 1333
    m = synth.ellipse_masks(8, H, W, seed=77, axes=(40, 160), margin=100)
    gt = np.zeros((H, W), np.int32)
    for k in range(8):
        gt[m[k] > 0] = k + 1
    gt[5:745, 5:8] = 20; gt[5:8, 5:1328] = 20; gt[742:745, 5:1328] = 20; gt[5:745, 1325:1328] = 20      # a frame-sized ring object
    pr = np.roll(gt, 3, 0)
    c = qe.eval_counts(pr, gt)[0]
    ref = oracle.eval_counts(pr, gt, int(qe.bound_pixels(gt.shape)))
    for k in ("inter", "bcount_pred", "bcount_gt", "btp"):
        np.testing.assert_array_equal(c[k], ref[k], err_msg=k)
