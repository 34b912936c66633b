"""Parity at the shapes of BASELINE.json's other configurations: C3 = OSD-like dense clutter
(60 overlapping masks per frame, bf16 MSDA values), C4 = armbench-class frames (1080x1920 resized to 750x1333,
network output 752x1336, Swin-T pixel-decoder levels 96x168 / 48x84 / 24x42, S = 21 168).  The CUDA path is
called through the C ABI and compared with the oracle on the same seeded inputs.  B200 only."""
import numpy as np
import pytest
import torch

import oracle
from quber_b200 import boundary as bd
from quber_b200 import ops, synth
from test_gpu_parity import _check_post, cu

pytestmark = pytest.mark.gpu


def _post_vs_oracle(masks, seed, max_instances):
    ho = synth.head_outputs(masks, seed=seed)
    low = np.concatenate([ho["foreground"], ho["center"], ho["offset"], ho["eee_boundary"]], 0)
    r = ops.post_process(cu(low[None, 0:1]), cu(low[None, 1:2]), cu(low[None, 2:4]), cu(low[None, 4:8]), scale=4,
                         max_instances=max_instances)
    full = oracle.upsample_bilinear(low, 4)
    np.testing.assert_array_equal(r["sem_seg"][0].cpu().numpy(), full[0:1])
    np.testing.assert_array_equal(r["eee_full"][0].cpu().numpy(), full[4:8])
    full[2:4] *= 4.0
    ref = oracle.post_process(full[0], full[1], full[2:4], full[4:8])
    _check_post(r, ref, masks.shape[1], masks.shape[2])
    return len(ref["labels"])


# ------------------------------------------------------------------ C3: dense clutter, K = 60
@pytest.mark.parametrize("seed", [300, 301])
def test_c3_clutter_encode_boundary_post(seed):
    K, H, W = 60, 480, 640
    masks = synth.ellipse_masks(K, H, W, seed=seed, axes=(10, 40), margin=40)          # overlaps allowed
    assert int((masks != 0).sum(0).max()) >= 2                                         # and present
    dm = cu(masks)[None]
    np.testing.assert_array_equal(ops.encode_initial_masks(dm)[0].cpu().numpy(), oracle.encode_initial_masks(masks))
    for ratio in (0.01, 0.02):
        d = ops.boundary_dilation(H, W, ratio)
        per, uni, fg = ops.mask_boundary(dm, d)
        u, p = oracle.masks_to_boundary(masks, ratio, return_per_instance=True)
        np.testing.assert_array_equal(per[0].cpu().numpy(), p)
        np.testing.assert_array_equal(uni[0].cpu().numpy(), u)
        np.testing.assert_array_equal(fg[0].cpu().numpy(), oracle.masks_to_fg_mask(masks))
    n = _post_vs_oracle(masks, seed, max_instances=128)
    assert n >= 10                                                                      # error-guided fusion stress: many instances


def test_c3_msda_bf16_values():
    """bf16 value / output, fp32 locations, weights and accumulation: |err| <= 1e-4 absolute against the fp32 oracle
    on values of magnitude <= 1e-2 (bf16 keeps 8 bits: 4e-5 per product, averaged by the convex weights), stated
    tolerance for the bf16 variant (BASELINE.json north_star)."""
    inp = synth.msda_inputs(B=2, mode="module", seed=31)
    v, loc, at = cu(inp["value"]), cu(inp["loc"]), cu(inp["attn"])
    sh, st = cu(inp["shapes"]), cu(inp["starts"])
    out = ops.msda_forward(v.bfloat16(), sh, st, loc, at)
    assert out.dtype == torch.bfloat16
    sel = np.r_[0:200, 4700:4900, 6200:6300]
    vq = v.bfloat16().float().cpu().numpy()                                             # the values the kernel actually reads
    for b in range(2):
        ref_q = oracle.msda_forward(vq[b:b + 1], inp["shapes"], inp["starts"], inp["loc"][b:b + 1, sel], inp["attn"][b:b + 1, sel])
        ref = oracle.msda_forward(inp["value"][b:b + 1], inp["shapes"], inp["starts"], inp["loc"][b:b + 1, sel], inp["attn"][b:b + 1, sel])
        got = out[b, sel].float().cpu().numpy()
        assert np.abs(got - ref[0]).max() <= 1e-4
        assert np.abs(got - ref_q[0]).max() <= 4e-5                                     # only the output rounding is left


# ------------------------------------------------------------------ C4: armbench-class frames
def test_c4_armbench_boundary_and_encode_750x1333():
    H, W, K = 750, 1333, 30                                                             # eval/refiner_model.py:228
    masks = synth.ellipse_masks(K, H, W, seed=40, axes=(40, 120), margin=120)
    for ratio in (0.01, 0.02):
        np.testing.assert_array_equal(bd.masks_to_boundary(masks, ratio), oracle.masks_to_boundary(masks, ratio))
    np.testing.assert_array_equal(bd.mask_to_boundary(masks[3], 0.02), oracle.mask_to_boundary(masks[3], 0.02))
    np.testing.assert_array_equal(bd.masks_to_fg_mask(masks), oracle.masks_to_fg_mask(masks))
    enc = ops.encode_initial_masks_any_width(cu(masks)[None])[0].cpu().numpy()
    ref = oracle.encode_initial_masks(masks)
    np.testing.assert_array_equal(enc[0], ref[0])                                       # centre heat-map
    np.testing.assert_array_equal(enc[1], ref[1])                                       # (c_y - y) / H
    # (c_x - x) / W: W % 16 != 0 -> computed on the padded width and rescaled, <= 1 ulp (ops.encode_initial_masks_any_width)
    ulp = np.spacing(np.abs(ref[2]).astype(np.float32))
    assert (np.abs(enc[2] - ref[2]) <= ulp).all()
    assert np.array_equal(enc[2] != 0, ref[2] != 0)


def test_c4_armbench_post_752x1336():
    """The network's x4 output for a 750x1333 frame is 752x1336 (188x334 at stride 4)."""
    H, W, K = 752, 1336, 30
    masks = synth.ellipse_masks(K, H, W, seed=41, axes=(40, 120), margin=120)
    n = _post_vs_oracle(masks, 41, max_instances=64)
    assert n >= 5


def test_c4_msda_swin_levels():
    """Swin-T pixel decoder at 750x1333 (size divisibility 32 -> 768x1344): levels 96x168, 48x84, 24x42; S = 21 168."""
    shapes = ((96, 168), (48, 84), (24, 42))
    inp = synth.msda_inputs(B=1, shapes=shapes, mode="module", seed=42)
    S = sum(h * w for h, w in shapes)
    assert inp["value"].shape[1] == S == 21168
    out = ops.msda_forward(cu(inp["value"]), cu(inp["shapes"]), cu(inp["starts"]), cu(inp["loc"]), cu(inp["attn"]))
    sel = np.r_[0:168, 8000:8100, 16128:16228, 20160:20260, S - 50:S]
    ref = oracle.msda_forward(inp["value"], inp["shapes"], inp["starts"], inp["loc"][:, sel], inp["attn"][:, sel])
    np.testing.assert_allclose(out[:, sel].cpu().numpy(), ref, rtol=0, atol=1e-6)
    # border-crossing taps at the same shapes
    inb = synth.msda_inputs(B=1, shapes=shapes, Q=512, mode="border", seed=43)
    outb = ops.msda_forward(cu(inb["value"]), cu(inb["shapes"]), cu(inb["starts"]), cu(inb["loc"]), cu(inb["attn"]))
    refb = oracle.msda_forward(inb["value"], inb["shapes"], inb["starts"], inb["loc"], inb["attn"])
    np.testing.assert_allclose(outb.cpu().numpy(), refb, rtol=0, atol=1e-6)
