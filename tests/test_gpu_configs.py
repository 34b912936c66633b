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


# ------------------------------------------------------------------ C1: the reference's own CPU case (one frame, K = 20)
def test_c1_single_frame_20_masks_whole_path():
    """BASELINE.json configs[0]: ONE 640x480 frame with 20 initial masks — the shape the reference runs on a CPU.  Encode,
    boundaries (with and without the encode pass' occupancy), the six sampling layers' operands and post-processing
    against the oracle at batch 1."""
    K, H, W = 20, 480, 640
    masks = synth.ellipse_masks(K, H, W, seed=100, axes=(20, 60), margin=60)
    dm = cu(masks)[None]
    enc, occ = ops.encode_initial_masks(dm, with_occupancy=True)
    np.testing.assert_array_equal(enc[0].cpu().numpy(), oracle.encode_initial_masks(masks))
    d = ops.boundary_dilation(H, W, 0.01)
    u, p = oracle.masks_to_boundary(masks, 0.01, return_per_instance=True)
    for oc in (None, occ):
        per, uni, fg = ops.mask_boundary(dm, d, occupancy=oc)
        np.testing.assert_array_equal(per[0].cpu().numpy(), p)
        np.testing.assert_array_equal(uni[0].cpu().numpy(), u)
    inp = synth.msda_inputs(B=1, mode="module", seed=100, operands=True)
    got = ops.msda_forward_fused(cu(inp["value"]), cu(inp["shapes"]), cu(inp["starts"]), cu(inp["oa"]), cu(inp["ref"]), 3, 4,
                                 levels=((60, 80), (30, 40), (15, 20)))
    sel = np.r_[0:200, 4700:4900, 6200:6300]
    want = oracle.msda_forward(inp["value"], inp["shapes"], inp["starts"], inp["loc"][:, sel], inp["attn"][:, sel])
    np.testing.assert_allclose(got[0, sel].cpu().numpy(), want[0], rtol=0, atol=1e-6)
    assert _post_vs_oracle(masks, 100, max_instances=200) >= 5


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
    # (c_x - x) / W with W % 16 != 0: zero-padded row pitch, normalised by the frame's own width inside the kernel (valid_w)
    np.testing.assert_array_equal(enc[2], ref[2])


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


# ------------------------------------------------------------------ A7: sem_seg_postprocess on kernels
@pytest.mark.parametrize("case", ["crop_only", "armbench", "down", "near_identity"])
def test_upsample_crop_resize_kernel(case):
    from conftest import golden
    g = golden("resize.npz")
    low = g["low"]
    crop, size = tuple(int(x) for x in g[f"{case}_crop"]), tuple(int(x) for x in g[f"{case}_size"])
    x = cu(np.stack([low, low[::-1].copy()]))
    got = ops.upsample_crop_resize(x, 4, crop, size, mul=4.0)
    assert tuple(got.shape) == (2, 3) + size
    np.testing.assert_array_equal(got[0].cpu().numpy(), oracle.upsample_crop_resize(low, 4, crop, size, mul=4.0))   # bit-exact
    np.testing.assert_allclose(got[0].cpu().numpy(), g[f"{case}_out"], rtol=2e-6, atol=2e-6)                       # torch's chain
    np.testing.assert_array_equal(got[1, 0].cpu().numpy(), got[0, 2].cpu().numpy())


def test_resize_nearest_kernel_matches_cv2_golden():
    from conftest import golden
    g = golden("resize.npz")
    shape = tuple(int(x) for x in g["nn_shape"])
    n = int(np.prod(shape))
    m = np.unpackbits(g["nn_masks"])[:n].reshape(shape).astype(np.uint8) * 255
    for i in range(4):
        size = tuple(int(x) for x in g[f"nn{i}_size"])
        want = np.unpackbits(g[f"nn{i}_out"])[: shape[0] * size[0] * size[1]].reshape((shape[0],) + size).astype(np.uint8) * 255
        np.testing.assert_array_equal(ops.resize_nearest(cu(m), size).cpu().numpy(), want)


@pytest.mark.parametrize("crop", [(186, 333), (750 // 4 * 4 - 2, 331), (188, 336)])
def test_post_process_cropped_frames_match_oracle(crop):
    """Frames that are the top-left corner of the up-sampled maps (sem_seg_postprocess' crop; armbench 188 x 334 ->
    750 x 1333): centre NMS, grouping, panoptic merge and instance extraction on the cropped frame, width not a
    multiple of 8, against the oracle run on the cropped oracle-up-sampled maps."""
    K, Hf, Wf = 6, 188, 336
    crop = (min(crop[0], Hf), min(crop[1], Wf))
    masks = synth.ellipse_masks(K, Hf, Wf, seed=61, axes=(12, 40), margin=20)
    masks[0, :, 300:] = 255                                  # an instance cut by the crop's right edge
    ho = synth.head_outputs(masks, stride=4, seed=61)
    low = np.concatenate([ho["foreground"], ho["center"], ho["offset"], ho["eee_boundary"]], 0)
    r = ops.post_process(cu(low[None, 0:1]), cu(low[None, 1:2]), cu(low[None, 2:4]), cu(low[None, 4:8]), scale=4, crop=crop)
    Hc, Wc = crop
    assert r["valid_w"] == Wc and r["panoptic"].shape == (1, Hc, (Wc + 7) // 8 * 8) and int(r["status"][0]) == 0
    full = oracle.upsample_bilinear(low, 4)
    full[2:4] *= 4.0
    full = np.ascontiguousarray(full[:, :Hc, :Wc])
    ref = oracle.post_process(full[0], full[1], full[2:4], full[4:8])
    n = int(r["inst_offsets"][1])
    assert n == len(ref["labels"]) and n > 0
    np.testing.assert_array_equal(r["centers_yx"][0, : int(r["n_centers"][0])].cpu().numpy(), ref["centres"])
    np.testing.assert_array_equal(r["panoptic"][0, :, :Wc].cpu().numpy(), ref["panoptic"])
    assert (r["panoptic"][0, :, Wc:] == -1).all()            # the pitch padding is void
    np.testing.assert_array_equal(r["masks"][:n, :, :Wc].cpu().numpy(), ref["masks"])
    assert not r["masks"][:n, :, Wc:].any()
    np.testing.assert_array_equal(r["labels"][:n].cpu().numpy(), ref["labels"])
    np.testing.assert_array_equal(r["boxes"][:n].cpu().numpy(), ref["boxes"])
    np.testing.assert_allclose(r["scores"][:n].cpu().numpy(), ref["scores"], rtol=1e-4, atol=1e-6)
    np.testing.assert_array_equal(r["eee_onehot"][0, :, :, :Wc].cpu().numpy(), ref["onehot"])
    np.testing.assert_array_equal(r["sem_seg"][0, 0, :, :Wc].cpu().numpy(), full[0])
    np.testing.assert_array_equal(r["eee_full"][0, :, :, :Wc].cpu().numpy(), full[4:8])


def test_maskrefiner_forward_cropped_and_resized_outputs_on_kernels():
    """MaskRefiner.forward for frames whose size is not the up-sampled map's (armbench-shaped 94 x 166 -> head 24 x 42 -> 96 x 168)
    and for a requested output size that differs from the image size: no torch interpolate on the way, outputs equal
    the kernels called by hand on the same logits."""
    import os
    from quber_b200.config import get_cfg
    from quber_b200.meta_arch import build_model
    cfg = get_cfg()
    cfg.merge_from_file(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "configs",
                                     "mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8.yaml"))
    torch.manual_seed(0)
    model = build_model(cfg).eval()
    Hi, Wi = 94, 166
    g = torch.Generator().manual_seed(3)
    inp = [{"image": torch.randint(0, 255, (6, Hi, Wi), dtype=torch.uint8, generator=g),
            "initial_pred_offset": torch.randn(3, Hi, Wi, generator=g) * 0.1, "height": Hi, "width": Wi} for _ in range(2)]
    inp.append(dict(inp[0], height=141, width=250))                   # same frame, a real resize of the outputs
    import torch.nn.functional as F
    called = []
    orig = F.interpolate
    try:
        F.interpolate = lambda *a, **k: (called.append(1), orig(*a, **k))[1]
        x, sizes = model._pack(inp)
        logits = model.network_logits(x)
        n_net = len(called)
        res = model(inp)
        assert len(called) == 2 * n_net                                # only the network's own (decoder) up-sampling calls
    finally:
        F.interpolate = orig
    assert tuple(res[0]["sem_seg"].shape) == (1, Hi, Wi) and tuple(res[0]["eee_boundary"].shape) == (4, Hi, Wi)
    assert tuple(res[2]["sem_seg"].shape) == (1, 141, 250) and tuple(res[2]["panoptic_seg"][0].shape) == (141, 250)
    h, w = logits["foreground"].shape[-2:]
    assert h * 4 >= Hi and w * 4 >= Wi and (h * 4, w * 4) != (Hi, Wi)
    want0 = ops.upsample_crop_resize(logits["foreground"], 4, (Hi, Wi), (Hi, Wi))
    assert torch.equal(res[0]["sem_seg"], want0[0]) and torch.equal(res[1]["sem_seg"], want0[1])
    want2 = ops.upsample_crop_resize(logits["foreground"][2:3], 4, (Hi, Wi), (141, 250))
    assert torch.equal(res[2]["sem_seg"], want2[0])
    ref = F.interpolate(F.interpolate(logits["eee_boundary"][2:3], scale_factor=4, mode="bilinear", align_corners=False)[:, :, :Hi, :Wi],
                        size=(141, 250), mode="bilinear", align_corners=False)[0]
    torch.testing.assert_close(res[2]["eee_boundary"], ref, rtol=0, atol=1e-5)
