"""Parity of the CUDA path (called through the C ABI) against the oracle, the reference-generated
golden vectors, and size-independent properties at BASELINE.json's full sizes.  B200 only."""
import hashlib

import numpy as np
import pytest
import torch

import oracle
from quber_b200 import ops, synth
from conftest import golden, unpack_masks

pytestmark = pytest.mark.gpu
DEV = "cuda"


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


# ------------------------------------------------------------------ K0 encode
@pytest.mark.parametrize("case", ["small", "edge", "c1"])
def test_encode_golden_bit_exact(case):
    g = golden("encode.npz")
    m = unpack_masks(g[f"{case}_masks_bits"], g[f"{case}_shape"], g[f"{case}_mask_values"])
    out = ops.encode_initial_masks(cu(m)[None])[0].cpu().numpy()
    np.testing.assert_array_equal(out, oracle.encode_initial_masks(m))
    np.testing.assert_array_equal(sha(out), g[f"{case}_sha"])           # the reference's own bytes


def test_encode_batch_overlap_and_empty():
    ms = []
    for s in range(3):
        m = synth.ellipse_masks(9, 240, 320, seed=20 + s, axes=(20, 70), margin=30, empty_tail=1)   # heavy overlaps
        ms.append(m[np.random.default_rng(s).permutation(len(m))])                                  # empty mask mid-list
    ms = np.stack(ms)
    ms[1, 4] = 0
    ms[2][:] = 0                                                                                    # image without any mask
    out = ops.encode_initial_masks(cu(ms)).cpu().numpy()
    for b in range(3):
        np.testing.assert_array_equal(out[b], oracle.encode_initial_masks(ms[b]))
    assert not out[2].any()


# ------------------------------------------------------------------ K2 boundary
@pytest.mark.parametrize("case", ["small_r01", "small_r02", "c1_r01", "c1_r02"])
def test_boundary_golden_bit_exact(case):
    g = golden("boundary.npz")
    m = unpack_masks(g[f"{case}_masks"], g[f"{case}_shape"], g[f"{case}_mask_values"])
    H, W = m.shape[1:]
    d = ops.boundary_dilation(H, W, float(g[f"{case}_ratio"]))
    per, uni, fg = ops.mask_boundary(cu(m)[None], d)
    per, uni, fg = per[0].cpu().numpy(), uni[0].cpu().numpy(), fg[0].cpu().numpy()
    np.testing.assert_array_equal(sha(per), g[f"{case}_per_sha"])       # cv2's own bytes
    np.testing.assert_array_equal(np.packbits(uni), g[f"{case}_union"])
    np.testing.assert_array_equal(np.packbits(fg), g[f"{case}_fg"])


@pytest.mark.parametrize("d", [1, 2, 3, 5, 8, 15, 16, 31])
def test_boundary_random_bytes_any_radius(d):
    """Dense random uint8 planes (not binary): every tile takes the full min-filter path."""
    rng = np.random.default_rng(d)
    m = rng.integers(0, 256, (2, 3, 136, 272), dtype=np.uint8)       # H not a multiple of the 64-row tile, W of 128
    m[0, 1, 20:90, 40:200] = 200                                     # a uniform patch inside noise
    m[1, 2] = 77                                                     # a uniform non-zero plane (border still erodes)
    per, uni, fg = ops.mask_boundary(cu(m), d)
    for b in range(2):
        p = np.stack([_oracle_boundary_d(x, d) for x in m[b]])
        u = (p.sum(0, dtype=np.uint8) > 0).astype(np.uint8)            # uint8 wrap-around sum, util.py:96-98
        np.testing.assert_array_equal(per[b].cpu().numpy(), p)
        np.testing.assert_array_equal(uni[b].cpu().numpy(), u)
        np.testing.assert_array_equal(fg[b].cpu().numpy(), oracle.masks_to_fg_mask(m[b]))


@pytest.mark.parametrize("d", [1, 7, 8, 16, 17, 33, 64])
def test_boundary_binary_route_values_and_mixed_planes(d):
    """Binary planes with different non-zero values (bit-packed route), overlapping so that the uint8
    wrap-around of the union sums matters (128 + 128 == 0), mixed with a noise plane (byte route) and an
    empty plane, on a frame whose size is not a multiple of the 128x64 tile."""
    rng = np.random.default_rng(100 + d)
    base = synth.ellipse_masks(6, 200, 272, seed=d, axes=(20, 80), margin=10, value=1)
    vals = [255, 128, 128, 1, 7, 254]
    m = np.zeros((2, 8, 200, 272), np.uint8)
    for k in range(6):
        m[0, k] = base[k] * vals[k]
        m[1, k] = base[5 - k] * vals[k]
    m[0, 2] = m[0, 1]                                                # identical 128-valued planes: sums wrap to 0
    m[0, 6] = rng.integers(0, 256, (200, 272), dtype=np.uint8)       # noise plane between binary ones
    m[1, 6, :, :] = 255                                              # full plane: only the image border erodes
    per, uni, fg = ops.mask_boundary(cu(m), d)
    for b in range(2):
        p = np.stack([_oracle_boundary_d(x, d) for x in m[b]])
        u = (p.sum(0, dtype=np.uint8) > 0).astype(np.uint8)
        np.testing.assert_array_equal(per[b].cpu().numpy(), p)
        np.testing.assert_array_equal(uni[b].cpu().numpy(), u)
        np.testing.assert_array_equal(fg[b].cpu().numpy(), oracle.masks_to_fg_mask(m[b]))
    from quber_b200 import _lib
    assert _lib.lib().quber_tma_timeouts() == 0


@pytest.mark.parametrize("d,H,W", [(1, 200, 272), (8, 480, 640), (17, 136, 400), (64, 200, 272)])
def test_boundary_with_encode_occupancy_is_identical(d, H, W):
    """The tile occupancy written by the encode pass (which reads every mask byte anyway) lets quber_mask_boundary skip
    the (tile, plane) pairs that have no mask pixel within d of the tile: same per-instance / union / foreground maps as
    without it and as the oracle; occupancy itself = any() over the 128 x 16 tiles.  Small masks far apart, a mask that
    touches only a tile corner, an empty plane, a noise plane (byte route) and a full plane."""
    rng = np.random.default_rng(900 + d)
    K = 9
    m = np.zeros((2, K, H, W), np.uint8)
    el = synth.ellipse_masks(5, H, W, seed=d, axes=(6, 30), margin=8, value=255)
    m[0, :5] = el
    m[1, 1:6] = el[::-1]
    m[0, 5, 63:65, 127:129] = 255                                    # four pixels around a tile corner
    m[0, 6] = (rng.random((H, W)) < 0.001) * np.uint8(255)           # sparse specks everywhere
    m[1, 6] = rng.integers(0, 256, (H, W), dtype=np.uint8)           # noise: the exact min-filter route
    m[1, 7] = 255                                                    # full plane; plane 8 (and 0 of frame 1) stay empty
    dm = cu(m)
    enc, occ = ops.encode_initial_masks(dm, 10, with_occupancy=True)
    assert torch.equal(enc, ops.encode_initial_masks(dm, 10))
    th, tw = ops.OCC_TILE
    want = np.zeros(tuple(occ.shape), np.uint8)
    for r in range(occ.shape[2]):
        for c in range(occ.shape[3]):
            want[:, :, r, c] = m[:, :, r * th:(r + 1) * th, c * tw:(c + 1) * tw].reshape(2, K, -1).any(-1)
    np.testing.assert_array_equal(occ.cpu().numpy(), want)
    per0, uni0, fg0 = ops.mask_boundary(dm, d)
    per1, uni1, fg1 = ops.mask_boundary(dm, d, occupancy=occ)
    assert torch.equal(per0, per1) and torch.equal(uni0, uni1) and torch.equal(fg0, fg1)
    _, uni2, fg2 = ops.mask_boundary(dm, d, per_instance=False, occupancy=occ)
    assert torch.equal(uni0, uni2) and torch.equal(fg0, fg2)
    for b in range(2):
        p = np.stack([_oracle_boundary_d(x, d) for x in m[b]])
        np.testing.assert_array_equal(per1[b].cpu().numpy(), p)
        np.testing.assert_array_equal(fg1[b].cpu().numpy(), oracle.masks_to_fg_mask(m[b]))
    assert int(ops.boundary_status(dm.device)[0]) == 0
    with pytest.raises(ValueError):
        ops.mask_boundary(dm, d, occupancy=occ[:, :-1].contiguous())


@pytest.mark.parametrize("d,H,W,value", [(1, 200, 288, 255), (8, 480, 640, 255), (16, 136, 416, 1), (17, 136, 416, 7), (40, 200, 288, 255),
                                         (64, 100, 160, 128)])
def test_boundary_from_bit_planes_equals_byte_kernel_and_oracle(d, H, W, value):
    """quber_mask_boundary_bits (the planes as one bit per pixel, what the host-buffer routes hold) against the byte kernel
    and the oracle: per-instance, union and foreground maps, with and without the encode pass' occupancy; overlapping
    planes so that the uint8 wrap-around of the union sums matters (value 128: two planes sum to 0), masks touching the
    image border and tile corners, an empty and a full plane; W a multiple of 32 but not of the 128-pixel tile."""
    rng = np.random.default_rng(950 + d)
    K = 8
    m = np.zeros((2, K, H, W), np.uint8)
    el = synth.ellipse_masks(5, H, W, seed=d, axes=(8, 60), margin=0, value=1)
    m[0, :5] = el
    m[1, 1:6] = el[::-1]
    m[0, 5] = m[0, 4]                                                # identical planes: value 128 wraps to 0 in the sums
    m[0, 6, 63:65, 127:129] = 1
    m[0, 6, :3, :] = 1                                               # along the top border
    m[1, 6] = rng.random((H, W)) < 0.3                               # speckle: every tile live, thin structures
    m[1, 7] = 1                                                      # full plane: only the image border erodes
    m *= np.uint8(value)
    dm = cu(m)
    words, v = ops.host_pack_masks(torch.from_numpy(m).pin_memory())
    assert v == value
    dw = words.cuda()
    enc_p, occ = ops.encode_initial_masks_packed(dw, value, None, 10, with_occupancy=True, shape=m.shape)
    enc_b, occ_b = ops.encode_initial_masks(dm, 10, with_occupancy=True)
    assert torch.equal(enc_p, enc_b) and torch.equal(occ, occ_b)
    per0, uni0, fg0 = ops.mask_boundary(dm, d)
    for oc in (None, occ):
        per1, uni1, fg1 = ops.mask_boundary_bits(dw, value, m.shape, d, occupancy=oc)
        assert torch.equal(per0, per1) and torch.equal(uni0, uni1) and torch.equal(fg0, fg1)
    _, uni2, fg2 = ops.mask_boundary_bits(dw, value, m.shape, d, per_instance=False, occupancy=occ)
    assert torch.equal(uni0, uni2) and torch.equal(fg0, fg2)
    for b in range(2):
        p = np.stack([_oracle_boundary_d(x, d) for x in m[b]])
        np.testing.assert_array_equal(per1[b].cpu().numpy(), p)
    with pytest.raises(ValueError):
        ops.mask_boundary_bits(dw[:-1], value, m.shape, d)


def _oracle_boundary_d(mask, d):
    import ctypes
    out = np.empty_like(mask)
    H, W = mask.shape
    oracle.lib().oracle_mask_to_boundary(mask.ctypes.data_as(ctypes.c_void_p), H, W, d, out.ctypes.data_as(ctypes.c_void_p))
    return out


def test_boundary_union_only_and_quadruple():
    gt = synth.ellipse_masks(6, 240, 320, seed=3, axes=(15, 50), margin=20)
    pr = synth.perturb_masks(gt, seed=4)
    d = ops.boundary_dilation(240, 320, 0.01)
    _, gb, gf = ops.mask_boundary(cu(gt)[None], d, per_instance=False)
    _, pb, pf = ops.mask_boundary(cu(pr)[None], d, per_instance=False)
    np.testing.assert_array_equal(gb[0].cpu().numpy(), oracle.masks_to_boundary(gt, 0.01))
    q = ops.quadruple_maps(gb, pb)[0].cpu().numpy()
    np.testing.assert_array_equal(q, oracle.quadruple_maps(gb[0].cpu().numpy(), pb[0].cpu().numpy()))
    assert (q.sum(0) == 1).all()                                      # the four maps partition the image
    qm = ops.quadruple_maps(gf, pf)[0].cpu().numpy()
    np.testing.assert_array_equal(qm, oracle.quadruple_maps(oracle.masks_to_fg_mask(gt), oracle.masks_to_fg_mask(pr)))


# ------------------------------------------------------------------ K3/K4 post
def _low(g, case):
    low = g[f"{case}_low"]
    return low[None, 0:1], low[None, 1:2], low[None, 2:4], low[None, 4:8]


def _check_post(r, ref, H, W, b=0, exact_scores=False):
    o0, o1 = int(r["inst_offsets"][b]), int(r["inst_offsets"][b + 1])
    assert int(r["status"][0]) == 0
    np.testing.assert_array_equal(r["centers_yx"][b, : int(r["n_centers"][b])].cpu().numpy(), ref["centres"])
    np.testing.assert_array_equal(r["panoptic"][b].cpu().numpy(), ref["panoptic"])
    np.testing.assert_array_equal(r["labels"][o0:o1].cpu().numpy(), ref["labels"])
    np.testing.assert_array_equal(r["masks"][o0:o1].cpu().numpy(), ref["masks"])
    np.testing.assert_array_equal(r["boxes"][o0:o1].cpu().numpy(), ref["boxes"])
    np.testing.assert_allclose(r["scores"][o0:o1].cpu().numpy(), ref["scores"], rtol=1e-4, atol=1e-6)
    if ref.get("onehot") is not None:
        np.testing.assert_array_equal(r["eee_onehot"][b].cpu().numpy(), ref["onehot"])
    lm = r["label_map"][b].cpu().numpy()
    for j in range(o1 - o0):
        np.testing.assert_array_equal(lm == j + 1, ref["masks"][j].astype(bool))


@pytest.mark.parametrize("case", ["small", "topk", "plateau", "nocentre", "c1"])
def test_post_fused_upsample_matches_oracle_and_reference(case):
    """Mode (ii): the kernels up-sample the 1/4-resolution head outputs themselves."""
    g = golden("post.npz")
    fg, ce, of, ee = _low(g, case)
    top_k = int(g[f"{case}_topk"])
    r = ops.post_process(cu(fg), cu(ce), cu(of), cu(ee), scale=4, top_k=top_k)
    full = oracle.upsample_bilinear(g[f"{case}_low"], 4)
    # float outputs: bit-identical to the oracle's up-sampler, <= 4e-6 from torch's (tests/test_oracle_golden.py)
    np.testing.assert_array_equal(r["sem_seg"][0].cpu().numpy(), full[0:1])
    np.testing.assert_array_equal(r["eee_full"][0].cpu().numpy(), full[4:8])
    full[2:4] *= 4.0
    ref = oracle.post_process(full[0], full[1], full[2:4], full[4:8], top_k=top_k)
    H, W = full.shape[1:]
    _check_post(r, ref, H, W)
    # and against the reference's own get_panoptic_segmentation output
    np.testing.assert_array_equal(r["panoptic"][0].cpu().numpy().astype(np.int16), g[f"{case}_panoptic"])
    np.testing.assert_array_equal(r["centers_yx"][0, : int(r["n_centers"][0])].cpu().numpy(), g[f"{case}_centres"].reshape(-1, 2))
    np.testing.assert_array_equal(r["labels"][: len(ref["labels"])].cpu().numpy(), g[f"{case}_labels"])
    np.testing.assert_array_equal(r["boxes"][: len(ref["labels"])].cpu().numpy(), g[f"{case}_boxes"])
    np.testing.assert_allclose(r["scores"][: len(ref["labels"])].cpu().numpy(), g[f"{case}_scores"], rtol=1e-4, atol=1e-6)
    np.testing.assert_array_equal(np.packbits(r["eee_onehot"][0].cpu().numpy().astype(bool)), g[f"{case}_onehot"])


@pytest.mark.parametrize("case", ["small", "topk", "c1"])
def test_post_fullres_inputs_bit_exact(case):
    """Mode (i): the kernels consume full-resolution float maps (scale=1), outputs strictly bit-exact."""
    g = golden("post.npz")
    full = oracle.upsample_bilinear(g[f"{case}_low"], 4)
    full[2:4] *= 4.0
    top_k = int(g[f"{case}_topk"])
    r = ops.post_process(cu(full[None, 0:1]), cu(full[None, 1:2]), cu(full[None, 2:4]), cu(full[None, 4:8]), scale=1,
                         top_k=top_k)
    ref = oracle.post_process(full[0], full[1], full[2:4], full[4:8], top_k=top_k)
    _check_post(r, ref, *full.shape[1:])
    np.testing.assert_array_equal(r["panoptic"][0].cpu().numpy().astype(np.int16), g[f"{case}_panoptic"])


def test_post_batch_mixed_and_capacity():
    """Images with many / zero / below-threshold centres in one launch; compact mask rows; capacity overflow flag."""
    lows, refs = [], []
    for s, K in [(31, 7), (32, 0), (33, 14), (34, 3)]:
        m = synth.ellipse_masks(max(K, 1), 240, 320, seed=s, axes=(12, 40), margin=30)
        if K == 0:
            m[:] = 0
        ho = synth.head_outputs(m, stride=4, seed=s)
        if s == 34:
            ho["center"] *= 0.25                       # no centre passes -> stuff branch (label 1000)
        low = np.concatenate([ho["foreground"], ho["center"], ho["offset"], ho["eee_boundary"]], 0)
        lows.append(low)
        full = oracle.upsample_bilinear(low, 4)
        full[2:4] *= 4.0
        refs.append(oracle.post_process(full[0], full[1], full[2:4], full[4:8]))
    lows = np.stack(lows)
    r = ops.post_process(cu(lows[:, 0:1]), cu(lows[:, 1:2]), cu(lows[:, 2:4]), cu(lows[:, 4:8]), scale=4, max_instances=64)
    for b, ref in enumerate(refs):
        _check_post(r, ref, 240, 320, b=b)
    total = sum(len(x["labels"]) for x in refs)
    assert int(r["inst_offsets"][-1]) == total
    r2 = ops.post_process(cu(lows[:, 0:1]), cu(lows[:, 1:2]), cu(lows[:, 2:4]), cu(lows[:, 4:8]), scale=4,
                          max_instances=max(total - 2, 1))
    assert int(r2["status"][0]) & 1
    np.testing.assert_array_equal(r2["panoptic"].cpu().numpy(), r["panoptic"].cpu().numpy())


def test_post_many_candidates_radix_paths():
    """> top_k and > 4096 NMS survivors (radix select in shared memory and from the bit mask)."""
    rng = np.random.default_rng(5)
    H, W = 240, 320
    for n_peaks, top_k in [(600, 50), (9000, 200)]:
        ce = np.zeros((H, W), np.float32)
        ys, xs = rng.integers(0, H, n_peaks), rng.integers(0, W, n_peaks)
        ce[ys, xs] = rng.uniform(0.31, 1.0, n_peaks).astype(np.float32)
        ce[ys[:40], xs[:40]] = 0.75                                # ties around the cut
        fg = np.full((H, W), 3.0, np.float32)
        of = rng.normal(0, 2, (2, H, W)).astype(np.float32)
        r = ops.post_process(cu(fg[None, None]), cu(ce[None, None]), cu(of[None]), None, scale=1, top_k=top_k, nms_kernel=3,
                             want=("panoptic", "ins_seg"))
        ref_c = oracle.find_instance_center(ce, 0.3, 3, top_k)
        np.testing.assert_array_equal(r["centers_yx"][0, : int(r["n_centers"][0])].cpu().numpy(), ref_c)
        if len(ref_c):
            np.testing.assert_array_equal(r["ins_seg"][0].cpu().numpy(), oracle.group_pixels(ref_c, of))


def test_center_nms_sparse_peaks_at_tile_borders():
    """Tiles whose low-resolution footprint stays under the threshold are skipped (csrc/post.cu: center_nms_kernel);
    single low-resolution peaks next to every kind of tile border, values straddling the threshold by a few ulp, must give
    the oracle's centres exactly."""
    h, w, scale, thr = 64, 96, 4, 0.3
    rng = np.random.default_rng(12)
    for trial in range(6):
        low = np.full((h, w), 0.29, np.float32) if trial % 2 else rng.uniform(-1.0, 0.25, (h, w)).astype(np.float32)
        # full-resolution NMS tiles are 64 x 32 -> low-resolution cells 16 x 8; put peaks on both sides of the borders
        for (y, x) in [(7, 15), (8, 16), (15, 31), (16, 32), (0, 0), (h - 1, w - 1), (31, 47), (40, 17), (23, 80)]:
            low[y, x] = np.float32(thr) + np.float32(rng.choice([-3e-8, 3e-8, 1e-3, 0.2]))
        full = oracle.upsample_bilinear(low[None], scale)[0]
        ref_c = oracle.find_instance_center(full, thr, 7, 200)
        fg = np.full((h, w), 3.0, np.float32)
        of = np.zeros((2, h, w), np.float32)
        r = ops.post_process(cu(fg[None, None]), cu(low[None, None]), cu(of[None]), None, scale=scale, top_k=200, nms_kernel=7,
                             want=("panoptic",))
        n = int(r["n_centers"][0])
        np.testing.assert_array_equal(r["centers_yx"][0, :n].cpu().numpy(), ref_c)
    # an all-quiet map: no centre at all
    low = np.full((h, w), 0.1, np.float32)
    r = ops.post_process(cu(np.full((1, 1, h, w), 3.0, np.float32)), cu(low[None, None]), cu(np.zeros((1, 2, h, w), np.float32)), None,
                         scale=scale, want=("panoptic",))
    assert int(r["n_centers"][0]) == 0


def test_reference_named_post_processing_functions():
    from quber_b200 import post_processing as pp
    g = golden("post.npz")
    full = oracle.upsample_bilinear(g["small_low"], 4)
    full[2:4] *= 4.0
    c, o = cu(full[1:2]), cu(full[2:4])
    ctr = pp.find_instance_center(c, threshold=0.3, nms_kernel=7, top_k=200)
    assert ctr.dtype == torch.int64
    np.testing.assert_array_equal(ctr.cpu().numpy(), g["small_centres"])
    np.testing.assert_array_equal(pp.find_instance_center(c, 0.3, 7, None).cpu().numpy(), g["small_centres"])
    grp = pp.group_pixels(ctr, o)
    assert grp.shape == (1,) + full.shape[1:] and grp.dtype == torch.int64
    np.testing.assert_array_equal(grp[0].cpu().numpy(), g["small_grouped"])
    fgseg = torch.sigmoid(cu(full[0:1])).round()
    pan, cc = pp.get_panoptic_segmentation(fgseg, c, o, thing_ids=dict({1: 0}).values(), label_divisor=1000, stuff_area=2048,
                                           void_label=-1, threshold=0.3, nms_kernel=7, top_k=200)
    assert pan.shape == (1,) + full.shape[1:] and pan.dtype == torch.float32 and cc.shape == (1, len(ctr), 2)
    np.testing.assert_array_equal(pan[0].cpu().numpy().astype(np.int16), g["small_panoptic"])
    ins, _ = pp.get_instance_segmentation(c, o, fgseg, threshold=0.3, nms_kernel=7, top_k=200)
    np.testing.assert_array_equal(ins[0].cpu().numpy(), g["small_grouped"] * fgseg[0].cpu().numpy())
    with pytest.raises(ValueError):
        pp.get_panoptic_segmentation(fgseg, c[0], o, dict({1: 0}).values(), 1000, 2048, -1)


def test_group_pixels_tie_band_first_centre_wins():
    """Pixels landing within a few ulp of the bisector of two centres: squared distances differ while the rounded
    roots are equal for ~10 % of them, a tie torch.argmin gives to the EARLIER centre (post_processing.py:71-75).
    The kernel compares squared distances and takes the roots only inside that band (csrc/post.cu: group_kernel)."""
    from quber_b200 import post_processing as pp
    H, W = 64, 128
    rng = np.random.default_rng(3)
    yy, xx = np.meshgrid(np.arange(H, dtype=np.float32), np.arange(W, dtype=np.float32), indexing="ij")
    steps = rng.integers(-6, 7, size=(H, W)).astype(np.float32)
    tly = np.float32(30.0) + steps * np.float32(1.9073486e-06)
    tlx = rng.uniform(0, 100, size=(H, W)).astype(np.float32)
    off = np.stack([tly - yy, tlx - xx]).astype(np.float32)
    ly, lx = (yy + off[0]).astype(np.float32), (xx + off[1]).astype(np.float32)

    def d2(c):
        dy, dx = np.float32(c[0]) - ly, np.float32(c[1]) - lx
        return (dx.astype(np.float64) ** 2 + (dy * dy).astype(np.float64)).astype(np.float32)

    a, b = d2((20, 30)), d2((40, 30))
    assert int(((a != b) & (np.sqrt(a) == np.sqrt(b))).sum()) > 300          # the band is exercised
    for ctr in ([[20, 30], [40, 30]], [[40, 30], [20, 30]], [[20, 30], [40, 30], [20, 30], [40, 30], [3, 3]]):
        ctr = np.asarray(ctr, np.int32)
        ref = oracle.group_pixels(ctr, off)
        got = pp.group_pixels(cu(ctr.astype(np.int64)), cu(off))[0].cpu().numpy()
        np.testing.assert_array_equal(got, ref)
        assert not np.isin(ref, (3, 4)).any()                                 # duplicates never beat their first copy


def test_reference_named_boundary_functions_numpy():
    from quber_b200 import boundary as bd
    g = golden("boundary.npz")
    m = unpack_masks(g["small_r01_masks"], g["small_r01_shape"], g["small_r01_mask_values"])
    np.testing.assert_array_equal(np.packbits(bd.masks_to_boundary(m, 0.01)), g["small_r01_union"])
    np.testing.assert_array_equal(np.packbits(bd.masks_to_fg_mask(m)), g["small_r01_fg"])
    np.testing.assert_array_equal(bd.mask_to_boundary(m[0], 0.02), oracle.mask_to_boundary(m[0], 0.02))
    odd = m[:, :, :150].copy()                                     # W % 16 != 0 -> padded inside the wrapper
    np.testing.assert_array_equal(bd.masks_to_boundary(odd, 0.01), oracle.masks_to_boundary(odd, 0.01))
    q = bd.quadruple_boundary_maps(m, synth.perturb_masks(m, seed=9), 0.01)
    assert set(q) == {f"{a}_{b}" for a in ("tp", "tn", "fp", "fn") for b in ("mask", "boundary")}


# ------------------------------------------------------------------ K1 MSDA
MSDA_CASES = {
    "opstest": dict(B=1, shapes=((6, 4), (3, 2)), M=2, D=2, P=2, Q=2, seed=3, mode="rand"),
    "medium": dict(B=2, shapes=((8, 10), (4, 5), (2, 3)), M=8, D=32, P=4, Q=None, seed=1, mode="rand"),
    "border": dict(B=1, shapes=((12, 16), (6, 8), (3, 4)), M=8, D=32, P=4, Q=100, seed=2, mode="border"),
    "d8": dict(B=1, shapes=((10, 14), (5, 7)), M=8, D=8, P=4, Q=64, seed=5, mode="border"),
}


@pytest.mark.parametrize("case", list(MSDA_CASES))
def test_msda_golden(case):
    g = golden("msda.npz")
    inp = synth.msda_inputs(dtype=np.float64, **MSDA_CASES[case])
    v, loc, at = (inp[k].astype(np.float32) for k in ("value", "loc", "attn"))
    out = ops.msda_forward(cu(v), cu(inp["shapes"]), cu(inp["starts"]), cu(loc), cu(at)).cpu().numpy()
    np.testing.assert_allclose(out, g[f"{case}_out_f32"], rtol=0, atol=1e-3)      # north-star tolerance, reference's fp32 output
    np.testing.assert_allclose(out, g[f"{case}_out_f64"], rtol=0, atol=1e-6)      # what it actually achieves
    np.testing.assert_allclose(out, oracle.msda_forward(v, inp["shapes"], inp["starts"], loc, at), rtol=0, atol=1e-7)


@pytest.mark.parametrize("D", [8, 16, 32, 64, 12])
def test_msda_oracle_shapes_and_bf16(D):
    inp = synth.msda_inputs(B=2, shapes=((15, 20), (8, 10), (4, 5)), M=4, D=D, P=4, Q=333, seed=D, mode="border")
    ref = oracle.msda_forward(inp["value"], inp["shapes"], inp["starts"], inp["loc"], inp["attn"])
    args = (cu(inp["shapes"]), cu(inp["starts"]), cu(inp["loc"]), cu(inp["attn"]))
    out = ops.msda_forward(cu(inp["value"]), *args).cpu().numpy()
    np.testing.assert_allclose(out, ref, rtol=0, atol=1e-7)
    if D % 8 == 0:
        ob = ops.msda_forward(cu(inp["value"]).bfloat16(), *args)
        assert ob.dtype == torch.bfloat16
        # bf16 value/out, fp32 weights and accumulation: values ~1e-2, bf16 eps 2^-8 -> stated tolerance 2e-2 relative
        np.testing.assert_allclose(ob.float().cpu().numpy(), ref, rtol=2e-2, atol=1e-5)


@pytest.mark.parametrize("P", [1, 2, 3, 8])
def test_msda_other_point_counts_take_the_generic_vector_kernel(P):
    """P != 4 goes to the two-record kernel (csrc/msda.cu: msda_forward_kernel), P == 4 to the one-record kernel."""
    inp = synth.msda_inputs(B=2, shapes=((15, 20), (8, 10), (4, 5)), M=4, D=32, P=P, Q=200, seed=40 + P, mode="border")
    ref = oracle.msda_forward(inp["value"], inp["shapes"], inp["starts"], inp["loc"], inp["attn"])
    args = (cu(inp["shapes"]), cu(inp["starts"]), cu(inp["loc"]), cu(inp["attn"]))
    np.testing.assert_allclose(ops.msda_forward(cu(inp["value"]), *args).cpu().numpy(), ref, rtol=0, atol=1e-7)
    ob = ops.msda_forward(cu(inp["value"]).bfloat16(), *args)
    np.testing.assert_allclose(ob.float().cpu().numpy(), ref, rtol=2e-2, atol=1e-5)


@pytest.mark.parametrize("shapes", [((1, 5), (4, 1), (1, 1)), ((2, 2), (1, 7)), ((9, 1),)])
def test_msda_degenerate_levels_single_row_or_column(shapes):
    """Levels one pixel wide or high: both x (or y) corners of a tap fall on the same pixel, one of them outside the map
    (ms_deform_im2col_cuda.cuh:58-81 reads v = 0 there) — the one-record kernel folds the surviving weight."""
    for D in (16, 32, 64):
        inp = synth.msda_inputs(B=2, shapes=shapes, M=2, D=D, P=4, Q=257, seed=7 + D, mode="border")
        ref = oracle.msda_forward(inp["value"], inp["shapes"], inp["starts"], inp["loc"], inp["attn"])
        args = (cu(inp["shapes"]), cu(inp["starts"]), cu(inp["loc"]), cu(inp["attn"]))
        np.testing.assert_allclose(ops.msda_forward(cu(inp["value"]), *args).cpu().numpy(), ref, rtol=0, atol=1e-7)
        ob = ops.msda_forward(cu(inp["value"]).bfloat16(), *args)
        np.testing.assert_allclose(ob.float().cpu().numpy(), ref, rtol=2e-2, atol=1e-5)


def test_msda_module_and_function_api():
    from quber_b200.msda import MSDeformAttn, MSDeformAttnFunction
    torch.manual_seed(0)
    mod = MSDeformAttn(d_model=64, n_levels=2, n_heads=4, n_points=4).to(DEV)
    shapes = torch.tensor([[10, 12], [5, 6]], dtype=torch.long, device=DEV)
    starts = torch.cat((shapes.new_zeros((1,)), shapes.prod(1).cumsum(0)[:-1]))
    S = int(shapes.prod(1).sum())
    src = torch.randn(2, S, 64, device=DEV)
    ref_pts = torch.rand(2, S, 2, 2, device=DEV)
    with torch.no_grad():
        y = mod(src, ref_pts, src, shapes, starts)
        # same module maths with the oracle in place of the op
        value = mod.value_proj(src).view(2, S, 4, 16)
        off = mod.sampling_offsets(src).view(2, S, 4, 2, 4, 2)
        aw = torch.softmax(mod.attention_weights(src).view(2, S, 4, 8), -1).view(2, S, 4, 2, 4)
        norm = torch.stack([shapes[..., 1], shapes[..., 0]], -1)
        loc = ref_pts[:, :, None, :, None, :] + off / norm[None, None, None, :, None, :]
        o = oracle.msda_forward(value.cpu().numpy(), shapes.cpu().numpy(), starts.cpu().numpy(), loc.cpu().numpy(), aw.cpu().numpy())
        y_ref = mod.output_proj(torch.from_numpy(o).to(DEV))
        np.testing.assert_allclose(y.cpu().numpy(), y_ref.cpu().numpy(), rtol=0, atol=1e-4)
        y2 = MSDeformAttnFunction.apply(value, shapes, starts, loc.contiguous(), aw, 128)
        np.testing.assert_allclose(y2.cpu().numpy(), o, rtol=0, atol=1e-5)
        y3 = torch.ops.quber.msda_forward(value.contiguous(), shapes, starts, loc.contiguous(), aw)
        assert torch.equal(y2, y3)
    assert set(mod.state_dict()) == {f"{n}.{p}" for n in ("sampling_offsets", "attention_weights", "value_proj", "output_proj")
                                     for p in ("weight", "bias")}


# ------------------------------------------------------------------ full BASELINE sizes: properties
def test_full_size_properties_c2():
    """configs[1]: B=32 frames 640x480, K=30 masks.  The oracle is too slow at this size for every frame,
    so: two frames against the oracle bit-exactly, the rest through size-independent invariants."""
    B, K, H, W = 32, 30, 480, 640
    masks = np.stack([synth.ellipse_masks(K, H, W, seed=100 + b) for b in range(B)])
    dm = cu(masks)
    enc = ops.encode_initial_masks(dm)
    d = ops.boundary_dilation(H, W, 0.01)
    per, uni, fg = ops.mask_boundary(dm, d)
    for b in (0, 31):
        np.testing.assert_array_equal(enc[b].cpu().numpy(), oracle.encode_initial_masks(masks[b]))
        u, p = oracle.masks_to_boundary(masks[b], 0.01, return_per_instance=True)
        np.testing.assert_array_equal(per[b].cpu().numpy(), p)
        np.testing.assert_array_equal(uni[b].cpu().numpy(), u)
    mb = dm != 0
    assert torch.equal(fg.bool(), mb.any(1))                                   # fg = OR of masks
    assert torch.equal(uni.bool(), (per != 0).any(1))                          # union = OR of per-instance boundaries
    assert not ((per != 0) & ~mb).any()                                        # boundary inside its mask
    assert torch.equal(per[mb], dm[mb] * (per[mb] != 0))                       # boundary pixels keep the mask value
    # erosion is anti-extensive and idempotent on its own complement: boundary of the eroded core touches no old boundary pixel's outside
    core = (dm - per)
    per2, _, _ = ops.mask_boundary(core, d, union=False, fg=False)
    assert not ((per2 != 0) & (core == 0)).any()
    # encode: offsets vanish exactly off the union, heat-map in [0,1] with 1.0 at every rounded centroid
    off_nz = (enc[:, 1] != 0) | (enc[:, 2] != 0)
    assert not (off_nz & ~fg.bool()).any()
    assert float(enc[:, 0].max()) == 1.0 and float(enc[:, 0].min()) >= 0.0
    # post at full size on synthetic head outputs
    lows = np.stack([np.concatenate(list(synth.head_outputs(masks[b], seed=b).values()), 0) for b in range(4)])
    lows = np.concatenate([lows] * 8)                                          # 32 frames
    r = ops.post_process(cu(lows[:, 0:1]), cu(lows[:, 1:2]), cu(lows[:, 2:4]), cu(lows[:, 4:8]), scale=4,
                         max_instances=B * 40)
    assert int(r["status"][0]) == 0
    offs = r["inst_offsets"].cpu().numpy()
    n = offs[-1]
    assert n > B * 10
    mk = r["masks"][:n]
    area = mk.flatten(1).sum(1)
    assert int(area.min()) >= 512                                              # post_processing.py:145
    pan = r["panoptic"]
    assert int((pan != -1).sum()) == int(area.sum())                           # instances are disjoint and cover the non-void map
    for b in (0, 5, 31):
        lm = r["label_map"][b]
        for j in range(offs[b + 1] - offs[b]):
            assert torch.equal(mk[offs[b] - 0 + j].bool(), lm == j + 1)
            assert torch.equal(pan[b] == float(r["labels"][offs[b] + j]), lm == j + 1)
    assert torch.equal(r["eee_onehot"].sum(1), torch.ones_like(r["eee_onehot"][:, 0]))   # exactly one class per pixel
    full = oracle.upsample_bilinear(lows[0], 4)
    full[2:4] *= 4.0
    _check_post(r, oracle.post_process(full[0], full[1], full[2:4], full[4:8]), H, W, b=0)
    # frames 0 and 4 carry the same inputs -> identical outputs (determinism, no cross-image leakage)
    assert torch.equal(pan[0], pan[4]) and torch.equal(r["scores"][offs[0]:offs[1]], r["scores"][offs[4]:offs[5]])


def test_msda_full_size_linearity_and_oracle_sample():
    """S=6300 (640x480 at strides 8/16/32), 8 heads x 32, 3 levels x 4 points, B=4."""
    inp = synth.msda_inputs(B=4, mode="module", seed=7)
    sh, st = cu(inp["shapes"]), cu(inp["starts"])
    v, loc, at = cu(inp["value"]), cu(inp["loc"]), cu(inp["attn"])
    out = ops.msda_forward(v, sh, st, loc, at)
    # linear in value and in the attention weights
    out2 = ops.msda_forward(v * 2.0, sh, st, loc, at * 0.5)
    np.testing.assert_allclose(out2.cpu().numpy(), out.cpu().numpy(), rtol=1e-5, atol=1e-8)
    # constant value field + weights summing to 1 + in-bounds taps -> the constant
    inner = ((loc > 0.1) & (loc < 0.9)).all(-1).all(-1).all(-1)               # [B,Q,M]
    outc = ops.msda_forward(torch.full_like(v, 0.5), sh, st, loc, at).view(4, -1, 8, 32)
    assert float((outc[inner] - 0.5).abs().max()) < 1e-5
    ref = oracle.msda_forward(inp["value"][:1], inp["shapes"], inp["starts"], inp["loc"][:1, :500], inp["attn"][:1, :500])
    np.testing.assert_allclose(out[:1, :500].cpu().numpy(), ref, rtol=0, atol=1e-7)


# ------------------------------------------------------------------ tcgen05 GEMM (Linear / 1x1 conv)
def _tf32(x):
    return (x.view(torch.int32) & ~0x1FFF).view(torch.float32)


@pytest.mark.parametrize("shape", [(128, 128, 64), (300, 256, 256), (1000, 192, 256), (1000, 96, 256), (515, 1024, 256),
                                   (515, 256, 1024), (77, 40, 168), (1000, 4, 32), (333, 100, 72), (6300, 256, 256),
                                   # the W-resident bf16 kernel's edges: K off the 64-element k-block, one k-block only, several
                                   # n-tiles with fewer m-tiles than CTAs per n-tile, many m-tiles per CTA
                                   (130, 256, 72), (1, 512, 64), (40000, 768, 200), (257, 256, 8)])
@pytest.mark.parametrize("dt", ["bf16", "tf32"])
def test_gemm_bias_act_matches_fp64_reference(shape, dt):
    """C = relu(A W^T + b) against a float64 torch reference on inputs that are exactly representable in the
    multiply precision (bf16, or fp32 with a 10-bit mantissa for TF32), so that only the fp32 accumulation
    order differs: tolerance 1e-4 relative to the row scale (measured <= 2e-5)."""
    from quber_b200 import _lib
    M, N, K = shape
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device=DEV, generator=g)
    w = torch.randn(N, K, device=DEV, generator=g) / K ** 0.5
    b = torch.randn(N, device=DEV, generator=g)
    a, w = (a.bfloat16(), w.bfloat16()) if dt == "bf16" else (_tf32(a), _tf32(w))
    ref = torch.relu(a.double() @ w.double().T + b.double())
    out = ops.gemm_bias_act(a, w, b, act="relu", out_dtype=torch.float32)
    assert out.shape == (M, N) and out.dtype == torch.float32
    torch.testing.assert_close(out.double(), ref, rtol=0, atol=1e-4)
    lin = ops.gemm_bias_act(a, w, None, act="none", out_dtype=torch.float32)
    torch.testing.assert_close(lin.double(), a.double() @ w.double().T, rtol=0, atol=1e-4)
    if N % 8 == 0:
        ob = ops.gemm_bias_act(a, w, b, act="relu", out_dtype=torch.bfloat16)
        torch.testing.assert_close(ob.double(), ref, rtol=2 ** -7, atol=1e-3)       # bf16 output rounding
    assert _lib.lib().quber_gemm_timeouts() == 0


def test_gemm_leading_dims_and_general_fp32_inputs():
    """[B,S,K] activations (flattened rows), and arbitrary fp32 inputs multiplied as TF32: error bounded by the
    10-bit mantissa (|err| <= 2^-10 * sum |a||w| per element)."""
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(3, 700, 256, device=DEV, generator=g)
    w = torch.randn(192, 256, device=DEV, generator=g) / 16
    b = torch.randn(192, device=DEV, generator=g)
    y = ops.gemm_bias_act(x, w, b)
    assert y.shape == (3, 700, 192)
    ref = x.double() @ w.double().T + b.double()
    bound = (x.abs().double() @ w.abs().double().T) * 2.0 ** -10
    assert ((y.double() - ref).abs() <= bound + 1e-5).all()
    with pytest.raises(ValueError):
        ops.gemm_bias_act(x, w.bfloat16(), b)


# ------------------------------------------------------------------ round 2: executed-reference goldens on the kernels
@pytest.mark.parametrize("case", ["small", "topk", "plateau", "nocentre", "c1"])
def test_post_mode_ii_reference_interpolate_chain(case):
    """SURVEY Appendix A mode (ii): the kernels up-sample the 1/4-resolution head outputs themselves and are compared
    with the reference's real F.interpolate -> sigmoid().round() -> get_panoptic_segmentation -> instance loop ->
    argmax chain (tests/golden/post_ii.npz); equality outside the margin bands, in-band pixel count printed."""
    from test_oracle_golden import check_mode_ii
    g = golden("post.npz")
    fg, ce, of, ee = _low(g, case)
    r = ops.post_process(cu(fg), cu(ce), cu(of), cu(ee), scale=4, top_k=int(g[f"{case}_topk"]))
    n = int(r["inst_offsets"][1])
    check_mode_ii(case, r["panoptic"][0].cpu().numpy(), r["centers_yx"][0, : int(r["n_centers"][0])].cpu().numpy(),
                  r["eee_onehot"][0].cpu().numpy(), r["labels"][:n].cpu().numpy(), r["boxes"][:n].cpu().numpy(),
                  r["scores"][:n].cpu().numpy(), sem_seg=r["sem_seg"][0, 0].cpu().numpy(),
                  eee_full=r["eee_full"][0].cpu().numpy(), who="kernels")


@pytest.mark.parametrize("C", [2, 3])
def test_error_decode_e2_e33_e32_heads(C):
    """2- and 3-channel error heads (ERROR_TYPE e2 / e32 / e33, model.py:215-226) through group_kernel's decode."""
    g = golden("decode_types.npz")
    low = g[f"c{C}_low"]
    h, w = low.shape[1:]
    pg = golden("post.npz")
    fg, ce, of, _ = _low(pg, "small")
    assert fg.shape[-2:] == (h, w)
    r = ops.post_process(cu(fg), cu(ce), cu(of), cu(low[None]), scale=4)
    H, W = 4 * h, 4 * w
    ref = np.unpackbits(g[f"c{C}_onehot"])[: C * H * W].reshape(C, H, W)
    assert r["eee_onehot"].shape == (1, C, H, W)
    np.testing.assert_array_equal(r["eee_onehot"][0].cpu().numpy(), ref)
    np.testing.assert_array_equal(r["eee_full"][0].cpu().numpy(), oracle.upsample_bilinear(low, 4))
    # the rest of the outputs does not depend on the error head
    np.testing.assert_array_equal(r["panoptic"][0].cpu().numpy().astype(np.int16), pg["small_panoptic"])


@pytest.mark.parametrize("case", ["small", "c1"])
def test_quadruple_maps_match_executed_panoptic2eee(case):
    """quber_mask_boundary + quber_quadruple_maps against tools/ours/panoptic2eee.py executed (tests/golden/quadruple.npz)."""
    g = golden("quadruple.npz")
    shape = tuple(int(x) for x in g[f"{case}_shape"])
    K, H, W = shape
    n = K * H * W
    gt = np.unpackbits(g[f"{case}_gt"])[:n].reshape(shape).astype(np.uint8) * 255
    pr = np.unpackbits(g[f"{case}_pred"])[:n].reshape(shape).astype(np.uint8) * 255
    d = ops.boundary_dilation(H, W, 0.01)
    _, gb, gf = ops.mask_boundary(cu(gt)[None], d, per_instance=False)
    _, pb, pf = ops.mask_boundary(cu(pr)[None], d, per_instance=False)
    qm, qb = ops.quadruple_maps(gf, pf)[0].cpu().numpy(), ops.quadruple_maps(gb, pb)[0].cpu().numpy()
    for i, name in enumerate(("tp", "tn", "fp", "fn")):
        np.testing.assert_array_equal(qm[i], np.unpackbits(g[f"{case}_{name}_mask"])[: H * W].reshape(H, W))
        np.testing.assert_array_equal(qb[i], np.unpackbits(g[f"{case}_{name}_boundary"])[: H * W].reshape(H, W))


def test_c2_all_32_frames_against_oracle():
    """Every frame of configs[1] (32 frames 640x480, 30 masks) against the oracle, bit for bit: encode, per-instance and
    union boundaries, foreground, and the whole post-processing output set."""
    from concurrent.futures import ThreadPoolExecutor
    B, K, H, W = 32, 30, 480, 640
    masks = np.stack([synth.ellipse_masks(K, H, W, seed=300 + b) for b in range(B)])
    lows = np.stack([np.concatenate(list(synth.head_outputs(masks[b], seed=300 + b).values()), 0) for b in range(B)])
    dm = cu(masks)
    enc = ops.encode_initial_masks(dm).cpu().numpy()
    per, uni, fg = (t.cpu().numpy() for t in ops.mask_boundary(dm, ops.boundary_dilation(H, W, 0.01)))
    r = ops.post_process(cu(lows[:, 0:1]), cu(lows[:, 1:2]), cu(lows[:, 2:4]), cu(lows[:, 4:8]), scale=4, max_instances=B * 64)
    r = {k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in r.items()}

    def one(b):                       # the oracle is C called through ctypes: the GIL is released, frames run in parallel
        u, p = oracle.masks_to_boundary(masks[b], 0.01, return_per_instance=True)
        full = oracle.upsample_bilinear(lows[b], 4)
        full[2:4] *= 4.0
        return (oracle.encode_initial_masks(masks[b]), u, p, oracle.masks_to_fg_mask(masks[b]),
                oracle.post_process(full[0], full[1], full[2:4], full[4:8]))

    with ThreadPoolExecutor(8) as ex:
        refs = list(ex.map(one, range(B)))
    for b, (e, u, p, f, post) in enumerate(refs):
        np.testing.assert_array_equal(enc[b], e)
        np.testing.assert_array_equal(per[b], p)
        np.testing.assert_array_equal(uni[b], u)
        np.testing.assert_array_equal(fg[b], f)
        _check_post(r, post, H, W, b=b)


# ------------------------------------------------------------------ round 2: fused MSDA, fp64 MSDA, 3xTF32 GEMM
@pytest.mark.parametrize("D,dt", [(32, "f32"), (16, "f32"), (64, "f32"), (32, "bf16"), (64, "bf16")])
@pytest.mark.parametrize("L", [1, 3, 4])
def test_msda_fused_equals_prepare_plus_forward_and_oracle(D, dt, L):
    """quber_msda_forward_fused (softmax + ref + off / (W,H) in the set-up phase) against the two-kernel route and, with
    torch's softmax / location arithmetic in between, the oracle.  OPS/modules/ms_deform_attn.py:102-117."""
    shapes_l = [(15, 20), (8, 10), (4, 5), (2, 3)][:L]
    g = torch.Generator(device=DEV).manual_seed(100 * D + L)
    N, M, P, Q = 2, 8, 4, 777
    shapes = torch.tensor(shapes_l, dtype=torch.long, device=DEV)
    starts = torch.cat((shapes.new_zeros((1,)), shapes.prod(1).cumsum(0)[:-1]))
    S = int(shapes.prod(1).sum())
    value = torch.rand(N, S, M, D, device=DEV, generator=g) * 0.01
    oa = torch.randn(N, Q, M * L * P * 3, device=DEV, generator=g) * 2.0            # offsets of a few pixels: border crossings
    ref_pts = torch.rand(N, Q, L, 2, device=DEV, generator=g) * 1.1 - 0.05
    vt = value.bfloat16() if dt == "bf16" else value
    fused = ops.msda_forward_fused(vt, shapes, starts, oa, ref_pts, L, P)
    # the level table handed over on the host as well (parameter bank instead of shared memory): the same bits
    assert torch.equal(fused, ops.msda_forward_fused(vt, shapes, starts, oa, ref_pts, L, P, levels=shapes_l))
    with pytest.raises(ValueError):
        ops.msda_forward_fused(vt, shapes, starts, oa, ref_pts, L, P, levels=shapes_l[:-1] + [(7, 7)])
    loc, aw = ops.msda_prepare(oa, ref_pts, shapes, M, L, P)
    two = ops.msda_forward(vt, shapes, starts, loc, aw)
    assert fused.dtype == vt.dtype and fused.shape == (N, Q, M * D)
    tol = 1e-4 if dt == "bf16" else 2e-8                                            # softmax summed in a different order
    np.testing.assert_allclose(fused.float().cpu().numpy(), two.float().cpu().numpy(), rtol=0, atol=tol)
    if dt == "f32":
        n_off = M * L * P * 2
        off = oa[..., :n_off].reshape(N, Q, M, L, P, 2)
        aw_t = torch.softmax(oa[..., n_off:].reshape(N, Q, M, L * P), -1).view(N, Q, M, L, P)
        wh = torch.stack([shapes[..., 1], shapes[..., 0]], -1)
        loc_t = ref_pts[:, :, None, :, None, :] + off / wh[None, None, None, :, None, :]
        want = oracle.msda_forward(value.cpu().numpy(), shapes.cpu().numpy(), starts.cpu().numpy(),
                                   loc_t.contiguous().cpu().numpy(), aw_t.contiguous().cpu().numpy())
        np.testing.assert_allclose(fused.cpu().numpy(), want, rtol=0, atol=1e-7)


@pytest.mark.parametrize("rows,N,K1,K2,dt", [(1000, 256, 256, 256, "bf16"), (6300, 512, 512, 512, "bf16"), (333, 256, 64, 192, "bf16"),
                                              (515, 128, 128, 96, "bf16"), (700, 256, 96, 160, "f32")])
def test_gemm_cat2_equals_gemm_of_the_concatenation(rows, N, K1, K2, dt):
    """quber_gemm_bias_act_cat2: the K-concatenation [A1 | A2] is never materialised; the k-blocks are issued in the same
    order as for the concatenated matrix, so the result equals quber_gemm_bias_act(cat) bit for bit; and float64 within
    the GEMM tests' tolerance."""
    g = torch.Generator(device=DEV).manual_seed(rows + K1)
    tdt = torch.bfloat16 if dt == "bf16" else torch.float32
    a1 = (torch.randn(rows, K1, device=DEV, generator=g) * 0.5).to(tdt)
    a2 = (torch.randn(rows, K2, device=DEV, generator=g) * 0.5).to(tdt)
    w = (torch.randn(N, K1 + K2, device=DEV, generator=g) * 0.1).to(tdt)
    b = torch.randn(N, device=DEV, generator=g)
    for out_dtype in (torch.float32, torch.bfloat16):
        got = ops.gemm_bias_act_cat2(a1, a2, w, b, act="relu", out_dtype=out_dtype)
        want = ops.gemm_bias_act(torch.cat([a1, a2], 1), w, b, act="relu", out_dtype=out_dtype)
        assert got.dtype == out_dtype and torch.equal(got, want)
    ref = torch.relu(torch.cat([a1, a2], 1).double() @ w.double().T + b.double())
    got = ops.gemm_bias_act_cat2(a1, a2, w, b, act="relu", out_dtype=torch.float32)
    tol = 1e-4 if dt == "bf16" else 2e-2                    # bf16 inputs are exact products; fp32 inputs multiply as TF32
    assert float((got.double() - ref).abs().max()) <= tol * max(1.0, float(ref.abs().max()))
    # strided views of channels-last activations: row pitch larger than K
    big = (torch.randn(rows, K1 + K2 + 64, device=DEV, generator=g) * 0.5).to(tdt)
    v1, v2 = big[:, :K1], big[:, K1 + 64:]
    assert torch.equal(ops.gemm_bias_act_cat2(v1, v2, w, b), ops.gemm_bias_act(torch.cat([v1, v2], 1).contiguous(), w, b))
    with pytest.raises(ValueError):
        ops.gemm_bias_act_cat2(a1[:, :-8], a2, w[:, :-8].contiguous(), b)                     # K1 off a k-block boundary


@pytest.mark.parametrize("case", list(MSDA_CASES))
def test_msda_float64_is_real_float64(case):
    """The reference dispatches fp64 (ms_deform_attn_cuda.cu:69, OPS/test.py:24-44 checks it with allclose at fp64
    tolerances): fp64 inputs are computed in fp64, not in fp32 and cast back."""
    from quber_b200.msda import MSDeformAttnFunction
    g = golden("msda.npz")
    inp = synth.msda_inputs(dtype=np.float64, **MSDA_CASES[case])
    out = MSDeformAttnFunction.apply(cu(inp["value"]), cu(inp["shapes"]), cu(inp["starts"]), cu(inp["loc"]), cu(inp["attn"]), 2)
    assert out.dtype == torch.float64
    np.testing.assert_allclose(out.cpu().numpy(), g[f"{case}_out_f64"], rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("shape", [(128, 128, 64), (300, 256, 256), (1000, 288, 256), (515, 1024, 256), (515, 256, 1024),
                                   (77, 40, 168), (333, 100, 72), (6300, 256, 256)])
def test_gemm_x3_fp32_accuracy_on_arbitrary_inputs(shape):
    """quber_gemm_bias_act_x3 (split-operand 3xTF32) on arbitrary fp32 inputs against float64: <= 1e-5 of the row scale
    sum|a||w| (a single TF32 product is at 1e-3); cuBLAS sgemm itself sits at ~1e-6."""
    from quber_b200 import _lib
    M, N, K = shape
    g = torch.Generator(device="cuda").manual_seed(M * 3 + N + K)
    a = torch.randn(M, K, device=DEV, generator=g) * torch.rand(M, 1, device=DEV, generator=g) * 10
    w = torch.randn(N, K, device=DEV, generator=g) / K ** 0.5
    b = torch.randn(N, device=DEV, generator=g)
    hi, lo = ops.split_tf32(w)
    assert torch.equal(hi + lo, w) and int((hi.view(torch.int32) & 0x1FFF).abs().max()) == 0
    ref = torch.relu(a.double() @ w.double().T + b.double())
    scale = a.abs().double() @ w.abs().double().T + b.abs().double()
    out = ops.gemm_bias_act_x3(a, hi, lo, b, act="relu")
    err = ((out.double() - ref).abs() / scale).max().item()
    one = ops.gemm_bias_act(a, w, b, act="relu")
    err1 = ((one.double() - ref).abs() / scale).max().item()
    print(f"[x3 gemm {shape}] max err / sum|a||w|: 3xTF32 {err:.2e}, 1xTF32 {err1:.2e}")
    assert err <= 1e-5 and err1 > err
    lin = ops.gemm_bias_act_x3(a, hi, lo, None, out_dtype=torch.bfloat16) if N % 8 == 0 else None
    if lin is not None:
        torch.testing.assert_close(lin.double(), a.double() @ w.double().T, rtol=2 ** -7, atol=1e-3 * float(scale.max()))
    assert _lib.lib().quber_gemm_timeouts() == 0


def test_extension_module_name_resolves_to_these_kernels():
    """`import MultiScaleDeformableAttention as MSDA` (OPS/functions/ms_deform_attn_func.py:21)."""
    import sys
    from quber_b200 import msda
    sys.modules.pop("MultiScaleDeformableAttention", None)
    msda.install_extension_module()
    import MultiScaleDeformableAttention as MSDA
    inp = synth.msda_inputs(**MSDA_CASES["medium"])
    out = MSDA.ms_deform_attn_forward(cu(inp["value"]), cu(inp["shapes"]), cu(inp["starts"]), cu(inp["loc"]), cu(inp["attn"]), 64)
    np.testing.assert_allclose(out.cpu().numpy(), golden("msda.npz")["medium_out_f32"], rtol=0, atol=1e-6)
    with pytest.raises(NotImplementedError):
        MSDA.ms_deform_attn_backward()
