"""The end-to-end host call (HotPath.refine_batch_host: chunked, copy/compute pipelined over three streams)
against the resident single-batch path and the oracle.  B200 only."""
import numpy as np
import pytest
import torch

import oracle
from quber_b200 import ops, synth
from quber_b200.pipeline import HotPath, RefinerConfig

pytestmark = pytest.mark.gpu


def _assert_staged(hp, hm):
    """Every frame of the last host call reached the device by one of the two routes: its bytes in the byte staging
    buffer, or its bits in the packed words (the packed route no longer expands them into bytes)."""
    st = hp._stage
    dm = hm.cuda()
    ok_bytes = (st["masks"] == dm).flatten(1).all(1)
    if dm[0].numel() % 32 == 0:
        bits = ops.unpack_masks(st["packed_dev"], tuple(dm.shape), value=1)
        ok_bits = ((bits != 0) == (dm != 0)).flatten(1).all(1)
    else:
        ok_bits = torch.zeros_like(ok_bytes)
    assert bool((ok_bytes | ok_bits).all())


def _inputs(B, K, H, W, seed):
    masks = np.stack([synth.ellipse_masks(K, H, W, seed=seed + b, axes=(14, 40), margin=30) for b in range(B)])
    heads = [synth.head_outputs(masks[b], seed=seed + b) for b in range(B)]
    head = {k: np.stack([h[k] for h in heads]) for k in heads[0]}
    return masks, head


@pytest.mark.parametrize("B,chunk", [(6, 4), (5, 2), (3, 8)])
def test_refine_batch_host_matches_resident_path_and_oracle(B, chunk):
    K, H, W = 7, 240, 320
    masks, head = _inputs(B, K, H, W, seed=40)
    hp = HotPath(RefinerConfig(), "cuda")
    hm = torch.from_numpy(masks).pin_memory()
    hh = {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}
    for rep in range(3):                                    # staging buffers and pinned results are reused between calls
        out, keep = hp.refine_batch_host(hm, hh, chunk_frames=chunk)
        torch.cuda.current_stream().synchronize()
    dm = hm.cuda()
    enc, per, uni, fg = hp.encode_and_boundary(dm)
    r = hp.post({k: v.cuda() for k, v in hh.items()})
    offs = r["inst_offsets"].cpu().numpy()
    assert np.array_equal(out["label_map"].numpy(), r["label_map"].cpu().numpy())
    assert np.array_equal(out["boundary"].numpy(), uni.cpu().numpy())
    assert np.array_equal(out["inst_counts"].numpy(), offs[1:] - offs[:-1])
    assert out["inst_counts"].sum() > 0
    for b in range(B):
        s, n = int(out["inst_start"][b]), int(out["inst_counts"][b])
        for key in ("labels", "scores", "boxes"):
            assert np.array_equal(out[key][s:s + n].numpy(), r[key][offs[b]:offs[b + 1]].cpu().numpy()), (key, b)
    # oracle on frame 0 and the last frame (which sits in a short tail chunk)
    for b in (0, B - 1):
        low = np.concatenate([head["foreground"][b], head["center"][b], head["offset"][b], head["eee_boundary"][b]], 0)
        full = oracle.upsample_bilinear(low, 4)
        full[2:4] *= 4.0
        ref = oracle.post_process(full[0], full[1], full[2:4], full[4:8])
        lm = out["label_map"][b].numpy()
        assert int(out["inst_counts"][b]) == len(ref["labels"])
        for j in range(len(ref["labels"])):
            assert np.array_equal(lm == j + 1, ref["masks"][j] != 0)
        u, _ = oracle.masks_to_boundary(masks[b], 0.01, return_per_instance=True)
        assert np.array_equal(out["boundary"][b].numpy(), u)
    assert keep[0][0].is_cuda                               # device results of every chunk stay alive


def test_refine_batch_host_packed_mask_transfer_is_identical():
    """One bit per pixel over the host link (quber_host_pack_masks + quber_unpack_masks) gives the same device masks,
    hence the same results; a chunk carrying more than one non-zero value falls back to the byte copy."""
    K, H, W, B = 6, 240, 320, 7
    sets = []
    for s in range(2):
        masks, head = _inputs(B, K, H, W, seed=300 + s)
        sets.append((torch.from_numpy(masks).pin_memory(), {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}))
    plain, packed = HotPath(RefinerConfig(), "cuda"), HotPath(RefinerConfig(), "cuda")
    plain.pack_masks, packed.pack_masks = False, True
    for it in range(4):                                     # alternate inputs: pinned packed words are reused between calls
        hm, hh = sets[it % 2]
        o0, k0 = plain.refine_batch_host(hm, hh, chunk_frames=2)
        o1, k1 = packed.refine_batch_host(hm, hh, chunk_frames=2)
        torch.cuda.current_stream().synchronize()
        _assert_staged(packed, hm)
        for key in ("label_map", "boundary", "inst_start", "inst_counts"):
            assert torch.equal(o0[key], o1[key]), key
        for b in range(B):                                  # rows past inst_counts[b] are never written
            s0, n = int(o0["inst_start"][b]), int(o0["inst_counts"][b])
            for key in ("labels", "scores", "boxes"):
                assert torch.equal(o0[key][s0:s0 + n], o1[key][s0:s0 + n]), (key, b)
        for i in (0, 1):                                    # encoding and per-instance boundaries (group cuts differ by route)
            assert torch.equal(torch.cat([k[i] for k in k0]), torch.cat([k[i] for k in k1]))
    hm = sets[0][0].clone().pin_memory()
    hm[3][hm[3] != 0] = 7                                   # frame 3 (chunk 2..3) carries a second value: byte copy
    hm[5, 2][hm[5, 2] != 0] = 9                             # two values inside one frame
    o0, k0 = plain.refine_batch_host(hm, sets[0][1], chunk_frames=2)
    o1, k1 = packed.refine_batch_host(hm, sets[0][1], chunk_frames=2)
    torch.cuda.current_stream().synchronize()
    _assert_staged(packed, hm)
    assert torch.equal(torch.cat([k[1] for k in k0]), torch.cat([k[1] for k in k1]))   # boundary values follow the mask's own value (util.py:72-90)
    for src in (sets[1][0], (sets[1][0] != 0).to(torch.uint8)):            # 0/255 as the reference's eval makes them, and 0/1
        words, value = ops.host_pack_masks(src)
        assert value == int(src.max())
        assert torch.equal(ops.unpack_masks(words.cuda(), src.shape, value=value).cpu(), src)


@pytest.mark.parametrize("K,value", [(7, 255), (4, 1), (30, 255)])
def test_encode_from_packed_words_equals_encode_from_bytes(K, value):
    """quber_encode_initial_masks_packed: same encoding, and the byte planes it writes on the way are the masks."""
    B, H, W = 3, 240, 320
    masks = np.stack([synth.ellipse_masks(K, H, W, seed=70 + b, axes=(14, 60), margin=20, empty_tail=1 if K > 4 else 0) for b in range(B)])
    masks = ((masks != 0) * value).astype(np.uint8)
    hm = torch.from_numpy(masks)
    words, v = ops.host_pack_masks(hm)
    assert v == value
    out_masks = torch.full(hm.shape, 77, dtype=torch.uint8, device="cuda")
    enc = ops.encode_initial_masks_packed(words.cuda(), v, out_masks)
    assert torch.equal(out_masks.cpu(), hm)
    assert torch.equal(enc, ops.encode_initial_masks(hm.cuda()))
    np.testing.assert_array_equal(enc[1].cpu().numpy(), oracle.encode_initial_masks(masks[1]))
    with pytest.raises(ValueError):
        ops.encode_initial_masks_packed(words.cuda()[:-1], v, out_masks)


def test_refine_batch_host_full_size_c2_packed_graphs_equal_resident():
    """BASELINE.json C2 shape (32 frames 640x480, 30 masks of 0/255) through the call bench.py times — packed transfer,
    one CUDA graph per chunk, 16 + 8 + 8 frames — against the resident path on the same inputs, twice (graph replays)."""
    B, K, H, W = 32, 30, 480, 640
    base_m, base_h = _inputs(4, K, H, W, seed=900)
    masks = np.concatenate([base_m[(np.arange(4) + s) % 4] for s in range(B // 4)])
    head = {k: np.concatenate([v[(np.arange(4) + s) % 4] for s in range(B // 4)]) for k, v in base_h.items()}
    hm = torch.from_numpy(masks).pin_memory()
    hh = {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}
    hp = HotPath(RefinerConfig(), "cuda")
    hp.pack_masks, hp.use_graphs = True, True
    ref = HotPath(RefinerConfig(), "cuda")
    enc, per, uni, fg = ref.encode_and_boundary(hm.cuda())
    r = ref.post({k: v.cuda() for k, v in hh.items()})
    offs = r["inst_offsets"].cpu().numpy()
    for rep in range(3):
        out, keep = hp.refine_batch_host(hm, hh, chunk_frames=16, min_tail=8, byte_frames=0)     # every frame bit-packed
        torch.cuda.current_stream().synchronize()
        assert [k[0].shape[0] for k in keep] == [16, 8, 8]
        _assert_staged(hp, hm)
        assert np.array_equal(out["label_map"].numpy(), r["label_map"].cpu().numpy())
        assert np.array_equal(out["boundary"].numpy(), uni.cpu().numpy())
        assert np.array_equal(out["inst_counts"].numpy(), offs[1:] - offs[:-1])
        assert torch.equal(torch.cat([k[0] for k in keep]), enc) and torch.equal(torch.cat([k[1] for k in keep]), per)
        for b in (0, 15, 16, 31):
            s0, n = int(out["inst_start"][b]), int(out["inst_counts"][b])
            assert n > 0
            for key in ("labels", "scores", "boxes"):
                assert np.array_equal(out[key][s0:s0 + n].numpy(), r[key][offs[b]:offs[b + 1]].cpu().numpy()), (key, b)
    assert int(out["inst_counts"].sum()) >= 5 * B          # size-independent sanity: every frame keeps most of its objects


def test_refine_batch_host_new_inputs_between_calls():
    """Back-to-back calls with different inputs must not race on the shared staging buffers."""
    K, H, W, B = 5, 240, 320, 4
    hp = HotPath(RefinerConfig(), "cuda")
    sets = []
    for s in range(3):
        masks, head = _inputs(B, K, H, W, seed=100 * s)
        sets.append((torch.from_numpy(masks).pin_memory(), {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}))
    expect = []
    for hm, hh in sets:
        r = hp.post({k: v.cuda() for k, v in hh.items()})
        expect.append(r["label_map"].cpu().numpy())
    for it in range(6):
        hm, hh = sets[it % 3]
        out, _ = hp.refine_batch_host(hm, hh, chunk_frames=2)
        torch.cuda.current_stream().synchronize()
        assert np.array_equal(out["label_map"].numpy(), expect[it % 3])


@pytest.mark.parametrize("use_fg,use_zd", [(True, True), (True, False), (False, True)])
def test_filter_instances_matches_reference_loop(use_fg, use_zd):
    """Foreground-overlap filter (> 0.3, float64 division) and OCID zero-depth clear (eval/refiner_model.py:271-288)
    against the numpy restatement, on label maps of disjoint instances incl. ratios at exactly 0.3."""
    rng = np.random.default_rng(5)
    B, Hh, Ww, n = 3, 120, 160, 12
    lm = np.zeros((B, Hh, Ww), np.int32)
    for b in range(B):
        for j in range(n):                                              # disjoint 10-row bands split in two widths
            lm[b, 10 * j:10 * j + 10, 5 * j:5 * j + 60] = j + 1
    fg = (rng.random((B, Hh, Ww)) < 0.35).astype(np.uint8) * 255
    fg[0, 0:10, :] = 0
    fg[0, 0:3, 0:60] = 1                                                # instance 1 of frame 0: overlap 180 / 600 = 0.3 exactly -> dropped
    fg[0, 10:20, :] = 0
    fg[0, 10:13, 5:65] = 7; fg[0, 13, 5] = 9                            # instance 2: 181 / 600 > 0.3 -> kept
    zd = (rng.random((B, Hh, Ww)) < 0.1).astype(np.uint8)
    counts = np.array([n, n - 2, n], np.int32)                          # frame 1 reports fewer instances: the last two labels are dropped
    out, keep, newc = ops.filter_instances(torch.from_numpy(lm).cuda(), torch.from_numpy(counts).cuda(),
                                           torch.from_numpy(fg).cuda() if use_fg else None,
                                           torch.from_numpy(zd).cuda() if use_zd else None, thr=0.3, max_per=16)
    out, keep, newc = out.cpu().numpy(), keep.cpu().numpy().astype(bool), newc.cpu().numpy()
    for b in range(B):
        masks = np.stack([lm[b] == j + 1 for j in range(counts[b])])
        ref, rkeep = oracle.filter_instances(masks, fg[b] if use_fg else None, zd[b] if use_zd else None, 0.3)
        assert np.array_equal(keep[b, :counts[b]], rkeep) and not keep[b, counts[b]:].any()
        assert newc[b] == len(ref)
        for i in range(len(ref)):
            assert np.array_equal(out[b] == i + 1, ref[i]), (b, i)
        assert out[b].max() <= len(ref)
    if use_fg:
        assert not keep[0, 0] and keep[0, 1]


def test_refiner_model_adapter_filters_like_the_reference():
    """quber_b200.refiner_model.MaskRefiner.predict (eval/refiner_model.py:224-297) around a stub predictor: OSD keeps the
    fg-overlapping instances, OCID also clears zero-depth pixels, armbench returns the instances untouched."""
    from quber_b200.meta_arch import Instances
    from quber_b200.refiner_model import MaskRefiner
    Hh, Ww = 240, 320
    masks = synth.ellipse_masks(7, Hh, Ww, seed=3, axes=(12, 30), margin=30) != 0
    taken = np.zeros((Hh, Ww), bool)
    for m in masks:                                                    # refined instances are disjoint (model.py:331)
        m &= ~taken
        taken |= m
    rng = np.random.default_rng(1)
    fg = np.zeros((Hh, Ww), bool)
    for j in (0, 2, 3, 6):
        fg |= masks[j]
    fg &= rng.random((Hh, Ww)) < 0.8
    depth = rng.integers(300, 1400, (Hh, Ww)).astype(np.uint16)
    depth[rng.random((Hh, Ww)) < 0.08] = 0

    class Stub:
        def predict(self, rgb, dep, init):
            assert init.dtype == np.uint8 and init.max() == 255        # bool masks become 0 / 255 (:267)
            assert dep is None or (dep.dtype == np.uint8 and dep.shape == (Hh, Ww, 3))
            return [{"instances": Instances((Hh, Ww), pred_masks=torch.from_numpy(masks).cuda())}]

    rgb = np.zeros((Hh, Ww, 3), np.uint8)
    for dataset in ("OSD", "OCID", "armbench"):
        r = MaskRefiner.__new__(MaskRefiner)
        r.refiner_predictor, r.dataset, r.fg_model, r.inpaint = Stub(), dataset, None, None
        got, output, secs, fgm = r.predict(rgb, None if dataset == "armbench" else depth, masks, fg_mask=fg)
        if dataset == "armbench":
            assert np.array_equal(got, masks) and fgm is None
            continue
        ref, keep = oracle.filter_instances(masks, fg, depth == 0 if dataset == "OCID" else None, 0.3)
        assert keep.sum() == 4 and got.dtype == bool
        assert np.array_equal(got, ref), dataset
        assert "instances" in output and secs >= 0


def test_post_process_uint8_label_map_equals_int32():
    """QUBER_POST_LABEL_U8: the emit kernel writes the same ids as bytes; more than 255 possible ids is refused."""
    masks, head = _inputs(3, 9, 240, 320, seed=510)
    h = {k: torch.from_numpy(v).cuda() for k, v in head.items()}
    a = ops.post_process(h["foreground"], h["center"], h["offset"], h["eee_boundary"], want=("label_map", "masks"))
    b = ops.post_process(h["foreground"], h["center"], h["offset"], h["eee_boundary"], want=("label_map", "masks"),
                         label_dtype=torch.uint8)
    assert b["label_map"].dtype == torch.uint8 and a["label_map"].dtype == torch.int32
    assert int(a["label_map"].max()) > 0
    assert torch.equal(a["label_map"], b["label_map"].to(torch.int32))
    for key in ("masks", "labels", "scores", "boxes", "inst_offsets"):
        n = int(a["inst_offsets"][-1])
        assert torch.equal(a[key][:n] if key != "inst_offsets" else a[key], b[key][:n] if key != "inst_offsets" else b[key]), key
    with pytest.raises(ValueError):
        ops.post_process(h["foreground"], h["center"], h["offset"], None, top_k=256, label_dtype=torch.uint8)


@pytest.mark.parametrize("B,chunk,tail,kf,sizes", [(7, 2, 1, 4, [4, 3]), (7, 2, 1, 100, [7]), (6, 4, 2, 0, [4, 2])])
def test_refine_batch_host_kernel_groups_and_uint8_label_map(B, chunk, tail, kf, sizes):
    """kernel_frames: the masks cross the link chunk by chunk, encode / boundary run once per group of chunks; with
    compact_label_map the host label map is uint8.  Same results as the per-chunk int32 call, packed or not; a group whose
    chunks took different routes (one of them carries two mask values) falls back to chunk-by-chunk launches."""
    K, H, W = 6, 240, 320
    masks, head = _inputs(B, K, H, W, seed=620)
    hm = torch.from_numpy(masks).pin_memory()
    hh = {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}
    ref = HotPath(RefinerConfig(), "cuda")
    ref.pack_masks = False
    for pack in (False, True):
        hp = HotPath(RefinerConfig(compact_label_map=True), "cuda")
        hp.pack_masks = pack
        for rep in range(3):
            o1, k1 = hp.refine_batch_host(hm, hh, chunk_frames=chunk, min_tail=tail, kernel_frames=kf, byte_frames=0 if pack else None)
            torch.cuda.current_stream().synchronize()
        o0, k0 = ref.refine_batch_host(hm, hh, chunk_frames=chunk, min_tail=tail)
        torch.cuda.current_stream().synchronize()
        assert [k[0].shape[0] for k in k1] == sizes
        assert o1["label_map"].dtype == torch.uint8 and o0["label_map"].dtype == torch.int32
        assert torch.equal(o1["label_map"].to(torch.int32), o0["label_map"])
        for key in ("boundary", "inst_start", "inst_counts"):
            assert torch.equal(o0[key], o1[key]), key
        assert torch.equal(torch.cat([k[0] for k in k1]), torch.cat([k[0] for k in k0]))
        assert torch.equal(torch.cat([k[1] for k in k1]), torch.cat([k[1] for k in k0]))
        _assert_staged(hp, hm)
    if kf:
        hm2 = hm.clone().pin_memory()
        hm2[1][hm2[1] != 0] = 7                              # chunk (0, 2) now carries two values: byte route for it
        hp = HotPath(RefinerConfig(compact_label_map=True), "cuda")
        hp.pack_masks = True
        o1, k1 = hp.refine_batch_host(hm2, hh, chunk_frames=chunk, min_tail=tail, kernel_frames=kf, byte_frames=0)
        o0, k0 = ref.refine_batch_host(hm2, hh, chunk_frames=chunk, min_tail=tail)
        torch.cuda.current_stream().synchronize()
        assert sum(k[0].shape[0] for k in k1) == B and len(k1) > len(sizes)
        assert torch.equal(o0["boundary"], o1["boundary"])
        assert torch.equal(torch.cat([k[1] for k in k1]), torch.cat([k[1] for k in k0]))
        _assert_staged(hp, hm2)


@pytest.mark.parametrize("B,K,H,W,kf", [(5, 7, 240, 320, 2), (8, 30, 480, 640, 0)])
def test_refine_batch_rle_equals_dense_host_call(B, K, H, W, kf):
    """The RLE route (counts over the host link, quber_rle_decode on the device) gives the dense call's results bit for
    bit; the RLEs come from the oracle's encoder as compressed strings, parsed by quber_rle_from_string."""
    from quber_b200 import rle
    masks, head = _inputs(B, K, H, W, seed=700)
    rles = [{"size": [H, W], "counts": oracle.rle_to_string(oracle.rle_encode(m))} for m in masks.reshape(-1, H, W)]
    counts, offsets, hw = rle.pack_rles(rles)
    assert hw == (H, W) and counts.nbytes < masks.nbytes // 100
    hc, ho = torch.from_numpy(counts).pin_memory(), torch.from_numpy(offsets).pin_memory()
    hm = torch.from_numpy(masks).pin_memory()
    hh = {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}
    dense, sparse = HotPath(RefinerConfig(), "cuda"), HotPath(RefinerConfig(), "cuda")
    o0, k0 = dense.refine_batch_host(hm, hh, chunk_frames=max(B // 2, 1))
    torch.cuda.current_stream().synchronize()
    o0 = {k: v.clone() for k, v in o0.items()}
    for rep in range(3):                                    # graph replays, staging reuse
        o1, k1 = sparse.refine_batch_rle(hc, ho, (B, K, H, W), hh, value=255, kernel_frames=kf)
        torch.cuda.current_stream().synchronize()
        assert int(o1["rle_status"][0]) == 0 and int(o1["status"][0]) == 0
        _assert_staged(sparse, hm)
        for key in ("label_map", "boundary", "inst_start", "inst_counts"):
            assert torch.equal(o0[key], o1[key]), key
        for b in range(B):
            s0, n = int(o0["inst_start"][b]), int(o0["inst_counts"][b])
            for key in ("labels", "scores", "boxes"):
                assert torch.equal(o0[key][s0:s0 + n], o1[key][s0:s0 + n]), (key, b)
    enc0 = torch.cat([k[0] for k in k0]); enc1 = torch.cat([k[0] for k in k1])
    assert torch.equal(enc0, enc1)
    np.testing.assert_array_equal(enc1[B - 1].cpu().numpy(), oracle.encode_initial_masks(masks[B - 1]))


def test_custom_ops_compile_fullgraph_and_opcheck():
    """torch.compile(fullgraph=True) traces the path's custom ops without a graph break (fake kernels) and the compiled
    function returns the eager kernels' bits; torch.library.opcheck validates schema and fake-tensor consistency."""
    K, H, W = 6, 240, 320
    masks, head = _inputs(2, K, H, W, seed=77)
    fg, ce, of, ee = (torch.from_numpy(head[k]).cuda() for k in ("foreground", "center", "offset", "eee_boundary"))
    dm = torch.from_numpy(masks).cuda()
    args = (fg, ce, of, ee, 4, 4.0, 0.3, 7, 200, 512, 2048, 1000, -1.0, 2 * 200, True)

    def f(dm, fg, ce, of, ee):
        enc = torch.ops.quber.encode_initial_masks(dm, 10)
        per, uni, fgm = torch.ops.quber.mask_boundary(dm, 3)
        post = torch.ops.quber.post_process(fg, ce, of, ee, 4, 4.0, 0.3, 7, 200, 512, 2048, 1000, -1.0, 2 * 200, True)
        return enc, uni, post[3], post[4], post[8]

    eager = f(dm, fg, ce, of, ee)
    comp = torch.compile(f, fullgraph=True, backend="eager")(dm, fg, ce, of, ee)
    for a, b in zip(eager, comp):
        assert torch.equal(a, b)
    ref = ops.post_process(fg, ce, of, ee, scale=4, max_instances=400, label_dtype=torch.uint8)
    assert torch.equal(eager[2], ref["panoptic"]) and torch.equal(eager[3], ref["label_map"])
    torch.library.opcheck(torch.ops.quber.post_process.default, args, test_utils=("test_schema", "test_faketensor"))
    torch.library.opcheck(torch.ops.quber.mask_boundary.default, (dm, 3), test_utils=("test_schema", "test_faketensor"))
    a = torch.randn(300, 256, device="cuda")
    w = torch.randn(192, 256, device="cuda")
    hi, lo = torch.ops.quber.split_tf32(w)
    torch.library.opcheck(torch.ops.quber.gemm_bias_act_x3.default, (a, hi, lo, None, True, False), test_utils=("test_schema", "test_faketensor"))


@pytest.mark.parametrize("B,nb,kf", [(7, 3, 0), (7, 5, 4), (8, 8, 4), (6, 0, 2), (9, None, 4)])
def test_refine_batch_host_hybrid_route_bytes_and_packed_side_by_side(B, nb, kf):
    """The dense call reads the host masks with two readers at once: the first `byte_frames` frames go through the copy
    engine as bytes while the pool threads bit-pack the others (HotPath.byte_frames_rule by default).  Same results as the
    all-bytes call for every split, groups returned in frame order."""
    K, H, W = 6, 240, 320
    masks, head = _inputs(B, K, H, W, seed=810)
    hm = torch.from_numpy(masks).pin_memory()
    hh = {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}
    ref = HotPath(RefinerConfig(compact_label_map=True), "cuda")
    ref.pack_masks = False
    o0, k0 = ref.refine_batch_host(hm, hh, chunk_frames=2)
    torch.cuda.current_stream().synchronize()
    o0 = {k: v.clone() for k, v in o0.items()}
    e0, p0 = torch.cat([k[0] for k in k0]).clone(), torch.cat([k[1] for k in k0]).clone()
    hp = HotPath(RefinerConfig(compact_label_map=True), "cuda")
    hp.pack_masks = True
    for rep in range(3):
        o1, k1 = hp.refine_batch_host(hm, hh, chunk_frames=2, min_tail=1, kernel_frames=kf, byte_frames=nb)
        torch.cuda.current_stream().synchronize()
        HotPath.check_status(o1)
        want_nb = hp.byte_frames_rule(B, hm[0].numel(), sum(v.numel() * v.element_size() for v in hh.values())) if nb is None else nb
        assert hp.last_route == dict(byte_frames=want_nb, packed_frames=B - want_nb if want_nb < B else 0)
        assert [k[5] for k in k1] == sorted(k[5] for k in k1) and sum(k[0].shape[0] for k in k1) == B
        for key in ("label_map", "boundary", "inst_start", "inst_counts"):
            assert torch.equal(o0[key], o1[key]), key
        assert torch.equal(torch.cat([k[0] for k in k1]), e0) and torch.equal(torch.cat([k[1] for k in k1]), p0)
        _assert_staged(hp, hm)
    assert 0 <= hp.byte_frames_rule(32) <= 32 and hp.byte_frames_rule(32, 30 * 480 * 640, 20 << 20) % 4 == 0


def test_refiner_adapter_takes_file_paths_like_the_reference(tmp_path):
    """eval/refiner_model.py:224-297 takes PATHS: frames are read with OpenCV, resized to 640x480 (depth: nearest), the
    depth normalised and in-painted (cv2 TELEA, the reference's own calls); armbench frames go to 750x1333 and the masks
    through the nearest-neighbour resize kernel."""
    import cv2
    from quber_b200.meta_arch import Instances
    from quber_b200.refiner_model import MaskRefiner, inpaint_depth, normalize_depth
    rng = np.random.default_rng(5)
    rgb = rng.integers(0, 255, (240, 320, 3), dtype=np.uint8)
    depth = rng.integers(300, 1400, (240, 320)).astype(np.uint16)
    depth[rng.random((240, 320)) < 0.05] = 0
    cv2.imwrite(str(tmp_path / "rgb.png"), rgb)
    cv2.imwrite(str(tmp_path / "depth.png"), depth)
    np.save(str(tmp_path / "depth_m.npy"), depth.astype(np.float32) / 1000.0)
    seen = {}

    class Stub:
        def predict(self, rgb_img, dep, init):
            seen.update(rgb=rgb_img, dep=dep, init=init)
            H_, W_ = rgb_img.shape[:2]
            return [{"instances": Instances((H_, W_), pred_masks=torch.zeros((1, H_, W_), dtype=torch.bool, device="cuda"))}]

    masks = np.zeros((2, 480, 640), bool); masks[0, 100:200, 100:200] = True
    r = MaskRefiner.__new__(MaskRefiner)
    r.refiner_predictor, r.dataset, r.fg_model, r.inpaint = Stub(), "OSD", None, inpaint_depth
    r.predict(str(tmp_path / "rgb.png"), str(tmp_path / "depth.png"), masks, fg_mask=None)
    assert seen["rgb"].shape == (480, 640, 3) and np.array_equal(seen["rgb"], cv2.resize(rgb, (640, 480)))
    want = cv2.resize(normalize_depth(depth.astype(np.float32)), (640, 480), interpolation=cv2.INTER_NEAREST)
    mask = cv2.dilate(np.all(want == 0, axis=2).astype(np.uint8), np.ones((3, 3), np.uint8), iterations=1)
    want = np.where(want == 0, cv2.inpaint(want, mask, 3, cv2.INPAINT_TELEA), want)
    assert np.array_equal(seen["dep"], want) and (seen["dep"] > 0).mean() > 0.99
    assert seen["init"].dtype == np.uint8 and seen["init"].max() == 255
    r.predict(str(tmp_path / "rgb.png"), str(tmp_path / "depth_m.npy"), masks, fg_mask=None)         # metres branch (:258-259)
    assert seen["dep"].shape == (480, 640, 3) and abs(int(seen["dep"][100, 100, 0]) - int(want[100, 100, 0])) <= 1
    # armbench: 1080 x 1920 -> 750 x 1333, masks by cv2.INTER_NEAREST on the GPU
    big = rng.integers(0, 255, (1080, 1920, 3), dtype=np.uint8)
    cv2.imwrite(str(tmp_path / "big.png"), big)
    bm = np.zeros((3, 1080, 1920), bool); bm[0, 200:700, 300:900] = True; bm[1, ::7, ::5] = True
    r.dataset = "armbench"
    r.predict(str(tmp_path / "big.png"), None, bm, None)
    assert seen["rgb"].shape == (750, 1333, 3) and seen["dep"] is None
    ref_m = np.array([cv2.resize(m, (1333, 750), interpolation=cv2.INTER_NEAREST) for m in np.uint8(bm) * 255])
    assert np.array_equal(seen["init"].cpu().numpy(), ref_m)


@pytest.mark.parametrize("route", ["dense", "rle"])
def test_two_host_calls_in_flight_on_alternating_slots(route):
    """slot / async_result: call i+1 is issued before the results of call i are read (its copies and packing run under
    call i's kernels).  Six calls on two different input sets, alternating slots; every call's host results equal the
    one-call-at-a-time results of the same input bit for bit."""
    from quber_b200 import rle
    B, K, H, W = 6, 9, 240, 320
    data = []
    for seed in (810, 811):
        masks, head = _inputs(B, K, H, W, seed=seed)
        hm = torch.from_numpy(masks).pin_memory()
        hh = {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}
        rles = [{"size": [H, W], "counts": oracle.rle_to_string(oracle.rle_encode(m))} for m in masks.reshape(-1, H, W)]
        counts, offsets, _ = rle.pack_rles(rles)
        data.append((hm, hh, torch.from_numpy(counts).pin_memory(), torch.from_numpy(offsets).pin_memory()))
    keys = ("label_map", "boundary", "inst_start", "inst_counts", "status", "boundary_status")
    one = HotPath(RefinerConfig(), "cuda")
    want = []
    for hm, hh, hc, ho in data:
        o, _ = one.refine_batch_host(hm, hh, chunk_frames=2, kernel_frames=4)
        torch.cuda.current_stream().synchronize()
        want.append({k: o[k].clone() for k in keys})
    two = HotPath(RefinerConfig(), "cuda")

    def submit(i):
        hm, hh, hc, ho = data[i % 2]
        if route == "dense":
            two.refine_batch_host(hm, hh, chunk_frames=2, kernel_frames=4, slot=i % 2, async_result=True)
        else:
            two.refine_batch_rle(hc, ho, (B, K, H, W), hh, value=255, kernel_frames=4, slot=i % 2, async_result=True)
    submit(0)
    for i in range(1, 7):
        if i < 6:
            submit(i)                                       # call i in flight before call i-1's results are read
        o = two.wait_result((i - 1) % 2)
        HotPath.check_status(o)
        for k in keys:
            assert torch.equal(o[k], want[(i - 1) % 2][k]), (i - 1, k)
    assert two._stages[0]["masks"].data_ptr() != two._stages[1]["masks"].data_ptr()


def test_post_process_graph_capture_with_and_without_forked_error_maps():
    """quber_post_process forks its error-map pass onto a side stream (event fork / join).  Captured into a CUDA graph it
    must replay to the eager results both when the side stream already exists (this thread has called it before) and on
    a host thread's very first call, which happens under capture and therefore stays on one stream."""
    import threading
    B, K, H, W = 3, 6, 240, 320
    _, head = _inputs(B, K, H, W, seed=1234)
    dh = {k: torch.from_numpy(v).cuda() for k, v in head.items()}
    hp = HotPath(RefinerConfig(), "cuda")
    want = hp.post(dh)
    torch.cuda.synchronize()
    keys = ("label_map", "eee_onehot", "eee_full", "masks", "inst_offsets")
    result = {}

    def capture_and_replay(tag):
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                r = hp.post(dh)
            for _ in range(2):
                g.replay()
            s.synchronize()
            n_valid = int(want["inst_offsets"][-1])                  # rows of the instance pool past it are never written
            same = all(torch.equal(r[k], want[k]) for k in keys if k != "masks" and k in want and want[k] is not None)
            result[tag] = same and torch.equal(r["masks"][:n_valid], want["masks"][:n_valid])
    capture_and_replay("forked")                            # this thread's side stream exists since the eager call above
    t = threading.Thread(target=capture_and_replay, args=("first_call_under_capture",))
    t.start(); t.join()
    assert result == {"forked": True, "first_call_under_capture": True}


def test_refine_batch_bits_equals_dense_host_call():
    """The bit-plane route (the caller holds one bit per pixel: an eighth of the bytes over the link, no host packing)
    gives the dense call's results bit for bit, on repeated calls (graph replays, staging reuse)."""
    B, K, H, W = 5, 7, 240, 320
    masks, head = _inputs(B, K, H, W, seed=910)
    hm = torch.from_numpy(masks).pin_memory()
    hh = {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}
    words, value = ops.host_pack_masks(hm)
    assert value == 255
    hw = words.pin_memory()
    dense, bits = HotPath(RefinerConfig(), "cuda"), HotPath(RefinerConfig(), "cuda")
    o0, k0 = dense.refine_batch_host(hm, hh, chunk_frames=2)
    torch.cuda.current_stream().synchronize()
    o0 = {k: v.clone() for k, v in o0.items()}
    for rep in range(3):
        o1, k1 = bits.refine_batch_bits(hw, (B, K, H, W), hh, value=value, kernel_frames=2)
        torch.cuda.current_stream().synchronize()
        HotPath.check_status(o1)
        for key in ("label_map", "boundary", "inst_start", "inst_counts"):
            assert torch.equal(o0[key], o1[key]), key
        assert torch.equal(torch.cat([k[0] for k in k0]), torch.cat([k[0] for k in k1]))
        assert torch.equal(torch.cat([k[1] for k in k0]), torch.cat([k[1] for k in k1]))
    with pytest.raises(ValueError):
        bits.refine_batch_bits(hw[:-1], (B, K, H, W), hh)


def test_predict_many_equals_frame_by_frame_predict(tmp_path):
    """MaskRefiner.predict_many: host preparation on worker threads, frames of one size through the predictor as ONE batch,
    the caller-side filter per frame — the results are those of `predict` called frame by frame, in order, for files and
    for arrays, with frames of two sizes in one call."""
    import cv2
    from quber_b200.meta_arch import Instances
    from quber_b200.refiner_model import MaskRefiner, inpaint_depth
    rng = np.random.default_rng(11)
    n = 5
    rgb_paths, depth_paths, masks, fgs = [], [], [], []
    for i in range(n):
        rgb = rng.integers(0, 255, (480, 640, 3), dtype=np.uint8)          # OCID frames are 640 x 480 on disk
        depth = rng.integers(300, 1400, (480, 640)).astype(np.uint16)
        depth[rng.random((480, 640)) < 0.05] = 0
        cv2.imwrite(str(tmp_path / f"rgb{i}.png"), rgb)
        cv2.imwrite(str(tmp_path / f"depth{i}.png"), depth)
        rgb_paths.append(str(tmp_path / f"rgb{i}.png")); depth_paths.append(str(tmp_path / f"depth{i}.png"))
        m = synth.ellipse_masks(3 + i, 480, 640, seed=20 + i, axes=(15, 50), margin=40) != 0
        masks.append(m)
        fgs.append(m[: 2].any(0))
    calls = []

    class Stub:                                   # "refined" instances = the initial masks made disjoint, scaled by the mean depth sign
        def _one(self, rgb_img, dep, init):
            init = np.asarray(init) != 0
            taken = np.zeros(init.shape[1:], bool)
            out = []
            for m in init:
                m = m & ~taken
                taken |= m
                out.append(m)
            return {"instances": Instances(init.shape[1:], pred_masks=torch.from_numpy(np.stack(out)).cuda()),
                    "probe": (int(rgb_img.sum()) % 9973, None if dep is None else int(dep.sum()) % 9973)}

        def predict(self, rgb_img, dep, init):
            calls.append(1)
            return [self._one(rgb_img, dep, init)]

        def predict_batch(self, rgbs, deps, inits):
            calls.append(len(rgbs))
            return [self._one(r, d, m) for r, d, m in zip(rgbs, deps, inits)]

    r = MaskRefiner.__new__(MaskRefiner)
    r.refiner_predictor, r.dataset, r.fg_model, r.inpaint = Stub(), "OCID", None, inpaint_depth
    want = [r.predict(a, b, m, fg_mask=f) for a, b, m, f in zip(rgb_paths, depth_paths, masks, fgs)]
    calls.clear()
    got = r.predict_many(rgb_paths, depth_paths, masks, fgs, workers=4)
    assert calls == [n]                                                   # one batch
    assert len(got) == n
    for g, w in zip(got, want):
        assert np.array_equal(g[0], w[0]) and g[1]["probe"] == w[1]["probe"] and np.array_equal(g[3], w[3])
    # arrays of two sizes: two batches, results in the caller's order
    arrs = [rng.integers(0, 255, (120 if i % 2 else 96, 160, 3), dtype=np.uint8) for i in range(4)]
    deps = [rng.integers(300, 1400, a.shape[:2]).astype(np.uint16) for a in arrs]
    ms = [synth.ellipse_masks(2, a.shape[0], 160, seed=40 + i, axes=(6, 20), margin=10) != 0 for i, a in enumerate(arrs)]
    r.dataset, r.inpaint = "OSD", None
    calls.clear()
    got = r.predict_many(arrs, deps, ms, [m.any(0) for m in ms], workers=2)
    assert sorted(calls) == [2, 2]
    for i, g in enumerate(got):
        w = r.predict(arrs[i], deps[i], ms[i], fg_mask=ms[i].any(0))
        assert np.array_equal(g[0], w[0]) and g[1]["probe"] == w[1]["probe"]
