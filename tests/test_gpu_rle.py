"""COCO RLE on the device against the oracle and tests/golden/rle.npz, through the C ABI.  B200 only."""
import numpy as np
import pytest
import torch

import oracle
from quber_b200 import ops, rle, synth
from test_oracle_golden import RLE_CASES, rle_case

pytestmark = pytest.mark.gpu
DEV = "cuda"


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.mark.parametrize("case", RLE_CASES)
def test_rle_decode_golden(case):
    masks, counts, offsets, _ = rle_case(case)
    n, H, W = masks.shape
    c, o = cu(counts.view(np.int32)), cu(offsets)
    out, status = ops.rle_decode(c, o, n, H, W, value=255)
    assert int(status[0]) == 0
    np.testing.assert_array_equal(out.cpu().numpy(), masks * 255)
    for i in range(n):                                            # and the oracle, mask by mask
        np.testing.assert_array_equal(oracle.rle_decode(counts[offsets[i]:offsets[i + 1]], H, W), masks[i])
    if W % 32 == 0:
        pk, status = ops.rle_decode(c, o, n, H, W, packed=True)
        ref, _ = ops.host_pack_masks(torch.from_numpy(masks).contiguous())
        np.testing.assert_array_equal(pk.cpu().numpy(), ref.numpy())
    else:
        with pytest.raises(ValueError):
            ops.rle_decode(c, o, n, H, W, packed=True)


@pytest.mark.parametrize("case", RLE_CASES)
def test_rle_encode_golden(case):
    masks, counts, offsets, strs = rle_case(case)
    c, o, status = ops.rle_encode(cu(masks * 7), cap=int(offsets[-1]))
    assert int(status[0]) == 0
    np.testing.assert_array_equal(o.cpu().numpy(), offsets)
    np.testing.assert_array_equal(c.cpu().numpy().view(np.uint32), counts)
    # capacity too small: flagged, offsets still complete
    c2, o2, st2 = ops.rle_encode(cu(masks), cap=max(int(offsets[-1]) - 3, 1))
    assert int(st2[0]) & 2 and int(o2[-1]) == int(offsets[-1])
    # the pycocotools-shaped module: compressed strings
    enc = rle.encode_from_device(cu(masks))
    assert [e["counts"] for e in enc] == strs and enc[0]["size"] == list(masks.shape[1:])


def test_rle_module_mirrors_pycocotools_api():
    m = synth.ellipse_masks(5, 96, 128, seed=9, axes=(6, 30), margin=5, empty_tail=1)
    rles = rle.encode(np.asfortranarray((m != 0).astype(np.uint8).transpose(1, 2, 0)))     # [h,w,n] as pycocotools takes it
    assert len(rles) == 6 and all(isinstance(r["counts"], bytes) for r in rles)
    back = rle.decode(rles)
    assert back.shape == (96, 128, 6) and back.dtype == np.uint8 and back.flags.f_contiguous
    np.testing.assert_array_equal(back.transpose(2, 0, 1), (m != 0).astype(np.uint8))
    one = rle.decode(rles[2])
    np.testing.assert_array_equal(one, (m[2] != 0).astype(np.uint8))
    np.testing.assert_array_equal(rle.area(rles), (m != 0).reshape(6, -1).sum(1))
    # frPyObjects on an uncompressed RLE dict (register_uoais_sim_instance.py:207-209)
    unc = {"size": [96, 128], "counts": oracle.rle_encode(m[1]).tolist()}
    assert rle.frPyObjects(unc, 96, 128)["counts"] == rles[1]["counts"]
    # str counts (what `rle['counts'].decode('ascii')` leaves in the reference's JSON files, panoptic2eee.py:63)
    js = [{"size": r["size"], "counts": r["counts"].decode("ascii")} for r in rles]
    np.testing.assert_array_equal(rle.decode(js), back)
    with pytest.raises(NotImplementedError):
        rle.frPyObjects([[0.0, 0.0, 4.0, 0.0, 4.0, 4.0]], 96, 128)


def test_rle_bad_counts_are_flagged_not_trusted():
    H, W = 64, 96
    good = oracle.rle_encode(synth.ellipse_masks(1, H, W, seed=3, axes=(5, 20), margin=5)[0])
    short = good[:-1].copy()                                        # runs end before the last pixel
    long_ = good.copy(); long_[-1] += 50                            # runs overrun the mask
    for bad in (short, long_):
        out, status = ops.rle_decode(cu(bad.view(np.int32)), cu(np.array([0, len(bad)], np.int64)), 1, H, W)
        assert int(status[0]) & 1
        ref = oracle.rle_decode(good, H, W)
        n_ok = int(bad[: len(good) - 1].sum())                      # pixels covered by the shared runs (column-major)
        got_cm = out[0].cpu().numpy().T.reshape(-1)
        np.testing.assert_array_equal(got_cm[:n_ok], ref.T.reshape(-1)[:n_ok])
    with pytest.raises(ValueError):
        ops.rle_decode(cu(short.view(np.int32)), cu(np.array([0, len(short)], np.int64)), 1, H, W, check=True)


def test_rle_to_encode_and_boundary_c2_shapes():
    """The end-to-end RLE route at configs[1] shapes: RLE -> packed words -> encode (+ byte planes) -> boundaries equals
    the dense-byte route bit for bit; 4 frames of 30 masks."""
    B, K, H, W = 4, 30, 480, 640
    masks = np.stack([synth.ellipse_masks(K, H, W, seed=500 + b) for b in range(B)])
    counts, offs = [], [0]
    for m in masks.reshape(-1, H, W):
        c = oracle.rle_encode(m)
        counts.append(c)
        offs.append(offs[-1] + len(c))
    counts = np.concatenate(counts)
    assert counts.nbytes < masks.nbytes // 200                      # the point of the format
    pk, st = ops.rle_decode(cu(counts.view(np.int32)), cu(np.array(offs, np.int64)), B * K, H, W, packed=True)
    assert int(st[0]) == 0
    planes = torch.empty((B, K, H, W), dtype=torch.uint8, device=DEV)
    enc = ops.encode_initial_masks_packed(pk, 255, planes)
    dm = cu(masks)
    assert torch.equal(planes, dm)
    assert torch.equal(enc, ops.encode_initial_masks(dm))
    d = ops.boundary_dilation(H, W, 0.01)
    for a, b in zip(ops.mask_boundary(planes, d), ops.mask_boundary(dm, d)):
        assert torch.equal(a, b)
