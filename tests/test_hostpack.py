"""quber_host_pack_masks is a HOST function of the C-ABI library (csrc/hostpack.cpp): checked on CPU against numpy."""
import numpy as np
import pytest
import torch

from quber_b200 import ops


@pytest.mark.parametrize("threads", [1, 3, 8])
@pytest.mark.parametrize("shape", [(1, 1, 2, 16), (2, 5, 48, 64), (4, 30, 240, 320)])
def test_host_pack_matches_numpy_packbits(shape, threads):
    rng = np.random.default_rng(sum(shape) + threads)
    m = (rng.random(shape) < 0.3).astype(np.uint8)
    out, value = ops.host_pack_masks(torch.from_numpy(m), threads=threads)
    ref = np.packbits(m.reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1)
    assert value == 1
    out255, value255 = ops.host_pack_masks(torch.from_numpy(m * 255), threads=threads)     # eval/refiner_model.py:230-231
    assert value255 == 255
    np.testing.assert_array_equal(out255.numpy().view(np.uint32), ref)
    np.testing.assert_array_equal(out.numpy().view(np.uint32), ref)
    # the worker pool is reused: a second buffer of another size right after
    m2 = np.ascontiguousarray(m.reshape(-1)[: 32 * max(1, m.size // 64)])
    out2, _ = ops.host_pack_masks(torch.from_numpy(m2), threads=threads)
    np.testing.assert_array_equal(out2.numpy().view(np.uint32), np.packbits(m2.reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1))


def test_host_pack_reports_mixed_values_and_bad_sizes():
    m = np.zeros((2, 3, 32, 64), np.uint8)
    assert ops.host_pack_masks(torch.from_numpy(m), threads=4)[1] == 0  # all-zero buffer
    m[1, 2, 7, 9] = 255
    assert ops.host_pack_masks(torch.from_numpy(m), threads=4)[1] == 255
    m[0, 0, 0, 0] = 1
    out, value = ops.host_pack_masks(torch.from_numpy(m), threads=4)
    assert value == -1                                                  # two different non-zero values: send the bytes
    ref = np.packbits((m != 0).reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1)
    np.testing.assert_array_equal(out.numpy().view(np.uint32), ref)     # the words are still "nonzero" bits
    with pytest.raises(ValueError):
        ops.host_pack_masks(torch.zeros(33, dtype=torch.uint8))
    with pytest.raises(ValueError):
        ops.host_pack_masks(torch.zeros(64, dtype=torch.float32))
    with pytest.raises(ValueError):
        ops.host_pack_masks(torch.zeros(64, dtype=torch.uint8), out=torch.zeros(1, dtype=torch.int32))


@pytest.mark.parametrize("threads", [1, 4, 12])
def test_host_pack_chunked_async_matches_numpy_and_reports_values_per_chunk(threads):
    """quber_host_pack_begin / _wait: chunks finish in order, each with its own value; jobs back to back reuse the pool."""
    rng = np.random.default_rng(threads)
    frames, fb = 11, 3 * 48 * 64
    m = (rng.random((frames, fb)) < 0.3).astype(np.uint8) * 255
    m[4] = (m[4] != 0)                       # chunk 2 (frames 4..5) carries 1 and 255 -> -1
    m[8:10] = 0                              # chunk 4 is empty -> 0
    sched = [(0, 2), (2, 4), (4, 6), (6, 8), (8, 10), (10, 11)]
    want = [255, 255, -1, 255, 0, 255]
    t = torch.from_numpy(m)
    ref = np.packbits((m != 0).reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1)
    for rep in range(3):
        out = torch.zeros(m.size // 32, dtype=torch.int32)
        ops.host_pack_begin(t, out, threads, [c1 * fb for _, c1 in sched])
        got = []
        for ci, (c0, c1) in enumerate(sched):
            got.append(ops.host_pack_wait(ci))
            w0, w1 = c0 * fb // 32, c1 * fb // 32
            np.testing.assert_array_equal(out.numpy().view(np.uint32)[w0:w1], ref[w0:w1])    # complete once waited for
        assert got == want
    big = torch.from_numpy((rng.random(8 * 1024 * 1024) < 0.5).astype(np.uint8))             # many 256 KB blocks
    out = torch.zeros(big.numel() // 32, dtype=torch.int32)
    ops.host_pack_begin(big, out, threads, [big.numel() // 4, big.numel()])
    assert ops.host_pack_wait(1) == 1 and ops.host_pack_wait(0) == 1
    np.testing.assert_array_equal(out.numpy().view(np.uint32),
                                  np.packbits(big.numpy().reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1))
    with pytest.raises(ValueError):
        ops.host_pack_begin(big, out, threads, [64, 32, big.numel()])                        # not ascending
