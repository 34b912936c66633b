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
