"""quber_encode_initial_masks_static: the encode entry point that allocates nothing (its scratch is a static device array
inside the library).  No counterpart in the reference — added for callers that
cannot call cudaMalloc (graph capture, serving loops).  Collected last on purpose (file name):
a failure here must not hide the parity tests."""
import ctypes

import numpy as np
import pytest
import torch

from quber_b200 import _lib, ops, synth


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def test_static_scratch_limit_is_checked_before_any_cuda_call():
    """Sizes beyond the static array are refused with QUBER_E_WORKSPACE and a message naming the sizes — on a CPU-only
    host too (the check precedes every CUDA call)."""
    L = _lib.lib()
    cap = L.quber_encode_static_scratch_bytes()
    assert cap == 16 << 20
    assert L.quber_encode_workspace_bytes(32, 30, 480, 640) <= cap          # the C2 batch fits
    B, K, H, W = 64, 30, 480, 640
    assert L.quber_encode_workspace_bytes(B, K, H, W) > cap
    one = ctypes.c_void_p(256)
    rc = L.quber_encode_initial_masks_static(one, B, K, H, W, 0, 10, one, one, None, None)
    assert rc == _lib.E_WORKSPACE
    msg = L.quber_last_cuda_error().decode()
    assert "quber_encode_initial_masks_static" in msg and str(cap) in msg
    assert L.quber_encode_initial_masks_static(one, 0, K, H, W, 0, 10, one, one, None, None) == _lib.E_INVALID


@pytest.mark.gpu
def test_encode_static_scratch_equals_workspace_entry_point_twice_in_a_row():
    """Through the C ABI: the allocation-free entry point against quber_encode_initial_masks on the same inputs — bit-exact
    (the kernels are the same, integer sums and float64 centroids), heat-map / offsets and the tile occupancy — on two
    calls in a row with different masks, shapes and mask counts, so that the array's left-over contents (sums of the
    previous call, a larger last-mask map) would show if they were not cleared / rewritten."""
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    g = ops.gaussian_table(10, dev)
    for call, (B, K, H, W, seed) in enumerate([(3, 12, 240, 320, 5), (2, 5, 136, 272, 6)]):
        m = np.stack([synth.ellipse_masks(K, H, W, seed=seed + b, axes=(8, 50), margin=5, value=255) for b in range(B)])
        m[0, K - 1] = 0                                                     # an empty mask: skipped by the reference
        masks = torch.from_numpy(m).to(dev)
        want, want_occ = ops.encode_initial_masks(masks, 10, with_occupancy=True)      # the existing entry point
        out = torch.full((B, 3, H, W), float("nan"), dtype=torch.float32, device=dev)
        occ = torch.full(tuple(want_occ.shape), 7, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize(dev)
        before = torch.cuda.memory_allocated(dev)
        with torch.cuda.device(dev):
            rc = L.quber_encode_initial_masks_static(_ptr(masks), B, K, H, W, 0, 10, _ptr(g), _ptr(out), _ptr(occ), stream)
        assert rc == _lib.QUBER_OK, (rc, L.quber_last_cuda_error().decode())
        torch.cuda.synchronize(dev)
        assert torch.cuda.memory_allocated(dev) == before
        assert torch.equal(out, want), f"call {call}: max abs diff {float((out - want).abs().max())}"
        assert torch.equal(occ, want_occ)
        # and without the optional occupancy output
        out2 = torch.empty_like(out)
        with torch.cuda.device(dev):
            rc = L.quber_encode_initial_masks_static(_ptr(masks), B, K, H, W, 0, 10, _ptr(g), _ptr(out2), None, stream)
        assert rc == _lib.QUBER_OK
        torch.cuda.synchronize(dev)
        assert torch.equal(out2, want)
