"""CPU-only: the C-ABI library loads and exports every symbol include/quber_b200.h declares; the
ctypes structs mirror the header; host-only entry points (no GPU needed) agree with the oracle."""
import ctypes
import os
import re

import numpy as np
import pytest

import oracle
from quber_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header():
    return open(os.path.join(ROOT, "include", "quber_b200.h")).read()


def test_library_present_and_versioned():
    assert os.path.exists(_lib.SO_PATH), "build it: python -c 'import __graft_entry__ as g; g.build()'"
    assert _lib.lib().quber_version() == _lib.ABI_VERSION == 3


def test_every_declared_symbol_is_exported_and_bound():
    h = re.sub(r"/\*.*?\*/", "", _header(), flags=re.S)
    declared = set(re.findall(r"\b(quber_[a-z_0-9]+)\s*\(", h))
    assert declared, "no declarations parsed"
    raw = ctypes.CDLL(_lib.SO_PATH)
    for name in declared:
        assert hasattr(raw, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)


def test_struct_layout_matches_header():
    h = _header()
    body = re.search(r"typedef struct quber_post_params \{(.*?)\} quber_post_params;", h, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if decl:
            names += [n.strip() for n in decl.split(None, 1)[1].split(",")]
    assert names == [f[0] for f in _lib.PostParams._fields_]
    body = re.search(r"typedef struct quber_post_io \{(.*?)\} quber_post_io;", h, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = [re.findall(r"\*\s*([a-z_]+)$", d.strip())[0] for d in body.split(";") if d.strip()]
    assert names == [f[0] for f in _lib.PostIO._fields_]


def test_error_strings_and_exception_mapping():
    L = _lib.lib()
    assert L.quber_error_string(0) == b"ok"
    for code in (-1, -2, -3, -4, -5):
        assert L.quber_error_string(code)
    with pytest.raises(ValueError):
        _lib.check(_lib.E_INVALID, "x")
    with pytest.raises(ValueError):
        _lib.check(_lib.E_ALIGN, "x")
    with pytest.raises(RuntimeError):
        _lib.check(_lib.E_WORKSPACE, "x")


def test_argument_validation_without_gpu():
    """Argument checks run before any CUDA call, so they are testable on a CPU-only host."""
    L = _lib.lib()
    assert L.quber_encode_initial_masks(None, 1, 1, 16, 16, 0, 10, None, None, None, None, 0, None) == _lib.E_INVALID
    assert L.quber_mask_boundary(None, 1, 1, 16, 16, 8, None, None, None, None, None, None) == _lib.E_INVALID
    assert L.quber_mask_boundary(ctypes.c_void_p(16), 1, 1, 16, 24, 8, ctypes.c_void_p(16), None, None, None, None, None) == _lib.E_ALIGN
    assert L.quber_mask_boundary(ctypes.c_void_p(16), 1, 1, 16, 16, 0, ctypes.c_void_p(16), None, None, None, None, None) == _lib.E_INVALID
    assert L.quber_msda_forward(None, None, None, None, None, 1, 1, 1, 1, 1, 1, 1, 0, None, None) == _lib.E_INVALID
    v16 = ctypes.c_void_p(16)
    assert L.quber_gemm_bias_act(None, v16, None, v16, 128, 128, 64, 64, 64, 128, 1, 0, 0, None) == _lib.E_INVALID
    assert L.quber_gemm_bias_act(v16, v16, None, v16, 128, 128, 64, 64, 64, 128, 3, 0, 0, None) == _lib.E_INVALID     # dtype
    assert L.quber_gemm_bias_act(v16, v16, None, v16, 128, 128, 64, 32, 64, 128, 1, 0, 0, None) == _lib.E_INVALID     # lda < K
    assert L.quber_gemm_bias_act(v16, v16, None, v16, 128, 128, 60, 60, 60, 128, 1, 0, 0, None) == _lib.E_ALIGN       # 120-byte pitch
    p = _lib.PostParams(1, 480, 640, 4, 4.0, 4, 0.3, 8, 200, 512, 2048, 1000, -1.0, 8, 0)    # even nms kernel
    assert L.quber_post_workspace_bytes(ctypes.byref(p)) == 0
    p.nms_kernel = 7
    assert L.quber_post_workspace_bytes(ctypes.byref(p)) > 480 * 640 * 2
    assert L.quber_encode_workspace_bytes(2, 30, 480, 640) >= 2 * 480 * 640


def test_host_helpers_match_oracle():
    L = _lib.lib()
    for (H, W, r) in [(480, 640, 0.01), (480, 640, 0.02), (750, 1333, 0.01), (10, 10, 0.01), (1080, 1920, 0.02)]:
        assert L.quber_boundary_dilation(H, W, r) == oracle.boundary_dilation(H, W, r)
    for sigma in (8, 10):
        n = 6 * sigma + 3
        t = np.empty((n, n), np.float32)
        L.quber_gaussian_table(sigma, t.ctypes.data_as(ctypes.c_void_p))
        np.testing.assert_array_equal(t, oracle.gaussian_table(sigma))


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under quber_b200/ may reference it."""
    pkg = os.path.join(ROOT, "quber_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, fn)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", src, re.M), f"{fn} imports oracle"
                assert "libquber_oracle" not in src, fn


def test_ops_fail_loudly_without_cuda():
    import torch
    from quber_b200 import ops
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(ValueError, match="CUDA tensor"):
        ops.encode_initial_masks(torch.zeros(1, 1, 16, 16, dtype=torch.uint8))
    with pytest.raises(ValueError, match="CUDA tensor"):
        ops.msda_forward(torch.zeros(1, 4, 1, 8), torch.zeros(1, 2, dtype=torch.long), torch.zeros(1, dtype=torch.long),
                         torch.zeros(1, 1, 1, 1, 1, 2), torch.zeros(1, 1, 1, 1, 1))
    from quber_b200 import msda
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):     # OPS/src/ms_deform_attn.h:43
        msda.ms_deform_attn_forward(torch.zeros(1, 4, 1, 8), torch.zeros(1, 2, dtype=torch.long),
                                    torch.zeros(1, dtype=torch.long), torch.zeros(1, 1, 1, 1, 1, 2),
                                    torch.zeros(1, 1, 1, 1, 1), 64)


def test_every_device_entry_point_is_a_custom_op_with_a_fake_kernel():
    """north_star: "registered as PyTorch custom ops".  Fake-tensor propagation (what torch.compile / export run) works
    without a GPU: fake CUDA tensors through torch.ops.quber.* give the shapes / dtypes the kernels produce."""
    import torch
    from torch._subclasses.fake_tensor import FakeTensorMode
    from quber_b200 import ops
    names = {n for n in dir(torch.ops.quber) if not n.startswith("_") and n != "name"}
    want = {"msda_forward", "msda_forward_fused", "msda_prepare", "encode_initial_masks", "encode_initial_masks_packed",
            "unpack_masks", "mask_boundary", "quadruple_maps", "post_process", "gemm_bias_act", "gemm_bias_act_x3", "split_tf32",
            "pack_input", "fuse_concat", "group_norm_relu", "upsample_concat", "maxpool3x3s2", "add_layer_norm",
            "filter_instances", "rle_decode", "rle_encode"}
    assert want <= names, want - names
    q = torch.ops.quber
    with FakeTensorMode():
        e = lambda *s, dt=torch.float32: torch.empty(*s, dtype=dt, device="cuda")
        fg, ce, of, ee = e(2, 1, 60, 80), e(2, 1, 60, 80), e(2, 2, 60, 80), e(2, 3, 60, 80)
        out = dict(zip(ops.POST_OUTPUTS, q.post_process(fg, ce, of, ee, 4, 4.0, 0.3, 7, 200, 512, 2048, 1000, -1.0, 400, True)))
        assert tuple(out["sem_seg"].shape) == (2, 1, 240, 320) and tuple(out["eee_onehot"].shape) == (2, 3, 240, 320)
        assert out["label_map"].dtype == torch.uint8 and tuple(out["masks"].shape) == (400, 240, 320)
        assert tuple(out["inst_offsets"].shape) == (3,) and out["eee_onehot"].dtype == torch.uint8
        out = q.post_process(fg, ce, of, None, 4, 4.0, 0.3, 7, 200, 512, 2048, 1000, -1.0, 400, False)
        assert tuple(out[1].shape) == (2, 0, 240, 320) and out[4].dtype == torch.int32
        assert q.gemm_bias_act(e(3, 100, 256), e(192, 256), None, True, True).shape == (3, 100, 192)
        assert q.gemm_bias_act_x3(e(100, 256), e(192, 256), e(192, 256), e(192), False, False).dtype == torch.float32
        v = e(2, 500, 8, 32, dt=torch.bfloat16)
        sh, st = e(3, 2, dt=torch.int64), e(3, dt=torch.int64)
        assert q.msda_forward_fused(v, sh, st, e(2, 77, 288), e(2, 77, 3, 2), 3, 4).shape == (2, 77, 256)
        loc, aw = q.msda_prepare(e(2, 77, 288), e(2, 77, 3, 2), sh, 8, 3, 4)
        assert loc.shape == (2, 77, 8, 3, 4, 2) and aw.shape == (2, 77, 8, 3, 4)
        assert q.msda_forward(v, sh, st, loc, aw).dtype == torch.bfloat16
        m = e(2, 5, 64, 96, dt=torch.uint8)
        per, uni, fgm = q.mask_boundary(m, 8)
        assert per.shape == m.shape and uni.shape == (2, 64, 96) and fgm.dtype == torch.uint8
        assert q.encode_initial_masks(m, 10).shape == (2, 3, 64, 96)
        pk, stt = q.rle_decode(e(100, dt=torch.int32), e(11, dt=torch.int64), 10, 64, 96, 255, True)
        assert pk.shape == (10 * 64 * 96 // 32,) and pk.dtype == torch.int32 and stt.shape == (1,)
        c, o, s2 = q.rle_encode(m[0], 1000)
        assert c.shape == (1000,) and o.shape == (6,) and o.dtype == torch.int64
        x = e(2, 64, 8, 12).to(memory_format=torch.channels_last)
        assert q.maxpool3x3s2(x).shape == (2, 64, 4, 6)
        assert q.upsample_concat(e(2, 32, 16, 24).to(memory_format=torch.channels_last), x).shape == (2, 96, 16, 24)
        assert q.fuse_concat([e(2, 8, 6, 6), e(2, 4, 6, 6)], [0, 2], 2, 16).shape == (72, 16)
        assert q.group_norm_relu(x, e(64), e(64), 32, 1e-5, True, None).shape == x.shape
        assert q.add_layer_norm(e(10, 256), e(10, 256), e(256), e(256), 1e-5).shape == (10, 256)
        assert q.pack_input(e(2, 6, 16, 16, dt=torch.uint8), e(2, 3, 16, 16), e(6), e(6)).shape == (2, 9, 16, 16)
        lm, keep, nc = q.filter_instances(e(2, 16, 16, dt=torch.int32), e(2, dt=torch.int32), None, None, 0.3, 64)
        assert lm.shape == (2, 16, 16) and keep.shape == (2, 64) and nc.shape == (2,)
