"""ctypes binding of libquber_b200.so (C ABI: include/quber_b200.h).

The product has no CPU fallback: if the shared library is missing this module raises
at import of the first op, telling the user how to build it.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libquber_b200.so")

ABI_VERSION = 3      # QUBER_ABI_VERSION of include/quber_b200.h
QUBER_OK, E_INVALID, E_WORKSPACE, E_CUDA, E_ALIGN, E_UNSUPPORTED = 0, -1, -2, -3, -4, -5
POST_GIVEN_CENTERS, POST_NO_TOPK, POST_LABEL_U8 = 1, 2, 4

c_vp, c_i, c_sz, c_f, c_d = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_float, ctypes.c_double


class PostParams(ctypes.Structure):
    """struct quber_post_params"""
    _fields_ = [("B", ctypes.c_int32), ("H", ctypes.c_int32), ("W", ctypes.c_int32), ("scale", ctypes.c_int32),
                ("offset_mul", c_f), ("eee_channels", ctypes.c_int32), ("threshold", c_f),
                ("nms_kernel", ctypes.c_int32), ("top_k", ctypes.c_int32), ("min_area", ctypes.c_int32),
                ("stuff_area", ctypes.c_int32), ("label_divisor", ctypes.c_int32), ("void_label", c_f),
                ("max_instances", ctypes.c_int32), ("flags", ctypes.c_int32),
                ("in_h", ctypes.c_int32), ("in_w", ctypes.c_int32), ("valid_w", ctypes.c_int32)]


class PostIO(ctypes.Structure):
    """struct quber_post_io"""
    _fields_ = [(n, c_vp) for n in (
        "foreground", "center", "offset", "eee", "sem_seg", "eee_full", "eee_onehot", "panoptic", "label_map",
        "ins_seg", "masks", "centers_yx", "n_centers", "inst_offsets", "labels", "scores", "boxes", "status")]


# every symbol include/quber_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "quber_version": (c_i, []),
    "quber_error_string": (ctypes.c_char_p, [c_i]),
    "quber_last_cuda_error": (ctypes.c_char_p, []),
    "quber_gaussian_table": (None, [c_i, c_vp]),
    "quber_encode_workspace_bytes": (c_sz, [c_i, c_i, c_i, c_i]),
    "quber_occupancy_bytes": (c_sz, [c_i, c_i, c_i, c_i]),
    "quber_encode_initial_masks": (c_i, [c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "quber_encode_static_scratch_bytes": (c_sz, []),
    "quber_encode_initial_masks_static": (c_i, [c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp, c_vp, c_vp]),
    "quber_encode_initial_masks_packed": (c_i, [c_vp, c_i, c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "quber_boundary_dilation": (c_i, [c_i, c_i, c_d]),
    "quber_mask_boundary": (c_i, [c_vp, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "quber_mask_boundary_bits": (c_i, [c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "quber_quadruple_maps": (c_i, [c_vp, c_vp, c_i, c_sz, c_vp, c_vp]),
    "quber_tma_timeouts": (c_i, []),
    "quber_gemm_bias_act": (c_i, [c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_vp]),
    "quber_gemm_timeouts": (c_i, []),
    "quber_split_tf32": (c_i, [c_vp, c_sz, c_vp, c_vp, c_vp]),
    "quber_gemm_bias_act_x3": (c_i, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_vp]),
    "quber_pack_input": (c_i, [c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp]),
    "quber_fuse_concat": (c_i, [c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp]),
    "quber_fuse_concat_rows": (c_i, [c_vp, c_vp, c_vp, c_vp, c_i, c_sz, c_i, c_i, c_vp, c_vp]),
    "quber_post_workspace_bytes": (c_sz, [ctypes.POINTER(PostParams)]),
    "quber_post_process": (c_i, [ctypes.POINTER(PostParams), ctypes.POINTER(PostIO), c_vp, c_sz, c_vp]),
    "quber_group_norm_workspace_bytes": (c_sz, [c_i, c_i, c_i, c_i, c_i]),
    "quber_group_norm_relu": (c_i, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, ctypes.c_float, c_i, c_i, c_i, c_vp, c_sz, c_vp]),
    "quber_upsample_concat": (c_i, [c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_vp]),
    "quber_filter_workspace_bytes": (c_sz, [c_i, c_i]),
    "quber_filter_instances": (c_i, [c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_d, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "quber_maxpool3x3s2": (c_i, [c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_vp]),
    "quber_add_layer_norm": (c_i, [c_vp, c_vp, c_vp, c_vp, c_sz, c_i, ctypes.c_float, c_vp, c_vp, c_vp, c_vp, c_i, c_vp]),
    "quber_host_pack_masks": (c_i, [c_vp, c_sz, c_vp, c_i, c_vp]),
    "quber_host_pack_begin": (c_i, [c_vp, c_sz, c_vp, c_i, c_vp, c_i]),
    "quber_host_pack_wait": (c_i, [c_i, c_vp]),
    "quber_unpack_masks": (c_i, [c_vp, c_sz, c_i, c_vp, c_vp]),
    "quber_msda_prepare": (c_i, [c_vp, c_i, c_vp, c_vp, c_sz, c_i, c_i, c_i, c_vp, c_vp, c_vp]),
    "quber_rle_decode_workspace_bytes": (c_sz, [c_sz]),
    "quber_rle_decode": (c_i, [c_vp, c_vp, c_sz, c_i, c_i, c_i, c_i, c_vp, c_vp, c_vp, c_sz, c_vp, c_vp]),
    "quber_rle_encode_workspace_bytes": (c_sz, [c_i, c_i, c_i, c_sz]),
    "quber_rle_encode": (c_i, [c_vp, c_i, c_i, c_i, c_vp, c_sz, c_vp, c_vp, c_sz, c_vp, c_vp]),
    "quber_rle_from_string": (ctypes.c_longlong, [ctypes.c_char_p, c_sz, c_vp, c_sz]),
    "quber_rle_to_string": (ctypes.c_longlong, [c_vp, c_sz, c_vp, c_sz]),
    "quber_eval_workspace_bytes": (c_sz, [c_i, c_i, c_i, c_i]),
    "quber_eval_counts": (c_i, [c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                c_sz, c_vp, c_vp]),
    "quber_upsample_crop_resize": (c_i, [c_vp, c_i, c_i, c_i, c_i, c_i, c_f, c_i, c_i, c_i, c_i, c_vp, c_vp]),
    "quber_resize_nearest_u8": (c_i, [c_vp, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp]),
    "quber_gemm_bias_act_cat2": (c_i, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_vp]),
    "quber_split_rgbd_input": (c_i, [c_vp, c_i, c_i, c_i, c_i, c_vp, c_vp, c_vp]),
    "quber_msda_forward_fused": (c_i, [c_vp, c_vp, c_vp, c_vp, c_i, c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp, c_vp]),
    "quber_msda_forward": (c_i, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_vp, c_vp]),
}

_lib = None


class QuberExtensionMissing(RuntimeError):
    pass


def lib():
    """The loaded shared library; raises QuberExtensionMissing (never falls back) if it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise QuberExtensionMissing(
                f"{SO_PATH} not found. quber_b200 has no CPU/PyTorch fallback: build the sm_100a library with "
                "`python -c 'import __graft_entry__ as g; g.build()'` (or `make -C quber_b200/csrc`).")
        l = ctypes.CDLL(SO_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)          # AttributeError if the .so does not export a declared symbol
            fn.restype, fn.argtypes = res, args
        if l.quber_version() != ABI_VERSION:
            raise QuberExtensionMissing(f"{SO_PATH}: ABI version {l.quber_version()} != {ABI_VERSION}, rebuild")
        _lib = l
    return _lib


def check(rc, what):
    """Map a QUBER_E_* code to the exception type the reference raises for the same condition."""
    if rc == QUBER_OK:
        return
    msg = f"{what}: {lib().quber_error_string(rc).decode()} (code {rc})"
    if rc == E_CUDA:
        raise RuntimeError(msg + " — " + lib().quber_last_cuda_error().decode())
    if rc == E_WORKSPACE:
        raise RuntimeError(msg)
    raise ValueError(msg)     # INVALID / ALIGN / UNSUPPORTED: the reference raises ValueError / AT_ASSERT
