"""Drop-in for the vendored MSDeformAttn op (OPS = ext_modules/UnseenObjectsWithMeanShift/MSMFormer/
meanshiftformer/modeling/pixel_decoder/ops): `MSDeformAttnFunction` (OPS/functions/ms_deform_attn_func.py:32-49),
the `MultiScaleDeformableAttention` extension module surface (OPS/src/vision.cpp:18-21) and the
`MSDeformAttn` nn.Module (OPS/modules/ms_deform_attn.py:34-125).  Forward only: training is out of scope.

The module keeps the upstream constructor signature, the four `nn.Linear` attribute names and the initialisation
formula, because the drop-in contract is "a reference state_dict loads unchanged and a fresh module initialises
identically".  Those parts derive from Deformable DETR (Copyright (c) 2020 SenseTime, Apache License 2.0) as vendored and
modified by Mask2Former (Copyright (c) Facebook, Inc. and its affiliates) under the path above; see the upstream files
for the full notices.  `forward` is NOT the upstream torch glue: every step runs on this repo's kernels —

  value_proj                               one tcgen05 GEMM (quber_gemm_bias_act / _x3)
  sampling_offsets + attention_weights     ONE tcgen05 GEMM on the concatenated [M*L*P*3, C] weight (both read `query`)
  softmax over L*P, ref + off / (W_l, H_l), bilinear sampling      quber_msda_forward_fused (set-up phase + gather)
  output_proj                              one tcgen05 GEMM

`precision`: "fp32" (default — the reference's own, split-operand 3xTF32 GEMMs + fp32 sampling), "tf32", "bf16"
(bf16 GEMM operands and bf16 `value`, fp32 weights / accumulation in the sampler).
"""
import math
import sys
import types
import warnings

import torch
from torch import nn
from torch.autograd import Function

from . import ops
from .linear import PRECISIONS, PackedLinear


def ms_deform_attn_forward(value, spatial_shapes, level_start_index, sampling_loc, attn_weight, im2col_step):
    """OPS/src/ms_deform_attn.h:25-46.  CPU tensors raise like the reference's AT_ERROR("Not implemented on the CPU");
    float64 runs in float64 as the reference dispatches it (ms_deform_attn_cuda.cu:69)."""
    if not value.is_cuda:
        raise RuntimeError("Not implemented on the CPU")
    aux = torch.float64 if value.dtype == torch.float64 else torch.float32
    return ops.msda_forward(value.contiguous(), spatial_shapes.contiguous(), level_start_index.contiguous(),
                            sampling_loc.to(aux).contiguous(), attn_weight.to(aux).contiguous(), im2col_step)


def ms_deform_attn_backward(*args, **kwargs):
    raise NotImplementedError("quber_b200 implements the inference path only (MSDeformAttn backward is out of scope)")


class MSDeformAttnFunction(Function):
    @staticmethod
    def forward(ctx, value, value_spatial_shapes, value_level_start_index, sampling_locations, attention_weights, im2col_step):
        ctx.im2col_step = im2col_step
        return ms_deform_attn_forward(value, value_spatial_shapes, value_level_start_index, sampling_locations,
                                      attention_weights, im2col_step)

    @staticmethod
    def backward(ctx, grad_output):
        ms_deform_attn_backward()


def install_extension_module(name="MultiScaleDeformableAttention"):
    """Puts a module with the compiled extension's name and surface (OPS/src/vision.cpp:18-21, OPS/setup.py:60) into
    sys.modules, so that the reference's `import MultiScaleDeformableAttention as MSDA`
    (OPS/functions/ms_deform_attn_func.py:21) resolves to these kernels.  An already importable module is left alone."""
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    m.__doc__ = "quber_b200 stand-in for the MultiScaleDeformableAttention CUDA extension (forward only)"
    m.ms_deform_attn_forward = ms_deform_attn_forward
    m.ms_deform_attn_backward = ms_deform_attn_backward
    sys.modules[name] = m
    return m


def _power_of_two(n):
    if not isinstance(n, int) or n < 0:
        raise ValueError("invalid input for _is_power_of_2: {} (type: {})".format(n, type(n)))
    return n != 0 and n & (n - 1) == 0


_is_power_of_2 = _power_of_two        # the upstream helper's name


class MSDeformAttn(nn.Module):
    def __init__(self, d_model=256, n_levels=4, n_heads=8, n_points=4):
        super().__init__()
        if d_model % n_heads != 0:
            raise ValueError('d_model must be divisible by n_heads, but got {} and {}'.format(d_model, n_heads))
        if not _power_of_two(d_model // n_heads):
            warnings.warn("You'd better set d_model in MSDeformAttn to make the dimension of each attention head a power of 2 "
                          "which is more efficient in our CUDA implementation.")
        self.im2col_step = 128
        self.d_model, self.n_levels, self.n_heads, self.n_points = d_model, n_levels, n_heads, n_points
        # creation order = the upstream's: a seeded fresh module draws the same random numbers
        self.sampling_offsets = nn.Linear(d_model, n_heads * n_levels * n_points * 2)
        self.attention_weights = nn.Linear(d_model, n_heads * n_levels * n_points)
        self.value_proj = nn.Linear(d_model, d_model)
        self.output_proj = nn.Linear(d_model, d_model)
        self.precision = "fp32"
        self._reset_parameters()

    def _reset_parameters(self):
        """The initial state of OPS/modules/ms_deform_attn.py:66-80: zero offset / attention weights, the offset bias a
        ring of unit steps per head scaled by the point index, Xavier-uniform value / output projections."""
        M, L, P = self.n_heads, self.n_levels, self.n_points
        with torch.no_grad():
            ang = torch.arange(M, dtype=torch.float32) * (2.0 * math.pi / M)
            ring = torch.stack([ang.cos(), ang.sin()], -1)
            ring = ring / ring.abs().max(-1, keepdim=True)[0]
            bias = ring.view(M, 1, 1, 2).repeat(1, L, P, 1)
            for i in range(P):
                bias[:, :, i, :] *= i + 1
            self.sampling_offsets.weight.zero_()
            self.sampling_offsets.bias = nn.Parameter(bias.view(-1))
            self.attention_weights.weight.zero_()
            self.attention_weights.bias.zero_()
            nn.init.xavier_uniform_(self.value_proj.weight)
            self.value_proj.bias.zero_()
            nn.init.xavier_uniform_(self.output_proj.weight)
            self.output_proj.bias.zero_()

    # ---- kernels ------------------------------------------------------------------------------------------
    def _packs(self):
        p = self.__dict__.get("_packed")
        if p is None:
            p = dict(v=PackedLinear(self.value_proj), qk=PackedLinear(self.sampling_offsets, self.attention_weights),
                     o=PackedLinear(self.output_proj))
            self.__dict__["_packed"] = p
        return p

    def sample(self, value, oa, reference_points, spatial_shapes, level_start_index, levels=None):
        """value [N,S,M,D], oa [N,Q,M*L*P*3] (offsets | logits, the qk GEMM's rows), reference_points [N,Q,L,2|4]
        -> [N,Q,M*D]: ms_deform_attn.py:102-117 + the sampling op."""
        N, Q, _ = oa.shape
        M, L, P = self.n_heads, self.n_levels, self.n_points
        rp = reference_points
        if rp.shape[-1] == 2 and rp.dtype == torch.float32 and ops.msda_fused_supported(value, L, P):
            # levels: the host copy of spatial_shapes when the caller has one (the pixel decoder builds the tensor from it)
            return ops.msda_forward_fused(value, spatial_shapes, level_start_index, oa, rp.contiguous(), L, P, levels=levels)
        if rp.shape[-1] == 2 and rp.dtype == torch.float32:
            loc, aw = ops.msda_prepare(oa, rp.contiguous(), spatial_shapes, M, L, P)
        elif rp.shape[-1] in (2, 4):
            # reference boxes (:110-112; never produced by the pixel decoder) and non-fp32 points: elementwise torch
            n_off = M * L * P * 2
            off = oa[..., :n_off].reshape(N, Q, M, L, P, 2)
            aw = torch.softmax(oa[..., n_off:].reshape(N, Q, M, L * P), -1).view(N, Q, M, L, P).contiguous()
            if rp.shape[-1] == 2:
                wh = torch.stack([spatial_shapes[..., 1], spatial_shapes[..., 0]], -1)
                loc = rp[:, :, None, :, None, :] + off / wh[None, None, None, :, None, :]
            else:
                loc = rp[:, :, None, :, None, :2] + off / P * rp[:, :, None, :, None, 2:] * 0.5
            loc = loc.float().contiguous()
        else:
            raise ValueError('Last dim of reference_points must be 2 or 4, but get {} instead.'.format(rp.shape[-1]))
        return ops.msda_forward(value, spatial_shapes, level_start_index, loc, aw)

    def forward(self, query, reference_points, input_flatten, input_spatial_shapes, input_level_start_index,
                input_padding_mask=None):
        if not query.is_cuda:
            raise RuntimeError("Not implemented on the CPU")         # OPS/src/ms_deform_attn.h:43 (no silent CPU fallback)
        if self.precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {PRECISIONS}")
        N, Len_q, _ = query.shape
        _, Len_in, _ = input_flatten.shape
        if int((input_spatial_shapes[:, 0] * input_spatial_shapes[:, 1]).sum()) != Len_in:
            raise AssertionError("input_spatial_shapes do not add up to the flattened input length")     # :96
        pk, prec = self._packs(), self.precision
        vdt = torch.bfloat16 if prec == "bf16" else torch.float32
        value = pk["v"](input_flatten, prec, out_dtype=vdt)                                             # :98
        if input_padding_mask is not None:
            value = value.masked_fill(input_padding_mask[..., None], float(0))                           # :99-100
        value = value.view(N, Len_in, self.n_heads, self.d_model // self.n_heads)
        oa = pk["qk"](query, prec, out_dtype=torch.float32)                                              # :102-103
        samp = self.sample(value, oa, reference_points, input_spatial_shapes, input_level_start_index)
        return pk["o"](samp, prec, out_dtype=torch.float32)                                              # :124
