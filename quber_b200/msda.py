"""Drop-in for the vendored MSDeformAttn op (OPS = ext_modules/UnseenObjectsWithMeanShift/MSMFormer/
meanshiftformer/modeling/pixel_decoder/ops): `MSDeformAttnFunction` (OPS/functions/ms_deform_attn_func.py:32-49),
the `MultiScaleDeformableAttention` extension module surface (OPS/src/vision.cpp:18-21) and the
`MSDeformAttn` nn.Module (OPS/modules/ms_deform_attn.py:34-125) with identical parameter names, so a
reference state_dict loads unchanged.  Forward only: training is out of scope.
"""
import math
import warnings

import torch
import torch.nn.functional as F
from torch import nn
from torch.autograd import Function
from torch.nn.init import constant_, xavier_uniform_

from . import ops


def ms_deform_attn_forward(value, spatial_shapes, level_start_index, sampling_loc, attn_weight, im2col_step):
    """OPS/src/ms_deform_attn.h:25-46.  CPU tensors raise like the reference's AT_ERROR("Not implemented on the CPU")."""
    if not value.is_cuda:
        raise RuntimeError("Not implemented on the CPU")
    out_dtype = value.dtype
    if value.dtype == torch.float64:       # the reference dispatches fp64 too (ms_deform_attn_cuda.cu:69); computed in fp32 here
        value, sampling_loc, attn_weight = value.float(), sampling_loc.float(), attn_weight.float()
    return ops.msda_forward(value.contiguous(), spatial_shapes.contiguous(), level_start_index.contiguous(),
                            sampling_loc.contiguous().float(), attn_weight.contiguous().float(), im2col_step).to(out_dtype)


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


def _is_power_of_2(n):
    if (not isinstance(n, int)) or (n < 0):
        raise ValueError("invalid input for _is_power_of_2: {} (type: {})".format(n, type(n)))
    return (n & (n - 1) == 0) and n != 0


class MSDeformAttn(nn.Module):
    def __init__(self, d_model=256, n_levels=4, n_heads=8, n_points=4):
        super().__init__()
        if d_model % n_heads != 0:
            raise ValueError('d_model must be divisible by n_heads, but got {} and {}'.format(d_model, n_heads))
        if not _is_power_of_2(d_model // n_heads):
            warnings.warn("You'd better set d_model in MSDeformAttn to make the dimension of each attention head a power of 2 "
                          "which is more efficient in our CUDA implementation.")
        self.im2col_step = 128
        self.d_model, self.n_levels, self.n_heads, self.n_points = d_model, n_levels, n_heads, n_points
        self.sampling_offsets = nn.Linear(d_model, n_heads * n_levels * n_points * 2)
        self.attention_weights = nn.Linear(d_model, n_heads * n_levels * n_points)
        self.value_proj = nn.Linear(d_model, d_model)
        self.output_proj = nn.Linear(d_model, d_model)
        self._reset_parameters()

    def _reset_parameters(self):
        # OPS/modules/ms_deform_attn.py:66-80
        constant_(self.sampling_offsets.weight.data, 0.)
        thetas = torch.arange(self.n_heads, dtype=torch.float32) * (2.0 * math.pi / self.n_heads)
        grid_init = torch.stack([thetas.cos(), thetas.sin()], -1)
        grid_init = (grid_init / grid_init.abs().max(-1, keepdim=True)[0]).view(self.n_heads, 1, 1, 2).repeat(
            1, self.n_levels, self.n_points, 1)
        for i in range(self.n_points):
            grid_init[:, :, i, :] *= i + 1
        with torch.no_grad():
            self.sampling_offsets.bias = nn.Parameter(grid_init.view(-1))
        constant_(self.attention_weights.weight.data, 0.)
        constant_(self.attention_weights.bias.data, 0.)
        xavier_uniform_(self.value_proj.weight.data)
        constant_(self.value_proj.bias.data, 0.)
        xavier_uniform_(self.output_proj.weight.data)
        constant_(self.output_proj.bias.data, 0.)

    def forward(self, query, reference_points, input_flatten, input_spatial_shapes, input_level_start_index,
                input_padding_mask=None):
        # OPS/modules/ms_deform_attn.py:82-125; no silent CPU fallback (the reference's bare `except` is not reproduced)
        N, Len_q, _ = query.shape
        N, Len_in, _ = input_flatten.shape
        assert (input_spatial_shapes[:, 0] * input_spatial_shapes[:, 1]).sum() == Len_in
        value = self.value_proj(input_flatten)
        if input_padding_mask is not None:
            value = value.masked_fill(input_padding_mask[..., None], float(0))
        value = value.view(N, Len_in, self.n_heads, self.d_model // self.n_heads)
        sampling_offsets = self.sampling_offsets(query).view(N, Len_q, self.n_heads, self.n_levels, self.n_points, 2)
        attention_weights = self.attention_weights(query).view(N, Len_q, self.n_heads, self.n_levels * self.n_points)
        attention_weights = F.softmax(attention_weights, -1).view(N, Len_q, self.n_heads, self.n_levels, self.n_points)
        if reference_points.shape[-1] == 2:
            offset_normalizer = torch.stack([input_spatial_shapes[..., 1], input_spatial_shapes[..., 0]], -1)
            sampling_locations = reference_points[:, :, None, :, None, :] \
                + sampling_offsets / offset_normalizer[None, None, None, :, None, :]
        elif reference_points.shape[-1] == 4:
            sampling_locations = reference_points[:, :, None, :, None, :2] \
                + sampling_offsets / self.n_points * reference_points[:, :, None, :, None, 2:] * 0.5
        else:
            raise ValueError('Last dim of reference_points must be 2 or 4, but get {} instead.'.format(
                reference_points.shape[-1]))
        output = MSDeformAttnFunction.apply(value, input_spatial_shapes, input_level_start_index, sampling_locations,
                                            attention_weights, self.im2col_step)
        return self.output_proj(output)
