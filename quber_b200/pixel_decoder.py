"""The optional multi-scale deformable-attention pixel decoder — drop-in for
MSMF/modeling/pixel_decoder/msdeformattn.py (MSMF = ext_modules/UnseenObjectsWithMeanShift/MSMFormer/meanshiftformer):
`MSDeformAttnTransformerEncoderLayer` (:95-131), `MSDeformAttnTransformerEncoder` (:134-161),
`MSDeformAttnTransformerEncoderOnly` (:21-92), `MSDeformAttnPixelDecoder` (:164-358) and
`PositionEmbeddingSine` (MSMF/modeling/transformer_decoder/position_encoding.py:12-52), with the same module and
parameter names (a reference state_dict loads unchanged).  Selected by
cfg.MODEL.SEM_SEG_HEAD.PIXEL_DECODER_NAME == "MSDeformAttnPixelDecoder" (maskrefiner/config.py:172).

What runs where, per encoder layer (rows = B * S tokens, S = sum of level pixels):
  sampling_offsets + attention_weights Linear (ONE GEMM on the concatenated [288, 256] weight — both read
  query = src + pos), value_proj, output_proj, FFN linear1 (+ReLU) / linear2   -> quber_gemm_bias_act (tcgen05)
  softmax over L*P, reference-point arithmetic, multi-scale bilinear sampling -> quber_msda_forward_fused (one kernel)
  residual + LayerNorm (+ the next layer's operands)                          -> quber_add_layer_norm
`precision`: "fp32" (the reference's, `autocast(enabled=False)` msdeformattn.py:314) = split-operand 3xTF32 tcgen05 GEMMs
(fp32 accuracy, quber_gemm_bias_act_x3) + fp32 sampling; "tf32" = one kind::tf32 product per GEMM; "bf16" = bf16 GEMMs,
bf16 value tensor and bf16 sampling (fp32 weights / accumulation inside the sampler).  No cuBLAS call on any of them.  Dropout layers are kept for state-dict / repr parity; inference only.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn

from . import ops
from .linear import PackedLinear
from .msda import MSDeformAttn
from .network import Conv2d, get_norm


class PositionEmbeddingSine(nn.Module):
    """position_encoding.py:12-52 with mask=None (the pixel decoder never passes a mask): depends only on the level
    shape, so it is cached per (H, W, device)."""

    def __init__(self, num_pos_feats=64, temperature=10000, normalize=False, scale=None):
        super().__init__()
        if scale is not None and normalize is False:
            raise ValueError("normalize should be True if scale is passed")
        self.num_pos_feats, self.temperature, self.normalize = num_pos_feats, temperature, normalize
        self.scale = 2 * math.pi if scale is None else scale
        self._cache = {}

    def forward(self, x, mask=None):
        key = (x.shape[2], x.shape[3], str(x.device))
        if mask is None and key in self._cache:
            return self._cache[key].expand(x.shape[0], -1, -1, -1)
        if mask is None:
            mask = torch.zeros((1, x.size(2), x.size(3)), device=x.device, dtype=torch.bool)
        not_mask = ~mask
        y_embed = not_mask.cumsum(1, dtype=torch.float32)
        x_embed = not_mask.cumsum(2, dtype=torch.float32)
        if self.normalize:
            eps = 1e-6
            y_embed = y_embed / (y_embed[:, -1:, :] + eps) * self.scale
            x_embed = x_embed / (x_embed[:, :, -1:] + eps) * self.scale
        dim_t = torch.arange(self.num_pos_feats, dtype=torch.float32, device=x.device)
        dim_t = self.temperature ** (2 * (dim_t // 2) / self.num_pos_feats)
        pos_x = x_embed[:, :, :, None] / dim_t
        pos_y = y_embed[:, :, :, None] / dim_t
        pos_x = torch.stack((pos_x[:, :, :, 0::2].sin(), pos_x[:, :, :, 1::2].cos()), dim=4).flatten(3)
        pos_y = torch.stack((pos_y[:, :, :, 0::2].sin(), pos_y[:, :, :, 1::2].cos()), dim=4).flatten(3)
        pos = torch.cat((pos_y, pos_x), dim=3).permute(0, 3, 1, 2)
        if pos.shape[0] == 1:
            self._cache[key] = pos
            return pos.expand(x.shape[0], -1, -1, -1)
        return pos


def _get_activation_fn(activation):
    if activation == "relu":
        return F.relu
    if activation == "gelu":
        return F.gelu
    if activation == "glu":
        return F.glu
    raise RuntimeError(f"activation should be relu/gelu, not {activation}.")


class MSDeformAttnTransformerEncoderLayer(nn.Module):
    def __init__(self, d_model=256, d_ffn=1024, dropout=0.1, activation="relu", n_levels=4, n_heads=8, n_points=4):
        super().__init__()
        self.self_attn = MSDeformAttn(d_model, n_levels, n_heads, n_points)
        self.dropout1 = nn.Dropout(dropout)
        self.norm1 = nn.LayerNorm(d_model)
        self.linear1 = nn.Linear(d_model, d_ffn)
        self.activation_name = activation
        self.activation = _get_activation_fn(activation)
        self.dropout2 = nn.Dropout(dropout)
        self.linear2 = nn.Linear(d_ffn, d_model)
        self.dropout3 = nn.Dropout(dropout)
        self.norm2 = nn.LayerNorm(d_model)
        self.precision = "fp32"

    @staticmethod
    def with_pos_embed(tensor, pos):
        return tensor if pos is None else tensor + pos

    def forward_ffn(self, src):
        src2 = self.linear2(self.dropout2(self.activation(self.linear1(src))))
        src = src + self.dropout3(src2)
        return self.norm2(src)

    # ---- kernels (every precision; there is no cuBLAS / torch route on the GPU) -----------------------------
    def _ffn(self):
        p = self.__dict__.get("_ffn_packed")
        if p is None:
            p = dict(l1=PackedLinear(self.linear1), l2=PackedLinear(self.linear2))
            self.__dict__["_ffn_packed"] = p
        return p

    def _forward_kernels(self, src, pos, reference_points, spatial_shapes, level_start_index, padding_mask, pre=None,
                         emit_next=False, host_shapes=None):
        """msdeformattn.py:122-131 on this repo's kernels.  `pre` = (x, q) already in the GEMM dtype, written by the
        previous layer's last LayerNorm pass; `emit_next` asks this layer's last LayerNorm for the same pair (returned
        as the second value)."""
        a = self.self_attn
        prec = self.precision
        a.precision = prec
        bf16 = prec == "bf16"
        dt = torch.bfloat16 if bf16 else torch.float32
        pk, ffn = a._packs(), self._ffn()
        N, S, C = src.shape
        M = a.n_heads
        if pre is not None:
            x, q = pre
        else:
            q = self.with_pos_embed(src, pos).to(dt)
            x = src.to(dt)
        value = pk["v"](x, prec, out_dtype=dt)                                                # ms_deform_attn.py:98
        if padding_mask is not None:
            value = value.masked_fill(padding_mask[..., None], float(0))
        oa = pk["qk"](q, prec, out_dtype=torch.float32)                                       # :102-103, one GEMM
        # softmax over L*P, ref + off / (W_l, H_l) and the sampling in ONE kernel (:104-117)
        samp = a.sample(value.view(N, S, M, C // M), oa.view(N, S, -1), reference_points, spatial_shapes, level_start_index,
                        levels=host_shapes)
        src2 = pk["o"](samp, prec, out_dtype=torch.float32)                                   # :124
        fused_ln = C % 128 == 0 and C <= 1024 and C // 128 not in (5, 7)
        if fused_ln:                                                                          # norm1(src + src2), one pass
            r = ops.add_layer_norm(src.contiguous(), src2.view(N, S, C), self.norm1.weight, self.norm1.bias, self.norm1.eps, want_bf16=bf16)
            src, src16 = r if bf16 else (r, None)
        else:
            src, src16 = self.norm1(src + src2), None
        if self.activation_name != "relu":
            out = self.forward_ffn(src)
            return (out, None) if emit_next else out
        hmid = ffn["l1"](src16 if src16 is not None else src, prec, act="relu", out_dtype=dt)  # msdeformattn.py:116-120
        src2 = ffn["l2"](hmid, prec, out_dtype=torch.float32)
        if not emit_next:
            if fused_ln:
                return ops.add_layer_norm(src, src2.view(N, S, C), self.norm2.weight, self.norm2.bias, self.norm2.eps)
            return self.norm2(src + src2)
        if fused_ln and pos is not None and pos.dtype == torch.float32 and pos.shape == src.shape:
            # norm2 also writes the next layer's operands: x (bf16 copy; the fp32 result itself otherwise) and q = x + pos
            r = ops.add_layer_norm(src, src2.view(N, S, C), self.norm2.weight, self.norm2.bias, self.norm2.eps, want_bf16=bf16,
                                   pos=pos.contiguous(), q_dtype=dt)
            return r[0], ((r[1], r[2]) if bf16 else (r[0], r[1]))
        out = ops.add_layer_norm(src, src2.view(N, S, C), self.norm2.weight, self.norm2.bias, self.norm2.eps) if fused_ln \
            else self.norm2(src + src2)
        return out, None

    def forward(self, src, pos, reference_points, spatial_shapes, level_start_index, padding_mask=None):
        if not src.is_cuda:
            raise RuntimeError("Not implemented on the CPU")          # as the op itself (OPS/src/ms_deform_attn.h:43)
        return self._forward_kernels(src, pos, reference_points, spatial_shapes, level_start_index, padding_mask)


class MSDeformAttnTransformerEncoder(nn.Module):
    def __init__(self, encoder_layer, num_layers):
        super().__init__()
        import copy
        self.layers = nn.ModuleList([copy.deepcopy(encoder_layer) for _ in range(num_layers)])
        self.num_layers = num_layers

    @staticmethod
    def get_reference_points(spatial_shapes, valid_ratios, device):
        ref_list = []
        for lvl, (H_, W_) in enumerate(spatial_shapes):
            H_, W_ = int(H_), int(W_)
            ref_y, ref_x = torch.meshgrid(torch.linspace(0.5, H_ - 0.5, H_, dtype=torch.float32, device=device),
                                          torch.linspace(0.5, W_ - 0.5, W_, dtype=torch.float32, device=device), indexing="ij")
            ref_y = ref_y.reshape(-1)[None] / (valid_ratios[:, None, lvl, 1] * H_)
            ref_x = ref_x.reshape(-1)[None] / (valid_ratios[:, None, lvl, 0] * W_)
            ref_list.append(torch.stack((ref_x, ref_y), -1))
        reference_points = torch.cat(ref_list, 1)
        return reference_points[:, :, None] * valid_ratios[:, None]

    def forward(self, src, spatial_shapes, level_start_index, valid_ratios, pos=None, padding_mask=None, host_shapes=None):
        output = src
        shapes = host_shapes if host_shapes is not None else spatial_shapes.tolist()
        reference_points = self.get_reference_points(shapes, valid_ratios, device=src.device)
        pre = None
        for i, layer in enumerate(self.layers):
            if output.is_cuda:
                nxt = i + 1 < len(self.layers) and self.layers[i + 1].precision == layer.precision
                r = layer._forward_kernels(output, pos, reference_points, spatial_shapes, level_start_index, padding_mask,
                                           pre=pre, emit_next=nxt, host_shapes=shapes)
                output, pre = r if nxt else (r, None)
            else:
                output, pre = layer(output, pos, reference_points, spatial_shapes, level_start_index, padding_mask), None
        return output


class MSDeformAttnTransformerEncoderOnly(nn.Module):
    def __init__(self, d_model=256, nhead=8, num_encoder_layers=6, dim_feedforward=1024, dropout=0.1, activation="relu",
                 num_feature_levels=4, enc_n_points=4):
        super().__init__()
        self.d_model, self.nhead = d_model, nhead
        layer = MSDeformAttnTransformerEncoderLayer(d_model, dim_feedforward, dropout, activation, num_feature_levels, nhead,
                                                    enc_n_points)
        self.encoder = MSDeformAttnTransformerEncoder(layer, num_encoder_layers)
        self.level_embed = nn.Parameter(torch.Tensor(num_feature_levels, d_model))
        self._reset_parameters()

    def _reset_parameters(self):
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.xavier_uniform_(p)
        for m in self.modules():
            if isinstance(m, MSDeformAttn):
                m._reset_parameters()
        nn.init.normal_(self.level_embed)

    def set_precision(self, precision):
        if precision not in ("fp32", "tf32", "bf16"):
            raise ValueError("precision must be fp32, tf32 or bf16")
        for l in self.encoder.layers:
            l.precision = precision

    def forward(self, srcs, pos_embeds):
        # masks are all-False in the reference (msdeformattn.py:62), so every valid ratio is exactly 1
        src_flatten, host_shapes = [], []
        for lvl, src in enumerate(srcs):
            bs, c, h, w = src.shape
            host_shapes.append((h, w))
            src_flatten.append(src.flatten(2).transpose(1, 2))       # a view for channels-last sources: the cat is one copy
        src_flatten = torch.cat(src_flatten, 1)
        # position + level embedding (msdeformattn.py:70-72) depends only on the level shapes and the level_embed
        # parameter: built once per (sources' identity, parameter version), not per forward
        key = tuple((p.data_ptr(), p._version, tuple(p.shape)) for p in pos_embeds) + ((self.level_embed.data_ptr(), self.level_embed._version),)
        cache = self.__dict__.get("_lvl_pos")
        if cache is None or cache[0] != key:
            lvl_pos = torch.cat([p.flatten(2).transpose(1, 2) + self.level_embed[l].view(1, 1, -1) for l, p in enumerate(pos_embeds)], 1)
            cache = self.__dict__["_lvl_pos"] = (key, lvl_pos.contiguous())
        lvl_pos = cache[1]
        spatial_shapes = torch.as_tensor(host_shapes, dtype=torch.long, device=src_flatten.device)
        level_start_index = torch.cat((spatial_shapes.new_zeros((1,)), spatial_shapes.prod(1).cumsum(0)[:-1]))
        valid_ratios = torch.ones((src_flatten.shape[0], len(srcs), 2), dtype=torch.float32, device=src_flatten.device)
        memory = self.encoder(src_flatten, spatial_shapes, level_start_index, valid_ratios, lvl_pos, None, host_shapes=host_shapes)
        return memory, spatial_shapes, level_start_index


class MSDeformAttnPixelDecoder(nn.Module):
    """msdeformattn.py:164-358.  `input_shape`: {name: (channels, stride)}."""

    def __init__(self, input_shape, *, transformer_dropout=0.0, transformer_nheads=8, transformer_dim_feedforward=1024,
                 transformer_enc_layers=6, conv_dim=256, mask_dim=256, norm="GN",
                 transformer_in_features=("res3", "res4", "res5"), common_stride=4, transformer_enc_n_points=4):
        super().__init__()
        tin = sorted(((k, v) for k, v in input_shape.items() if k in transformer_in_features), key=lambda kv: kv[1][1])
        items = sorted(input_shape.items(), key=lambda kv: kv[1][1])
        self.in_features = [k for k, _ in items]
        self.feature_strides = [v[1] for _, v in items]
        self.feature_channels = [v[0] for _, v in items]
        self.transformer_in_features = [k for k, _ in tin]
        tch = [v[0] for _, v in tin]
        self.transformer_feature_strides = [v[1] for _, v in tin]
        self.transformer_num_feature_levels = len(self.transformer_in_features)
        chans = tch[::-1] if self.transformer_num_feature_levels > 1 else [tch[-1]]
        self.input_proj = nn.ModuleList([nn.Sequential(nn.Conv2d(c, conv_dim, kernel_size=1), nn.GroupNorm(32, conv_dim))
                                         for c in chans])
        for proj in self.input_proj:
            nn.init.xavier_uniform_(proj[0].weight, gain=1)
            nn.init.constant_(proj[0].bias, 0)
        self.transformer = MSDeformAttnTransformerEncoderOnly(
            d_model=conv_dim, dropout=transformer_dropout, nhead=transformer_nheads, dim_feedforward=transformer_dim_feedforward,
            num_encoder_layers=transformer_enc_layers, num_feature_levels=self.transformer_num_feature_levels,
            enc_n_points=transformer_enc_n_points)
        self.pe_layer = PositionEmbeddingSine(conv_dim // 2, normalize=True)
        self.mask_dim = mask_dim
        self.mask_features = Conv2d(conv_dim, mask_dim, kernel_size=1, stride=1, padding=0)
        self.maskformer_num_feature_levels = 3
        self.common_stride = common_stride
        stride = min(self.transformer_feature_strides)
        self.num_fpn_levels = int(np.log2(stride) - np.log2(self.common_stride))
        lateral_convs, output_convs = [], []
        use_bias = norm == ""
        for idx, cin in enumerate(self.feature_channels[: self.num_fpn_levels]):
            lateral = Conv2d(cin, conv_dim, kernel_size=1, bias=use_bias, norm=get_norm(norm, conv_dim))
            output = Conv2d(conv_dim, conv_dim, kernel_size=3, stride=1, padding=1, bias=use_bias, norm=get_norm(norm, conv_dim),
                            activation=F.relu)
            self.add_module("adapter_{}".format(idx + 1), lateral)
            self.add_module("layer_{}".format(idx + 1), output)
            lateral_convs.append(lateral)
            output_convs.append(output)
        self.lateral_convs, self.output_convs = lateral_convs[::-1], output_convs[::-1]

    @classmethod
    def from_config(cls, cfg, input_shape, mask_dim=None, in_features=None):
        """msdeformattn.py:290-311 (`from_config`): the keys of maskrefiner/config.py:172,186,202-205.  `input_shape`
        {name: (channels, stride)} or detectron2 ShapeSpecs; `in_features` / `mask_dim` override SEM_SEG_HEAD.IN_FEATURES /
        MASK_DIM (the maskrefiner head feeds it the RGB-D fusion backbone's levels and wants CONVS_DIM channels back)."""
        h = cfg.MODEL.SEM_SEG_HEAD
        mf = cfg.MODEL.get("MASK_FORMER", {}) or {}
        shapes = {k: (v if isinstance(v, tuple) else (v.channels, v.stride)) for k, v in input_shape.items()}
        feats = list(h.IN_FEATURES) if in_features is None else list(in_features)
        shapes = {k: v for k, v in shapes.items() if k in feats}
        tin = [f for f in h.DEFORMABLE_TRANSFORMER_ENCODER_IN_FEATURES if f in shapes]
        if not tin:
            raise ValueError("DEFORMABLE_TRANSFORMER_ENCODER_IN_FEATURES names no feature of the backbone: "
                             f"{list(h.DEFORMABLE_TRANSFORMER_ENCODER_IN_FEATURES)} vs {sorted(shapes)}")
        return cls(shapes, transformer_dropout=mf.get("DROPOUT", 0.0),
                   transformer_nheads=h.get("DEFORMABLE_TRANSFORMER_ENCODER_N_HEADS", mf.get("NHEADS", 8)),
                   transformer_dim_feedforward=1024, transformer_enc_layers=h.TRANSFORMER_ENC_LAYERS or 6,
                   conv_dim=h.CONVS_DIM, mask_dim=h.MASK_DIM if mask_dim is None else mask_dim, norm=h.NORM,
                   transformer_in_features=tin, common_stride=h.COMMON_STRIDE,
                   transformer_enc_n_points=h.get("DEFORMABLE_TRANSFORMER_ENCODER_N_POINTS", 4))

    def set_precision(self, precision):
        self.transformer.set_precision(precision)

    @torch.no_grad()
    def forward_features(self, features):
        srcs, pos = [], []
        for idx, f in enumerate(self.transformer_in_features[::-1]):
            x = features[f].float()
            srcs.append(self.input_proj[idx](x))
            pos.append(self.pe_layer(x))
        y, spatial_shapes, level_start_index = self.transformer(srcs, pos)
        bs = y.shape[0]
        sizes = [int(s.shape[2] * s.shape[3]) for s in srcs]
        out = [z.transpose(1, 2).reshape(bs, -1, s.shape[2], s.shape[3]) for z, s in zip(torch.split(y, sizes, dim=1), srcs)]
        for idx, f in enumerate(self.in_features[: self.num_fpn_levels][::-1]):
            cur_fpn = self.lateral_convs[idx](features[f].float())
            yy = cur_fpn + F.interpolate(out[-1], size=cur_fpn.shape[-2:], mode="bilinear", align_corners=False)
            out.append(self.output_convs[idx](yy))
        multi_scale_features = out[: self.maskformer_num_feature_levels]
        return self.mask_features(out[-1]), out[0], multi_scale_features


SEM_SEG_HEADS_REGISTRY = {"MSDeformAttnPixelDecoder": MSDeformAttnPixelDecoder}
