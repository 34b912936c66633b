"""The MaskRefiner network between the initial-mask encoding and the post-processing, restated in plain
PyTorch with the reference's module / parameter names so that a reference `state_dict` loads unchanged.

Reference: maskrefiner/modeling/backbone/resnet.py:24-76, 358-519 (DeepLabStem, ResNet builders,
RGBDFusionBackbone), maskrefiner/modeling/mask_refiner/model.py:369-458, 462-764 (SinglePredictionHead,
SinglePredictor, FusionLayers, MaskRefinerInsEmbedHead).  The detectron2 pieces those files import
(Conv2d wrapper, get_norm, FrozenBatchNorm2d, BottleneckBlock, ASPP, DeepLabV3PlusHead) are NOT vendored by
the reference; they are restated here from upstream detectron2 v0.6 — parity at
that boundary is unpinned.

SURVEY §8-a rows A3-A5 keep the 3x3 / dilated convolutions in cuDNN ("stays PyTorch"); the parts of the
network that are pure bandwidth or plain GEMMs go to the hand-written kernels when the module runs on a
GPU with `use_kernels=True`:
  * the softmax / sigmoid / channel-concat in front of every hierarchy stage (model.py:745-759)
    -> ops.fuse_concat (one pass, csrc/fusion.cu)
  * the 1x1 convolution + eval BatchNorm + ReLU that opens FusionLayers (model.py:430-440), in bf16 mode
    -> ops.gemm_bias_act on channels-last rows (tcgen05, csrc/gemm.cu) with the BatchNorm folded into W, b
On CPU (the test oracle for this file) the same modules run in plain torch fp32.
"""
import os

import torch
import torch.nn.functional as F
from torch import nn


# ------------------------------------------------------------------------------------------------ detectron2 layers
class FrozenBatchNorm2d(nn.Module):
    """detectron2.layers.FrozenBatchNorm2d: affine from buffers, eps 1e-5."""

    def __init__(self, num_features, eps=1e-5):
        super().__init__()
        self.num_features, self.eps = num_features, eps
        self.register_buffer("weight", torch.ones(num_features))
        self.register_buffer("bias", torch.zeros(num_features))
        self.register_buffer("running_mean", torch.zeros(num_features))
        self.register_buffer("running_var", torch.ones(num_features) - eps)

    def forward(self, x):
        scale = self.weight * (self.running_var + self.eps).rsqrt()
        bias = self.bias - self.running_mean * scale
        return x * scale.reshape(1, -1, 1, 1).to(x.dtype) + bias.reshape(1, -1, 1, 1).to(x.dtype)


def get_norm(norm, out_channels):
    """detectron2.layers.get_norm for the names the configs use."""
    if norm is None or norm == "":
        return None
    if norm in ("BN", "SyncBN"):
        return nn.BatchNorm2d(out_channels)
    if norm == "FrozenBN":
        return FrozenBatchNorm2d(out_channels)
    if norm == "GN":
        return nn.GroupNorm(32, out_channels)
    raise ValueError(f"unsupported norm {norm!r}")


def _fused_conv(x, w, b, m, relu, add=None):
    """conv (+ bias) (+ residual) (+ ReLU) as ONE cuDNN call where cuDNN has the fused engine (CUDA, bias present),
    the plain sequence otherwise.  Library convolution either way (the 3x3s stay cuDNN)."""
    if x.dtype != w.dtype:
        x = x.to(w.dtype)
    if add is not None and add.dtype != w.dtype:
        add = add.to(w.dtype)
    if relu and x.is_cuda and b is not None:
        if add is None:
            return torch.cudnn_convolution_relu(x, w, b, m.stride, m.padding, m.dilation, m.groups)
        return torch.cudnn_convolution_add_relu(x, w, add, 1.0, b, m.stride, m.padding, m.dilation, m.groups)
    y = F.conv2d(x, w, b, m.stride, m.padding, m.dilation, m.groups)
    if add is not None:
        y = y + add
    return F.relu_(y) if relu else y


def _gn_relu(gn, x, relu, pre_bias=None):
    """GroupNorm (+ ReLU) in place through quber_group_norm_relu (csrc/norm.cu) when the tensor qualifies, else None.
    `pre_bias`: the bias of the convolution that produced x, not yet added."""
    from . import ops
    if not x.is_cuda or not ops.group_norm_supported(x, gn.num_groups):
        return None
    return ops.group_norm_relu(x, gn.weight, gn.bias, gn.num_groups, gn.eps, relu=relu, inplace=True, pre_bias=pre_bias)


def _maxpool(conv, x):
    """The stems' 3x3 / 2 max-pool: quber_maxpool3x3s2 on the inference plan, F.max_pool2d otherwise."""
    if conv.__dict__.get("_plan") is not None and not conv.training and x.is_cuda:
        from . import ops
        if ops.maxpool3x3s2_supported(x):
            return ops.maxpool3x3s2(x)
    return F.max_pool2d(x, kernel_size=3, stride=2, padding=1)


def run_conv_gn_relu(seq, x, fast, skip_conv=False):
    """An nn.Sequential of nn.Conv2d / nn.GroupNorm / nn.ReLU (resnet.py:466-479); with `fast`, GroupNorm + ReLU pairs
    go through the fused kernel.  skip_conv: `x` already is the first convolution's output WITHOUT its bias (the K-split
    GEMM of the concat fusion); the bias still rides in the GroupNorm kernel."""
    if not fast:
        return seq(x)
    from . import ops
    mods = list(seq.children())
    i = 0
    while i < len(mods):
        m = mods[i]
        if type(m) is nn.Conv2d and m.bias is not None and i + 1 < len(mods) and isinstance(mods[i + 1], nn.GroupNorm) and x.is_cuda:
            # conv bias rides in the GroupNorm kernel's loads instead of a separate pass over the activation
            y = x if (skip_conv and i == 0) else F.conv2d(x, m.weight, None, m.stride, m.padding, m.dilation, m.groups)
            relu = i + 2 < len(mods) and isinstance(mods[i + 2], nn.ReLU)
            z = _gn_relu(mods[i + 1], y, relu, pre_bias=m.bias) if ops.group_norm_supported(y, mods[i + 1].num_groups) else None
            if z is not None:
                x = z
                i += 3 if relu else 2
                continue
            x = y + m.bias.reshape(1, -1, 1, 1).to(y.dtype)
            i += 1
            continue
        if isinstance(m, nn.GroupNorm):
            relu = i + 1 < len(mods) and isinstance(mods[i + 1], nn.ReLU)
            y = _gn_relu(m, x, relu)
            if y is not None:
                x = y
                i += 2 if relu else 1
                continue
        x = m(x)
        i += 1
    return x


class Conv2d(nn.Conv2d):
    """detectron2.layers.Conv2d: conv -> norm -> activation, `norm` a child module named "norm".

    Inference plan (`prepare_inference`): a frozen / eval-mode BatchNorm is folded into the convolution
    (W' = W * s, b' = beta + (b - mean) * s, s = gamma / sqrt(var + eps)), weights are cached in the compute dtype and
    memory format, and bias + ReLU (+ the bottleneck's residual add) ride in the convolution's epilogue."""

    def __init__(self, *args, norm=None, activation=None, **kwargs):
        super().__init__(*args, **kwargs)
        self.norm = norm
        self.activation = activation

    def _sig(self):
        """Identity + version of every tensor the plan was derived from (in-place edits, checkpoint loads, .to())."""
        ts = [self.weight, self.bias]
        if isinstance(self.norm, (FrozenBatchNorm2d, nn.BatchNorm2d)):
            ts += [self.norm.weight, self.norm.bias, self.norm.running_mean, self.norm.running_var]
        return tuple((t.data_ptr(), t._version) for t in ts if t is not None)

    def plan(self, dtype=torch.float32, channels_last=False):
        self.__dict__["_plan_args"] = (dtype, channels_last)
        n = self.norm
        w, b = self.weight.detach().float(), None if self.bias is None else self.bias.detach().float()
        folded = False
        if isinstance(n, FrozenBatchNorm2d) or (isinstance(n, nn.BatchNorm2d) and not n.training):
            s = n.weight.detach().float() * (n.running_var.float() + n.eps).rsqrt()
            b = n.bias.detach().float() + ((0 if b is None else b) - n.running_mean.float()) * s
            w = w * s.reshape(-1, 1, 1, 1)
            folded = True
        w = w.to(dtype)
        # cuDNN 9 on B200: a 3x3 whose dilation exceeds half the map (ASPP dilation 18 on 30x40) falls to
        # conv2d_grouped_direct_kernel with channels-last operands — 298 ms instead of 1.1 ms for 32 frames in bf16
        # (scripts/bench_dilated_conv.py); such calls get NCHW operands (forward() decides per input shape).
        self.__dict__["_w_nchw"] = w.contiguous() if (channels_last and max(self.dilation) > 1) else None
        if channels_last:
            w = w.contiguous(memory_format=torch.channels_last)
        self.__dict__["_plan"] = (w.contiguous() if not channels_last else w, None if b is None else b.to(dtype).contiguous(), folded,
                                  self._sig())

    def _padded_weight(self, p, cin):
        """The plan's weight with zero input channels appended up to `cin` (same sums), cached per plan."""
        c = self.__dict__.get("_w_pad")
        if c is None or c[0] is not p[0] or c[1] != cin:
            w = F.pad(p[0], (0, 0, 0, 0, 0, cin - p[0].shape[1]))
            if p[0].is_contiguous(memory_format=torch.channels_last):
                w = w.contiguous(memory_format=torch.channels_last)
            c = (p[0], cin, w)
            self.__dict__["_w_pad"] = c
        return c[2]

    def drop_plan(self):
        self.__dict__.pop("_plan", None)
        self.__dict__.pop("_w_pad", None)

    def forward(self, x, relu=False, add=None):
        """`relu` / `add`: the caller's trailing F.relu_ / residual add, handed over so that they can be fused."""
        p = self.__dict__.get("_plan")
        if p is None or self.training:
            x = F.conv2d(x, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups)
            if self.norm is not None:
                x = self.norm(x)
            if self.activation is not None:
                x = self.activation(x)
            if add is not None:
                x = x + add
            return F.relu_(x) if relu else x
        if p[3] != self._sig():                      # parameters changed under the plan: rebuild it
            self.plan(*self.__dict__["_plan_args"])
            p = self.__dict__["_plan"]
        w, b, folded = p[:3]
        if x.shape[1] > w.shape[1] and self.groups == 1:          # input with zero channels appended (ops.split_rgbd_input(pad8))
            w = self._padded_weight(p, x.shape[1])
        wn = self.__dict__.get("_w_nchw")
        if wn is not None and 2 * max(self.dilation) > min(x.shape[-2:]):
            w, x = wn, x.contiguous()
        if self.norm is None or folded:
            act = relu or self.activation is F.relu
            assert self.activation in (None, F.relu)
            return _fused_conv(x, w, b, self, act, add)
        xin = x if x.dtype == w.dtype else x.to(w.dtype)
        if isinstance(self.norm, nn.GroupNorm) and add is None and self.activation in (None, F.relu) and x.is_cuda:
            x = F.conv2d(xin, w, None, self.stride, self.padding, self.dilation, self.groups)
            y = _gn_relu(self.norm, x, relu or self.activation is F.relu, pre_bias=b)
            if y is not None:
                return y
            if b is not None:
                x = x + b.reshape(1, -1, 1, 1)
        else:
            x = F.conv2d(xin, w, b, self.stride, self.padding, self.dilation, self.groups)
        x = self.norm(x)
        if self.activation is not None:
            x = self.activation(x)
        if add is not None:
            x = x + add
        return F.relu_(x) if relu else x


def prepare_inference(module, dtype=torch.float32, channels_last=False):
    """Build (or rebuild, after a state_dict load) the inference plan of every Conv2d below `module`."""
    for m in module.modules():
        if isinstance(m, Conv2d):
            m.plan(dtype, channels_last)
        if isinstance(m, FusionLayers):
            m._folded = None
        if isinstance(m, (RGBDFusionBackbone, DeepLabV3PlusDecoder)):
            m.__dict__["_fast"] = True


def drop_inference_plan(module):
    for m in module.modules():
        if isinstance(m, Conv2d):
            m.drop_plan()
        if isinstance(m, FusionLayers):
            m._folded = None
        if isinstance(m, (RGBDFusionBackbone, DeepLabV3PlusDecoder)):
            m.__dict__["_fast"] = False


class BottleneckBlock(nn.Module):
    """detectron2.modeling.backbone.resnet.BottleneckBlock."""

    def __init__(self, in_channels, out_channels, *, bottleneck_channels, stride=1, num_groups=1, norm="BN",
                 stride_in_1x1=False, dilation=1):
        super().__init__()
        self.in_channels, self.out_channels, self.stride = in_channels, out_channels, stride
        if in_channels != out_channels:
            self.shortcut = Conv2d(in_channels, out_channels, kernel_size=1, stride=stride, bias=False,
                                   norm=get_norm(norm, out_channels))
        else:
            self.shortcut = None
        stride_1x1, stride_3x3 = (stride, 1) if stride_in_1x1 else (1, stride)
        self.conv1 = Conv2d(in_channels, bottleneck_channels, kernel_size=1, stride=stride_1x1, bias=False,
                            norm=get_norm(norm, bottleneck_channels))
        self.conv2 = Conv2d(bottleneck_channels, bottleneck_channels, kernel_size=3, stride=stride_3x3, padding=1 * dilation,
                            bias=False, groups=num_groups, dilation=dilation, norm=get_norm(norm, bottleneck_channels))
        self.conv3 = Conv2d(bottleneck_channels, out_channels, kernel_size=1, bias=False, norm=get_norm(norm, out_channels))

    def forward(self, x):
        out = self.conv1(x, relu=True)
        out = self.conv2(out, relu=True)
        ps = self.shortcut.__dict__.get("_plan") if self.shortcut is not None else None
        p3 = self.conv3.__dict__.get("_plan")
        if ps is not None and p3 is not None and not self.training and ps[2] and p3[2] and x.is_cuda \
                and ps[3] == self.shortcut._sig() and p3[3] == self.conv3._sig():
            # both BatchNorms folded: the shortcut's bias moves into conv3's epilogue (one bias pass less)
            sc = F.conv2d(x if x.dtype == ps[0].dtype else x.to(ps[0].dtype), ps[0], None, self.shortcut.stride,
                          self.shortcut.padding, self.shortcut.dilation, self.shortcut.groups)
            return _fused_conv(out, p3[0], p3[1] + ps[1], self.conv3, True, sc)
        shortcut = self.shortcut(x) if self.shortcut is not None else x
        return self.conv3(out, relu=True, add=shortcut)


# ------------------------------------------------------------------------------------------------ backbone
class DeepLabStem(nn.Module):
    """maskrefiner/modeling/backbone/resnet.py:24-76."""

    def __init__(self, in_channels=3, out_channels=128, norm="BN"):
        super().__init__()
        self.in_channels, self.out_channels, self.stride = in_channels, out_channels, 4
        h = out_channels // 2
        self.conv1 = Conv2d(in_channels, h, kernel_size=3, stride=2, padding=1, bias=False, norm=get_norm(norm, h))
        self.conv2 = Conv2d(h, h, kernel_size=3, stride=1, padding=1, bias=False, norm=get_norm(norm, h))
        self.conv3 = Conv2d(h, out_channels, kernel_size=3, stride=1, padding=1, bias=False, norm=get_norm(norm, out_channels))

    def forward(self, x):
        x = self.conv1(x, relu=True)
        x = self.conv2(x, relu=True)
        x = self.conv3(x, relu=True)
        return _maxpool(self.conv3, x)


class BasicStem(nn.Module):
    """maskrefiner/modeling/backbone/resnet.py:98-125 (note the reference's `conv_1` spelling)."""

    def __init__(self, in_channels=3, out_channels=64, norm="BN"):
        super().__init__()
        self.in_channels, self.out_channels, self.stride = in_channels, out_channels, 4
        self.conv_1 = Conv2d(in_channels, out_channels, kernel_size=7, stride=2, padding=3, bias=False,
                             norm=get_norm(norm, out_channels))

    def forward(self, x):
        return _maxpool(self.conv_1, self.conv_1(x, relu=True))


class ResNet(nn.Module):
    """maskrefiner/modeling/backbone/resnet.py:127-233: stages registered as `<prefix>res<i>`."""

    def __init__(self, stem, stages, out_features, prefix=""):
        super().__init__()
        self.stem = stem
        self.stage_names, self.stages = [], []
        num_stages = max({"res2": 1, "res3": 2, "res4": 3, "res5": 4}.get(f, 0) for f in out_features)
        stride = stem.stride
        self._out_feature_channels, self._out_feature_strides = {}, {}
        for i, blocks in enumerate(stages[:num_stages]):
            name = prefix + "res" + str(i + 2)
            stage = nn.Sequential(*blocks)
            self.add_module(name, stage)
            self.stage_names.append(name)
            self.stages.append(stage)
            for b in blocks:
                stride *= b.stride
            self._out_feature_strides[name] = stride
            self._out_feature_channels[name] = blocks[-1].out_channels
        self._out_features = [prefix + f for f in out_features]

    def forward(self, x):
        assert x.dim() == 4, f"ResNet takes an input of shape (N, C, H, W). Got {x.shape} instead!"
        out = {}
        x = self.stem(x)
        for name, stage in zip(self.stage_names, self.stages):
            x = stage(x)
            if name in self._out_features:
                out[name] = x
        return out


def build_resnet_deeplab_fusion_backbone(cfg, input_shape=None, prefix=""):
    """maskrefiner/modeling/backbone/resnet.py:358-449."""
    r = cfg.MODEL.RESNETS
    norm = r.NORM
    in_channels = 6 if cfg.INPUT.OFFSET_INPUT_ON else 3
    if r.STEM_TYPE == "basic":
        stem = BasicStem(in_channels, r.STEM_OUT_CHANNELS, norm)
    elif r.STEM_TYPE == "deeplab":
        stem = DeepLabStem(in_channels, r.STEM_OUT_CHANNELS, norm)
    else:
        raise ValueError("Unknown stem type: {}".format(r.STEM_TYPE))
    assert r.RES4_DILATION in {1, 2}, "res4_dilation cannot be {}.".format(r.RES4_DILATION)
    assert r.RES5_DILATION in {1, 2, 4}, "res5_dilation cannot be {}.".format(r.RES5_DILATION)
    if r.RES4_DILATION == 2:
        assert r.RES5_DILATION == 4
    if any(r.DEFORM_ON_PER_STAGE):
        raise ValueError("deformable-conv ResNet stages are not part of the maskrefiner configs")
    num_blocks_per_stage = {50: [3, 4, 6, 3], 101: [3, 4, 23, 3], 152: [3, 8, 36, 3]}[r.DEPTH]
    bottleneck = r.NUM_GROUPS * r.WIDTH_PER_GROUP
    cin, cout = r.STEM_OUT_CHANNELS, r.RES2_OUT_CHANNELS
    max_stage = max({"res2": 2, "res3": 3, "res4": 4, "res5": 5}[f] for f in r.OUT_FEATURES)
    stages = []
    for idx, stage_idx in enumerate(range(2, max_stage + 1)):
        dilation = r.RES4_DILATION if stage_idx == 4 else r.RES5_DILATION if stage_idx == 5 else 1
        first_stride = 1 if idx == 0 or dilation > 1 else 2
        n = num_blocks_per_stage[idx]
        dil = [dilation * mg for mg in r.RES5_MULTI_GRID] if stage_idx == 5 else [dilation] * n
        blocks = []
        for i in range(n):
            blocks.append(BottleneckBlock(cin if i == 0 else cout, cout, bottleneck_channels=bottleneck,
                                          stride=first_stride if i == 0 else 1, num_groups=r.NUM_GROUPS, norm=norm,
                                          stride_in_1x1=r.STRIDE_IN_1X1, dilation=dil[i]))
        stages.append(blocks)
        cin, cout, bottleneck = cout, cout * 2, bottleneck * 2
    return ResNet(stem, stages, out_features=list(r.OUT_FEATURES), prefix=prefix)


class RGBDFusionBackbone(nn.Module):
    """maskrefiner/modeling/backbone/resnet.py:453-507."""

    def __init__(self, cfg, rgb_backbone, depth_backbone):
        super().__init__()
        self.rgb_backbone, self.depth_backbone = rgb_backbone, depth_backbone
        self.fusion_strategy = cfg.MODEL.BACKBONE.FUSION_STRATEGY
        self.num_fusion_layers = cfg.MODEL.BACKBONE.NUM_FUSION_LAYERS
        self.out_features = list(cfg.MODEL.RESNETS.OUT_FEATURES)
        channels = {"res2": 256, "res3": 512, "res4": 1024, "res5": 2048}
        self.fusion_layers = {}
        for f in self.out_features:
            seq, c = nn.Sequential(), channels[f]
            if self.fusion_strategy == "concat":
                seq.add_module("conv", nn.Conv2d(c * 2, c, kernel_size=1, stride=1, padding=0))
                seq.add_module("gn", nn.GroupNorm(32, c))
                seq.add_module("relu", nn.ReLU(inplace=True))
            for i in range(self.num_fusion_layers):
                if f == "res5":
                    continue
                seq.add_module("conv{}".format(i), nn.Conv2d(c, c, kernel_size=3, stride=1, padding=1))
                seq.add_module("gn{}".format(i), nn.GroupNorm(32, c))
                seq.add_module("relu{}".format(i), nn.ReLU(inplace=True))
            self.fusion_layers[f] = seq
            self.add_module("fusion_{}".format(f), seq)
        self._out_features = rgb_backbone._out_features
        self._out_feature_channels = rgb_backbone._out_feature_channels
        self._out_feature_strides = rgb_backbone._out_feature_strides
        self.size_divisibility = 0
        self.splits_packed_input = True              # forward takes the packed fp32 [B,9,H,W] input as it is (bf16 plan)

    def forward(self, x):
        fast = self.__dict__.get("_fast", False)
        stem = self.rgb_backbone.stem
        first = getattr(stem, "conv_1", None) if hasattr(stem, "conv_1") else getattr(stem, "conv1", None)     # BasicStem / DeepLabStem
        stem_plan = first.__dict__.get("_plan") if first is not None else None
        if (fast and x.is_cuda and x.dtype == torch.float32 and x.shape[1] == 9 and x.is_contiguous() and (x.shape[2] * x.shape[3]) % 4 == 0
                and stem_plan is not None and stem_plan[0].dtype == torch.bfloat16):
            from . import ops
            rgb_in, depth_in = ops.split_rgbd_input(x, pad8=True)    # both cats, the channels-last layout and the bf16 cast in one pass
        else:
            if fast and x.is_cuda and stem_plan is not None and x.dtype != stem_plan[0].dtype:
                x = x.to(stem_plan[0].dtype).contiguous(memory_format=torch.channels_last)     # what the caller used to do
            rgb_x, depth_x, offset_x = x[:, :3], x[:, 3:6], x[:, 6:]
            rgb_in, depth_in = torch.cat([rgb_x, offset_x], dim=1), torch.cat([depth_x, offset_x], dim=1)
        rgb_out = self.rgb_backbone(rgb_in)
        depth_out = self.depth_backbone(depth_in)
        for k in rgb_out.keys():
            if self.fusion_strategy == "concat":
                y = self._concat_1x1(self.fusion_layers[k], rgb_out[k], depth_out["depth_" + k]) if fast else None
                if y is not None:                               # the concatenation never exists: K-split GEMM over both streams
                    rgb_out[k] = run_conv_gn_relu(self.fusion_layers[k], y, True, skip_conv=True)
                    continue
                rgb_out[k] = torch.cat([rgb_out[k], depth_out["depth_" + k]], dim=1)
            elif self.fusion_strategy == "add":
                rgb_out[k] = rgb_out[k] + depth_out["depth_" + k]
            rgb_out[k] = run_conv_gn_relu(self.fusion_layers[k], rgb_out[k], fast)
        return rgb_out

    @staticmethod
    def _concat_1x1(seq, a, b):
        """Conv2d(2c, c, 1) of torch.cat([a, b], 1) (resnet.py:466-468, 497-501) as ONE quber_gemm_bias_act_cat2 call on the
        channels-last rows of both streams, bias left to the GroupNorm kernel behind it (`pre_bias`); None when the
        tensors do not qualify (CPU, fp32 plan, NCHW-contiguous memory): the caller then concatenates."""
        conv = seq[0]
        if len(seq) < 2 or conv.bias is None or not isinstance(seq[1], nn.GroupNorm):
            return None
        if not (type(conv) is nn.Conv2d and conv.kernel_size == (1, 1) and a.is_cuda and a.dtype == b.dtype == torch.bfloat16
                and a.shape == b.shape and conv.in_channels == 2 * a.shape[1] and a.shape[1] % 64 == 0):
            return None
        if not (a.is_contiguous(memory_format=torch.channels_last) and b.is_contiguous(memory_format=torch.channels_last)):
            return None
        from . import ops
        B, C, h, w = a.shape
        wkey = (conv.weight.data_ptr(), conv.weight._version)
        cache = conv.__dict__.get("_w_rows")
        if cache is None or cache[0] != wkey:
            cache = (wkey, conv.weight.detach().reshape(conv.out_channels, 2 * C).to(torch.bfloat16).contiguous())
            conv.__dict__["_w_rows"] = cache
        y = ops.gemm_bias_act_cat2(a.permute(0, 2, 3, 1).reshape(-1, C), b.permute(0, 2, 3, 1).reshape(-1, C), cache[1], None,
                                   out_dtype=torch.bfloat16)
        return y.view(B, h, w, conv.out_channels).permute(0, 3, 1, 2)             # NCHW view of channels-last memory


def build_resnet_deeplab_rgbd_fusion_backbone(cfg, input_shape=None):
    """maskrefiner/modeling/backbone/resnet.py:510-519."""
    assert cfg.INPUT.DEPTH_ON, "depth_on must be True for RGBD fusion backbone"
    rgb = build_resnet_deeplab_fusion_backbone(cfg, input_shape)
    depth = build_resnet_deeplab_fusion_backbone(cfg, input_shape, prefix="depth_")
    return RGBDFusionBackbone(cfg, rgb, depth)


BACKBONE_REGISTRY = {
    "build_resnet_deeplab_fusion_backbone": build_resnet_deeplab_fusion_backbone,
    "build_resnet_deeplab_rgbd_fusion_backbone": build_resnet_deeplab_rgbd_fusion_backbone,
}


# ------------------------------------------------------------------------------------------------ DeepLabV3+ decoder
class ASPP(nn.Module):
    """detectron2.layers.ASPP (non depthwise-separable branch; pool_kernel_size None = global average pooling)."""

    def __init__(self, in_channels, out_channels, dilations, *, norm, activation, dropout=0.0):
        super().__init__()
        assert len(dilations) == 3, "ASPP expects 3 dilations, got {}".format(len(dilations))
        use_bias = norm == ""
        self.dropout = dropout
        self.convs = nn.ModuleList()
        self.convs.append(Conv2d(in_channels, out_channels, kernel_size=1, bias=use_bias, norm=get_norm(norm, out_channels),
                                 activation=activation))
        for d in dilations:
            self.convs.append(Conv2d(in_channels, out_channels, kernel_size=3, padding=d, dilation=d, bias=use_bias,
                                     norm=get_norm(norm, out_channels), activation=activation))
        self.convs.append(nn.Sequential(nn.AdaptiveAvgPool2d(1),
                                        Conv2d(in_channels, out_channels, 1, bias=True, activation=activation)))
        self.project = Conv2d(5 * out_channels, out_channels, kernel_size=1, bias=use_bias, norm=get_norm(norm, out_channels),
                              activation=activation)

    def forward(self, x):
        size = x.shape[-2:]
        res = [conv(x) for conv in self.convs]
        if self.convs[0].__dict__.get("_plan") is not None and not self.training:
            res[-1] = res[-1].expand(-1, -1, size[0], size[1])     # bilinear up-sampling of a 1x1 map is a broadcast
        else:
            res[-1] = F.interpolate(res[-1], size=size, mode="bilinear", align_corners=False)
        res = self.project(torch.cat(res, dim=1))
        return F.dropout(res, self.dropout, training=self.training) if self.dropout > 0 else res


class DeepLabV3PlusDecoder(nn.Module):
    """detectron2.projects.deeplab.DeepLabV3PlusHead with num_classes=None (decoder only), non-DSC fuse convs."""

    def __init__(self, input_shape, *, project_channels, aspp_dilations, aspp_dropout, decoder_channels, common_stride, norm):
        super().__init__()
        items = sorted(input_shape.items(), key=lambda kv: kv[1][1])            # by stride
        self.in_features = [k for k, _ in items]
        in_channels = [v[0] for _, v in items]
        aspp_channels = decoder_channels[-1]
        self.common_stride = common_stride
        assert len(project_channels) == len(self.in_features) - 1
        assert len(decoder_channels) == len(self.in_features)
        use_bias = norm == ""
        self.decoder = nn.ModuleDict()
        for idx, c in enumerate(in_channels):
            stage = nn.ModuleDict()
            if idx == len(self.in_features) - 1:
                stage["project_conv"] = ASPP(c, aspp_channels, aspp_dilations, norm=norm, activation=F.relu, dropout=aspp_dropout)
                stage["fuse_conv"] = None
            else:
                stage["project_conv"] = Conv2d(c, project_channels[idx], kernel_size=1, bias=use_bias,
                                               norm=get_norm(norm, project_channels[idx]), activation=F.relu)
                stage["fuse_conv"] = nn.Sequential(
                    Conv2d(project_channels[idx] + decoder_channels[idx + 1], decoder_channels[idx], kernel_size=3, padding=1,
                           bias=use_bias, norm=get_norm(norm, decoder_channels[idx]), activation=F.relu),
                    Conv2d(decoder_channels[idx], decoder_channels[idx], kernel_size=3, padding=1, bias=use_bias,
                           norm=get_norm(norm, decoder_channels[idx]), activation=F.relu))
            self.decoder[self.in_features[idx]] = stage

    def decode(self, features):
        y = None
        for f in self.in_features[::-1]:
            proj_x = self.decoder[f]["project_conv"](features[f])
            if self.decoder[f]["fuse_conv"] is None:
                y = proj_x
            else:
                y = self.decoder[f]["fuse_conv"](self._up_cat(proj_x, y))
        return y

    def _up_cat(self, proj_x, y):
        """cat([proj_x, bilinear(y -> proj_x's size)], 1): one quber_upsample_concat pass on the inference plan."""
        if self.__dict__.get("_fast") and proj_x.is_cuda:
            from . import ops
            a = proj_x.contiguous(memory_format=torch.channels_last)
            b = y.to(a.dtype).contiguous(memory_format=torch.channels_last)
            if ops.upsample_concat_supported(a, b):
                return ops.upsample_concat(a, b)
            if os.environ.get("QUBER_DEBUG"):
                print("upsample_concat fallback:", a.shape, a.dtype, a.stride(), b.shape, b.dtype, b.stride())
        y = F.interpolate(y, size=proj_x.size()[2:], mode="bilinear", align_corners=False)
        return torch.cat([proj_x, y], dim=1)


# ------------------------------------------------------------------------------------------------ heads
class SinglePredictionHead(nn.Module):
    """model.py:369-411 (the 3x3 branch; the 5x5 depthwise-separable branch is unreachable from the reference's
    YAMLs because the head reads MODEL.SEM_SEG_HEAD.USE_DEPTHWISE_SEPARABLE_CONV — model.py:637)."""

    def __init__(self, in_channels, head_channels, use_bias, norm, use_depthwise_separable_conv=False):
        super().__init__()
        if use_depthwise_separable_conv:
            raise ValueError("depthwise-separable prediction heads are not used by any maskrefiner config")
        self.head = nn.Sequential(
            Conv2d(in_channels, in_channels, kernel_size=3, padding=1, bias=use_bias, norm=get_norm(norm, in_channels),
                   activation=F.relu),
            Conv2d(in_channels, head_channels, kernel_size=3, padding=1, bias=use_bias, norm=get_norm(norm, head_channels),
                   activation=F.relu))

    def forward(self, x):
        return self.head(x)


class SinglePredictor(nn.Module):
    """model.py:413-422."""

    def __init__(self, head_channels, out_channels):
        super().__init__()
        self.predictor = Conv2d(head_channels, out_channels, kernel_size=1)
        nn.init.normal_(self.predictor.weight, 0, 0.001)
        nn.init.constant_(self.predictor.bias, 0)

    def forward(self, x):
        return self.predictor(x)


class FusionLayers(nn.Module):
    """model.py:424-458: 1x1 conv(+bias)+BatchNorm+ReLU, then n x (3x3 conv+BN+ReLU); BN is hard-coded."""

    def __init__(self, in_channels, out_channels, num_fusion_layers, fusion_strategy="concat", norm="GN"):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.fusion_layers = nn.ModuleList()
        self.fusion_layers.append(Conv2d(in_channels, out_channels, kernel_size=1, padding=0, bias=True,
                                         norm=get_norm("BN", out_channels), activation=F.relu))
        for _ in range(num_fusion_layers):
            self.fusion_layers.append(Conv2d(out_channels, out_channels, kernel_size=3, padding=1, bias=True,
                                             norm=get_norm("BN", out_channels), activation=F.relu))
        self._folded = None

    def forward(self, x):
        for layer in self.fusion_layers:
            x = layer(x)
        return x

    # ---- tcgen05 route for the opening 1x1: eval-mode BatchNorm folded into (W, b) -----------------
    def folded_1x1(self, pitch, dtype):
        """(W' [out, pitch], b' [out]) with BN(conv(x)) == x W'^T + b' in eval mode; W' zero-padded to `pitch`."""
        l0 = self.fusion_layers[0]
        # keyed on identity + version of every tensor it derives from (Conv2d._sig), like the convolutions' own plan:
        # a checkpoint load, an in-place edit or .to() after a bf16 forward must not leave stale folded weights behind
        key = (pitch, dtype, l0._sig())
        if self._folded is None or self._folded[0] != key:
            bn = l0.norm
            scale = bn.weight * (bn.running_var + bn.eps).rsqrt()
            w = l0.weight.reshape(self.out_channels, self.in_channels) * scale[:, None]
            b = (l0.bias - bn.running_mean) * scale + bn.bias
            wp = torch.zeros((self.out_channels, pitch), dtype=torch.float32, device=w.device)
            wp[:, : self.in_channels] = w
            self._folded = (key, wp.to(dtype).contiguous(), b.float().contiguous())
        return self._folded[1], self._folded[2]

    def forward_rows(self, rows, B, h, w):
        """rows: channels-last [B*h*w, pitch] (ops.fuse_concat) -> NCHW-shaped, channels-last-strided output of the
        whole block; the opening 1x1 + BN + ReLU is ONE quber_gemm_bias_act call."""
        from . import ops
        wp, bp = self.folded_1x1(rows.shape[1], rows.dtype)
        y = ops.gemm_bias_act(rows, wp, bp, act="relu", out_dtype=torch.float32)
        x = y.view(B, h, w, self.out_channels).permute(0, 3, 1, 2)               # NCHW view of channels-last memory
        for layer in list(self.fusion_layers)[1:]:
            x = layer(x)
        return x


_EEE_CHANNELS = {"e3": 4, "e33": 3, "e2": 2, "e32": 2}        # model.py:549-568


class MaskRefinerInsEmbedHead(DeepLabV3PlusDecoder):
    """model.py:461-764, inference branches.  `layers()` returns the 1/COMMON_STRIDE-resolution logits; the x4
    up-sampling of model.py:688-708 is NOT done here — quber_post_process up-samples on the fly (A6)."""

    def __init__(self, cfg, input_shape):
        h = cfg.MODEL.INS_EMBED_HEAD
        in_features = list(h.IN_FEATURES)
        decoder_channels = [h.CONVS_DIM] * (len(in_features) - 1) + [h.ASPP_CHANNELS]          # model.py:617-619
        super().__init__({k: v for k, v in input_shape.items() if k in in_features}, project_channels=list(h.PROJECT_CHANNELS),
                         aspp_dilations=list(h.ASPP_DILATIONS), aspp_dropout=h.ASPP_DROPOUT, decoder_channels=decoder_channels,
                         common_stride=h.COMMON_STRIDE, norm=h.NORM)
        norm, hc = h.NORM, h.HEAD_CHANNELS
        use_bias = norm == ""
        dsc = cfg.MODEL.SEM_SEG_HEAD.USE_DEPTHWISE_SEPARABLE_CONV                               # model.py:637 (sic)
        self.error_type = h.ERROR_TYPE
        self.eee_mask_on, self.eee_boundary_on = h.EEE_MASK_ON, h.EEE_BOUNDARY_ON
        d0 = decoder_channels[0]
        self.foreground_pred_head = SinglePredictionHead(d0, hc, use_bias, norm, dsc)
        self.foreground_predictor = SinglePredictor(hc, 1)
        self.center_pred_head = SinglePredictionHead(d0, hc, use_bias, norm, dsc)
        self.center_predictor = SinglePredictor(hc, 1)
        self.offset_pred_head = SinglePredictionHead(d0, hc, use_bias, norm, dsc)
        self.offset_predictor = SinglePredictor(hc, 2)
        if self.eee_mask_on:
            self.eee_mask_pred_head = SinglePredictionHead(d0, hc, use_bias, norm, dsc)
            self.eee_mask_predictor = SinglePredictor(hc, _EEE_CHANNELS[self.error_type])
        if self.eee_boundary_on:
            self.eee_boundary_pred_head = SinglePredictionHead(d0, hc, use_bias, norm, dsc)
            self.eee_boundary_predictor = SinglePredictor(hc, _EEE_CHANNELS[self.error_type])
        self.hierarchical_fusion_on = h.HIERARCHICAL_FUSION_ON
        self.hierarchy = [list(x) for x in h.HIERARCHY]
        self.fusion_target = list(h.FUSION_TARGET)
        self.fusion_layers = {}
        if self.hierarchical_fusion_on:
            for i in range(1, len(self.hierarchy)):
                cin = d0
                if "feat" in self.fusion_target:
                    cin += hc * len(self.hierarchy[i - 1])
                if "pred" in self.fusion_target:
                    for name in self.hierarchy[i - 1]:
                        cin += _EEE_CHANNELS[self.error_type] if "eee" in name else 2 if "offset" in name else 1
                self.fusion_layers[i] = FusionLayers(cin, d0, h.NUM_FUSION_LAYERS, h.FUSION_STRATEGY, norm)
                self.add_module("fusion_layers_{}".format(i), self.fusion_layers[i])
        self.use_kernels = False           # set by MaskRefiner: route softmax/sigmoid/cat (+1x1) through the CUDA kernels
        self.gemm_1x1 = False              # bf16 mode: opening 1x1 of FusionLayers on tcgen05
        # Optional pixel decoder, selected by the reference's own key (maskrefiner/config.py:172, :202-205):
        # MODEL.SEM_SEG_HEAD.PIXEL_DECODER_NAME == "MSDeformAttnPixelDecoder" replaces the DeepLabV3+ decoder by the
        # multi-scale deformable-attention pixel decoder (MSMF/modeling/pixel_decoder/msdeformattn.py:164-358); its
        # stride-COMMON_STRIDE `mask_features` (CONVS_DIM channels) are what the hierarchy of heads consumes.
        self.pixel_decoder = None
        pd_name = cfg.MODEL.SEM_SEG_HEAD.get("PIXEL_DECODER_NAME", "BasePixelDecoder")
        if pd_name in ("MSDeformAttnPixelDecoder",):
            from .pixel_decoder import SEM_SEG_HEADS_REGISTRY
            feats = {k: v for k, v in input_shape.items() if k in in_features}
            self.pixel_decoder = SEM_SEG_HEADS_REGISTRY[pd_name].from_config(cfg, feats, mask_dim=d0, in_features=in_features)
            del self.decoder               # the DeepLabV3+ decoder is not built into this variant's state_dict

    def _stage_input(self, y, feature_dict, output_dict, prev):
        """model.py:745-759 -> (NCHW tensor | None, channels-last rows | None)."""
        srcs, ops_ = [y], ["copy"]
        if "feat" in self.fusion_target:
            for k in prev:
                srcs.append(feature_dict[k]); ops_.append("copy")
        if "pred" in self.fusion_target:
            for k in prev:
                srcs.append(output_dict[k]); ops_.append("softmax" if "eee" in k else "sigmoid")
        if self.use_kernels and y.is_cuda:
            from . import ops
            cl = torch.channels_last
            if self.gemm_1x1 and all(s.dtype in (torch.float32, torch.bfloat16) and s.is_contiguous(memory_format=cl) for s in srcs):
                ctot = sum(s.shape[1] for s in srcs)        # channels-last already: no NCHW fp32 copies
                return None, ops.fuse_concat_rows(srcs, ops_, pitch=(ctot + 7) // 8 * 8)
            srcs = [s.float().contiguous() for s in srcs]
            if self.gemm_1x1:
                ctot = sum(s.shape[1] for s in srcs)
                return None, ops.fuse_concat(srcs, ops_, layout="nhwc_bf16", pitch=(ctot + 7) // 8 * 8)
            return ops.fuse_concat(srcs, ops_, layout="nchw"), None
        parts = [s if o == "copy" else s.softmax(dim=1) if o == "softmax" else s.sigmoid() for s, o in zip(srcs, ops_)]
        return torch.cat(parts, dim=1), None

    def layers(self, features):
        if self.pixel_decoder is not None:
            y = self.pixel_decoder.forward_features({k: v for k, v in features.items()})[0].to(next(iter(features.values())).dtype)
        else:
            y = self.decode(features)
        feature_dict, output_dict = {}, {}
        if not self.hierarchical_fusion_on:
            keys = (["eee_mask"] if self.eee_mask_on else []) + (["eee_boundary"] if self.eee_boundary_on else []) + \
                   ["foreground", "center", "offset"]
            for k in keys:
                feature_dict[k] = getattr(self, k + "_pred_head")(y)
                output_dict[k] = getattr(self, k + "_predictor")(feature_dict[k])
            return output_dict
        for i, hkeys in enumerate(self.hierarchy):
            if i == 0:
                x = y
            else:
                nchw, rows = self._stage_input(y, feature_dict, output_dict, self.hierarchy[i - 1])
                # the reference recomputes fusion_layers[i](y') once per key of the stage (model.py:761): same value
                if rows is not None:
                    B, _, hh, ww = y.shape
                    x = self.fusion_layers[i].forward_rows(rows, B, hh, ww)
                else:
                    x = self.fusion_layers[i](nchw)
            for k in hkeys:
                feature_dict[k] = getattr(self, k + "_pred_head")(x)
                output_dict[k] = getattr(self, k + "_predictor")(feature_dict[k])
        return output_dict

    def forward(self, features):
        """Inference branch of model.py:688-709 for callers that want the reference's full-resolution dict."""
        out = self.layers(features)
        s = self.common_stride
        up = lambda t: F.interpolate(t, scale_factor=s, mode="bilinear", align_corners=False)
        res = {k: up(v) for k, v in out.items()}
        res["offset"] = res["offset"] * s
        return res, {}


INS_EMBED_BRANCHES_REGISTRY = {"MaskRefinerInsEmbedHead": MaskRefinerInsEmbedHead}


def sem_seg_postprocess(result, img_size, output_height, output_width):
    """detectron2.modeling.postprocessing.sem_seg_postprocess (model.py:266-289)."""
    result = result[:, : img_size[0], : img_size[1]].expand(1, -1, -1, -1)
    return F.interpolate(result, size=(output_height, output_width), mode="bilinear", align_corners=False)[0]
