"""GPU parity of the network-side kernels and of the drop-in meta-architecture / predictor.  B200 only."""
import os

import numpy as np
import pytest
import torch

import oracle
from quber_b200 import config as qcfg
from quber_b200 import ops, synth

pytestmark = pytest.mark.gpu
DEV = "cuda"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = os.path.join(ROOT, "configs", "mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8.yaml")


def _cfg(device):
    c = qcfg.get_cfg()
    c.merge_from_file(CFG)
    c.MODEL.DEVICE = device
    return c


def _randomise(model, seed=0):
    """Random-init predictors are N(0, 0.001) (model.py:418): logits would be ~0 and every comparison vacuous.
    Give predictors, norms and BatchNorm statistics non-trivial seeded values."""
    g = torch.Generator().manual_seed(seed)
    for n, m in model.named_modules():
        if isinstance(m, torch.nn.BatchNorm2d) or type(m).__name__ == "FrozenBatchNorm2d":
            m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=g) * 0.1)
            m.running_var.copy_(torch.rand(m.running_var.shape, generator=g) + 0.5)
        if n.endswith("_predictor.predictor"):
            m.weight.data.copy_(torch.randn(m.weight.shape, generator=g) * 0.05)
            m.bias.data.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)


def test_pack_input_bit_exact_with_torch():
    g = torch.Generator(device=DEV).manual_seed(0)
    img = torch.randint(0, 256, (3, 6, 48, 80), dtype=torch.uint8, device=DEV, generator=g)
    off = torch.randn(3, 3, 48, 80, device=DEV, generator=g)
    mean = torch.tensor([103.530, 116.280, 123.675, 127.5, 127.5, 127.5], device=DEV)
    std = torch.tensor([1.0, 57.375, 1.0, 2.0, 0.3, 1.0], device=DEV)
    out = ops.pack_input(img, off, mean, std)
    ref = torch.cat([(img - mean.view(-1, 1, 1)) / std.view(-1, 1, 1), off], dim=1)     # model.py:138,153
    assert torch.equal(out, ref)
    assert torch.equal(ops.pack_input(img, None, mean, std), ref[:, :6])


@pytest.mark.parametrize("hw", [(120, 160), (30, 37)])
def test_fuse_concat_matches_torch_chain(hw):
    h, w = hw
    g = torch.Generator(device=DEV).manual_seed(1)
    y = torch.randn(2, 128, h, w, device=DEV, generator=g)
    feat = torch.randn(2, 32, h, w, device=DEV, generator=g)
    eee = torch.randn(2, 4, h, w, device=DEV, generator=g) * 3
    fg = torch.randn(2, 1, h, w, device=DEV, generator=g) * 3
    ref = torch.cat([y, feat, eee.softmax(dim=1), fg.sigmoid()], dim=1)                 # model.py:745-759
    out = ops.fuse_concat([y, feat, eee, fg], ["copy", "copy", "softmax", "sigmoid"])
    assert out.shape == ref.shape
    assert torch.equal(out[:, :160], ref[:, :160])                                      # copies are exact
    torch.testing.assert_close(out, ref, rtol=0, atol=2e-6)                             # __expf vs expf
    rows = ops.fuse_concat([y, feat, eee, fg], ["copy", "copy", "softmax", "sigmoid"], layout="nhwc_f32", pitch=168)
    assert rows.shape == (2 * h * w, 168) and not rows[:, 165:].any()
    torch.testing.assert_close(rows[:, :165], ref.permute(0, 2, 3, 1).reshape(-1, 165), rtol=0, atol=2e-6)
    rb = ops.fuse_concat([y, feat, eee, fg], ["copy", "copy", "softmax", "sigmoid"], layout="nhwc_bf16", pitch=168)
    torch.testing.assert_close(rb[:, :165].float(), ref.permute(0, 2, 3, 1).reshape(-1, 165), rtol=2 ** -8, atol=1e-6)
    with pytest.raises(ValueError):
        ops.fuse_concat([y, feat[:, :, :-1]], ["copy", "copy"])


@pytest.fixture(scope="module")
def models():
    from quber_b200.meta_arch import build_model
    torch.manual_seed(0)
    cpu = build_model(_cfg("cpu")).eval()
    _randomise(cpu)
    gpu = build_model(_cfg("cuda")).eval()
    gpu.load_state_dict(cpu.state_dict())                    # never re-initialise independently
    return cpu, gpu


def test_network_logits_fp32_within_1e3_of_cpu_restatement(models):
    """North-star tolerance: float error logits within 1e-3 absolute in fp32.  Oracle = the same restated modules
    in plain torch fp32 on the CPU (no kernels); product = GPU with fuse_concat in the loop, TF32 off."""
    cpu, gpu = models
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.Generator().manual_seed(3)
    img = torch.randint(0, 256, (2, 6, 96, 128), dtype=torch.uint8, generator=g)
    m = np.stack([synth.ellipse_masks(4, 96, 128, seed=s, axes=(8, 24), margin=10) for s in range(2)])
    enc = torch.from_numpy(np.stack([oracle.encode_initial_masks(x) for x in m]))
    x_cpu = torch.cat([(img - cpu.pixel_mean) / cpu.pixel_std, enc], dim=1)
    with torch.no_grad():
        ref = cpu.network_logits(x_cpu)
        gpu.set_precision("fp32")
        x_gpu = ops.pack_input(img.to(DEV), enc.to(DEV), gpu.pixel_mean, gpu.pixel_std)
        assert torch.equal(x_gpu.cpu(), x_cpu)
        out = gpu.network_logits(x_gpu)
    for k in ref:
        assert ref[k].abs().max() > 0.05, k                                             # not a vacuous comparison
        torch.testing.assert_close(out[k].cpu(), ref[k], rtol=0, atol=1e-3)


def test_network_logits_bf16_mode_stated_tolerance(models):
    """C3: channels-last bf16 autocast for cuDNN + tcgen05 bf16 GEMM for the FusionLayers 1x1.  Stated tolerance:
    max abs error <= 6 % of the logit range of each head (bf16 has 8 mantissa bits; ~60 layers deep)."""
    cpu, gpu = models
    g = torch.Generator().manual_seed(4)
    x = torch.randn(2, 9, 96, 128, generator=g)
    x[:, :6] *= 60
    with torch.no_grad():
        ref = cpu.network_logits(x)
        gpu.set_precision("bf16")
        out = gpu.network_logits(x.to(DEV))
        gpu.set_precision("fp32")
    from quber_b200 import _lib
    assert _lib.lib().quber_gemm_timeouts() == 0
    for k in ref:
        rng = (ref[k].max() - ref[k].min()).item()
        err = (out[k].cpu() - ref[k]).abs().max().item()
        assert err <= 0.06 * rng, (k, err, rng)


def test_meta_arch_forward_matches_oracle_on_synthetic_heads(models, monkeypatch):
    """MaskRefiner.forward end to end (pack -> network -> fused post-processing -> output dicts) with the head
    outputs replaced by synthetic ones that contain instances (random-init weights give none)."""
    _, gpu = models
    H, W, B = 240, 320, 2
    ms = [synth.ellipse_masks(5, H, W, seed=10 + b, axes=(14, 40), margin=30) for b in range(B)]
    heads = [synth.head_outputs(m, seed=b) for b, m in enumerate(ms)]
    low = {k: torch.from_numpy(np.stack([h[k] for h in heads])).to(DEV) for k in heads[0]}
    monkeypatch.setattr(gpu, "network_logits", lambda x: low)
    inputs = [{"image": torch.zeros(6, H, W, dtype=torch.uint8), "initial_pred_offset": torch.zeros(3, H, W), "height": H, "width": W}
              for _ in range(B)]
    res = gpu(inputs)
    assert len(res) == B
    for b in range(B):
        lowb = np.concatenate([heads[b]["foreground"], heads[b]["center"], heads[b]["offset"], heads[b]["eee_boundary"]], 0)
        full = oracle.upsample_bilinear(lowb, 4)
        full[2:4] *= 4.0
        ref = oracle.post_process(full[0], full[1], full[2:4], full[4:8])
        r = res[b]
        assert set(r) == {"sem_seg", "eee_boundary", "panoptic_seg", "instances"}
        np.testing.assert_array_equal(r["sem_seg"].cpu().numpy()[0], full[0])
        np.testing.assert_array_equal(r["eee_boundary"].cpu().numpy(), full[4:8])
        np.testing.assert_array_equal(r["panoptic_seg"][0].cpu().numpy(), ref["panoptic"])
        inst = r["instances"]
        assert len(inst) == len(ref["labels"]) > 0 and inst.image_size == (H, W)
        np.testing.assert_array_equal(inst.pred_masks.cpu().numpy(), ref["masks"].astype(bool))
        np.testing.assert_array_equal(inst.pred_boxes.tensor.cpu().numpy(), ref["boxes"])
        np.testing.assert_allclose(inst.scores.cpu().numpy(), ref["scores"], rtol=1e-4, atol=1e-6)
        assert (inst.pred_classes == 0).all()


def test_predictor_predict_api_and_encoding():
    from quber_b200.predictor import MaskRefinerPredictor
    torch.manual_seed(0)
    p = MaskRefinerPredictor(CFG)
    H, W = 480, 640
    rgb, depth = synth.rgbd_frame(H, W, seed=0)
    masks = synth.ellipse_masks(20, H, W, seed=0)                                       # config C1
    enc = p.encode(masks)[0].cpu().numpy()
    np.testing.assert_array_equal(enc, oracle.encode_initial_masks(masks))
    out = p.predict(rgb, depth, masks)
    assert isinstance(out, list) and len(out) == 1
    o = out[0]
    assert tuple(o["sem_seg"].shape) == (1, H, W) and tuple(o["eee_boundary"].shape) == (4, H, W)
    assert tuple(o["panoptic_seg"][0].shape) == (H, W) and o["panoptic_seg"][1] is None
    # random-init predictors: centre map ~0 < threshold -> no centre.  The only possible "instance" is then the
    # reference's stuff paste of a large enough foreground region (post_processing.py:152-160, label 1000 -> class 0)
    if "instances" in o:
        assert len(o["instances"]) == 1 and int(o["instances"].pred_classes[0]) == 0
        assert set(np.unique(o["panoptic_seg"][0].cpu().numpy())) <= {-1.0, 1000.0}
    outs = p.predict_batch([rgb, rgb], [depth, depth], [masks, masks[:3]])
    assert len(outs) == 2
    torch.testing.assert_close(outs[0]["sem_seg"], o["sem_seg"], rtol=0, atol=1e-4)
    # armbench-shaped frame: width not a multiple of 16 / 8
    m2 = synth.ellipse_masks(3, 120, 150, seed=2, axes=(8, 20), margin=12)
    e2 = p.encode(m2)[0].cpu().numpy()
    r2 = oracle.encode_initial_masks(m2)
    np.testing.assert_array_equal(e2[:2], r2[:2])
    np.testing.assert_allclose(e2[2], r2[2], rtol=2e-7, atol=0)


# ------------------------------------------------------------------ A15: MSDeformAttn pixel decoder
def _cpu_encoder_layer(layer, src, pos, ref, shapes, starts):
    """msdeformattn.py:122-131 + ms_deform_attn.py:82-125 on the CPU, the sampling done by the oracle."""
    a = layer.self_attn
    N, S, C = src.shape
    M, L, P = a.n_heads, a.n_levels, a.n_points
    q = src + pos
    value = a.value_proj(src).view(N, S, M, C // M)
    off = a.sampling_offsets(q).view(N, S, M, L, P, 2)
    aw = torch.softmax(a.attention_weights(q).view(N, S, M, L * P), -1).view(N, S, M, L, P)
    norm = torch.stack([shapes[..., 1], shapes[..., 0]], -1)
    loc = ref[:, :, None, :, None, :] + off / norm[None, None, None, :, None, :]
    samp = torch.from_numpy(oracle.msda_forward(value.numpy(), shapes.numpy(), starts.numpy(), loc.contiguous().numpy(),
                                                aw.contiguous().numpy()))
    src = layer.norm1(src + a.output_proj(samp))
    return layer.norm2(src + layer.linear2(torch.relu(layer.linear1(src))))


def test_encoder_layer_fp32_tf32_bf16_against_cpu_restatement():
    from quber_b200.pixel_decoder import MSDeformAttnTransformerEncoder, MSDeformAttnTransformerEncoderLayer
    torch.manual_seed(0)
    layer = MSDeformAttnTransformerEncoderLayer(256, 1024, 0.0, "relu", 3, 8, 4).eval()
    with torch.no_grad():                                    # non-trivial offsets / attention (both weights are 0 at init)
        layer.self_attn.sampling_offsets.weight.normal_(0, 0.02)
        layer.self_attn.attention_weights.weight.normal_(0, 0.05)
    hs = [(15, 20), (8, 10), (4, 5)]
    S = sum(h * w for h, w in hs)
    shapes = torch.tensor(hs, dtype=torch.long)
    starts = torch.cat((shapes.new_zeros((1,)), shapes.prod(1).cumsum(0)[:-1]))
    g = torch.Generator().manual_seed(1)
    src, pos = torch.randn(2, S, 256, generator=g), torch.randn(2, S, 256, generator=g) * 0.5
    ref = MSDeformAttnTransformerEncoder.get_reference_points(hs, torch.ones(2, 3, 2), "cpu")
    torch.backends.cuda.matmul.allow_tf32 = False
    with torch.no_grad():
        want = _cpu_encoder_layer(layer, src, pos, ref, shapes, starts)
        gl = layer.to(DEV)
        args = (src.to(DEV), pos.to(DEV), ref.to(DEV), shapes.to(DEV), starts.to(DEV))
        got = {}
        for prec in ("fp32", "tf32", "bf16"):
            gl.precision = prec
            got[prec] = gl(*args).cpu()
    torch.testing.assert_close(got["fp32"], want, rtol=0, atol=1e-3)        # north-star fp32 tolerance (measured ~1e-5)
    torch.testing.assert_close(got["tf32"], want, rtol=0, atol=2e-2)        # 10-bit mantissa products, LayerNorm output O(1)
    torch.testing.assert_close(got["bf16"], want, rtol=0, atol=1e-1)        # stated bf16 tolerance (8-bit mantissa)
    assert (got["bf16"] - want).abs().mean() < 1.5e-2
    from quber_b200 import _lib
    assert _lib.lib().quber_gemm_timeouts() == 0


def test_pixel_decoder_forward_features_shapes_and_precisions():
    from quber_b200.pixel_decoder import MSDeformAttnPixelDecoder
    shapes = {"res2": (256, 4), "res3": (512, 8), "res4": (1024, 16), "res5": (2048, 32)}
    torch.manual_seed(0)
    pd = MSDeformAttnPixelDecoder(shapes, transformer_enc_layers=6).eval().to(DEV)
    g = torch.Generator(device=DEV).manual_seed(2)
    feats = {k: torch.randn(2, c, 480 // s, 640 // s, device=DEV, generator=g) for k, (c, s) in shapes.items()}
    outs = {}
    for prec in ("fp32", "tf32", "bf16"):
        pd.set_precision(prec)
        mask_features, low, multi = pd.forward_features(feats)
        assert tuple(mask_features.shape) == (2, 256, 120, 160) and tuple(low.shape) == (2, 256, 15, 20)
        assert [tuple(m.shape[2:]) for m in multi] == [(15, 20), (30, 40), (60, 80)]       # S = 6300 tokens
        outs[prec] = mask_features
    torch.testing.assert_close(outs["tf32"], outs["fp32"], rtol=0, atol=5e-2)
    assert (outs["bf16"] - outs["fp32"]).abs().mean() < 5e-2


# ------------------------------------------------------------------ GroupNorm + ReLU (csrc/norm.cu)
@pytest.mark.parametrize("C,H,W", [(32, 24, 40), (64, 30, 40), (128, 24, 40), (256, 12, 20), (512, 15, 20), (2048, 6, 8), (96, 8, 12)])
@pytest.mark.parametrize("cl", [False, True])
def test_group_norm_relu_matches_float64_reference(C, H, W, cl):
    """nn.GroupNorm(32, C) + ReLU against a float64 torch reference: fp32 <= 2e-5 abs on unit-scale activations (torch's own
    fp32 kernel is at the same distance), NCHW and channels-last, with a mean far from zero in half of the channels."""
    g = torch.Generator(device=DEV).manual_seed(C + H)
    x = torch.randn(3, C, H, W, device=DEV, generator=g) * 1.7
    x[:, ::2] += 25.0                                                   # |mean| >> std inside every group
    gn = torch.nn.GroupNorm(32, C).to(DEV)
    gn.weight.data.uniform_(0.5, 1.5, generator=g); gn.bias.data.normal_(0, 0.5, generator=g)
    if cl:
        x = x.contiguous(memory_format=torch.channels_last)
    if not ops.group_norm_supported(x, 32):
        assert C == 96 and cl is True or (H * W) % 4 != 0
        with pytest.raises(ValueError):
            ops.group_norm_relu(x, gn.weight, gn.bias, 32)
        return
    ref = torch.relu(torch.nn.functional.group_norm(x.double(), 32, gn.weight.double(), gn.bias.double(), gn.eps))
    out = ops.group_norm_relu(x, gn.weight, gn.bias, 32, gn.eps)
    assert out.dtype == x.dtype and out.is_contiguous(memory_format=torch.channels_last if cl else torch.contiguous_format)
    assert float((out.double() - ref).abs().max()) <= 2e-5
    lin = ops.group_norm_relu(x, gn.weight, gn.bias, 32, gn.eps, relu=False)
    ref_lin = torch.nn.functional.group_norm(x.double(), 32, gn.weight.double(), gn.bias.double(), gn.eps)
    assert float((lin.double() - ref_lin).abs().max()) <= 2e-5
    assert torch.equal(out, ops.group_norm_relu(x, gn.weight, gn.bias, 32, gn.eps))              # deterministic
    y = x.clone(memory_format=torch.preserve_format)
    assert ops.group_norm_relu(y, gn.weight, gn.bias, 32, gn.eps, inplace=True) is y and torch.equal(y, out)
    # bf16 storage: same arithmetic on the bf16-rounded input, one bf16 rounding of the result
    xb = x.bfloat16()
    refb = torch.relu(torch.nn.functional.group_norm(xb.double(), 32, gn.weight.double(), gn.bias.double(), gn.eps))
    if ops.group_norm_supported(xb, 32):
        outb = ops.group_norm_relu(xb, gn.weight, gn.bias, 32, gn.eps)
        assert outb.dtype == torch.bfloat16
        assert float((outb.double() - refb).abs().max()) <= 2.0 ** -8 * float(refb.abs().max()) + 1e-3


def test_group_norm_relu_full_size_and_many_tiles():
    """C2 shapes (several tiles per sample): 128 x 120 x 160 channels-last bf16 and 256 x 120 x 160 NCHW fp32."""
    g = torch.Generator(device=DEV).manual_seed(9)
    for C, cl, dt, tol in [(128, True, torch.bfloat16, 2.5e-2), (256, False, torch.float32, 2e-5)]:
        x = (torch.randn(2, C, 120, 160, device=DEV, generator=g) + 0.7).to(dt)
        if cl:
            x = x.contiguous(memory_format=torch.channels_last)
        w = torch.rand(C, device=DEV, generator=g) + 0.5
        b = torch.randn(C, device=DEV, generator=g)
        ref = torch.relu(torch.nn.functional.group_norm(x.double(), 32, w.double(), b.double(), 1e-5))
        out = ops.group_norm_relu(x, w, b, 32, 1e-5)
        assert float((out.double() - ref).abs().max()) <= tol


def test_network_inference_plan_matches_layer_by_layer():
    """Folded BN + fused conv epilogues + the GroupNorm kernel against the layer-by-layer network (same weights), fp32."""
    from quber_b200.meta_arch import build_model
    torch.manual_seed(0)
    model = build_model(_cfg(DEV), "fp32").eval()
    _randomise(model, seed=3)
    x = torch.randn(2, 9, 96, 128, device=DEV)
    with torch.no_grad():
        model.use_inference_plan = False
        ref = model.network_logits(x)
        model.use_inference_plan = True
        out = model.network_logits(x)
    for k in ref:
        assert float((out[k] - ref[k]).abs().max()) <= 1e-3 * max(1.0, float(ref[k].abs().max())), k


def test_group_norm_pre_bias_equals_bias_then_norm():
    g = torch.Generator(device=DEV).manual_seed(21)
    for cl in (False, True):
        x = torch.randn(2, 256, 12, 20, device=DEV, generator=g)
        if cl:
            x = x.contiguous(memory_format=torch.channels_last)
        pb = torch.randn(256, device=DEV, generator=g) * 3
        w = torch.rand(256, device=DEV, generator=g) + 0.5
        b = torch.randn(256, device=DEV, generator=g)
        ref = torch.relu(torch.nn.functional.group_norm((x + pb.view(1, -1, 1, 1)).double(), 32, w.double(), b.double(), 1e-5))
        out = ops.group_norm_relu(x, w, b, 32, 1e-5, pre_bias=pb)
        assert float((out.double() - ref).abs().max()) <= 2e-5


@pytest.mark.parametrize("dt,tol", [(torch.float32, 2e-6), (torch.bfloat16, 2.0 ** -7)])
@pytest.mark.parametrize("shape", [((2, 64, 60, 80), (2, 256, 30, 40)), ((1, 32, 24, 40), (1, 128, 12, 20)), ((2, 0, 47, 83), (2, 16, 12, 21))])
def test_upsample_concat_matches_torch(dt, tol, shape):
    """cat([proj_x, F.interpolate(y, size=proj_x.shape[-2:], bilinear, align_corners=False)], 1) — the decoder step
    (DeepLabV3PlusHead.layers) — against torch on the same channels-last tensors; non-integer scale in the last case."""
    (N, Ca, H, W), (_, Cb, h, w) = shape
    g = torch.Generator(device=DEV).manual_seed(H + Cb)
    b = torch.randn(N, Cb, h, w, device=DEV, generator=g).to(dt).contiguous(memory_format=torch.channels_last)
    a = torch.randn(N, Ca, H, W, device=DEV, generator=g).to(dt).contiguous(memory_format=torch.channels_last) if Ca else None
    up = torch.nn.functional.interpolate(b.float(), size=(H, W), mode="bilinear", align_corners=False)
    ref = up if a is None else torch.cat([a.float(), up], dim=1)
    out = ops.upsample_concat(a, b, size=(H, W))
    assert out.shape == ref.shape and out.dtype == dt and out.is_contiguous(memory_format=torch.channels_last)
    if a is not None:
        assert torch.equal(out[:, :Ca], a)
    assert float((out.float() - ref).abs().max()) <= tol * max(1.0, float(ref.abs().max()))
    with pytest.raises(ValueError):
        ops.upsample_concat(a, b.contiguous(), size=(H, W))              # NCHW operand


def test_network_cuda_graph_replay_matches_eager():
    from quber_b200.meta_arch import build_model
    torch.manual_seed(0)
    model = build_model(_cfg(DEV), "bf16").eval()
    _randomise(model, seed=4)
    xs = [torch.randn(2, 9, 64, 96, device=DEV) for _ in range(3)]
    with torch.no_grad():
        eager = [model.network_logits(x) for x in xs]
        model.use_cuda_graph = True
        graphed = [model.network_logits(x) for x in xs]          # first call captures, the others replay
        again = model.network_logits(xs[0])
    for e, g in zip(eager, graphed):
        for k in e:
            assert torch.equal(e[k], g[k]), k
    for k in again:
        assert torch.equal(again[k], eager[0][k])


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(2, 64, 48, 64), (1, 128, 37, 51), (2, 8, 5, 7)])
def test_maxpool3x3s2_bit_exact_with_torch(dt, shape):
    g = torch.Generator(device=DEV).manual_seed(sum(shape))
    x = torch.randn(*shape, device=DEV, generator=g).to(dt).contiguous(memory_format=torch.channels_last)
    ref = torch.nn.functional.max_pool2d(x, kernel_size=3, stride=2, padding=1)
    out = ops.maxpool3x3s2(x)
    assert out.shape == ref.shape and out.is_contiguous(memory_format=torch.channels_last)
    assert torch.equal(out, ref)
    with pytest.raises(ValueError):
        ops.maxpool3x3s2(x.contiguous())


def test_fuse_concat_rows_matches_torch_chain_on_channels_last_sources():
    g = torch.Generator(device=DEV).manual_seed(11)
    cl = lambda t: t.contiguous(memory_format=torch.channels_last)
    y = cl(torch.randn(2, 128, 30, 37, device=DEV, generator=g).bfloat16())
    feat = cl(torch.randn(2, 32, 30, 37, device=DEV, generator=g).bfloat16())
    eee = cl(torch.randn(2, 4, 30, 37, device=DEV, generator=g) * 3)                     # float32 logits
    fg = cl(torch.randn(2, 1, 30, 37, device=DEV, generator=g).bfloat16() * 3)
    ref = torch.cat([y.float(), feat.float(), eee.softmax(dim=1), fg.float().sigmoid()], dim=1)      # model.py:745-759
    ref_rows = ref.permute(0, 2, 3, 1).reshape(-1, 165)
    rows = ops.fuse_concat_rows([y, feat, eee, fg], ["copy", "copy", "softmax", "sigmoid"], pitch=168)
    assert rows.shape == (2 * 30 * 37, 168) and rows.dtype == torch.bfloat16
    torch.testing.assert_close(rows[:, :165].float(), ref_rows, rtol=2 ** -8, atol=1e-6)
    assert torch.equal(rows[:, :160], torch.cat([y, feat], 1).permute(0, 2, 3, 1).reshape(-1, 160))   # copies are exact
    assert not rows[:, 165:].any()
    r32 = ops.fuse_concat_rows([y, feat, eee, fg], ["copy", "copy", "softmax", "sigmoid"], out_dtype=torch.float32)
    torch.testing.assert_close(r32, ref_rows, rtol=1e-5, atol=2e-6)
    with pytest.raises(ValueError):
        ops.fuse_concat_rows([y.contiguous(), feat], ["copy", "copy"])


@pytest.mark.parametrize("C", [128, 256, 768, 1024])
def test_add_layer_norm_matches_float64_reference(C):
    g = torch.Generator(device=DEV).manual_seed(C)
    a = torch.randn(3, 701, C, device=DEV, generator=g) * 2 + 5.0
    b = torch.randn(3, 701, C, device=DEV, generator=g)
    ln = torch.nn.LayerNorm(C).to(DEV)
    ln.weight.data.uniform_(0.5, 1.5, generator=g); ln.bias.data.normal_(0, 0.5, generator=g)
    ref = torch.nn.functional.layer_norm((a + b).double(), (C,), ln.weight.double(), ln.bias.double(), ln.eps)
    y, y16 = ops.add_layer_norm(a, b, ln.weight, ln.bias, ln.eps, want_bf16=True)
    assert float((y.double() - ref).abs().max()) <= 5e-6
    assert torch.equal(y16, y.bfloat16())
    y1 = ops.add_layer_norm(a, None, ln.weight, ln.bias, ln.eps)
    ref1 = torch.nn.functional.layer_norm(a.double(), (C,), ln.weight.double(), ln.bias.double(), ln.eps)
    assert float((y1.double() - ref1).abs().max()) <= 5e-6
    torch_y = ln(a + b)
    assert float((y - torch_y).abs().max()) <= 5e-6                          # and as close to torch's own kernel
    with pytest.raises(ValueError):
        ops.add_layer_norm(a[..., :100].contiguous(), None, ln.weight[:100], ln.bias[:100])


@pytest.mark.parametrize("qd", [torch.float32, torch.bfloat16])
def test_add_layer_norm_emits_next_query(qd):
    """q = LayerNorm(a + b) + pos (with_pos_embed of the next layer, msdeformattn.py:124) from the same pass: the fp32 sum
    is the one torch makes, the bf16 variant its rounding."""
    g = torch.Generator(device=DEV).manual_seed(5)
    a, b = torch.randn(2, 333, 256, device=DEV, generator=g), torch.randn(2, 333, 256, device=DEV, generator=g)
    pos = torch.randn(2, 333, 256, device=DEV, generator=g) * 0.5
    ln = torch.nn.LayerNorm(256).to(DEV)
    y0 = ops.add_layer_norm(a, b, ln.weight, ln.bias, ln.eps)
    y, y16, q = ops.add_layer_norm(a, b, ln.weight, ln.bias, ln.eps, want_bf16=True, pos=pos, q_dtype=qd)
    assert torch.equal(y, y0) and torch.equal(y16, y.bfloat16())
    assert q.dtype == qd and torch.equal(q, (y + pos).to(qd))
    y2, q2 = ops.add_layer_norm(a, b, ln.weight, ln.bias, ln.eps, pos=pos, q_dtype=qd)
    assert torch.equal(y2, y0) and torch.equal(q2, q)
    with pytest.raises(ValueError):
        ops.add_layer_norm(a, b, ln.weight, ln.bias, ln.eps, pos=pos[:, :10].contiguous(), q_dtype=qd)


@pytest.mark.parametrize("prec", ["tf32", "bf16"])
def test_encoder_chained_operands_equal_layer_by_layer(prec):
    """The encoder hands each layer the (x, q) pair its predecessor's last LayerNorm wrote; the result is bit-identical to
    calling the layers one by one (each making its own casts)."""
    from quber_b200.pixel_decoder import MSDeformAttnTransformerEncoder, MSDeformAttnTransformerEncoderLayer
    torch.manual_seed(3)
    layer = MSDeformAttnTransformerEncoderLayer(256, 1024, 0.0, "relu", 3, 8, 4).eval()
    enc = MSDeformAttnTransformerEncoder(layer, 3).eval()
    with torch.no_grad():
        for l in enc.layers:
            l.self_attn.sampling_offsets.weight.normal_(0, 0.02)
            l.self_attn.attention_weights.weight.normal_(0, 0.05)
            l.precision = prec
    enc = enc.to(DEV)
    hs = [(15, 20), (8, 10), (4, 5)]
    S = sum(h * w for h, w in hs)
    shapes = torch.tensor(hs, dtype=torch.long, device=DEV)
    starts = torch.cat((shapes.new_zeros((1,)), shapes.prod(1).cumsum(0)[:-1]))
    g = torch.Generator(device=DEV).manual_seed(4)
    src, pos = torch.randn(2, S, 256, device=DEV, generator=g), torch.randn(2, S, 256, device=DEV, generator=g) * 0.5
    ones = torch.ones(2, 3, 2, device=DEV)
    with torch.no_grad():
        got = enc(src, shapes, starts, ones, pos, None, host_shapes=hs)
        ref_pts = enc.get_reference_points(hs, ones, DEV)
        out = src
        for l in enc.layers:
            out = l(out, pos, ref_pts, shapes, starts, None)
    assert torch.equal(got, out)


def test_msda_prepare_matches_module_arithmetic():
    """OPS/modules/ms_deform_attn.py:102-109 restated with torch on the same GEMM output."""
    g = torch.Generator(device=DEV).manual_seed(8)
    N, Q, M, L, P = 2, 333, 8, 3, 4
    oa = torch.randn(N, Q, M * L * P * 3, device=DEV, generator=g)
    ref_pts = torch.rand(N, Q, L, 2, device=DEV, generator=g)
    shapes = torch.tensor([[15, 20], [30, 40], [60, 80]], dtype=torch.long, device=DEV)
    loc, aw = ops.msda_prepare(oa, ref_pts, shapes, M, L, P)
    off = oa[..., : M * L * P * 2].reshape(N, Q, M, L, P, 2)
    aw_ref = torch.softmax(oa[..., M * L * P * 2:].reshape(N, Q, M, L * P), -1).view(N, Q, M, L, P)
    norm = torch.stack([shapes[..., 1], shapes[..., 0]], -1)
    loc_ref = ref_pts[:, :, None, :, None, :] + off / norm[None, None, None, :, None, :]
    assert torch.equal(loc, loc_ref)                                          # one division and one add per element: exact
    torch.testing.assert_close(aw, aw_ref, rtol=2e-6, atol=1e-8)
    torch.testing.assert_close(aw.sum((-1, -2)), torch.ones(N, Q, M, device=DEV), rtol=0, atol=1e-6)
    # generic (scalar) route: P != 4
    P2 = 3
    oa2 = torch.randn(N, Q, M * L * P2 * 3, device=DEV, generator=g)
    loc2, aw2 = ops.msda_prepare(oa2, ref_pts, shapes, M, L, P2)
    off2 = oa2[..., : M * L * P2 * 2].reshape(N, Q, M, L, P2, 2)
    assert torch.equal(loc2, ref_pts[:, :, None, :, None, :] + off2 / norm[None, None, None, :, None, :])
    torch.testing.assert_close(aw2, torch.softmax(oa2[..., M * L * P2 * 2:].reshape(N, Q, M, L * P2), -1).view(N, Q, M, L, P2),
                               rtol=2e-6, atol=1e-8)


def test_maskrefiner_with_msdeformattn_pixel_decoder_end_to_end(tmp_path):
    """PIXEL_DECODER_NAME: "MSDeformAttnPixelDecoder" (maskrefiner/config.py:172): the head consumes the deformable-
    attention decoder's stride-4 mask features; forward runs on the repo's kernels in fp32 (3xTF32) and bf16."""
    import os
    from quber_b200.config import get_cfg
    from quber_b200.meta_arch import build_model
    base = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "configs",
                        "mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8.yaml")
    y = tmp_path / "msda.yaml"
    y.write_text(f"""_BASE_: "{base}"
MODEL:
  SEM_SEG_HEAD:
    PIXEL_DECODER_NAME: "MSDeformAttnPixelDecoder"
    CONVS_DIM: 256
    TRANSFORMER_ENC_LAYERS: 2
""")
    cfg = get_cfg()
    cfg.merge_from_file(str(y))
    torch.manual_seed(0)
    model = build_model(cfg).eval()
    with torch.no_grad():                                     # non-trivial offsets (zero weights at init)
        for l in model.ins_embed_head.pixel_decoder.transformer.encoder.layers:
            l.self_attn.sampling_offsets.weight.normal_(0, 0.02)
            l.self_attn.attention_weights.weight.normal_(0, 0.05)
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.randn(2, 9, 256, 320, device=DEV, generator=g)
    outs = {}
    for prec in ("fp32", "bf16"):
        model.set_precision(prec)
        with torch.no_grad():
            o = model.network_logits(x)
        assert set(o) == {"eee_boundary", "foreground", "center", "offset"}
        assert tuple(o["eee_boundary"].shape) == (2, 4, 64, 80) and tuple(o["offset"].shape) == (2, 2, 64, 80)
        assert all(torch.isfinite(v).all() for v in o.values())
        outs[prec] = o
    for k in outs["fp32"]:
        scale = float(outs["fp32"][k].abs().max()) + 1e-6
        assert float((outs["bf16"][k] - outs["fp32"][k]).abs().max()) <= 0.15 * scale + 5e-3, k
    # the whole predictor-level forward (post-processing included) runs too
    inp = [{"image": torch.randint(0, 255, (6, 256, 320), dtype=torch.uint8), "initial_pred_offset": torch.zeros(3, 256, 320),
            "height": 256, "width": 320} for _ in range(2)]
    res = model(inp)
    assert len(res) == 2 and tuple(res[0]["sem_seg"].shape) == (1, 256, 320) and "panoptic_seg" in res[0]


def test_split_rgbd_input_equals_the_two_cats():
    """quber_split_rgbd_input: cat([rgb, offsets], 1) / cat([depth, offsets], 1) of RGBDFusionBackbone.forward
    (resnet.py:489-493) as channels-last bfloat16 in one pass — the same bytes as cat + cast + layout change."""
    from quber_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(77)
    for B, H, W in ((2, 48, 64), (3, 30, 44), (1, 480, 640)):
        x = torch.randn(B, 9, H, W, device="cuda", generator=g) * 3
        a, b = ops.split_rgbd_input(x)
        for got, img in ((a, x[:, :3]), (b, x[:, 3:6])):
            want = torch.cat([img, x[:, 6:]], 1).to(torch.bfloat16)
            assert got.shape == want.shape and got.dtype == torch.bfloat16
            assert got.is_contiguous(memory_format=torch.channels_last)
            assert torch.equal(got, want)
        a8, b8 = ops.split_rgbd_input(x, pad8=True)
        assert a8.shape == (B, 8, H, W) and torch.equal(a8[:, :6], a) and torch.equal(b8[:, :6], b)
        assert float(a8[:, 6:].abs().max()) == 0.0 and float(b8[:, 6:].abs().max()) == 0.0
    with pytest.raises(ValueError):
        ops.split_rgbd_input(torch.zeros(1, 8, 4, 4, device="cuda"))
