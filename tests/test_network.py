"""Host-side mirror of the reference's config / registry / module layout (CPU): the YAML loader, the
state_dict key layout a reference checkpoint has, and the shapes of the head outputs."""
import os

import pytest
import torch

from quber_b200 import config as qcfg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = os.path.join(ROOT, "configs", "mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8.yaml")
REF_CFG = "/root/reference/configs/uoais-sim/instance-segmentation/seed777/mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8_new.yaml"


def _cfg():
    c = qcfg.get_cfg()
    c.merge_from_file(CFG)
    c.MODEL.DEVICE = "cpu"
    return c


def test_config_defaults_and_merge():
    c = qcfg.get_cfg()
    assert c.MODEL.PANOPTIC_DEEPLAB.CENTER_THRESHOLD == 0.1 and c.MODEL.INS_EMBED_HEAD.NORM == "SyncBN"   # config.py:79,93
    assert c.MODEL.RESNETS.STEM_TYPE == "deeplab" and c.MODEL.RESNETS.RES5_MULTI_GRID == [1, 2, 4]
    c = _cfg()
    assert c.MODEL.PANOPTIC_DEEPLAB.CENTER_THRESHOLD == 0.3 and c.MODEL.INS_EMBED_HEAD.NORM == "GN"
    assert c.MODEL.INS_EMBED_HEAD.HIERARCHY == [["eee_boundary"], ["foreground", "center", "offset"]]
    assert c.MODEL.SEM_SEG_HEAD.USE_DEPTHWISE_SEPARABLE_CONV is False      # the key the head actually reads (model.py:637)
    c.merge_from_list(["MODEL.PANOPTIC_DEEPLAB.TOP_K_INSTANCE", 50])
    assert c.MODEL.PANOPTIC_DEEPLAB.TOP_K_INSTANCE == 50


@pytest.mark.skipif(not os.path.exists(REF_CFG), reason="reference checkout not present (GPU box)")
def test_reference_yaml_with_base_loads_unchanged():
    ours, ref = _cfg(), qcfg.get_cfg()
    ref.merge_from_file(REF_CFG)                                            # _BASE_: ../Base-Mask-Refiner.yaml
    for sect in ("BACKBONE", "RESNETS", "INS_EMBED_HEAD", "PANOPTIC_DEEPLAB"):
        for k, v in ours.MODEL[sect].items():
            assert ref.MODEL[sect][k] == v, (sect, k, ref.MODEL[sect][k], v)
    assert list(ref.MODEL.PIXEL_MEAN) == list(ours.MODEL.PIXEL_MEAN)
    assert ref.INPUT.OFFSET_INPUT_ON and ref.INPUT.DEPTH_ON and ref.SEED == 777


@pytest.fixture(scope="module")
def model():
    from quber_b200.meta_arch import build_model
    torch.manual_seed(0)
    return build_model(_cfg()).eval()


def test_state_dict_layout_matches_reference_checkpoints(model):
    sd = model.state_dict()
    expect = {
        "backbone.rgb_backbone.stem.conv1.weight": (32, 6, 3, 3),
        "backbone.rgb_backbone.stem.conv3.norm.running_var": (64,),
        "backbone.rgb_backbone.res2.0.shortcut.weight": (256, 64, 1, 1),
        "backbone.rgb_backbone.res5.2.conv2.weight": (512, 512, 3, 3),
        "backbone.depth_backbone.stem.conv1.weight": (32, 6, 3, 3),
        "backbone.depth_backbone.depth_res2.0.conv1.norm.running_mean": (64,),          # resnet.py:169 prefix
        "backbone.depth_backbone.depth_res5.0.shortcut.norm.weight": (2048,),
        "backbone.fusion_res2.conv.weight": (256, 512, 1, 1),                           # resnet.py:473-485
        "backbone.fusion_res2.gn1.weight": (256,),
        "backbone.fusion_res3.conv1.weight": (512, 512, 3, 3),
        "backbone.fusion_res5.conv.bias": (2048,),
        "ins_embed_head.decoder.res5.project_conv.convs.0.weight": (256, 2048, 1, 1),
        "ins_embed_head.decoder.res5.project_conv.convs.3.norm.weight": (256,),
        "ins_embed_head.decoder.res5.project_conv.convs.4.1.bias": (256,),
        "ins_embed_head.decoder.res5.project_conv.project.weight": (256, 1280, 1, 1),
        "ins_embed_head.decoder.res3.project_conv.weight": (64, 512, 1, 1),
        "ins_embed_head.decoder.res3.fuse_conv.0.weight": (128, 320, 3, 3),
        "ins_embed_head.decoder.res2.fuse_conv.1.norm.bias": (128,),
        "ins_embed_head.eee_boundary_pred_head.head.1.weight": (32, 128, 3, 3),
        "ins_embed_head.eee_boundary_predictor.predictor.weight": (4, 32, 1, 1),
        "ins_embed_head.offset_predictor.predictor.bias": (2,),
        "ins_embed_head.fusion_layers_1.fusion_layers.0.weight": (128, 164, 1, 1),      # 128 + 32 + 4 (model.py:595-608)
        "ins_embed_head.fusion_layers_1.fusion_layers.3.norm.running_mean": (128,),
    }
    for k, shape in expect.items():
        assert k in sd, k
        assert tuple(sd[k].shape) == shape, (k, tuple(sd[k].shape))
    assert "backbone.fusion_res5.conv0.weight" not in sd                                # resnet.py:480-481
    assert not any(k.startswith("ins_embed_head.eee_mask") for k in sd)                 # EEE_MASK_ON False
    assert "pixel_mean" not in sd                                                       # non-persistent buffer (model.py:90)


def test_head_output_shapes_and_res5_stride(model):
    x = torch.randn(1, 9, 64, 96)
    feats = model.backbone(x)
    assert {k: tuple(v.shape[1:]) for k, v in feats.items()} == {"res2": (256, 16, 24), "res3": (512, 8, 12), "res5": (2048, 4, 6)}
    out = model.network_logits(x)
    assert {k: tuple(v.shape) for k, v in out.items()} == {"eee_boundary": (1, 4, 16, 24), "foreground": (1, 1, 16, 24),
                                                           "center": (1, 1, 16, 24), "offset": (1, 2, 16, 24)}
    full, _ = model.ins_embed_head(model.backbone(x))
    assert tuple(full["offset"].shape) == (1, 2, 64, 96)
    torch.testing.assert_close(full["offset"], torch.nn.functional.interpolate(out["offset"], scale_factor=4, mode="bilinear",
                                                                               align_corners=False) * 4)


def test_fusionlayers_bn_folding_is_exact_algebra(model):
    fl = model.ins_embed_head.fusion_layers[1]
    torch.manual_seed(1)
    bn = fl.fusion_layers[0].norm
    bn.running_mean.normal_(0, 0.1); bn.running_var.uniform_(0.5, 1.5); bn.weight.data.uniform_(0.5, 1.5); bn.bias.data.normal_()
    x = torch.randn(2, 164, 5, 7)
    ref = fl.fusion_layers[0](x)
    w, b = fl.folded_1x1(168, torch.float32)
    rows = torch.zeros(2 * 5 * 7, 168)
    rows[:, :164] = x.permute(0, 2, 3, 1).reshape(-1, 164)
    got = torch.relu(rows @ w.T + b).view(2, 5, 7, 128).permute(0, 3, 1, 2)
    torch.testing.assert_close(got, ref, rtol=1e-4, atol=1e-5)


def test_cpu_forward_refuses_postprocessing(model):
    inp = {"image": torch.zeros(6, 32, 48, dtype=torch.uint8), "initial_pred_offset": torch.zeros(3, 32, 48), "height": 32, "width": 48}
    with pytest.raises(RuntimeError, match="no CPU post-processing"):
        model([inp])


def test_pixel_decoder_layout_and_position_embedding():
    from quber_b200.pixel_decoder import MSDeformAttnPixelDecoder, PositionEmbeddingSine
    shapes = {"res2": (256, 4), "res3": (512, 8), "res4": (1024, 16), "res5": (2048, 32)}
    torch.manual_seed(0)
    pd = MSDeformAttnPixelDecoder(shapes, transformer_enc_layers=2)
    sd = pd.state_dict()
    expect = {                                                                           # msdeformattn.py:205-287
        "input_proj.0.0.weight": (256, 2048, 1, 1), "input_proj.2.1.weight": (256,),
        "transformer.level_embed": (3, 256),
        "transformer.encoder.layers.1.self_attn.sampling_offsets.weight": (192, 256),
        "transformer.encoder.layers.0.self_attn.attention_weights.bias": (96,),
        "transformer.encoder.layers.0.self_attn.value_proj.weight": (256, 256),
        "transformer.encoder.layers.0.linear1.weight": (1024, 256), "transformer.encoder.layers.0.norm2.bias": (256,),
        "adapter_1.weight": (256, 256, 1, 1), "layer_1.norm.weight": (256,), "mask_features.bias": (256,),
    }
    for k, s in expect.items():
        assert k in sd and tuple(sd[k].shape) == s, k
    assert pd.num_fpn_levels == 1 and pd.transformer_in_features == ["res3", "res4", "res5"]
    # cached position embedding == the reference formula evaluated with an explicit all-False mask
    pe = PositionEmbeddingSine(128, normalize=True)
    x = torch.zeros(2, 8, 15, 20)
    a = pe(x)
    b = pe(x, mask=torch.zeros(2, 15, 20, dtype=torch.bool))
    assert a.shape == (2, 256, 15, 20) and torch.equal(a, b) and torch.equal(pe(x), a)
    # reference points of the encoder: (i + 0.5) / dim, x first (msdeformattn.py:141-153)
    from quber_b200.pixel_decoder import MSDeformAttnTransformerEncoder
    rp = MSDeformAttnTransformerEncoder.get_reference_points([(2, 3), (1, 2)], torch.ones(1, 2, 2), "cpu")
    assert rp.shape == (1, 8, 2, 2)
    torch.testing.assert_close(rp[0, 4, 0], torch.tensor([1.5 / 3, 1.5 / 2]))
    # the CUDA-only op refuses CPU tensors like the reference's AT_ERROR
    feats = {k: torch.randn(1, c, 64 // s, 96 // s) for k, (c, s) in shapes.items()}
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        pd.forward_features(feats)


def test_caller_side_helpers_match_reference_values():
    """normalize_depth (eval/preprocess_utils.py:12-28) on values whose expected bytes were produced by the reference's
    own function in the build container, and the armbench resize rule (eval/refiner_model.py:228)."""
    import numpy as np
    from quber_b200.refiner_model import normalize_depth, resize_shortest_edge_shape
    d = np.array([[0, 100, 250, 251.3, 600.7, 875, 1234.56, 1499.9, 1500, 1501, 65535]], np.float32)
    out = normalize_depth(d)
    assert out.shape == (1, 11, 3) and out.dtype == np.uint8
    assert out[0, :, 0].tolist() == [0, 0, 0, 0, 71, 127, 200, 254, 255, 255, 255] and (out[..., 0] == out[..., 2]).all()
    assert d[0, 0] == 0                                                   # the input is left untouched
    d2 = np.array([[0, 0.1, 0.25, 0.4, 0.9, 1.2, 1.4999, 1.5, 9]], np.float32)
    assert normalize_depth(d2, 0.25, 1.5)[0, :, 0].tolist() == [0, 0, 0, 30, 132, 193, 254, 255, 255]
    assert resize_shortest_edge_shape(1080, 1920) == (750, 1333)
    assert resize_shortest_edge_shape(480, 640) == (800, 1067)


def test_msdeformattn_fresh_module_initialises_like_the_reference():
    """A seeded fresh MSDeformAttn equals the reference module built under the same seed, parameter for parameter
    (OPS/modules/ms_deform_attn.py:34-80, executed from /root/reference when it is there — the build container)."""
    import os
    import sys
    import types
    ops_dir = "/root/reference/ext_modules/UnseenObjectsWithMeanShift/MSMFormer/meanshiftformer/modeling/pixel_decoder/ops"
    if not os.path.exists(ops_dir):
        pytest.skip("reference checkout not present (GPU box)")
    from quber_b200.msda import MSDeformAttn
    sys.modules.setdefault("MultiScaleDeformableAttention", types.ModuleType("MultiScaleDeformableAttention"))
    src = open(ops_dir + "/modules/ms_deform_attn.py").read().replace(
        "from ..functions import MSDeformAttnFunction", "MSDeformAttnFunction = None").replace(
        "from ..functions.ms_deform_attn_func import ms_deform_attn_core_pytorch", "ms_deform_attn_core_pytorch = None")
    ns = {}
    exec(compile(src, ops_dir + "/modules/ms_deform_attn.py", "exec"), ns)
    for args in ((256, 3, 8, 4), (64, 2, 4, 4), (256, 4, 8, 4)):
        torch.manual_seed(5)
        ref = ns["MSDeformAttn"](*args)
        torch.manual_seed(5)
        ours = MSDeformAttn(*args)
        assert list(ref.state_dict()) == list(ours.state_dict())
        for k, v in ref.state_dict().items():
            assert torch.equal(v, ours.state_dict()[k]), k


def test_registers_with_detectron2_when_importable(monkeypatch):
    """registration 'iff import detectron2 succeeds'.  detectron2 is absent from this image, so a
    stand-in package with fvcore-style registries is put on sys.modules; the reference's names must resolve to this
    implementation, and a name that is already taken is left alone unless override=True."""
    import sys
    import types
    import quber_b200
    from quber_b200 import meta_arch, network, pixel_decoder

    class Reg:                                       # fvcore.common.registry.Registry, as far as registration goes
        def __init__(self, name):
            self._name, self._obj_map = name, {}

        def register(self, obj=None):
            assert obj.__name__ not in self._obj_map, "An object named '{}' was already registered".format(obj.__name__)
            self._obj_map[obj.__name__] = obj
            return obj

        def get(self, name):
            return self._obj_map[name]

    d2, mod = types.ModuleType("detectron2"), types.ModuleType("detectron2.modeling")
    mod.META_ARCH_REGISTRY, mod.BACKBONE_REGISTRY, mod.SEM_SEG_HEADS_REGISTRY = Reg("META_ARCH"), Reg("BACKBONE"), Reg("SEM_SEG_HEADS")
    d2.modeling = mod
    taken = object()
    mod.BACKBONE_REGISTRY._obj_map["build_resnet_deeplab_fusion_backbone"] = taken
    monkeypatch.setitem(sys.modules, "detectron2", d2)
    monkeypatch.setitem(sys.modules, "detectron2.modeling", mod)
    monkeypatch.delitem(sys.modules, "MultiScaleDeformableAttention", raising=False)
    done = quber_b200.register()
    assert done["META_ARCH_REGISTRY"] == ["MaskRefiner"] and mod.META_ARCH_REGISTRY.get("MaskRefiner") is meta_arch.MaskRefiner
    assert mod.BACKBONE_REGISTRY.get("build_resnet_deeplab_rgbd_fusion_backbone") is network.build_resnet_deeplab_rgbd_fusion_backbone
    assert mod.BACKBONE_REGISTRY.get("build_resnet_deeplab_fusion_backbone") is taken          # not clobbered
    assert mod.SEM_SEG_HEADS_REGISTRY.get("MSDeformAttnPixelDecoder") is pixel_decoder.MSDeformAttnPixelDecoder
    assert quber_b200.register()["META_ARCH_REGISTRY"] == []                                   # idempotent
    quber_b200.register(override=True)
    assert mod.BACKBONE_REGISTRY.get("build_resnet_deeplab_fusion_backbone") is network.build_resnet_deeplab_fusion_backbone
    import MultiScaleDeformableAttention as MSDA                                               # OPS/functions/ms_deform_attn_func.py:21
    assert callable(MSDA.ms_deform_attn_forward) and callable(MSDA.ms_deform_attn_backward)


def test_model_builds_with_msdeformattn_pixel_decoder_from_yaml(tmp_path):
    """MODEL.SEM_SEG_HEAD.PIXEL_DECODER_NAME: "MSDeformAttnPixelDecoder" (maskrefiner/config.py:172, :202-205) selects the
    deformable-attention pixel decoder in the head builder."""
    import os
    from quber_b200.config import get_cfg
    from quber_b200.meta_arch import build_model
    from quber_b200.pixel_decoder import MSDeformAttnPixelDecoder
    base = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "configs",
                        "mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8.yaml")
    y = tmp_path / "msda.yaml"
    y.write_text(f"""_BASE_: "{base}"
MODEL:
  DEVICE: "cpu"
  SEM_SEG_HEAD:
    PIXEL_DECODER_NAME: "MSDeformAttnPixelDecoder"
    CONVS_DIM: 256
    TRANSFORMER_ENC_LAYERS: 2
    DEFORMABLE_TRANSFORMER_ENCODER_IN_FEATURES: ["res3", "res4", "res5"]
    DEFORMABLE_TRANSFORMER_ENCODER_N_POINTS: 4
    DEFORMABLE_TRANSFORMER_ENCODER_N_HEADS: 8
""")
    cfg = get_cfg()
    cfg.merge_from_file(str(y))
    model = build_model(cfg)
    pd = model.ins_embed_head.pixel_decoder
    assert isinstance(pd, MSDeformAttnPixelDecoder)
    assert pd.transformer_in_features == ["res3", "res5"]                  # the RGB-D fusion backbone has no res4
    assert pd.mask_dim == cfg.MODEL.INS_EMBED_HEAD.CONVS_DIM and len(pd.transformer.encoder.layers) == 2
    sd = model.state_dict()
    assert "ins_embed_head.pixel_decoder.transformer.encoder.layers.1.self_attn.sampling_offsets.weight" in sd
    assert not any(k.startswith("ins_embed_head.decoder.") for k in sd)    # the DeepLabV3+ decoder is not built
    assert "ins_embed_head.fusion_layers_1.fusion_layers.0.weight" in sd   # the hierarchy of heads is unchanged
    # default config: DeepLabV3+ decoder, no pixel decoder
    cfg2 = get_cfg()
    cfg2.merge_from_file(base)
    cfg2.MODEL.DEVICE = "cpu"
    m2 = build_model(cfg2)
    assert m2.ins_embed_head.pixel_decoder is None and any(k.startswith("ins_embed_head.decoder.") for k in m2.state_dict())


def test_inpaint_depth_equals_the_reference_function():
    """eval/preprocess_utils.py:44-64 executed from /root/reference (build container) against refiner_model.inpaint_depth."""
    import os
    import importlib.util
    path = "/root/reference/eval/preprocess_utils.py"
    if not os.path.exists(path):
        pytest.skip("reference checkout not present (GPU box)")
    import numpy as np
    from quber_b200.refiner_model import inpaint_depth, normalize_depth
    spec = importlib.util.spec_from_file_location("ref_preprocess_utils", path)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    rng = np.random.default_rng(2)
    d = rng.uniform(200, 1600, (120, 160)).astype(np.float32)
    d[rng.random((120, 160)) < 0.06] = 0
    a = normalize_depth(d)
    assert np.array_equal(a, ref.normalize_depth(d.copy()))
    assert np.array_equal(inpaint_depth(a), ref.inpaint_depth(a.copy()))
