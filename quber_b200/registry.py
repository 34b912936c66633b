"""Registration of the path's entry points under the names the reference registers them with.

The reference reaches its modules through detectron2's registries:
  META_ARCH_REGISTRY       "MaskRefiner"                       maskrefiner/modeling/mask_refiner/model.py:74-75
  BACKBONE_REGISTRY        "build_resnet_deeplab_fusion_backbone", "build_resnet_deeplab_rgbd_fusion_backbone"
                                                               maskrefiner/modeling/backbone/resnet.py:358-359, 510-511
  INS_EMBED_BRANCHES       "MaskRefinerInsEmbedHead"           model.py:30, 461-462 (a registry local to the reference)
  SEM_SEG_HEADS_REGISTRY   "MSDeformAttnPixelDecoder"          MSMF/modeling/pixel_decoder/msdeformattn.py:164-165
and its compiled op through a module named `MultiScaleDeformableAttention` (OPS/src/vision.cpp:18-21,
OPS/functions/ms_deform_attn_func.py:21).

detectron2 is not vendored by the reference and is not installed in this image, so quber_b200 carries plain dict
registries of its own (`network.BACKBONE_REGISTRY`, `network.INS_EMBED_BRANCHES_REGISTRY`,
`meta_arch.META_ARCH_REGISTRY`, `pixel_decoder.SEM_SEG_HEADS_REGISTRY`) and `register()` ADDITIONALLY puts the same
objects into detectron2's registries iff `import detectron2` succeeds — `build_model(cfg)` with
`cfg.MODEL.META_ARCHITECTURE == "MaskRefiner"` then resolves to this implementation with no edit to the caller.
A name somebody registered before (the reference's own package imported in the same process) is left alone unless
`override=True`.
"""
import importlib


def _put(reg, name, obj, override):
    """detectron2 / fvcore Registry: `_obj_map` holds the entries and `register` asserts on duplicates."""
    existing = getattr(reg, "_obj_map", None)
    if existing is not None and name in existing:
        if not override:
            return False
        existing[name] = obj
        return True
    try:
        reg.register(obj) if getattr(obj, "__name__", None) == name else existing.__setitem__(name, obj)
    except AssertionError:
        return False
    return True


def register(override=False):
    """-> {registry name: [names registered]}; {} when detectron2 is not importable.  Idempotent."""
    from . import meta_arch, network, pixel_decoder
    from .msda import install_extension_module
    install_extension_module()
    try:
        d2 = importlib.import_module("detectron2.modeling")
    except Exception:
        return {}
    done = {}
    plan = [
        ("META_ARCH_REGISTRY", meta_arch.META_ARCH_REGISTRY),
        ("BACKBONE_REGISTRY", network.BACKBONE_REGISTRY),
        ("SEM_SEG_HEADS_REGISTRY", pixel_decoder.SEM_SEG_HEADS_REGISTRY),
    ]
    for reg_name, ours in plan:
        reg = getattr(d2, reg_name, None)
        if reg is None:
            continue
        done[reg_name] = [n for n, obj in ours.items() if _put(reg, n, obj, override)]
    # the reference's local INS_EMBED_BRANCHES registry lives in its own module; when that module is importable its
    # registry gets the entry too (model.py:30)
    try:
        ref = importlib.import_module("maskrefiner.modeling.mask_refiner.model")
        reg = getattr(ref, "INS_EMBED_BRANCHES_REGISTRY", None)
        if reg is not None:
            done["INS_EMBED_BRANCHES"] = [n for n, obj in network.INS_EMBED_BRANCHES_REGISTRY.items() if _put(reg, n, obj, override)]
    except Exception:
        pass
    return done
