"""Config for the maskrefiner path: the reference's yacs keys, verbatim, without detectron2.

The reference assembles its config with detectron2's `get_cfg()` + `add_panoptic_deeplab_config` +
`add_mask_refiner_config` (maskrefiner/config.py:6-102, maskrefiner/predictor.py:211-215) and merges a YAML
that may inherit from another through `_BASE_`.  detectron2 is not vendored by the reference and is not
installable here, so this module restates only the keys the inference path reads, with the defaults those
three functions give them (detectron2 defaults quoted in SURVEY.md Appendix C), and reads the SAME YAML
files (e.g. configs/uoais-sim/instance-segmentation/seed777/mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8_new.yaml).
Unknown keys in a YAML are kept, so a config written for the reference loads unchanged.
"""
import copy
import os

import yaml


class CfgNode(dict):
    """Attribute-style nested dict (the subset of yacs.CfgNode the path uses)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    def __setattr__(self, k, v):
        self[k] = v

    @staticmethod
    def wrap(d):
        if isinstance(d, dict):
            return CfgNode({k: CfgNode.wrap(v) for k, v in d.items()})
        return d

    def merge(self, other):
        for k, v in other.items():
            if isinstance(v, dict) and isinstance(self.get(k), dict):
                self[k].merge(v)
            else:
                self[k] = CfgNode.wrap(copy.deepcopy(v))
        return self

    def merge_from_file(self, path):
        self.merge(_load_yaml_with_base(path))
        return self

    def merge_from_list(self, kv):
        for key, v in zip(kv[0::2], kv[1::2]):
            node = self
            parts = key.split(".")
            for p in parts[:-1]:
                node = node[p]
            node[parts[-1]] = v
        return self

    def clone(self):
        return CfgNode.wrap(copy.deepcopy(dict(self)))


def _load_yaml_with_base(path):
    with open(path) as f:
        d = yaml.safe_load(f) or {}
    base = d.pop("_BASE_", None)
    if base is not None:
        if not os.path.isabs(base):
            base = os.path.join(os.path.dirname(path), base)
        out = CfgNode.wrap(_load_yaml_with_base(base))
        out.merge(d)
        return out
    return CfgNode.wrap(d)


def get_cfg():
    """detectron2 defaults for the keys read on the path, then add_deeplab_config /
    add_panoptic_deeplab_config / add_mask_refiner_config (maskrefiner/config.py:6-102)."""
    c = CfgNode.wrap({
        "VERSION": 2,
        "SEED": -1,
        "MODEL": {
            "META_ARCHITECTURE": "GeneralizedRCNN",
            "DEVICE": "cuda",
            "WEIGHTS": "",
            "PIXEL_MEAN": [103.530, 116.280, 123.675],
            "PIXEL_STD": [1.0, 1.0, 1.0],
            "BACKBONE": {"NAME": "build_resnet_backbone", "FREEZE_AT": 2,
                         # add_mask_refiner_config
                         "FUSION_STRATEGY": "concat", "NUM_FUSION_LAYERS": 3, "WEIGHTS": "", "FREEZE_LAYERS": False},
            "RESNETS": {"DEPTH": 50, "OUT_FEATURES": ["res4"], "NUM_GROUPS": 1, "NORM": "FrozenBN", "WIDTH_PER_GROUP": 64,
                        "STRIDE_IN_1X1": True, "RES5_DILATION": 1, "RES2_OUT_CHANNELS": 256, "STEM_OUT_CHANNELS": 64,
                        "DEFORM_ON_PER_STAGE": [False, False, False, False], "DEFORM_MODULATED": False,
                        "DEFORM_NUM_GROUPS": 1,
                        # add_deeplab_config
                        "RES4_DILATION": 1, "RES5_MULTI_GRID": [1, 2, 4], "STEM_TYPE": "deeplab"},
            "SEM_SEG_HEAD": {"NAME": "SemSegFPNHead", "IN_FEATURES": ["p2", "p3", "p4", "p5"], "IGNORE_VALUE": 255,
                             "NUM_CLASSES": 54, "CONVS_DIM": 128, "COMMON_STRIDE": 4, "NORM": "GN", "LOSS_WEIGHT": 1.0,
                             # add_deeplab_config
                             "LOSS_TYPE": "hard_pixel_mining", "PROJECT_FEATURES": ["res2"], "PROJECT_CHANNELS": [48],
                             "ASPP_CHANNELS": 256, "ASPP_DILATIONS": [6, 12, 18], "ASPP_DROPOUT": 0.1,
                             "USE_DEPTHWISE_SEPARABLE_CONV": False,
                             # add_panoptic_deeplab_config
                             "HEAD_CHANNELS": 256, "LOSS_TOP_K": 0.2,
                             # add_maskformer2_config (maskrefiner/config.py:172,202-205): only consulted by the
                             # optional MSDeformAttn pixel decoder
                             "PIXEL_DECODER_NAME": "BasePixelDecoder", "MASK_DIM": 256, "TRANSFORMER_ENC_LAYERS": 0,
                             "DEFORMABLE_TRANSFORMER_ENCODER_IN_FEATURES": ["res3", "res4", "res5"],
                             "DEFORMABLE_TRANSFORMER_ENCODER_N_POINTS": 4, "DEFORMABLE_TRANSFORMER_ENCODER_N_HEADS": 8},
            # add_panoptic_deeplab_config
            "INS_EMBED_HEAD": {"NAME": "PanopticDeepLabInsEmbedHead", "IN_FEATURES": ["res2", "res3", "res5"],
                               "PROJECT_FEATURES": ["res2", "res3"], "PROJECT_CHANNELS": [32, 64], "ASPP_CHANNELS": 256,
                               "ASPP_DILATIONS": [6, 12, 18], "ASPP_DROPOUT": 0.1, "HEAD_CHANNELS": 32, "CONVS_DIM": 128,
                               "COMMON_STRIDE": 4, "NORM": "SyncBN", "CENTER_LOSS_WEIGHT": 200.0, "OFFSET_LOSS_WEIGHT": 0.01,
                               # add_mask_refiner_config
                               "FOREGROUND_LOSS_WEIGHT": 1.0, "FOREGROUND_LOSS_TYPE": "hard_pixel_mining",
                               "FOREGROUND_LOSS_TOP_K": 1.0, "EEE_MASK_ON": False, "EEE_POST_PROCESS_ON": False,
                               "EEE_MASK_LOSS_TYPE": "dice", "EEE_MASK_LOSS_WEIGHT": 1.0, "EEE_BOUNDARY_ON": True,
                               "EEE_BOUNDARY_LOSS_TYPE": "dice", "EEE_BOUNDARY_LOSS_WEIGHT": 1.0,
                               "HIERARCHICAL_FUSION_ON": False,
                               "HIERARCHY": [["eee_mask", "eee_boundary"], ["foreground", "center", "offset"]],
                               "NUM_FUSION_LAYERS": 3, "FUSION_STRATEGY": "concat", "FUSION_TARGET": ["feat", "pred"],
                               "ERROR_TYPE": "e3"},
            "PANOPTIC_DEEPLAB": {"STUFF_AREA": 2048, "CENTER_THRESHOLD": 0.1, "NMS_KERNEL": 7, "TOP_K_INSTANCE": 200,
                                 "PREDICT_INSTANCES": True, "USE_DEPTHWISE_SEPARABLE_CONV": False, "SIZE_DIVISIBILITY": -1,
                                 "BENCHMARK_NETWORK_SPEED": False},
        },
        "INPUT": {"MIN_SIZE_TEST": 800, "MAX_SIZE_TEST": 1333, "FORMAT": "BGR", "CROP": {"ENABLED": False},
                  "GAUSSIAN_SIGMA": 10,
                  # add_mask_refiner_config
                  "OFFSET_INPUT_ON": False, "DEPTH_ON": False, "RGB_ON": True, "DEPTH_RANGE": [2500, 15000]},
        "DATASETS": {"TRAIN": (), "TEST": ()},
    })
    return c


# uoais-sim metadata the post-processing reads (maskrefiner/data/datasets/register_uoais_sim_panoptic.py:184-198):
# one "thing" class (dataset id 1 -> contiguous id 0), label_divisor 1000, no stuff classes.
UOAIS_SIM_META = {"label_divisor": 1000, "thing_dataset_id_to_contiguous_id": {1: 0}}
