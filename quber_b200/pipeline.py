"""Batched driver of the maskrefiner hot path: the call a user of the reference's
`MaskRefinerPredictor.predict` / `eval/refiner_model.py::MaskRefiner.predict` makes, for a whole batch
of frames, with every stage on the GPU and no host synchronisation between stages.

Stages (SURVEY.md §8-a row ids):
  A1  encode initial masks        -> ops.encode_initial_masks      (predictor.py:304-357)
  A13 per-instance boundaries      -> ops.mask_boundary             (explicit_error_estimation/util.py:62-99)
  A14 MSDeformAttn sampling        -> ops.msda_forward              (OPS ms_deform_attn_forward), optional decoder
  A6-A12 up-sample, decode, group, fuse -> ops.post_process         (model.py:260-356, post_processing.py)
The network between A1 and A6 (backbone / DeepLabV3+ decoder / heads: cuDNN) is a
callable the caller supplies (`head_fn`); bench.py feeds synthetic head outputs instead because
random-init predictors give no instance (SURVEY §8-d hazard).
"""
from dataclasses import dataclass, field

import numpy as np
import torch

from . import ops


@dataclass
class RefinerConfig:
    """The config keys the path reads (maskrefiner/config.py:51-102, Base-Mask-Refiner.yaml)."""
    center_threshold: float = 0.3      # MODEL.PANOPTIC_DEEPLAB.CENTER_THRESHOLD
    nms_kernel: int = 7                # MODEL.PANOPTIC_DEEPLAB.NMS_KERNEL
    top_k: int = 200                   # MODEL.PANOPTIC_DEEPLAB.TOP_K_INSTANCE
    stuff_area: int = 2048             # MODEL.PANOPTIC_DEEPLAB.STUFF_AREA
    min_area: int = 512                # post_processing.py:145 (hard-coded in the reference)
    label_divisor: int = 1000          # register_uoais_sim_panoptic.py:184
    common_stride: int = 4             # MODEL.SEM_SEG_HEAD.COMMON_STRIDE
    sigma: int = 10                    # predictor.py:246
    boundary_ratio: float = 0.01       # masks_to_boundary default (util.py:92)
    max_instances_per_image: int = 64
    want: tuple = ("label_map", "masks", "eee_onehot")


class HotPath:
    def __init__(self, cfg=None, device="cuda"):
        self.cfg = cfg or RefinerConfig()
        self.device = torch.device(device)
        self.launches_per_step = 0

    # --- resident: every input already on the device -------------------------------------------
    def encode_and_boundary(self, masks):
        c = self.cfg
        B, K, H, W = masks.shape
        enc = ops.encode_initial_masks(masks, c.sigma)
        d = ops.boundary_dilation(H, W, c.boundary_ratio)
        per, uni, fg = ops.mask_boundary(masks, d)
        return enc, per, uni, fg

    def post(self, head):
        c = self.cfg
        B = head["foreground"].shape[0]
        return ops.post_process(head["foreground"], head["center"], head["offset"], head.get("eee_boundary"),
                                scale=c.common_stride, threshold=c.center_threshold, nms_kernel=c.nms_kernel,
                                top_k=c.top_k, min_area=c.min_area, stuff_area=c.stuff_area,
                                label_divisor=c.label_divisor, max_instances=B * c.max_instances_per_image, want=c.want)

    def msda_layers(self, msda, n_layers=6):
        out = None
        for _ in range(n_layers):
            out = ops.msda_forward(msda["value"], msda["shapes"], msda["starts"], msda["loc"], msda["attn"])
        return out

    # kernels launched per call (for bench.py's gpu_launches; memsets are not kernels)
    LAUNCHES = dict(encode=2, boundary=1, post=7, msda_layer=1)

    # --- end to end: host buffers in, host results out ------------------------------------------
    def refine_batch_host(self, masks_host, head_host, out_host=None):
        """The end-to-end call: host buffers in, host results out, everything on the current stream.

        masks_host: pinned uint8 [B,K,H,W]; head_host: dict of pinned float32 head outputs (what the
        network hands to post-processing).  Returns (host dict, device results): label_map [B,H,W] int32
        (0 = void, j+1 = j-th refined instance of the frame — the lossless compact form of the
        reference's pred_masks), per-instance labels / scores / boxes / offsets and the union
        boundary [B,H,W] uint8 of the initial masks.  The host tensors are pinned and reused between
        calls; they are valid after the caller synchronises the stream.
        """
        dev = self.device
        masks = masks_host.to(dev, non_blocking=True)
        head = {k: v.to(dev, non_blocking=True) for k, v in head_host.items()}
        enc, per, uni, fg = self.encode_and_boundary(masks)
        r = self.post(head)
        src = dict(label_map=r["label_map"], inst_offsets=r["inst_offsets"], labels=r["labels"], scores=r["scores"],
                   boxes=r["boxes"], boundary=uni)
        if out_host is None:
            out_host = getattr(self, "_out_host", None)
        if out_host is None or any(out_host[k].shape != v.shape for k, v in src.items()):
            out_host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True) for k, v in src.items()}
            self._out_host = out_host
        for k, v in src.items():
            out_host[k].copy_(v, non_blocking=True)
        return out_host, (enc, per, r)
