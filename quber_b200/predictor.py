"""`MaskRefinerPredictor` — drop-in for maskrefiner/predictor.py:207-359.

Same constructor (`config_file, dataset_name, weights_file`) and the same
`predict(rgb_img HxWx3 uint8, depth_img HxWx3 uint8, perturbed_masks [N,H,W] uint8) -> list[dict]`
(one dict, the meta-architecture's output for the frame).  Differences, all deliberate:
  * the per-mask NumPy loop that encodes the initial masks (predictor.py:304-357) runs as
    quber_encode_initial_masks on the GPU (bit-exact with the loop: tests/test_gpu_parity.py);
  * the model is run under torch.no_grad() (the reference keeps autograd on, predictor.py:358 — wasted memory);
  * MODEL.WEIGHTS is honoured; the reference overwrites it with a hard-coded absolute path
    (predictor.py:226) and builds an unused dataset loader (predictor.py:237-241) — neither is reproduced;
  * `predict_batch` refines many frames in one pass (the reference is batch 1 by construction).
"""
import os

import numpy as np
import torch

from . import ops
from .config import get_cfg
from .meta_arch import build_model


class MaskRefinerPredictor:
    def __init__(self, config_file, dataset_name="uoais_sim_val_panoptic", weights_file=None, precision="fp32", device=None):
        self.cfg = get_cfg()
        self.cfg.merge_from_file(config_file)
        if device is not None:
            self.cfg.MODEL.DEVICE = device
        self.model = build_model(self.cfg, precision)
        self.model.eval()
        self.dataset_name = dataset_name
        if weights_file is not None:
            path = weights_file if os.path.isabs(weights_file) or os.path.exists(weights_file) else \
                config_file.replace(".yaml", "/{}".format(weights_file)).replace("configs", "output")   # predictor.py:221-224
            self.load_weights(path)
        self.input_format = self.cfg.INPUT.FORMAT
        assert self.input_format in ["RGB", "BGR"], self.input_format
        self.sigma = 10                                                  # predictor.py:246
        self.depth_on = self.cfg.INPUT.DEPTH_ON
        self.rgb_on = self.cfg.INPUT.RGB_ON

    def load_weights(self, path):
        """A detectron2 checkpoint ({"model": state_dict, ...}) or a bare state_dict; keys as the reference saves them."""
        ck = torch.load(path, map_location="cpu", weights_only=False)
        sd = ck["model"] if isinstance(ck, dict) and "model" in ck else ck
        sd = {k: (torch.as_tensor(v) if not isinstance(v, torch.Tensor) else v) for k, v in sd.items()}
        missing, unexpected = self.model.load_state_dict(sd, strict=False)
        missing = [k for k in missing if not k.endswith("num_batches_tracked")]
        if missing or unexpected:
            raise RuntimeError(f"checkpoint does not match the reference layout: missing {missing[:5]}, unexpected {unexpected[:5]}")

    # ---- predictor.py:290-302: pack the frame ---------------------------------------------------------
    def _image(self, rgb_img, depth_img):
        img = torch.as_tensor(np.ascontiguousarray(rgb_img.transpose(2, 0, 1)))
        if not self.rgb_on and self.depth_on:
            img = torch.as_tensor(np.ascontiguousarray(depth_img.transpose(2, 0, 1)))
        if self.rgb_on and self.depth_on:
            img = torch.cat((img, torch.as_tensor(np.ascontiguousarray(depth_img.transpose(2, 0, 1)))), dim=0)
        return img

    # ---- predictor.py:304-357 on the GPU ---------------------------------------------------------------
    def encode(self, perturbed_masks):
        """[B,K,H,W] (or [K,H,W]) uint8 masks -> [B,3,H,W] float32 (centre heat-map, offset_y / H, offset_x / W)."""
        m = torch.as_tensor(np.ascontiguousarray(perturbed_masks)) if not isinstance(perturbed_masks, torch.Tensor) else perturbed_masks
        if m.dim() == 3:
            m = m[None]
        m = m.to(self.model.device).to(torch.uint8)
        B, K, H, W = m.shape
        if K == 0:
            return torch.zeros((B, 3, H, W), dtype=torch.float32, device=m.device)
        return ops.encode_initial_masks_any_width(m.contiguous(), self.sigma)

    def predict_batch(self, rgb_imgs, depth_imgs, perturbed_masks_list):
        """Frames of equal size: lists of HxWx3 uint8 RGB / depth images and [N_i,H,W] masks -> list of output dicts."""
        K = max((len(m) for m in perturbed_masks_list), default=0)
        H, W = rgb_imgs[0].shape[:2]
        masks = np.zeros((len(rgb_imgs), max(K, 1), H, W), np.uint8)
        for b, m in enumerate(perturbed_masks_list):
            if len(m):
                masks[b, : len(m)] = np.asarray(m) != 0                  # empty (padding) masks are skipped by the encoder
        enc = self.encode(masks)
        inputs = []
        for b, (rgb, dep) in enumerate(zip(rgb_imgs, depth_imgs)):
            inputs.append({"height": H, "width": W, "image": self._image(rgb, dep), "initial_pred_offset": enc[b]})
        return self.model(inputs)

    def predict(self, rgb_img, depth_img=None, perturbed_masks=None):
        height, width = rgb_img.shape[:2]
        inputs = {"height": height, "width": width, "image": self._image(rgb_img, depth_img)}
        pm = np.zeros((0, height, width), np.uint8) if perturbed_masks is None else perturbed_masks
        inputs["initial_pred_offset"] = self.encode(pm)[0]
        return self.model([inputs])
