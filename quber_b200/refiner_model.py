"""Drop-in for the caller of the path, `eval/refiner_model.py::MaskRefiner` (:213-297): same constructor and the same
`predict(...) -> (refined_masks [N,H,W] bool ndarray, output dict, seconds, fg_mask)` contract, with arrays instead of
file paths (reading the PNG / NPY files stays with the caller, as does the optional LMFFNet foreground model).

What runs where:
  * depth normalisation (`eval/preprocess_utils.py:12-28`) — `normalize_depth` below, numpy, as the reference;
  * depth in-painting (`preprocess_utils.py:44-64`, cv2.inpaint TELEA) — NOT rebuilt: pass an in-painted depth image or a
    callable `inpaint=`; it is CPU OpenCV code on the caller's side of the boundary in the reference as well;
  * armbench resize rule `ResizeShortestEdge.get_output_shape(h, w, 800, 1333)` (:228) — `resize_shortest_edge_shape`;
  * the network + post-processing — `MaskRefinerPredictor.predict` (GPU);
  * the foreground-overlap filter `> 0.3` and the OCID zero-depth clear (:271-288) — `ops.filter_instances` (GPU,
    quber_filter_instances), bit-exact with the numpy loop.
"""
import time

import numpy as np
import torch

from . import ops
from .predictor import MaskRefinerPredictor

W, H = 640, 480                                   # eval/refiner_model.py:30 (module constants of the reference)


def normalize_depth(depth, min_val=250.0, max_val=1500.0):
    """eval/preprocess_utils.py:12-28: clip to [min, max] mm, scale to 0..255, uint8 (truncation), 3 channels.
    Unlike the reference it does not modify `depth` in place."""
    d = np.clip(np.asarray(depth), min_val, max_val)
    d = (d - min_val) / (max_val - min_val) * 255
    if d.ndim == 2:
        d = d[..., None]
    return np.uint8(np.repeat(d, 3, -1))


def resize_shortest_edge_shape(h, w, short=800, max_size=1333):
    """detectron2 ResizeShortestEdge.get_output_shape (eval/refiner_model.py:228): 1080x1920 -> 750x1333."""
    scale = short * 1.0 / min(h, w)
    newh, neww = (short, scale * w) if h < w else (scale * h, short)
    if max(newh, neww) > max_size:
        s = max_size * 1.0 / max(newh, neww)
        newh, neww = newh * s, neww * s
    return int(newh + 0.5), int(neww + 0.5)


class MaskRefiner:
    def __init__(self, config_file, weights_file=None, dataset="OSD", precision="fp32", fg_model=None, inpaint=None):
        """`fg_model(rgb, depth) -> [H,W] bool` replaces the reference's LMFFNet (:219, :269); `inpaint(depth_u8x3)`
        replaces preprocess_utils.inpaint_depth (:264)."""
        self.refiner_predictor = MaskRefinerPredictor(config_file, weights_file=weights_file, precision=precision)
        self.dataset = dataset
        self.fg_model = fg_model
        self.inpaint = inpaint

    def predict(self, rgb_img, depth_img, initial_masks, fg_mask=None, depth_is_metres=False):
        """rgb_img [H,W,3] uint8 (BGR as cv2.imread gives it); depth_img raw depth [H,W] (mm, or metres with
        `depth_is_metres`, the reference's '.npy' branch :258-259) or None (armbench, RGB only);
        initial_masks [N,H,W] bool | uint8.  Frames are expected at the working resolution (640x480, or the armbench
        shape from `resize_shortest_edge_shape`) — resizing image files is the caller's job, as in the reference."""
        initial_masks = np.asarray(initial_masks)
        if initial_masks.dtype == bool:
            initial_masks = np.uint8(initial_masks) * 255                                    # :236, :267
        zero_depth = None
        depth_u8 = None
        if depth_img is not None and self.dataset != "armbench":
            raw = np.asarray(depth_img)
            zero_depth = raw == 0                                                            # :255
            depth_u8 = normalize_depth(raw.astype(np.float32), 0.25, 1.5) if depth_is_metres else normalize_depth(raw.astype(np.float32))
            if self.inpaint is not None:
                depth_u8 = self.inpaint(depth_u8)
        start = time.time()
        output = self.refiner_predictor.predict(rgb_img, depth_u8, initial_masks)[0]
        if "instances" not in output:
            return [], output, time.time() - start, fg_mask
        masks = output["instances"].pred_masks                                               # [N,H,W] bool, device
        if self.dataset == "armbench":
            return masks.cpu().numpy(), output, time.time() - start, None                    # :239-246
        if self.fg_model is not None:
            fg_mask = self.fg_model(rgb_img, depth_img)                                      # :269
        if fg_mask is None and self.dataset != "OCID":
            return masks.cpu().numpy(), output, time.time() - start, fg_mask
        n = masks.shape[0]
        dev = masks.device
        ids = torch.arange(1, n + 1, device=dev, dtype=torch.int32)
        label_map = (masks.to(torch.int32) * ids[:, None, None]).amax(0, keepdim=True).contiguous()      # instances are disjoint
        counts = torch.tensor([n], dtype=torch.int32, device=dev)
        fg = None if fg_mask is None else torch.from_numpy(np.ascontiguousarray(fg_mask, dtype=np.uint8))[None].to(dev)
        zd = None
        if self.dataset == "OCID" and zero_depth is not None:
            zd = torch.from_numpy(np.ascontiguousarray(zero_depth, dtype=np.uint8))[None].to(dev)
        out_map, keep, new_counts = ops.filter_instances(label_map, counts, fg, zd, thr=0.3, max_per=max(n, 1))
        m = int(new_counts[0])
        refined = ops_masks_from_label_map(out_map[0], m).cpu().numpy()
        return refined, output, time.time() - start, fg_mask


def ops_masks_from_label_map(label_map, n):
    ids = torch.arange(1, int(n) + 1, device=label_map.device, dtype=label_map.dtype)
    return label_map[None] == ids[:, None, None]
