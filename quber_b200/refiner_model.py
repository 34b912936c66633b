"""Drop-in for the caller of the path, `eval/refiner_model.py::MaskRefiner` (:213-297): same constructor and the same
`predict(rgb_path, depth_path, initial_masks, fg_mask) -> (refined_masks [N,H,W] bool ndarray, output dict, seconds,
fg_mask)` contract.  `rgb_path` / `depth_path` may be file paths (as in the reference: cv2.imread, np.load for '.npy',
16-bit PNG otherwise) or the arrays themselves.

What runs where:
  * file reading, cv2.resize of the frame (:247-262) — host, OpenCV, exactly the reference's calls;
  * depth normalisation (`eval/preprocess_utils.py:12-28`) — `normalize_depth` below, numpy, as the reference;
  * depth in-painting (`preprocess_utils.py:44-64`, cv2.inpaint TELEA) — `inpaint_depth` below: the reference's own
    sequence of OpenCV calls on the host (fast-marching in-painting is a serial priority-queue algorithm on the caller's
    side of the boundary; not rebuilt as a kernel).  `inpaint=` replaces it, `inpaint=False` skips it;
  * the armbench mask resize, cv2.INTER_NEAREST per mask (:231) — `ops.resize_nearest` (GPU, bit-exact with OpenCV);
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


def inpaint_depth(depth, factor=1, kernel_size=3, dilate=True):
    """eval/preprocess_utils.py:44-64: fill the zero pixels of a normalised [H,W,3] uint8 depth image with OpenCV's TELEA
    in-painting (host side, the reference's own call sequence)."""
    import cv2
    Hd, Wd, _ = depth.shape
    small = cv2.resize(depth, (Wd // factor, Hd // factor))
    mask = np.all(small == 0, axis=2).astype(np.uint8)
    if dilate:
        mask = cv2.dilate(mask, np.ones((kernel_size, kernel_size), np.uint8), iterations=1)
    filled = cv2.resize(cv2.inpaint(small, mask, kernel_size, cv2.INPAINT_TELEA), (Wd, Hd))
    return np.where(depth == 0, filled, depth)


def _read_depth(path):
    """:249-252: np.load for '.npy' (metres), imageio.imread otherwise (16-bit PNG, millimetres); OpenCV reads the same
    PNG samples with IMREAD_UNCHANGED."""
    if "npy" in path:
        return np.load(path), True
    import cv2
    d = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if d is None:
        raise FileNotFoundError(path)
    return d, False


class MaskRefiner:
    def __init__(self, config_file, weights_file=None, dataset="OSD", precision="fp32", fg_model=None, inpaint=None):
        """`fg_model(rgb, depth) -> [H,W] bool` replaces the reference's LMFFNet (:219, :269; a separate foreground
        network, out of this path's scope); `inpaint(depth_u8x3)` replaces `inpaint_depth` (:264), False skips it."""
        self.refiner_predictor = MaskRefinerPredictor(config_file, weights_file=weights_file, precision=precision)
        self.dataset = dataset
        self.fg_model = fg_model
        self.inpaint = inpaint_depth if inpaint is None else (inpaint or None)

    def _prepare(self, rgb_path, depth_path, initial_masks, depth_is_metres=False):
        """The host side in front of the predictor (eval/refiner_model.py:226-236, :247-267): read, resize, normalise and
        in-paint with the reference's own OpenCV calls.  -> (rgb_img, depth_u8 | None, initial_masks, zero_depth | None)."""
        rgb_img, depth_img = rgb_path, depth_path
        from_files = isinstance(rgb_path, str)
        initial_masks = np.asarray(initial_masks)
        if initial_masks.dtype == bool:
            initial_masks = np.uint8(initial_masks) * 255                                    # :229-230, :266-267
        if from_files:
            import cv2
            rgb_img = cv2.imread(rgb_path)
            if rgb_img is None:
                raise FileNotFoundError(rgb_path)
            if self.dataset == "armbench":
                h, w = resize_shortest_edge_shape(rgb_img.shape[0], rgb_img.shape[1], 800, 1333)          # :228
                rgb_img = cv2.resize(rgb_img, (w, h))
                if len(initial_masks) and initial_masks.shape[1:] != (h, w):                  # :231, cv2.INTER_NEAREST per mask
                    initial_masks = ops.resize_nearest(torch.from_numpy(np.ascontiguousarray(initial_masks)).cuda(), (h, w))
                depth_img = None
            else:
                rgb_img = cv2.resize(rgb_img, (W, H))                                        # :253
                if isinstance(depth_path, str):
                    depth_img, depth_is_metres = _read_depth(depth_path)
        zero_depth = None
        depth_u8 = None
        if depth_img is not None and self.dataset != "armbench":
            raw = np.asarray(depth_img)
            zero_depth = raw == 0                                                            # :255
            depth_u8 = normalize_depth(raw.astype(np.float32), 0.25, 1.5) if depth_is_metres else normalize_depth(raw.astype(np.float32))
            if from_files:
                import cv2
                depth_u8 = cv2.resize(depth_u8, (W, H), interpolation=cv2.INTER_NEAREST)     # :262
            if self.inpaint is not None:
                depth_u8 = self.inpaint(depth_u8)                                            # :264
        return rgb_img, depth_u8, initial_masks, zero_depth

    def _finish(self, output, start, rgb_path, depth_path, fg_mask, zero_depth):
        """The host side behind the predictor (:239-246, :269-297): foreground-overlap filter and OCID zero-depth clear on
        the device."""
        if "instances" not in output:
            return [], output, time.time() - start, fg_mask
        masks = output["instances"].pred_masks                                               # [N,H,W] bool, device
        if self.dataset == "armbench":
            return masks.cpu().numpy(), output, time.time() - start, None                    # :239-246
        if self.fg_model is not None:
            fg_mask = self.fg_model(rgb_path, depth_path)                                    # :269 (lmffnet.predict takes the paths)
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

    def predict(self, rgb_path, depth_path, initial_masks, fg_mask=None, depth_is_metres=False):
        """eval/refiner_model.py:224-297.  rgb_path: file path or [H,W,3] uint8 BGR array; depth_path: file path, raw
        depth array [H,W] (mm; metres with `depth_is_metres` or a '.npy' path, :258-259) or None (armbench, RGB only);
        initial_masks [N,h,w] bool | uint8.  Files are read and the frame resized as the reference does (:226-236,
        :247-262); arrays are taken as they come (already at the working resolution)."""
        rgb_img, depth_u8, initial_masks, zero_depth = self._prepare(rgb_path, depth_path, initial_masks, depth_is_metres)
        start = time.time()
        output = self.refiner_predictor.predict(rgb_img, depth_u8, initial_masks)[0]
        return self._finish(output, start, rgb_path, depth_path, fg_mask, zero_depth)

    def predict_many(self, rgb_paths, depth_paths, initial_masks_list, fg_masks=None, depth_is_metres=False, workers=8):
        """`predict` for a list of frames — what the reference's evaluation loop does one frame at a time
        (eval/eval_utils.py:360-392).  The host work in front of the network (file reads, resizes, depth normalisation and
        TELEA in-painting: OpenCV releases the GIL) runs on `workers` threads, frames of one size go through the network and
        the hot path as ONE batch (`MaskRefinerPredictor.predict_batch`), the caller-side filter runs per frame.  Returns
        the list of `predict`'s tuples, in order; the results are the per-frame call's."""
        n = len(rgb_paths)
        depth_paths = [None] * n if depth_paths is None else depth_paths
        fg_masks = [None] * n if fg_masks is None else fg_masks
        if self.dataset == "armbench":                          # masks come back as device tensors of per-frame size: frame by frame
            return [self.predict(r, d, m, f, depth_is_metres) for r, d, m, f in zip(rgb_paths, depth_paths, initial_masks_list, fg_masks)]
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max(1, int(workers))) as ex:
            prep = list(ex.map(lambda a: self._prepare(a[0], a[1], a[2], depth_is_metres), zip(rgb_paths, depth_paths, initial_masks_list)))
        results = [None] * n
        groups = {}
        for i, (rgb, dep, _, _) in enumerate(prep):
            groups.setdefault((rgb.shape[:2], dep is None), []).append(i)
        for idx in groups.values():
            start = time.time()
            outs = self.refiner_predictor.predict_batch([prep[i][0] for i in idx], [prep[i][1] for i in idx], [prep[i][2] for i in idx])
            for i, out in zip(idx, outs):
                results[i] = self._finish(out, start, rgb_paths[i], depth_paths[i], fg_masks[i], prep[i][3])
        return results


def ops_masks_from_label_map(label_map, n):
    ids = torch.arange(1, int(n) + 1, device=label_map.device, dtype=label_map.dtype)
    return label_map[None] == ids[:, None, None]
