"""`MaskRefiner` meta-architecture — drop-in for maskrefiner/modeling/mask_refiner/model.py:74-358 (inference).

Same constructor argument (a cfg with the reference's keys), same `forward(batched_inputs) -> list[dict]`
contract and output keys ("sem_seg", "eee_boundary" / "eee_mask" when enabled, "panoptic_seg", and
"instances" only when an instance survives — callers test `"instances" not in output`,
eval/refiner_model.py:267).  What differs is what runs behind it:

  reference                                               here
  ------------------------------------------------------  ---------------------------------------------------
  per image: .to(device), (x-mean)/std, ImageList, cat    ops.pack_input: one kernel for the whole batch (A2)
  backbone / decoder / heads (cuDNN)                      network.py (cuDNN) + ops.fuse_concat / gemm_bias_act (A3-A5)
  5 x F.interpolate(x4), sem_seg_postprocess, sigmoid,    ops.post_process: ONE launch sequence for the batch, no host
  find_instance_center, group_pixels, merge, per-label    round trip (A6-A12); up-sampling is recomputed where it is
  Python loop with .item() syncs, D2H of the map          consumed
"""
import torch
from torch import nn

from . import network, ops
from .config import UOAIS_SIM_META


class Boxes:
    """detectron2.structures.Boxes, as far as consumers of `pred_boxes` use it."""

    def __init__(self, tensor):
        self.tensor = tensor

    def __len__(self):
        return self.tensor.shape[0]

    def to(self, *a, **k):
        return Boxes(self.tensor.to(*a, **k))


class Instances:
    """detectron2.structures.Instances, as far as the path's callers use it (attribute bag keyed by image size;
    consumers read `.pred_masks`, eval/refiner_model.py:270-271)."""

    def __init__(self, image_size, **fields):
        self._image_size = image_size
        self._fields = dict(fields)

    @property
    def image_size(self):
        return self._image_size

    def __getattr__(self, name):
        f = self.__dict__.get("_fields", {})
        if name in f:
            return f[name]
        raise AttributeError("Cannot find field '{}' in the given Instances!".format(name))

    def __setattr__(self, name, val):
        if name.startswith("_"):
            super().__setattr__(name, val)
        else:
            self._fields[name] = val

    def has(self, name):
        return name in self._fields

    def get_fields(self):
        return self._fields

    def __len__(self):
        for v in self._fields.values():
            return len(v)
        raise NotImplementedError("Empty Instances does not support __len__!")

    def to(self, *a, **k):
        return Instances(self._image_size, **{n: (v.to(*a, **k) if hasattr(v, "to") else v) for n, v in self._fields.items()})


class MaskRefiner(nn.Module):
    def __init__(self, cfg, precision="fp32"):
        """precision "fp32": every convolution / GEMM in fp32 (parity mode, config C2);
        "bf16": channels-last bf16 autocast for the cuDNN part and tcgen05 bf16 GEMMs for the 1x1 of FusionLayers (C3)."""
        super().__init__()
        self.backbone = network.BACKBONE_REGISTRY[cfg.MODEL.BACKBONE.NAME](cfg, None)
        shapes = {k: (self.backbone._out_feature_channels[k], self.backbone._out_feature_strides[k])
                  for k in self.backbone._out_features}
        self.offset_input_on = cfg.INPUT.OFFSET_INPUT_ON
        self.depth_on = cfg.INPUT.DEPTH_ON
        self.ins_embed_head = network.INS_EMBED_BRANCHES_REGISTRY[cfg.MODEL.INS_EMBED_HEAD.NAME](cfg, shapes)
        self.register_buffer("pixel_mean", torch.tensor(cfg.MODEL.PIXEL_MEAN, dtype=torch.float32).view(-1, 1, 1), False)
        self.register_buffer("pixel_std", torch.tensor(cfg.MODEL.PIXEL_STD, dtype=torch.float32).view(-1, 1, 1), False)
        self.meta = UOAIS_SIM_META
        p = cfg.MODEL.PANOPTIC_DEEPLAB
        self.stuff_area, self.threshold, self.nms_kernel, self.top_k = p.STUFF_AREA, p.CENTER_THRESHOLD, p.NMS_KERNEL, p.TOP_K_INSTANCE
        self.predict_instances = p.PREDICT_INSTANCES
        self.size_divisibility = p.SIZE_DIVISIBILITY
        self.benchmark_network_speed = p.BENCHMARK_NETWORK_SPEED
        self.eee_mask_on = cfg.MODEL.INS_EMBED_HEAD.EEE_MASK_ON
        self.eee_boundary_on = cfg.MODEL.INS_EMBED_HEAD.EEE_BOUNDARY_ON
        self.error_type = cfg.MODEL.INS_EMBED_HEAD.ERROR_TYPE
        self.common_stride = cfg.MODEL.INS_EMBED_HEAD.COMMON_STRIDE
        self.max_instances_per_image = max(int(self.top_k), 1)   # instances <= centres <= top_k: the pool can not overflow
        self.use_cuda_graph = False         # capture / replay network_logits (fixed shapes; see network_logits)
        self.use_inference_plan = True      # folded BN + fused conv epilogues (network.Conv2d.plan); False = layer by layer
        self.set_precision(precision)

    def set_precision(self, precision):
        if precision not in ("fp32", "bf16"):
            raise ValueError("precision must be 'fp32' or 'bf16'")
        self.precision = precision
        self.ins_embed_head.use_kernels = True
        self.ins_embed_head.gemm_1x1 = precision == "bf16"
        if getattr(self.ins_embed_head, "pixel_decoder", None) is not None:
            self.ins_embed_head.pixel_decoder.set_precision(precision)

    @property
    def device(self):
        return self.pixel_mean.device

    # ---- A2: batch assembly --------------------------------------------------------------------------
    def _pack(self, batched_inputs):
        imgs = [x["image"] for x in batched_inputs]
        sizes = [tuple(i.shape[-2:]) for i in imgs]
        div = self.size_divisibility if self.size_divisibility > 0 else getattr(self.backbone, "size_divisibility", 0)
        Hm, Wm = max(s[0] for s in sizes), max(s[1] for s in sizes)
        if div > 1:
            Hm, Wm = (Hm + div - 1) // div * div, (Wm + div - 1) // div * div
        same = all(s == (Hm, Wm) for s in sizes)
        dev = self.device
        offs = [x["initial_pred_offset"] for x in batched_inputs] if self.offset_input_on else None
        if dev.type == "cuda" and same and all(i.dtype == torch.uint8 for i in imgs) and (Hm * Wm) % 16 == 0:
            image = torch.stack([i.to(dev, non_blocking=True) for i in imgs]).contiguous()
            off = torch.stack([o.to(dev, non_blocking=True) for o in offs]).float().contiguous() if offs is not None else None
            return ops.pack_input(image, off, self.pixel_mean, self.pixel_std), sizes
        # ImageList.from_tensors semantics (pad right / bottom with 0 AFTER normalisation), model.py:137-153
        out = []
        for n, i in enumerate(imgs):
            t = (i.to(dev) - self.pixel_mean) / self.pixel_std
            if offs is not None:
                t = torch.cat([t, offs[n].to(dev).float()], dim=0)
            out.append(torch.nn.functional.pad(t, (0, Wm - t.shape[-1], 0, Hm - t.shape[-2])))
        return torch.stack(out), sizes

    # ---- A3-A5: network ------------------------------------------------------------------------------
    def _ensure_plan(self, device):
        """(Re)build the convolutions' inference plan (network.prepare_inference) when the precision, the device or any
        parameter / buffer changed since it was built (a checkpoint load, an in-place edit)."""
        ver = sum(t._version for t in self.state_dict(keep_vars=True).values())
        key = (self.precision, str(device), ver, self.training)
        if getattr(self, "_plan_key", None) != key:
            if self.training or not self.use_inference_plan:
                network.drop_inference_plan(self)
            else:
                bf16 = self.precision == "bf16" and device.type == "cuda"
                # channels-last on CUDA in both precisions: cuDNN's tensor-core kernels (bf16, and TF32 under PyTorch's
                # default for fp32 convolutions) are NHWC and otherwise transpose around every call (25 of 68 ms at C2)
                network.prepare_inference(self, torch.bfloat16 if bf16 else torch.float32, channels_last=device.type == "cuda")
            self._plan_key = key

    def network_logits(self, x):
        """[B,9,H,W] normalised input -> dict of 1/COMMON_STRIDE-resolution float32 logits (the head outputs
        before model.py:688-708's up-sampling).  With `use_cuda_graph` the ~400 launches of the forward are captured
        once per (shape, precision, parameter version) and replayed (launch gaps: 2 of 23 ms at C2 in bf16)."""
        self._ensure_plan(x.device)
        if self.use_cuda_graph and x.is_cuda and not self.training:
            return self._graphed_logits(x)
        return self._network_logits(x)

    def _graphed_logits(self, x):
        key = (tuple(x.shape), x.dtype, self._plan_key)
        g = self.__dict__.get("_graph")
        if g is None or g["key"] != key:
            static_x = x.clone()
            side = torch.cuda.Stream(device=x.device)
            side.wait_stream(torch.cuda.current_stream(x.device))
            with torch.cuda.stream(side):                          # warm-up off the capture: cuDNN plans, autocast caches
                for _ in range(2):
                    self._network_logits(static_x)
            torch.cuda.current_stream(x.device).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                static_out = self._network_logits(static_x)
            g = dict(key=key, graph=graph, x=static_x, out=static_out)
            self.__dict__["_graph"] = g
        g["x"].copy_(x)
        g["graph"].replay()
        return {k: v.clone() for k, v in g["out"].items()}

    def _network_logits(self, x):
        if self.precision == "bf16" and x.is_cuda:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                # the RGB-D backbone splits / casts / re-lays-out the packed fp32 input itself in one pass (ops.split_rgbd_input)
                own_split = getattr(self.backbone, "splits_packed_input", False) and x.dtype == torch.float32 and x.is_contiguous()
                feats = self.backbone(x if own_split else x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last))
                out = self.ins_embed_head.layers({k: v for k, v in feats.items()})
        else:
            if x.is_cuda and self.use_inference_plan and not self.training:
                x = x.contiguous(memory_format=torch.channels_last)
            feats = self.backbone(x)
            out = self.ins_embed_head.layers(feats)
        return {k: v.float().contiguous() for k, v in out.items()}

    # ---- forward -------------------------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, batched_inputs):
        x, sizes = self._pack(batched_inputs)
        out = self.network_logits(x)
        if self.benchmark_network_speed:
            return []
        if not x.is_cuda:
            raise RuntimeError("quber_b200 has no CPU post-processing: MaskRefiner.forward needs a CUDA device "
                               "(network_logits() runs anywhere)")
        B = x.shape[0]
        s = self.common_stride
        heights = [bi.get("height", sz[0]) for bi, sz in zip(batched_inputs, sizes)]
        widths = [bi.get("width", sz[1]) for bi, sz in zip(batched_inputs, sizes)]
        eee = out.get("eee_boundary") if self.eee_boundary_on else None
        want = ("sem_seg", "panoptic", "masks") + (("eee_full",) if eee is not None else ())
        kw = dict(threshold=self.threshold, nms_kernel=self.nms_kernel, top_k=self.top_k, stuff_area=self.stuff_area,
                  label_divisor=self.meta["label_divisor"], want=want)
        results = [None] * B
        # sem_seg_postprocess (model.py:266-289) = crop to the image size, bilinear resize to (height, width).  Frames that
        # share (image size, output size) go through the kernels together.
        groups = {}
        for b in range(B):
            groups.setdefault((tuple(sizes[b]), (int(heights[b]), int(widths[b]))), []).append(b)
        full_hw = (out["foreground"].shape[2] * s, out["foreground"].shape[3] * s)
        for (size, out_hw), idx in groups.items():
            sel = (lambda t: t) if len(idx) == B else (lambda t: t[idx].contiguous())
            fg, ce, of = sel(out["foreground"]), sel(out["center"]), sel(out["offset"])
            ee = None if eee is None else sel(eee)
            n = len(idx)
            if out_hw == size:
                # the resize is the identity: up-sample inside the fused kernels; a frame smaller than
                # the up-sampled maps (armbench 750 x 1333 of 752 x 1336) is their top-left corner
                crop = None if (size == full_hw and size[1] % 8 == 0) else size
                r = ops.post_process(fg, ce, of, ee, scale=s, max_instances=n * self.max_instances_per_image, crop=crop, **kw)
            else:
                # a real resize: full-resolution maps in one pass per head (up-sample, crop, resize), then the same kernels
                rs = lambda t, mul=1.0: ops.upsample_crop_resize(t, s, size, out_hw, mul)
                r = ops.post_process(rs(fg), rs(ce), rs(of, float(s)), None if ee is None else rs(ee), scale=1, offset_mul=1.0,
                                     max_instances=n * self.max_instances_per_image, crop=out_hw, **kw)
            self._check_pool(r)
            mfull = ops.upsample_crop_resize(sel(out["eee_mask"]), s, size, out_hw) if self.eee_mask_on else None   # model.py:282-284
            for j, b in enumerate(idx):
                results[b] = self._result(r, j, b, out_hw[0], out_hw[1], None if mfull is None else mfull[j],
                                          crop=out_hw[1] if r["valid_w"] != r["panoptic"].shape[-1] else None)
        return results

    @staticmethod
    def _check_pool(r):
        """Never clamp silently: quber_post_process sets status bit 0 when the instance pool was too small."""
        if int(r["status"][0]) & 1:
            raise RuntimeError(f"quber_post_process: more refined instances than the output pool holds "
                               f"({r['max_instances']} rows); raise MaskRefiner.max_instances_per_image")

    def _result(self, r, rb, b, height, width, eee_mask, crop=None):
        cw = (lambda t: t[..., :crop]) if crop else (lambda t: t)
        res = {"sem_seg": cw(r["sem_seg"][rb])}
        if self.eee_boundary_on and "eee_full" in r:
            res["eee_boundary"] = cw(r["eee_full"][rb])
        if eee_mask is not None:
            res["eee_mask"] = eee_mask
        res["panoptic_seg"] = (cw(r["panoptic"][rb]), None)
        if self.predict_instances:
            o0, o1 = int(r["inst_offsets"][rb]), int(r["inst_offsets"][rb + 1])     # the single host read-back per image
            if o1 > o0:
                res["instances"] = Instances(
                    (height, width), pred_masks=cw(r["masks"][o0:o1]).bool(), scores=r["scores"][o0:o1],
                    pred_boxes=Boxes(r["boxes"][o0:o1]),
                    pred_classes=(r["labels"][o0:o1] // self.meta["label_divisor"] - 1).long())
        return res


META_ARCH_REGISTRY = {"MaskRefiner": MaskRefiner}


def build_model(cfg, precision="fp32"):
    """detectron2.modeling.build_model for the registered meta-architecture (predictor.py:217).  Also registers the
    path's entry points with detectron2 when it is importable (quber_b200/registry.py)."""
    from .registry import register
    register()
    model = META_ARCH_REGISTRY[cfg.MODEL.META_ARCHITECTURE](cfg, precision)
    return model.to(torch.device(cfg.MODEL.DEVICE if torch.cuda.is_available() or cfg.MODEL.DEVICE == "cpu" else "cpu"))
