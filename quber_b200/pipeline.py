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
import os
from dataclasses import dataclass, field

import numpy as np
import torch

from . import ops

# Host-buffer path: pack the masks to one bit per pixel before the host->device copy?  None = time both routes on the
# first call (HotPath._calibrate_pack).  On the 16-core B200 boxes of this pool 12 threads pack 8 cold frames in 0.63 ms
# against 1.34 ms for their byte copy (scripts/bench_hostpack.py, profiles/r01_hostpack_v12.txt).
PACK_MIN_THREADS = 4


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
    max_instances_per_image: int = 0   # rows of the output pool per frame; 0 = top_k, the worst case (instances <= centres
                                       # <= top_k), so the pool can not overflow; a smaller value sets status bit 0 on overflow
    compact_label_map: bool = False    # label maps as uint8 (QUBER_POST_LABEL_U8; needs top_k <= 255): 1/4 of the bytes
    # what quber_post_process writes: everything MaskRefiner.forward returns ("sem_seg", "eee_boundary" = eee_full,
    # "panoptic_seg", "instances".pred_masks, model.py:303-356), the caller-side error decode (eval_utils.py:310-313) and the
    # label map (the collective / host payload)
    want: tuple = ("sem_seg", "eee_full", "panoptic", "masks", "eee_onehot", "label_map")
    crop: tuple = None                 # (H, W) of the frames when they are a crop of the up-sampled maps (armbench 750 x 1333)
    valid_w: int = 0                   # the frames' own width when the mask planes carry a zero-padded row pitch


class HotPath:
    def __init__(self, cfg=None, device="cuda"):
        self.cfg = cfg or RefinerConfig()
        self.device = torch.device(device)
        self.launches_per_step = 0
        # host-buffer path: send the 0/1 masks as one bit per pixel (csrc/hostpack.cpp).  QUBER_PACK_MASKS=0|1 overrides.
        # None ("auto") = decide on the first call by timing both routes on this host (_calibrate_pack).
        self.use_graphs = os.environ.get("QUBER_E2E_GRAPHS", "1") == "1"   # host-buffer path: one CUDA graph per chunk
        # this rank's share of the box's cores (torchrun exports LOCAL_WORLD_SIZE), two left for the launching threads
        cores = len(os.sched_getaffinity(0)) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
        self.pack_threads = max(1, min(cores - 2, 16))      # the caller packs too while it waits
        # Deterministic route (no timing race inside the call): packing 8 frames takes 12 threads 0.63 ms against 1.34 ms
        # for their DMA, 2 threads 5.0 ms against 3.7 ms (profiles/r01_hostpack_v12.txt, r01_bench_n8_v14.json) — the
        # break-even is at about 4 threads, so that is the rule.  QUBER_PACK_MASKS=0|1 overrides; "auto" times both routes
        # on the first call (_calibrate_pack, kept as a tool).
        env = os.environ.get("QUBER_PACK_MASKS")
        self.pack_rule = f"pack iff >= {PACK_MIN_THREADS} host threads per rank ({self.pack_threads} here)"
        self.pack_masks = (self.pack_threads >= PACK_MIN_THREADS) if env is None else (None if env == "auto" else env == "1")

    # --- resident: every input already on the device -------------------------------------------
    def encode_and_boundary(self, masks, status=None):
        c = self.cfg
        B, K, H, W = masks.shape
        # the encode pass reads every mask byte: its tile occupancy lets the boundary pass skip the empty (tile, plane) pairs
        enc, occ = ops.encode_initial_masks(masks, c.sigma, c.valid_w, with_occupancy=True)
        d = ops.boundary_dilation(H, c.valid_w or W, c.boundary_ratio)
        per, uni, fg = ops.mask_boundary(masks, d, status=status, occupancy=occ)
        return enc, per, uni, fg

    def encode_and_boundary_packed(self, st, c0, c1, w0, w1, value, status=None):
        """encode_and_boundary for a chunk whose masks arrived bit-packed (refine_batch_host / refine_batch_rle): both kernels
        work from the bit planes, no byte planes are written on the device (W % 32 == 0; other widths expand the bits and
        take the byte kernels)."""
        c = self.cfg
        masks = st["masks"][c0:c1]
        B, K, H, W = masks.shape
        if W % 32:
            ops.unpack_masks(st["packed_dev"][w0:w1], masks.shape, value=value, out=masks)
            return self.encode_and_boundary(masks, status)
        words = st["packed_dev"][w0:w1]
        enc, occ = ops.encode_initial_masks_packed(words, value, None, c.sigma, c.valid_w, with_occupancy=True, shape=(B, K, H, W))
        per, uni, fg = ops.mask_boundary_bits(words, value, (B, K, H, W), ops.boundary_dilation(H, c.valid_w or W, c.boundary_ratio),
                                              occupancy=occ)
        return enc, per, uni, fg

    def post(self, head, label_map_out=None):
        c = self.cfg
        B = head["foreground"].shape[0]
        return ops.post_process(head["foreground"], head["center"], head["offset"], head.get("eee_boundary"),
                                scale=c.common_stride, threshold=c.center_threshold, nms_kernel=c.nms_kernel,
                                top_k=c.top_k, min_area=c.min_area, stuff_area=c.stuff_area,
                                label_divisor=c.label_divisor, max_instances=B * self.slab, want=c.want,
                                label_dtype=torch.uint8 if c.compact_label_map and c.top_k <= 255 else torch.int32,
                                crop=c.crop, label_map_out=label_map_out)

    @property
    def slab(self):
        """Rows of the instance pool per frame."""
        return self.cfg.max_instances_per_image or max(self.cfg.top_k, 1)

    def msda_layers(self, msda, n_layers=6):
        """The sampling of `n_layers` encoder layers.  With the module's operands (`oa` = the sampling_offsets |
        attention_weights projection rows, `ref` = reference points) the fused kernel does ms_deform_attn.py:102-117 and
        the sampling in one launch; with precomputed loc / attn the op alone (ms_deform_attn_forward)."""
        out = None
        for _ in range(n_layers):
            if "oa" in msda:
                L = msda["ref"].shape[2]
                out = ops.msda_forward_fused(msda["value"], msda["shapes"], msda["starts"], msda["oa"], msda["ref"], L, 4,
                                             levels=msda.get("levels"))
            else:
                out = ops.msda_forward(msda["value"], msda["shapes"], msda["starts"], msda["loc"], msda["attn"])
        return out

    # kernels launched per call (for bench.py's gpu_launches; memsets are not kernels)
    LAUNCHES = dict(encode=2, boundary=1, post=8, msda_layer=1)

    # --- end to end: host buffers in, host results out ------------------------------------------
    def _staging(self, masks_host, head_host, slot=0):
        return self._staging_shape(tuple(masks_host.shape), head_host, slot)

    def _staging_shape(self, shape, head_host, slot=0):
        """Persistent device staging buffers, pinned result buffers, copy streams and events, keyed by shape; one set per
        `slot` (a caller that keeps two calls in flight alternates between two slots)."""
        ldt = torch.uint8 if self.cfg.compact_label_map and self.cfg.top_k <= 255 else torch.int32
        key = (tuple(shape), ldt) + tuple((k, tuple(v.shape)) for k, v in sorted(head_host.items()))
        stages = self.__dict__.setdefault("_stages", {})
        st = stages.get(slot)
        if st is None or st["key"] != key:
            dev = self.device
            B, K, H, W = shape
            numel = B * K * H * W
            slab = self.slab
            Ho, Wo = (self.cfg.crop[0], (self.cfg.crop[1] + 7) // 8 * 8) if self.cfg.crop else (H, W)   # post outputs' rows / pitch
            pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)
            st = dict(key=key, masks=torch.empty(shape, dtype=torch.uint8, device=dev),
                      head={k: torch.empty(v.shape, dtype=v.dtype, device=dev) for k, v in head_host.items()},
                      h2d=torch.cuda.Stream(device=dev), h2d_bytes=torch.cuda.Stream(device=dev), d2h=torch.cuda.Stream(device=dev), free=None,
                      packed_host=pin((numel + 31) // 32, torch.int32),
                      packed_dev=torch.empty((numel + 31) // 32, dtype=torch.int32, device=dev), copied={},
                      graphs={}, bstatus=torch.zeros(B + 1, dtype=torch.int32, device=dev),
                      out=dict(label_map=pin((B, Ho, Wo), ldt), boundary=pin((B, H, W), torch.uint8),
                               inst_start=pin((B,), torch.int32), inst_counts=pin((B,), torch.int32),
                               status=pin((1,), torch.int32), rle_status=pin((1,), torch.int32),
                               boundary_status=pin((B + 1,), torch.int32), labels=pin((B * slab,), torch.int32), scores=pin((B * slab,), torch.float32),
                               boxes=pin((B * slab, 4), torch.float32)))
            for k in ("status", "rle_status", "boundary_status"):
                st["out"][k].zero_()
            st["ready"] = None
            stages[slot] = st
        return st

    @property
    def _stage(self):
        return self._stages[0]

    def wait_result(self, slot=0):
        """Block until the host results of the last call on `slot` are complete (all its device->host copies done) and
        return them — the counterpart of `async_result=True`."""
        st = self._stages[slot]
        if st["ready"] is not None:
            st["ready"].synchronize()
        return st["out"]

    def _run_graphed(self, st, key, body):
        """`body()` (kernel launches on the current stream, inputs = the persistent staging buffers) eagerly, or with
        `use_graphs` captured once per `key` into a CUDA graph and replayed (outputs live in the graph's pool and are
        overwritten by the next replay) — the host then spends one graph launch instead of 0.1-0.3 ms of enqueueing,
        which is what bounds the call once the masks travel bit-packed."""
        if not self.use_graphs:
            return body()
        ent = st["graphs"].get(key)
        if ent is None:
            body()                                # eager once: lazy tables, function attributes, allocator warm-up
            try:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    outs = body()
            except RuntimeError as e:             # capture refused (another thread allocating, driver quirk): stay eager
                import warnings
                warnings.warn(f"quber_b200: CUDA graph capture failed, launching eagerly ({e})")
                self.use_graphs = False
                torch.cuda.synchronize(self.device)
                return body()
            ent = st["graphs"][key] = (g, outs)
        ent[0].replay()
        return ent[1]

    def _calibrate_pack(self, masks_host, st, sched):
        """True when packing a chunk on the host cores is clearly cheaper than copying its bytes: both are timed once on
        the second chunk (cold in the host caches for the cores, a repeated DMA for the copy engine), after a warm-up.
        Measured on a 24-core box with two ranks of 10 threads: packing at 0.66 of the copy time still wins end to end
        (6.07 against 6.77 ms per step), hence the 0.85."""
        import time
        fb = masks_host[0].numel()
        if fb % 32 or not masks_host.is_contiguous() or self.pack_threads < 2:
            return False
        (a0, a1), (b0, b1) = sched[0], sched[min(1, len(sched) - 1)]
        ops.host_pack_masks(masks_host[a0:a1], st["packed_host"][a0 * fb // 32:a1 * fb // 32], self.pack_threads)
        t = time.perf_counter()
        _, value = ops.host_pack_masks(masks_host[b0:b1], st["packed_host"][b0 * fb // 32:b1 * fb // 32], self.pack_threads)
        t_pack = time.perf_counter() - t
        h2d = st["h2d"]
        torch.cuda.synchronize(self.device)                 # nothing else on the link or the copy engines while timing
        with torch.cuda.stream(h2d):
            st["masks"][b0:b1].copy_(masks_host[b0:b1], non_blocking=True)    # the first DMA of fresh pinned pages is slow
            e0 = h2d.record_event(torch.cuda.Event(enable_timing=True))
            st["masks"][b0:b1].copy_(masks_host[b0:b1], non_blocking=True)
            e1 = h2d.record_event(torch.cuda.Event(enable_timing=True))
        e1.synchronize()
        t_copy = e0.elapsed_time(e1) * 1e-3
        self.pack_calibration = dict(pack_s=t_pack, copy_s=t_copy, frames=b1 - b0, threads=self.pack_threads)
        return value >= 0 and t_pack < 0.85 * t_copy

    @staticmethod
    def chunk_schedule(B, chunk_frames, min_tail=2):
        """[(start, stop), ...]: chunks of `chunk_frames` frames, the last one halved repeatedly down to `min_tail`
        frames — what runs after the final host->device copy (kernels + device->host copy of the last chunk) is the
        only part of the call the copies cannot hide, so the last chunk is kept small."""
        sizes = [chunk_frames] * (B // chunk_frames)
        if B % chunk_frames:
            sizes.append(B % chunk_frames)
        if sizes:
            last = sizes.pop()
            while last > min_tail:
                half = (last + 1) // 2
                sizes.append(half)
                last -= half
            sizes.append(last)
        out, c0 = [], 0
        for n in sizes:
            if n > 0:
                out.append((c0, c0 + n))
                c0 += n
        return out

    def byte_frames_rule(self, B, frame_bytes=0, head_bytes=0):
        """How many of the B frames travel as bytes through the copy engine WHILE the host threads bit-pack the others:
        the dense masks have to be read from host DRAM once either way, so the two readers (DMA at ~55 GB/s, packing at
        ~5 GB/s per thread up to ~60 GB/s) are run side by side.  The split (a multiple of 4 frames) minimises the longer
        of the two: the link carries the byte frames AND an eighth of the packed ones AND the head outputs, the threads
        read the packed frames.  Deterministic (a function of the thread count and the sizes), nominal rates from
        profiles/r01_hostpack_v12.txt and the bench's host_link figures; measured at C2 with 14 threads: 12 byte frames
        3.73 ms per call, 16: 3.90, 8: 4.44."""
        if not self.pack_masks:
            return B
        copy_gbs, pack_gbs = 55.0, min(5.0 * self.pack_threads, 60.0)
        if frame_bytes <= 0:                      # sizes unknown: proportional to the rates
            nb = int(round(B * copy_gbs / (copy_gbs + pack_gbs) / 4.0)) * 4 if B >= 8 else int(round(B * copy_gbs / (copy_gbs + pack_gbs)))
            return max(0, min(B, nb))
        step = 4 if B >= 8 else 1
        best, best_t = B, None
        for nb in range(0, B + 1, step):
            t_copy = (nb * frame_bytes + (B - nb) * frame_bytes / 8.0 + head_bytes) / copy_gbs
            t_pack = (B - nb) * frame_bytes / pack_gbs
            t = max(t_copy, t_pack)
            if best_t is None or t < best_t - 1e-9:
                best, best_t = nb, t
        return best

    def refine_batch_host(self, masks_host, head_host, out_host=None, chunk_frames=8, min_tail=2, kernel_frames=0, byte_frames=None,
                          slot=0, async_result=False):
        """The end-to-end call: host buffers in, host results out.

        masks_host: pinned uint8 [B,K,H,W]; head_host: dict of pinned float32 head outputs (what the
        network hands to post-processing).  The batch is cut into chunks of `chunk_frames` frames (frames are
        independent — the reference itself is batch 1, eval/run_eval.py:72) and pipelined over three streams:
        host->device copies of chunk c+1 (copy stream) run under the kernels of chunk c (the caller's current
        stream) and the device->host copies of chunk c-1 (second copy stream), so the call costs about the PCIe
        time of the masks instead of copies + kernels.  Returns (host dict, device results per chunk):
        label_map [B,H,W] int32 (0 = void, j+1 = j-th refined instance of the frame — the lossless compact
        form of the reference's pred_masks), boundary [B,H,W] uint8 (union boundary of the initial
        masks), inst_start / inst_counts [B] int32 (frame b's instances are rows inst_start[b] ... +inst_counts[b]
        of labels / scores / boxes).  Post-processing runs once for the whole batch as soon as the head outputs have
        arrived; the chunks carry the masks (encode, boundaries).  `kernel_frames` > 0 decouples the two granularities:
        the masks still cross the link chunk by chunk (fine chunks keep the copies right behind the packing), but encode /
        boundary kernels are launched once per group of consecutive chunks holding at least `kernel_frames` frames — full
        grids and one graph launch instead of one small launch sequence per chunk.  `keep` = per kernel group (encoding, per-instance boundaries,
        the batch's post-processing result, inst_start slice, inst_counts slice).
        `byte_frames` (default: byte_frames_rule): the first that many frames go over the link as bytes on a second copy
        stream while the host threads pack the rest — two readers of host DRAM instead of one.
        The host tensors are pinned and reused between calls; they are valid once the caller has synchronised
        the current stream (which this call makes wait for both copy streams).
        Two calls in flight: with `async_result=True` the current stream is NOT made to wait for the device->host copies
        (the next call's kernels start right behind this call's); the results are complete after `wait_result(slot)`.
        A caller that alternates `slot` between 0 and 1 has the copies and the host-side packing of call i+1 running under
        the kernels of call i; each slot has its own staging buffers, pinned results and graphs, and call i+2 on a slot
        must not be issued before the results of call i on it have been read.
        """
        dev = self.device
        st = self._staging(masks_host, head_host, slot)
        out = st["out"] if out_host is None else out_host
        cur = torch.cuda.current_stream(dev)
        h2d, d2h = st["h2d"], st["d2h"]
        B = masks_host.shape[0]
        slab = self.slab
        if st["free"] is not None:
            h2d.wait_event(st["free"])            # the previous call's kernels have finished reading the staging buffers
        d2h.wait_stream(cur)                      # ... and whoever consumed the pinned results was ordered on `cur`
        keep = []
        frame_bytes = masks_host[0].numel()
        if self.pack_masks is None:
            if st["free"] is not None:
                st["free"].synchronize()
            self.pack_masks = self._calibrate_pack(masks_host, st, self.chunk_schedule(B, chunk_frames, min_tail))
        pack = bool(self.pack_masks) and frame_bytes % 32 == 0 and masks_host.is_contiguous()
        head_bytes = sum(v.numel() * v.element_size() for v in head_host.values())
        nb = B if not pack else (self.byte_frames_rule(B, frame_bytes, head_bytes) if byte_frames is None else max(0, min(B, int(byte_frames))))
        pack = pack and nb < B
        byte_sched = self.chunk_schedule(nb, chunk_frames, chunk_frames) if nb else []
        sched = [(nb + a, nb + b) for a, b in self.chunk_schedule(B - nb, chunk_frames, min_tail)] if nb < B else []
        self.last_route = dict(byte_frames=nb, packed_frames=B - nb if pack else 0)
        if pack:
            # One bit per pixel over the host link (csrc/hostpack.cpp): the pool threads pack the chunks in order while
            # this thread enqueues copies and kernels of the chunks already packed.  The pinned words may still be read
            # by the previous call's copies: wait for those first.
            for ev in st["copied"].values():
                ev.synchronize()
            ops.host_pack_begin(masks_host[nb:], st["packed_host"][nb * frame_bytes // 32:], self.pack_threads,
                                [(c1 - nb) * frame_bytes for _, c1 in sched])
        try:
            # Post-processing needs the head outputs only (20 MB): they go first and the whole batch is post-processed in
            # one launch sequence (full grids, one set of the small serial kernels) while the masks are packed / copied.
            with torch.cuda.stream(h2d):
                for k, v in head_host.items():
                    st["head"][k].copy_(v, non_blocking=True)
                heads_ready = h2d.record_event()
            st["copied"]["heads"] = heads_ready
            bytes_ready = None
            if byte_sched:
                # the byte share of the masks: queued on its own copy stream at once, the copy engine reads host DRAM
                # while the pool threads read (and pack) the other frames
                h2b = st["h2d_bytes"]
                if st["free"] is not None:
                    h2b.wait_event(st["free"])
                bytes_ready = []
                with torch.cuda.stream(h2b):
                    for (c0, c1) in byte_sched:
                        st["masks"][c0:c1].copy_(masks_host[c0:c1], non_blocking=True)
                        bytes_ready.append((c1, h2b.record_event()))
                st["copied"]["bytes"] = bytes_ready[-1][1]
            cur.wait_event(heads_ready)
            r = self._run_graphed(st, "post", lambda: self.post(st["head"]))
            off = r["inst_offsets"]
            start, counts = off[:-1], off[1:] - off[:-1]
            post_done = cur.record_event()
            with torch.cuda.stream(d2h):
                d2h.wait_event(post_done)
                out["label_map"].copy_(r["label_map"], non_blocking=True)
                out["inst_start"].copy_(start, non_blocking=True)
                out["inst_counts"].copy_(counts, non_blocking=True)
                out["status"].copy_(r["status"], non_blocking=True)     # bit 0: instance pool overflow (never with slab = top_k)
                out["labels"].copy_(r["labels"][:B * slab], non_blocking=True)
                out["scores"].copy_(r["scores"][:B * slab], non_blocking=True)
                out["boxes"].copy_(r["boxes"][:B * slab], non_blocking=True)

            def byte_groups():
                """encode / boundary kernels of the byte share, in groups of `kernel_frames` frames (one group if 0)."""
                step = kernel_frames if kernel_frames else chunk_frames
                for g0 in range(0, nb, step):
                    g1 = min(nb, g0 + step)
                    cur.wait_event(next(ev for c1, ev in bytes_ready if c1 >= g1))   # copies are in stream order
                    bs = st["bstatus"][g0:g0 + 1]
                    enc, per, uni, fg = self._run_graphed(st, (g0, g1), lambda: self.encode_and_boundary(st["masks"][g0:g1], bs))
                    done = cur.record_event()
                    with torch.cuda.stream(d2h):
                        d2h.wait_event(done)
                        out["boundary"][g0:g1].copy_(uni, non_blocking=True)
                    keep.append((enc, per, r, start[g0:g1], counts[g0:g1], g0))

            if byte_sched and not sched:
                byte_groups()
            pend = []                                 # chunks copied but not yet processed: (c0, c1, packed, value)
            for ci, (c0, c1) in enumerate(sched):
                packed, value = False, -1
                if pack:
                    w0, w1 = c0 * frame_bytes // 32, c1 * frame_bytes // 32
                    value = ops.host_pack_wait(ci)
                    packed = value >= 0                                         # one value on every set pixel (1 or 255)
                with torch.cuda.stream(h2d):
                    if packed:
                        st["packed_dev"][w0:w1].copy_(st["packed_host"][w0:w1], non_blocking=True)
                    else:                                                       # several non-zero values: the bytes as they are
                        st["masks"][c0:c1].copy_(masks_host[c0:c1], non_blocking=True)
                    ready = h2d.record_event()
                st["copied"][(c0, c1)] = ready
                pend.append((c0, c1, packed, max(value, 1)))
                if ci + 1 < len(sched) and kernel_frames and pend[-1][1] - pend[0][0] < kernel_frames:
                    continue                      # copy-only chunk: its kernels run with the following ones
                cur.wait_event(ready)             # the copies are in stream order: this one covers the whole group
                if len({(pk, v) for _, _, pk, v in pend}) > 1:
                    groups = [[q] for q in pend]  # mixed routes / values inside the group: chunk by chunk
                else:
                    groups = [pend]
                for grp in groups:
                    g0, g1, pk, v = grp[0][0], grp[-1][1], grp[0][2], grp[0][3]
                    bs = st["bstatus"][g0:g0 + 1]   # this group's status word (quber_mask_boundary clears and sets it)
                    if pk:                        # the encode pass expands the bits into the staging bytes on its way
                        enc, per, uni, fg = self._run_graphed(
                            st, (g0, g1, v), lambda: self.encode_and_boundary_packed(
                                st, g0, g1, g0 * frame_bytes // 32, g1 * frame_bytes // 32, v, bs))
                    else:
                        enc, per, uni, fg = self._run_graphed(st, (g0, g1), lambda: self.encode_and_boundary(st["masks"][g0:g1], bs))
                    done = cur.record_event()
                    with torch.cuda.stream(d2h):
                        d2h.wait_event(done)
                        out["boundary"][g0:g1].copy_(uni, non_blocking=True)
                    keep.append((enc, per, r, start[g0:g1], counts[g0:g1], g0))
                pend = []
            if byte_sched and sched:
                byte_groups()
            keep.sort(key=lambda k: k[5])             # frame order, whatever order the groups were launched in
        except BaseException:
            if pack:
                for c in range(len(sched)):           # the pool must not go on reading the caller's buffers
                    ops.host_pack_wait(c)
            raise
        with torch.cuda.stream(d2h):
            d2h.wait_event(cur.record_event())
            out["boundary_status"].copy_(st["bstatus"], non_blocking=True)     # non-zero: a TMA load was lost, see check_status
            st["ready"] = d2h.record_event()
        st["free"] = cur.record_event()
        if not async_result:
            cur.wait_stream(d2h)
        self.__dict__.setdefault("_keeps", {})[slot] = keep    # device results stay alive until the slot's next call replaces them
        return out, keep

    @staticmethod
    def check_status(out):
        """Call after synchronising: raises if the instance pool overflowed, a boundary TMA load was lost or an RLE did
        not sum to H*W — the host results of that call are then not to be used."""
        if int(out["status"][0]) & 1:
            raise RuntimeError("quber_post_process: instance pool overflow (RefinerConfig.max_instances_per_image too small)")
        if int(out["boundary_status"].max()) != 0:
            raise RuntimeError("quber_mask_boundary: a TMA tile load timed out; boundary outputs invalid")
        if "rle_status" in out and int(out["rle_status"][0]) & 1:
            raise ValueError("quber_rle_decode: RLE counts do not sum to H*W")

    def refine_batch_rle(self, counts_host, offsets_host, shape, head_host, value=255, out_host=None, kernel_frames=16,
                         slot=0, async_result=False):
        """The end-to-end call with the initial masks in the form the reference's datasets hold them: COCO run-length
        encodings (register_uoais_sim_panoptic.py:121-144 -> mask_util.decode at
        perturbed_panoptic_dataset_mapper.py:195-214).  counts_host: pinned int32 [total] (uint32 run lengths of the
        B*K masks back to back, `rle.pack_rles`), offsets_host: pinned int64 [B*K+1], shape = (B, K, H, W); head_host as
        refine_batch_host.  ~0.6 MB cross the host link instead of 295 MB (C2): quber_rle_decode writes the bit-packed
        planes that the encode pass reads anyway.  Same results and result layout as refine_batch_host; `slot` /
        `async_result` as there."""
        B, K, H, W = shape
        n = B * K
        total = int(counts_host.numel())

        def copy_in(st):                          # on the copy stream
            if st.get("rle_cap", 0) < total:
                st["rle_counts"] = torch.empty(max(total * 2, 1 << 16), dtype=torch.int32, device=self.device)
                st["rle_offsets"] = torch.empty(n + 1, dtype=torch.int64, device=self.device)
                st["rle_cap"] = st["rle_counts"].numel()
            st["rle_counts"][:total].copy_(counts_host, non_blocking=True)
            st["rle_offsets"].copy_(offsets_host, non_blocking=True)

        def make_planes(st):                      # on the caller's stream, after the copies: -> status word or None
            return ops.rle_decode(st["rle_counts"][:total], st["rle_offsets"], n, H, W, value=value, packed=True,
                                  out=st["packed_dev"])[1]
        return self._refine_planes("refine_batch_rle", shape, head_host, value, out_host, kernel_frames, slot, async_result,
                                   copy_in, make_planes)

    def refine_batch_bits(self, packed_host, shape, head_host, value=255, out_host=None, kernel_frames=16, slot=0,
                          async_result=False):
        """The end-to-end call for a caller that already holds the initial masks as bit planes: packed_host pinned int32
        [B*K*H*W / 32] (bit i of word w = pixel 32 w + i of the flattened [B,K,H,W] array — what ops.host_pack_masks
        writes), set pixels carry `value`.  An eighth of the dense call's bytes cross the link and no host thread packs.
        Same results and result layout as refine_batch_host."""
        B, K, H, W = shape
        if packed_host.dtype != torch.int32 or packed_host.numel() * 32 != B * K * H * W:
            raise ValueError("packed_host must hold B*K*H*W / 32 int32 words")

        def copy_in(st):
            st["packed_dev"].copy_(packed_host, non_blocking=True)
        return self._refine_planes("refine_batch_bits", shape, head_host, value, out_host, kernel_frames, slot, async_result,
                                   copy_in, lambda st: None)

    def _refine_planes(self, what, shape, head_host, value, out_host, kernel_frames, slot, async_result, copy_in, make_planes):
        """Shared body of the routes whose masks reach the device as bit planes (st["packed_dev"])."""
        dev = self.device
        B, K, H, W = shape
        if W % 32:
            raise ValueError(f"{what} needs W % 32 == 0 (use the dense call)")
        st = self._staging_shape((B, K, H, W), head_host, slot)
        out = st["out"] if out_host is None else out_host
        cur = torch.cuda.current_stream(dev)
        h2d, d2h = st["h2d"], st["d2h"]
        slab = self.slab
        if st["free"] is not None:
            h2d.wait_event(st["free"])
        d2h.wait_stream(cur)
        with torch.cuda.stream(h2d):
            for k, v in head_host.items():
                st["head"][k].copy_(v, non_blocking=True)
            heads_ready = h2d.record_event()
            copy_in(st)
            planes_ready = h2d.record_event()
        st["copied"]["heads"] = heads_ready
        cur.wait_event(heads_ready)
        r = self._run_graphed(st, "post", lambda: self.post(st["head"]))
        off = r["inst_offsets"]
        start, counts = off[:-1], off[1:] - off[:-1]
        post_done = cur.record_event()
        with torch.cuda.stream(d2h):
            d2h.wait_event(post_done)
            out["label_map"].copy_(r["label_map"], non_blocking=True)
            out["inst_start"].copy_(start, non_blocking=True)
            out["inst_counts"].copy_(counts, non_blocking=True)
            out["status"].copy_(r["status"], non_blocking=True)
            out["labels"].copy_(r["labels"][:B * slab], non_blocking=True)
            out["scores"].copy_(r["scores"][:B * slab], non_blocking=True)
            out["boxes"].copy_(r["boxes"][:B * slab], non_blocking=True)
        cur.wait_event(planes_ready)
        plane_status = make_planes(st)
        keep = []
        fw = K * H * W // 32
        step = kernel_frames if kernel_frames and kernel_frames < B else B
        for g0 in range(0, B, step):
            g1 = min(B, g0 + step)
            bs = st["bstatus"][g0:g0 + 1]
            enc, per, uni, fg = self._run_graphed(
                st, (g0, g1, value), lambda: self.encode_and_boundary_packed(st, g0, g1, g0 * fw, g1 * fw, value, bs))
            done = cur.record_event()
            with torch.cuda.stream(d2h):
                d2h.wait_event(done)
                out["boundary"][g0:g1].copy_(uni, non_blocking=True)
            keep.append((enc, per, r, start[g0:g1], counts[g0:g1], g0))
        with torch.cuda.stream(d2h):
            d2h.wait_event(cur.record_event())
            if plane_status is not None:
                out["rle_status"].copy_(plane_status, non_blocking=True)
            else:
                out["rle_status"].zero_()
            out["boundary_status"].copy_(st["bstatus"], non_blocking=True)
            st["ready"] = d2h.record_event()
        st["free"] = cur.record_event()
        if not async_result:
            cur.wait_stream(d2h)
        self.__dict__.setdefault("_keeps", {})[slot] = keep
        return out, keep
