#!/usr/bin/env python
"""bench.py — refined instances/s of the maskrefiner hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c1|c2|c3|c4|c5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of synthetic input per GPU.  Default workload = C2 of
BASELINE.json (`configs[1]`, the configuration the metric is quoted on): 32 OCID-shaped 640x480 frames, 30 initial
masks each, fp32:
    K0 encode initial masks -> K2 per-instance boundaries -> K1 MSDeformAttn forward between its Linear layers (softmax over
    L*P, reference-point arithmetic and the multi-scale sampling in one kernel; 6 encoder layers of the optional pixel
    decoder, S=6300, 8 heads x 32, 3 levels x 4 points) -> K3/K4 fused up-sample / error-map decode / centre NMS /
    grouping / panoptic merge / per-instance masks on synthetic head outputs, writing everything the predictor returns.
The other BASELINE.json configs have their own lines (--config): c3 = OSD clutter, 60 masks per frame, bf16 value tensor;
c4 = armbench 750x1333 frames, Swin-T levels (S = 21 168), 8 frames; c5 = the 4 096-frame sweep (total work fixed).
The cuDNN network between encode and decode (backbone, DeepLabV3+ decoder, 3x3 heads: SURVEY §8-a A3-A5, "stays PyTorch")
and the decoder's GEMMs are NOT part of this step — neither is in the reference arm, whose definition in BASELINE.json is
"ms_deform_attn_core_pytorch plus the torch/numpy boundary and fusion code" — they are reported as the `network` and
`encoder` asides of the same line.  Random-init predictors give no instance (SURVEY §8-d hazard), hence synthetic head
outputs.

Prints ONE JSON line (rank 0).  `value` = instances all ranks produced / max-over-ranks device time with inputs
resident in HBM; `e2e` = the same through HotPath.refine_batch_host (pinned host byte masks, H2D + D2H inside the timed
region), `e2e_rle` = through HotPath.refine_batch_rle (the masks as COCO RLE, the form the reference's datasets hold).
`--impl reference` times the CPU implementation of the path (the reference's own Python when /root/reference is readable,
else the C port oracle/quber_oracle.c) on all host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "refined_instances_per_sec"
UNIT = "instances/s"
N_LAYERS = 6
NSETS = 2      # distinct input sets alternated between steps (each far larger than L2)
M_HEADS, D_HEAD, N_POINTS = 8, 32, 4

# BASELINE.json configs (SURVEY §8-d "Config -> concrete run").  H x W = the frames; Wp = zero-padded row pitch of the
# mask planes (W % 16 == 0 for the kernels); head = the 1/4-resolution head outputs' size.
CONFIGS = {
    "c1": dict(B=1, K=20, H=480, W=640, axes=(20, 60), margin=60, levels=((60, 80), (30, 40), (15, 20)), value="f32",
               name="C1: ONE 640x480 RGB-D frame per step, 20 initial masks, fp32 — the reference's own CPU-runnable case on "
                    "the GPU at batch 1 (BASELINE.json configs[0])"),
    "c2": dict(B=32, K=30, H=480, W=640, axes=(20, 60), margin=60, levels=((60, 80), (30, 40), (15, 20)), value="f32",
               name="C2: 32 frames 640x480 RGB-D per GPU per step, 30 initial masks each, fp32 (BASELINE.json configs[1])"),
    "c3": dict(B=32, K=60, H=480, W=640, axes=(10, 40), margin=40, levels=((60, 80), (30, 40), (15, 20)), value="bf16",
               name="C3: 32 OSD-shaped 640x480 frames per GPU per step, 60 initial masks each (dense clutter), bf16 value "
                    "tensor in the sampler (BASELINE.json configs[2])"),
    "c4": dict(B=8, K=30, H=750, W=1333, axes=(40, 120), margin=120, levels=((96, 168), (48, 84), (24, 42)), value="f32",
               name="C4: 8 armbench frames 750x1333 (1080x1920 resized, eval/refiner_model.py:228) per step, 30 masks each, "
                    "Swin-T pixel-decoder levels S = 21 168, fp32 (BASELINE.json configs[3])"),
    "c5": dict(B=32, K=30, H=480, W=640, axes=(20, 60), margin=60, levels=((60, 80), (30, 40), (15, 20)), value="f32",
               total_frames=4096,
               name="C5: 4 096-frame sweep, 640x480, 30 masks each, blocks of 32 frames dealt round-robin to the ranks "
                    "(sharding.frame_blocks), one all-gather of the refined label maps per block (BASELINE.json configs[4])"),
}


def geometry(c):
    """Derived sizes of a config: mask-plane pitch, head size, crop of the post outputs."""
    H, W = c["H"], c["W"]
    Wp = (W + 15) // 16 * 16
    h, w = (H + 3) // 4, (W + 3) // 4
    crop = None if (h * 4, w * 4) == (H, W) else (H, W)
    return dict(Wp=Wp, h=h, w=w, crop=crop, Hf=h * 4, Wf=w * 4, S=sum(a * b for a, b in c["levels"]),
                Ho=H, Wo=(W + 7) // 8 * 8)


def workload_config(name, n_gpus, extra=None):
    c, g = CONFIGS[name], geometry(CONFIGS[name])
    cfg = {
        "workload": c["name"], "config_id": name,
        "frames_per_gpu_per_step": c["B"], "masks_per_frame": c["K"], "height": c["H"], "width": c["W"],
        "stages": ["K0 encode", f"K2 boundary (ratio 0.01)",
                   f"K1 msda fused x{N_LAYERS} (S={g['S']},M={M_HEADS},D={D_HEAD},L={len(c['levels'])},P={N_POINTS},value {c['value']})",
                   "K3/K4 post (x4 upsample+decode+nms+group+merge+masks), writes sem_seg, eee_boundary, panoptic, masks, "
                   "one-hot error maps, label map"],
        "excluded": "cuDNN backbone/decoder/heads (SURVEY 8a A3-A5) and the decoder's GEMMs (asides `network`, `encoder`); "
                    "head outputs are synthetic",
        "l2": f"inputs larger than L2 ({NSETS} input sets alternated)",
        "sharding": f"{n_gpus} x independent image batches" + (", ONE all-gather (label maps + counts in one buffer) per step" if n_gpus > 1 else ""),
    }
    if extra:
        cfg.update(extra)
    return cfg


def make_inputs(name, seed0):
    """Synthetic step inputs as numpy: masks [B,K,H,Wp] (zero-padded pitch), head outputs at 1/4 resolution,
    MSDA operands (value, the module's offsets | logits rows, reference points; loc / attn for the CPU arm)."""
    from quber_b200 import synth
    c, g = CONFIGS[name], geometry(CONFIGS[name])
    B, K, H, W = c["B"], c["K"], c["H"], c["W"]
    masks = np.zeros((B, K, H, g["Wp"]), np.uint8)
    heads = []
    for b in range(B):
        m = synth.ellipse_masks(K, g["Hf"], g["Wf"], seed=seed0 + b, axes=c["axes"], margin=c["margin"])
        m[:, H:, :] = 0
        m[:, :, W:] = 0
        masks[b, :, :, :W] = m[:, :H, :W]
        heads.append(synth.head_outputs(m, seed=seed0 + b))
    head = {k: np.stack([h[k] for h in heads]) for k in heads[0]}
    msda = synth.msda_inputs(B=B, shapes=c["levels"], M=M_HEADS, D=D_HEAD, P=N_POINTS, mode="module", seed=seed0, operands=True)
    return masks, head, msda


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for t, line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                clk = float(f[1]); mx = float(f[2])
            except ValueError:
                continue
            if t0 - 0.05 <= t <= t1 + 0.05:
                sm.append(clk)
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        if not sm:      # timed region shorter than the sampling period: take every sample we have
            for t, line in self.lines:
                f = [x.strip() for x in line.split(",")]
                try:
                    sm.append(float(f[1])); mx = float(f[2])
                except (ValueError, IndexError):
                    pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------- ours
def ncu_metric(kernel_substr, metrics):
    """Per-launch averages of `metrics` for a kernel, from the newest committed `ncu --set full` summary under profiles/
    (scripts/ncu_summary.py; the capture is a separate run under the profiler, never this one).
    -> ({metric: value in base units} | {}, file name | None)."""
    import csv, glob, re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_summary.csv")),
                   key=lambda f: [int(x) for x in re.findall(r"\d+", os.path.basename(f))])
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for f in reversed(files):
        try:
            rows = list(csv.reader(open(f)))
            h, units = rows[0], rows[1]
            sel = [r for r in rows[2:] if kernel_substr in r[0]]
            if not sel:
                continue
            out = {}
            for m in metrics:
                if m in h:
                    i = h.index(m)
                    out[m] = sum(float(r[i]) * scale.get(units[i], 1.0) for r in sel) / len(sel)
            if out:
                return out, os.path.basename(f)
        except Exception:
            continue
    return {}, None


def algorithmic_bytes(name, inst_per_step):
    """Per launch, B frames (SURVEY §8-d formulas: inputs read once + outputs written once)."""
    c, g = CONFIGS[name], geometry(CONFIGS[name])
    B, K = c["B"], c["K"]
    HWp = c["H"] * g["Wp"]                    # mask planes (pitch-padded)
    HWo = g["Ho"] * g["Wo"]                   # post outputs
    hw = g["h"] * g["w"]
    S, L = g["S"], len(c["levels"])
    vb = 4 if c["value"] == "f32" else 2
    rows = B * S
    return {
        "encode": B * (K * HWp + 12 * HWp),                                   # masks in, 3 float planes out (12.9 MB/frame at C2)
        "boundary": B * (2 * K * HWp + 2 * HWp),                              # masks in, per-instance out, union + fg out
        # value + out (M*D channels) + the module's offsets | logits rows (M*L*P*3 floats) + reference points
        "msda_layer": rows * (M_HEADS * D_HEAD * vb * 2 + M_HEADS * L * N_POINTS * 3 * 4 + L * 2 * 4),
        # 8 low-resolution planes in; sem_seg + 4 eee planes + panoptic (float), one-hots (4 bytes), label map (1 byte),
        # one byte plane per refined instance
        "post": B * 8 * hw * 4 + B * HWo * (4 + 16 + 4 + 4 + 1) + int(inst_per_step) * HWo,
    }


def run_ours(args):
    import torch
    import torch.distributed as dist
    from quber_b200 import _lib, ops, sharding
    from quber_b200.pipeline import HotPath, RefinerConfig
    _lib.lib()     # fail loudly if the CUDA library is missing
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    name = args.config
    c, g = CONFIGS[name], geometry(CONFIGS[name])
    B, K, H, W, Wp = c["B"], c["K"], c["H"], c["W"], g["Wp"]
    L = len(c["levels"])
    # label maps leave the emit kernel as uint8 (at most top_k = 200 ids per frame; QUBER_POST_LABEL_U8): a quarter of the
    # bytes to write, to all-gather and to send to the host
    hp = HotPath(RefinerConfig(compact_label_map=True, crop=g["crop"], valid_w=W if Wp != W else 0), dev)
    steps = args.steps
    if "total_frames" in c:                                   # C5: the total work is fixed, the ranks share its blocks
        blocks = sharding.frame_blocks(c["total_frames"], B, rank, world)
        steps = len(blocks)

    sets = []
    for s_ in range(NSETS):
        masks, head, msda = make_inputs(name, 1000 * rank + 100 * s_)
        hm = torch.from_numpy(masks).pin_memory()
        hh = {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}
        md = {k: torch.from_numpy(v).to(dev) for k, v in msda.items() if k in ("value", "shapes", "starts", "oa", "ref")}
        if c["value"] == "bf16":
            md["value"] = md["value"].bfloat16()
        md["levels"] = tuple(tuple(int(v) for v in hw) for hw in c["levels"])   # the host copy of the level table
        d = dict(h_masks=hm, h_head=hh, masks=hm.to(dev), head={k: v.to(dev) for k, v in hh.items()}, msda=md)
        # the same masks as COCO RLE (the form the reference's datasets hold), encoded once by the product's own encoder
        cnt, off, st_ = ops.rle_encode(d["masks"].view(B * K, H, Wp))
        assert int(st_[0]) == 0
        n_runs = int(off[-1])
        d["h_rle_counts"] = cnt[:n_runs].cpu().pin_memory()
        d["h_rle_offsets"] = off.cpu().pin_memory()
        sets.append(d)
    payload = sharding.Payload(B, g["Ho"], g["Wo"], dev, world)
    side = torch.cuda.Stream(device=dev) if world > 1 else None
    d_bnd = ops.boundary_dilation(H, W, hp.cfg.boundary_ratio)

    stage_names = ["encode", "boundary", "msda", "post"]
    stage_ms = {k: 0.0 for k in stage_names}

    def step(i, evs=None):
        d = sets[i % NSETS]
        mark = (lambda: None) if evs is None else (lambda: evs.append(_ev()))
        mark()
        enc, occ = ops.encode_initial_masks(d["masks"], hp.cfg.sigma, hp.cfg.valid_w, with_occupancy=True)
        mark()
        per, uni, fg = ops.mask_boundary(d["masks"], d_bnd, occupancy=occ)
        mark()
        hp.msda_layers(d["msda"], N_LAYERS)
        mark()
        if world > 1:
            torch.cuda.current_stream().wait_stream(side)   # the previous step's gather has read the payload buffer
        r = hp.post(d["head"], label_map_out=payload.maps.view(-1))
        mark()
        if world > 1:
            # the ONE collective of the path (maps + counts in one buffer), on a side stream under the next step's kernels
            payload.counts.copy_(r["inst_offsets"][1:] - r["inst_offsets"][:-1])
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                payload.gather()
        return r

    def _ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident (value) ----
    for i in range(args.warmup):
        r = step(i)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    inst_counts, all_evs = [], []
    barrier()
    t0w = time.time()
    e0 = _ev()
    for i in range(steps):
        evs = []
        r = step(i, evs)
        all_evs.append(evs)
        inst_counts.append(r["inst_offsets"][-1:].clone())    # device tensor, read after the timed region
    if world > 1:
        torch.cuda.current_stream().wait_stream(side)         # the last gather is inside the timed region
    e1 = _ev()
    barrier()
    t1w = time.time()
    clocks = sampler.stop(t0w, t1w) if rank == 0 else None
    ms = e0.elapsed_time(e1)
    for evs in all_evs:
        for j, k in enumerate(stage_names):
            stage_ms[k] += evs[j].elapsed_time(evs[j + 1])
    n_inst = int(torch.stack(inst_counts).sum()) if inst_counts else 0
    status = int(r["status"][0]) | (int(ops.boundary_status(dev)[0]) << 4)
    max_steps = total_steps = steps
    if world > 1:
        t = torch.tensor([ms, float(n_inst), float(steps)], device=dev, dtype=torch.float64)
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, n_inst_total, max_steps, total_steps = float(tmax[0]), float(tsum[1]), int(tmax[2]), int(tsum[2])
    else:
        n_inst_total = float(n_inst)
    value = n_inst_total / (ms * 1e-3)

    # ---- end to end (host buffers in, host results out) ----
    # frames per pipeline chunk and smallest chunk of the halving tail: 8 / 2 -> 8 + 8 + 8 + 4 + 2 + 2.  encode / boundary
    # kernels once per group of chunks holding `kernel_frames` frames (profiles/r01_e2e_sweep_v16.txt: 16 is best at C2).
    E2E_CHUNK = int(os.environ.get("QUBER_E2E_CHUNK", "8"))
    E2E_TAIL = int(os.environ.get("QUBER_E2E_TAIL", "2"))
    E2E_KERNEL = int(os.environ.get("QUBER_E2E_KERNEL_FRAMES", "16"))
    E2E_BYTE_FRAMES = int(os.environ["QUBER_E2E_BYTE_FRAMES"]) if "QUBER_E2E_BYTE_FRAMES" in os.environ else None   # None: the rule

    # host calls: one payload per slot, the collective on the side stream (ranks are not made to meet on the main stream;
    # a slot's payload is rewritten two calls later, after its gather has read it)
    payloads_e2e = [sharding.Payload(B, g["Ho"], g["Wo"], dev, world) for _ in range(2)] if world > 1 else []
    gathered = [None, None]

    def gather_host_call(keep, slot):
        if world > 1:
            p = payloads_e2e[slot]
            cur = torch.cuda.current_stream()
            if gathered[slot] is not None:
                cur.wait_event(gathered[slot])
            p.maps.copy_(keep[0][2]["label_map"])              # the call's label maps are the graph pool's: one D2D copy
            p.counts.copy_(torch.cat([k[4] for k in keep]))
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                p.gather()
                gathered[slot] = side.record_event()

    def e2e_dense(i, slot=0, async_result=False):
        d = sets[i % NSETS]
        # decoder activations never live on the host (resident inputs); launched first so the sampling kernels run
        # under the host-side packing / host->device copies of refine_batch_host's chunk pipeline
        hp.msda_layers(d["msda"], N_LAYERS)
        o, keep = hp.refine_batch_host(d["h_masks"], d["h_head"], chunk_frames=E2E_CHUNK, min_tail=E2E_TAIL, kernel_frames=E2E_KERNEL,
                                       slot=slot, async_result=async_result, byte_frames=E2E_BYTE_FRAMES)
        hp.last_route_dense = dict(hp.last_route)
        gather_host_call(keep, slot)
        return o

    def e2e_rle(i, slot=0, async_result=False):
        d = sets[i % NSETS]
        hp.msda_layers(d["msda"], N_LAYERS)
        o, keep = hp.refine_batch_rle(d["h_rle_counts"], d["h_rle_offsets"], (B, K, H, Wp), d["h_head"], value=255,
                                      kernel_frames=E2E_KERNEL, slot=slot, async_result=async_result)
        gather_host_call(keep, slot)
        return o

    E2E_DEPTH = 2 if os.environ.get("QUBER_E2E_DEPTH", "2") != "1" else 1

    def time_e2e(fn, depth=E2E_DEPTH):
        """K host calls, every one with its host->device copies and the device->host read of its results inside the timed
        region.  depth 1: the host reads call i's results before it issues call i+1.  depth 2 (default): the caller keeps
        two calls in flight on alternating slots — call i+1 is issued (its packing, copies and kernels queued) before the
        results of call i are read, so the copies of one call run under the kernels of the other; every call's results
        are still read and checked on the host inside the timed region."""
        for i in range(max(2, min(args.warmup, 3))):
            fn(i, i % depth, depth > 1)
            o = hp.wait_result(i % depth)
        torch.cuda.synchronize()
        barrier()
        n_steps = max(2, min(steps, 10))
        e0 = _ev()
        inst = 0
        pending = []
        host = dict(issue=0.0, wait=0.0)

        def collect():
            nonlocal inst
            t = time.perf_counter()
            o = hp.wait_result(pending.pop(0))                 # the caller reads the host result of every step
            host["wait"] += time.perf_counter() - t
            HotPath.check_status(o)
            inst += int(o["inst_counts"].sum())
            return o
        for i in range(n_steps):
            t = time.perf_counter()
            fn(i, i % depth, depth > 1)
            host["issue"] += time.perf_counter() - t
            pending.append(i % depth)
            if len(pending) >= depth:
                o = collect()
        while pending:
            o = collect()
        if world > 1:
            torch.cuda.current_stream().wait_stream(side)      # the last gather is inside the timed region
        torch.cuda.current_stream().synchronize()
        e1 = _ev()
        barrier()
        time_e2e.host_ms = {k: v * 1e3 / n_steps for k, v in host.items()}
        t_ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([t_ms, float(inst)], device=dev, dtype=torch.float64)
            tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
            t_ms, inst = float(tmax[0]), float(tsum[1])
        d2h = sum(v.numel() * v.element_size() for v in o.values())
        return inst / (t_ms * 1e-3), t_ms / n_steps, n_steps, d2h

    # the collective's transport, and (outside every timed region) its result checked against a plain NCCL all-gather
    gather_info = None
    if world > 1:
        torch.cuda.synchronize()
        want = torch.empty((world, payload.nbytes), dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(want.view(-1), payload.buf)
        payload.gather()
        torch.cuda.synchronize()
        same = bool(torch.equal(payload.all, want))
        gather_info = {"gather_transport": payload.transport, "gather_verified_against_nccl": same,
                       "gather_bytes_per_rank": int(payload.nbytes)}
        if getattr(payload, "peer_copy_error", None):
            gather_info["peer_copy_unavailable"] = payload.peer_copy_error[:200]
        if not same:
            raise RuntimeError("the gathered payload differs from an NCCL all-gather of the same buffers")
    e2e_val, e2e_ms, e2e_steps, d2h = time_e2e(e2e_dense)
    e2e_host = dict(time_e2e.host_ms)
    rle_val, rle_ms, _, _ = time_e2e(e2e_rle)
    rle_host = dict(time_e2e.host_ms)
    e2e_val1, e2e_ms_sync, _, _ = time_e2e(e2e_dense, 1) if E2E_DEPTH > 1 else (e2e_val, e2e_ms, 0, 0)
    rle_val1, rle_ms_sync, _, _ = time_e2e(e2e_rle, 1) if E2E_DEPTH > 1 else (rle_val, rle_ms, 0, 0)
    # Both ways of driving the same public call are measured (two calls in flight on alternating slots / one at a time);
    # the line's `e2e` is the faster one on this host and says which (`calls_in_flight`), the other figure stays beside it.
    e2e_ms2, rle_ms2 = e2e_ms, rle_ms
    e2e_depth = rle_depth = E2E_DEPTH
    if e2e_val1 > e2e_val:
        e2e_val, e2e_ms, e2e_depth = e2e_val1, e2e_ms_sync, 1
    if rle_val1 > rle_val:
        rle_val, rle_ms, rle_depth = rle_val1, rle_ms_sync, 1
    packed = bool(hp.pack_masks)
    head_bytes = sum(v.numel() * 4 for v in sets[0]["h_head"].values())
    route = getattr(hp, "last_route_dense", dict(byte_frames=B, packed_frames=0))
    fb = sets[0]["h_masks"][0].numel()
    h2d = route["byte_frames"] * fb + route["packed_frames"] * fb // 8 + head_bytes
    h2d_rle = sets[0]["h_rle_counts"].numel() * 4 + sets[0]["h_rle_offsets"].numel() * 8 + head_bytes
    # what this box's host link gives a plain pinned copy of the same mask bytes (the floor under the dense call; it differs
    # between boxes of the pool by NUMA placement of the pinned pages)
    pcie = {}
    if rank == 0:
        hm = sets[0]["h_masks"]
        dm = torch.empty(hm.shape, dtype=hm.dtype, device=dev)
        back = torch.empty(hm.shape, dtype=hm.dtype).pin_memory()
        for nm, fn in (("h2d_gbs", lambda: dm.copy_(hm, non_blocking=True)), ("d2h_gbs", lambda: back.copy_(dm, non_blocking=True))):
            fn(); torch.cuda.synchronize()
            c0 = _ev()
            for _ in range(3):
                fn()
            c1 = _ev()
            torch.cuda.synchronize()
            pcie[nm] = 3 * hm.numel() / (c0.elapsed_time(c1) * 1e-3) / 1e9
        # host DRAM rate seen by the packing threads (one read of the mask bytes): the other floor of the dense call
        t0 = time.perf_counter()
        ops.host_pack_masks(hm, threads=hp.pack_threads)
        pcie["host_pack_gbs"] = hm.numel() / (time.perf_counter() - t0) / 1e9
        pcie["h2d_floor_ms_per_step"] = h2d / (pcie["h2d_gbs"] * 1e9) * 1e3
        pcie["dense_bytes_floor_ms_per_step"] = hm.numel() / (pcie["h2d_gbs"] * 1e9) * 1e3
        del dm, back

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs, peak_kind = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    inst_per_step = n_inst_total / max(1, total_steps)        # per GPU and step
    ab = algorithmic_bytes(name, inst_per_step)
    per_stage = {k: stage_ms[k] / steps for k in stage_names}
    msda_kernel = "msda_fused_p4_kernel"
    roof_all = {
        "encode(reduce+emit)": {"ms": per_stage["encode"], "bytes": ab["encode"], "launches": 2},
        "boundary_tma_kernel": {"ms": per_stage["boundary"], "bytes": ab["boundary"], "launches": 1},
        msda_kernel: {"ms": per_stage["msda"] / N_LAYERS, "bytes": ab["msda_layer"], "launches": 1},
        "post(nms+select+group+eee+relabel+offsets+emit+finalize)": {"ms": per_stage["post"], "bytes": ab["post"], "launches": 8},
    }
    for v in roof_all.values():
        v["achieved_gbs"] = v["bytes"] / (v["ms"] * 1e-3) / 1e9
        v["frac"] = v["achieved_gbs"] / peak_gbs
    dominant = max(roof_all, key=lambda k: roof_all[k]["ms"] * (N_LAYERS if k == msda_kernel else 1))
    dv = roof_all[dominant]
    ncu, ncu_src = ncu_metric("msda_f", ["dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_pipe_lsu_wavefronts.sum",
                                         "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"])
    traffic = (ncu.get("dram__bytes_read.sum", 0) + ncu.get("dram__bytes_write.sum", 0)) or None
    # The L1 floor of the gather: every (query, head, level, point) pulls 4 corner lines of 128 bytes (D = 32 fp32; one line
    # per corner for bf16 as well) through an L1 data stage that returns one 128-byte wavefront per clock and SM.
    sm_mhz = (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
    n_lines = B * g["S"] * M_HEADS * L * N_POINTS * 4
    l1_floor_us = n_lines / 148.0 / (sm_mhz * 1e6) * 1e6
    launches = (2 + 1 + N_LAYERS + 8) * steps
    bound = "l1" if dominant == msda_kernel else "hbm"
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": args.warmup,
        "ms_per_step": ms / max_steps, "higher_is_better": True, "scaling": "strong" if "total_frames" in c else "weak",
        "vs_baseline": None, "dtype": "f32" if c["value"] == "f32" else "f32 (bf16 value tensor)",
        "data": "synthetic", "config": workload_config(name, world, gather_info), "clocks": clocks,
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms, "steps": e2e_steps, "host_link": pcie,
                "calls_in_flight": e2e_depth, "ms_per_step_one_call_in_flight": e2e_ms_sync, "ms_per_step_two_calls_in_flight": e2e_ms2,
                "host_ms_per_step": e2e_host,
                "input": "dense uint8 masks [B,K,H,W] in pinned host memory (MaskRefinerPredictor.predict's contract)",
                "mask_transfer": ("%d frames as bytes through the copy engine while %d host threads bit-pack the other %d (one bit per pixel), "
                                  "inside the timed region" % (route["byte_frames"], hp.pack_threads, route["packed_frames"])) if packed
                                 else "bytes",
                "route_rule": hp.pack_rule, "route": getattr(hp, "last_route_dense", None), "graphs": bool(hp.use_graphs),
                # the dense call has to read every mask byte on the host once (packing) or push it over the link (bytes)
                "bound": "host (reading the dense mask bytes: DRAM rate of the packing threads / host link)"
                         if pcie and e2e_ms < 2.0 * max(pcie.get("h2d_floor_ms_per_step", 0.0),
                                                        sets[0]["h_masks"].numel() / (pcie.get("host_pack_gbs", 1e9) * 1e9) * 1e3)
                         else "kernels",
                "schedule": {"chunk_frames": E2E_CHUNK, "min_tail": E2E_TAIL, "kernel_frames": E2E_KERNEL, "label_map": "uint8"}},
        "e2e_rle": {"value": rle_val, "unit": UNIT, "ms_per_step": rle_ms, "calls_in_flight": rle_depth,
                    "ms_per_step_one_call_in_flight": rle_ms_sync, "ms_per_step_two_calls_in_flight": rle_ms2, "host_ms_per_step": rle_host, "h2d_bytes_per_step": int(h2d_rle),
                    "d2h_bytes_per_step": int(d2h), "graphs": bool(hp.use_graphs),
                    "input": "the same masks as COCO RLE counts (register_uoais_sim_panoptic.py:121-144) in pinned host memory, "
                             "decoded on the device by quber_rle_decode",
                    "rle_runs_per_step": int(sets[0]["h_rle_counts"].numel())},
        "gpu_launches": launches,
        "instances_per_step": inst_per_step * world, "status": status,
        "stage_ms": per_stage,
        "roofline": {"kernel": dominant, "bound": bound, "achieved": dv["achieved_gbs"], "peak": peak_gbs, "unit": "GB/s",
                     "frac": dv["frac"], "traffic": traffic if dominant == msda_kernel else None, "traffic_source": ncu_src,
                     "peak_kind": peak_kind, "algorithmic_bytes_per_launch": dv["bytes"], "avg_launch_ms": dv["ms"],
                     "l1_floor": {"lines_per_launch": n_lines, "floor_us": l1_floor_us,
                                  "frac_of_l1_floor": l1_floor_us / (dv["ms"] * 1e3) if dominant == msda_kernel else None,
                                  "ncu_lsu_wavefronts_per_launch": ncu.get("l1tex__data_pipe_lsu_wavefronts.sum"),
                                  "ncu_shared_wavefronts_per_launch": ncu.get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"),
                                  "sm_mhz": sm_mhz},
                     "note": ("the sampling kernel is bounded by the L1 data stage, not by HBM: 48 gathered 128-byte lines per "
                              "(query, head) through 128 B/clk/SM (DESIGN.md 3.1); `frac` is the HBM fraction the spec asks for, "
                              "`l1_floor.frac_of_l1_floor` the fraction of the bound that applies") if dominant == msda_kernel else None},
        "roofline_all": roof_all,
    }
    if world == 1 and not args.no_network and name in ("c2", "c3"):
        enc0 = ops.encode_initial_masks(sets[0]["masks"], hp.cfg.sigma)
        line["network"] = network_side_measurement(dev, enc0, inst_per_step, ms / steps, c)
        line["encoder"] = encoder_side_measurement(dev, c, peaks)
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(name, budget_s=args.cpu_budget)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------- network (disclosed aside)
def network_side_measurement(dev, enc, inst_per_step, hot_ms, c):
    """SURVEY §8-d's "Net" stage, reported NEXT TO the headline and never inside it: the cuDNN network between
    encode and decode (two ResNet-50 streams, DeepLabV3+ decoder, hierarchical fusion heads; random-init weights of the
    uoais-sim config) on the same batch, fp32 and bf16-autocast, and the instances/s the step would give
    with it added (random-init predictors yield no instance, so the post stage keeps its synthetic head outputs)."""
    import torch
    from quber_b200 import config as qcfg, ops
    from quber_b200.meta_arch import build_model
    out = {}
    try:
        cfg = qcfg.get_cfg()
        cfg.merge_from_file(os.path.join(ROOT, "configs", "mask-refiner-rgbd-concat-l2-gn-hf-b-fco-l3-b8.yaml"))
        cfg.MODEL.DEVICE = str(dev)
        torch.manual_seed(0)
        model = build_model(cfg, "fp32").eval()
        g = torch.Generator(device=dev).manual_seed(0)
        img = torch.randint(0, 256, (c["B"], 6, c["H"], c["W"]), dtype=torch.uint8, device=dev, generator=g)
        x = ops.pack_input(img, enc, model.pixel_mean, model.pixel_std)
        model.use_cuda_graph = True
        for prec in ("fp32", "bf16"):
            model.set_precision(prec)
            with torch.no_grad():
                for _ in range(2):
                    model.network_logits(x)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(3):
                    model.network_logits(x)
                e1.record()
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 3
            out[f"ms_per_step_{prec}"] = ms
            out[f"instances_per_sec_with_network_{prec}"] = inst_per_step / ((hot_ms + ms) * 1e-3)
        out["note"] = ("network forward for the same frames (cuDNN convolutions with folded BN and fused epilogues, own "
                       "GroupNorm+ReLU / upsample+concat / fuse-concat / tcgen05 1x1 kernels, CUDA graph), outside `value` and outside the reference arm "
                       "(BASELINE.json times the CPU path without the network); fp32 = PyTorch defaults (cuDNN convolutions may use TF32)")
    except Exception as e:      # the aside must never cost the headline line
        out["error"] = f"{type(e).__name__}: {e}"[:300]
    return out


# --------------------------------------------------------------------------------------------- encoder (disclosed aside)
def encoder_side_measurement(dev, c, peaks):
    """The caller of the sampling kernel: the 6-layer MSDeformAttn transformer encoder of the optional
    pixel decoder on the same frames (d_model 256, FFN 1024) — tcgen05 GEMMs, the fused sampling kernel, fused residual +
    LayerNorm — in the reference's precision (fp32 accuracy: split-operand 3xTF32 GEMMs), TF32 and bf16, with the GEMMs'
    share against the measured tensor peak."""
    import torch
    out = {}
    try:
        from torch.profiler import profile, ProfilerActivity
        from quber_b200.pixel_decoder import MSDeformAttnTransformerEncoderOnly
        torch.manual_seed(0)
        L = len(c["levels"])
        enc = MSDeformAttnTransformerEncoderOnly(256, M_HEADS, N_LAYERS, 1024, 0.0, "relu", L, N_POINTS).to(dev).eval()
        shapes = list(c["levels"])[::-1]                      # res5 -> res3 order (msdeformattn.py:318-324)
        B = c["B"]
        # channels-last memory, as the network's input_proj convolutions hand the level maps over (the flatten is then a view)
        srcs = [torch.randn(B, 256, h, w, device=dev).contiguous(memory_format=torch.channels_last) for h, w in shapes]
        pos = [(torch.randn(B, 256, h, w, device=dev) * 0.1).contiguous(memory_format=torch.channels_last) for h, w in shapes]
        rows = B * sum(h * w for h, w in shapes)
        flops = 2.0 * rows * 256 * (256 + 288 + 256 + 1024 + 1024) * N_LAYERS      # value, offsets|weights, output, FFN1, FFN2
        with torch.no_grad():
            for prec in ("fp32", "tf32", "bf16"):
                enc.set_precision(prec)
                for _ in range(2):
                    enc(srcs, pos)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(3):
                    enc(srcs, pos)
                e1.record()
                torch.cuda.synchronize()
                out[f"ms_per_step_{prec}"] = e0.elapsed_time(e1) / 3
                with profile(activities=[ProfilerActivity.CUDA]) as prof:
                    enc(srcs, pos)
                    torch.cuda.synchronize()
                by = {}
                for ev in prof.key_averages():
                    t_us = getattr(ev, "device_time_total", None)
                    if t_us is None:
                        t_us = getattr(ev, "cuda_time_total", 0.0)
                    for key, names in (("gemm_bias_act", ("gemm_bias_act", "gemm_wres")), ("msda_fused", ("msda_fused",)),
                                       ("add_layer_norm", ("add_layer_norm",))):
                        if any(nm in ev.key for nm in names):            # both GEMM kernels (csrc/gemm.cu) under one key
                            by[key] = by.get(key, 0.0) + t_us * 1e-3
                gemm_ms = by.get("gemm_bias_act", 0.0)
                issued = flops * (3 if prec == "fp32" else 1)
                out[f"kernels_ms_{prec}"] = {k: round(v, 3) for k, v in by.items()}
                if gemm_ms > 0:
                    out[f"gemm_tflops_{prec}"] = {"algorithmic": flops / (gemm_ms * 1e-3) / 1e12, "issued": issued / (gemm_ms * 1e-3) / 1e12}
        if "bf16_tflops_sustained" in peaks and "gemm_tflops_bf16" in out:
            out["gemm_frac_of_measured_bf16_peak"] = out["gemm_tflops_bf16"]["algorithmic"] / peaks["bf16_tflops_sustained"]
        out["note"] = ("6 encoder layers for the config's frames, every Linear / FFN GEMM on quber_gemm_bias_act(_x3), sampling on "
                       "quber_msda_forward_fused, residual + LayerNorm on quber_add_layer_norm; no cuBLAS call; level maps in channels-last memory "
                       "(what the network hands over); outside `value`")
    except Exception as e:
        out["error"] = f"{type(e).__name__}: {e}"[:300]
    return out


# --------------------------------------------------------------------------------------------- CPU legs
def _oracle_frame(a):
    """The whole hot path for ONE frame on ONE core through the oracle port."""
    import oracle
    masks, low, msda, crop = a
    oracle.encode_initial_masks(masks)
    oracle.masks_to_boundary(masks, 0.01, return_per_instance=True)
    for _ in range(N_LAYERS):
        oracle.msda_forward(msda["value"], msda["shapes"], msda["starts"], msda["loc"], msda["attn"])
    full = oracle.upsample_bilinear(low, 4)
    full[2:4] *= 4.0
    if crop is not None:
        full = np.ascontiguousarray(full[:, :crop[0], :crop[1]])
    r = oracle.post_process(full[0], full[1], full[2:4], full[4:8])
    return r["n_instances"]


def _cpu_frames(name, n):
    from quber_b200 import synth
    c, g = CONFIGS[name], geometry(CONFIGS[name])
    frames = []
    for b in range(n):
        m = synth.ellipse_masks(c["K"], g["Hf"], g["Wf"], seed=b, axes=c["axes"], margin=c["margin"])
        m[:, c["H"]:, :] = 0
        m[:, :, c["W"]:] = 0
        ho = synth.head_outputs(m, seed=b)
        low = np.concatenate([ho["foreground"], ho["center"], ho["offset"], ho["eee_boundary"]], 0)
        msda = synth.msda_inputs(B=1, shapes=c["levels"], M=M_HEADS, D=D_HEAD, P=N_POINTS, mode="module", seed=b)
        frames.append((np.ascontiguousarray(m[:, :c["H"], :c["W"]]), low, msda, g["crop"]))
    return frames


def _run_cpu(frames, threads):
    from concurrent.futures import ThreadPoolExecutor
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:          # ctypes releases the GIL: real parallelism over frames
        n = sum(ex.map(_oracle_frame, frames))
    return n, time.perf_counter() - t0


def cpu_baseline(name, budget_s=20.0):
    import oracle
    oracle.lib()
    c = CONFIGS[name]
    cores = len(os.sched_getaffinity(0))
    frames = _cpu_frames(name, min(cores, c["B"]))
    _run_cpu(frames[:1], 1)                                  # warm-up (page-in)
    n1, t1 = _run_cpu(frames[:1], 1)
    reps = max(1, int(budget_s / max(t1 * max(1.0, len(frames) / cores), 1e-3)) // 2)
    n, t = 0, 0.0
    for _ in range(min(reps, 3)):
        a, b_ = _run_cpu(frames, cores)
        n += a; t += b_
    out = {"value": n / t, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"{len(frames)} frames of the same workload ({name}: K={c['K']}, {c['W']}x{c['H']}, msda x{N_LAYERS}), one frame per "
                     f"thread, {min(reps, 3)} repetition(s), oracle/quber_oracle.c; single-core: {n1 / t1:.2f} instances/s"}
    # the reference's OWN Python cannot run on the GPU box (no /root/reference there): quote the line recorded where it can
    rec = os.path.join(ROOT, "profiles", "r02_bench_ref_reference_python_build_container.json")
    if name == "c2" and os.path.exists(rec):
        try:
            txt = open(rec).read()
            r = json.loads(txt[txt.index("{"):])
            out["reference_python_recorded"] = {"value": r["value"], "unit": r["unit"], "cores": r["cpu_baseline"]["cores"],
                                                "where": "build container (bench.py --impl reference with /root/reference readable)",
                                                "source": os.path.basename(rec)}
        except Exception:
            pass
    return out


def _reference_python_available():
    return os.path.isdir("/root/reference/maskrefiner") and os.path.exists(os.path.join(ROOT, "oracle", "time_reference_cpu.py"))


def run_reference(args):
    """--impl reference: the CPU implementation of the path on the box's host cores.  When /root/reference is readable
    (the build container) this is the reference's OWN Python (numpy / cv2 / torch: predictor encode block, cv2 erosions,
    ms_deform_attn_core_pytorch, F.interpolate + get_panoptic_segmentation + instance loop + argmax decode), imported where
    it lies by oracle/time_reference_cpu.py; on the GPU box, which has no /root/reference, the C port
    (oracle/quber_oracle.c — pinned to the reference's outputs by tests/test_oracle_golden.py)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    name = args.config
    c = CONFIGS[name]
    cores = len(os.sched_getaffinity(0))
    if _reference_python_available() and not args.port:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import time_reference_cpu as trc
        for _ in range(args.warmup):
            trc.reference_frame(c, geometry(c), 0, cores, N_LAYERS)
        n, t = 0, 0.0
        for i in range(args.steps):
            a, b_ = trc.reference_frame(c, geometry(c), i, cores, N_LAYERS)
            n += a; t += b_
        value, kind, frames_per_step = n / t, "reference_python", 1
        sample = (f"each step = ONE frame of the {name} workload through the reference's own Python (numpy / cv2 / torch ops "
                  f"threaded over {cores} host threads); /root/reference imported where it lies")
    else:
        import oracle
        oracle.lib()
        frames = _cpu_frames(name, min(cores, c["B"]))
        _run_cpu(frames[:1], 1)
        n1, t1 = _run_cpu(frames[:1], 1)
        est_step = t1 * max(1.0, len(frames) / cores) * 1.3
        total_steps = args.steps + args.warmup
        if est_step * total_steps > 150.0:                        # keep the whole run within a few minutes
            keep = max(1, int(len(frames) * 150.0 / (est_step * total_steps)))
            frames = frames[:keep]
        for _ in range(args.warmup):
            _run_cpu(frames, cores)
        n, t = 0, 0.0
        for _ in range(args.steps):
            a, b_ = _run_cpu(frames, cores)
            n += a; t += b_
        value, kind, frames_per_step = n / t, "port", len(frames)
        sample = (f"each step = {len(frames)} frames of the {name} workload (K={c['K']}, {c['W']}x{c['H']}, msda x{N_LAYERS}), one "
                  f"frame per thread on {cores} host threads; C port of the reference's numpy/cv2/torch path (the reference's "
                  "Python is not on this box)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True,
            "scaling": "strong" if "total_frames" in c else "weak",
            "vs_baseline": None, "dtype": "f32" if c["value"] == "f32" else "f32 (bf16 value tensor)", "data": "synthetic",
            "config": workload_config(name, args.gpus, {"reference_frames_per_step": frames_per_step,
                                                        "reference_note": "the metric is normalised per instance; the CPU arm's step "
                                                                          "is a bounded sample of the GPU arm's frames_per_gpu_per_step"}),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS), help="BASELINE.json workload (default: the one the metric is quoted on)")
    ap.add_argument("--port", action="store_true", help="--impl reference: time the C port even when /root/reference is readable")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-network", action="store_true", help="skip the disclosed network / encoder asides")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
