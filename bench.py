#!/usr/bin/env python
"""bench.py — refined instances/s of the maskrefiner hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of synthetic input per GPU — workload C2 of
BASELINE.json (`configs[1]`): 32 OCID-shaped 640x480 frames, 30 initial masks each, fp32:
    K0 encode initial masks -> K2 per-instance boundaries -> K1 MSDeformAttn sampling (6 encoder layers of the
    optional pixel decoder, S=6300, 8 heads x 32, 3 levels x 4 points) -> K3/K4 fused up-sample / error-map
    decode / centre NMS / grouping / panoptic merge / per-instance masks on synthetic head outputs.
The cuDNN network between encode and decode (backbone, DeepLabV3+ decoder, 3x3 heads: SURVEY §8-a A3-A5,
"stays PyTorch") is NOT part of this step; it is named in config.excluded.  Random-init predictors give no
instance (SURVEY §8-d hazard), hence synthetic head outputs.

Prints ONE JSON line (rank 0).  `value` = instances all ranks produced / max-over-ranks device time with inputs
resident in HBM; `e2e` = the same through HotPath.refine_batch_host (pinned host buffers, H2D + D2H inside the
timed region).  `--impl reference` times the CPU oracle port (oracle/quber_oracle.c, all host cores) on a
bounded sample of the same workload.
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
B, K, H, W = 32, 30, 480, 640
N_LAYERS = 6
NSETS = 2      # distinct input sets alternated between steps (each far larger than L2)


def workload_config(n_gpus):
    return {
        "workload": "C2: 32 frames 640x480 RGB-D per GPU per step, 30 initial masks each, fp32 "
                    "(BASELINE.json configs[1])",
        "frames_per_gpu_per_step": B, "masks_per_frame": K, "height": H, "width": W,
        "stages": ["K0 encode", "K2 boundary d=8", f"K1 msda x{N_LAYERS} (S=6300,M=8,D=32,L=3,P=4)",
                   "K3/K4 post (x4 upsample+decode+nms+group+merge+masks)"],
        "excluded": "cuDNN backbone/decoder/heads (SURVEY 8a A3-A5); head outputs are synthetic",
        "l2": f"inputs larger than L2: masks 295 MB + msda 440 MB per step, {NSETS} input sets alternated",
        "sharding": f"{n_gpus} x independent image batches" + (", one all_gather_into_tensor of label maps per step" if n_gpus > 1 else ""),
    }


def make_inputs(seed0):
    """Synthetic step inputs as numpy."""
    from quber_b200 import synth
    masks = np.stack([synth.ellipse_masks(K, H, W, seed=seed0 + b) for b in range(B)])
    heads = [synth.head_outputs(masks[b], seed=seed0 + b) for b in range(B)]
    head = {k: np.stack([h[k] for h in heads]) for k in heads[0]}
    msda = synth.msda_inputs(B=B, mode="module", seed=seed0)
    return masks, head, msda


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
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
def ncu_dram_bytes(kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of a kernel, from the newest committed `ncu --set full`
    summary under profiles/ (scripts/ncu_summary.py; the capture is a separate run under the profiler, never this one).
    -> (bytes | None, file name | None)."""
    import csv, glob, re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_v*_ncu_full_summary.csv")),
                   key=lambda f: [int(x) for x in re.findall(r"\d+", os.path.basename(f))])
    for f in reversed(files):
        try:
            rows = list(csv.reader(open(f)))
            h, units = rows[0], rows[1]
            ir, iw = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            vals = [float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]] for r in rows[2:] if kernel_substr in r[0]]
            if vals:
                return sum(vals) / len(vals), os.path.basename(f)
        except Exception:
            continue
    return None, None


def algorithmic_bytes():
    """Per launch, B frames (SURVEY §8-d formulas)."""
    HW = H * W
    S = 60 * 80 + 30 * 40 + 15 * 20
    return {
        "encode_reduce": B * (K * HW + HW),                       # masks in, last-mask map out
        "encode_emit": B * (HW + 12 * HW),                        # last-mask map in, 3 float planes out
        "encode": B * (K * HW + 12 * HW),                         # SURVEY figure for the pair (12.9 MB/frame)
        "boundary": B * (2 * K * HW + 2 * HW),                    # masks in, per-instance out, union + fg out
        "msda_layer": B * (S * 256 * 4 * 2 + S * 8 * 12 * 2 * 4 + S * 8 * 12 * 4),   # value + out + loc + attn (20.2 MB/frame)
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
    # label maps leave the emit kernel as uint8 (at most top_k = 200 ids per frame; QUBER_POST_LABEL_U8): 9.8 MB per step
    # instead of 39 MB to write, to all-gather and to send to the host
    hp = HotPath(RefinerConfig(compact_label_map=os.environ.get("QUBER_LABEL_U8", "1") == "1"), dev)

    sets = []
    for s in range(NSETS):
        masks, head, msda = make_inputs(1000 * rank + 100 * s)
        hm = torch.from_numpy(masks).pin_memory()
        hh = {k: torch.from_numpy(v).pin_memory() for k, v in head.items()}
        sets.append(dict(h_masks=hm, h_head=hh, masks=hm.to(dev), head={k: v.to(dev) for k, v in hh.items()},
                         msda={k: torch.from_numpy(v).to(dev) for k, v in msda.items()}))
    compact = hp.cfg.max_instances_per_image <= 255          # int32 maps (QUBER_LABEL_U8=0) are narrowed before the gather
    gather_buf = torch.empty((world, B, H, W), dtype=torch.uint8 if compact else torch.int32, device=dev) if world > 1 else None
    side = torch.cuda.Stream(device=dev) if world > 1 else None

    stage_names = ["encode", "boundary", "msda", "post"]
    stage_ms = {k: 0.0 for k in stage_names}

    def step(s, timed, evs=None):
        d = sets[s % NSETS]
        mark = (lambda: None) if evs is None else (lambda: evs.append(_ev()))
        mark()
        enc = ops.encode_initial_masks(d["masks"], hp.cfg.sigma)
        mark()
        per, uni, fg = ops.mask_boundary(d["masks"], ops.boundary_dilation(H, W, hp.cfg.boundary_ratio))
        mark()
        hp.msda_layers(d["msda"], N_LAYERS)
        mark()
        r = hp.post(d["head"])
        mark()
        if world > 1:
            # the one collective of the path, on a side stream so that it runs under the next step's kernels
            counts = r["inst_offsets"][1:] - r["inst_offsets"][:-1]
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                sharding.gather_label_maps(r["label_map"], counts, out=gather_buf, compact=compact)
            r["label_map"].record_stream(side)
            counts.record_stream(side)
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
        r = step(i, False)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    inst_counts = []
    all_evs = []
    barrier()
    t0w = time.time()
    e0 = _ev()
    for i in range(args.steps):
        evs = []
        r = step(i, True, evs)
        all_evs.append(evs)
        inst_counts.append(r["inst_offsets"][-1:])       # device tensor, read after the timed region
    if world > 1:
        torch.cuda.current_stream().wait_stream(side)       # the last gather is inside the timed region
    e1 = _ev()
    barrier()
    t1w = time.time()
    clocks = sampler.stop(t0w, t1w) if rank == 0 else None
    ms = e0.elapsed_time(e1)
    for evs in all_evs:
        for j, k in enumerate(stage_names):
            stage_ms[k] += evs[j].elapsed_time(evs[j + 1])
    n_inst = int(torch.stack(inst_counts).sum())
    status = int(r["status"][0])
    if world > 1:
        t = torch.tensor([ms, float(n_inst)], device=dev, dtype=torch.float64)
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, n_inst_total = float(tmax[0]), float(tsum[1])
    else:
        n_inst_total = float(n_inst)
    value = n_inst_total / (ms * 1e-3)

    # ---- end to end (host buffers in, host results out) ----
    outs_host = None
    # frames per pipeline chunk and smallest chunk of the halving tail: 8 / 2 -> 8 + 8 + 8 + 4 + 2 + 2.  On a quiet host
    # 16 / 8, 8 / 4 and 8 / 2 all give 4.4-4.5 ms per step; on a box with a slow host (device->host at 25 GB/s, packing at
    # half speed) 8 / 2 gave 5.0 ms where 16 / 8 gave 7.1 ms, so the fine schedule is the default.
    E2E_CHUNK = int(os.environ.get("QUBER_E2E_CHUNK", "8"))
    E2E_TAIL = int(os.environ.get("QUBER_E2E_TAIL", "2"))
    # encode / boundary kernels once per group of chunks holding this many frames (0 = per chunk): the masks still cross
    # the link chunk by chunk behind the packing, the kernels run on full grids after the sampling kernels.  v16 sweep
    # (profiles/r01_e2e_sweep_v16.txt): 0 -> 3.98, 16 -> 3.86, 32 -> 3.92 ms per step
    E2E_KERNEL = int(os.environ.get("QUBER_E2E_KERNEL_FRAMES", "16"))
    def e2e_step(s):
        d = sets[s % NSETS]
        # decoder activations never live on the host (resident inputs); launched first so the sampling kernels run
        # under the host-side packing / host->device copies of refine_batch_host's chunk pipeline.  (Putting the chunk
        # pipeline on a second, high-priority stream next to the sampling kernels was measured: 4.36-4.46 against
        # 4.43-4.47 ms per step — the sampling kernels fill the register files, the two only time-share.)
        hp.msda_layers(d["msda"], N_LAYERS)
        o, keep = hp.refine_batch_host(d["h_masks"], d["h_head"], chunk_frames=E2E_CHUNK, min_tail=E2E_TAIL, kernel_frames=E2E_KERNEL)
        if world > 1:
            sharding.gather_label_maps(keep[0][2]["label_map"], torch.cat([k[4] for k in keep]), out=gather_buf, compact=compact)
        return o
    for i in range(max(1, min(args.warmup, 3))):
        o = e2e_step(i)
    barrier()
    e2e_steps = max(2, min(args.steps, 10))
    e0 = _ev()
    e2e_inst = 0
    for i in range(e2e_steps):
        o = e2e_step(i)
        torch.cuda.current_stream().synchronize()            # the caller reads the host result of every step
        e2e_inst += int(o["inst_counts"].sum())
    e1 = _ev()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([e2e_ms, float(e2e_inst)], device=dev, dtype=torch.float64)
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        e2e_ms, e2e_inst = float(tmax[0]), float(tsum[1])
    packed = bool(hp.pack_masks)                              # decided on the first call (HotPath._calibrate_pack)
    h2d = sets[0]["h_masks"].numel() // (8 if packed else 1) + sum(v.numel() * 4 for v in sets[0]["h_head"].values())
    d2h = sum(v.numel() * v.element_size() for v in o.values())
    # what this box's host link gives a plain pinned copy of the same mask bytes (the floor under e2e; it differs between
    # boxes of the pool by NUMA placement of the pinned pages)
    pcie = {}
    if rank == 0:
        hm = sets[0]["h_masks"]
        dm = torch.empty(hm.shape, dtype=hm.dtype, device=dev)
        back = torch.empty(hm.shape, dtype=hm.dtype).pin_memory()
        for name, fn in (("h2d_gbs", lambda: dm.copy_(hm, non_blocking=True)), ("d2h_gbs", lambda: back.copy_(dm, non_blocking=True))):
            fn(); torch.cuda.synchronize()
            c0 = _ev()
            for _ in range(3):
                fn()
            c1 = _ev()
            torch.cuda.synchronize()
            pcie[name] = 3 * hm.numel() / (c0.elapsed_time(c1) * 1e-3) / 1e9
        pcie["h2d_floor_ms_per_step"] = h2d / (pcie["h2d_gbs"] * 1e9) * 1e3
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
    ab = algorithmic_bytes()
    per_stage = {k: stage_ms[k] / args.steps for k in stage_names}
    # per-launch time of each kernel family = stage time / launches in the stage (msda: 6 identical launches)
    roof_all = {
        "encode(reduce+emit)": {"ms": per_stage["encode"], "bytes": ab["encode"]},
        "boundary_kernel": {"ms": per_stage["boundary"], "bytes": ab["boundary"]},
        "msda_forward_p4_kernel": {"ms": per_stage["msda"] / N_LAYERS, "bytes": ab["msda_layer"]},
    }
    for v in roof_all.values():
        v["achieved_gbs"] = v["bytes"] / (v["ms"] * 1e-3) / 1e9
        v["frac"] = v["achieved_gbs"] / peak_gbs
    ncu_traffic, ncu_src = ncu_dram_bytes("msda_forward")
    dominant = max(("boundary_kernel", "msda_forward_p4_kernel", "encode(reduce+emit)"),
                   key=lambda k: roof_all[k]["ms"] * (N_LAYERS if k.startswith("msda") else 1))
    dv = roof_all[dominant]
    launches = (2 + 1 + N_LAYERS + 7) * args.steps
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(world), "clocks": clocks,
        "e2e": {"value": e2e_inst / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps, "host_link": pcie,
                "mask_transfer": ("one bit per pixel, packed on %d host threads inside the timed region" % hp.pack_threads) if packed
                                 else "bytes",
                "pack_calibration": getattr(hp, "pack_calibration", None),
                "schedule": {"chunk_frames": E2E_CHUNK, "min_tail": E2E_TAIL, "kernel_frames": E2E_KERNEL,
                             "label_map": "uint8" if hp.cfg.compact_label_map else "int32"}},
        "gpu_launches": launches,
        "instances_per_step": n_inst_total / args.steps, "post_status": status,
        "stage_ms": per_stage,
        "roofline": {"kernel": dominant, "bound": "hbm", "achieved": dv["achieved_gbs"], "peak": peak_gbs, "unit": "GB/s",
                     "frac": dv["frac"], "traffic": ncu_traffic if dominant.startswith("msda") else None, "traffic_source": ncu_src,
                     "peak_kind": peak_kind, "algorithmic_bytes_per_launch": dv["bytes"],
                     "avg_launch_ms": dv["ms"],
                     "note": ("msda fp32 is bounded by the L1 data pipe, not HBM: 48 gathered 128-byte lines per (query, head) = "
                              "266 us/launch at 128 B/clk/SM, i.e. frac <= 0.37 for any L1/shared-memory gather (DESIGN.md 3.1); "
                              "ncu dram bytes per launch in profiles/") if dominant.startswith("msda") else None},
        "roofline_all": roof_all,
    }
    if world == 1 and not args.no_network:
        enc0 = ops.encode_initial_masks(sets[0]["masks"], hp.cfg.sigma)
        line["network"] = network_side_measurement(dev, enc0, n_inst_total / args.steps, ms / args.steps)
        line["encoder"] = encoder_side_measurement(dev)
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(budget_s=args.cpu_budget)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------- network (disclosed aside)
def network_side_measurement(dev, enc, inst_per_step, hot_ms):
    """SURVEY §8-d's "Net" stage, reported NEXT TO the headline and never inside it: the cuDNN network between
    encode and decode (two ResNet-50 streams, DeepLabV3+ decoder, hierarchical fusion heads; random-init weights of the
    uoais-sim config) on the same 32-frame batch, fp32 and bf16-autocast, and the instances/s the step would give
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
        img = torch.randint(0, 256, (B, 6, H, W), dtype=torch.uint8, device=dev, generator=g)
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
        out["note"] = ("network forward for the same 32 frames (cuDNN convolutions with folded BN and fused epilogues, own "
                       "GroupNorm+ReLU / upsample+concat / fuse-concat / tcgen05 1x1 kernels, CUDA graph), outside `value` and outside the reference arm "
                       "(BASELINE.json times the CPU path without the network); fp32 = PyTorch defaults (cuDNN convolutions may use TF32)")
    except Exception as e:      # the aside must never cost the headline line
        out["error"] = f"{type(e).__name__}: {e}"[:300]
    return out


# --------------------------------------------------------------------------------------------- encoder (disclosed aside)
def encoder_side_measurement(dev):
    """The caller of the sampling kernel: the 6-layer MSDeformAttn transformer encoder of the optional
    pixel decoder on the same 32 frames (S = 6300, d_model 256, FFN 1024) — tcgen05 GEMMs, quber_msda_prepare,
    quber_msda_forward, fused residual + LayerNorm — next to torch's fp32 layer-by-layer module."""
    import torch
    out = {}
    try:
        from quber_b200.pixel_decoder import MSDeformAttnTransformerEncoderOnly
        torch.manual_seed(0)
        enc = MSDeformAttnTransformerEncoderOnly(256, 8, N_LAYERS, 1024, 0.0, "relu", 3, 4).to(dev).eval()
        shapes = [(15, 20), (30, 40), (60, 80)]
        srcs = [torch.randn(B, 256, h, w, device=dev) for h, w in shapes]
        pos = [torch.randn(B, 256, h, w, device=dev) * 0.1 for h, w in shapes]
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
        out["note"] = ("6 encoder layers for 32 frames; fp32 = torch nn.Linear / LayerNorm modules around quber_msda_forward, "
                       "tf32 / bf16 = tcgen05 GEMMs + fused glue kernels; outside `value`")
    except Exception as e:
        out["error"] = f"{type(e).__name__}: {e}"[:300]
    return out


# --------------------------------------------------------------------------------------------- CPU legs
def _oracle_frame(args):
    """The whole hot path for ONE frame on ONE core through the oracle port."""
    import oracle
    masks, low, msda = args
    oracle.encode_initial_masks(masks)
    oracle.masks_to_boundary(masks, 0.01, return_per_instance=True)
    for _ in range(N_LAYERS):
        oracle.msda_forward(msda["value"], msda["shapes"], msda["starts"], msda["loc"], msda["attn"])
    full = oracle.upsample_bilinear(low, 4)
    full[2:4] *= 4.0
    r = oracle.post_process(full[0], full[1], full[2:4], full[4:8])
    return r["n_instances"]


def _cpu_frames(n):
    from quber_b200 import synth
    frames = []
    for b in range(n):
        m = synth.ellipse_masks(K, H, W, seed=b)
        ho = synth.head_outputs(m, seed=b)
        low = np.concatenate([ho["foreground"], ho["center"], ho["offset"], ho["eee_boundary"]], 0)
        frames.append((m, low, synth.msda_inputs(B=1, mode="module", seed=b)))
    return frames


def _run_cpu(frames, threads):
    from concurrent.futures import ThreadPoolExecutor
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:          # ctypes releases the GIL: real parallelism over frames
        n = sum(ex.map(_oracle_frame, frames))
    return n, time.perf_counter() - t0


def cpu_baseline(budget_s=20.0):
    import oracle
    oracle.lib()
    cores = len(os.sched_getaffinity(0))
    frames = _cpu_frames(min(cores, B))
    _run_cpu(frames[:1], 1)                                  # warm-up (page-in)
    n1, t1 = _run_cpu(frames[:1], 1)
    reps = max(1, int(budget_s / max(t1 * max(1.0, len(frames) / cores), 1e-3)) // 2)
    n, t = 0, 0.0
    for _ in range(min(reps, 3)):
        a, b_ = _run_cpu(frames, cores)
        n += a; t += b_
    return {"value": n / t, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{len(frames)} frames of the same workload (K={K}, 640x480, msda x{N_LAYERS}), one frame per thread, "
                      f"{min(reps, 3)} repetition(s), oracle/quber_oracle.c; single-core: {n1 / t1:.2f} instances/s"}


def run_reference(args):
    """--impl reference: the CPU implementation of the path on the box's host cores.  The reference is
    Python (numpy/cv2/torch) and does not exist on the GPU box, so this is the oracle port
    (oracle/quber_oracle.c — pinned to the reference's outputs by tests/test_oracle_golden.py)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    import oracle
    oracle.lib()
    cores = len(os.sched_getaffinity(0))
    frames = _cpu_frames(min(cores, B))
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
    value = n / t
    sample = (f"each step = {len(frames)} frames of the C2 workload (K={K}, 640x480, msda x{N_LAYERS}), one frame per "
              f"thread on {cores} host threads; oracle port of the reference's numpy/cv2/torch path")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-network", action="store_true", help="skip the disclosed network-stage aside")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
