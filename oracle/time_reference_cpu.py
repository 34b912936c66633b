"""Time the REFERENCE'S OWN PYTHON (numpy / cv2 / torch CPU) for the hot path on one C2-shaped frame, in the build
container (the GPU box has no /root/reference), next to the C oracle port on the same frame and core count.
Writes profiles/r01_reference_python_cpu.json.  Test infrastructure, not product.

    python oracle/time_reference_cpu.py [threads]
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import gen_golden as gg  # noqa: E402  (import machinery for the reference modules)
import oracle  # noqa: E402
from quber_b200 import synth  # noqa: E402


def med(f, n=3):
    f()
    ts = []
    for _ in range(n):
        t = time.perf_counter(); f(); ts.append(time.perf_counter() - t)
    return float(np.median(ts))


_REF = {}


def reference_frame(c, g, seed, threads, n_layers=6):
    """ONE frame of a bench.py config through the reference's OWN Python, timed: the predictor's encode block
    (predictor.py:304-357), the cv2 erosions (util.py:72-99), `n_layers` x ms_deform_attn_core_pytorch, F.interpolate x4 +
    sem_seg_postprocess' crop + sigmoid().round() + get_panoptic_segmentation + the instance loop's masks + the argmax
    decode (model.py:689-708, :266-336; eval_utils.py:310-313).  -> (instances, seconds).  bench.py --impl reference."""
    import cv2
    torch.set_num_threads(threads)
    cv2.setNumThreads(threads)
    if not _REF:
        _REF.update(bfn=gg.ref_boundary_fns(), pp=gg.ref_post_processing(), mf=gg.ref_msda())
    K, H, W = c["K"], c["H"], c["W"]
    m = synth.ellipse_masks(K, g["Hf"], g["Wf"], seed=seed, axes=c["axes"], margin=c["margin"])
    m[:, H:, :] = 0
    m[:, :, W:] = 0
    ho = synth.head_outputs(m, seed=seed)
    low = torch.from_numpy(np.concatenate([ho["foreground"], ho["center"], ho["offset"], ho["eee_boundary"]], 0))[None]
    masks = np.ascontiguousarray(m[:, :H, :W])
    msda = synth.msda_inputs(B=1, shapes=c["levels"], mode="module", seed=seed)
    v, sh = torch.from_numpy(msda["value"]), torch.from_numpy(msda["shapes"])
    loc, at = torch.from_numpy(msda["loc"]), torch.from_numpy(msda["attn"])
    bfn, pp, mf = _REF["bfn"], _REF["pp"], _REF["mf"]
    t0 = time.perf_counter()
    gg.ref_encode(masks, H, W)
    per = [bfn["mask_to_boundary"](x, 0.01) for x in masks]
    bfn["masks_to_boundary"](masks, 0.01)
    for _ in range(n_layers):
        mf.ms_deform_attn_core_pytorch(v, sh, loc, at)
    up = torch.nn.functional.interpolate(low, scale_factor=4, mode="bilinear", align_corners=False)[0][:, :H, :W]
    fg = up[0:1].sigmoid().round()
    pan, _ = pp.get_panoptic_segmentation(fg, up[1:2].clone(), up[2:4] * 4, thing_ids={1}, label_divisor=1000, stuff_area=2048,
                                          void_label=-1, threshold=0.3, nms_kernel=7, top_k=200)
    pan = pan[0].numpy()
    masks_out = [pan == l for l in np.unique(pan) if l != -1]
    am = up[4:8].argmax(0)
    torch.stack([am == i for i in range(4)])
    return len(masks_out), time.perf_counter() - t0


def main():
    threads = int(sys.argv[1]) if len(sys.argv) > 1 else len(os.sched_getaffinity(0))
    torch.set_num_threads(threads)
    import cv2
    cv2.setNumThreads(threads)
    K, H, W = 30, 480, 640
    masks = synth.ellipse_masks(K, H, W, seed=0)
    ho = synth.head_outputs(masks, seed=0)
    low = np.concatenate([ho["foreground"], ho["center"], ho["offset"], ho["eee_boundary"]], 0)
    msda = synth.msda_inputs(B=1, mode="module", seed=0)
    bfn = gg.ref_boundary_fns()
    pp = gg.ref_post_processing()
    mf = gg.ref_msda()
    tl = torch.from_numpy(low)[None]

    def ref_post():
        up = torch.nn.functional.interpolate(tl, scale_factor=4, mode="bilinear", align_corners=False)[0]   # model.py:688-709
        fg = up[0:1].sigmoid().round()                                                                       # model.py:292
        pan, _ = pp.get_panoptic_segmentation(fg, up[1:2], up[2:4] * 4, thing_ids={1}, label_divisor=1000, stuff_area=2048,
                                              void_label=-1, threshold=0.3, nms_kernel=7, top_k=200)
        pan = pan[0].numpy()
        masks_out = [pan == l for l in np.unique(pan) if l != -1]                                            # model.py:313-336
        am = up[4:8].argmax(0)
        onehot = torch.stack([am == i for i in range(4)])                                                    # eval_utils.py:311-313
        return len(masks_out), onehot

    v = torch.from_numpy(msda["value"]); sh = torch.from_numpy(msda["shapes"])
    loc = torch.from_numpy(msda["loc"]); at = torch.from_numpy(msda["attn"])
    t_enc = med(lambda: gg.ref_encode(masks, H, W))
    t_bnd = med(lambda: [bfn["mask_to_boundary"](m, 0.01) for m in masks] + [bfn["masks_to_boundary"](masks, 0.01)])
    t_msda = med(lambda: mf.ms_deform_attn_core_pytorch(v, sh, loc, at))
    t_post = med(ref_post)
    n_inst = ref_post()[0]

    def port():
        oracle.encode_initial_masks(masks)
        oracle.masks_to_boundary(masks, 0.01, return_per_instance=True)
        for _ in range(6):
            oracle.msda_forward(msda["value"], msda["shapes"], msda["starts"], msda["loc"], msda["attn"])
        full = oracle.upsample_bilinear(low, 4); full[2:4] *= 4.0
        return oracle.post_process(full[0], full[1], full[2:4], full[4:8])["n_instances"]
    t_port = med(port)
    total = t_enc + t_bnd + 6 * t_msda + t_post
    out = {"where": "build container (no GPU)", "threads": threads, "frame": "C2 shape: 640x480, K=30, MSDA S=6300 M=8 D=32 L=3 P=4",
           "reference_python_ms": {"encode (predictor.py:304-357)": t_enc * 1e3,
                                   "boundary per-instance + union (util.py:72-99, cv2)": t_bnd * 1e3,
                                   "msda one layer (ms_deform_attn_core_pytorch)": t_msda * 1e3,
                                   "upsample + panoptic post-processing + instance masks + error one-hot": t_post * 1e3,
                                   "frame total (6 MSDA layers)": total * 1e3},
           "reference_python_instances_per_s": n_inst / total,
           "oracle_port_single_thread_ms_per_frame": t_port * 1e3,
           "oracle_port_single_thread_instances_per_s": port() / t_port,
           "instances_per_frame": n_inst}
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", "r02_reference_python_cpu.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
