"""Generate tests/golden/*.npz by running the REFERENCE'S OWN PYTHON on seeded inputs.

Runs only in the build container (needs /root/reference, which the GPU box does
not have).  The reference is imported/executed where it lies — nothing is copied
into this repo:

  * maskrefiner/modeling/mask_refiner/post_processing.py  — importlib by file path
  * explicit_error_estimation/util.py:62-99               — the three FunctionDefs are
    AST-extracted (the module import needs segmentation_models_pytorch) and exec'd with cv2/numpy
  * maskrefiner/predictor.py:287-359 (MaskRefinerPredictor.predict) — the method's FunctionDef is
    AST-extracted and run with a stub `self` whose `model` echoes its input (the package import
    needs detectron2)
  * maskrefiner/data/dataset_mappers/target_generator.py (PerturbedInputGenerator) — second caller
    of the same encode math, AST-extracted class
  * OPS/functions/ms_deform_attn_func.py — imported after stubbing the compiled
    `MultiScaleDeformableAttention` module; `ms_deform_attn_core_pytorch` runs on CPU

Usage:  python oracle/gen_golden.py   (writes tests/golden/; commit the result)
"""
import ast
import hashlib
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)

from quber_b200 import synth  # noqa: E402
import oracle  # noqa: E402  (only its deterministic up-sampler, to fix the full-res inputs)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load_by_path(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def extract_defs(path, names, kind=(ast.FunctionDef, ast.ClassDef)):
    src = open(path).read()
    tree = ast.parse(src)
    out = []
    for node in ast.walk(tree):
        if isinstance(node, kind) and node.name in names:
            out.append(ast.get_source_segment(src, node))
    return out


def ref_post_processing():
    return load_by_path("ref_post_processing", f"{REF}/maskrefiner/modeling/mask_refiner/post_processing.py")


def ref_boundary_fns():
    import cv2
    ns = {"np": np, "cv2": cv2}
    for s in extract_defs(f"{REF}/explicit_error_estimation/util.py",
                          {"masks_to_fg_mask", "mask_to_boundary", "masks_to_boundary"}, ast.FunctionDef):
        exec(s, ns)
    return ns


def ref_predict_fn():
    """MaskRefinerPredictor.predict (predictor.py:287-359) as a free function of a stub self."""
    src = open(f"{REF}/maskrefiner/predictor.py").read()
    tree = ast.parse(src)
    for node in ast.walk(tree):
        if isinstance(node, ast.ClassDef) and node.name == "MaskRefinerPredictor":
            for f in node.body:
                if isinstance(f, ast.FunctionDef) and f.name == "predict":
                    import textwrap
                    ns = {"np": np, "torch": torch}
                    exec(textwrap.dedent(ast.get_source_segment(src, f)), ns)
                    return ns["predict"]
    raise RuntimeError("predict not found")


def ref_encode(masks, H, W, sigma=10):
    predict = ref_predict_fn()
    stub = types.SimpleNamespace()
    stub.sigma = sigma
    size = 6 * sigma + 3                      # predictor.py:246-251, evaluated here with numpy as there
    x = np.arange(0, size, 1, float)
    y = x[:, np.newaxis]
    x0 = y0 = 3 * sigma + 1
    stub.g = np.exp(-((x - x0) ** 2 + (y - y0) ** 2) / (2 * sigma ** 2))
    stub.rgb_on, stub.depth_on = True, False
    stub.model = lambda batched: batched
    rgb = np.zeros((H, W, 3), np.uint8)
    out = predict(stub, rgb, None, masks)
    return out[0]["initial_pred_offset"].numpy(), stub.g


def ref_msda():
    sys.modules.setdefault("MultiScaleDeformableAttention", types.ModuleType("MultiScaleDeformableAttention"))
    ops = f"{REF}/ext_modules/UnseenObjectsWithMeanShift/MSMFormer/meanshiftformer/modeling/pixel_decoder/ops"
    return load_by_path("ref_msda_func", f"{ops}/functions/ms_deform_attn_func.py")


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(1)

    # ---------------- A1 encode ----------------
    cases = {
        "small": dict(K=6, H=120, W=160, axes=(6, 20), margin=10, seed=3, empty_tail=1),
        "edge": dict(K=5, H=96, W=128, axes=(20, 50), margin=2, seed=4, empty_tail=0),   # centroids near borders
        "c1": dict(K=20, H=480, W=640, axes=(20, 60), margin=60, seed=0, empty_tail=1),
    }
    enc = {}
    for name, c in cases.items():
        m = synth.ellipse_masks(c["K"], c["H"], c["W"], seed=c["seed"], axes=c["axes"], margin=c["margin"],
                                empty_tail=c["empty_tail"])
        if name == "small":
            m[2][m[2] > 0] = 1       # non-255 "inside" value
        out, g = ref_encode(m, c["H"], c["W"])
        enc[f"{name}_masks_bits"] = np.packbits(m != 0)
        enc[f"{name}_mask_values"] = np.array([int(x.max()) for x in m], np.uint8)
        enc[f"{name}_shape"] = np.array(m.shape)
        enc[f"{name}_sha"] = np.frombuffer(bytes.fromhex(sha(out)), np.uint8)
        if name != "c1":
            enc[f"{name}_out"] = out
        else:
            enc[f"{name}_out_center_q"] = out[0, ::4, ::4].copy()
            enc[f"{name}_out_sum"] = out.astype(np.float64).sum((1, 2))
    enc["gauss_table_f32"] = g.astype(np.float32)
    np.savez_compressed(os.path.join(OUT, "encode.npz"), **enc)
    print("encode.npz", {k: v.shape for k, v in enc.items() if "sha" not in k})

    # ---------------- A13 boundary ----------------
    bfn = ref_boundary_fns()
    bnd = {}
    bcases = {
        "small_r01": dict(K=6, H=120, W=160, axes=(6, 30), margin=4, seed=5, ratio=0.01),
        "small_r02": dict(K=6, H=120, W=160, axes=(6, 30), margin=4, seed=5, ratio=0.02),
        "c1_r01": dict(K=20, H=480, W=640, axes=(20, 60), margin=60, seed=0, ratio=0.01),
        "c1_r02": dict(K=8, H=480, W=640, axes=(20, 60), margin=20, seed=7, ratio=0.02),
    }
    for name, c in bcases.items():
        m = synth.ellipse_masks(c["K"], c["H"], c["W"], seed=c["seed"], axes=c["axes"], margin=c["margin"])
        rng = np.random.default_rng(c["seed"] + 100)
        for k in range(len(m)):                       # pinholes and ragged edges
            holes = rng.random(m[k].shape) < 0.002
            m[k][holes] = 0
        if name.startswith("small"):
            m[1][m[1] > 0] = 1
            m[3][m[3] > 0] = 7
        per = np.stack([bfn["mask_to_boundary"](x, dilation_ratio=c["ratio"]) for x in m])
        uni = bfn["masks_to_boundary"](m, dilation_ratio=c["ratio"])
        fg = bfn["masks_to_fg_mask"](m)
        bnd[f"{name}_masks"] = np.packbits(m != 0)
        bnd[f"{name}_mask_values"] = np.array([int(x.max()) for x in m], np.uint8)
        bnd[f"{name}_shape"] = np.array(m.shape)
        bnd[f"{name}_ratio"] = np.array(c["ratio"])
        bnd[f"{name}_per_bits"] = np.packbits(per != 0)
        bnd[f"{name}_per_sha"] = np.frombuffer(bytes.fromhex(sha(per)), np.uint8)
        bnd[f"{name}_union"] = np.packbits(uni)
        bnd[f"{name}_fg"] = np.packbits(fg)
    np.savez_compressed(os.path.join(OUT, "boundary.npz"), **bnd)
    print("boundary.npz ok")

    # ---------------- A8-A12 post-processing ----------------
    pp = ref_post_processing()
    post = {}
    pcases = {
        "small": dict(K=5, H=240, W=320, axes=(18, 40), margin=40, seed=11, top_k=200),
        "topk": dict(K=12, H=240, W=320, axes=(14, 30), margin=30, seed=12, top_k=8),
        "plateau": dict(K=12, H=240, W=320, axes=(14, 30), margin=30, seed=12, top_k=8),
        "nocentre": dict(K=3, H=240, W=320, axes=(30, 50), margin=60, seed=13, top_k=200),
        "c1": dict(K=20, H=480, W=640, axes=(20, 60), margin=60, seed=0, top_k=200),
    }
    for name, c in pcases.items():
        m = synth.ellipse_masks(c["K"], c["H"], c["W"], seed=c["seed"], axes=c["axes"], margin=c["margin"])
        ho = synth.head_outputs(m, stride=4, seed=c["seed"])
        if name == "nocentre":
            ho["center"] = ho["center"] * 0.2          # below CENTER_THRESHOLD everywhere -> stuff branch
        if name == "plateau":
            # equal-valued plateaus: every plateau pixel survives the NMS, the k-th score ties the
            # maximum and the strict `>` at post_processing.py:41 leaves nothing
            ho["center"] = np.round(ho["center"] * 8) / 8
        low = np.concatenate([ho["foreground"], ho["center"], ho["offset"], ho["eee_boundary"]], 0)
        full = oracle.upsample_bilinear(low, 4)
        full[2:4] *= 4.0                                # model.py:698
        r, cmap, o, b = full[0:1], full[1:2], full[2:4], full[4:8]
        rt, ct, ot = torch.from_numpy(r), torch.from_numpy(cmap.copy()), torch.from_numpy(o)
        fgseg = rt.sigmoid().round()                    # model.py:292
        pan, ctr = pp.get_panoptic_segmentation(
            fgseg, ct.clone(), ot, thing_ids=dict({1: 0}).values(), label_divisor=1000, stuff_area=2048,
            void_label=-1, threshold=0.3, nms_kernel=7, top_k=c["top_k"])
        centres = pp.find_instance_center(ct.clone(), threshold=0.3, nms_kernel=7, top_k=c["top_k"])
        grouped = pp.group_pixels(centres, ot) if len(centres) else torch.zeros(1, c["H"], c["W"], dtype=torch.long)
        pan = pan.squeeze(0)
        # instance extraction exactly as model.py:318-356 words it (torch ops; Instances/BitMasks are
        # detectron2 containers, un-vendored — only their arithmetic is reproduced)
        sem_prob = rt.sigmoid().squeeze(0)
        labels, scores, boxes, cyx = [], [], [], []
        for lab in np.unique(pan.numpy()):
            if lab == -1:
                continue
            mask = pan == lab
            s = torch.mean(sem_prob.clone()[mask])
            idx = torch.nonzero(mask).float()
            cy, cx = torch.mean(idx[:, 0]), torch.mean(idx[:, 1])
            cs = ct[0, int(cy.item()), int(cx.item())]
            labels.append(int(lab)); scores.append(float(s * cs)); cyx.append([float(cy), float(cx)])
            ys, xs = torch.nonzero(mask.any(1)).flatten(), torch.nonzero(mask.any(0)).flatten()
            boxes.append([float(xs.min()), float(ys.min()), float(xs.max() + 1), float(ys.max() + 1)])
        bt = torch.from_numpy(b)
        am = bt.argmax(0, keepdim=True)                 # eval/eval_utils.py:311-313
        onehot = torch.cat([am == i for i in range(4)], 0).numpy()
        post[f"{name}_low"] = low
        post[f"{name}_topk"] = np.array(c["top_k"])
        post[f"{name}_fgseg"] = np.packbits(fgseg.numpy().astype(bool))
        post[f"{name}_panoptic"] = pan.numpy().astype(np.int16)
        post[f"{name}_centres"] = centres.numpy().astype(np.int32)
        post[f"{name}_grouped"] = grouped.numpy().astype(np.uint8)[0]
        post[f"{name}_labels"] = np.array(labels, np.int32)
        post[f"{name}_scores"] = np.array(scores, np.float32)
        post[f"{name}_boxes"] = np.array(boxes, np.float32).reshape(-1, 4)
        post[f"{name}_centroid_f32"] = np.array(cyx, np.float32).reshape(-1, 2)
        post[f"{name}_onehot"] = np.packbits(onehot)
        # torch's own x4 interpolate of the same low-res maps, to bound the up-sampler's tolerance
        tfull = torch.nn.functional.interpolate(torch.from_numpy(low)[None], scale_factor=4, mode="bilinear",
                                                align_corners=False)[0].numpy()
        post[f"{name}_torch_up_maxabs"] = np.array(np.abs(tfull - oracle.upsample_bilinear(low, 4)).max())
        if name == "small":
            post[f"{name}_torch_up_ch1_q"] = tfull[1, ::2, ::2].copy()
        print(name, "centres", len(centres), "instances", len(labels), "up diff", post[f"{name}_torch_up_maxabs"])
    np.savez_compressed(os.path.join(OUT, "post.npz"), **post)

    # ---------------- A14 MSDA ----------------
    msda = ref_msda()
    md = {}
    mcases = {
        "opstest": dict(B=1, shapes=((6, 4), (3, 2)), M=2, D=2, P=2, Q=2, seed=3, mode="rand"),      # OPS/test.py:24-28
        "medium": dict(B=2, shapes=((8, 10), (4, 5), (2, 3)), M=8, D=32, P=4, Q=None, seed=1, mode="rand"),
        "border": dict(B=1, shapes=((12, 16), (6, 8), (3, 4)), M=8, D=32, P=4, Q=100, seed=2, mode="border"),
        "d8": dict(B=1, shapes=((10, 14), (5, 7)), M=8, D=8, P=4, Q=64, seed=5, mode="border"),      # CONVS_DIM 64 variant
    }
    for name, c in mcases.items():
        inp = synth.msda_inputs(B=c["B"], shapes=c["shapes"], M=c["M"], D=c["D"], P=c["P"], Q=c["Q"], seed=c["seed"],
                                mode=c["mode"], dtype=np.float64)
        t = {k: torch.from_numpy(v) for k, v in inp.items()}
        o64 = msda.ms_deform_attn_core_pytorch(t["value"], t["shapes"], t["loc"], t["attn"]).numpy()
        o32 = msda.ms_deform_attn_core_pytorch(t["value"].float(), t["shapes"], t["loc"].float(), t["attn"].float()).numpy()
        # inputs are NOT stored: synth.msda_inputs(seed) uses only Generator.random(), which numpy
        # guarantees bit-stable across platforms; the tests regenerate them from `cfg`.
        md[f"{name}_in_sha"] = np.frombuffer(bytes.fromhex(sha(np.concatenate([inp["value"].ravel(), inp["loc"].ravel(), inp["attn"].ravel()]))), np.uint8)
        md[f"{name}_out_f64"] = o64     # computed from the float64 inputs
        md[f"{name}_out_f32"] = o32     # computed from their float32 roundings
    np.savez_compressed(os.path.join(OUT, "msda.npz"), **md)
    print("msda.npz ok")
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
