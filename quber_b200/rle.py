"""`pycocotools.mask` for the reference's call sites, with the run-length work on the GPU.

The reference keeps every dataset mask as a COCO RLE dict `{"size": [h, w], "counts": <compressed ascii>}` and calls
`mask_util.decode` / `mask_util.encode` / `mask_util.frPyObjects` on it
(maskrefiner/data/dataset_mappers/perturbed_panoptic_dataset_mapper.py:195-214,
maskrefiner/data/datasets/register_uoais_sim_instance.py:207-209, tools/ours/panoptic2eee.py:61-64,95-97,125-132).
`import quber_b200.rle as mask_util` keeps those lines working; the same module gives the batched device forms the hot
path uses (`decode_to_device`, `encode_from_device`, `pack_rles`): a frame's 30 initial masks are ~20 KB of run lengths
instead of 9.2 MB of bytes on the host link.

The string <-> counts codec runs on the host (a few hundred bytes per mask, csrc/rle_host.cpp); counts -> pixels and
pixels -> counts are CUDA kernels (csrc/rle.cu).  There is no CPU pixel path here.
"""
import numpy as np
import torch

from . import ops


def _counts_of(rle):
    c = rle["counts"]
    if isinstance(c, (bytes, str)):
        return ops.rle_from_string(c)
    return np.ascontiguousarray(c, dtype=np.uint32)          # uncompressed RLE: a list of ints


def pack_rles(rles):
    """list of RLE dicts (compressed or uncompressed, all of one size) -> (counts int32 numpy holding the uint32 run
    lengths back to back, offsets int64 numpy [n+1], (h, w)): the host-side form quber_rle_decode takes."""
    if isinstance(rles, dict):
        rles = [rles]
    if not rles:
        raise ValueError("no RLE given")
    h, w = (int(v) for v in rles[0]["size"])
    cs = []
    for r in rles:
        if tuple(int(v) for v in r["size"]) != (h, w):
            raise ValueError("all RLEs of one call must have the same size")
        cs.append(_counts_of(r))
    offsets = np.zeros(len(cs) + 1, np.int64)
    np.cumsum([len(c) for c in cs], out=offsets[1:])
    counts = np.concatenate(cs) if cs else np.zeros(0, np.uint32)
    return counts.view(np.int32), offsets, (h, w)


def decode_to_device(rles, device="cuda", packed=False, value=1, check=True):
    """mask_util.decode for a list of RLEs, result left on the GPU: uint8 CUDA [n,h,w] (0 / value), row-major — or with
    packed=True the bit-packed words ops.encode_initial_masks_packed takes."""
    counts, offsets, (h, w) = pack_rles(rles)
    c = torch.from_numpy(counts).to(device, non_blocking=True)
    o = torch.from_numpy(offsets).to(device, non_blocking=True)
    out, _ = ops.rle_decode(c, o, len(offsets) - 1, h, w, value=value, packed=packed, check=check)
    return out


def decode(rleObjs):
    """pycocotools.mask.decode: one RLE dict -> uint8 [h,w]; a list of n -> uint8 [h,w,n] (Fortran order, as
    pycocotools returns it)."""
    single = isinstance(rleObjs, dict)
    m = decode_to_device(rleObjs).permute(1, 2, 0).cpu().numpy()          # [h,w,n]
    m = np.asfortranarray(m)
    return m[:, :, 0] if single else m


def encode_from_device(masks):
    """uint8 / bool CUDA [n,h,w] -> list of n compressed RLE dicts (counts as bytes, like pycocotools)."""
    if masks.dtype == torch.bool:
        masks = masks.to(torch.uint8)
    masks = masks.contiguous()
    n, h, w = masks.shape
    counts, offsets, status = ops.rle_encode(masks)
    off = offsets.cpu().numpy()
    if int(status[0]) & 2:                                               # capacity: retry with what is needed
        counts, offsets, status = ops.rle_encode(masks, cap=int(off[-1]))
        off = offsets.cpu().numpy()
    c = counts[: int(off[-1])].cpu().numpy().view(np.uint32)
    return [{"size": [h, w], "counts": ops.rle_to_string(c[off[i]:off[i + 1]])} for i in range(n)]


def encode(bimask):
    """pycocotools.mask.encode: uint8 [h,w] -> one RLE dict; [h,w,n] -> list of n.  (pycocotools insists on a
    Fortran-ordered array; any layout is taken here.)  CUDA tensors [n,h,w] go to encode_from_device."""
    if isinstance(bimask, torch.Tensor):
        return encode_from_device(bimask if bimask.dim() == 3 else bimask[None])[slice(None) if bimask.dim() == 3 else 0]
    a = np.asarray(bimask)
    if a.ndim == 2:
        return encode_from_device(torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint8))[None].cuda())[0]
    if a.ndim != 3:
        raise ValueError("encode expects a [h,w] or [h,w,n] array")
    return encode_from_device(torch.from_numpy(np.ascontiguousarray(a.transpose(2, 0, 1), dtype=np.uint8)).cuda())


def frPyObjects(pyobj, h, w):
    """pycocotools.mask.frPyObjects for the branch the reference uses (register_uoais_sim_instance.py:207-209): an
    UNCOMPRESSED RLE dict (counts = list of ints), or a list of them -> compressed RLE(s).  Polygons and boxes are
    rasterised by pycocotools on the CPU and are outside this path."""
    def one(o):
        if not (isinstance(o, dict) and "counts" in o and "size" in o):
            raise NotImplementedError("frPyObjects: only uncompressed RLE dicts are supported (polygons / boxes are out of scope)")
        if isinstance(o["counts"], (bytes, str)):
            return {"size": [int(h), int(w)], "counts": o["counts"] if isinstance(o["counts"], bytes) else o["counts"].encode("ascii")}
        return {"size": [int(h), int(w)], "counts": ops.rle_to_string(np.asarray(o["counts"], np.uint32))}
    return one(pyobj) if isinstance(pyobj, dict) else [one(o) for o in pyobj]


def area(rleObjs):
    """pycocotools.mask.area: number of set pixels = the sum of the odd-indexed run lengths."""
    single = isinstance(rleObjs, dict)
    out = np.array([int(_counts_of(r)[1::2].astype(np.uint64).sum()) for r in ([rleObjs] if single else rleObjs)], np.uint32)
    return out[0] if single else out
