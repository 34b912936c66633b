"""Multi-GPU plumbing of the path: frames are independent, so the batch is sharded across ranks with
no data-path collective; the only exchange is ONE all-gather per step of the refined label maps.

The reference runs inference as one hand-started process per GPU (CUDA_VISIBLE_DEVICES, eval/run_eval.py:72,
bash/eval_*.sh) at batch size 1 and never collects anything; here one process per GPU is launched by torchrun
and rank r refines the frame blocks {r, r + G, r + 2G, ...}.  The payload is the fixed-shape int32 [B,H,W] label map
(0 = void, j + 1 = j-th refined instance of the frame) — lossless, because the reference's instances are disjoint
by construction (model.py:331) — plus the per-frame instance counts; per-instance bool masks [K,H,W] have a
data-dependent K and are re-materialised by the consumer with `masks_from_label_map`.
"""
import torch
import torch.distributed as dist


def frame_blocks(n_frames, block, rank, world):
    """[(start, stop), ...] of the contiguous frame blocks of size `block` that rank `rank` refines: block i goes to
    rank i % world.  Every frame belongs to exactly one rank."""
    if block <= 0 or world <= 0 or not 0 <= rank < world:
        raise ValueError("bad sharding arguments")
    n_blocks = (n_frames + block - 1) // block
    return [(i * block, min((i + 1) * block, n_frames)) for i in range(rank, n_blocks, world)]


def gather_label_maps(label_map, counts, group=None, out=None, compact=False):
    """label_map int32 [B,H,W], counts int32 [B] on this rank -> (maps [G,B,H,W], counts [G,B]) on every rank.
    One collective for the maps (NCCL all_gather_into_tensor over NVLink; gloo falls back to all_gather).
    compact: the maps travel (and are returned) as uint8 — a quarter of the bytes; the caller guarantees at most 255
    instances per frame (RefinerConfig.max_instances_per_image), checked here against `counts` only in debug runs."""
    if label_map.dtype not in (torch.int32, torch.uint8) or counts.dtype != torch.int32:
        raise ValueError("label_map must be int32 (or uint8, already compact) and counts int32")
    if compact and label_map.dtype != torch.uint8:
        label_map = label_map.to(torch.uint8)
    if not dist.is_available() or not dist.is_initialized():
        return label_map[None], counts[None]
    world = dist.get_world_size(group)
    if out is not None and (out.dtype != label_map.dtype or tuple(out.shape) != (world,) + tuple(label_map.shape)):
        raise ValueError("out must be [world, B, H, W] of the dtype the maps travel in")
    maps = out if out is not None else torch.empty((world,) + tuple(label_map.shape), dtype=label_map.dtype, device=label_map.device)
    cnts = torch.empty((world,) + tuple(counts.shape), dtype=counts.dtype, device=counts.device)
    if dist.get_backend(group) == "nccl":
        dist.all_gather_into_tensor(maps.view(-1), label_map.contiguous().view(-1), group=group)
        dist.all_gather_into_tensor(cnts.view(-1), counts.contiguous().view(-1), group=group)
    else:
        dist.all_gather(list(maps.unbind(0)), label_map.contiguous(), group=group)
        dist.all_gather(list(cnts.unbind(0)), counts.contiguous(), group=group)
    return maps, cnts


class Payload:
    """The ONE buffer a rank contributes to the step's collective: the uint8 label maps [B,H,W] followed (16-byte
    aligned) by the int32 instance counts [B].  quber_post_process writes the maps straight into `maps`
    (ops.post_process(label_map_out=...)), the counts are a B-element copy; `gather()` is then a single
    all-gather — north_star: "results are collected with a single NCCL all-gather".

    Transport of that all-gather (`self.transport`):
      "peer_copy"  one process per GPU on an NVLink / NVSwitch box: the gathered buffer of every rank is symmetric memory
                   (torch.distributed._symmetric_memory), each rank PUSHES its payload into row `rank` of every peer's
                   buffer with plain device-to-device copies — copy engines over NVLink, no SM and no NCCL kernel under
                   the memory-bound kernels of the next step — between two device-side barriers (nobody still reads the
                   previous contents / everybody's rows have landed).
      "nccl"       all_gather_into_tensor, when symmetric memory cannot be set up on every rank.
      "gloo"       all_gather (CPU tests).
    """

    def __init__(self, B, H, W, device, world=1, group=None, peer_copy=True):
        self.B, self.H, self.W, self.world = B, H, W, world
        self.map_bytes = (B * H * W + 15) // 16 * 16
        self.nbytes = self.map_bytes + 4 * B
        self.buf = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)
        self.maps = self.buf[: B * H * W].view(B, H, W)
        self.counts = self.buf[self.map_bytes:].view(torch.int32)
        self.transport = "local"
        self._peer_rows = None
        if world > 1:
            self.transport = "nccl" if dist.get_backend(group) == "nccl" else "gloo"
            if self.transport == "nccl" and peer_copy:
                self._setup_peer_copy(device, group)
            if self._peer_rows is None:
                self.all = torch.empty((world, self.nbytes), dtype=torch.uint8, device=device)
        else:
            self.all = self.buf[None]

    def _setup_peer_copy(self, device, group):
        """Symmetric `all` buffer + this rank's row in every peer's copy of it.  Every rank must end up on the same
        transport: the outcome is agreed with one all-reduce."""
        ok, rows, all_, hdl = 1, None, None, None
        try:
            import torch.distributed._symmetric_memory as symm
            g = group if group is not None else dist.group.WORLD
            rank = dist.get_rank(g)
            all_ = symm.empty((self.world, self.nbytes), dtype=torch.uint8, device=device)
            hdl = symm.rendezvous(all_, g)
            rows = [hdl.get_buffer(r, (self.world, self.nbytes), torch.uint8)[rank] for r in range(self.world)]
            order = [(rank + 1 + i) % self.world for i in range(self.world)]       # start at the neighbour: no hot peer
            rows = [rows[r] for r in order]
        except Exception as e:                                   # noqa: BLE001 — any failure means "use NCCL"
            ok, self.peer_copy_error = 0, f"{type(e).__name__}: {e}"
        flag = torch.tensor([ok], dtype=torch.int32, device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag[0]) == 1:
            self.all, self._peer_rows, self._hdl, self.transport = all_, rows, hdl, "peer_copy"

    def gather(self, group=None):
        """-> (maps [G,B,H,W] uint8, counts [G,B] int32) on every rank; one collective, on the current stream."""
        if self._peer_rows is not None:
            self._hdl.barrier(channel=0)                          # every rank is done with the previous contents of its `all`
            for row in self._peer_rows:
                row.copy_(self.buf, non_blocking=True)
            self._hdl.barrier(channel=1)                          # every rank's pushes (stream-ordered before its arrival) have landed
        elif self.world > 1:
            if self.transport == "nccl":
                dist.all_gather_into_tensor(self.all.view(-1), self.buf, group=group)
            else:
                dist.all_gather(list(self.all.unbind(0)), self.buf, group=group)
        return self.unpack(self.all)

    def unpack(self, gathered):
        G = gathered.shape[0]
        maps = gathered[:, : self.B * self.H * self.W].view(G, self.B, self.H, self.W)
        counts = gathered[:, self.map_bytes:].contiguous().view(torch.int32).view(G, self.B)
        return maps, counts


def masks_from_label_map(label_map, n_instances):
    """[H,W] int32 label map -> bool [n,H,W], the reference's `pred_masks` (model.py:331-336)."""
    ids = torch.arange(1, int(n_instances) + 1, device=label_map.device, dtype=torch.int32).to(label_map.dtype)
    return label_map[None] == ids[:, None, None]
