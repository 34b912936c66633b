"""The N > 1 path on CPU: world_size-2 gloo processes shard frames and all-gather label maps."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from quber_b200 import sharding


def test_frame_blocks_partition_every_frame_once():
    for n, block, world in [(4096, 32, 8), (100, 32, 3), (5, 8, 2), (64, 32, 2)]:
        seen = []
        for r in range(world):
            for a, b in sharding.frame_blocks(n, block, r, world):
                assert 0 <= a < b <= n and b - a <= block
                seen.extend(range(a, b))
        assert sorted(seen) == list(range(n))
    assert sharding.frame_blocks(4096, 32, 3, 8)[:2] == [(96, 128), (352, 384)]
    with pytest.raises(ValueError):
        sharding.frame_blocks(10, 4, 2, 2)


def _label_maps(rank, B=3, H=12, W=16):
    g = torch.Generator().manual_seed(100 + rank)
    counts = torch.randint(1, 5, (B,), generator=g, dtype=torch.int32)
    maps = torch.stack([torch.randint(0, int(c) + 1, (H, W), generator=g, dtype=torch.int32) for c in counts])
    return maps, counts


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        maps, counts = _label_maps(rank)
        allm, allc = sharding.gather_label_maps(maps, counts)
        ok = allm.shape == (world, 3, 12, 16) and allc.shape == (world, 3)
        for r in range(world):
            em, ec = _label_maps(r)
            ok = ok and torch.equal(allm[r], em) and torch.equal(allc[r], ec)
        # a consumer on any rank re-materialises the per-instance masks of another rank's frame
        other = (rank + 1) % world
        m = sharding.masks_from_label_map(allm[other, 0], allc[other, 0])
        ok = ok and m.shape[0] == int(allc[other, 0]) and bool((m.sum(0) == (allm[other, 0] > 0)).all())
        # compact transport: uint8 maps (<= 255 instances per frame), same content
        cm, cc = sharding.gather_label_maps(maps, counts, compact=True)
        ok = ok and cm.dtype == torch.uint8 and torch.equal(cm.to(torch.int32), allm) and torch.equal(cc, allc)
        m8 = sharding.masks_from_label_map(cm[other, 0], cc[other, 0])
        ok = ok and torch.equal(m8, m)
        # the single-collective payload (maps + counts in ONE buffer, one all-gather per step)
        pl = sharding.Payload(3, 12, 16, "cpu", world)
        pl.maps.copy_(maps.to(torch.uint8))
        pl.counts.copy_(counts)
        pm, pc = pl.gather()
        ok = ok and pm.shape == (world, 3, 12, 16) and torch.equal(pm.to(torch.int32), allm) and torch.equal(pc, allc)
        # whole-job instance count = sum over ranks, as bench.py aggregates it
        tot = allc.sum().clone()
        mine = counts.sum().clone()
        dist.all_reduce(mine)
        ok = ok and int(tot) == int(mine)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_gather():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_payload_single_process_layout():
    pl = sharding.Payload(2, 5, 7, "cpu")                  # 70 map bytes -> counts start at the next 16-byte boundary
    assert pl.map_bytes == 80 and pl.nbytes == 88 and pl.counts.dtype == torch.int32 and pl.counts.numel() == 2
    pl.maps.fill_(3); pl.counts.copy_(torch.tensor([4, 9], dtype=torch.int32))
    m, c = pl.gather()
    assert m.shape == (1, 2, 5, 7) and int(m.sum()) == 210 and c.tolist() == [[4, 9]]


def test_single_process_passthrough():
    maps, counts = _label_maps(0)
    a, c = sharding.gather_label_maps(maps, counts)
    assert a.shape[0] == 1 and torch.equal(a[0], maps) and torch.equal(c[0], counts)
    a8, _ = sharding.gather_label_maps(maps, counts, compact=True)
    assert a8.dtype == torch.uint8 and torch.equal(a8[0].to(torch.int32), maps)
