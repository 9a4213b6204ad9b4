"""Host-side multi-GPU logic on CPU: partition functions, and the gather fallback over gloo with world_size 2 and 3.
(The NVLink peer-store path itself needs GPUs: tests/test_multigpu_gpu.py, bench.py --workload frame4k_tiles.)"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cudaneuralrender_b200 import multigpu as mg


@pytest.mark.parametrize("W,H", [(3840, 2160), (1920, 1080), (512, 512), (100, 70), (64, 64), (1, 1)])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_tiles_partition_the_frame(W, H, world):
    seen = np.zeros(W * H, dtype=np.int32)
    counts = []
    for r in range(world):
        idx = mg.tile_pixel_indices(W, H, mg.tiles_of_rank(W, H, r, world))
        seen[idx] += 1
        counts.append(idx.size)
    assert np.all(seen == 1)                      # every pixel exactly once
    tx, ty = mg.tile_grid(W, H)
    assert sum(len(mg.tiles_of_rank(W, H, r, world)) for r in range(world)) == tx * ty


def test_tile_rect_matches_kernel_enumeration():
    # csrc/render_kernels.cuh decode_pixel: tile t = rank + j*world, tx = t % tiles_x, ty = t / tiles_x
    W, H = 200, 130
    tx, ty = mg.tile_grid(W, H)
    assert (tx, ty) == (4, 3)
    assert mg.tile_rect(W, H, 0) == (0, 0, 64, 64)
    assert mg.tile_rect(W, H, 3) == (192, 0, 200, 64)       # clipped in x
    assert mg.tile_rect(W, H, 11) == (192, 128, 200, 130)   # clipped in both
    assert mg.tiles_of_rank(W, H, 1, 4) == [1, 5, 9]


def test_views_and_slabs():
    assert mg.views_of_rank(64, 3, 8) == list(range(3, 64, 8))
    assert sum(len(mg.views_of_rank(5, r, 8)) for r in range(8)) == 5
    for n in (0, 1, 7, 16777216, 1000003):
        for world in (1, 2, 3, 8):
            slabs = [mg.slab_of_rank(n, r, world) for r in range(world)]
            assert slabs[0][0] == 0 and sum(c for _, c in slabs) == n
            for (f0, c0), (f1, _) in zip(slabs, slabs[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in slabs) - min(c for _, c in slabs) <= 1
    with pytest.raises(ValueError):
        mg.tiles_of_rank(64, 64, 2, 2)


def test_round_robin_balances_better_than_bands():
    # a centred blob of "work": round-robin tiles must beat contiguous bands
    W, H, G = 1920, 1080, 8
    ys, xs = np.mgrid[0:H, 0:W]
    work = np.exp(-(((xs - W / 2) / 300.0) ** 2 + ((ys - H / 2) / 300.0) ** 2)).ravel()
    rr = [work[mg.tile_pixel_indices(W, H, mg.tiles_of_rank(W, H, r, G))].sum() for r in range(G)]
    bands = [work[r * (W * H // G):(r + 1) * (W * H // G)].sum() for r in range(G)]
    assert mg.tile_balance(rr) > 0.9 > mg.tile_balance(bands)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, W, H, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = (torch.arange(W * H, dtype=torch.int64) * 2654435761 % 4294967291).to(torch.int32)
        frame = torch.full((W * H,), -1, dtype=torch.int32)
        mine = torch.from_numpy(mg.tile_pixel_indices(W, H, mg.tiles_of_rank(W, H, rank, world)))
        frame[mine] = full[mine]                       # this rank "rendered" only its tiles
        mg.gather_tiles(frame, W, H, rank, world, root=0)
        ok = bool(torch.equal(frame, full)) if rank == 0 else True
        # timing plumbing used by bench.py: max over ranks
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        q.put((rank, ok, float(t.item())))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,W,H", [(2, 200, 130), (3, 257, 129)])
def test_gather_tiles_gloo(world, W, H):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, W, H, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    assert all(t == float(world) for _, _, t in res)
