"""World-size-2 gloo test of the multi-GPU host logic (tile partition, packed gather, scene
broadcast) with CPU tensors standing in for the device frames."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank, world, port, w, h, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cs378hgraphics_raytracer_b200 import parallel as par
    rng = np.random.default_rng(5)
    truth = torch.from_numpy(rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8))
    # what rt_render_device(part=rank) leaves behind: own tiles rendered, the rest zero
    mine = torch.zeros_like(truth)
    tx, ty = par.tile_grid(w, h)
    for t in par.tiles_of_rank(w, h, rank, world):
        x0, y0 = (t % tx) * par.TILE, (t // tx) * par.TILE
        mine[y0:y0 + par.TILE, x0:x0 + par.TILE] = truth[y0:y0 + par.TILE, x0:x0 + par.TILE]
    full = par.gather_frame(mine, rank, world)
    arena = bytes(range(256)) * 37 if rank == 0 else b""
    got = par.broadcast_arena(arena, rank, world, torch.device("cpu"))
    ok = got == bytes(range(256)) * 37
    if rank == 0:
        ok = ok and torch.equal(full, truth)
    else:
        ok = ok and full is None
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


import pytest


@pytest.mark.parametrize("world", [2, 4])  # 4: 35 tiles of the 200x150 frame do not divide evenly, ranks pad
def test_gather_and_broadcast_world2(world):
    ctx = mp.get_context("spawn")
    for (w, h) in ((200, 150), (64, 32)):
        q = ctx.Queue()
        port = 29500 + (os.getpid() % 2000) + world
        procs = [ctx.Process(target=_worker, args=(r, world, port, w, h, q)) for r in range(world)]
        for p in procs:
            p.start()
        res = [q.get(timeout=120) for _ in procs]
        for p in procs:
            p.join(60)
        assert all(ok for _, ok in res), res


def test_pack_unpack_roundtrip_single():
    from cs378hgraphics_raytracer_b200 import parallel as par
    rng = np.random.default_rng(1)
    for (w, h, world) in ((100, 70, 3), (32, 32, 1), (33, 1, 8)):
        truth = torch.from_numpy(rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8))
        packed = [par.pack_tiles(truth, r, world) for r in range(world)]
        assert sum(p.shape[0] for p in packed) == np.prod(par.tile_grid(w, h))
        assert torch.equal(par.unpack_tiles(packed, w, h, world), truth)
