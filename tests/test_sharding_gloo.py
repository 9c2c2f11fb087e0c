"""world_size-2 gloo test (CPU) of the multi-GPU host logic: shard ranges cover the work exactly
once and the in-place all-gather reassembles the irradiance arrays on every rank."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n, q):
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    pkg = ge.load_package()
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    first, count, per = pkg.shard_range(n, rank, world)
    pos = torch.zeros(world * per, 4); E = torch.zeros(world * per, 32)
    idx = torch.arange(first, first + count, dtype=torch.float32)
    pos[first:first + count] = idx[:, None] * torch.tensor([1.0, 2.0, 3.0, 0.5])        # this rank's "irradiance pass"
    E[first:first + count] = idx[:, None] + torch.arange(32, dtype=torch.float32)[None] / 64
    pkg.allgather_inplace(dist, pos, rank, per); pkg.allgather_inplace(dist, E, rank, per)
    full = torch.arange(n, dtype=torch.float32)
    ok = torch.equal(pos[:n], full[:, None] * torch.tensor([1.0, 2.0, 3.0, 0.5])) and torch.equal(E[:n], full[:, None] + torch.arange(32, dtype=torch.float32)[None] / 64)
    ok = ok and bool((pos[n:] == 0).all())
    q.put((rank, first, count, ok))
    dist.destroy_process_group()


def test_shard_ranges():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    pkg = ge.load_package()
    for n in (0, 1, 7, 8, 24648458, 3686400):
        for world in (1, 2, 4, 8):
            cover = []
            for r in range(world):
                f, c, per = pkg.shard_range(n, r, world)
                assert c <= per and f + c <= n
                cover += list(range(f, f + c)) if n < 100 else [(f, c)]
            if n < 100:
                assert cover == list(range(n))
            else:
                assert sum(c for _, c in cover) == n and all(cover[i][0] + cover[i][1] == cover[i + 1][0] for i in range(world - 1))


def test_inplace_allgather_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    n, world, port = 1001, 2, 29731 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
    assert [r[3] for r in res] == [True, True]
    assert res[0][1:3] == (0, 501) and res[1][1:3] == (501, 500)


def test_tile_round_robin_partitions_all_samples(pkg):
    n, tile = 10_000, 1024
    for world in (1, 2, 3, 8):
        parts = [pkg.tile_round_robin(n, tile, r, world) for r in range(world)]
        allidx = np.sort(np.concatenate(parts))
        assert np.array_equal(allidx, np.arange(n))
        # whole tiles stay together and consecutive tiles go to consecutive ranks
        for r, p in enumerate(parts):
            assert np.all((p // tile) % world == r)
