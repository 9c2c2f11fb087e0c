"""world_size-2 gloo test (CPU) of the multi-GPU host logic (pbrt-v2-skin_b200/multigpu.py): the shard ranges cover the points
exactly once; rank 0's communicator id reaches every rank; and an in-place all-gather over arrays laid out as
skc_comm_allgather_positions / _rows expect them (world x padded shard, rank r's block at r x shard) reassembles the set on
every rank -- gloo's all_gather_into_tensor standing in for the ncclAllGather of csrc/skc_comm.cu, same layout, same call shape."""
import os
import sys

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
    ident = pkg.exchange_unique_id(dist, rank, lambda: bytes((7 * i + 3) % 256 for i in range(128)))
    first, count, per = pkg.shard_range(n, rank, world)
    pos = torch.zeros(world * per, 4); E = torch.zeros(world * per, 32)
    idx = torch.arange(first, first + count, dtype=torch.float32)
    pos[first:first + count] = idx[:, None] * torch.tensor([1.0, 2.0, 3.0, 0.5])        # this rank's "irradiance pass"
    E[first:first + count] = idx[:, None] + torch.arange(32, dtype=torch.float32)[None] / 64
    dist.all_gather_into_tensor(pos, pos[rank * per:(rank + 1) * per])                  # in place: send = own block of recv
    dist.all_gather_into_tensor(E, E[rank * per:(rank + 1) * per])
    full = torch.arange(n, dtype=torch.float32)
    ok = torch.equal(pos[:n], full[:, None] * torch.tensor([1.0, 2.0, 3.0, 0.5])) and torch.equal(E[:n], full[:, None] + torch.arange(32, dtype=torch.float32)[None] / 64)
    ok = ok and bool((pos[n:] == 0).all())
    q.put((rank, first, count, ok, ident))
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


def test_id_exchange_and_inplace_allgather_world2():
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
    assert res[0][4] == res[1][4] == bytes((7 * i + 3) % 256 for i in range(128))


def test_library_exports_the_comm_entry_points(pkg):
    lib = pkg.load_library()
    for s in ("skc_comm_unique_id", "skc_comm_create", "skc_comm_create_local", "skc_comm_allgather_positions", "skc_comm_allgather_rows_begin",
              "skc_comm_allgather_rows_wait", "skc_comm_reduce_film", "skc_comm_max_f32", "skc_comm_barrier", "skc_comm_allgather_host", "skc_comm_destroy"):
        assert hasattr(lib, s), s
