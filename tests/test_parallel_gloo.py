"""World-size-2 gloo tests of the multi-GPU host logic (renderengine_b200/parallel.py), run on the CPU.
The per-rank rendering is done by the CPU oracle here (the checker standing in for the GPU pass), so what is under
test is the partition / broadcast / gather logic the N>1 bench path uses."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from renderengine_b200 import parallel


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import orc
        from renderengine_b200 import scene as sc
        s = sc.make_scene("tiny")
        # shadow map: rendered on rank 0 only, broadcast to the others
        S = s["shadow_size"]
        shadow = torch.zeros((S, S), dtype=torch.float32)
        if rank == 0:
            shadow.copy_(torch.from_numpy(orc.Oracle(s, n_threads=2).render_shadow()))
        parallel.broadcast_shadow(shadow, src=0)
        # sort-first: every rank renders only its row bands, then the frame is composed by all-gather
        o = orc.Oracle(s, n_threads=2, rank=rank, n_ranks=world, band_h=16)
        part = o.render_forward(shadow.numpy())
        tri = parallel.gather_bands(torch.from_numpy(part["tri_id"].astype(np.int64)), band_h=16)
        col = parallel.gather_bands(torch.from_numpy(part["color_srgb8"]), band_h=16)
        # view-parallel: each rank renders its own views with the shared shadow map
        views = parallel.views_for_rank(5, rank, world)
        hashes = []
        full = orc.Oracle(s, n_threads=2)
        for v in views:
            full.set_view(sc.view_for_azimuth(s, 2 * np.pi * v / 5))
            hashes.append((v, int(full.render_forward(shadow.numpy())["tri_id"].astype(np.uint64).sum())))
        q.put((rank, tri.numpy(), col.numpy(), shadow.numpy().copy(), hashes))
    finally:
        dist.destroy_process_group()


def test_partition_helpers():
    assert parallel.views_for_rank(512, 3, 8) == list(range(192, 256))
    assert sum(len(parallel.views_for_rank(5, r, 2)) for r in range(2)) == 5
    m0, m1 = parallel.owned_rows(96, 0, 2, 16), parallel.owned_rows(96, 1, 2, 16)
    assert not (m0 & m1).any() and (m0 | m1).all()
    assert parallel.owned_rows(10, 0, 1).all()


def test_two_rank_sort_first_and_view_parallel():
    from oracle import orc
    from renderengine_b200 import scene as sc
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=240) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    s = sc.make_scene("tiny")
    o = orc.Oracle(s, n_threads=2)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    for rank, tri, col, shadow, hashes in res:
        assert np.array_equal(shadow.view(np.uint32), ref_shadow.view(np.uint32)), "broadcast shadow map differs"
        assert np.array_equal(tri.astype(np.uint32), ref["tri_id"]), "composed sort-first frame differs from the single-rank frame"
        assert np.array_equal(col, ref["color_srgb8"])
    all_views = sorted(v for r in res for v, _ in r[4])
    assert all_views == list(range(5))
    for r in res:
        for v, h in r[4]:
            o.set_view(sc.view_for_azimuth(s, 2 * np.pi * v / 5))
            assert h == int(o.render_forward(ref_shadow)["tri_id"].astype(np.uint64).sum())


def _slab_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        S = 128
        rows = parallel.slab_rows(S, world)
        full = torch.arange(S * S, dtype=torch.float32).reshape(S, S)
        mine = torch.ones((S, S), dtype=torch.float32)                      # rows outside the slab hold the clear value
        mine[rank * rows:(rank + 1) * rows] = full[rank * rows:(rank + 1) * rows]
        parallel.allgather_shadow_slabs(mine)
        q.put((rank, bool(torch.equal(mine, full))))
    finally:
        dist.destroy_process_group()


def test_two_rank_shadow_slab_allgather():
    """Shadow pass split by rows (reb_set_shadow_partition): the in-place all-gather of the slabs completes the map."""
    assert parallel.slab_rows(2048, 8) == 256 and parallel.slab_rows(512, 4) == 128 and parallel.slab_rows(2048, 3) is None
    assert parallel.contiguous_band(4320, 8) == 544 and parallel.contiguous_band(360, 4) == 96 and parallel.contiguous_band(1080, 1) == 1080
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_slab_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]


def _min_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # every rank holds a whole map with its share of the "triangles" rasterised into it (here: every world-th texel) over 1.0
        S = 64
        g = torch.Generator().manual_seed(7)
        full = torch.rand((S, S), generator=g, dtype=torch.float32)
        mine = torch.ones((S, S), dtype=torch.float32)
        sel = (torch.arange(S * S).reshape(S, S) % world) == rank
        mine[sel] = full[sel]
        parallel.allreduce_shadow_min(mine)
        q.put((rank, bool(torch.equal(mine, full))))
    finally:
        dist.destroy_process_group()


def test_two_rank_shadow_split_by_triangles_allreduce_min():
    """Shadow pass split by triangles (reb_set_shadow_split): the element-wise minimum of the ranks' whole maps is the map."""
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_min_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]
