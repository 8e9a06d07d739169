"""Host-side multi-GPU logic (tiler_b200/shard.py) on CPU: world_size 2, gloo backend.

The search itself is replaced by the CPU oracle through ShardedFrameTiling's injection points
(tests may use the oracle; the product default is libtmann).  What is covered: whole-frame query
sharding, the one-time tileset broadcast (rank 0 wins), ragged result gather, and that the
gathered tilemap equals the single-process answer.
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_frame_shards_partition():
    from tiler_b200.shard import frame_shards, point_shards
    for frames, pos, world in [(64, 1200, 1), (64, 1200, 2), (7, 10, 4), (3, 5, 8), (0, 5, 2), (300, 14400, 8)]:
        sh = frame_shards(frames, pos, world)
        assert len(sh) == world and sh[0][0] == 0 and sh[-1][1] == frames * pos
        for (a0, a1), (b0, b1) in zip(sh, sh[1:]):
            assert a1 == b0
        sizes = [(b - a) // max(pos, 1) for a, b in sh]
        assert all((b - a) % max(pos, 1) == 0 for a, b in sh)  # whole frames only
        assert max(sizes) - min(sizes) <= 1
    assert point_shards(10, 3) == [(0, 4), (4, 7), (7, 10)]
    with pytest.raises(ValueError):
        frame_shards(4, 4, 0)


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import tmoracle as o
    from tiler_b200.shard import ShardedFrameTiling, frame_shards
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frames, pos = 5, 40  # 5 frames over 2 ranks: ragged shards (3 + 2 frames)
        pal, s = o.gen_palette(o.BASE_SEED + 11)
        tiles, s = o.gen_tileset(s, 64)
        rgb, s = o.gen_queries(s, frames * pos, tiles, pal)
        t_tiles = torch.from_numpy(tiles.copy())
        t_pal = torch.from_numpy(pal.copy()).view(1, -1)
        if rank != 0:  # other ranks start with garbage: the broadcast must overwrite it
            t_tiles.zero_()
            t_pal.zero_()

        def build(palpix, pals, flags):
            return o.psyv_tileset_f32(palpix.numpy(), pals.numpy()[0])

        def search(ds, rgb_local, out_idx, out_err):
            i, d = o.knn_f32(ds, o.psyv_queries_f32(rgb_local.numpy()), 1)
            return torch.from_numpy(i[:, 0].copy()), torch.from_numpy(d[:, 0].copy())

        sft = ShardedFrameTiling(search_fn=search, build_fn=build)
        ds = sft.broadcast_tileset(t_tiles, t_pal)
        assert np.array_equal(t_tiles.numpy(), tiles) and np.array_equal(t_pal.numpy()[0], pal)
        shards = frame_shards(frames, pos, world)
        q0, q1 = shards[rank]
        idx, err = sft.search_local(torch.from_numpy(rgb[q0:q1].copy()))
        gi, ge = sft.gather_results(idx, err, [b - a for a, b in shards])
        if rank == 0:
            wi, wd = o.knn_f32(ds, o.psyv_queries_f32(rgb), 1)
            assert np.array_equal(gi.numpy(), wi[:, 0]) and np.array_equal(ge.numpy(), wd[:, 0])
        else:
            assert gi is None and ge is None
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


def _kmodes_worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import tmoracle as o
    from tiler_b200.shard import ShardedKModesAssign, point_shards
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tiles, _ = o.gen_tileset(o.BASE_SEED + 31, 333 + 60)
        rows = o.kmodes_rows(tiles)
        X, cent = rows[:333], rows[333:].copy()       # 333 points over 2 ranks: ragged (167 + 166)
        t_cent = torch.from_numpy(cent.copy())
        if rank != 0:
            t_cent.zero_()                            # the broadcast must overwrite it

        def assign(xl, c):
            cl, di = o.kmodes_assign(xl.numpy(), c.numpy())
            return torch.from_numpy(cl.copy()), torch.from_numpy(di.astype(np.int64))

        ska = ShardedKModesAssign(assign_fn=assign)
        shards = point_shards(len(X), world)
        a, b = shards[rank]
        gc, gd = ska.assign(torch.from_numpy(X[a:b].copy()), t_cent, [q - p for p, q in shards])
        assert np.array_equal(t_cent.numpy(), cent)
        if rank == 0:
            wc, wd = o.kmodes_assign(X, cent)
            assert np.array_equal(gc.numpy(), wc) and np.array_equal(gd.numpy(), wd.astype(np.int64))
        else:
            assert gc is None and gd is None
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


def _run_world2(target):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=target, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    for p in procs:
        if p.is_alive():
            p.terminate()
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(ret) == {0: "ok", 1: "ok"}


def test_sharded_kmodes_assign_world2_gloo():
    _run_world2(_kmodes_worker)


def test_sharded_frametiling_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    for p in procs:
        if p.is_alive():
            p.terminate()
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(ret) == {0: "ok", 1: "ok"}
