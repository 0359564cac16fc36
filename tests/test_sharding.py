"""Multi-GPU host logic on CPU: the shard plan, and a world_size-2 `gloo` run in which each
rank coarse-grains its shards (the oracle stands in for the kernels here: tests may use it)
and the gathered, re-assembled result equals the unsharded one bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from fv3net_b200 import sharding as S


def test_split_range_balanced_and_aligned():
    assert [S.split_range(10, 3, i) for i in range(3)] == [(0, 4), (4, 7), (7, 10)]
    assert [S.split_range(384, 5, i, multiple=8) for i in range(5)] == [
        (0, 80), (80, 160), (160, 240), (240, 312), (312, 384)]
    with pytest.raises(ValueError):
        S.split_range(10, 2, 0, multiple=4)


@pytest.mark.parametrize("n_time,world", [(8, 8), (12, 8), (1, 1), (3, 2), (1, 2), (1, 6), (2, 8), (1, 8), (1, 12)])
def test_plan_covers_everything_once(n_time, world):
    n_tile, ny, f = 6, 48, 8
    shards = S.plan(n_time, world, n_tile=n_tile, ny=ny, factor=f)
    assert len(shards) == world
    cover = np.zeros((n_time, n_tile, ny), int)
    for rank_shards in shards:
        for s in rank_shards:
            r1 = ny if s.row1 is None else s.row1
            assert s.row0 % f == 0 and r1 % f == 0
            cover[s.time, s.tile0:s.tile1, s.row0:r1] += 1
    assert (cover == 1).all()
    if n_time >= world:                                   # the reference's partition: whole timesteps
        assert all(s.tile0 == 0 and s.tile1 == n_tile for r in shards for s in r)
        counts = [len(r) for r in shards]
        assert max(counts) - min(counts) <= 1
    else:
        assert all(len(r) >= 1 for r in shards)


@pytest.mark.parametrize("n_time,world,n_tile", [(1, 8, 6), (1, 5, 6), (3, 4, 6), (1, 1, 6), (2, 3, 1)])
def test_coarse_buffers_layout_is_the_same_on_every_rank(n_time, world, n_tile):
    """The exchange buffer of sharding.CoarseBuffers: every rank computes the same layout from the plan alone; the slots
    are 16-byte aligned, padded to the longest rank, and the views of all ranks' shards tile them without overlap."""
    nz, ny, nx, f = 3, 32, 16, 4
    shards = S.plan(n_time, world, n_tile=n_tile, ny=ny, factor=f)
    shapes = {"q": ((nz,), ny, nx, f), "area": ((), ny, nx, f)}
    layouts = []
    for rank in range(world):
        b = S.CoarseBuffers(shards, shapes, rank, world, "cpu")
        assert b.whole.numel() == world * b.longest and (b.longest * 4) % 16 == 0
        assert [len(b.local[k]) for k in shapes] == [len(shards[rank])] * 2
        spans = []
        for name, per_shard in b.views().items():
            for s_, v in per_shard.items():
                start = (v.data_ptr() - b.whole.data_ptr()) // 4
                spans.append((start, start + v.numel(), name, s_))
                assert tuple(v.shape) == tuple(S.coarse_shape(s_, shapes[name][0], ny, nx, f))
        spans.sort()
        for (a0, a1, *_), (b0, b1, *_) in zip(spans, spans[1:]):
            assert a1 <= b0                                            # no two views overlap
        for r in range(world):                                        # every view of rank r lies inside slot r
            mine = [(a, e) for a, e, name, s_ in spans if s_ in shards[r]]
            assert all(r * b.longest <= a and e <= (r + 1) * b.longest for a, e in mine)
        for name in shapes:                                           # this rank's own views are the plan's, in order
            for v, s_ in zip(b.local[name], shards[rank]):
                assert v.data_ptr() == b.views()[name][s_].data_ptr()
        layouts.append([(a, e, name, tuple(s_)) for a, e, name, s_ in spans])
        assert b.peer_offsets() == [] and b.multicast_offset() is None     # not symmetric memory
    assert all(l == layouts[0] for l in layouts)
    total = sum(e - a for a, e, *_ in layouts[0])
    every = sum(int(np.prod(S.coarse_shape(s_, shapes[k][0], ny, nx, f))) for k in shapes for r in shards for s_ in r)
    assert total == every


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_time, n_tile, case, result_file):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nz, ny, nx, f = 3, 16, 16, 4
        rng = np.random.RandomState(7)
        field = rng.uniform(-1000, 1000, (n_time, n_tile, nz, ny, nx)).astype(np.float32)
        area = rng.uniform(0.5, 1, (n_tile, ny, nx)).astype(np.float32)
        shards = S.plan(n_time, world, n_tile=n_tile, ny=ny, factor=f)
        local = {"q": [], "area": []}
        for s in shards[rank]:
            q = S.take(s, field)                                           # (tiles, nz, rows, nx)
            a = S.take(s, area, time_axis=None, tile_axis=0)               # same tiles / rows of area
            local["q"].append(torch.from_numpy(oracle.weighted_block_average(q, a[:, None], f)))
            local["area"].append(torch.from_numpy(oracle.block_sum(a, f)))
        shapes = {"q": ((nz,), ny, nx, f), "area": ((), ny, nx, f)}
        everywhere = case.endswith("(all ranks)") or case.endswith("(in place)")
        if case.endswith("(in place)"):
            # the producers write into their slot of the receive buffer, the all-gather runs in place
            bufs = S.CoarseBuffers(shards, shapes, rank, world, "cpu")
            assert bufs.whole.numel() == world * bufs.longest and bufs.longest % 4 == 0
            for name, parts in local.items():
                assert len(parts) == len(bufs.local[name])
                for t, view in zip(parts, bufs.local[name]):
                    assert view.shape == t.shape and view.data_ptr() >= bufs.slot().data_ptr()
                    view.copy_(t)
            gathered = bufs.all_gather()
        else:
            gathered = S.gather_coarse(local, shards, shapes, dst=None if everywhere else 0)
        if everywhere:
            assert gathered is not None                     # one all-gather: every rank holds every shard
        if rank == 0:
            got_q = S.assemble(gathered["q"], n_time, n_tile, (nz,), ny, nx, f).numpy()
            got_a = S.assemble(gathered["area"], n_time, n_tile, (), ny, nx, f).numpy()
            exp_q = np.stack([oracle.weighted_block_average(field[t], area[:, None], f) for t in range(n_time)])
            exp_a = np.stack([oracle.block_sum(area, f)] * n_time)
            np.save(result_file, np.array([np.array_equal(got_q, exp_q), np.array_equal(got_a, exp_a)]))
        elif not everywhere:
            assert gathered is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_time,n_tile,case", [(4, 6, "by time"), (3, 6, "ragged by time"),
                                                (1, 6, "by tile"), (1, 1, "by row band"),
                                                (1, 6, "by tile (all ranks)"), (1, 5, "by row band across tiles (all ranks)"),
                                                (1, 5, "by row band across tiles (in place)"), (3, 6, "ragged by time (in place)")])
def test_world_size_2_gloo_matches_unsharded(tmp_path, n_time, n_tile, case):
    result_file = str(tmp_path / "ok.npy")
    mp.spawn(_worker, args=(2, _free_port(), n_time, n_tile, case, result_file), nprocs=2, join=True)
    assert np.load(result_file).all(), case
