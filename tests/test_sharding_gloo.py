"""N>1 host logic on CPU: two gloo ranks evaluate their shards of the grid points (with the oracle standing in for the
device kernels), rank 0 gathers the per-point blocks exactly as mocob200's NCCL gather lays them out, and the result
must equal the single-rank evaluation bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, golden


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fixture, opts, out_path):
    os.environ["MOCOB200_MIN_POINTS_PER_RANK"] = "1"  # split even these small grids
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ctypes as C

    import __graft_entry__ as ge
    import oracle_py as O

    m = ge.package()
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        p = m.Problem.load(golden(fixture), opts)
        o = O.Oracle(p)
        n_grid, n_out = o.n_grid, o.n_out
        # rank 0 owns x; the others receive it (mocob200: ncclBroadcast of x)
        x = torch.from_numpy(p.make_iterate(seed=4, spread=0.3)) if rank == 0 else torch.empty(p.dims().n_var, dtype=torch.float64)
        dist.broadcast(x, 0)
        a, b = C.c_int(), C.c_int()
        assert m.lib().mocob200_shard_range(n_grid, rank, world, C.byref(a), C.byref(b)) == 0
        p0, p1 = a.value, b.value
        mine = np.stack([o.eval_point(o.gather_inputs(x.numpy(), j)) for j in range(p0, p1)]) if p1 > p0 else np.empty((0, n_out))
        # gather of contiguous row ranges to rank 0 at offset p0 * stride (mocob200: grouped ncclSend/ncclRecv)
        if rank == 0:
            Y = np.full((n_grid, n_out), np.nan)
            Y[p0:p1] = mine
            covered = np.zeros(n_grid, dtype=int)
            covered[p0:p1] += 1
            for r in range(1, world):
                ra, rb = C.c_int(), C.c_int()
                m.lib().mocob200_shard_range(n_grid, r, world, C.byref(ra), C.byref(rb))
                buf = torch.empty((rb.value - ra.value, n_out), dtype=torch.float64)
                if rb.value > ra.value:
                    dist.recv(buf, src=r)
                Y[ra.value:rb.value] = buf.numpy()
                covered[ra.value:rb.value] += 1
            full = o.eval_points(x.numpy())
            np.save(out_path, np.array([float(np.array_equal(Y, full)), float((covered == 1).all())]))
        elif p1 > p0:
            dist.send(torch.from_numpy(np.ascontiguousarray(mine)), dst=0)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("fixture,opts,world", [("gait2d_tracking.mb20", "num_mesh_intervals=5", 2),
                                                ("pendulum2.mb20", None, 2),
                                                ("sliding_mass_trap.mb20", "num_mesh_intervals=2", 3)])
def test_sharded_point_blocks_gather_to_the_single_rank_result(pkg, orc, tmp_path, fixture, opts, world):
    out = str(tmp_path / "ok.npy")
    mp.spawn(_worker, args=(world, _free_port(), fixture, opts, out), nprocs=world, join=True)
    same, tiled = np.load(out)
    assert tiled == 1.0, "shards do not tile the grid exactly once"
    assert same == 1.0, "gathered blocks differ from the single-rank evaluation"


def test_shard_ranges_tile_any_grid(pkg, monkeypatch):
    import ctypes as C

    assert pkg.lib().mocob200_effective_world(101, 8) == 1 and pkg.lib().mocob200_effective_world(1601, 8) == 4
    monkeypatch.setenv("MOCOB200_MIN_POINTS_PER_RANK", "1")

    for n_grid in (0, 1, 3, 20, 101, 1601):
        for world in (1, 2, 3, 4, 8):
            edges = []
            for r in range(world):
                a, b = C.c_int(), C.c_int()
                assert pkg.lib().mocob200_shard_range(n_grid, r, world, C.byref(a), C.byref(b)) == 0
                edges.append((a.value, b.value))
            eff = pkg.lib().mocob200_effective_world(n_grid, world)
            assert 1 <= eff <= world
            assert edges[0][0] == 0 and edges[-1][1] == n_grid
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes[:eff]) - min(sizes[:eff]) <= 1 and all(sz == 0 for sz in sizes[eff:])
    a, b = C.c_int(), C.c_int()
    assert pkg.lib().mocob200_shard_range(10, 2, 2, C.byref(a), C.byref(b)) != 0
