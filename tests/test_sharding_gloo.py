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
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ctypes as C

    import __graft_entry__ as ge
    import oracle_py as O

    m = ge.package()
    m.lib().mocob200_set_min_points_per_rank(1)  # split even these small grids (every rank makes the same call)
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
            ok_blocks = float(np.array_equal(Y, full))
        elif p1 > p0:
            dist.send(torch.from_numpy(np.ascontiguousarray(mine)), dst=0)
        dist.barrier()
        # the Jacobian exchange (jacPipelineSharded): every rank owns the CSC value runs of the state / control / derivative
        # columns of its grid points (three contiguous runs, CasOCTranscription.h:192-199 column order), rank 0 the t0 / tf columns
        d = p.dims()
        colptr, rowidx = p.structure()
        vals = o.eval_jac_g(x.numpy())  # stands in for the rank's own assembly of its runs
        n_var, ns, nc, nd = d.n_var, d.ns, d.nc, d.nd
        off_s, off_c = 2, 2 + n_grid * ns
        off_d = off_c + n_grid * nc

        def runs_of(r):
            ra, rb = C.c_int(), C.c_int()
            m.lib().mocob200_shard_range(n_grid, r, world, C.byref(ra), C.byref(rb))
            return [(int(colptr[o0 + ra.value * w]), int(colptr[o0 + rb.value * w])) for o0, w in ((off_s, ns), (off_c, nc), (off_d, nd))]

        if rank == 0:
            out = np.full(vals.size, np.nan)
            count = np.zeros(vals.size, dtype=int)
            out[:colptr[2]] = vals[:colptr[2]]
            count[:colptr[2]] += 1
            for k0, k1 in runs_of(0):
                out[k0:k1] = vals[k0:k1]
                count[k0:k1] += 1
            for r in range(1, world):
                for k0, k1 in runs_of(r):
                    if k1 > k0:
                        buf = torch.empty(k1 - k0, dtype=torch.float64)
                        dist.recv(buf, src=r)
                        out[k0:k1] = buf.numpy()
                        count[k0:k1] += 1
            np.save(out_path, np.array([ok_blocks, float((covered == 1).all()), float(np.array_equal(out, vals, equal_nan=True)),
                                        float((count == 1).all())]))
        else:
            for k0, k1 in runs_of(rank):
                if k1 > k0:
                    dist.send(torch.from_numpy(np.ascontiguousarray(vals[k0:k1])), dst=0)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("fixture,opts,world", [("gait2d_tracking.mb20", "num_mesh_intervals=5", 2),
                                                ("pendulum2.mb20", None, 2),
                                                ("sliding_mass_trap.mb20", "num_mesh_intervals=2", 3)])
def test_sharded_point_blocks_gather_to_the_single_rank_result(pkg, orc, tmp_path, fixture, opts, world):
    out = str(tmp_path / "ok.npy")
    mp.spawn(_worker, args=(world, _free_port(), fixture, opts, out), nprocs=world, join=True)
    same, tiled, same_vals, tiled_vals = np.load(out)
    assert tiled == 1.0, "shards do not tile the grid exactly once"
    assert same == 1.0, "gathered blocks differ from the single-rank evaluation"
    assert tiled_vals == 1.0, "the ranks' value runs and the time columns do not tile the CSC value array exactly once"
    assert same_vals == 1.0, "the exchanged value runs differ from the single-rank Jacobian"


def test_shard_ranges_tile_any_grid(pkg, monkeypatch):
    import ctypes as C

    assert pkg.lib().mocob200_effective_world(101, 8) == 1 and pkg.lib().mocob200_effective_world(1601, 8) == 8
    assert pkg.lib().mocob200_effective_world(1601, 32) == 16
    pkg.lib().mocob200_set_min_points_per_rank(1)
    try:
        _check_tiling(pkg)
    finally:
        pkg.lib().mocob200_set_min_points_per_rank(0)


def _check_tiling(pkg):
    import ctypes as C

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
