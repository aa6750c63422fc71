"""Where do the warp-per-evaluation kernels overtake the thread-per-evaluation ones on the config-5 problem (central FD) as the
number of grid points per GPU shrinks (what sharding over 4 / 8 ranks leaves each rank)?  prints ms per Jacobian."""
import os, sys, time
os.environ["MOCOB200_DEVELOPER_SWITCHES"] = "1"
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
m = ge.package()
for n in (50, 100, 200, 400, 800):
    row = []
    for kern in ("thread", "warp", ""):
        if kern: os.environ["MOCOB200_KERNEL"] = kern
        else: os.environ.pop("MOCOB200_KERNEL", None)
        p = m.Problem.load(os.path.join(ROOT, "opensim-moco_b200/data/gait2d_prediction.mb20"), "num_mesh_intervals=%d" % n)
        ev = m.Evaluator(p)
        ev.set_fd(1)
        xs = [torch.from_numpy(p.make_iterate(seed=k, spread=0.1)).cuda() for k in range(8)]
        vals = torch.empty(ev.dims.nnz_jac, dtype=torch.float64, device="cuda")
        for k in range(5): ev.eval_jac_g_device(xs[k % 8].data_ptr(), vals.data_ptr())
        ev.sync()
        t0 = time.perf_counter()
        K = 30
        for k in range(K): ev.eval_jac_g_device(xs[k % 8].data_ptr(), vals.data_ptr())
        ev.sync()
        row.append(1e3 * (time.perf_counter() - t0) / K)
        ev.close()
    print("intervals %4d grid points %5d   thread %.3f ms   warp %.3f ms   default %.3f ms" % (n, 2 * n + 1, row[0], row[1], row[2]), flush=True)
