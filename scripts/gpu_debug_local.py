import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
m = ge.package()
gold = os.path.join(ROOT, "tests", "golden")
os.environ["MOCOB200_KERNEL"] = "warp"
for name in ("gait2d_tracking", "hanging_muscle", "pendulum3_implicit", "sliding_mass"):
    opts = "num_mesh_intervals=10" if name.startswith("gait") else None
    p = m.Problem.load(os.path.join(gold, name + ".mb20"), opts)
    x = p.make_iterate(seed=9, spread=0.4)
    os.environ.pop("MOCOB200_FD_LOCAL", None)
    ev = m.Evaluator(p); ev.set_fd(0, 1e-4, 0)
    J = ev.eval_points(x, jacobian=True); Y = ev.eval_points(x)
    os.environ["MOCOB200_FD_LOCAL"] = "0"
    ev0 = m.Evaluator(p); ev0.set_fd(0, 1e-4, 0)
    J0 = ev0.eval_points(x, jacobian=True)
    scale = np.abs(Y)[:, :, None] + np.abs(J0) + 1e-300
    err = np.abs(J - J0) / scale
    print(name, "max err", np.nanmax(err), "nan mismatch", (np.isnan(J) != np.isnan(J0)).sum())
    idx = np.argsort(-np.nan_to_num(err).ravel())[:12]
    for i in idx:
        j, o, d = np.unravel_index(i, err.shape)
        print("   point %d row %d dir %d: local %.12g generic %.12g err %.3g" % (j, o, d, J[j, o, d], J0[j, o, d], err[j, o, d]))
