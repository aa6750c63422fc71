import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as ge
import oracle_py as O
m = ge.package()
gold = os.path.join(ROOT, "tests", "golden")
def report(name, opts, scheme, h, seed, spread):
    p = m.Problem.load(os.path.join(gold, name + ".mb20"), opts)
    ev, o = m.Evaluator(p), O.Oracle(p)
    ev.set_fd(scheme, h, 0); o.set_fd(scheme, h, 0)
    x = p.make_iterate(seed=seed, spread=spread)
    J = ev.eval_points(x, jacobian=True)
    Yo, Jo = o.eval_point_jacobians(x)
    scale = np.abs(Yo)[:, :, None] + np.abs(Jo) + 1e-300
    e = np.abs(J - Jo) / scale
    idx = np.unravel_index(np.nanargmax(e), e.shape)
    print(name, opts, "scheme", scheme, "max", np.nanmax(e), "at (pt,out,in)", idx, "J", J[idx], "Jo", Jo[idx], "Yo", Yo[idx[0], idx[1]])
    worst_out = np.nanmax(e, axis=(0, 2)); print("  per-output max:", np.array2string(worst_out, precision=1))
    worst_in = np.nanmax(e, axis=(0, 1)); print("  per-input max:", np.array2string(worst_in, precision=1))
    v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
    ee = np.abs(v - vo) / (np.abs(vo) + 1e-6 * np.abs(vo).max())
    k = np.argmax(ee); colptr, rowidx = o.structure()
    col = np.searchsorted(colptr, k, side="right") - 1
    print("  jac max rel", ee[k], "k", k, "row", rowidx[k], "col", col, v[k], vo[k])
report("gait2d_tracking", "num_mesh_intervals=10", 1, 1e-3, 5, 0.3)
for name in ("sliding_mass_trap", "hanging_muscle", "gait2d_prediction"):
    report(name, "num_mesh_intervals=1", 1, 1e-3, 3, 0.4)
