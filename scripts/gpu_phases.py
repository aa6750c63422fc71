"""Development: phase stopwatch of one warp evaluation (library built with MOCOB200_PROFILE_PHASES=1) and the NaN hunt."""
import os as _os; _os.environ.setdefault("MOCOB200_DEVELOPER_SWITCHES", "1")
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
m = ge.package()
gold = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
names = ["t0", "kin", "inertia", "musgeom", "forces", "gather", "inward", "mass", "ldlt", "solve", "integrands", "lookups"]
def clocks():
    buf = (C.c_longlong * 32)()
    assert m.lib().mocob200_debug_phase_clocks(buf) == 0
    a = np.array(buf[:12])
    return a
if hasattr(m.lib(), "mocob200_debug_phase_clocks"):
    p = m.Problem.load(os.path.join(gold, "gait2d_tracking.mb20"), os.environ.get("PH_OPTS"))
    ev = m.Evaluator(p)
    x = p.make_iterate(seed=1, spread=0.1)
    for what in ("nominal", "jac"):
        for rep in range(3):
            if what == "nominal": ev.eval_points(x)
            else: ev.eval_jac_g(x)
            a = clocks()
            d = dict(zip(names[1:11], (a[1:11] - a[:10]).tolist()))
            d["lookups"] = int(a[11] - a[0]); d["kin"] = int(a[1] - a[11])
            print(what, rep, "total", a[10] - a[0], d)
            if what == "nominal" and hasattr(m.lib(), "mocob200_debug_item_clocks"):
                n = min(ev.dims.n_grid, 16384)
                buf = (C.c_longlong * n)()
                m.lib().mocob200_debug_item_clocks(buf, 0, n)
                b = np.array(buf[:n])
                print("   nominal kernel item cycles: min %d median %d max %d (n=%d)" % (b.min(), np.median(b), b.max(), n))
if hasattr(m.lib(), "mocob200_debug_item_clocks") and not os.environ.get("PH_OPTS"):
    ev.eval_jac_g(x); ev.eval_jac_g(x)
    for phase, per in ((0, 11), (1, 10), (2, 1), (3, 1)):
        n = 101 * per
        buf = (C.c_longlong * n)()
        assert m.lib().mocob200_debug_item_clocks(buf, phase, n) == 0
        a = np.array(buf[:n]).reshape(101, per)
        print("phase", phase, "item cycles by type: median", np.median(a, axis=0).astype(int).tolist(), "max", a.max(axis=0).tolist())
sys.exit(0)
# NaN hunt: central h=1e-3 on gait2d_tracking
sys.path.insert(0, os.path.join(os.path.dirname(gold), "..", "oracle"))
import oracle_py as O
p = m.Problem.load(os.path.join(gold, "gait2d_tracking.mb20"), "num_mesh_intervals=10")
ev, o = m.Evaluator(p), O.Oracle(p)
ev.set_fd(1, 1e-3, 0); o.set_fd(1, 1e-3, 0)
x = p.make_iterate(seed=5, spread=0.3)
J = ev.eval_points(x, jacobian=True)
Yo, Jo = o.eval_point_jacobians(x)
print("nan gpu", np.argwhere(np.isnan(J))[:20].tolist(), "nan oracle", np.argwhere(np.isnan(Jo))[:20].tolist())
print("n nan", np.isnan(J).sum(), np.isnan(Jo).sum())
