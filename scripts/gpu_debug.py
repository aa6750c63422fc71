import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as ge
import oracle_py as O
m = ge.package()
gold = os.path.join(ROOT, "tests", "golden")
p = m.Problem.load(os.path.join(gold, "gait2d_tracking.mb20"), "num_mesh_intervals=10")
ev, o = m.Evaluator(p), O.Oracle(p)
for scheme, h in ((0, 1e-4), (1, 1e-3)):
    ev.set_fd(scheme, h, 0); o.set_fd(scheme, h, 0)
    x = p.make_iterate(seed=5, spread=0.3)
    gr, gro = ev.eval_grad_f(x), o.eval_grad_f(x)
    d = np.abs(gr - gro)
    idx = np.argsort(-d)[:8]
    print("scheme", scheme, "f", o.eval_f(x, terms=True), "max|gro|", np.abs(gro).max())
    for i in idx: print("   var", i, "gpu", gr[i], "orc", gro[i], "diff", d[i])
    J = ev.eval_points(x, jacobian=True); Yo, Jo = o.eval_point_jacobians(x)
    for row in range(28, 31):
        e = np.abs(J[:, row, :] - Jo[:, row, :]); k = np.unravel_index(np.argmax(e), e.shape)
        print("   integrand row", row, "max abs diff", e.max(), "at", k, J[k[0], row, k[1]], Jo[k[0], row, k[1]], "Y", Yo[k[0], row])
