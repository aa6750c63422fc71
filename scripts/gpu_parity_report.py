"""Print device-vs-oracle error statistics for every fixture (used to calibrate the tolerances in tests/)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as ge
import oracle_py as O
m = ge.package()
gold = os.path.join(ROOT, "tests", "golden")
def rel(a, b, floor):
    return np.max(np.abs(a - b) / (np.abs(b) + floor))
for name in sorted(os.listdir(gold)):
    if not name.endswith(".mb20"): continue
    opts = "num_mesh_intervals=10" if name.startswith("gait2d") else None
    p = m.Problem.load(os.path.join(gold, name), opts)
    ev = m.Evaluator(p); o = O.Oracle(p)
    for seed, spread in ((1, 0.1), (2, 0.5), (3, 1.0)):
        x = p.make_iterate(seed=seed, spread=spread)
        Y, Yo = ev.eval_points(x), o.eval_points(x)
        g, go = ev.eval_g(x), o.eval_g(x)
        f, fo = ev.eval_f(x), o.eval_f(x)
        line = "%s seed%d Y %.2e g %.2e (gmax %.1e) f %.2e" % (name, seed, rel(Y, Yo, 1e-6 * np.abs(Yo).max() + 1e-300),
              rel(g, go, 1e-6 * np.abs(go).max() + 1e-300), np.abs(go).max(), abs(f - fo) / (abs(fo) + 1e-300))
        for fd in (0, 1):
            ev.set_fd(fd); o.set_fd(fd)
            J, (_, Jo) = ev.eval_points(x, jacobian=True), o.eval_point_jacobians(x)
            v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
            gr, gro = ev.eval_grad_f(x), o.eval_grad_f(x)
            # scale per output row of each block: FD noise is proportional to |y| eps / h
            scale = np.abs(Yo)[:, :, None] + np.abs(Jo)
            line += " | fd%d J %.2e (abs/rowscale %.2e) jac %.2e grad %.2e" % (fd, rel(J, Jo, 1e-6 * np.abs(Jo).max() + 1e-300),
                    np.nanmax(np.abs(J - Jo) / (scale + 1e-300)), rel(v, vo, 1e-6 * np.abs(vo).max() + 1e-300),
                    rel(gr, gro, 1e-6 * np.abs(gro).max() + 1e-300))
        print(line, flush=True)
