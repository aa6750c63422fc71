"""First GPU contact: run every fixture through the device path, report finiteness and rough timings."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
m = ge.package()
gold = os.path.join(ROOT, "tests", "golden")
print("fp64 peak TFLOP/s", m.measure_fp64_peak(0))
for name in sorted(os.listdir(gold)):
    if not name.endswith(".mb20"):
        continue
    p = m.Problem.load(os.path.join(gold, name))
    ev = m.Evaluator(p)
    x = p.make_iterate(seed=1, spread=0.2)
    Y = ev.eval_points(x)
    g = ev.eval_g(x)
    f = ev.eval_f(x)
    v = ev.eval_jac_g(x)
    gr = ev.eval_grad_f(x)
    print(name, "Y finite", np.isfinite(Y).all(), "g", np.isfinite(g).all(), "f", f, "jac", np.isfinite(v).all(),
          "grad", np.isfinite(gr).all(), "|Y|max", np.abs(Y).max())
    for fd in (0, 1):
        ev.set_fd(fd)
        ev.eval_jac_g(x)
        t0 = time.time(); n = 5
        for _ in range(n):
            ev.set_fd(fd)  # clears the x memo
            ev.eval_jac_g(x)
        dt = (time.time() - t0) / n
        print("   fd", fd, "jac ms (host api)", dt * 1e3, "kernel ms", ev.last_kernel_ms(), "evals/jac", ev.point_evals_per_jacobian())
for opts in ("num_mesh_intervals=800;optim_finite_difference_scheme=central",):
    p = m.Problem.load(os.path.join(gold, "gait2d_prediction.mb20"), opts)
    ev = m.Evaluator(p)
    x = p.make_iterate(seed=1, spread=0.2)
    for fd in (0, 1):
        ev.set_fd(fd); ev.eval_jac_g(x)
        t0 = time.time(); ev.set_fd(fd); ev.eval_jac_g(x); dt = time.time() - t0
        print("config5 fd", fd, "jac ms", dt * 1e3, "kernel ms", ev.last_kernel_ms(), "evals", ev.point_evals_per_jacobian(), p.dims().asdict())
