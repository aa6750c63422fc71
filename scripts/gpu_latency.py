"""Device-side latency of the stages of one Jacobian (config 3 / config 5), warp vs thread kernels."""
import os as _os; _os.environ.setdefault("MOCOB200_DEVELOPER_SWITCHES", "1")
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
m = ge.package()
gold = os.path.join(ROOT, "tests", "golden")
for fixture, opts, fdm in (("gait2d_tracking.mb20", None, 0), ("gait2d_prediction.mb20", "num_mesh_intervals=800", 0), ("gait2d_prediction.mb20", "num_mesh_intervals=800", 1)):
    for mode in ("warp", "thread"):
        os.environ["MOCOB200_KERNEL"] = mode
        p = m.Problem.load(os.path.join(gold, fixture), opts)
        ev = m.Evaluator(p); ev.set_fd(fdm)
        x = torch.from_numpy(p.make_iterate(seed=1, spread=0.1)).cuda()
        out = {}
        for name, fn in (("eval_g", lambda: ev.eval_g_device(x.data_ptr())), ("eval_jac_g", lambda: ev.eval_jac_g_device(x.data_ptr()))):
            for _ in range(3): fn()
            ev.sync(); ts = []
            for _ in range(10):
                ev.mark(0); fn(); ev.mark(1); ev.sync(); ts.append(ev.elapsed_ms())
            out[name] = min(ts)
        print(fixture, opts, "fd", fdm, mode, {k: round(v, 4) for k, v in out.items()}, "points/asm kernel ms", ev.last_kernel_ms(), flush=True)
