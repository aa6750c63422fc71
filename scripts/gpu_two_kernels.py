"""Development: run the nominal kernel and the one-sided FD kernels on the same problem (for ncu captures)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
m = ge.package()
p = m.Problem.load(os.path.join(ROOT, "tests", "golden", "gait2d_tracking.mb20"), os.environ.get("PH_OPTS"))
ev = m.Evaluator(p)
x = p.make_iterate(seed=1, spread=0.1)
for _ in range(4):
    ev.eval_g(x)
    ev.eval_jac_g(x + 1e-9)
