"""BASELINE metric (ii) stand-in: time-to-solve of 2D gait tracking with the GPU callbacks and with the CPU oracle's callbacks
inside the same host interior-point loop (scripts/ip_solve.py — NOT IPOPT, which is absent from this image; the reference's
loop is CasOCTranscription.cpp:630-721). Both solves start from the same guess: the reference walking trajectory
(referenceCoordinates.sto, as the tracking goal's splines hold it) with activations and excitations at 0.05.

usage: python scripts/solve_standin.py [--intervals 10] [--maxiter 1500] [--arms gpu,oracle] [--out profiles/solve_standin.json]
Prints one JSON record: per arm iterations, final objective, constraint violation, callback counts, callback seconds, total
seconds; and the two comparisons north_star asks for (objective difference, iteration-count ratio)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def quiet_guess(bench, m, prob):
    d = prob.dims()
    N, ns, nc = d.n_grid, d.ns, d.nc
    xs = bench.make_iterates(m, prob, 1, "trajectory")
    x0 = xs[0].copy()
    # make_iterates adds seeded noise to the coordinates and speeds (kept: both arms get the same vector) and draws the
    # muscle states / controls at random: replace those by a flat 0.05
    nq2 = 20
    st = x0[2:2 + N * ns].reshape(N, ns)
    st[:, nq2:] = 0.05
    x0[2 + N * ns:2 + N * ns + N * nc] = 0.05
    xl, xh, _, _ = prob.bounds()
    return np.clip(x0, xl, xh)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--intervals", type=int, default=10)
    ap.add_argument("--maxiter", type=int, default=1500)
    ap.add_argument("--tol", type=float, default=1e-6)
    ap.add_argument("--arms", default="gpu,oracle")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import __graft_entry__ as ge
    import bench
    from ip_solve import solve

    m = ge.package()
    prob = m.Problem.load(os.path.join(ROOT, "opensim-moco_b200", "data", "gait2d_tracking.mb20"), "num_mesh_intervals=%d" % a.intervals)
    d = prob.dims()
    x0 = quiet_guess(bench, m, prob)
    rec = {"problem": "example2DWalking gaitTracking, Hermite-Simpson, %d intervals (%d grid points), n_var %d, n_con %d, forward FD"
                      % (a.intervals, d.n_grid, d.n_var, d.n_con),
           "solver": "scripts/ip_solve.py: primal-dual barrier method, damped BFGS, filter line search - a stand-in, not IPOPT",
           "tol": a.tol, "maxiter": a.maxiter, "arms": {}}
    sols = {}
    for arm in a.arms.split(","):
        if arm == "gpu":
            ev = m.Evaluator(prob, device=0)
        else:
            import oracle_py

            oracle_py.set_threads(os.cpu_count() or 1)
            ev = oracle_py.Oracle(prob)
        t0 = time.perf_counter()
        res, st = solve(prob, ev, x0=x0, tol=a.tol, maxiter=a.maxiter)
        wall = time.perf_counter() - t0
        sols[arm] = res
        rec["arms"][arm] = {"status": st["status"], "iterations": st["iterations"], "objective": res.fun,
                            "constraint_violation": res.constraint_violation, "restorations": st.get("restorations", 0),
                            "calls": {k: st[k] for k in ("f", "grad", "g", "jac")}, "callback_seconds": st["t_eval"],
                            "total_seconds": wall, "host_threads": os.cpu_count(),
                            "objective_at_iteration": {str(h[0]): h[1] for h in st["history"] if h[0] in (0, 1, 2, 5, 10, 20, 50, 100, 150, 200, 300, 500, 1000)}}
        print(arm, json.dumps(rec["arms"][arm]), file=sys.stderr, flush=True)
    if len(sols) == 2:
        g, o = rec["arms"]["gpu"], rec["arms"]["oracle"]
        rec["compare"] = {"objective_rel_diff": abs(g["objective"] - o["objective"]) / max(1e-300, abs(o["objective"])),
                          "iteration_ratio": g["iterations"] / max(1, o["iterations"]),
                          "callback_speedup": o["callback_seconds"] / max(1e-12, g["callback_seconds"]),
                          "time_to_solve_speedup": o["total_seconds"] / max(1e-12, g["total_seconds"]),
                          "objective_path_rel_diff": {k: abs(v - o["objective_at_iteration"][k]) / max(1e-300, abs(o["objective_at_iteration"][k]))
                                                      for k, v in g["objective_at_iteration"].items() if k in o["objective_at_iteration"]},
                          "max_abs_solution_diff": float(np.abs(sols["gpu"].x - sols["oracle"].x).max())}
    line = json.dumps(rec)
    print(line)
    if a.out:
        with open(a.out, "w") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
