"""Sliding-mass NLP solved twice with the same host SQP loop (tests/nlp_solve.py): callbacks from the GPU evaluator and from the
CPU oracle. Prints iterations, objective and time spent inside the callbacks -> profiles/r01/solve_r01.json. A 65-variable problem
is launch-latency territory: the point of the report is identical convergence, not speed."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import __graft_entry__ as ge  # noqa: E402
import oracle_py as O  # noqa: E402
from nlp_solve import solve  # noqa: E402

m = ge.package()
O.set_threads(1)  # a few hundred evaluations of a 20-point problem: OpenMP fork/join would dominate the oracle's time
out = []
for prob, scheme, n in (("sliding_mass", "trapezoidal", 20), ("sliding_mass", "hermite-simpson", 20),
                        ("hanging_muscle_rigidtendon", "hermite-simpson", 10), ("hanging_muscle", "hermite-simpson", 10),
                        ("pendulum:2", "hermite-simpson", 10)):
    p = m.Problem.example(prob, ".", "num_mesh_intervals=%d;transcription_scheme=%s" % (n, scheme))
    row = {"problem": prob, "scheme": scheme, "intervals": n, "n_var": p.dims().n_var}
    if prob == "sliding_mass":
        row["exact_final_time"] = 0.4
    for name, ev in (("gpu", m.Evaluator(p)), ("cpu_oracle", O.Oracle(p))):
        res, st = solve(p, ev)
        row[name] = {"success": bool(res.success), "objective": float(res.fun), "iterations": st["iterations"],
                     "callbacks": {k: st[k] for k in ("f", "grad", "g", "jac")}, "callback_seconds": st["t_eval"],
                     "total_seconds": st["t_total"]}
    row["objective_difference"] = abs(row["gpu"]["objective"] - row["cpu_oracle"]["objective"])
    out.append(row)
    print(json.dumps(row), flush=True)
