"""Algorithmic FLOP / libm-call counts per point-function evaluation (SURVEY.md §8(d) convention), from the oracle
instantiated with its operation-counting scalar. Averages over the grid points of 3 seeded iterates (branches such as
conditional path points make the count iterate-dependent by a few operations). Writes profiles/flop_counts.json."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O
gold = os.path.join(ROOT, "tests", "golden")
out = {"convention": "+,-,*,/,sqrt,abs,min,max = 1; each libm call = 1 (also listed as transc); oracle formulation "
                     "(Newton-Euler + composite-rigid-body mass matrix + Cholesky)", "problems": {}}
for name in sorted(os.listdir(gold)):
    if not name.endswith(".mb20"): continue
    o = O.Oracle(os.path.join(gold, name))
    rows = {}
    masks = {"all": 0xFFFFFFFF, "dynamics": 1}
    for k in range(o.n_integrand): masks["integrand%d" % k] = 1 << (1 + k)
    for label, mask in masks.items():
        fl, tr = [], []
        for seed in (0, 1, 2):
            x = o.problem.make_iterate(seed=seed, spread=0.3)
            for j in range(0, o.n_grid, max(1, o.n_grid // 10)):
                f, t = o.count_ops(o.gather_inputs(x, j), mask)
                fl.append(f); tr.append(t)
        rows[label] = {"flop": float(np.mean(fl)), "transc": float(np.mean(tr))}
    rows["bytes_io"] = 8 * (o.n_in + o.n_out)
    rows["n_in"], rows["n_out"] = o.n_in, o.n_out
    out["problems"][name[:-5]] = rows
    print(name, rows)
json.dump(out, open(os.path.join(ROOT, "profiles", "flop_counts.json"), "w"), indent=1)
