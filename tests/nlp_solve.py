"""A host NLP loop for tests: scipy's SLSQP driving the four TNLP callbacks (eval_f, eval_grad_f, eval_g, eval_jac_g with the CSC
structure) of an evaluator — the calls IPOPT makes through CasADi's nlpsol (CasOCTranscription.cpp:678-689). IPOPT and CasADi
are absent from this image; scipy stands in for them, in tests only, on problems small enough for a dense SQP."""
import time

import numpy as np


def solve(problem, evaluator, x0=None, maxiter=200, ftol=1e-10):
    from scipy.optimize import minimize
    from scipy.sparse import csc_matrix

    xl, xh, gl, gh = problem.bounds()
    colptr, rowidx = evaluator.structure()
    n_con, n_var = gl.size, xl.size
    stats = {"f": 0, "grad": 0, "g": 0, "jac": 0, "t_eval": 0.0}

    def timed(key, fn, x):
        t0 = time.perf_counter()
        r = fn(np.ascontiguousarray(x, np.float64))
        stats["t_eval"] += time.perf_counter() - t0
        stats[key] += 1
        return r

    def g(x):
        return np.array(timed("g", evaluator.eval_g, x))

    def jac(x):
        v = timed("jac", evaluator.eval_jac_g, x)
        return csc_matrix((np.array(v), rowidx, colptr), shape=(n_con, n_var)).toarray()

    eq = gl == gh
    lo, hi = (~eq) & np.isfinite(gl), (~eq) & np.isfinite(gh)
    cons = [{"type": "eq", "fun": lambda x: g(x)[eq] - gl[eq], "jac": lambda x: jac(x)[eq]}]
    if lo.any():
        cons.append({"type": "ineq", "fun": lambda x: g(x)[lo] - gl[lo], "jac": lambda x: jac(x)[lo]})
    if hi.any():
        cons.append({"type": "ineq", "fun": lambda x: gh[hi] - g(x)[hi], "jac": lambda x: -jac(x)[hi]})
    if x0 is None:
        x0 = problem.make_iterate(seed=-1, spread=0.0)  # bounds midpoint, as createGuess("bounds")
    x0 = np.clip(x0, xl, xh)
    t0 = time.perf_counter()
    res = minimize(lambda x: float(timed("f", evaluator.eval_f, x)), x0, jac=lambda x: np.array(timed("grad", evaluator.eval_grad_f, x)),
                   method="SLSQP", bounds=list(zip(xl, xh)), constraints=cons, options={"maxiter": maxiter, "ftol": ftol})
    stats["t_total"] = time.perf_counter() - t0
    stats["iterations"] = int(res.nit)
    return res, stats
