"""A primal-dual interior-point loop for tests and for the time-to-solve stand-in (scripts/solve_standin.py): the algorithm
family of the solver the reference drives (IPOPT through CasADi's nlpsol, CasOCTranscription.cpp:630-689, with
optim_hessian_approximation = "limited-memory", MocoCasADiSolver.h:166-170) written against the four TNLP callbacks
(eval_f, eval_grad_f, eval_g, eval_jac_g + CSC structure). NOT IPOPT: IPOPT, MUMPS and CasADi are absent from this image; this
is a plain barrier method (log barrier on the variable bounds and on slacks of inequality rows, damped-BFGS Lagrangian Hessian
held as a dense matrix, filter line search with a minimum-norm feasibility step in place of the restoration phase,
Fiacco-McCormick barrier schedule, gradient-based scaling), good for grids of a few
dozen intervals. What it shares with the reference's loop is what the metric needs: one eval_jac_g + eval_grad_f per
iteration, eval_f + eval_g per line-search trial, and sensitivity of the iteration path to the callbacks' values."""
import time

import numpy as np


def solve(problem, evaluator, x0=None, tol=1e-6, maxiter=500, mu0=0.1, verbose=False, scale_hessian=False):
    from scipy.sparse import csc_matrix

    xl, xh, gl, gh = [np.array(a, np.float64) for a in problem.bounds()]
    colptr, rowidx = evaluator.structure()
    m, n = gl.size, xl.size
    stats = {"f": 0, "grad": 0, "g": 0, "jac": 0, "t_eval": 0.0}

    def timed(key, fn, x):
        t0 = time.perf_counter()
        r = np.array(fn(np.ascontiguousarray(x, np.float64)), np.float64)
        stats["t_eval"] += time.perf_counter() - t0
        stats[key] += 1
        return r

    sc = {"f": 1.0, "g": np.ones(m)}  # gradient-based scaling fixed at the starting point (IPOPT's nlp_scaling_max_gradient = 100)
    f_ = lambda x: sc["f"] * float(timed("f", evaluator.eval_f, x))
    df_ = lambda x: sc["f"] * timed("grad", evaluator.eval_grad_f, x)
    g_ = lambda x: sc["g"] * timed("g", evaluator.eval_g, x)
    J_ = lambda x: sc["g"][:, None] * csc_matrix((timed("jac", evaluator.eval_jac_g, x), rowidx, colptr), shape=(m, n)).toarray()

    # z = (x_free, s): fixed variables leave the problem, inequality rows get slacks with their own bounds
    fixed = xl == xh
    free = np.flatnonzero(~fixed)
    ineq = np.flatnonzero(gl != gh)
    nf, ni = free.size, ineq.size
    nz = nf + ni
    zl = np.concatenate([xl[free], gl[ineq]])
    zh = np.concatenate([xh[free], gh[ineq]])
    hasl, hash_ = np.isfinite(zl), np.isfinite(zh)
    rhs = np.where(gl == gh, gl, 0.0)

    if x0 is None:
        x0 = problem.make_iterate(seed=-1, spread=0.0)
    x = np.clip(np.array(x0, np.float64), xl, xh)

    def push(z):  # IPOPT's bound_push / bound_frac = 1e-2
        z = z.copy()
        span = np.where(hasl & hash_, zh - zl, np.inf)
        pl = np.minimum(1e-2 * np.maximum(1.0, np.abs(np.where(hasl, zl, 0.0))), 1e-2 * span)
        ph = np.minimum(1e-2 * np.maximum(1.0, np.abs(np.where(hash_, zh, 0.0))), 1e-2 * span)
        z = np.where(hasl, np.maximum(z, zl + pl), z)
        z = np.where(hash_, np.minimum(z, zh - ph), z)
        return z

    def full(z):
        xx = x.copy()
        xx[free] = z[:nf]
        return xx

    sc["g"] = np.minimum(1.0, 100.0 / np.maximum(np.abs(J_(x)).max(axis=1), 1e-8))
    sc["f"] = min(1.0, 100.0 / max(np.abs(df_(x)).max(), 1e-8))
    gl, gh = gl * sc["g"], gh * sc["g"]
    zl[nf:], zh[nf:] = gl[ineq], gh[ineq]
    rhs = np.where(gl == gh, gl, 0.0)
    g0 = g_(x)
    z = push(np.concatenate([x[free], g0[ineq]]))

    def cons(z, gx):
        c = gx - rhs
        c[ineq] -= z[nf:]
        return c

    def barrier(z, fx, mu):
        b = fx
        b -= mu * np.log(z[hasl] - zl[hasl]).sum()
        b -= mu * np.log(zh[hash_] - z[hash_]).sum()
        return b

    mu = mu0
    lam = np.zeros(m)
    vl = np.where(hasl, 1.0, 0.0)
    vh = np.where(hash_, 1.0, 0.0)
    B = np.eye(nz)
    B[nf:, nf:] = 0.0  # slacks enter linearly
    filt, filt_mu, th_max, th_min = [], None, np.inf, 0.0
    scaled = not scale_hessian
    t_start = time.perf_counter()
    xx = full(z)
    fx, gx = f_(xx), g_(xx)
    dfx, Jx = df_(xx), J_(xx)
    it = 0
    status = "maxiter"
    hist = []
    while it < maxiter:
        A = np.zeros((m, nz))
        A[:, :nf] = Jx[:, free]
        A[ineq, nf + np.arange(ni)] = -1.0
        grad = np.concatenate([dfx[free], np.zeros(ni)])
        c = cons(z, gx)
        dl = np.where(hasl, z - zl, 1.0)
        dh = np.where(hash_, zh - z, 1.0)
        rd = grad + A.T @ lam - vl + vh
        sd = max(100.0, (np.abs(lam).sum() + vl.sum() + vh.sum()) / max(1, m + 2 * nz)) / 100.0

        def err(mu_):
            cl = np.abs(np.where(hasl, dl * vl - mu_, 0.0)).max(initial=0.0)
            ch = np.abs(np.where(hash_, dh * vh - mu_, 0.0)).max(initial=0.0)
            return max(np.abs(rd).max() / sd, np.abs(c).max(initial=0.0), cl / sd, ch / sd)

        e0 = err(0.0)
        hist.append((it, fx / sc["f"], np.abs(c).max(initial=0.0), np.abs(rd).max(), mu))
        if verbose:
            print("%4d f=%.8e |c|=%.2e |rd|=%.2e mu=%.1e" % hist[-1], flush=True)
        if e0 <= tol:
            status = "converged"
            break
        while mu > tol / 10.0 and err(mu) <= 10.0 * mu:
            mu = max(tol / 10.0, min(0.2 * mu, mu ** 1.5))
        # Newton step on the barrier problem, primal-dual Hessian
        sig = np.where(hasl, vl / dl, 0.0) + np.where(hash_, vh / dh, 0.0)
        gb = grad - np.where(hasl, mu / dl, 0.0) + np.where(hash_, mu / dh, 0.0)
        delta = 0.0
        for _try in range(12):
            H = B + np.diag(sig + delta)
            K = np.zeros((nz + m, nz + m))
            K[:nz, :nz] = H
            K[:nz, nz:] = A.T
            K[nz:, :nz] = A
            K[nz:, nz:] = -1e-10 * np.eye(m)
            try:
                sol = np.linalg.solve(K, -np.concatenate([gb + A.T @ lam, c]))
            except np.linalg.LinAlgError:
                sol = None
            if sol is not None and np.all(np.isfinite(sol)):
                dz, dlam = sol[:nz], sol[nz:]
                if dz @ (H @ dz) > 1e-12 * (dz @ dz) or np.abs(dz).max() < 1e-14:  # curvature along the step: descent for the merit function
                    break
            delta = 1e-4 if delta == 0.0 else delta * 10.0
        dvl = np.where(hasl, (mu - vl * dz) / dl - vl, 0.0)
        dvh = np.where(hash_, (mu + vh * dz) / dh - vh, 0.0)
        tau = max(0.99, 1.0 - mu)
        neg, pos = (dz < 0) & hasl, (dz > 0) & hash_
        amax = min(1.0, (tau * dl[neg] / -dz[neg]).min(initial=1.0), (tau * dh[pos] / dz[pos]).min(initial=1.0))
        nl, nh = (dvl < 0) & hasl, (dvh < 0) & hash_
        adual = min(1.0, (tau * vl[nl] / -dvl[nl]).min(initial=1.0), (tau * vh[nh] / -dvh[nh]).min(initial=1.0))
        # filter line search (Waechter & Biegler 2006, Algorithm A without the restoration phase and second-order correction)
        th0 = np.abs(c).sum()
        phi0 = barrier(z, fx, mu)
        dphi = gb @ dz
        if mu != filt_mu:
            filt, filt_mu = [], mu
            th_max, th_min = 1e4 * max(1.0, th0), 1e-4 * max(1.0, th0)
        alpha = amax
        accepted = False
        for _ls in range(40):
            zt = z + alpha * dz
            xt = full(zt)
            ft, gt = f_(xt), g_(xt)
            if np.isfinite(ft) and np.all(np.isfinite(gt)):
                tht = np.abs(cons(zt, gt)).sum()
                phit = barrier(zt, ft, mu)
                ok = tht <= th_max and all(tht < (1 - 1e-5) * a or phit < b - 1e-5 * a for a, b in filt)
                if ok:
                    switching = dphi < 0 and alpha * (-dphi) ** 2.3 > th0 ** 1.1 and th0 <= th_min
                    if switching:
                        if phit <= phi0 + 1e-8 * alpha * dphi + 1e-13 * abs(phi0):
                            accepted = True
                            break
                    elif tht <= (1 - 1e-5) * th0 or phit <= phi0 - 1e-5 * th0 + 1e-13 * abs(phi0):
                        filt.append(((1 - 1e-5) * th0, phi0 - 1e-5 * th0))
                        accepted = True
                        break
            alpha *= 0.5
        if not accepted:
            # stand-in for IPOPT's restoration phase: the minimum-norm step towards feasibility of the linearised rows,
            # accepted on a decrease of the violation alone; its starting point enters the filter
            filt.append(((1 - 1e-5) * th0, phi0 - 1e-5 * th0))
            dz = -np.linalg.lstsq(A, c, rcond=None)[0]
            neg, pos = (dz < 0) & hasl, (dz > 0) & hash_
            alpha = min(1.0, (tau * dl[neg] / -dz[neg]).min(initial=1.0), (tau * dh[pos] / dz[pos]).min(initial=1.0))
            dlam = np.zeros(m)
            stats["restorations"] = stats.get("restorations", 0) + 1
            for _ls in range(40):
                zt = z + alpha * dz
                xt = full(zt)
                ft, gt = f_(xt), g_(xt)
                if np.isfinite(ft) and np.all(np.isfinite(gt)) and np.abs(cons(zt, gt)).sum() <= (1 - 1e-4 * alpha) * th0:
                    accepted = True
                    break
                alpha *= 0.5
        if not accepted:
            status = "line search failed"
            break
        # accept; duals, then the damped BFGS update of the Lagrangian Hessian
        lam_new = lam + alpha * dlam
        vl = vl + adual * dvl
        vh = vh + adual * dvh
        dl_n = np.where(hasl, zt - zl, 1.0)
        dh_n = np.where(hash_, zh - zt, 1.0)
        vl = np.where(hasl, np.clip(vl, mu / (1e10 * dl_n), 1e10 * mu / dl_n), 0.0)
        vh = np.where(hash_, np.clip(vh, mu / (1e10 * dh_n), 1e10 * mu / dh_n), 0.0)
        dft, Jt = df_(xt), J_(xt)
        s = (zt - z)[:nf]
        y = (dft[free] + Jt[:, free].T @ lam_new) - (dfx[free] + Jx[:, free].T @ lam_new)
        Bx = B[:nf, :nf]
        Bs = Bx @ s
        sBs = s @ Bs
        if sBs > 0 and s @ s > 0:
            sy = s @ y
            if not scaled and sy > 0:  # first pair: size the initial matrix (Nocedal & Wright (6.20)), as limited-memory updates do
                B[:nf, :nf] = np.eye(nf) * min(1e8, max(1e-8, (y @ y) / sy))
                Bx = B[:nf, :nf]
                Bs = Bx @ s
                sBs = s @ Bs
                scaled = True
            th = 1.0 if sy >= 0.2 * sBs else 0.8 * sBs / (sBs - sy)
            r = th * y + (1.0 - th) * Bs
            B[:nf, :nf] = Bx - np.outer(Bs, Bs) / sBs + np.outer(r, r) / (s @ r)
        z, lam, fx, gx, dfx, Jx = zt, lam_new, ft, gt, dft, Jt
        it += 1
    stats["t_total"] = time.perf_counter() - t_start
    stats["iterations"] = it
    stats["status"] = status
    stats["history"] = hist

    class R:
        pass

    res = R()
    fx = fx / sc["f"]
    res.x, res.fun, res.success, res.message, res.nit = full(z), fx, status == "converged", status, it
    res.constraint_violation = float(np.abs(cons(z, gx) / sc["g"]).max(initial=0.0))
    return res, stats
