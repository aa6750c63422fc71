"""Prescribed kinematics (Components/PositionMotion.cpp, MocoInverse.cpp:44-140; the prescribed-kinematics variant
of MultibodySystemImplicit, MocoCasOCProblem.h:247-297 with CasOCProblem.h:223-233,485-517).

Fixtures: gait2d_inverse.mb20 = MocoInverse on example2DWalking's model (kinematics of referenceCoordinates.sto prescribed,
reserves, implicit tendons on the calf / shin muscles), gait2d_inverse_free.mb20 = the same problem with the coordinates left
as states (scripts/make_problem_fixtures.py).  The key check is implementation-independent: the prescribed point function must
equal the ordinary implicit one evaluated at q(t), qdot(t), qddot(t), with the three evaluated HERE (numpy) from the tables."""
import ctypes as C

import numpy as np
import pytest

from conftest import golden


def pools(p):
    t = p.table_sizes()
    dbl = np.ctypeslib.as_array(C.cast(t.dbl, C.POINTER(C.c_double)), shape=(t.n_dbl,))
    idx = np.ctypeslib.as_array(C.cast(t.idx, C.POINTER(C.c_int)), shape=(t.n_int,))
    return t, dbl, idx


def ppoly(t, dbl, idx, pid, times, deriv):
    """value / derivative of piecewise polynomial `pid` at `times` (numpy.polynomial, not Horner on derivatives)"""
    off, npc, order = (int(v) for v in idx[t.o_ppoly_i + 3 * pid:t.o_ppoly_i + 3 * pid + 3])
    br = dbl[t.o_ppoly_data + off:t.o_ppoly_data + off + npc + 1]
    co = dbl[t.o_ppoly_data + off + npc + 1:t.o_ppoly_data + off + npc + 1 + npc * order].reshape(npc, order)
    out = np.zeros(len(times))
    for i, tt in enumerate(times):
        k = int(np.clip(np.searchsorted(br, tt, side="right") - 1, 0, npc - 1))
        pol = np.polynomial.Polynomial(co[k])
        out[i] = (pol.deriv(deriv) if deriv else pol)(tt - br[k])
    return out


def twin_iterates(pp, pf, seed):
    """An iterate of the prescribed problem and the iterate of the free problem that holds the same auxiliary states, controls
    and auxiliary derivatives plus q, u, udot from the prescribed functions at the grid times."""
    dp, df = pp.dims(), pf.dims()
    assert dp.n_grid == df.n_grid and dp.nc == df.nc
    t, dbl, idx = pools(pp)
    nq = t.nq
    assert t.prescribed == 1 and df.ns == dp.ns + 2 * nq and df.nd == dp.nd + nq
    xp = pp.make_iterate(seed=seed, spread=0.3)
    N = dp.n_grid
    hs = dp.n_grid == 2 * dp.n_mesh - 1
    grid = np.linspace(0, 1, N) if True else None  # uniform mesh: mesh points and midpoints are equally spaced
    times = xp[0] + (xp[1] - xp[0]) * grid
    xf = np.zeros(df.n_var)
    xf[:2] = xp[:2]
    sp = xp[2:2 + N * dp.ns].reshape(N, dp.ns)
    cp = xp[2 + N * dp.ns:2 + N * (dp.ns + dp.nc)].reshape(N, dp.nc)
    ddp = xp[2 + N * (dp.ns + dp.nc):].reshape(N, dp.nd)
    sf = np.zeros((N, df.ns))
    ddf = np.zeros((N, df.nd))
    for k in range(nq):
        pid = int(idx[t.o_presc_i + k])
        sf[:, k] = ppoly(t, dbl, idx, pid, times, 0)
        sf[:, nq + k] = ppoly(t, dbl, idx, pid, times, 1)
        ddf[:, k] = ppoly(t, dbl, idx, pid, times, 2)
    sf[:, 2 * nq:] = sp
    ddf[:, nq:] = ddp
    xf[2:] = np.concatenate([sf.ravel(), cp.ravel(), ddf.ravel()])
    assert hs or dp.n_grid == dp.n_mesh
    return xp, xf


def test_layout_follows_the_reference_counting_rules(pkg):
    """MocoProblemRep.cpp:546-560: no /value and /speed states; CasOCProblem.h:485-517: no acceleration variables, nq multibody
    residual rows per grid point, no kinematic-constraint rows."""
    p = pkg.Problem.load(golden("gait2d_inverse.mb20"))
    f = pkg.Problem.load(golden("gait2d_inverse_free.mb20"))
    d, e = p.dims(), f.dims()
    t = p.table_sizes()
    nq, nz, naux = t.nq, t.nz, t.nderiv_aux
    assert (nq, nz, naux) == (10, 24, 6)          # 18 activations + 6 normalized tendon forces; 6 implicit tendons
    assert d.ns == nz and d.nd == naux and d.n_in == 1 + nz + d.nc + naux
    assert e.ns == 2 * nq + nz and e.nd == nq + naux
    assert d.n_out == e.n_out == nq + nz + naux + 1  # residuals, zdot, auxiliary residuals, the effort integrand
    N = d.n_grid
    assert d.n_var == 2 + N * (d.ns + d.nc + d.nd)
    # 18 initial-activation rows, (nq + naux) residual rows per grid point, two defects per state and interval
    assert d.n_con == 18 + N * (nq + naux) + (d.n_mesh - 1) * 2 * d.ns
    names = p.names(0) if hasattr(p, "names") else None
    if names:
        assert not any(n.endswith("/value") or n.endswith("/speed") for n in names)


def test_prescribed_requires_implicit_mode(pkg):
    with pytest.raises(Exception, match="implicit"):
        pkg.Problem.load(golden("gait2d_inverse.mb20"), "multibody_dynamics_mode=explicit")


@pytest.mark.parametrize("seed", [1, 2])
def test_oracle_prescribed_point_function_equals_the_implicit_one_at_the_prescribed_values(pkg, orc, seed):
    pp, pf = pkg.Problem.load(golden("gait2d_inverse.mb20")), pkg.Problem.load(golden("gait2d_inverse_free.mb20"))
    xp, xf = twin_iterates(pp, pf, seed)
    Yp, Yf = orc.Oracle(pp).eval_points(xp), orc.Oracle(pf).eval_points(xf)
    assert np.isfinite(Yf).all()
    scale = np.abs(Yf) + 1e-6 * np.abs(Yf).max()
    assert np.max(np.abs(Yp - Yf) / scale) < 1e-9  # spline derivatives evaluated two ways


def test_oracle_jacobian_is_the_derivative_of_g(pkg, orc):
    p = pkg.Problem.load(golden("gait2d_inverse_trap_central.mb20"), "num_mesh_intervals=4")
    o = orc.Oracle(p)
    o.set_fd(1, 1e-5, 0)
    x = p.make_iterate(seed=3, spread=0.3)
    colptr, rowidx = o.structure()
    vals = o.eval_jac_g(x)
    rng = np.random.default_rng(0)
    cols = [0, 1] + rng.choice(np.arange(2, o.n_var), 40, replace=False).tolist()
    g0 = o.eval_g(x)
    for c in cols:
        h = 1e-5
        xa, xb = x.copy(), x.copy()
        xa[c] += h; xb[c] -= h
        dg = (o.eval_g(xa) - o.eval_g(xb)) / (2 * h)
        dense = np.zeros(o.n_con)
        dense[rowidx[colptr[c]:colptr[c + 1]]] = vals[colptr[c]:colptr[c + 1]]
        assert np.max(np.abs(dense - dg)) < 2e-5 * max(1.0, np.abs(dg).max(), np.abs(g0).max() * 1e-3), c


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["warp", "thread"])
def test_gpu_prescribed_point_function_equals_the_implicit_one_at_the_prescribed_values(pkg, kernel, monkeypatch):
    monkeypatch.setenv("MOCOB200_KERNEL", kernel)
    pp, pf = pkg.Problem.load(golden("gait2d_inverse.mb20")), pkg.Problem.load(golden("gait2d_inverse_free.mb20"))
    xp, xf = twin_iterates(pp, pf, 4)
    Yp, Yf = pkg.Evaluator(pp).eval_points(xp), pkg.Evaluator(pf).eval_points(xf)
    scale = np.abs(Yf) + 1e-6 * np.abs(Yf).max()
    assert np.max(np.abs(Yp - Yf) / scale) < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["warp", "thread"])
@pytest.mark.parametrize("name,fd,h", [("gait2d_inverse", 0, 1e-4), ("gait2d_inverse_trap_central", 1, 1e-3)])
def test_gpu_jacobian_matches_oracle_with_prescribed_kinematics(pkg, orc, name, fd, h, kernel, monkeypatch):
    monkeypatch.setenv("MOCOB200_KERNEL", kernel)
    p = pkg.Problem.load(golden(name + ".mb20"))
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    ev.set_fd(fd, h, 0)
    o.set_fd(fd, h, 0)
    x = p.make_iterate(seed=7, spread=0.3)
    assert np.max(np.abs(ev.eval_g(x) - o.eval_g(x))) < 1e-10 * max(1.0, np.abs(o.eval_g(x)).max())
    v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
    _, rowidx = o.structure()
    rowmax = np.zeros(o.n_con)
    np.maximum.at(rowmax, rowidx, np.abs(vo))
    assert np.max(np.abs(v - vo) / (np.abs(vo) + rowmax[rowidx] + 1e-300)) < 5e-7
    gr, gro = ev.eval_grad_f(x), o.eval_grad_f(x)
    assert np.max(np.abs(gr - gro)) <= 1e-7 * max(1.0, np.abs(gro).max(), abs(o.eval_f(x)))
