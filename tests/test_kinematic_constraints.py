"""Kinematic constraints: CoordinateCouplerConstraints with Lagrange multipliers, velocity-correction slacks and
the constraint-error rows, as MocoCasADiSolver transcribes them (MocoCasOCProblem.h:298-332, 643-732; CasOCTranscription.cpp:
83-250, 253-446, 513-521; CasOCTranscription.h:219-313). Layout facts are checked against the counts the reference's own
rules give; the physics against closed forms; the solve against the reference's known answer (testConstraints.cpp:1371-1416)."""
import numpy as np
import pytest

from conftest import golden
from nlp_solve import solve


def load(pkg, name, opts=None):
    return pkg.Problem.load(golden(name + ".mb20"), opts)


def blocks(p, x):
    d = p.dims()
    N, ni = d.n_grid, d.n_grid - d.n_mesh
    o = 2
    out = {}
    for key, w, n in (("states", d.ns, N), ("controls", d.nc, N), ("multipliers", d.nm, N), ("slacks", d.nsl, ni), ("derivatives", d.nd, N)):
        out[key] = x[o:o + w * n].reshape(n, w)
        o += w * n
    assert o == d.n_var
    return out


def test_layout_follows_the_reference_counting_rules(pkg):
    # Hermite-Simpson, 10 intervals, derivatives enforced: 1 multiplier per grid point, 1 slack per midpoint, 3 rows per mesh point
    d = load(pkg, "point_mass_coupler").dims()
    assert (d.nm, d.nsl, d.n_kc) == (1, 1, 3)
    assert d.n_var == 2 + (4 + 2 + 1) * 21 + 1 * 10
    assert d.n_con == 3 * 11 + 2 * 4 * 10 + 2 * 10
    assert d.n_in == 1 + 4 + 2 + 1 and d.n_out == 2 + 0 + 0 + 3 + 1
    # implicit mode adds the accelerations (derivatives) and their residual rows
    di = load(pkg, "point_mass_coupler_implicit").dims()
    assert di.nd == 2 and di.n_var == d.n_var + 2 * 21 and di.n_con == d.n_con + 2 * 21
    # trapezoidal without derivatives: no slacks, one row per mesh point (CasOCProblem.h:473-524)
    dt = load(pkg, "double_pendulum_coupler_noderiv_trap").dims()
    assert (dt.nm, dt.nsl, dt.n_kc) == (1, 0, 1) and dt.n_var == 2 + (4 + 2 + 1) * 9 and dt.n_con == 1 * 9 + 4 * 8
    # minimize_lagrange_multipliers adds an objective term (CasOCTranscription.cpp:513-521)
    assert load(pkg, "double_pendulum_coupler").dims().n_cost_terms == 2


def test_bounds_of_multipliers_and_slacks(pkg):
    p = load(pkg, "point_mass_coupler")
    xl, xh, gl, gh = p.bounds()
    b_lo, b_hi = blocks(p, xl), blocks(p, xh)
    assert np.all(b_lo["multipliers"] == -1000) and np.all(b_hi["multipliers"] == 1000)      # MocoPhase multiplier_bounds default
    assert np.all(b_lo["slacks"] == -0.1) and np.all(b_hi["slacks"] == 0.1)                  # velocity_correction_bounds default
    assert np.all(gl == 0) and np.all(gh == 0)


def test_constraint_rows_come_first_at_every_mesh_point(pkg, orc):
    """flattenConstraints (CasOCTranscription.h:279-283): kinematic rows of a mesh point sit in front of its other rows."""
    p = load(pkg, "point_mass_coupler")
    o = orc.Oracle(p)
    x = p.make_iterate(seed=3, spread=0.5)
    g, Y = o.eval_g(x), o.eval_points(x)
    d = p.dims()
    per_interval = d.n_kc + 2 * d.ns + d.nc
    for im in range(d.n_mesh):
        row = im * per_interval
        assert np.array_equal(g[row:row + d.n_kc], Y[2 * im, 2:2 + d.n_kc])


def test_point_function_with_a_coupler_obeys_newton(pkg, orc):
    """Planar point mass, ty = tx (perr = tx - ty, G = [1, -1]): m ax = Fx - lambda, m ay = Fy + lambda; errors
    perr = tx - ty, verr = vx - vy, aerr = ax - ay (MocoCasOCProblem.h:643-732)."""
    p = load(pkg, "point_mass_coupler")
    o = orc.Oracle(p)
    rng = np.random.default_rng(0)
    for _ in range(5):
        t, tx, ty, vx, vy, fx, fy, lam = rng.normal(size=8)
        out = o.eval_point(np.array([t, tx, ty, vx, vy, fx, fy, lam]))
        ax, ay = fx - lam, fy + lam
        assert np.allclose(out[:2], [ax, ay], atol=1e-13)
        assert np.allclose(out[2:5], [tx - ty, vx - vy, ax - ay], atol=1e-13)


def test_velocity_correction_enters_the_simpson_defects(pkg, orc):
    """qdot_mid = u_mid + G^T gamma (MocoCasOCProblem.h:298-332): moving one slack by d changes the Simpson defect of tx by
    -4h/6 d and that of ty by +4h/6 d, and nothing else."""
    p = load(pkg, "point_mass_coupler")
    o = orc.Oracle(p)
    d = p.dims()
    x = p.make_iterate(seed=5, spread=0.5)
    g0 = o.eval_g(x)
    off_slacks = 2 + (d.ns + d.nc + d.nm) * d.n_grid
    m, delta = 4, 0.01
    x2 = x.copy()
    x2[off_slacks + m] += delta
    dg = o.eval_g(x2) - g0
    h = (x[1] - x[0]) / (d.n_mesh - 1)
    per_interval = d.n_kc + 2 * d.ns + d.nc
    simpson = m * per_interval + d.n_kc + d.ns
    expect = np.zeros_like(dg)
    expect[simpson + 0] = -4.0 * h / 6.0 * delta
    expect[simpson + 1] = +4.0 * h / 6.0 * delta
    assert np.allclose(dg, expect, atol=1e-14)


def test_oracle_jacobian_is_the_derivative_of_g_with_constraints(pkg, orc):
    for name in ("point_mass_coupler", "double_pendulum_coupler", "point_mass_coupler_implicit", "double_pendulum_coupler_noderiv_trap"):
        p = load(pkg, name)
        o = orc.Oracle(p)
        o.set_fd(1, 1e-4, 0)
        x = p.make_iterate(seed=7, spread=0.3)
        colptr, rowidx = o.structure()
        v = o.eval_jac_g(x)
        from scipy.sparse import csc_matrix

        Jd = csc_matrix((v, rowidx, colptr), shape=(o.n_con, o.n_var)).toarray()
        g0 = o.eval_g(x)
        rng = np.random.default_rng(1)
        for _ in range(3):
            dx = rng.normal(size=o.n_var)
            eps = 1e-6
            num = (o.eval_g(x + eps * dx) - o.eval_g(x - eps * dx)) / (2 * eps)
            assert np.allclose(Jd @ dx, num, rtol=1e-5, atol=1e-6 * (1 + np.abs(num).max())), name


def test_oracle_solve_reproduces_the_reference_multiplier(pkg, orc):
    """testConstraints.cpp:1371-1416 'Planar point mass with CoordinateCouplerConstraint': lambda = 0.5 (Fx - Fy) along the
    solution (the test's tolerance: 1e-5)."""
    p = load(pkg, "point_mass_coupler")
    nthreads = orc.max_threads()
    orc.set_threads(1)
    try:
        res, stats = solve(p, orc.Oracle(p), maxiter=300)
    finally:
        orc.set_threads(nthreads)
    assert res.success, res.message
    b = blocks(p, res.x)
    lam, fx, fy = b["multipliers"][:, 0], b["controls"][:, 0], b["controls"][:, 1]
    assert np.abs(lam - 0.5 * (fx - fy)).max() < 1e-5
    assert np.abs(b["states"][:, 0] - b["states"][:, 1]).max() < 1e-6   # ty = tx at every grid point the constraint rows cover
    assert abs(b["states"][-1, 0] - 3.0) < 1e-9 and np.allclose(fx, 0.5)


@pytest.mark.gpu
def test_gpu_solve_with_constraints_converges_like_the_cpu_solve(pkg, orc):
    p = load(pkg, "point_mass_coupler")
    nthreads = orc.max_threads()
    orc.set_threads(1)
    try:
        res_o, st_o = solve(p, orc.Oracle(p), maxiter=300)
    finally:
        orc.set_threads(nthreads)
    res_g, st_g = solve(p, pkg.Evaluator(p), maxiter=300)
    assert res_g.success, res_g.message
    assert abs(res_g.fun - res_o.fun) < 1e-6 * max(1.0, abs(res_o.fun))
    assert abs(st_g["iterations"] - st_o["iterations"]) <= max(1, round(0.05 * st_o["iterations"]))
    b = blocks(p, res_g.x)
    assert np.abs(b["multipliers"][:, 0] - 0.5 * (b["controls"][:, 0] - b["controls"][:, 1])).max() < 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["point_mass_coupler", "point_mass_coupler_implicit", "double_pendulum_coupler", "double_pendulum_coupler_noderiv_trap"])
@pytest.mark.parametrize("kernel", ["warp", "thread"])
def test_gpu_constraint_rows_and_columns_match_the_oracle(pkg, orc, name, kernel, monkeypatch):
    monkeypatch.setenv("MOCOB200_KERNEL", kernel)
    p = load(pkg, name)
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    x = p.make_iterate(seed=11, spread=0.6)
    assert np.allclose(ev.eval_g(x), o.eval_g(x), rtol=1e-10, atol=1e-12)
    f, terms = ev.eval_f(x, terms=True)
    fo, termso = o.eval_f(x, terms=True)
    assert abs(f - fo) <= 1e-10 * max(1.0, abs(fo)) and np.allclose(terms, termso, rtol=1e-10, atol=1e-12)
    for fd, h in ((0, 1e-4), (1, 1e-3)):
        ev.set_fd(fd, h, 0)
        o.set_fd(fd, h, 0)
        v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
        assert np.array_equal(ev.structure()[1], o.structure()[1])
        assert np.abs(v - vo).max() < 1e-7 * (1.0 + np.abs(vo).max())
        gr, gro = ev.eval_grad_f(x), o.eval_grad_f(x)
        assert np.abs(gr - gro).max() < 1e-7 * (1.0 + np.abs(gro).max())
