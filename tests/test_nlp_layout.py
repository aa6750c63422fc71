"""Host logic (CPU): NLP shapes, variable/constraint ordering and the CSC pattern of jac_g handed to IPOPT
(CasOCTranscription.h:181-313), product layout vs the oracle's independent row-wise derivation."""
import numpy as np
import pytest

from conftest import FIXTURES, golden

# (n_var, n_con) of the BASELINE configs
APPENDIX_A = {
    "sliding_mass": (305, 250),
    "sliding_mass_trap": (62, 38),
    "hanging_muscle": (359, 328),
    "gait2d_tracking": (5759, 4788),
    "gait2d_prediction": (5759, 4789),
}


@pytest.mark.parametrize("name", sorted(APPENDIX_A))
def test_problem_shapes(pkg, name):
    d = pkg.Problem.load(golden(name + ".mb20")).dims()
    assert (d.n_var, d.n_con) == APPENDIX_A[name]


def test_config5_shape(pkg):
    d = pkg.Problem.load(golden("gait2d_prediction.mb20"), "num_mesh_intervals=800").dims()
    assert (d.n_grid, d.n_var, d.n_con) == (1601, 91259, 76039)
    assert d.n_in == 58 and d.n_dyn_out == 28


@pytest.mark.parametrize("name", FIXTURES)
def test_csc_structure_matches_oracle(pkg, orc, name):
    opts = "num_mesh_intervals=6" if name.startswith("gait2d") else None
    p = pkg.Problem.load(golden(name + ".mb20"), opts)
    colptr, rowidx = p.structure()
    o = orc.Oracle(p)
    ocol, orow = o.structure()
    assert np.array_equal(colptr, ocol)
    assert np.array_equal(rowidx, orow)
    # CSC invariants: rows strictly ascending within each column
    for c in range(len(colptr) - 1):
        r = rowidx[colptr[c]:colptr[c + 1]]
        assert np.all(np.diff(r) > 0)


@pytest.mark.parametrize("name", ["sliding_mass_trap", "hanging_muscle", "pendulum3_implicit"])
def test_oracle_jacobian_is_derivative_of_oracle_g(pkg, orc, name):
    """The chained FD Jacobian must be the derivative of g as a whole: J v ~ (g(x + e v) - g(x - e v)) / 2e."""
    p = pkg.Problem.load(golden(name + ".mb20"), "num_mesh_intervals=5")
    o = orc.Oracle(p)
    d = p.dims()
    x = p.make_iterate(seed=4, spread=0.2)
    colptr, rowidx = o.structure()
    vals = o.eval_jac_g(x)
    rng = np.random.default_rng(0)
    v = rng.uniform(-1, 1, d.n_var)
    Jv = np.zeros(d.n_con)
    for c in range(d.n_var):
        sl = slice(colptr[c], colptr[c + 1])
        Jv[rowidx[sl]] += vals[sl] * v[c]
    e = 1e-6
    fd = (o.eval_g(x + e * v) - o.eval_g(x - e * v)) / (2 * e)
    assert np.allclose(Jv, fd, rtol=2e-5, atol=2e-5 * max(1.0, np.abs(fd).max()))
    # gradient of the objective likewise
    g = o.eval_grad_f(x)
    fdf = (o.eval_f(x + e * v) - o.eval_f(x - e * v)) / (2 * e)
    assert g @ v == pytest.approx(fdf, rel=2e-5, abs=1e-6)


def test_bounds_follow_reference_rules(pkg):
    """Interior columns get `bounds`, first/last columns the (clipped) initial/final bounds
    (CasOCTranscription.cpp:186-220); defect/residual/interp rows are equalities."""
    p = pkg.Problem.load(golden("sliding_mass.mb20"))
    xl, xu, gl, gu = p.bounds()
    d = p.dims()
    assert (xl[0], xu[0], xl[1], xu[1]) == (0, 0, 0, 5)
    # position: bounds [-5,5], initial 0, final 1 ; speed: [-50,50], initial 0, final 0
    assert (xl[2], xu[2]) == (0, 0) and (xl[3], xu[3]) == (0, 0)
    assert (xl[4], xu[4]) == (-5, 5) and (xl[5], xu[5]) == (-50, 50)
    last = 2 + 2 * (d.n_grid - 1)
    assert (xl[last], xu[last]) == (1, 1) and (xl[last + 1], xu[last + 1]) == (0, 0)
    assert np.all(gl == 0) and np.all(gu == 0)


def test_make_iterate_is_deterministic_and_within_bounds(pkg):
    p = pkg.Problem.load(golden("gait2d_tracking.mb20"))
    xl, xu, _, _ = p.bounds()
    a, b = p.make_iterate(seed=7, spread=0.5), p.make_iterate(seed=7, spread=0.5)
    assert np.array_equal(a, b)
    assert np.all(a >= xl) and np.all(a <= xu)
    assert not np.array_equal(a, p.make_iterate(seed=8, spread=0.5))


# ---- path constraints: MocoControlBoundConstraint on the pendulum of testConstraints.cpp:1458-1541
CB_CASES = ["pendulum_cb_lower", "pendulum_cb_upper_trap", "pendulum_cb_equality", "pendulum_cb_both"]


def test_control_bound_rows_and_bounds(pkg, orc):
    """Section 'Lower bound only': one row per mesh point, bounds [0, inf), value = control - 0.1318
    (MocoControlBoundConstraint.cpp:93-119,121-148); rows sit at the head of each mesh point's block."""
    p = pkg.Problem.load(golden("pendulum_cb_lower.mb20"))
    d = p.dims()
    xl, xu, gl, gu = p.bounds()
    rows = np.flatnonzero(np.isinf(gu))
    assert len(rows) == d.n_mesh and np.all(gl[rows] == 0)
    # HS, ns = 2, nc = 1: per interval [path(1) | defects(4) | interp(1)] and the last mesh point's path row at the end
    assert np.array_equal(rows, np.arange(d.n_mesh) * 6)
    o = orc.Oracle(p)
    x = p.make_iterate(seed=2, spread=0.5)
    g = o.eval_g(x)
    ctl = x[2 + 2 * d.n_grid:][: d.n_grid]
    assert np.array_equal(g[rows], ctl[::2] - 0.1318)


def test_control_bound_equality_and_both(pkg, orc):
    """'Upper and lower bounds are the same' -> one equality row per mesh point following the piecewise-linear function;
    both bounds -> (lower, upper) row pairs with bounds [0, inf), (-inf, 0]."""
    p = pkg.Problem.load(golden("pendulum_cb_equality.mb20"))
    d = p.dims()
    xl, xu, gl, gu = p.bounds()
    assert np.all(gl == 0) and np.all(gu == 0) and d.n_con == 4 * 25 + 25 + 26
    o = orc.Oracle(p)
    x = p.make_iterate(seed=1, spread=0.3)
    g = o.eval_g(x)
    t = x[0] + (x[1] - x[0]) * np.linspace(0, 1, d.n_mesh)
    want = x[2 + 2 * d.n_grid:][: d.n_grid][::2] - np.interp(t, [0, 0.2, 0.7, 1], [0, 0.5316, -0.3137, .0319])
    assert np.allclose(g[np.arange(d.n_mesh) * 6], want, rtol=0, atol=1e-15)
    p = pkg.Problem.load(golden("pendulum_cb_both.mb20"))
    xl, xu, gl, gu = p.bounds()
    lo_rows, up_rows = np.flatnonzero(np.isinf(gu)), np.flatnonzero(np.isinf(gl))
    assert np.array_equal(up_rows, lo_rows + 1) and len(lo_rows) == p.dims().n_mesh
    x = p.make_iterate(seed=1, spread=0.3)
    g = orc.Oracle(p).eval_g(x)
    ctl = x[2 + 2 * p.dims().n_grid:][: p.dims().n_grid][::2]
    assert np.array_equal(g[up_rows], ctl - 0.75)


def test_control_bound_refusals(pkg):
    """Error behaviour of MocoControlBoundConstraint::initializeOnModelImpl (.cpp:37-92)."""
    assert pkg.Problem.example("pendulum_control_bound:none").dims().n_con == 125  # 'Can omit both bounds.'
    with pytest.raises(Exception):
        pkg.Problem.example("pendulum_control_bound:bogus")


@pytest.mark.parametrize("name", CB_CASES)
def test_control_bound_oracle_jacobian_is_derivative_of_g(pkg, orc, name):
    p = pkg.Problem.load(golden(name + ".mb20"), "num_mesh_intervals=7")  # no mesh point on a knot of the bound function
    o = orc.Oracle(p)
    d = p.dims()
    x = p.make_iterate(seed=4, spread=0.2)
    colptr, rowidx = o.structure()
    vals = o.eval_jac_g(x)
    v = np.random.default_rng(0).uniform(-1, 1, d.n_var)
    Jv = np.zeros(d.n_con)
    for c in range(d.n_var):
        sl = slice(colptr[c], colptr[c + 1])
        Jv[rowidx[sl]] += vals[sl] * v[c]
    e = 1e-6
    fd = (o.eval_g(x + e * v) - o.eval_g(x - e * v)) / (2 * e)
    assert np.allclose(Jv, fd, rtol=2e-5, atol=2e-5 * max(1.0, np.abs(fd).max()))
