"""Parity tests proper (need a B200): the CUDA path, called through the C ABI, against the CPU oracle on the same seeded
inputs, plus size-independent properties at BASELINE.json's full sizes.

Tolerances (also in DESIGN.md):
  * function values (point outputs, g, f): 1e-10 relative (north_star) — measured ~1e-12.
  * FD Jacobians: two implementations that agree to ~1e-13 in the function produce finite differences that differ by
    (that round-off)/h. With CasADi's default steps (h = 1.5e-8 forward) this noise is ~1e-4 of the row scale for ANY
    implementation that is not bit-identical, so the 1e-7 criterion is checked with round-off-negligible steps
    (forward h = 1e-4, central h = 1e-3: identical truncation error on both sides), and the default steps get a
    noise-model bound.
"""
import numpy as np
import pytest

from conftest import FIXTURES, golden

pytestmark = pytest.mark.gpu

FD_FORWARD, FD_CENTRAL, FD_BACKWARD = 0, 1, 2


def relerr(a, b):
    """max |a-b| / (|b| + 1e-6 max|b|): relative, with a floor so that structural zeros do not divide by zero."""
    return float(np.max(np.abs(a - b) / (np.abs(b) + 1e-6 * np.abs(b).max() + 1e-300)))


def rowscaled_err(v, vo, rowidx, n_con):
    """max |v-vo| / (|vo| + largest |entry| of the same constraint row): FD error scales with the row's output."""
    rowmax = np.zeros(n_con)
    np.maximum.at(rowmax, rowidx, np.abs(vo))
    return float(np.max(np.abs(v - vo) / (np.abs(vo) + rowmax[rowidx] + 1e-300)))


def load(pkg, name, intervals=None):
    opts = None
    if intervals is not None:
        opts = "num_mesh_intervals=%d" % intervals
    elif name.startswith("gait2d"):
        opts = "num_mesh_intervals=10"
    return pkg.Problem.load(golden(name + ".mb20"), opts)


@pytest.mark.parametrize("name", FIXTURES)
@pytest.mark.parametrize("seed,spread", [(1, 0.1), (2, 0.5), (3, 1.0)])
def test_values_match_oracle(pkg, orc, name, seed, spread):
    p = load(pkg, name)
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    x = p.make_iterate(seed=seed, spread=spread)
    Y, Yo = ev.eval_points(x), o.eval_points(x)
    assert np.isfinite(Yo).all()
    assert relerr(Y, Yo) < 1e-10
    assert relerr(ev.eval_g(x), o.eval_g(x)) < 1e-10
    f, terms = ev.eval_f(x, terms=True)
    fo, termso = o.eval_f(x, terms=True)
    assert abs(f - fo) <= 1e-10 * max(1.0, abs(fo))
    assert np.allclose(terms, termso, rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize("name", FIXTURES)
def test_structure_matches_oracle(pkg, orc, name):
    p = load(pkg, name)
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    c, r = ev.structure()
    co, ro = o.structure()
    assert np.array_equal(c, co) and np.array_equal(r, ro)


@pytest.mark.parametrize("name", FIXTURES)
@pytest.mark.parametrize("scheme,h", [(FD_FORWARD, 1e-4), (FD_BACKWARD, 1e-4), (FD_CENTRAL, 1e-3)])
def test_fd_jacobian_pipeline_1e7(pkg, orc, name, scheme, h):
    """north_star: Jacobian entries within 1e-7 relative — with steps where round-off/h is below that bar."""
    p = load(pkg, name)
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    ev.set_fd(scheme, h, 0)
    o.set_fd(scheme, h, 0)
    x = p.make_iterate(seed=5, spread=0.3)
    J = ev.eval_points(x, jacobian=True)
    Yo, Jo = o.eval_point_jacobians(x)
    assert np.isfinite(Jo).all(), "oracle FD block not finite at %s" % np.argwhere(~np.isfinite(Jo))[:8].tolist()
    assert np.isfinite(J).all(), "device FD block not finite at %s" % np.argwhere(~np.isfinite(J))[:8].tolist()
    # floor of the scale: an output that is small only because large terms cancel (a multibody residual of 0.03 made of
    # contact and gravity terms of 1e4) still carries eps * |terms| / h of difference noise in the whole-function quotient,
    # where the factored local columns are exactly zero; 1e-4 of the largest output of the problem bounds that noise
    scale = np.abs(Yo)[:, :, None] + np.abs(Jo) + 1e-4 * np.abs(Yo).max()
    assert np.max(np.abs(J - Jo) / scale) < 1e-7
    v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
    # CSC entries: FD error scales with the magnitude of the row's callback output, so compare against the row scale
    _, rowidx = o.structure()
    assert rowscaled_err(v, vo, rowidx, o.n_con) < 5e-7
    gr, gro = ev.eval_grad_f(x), o.eval_grad_f(x)
    # FD error of the integrand blocks scales with the integrand values themselves (~ the objective)
    gerr = np.max(np.abs(gr - gro)) / max(1.0, np.abs(gro).max(), abs(o.eval_f(x)))
    # the endpoint functions are differenced w.r.t. their `integral` input with an ABSOLUTE step (CasADi does not scale
    # h): for a huge integral I the difference (I+h)-(I-h) itself carries ulp(I)/h relative round-off on either side
    imax = (x[1] - x[0]) * np.abs(Yo[:, o.n_dyn_out:]).mean(axis=0).max() if o.n_integrand else 0.0
    assert gerr <= 1e-7 + 4 * 2.2e-16 * imax / h, gerr


@pytest.mark.parametrize("name", FIXTURES)
@pytest.mark.parametrize("scheme", [FD_FORWARD, FD_CENTRAL])
def test_fd_jacobian_default_steps_noise_bound(pkg, orc, name, scheme):
    """CasADi's default steps (forward sqrt(eps); central cbrt(eps) + one refinement): agreement up to the FD noise
    floor eps*cond/h of two non-bit-identical implementations."""
    p = load(pkg, name)
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    ev.set_fd(scheme)
    o.set_fd(scheme)
    x = p.make_iterate(seed=6, spread=0.3)
    J = ev.eval_points(x, jacobian=True)
    Yo, Jo = o.eval_point_jacobians(x)
    scale = np.abs(Yo)[:, :, None] + np.abs(Jo) + 1e-4 * np.abs(Yo).max()  # floor: see test_fd_jacobian_pipeline_1e7
    assert np.max(np.abs(J - Jo) / scale) < 2e-3
    # and the median entry is far better than the worst one
    assert np.median(np.abs(J - Jo) / scale) < 1e-6


@pytest.mark.parametrize("name", ["gait2d_tracking", "gait2d_prediction", "hanging_muscle"])
def test_no_read_of_uninitialised_scratch(pkg, orc, name, monkeypatch):
    """MOCOB200_POISON_SCRATCH=1 starts every warp evaluation from NaN-filled shared memory: a quantity read before it
    is written (e.g. the position of an inactive ConditionalPathPoint multiplied by its zero force) shows up as NaN."""
    monkeypatch.setenv("MOCOB200_POISON_SCRATCH", "1")
    monkeypatch.setenv("MOCOB200_KERNEL", "warp")
    p = load(pkg, name)
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    for seed, spread in ((1, 0.2), (2, 1.0)):
        x = p.make_iterate(seed=seed, spread=spread)
        Yo = o.eval_points(x)
        assert relerr(ev.eval_points(x), Yo) < 1e-10
        for scheme in (FD_FORWARD, FD_CENTRAL):
            ev.set_fd(scheme)
            o.set_fd(scheme)
            J = ev.eval_points(x, jacobian=True)
            _, Jo = o.eval_point_jacobians(x)
            assert np.array_equal(np.isfinite(J), np.isfinite(Jo))


@pytest.mark.parametrize("name", FIXTURES)
@pytest.mark.parametrize("scheme", [FD_FORWARD, FD_BACKWARD])
def test_local_columns_equal_whole_function_differences(pkg, name, scheme, monkeypatch):
    """Inputs that act through one muscle / one coordinate actuator: the factored difference written by the nominal
    warp (point_warp_local.cuh) against differencing the whole point function (MOCOB200_FD_LOCAL=0), same step."""
    monkeypatch.setenv("MOCOB200_KERNEL", "warp")
    p = load(pkg, name)
    x = p.make_iterate(seed=9, spread=0.4)
    ev = pkg.Evaluator(p)
    ev.set_fd(scheme, 1e-4, 0)
    J = ev.eval_points(x, jacobian=True)
    Y = ev.eval_points(x)
    monkeypatch.setenv("MOCOB200_FD_LOCAL", "0")
    ev0 = pkg.Evaluator(p)
    ev0.set_fd(scheme, 1e-4, 0)
    J0 = ev0.eval_points(x, jacobian=True)
    scale = np.abs(Y)[:, :, None] + np.abs(J0) + 1e-4 * np.abs(Y).max()  # floor: see test_fd_jacobian_pipeline_1e7
    # the whole-function difference carries eps*|terms|/h of cancellation noise (2e-12 |terms| here, and the terms of a gait
    # residual - contact, gravity - reach 1e4 times the scale floor); the factored one does not
    assert np.max(np.abs(J - J0) / scale) < 5e-8


@pytest.mark.parametrize("name", FIXTURES)
@pytest.mark.parametrize("kernel", ["warp", "thread"])
def test_local_columns_central_scheme(pkg, name, kernel, monkeypatch):
    """Central differences: factored local columns (with CasADi's step refinement applied to the factored values) against
    whole-function central differences, fixed step (no refinement) so that both use the same h."""
    monkeypatch.setenv("MOCOB200_KERNEL", kernel)
    p = load(pkg, name)
    if kernel == "thread" and p.table_sizes().ncustom > 0:
        pytest.skip("function-based joints run on the warp kernels only (the library refuses the forced thread mode)")
    x = p.make_iterate(seed=10, spread=0.4)
    ev = pkg.Evaluator(p)
    ev.set_fd(FD_CENTRAL, 1e-3, 0)
    J = ev.eval_points(x, jacobian=True)
    Y = ev.eval_points(x)
    monkeypatch.setenv("MOCOB200_FD_LOCAL", "0")
    ev0 = pkg.Evaluator(p)
    ev0.set_fd(FD_CENTRAL, 1e-3, 0)
    J0 = ev0.eval_points(x, jacobian=True)
    scale = np.abs(Y)[:, :, None] + np.abs(J0) + 1e-4 * np.abs(Y).max()  # floor: see test_fd_jacobian_pipeline_1e7
    assert np.max(np.abs(J - J0) / scale) < 1e-8
    # default refinement: the refined steps differ (they are driven by round-off in the linear rows), the values agree
    ev.set_fd(FD_CENTRAL)
    ev0.set_fd(FD_CENTRAL)
    Jr, Jr0 = ev.eval_points(x, jacobian=True), ev0.eval_points(x, jacobian=True)
    # (prescribed kinematics: the time column is the derivative of everything along the motion, contact forces included, and
    # its value moves with the refined step far more than the other columns do)
    assert np.max(np.abs(Jr - Jr0) / scale) < (2e-4 if p.table_sizes().prescribed else 1e-5)


@pytest.mark.parametrize("name", ["gait2d_prediction", "hanging_muscle", "pendulum3_implicit", "pendulum_cb_both"])
def test_paired_central_differences_agree_with_the_one_thread_variant(pkg, name, monkeypatch):
    """The thread kernel puts the two sides of a central difference on adjacent lanes (k_points_fd_flat_pair): the same
    formula as one thread doing both sides. The two are separately compiled kernels (the compiler contracts multiply-adds
    differently), so with a fixed step they agree to round-off / h, and with CasADi's step refinement - whose refined step is
    itself driven by round-off in the near-linear rows - to the noise bound of the refinement test above. An odd number of
    (point, direction) items exercises the shadow lanes of the last warp."""
    monkeypatch.setenv("MOCOB200_KERNEL", "thread")
    monkeypatch.setenv("MOCOB200_GRAPH", "0")
    p = load(pkg, name, 3)
    x = p.make_iterate(seed=12, spread=0.5)
    out = {}
    for pair in ("1", "0"):
        monkeypatch.setenv("MOCOB200_FLAT_PAIR", pair)
        ev = pkg.Evaluator(p)
        ev.set_fd(FD_CENTRAL)
        a = ev.eval_points(x, jacobian=True)
        ev.set_fd(FD_CENTRAL, 1e-3, 0)
        out[pair] = (a, ev.eval_points(x, jacobian=True), ev.eval_jac_g(x), ev.eval_points(x))
    Y = out["1"][3]
    scale = np.abs(Y)[:, :, None] + np.abs(out["0"][1]) + 1e-4 * np.abs(Y).max()  # floor: see test_fd_jacobian_pipeline_1e7
    assert np.array_equal(np.isnan(out["1"][0]), np.isnan(out["0"][0]))
    assert np.nanmax(np.abs(out["1"][1] - out["0"][1]) / scale) < 1e-9   # fixed step
    assert np.nanmax(np.abs(out["1"][0] - out["0"][0]) / scale) < 1e-5   # refined step
    assert np.allclose(out["1"][2], out["0"][2], rtol=1e-7, atol=1e-9 * np.abs(out["0"][2]).max(), equal_nan=True)


@pytest.mark.parametrize("name", ["gait2d_tracking", "hanging_muscle", "pendulum_cb_both"])
def test_same_x_reuses_the_last_evaluation(pkg, name):
    """IPOPT's call pattern: f, g, grad_f, jac_g at one x. The point outputs and endpoint values of the last evaluation are
    reused while x is unchanged (fewer launches, identical results); a different x recomputes."""
    p = load(pkg, name)
    x, x2 = p.make_iterate(seed=21, spread=0.4), p.make_iterate(seed=22, spread=0.4)
    fresh = pkg.Evaluator(p)
    want = dict(f=fresh.eval_f(x), g=pkg.Evaluator(p).eval_g(x), grad=pkg.Evaluator(p).eval_grad_f(x), jac=pkg.Evaluator(p).eval_jac_g(x))
    ev = pkg.Evaluator(p)
    n0 = ev.launch_count()
    g = ev.eval_g(x)
    n1 = ev.launch_count()
    f = ev.eval_f(x)
    g_again = ev.eval_g(x)
    n2 = ev.launch_count()
    assert n2 - n1 < n1 - n0, "same x: only the assembly kernels run"
    assert f == want["f"] and np.array_equal(g, want["g"]) and np.array_equal(g_again, want["g"])
    jac = ev.eval_jac_g(x)
    n3 = ev.launch_count()
    grad, f3, g3 = ev.eval_grad_f(x), ev.eval_f(x), ev.eval_g(x)
    assert ev.launch_count() - n3 <= 5
    assert np.array_equal(jac, want["jac"], equal_nan=True) and np.array_equal(grad, want["grad"], equal_nan=True)
    assert f3 == want["f"] and np.array_equal(g3, want["g"])
    # a new x invalidates everything
    assert np.array_equal(ev.eval_g(x2), pkg.Evaluator(p).eval_g(x2))
    assert np.array_equal(ev.eval_jac_g(x2), pkg.Evaluator(p).eval_jac_g(x2), equal_nan=True)
    assert ev.eval_f(x) == want["f"]


@pytest.mark.parametrize("name", ["gait2d_tracking", "hanging_muscle", "pendulum3_implicit"])
def test_graph_replay_equals_kernel_by_kernel_launches(pkg, name, monkeypatch):
    """The captured CUDA graph of the Jacobian pipeline, replayed on changing iterates and after a change of the FD
    scheme, gives bit-identical values to launching the kernels one by one (MOCOB200_GRAPH=0)."""
    p = load(pkg, name)
    ev = pkg.Evaluator(p)
    monkeypatch.setenv("MOCOB200_GRAPH", "0")
    ev0 = pkg.Evaluator(p)
    for scheme in (FD_FORWARD, FD_CENTRAL, FD_FORWARD):
        ev.set_fd(scheme)
        ev0.set_fd(scheme)
        for seed in (21, 22, 23):
            x = p.make_iterate(seed=seed, spread=0.3)
            v, v0 = ev.eval_jac_g(x), ev0.eval_jac_g(x)
            assert np.array_equal(v, v0, equal_nan=True)
            assert np.array_equal(ev.eval_grad_f(x), ev0.eval_grad_f(x), equal_nan=True)


@pytest.mark.parametrize("scheme", [FD_FORWARD, FD_BACKWARD, FD_CENTRAL])
def test_control_bound_path_constraint_rows(pkg, orc, scheme):
    """SURVEY §8 a10 — MocoControlBoundConstraint (MocoControlBoundConstraint.cpp:121-148) through the C ABI: g rows are
    control - bound(t) at the mesh points; their Jacobian rows are the callback's dense FD block (1 in the control's
    column up to FD round-off, exact zeros elsewhere) and agree with the oracle bit for bit (same IEEE operations)."""
    p = load(pkg, "pendulum_cb_lower")
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    ev.set_fd(scheme)
    o.set_fd(scheme)
    d = ev.dims
    x = p.make_iterate(seed=8, spread=0.7)
    rows = np.arange(d.n_mesh) * 6
    ctl = x[2 + 2 * d.n_grid:][: d.n_grid][::2]
    g = ev.eval_g(x)
    assert np.array_equal(g[rows], ctl - 0.1318)
    colptr, rowidx = ev.structure()
    v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
    cols = np.repeat(np.arange(d.n_var), np.diff(colptr))
    on_path = np.isin(rowidx, rows)
    assert on_path.sum() == d.n_mesh * (2 + d.n_in - 1)  # t0, tf and the mesh point's states + control
    assert np.array_equal(v[on_path], vo[on_path])
    is_ctl = on_path & (cols >= 2 + 2 * d.n_grid)
    assert is_ctl.sum() == d.n_mesh and np.max(np.abs(v[is_ctl] - 1.0)) < 1e-7
    assert np.all(v[on_path & ~is_ctl] == 0.0)  # constant bound: no dependence on time or the states
    # a time-varying bound: the time column chains into t0 and tf
    p = load(pkg, "pendulum_cb_both")
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    ev.set_fd(scheme)
    o.set_fd(scheme)
    x = p.make_iterate(seed=9, spread=0.4)
    assert relerr(ev.eval_g(x), o.eval_g(x)) < 1e-12
    v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
    assert np.isfinite(vo).all() and np.max(np.abs(v - vo)) < 1e-7 * max(1.0, np.abs(vo).max())


@pytest.mark.parametrize("name", ["gait2d_tracking", "gait2d_prediction", "hanging_muscle", "pendulum3_implicit", "sliding_mass",
                                  "pendulum_cb_lower", "pendulum_cb_equality", "pendulum_cb_both"])
def test_random_sparsity_detection_keeps_every_nonzero(pkg, name):
    """optim_sparsity_detection = "random" (CasOCFunction.cpp:25-105): the detected structure is a subset of the dense-
    block structure, the surviving entries carry bit-identical values, and every dropped entry is exactly zero."""
    p = load(pkg, name)
    ev = pkg.Evaluator(p)
    colptr, rowidx = ev.structure()
    n_var, n_con = ev.dims.n_var, ev.dims.n_con
    cols = np.repeat(np.arange(n_var), np.diff(colptr))
    key = cols.astype(np.int64) * n_con + rowidx
    xs = [p.make_iterate(seed=s, spread=sp) for s, sp in ((31, 0.2), (32, 0.6), (33, 1.0))]
    dense = [ev.eval_jac_g(x) for x in xs]
    g = [ev.eval_g(x) for x in xs]
    ev.set_sparsity_detection("random")
    c2, r2 = ev.structure()
    assert ev.dims.nnz_jac == r2.size <= rowidx.size
    cols2 = np.repeat(np.arange(n_var), np.diff(c2))
    key2 = cols2.astype(np.int64) * n_con + r2
    assert np.all(np.diff(key2) > 0), "CSC order: columns ascending, rows ascending inside a column"
    pos = np.searchsorted(key, key2)
    assert np.array_equal(key[pos], key2), "detected structure is not a subset of the dense-block structure"
    dropped = np.ones(key.size, bool)
    dropped[pos] = False
    for x, vd, gd in zip(xs, dense, g):
        vs = ev.eval_jac_g(x)
        assert np.array_equal(vs, vd[pos], equal_nan=True)
        assert np.all(vd[dropped] == 0.0)
        assert np.array_equal(ev.eval_g(x), gd)
    if name.startswith("gait2d"):
        assert r2.size < 0.8 * rowidx.size  # the muscle blocks are far from dense
    ev.set_sparsity_detection("none")
    c3, r3 = ev.structure()
    assert np.array_equal(c3, colptr) and np.array_equal(r3, rowidx)


def test_fd_structural_zeros_are_exact(pkg):
    """Perturbing an input that an output does not depend on must give an exactly zero difference (the evaluations that
    reuse the nominal poses / velocities are bit-identical to full ones): muscle excitations do not enter udot, and an
    activation only enters its own muscle's activation dynamics."""
    p = load(pkg, "gait2d_tracking", 6)
    ev = pkg.Evaluator(p)
    d = ev.dims
    x = p.make_iterate(seed=21, spread=0.3)
    for scheme in (FD_FORWARD, FD_CENTRAL):
        ev.set_fd(scheme)
        J = ev.eval_points(x, jacobian=True)
        nq, nz = 10, 18
        ctl0 = 1 + d.ns  # first control input; control 0 is lumbarAct, 1..18 the muscle excitations
        assert np.all(J[:, :nq, ctl0 + 1:ctl0 + 19] == 0.0)
        zrows = J[:, nq:nq + nz, 1 + 2 * nq:1 + 2 * nq + nz]  # d zdot_m / d a_n
        off = ~np.eye(nz, dtype=bool)
        assert np.all(zrows[:, off] == 0.0)
        assert np.all(J[:, :nq + nz, 0] == 0.0)  # the dynamics do not depend on time


def test_exact_linear_entries(pkg):
    """Entries that do not pass through a callback are exact: interpolation rows (+1, -1/2, -1/2) and d(defect)/dx."""
    p = load(pkg, "gait2d_tracking", 8)
    ev = pkg.Evaluator(p)
    d = ev.dims
    x = p.make_iterate(seed=9, spread=0.2)
    colptr, rowidx = ev.structure()
    vals = ev.eval_jac_g(x)
    # rows of the interpolating-control constraints: the last nc rows of every interval block
    rows_per_interval = 2 * d.ns + d.nc
    first_interp = 2 * d.ns  # after the defects (no residual rows in explicit mode, no endpoint... skip endpoint rows)
    n_endpoint = d.n_con - rows_per_interval * (d.n_mesh - 1)
    seen = 0
    for c in range(d.n_var):
        for k in range(colptr[c], colptr[c + 1]):
            r = rowidx[k] - n_endpoint
            if r >= 0 and (r % rows_per_interval) >= first_interp:
                assert vals[k] in (1.0, -0.5)
                seen += 1
    assert seen == 3 * d.nc * (d.n_mesh - 1)


def test_nan_passes_through(pkg):
    """NaN/Inf in the iterate must come back as NaN in the outputs with status OK (IPOPT handles them)."""
    p = load(pkg, "pendulum2")
    ev = pkg.Evaluator(p)
    x = p.make_iterate(seed=1, spread=0.1)
    x[2 + 4 * 3] = np.nan  # a state of grid point 3
    g = ev.eval_g(x)
    assert np.isnan(g).any() and np.isfinite(g).any()
    v = ev.eval_jac_g(x)
    assert np.isnan(v).any()


@pytest.mark.parametrize("intervals", [1, 2, 7])
def test_small_and_odd_meshes(pkg, orc, intervals):
    for name in ("sliding_mass_trap", "hanging_muscle", "gait2d_prediction"):
        p = load(pkg, name, intervals)
        ev, o = pkg.Evaluator(p), orc.Oracle(p)
        x = p.make_iterate(seed=3, spread=0.4)
        assert relerr(ev.eval_g(x), o.eval_g(x)) < 1e-10
        ev.set_fd(FD_CENTRAL, 1e-3, 0)
        o.set_fd(FD_CENTRAL, 1e-3, 0)
        assert rowscaled_err(ev.eval_jac_g(x), o.eval_jac_g(x), o.structure()[1], o.n_con) < 5e-7


def test_nonuniform_mesh(pkg, orc):
    p = pkg.Problem.load(golden("pendulum2.mb20"), "mesh=0,0.1,0.15,0.5,0.9,1")
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    assert ev.dims.n_mesh == 6
    x = p.make_iterate(seed=2, spread=0.5)
    assert relerr(ev.eval_g(x), o.eval_g(x)) < 1e-10
    assert abs(ev.eval_f(x) - o.eval_f(x)) < 1e-10 * max(1, abs(o.eval_f(x)))


# ---- BASELINE config 3 at its full size (50 intervals, 101 grid points): everything against the oracle ------------------
def test_config3_full_size_against_oracle(pkg, orc):
    p = pkg.Problem.load(golden("gait2d_tracking.mb20"))
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    assert (ev.dims.n_grid, ev.dims.n_var, ev.dims.n_con, ev.dims.nnz_jac) == (101, 5759, 4788, 420258)
    c, r = ev.structure()
    co, ro = o.structure()
    assert np.array_equal(c, co) and np.array_equal(r, ro)
    for seed, spread in ((41, 0.1), (42, 0.4)):
        x = p.make_iterate(seed=seed, spread=spread)
        assert relerr(ev.eval_points(x), o.eval_points(x)) < 1e-10
        assert relerr(ev.eval_g(x), o.eval_g(x)) < 1e-10
        assert abs(ev.eval_f(x) - o.eval_f(x)) <= 1e-10 * max(1.0, abs(o.eval_f(x)))
        ev.set_fd(FD_FORWARD, 1e-4, 0)
        o.set_fd(FD_FORWARD, 1e-4, 0)
        assert rowscaled_err(ev.eval_jac_g(x), o.eval_jac_g(x), ro, o.n_con) < 5e-7
        ev.set_fd(FD_FORWARD)  # CasADi's default step, as the example uses: noise-model bound
        o.set_fd(FD_FORWARD)
        v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
        assert rowscaled_err(v, vo, ro, o.n_con) < 2e-3
        rowmax = np.zeros(o.n_con)
        np.maximum.at(rowmax, ro, np.abs(vo))
        assert np.median(np.abs(v - vo) / (np.abs(vo) + rowmax[ro] + 1e-300)) < 1e-8


# ---- full-size (BASELINE.json) properties that need no oracle run ---------------------------------------------
@pytest.fixture(scope="module")
def config5(pkg):
    p = pkg.Problem.load(golden("gait2d_prediction.mb20"), "num_mesh_intervals=800")
    return p, pkg.Evaluator(p)


def test_config5_point_sample_matches_oracle(config5, orc):
    """At 1601 grid points the oracle checks a sample of the points (each point function is independent)."""
    p, ev = config5
    o = orc.Oracle(p)
    x = p.make_iterate(seed=12, spread=0.3)
    Y = ev.eval_points(x)
    for j in (0, 1, 2, 799, 800, 1599, 1600):
        yo = o.eval_point(o.gather_inputs(x, j))
        assert np.max(np.abs(Y[j] - yo) / (np.abs(yo) + 1e-6 * np.abs(yo).max())) < 1e-10


def test_config5_g_is_the_transcription_of_the_point_outputs(config5):
    """eval_g at full size against a vectorised numpy restatement of the Hermite–Simpson defects
    (CasOCHermiteSimpson.cpp:53-105) fed with the device's own point outputs."""
    p, ev = config5
    d = ev.dims
    x = p.make_iterate(seed=13, spread=0.3)
    Y, g = ev.eval_points(x), ev.eval_g(x)
    N, ns, nc, nq = d.n_grid, d.ns, d.nc, 10
    X = x[2:2 + ns * N].reshape(N, ns)
    U = x[2 + ns * N:2 + (ns + nc) * N].reshape(N, nc)
    xdot = np.concatenate([X[:, nq:2 * nq], Y[:, :ns - nq]], axis=1)
    h = (x[1] - x[0]) / 800.0
    i, m, ip = slice(0, N - 1, 2), slice(1, N, 2), slice(2, N, 2)
    herm = X[m] - 0.5 * (X[ip] + X[i]) - (h / 8.0) * (xdot[i] - xdot[ip])
    simp = X[ip] - X[i] - (h / 6.0) * (xdot[ip] + 4 * xdot[m] + xdot[i])
    interp = U[m] - 0.5 * (U[ip] + U[i])
    want = np.concatenate([herm, simp, interp], axis=1).ravel()
    n_endpoint = d.n_con - want.size
    assert n_endpoint == 39
    assert np.allclose(g[n_endpoint:], want, rtol=1e-12, atol=1e-9 * np.abs(want).max())


def test_config5_jacobian_is_derivative_of_g(config5):
    p, ev = config5
    d = ev.dims
    ev.set_fd(FD_CENTRAL, 1e-3, 0)
    x = p.make_iterate(seed=14, spread=0.1)
    colptr, rowidx = ev.structure()
    vals = ev.eval_jac_g(x)
    rng = np.random.default_rng(1)
    v = rng.uniform(-1, 1, d.n_var)
    cols = np.repeat(np.arange(d.n_var), np.diff(colptr))
    Jv = np.bincount(rowidx, weights=vals * v[cols], minlength=d.n_con)
    e = 1e-6
    fd = (ev.eval_g(x + e * v) - ev.eval_g(x - e * v)) / (2 * e)
    assert np.max(np.abs(Jv - fd)) < 1e-4 * np.abs(fd).max()


def test_config5_invariances(config5):
    """Physics the evaluator must respect at any size: the prediction problem has no explicit time dependence, and
    moving the whole model forward (pelvis_tx) changes nothing because the ground is a horizontal half space."""
    p, ev = config5
    d = ev.dims
    x = p.make_iterate(seed=15, spread=0.2)
    Y = ev.eval_points(x)
    x2 = x.copy()
    x2[0] += 0.123
    x2[1] += 0.123
    assert np.max(np.abs(ev.eval_points(x2)[:, :28] - Y[:, :28])) == 0
    x3 = x.copy()
    x3[2 + 1:2 + d.ns * d.n_grid:d.ns] += 0.37  # state index 1 = pelvis_tx of every grid point
    Y3 = ev.eval_points(x3)
    assert np.max(np.abs(Y3 - Y) / (np.abs(Y) + 1e-6 * np.abs(Y).max())) < 1e-9


def test_config5_left_right_mirror(config5):
    """2D_gait.osim is left/right symmetric: swapping the legs' states and controls swaps the outputs."""
    p, ev = config5
    d = ev.dims
    N, ns, nc = d.n_grid, d.ns, d.nc
    x = p.make_iterate(seed=16, spread=0.2)
    # state order: [tilt tx ty hip_l hip_r lumbar knee_l knee_r ankle_l ankle_r] values, same speeds, then 18 activations
    # (9 right then 9 left); controls: lumbarAct, 9 right muscles, 9 left (Problem.names of the example)
    qperm = np.array([0, 1, 2, 4, 3, 5, 7, 6, 9, 8])
    sperm = np.concatenate([qperm, 10 + qperm, 20 + np.concatenate([np.arange(9, 18), np.arange(0, 9)])])
    cperm = np.concatenate([[0], np.arange(10, 19), np.arange(1, 10)])
    X = x[2:2 + ns * N].reshape(N, ns)
    U = x[2 + ns * N:2 + (ns + nc) * N].reshape(N, nc)
    xm = x.copy()
    xm[2:2 + ns * N] = X[:, sperm].ravel()
    xm[2 + ns * N:2 + (ns + nc) * N] = U[:, cperm].ravel()
    Y, Ym = ev.eval_points(x), ev.eval_points(xm)
    operm = np.concatenate([qperm, 10 + np.concatenate([np.arange(9, 18), np.arange(0, 9)])])
    a, b = Ym[:, :28], Y[:, operm]
    assert np.max(np.abs(a - b) / (np.abs(b) + 1e-6 * np.abs(b).max())) < 1e-8


def test_batch_of_independent_problems_is_bit_identical(pkg):
    """mocob200_eval_jac_g_batch: different problems side by side on their own streams give exactly the values each
    handle gives alone, whatever the order of the handles; results land in the caller's buffers or stay in the handle."""
    import torch

    names = ["gait2d_hs", "double_pendulum_trap", "gait2d_hs", "sliding_mass"]
    names = [n for n in names if n in FIXTURES] or list(FIXTURES)[:3]
    probs = [load(pkg, n, intervals=6 + 2 * k) for k, n in enumerate(names)]
    evs = [pkg.Evaluator(p) for p in probs]
    xs = [p.make_iterate(seed=10 + k, spread=0.3) for k, p in enumerate(probs)]
    alone = [ev.eval_jac_g(x).copy() for ev, x in zip(evs, xs)]
    dx = [torch.from_numpy(x).cuda() for x in xs]
    dv = [torch.full((ev.dims.nnz_jac,), float("nan"), dtype=torch.float64, device="cuda") for ev in evs]
    for order in (list(range(len(evs))), list(reversed(range(len(evs))))):
        for v in dv:
            v.fill_(float("nan"))
        for rep in range(3):  # the second and third pass replay the handles' CUDA graphs
            pkg.eval_jac_g_batch([evs[k] for k in order], [dx[k].data_ptr() for k in order], [dv[k].data_ptr() for k in order])
        evs[order[0]].sync()  # the join makes the first handle's stream wait for every other handle
        for k in range(len(evs)):
            assert np.array_equal(dv[k].cpu().numpy(), alone[k]), (order, k)
    with pytest.raises(RuntimeError):
        pkg.eval_jac_g_batch([evs[0], evs[0]], [dx[0].data_ptr()] * 2, None)


@pytest.mark.parametrize("name", ["gait2d_tracking", "pendulum_cb_both", "pendulum3_implicit"])
def test_pinned_host_buffer_receives_the_same_values(pkg, name):
    """Pinned and pageable caller buffers receive the same bits: on the first call (graph capture), on replays, when the
    same x comes again (re-assembly only) and when the buffer changes."""
    import torch

    p = load(pkg, name)
    ev = pkg.Evaluator(p)
    n = ev.dims.nnz_jac
    xs = [p.make_iterate(seed=s, spread=0.3) for s in (1, 2, 3)]
    pageable = [ev.eval_jac_g(x).copy() for x in xs]
    xp = [torch.from_numpy(x).pin_memory() for x in xs]
    bufs = [torch.full((n,), float("nan"), dtype=torch.float64).pin_memory() for _ in range(2)]
    for k in (0, 1, 2, 2, 0):  # 2, 2: same x twice
        for b in bufs:
            b.fill_(float("nan"))
            ev.eval_jac_g_ptr(xp[k].data_ptr(), b.data_ptr())
            assert np.array_equal(b.numpy(), pageable[k]), (k,)


@pytest.mark.parametrize("name", ["gait2d_prediction", "gait2d_tracking", "hanging_muscle", "pendulum3_implicit", "pendulum2"])
@pytest.mark.parametrize("fd", [FD_FORWARD, FD_CENTRAL])
def test_pose_reusing_thread_evaluation_matches_the_whole_evaluation(pkg, orc, name, fd, monkeypatch):
    """Thread path: the speed (and implicit-mode acceleration) directions go through eval_point_vel, which reads the nominal
    poses, path geometry and mass-matrix factor of the grid point from the blob of the nominal warp evaluation instead of
    recomputing them. Same function, so with a round-off-negligible fixed step the columns agree with whole evaluations
    (MOCOB200_VEL_KERNEL=0) and with the oracle to 1e-7 of the row scale."""
    monkeypatch.setenv("MOCOB200_KERNEL", "thread")
    monkeypatch.setenv("MOCOB200_GRAPH", "0")
    p = load(pkg, name, 3)
    x = p.make_iterate(seed=21, spread=0.4)
    h = 1e-4 if fd == FD_FORWARD else 1e-3
    out = {}
    for vel in ("1", "0"):
        monkeypatch.setenv("MOCOB200_VEL_KERNEL", vel)
        ev = pkg.Evaluator(p)
        ev.set_fd(fd, h, 0)
        out[vel] = ev.eval_points(x, jacobian=True)
    o = orc.Oracle(p)
    o.set_fd(fd, h, 0)
    _, Jo = o.eval_point_jacobians(x)
    scale = np.abs(Jo).max(axis=2, keepdims=True) + 1e-300
    assert np.array_equal(np.isnan(out["1"]), np.isnan(out["0"]))
    assert np.nanmax(np.abs(out["1"] - out["0"]) / scale) < 1e-8
    assert np.nanmax(np.abs(out["1"] - Jo) / scale) < 1e-7


@pytest.mark.parametrize("name", ["gait2d_tracking", "gait2d_prediction"])
def test_gait2d_implicit_mode_matches_oracle(pkg, orc, name):
    """MultibodySystemImplicit on the gait model (CasOCFunction.cpp:317-331): residuals, zdot and the FD blocks with the
    acceleration variables as inputs, values 1e-10, Jacobian entries 1e-7 of the row scale at round-off-negligible steps."""
    p = pkg.Problem.load(golden(name + ".mb20"), "multibody_dynamics_mode=implicit;num_mesh_intervals=6")
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    assert ev.dims.nd == 10 and ev.dims.n_in == 1 + 38 + 19 + 10
    x = p.make_iterate(seed=9, spread=0.3)
    Y, Yo = ev.eval_points(x), o.eval_points(x)
    assert relerr(Y, Yo) < 1e-10
    assert relerr(ev.eval_g(x), o.eval_g(x)) < 1e-10
    for fd, h in ((FD_FORWARD, 1e-4), (FD_CENTRAL, 1e-3)):
        ev.set_fd(fd, h, 0)
        o.set_fd(fd, h, 0)
        v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
        colptr, rowidx = ev.structure()
        assert np.array_equal(rowidx, o.structure()[1])
        assert rowscaled_err(v, vo, rowidx, ev.dims.n_con) < 1e-7


def test_config5_full_size_jacobian_matches_oracle(pkg, orc):
    """BASELINE config 5 (gaitPrediction, 800 intervals = 1601 grid points, central differences) at full size: every CSC value
    against the oracle's — 1e-7 of the row scale with a fixed round-off-negligible step, the refinement noise bound at
    CasADi's default steps. One oracle Jacobian of this size costs a fraction of a second on the host cores."""
    p = pkg.Problem.load(golden("gait2d_prediction.mb20"), "num_mesh_intervals=800")
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    assert ev.dims.n_grid == 1601
    x = p.make_iterate(seed=3, spread=0.1)
    colptr, rowidx = ev.structure()
    assert np.array_equal(rowidx, o.structure()[1])
    ev.set_fd(FD_CENTRAL, 1e-3, 0)
    o.set_fd(FD_CENTRAL, 1e-3, 0)
    assert rowscaled_err(ev.eval_jac_g(x), o.eval_jac_g(x), rowidx, ev.dims.n_con) < 1e-7
    ev.set_fd(FD_CENTRAL)
    o.set_fd(FD_CENTRAL)
    v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
    rowmax = np.zeros(ev.dims.n_con)
    np.maximum.at(rowmax, rowidx, np.abs(vo))
    rel = np.abs(v - vo) / (np.abs(vo) + rowmax[rowidx] + 1e-300)
    assert rel.max() < 2e-3 and np.median(rel) < 1e-6


@pytest.mark.parametrize("workload", ["config3", "config5"])
def test_bench_trajectory_iterates_match_oracle(pkg, orc, workload):
    """The bench's default iterates (reference walking trajectory + bounded noise, feet in contact) at the bench's
    full sizes: point outputs and g against the oracle to 1e-10."""
    import bench

    fixture, opts, fd, _ = bench.WORKLOADS[workload]
    p = pkg.Problem.load(golden(fixture), opts)
    x = bench.make_iterates(pkg, p, 2, "trajectory")[1]
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    Y, Yo = ev.eval_points(x), o.eval_points(x)
    assert np.isfinite(Yo).all()
    # with the feet on the ground the contact stiffness (3e6 N/m^1.5 smoothed Hertz) multiplies the 1e-16 pose differences of
    # the two formulations: measured 1.0e-10 on the raw point outputs at 1601 points; the constraint values keep 1e-10
    assert relerr(Y, Yo) < 5e-10
    assert relerr(ev.eval_g(x), o.eval_g(x)) < 1e-10
    # the feet are on the ground: the contact forces are not negligible in these iterates
    nq = 10
    assert np.abs(Yo[:, :nq]).max() > 10.0
