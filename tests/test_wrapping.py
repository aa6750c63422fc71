"""Cylinder wrapping (Simulation/Wrap/WrapCylinder.cpp, GeometryPath.cpp:868-1094) pinned to numbers the
REFERENCE ITSELF produced: opensim-models/Pipelines/Rajagopal/outputReference/Results_CMC — the reference's CMC run on the
full-body Rajagopal model (80 muscles, 38 of its wrap cylinders in use through 46 PathWraps, 4 muscles with two of them;
CustomJoints, UniversalJoints, document Version 30000). cmc_states.sto holds q and u, cmc_MuscleAnalysis_Length.sto the
musculotendon path lengths and cmc_Actuation_speed.sto the path lengthening speeds, all to 8 decimals. Fixtures (every 4th
states row; all 47 Actuation rows): tests/golden/rajagopal_cmc_rows.npz + rajagopal_replay.mb20, scripts/make_wrapping_pins.py.

End to end: .osim -> flattener (wrap frames, quadrants, ranges) -> tables -> path evaluation with wrapping."""
import ctypes as C

import numpy as np
import pytest

from conftest import golden

# the files round to 1e-8 (half a unit = 5e-9) and the states they are evaluated at are rounded to 1e-8 as well
LENGTH_TOL = 1.2e-8  # m
SPEED_TOL = 1.5e-8   # m/s


@pytest.fixture(scope="module")
def rows():
    return np.load(golden("rajagopal_cmc_rows.npz"))


def wraps_per_muscle(p):
    t = p.table_sizes()
    idx = np.ctypeslib.as_array(C.cast(t.idx, C.POINTER(C.c_int)), shape=(t.n_int,))
    return np.array([int(idx[t.o_pw_i + 2 * k + 1]) for k in range(t.nmus)])


def test_flattened_model_has_the_wrap_objects(pkg):
    p = pkg.Problem.load(golden("rajagopal_replay.mb20"))
    t = p.table_sizes()
    assert (t.nb, t.nq, t.nmus, t.npp) == (22, 39, 80, 288)
    assert (t.nwrapobj, t.npathwrap) == (38, 46)
    nw = wraps_per_muscle(p)
    assert (nw == 2).sum() == 4 and (nw == 1).sum() == 38 and (nw == 0).sum() == 38


def oracle_paths(pkg, orc, q, u):
    p = pkg.Problem.load(golden("rajagopal_replay.mb20"))
    o = orc.Oracle(p)
    out = np.zeros(2)
    res = np.zeros((len(q), p.table_sizes().nmus, 2))
    for r in range(len(q)):
        qq, uu = np.ascontiguousarray(q[r]), np.ascontiguousarray(u[r])
        for m in range(res.shape[1]):
            orc.lib().orc_path(o._h, m, orc._d(qq), orc._d(uu), orc._d(out))
            res[r, m] = out
    return p, res


def test_oracle_path_lengths_match_the_reference_muscle_analysis(pkg, orc, rows):
    p, res = oracle_paths(pkg, orc, rows["q"], rows["u"])
    err = np.abs(res[:, :, 0] - rows["lengths"]).max(axis=0)
    nw = wraps_per_muscle(p)
    assert err[nw == 0].max() < LENGTH_TOL                      # straight paths: joints, frames, path points
    assert err[nw > 0].max() < LENGTH_TOL, rows["muscles"][np.argmax(err)]  # paths over one or two cylinders


def test_oracle_path_speeds_match_the_reference_actuation_analysis(pkg, orc, rows):
    _, res = oracle_paths(pkg, orc, rows["q_v"], rows["u_v"])
    err = np.abs(res[:, :, 1] - rows["speeds"])
    assert err.max() < SPEED_TOL, rows["muscles"][np.argmax(err.max(axis=0))]


def test_wrapping_changes_the_lengths_it_should(pkg, orc, rows):
    """Guard against a vacuous pin: with the cylinders switched off the wrapped muscles come out shorter by millimetres."""
    p, res = oracle_paths(pkg, orc, rows["q"][:8], rows["u"][:8])
    nw = wraps_per_muscle(p)
    t = p.table_sizes()
    idx = np.ctypeslib.as_array(C.cast(t.idx, C.POINTER(C.c_int)), shape=(t.n_int,))
    saved = idx[t.o_wo_i + 3:t.o_wo_i + 4 * t.nwrapobj:4].copy()
    idx[t.o_wo_i + 3:t.o_wo_i + 4 * t.nwrapobj:4] = 0  # active = false
    try:
        o = orc.Oracle(p)
        out = np.zeros(2)
        off = np.zeros((8, t.nmus))
        for r in range(8):
            qq, uu = np.ascontiguousarray(rows["q"][r]), np.ascontiguousarray(rows["u"][r])
            for m in range(t.nmus):
                orc.lib().orc_path(o._h, m, orc._d(qq), orc._d(uu), orc._d(out))
                off[r, m] = out[0]
    finally:
        idx[t.o_wo_i + 3:t.o_wo_i + 4 * t.nwrapobj:4] = saved
    d = res[:, :, 0] - off
    assert np.abs(d[:, nw == 0]).max() == 0.0
    assert d[:, nw > 0].max() > 5e-3 and d.min() > -1e-12  # wrapping never shortens a path
    assert (np.abs(d[:, nw > 0]).max(axis=0) > 1e-6).sum() >= 20  # most wrapped muscles do wrap somewhere in the trial


def test_tension_through_the_tangent_points_is_consistent_with_the_path_length(pkg, orc, rows):
    """Forces are applied through the tangent points as stations of the cylinders' bodies (GeometryPath.cpp:306-432 with
    PathWrapPoints). Implementation-independent check: the generalized force of a unit tension equals -dL/dq - exactly for
    straight paths, and to the 0.1 degree tangency tolerance of the reference's heuristic (about 1e-4 m of moment arm) for
    wrapped ones."""
    p = pkg.Problem.load(golden("rajagopal_replay.mb20"))
    o = orc.Oracle(p)
    nw = wraps_per_muscle(p)
    nq = p.table_sizes().nq
    out, tau, u = np.zeros(2), np.zeros(nq), np.zeros(nq)
    worst = {0: 0.0, 1: 0.0, 2: 0.0}
    q = np.ascontiguousarray(rows["q"][40])
    for m in range(len(nw)):
        orc.lib().orc_path_generalized_force(o._h, m, orc._d(q), 1.0, orc._d(tau))
        g = np.zeros(nq)
        for k in range(nq):
            h = 1e-6
            qa, qb = q.copy(), q.copy()
            qa[k] += h
            qb[k] -= h
            orc.lib().orc_path(o._h, m, orc._d(qa), orc._d(u), orc._d(out))
            la = out[0]
            orc.lib().orc_path(o._h, m, orc._d(qb), orc._d(u), orc._d(out))
            g[k] = -(la - out[0]) / (2 * h)
        worst[int(nw[m])] = max(worst[int(nw[m])], float(np.abs(g - tau).max()))
    assert worst[0] < 2e-9
    assert 1e-7 < max(worst[1], worst[2]) < 5e-4  # moment arms are centimetres: a few parts in a thousand


def replay_problem(pkg, times):
    """The replay problem on a mesh whose grid points sit at the given times (trapezoidal: grid = mesh)."""
    mesh = (times - times[0]) / (times[-1] - times[0])
    mesh[0], mesh[-1] = 0.0, 1.0
    opts = "transcription_scheme=trapezoidal;mesh=" + " ".join("%.17g" % v for v in mesh)
    return pkg.Problem.load(golden("rajagopal_replay.mb20"), opts)


def iterate_at(p, times, q, u):
    d = p.dims()
    nq = q.shape[1]
    x = np.zeros(d.n_var)
    x[0], x[1] = times[0], times[-1]
    st = np.zeros((d.n_grid, d.ns))
    st[:, :nq] = q
    st[:, nq:2 * nq] = u
    x[2:2 + d.n_grid * d.ns] = st.ravel()
    return x


@pytest.mark.gpu
def test_gpu_path_lengths_and_speeds_match_the_reference(pkg, orc, rows):
    p = replay_problem(pkg, rows["time"].copy())
    ev = pkg.Evaluator(p)
    paths = ev.eval_paths(iterate_at(p, rows["time"], rows["q"], rows["u"]))
    assert np.abs(paths[:, :, 0] - rows["lengths"]).max() < LENGTH_TOL
    pv = replay_problem(pkg, rows["time_v"].copy())
    evv = pkg.Evaluator(pv)
    xv = iterate_at(pv, rows["time_v"], rows["q_v"], rows["u_v"])
    pathsv = evv.eval_paths(xv)
    assert np.abs(pathsv[:, :, 1] - rows["speeds"]).max() < SPEED_TOL
    # and the kernel against the oracle (same procedure, separately written): 1e-10
    o = orc.Oracle(pv)
    out = np.zeros(2)
    worst = 0.0
    for r in range(0, len(rows["time_v"]), 6):
        qq, uu = np.ascontiguousarray(rows["q_v"][r]), np.ascontiguousarray(rows["u_v"][r])
        for m in range(pathsv.shape[1]):
            orc.lib().orc_path(o._h, m, orc._d(qq), orc._d(uu), orc._d(out))
            worst = max(worst, abs(out[0] - pathsv[r, m, 0]) / out[0], abs(out[1] - pathsv[r, m, 1]))
    assert worst < 1e-10, worst


@pytest.mark.gpu
def test_gpu_dynamics_with_wrap_cylinders_match_the_oracle(pkg, orc, rows):
    """The full-body model through the point functions (explicit forward dynamics of 39 coordinates, 80 path actuators over
    their wrap cylinders, 2 coupler constraints): kernels against the oracle, values 1e-10 and FD Jacobian blocks 1e-7."""
    sel = np.arange(0, len(rows["time_v"]), 8)
    p = replay_problem(pkg, rows["time_v"][sel].copy())
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    x = iterate_at(p, rows["time_v"][sel], rows["q_v"][sel], rows["u_v"][sel])
    d = p.dims()
    rng = np.random.default_rng(0)
    N = d.n_grid
    ctl = x[2 + N * d.ns:2 + N * (d.ns + d.nc)].reshape(N, d.nc)
    ctl[:] = rng.uniform(0.0, 300.0, ctl.shape)  # tensions (N) and actuator controls
    x[2 + N * (d.ns + d.nc):2 + N * (d.ns + d.nc + d.nm)] = rng.uniform(-50, 50, N * d.nm)
    Y, Yo = ev.eval_points(x), o.eval_points(x)
    assert np.isfinite(Yo).all()
    assert np.max(np.abs(Y - Yo) / (np.abs(Yo) + 1e-6 * np.abs(Yo).max())) < 1e-10
    assert np.max(np.abs(ev.eval_g(x) - o.eval_g(x))) < 1e-10 * np.abs(o.eval_g(x)).max()
    # the mass matrix of the full-body model has a condition number of 7e5 (wrists against the trunk), so the accelerations of
    # the light segments carry 1e-10 of round-off and a forward quotient with h = 1e-4 differs by 1e-6 between ANY two
    # implementations (measured: the mismatch scales with 1/h and sits in one wrist row); h = 1e-3 keeps it below the bar
    ev.set_fd(0, 1e-3, 0)
    o.set_fd(0, 1e-3, 0)
    J = ev.eval_points(x, jacobian=True)
    _, Jo = o.eval_point_jacobians(x)
    scale = np.abs(Yo)[:, :, None] + np.abs(Jo) + 1e-4 * np.abs(Yo).max()
    assert np.max(np.abs(J - Jo) / scale) < 1e-7



# ---- BASELINE config 4: exampleMocoInverse on the armless Rajagopal model (80 DeGrooteFregly2016 muscles, prescribed kinematics)
def test_config4_layout(pkg):
    """exampleMocoInverse.cpp:34-77 through MocoInverse.cpp:44-140: 80 activation states, 80 excitations + 6 residual and 10
    reserve actuators, 2 multipliers (the patella couplers; no constraint rows with prescribed kinematics), no derivative
    variables (rigid tendons), 18 multibody residual rows per grid point, 80 initial-activation rows."""
    p = pkg.Problem.load(golden("rajagopal_inverse.mb20"))
    d, t = p.dims(), p.table_sizes()
    assert (t.nb, t.nq, t.nmus, t.prescribed, t.npathwrap, t.ncoupler, t.nxforce) == (14, 18, 80, 1, 0, 2, 2)
    assert (d.n_mesh, d.n_grid, d.ns, d.nc, d.nd, d.nm, d.n_kc) == (50, 99, 80, 96, 0, 2, 0)
    assert d.n_in == 1 + 80 + 96 + 2 and d.n_out == 18 + 80 + 1
    assert d.n_var == 2 + 99 * (80 + 96 + 2)
    assert d.n_con == 80 + 99 * 18 + 49 * 2 * 80


@pytest.mark.gpu
def test_config4_gpu_matches_oracle(pkg, orc):
    p = pkg.Problem.load(golden("rajagopal_inverse.mb20"), "num_mesh_intervals=6")
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    x = p.make_iterate(seed=2, spread=0.3)
    Y, Yo = ev.eval_points(x), o.eval_points(x)
    assert np.isfinite(Yo).all()
    assert np.max(np.abs(Y - Yo) / (np.abs(Yo) + 1e-6 * np.abs(Yo).max())) < 1e-10
    g, go = ev.eval_g(x), o.eval_g(x)
    assert np.max(np.abs(g - go)) < 1e-10 * np.abs(go).max()
    f, fo = ev.eval_f(x), o.eval_f(x)
    assert abs(f - fo) < 1e-10 * abs(fo)
    ev.set_fd(0, 1e-4, 0)
    o.set_fd(0, 1e-4, 0)
    v, vo = ev.eval_jac_g(x), o.eval_jac_g(x)
    _, rowidx = o.structure()
    rowmax = np.zeros(o.n_con)
    np.maximum.at(rowmax, rowidx, np.abs(vo))
    assert np.max(np.abs(v - vo) / (np.abs(vo) + rowmax[rowidx] + 1e-300)) < 5e-7
    gr, gro = ev.eval_grad_f(x), o.eval_grad_f(x)
    assert np.max(np.abs(gr - gro)) <= 1e-7 * max(1.0, np.abs(gro).max(), abs(fo))
