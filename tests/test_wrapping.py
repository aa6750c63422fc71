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
def test_gpu_refuses_dynamics_with_wrap_cylinders_loudly(pkg, rows):
    """Until the warp kernels carry the wrap points, a model with wrap cylinders must not silently evaluate without them."""
    p = replay_problem(pkg, rows["time_v"][:4].copy())
    ev = pkg.Evaluator(p)
    x = iterate_at(p, rows["time_v"][:4], rows["q_v"][:4], rows["u_v"][:4])
    with pytest.raises(RuntimeError, match="wrap cylinders"):
        ev.eval_points(x)
    with pytest.raises(RuntimeError, match="wrap cylinders"):
        ev.eval_jac_g(x)
