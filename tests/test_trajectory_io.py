"""Guess / solution files: MocoTrajectory tables written and read by the host library
(Moco/Moco/MocoTrajectory.cpp:610-860; setGuessFile, MocoCasADiSolver.h:162-202). CPU only."""
import os

import numpy as np
import pytest

from conftest import golden


def _example(pkg, intervals):
    return pkg.Problem.example("pendulum:3", ".", "multibody_dynamics_mode=implicit;num_mesh_intervals=%d" % intervals)


def test_write_then_read_is_the_identity(pkg, tmp_path):
    p = _example(pkg, 12)
    x = p.make_iterate(seed=3, spread=0.5)
    path = tmp_path / "iterate.sto"
    p.write_sto(x, path)
    head = open(path).read().split("endheader")[0]
    # the metadata MocoTrajectory's file constructor reads (MocoTrajectory.cpp:664-692)
    for key in ("num_states=6", "num_controls=3", "num_derivatives=3", "num_multipliers=0", "num_slacks=0", "num_parameters=0",
                "inDegrees=no", "DataType=double", "version=3"):
        assert key in head
    labels = open(path).read().split("endheader\n")[1].split("\n")[0].split("\t")
    assert labels[0] == "time" and labels[1] == "/jointset/j0/q0/value" and labels[7] == "/tau0"
    assert labels[10] == "/jointset/j0/q0/accel"  # implicit mode: acceleration variables named after the speeds
    assert np.array_equal(p.read_sto(path), x)


def test_guess_is_resampled_onto_the_grid(pkg, tmp_path):
    """MocoTrajectory::resample: interpolating spline of degree min(rows - 1, 5) through the table, evaluated at the grid
    times between the first and the last row (CasOCTranscription.cpp:595-597)."""
    p = _example(pkg, 12)
    d = p.dims()
    x = p.make_iterate(seed=3, spread=0.5)
    t = x[0] + (x[1] - x[0]) * np.linspace(0, 1, d.n_grid)
    for k in range(d.ns):
        x[2 + np.arange(d.n_grid) * d.ns + k] = np.sin(1.3 * t + k)
    off_c = 2 + d.ns * d.n_grid
    for k in range(d.nc):
        x[off_c + np.arange(d.n_grid) * d.nc + k] = np.cos(0.7 * t + k)
    path = tmp_path / "guess.sto"
    p.write_sto(x, path)
    q = _example(pkg, 17)
    dq = q.dims()
    xr = q.read_sto(path)
    assert (xr[0], xr[1]) == (x[0], x[1])
    tq = x[0] + (x[1] - x[0]) * np.linspace(0, 1, dq.n_grid)
    for k in range(d.ns):
        assert np.abs(xr[2 + np.arange(dq.n_grid) * d.ns + k] - np.sin(1.3 * tq + k)).max() < 1e-4
    off_cq = 2 + d.ns * dq.n_grid
    for k in range(d.nc):
        assert np.abs(xr[off_cq + np.arange(dq.n_grid) * d.nc + k] - np.cos(0.7 * tq + k)).max() < 1e-4
    # the table's own rows are interpolated exactly
    assert xr[2] == pytest.approx(x[2], abs=1e-12)


def _sine_guess(p, tmp_path, name):
    d = p.dims()
    x = p.make_iterate(seed=3, spread=0.5)
    t = x[0] + (x[1] - x[0]) * np.linspace(0, 1, d.n_grid)
    for k in range(d.ns):
        x[2 + np.arange(d.n_grid) * d.ns + k] = np.sin(1.3 * t + k)
    off_c = 2 + d.ns * d.n_grid
    for k in range(d.nc):
        x[off_c + np.arange(d.n_grid) * d.nc + k] = np.cos(0.7 * t + k)
    path = tmp_path / name
    p.write_sto(x, path)
    return x, path


def test_banded_spline_fit_equals_the_dense_one(pkg, tmp_path, monkeypatch):
    """Long tables go through a banded factorisation of the same spline equations (piece-by-piece ordering); forced onto a
    short table it reproduces the dense solve to round-off."""
    x, path = _sine_guess(_example(pkg, 30), tmp_path, "guess.sto")
    q = _example(pkg, 17)
    dense = q.read_sto(path)
    monkeypatch.setenv("MOCOB200_SPLINE_DENSE_MAX", "0")
    banded = q.read_sto(path)
    assert not np.array_equal(dense, banded)  # another elimination order
    assert np.abs(dense - banded).max() < 1e-11


def test_long_guess_is_resampled(pkg, tmp_path):
    """A solution on a fine mesh as the guess of another mesh (mesh refinement): 1001 rows -> 401 and -> 1201 grid
    points; linear cost in the number of rows."""
    import time

    x, path = _sine_guess(_example(pkg, 500), tmp_path, "fine.sto")
    for intervals in (200, 600):
        q = _example(pkg, intervals)
        dq = q.dims()
        t0 = time.perf_counter()
        xr = q.read_sto(path)
        assert time.perf_counter() - t0 < 5.0
        tq = x[0] + (x[1] - x[0]) * np.linspace(0, 1, dq.n_grid)
        for k in range(dq.ns):
            assert np.abs(xr[2 + np.arange(dq.n_grid) * dq.ns + k] - np.sin(1.3 * tq + k)).max() < 1e-9
        off_cq = 2 + dq.ns * dq.n_grid
        for k in range(dq.nc):
            assert np.abs(xr[off_cq + np.arange(dq.n_grid) * dq.nc + k] - np.cos(0.7 * tq + k)).max() < 1e-9


def test_incompatible_or_nameless_guesses_are_refused(pkg, tmp_path):
    """MocoTrajectory::isCompatible: every variable of the problem needs a column of its name. The file here is the head of
    a solution written by the reference itself (Moco/Tests/std_testMocoInverse_subject_18musc_solution.sto), so the header
    parser also sees the reference's metadata lines (objective=, status=, num_iterations= ...)."""
    p = _example(pkg, 5)
    with pytest.raises(RuntimeError, match="no column '/jointset/j0/q0/value'"):
        p.read_sto(golden("moco_trajectory_head.sto"))
    loaded = pkg.Problem.load(golden("pendulum3_implicit.mb20"))  # .mb20 files hold numbers only
    with pytest.raises(RuntimeError, match="no variable names"):
        loaded.read_sto(golden("moco_trajectory_head.sto"))
    with pytest.raises(RuntimeError):
        p.read_sto(os.path.join(str(tmp_path), "does_not_exist.sto"))


def test_multiplier_and_slack_columns(pkg, tmp_path):
    """Kinematic constraints: the table carries the multipliers (one row per grid time) between the controls and the
    derivatives and the velocity-correction slacks last (MocoTrajectory.cpp:793-829). The slacks exist on the interval
    midpoints only; the reference stretches those values over linspace(t0, tf) and interpolates them piecewise-linearly onto
    every grid time when it builds the table (MocoCasOCProblem.h:170-185), and a guess whose slack column has one value per
    grid time keeps the entries of the midpoints (CasOCTranscription.cpp:601-618)."""
    p = pkg.Problem.example("double_pendulum_coupler", ".", "num_mesh_intervals=6")
    d = p.dims()
    assert d.nm == 1 and d.nsl == 1 and d.n_grid == 13
    x = p.make_iterate(seed=5, spread=0.5)
    n_int = d.n_grid - d.n_mesh
    off_m = 2 + (d.ns + d.nc) * d.n_grid
    off_s = off_m + d.nm * d.n_grid
    off_d = off_s + d.nsl * n_int
    assert off_d + d.nd * d.n_grid == d.n_var
    path = tmp_path / "constrained.sto"
    p.write_sto(x, path)
    head, body = open(path).read().split("endheader\n")
    assert "num_multipliers=1" in head and "num_slacks=1" in head
    labels = body.split("\n")[0].split("\t")
    assert labels[1 + d.ns + d.nc] == p.names(3)[0] == "lambda_cid0_p0"
    assert labels[-1] == p.names(4)[0] == "gamma_cid0_p0"
    assert len(labels) == 1 + d.ns + d.nc + d.nm + d.nd + d.nsl
    table = np.array([[float(v) for v in row.split("\t")] for row in body.split("\n")[1:] if row])
    assert table.shape == (d.n_grid, len(labels))
    assert np.array_equal(table[:, 1 + d.ns + d.nc], x[off_m:off_m + d.n_grid])
    slack = x[off_s:off_s + n_int]
    expect = np.interp(np.linspace(0, 1, d.n_grid), np.linspace(0, 1, n_int), slack)
    assert np.abs(table[:, -1] - expect).max() < 1e-15
    # read back: everything but the slacks is the identity; the slacks are the file's column at the midpoints
    xr = p.read_sto(path)
    keep = np.ones(d.n_var, bool)
    keep[off_s:off_d] = False
    assert np.array_equal(xr[keep], x[keep])
    assert np.array_equal(xr[off_s:off_d], table[1::2, -1])
    # a file without the multiplier column is not a compatible guess
    q = pkg.Problem.example("pendulum:2", ".", "num_mesh_intervals=6")
    plain = tmp_path / "plain.sto"
    q.write_sto(q.make_iterate(seed=1, spread=0.2), plain)
    with pytest.raises(Exception):
        p.read_sto(plain)
