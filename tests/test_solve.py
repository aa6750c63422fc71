"""The evaluator inside an optimisation loop: the reference's "Sliding mass" known answer (Moco/Tests/testMocoInterface.cpp:
1701-1742: minimum-time rest-to-rest move, bang-bang force, quadratic position, triangular speed), solved with the four TNLP
callbacks. The problem is exampleSlidingMass.cpp's (mass 2 kg, |force| <= 50 N, 1 m): final time 2*sqrt(m*d/F) = 0.4 s.
BASELINE.json's second metric asks for identical convergence of the GPU and the CPU path: final objective within 1e-6 and
iteration count within 5 %."""
import os
import sys

import numpy as np
import pytest

from nlp_solve import solve

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))  # ip_solve.py

T_EXACT = 2.0 * np.sqrt(2.0 * 1.0 / 50.0)


def _problem(pkg, scheme, intervals=20):
    return pkg.Problem.example("sliding_mass", ".", "num_mesh_intervals=%d;transcription_scheme=%s" % (intervals, scheme))


def _check_bang_bang(p, x, tol):
    d = p.dims()
    t0, tf = x[0], x[1]
    assert t0 == 0.0 and abs(tf - T_EXACT) < tol * T_EXACT
    states = x[2:2 + d.ns * d.n_grid].reshape(d.n_grid, d.ns)
    controls = x[2 + d.ns * d.n_grid:2 + (d.ns + d.nc) * d.n_grid].reshape(d.n_grid, d.nc)
    grid = np.linspace(0, 1, d.n_grid) * tf
    half, a = 0.5 * tf, 25.0
    pos = np.where(grid < half, 0.5 * a * grid ** 2, 1.0 - 0.5 * a * (tf - grid) ** 2)
    speed = np.where(grid < half, a * grid, a * (tf - grid))
    assert np.abs(states[:, 0] - pos).max() < tol
    away = np.abs(grid - half) > 1.5 * tf / (d.n_grid - 1)  # the switch is smeared over the grid points next to it
    assert np.abs(states[away, 1] - speed[away]).max() < tol * 5.0  # peak speed 5 m/s
    assert np.abs(controls[away, 0] - np.where(grid[away] < half, 50.0, -50.0)).max() < tol * 50.0


@pytest.mark.parametrize("scheme", ["trapezoidal", "hermite-simpson"])
def test_oracle_solves_the_sliding_mass(pkg, orc, scheme):
    p = _problem(pkg, scheme)
    nthreads = orc.max_threads()
    orc.set_threads(1)  # a 21-point problem: OpenMP fork/join would dominate
    try:
        res, stats = solve(p, orc.Oracle(p))
    finally:
        orc.set_threads(nthreads)
    assert res.success, res.message
    _check_bang_bang(p, res.x, 1e-2)
    assert abs(res.fun - res.x[1]) < 1e-12  # MocoFinalTimeGoal


@pytest.mark.gpu
@pytest.mark.parametrize("scheme", ["trapezoidal", "hermite-simpson"])
def test_gpu_solve_converges_like_the_cpu_solve(pkg, orc, scheme):
    p = _problem(pkg, scheme)
    nthreads = orc.max_threads()
    orc.set_threads(1)
    try:
        res_o, st_o = solve(p, orc.Oracle(p))
    finally:
        orc.set_threads(nthreads)
    res_g, st_g = solve(p, pkg.Evaluator(p))
    assert res_g.success, res_g.message
    _check_bang_bang(p, res_g.x, 1e-2)
    assert abs(res_g.fun - res_o.fun) < 1e-6
    assert abs(st_g["iterations"] - st_o["iterations"]) <= max(1, round(0.05 * st_o["iterations"]))
    assert np.abs(res_g.x - res_o.x).max() < 1e-4 * 50.0


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["hanging_muscle_rigidtendon", "hanging_muscle", "pendulum:2"])
def test_gpu_solve_of_muscle_and_pendulum_problems_converges_like_the_cpu_solve(pkg, orc, name):
    """Minimum-time lift with a DeGrooteFregly2016 muscle (testMocoActuators.cpp:1088-1160; rigid tendon, and compliant tendon
    in implicit mode) and the double pendulum: the finite-difference Jacobians of the two paths differ by round-off / h, the
    optimisation paths must not."""
    p = pkg.Problem.example(name, ".", "num_mesh_intervals=10")
    nthreads = orc.max_threads()
    orc.set_threads(1)
    try:
        res_o, st_o = solve(p, orc.Oracle(p), maxiter=150)
    finally:
        orc.set_threads(nthreads)
    res_g, st_g = solve(p, pkg.Evaluator(p), maxiter=150)
    assert res_o.success and res_g.success, (res_o.message, res_g.message)
    assert abs(res_g.fun - res_o.fun) < 1e-6 * max(1.0, abs(res_o.fun))
    assert abs(st_g["iterations"] - st_o["iterations"]) <= max(1, round(0.05 * st_o["iterations"]))


def test_interior_point_loop_solves_the_sliding_mass(pkg, orc):
    """The barrier-method stand-in (scripts/ip_solve.py: the algorithm family of IPOPT, which the reference drives through
    CasADi's nlpsol, CasOCTranscription.cpp:630-689) on the reference's known answer."""
    from ip_solve import solve as ip_solve

    p = _problem(pkg, "trapezoidal")
    nthreads = orc.max_threads()
    orc.set_threads(1)
    try:
        res, stats = ip_solve(p, orc.Oracle(p), tol=1e-6, maxiter=400)
    finally:
        orc.set_threads(nthreads)
    assert res.success, res.message
    assert res.constraint_violation < 1e-8
    _check_bang_bang(p, res.x, 1e-2)
    assert abs(res.fun - res.x[1]) < 1e-12
    assert stats["jac"] == stats["iterations"] + 2  # one eval_jac_g per iterate (as in the reference's loop) + one for the scaling


@pytest.mark.gpu
def test_gpu_interior_point_solve_converges_like_the_cpu_solve(pkg, orc):
    from ip_solve import solve as ip_solve

    p = _problem(pkg, "trapezoidal")
    nthreads = orc.max_threads()
    orc.set_threads(1)
    try:
        res_o, st_o = ip_solve(p, orc.Oracle(p), tol=1e-6, maxiter=400)
    finally:
        orc.set_threads(nthreads)
    res_g, st_g = ip_solve(p, pkg.Evaluator(p), tol=1e-6, maxiter=400)
    assert res_o.success and res_g.success, (res_o.message, res_g.message)
    assert abs(res_g.fun - res_o.fun) < 1e-6
    assert abs(st_g["iterations"] - st_o["iterations"]) <= max(1, round(0.05 * st_o["iterations"]))
