"""Pins against numbers the REFERENCE ITSELF produced (real OpenSim + Simbody), not against this repo's own restatement.

Data: opensim-models/Pipelines/Gait10dof18musc/OutputReference/CMC/ResultsCMC (the reference's CMC tool on a walking trial):
states q, u (walk_subject_Kinematics_{q,u}.sto), accelerations (…_dudt.sto), actuator forces (…_Actuation_force.sto) and
path lengthening speeds (…_Actuation_speed.sto), 8 decimals, 4416 rows of which every 8th is committed as
tests/golden/gait10dof_cmc_rows.npz by scripts/make_reference_pins.py (which also flattens the replay model
tests/golden/gait10dof_replay.mb20 from CMC/subject01_metabolics.osim, document Version 30000: CustomJoint pelvis and
knees with SimmSpline translations, conditional and moving path points).

What each test pins, end to end (.osim -> flattener -> tables -> evaluation):
  * path speeds:  joint frame conventions, CustomJoint transform + motion axes, path-point kinds, SimmSpline, Point::calcSpeedBetween
  * accelerations: all of the above + mass properties, gravity, GeometryPath::addInEquivalentForces, CoordinateActuator,
                   ExternalForce (GCVSpline of the measured ground reactions), and the multibody forward dynamics itself
Tolerances are what the files' 8 decimals allow (stated per test).
"""
import numpy as np
import pytest

from conftest import golden


@pytest.fixture(scope="module")
def rows():
    return np.load(golden("gait10dof_cmc_rows.npz"))


def replay_problem(pkg, rows, sel):
    """The replay problem on a mesh whose grid points sit at the selected rows' times (trapezoidal: grid = mesh)."""
    t = rows["time"][sel]
    mesh = (t - t[0]) / (t[-1] - t[0])
    mesh[0], mesh[-1] = 0.0, 1.0
    opts = "transcription_scheme=trapezoidal;mesh=" + " ".join("%.17g" % v for v in mesh)
    return pkg.Problem.load(golden("gait10dof_replay.mb20"), opts)


def iterate_from_rows(p, rows, sel):
    d = p.dims()
    x = np.zeros(d.n_var)
    t = rows["time"][sel]
    x[0], x[1] = t[0], t[-1]
    n = len(sel)
    ns, nc = d.ns, d.nc
    states = np.hstack([rows["q"][sel], rows["u"][sel]])
    assert states.shape == (n, ns) and d.n_grid == n
    x[2:2 + n * ns] = states.reshape(-1)
    x[2 + n * ns:2 + n * ns + n * nc] = rows["controls"][sel].reshape(-1)
    return x


# accelerations: the recorded forces and ground-reaction points carry 8 decimals; 800 N of ground reaction times 1e-8 m of
# centre-of-pressure rounding is 8e-6 N m, which the 1.2 kg foot turns into ~1e-3 rad/s^2 at the ankle. Everywhere else the
# agreement is at the 1e-5 rad/s^2 level (1e-6 of the rms acceleration).
UDOT_TOL = {"ankle_angle_r": 2e-3, "ankle_angle_l": 2e-3}
UDOT_TOL_DEFAULT = 6e-5


def check_udot(udot, rows, sel):
    coords = [str(c) for c in rows["coords"]]
    ref = rows["udot"][sel]
    for k, c in enumerate(coords):
        err = np.abs(udot[:, k] - ref[:, k]).max()
        tol = UDOT_TOL.get(c, UDOT_TOL_DEFAULT)
        assert err < tol, "%s: max |udot - reference| = %.3g (tolerance %.3g)" % (c, err, tol)
        assert err < 3e-5 * np.sqrt((ref[:, k] ** 2).mean()), c  # relative to the rms acceleration of that coordinate


def test_oracle_path_speeds_match_the_reference_cmc_results(pkg, orc, rows):
    sel = np.arange(len(rows["time"]))
    p = replay_problem(pkg, rows, sel[:2])
    o = orc.Oracle(p)
    out = np.zeros(2)
    worst = 0.0
    for r in sel:
        q, u = np.ascontiguousarray(rows["q"][r]), np.ascontiguousarray(rows["u"][r])
        for m in range(rows["speeds"].shape[1]):
            orc.lib().orc_path(o._h, m, orc._d(q), orc._d(u), orc._d(out))
            worst = max(worst, abs(out[1] - rows["speeds"][r, m]))
    # the file rounds to 1e-8 m/s (half a unit = 5e-9); states are rounded to 1e-8 deg
    assert worst < 1.2e-8, worst


def test_oracle_forward_dynamics_matches_the_reference_cmc_results(pkg, orc, rows):
    sel = np.arange(len(rows["time"]))
    p = replay_problem(pkg, rows, sel[:2])
    o = orc.Oracle(p)
    nq = rows["q"].shape[1]
    udot = np.zeros((len(sel), nq))
    for k, r in enumerate(sel):
        x = np.concatenate([[rows["time"][r]], rows["q"][r], rows["u"][r], rows["controls"][r]])
        udot[k] = o.eval_point(x, 1)[:nq]
    check_udot(udot, rows, sel)


def test_oracle_custom_joint_power_balance(pkg, orc, rows):
    """Implementation-independent check of the function-based joint's velocity-product terms: d(KE + PE)/dt along a
    trajectory with no applied force is zero, i.e. u . (M udot + bias - gravity) = 0 identically."""
    sel = np.arange(0, len(rows["time"]), 40)
    p = replay_problem(pkg, rows, sel[:2])
    o = orc.Oracle(p)
    nq = rows["q"].shape[1]
    rng = np.random.default_rng(0)
    for r in sel:
        q = np.ascontiguousarray(rows["q"][r])
        u = np.ascontiguousarray(rows["u"][r] + rng.normal(size=nq))
        h = 1e-6
        e = np.zeros(2)
        com = np.zeros(3)

        def energy(qq, uu):
            orc.lib().orc_energy(o._h, orc._d(np.ascontiguousarray(qq)), orc._d(np.ascontiguousarray(uu)), orc._d(e), orc._d(com))
            return e[0] + e[1]

        # free motion: udot from inverse dynamics with zero applied force (tau = 0) -> solve M udot = -(bias - gravity)
        M = np.zeros((nq, nq))
        orc.lib().orc_mass_matrix(o._h, orc._d(q), orc._d(M))
        tau0 = np.zeros(nq)
        orc.lib().orc_inverse_dynamics(o._h, orc._d(q), orc._d(u), orc._d(np.zeros(nq)), 1, orc._d(tau0))
        udot = np.linalg.solve(M, -tau0)
        dE = (energy(q + h * u, u + h * udot) - energy(q - h * u, u - h * udot)) / (2 * h)
        scale = abs(e[0]) + abs(e[1])
        assert abs(dE) < 2e-6 * scale, (r, dE, scale)


@pytest.mark.gpu
def test_gpu_path_speeds_match_the_reference_cmc_results(pkg, rows):
    sel = np.arange(len(rows["time"]))
    p = replay_problem(pkg, rows, sel)
    ev = pkg.Evaluator(p)
    x = iterate_from_rows(p, rows, sel)
    paths = ev.eval_paths(x)
    assert np.abs(paths[:, :, 1] - rows["speeds"][sel]).max() < 1.2e-8


@pytest.mark.gpu
def test_gpu_forward_dynamics_matches_the_reference_cmc_results(pkg, orc, rows):
    sel = np.arange(len(rows["time"]))
    p = replay_problem(pkg, rows, sel)
    ev = pkg.Evaluator(p)
    x = iterate_from_rows(p, rows, sel)
    Y = ev.eval_points(x)
    nq = rows["q"].shape[1]
    check_udot(Y[:, :nq], rows, sel)
    # and the kernels against the oracle on this model (custom joints, path actuators, external forces): 1e-10
    Yo = orc.Oracle(p).eval_points(x)
    assert np.max(np.abs(Y - Yo) / (np.abs(Yo) + 1e-6 * np.abs(Yo).max())) < 1e-10


@pytest.mark.gpu
@pytest.mark.parametrize("fd", [0, 1])
def test_gpu_jacobian_matches_oracle_on_the_custom_joint_model(pkg, orc, rows, fd):
    sel = np.arange(0, len(rows["time"]), 24)
    p = replay_problem(pkg, rows, sel)
    ev, o = pkg.Evaluator(p), orc.Oracle(p)
    h = 1e-4 if fd == 0 else 1e-3
    ev.set_fd(fd, h, 0)
    o.set_fd(fd, h, 0)
    x = iterate_from_rows(p, rows, sel)
    J = ev.eval_points(x, jacobian=True)
    _, Jo = o.eval_point_jacobians(x)
    rowscale = np.abs(Jo).max(axis=2, keepdims=True) + 1e-300
    assert np.max(np.abs(J - Jo) / rowscale) < 1e-7
