"""Pins the CPU oracle against the known-answer facts the reference's own tests hold for this path
(SURVEY.md §8(c)), and against physics that does not depend on any implementation.  CPU only."""
import math

import numpy as np
import pytest

from conftest import golden

# DeGrooteFregly2016Muscle defaults (Moco/Moco/Components/DeGrooteFregly2016Muscle.cpp:46-63) in MOCOB200_MUS_DSTRIDE order
DEFAULTS = dict(Fmax=1000.0, lopt=0.1, lslack=0.2, alpha=0.0, vmax=10.0, tau_a=0.015, tau_d=0.060, damping=0.0, eT=0.049,
                e0=0.6, wscale=1.0, minc=0.0, maxc=1.0)
RIGID, NO_ACT_DYN, NO_PASSIVE, IMPLICIT = 1, 2, 4, 8


def params(**kw):
    d = dict(DEFAULTS)
    d.update(kw)
    kT = math.log((1.0 + 0.2) / 0.2) / (1.0 + d["eT"] - 1.0)
    return [d[k] for k in "Fmax lopt lslack alpha vmax tau_a tau_d damping eT e0 wscale minc maxc".split()] + [kT]


def test_curve_anchor_values(orc):
    """Moco/Tests/testMocoActuators.cpp:199-222 ("Curve values")."""
    eT, e0 = DEFAULTS["eT"], DEFAULTS["e0"]
    assert orc.dgf_curve("tendon", 1.0, eT) == 0
    assert orc.dgf_curve("tendon", 1.0 + eT, eT) == pytest.approx(1, rel=1e-10)
    assert orc.dgf_curve("passive", 1.0, e0) == pytest.approx(0.0182288, rel=1e-4)
    assert orc.dgf_curve("passive", 0.2, e0) == pytest.approx(0, abs=1e-12)
    assert orc.dgf_curve("passive", 1.0 + e0, e0) == pytest.approx(1, rel=1e-4)
    assert orc.dgf_curve("active", 1.0, 1.0) == pytest.approx(1)
    assert orc.dgf_curve("fv", -1.0) == pytest.approx(0, abs=1e-15)
    assert orc.dgf_curve("fv", 0.0) == pytest.approx(1)
    assert orc.dgf_curve("fv", 1.0) == pytest.approx(1.794, rel=1e-3)


def test_force_velocity_inverse_round_trip(orc):
    """testMocoActuators.cpp:1026-1038 (f-v inverse curve)."""
    for v in np.linspace(-1, 1, 41):
        assert orc.dgf_curve("fv_inv", orc.dgf_curve("fv", v)) == pytest.approx(v, abs=1e-12)
    for f in np.linspace(0.0, 1.0, 11):
        assert orc.dgf_curve("tendon", orc.dgf_curve("tendon_inv", f, 0.049), 0.049) == pytest.approx(f, abs=1e-13)


def test_curve_derivatives_match_finite_differences(orc):
    h = 1e-6
    for x in np.linspace(0.4, 1.6, 13):
        for name, p0 in (("active", 1.3), ("passive", 0.6), ("tendon", 0.049)):
            xx = x if name != "tendon" else 1.0 + 0.04 * x
            fd = (orc.dgf_curve(name, xx + h, p0) - orc.dgf_curve(name, xx - h, p0)) / (2 * h)
            assert orc.dgf_curve(name + "_d", xx, p0) == pytest.approx(fd, rel=2e-7, abs=1e-8)


@pytest.mark.parametrize("nfl,act", [(1.0, 1.0), (1.0, 0.38), (0.5, 1.0)])
def test_rigid_tendon_operating_points_static(orc, nfl, act):
    """testMocoActuators.cpp:224-437: length = nfl*lopt + lslack, speed 0, rigid tendon, no damping."""
    P = params()
    r = orc.muscle(P, RIGID, nfl * 0.1 + 0.2, 0.0, act)
    fl, fp = orc.dgf_curve("active", nfl, 1.0), orc.dgf_curve("passive", nfl, 0.6)
    assert r["fiberLength"] == pytest.approx(nfl * 0.1)
    assert r["normFiberLength"] == pytest.approx(nfl)
    assert r["pennationAngle"] == pytest.approx(0.0, abs=1e-12)
    assert r["cosPennation"] == pytest.approx(1.0)
    assert r["tendonLength"] == 0.2
    assert r["fiberVelocity"] == 0 and r["normFiberVelocity"] == 0 and r["tendonVelocity"] == 0
    assert r["fv"] == pytest.approx(1.0)
    assert r["activeFiberForce"] == pytest.approx(1000 * act * fl)
    assert r["passiveFiberForce"] == pytest.approx(1000 * fp)
    assert r["fiberForce"] == pytest.approx(1000 * (act * fl + fp))
    assert r["tendonForce"] == pytest.approx(1000 * (act * fl + fp))
    assert math.isinf(r["tendonStiffness"])
    # stiffness against a numerical derivative of fiber force w.r.t. fiber length (testMocoActuators.cpp:296-311)
    h = 1e-7
    fp_ = orc.muscle(P, RIGID, nfl * 0.1 + 0.2 + h, 0.0, act)["fiberForce"]
    fm_ = orc.muscle(P, RIGID, nfl * 0.1 + 0.2 - h, 0.0, act)["fiberForce"]
    assert r["fiberStiffness"] == pytest.approx((fp_ - fm_) / (2 * h), rel=1e-6)
    assert r["fiberStiffnessAlongTendon"] == pytest.approx(r["fiberStiffness"], rel=1e-9)
    assert r["muscleStiffness"] == pytest.approx(r["fiberStiffnessAlongTendon"])


def test_rigid_tendon_velocity_operating_point(orc):
    """testMocoActuators.cpp:439-531: normalized fiber velocity 0.21."""
    vmax = 0.1 * 10.0
    r = orc.muscle(params(), RIGID, 0.3, 0.21 * vmax, 1.0)
    fv, fp = orc.dgf_curve("fv", 0.21), orc.dgf_curve("passive", 1.0, 0.6)
    assert r["fiberVelocity"] == pytest.approx(0.21 * vmax, rel=1e-15)
    assert r["normFiberVelocity"] == pytest.approx(0.21, rel=1e-15)
    assert r["fv"] == fv
    assert r["activeFiberForce"] == pytest.approx(1000 * fv)
    assert r["tendonForce"] == pytest.approx(1000 * (fv + fp))


def test_pennation_and_damping(orc):
    """testMocoActuators.cpp:533-760 (pennation angle 0.5 rad at optimal; fiber damping)."""
    alpha = 0.5
    P = params(alpha=alpha, damping=0.1)
    lopt = 0.1
    # fiber at optimal length: along-tendon length = lopt*cos(alpha)
    mt = lopt * math.cos(alpha) + 0.2
    v = 0.3
    r = orc.muscle(P, RIGID, mt, v, 0.7)
    assert r["normFiberLength"] == pytest.approx(1.0)
    assert r["pennationAngle"] == pytest.approx(alpha)
    assert r["cosPennation"] == pytest.approx(math.cos(alpha))
    nfv = v * math.cos(alpha) / (lopt * 10.0)
    assert r["normFiberVelocity"] == pytest.approx(nfv)
    fv, fp = orc.dgf_curve("fv", nfv), orc.dgf_curve("passive", 1.0, 0.6)
    assert r["fiberForce"] == pytest.approx(1000 * (0.7 * fv + fp + 0.1 * nfv))
    assert r["tendonForce"] == pytest.approx(r["fiberForce"] * math.cos(alpha))


def test_compliant_tendon_equilibrium_residual(orc):
    """testMocoActuators.cpp:832-862: at a (normalized tendon force, length) pair in equilibrium the residual is 0
    and explicit / implicit tendon modes agree on the tendon-force derivative."""
    P = params(eT=0.10, damping=0.0, lopt=0.10, lslack=0.05, Fmax=30.0, alpha=0.10)  # explicit mode inverts f-v without damping
    act, mt, mtv = 0.6, 0.16, 0.05
    # find ntf such that residual(ntf, rate=rateExplicit) = 0 by construction: explicit mode computes the rate
    ntf = 0.4
    e = orc.muscle(P, 0, mt, mtv, act, ntf)  # explicit tendon dynamics
    i = orc.muscle(P, IMPLICIT, mt, mtv, act, ntf, e["normTendonForceRate"])
    assert i["equilibriumResidual"] == pytest.approx(0.0, abs=1e-10 * 30)
    assert i["normFiberVelocity"] == pytest.approx(e["normFiberVelocity"], rel=1e-10)
    assert i["tendonForce"] == pytest.approx(30.0 * ntf)


def test_activation_dynamics(orc):
    """DeGrooteFregly2016Muscle.cpp:190-210: rate is zero at e == a, positive for e > a, and the time constants bound it."""
    assert orc.activation_derivative(0.5, 0.5) == 0
    assert orc.activation_derivative(1.0, 0.2) > 0 > orc.activation_derivative(0.0, 0.8)
    e, a = 0.9, 0.3
    z = 0.5 + 1.5 * a
    f = 0.5 * math.tanh(0.1 * (e - a))
    want = ((f + 0.5) / (0.015 * z) + (0.5 - f) * z / 0.060) * (e - a)
    assert orc.activation_derivative(e, a) == pytest.approx(want, rel=1e-15)


# ---------------------------------------------------------------------------------------------
def test_sliding_mass_is_f_equals_ma(orc):
    """testMocoInterface.cpp:66-83 / 1701-1742: 2 kg mass on a slider, force = control => udot = F/m."""
    o = orc.Oracle(golden("sliding_mass.mb20"))
    for F in (-50.0, 0.0, 3.7, 50.0):
        out = o.eval_point([0.3, 0.1, -2.0, F])
        assert out[0] == pytest.approx(F / 2.0, rel=1e-15, abs=1e-15)


def test_pendulum_matches_closed_form(orc):
    """ModelFactory::createNLinkPendulum (Components/ModelFactory.cpp:33-89): link k has mass 1 at its origin, 1 m from
    the joint; inertia (1,1,1) about the mass centre. Single link: (I + m l^2) udot = tau - m g l cos(q)."""
    import __graft_entry__ as ge

    m = ge.package()
    p = m.Problem.example("pendulum:1", ".", None)
    o = orc.Oracle(p)
    g = 9.80665
    for q, u, tau in ((0.0, 0.0, 0.0), (0.7, -1.3, 2.0), (-2.1, 0.4, -5.0)):
        out = o.eval_point([0.0, q, u, tau])
        assert out[0] == pytest.approx((tau - g * math.cos(q)) / 2.0, rel=1e-13)


@pytest.mark.parametrize("name", ["pendulum2", "pendulum3_implicit", "gait2d_tracking", "hanging_muscle"])
def test_multibody_physics_invariants(orc, name):
    """Implementation-independent checks of the oracle's multibody core: the mass matrix is symmetric positive
    definite and equals the inverse-dynamics columns; the power of the required generalized forces equals the rate of
    change of (kinetic + gravitational) energy along q(t) = q + u t + udot t^2/2."""
    o = orc.Oracle(golden(name + ".mb20"))
    rng = np.random.default_rng(5)
    nq = o.nq
    for trial in range(3):
        q, u, ud = rng.uniform(-0.5, 0.5, nq), rng.uniform(-2, 2, nq), rng.uniform(-5, 5, nq)
        M = o.mass_matrix(q)
        assert np.abs(M - M.T).max() == 0
        assert np.linalg.eigvalsh(M).min() > 0
        cols = np.stack([o.inverse_dynamics(q, np.zeros(nq), np.eye(nq)[k], gravity=False) for k in range(nq)], 1)
        assert np.abs(M - cols).max() < 1e-12 * max(1.0, np.abs(M).max())
        tau = o.inverse_dynamics(q, u, ud, gravity=True)
        h = 1e-5

        def energy(t):
            ke, pe, _ = o.energy(q + u * t + 0.5 * ud * t * t, u + ud * t)
            return ke + pe

        dE = (energy(h) - energy(-h)) / (2 * h)
        assert tau @ u == pytest.approx(dE, rel=1e-7, abs=1e-7)


def test_explicit_and_implicit_modes_agree(orc, pkg):
    """testImplicit.cpp:114-166 in point form: the implicit residual vanishes at the explicit-mode accelerations."""
    for name in ("pendulum:3", "gait2d"):
        if name == "gait2d":
            pe = pkg.Problem.load(golden("gait2d_prediction.mb20"))
            pi = pkg.Problem.load(golden("gait2d_prediction.mb20"), "multibody_dynamics_mode=implicit")
        else:
            pe = pkg.Problem.example(name, ".", None)
            pi = pkg.Problem.example(name, ".", "multibody_dynamics_mode=implicit")
        oe, oi = orc.Oracle(pe), orc.Oracle(pi)
        x = pe.make_iterate(seed=11, spread=0.3)
        for j in (0, 3, oe.n_grid - 1):
            ine = oe.gather_inputs(x, j)
            ye = oe.eval_point(ine)
            ini = np.concatenate([ine, ye[:oe.nq]])
            yi = oi.eval_point(ini)
            scale = max(1.0, np.abs(ye[:oe.nq]).max())
            assert np.abs(yi[:oe.nq]).max() < 1e-9 * scale * 100
            assert np.allclose(yi[oe.nq:oe.nq + oe.nz], ye[oe.nq:oe.nq + oe.nz], rtol=1e-13, atol=0)


def test_path_speed_is_time_derivative_of_length(orc):
    """GeometryPath lengthening speed (GeometryPath.cpp:848-864) must be d(length)/dt; moment arms are -dL/dq."""
    o = orc.Oracle(golden("gait2d_tracking.mb20"))
    rng = np.random.default_rng(2)
    # [tilt, tx, ty, hip_l, hip_r, lumbar, knee_l, knee_r, ankle_l, ankle_r]; knees inside their range: outside it the
    # reference clamps MovingPathPoint locations but not their derivatives (MovingPathPoint.cpp:216-315), see next test
    q = np.array([-0.2, 0.5, 0.95, 0.3, -0.4, 0.1, -0.25, -0.6, -0.05, 0.1]) + rng.uniform(-0.05, 0.05, 10)
    u = rng.uniform(-2, 2, 10)
    h = 1e-6
    for m in range(o.nmus):
        L, Ldot = o.path(m, q, u)
        Lp, _ = o.path(m, q + h * u, u)
        Lm, _ = o.path(m, q - h * u, u)
        assert Ldot == pytest.approx((Lp - Lm) / (2 * h), rel=1e-6, abs=1e-8)
        tau = o.path_generalized_force(m, q, 1.0)
        dLdq = np.array([(o.path(m, q + h * np.eye(10)[k], u)[0] - o.path(m, q - h * np.eye(10)[k], u)[0]) / (2 * h)
                         for k in range(10)])
        assert np.allclose(tau, -dLdq, rtol=1e-6, atol=1e-8)


def test_moving_path_point_clamp_quirk(orc):
    """beyond the knee range the point location freezes (length stops changing with the knee) while
    the reported speed still uses the unclamped spline derivative."""
    o = orc.Oracle(golden("gait2d_tracking.mb20"))
    q = np.array([-0.2, 0.5, 0.95, 0.3, -0.4, 0.1, 0.30, -0.6, -0.05, 0.1])  # knee_l = 0.30 > 0.1745 (range max)
    u = np.zeros(10)
    u[6] = 1.0
    vasti_l = 14
    L0, Ld = o.path(vasti_l, q, u)
    q2 = q.copy()
    q2[6] += 1e-3
    L1, _ = o.path(vasti_l, q2, u)
    dLdq = (L1 - L0) / 1e-3
    assert abs(Ld - dLdq) > 1e-3  # speed is NOT the derivative of length here, by the reference's construction


def test_contact_force_behaviour(orc):
    """testContact.cpp:309-350 style facts for SmoothSphereHalfSpaceForce: a resting sphere pressed into the plane is pushed
    up along the plane normal; far above the plane only the constant contact force (~1e-5 N) remains; sliding produces
    friction opposing the tangential velocity bounded by mu * normal force."""
    o = orc.Oracle(golden("gait2d_tracking.mb20"))
    q = np.zeros(10)
    u = np.zeros(10)
    q[2] = 2.0  # pelvis 2 m up: no contact
    f, _ = o.contact(0, q, u)
    assert 0 < f[1] < 1e-3 and abs(f[0]) < 1e-3
    q[2] = 0.90  # feet pressed into the ground
    f, p = o.contact(0, q, u)
    assert f[1] > 10 and abs(f[0]) < 1e-6 * f[1] and f[2] == 0
    u[1] = 0.5  # sliding forward well above the transition speed (0.2): friction = (mu_d + mu_v v) fn, backwards
    f, _ = o.contact(0, q, u)
    assert f[0] < 0 and abs(f[0]) == pytest.approx((0.8 + 0.5 * math.sqrt(0.25 + 1e-5)) * f[1] * 0.5 / math.sqrt(0.25 + 1e-5), rel=1e-9)


def test_contact_forces_reproduce_the_reference_grf_file(orc):
    """Golden data produced by the reference itself: example2DWalking's referenceGRF.sto holds the ground-reaction forces
    that Simbody's SmoothSphereHalfSpaceForce gave along referenceCoordinates.sto in a Moco simulation of 2D_gait.osim
    (both files in Moco/Examples/C++/example2DWalking; fixture made by scripts/make_grf_fixture.py). The files carry no
    speeds, so they are taken from a cubic spline through the 50 coordinate rows: the normal (vertical) force, which is set
    by the penetration depth, is reproduced to 0.25 % rms of its 1138 N peak; the friction force depends on the tangential
    velocity through tanh(v / v_t) and is only loosely determined at the rows around heel strike."""
    import json

    from scipy.interpolate import CubicSpline

    d = json.load(open(golden("gait2d_reference_grf.json")))
    t, q, G = np.array(d["time"]), np.array(d["q"]), np.array(d["grf_r_xy_l_xy"])
    u = CubicSpline(t, q)(t, 1)
    o = orc.Oracle(golden("gait2d_tracking.mb20"))
    # 2D_gait.osim order: contactHeel_r, contactHeel_l, contactFront_r, contactFront_l ; force on the sphere (= on the foot)
    F = np.array([[o.contact(c, q[i], u[i])[0][:2] for c in range(4)] for i in range(len(t))])
    right, left = F[:, 0] + F[:, 2], F[:, 1] + F[:, 3]
    for mine, ref in ((right, G[:, 0:2]), (left, G[:, 2:4])):
        peak = np.abs(ref[:, 1]).max()
        dv = mine[:, 1] - ref[:, 1]
        assert np.sqrt(np.mean(dv ** 2)) < 0.005 * peak
        assert np.abs(dv).max() < 0.02 * peak
        # friction: same sign and size away from the impact transients
        dh = mine[:, 0] - ref[:, 0]
        assert np.median(np.abs(dh)) < 0.01 * np.abs(ref[:, 0]).max()
    # mid-stance rows, where the speeds from the spline are good: both components to better than 1 N
    for i in (10, 30, 40, 49):
        assert np.allclose(right[i], G[i, 0:2], atol=1.5, rtol=5e-3)
        assert np.allclose(left[i], G[i, 2:4], atol=1.5, rtol=5e-3)
