// TEST INFRASTRUCTURE ONLY — CPU oracle: the CasOC point functions restated for the CPU.
//
// What the reference computes per grid point through SimTK::State + realizeAcceleration
// (Moco/Moco/MocoCasADiSolver/MocoCasOCProblem.h:203-356), written here from the equations in a
// formulation deliberately different from the CUDA kernels (which use an articulated-body
// recursion on origin-referenced spatial quantities):
//   * kinematics by classical point/frame relations in the ground frame;
//   * Newton–Euler about every body's MASS CENTRE, wrenches summed about the GROUND ORIGIN;
//   * explicit mode: joint-space mass matrix by the composite-rigid-body method + Cholesky solve;
//   * implicit mode: the same Newton–Euler pass with the given accelerations.
// Model data comes in the POD tables of include/mocob200.h (the ABI's own input format).
#pragma once
#include <vector>

#include "../include/mocob200.h"
#include "orc_math.hpp"

namespace orc {

struct Model {
    const mocob200_tables* T = nullptr;
    bool implicit = false;
    int nq = 0, nz = 0, nc = 0, nb = 0, n_auxres = 0, n_integrand = 0, nd = 0, n_in = 0, n_dyn_out = 0, n_out = 0;
    bool prescribed = false;  // PositionMotion: every coordinate is a function of time (no coordinate states, no acceleration variables)
    int nqs = 0;              // coordinate value states among the inputs: nq, or 0 when prescribed (MocoProblemRep.cpp:546-560)
    int nm = 0, n_kc = 0;     // Lagrange multipliers per grid point ; kinematic-constraint equations (3 nm with their derivatives, else nm)
    bool enforceDerivs = false;
    std::vector<int> integrandGoals; // goal index of each integrand slot
    void init(const mocob200_tables* t, bool implicitMode, bool enforceConstraintDerivatives = false) {
        T = t; implicit = implicitMode;
        nq = t->nq; nz = t->nz; nc = t->nc; nb = t->nb; n_auxres = t->nderiv_aux;
        prescribed = t->prescribed != 0;
        nqs = prescribed ? 0 : nq;
        nd = ((implicit && !prescribed) ? nq : 0) + t->nderiv_aux; // CasOCProblem.h:485-493
        // CasOCProblem.h:473-524: one multiplier per holonomic equation; 3 rows per equation when the derivatives are enforced
        nm = t->ncoupler;
        enforceDerivs = nm > 0 && enforceConstraintDerivatives && !prescribed;
        n_kc = prescribed ? 0 : (enforceDerivs ? 3 * nm : nm); // prescribed kinematics are taken to obey the constraints (CasOCProblem.h:511-515)
        n_in = 1 + 2 * nqs + nz + nc + nm + nd;
        integrandGoals.clear();
        for (int g = 0; g < t->ngoal; ++g) if (goalRec(g)[3]) integrandGoals.push_back(g);
        n_integrand = (int)integrandGoals.size();
        n_dyn_out = nq + nz + n_auxres + n_kc;
        n_out = n_dyn_out + n_integrand;
    }
    const double* bodyD(int b) const { return T->dbl + T->o_body_d + b * MOCOB200_BODY_DSTRIDE; }
    const int* bodyI(int b) const { return T->idx + T->o_body_i + b * MOCOB200_BODY_ISTRIDE; }
    const double* ppD(int k) const { return T->dbl + T->o_pp_d + k * MOCOB200_PP_DSTRIDE; }
    const int* ppI(int k) const { return T->idx + T->o_pp_i + k * MOCOB200_PP_ISTRIDE; }
    const double* musD(int m) const { return T->dbl + T->o_mus_d + m * MOCOB200_MUS_DSTRIDE; }
    const int* musI(int m) const { return T->idx + T->o_mus_i + m * MOCOB200_MUS_ISTRIDE; }
    const double* contactD(int c) const { return T->dbl + T->o_contact_d + c * MOCOB200_CONTACT_DSTRIDE; }
    const int* contactI(int c) const { return T->idx + T->o_contact_i + c * MOCOB200_CONTACT_ISTRIDE; }
    const int* goalRec(int g) const { return T->idx + T->o_goal_i + g * MOCOB200_GOAL_ISTRIDE; }
    // custom-joint record of body b (null for the built-in joint kinds)
    int customOf(int b) const { return T->ncustom > 0 ? T->idx[T->o_body_cj + b] : -1; }
    const double* cjD(int cj) const { return T->dbl + T->o_cj_d + cj * MOCOB200_CJ_DSTRIDE; }
    const int* cjI(int cj) const { return T->idx + T->o_cj_i + cj * MOCOB200_CJ_ISTRIDE; }
    const int* xforceI(int f) const { return T->idx + T->o_xforce_i + f * MOCOB200_XFORCE_ISTRIDE; }
    const int* couplerI(int k) const { return T->idx + T->o_cc_i + k * MOCOB200_CC_ISTRIDE; }
    // wrap cylinders and the (muscle, cylinder) pairings
    const double* wrapObjD(int w) const { return T->dbl + T->o_wo_d + w * MOCOB200_WO_DSTRIDE; }
    const int* wrapObjI(int w) const { return T->idx + T->o_wo_i + w * MOCOB200_WO_ISTRIDE; }
    int pathWrapCount(int m) const { return T->idx[T->o_pw_i + 2 * m + 1]; }
    const int* pathWrapI(int m, int k) const { return T->idx + T->o_pw_i + 2 * T->nmus + (T->idx[T->o_pw_i + 2 * m] + k) * MOCOB200_PW_ISTRIDE; }
    const double* couplerD(int k) const { return T->dbl + T->o_cc_d + k * MOCOB200_CC_DSTRIDE; }
};

// ---- scalar functions of the path ------------------------------------------------------------

// SimmSpline value / first derivative (opensim-core/OpenSim/Common/SimmSpline.cpp:511-576, 578-672),
// scaled by the enclosing MultiplierFunction (Common/MultiplierFunction.cpp).
template <class S>
S simmSpline(const Model& M, int id, S arg, bool derivative) {
    const int* rec = M.T->idx + M.T->o_spline_i + id * MOCOB200_SPLINE_ISTRIDE;
    const int n = rec[1];
    const double* X = M.T->dbl + M.T->o_spline_data + rec[0];
    const double *Y = X + n, *B = Y + n, *Cc = B + n, *D = Cc + n;
    const S scale(D[n]);
    const double a = val(arg);
    auto lin = [&](int e) { return derivative ? S(B[e]) : S(Y[e]) + (arg - S(X[e])) * S(B[e]); };
    if (a < X[0]) return scale * lin(0);
    if (a > X[n - 1]) return scale * lin(n - 1);
    if (std::fabs(a - X[0]) <= 0.0000000000002) return scale * (derivative ? S(B[0]) : S(Y[0]));
    if (std::fabs(a - X[n - 1]) <= 0.0000000000002) return scale * (derivative ? S(B[n - 1]) : S(Y[n - 1]));
    int k = 0;
    if (n >= 3) {
        int i = 0, j = n;
        while (true) {
            k = (i + j) / 2;
            if (a < X[k]) j = k;
            else if (a > X[k + 1]) i = k;
            else break;
        }
    }
    const S dx = arg - S(X[k]);
    if (derivative) return scale * (S(B[k]) + dx * (S(2.0) * S(Cc[k]) + S(3.0) * dx * S(D[k])));
    return scale * (S(Y[k]) + dx * (S(B[k]) + dx * (S(Cc[k]) + dx * S(D[k]))));
}

// Tracking reference (a GCVSpline in the reference, MocoStateTrackingGoal.cpp:103-117; here the
// host has already converted it to a piecewise polynomial in (t - break)).
template <class S>
S piecewisePoly(const Model& M, int id, S t) {
    const int* rec = M.T->idx + M.T->o_ppoly_i + id * MOCOB200_PPOLY_ISTRIDE;
    const int np = rec[1], order = rec[2];
    const double* br = M.T->dbl + M.T->o_ppoly_data + rec[0];
    const double* co = br + np + 1;
    const double tv = val(t);
    int k = 0;
    while (k + 1 < np && tv >= br[k + 1]) ++k; // piece containing t (last piece extrapolates)
    if (tv < br[0]) k = 0;
    const S s = t - S(br[k]);
    S v(co[(size_t)k * order + order - 1]);
    for (int j = order - 2; j >= 0; --j) v = v * s + S(co[(size_t)k * order + j]);
    return v;
}

// k-th derivative (k = 1, 2) of the same function, term by term (PositionMotion's calcPrescribedPositionDot / DotDot,
// Components/PositionMotion.cpp:53-75)
template <class S>
S piecewisePolyDeriv(const Model& M, int id, S t, int deriv) {
    const int* rec = M.T->idx + M.T->o_ppoly_i + id * MOCOB200_PPOLY_ISTRIDE;
    const int np = rec[1], order = rec[2];
    const double* br = M.T->dbl + M.T->o_ppoly_data + rec[0];
    const double* co = br + np + 1;
    const double tv = val(t);
    int k = 0;
    while (k + 1 < np && tv >= br[k + 1]) ++k;
    if (tv < br[0]) k = 0;
    const S s = t - S(br[k]);
    S v(0.0), p(1.0); // p = s^(j - deriv)
    for (int j = deriv; j < order; ++j) {
        double f = 1.0;
        for (int i = 0; i < deriv; ++i) f *= (double)(j - i);
        v = v + S(f * co[(size_t)k * order + j]) * p;
        p = p * s;
    }
    return v;
}

// De Groote–Fregly 2016 curves (Moco/Moco/Components/DeGrooteFregly2016Muscle.h:332-474, 720-738, 769-815)
namespace dgf {
constexpr double b11 = 0.8150671134243542, b21 = 1.055033428970575, b31 = 0.162384573599574, b41 = 0.063303448465465;
constexpr double b12 = 0.433004984392647, b22 = 0.716775413397760, b32 = -0.029947116970696, b42 = 0.200356847296188;
constexpr double b13 = 0.1, b23 = 1.0, b33 = 0.353553390593274, b43 = 0.0;
constexpr double kPE = 4.0, c1 = 0.200, c2 = 1.0, c3 = 0.200;
constexpr double d1 = -0.3211346127989808, d2 = -8.149, d3 = -0.374, d4 = 0.8825327733249912;
constexpr double minNormFiberLength = 0.2;

template <class S> S gaussianLike(S x, double b1, double b2, double b3, double b4) {
    const S a = x - S(b2), d = S(b3) + S(b4) * x;
    return S(b1) * exp(S(-0.5) * (a * a) / (d * d));
}
template <class S> S gaussianLikeDeriv(S x, double b1, double b2, double b3, double b4) {
    const S d = S(b3) + S(b4) * x, a = S(b2) - x;
    return (S(b1) * exp(-(a * a) / (S(2.0) * (d * d))) * a * S(b3 + b2 * b4)) / (d * d * d);
}
template <class S> S activeForceLength(S normFiberLength, double widthScale) {
    const S x = (normFiberLength - S(1.0)) / S(widthScale) + S(1.0);
    return gaussianLike(x, b11, b21, b31, b41) + gaussianLike(x, b12, b22, b32, b42) + gaussianLike(x, b13, b23, b33, b43);
}
template <class S> S activeForceLengthDeriv(S normFiberLength, double widthScale) {
    const S x = (normFiberLength - S(1.0)) / S(widthScale) + S(1.0);
    return S(1.0 / widthScale) * (gaussianLikeDeriv(x, b11, b21, b31, b41) + gaussianLikeDeriv(x, b12, b22, b32, b42) +
                                  gaussianLikeDeriv(x, b13, b23, b33, b43));
}
template <class S> S passiveForce(S normFiberLength, double e0) {
    const S offset = exp(S(kPE * (minNormFiberLength - 1.0) / e0));
    const S denom = exp(S(kPE)) - offset;
    return (exp(S(kPE) * (normFiberLength - S(1.0)) / S(e0)) - offset) / denom;
}
template <class S> S passiveForceDeriv(S normFiberLength, double e0) {
    const S offset = exp(S(kPE * (minNormFiberLength - 1.0) / e0));
    return (S(kPE) * exp((S(kPE) * (normFiberLength - S(1.0))) / S(e0))) / (S(e0) * (exp(S(kPE)) - offset));
}
template <class S> S forceVelocity(S normFiberVelocity) {
    const S t = S(d2) * normFiberVelocity + S(d3);
    return S(d1) * log(t + sqrt(t * t + S(1.0))) + S(d4);
}
template <class S> S forceVelocityInverse(S fv) { return (sinh(S(1.0 / d1) * (fv - S(d4))) - S(d3)) / S(d2); }
template <class S> S tendonForce(S normTendonLength, double kT) { return S(c1) * exp(S(kT) * (normTendonLength - S(c2))) - S(c3); }
template <class S> S tendonForceDeriv(S normTendonLength, double kT) { return S(c1 * kT) * exp(S(kT) * (normTendonLength - S(c2))); }
template <class S> S tendonForceInverse(S normTendonForce, double kT) {
    return log(S(1.0 / c1) * (normTendonForce + S(c3))) / S(kT) + S(c2);
}
inline double tendonKT(double strainAtOneNormForce) { return std::log((1.0 + c3) / c1) / (1.0 + strainAtOneNormForce - c2); }

// activation dynamics, DeGrooteFregly2016Muscle.cpp:190-210
template <class S> S activationDerivative(S excitation, S activation, double tauAct, double tauDeact) {
    const S z = S(0.5) + S(1.5) * activation;
    const S tempAct = S(1.0) / (S(tauAct) * z);
    const S tempDeact = z / S(tauDeact);
    const S f = S(0.5) * tanh(S(0.1) * (excitation - activation));
    return (tempAct * (f + S(0.5)) + tempDeact * (-f + S(0.5))) * (excitation - activation);
}
} // namespace dgf

template <class S>
struct MuscleResult {
    S tendonForce, equilibriumResidual, normTendonForceRate, linearizedResidualRate;
    // extra getters, used by the known-answer tests (testMocoActuators.cpp:224-1047)
    S normFiberLength, cosPennation, normFiberVelocity, fiberForce, activeFiberForce, passiveFiberForce,
      fiberStiffness, tendonStiffness, fiberStiffnessAlongTendon, muscleStiffness, fiberVelocity, tendonVelocity,
      fal, fpe, fv, pennationAngle, fiberLength, tendonLength;
};

// The muscle model given path length and lengthening speed
// (calcMuscleLengthInfoHelper / calcFiberVelocityInfoHelper / calcMuscleDynamicsInfoHelper,
//  DeGrooteFregly2016Muscle.cpp:240-426; residual h:639-643; linearized residual rate h:645-661).
template <class S>
MuscleResult<S> muscleModel(const double* p, int flags, S mtLength, S mtVelocity, S activation, S normTendonForce,
                            S normTendonForceRate, bool extras) {
    const double Fmax = p[0], lopt = p[1], lslack = p[2], alpha0 = p[3], vmaxOpt = p[4], damping = p[7], e0 = p[9],
                 wscale = p[10];
    const double kT = dgf::tendonKT(p[8]);
    const bool rigid = (flags & MOCOB200_MUS_FLAG_IGNORE_TENDON_COMPLIANCE) != 0;
    const bool explicitTendon = !rigid && !(flags & MOCOB200_MUS_FLAG_TENDON_IMPLICIT);
    const bool noPassive = (flags & MOCOB200_MUS_FLAG_IGNORE_PASSIVE) != 0;
    const double fiberWidth = lopt * std::sin(alpha0);
    const double vmax = vmaxOpt * lopt;
    MuscleResult<S> r;
    // length info
    const S normTendonLength = rigid ? S(1.0) : dgf::tendonForceInverse(normTendonForce, kT);
    const S tendonLength = S(lslack) * normTendonLength;
    const S along = mtLength - tendonLength;
    const S fiberLength = sqrt(along * along + S(fiberWidth * fiberWidth));
    const S nfl = fiberLength / S(lopt);
    const S cosA = along / fiberLength;
    const S sinA = S(fiberWidth) / fiberLength;
    const S fpe = noPassive ? S(0.0) : dgf::passiveForce(nfl, e0);
    const S fal = dgf::activeForceLength(nfl, wscale);
    // velocity info
    S nfv, fv, fiberVelAlong, tendonVel, fiberVel;
    if (explicitTendon) {
        fv = (normTendonForce / cosA - fpe) / (activation * fal);
        nfv = dgf::forceVelocityInverse(fv);
        fiberVel = nfv * S(vmax);
        fiberVelAlong = fiberVel / cosA;
        tendonVel = mtVelocity - fiberVelAlong;
    } else {
        const S normTendonVel = rigid ? S(0.0) : normTendonForceRate / dgf::tendonForceDeriv(normTendonLength, kT);
        tendonVel = S(lslack) * normTendonVel;
        fiberVelAlong = mtVelocity - tendonVel;
        fiberVel = fiberVelAlong * cosA;
        nfv = fiberVel / S(vmax);
        fv = dgf::forceVelocity(nfv);
    }
    // dynamics info
    const S activeF = S(Fmax) * (activation * fal * fv);
    const S conPassive = S(Fmax) * fpe;
    const S nonConPassive = S(Fmax) * S(damping) * nfv;
    const S fiberForce = activeF + conPassive + nonConPassive;
    const S fiberForceAlong = fiberForce * cosA;
    r.tendonForce = rigid ? fiberForceAlong : S(Fmax) * normTendonForce;
    r.equilibriumResidual = rigid ? S(0.0) : r.tendonForce - fiberForceAlong;
    if (rigid) r.normTendonForceRate = S(0.0);
    else if (explicitTendon) r.normTendonForceRate = (tendonVel / S(lslack)) * dgf::tendonForceDeriv(normTendonLength, kT);
    else r.normTendonForceRate = normTendonForceRate;
    r.linearizedResidualRate = S(0.0);
    if (extras) {
        // stiffnesses (DeGrooteFregly2016Muscle.h:506-618)
        const S dfal = S(1.0 / lopt) * dgf::activeForceLengthDeriv(nfl, wscale);
        const S dfpe = noPassive ? S(0.0) : S(1.0 / lopt) * dgf::passiveForceDeriv(nfl, e0);
        const S fiberStiffness = S(Fmax) * (activation * dfal * fv + dfpe);
        const S ratio = S(fiberWidth) / fiberLength;
        const S dAlpha_dL = -ratio / fiberLength / sqrt(S(1.0) - ratio * ratio);
        const S dFalong_dL = fiberStiffness * cosA - fiberForce * sinA * dAlpha_dL;
        const S dLalong_dL = cosA - fiberLength * sinA * dAlpha_dL;
        const S kFiberAlong = dFalong_dL * (S(1.0) / dLalong_dL);
        const S kTendon = rigid ? S(INFINITY) : S(Fmax / lslack) * dgf::tendonForceDeriv(normTendonLength, kT);
        // getLinearizedEquilibriumResidualDerivative: d/dt of (tendon force - fiber force along tendon)
        // linearized: fiberStiffnessAlongTendon * fiberVelocityAlongTendon
        //             - tendonStiffness * (muscleTendonVelocity - fiberVelocityAlongTendon)
        r.linearizedResidualRate = kFiberAlong * fiberVelAlong - kTendon * (mtVelocity - fiberVelAlong);
        r.fiberStiffness = fiberStiffness; r.tendonStiffness = kTendon; r.fiberStiffnessAlongTendon = kFiberAlong;
        r.muscleStiffness = rigid ? kFiberAlong : (kFiberAlong * kTendon) / (kFiberAlong + kTendon);
    }
    r.normFiberLength = nfl; r.cosPennation = cosA; r.normFiberVelocity = nfv; r.fiberForce = fiberForce;
    r.activeFiberForce = activeF; r.passiveFiberForce = conPassive + nonConPassive; r.fiberVelocity = fiberVel;
    r.tendonVelocity = tendonVel; r.fal = fal; r.fpe = fpe; r.fv = fv; r.pennationAngle = asin(sinA);
    r.fiberLength = fiberLength; r.tendonLength = tendonLength;
    return r;
}

} // namespace orc
