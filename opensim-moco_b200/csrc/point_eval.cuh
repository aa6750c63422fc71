// Device-side point function of the direct-collocation hot path (FP64, sm_100a).
//
// One call = one evaluation of the fused CasOC point functions at one grid point:
//   (time, states, controls, derivatives) -> [udot | residual, zdot, aux residuals, integrands...]
// i.e. what MultibodySystemExplicit/Implicit::eval + every CostIntegrand::eval compute in the
// reference through SimTK::State + realizeAcceleration (MocoCasOCProblem.h:203-356).
//
// Formulation (not Simbody's code; written from the equations): ground-frame spatial quantities
// referenced to each body's origin, spatial vectors [angular; linear]; articulated-body forward
// dynamics (explicit mode) or recursive Newton–Euler residual (implicit mode); straight-line
// muscle paths with conditional / moving points; De Groote–Fregly 2016 muscle curves; smooth
// sphere–half-space contact; goal integrands.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#include "../../include/mocob200.h"

namespace mocob200 {

#ifndef MOCOB200_MAXB
#define MOCOB200_MAXB 16 /* mobilized bodies */
#endif
#ifndef MOCOB200_MAXQ
#define MOCOB200_MAXQ 16 /* coordinates */
#endif
#define MOCOB200_MAXCONTACT 8
#define MOCOB200_MAXPPM 12 /* path points per muscle */
// The thread-per-evaluation code keeps its scratch in thread-local memory, whose traffic bounds it: every array
// is sized by a size class chosen at launch, so that a 8-body / 58-input model does not carry 16-body / 160-input frames.
// WRAP: the muscle code of this instantiation evaluates wrap cylinders (kept out of the FD kernels' instantiations: the mere
// presence of the call costs the config-5 kernels 10 % through register allocation; models with wrap cylinders do not run there)
struct SizeLarge { static constexpr int NB = MOCOB200_MAXB, NQ = MOCOB200_MAXQ, NPPM = MOCOB200_MAXPPM, NIN = 160, NOUT = 72; static constexpr bool WRAP = false; };
struct SizeSmall { static constexpr int NB = 8, NQ = 12, NPPM = 8, NIN = 64, NOUT = 32; static constexpr bool WRAP = false; };
// full-body models (Rajagopal: 22 bodies, 39 coordinates): kinematics / path kernels only; their dynamics run on the warp kernels
struct SizeHuge { static constexpr int NB = 24, NQ = 40, NPPM = MOCOB200_MAXPPM, NIN = 240, NOUT = 72; static constexpr bool WRAP = true; };

struct DevTables { // mocob200_tables with device (or shared-memory) pool pointers
    const double* dbl;
    const int* idx;
    // The warp kernels stage the model part of the pools in shared memory; the (large) tracking-reference data stay in
    // global memory behind dblP, and the goal payload is addressed through dblG (same absolute offsets as in dbl).
    const double* dblP;
    const double* dblG;
    int nb, nq, nz, nc, nderiv_aux, nmus, npp, nspline, ncontact, ncact, ngoal, nppoly;
    int o_body_d, o_gravity, o_pp_d, o_mus_d, o_contact_d, o_cact_d, o_spline_data, o_ppoly_data;
    int o_body_i, o_pp_i, o_mus_i, o_contact_i, o_cact_i, o_spline_i, o_goal_i, o_ppoly_i;
    int ncustom, o_cj_d, o_cj_i, o_body_cj; // custom (function-based) joints
    int nxforce, o_xforce_i;                // external forces (functions of time)
    int ncoupler, o_cc_i, o_cc_d;           // coordinate-coupler kinematic constraints
    int nm, n_kc, enforce_derivs;           // multipliers per point, kinematic-constraint outputs, their derivatives enforced
    int implicit; // multibody dynamics mode
    int prescribed, o_presc_i; // prescribed kinematics: idx[o_presc_i + k] = piecewise polynomial of coordinate k
    int nwrapobj, o_wo_d, o_wo_i, npathwrap, o_pw_i; // wrap cylinders and (muscle, cylinder) pairings
    int n_auxres;
    int n_integrand; // goals with an integrand, in goal order
};

struct V3 {
    double x, y, z;
};
__device__ __forceinline__ V3 mk(double x, double y, double z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(double s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 ld3(const double* p) { return mk(p[0], p[1], p[2]); }
// R row-major 3x3
__device__ __forceinline__ V3 rmul(const double* R, V3 v) {
    return mk(R[0] * v.x + R[1] * v.y + R[2] * v.z, R[3] * v.x + R[4] * v.y + R[5] * v.z, R[6] * v.x + R[7] * v.y + R[8] * v.z);
}
__device__ __forceinline__ V3 rtmul(const double* R, V3 v) {
    return mk(R[0] * v.x + R[3] * v.y + R[6] * v.z, R[1] * v.x + R[4] * v.y + R[7] * v.z, R[2] * v.x + R[5] * v.y + R[8] * v.z);
}
} // namespace mocob200
#include "point_wrap.cuh"
namespace mocob200 {
__device__ __forceinline__ void mm3(const double* A, const double* B, double* C) { // C = A B
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
__device__ __forceinline__ void mmt3(const double* A, const double* B, double* C) { // C = A B^T
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[3 * j] + A[3 * i + 1] * B[3 * j + 1] + A[3 * i + 2] * B[3 * j + 2];
}

// ---- SimmSpline evaluation (value / first derivative), semantics of
// opensim-core/OpenSim/Common/SimmSpline.cpp:511-685: linear extrapolation outside the knots,
// end-knot snapping within 2e-13, interval search, cubic in (x - x_k).
__device__ inline double simm_spline(const DevTables& T, int id, double ax, int deriv) {
    const int* rec = T.idx + T.o_spline_i + id * MOCOB200_SPLINE_ISTRIDE;
    const int n = rec[1];
    const double* x = T.dbl + T.o_spline_data + rec[0];
    const double* y = x + n;
    const double* b = y + n;
    const double* c = b + n;
    const double* d = c + n;
    const double scale = d[n];
    if (deriv == 2) { // second derivative (SimmSpline.cpp:578-685): zero outside the knots, 2 c at the end knots
        if (ax < x[0] || ax > x[n - 1]) return 0.0;
        if (fabs(ax - x[0]) <= 2e-13) return scale * 2.0 * c[0];
        if (fabs(ax - x[n - 1]) <= 2e-13) return scale * 2.0 * c[n - 1];
    } else {
    if (ax < x[0]) return scale * (deriv ? b[0] : y[0] + (ax - x[0]) * b[0]);
    if (ax > x[n - 1]) return scale * (deriv ? b[n - 1] : y[n - 1] + (ax - x[n - 1]) * b[n - 1]);
    if (fabs(ax - x[0]) <= 2e-13) return scale * (deriv ? b[0] : y[0]);
    if (fabs(ax - x[n - 1]) <= 2e-13) return scale * (deriv ? b[n - 1] : y[n - 1]);
    }
    int k = 0;
    if (n >= 3) { // x[k] <= ax <= x[k+1]: interpolation guess, then walk (the knots are sorted)
        k = (int)((ax - x[0]) / (x[n - 1] - x[0]) * (double)(n - 1));
        k = max(0, min(k, n - 2));
        while (k > 0 && ax < x[k]) --k;
        while (k < n - 2 && ax > x[k + 1]) ++k;
    }
    const double dx = ax - x[k];
    if (deriv == 2) return scale * (2.0 * c[k] + 6.0 * dx * d[k]);
    if (deriv) return scale * (b[k] + dx * (2.0 * c[k] + 3.0 * dx * d[k]));
    return scale * (y[k] + dx * (b[k] + dx * (c[k] + dx * d[k])));
}

// ---- custom (function-based) joint: SimbodyEngine/CustomJoint.cpp:205-273 -> SimTK::MobilizedBody::FunctionBased.
// X_FM(q) = [ R(A0, f0) R(A1, f1) R(A2, f2) , A3 f3 + A4 f4 + A5 f5 ], u = qdot.
// Everything below is written from that definition: the joint's spatial velocity in F is
//   omega = th0' A0 + th1' R0 A1 + th2' R0 R1 A2 ,  v = sum_i Ai fi'          (thi' = dfi/dq u)
// so the motion axis of coordinate k collects the terms whose function reads q_k, and the velocity-product acceleration
// (what remains of d/dt(omega, v) at zero udot) is
//   b_omega = w0 x w1 + (w0 + w1) x w2 + sum_i Wi fi'' u^2 ,  b_v = sum_i Ai fi'' u^2     (wi = thi' Wi)
// - except that one-coordinate joints follow the reference in dropping w1 x w2 (see custom_joint_bias).
struct CustomJointEval {
    double R[9];     // R_FM
    V3 p;            // p_FM (in F)
    V3 g[6];         // W0, W1, W2 (rotation axes carried by the preceding rotations), A3, A4, A5 — all in F
    double df[6], d2f[6];
    int qi[6];       // coordinate driving axis i, or -1
};
__device__ inline void rot_axis(double angle, V3 a, double* R) { // cos 1 + (1 - cos) a a^T + sin [a]x, row-major
    double s, c;
    sincos(angle, &s, &c);
    const double t = 1.0 - c;
    R[0] = t * a.x * a.x + c;       R[1] = t * a.x * a.y - s * a.z; R[2] = t * a.x * a.z + s * a.y;
    R[3] = t * a.y * a.x + s * a.z; R[4] = t * a.y * a.y + c;       R[5] = t * a.y * a.z - s * a.x;
    R[6] = t * a.z * a.x - s * a.y; R[7] = t * a.z * a.y + s * a.x; R[8] = t * a.z * a.z + c;
}
__device__ inline void custom_joint_eval(const DevTables& T, int cj, const double* q, CustomJointEval& E) {
    const double* d = T.dbl + T.o_cj_d + cj * MOCOB200_CJ_DSTRIDE;
    const int* rec = T.idx + T.o_cj_i + cj * MOCOB200_CJ_ISTRIDE;
    double f[6];
    for (int a = 0; a < 6; ++a) {
        const int kind = rec[3 * a], qi = rec[3 * a + 1];
        E.qi[a] = kind == MOCOB200_FN_CONSTANT ? -1 : qi;
        if (kind == MOCOB200_FN_LINEAR) { f[a] = d[5 * a + 3] * q[qi] + d[5 * a + 4]; E.df[a] = d[5 * a + 3]; E.d2f[a] = 0.0; }
        else if (kind == MOCOB200_FN_SIMMSPLINE) {
            f[a] = simm_spline(T, rec[3 * a + 2], q[qi], 0);
            E.df[a] = simm_spline(T, rec[3 * a + 2], q[qi], 1);
            E.d2f[a] = simm_spline(T, rec[3 * a + 2], q[qi], 2);
        } else { f[a] = d[5 * a + 4]; E.df[a] = 0.0; E.d2f[a] = 0.0; }
    }
    double R0[9], R1[9], R2[9], R01[9];
    rot_axis(f[0], ld3(d), R0);
    rot_axis(f[1], ld3(d + 5), R1);
    rot_axis(f[2], ld3(d + 10), R2);
    mm3(R0, R1, R01);
    mm3(R01, R2, E.R);
    E.g[0] = ld3(d);
    E.g[1] = rmul(R0, ld3(d + 5));
    E.g[2] = rmul(R01, ld3(d + 10));
    E.p = mk(0, 0, 0);
    for (int a = 3; a < 6; ++a) { E.g[a] = ld3(d + 5 * a); E.p = E.p + f[a] * E.g[a]; }
}
// One-coordinate joints (every axis function reads the same coordinate, e.g. a knee whose secondary rotations are splines
// of the flexion angle): the reference's mobilizer (SimTK::MobilizedBody::FunctionBased, Simbody 3.7 - a dependency whose
// source is not under /root/reference) leaves w1 x w2 out of the velocity-product term. Pinned on the reference's own RRA
// and CMC records of the Rajagopal model (tests/test_reference_dynamics_pins.py: with the term 0.8 N m off at the hip, without
// it 1e-4 N m); joints whose rotation axes have a coordinate each (hips, shoulders) match those records with the full
// expression, and no record here covers other sharing patterns, which keep it too.
__device__ inline bool cj_one_coordinate(const int* rec) { // rec: 6 x { function kind | q index | spline id }
    int q = -1;
    for (int i = 0; i < 6; ++i) {
        if (rec[3 * i] == MOCOB200_FN_CONSTANT) continue;
        if (q >= 0 && rec[3 * i + 1] != q) return false;
        q = rec[3 * i + 1];
    }
    return q >= 0;
}
// velocity-product acceleration of the joint (in F) for the speeds u
__device__ inline void custom_joint_bias(const CustomJointEval& E, const double* u, V3& bw, V3& bv) {
    V3 w[3];
    bw = mk(0, 0, 0);
    bv = mk(0, 0, 0);
    for (int i = 0; i < 3; ++i) {
        const double ui = E.qi[i] >= 0 ? u[E.qi[i]] : 0.0;
        w[i] = (E.df[i] * ui) * E.g[i];
        bw = bw + (E.d2f[i] * ui * ui) * E.g[i];
    }
    bool one = true; // E.qi is -1 for constant functions
    for (int i = 0, q = -1; i < 6; ++i) if (E.qi[i] >= 0) { one = one && (q < 0 || q == E.qi[i]); q = E.qi[i]; }
    bw = bw + cross(w[0], w[1]) + cross(one ? w[0] : w[0] + w[1], w[2]);
    for (int i = 3; i < 6; ++i) {
        const double ui = E.qi[i] >= 0 ? u[E.qi[i]] : 0.0;
        bv = bv + (E.d2f[i] * ui * ui) * E.g[i];
    }
}
// motion axis of the joint's coordinate qk (absolute index) in F: angular a, linear l (acting at the M origin)
__device__ inline void custom_joint_axis(const CustomJointEval& E, int qk, V3& a, V3& l) {
    a = mk(0, 0, 0);
    l = mk(0, 0, 0);
    for (int i = 0; i < 3; ++i) if (E.qi[i] == qk) a = a + E.df[i] * E.g[i];
    for (int i = 3; i < 6; ++i) if (E.qi[i] == qk) l = l + E.df[i] * E.g[i];
}

// ---- piecewise polynomial (tracking references; natural spline fitted on the host)
__device__ inline double ppoly_eval(const DevTables& T, int id, double t) {
    const int* rec = T.idx + T.o_ppoly_i + id * MOCOB200_PPOLY_ISTRIDE;
    const int np = rec[1], order = rec[2];
    const double* br = T.dblP + T.o_ppoly_data + rec[0];
    const double* co = br + np + 1;
    int k = 0;
    if (t >= br[np]) k = np - 1;
    else if (t > br[0]) { // br[k] <= t < br[k+1]: interpolation guess, then walk
        k = (int)((t - br[0]) / (br[np] - br[0]) * (double)np);
        k = max(0, min(k, np - 1));
        while (k > 0 && t < br[k]) --k;
        while (k < np - 1 && t >= br[k + 1]) ++k;
    }
    const double s = t - br[k];
    const double* c = co + (size_t)k * order;
    double v = c[order - 1];
    for (int j = order - 2; j >= 0; --j) v = v * s + c[j];
    return v;
}

// value and first two derivatives (prescribed kinematics: Function::calcValue / calcDerivative({0}) / ({0,0}),
// Components/PositionMotion.cpp:44-75)
__device__ inline void ppoly_eval2(const DevTables& T, int id, double t, double& v, double& d1, double& d2) {
    const int* rec = T.idx + T.o_ppoly_i + id * MOCOB200_PPOLY_ISTRIDE;
    const int np = rec[1], order = rec[2];
    const double* br = T.dblP + T.o_ppoly_data + rec[0];
    const double* co = br + np + 1;
    int k = 0;
    if (t >= br[np]) k = np - 1;
    else if (t > br[0]) {
        k = (int)((t - br[0]) / (br[np] - br[0]) * (double)np);
        k = max(0, min(k, np - 1));
        while (k > 0 && t < br[k]) --k;
        while (k < np - 1 && t >= br[k + 1]) ++k;
    }
    const double s = t - br[k];
    const double* c = co + (size_t)k * order;
    v = 0; d1 = 0; d2 = 0;
    for (int j = order - 1; j >= 0; --j) { // Horner on the value and both derivatives
        d2 = d2 * s + 2.0 * d1;
        d1 = d1 * s + v;
        v = v * s + c[j];
    }
}

// The inputs of one point-function evaluation, in = [time, q(nqs), u(nqs), z(nz), controls(nc), multipliers(nm), udot, auxder]:
// with prescribed kinematics (nqs = 0, no udot inputs) q, u and udot are the coordinates' functions of time, evaluated into pk.
struct PointInputs {
    const double *q, *u, *z, *ctl, *mult, *udot, *auxder;
};
__device__ inline PointInputs point_inputs(const DevTables& T, const double* in, double* pk /* [3 nq], used when prescribed */) {
    PointInputs P;
    const int nq = T.nq, nqs = T.prescribed ? 0 : nq;
    P.q = in + 1;
    P.u = P.q + nqs;
    P.z = P.u + nqs;
    P.ctl = P.z + T.nz;
    P.mult = P.ctl + T.nc;
    const double* der = P.mult + T.nm;
    P.udot = (T.implicit && !T.prescribed) ? der : nullptr;
    P.auxder = der + ((T.implicit && !T.prescribed) ? nq : 0);
    if (T.prescribed) {
        for (int k = 0; k < nq; ++k) ppoly_eval2(T, T.idx[T.o_presc_i + k], in[0], pk[k], pk[nq + k], pk[2 * nq + k]);
        P.q = pk; P.u = pk + nq; P.udot = pk + 2 * nq;
    }
    return P;
}

// ---- De Groote–Fregly 2016 curves (Components/DeGrooteFregly2016Muscle.h:332-474, constants :769-815)
__device__ __forceinline__ double dgf_gauss(double x, double b1, double b2, double b3, double b4) {
    const double num = x - b2, den = b3 + b4 * x;
    return b1 * exp(-0.5 * (num * num) / (den * den));
}
__device__ __forceinline__ double dgf_active_fl(double nfl, double scale) {
    const double x = (nfl - 1.0) / scale + 1.0;
    return dgf_gauss(x, 0.8150671134243542, 1.055033428970575, 0.162384573599574, 0.063303448465465) +
           dgf_gauss(x, 0.433004984392647, 0.716775413397760, -0.029947116970696, 0.200356847296188) +
           dgf_gauss(x, 0.1, 1.0, 0.353553390593274, 0.0);
}
__device__ __forceinline__ double dgf_gauss_d(double x, double b1, double b2, double b3, double b4) {
    const double den = b3 + b4 * x;
    return (b1 * exp(-((b2 - x) * (b2 - x)) / (2.0 * den * den)) * (b2 - x) * (b3 + b2 * b4)) / (den * den * den);
}
__device__ __forceinline__ double dgf_active_fl_d(double nfl, double scale) {
    const double x = (nfl - 1.0) / scale + 1.0;
    return (1.0 / scale) * (dgf_gauss_d(x, 0.8150671134243542, 1.055033428970575, 0.162384573599574, 0.063303448465465) +
                            dgf_gauss_d(x, 0.433004984392647, 0.716775413397760, -0.029947116970696, 0.200356847296188) +
                            dgf_gauss_d(x, 0.1, 1.0, 0.353553390593274, 0.0));
}
// offset = exp(kPE (minNormFiberLength - 1) / e0) and denom = exp(kPE) - offset depend on the muscle's parameters only: the device
// copy of the muscle record carries them (mocob200_create, DEV_MUS_* slots below)
__device__ __forceinline__ double dgf_passive_fl(double nfl, double e0, double offset, double denom) {
    const double kPE = 4.0;
    return (exp(kPE * (nfl - 1.0) / e0) - offset) / denom;
}
__device__ __forceinline__ double dgf_passive_fl_d(double nfl, double e0, double denom) {
    const double kPE = 4.0;
    return (kPE * exp((kPE * (nfl - 1.0)) / e0)) / (e0 * denom);
}
__device__ __forceinline__ double dgf_fv(double nv) {
    const double d1 = -0.3211346127989808, d2 = -8.149, d3 = -0.374, d4 = 0.8825327733249912;
    const double tv = d2 * nv + d3;
    return d1 * log(tv + sqrt(tv * tv + 1.0)) + d4;
}
__device__ __forceinline__ double dgf_fv_inv(double fv) {
    const double d1 = -0.3211346127989808, d2 = -8.149, d3 = -0.374, d4 = 0.8825327733249912;
    return (sinh(1.0 / d1 * (fv - d4)) - d3) / d2;
}

// Slots of the DEVICE copy of a muscle record that hold constants derived at mocob200_create from the record's own parameters
// (they replace fields only the host reads: tendon strain -> kT is already in slot 13, the control bounds live in the bounds table).
#define DEV_MUS_PE_DENOM 8     /* exp(kPE) - offset                        (was tendon_strain_at_one_norm_force) */
#define DEV_MUS_FIBER_WIDTH 11 /* optimal_fiber_length sin(pennation)      (was min_control) */
#define DEV_MUS_PE_OFFSET 12   /* exp(kPE (0.2 - 1) / passive_fiber_strain) (was max_control) */
#define DEV_CONTACT_HERTZ 25   /* (4/3) k sqrt(radius k), k = stiffness^(2/3) / 2 (a pad slot of the contact record) */
// |c|^n for the integer exponents of MocoControlGoal (n >= 2)
__device__ __forceinline__ double abs_ipow(double c, int n) {
    const double a = fabs(c);
    double r = a * a;
    for (int k = 2; k < n; ++k) r *= a;
    return r;
}

// CoordinateCouplerConstraint k: position error scale f(q_i) - q_d and the first two derivatives of scale f at q_i
// (SimbodyEngine/CoordinateCouplerConstraint.cpp:40-90 CompoundFunction -> SimTK::Constraint::CoordinateCoupler)
__device__ inline void coupler_function(const DevTables& T, int k, const double* q, double& err, double& d1, double& d2) {
    const int* ci = T.idx + T.o_cc_i + k * MOCOB200_CC_ISTRIDE;
    const double* cd = T.dbl + T.o_cc_d + k * MOCOB200_CC_DSTRIDE;
    const double qi = q[ci[1]];
    double f;
    if (ci[0] == MOCOB200_FN_LINEAR) { f = cd[0] * qi + cd[1]; d1 = cd[0]; d2 = 0.0; }
    else if (ci[0] == MOCOB200_FN_SIMMSPLINE) { f = simm_spline(T, ci[3], qi, 0); d1 = simm_spline(T, ci[3], qi, 1); d2 = simm_spline(T, ci[3], qi, 2); }
    else { f = cd[1]; d1 = 0.0; d2 = 0.0; }
    err = f - q[ci[2]];
}
// constraint forces of the solver's multipliers as applied mobility forces: G^T (-lambda) (MocoCasOCProblem.h:643-662)
__device__ inline void coupler_forces(const DevTables& T, const double* q, const double* mult, double* tau) {
    for (int k = 0; k < T.ncoupler; ++k) {
        double e, d1, d2;
        coupler_function(T, k, q, e, d1, d2);
        const int* ci = T.idx + T.o_cc_i + k * MOCOB200_CC_ISTRIDE;
        tau[ci[1]] -= d1 * mult[k];
        tau[ci[2]] += mult[k];
    }
}
// kinematic-constraint errors (MocoCasOCProblem.h:664-732): position level; with enforced derivatives also G u and
// G udot + Gdot u, udot being this model's (explicit: just solved; implicit: the acceleration variables)
__device__ inline void coupler_errors(const DevTables& T, int k, const double* q, const double* u, const double* udot, double* kc) {
    double e, d1, d2;
    coupler_function(T, k, q, e, d1, d2);
    const int* ci = T.idx + T.o_cc_i + k * MOCOB200_CC_ISTRIDE;
    kc[k] = e;
    if (T.enforce_derivs) {
        kc[T.nm + k] = d1 * u[ci[1]] - u[ci[2]];
        kc[2 * T.nm + k] = d1 * udot[ci[1]] + d2 * u[ci[1]] * u[ci[1]] - udot[ci[2]];
    }
}
// velocity correction of a Hermite-Simpson midpoint (MocoCasOCProblem.h:298-332): corr = G(q)^T gamma
__device__ inline double coupler_velocity_correction(const DevTables& T, const double* q, const double* gamma, int s) {
    double c = 0.0;
    for (int k = 0; k < T.ncoupler; ++k) {
        const int* ci = T.idx + T.o_cc_i + k * MOCOB200_CC_ISTRIDE;
        if (ci[1] == s) { double e, d1, d2; coupler_function(T, k, q, e, d1, d2); c += d1 * gamma[k]; }
        if (ci[2] == s) c -= gamma[k];
    }
    return c;
}

struct MuscleOut {
    double tendonForce;  // actuation
    double residual;     // implicit tendon: F_T - F_fiber cos(alpha)
    double nfDeriv;      // d(normalized tendon force)/dt
    double linResidDeriv; // getLinearizedEquilibriumResidualDerivative (for InitialVelocityEquilibriumDGF)
};

// DeGrooteFregly2016Muscle::calcMuscleLengthInfoHelper / calcFiberVelocityInfoHelper /
// calcMuscleDynamicsInfoHelper (Components/DeGrooteFregly2016Muscle.cpp:240-372) reduced to what
// reaches the outputs; wantStiffness adds the stiffness terms needed by the linearized residual.
__device__ inline MuscleOut dgf_muscle(const double* p, int flags, double mtLength, double mtVelocity, double activation,
                                      double normTendonForce, double normTendonForceDeriv, bool wantStiffness) {
    const double Fmax = p[0], lopt = p[1], lslack = p[2], vmaxN = p[4], damping = p[7], e0 = p[9], wscale = p[10], kT = p[13];
    const double c1 = 0.200, c2 = 1.0, c3 = 0.200;
    const bool rigid = flags & MOCOB200_MUS_FLAG_IGNORE_TENDON_COMPLIANCE;
    const bool implicitT = flags & MOCOB200_MUS_FLAG_TENDON_IMPLICIT;
    MuscleOut o;
    if (flags & MOCOB200_MUS_FLAG_PATH_ACTUATOR) { // PathActuator::computeActuation: optimal_force x control (PathActuator.cpp:156-165)
        o.tendonForce = Fmax * activation;
        o.residual = 0.0; o.nfDeriv = 0.0; o.linResidDeriv = 0.0;
        return o;
    }
    const double fiberWidth = p[DEV_MUS_FIBER_WIDTH]; // lopt * sin(alpha0)
    const double peOffset = p[DEV_MUS_PE_OFFSET], peDenom = p[DEV_MUS_PE_DENOM];
    const double vmax = vmaxN * lopt;
    // length info
    const double normTendonLength = rigid ? 1.0 : log((1.0 / c1) * (normTendonForce + c3)) / kT + c2;
    const double tendonLength = lslack * normTendonLength;
    const double fibAlong = mtLength - tendonLength;
    const double fiberLength = sqrt(fibAlong * fibAlong + fiberWidth * fiberWidth);
    const double nfl = fiberLength / lopt;
    const double cosA = fibAlong / fiberLength;
    const double sinA = fiberWidth / fiberLength;
    const double fpe = (flags & MOCOB200_MUS_FLAG_IGNORE_PASSIVE) ? 0.0 : dgf_passive_fl(nfl, e0, peOffset, peDenom);
    const double fal = dgf_active_fl(nfl, wscale);
    // velocity info
    double normFiberVelocity, fv, fiberVelAlong, tendonVelocity;
    if (!rigid && !implicitT) {
        const double normFiberForce = normTendonForce / cosA;
        fv = (normFiberForce - fpe) / (activation * fal);
        normFiberVelocity = dgf_fv_inv(fv);
        const double fiberVelocity = normFiberVelocity * vmax;
        fiberVelAlong = fiberVelocity / cosA;
        tendonVelocity = mtVelocity - fiberVelAlong;
    } else {
        const double normTendonVelocity =
                rigid ? 0.0 : normTendonForceDeriv / (c1 * kT * exp(kT * (normTendonLength - c2)));
        tendonVelocity = lslack * normTendonVelocity;
        fiberVelAlong = mtVelocity - tendonVelocity;
        const double fiberVelocity = fiberVelAlong * cosA;
        normFiberVelocity = fiberVelocity / vmax;
        fv = dgf_fv(normFiberVelocity);
    }
    // dynamics info
    const double activeF = Fmax * (activation * fal * fv);
    const double conPassiveF = Fmax * fpe;
    const double nonConPassiveF = Fmax * damping * normFiberVelocity;
    const double fiberForce = activeF + conPassiveF + nonConPassiveF;
    const double fiberForceAlong = fiberForce * cosA;
    o.tendonForce = rigid ? fiberForceAlong : Fmax * normTendonForce;
    o.residual = rigid ? 0.0 : (implicitT ? o.tendonForce - fiberForceAlong : NAN);
    if (rigid) o.nfDeriv = 0.0;
    else if (implicitT) o.nfDeriv = normTendonForceDeriv;
    else o.nfDeriv = (tendonVelocity / lslack) * (c1 * kT * exp(kT * (normTendonLength - c2)));
    o.linResidDeriv = 0.0;
    if (wantStiffness) {
        // calcFiberStiffness, calcPartialPennationAnglePartialFiberLength,
        // calcPartialFiberForceAlongTendonPartialFiberLength, calcFiberStiffnessAlongTendon,
        // calcTendonStiffness (DeGrooteFregly2016Muscle.h:506-618)
        const double fiberStiffness = Fmax * (activation * (1.0 / lopt) * dgf_active_fl_d(nfl, wscale) * fv +
                                              ((flags & MOCOB200_MUS_FLAG_IGNORE_PASSIVE) ? 0.0 : (1.0 / lopt) * dgf_passive_fl_d(nfl, e0, peDenom)));
        const double dAlpha_dL = -fiberWidth / (fiberLength * fiberLength) / sqrt(1.0 - (fiberWidth / fiberLength) * (fiberWidth / fiberLength));
        const double dFalong_dL = fiberStiffness * cosA - fiberForce * sinA * dAlpha_dL;
        const double dLalong_dL = cosA - fiberLength * sinA * dAlpha_dL;
        const double fiberStiffAlong = dFalong_dL * (1.0 / dLalong_dL);
        const double tendonStiffness = rigid ? INFINITY : (Fmax / lslack) * (c1 * kT * exp(kT * (normTendonLength - c2)));
        o.linResidDeriv = fiberStiffAlong * fiberVelAlong - tendonStiffness * (mtVelocity - fiberVelAlong);
    }
    return o;
}

// Scratch for one evaluation. Lives in thread-local memory.
template <class SZ>
struct ScratchT {
    double R[SZ::NB][9]; // R_GB
    V3 p[SZ::NB];        // body origin in ground
    V3 w[SZ::NB], v[SZ::NB]; // spatial velocity of the body origin
    V3 Fm[SZ::NB], Ff[SZ::NB]; // applied spatial force (moment about origin ; force)
    double tau[SZ::NQ];  // applied mobility forces
    V3 contactF[MOCOB200_MAXCONTACT]; // force on the sphere body of each contact (ground frame)
};
typedef ScratchT<SizeLarge> Scratch;

// position + velocity kinematics, base to tip.
template <class SZ>
__device__ inline void kinematics(const DevTables& T, const double* q, const double* u, ScratchT<SZ>& S) {
    for (int b = 0; b < T.nb; ++b) {
        const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
        const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
        const int par = bi[0], kind = bi[1], q0 = bi[2];
        double RGF[9];
        V3 oF, wP = mk(0, 0, 0), vP = mk(0, 0, 0), pP = mk(0, 0, 0);
        if (par >= 0) {
            mm3(S.R[par], bd, RGF);
            pP = S.p[par];
            oF = pP + rmul(S.R[par], ld3(bd + 9));
            wP = S.w[par];
            vP = S.v[par];
        } else {
#pragma unroll
            for (int k = 0; k < 9; ++k) RGF[k] = bd[k];
            oF = ld3(bd + 9);
        }
        double RGM[9];
        V3 oM = oF;
        const V3 xF = mk(RGF[0], RGF[3], RGF[6]), yF = mk(RGF[1], RGF[4], RGF[7]), zF = mk(RGF[2], RGF[5], RGF[8]);
        V3 wrel = mk(0, 0, 0), vtr = mk(0, 0, 0);
        if (kind == MOCOB200_JOINT_CUSTOM) {
            CustomJointEval E;
            custom_joint_eval(T, T.idx[T.o_body_cj + b], q, E);
            mm3(RGF, E.R, RGM);
            oM = oF + rmul(RGF, E.p);
            for (int k = 0; k < bi[3]; ++k) {
                V3 a, l;
                custom_joint_axis(E, q0 + k, a, l);
                wrel = wrel + u[q0 + k] * rmul(RGF, a);
                vtr = vtr + u[q0 + k] * rmul(RGF, l);
            }
        } else if (kind == MOCOB200_JOINT_PIN || kind == MOCOB200_JOINT_PLANAR) {
            double s, c;
            sincos(q[q0], &s, &c);
            // R_GM = R_GF * Rz(theta)
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                RGM[3 * i] = RGF[3 * i] * c + RGF[3 * i + 1] * s;
                RGM[3 * i + 1] = -RGF[3 * i] * s + RGF[3 * i + 1] * c;
                RGM[3 * i + 2] = RGF[3 * i + 2];
            }
            wrel = u[q0] * zF;
            if (kind == MOCOB200_JOINT_PLANAR) {
                oM = oF + q[q0 + 1] * xF + q[q0 + 2] * yF;
                vtr = u[q0 + 1] * xF + u[q0 + 2] * yF;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 9; ++k) RGM[k] = RGF[k];
            if (kind == MOCOB200_JOINT_SLIDER) {
                oM = oF + q[q0] * xF;
                vtr = u[q0] * xF;
            }
        }
        mmt3(RGM, bd + 12, S.R[b]); // R_GB = R_GM * R_BM^T
        const V3 pB = oM - rmul(S.R[b], ld3(bd + 21));
        S.p[b] = pB;
        const V3 sv = pB - oM;
        S.w[b] = wP + wrel;
        S.v[b] = vP + cross(wP, pB - pP) + cross(wrel, sv) + vtr;
        S.Fm[b] = mk(0, 0, 0);
        S.Ff[b] = mk(0, 0, 0);
    }
}

template <class SZ>
__device__ __forceinline__ void apply_point_force(ScratchT<SZ>& S, int b, V3 pointG, V3 f) {
    if (b < 0) return; // ground
    S.Ff[b] = S.Ff[b] + f;
    S.Fm[b] = S.Fm[b] + cross(pointG - S.p[b], f);
}
template <class SZ>
__device__ __forceinline__ V3 point_velocity(const ScratchT<SZ>& S, int b, V3 pointG) {
    if (b < 0) return mk(0, 0, 0);
    return S.v[b] + cross(S.w[b], pointG - S.p[b]);
}

// SimTK::SmoothSphereHalfSpaceForce (Simbody 3.7; not in the reference tree). Written from the
// model description in opensim-core/OpenSim/Simulation/Model/SmoothSphereHalfSpaceForce.h:28-66 and
// Serrancoli et al. 2019: smoothed Hertz force, Hunt–Crossley dissipation,
// smooth stick–slip friction. Returns the force on the sphere body; applies both sides.
template <class SZ>
__device__ inline V3 contact_force(const DevTables& T, int ci, ScratchT<SZ>& S, bool apply) {
    const double* cd = T.dbl + T.o_contact_d + ci * MOCOB200_CONTACT_DSTRIDE;
    const int* cidx = T.idx + T.o_contact_i + ci * MOCOB200_CONTACT_ISTRIDE;
    const int bs = cidx[0], bh = cidx[1];
    const double radius = cd[3];
    const double* pr = cd + 16;
    const double diss = pr[1], us = pr[2], ud = pr[3], uv = pr[4], vt = pr[5], cf = pr[6], bd = pr[7], bv = pr[8];
    // half-space frame in ground; the half space fills +x of its frame, so the outward normal is -x
    V3 nrm, oH;
    {
        const V3 xh = mk(cd[4], cd[7], cd[10]);
        V3 ph = ld3(cd + 13);
        if (bh >= 0) { nrm = -1.0 * rmul(S.R[bh], xh); oH = S.p[bh] + rmul(S.R[bh], ph); }
        else { nrm = -1.0 * xh; oH = ph; }
    }
    const V3 centre = (bs >= 0) ? S.p[bs] + rmul(S.R[bs], ld3(cd)) : ld3(cd);
    const double indentation = radius - dot(centre - oH, nrm);
    // contact point: sphere surface point nearest the plane, moved half the indentation back in
    const V3 cp = centre - (radius - 0.5 * indentation) * nrm;
    const V3 vrel = point_velocity(S, bs, cp) - point_velocity(S, bh, cp);
    const double vnormal = dot(vrel, nrm);
    const V3 vtan = vrel - vnormal * nrm;
    const double indVel = -vnormal;
    const double ind = sqrt(indentation * indentation + cf);
    const double fH = cd[DEV_CONTACT_HERTZ] * (ind * sqrt(ind)); // (4/3) k sqrt(radius k) ind^(3/2)
    const double fHs = fH * (0.5 + 0.5 * tanh(bd * indentation)) + cf;
    const double fHC = fHs * (1.0 + 1.5 * diss * indVel);
    const double fn = fHC * (0.5 + 0.5 * tanh(bv * (indVel + 2.0 / (3.0 * diss))));
    V3 force = fn * nrm;
    const double vslip = sqrt(dot(vtan, vtan) + cf);
    const double vr = vslip / vt;
    const double ff = fn * (fmin(vr, 1.0) * (ud + 2.0 * (us - ud) / (1.0 + vr * vr)) + uv * vslip);
    force = force + (-ff / vslip) * vtan;
    if (apply) {
        apply_point_force(S, bs, cp, force);
        apply_point_force(S, bh, cp, -1.0 * force);
    }
    return force;
}

// Muscle path geometry + muscle model + equivalent forces for muscle m
// (GeometryPath.cpp:306-432, 817-864, 1117-1145; Point.cpp:130-140; ConditionalPathPoint.cpp:160-169;
//  MovingPathPoint.cpp:216-315, 374-392; PathActuator.cpp:170-197).
// ---- a muscle path with wrap cylinders (GeometryPath.cpp:868-1094 applyWrapObjects, :1095-1110 calcPathLengthChange,
// :1117-1145): the active path points in `pos/body` (ground positions), `orig` their numbers in the PathPointSet. Builds the
// current path in `path` (entries >= 0: active point; -(2 w + 1), -(2 w + 2): the two tangent points of the muscle's pairing w,
// stations of the wrap object's body held in wr1 / wr2 in that body's frame, with the helix length in warc).
#define MOCOB200_MAXWRAP 4 /* pairings per muscle */
// FR: where the body frames live (R(b): row-major rotation, p(b): origin) - the thread kernels' scratch or the warp kernels'
// shared-memory scratch
template <class FR>
__device__ __noinline__ void apply_wrap_objects(const DevTables& T, int m, const FR& S, int nA, int nSet, const int* orig,
                                                const V3* pos, int* path, int& np, V3* wr1, V3* wr2, double* warc) {
    const int pw0 = T.idx[T.o_pw_i + 2 * m], nW = min(T.idx[T.o_pw_i + 2 * m + 1], MOCOB200_MAXWRAP);
    const int* pwBase = T.idx + T.o_pw_i + 2 * T.nmus + pw0 * MOCOB200_PW_ISTRIDE;
    auto wobj = [&](int e) { return e >= 0 ? -1 : pwBase[((-e - 1) >> 1) * MOCOB200_PW_ISTRIDE]; }; // wrap object of a path entry
    auto ground = [&](int e) -> V3 {
        if (e >= 0) return pos[e];
        const int w = (-e - 1) >> 1;
        const int b = T.idx[T.o_wo_i + wobj(e) * MOCOB200_WO_ISTRIDE];
        const V3 r = ((-e - 1) & 1) ? wr2[w] : wr1[w];
        return b >= 0 ? S.p(b) + rmul(S.R(b), r) : r;
    };
    auto removeWrap = [&](int w) {
        for (int j = 0; j < np; ++j)
            if (path[j] == -(2 * w + 1)) { for (int k = j; k + 2 < np; ++k) path[k] = path[k + 2]; np -= 2; return; }
    };
    int order[MOCOB200_MAXWRAP], result[MOCOB200_MAXWRAP];
    for (int i = 0; i < nW; ++i) order[i] = i;
    const int maxIterations = nW < 2 ? 1 : 8;
    double lastLength = INFINITY;
    for (int kk = 0; kk < maxIterations; ++kk) {
        for (int i = 0; i < nW; ++i) {
            result[i] = WRAP_NONE;
            const int w = order[i];
            const int* pw = pwBase + w * MOCOB200_PW_ISTRIDE;
            const int* wi = T.idx + T.o_wo_i + pw[0] * MOCOB200_WO_ISTRIDE;
            const double* wd = T.dbl + T.o_wo_d + pw[0] * MOCOB200_WO_DSTRIDE;
            removeWrap(w);
            if (!wi[3]) continue;
            const int wrapStart = pw[1] < 1 ? 0 : pw[1] - 1, wrapEnd = pw[2] < 1 ? nSet - 1 : pw[2] - 1;
            int first = -1, last = -1;
            for (int k = 0; k < nA; ++k) if (orig[k] >= wrapStart && orig[k] <= wrapEnd) { if (first < 0) first = k; last = k; }
            if (first < 0) return;
            int start = -1, end = -1;
            for (int j = 0; j < np; ++j) { if (path[j] == first) start = j; if (path[j] == last) end = j; }
            if (start < 0 || end < 0) return;
            const int wb = wi[0];
            auto toWrapFrame = [&](V3 g) { // ground -> wrap object's body -> wrap frame
                const V3 b = wb >= 0 ? rtmul(S.R(wb), g - S.p(wb)) : g;
                return rtmul(wd, b - ld3(wd + 9));
            };
            bool haveBest = false;
            CylWrap best;
            int bestEnd = -1;
            double minChange = INFINITY;
            for (int pt1 = start; pt1 < end; ++pt1) {
                const int pt2 = pt1 + 1;
                const int wa = wobj(path[pt1]), wc = wobj(path[pt2]);
                if (wa >= 0 && wc >= 0 && wa == wc) continue;
                const V3 g1 = ground(path[pt1]), g2 = ground(path[pt2]);
                const V3 a1 = toWrapFrame(g1), a2 = toWrapFrame(g2);
                CylWrap cw;
                result[i] = wr_cylinder(wd[12], wd[13], wi[1], wi[2], a1, a2, cw);
                if (result[i] == WRAP_MANDATORY) { best = cw; bestEnd = pt2; haveBest = true; break; }
                if (result[i] == WRAP_WRAPPED) {
                    const double change = wr_norm(a1 - cw.r1) + cw.arc + wr_norm(a2 - cw.r2) - wr_norm(g1 - g2);
                    if (change < minChange) { best = cw; bestEnd = pt2; haveBest = true; minChange = change; }
                }
            }
            if (!haveBest) continue;
            wr1[w] = rmul(wd, best.r1) + ld3(wd + 9);
            wr2[w] = rmul(wd, best.r2) + ld3(wd + 9);
            warc[w] = best.arc;
            for (int k = np - 1; k >= bestEnd; --k) path[k + 2] = path[k];
            path[bestEnd] = -(2 * w + 1);
            path[bestEnd + 1] = -(2 * w + 2);
            np += 2;
        }
        double len = 0;
        for (int k = 0; k + 1 < np; ++k) {
            const int a = path[k], b = path[k + 1];
            if (a < 0 && b < 0 && wobj(a) == wobj(b)) len += warc[(-b - 1) >> 1];
            else len += wr_norm(ground(a) - ground(b));
        }
        if (fabs(len - lastLength) < 0.0005) break;
        lastLength = len;
        if (kk == 0 && nW > 1 && result[0] == WRAP_NONE && result[1] == WRAP_INSIDE_RADIUS) {
            order[0] = 1; order[1] = 0;
            removeWrap(0);
        }
    }
}

// muscle_force for a muscle that has wrap-cylinder pairings (same contract)
template <class SZ>
__device__ __noinline__ MuscleOut muscle_force_wrapped(const DevTables& T, int m, const double* q, const double* u, const double* z,
                                                       const double* ctl, const double* deriv, ScratchT<SZ>& S, bool apply,
                                                       bool wantStiffness, double* lengthSpeed) {
    const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
    const double* md = T.dbl + T.o_mus_d + m * MOCOB200_MUS_DSTRIDE;
    const int first = mi[0], count = mi[1], flags = mi[5];
    constexpr int NP = SZ::NPPM, NPATH = SZ::NPPM + 2 * MOCOB200_MAXWRAP;
    V3 pos[NP], vel[NP], dPdq[NP];
    int body[NP], mcoord[NP], orig[NP];
    int na = 0;
    for (int k = 0; k < count && na < NP; ++k) {
        const int* pi = T.idx + T.o_pp_i + (first + k) * MOCOB200_PP_ISTRIDE;
        const double* pd = T.dbl + T.o_pp_d + (first + k) * MOCOB200_PP_DSTRIDE;
        const int b = pi[0], kind = pi[1];
        V3 loc = ld3(pd), vloc = mk(0, 0, 0), dp = mk(0, 0, 0);
        int mc = -1;
        if (kind == MOCOB200_PP_CONDITIONAL) {
            const double val = q[pi[2]];
            if (!(val >= pd[3] - 1e-5 && val <= pd[4] + 1e-5)) continue;
        } else if (kind == MOCOB200_PP_MOVING) {
            mc = pi[6];
            const double qv = q[mc];
            const double qc = fmin(fmax(qv, pd[3]), pd[4]);
            loc = mk(simm_spline(T, pi[3], qc, 0), simm_spline(T, pi[4], qc, 0), simm_spline(T, pi[5], qc, 0));
            dp = mk(simm_spline(T, pi[3], qv, 1), simm_spline(T, pi[4], qv, 1), simm_spline(T, pi[5], qv, 1));
            vloc = u[mc] * dp;
        }
        V3 pG, vG;
        if (b >= 0) {
            const V3 r = rmul(S.R[b], loc);
            pG = S.p[b] + r;
            vG = S.v[b] + cross(S.w[b], r);
            if (mc >= 0) { vG = vG + rmul(S.R[b], vloc); dp = rmul(S.R[b], dp); }
        } else {
            pG = loc;
            vG = vloc;
        }
        pos[na] = pG; vel[na] = vG; body[na] = b; mcoord[na] = mc; dPdq[na] = dp; orig[na] = k;
        ++na;
    }
    int path[NPATH], np = na;
    for (int k = 0; k < na; ++k) path[k] = k;
    V3 wr1[MOCOB200_MAXWRAP], wr2[MOCOB200_MAXWRAP];
    double warc[MOCOB200_MAXWRAP];
    struct Frames {
        const ScratchT<SZ>& S;
        __device__ const double* R(int b) const { return S.R[b]; }
        __device__ V3 p(int b) const { return S.p[b]; }
    };
    apply_wrap_objects(T, m, Frames{S}, na, count, orig, pos, path, np, wr1, wr2, warc);
    const int pw0 = T.idx[T.o_pw_i + 2 * m];
    const int* pwBase = T.idx + T.o_pw_i + 2 * T.nmus + pw0 * MOCOB200_PW_ISTRIDE;
    // the entries of the current path as (ground position, velocity, body); tangent points are stations of the wrap object's body
    auto entry = [&](int e, V3& pG, V3& vG, int& b, int& wo) {
        if (e >= 0) { pG = pos[e]; vG = vel[e]; b = body[e]; wo = -1; return; }
        const int w = (-e - 1) >> 1;
        wo = pwBase[w * MOCOB200_PW_ISTRIDE];
        b = T.idx[T.o_wo_i + wo * MOCOB200_WO_ISTRIDE];
        const V3 r = ((-e - 1) & 1) ? wr2[w] : wr1[w];
        if (b >= 0) { const V3 rg = rmul(S.R[b], r); pG = S.p[b] + rg; vG = S.v[b] + cross(S.w[b], rg); }
        else { pG = r; vG = mk(0, 0, 0); }
    };
    double length = 0, speed = 0;
    {
        V3 prevP = mk(0, 0, 0), prevV = mk(0, 0, 0);
        int prevWo = -1;
        for (int j = 0; j < np; ++j) {
            V3 pG, vG;
            int b, wo;
            entry(path[j], pG, vG, b, wo);
            if (j > 0) {
                const V3 r = prevP - pG, dv = prevV - vG;
                const double d = sqrt(dot(r, r));
                length += (wo >= 0 && wo == prevWo) ? warc[(-path[j] - 1) >> 1] : d; // over the surface: the helix length
                speed += (d < 2.220446049250313e-16) ? sqrt(dot(dv, dv)) : dot(dv, (1.0 / d) * r);
            }
            prevP = pG; prevV = vG; prevWo = wo;
        }
    }
    if (lengthSpeed) { lengthSpeed[0] = length; lengthSpeed[1] = speed; }
    double activation = (flags & MOCOB200_MUS_FLAG_IGNORE_ACTIVATION_DYNAMICS) ? ctl[mi[2]] : z[mi[3]];
    const double ntf = (mi[4] >= 0) ? z[mi[4]] : NAN;
    const double ntfd = (mi[6] >= 0) ? deriv[mi[6]] : NAN;
    MuscleOut o = dgf_muscle(md, flags, length, speed, activation, ntf, ntfd, wantStiffness);
    if (apply) {
        const double tension = o.tendonForce;
        for (int k = 0; k + 1 < np; ++k) {
            V3 pa, va, pb, vb;
            int ba, bb, woa, wob;
            entry(path[k], pa, va, ba, woa);
            entry(path[k + 1], pb, vb, bb, wob);
            if (ba == bb) continue;
            V3 dir = pb - pa;
            const double nn = sqrt(dot(dir, dir));
            dir = (nn < 1.9984014443252818e-14) ? (NAN * dir) : ((1.0 / nn) * dir);
            const V3 f = tension * dir;
            apply_point_force(S, ba, pa, f);
            apply_point_force(S, bb, pb, -1.0 * f);
            const int ea = path[k], eb = path[k + 1];
            if (ea >= 0 && mcoord[ea] >= 0) S.tau[mcoord[ea]] += dot(dPdq[ea], f);
            if (eb >= 0 && mcoord[eb] >= 0) S.tau[mcoord[eb]] -= dot(dPdq[eb], f);
        }
    }
    return o;
}

template <class SZ>
__device__ inline MuscleOut muscle_force(const DevTables& T, int m, const double* q, const double* u, const double* z,
                                        const double* ctl, const double* deriv, ScratchT<SZ>& S, bool apply, bool wantStiffness,
                                        double* lengthSpeed = nullptr) {
    if constexpr (SZ::WRAP) {
        if (T.npathwrap > 0 && T.idx[T.o_pw_i + 2 * m + 1] > 0) // the path runs over wrap cylinders
            return muscle_force_wrapped<SZ>(T, m, q, u, z, ctl, deriv, S, apply, wantStiffness, lengthSpeed);
    }
    const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
    const double* md = T.dbl + T.o_mus_d + m * MOCOB200_MUS_DSTRIDE;
    const int first = mi[0], count = mi[1], flags = mi[5];
    V3 pos[SZ::NPPM], dPdq[SZ::NPPM];
    int body[SZ::NPPM], mcoord[SZ::NPPM];
    int na = 0;
    double length = 0, speed = 0;
    V3 prevP = mk(0, 0, 0), prevV = mk(0, 0, 0);
    for (int k = 0; k < count; ++k) {
        const int* pi = T.idx + T.o_pp_i + (first + k) * MOCOB200_PP_ISTRIDE;
        const double* pd = T.dbl + T.o_pp_d + (first + k) * MOCOB200_PP_DSTRIDE;
        const int b = pi[0], kind = pi[1];
        V3 loc = ld3(pd), vloc = mk(0, 0, 0), dp = mk(0, 0, 0);
        int mc = -1;
        if (kind == MOCOB200_PP_CONDITIONAL) {
            const double val = q[pi[2]];
            if (!(val >= pd[3] - 1e-5 && val <= pd[4] + 1e-5)) continue;
        } else if (kind == MOCOB200_PP_MOVING) {
            mc = pi[6];
            const double qv = q[mc];
            const double qc = fmin(fmax(qv, pd[3]), pd[4]); // location clamps the coordinate, derivatives do not
            loc = mk(simm_spline(T, pi[3], qc, 0), simm_spline(T, pi[4], qc, 0), simm_spline(T, pi[5], qc, 0));
            dp = mk(simm_spline(T, pi[3], qv, 1), simm_spline(T, pi[4], qv, 1), simm_spline(T, pi[5], qv, 1));
            vloc = u[mc] * dp;
        }
        V3 pG, vG;
        if (b >= 0) {
            const V3 r = rmul(S.R[b], loc);
            pG = S.p[b] + r;
            vG = S.v[b] + cross(S.w[b], r);
            if (mc >= 0) { vG = vG + rmul(S.R[b], vloc); dp = rmul(S.R[b], dp); }
        } else {
            pG = loc;
            vG = vloc;
        }
        if (na > 0) {
            const V3 r = prevP - pG, dv = prevV - vG;
            const double d = sqrt(dot(r, r));
            length += d;
            speed += (d < 2.220446049250313e-16) ? sqrt(dot(dv, dv)) : dot(dv, (1.0 / d) * r);
        }
        prevP = pG;
        prevV = vG;
        if (na < SZ::NPPM) { pos[na] = pG; body[na] = b; mcoord[na] = mc; dPdq[na] = dp; }
        ++na;
    }
    if (lengthSpeed) { lengthSpeed[0] = length; lengthSpeed[1] = speed; }
    double activation = (flags & MOCOB200_MUS_FLAG_IGNORE_ACTIVATION_DYNAMICS) ? ctl[mi[2]] : z[mi[3]];
    const double ntf = (mi[4] >= 0) ? z[mi[4]] : NAN;
    const double ntfd = (mi[6] >= 0) ? deriv[mi[6]] : NAN;
    MuscleOut o = dgf_muscle(md, flags, length, speed, activation, ntf, ntfd, wantStiffness);
    if (apply) {
        const double tension = o.tendonForce;
        for (int k = 0; k + 1 < na; ++k) {
            if (body[k] == body[k + 1]) continue;
            V3 dir = pos[k + 1] - pos[k];
            const double nn = sqrt(dot(dir, dir));
            // SimTK::SignificantReal = eps^(7/8)
            dir = (nn < 1.9984014443252818e-14) ? (NAN * dir) : ((1.0 / nn) * dir);
            const V3 f = tension * dir;
            apply_point_force(S, body[k], pos[k], f);
            apply_point_force(S, body[k + 1], pos[k + 1], -1.0 * f);
            if (mcoord[k] >= 0) S.tau[mcoord[k]] += dot(dPdq[k], f);
            if (mcoord[k + 1] >= 0) S.tau[mcoord[k + 1]] -= dot(dPdq[k + 1], f);
        }
    }
    return o;
}

// activation dynamics (DeGrooteFregly2016Muscle.cpp:190-210)
__device__ __forceinline__ double activation_dot(double e, double a, double tauA, double tauD) {
    const double zf = 0.5 + 1.5 * a;
    const double tempAct = 1.0 / (tauA * zf);
    const double tempDeact = zf / tauD;
    const double f = 0.5 * tanh(0.1 * (e - a));
    const double timeConst = tempAct * (f + 0.5) + tempDeact * (-f + 0.5);
    return timeConst * (e - a);
}

// 6x6 symmetric articulated inertia in blocks: A (ang-ang, sym, 6), B (ang-lin, full 9), C (lin-lin, sym, 6)
struct ArtI {
    double A[6], B[9], C[6]; // sym order: xx yy zz xy xz yz
};
__device__ __forceinline__ V3 symmul(const double* s, V3 v) {
    return mk(s[0] * v.x + s[3] * v.y + s[4] * v.z, s[3] * v.x + s[1] * v.y + s[5] * v.z, s[4] * v.x + s[5] * v.y + s[2] * v.z);
}
__device__ __forceinline__ V3 bmul(const double* B, V3 v) { return rmul(B, v); }
__device__ __forceinline__ V3 btmul(const double* B, V3 v) { return rtmul(B, v); }

// spatial inertia of body b about its origin, in ground: I_o = R I R^T, c = R com
template <class SZ>
__device__ inline void body_inertia_ground(const DevTables& T, int b, const ScratchT<SZ>& S, double* Io /*sym6*/, V3& mc, double& mass) {
    const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
    mass = bd[24];
    const V3 c = rmul(S.R[b], ld3(bd + 25));
    mc = mass * c;
    const double* in = bd + 28;
    const double I[9] = {in[0], in[3], in[4], in[3], in[1], in[5], in[4], in[5], in[2]};
    double RI[9], RIRt[9];
    mm3(S.R[b], I, RI);
    mmt3(RI, S.R[b], RIRt);
    Io[0] = RIRt[0]; Io[1] = RIRt[4]; Io[2] = RIRt[8];
    Io[3] = 0.5 * (RIRt[1] + RIRt[3]); Io[4] = 0.5 * (RIRt[2] + RIRt[6]); Io[5] = 0.5 * (RIRt[5] + RIRt[7]);
}

// hinge data of body b in ground: returns dof; axes as (angular, linear-at-B-origin) pairs
template <class SZ>
__device__ inline int hinge(const DevTables& T, int b, const ScratchT<SZ>& S, V3* Ha, V3* Hl, V3& oMrel /* pB - oM */) {
    const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
    const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
    const int par = bi[0], kind = bi[1];
    double RGF[9];
    if (par >= 0) mm3(S.R[par], bd, RGF);
    else {
#pragma unroll
        for (int k = 0; k < 9; ++k) RGF[k] = bd[k];
    }
    const V3 xF = mk(RGF[0], RGF[3], RGF[6]), yF = mk(RGF[1], RGF[4], RGF[7]), zF = mk(RGF[2], RGF[5], RGF[8]);
    oMrel = rmul(S.R[b], ld3(bd + 21)); // = pB - oM with a minus sign: pB = oM - R_GB p_BM
    oMrel = -1.0 * oMrel;
    if (kind == MOCOB200_JOINT_PIN) {
        Ha[0] = zF; Hl[0] = cross(zF, oMrel);
        return 1;
    }
    if (kind == MOCOB200_JOINT_SLIDER) {
        Ha[0] = mk(0, 0, 0); Hl[0] = xF;
        return 1;
    }
    if (kind == MOCOB200_JOINT_PLANAR) {
        Ha[0] = zF; Hl[0] = cross(zF, oMrel);
        Ha[1] = mk(0, 0, 0); Hl[1] = xF;
        Ha[2] = mk(0, 0, 0); Hl[2] = yF;
        return 3;
    }
    return 0;
}

// velocity-dependent (Coriolis) acceleration of body b relative to its parent, ground frame
template <class SZ>
__device__ inline void coriolis(const DevTables& T, int b, const ScratchT<SZ>& S, const double* u, V3& ca, V3& cl) {
    const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
    const int par = bi[0], q0 = bi[2];
    V3 Ha[3], Hl[3], sv;
    const int nd = hinge(T, b, S, Ha, Hl, sv);
    V3 wP = mk(0, 0, 0), pP = mk(0, 0, 0);
    if (par >= 0) { wP = S.w[par]; pP = S.p[par]; }
    V3 wrel = mk(0, 0, 0), vrel = mk(0, 0, 0);
    for (int k = 0; k < nd; ++k) { wrel = wrel + u[q0 + k] * Ha[k]; vrel = vrel + u[q0 + k] * Hl[k]; }
    const V3 r = S.p[b] - pP;
    ca = cross(wP, wrel);
    cl = cross(wP, cross(wP, r)) + 2.0 * cross(wP, vrel) + cross(wrel, cross(wrel, sv));
}

// Explicit mode: articulated-body algorithm. udot out.
template <class SZ>
__device__ inline void forward_dynamics(const DevTables& T, const double* u, ScratchT<SZ>& S, double* udot) {
    ArtI P[SZ::NB];
    V3 zm[SZ::NB], zf[SZ::NB]; // z (moment; force)
    V3 Ga[SZ::NQ], Gl[SZ::NQ]; // G = P H D^-1 columns
    double nu[SZ::NQ];          // D^-1 eps
    const int nb = T.nb;
    // initialise with rigid-body inertias and bias forces
    for (int b = 0; b < nb; ++b) {
        double Io[6], mass;
        V3 mc;
        body_inertia_ground(T, b, S, Io, mc, mass);
        ArtI& a = P[b];
#pragma unroll
        for (int k = 0; k < 6; ++k) a.A[k] = Io[k];
        // B = m c× ; C = m 1
        a.B[0] = 0; a.B[1] = -mc.z; a.B[2] = mc.y;
        a.B[3] = mc.z; a.B[4] = 0; a.B[5] = -mc.x;
        a.B[6] = -mc.y; a.B[7] = mc.x; a.B[8] = 0;
        a.C[0] = a.C[1] = a.C[2] = mass; a.C[3] = a.C[4] = a.C[5] = 0;
        const V3 w = S.w[b];
        // gyroscopic moment and centrifugal force, then minus applied force
        zm[b] = cross(w, symmul(Io, w)) - S.Fm[b];
        zf[b] = cross(w, cross(w, mc)) - S.Ff[b];
    }
    // tip to base
    for (int b = nb - 1; b >= 0; --b) {
        const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
        const int par = bi[0], q0 = bi[2];
        V3 Ha[3], Hl[3], sv, ca, cl;
        const int nd = hinge(T, b, S, Ha, Hl, sv);
        coriolis(T, b, S, u, ca, cl);
        ArtI& a = P[b];
        // z += P * coriolis
        zm[b] = zm[b] + symmul(a.A, ca) + bmul(a.B, cl);
        zf[b] = zf[b] + btmul(a.B, ca) + symmul(a.C, cl);
        // joint: U = P H (6 x nd), D = H^T U
        V3 Ua[3], Ul[3];
        double D[9], Di[9], eps[3];
        for (int k = 0; k < nd; ++k) {
            Ua[k] = symmul(a.A, Ha[k]) + bmul(a.B, Hl[k]);
            Ul[k] = btmul(a.B, Ha[k]) + symmul(a.C, Hl[k]);
            eps[k] = S.tau[q0 + k] - (dot(Ha[k], zm[b]) + dot(Hl[k], zf[b]));
        }
        for (int i = 0; i < nd; ++i)
            for (int j = 0; j < nd; ++j) D[3 * i + j] = dot(Ha[i], Ua[j]) + dot(Hl[i], Ul[j]);
        if (nd == 1) Di[0] = 1.0 / D[0];
        else if (nd == 3) {
            const double c00 = D[4] * D[8] - D[5] * D[7], c01 = D[5] * D[6] - D[3] * D[8], c02 = D[3] * D[7] - D[4] * D[6];
            const double det = D[0] * c00 + D[1] * c01 + D[2] * c02, id = 1.0 / det;
            Di[0] = c00 * id; Di[1] = (D[2] * D[7] - D[1] * D[8]) * id; Di[2] = (D[1] * D[5] - D[2] * D[4]) * id;
            Di[3] = c01 * id; Di[4] = (D[0] * D[8] - D[2] * D[6]) * id; Di[5] = (D[2] * D[3] - D[0] * D[5]) * id;
            Di[6] = c02 * id; Di[7] = (D[1] * D[6] - D[0] * D[7]) * id; Di[8] = (D[0] * D[4] - D[1] * D[3]) * id;
        }
        V3 zpm = zm[b], zpf = zf[b];
        for (int i = 0; i < nd; ++i) {
            V3 ga = mk(0, 0, 0), gl = mk(0, 0, 0);
            double n = 0;
            for (int j = 0; j < nd; ++j) {
                ga = ga + Di[3 * j + i] * Ua[j];
                gl = gl + Di[3 * j + i] * Ul[j];
                n += Di[3 * i + j] * eps[j];
            }
            Ga[q0 + i] = ga; Gl[q0 + i] = gl; nu[q0 + i] = n;
        }
        for (int i = 0; i < nd; ++i) { zpm = zpm + eps[i] * Ga[q0 + i]; zpf = zpf + eps[i] * Gl[q0 + i]; }
        if (par < 0) continue;
        // P+ = P - G D G^T = P - sum_i G_i U_i^T ; then shift to the parent origin and accumulate
        double A[9], B[9], C[9];
        A[0] = a.A[0]; A[1] = a.A[3]; A[2] = a.A[4]; A[3] = a.A[3]; A[4] = a.A[1]; A[5] = a.A[5]; A[6] = a.A[4]; A[7] = a.A[5]; A[8] = a.A[2];
        C[0] = a.C[0]; C[1] = a.C[3]; C[2] = a.C[4]; C[3] = a.C[3]; C[4] = a.C[1]; C[5] = a.C[5]; C[6] = a.C[4]; C[7] = a.C[5]; C[8] = a.C[2];
#pragma unroll
        for (int k = 0; k < 9; ++k) B[k] = a.B[k];
        for (int i = 0; i < nd; ++i) {
            const double ga[3] = {Ga[q0 + i].x, Ga[q0 + i].y, Ga[q0 + i].z}, gl[3] = {Gl[q0 + i].x, Gl[q0 + i].y, Gl[q0 + i].z};
            const double ua[3] = {Ua[i].x, Ua[i].y, Ua[i].z}, ul[3] = {Ul[i].x, Ul[i].y, Ul[i].z};
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    A[3 * r + c] -= ga[r] * ua[c];
                    B[3 * r + c] -= ga[r] * ul[c];
                    C[3 * r + c] -= gl[r] * ul[c];
                }
        }
        // shift by r = pB - pP : A' = A + rx B^T + B rx^T + rx C rx^T ; B' = B + rx C ; C' = C
        const V3 r = S.p[b] - S.p[par];
        const double rx[9] = {0, -r.z, r.y, r.z, 0, -r.x, -r.y, r.x, 0};
        double rxC[9], Bp[9], t1[9], t2[9];
        mm3(rx, C, rxC);
#pragma unroll
        for (int k = 0; k < 9; ++k) Bp[k] = B[k] + rxC[k];
        mmt3(rx, B, t1);  // rx B^T
        mmt3(Bp, rx, t2); // (B + rx C) rx^T
        ArtI& pa = P[par];
        double Ap[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) Ap[k] = A[k] + t1[k] + t2[k];
        pa.A[0] += Ap[0]; pa.A[1] += Ap[4]; pa.A[2] += Ap[8];
        pa.A[3] += 0.5 * (Ap[1] + Ap[3]); pa.A[4] += 0.5 * (Ap[2] + Ap[6]); pa.A[5] += 0.5 * (Ap[5] + Ap[7]);
#pragma unroll
        for (int k = 0; k < 9; ++k) pa.B[k] += Bp[k];
        pa.C[0] += C[0]; pa.C[1] += C[4]; pa.C[2] += C[8];
        pa.C[3] += 0.5 * (C[1] + C[3]); pa.C[4] += 0.5 * (C[2] + C[6]); pa.C[5] += 0.5 * (C[5] + C[7]);
        zm[par] = zm[par] + zpm + cross(r, zpf);
        zf[par] = zf[par] + zpf;
    }
    // base to tip: accelerations. Reuse zm/zf as the spatial acceleration (alpha ; a_origin).
    for (int b = 0; b < nb; ++b) {
        const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
        const int par = bi[0], q0 = bi[2];
        V3 Ha[3], Hl[3], sv, ca, cl;
        const int nd = hinge(T, b, S, Ha, Hl, sv);
        coriolis(T, b, S, u, ca, cl);
        V3 aa = mk(0, 0, 0), al = mk(0, 0, 0);
        if (par >= 0) {
            const V3 r = S.p[b] - S.p[par];
            aa = zm[par];
            al = zf[par] + cross(zm[par], r);
        }
        for (int i = 0; i < nd; ++i) udot[q0 + i] = nu[q0 + i] - (dot(Ga[q0 + i], aa) + dot(Gl[q0 + i], al));
        aa = aa + ca;
        al = al + cl;
        for (int i = 0; i < nd; ++i) { aa = aa + udot[q0 + i] * Ha[i]; al = al + udot[q0 + i] * Hl[i]; }
        zm[b] = aa;
        zf[b] = al;
    }
}

// Implicit mode: residual = f_app - M udot - f_inertial by recursive Newton–Euler
// (what SimbodyMatterSubsystem::findMotionForces reports under AccelerationMotion,
//  MocoCasOCProblem.h:245-297; sign per Moco/doc/MocoDeveloperGuide.dox:175-208).
template <class SZ>
__device__ inline void inverse_dynamics_residual(const DevTables& T, const double* u, const double* udot, ScratchT<SZ>& S, double* resid) {
    V3 fm[SZ::NB], ff[SZ::NB], aa[SZ::NB], al[SZ::NB];
    const int nb = T.nb;
    for (int b = 0; b < nb; ++b) {
        const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
        const int par = bi[0], q0 = bi[2];
        V3 Ha[3], Hl[3], sv, ca, cl;
        const int nd = hinge(T, b, S, Ha, Hl, sv);
        coriolis(T, b, S, u, ca, cl);
        V3 a = ca, l = cl;
        if (par >= 0) {
            const V3 r = S.p[b] - S.p[par];
            a = a + aa[par];
            l = l + al[par] + cross(aa[par], r);
        }
        for (int i = 0; i < nd; ++i) { a = a + udot[q0 + i] * Ha[i]; l = l + udot[q0 + i] * Hl[i]; }
        aa[b] = a; al[b] = l;
        double Io[6], mass;
        V3 mc;
        body_inertia_ground(T, b, S, Io, mc, mass);
        const V3 w = S.w[b];
        fm[b] = symmul(Io, a) + cross(mc, l) + cross(w, symmul(Io, w)) - S.Fm[b];
        ff[b] = mass * l + cross(a, mc) + cross(w, cross(w, mc)) - S.Ff[b];
    }
    for (int b = nb - 1; b >= 0; --b) {
        const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
        const int par = bi[0], q0 = bi[2];
        V3 Ha[3], Hl[3], sv;
        const int nd = hinge(T, b, S, Ha, Hl, sv);
        for (int i = 0; i < nd; ++i) resid[q0 + i] = S.tau[q0 + i] - (dot(Ha[i], fm[b]) + dot(Hl[i], ff[b]));
        if (par >= 0) {
            const V3 r = S.p[b] - S.p[par];
            fm[par] = fm[par] + fm[b] + cross(r, ff[b]);
            ff[par] = ff[par] + ff[b];
        }
    }
}

// whole-model mass centre in ground (Model::calcMassCenterPosition), needs kinematics()
template <class SZ>
__device__ inline V3 mass_center(const DevTables& T, const ScratchT<SZ>& S) {
    V3 acc = mk(0, 0, 0);
    double mt = 0;
    for (int b = 0; b < T.nb; ++b) {
        const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
        const double m = bd[24];
        acc = acc + m * (S.p[b] + rmul(S.R[b], ld3(bd + 25)));
        mt += m;
    }
    return (1.0 / mt) * acc;
}

// The fused point function. in = [time, states(2nq+nz), controls(nc), derivatives(nd)], out[n_out].
// mask selects output groups (one group per reference callback): bit 0 = dynamics
// (udot|residual, zdot, aux residuals), bit 1+k = integrand k. Unselected outputs are left untouched.
template <class SZ = SizeLarge>
__device__ inline void eval_point(const DevTables& T, const double* in, double* out, unsigned mask) {
    const int nq = T.nq, nz = T.nz;
    const double time = in[0];
    double pk[3 * SZ::NQ]; // q, u, udot of prescribed kinematics
    const PointInputs P = point_inputs(T, in, pk);
    const double* states = in + 1; // the NLP's state vector at this point (goals index into it)
    const double *q = P.q, *u = P.u, *z = P.z, *ctl = P.ctl;
    const double* mult = P.mult;     // Lagrange multipliers of the kinematic constraints
    const double* der = P.udot;      // implicit: udot[nq]
    const double* auxder = P.auxder; // implicit auxiliary derivatives
    const bool dyn_on = mask & 1u;
    // which integrands are wanted, and do any of them need kinematics + contact forces?
    bool need_contacts = dyn_on;
    {
        int gi = 0;
        for (int g2 = 0; g2 < T.ngoal; ++g2) {
            const int* gr = T.idx + T.o_goal_i + g2 * MOCOB200_GOAL_ISTRIDE;
            if (!gr[3]) continue;
            if (((mask >> (1 + gi)) & 1u) && gr[0] == MOCOB200_GOAL_CONTACT_TRACKING) need_contacts = true;
            ++gi;
        }
    }
    ScratchT<SZ> S;
    if (need_contacts) {
        kinematics(T, q, u, S);
        for (int c = 0; c < T.ncontact; ++c) {
            const V3 f = contact_force(T, c, S, dyn_on);
            if (c < MOCOB200_MAXCONTACT) S.contactF[c] = f;
        }
    }
    double* dyn = out;            // udot or residual [nq]
    double* zdot = out + nq;      // [nz]
    double* auxres = zdot + nz;   // [n_auxres]
    if (dyn_on) {
        for (int k = 0; k < nq; ++k) S.tau[k] = 0;
        // gravity (Model.cpp:587-590)
        const V3 g = ld3(T.dbl + T.o_gravity);
        for (int b = 0; b < T.nb; ++b) {
            const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
            const double m = bd[24];
            const V3 c = rmul(S.R[b], ld3(bd + 25));
            S.Ff[b] = S.Ff[b] + m * g;
            S.Fm[b] = S.Fm[b] + cross(c, m * g);
        }
        // external forces (ExternalForce::computeForce, Model/ExternalForce.cpp:330-355)
        for (int f = 0; f < T.nxforce; ++f) {
            const int* xi = T.idx + T.o_xforce_i + f * MOCOB200_XFORCE_ISTRIDE;
            const int body = xi[0];
            if (xi[3] >= 0) {
                V3 force = mk(ppoly_eval(T, xi[3], time), ppoly_eval(T, xi[4], time), ppoly_eval(T, xi[5], time));
                if (xi[1] >= 0) force = rmul(S.R[xi[1]], force);
                V3 point = S.p[body];
                if (xi[6] >= 0) {
                    point = mk(ppoly_eval(T, xi[6], time), ppoly_eval(T, xi[7], time), ppoly_eval(T, xi[8], time));
                    if (xi[2] >= 0) point = S.p[xi[2]] + rmul(S.R[xi[2]], point);
                }
                apply_point_force(S, body, point, force);
            }
            if (xi[9] >= 0) {
                V3 torque = mk(ppoly_eval(T, xi[9], time), ppoly_eval(T, xi[10], time), ppoly_eval(T, xi[11], time));
                if (xi[1] >= 0) torque = rmul(S.R[xi[1]], torque);
                S.Fm[body] = S.Fm[body] + torque;
            }
        }
        for (int m = 0; m < T.nmus; ++m) {
            const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
            const double* md = T.dbl + T.o_mus_d + m * MOCOB200_MUS_DSTRIDE;
            MuscleOut mo = muscle_force(T, m, q, u, z, ctl, auxder, S, true, false);
            if (mi[3] >= 0) zdot[mi[3]] = activation_dot(ctl[mi[2]], z[mi[3]], md[5], md[6]);
            if (mi[4] >= 0) zdot[mi[4]] = mo.nfDeriv;
            if (mi[7] >= 0) auxres[mi[7]] = mo.residual;
        }
        // coordinate actuators (CoordinateActuator.cpp:185-204)
        for (int a = 0; a < T.ncact; ++a) {
            const int* ai = T.idx + T.o_cact_i + a * MOCOB200_CACT_ISTRIDE;
            S.tau[ai[0]] += T.dbl[T.o_cact_d + a * MOCOB200_CACT_DSTRIDE] * ctl[ai[1]];
        }
        coupler_forces(T, q, mult, S.tau);
        if (T.implicit) inverse_dynamics_residual(T, u, der, S, dyn);
        else forward_dynamics(T, u, S, dyn);
        double* kc = auxres + T.n_auxres;
        if (T.n_kc > 0)
            for (int k = 0; k < T.ncoupler; ++k) coupler_errors(T, k, q, u, T.implicit ? der : dyn, kc);
    }
    // goal integrands (MocoGoal::calcIntegrand; MocoControlGoal.cpp:121-131,
    // MocoStateTrackingGoal.cpp:103-117, MocoContactTrackingGoal.cpp:240-287)
    double* integ = auxres + T.n_auxres + T.n_kc;
    int gi = 0;
    for (int g2 = 0; g2 < T.ngoal; ++g2) {
        const int* gr = T.idx + T.o_goal_i + g2 * MOCOB200_GOAL_ISTRIDE;
        if (!gr[3]) continue;
        const int slot = gi++;
        if (!((mask >> (1 + slot)) & 1u)) continue;
        const int* ip = T.idx + gr[4];
        const double* dp = T.dblG + gr[5];
        double val = 0;
        switch (gr[0]) {
        case MOCOB200_GOAL_CONTROL: {
            const int expo = ip[0], n = ip[2];
            for (int k = 0; k < n; ++k) {
                const double c = ctl[ip[3 + k]];
                val += dp[3 + k] * (expo == 2 ? c * c : abs_ipow(c, expo));
            }
        } break;
        case MOCOB200_GOAL_STATE_TRACKING: {
            const int n = ip[0];
            for (int k = 0; k < n; ++k) {
                const double e = states[ip[1 + 2 * k]] - ppoly_eval(T, ip[2 + 2 * k], time);
                val += dp[3 + k] * (e * e);
            }
        } break;
        case MOCOB200_GOAL_CONTROL_TRACKING: {
            const int n = ip[0];
            for (int k = 0; k < n; ++k) {
                const double e = ctl[ip[1 + 2 * k]] - ppoly_eval(T, ip[2 + 2 * k], time);
                val += dp[3 + k] * e * e;
            }
        } break;
        case MOCOB200_GOAL_SUM_SQUARED_STATE: {
            const int n = ip[0];
            for (int k = 0; k < n; ++k) {
                const double e = states[ip[1 + 2 * k]];
                val += dp[3 + k] * e * e;
            }
        } break;
        case MOCOB200_GOAL_CONTACT_TRACKING: {
            const int proj = ip[0], ng = ip[1];
            const V3 pv = ld3(dp + 3);
            const int* gp = ip + 2;
            for (int k = 0; k < ng; ++k) {
                const int ncon = gp[0];
                V3 fm = mk(0, 0, 0);
                for (int c = 0; c < ncon; ++c) {
                    const V3 f = S.contactF[gp[1 + 2 * c]];
                    fm = (gp[2 + 2 * c] == 0) ? fm + f : fm - f;
                }
                gp += 1 + 2 * ncon;
                const V3 fr = mk(ppoly_eval(T, gp[0], time), ppoly_eval(T, gp[1], time), ppoly_eval(T, gp[2], time));
                gp += 3;
                const V3 e3 = fm - fr;
                V3 e = e3;
                if (proj == 1) e = dot(e3, pv) * pv;
                else if (proj == 2) e = e3 - dot(e3, pv) * pv;
                val += dot(e, e);
            }
        } break;
        default: break;
        }
        integ[slot] = val;
    }
}

// Endpoint function of goal g (MocoGoal::calcGoal, MocoGoal.h:215-247 and the calcGoalImpl of each
// goal class): ein = [t0, x0(ns), c0(nc), d0(nd), tf, xf(ns), cf(nc), df(nd), integral] -> vals[n_outputs].
__device__ inline void eval_endpoint(const DevTables& T, int g, const double* ein, int ns, int nc, int nd, double* vals) {
    const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
    const int* ip = T.idx + gr[4];
    const double* dp = T.dblG + gr[5];
    const int half = 1 + ns + nc + T.nm + nd;
    const double t0 = ein[0], tf = ein[half];
    const double* x0 = ein + 1;
    const double* c0 = x0 + ns;
    const double* d0 = c0 + nc + T.nm;
    const double* xf = ein + half + 1;
    const double* cf = xf + ns;
    const double integral = ein[2 * half];
    const double w = dp[0];
    const bool cost = gr[1] == MOCOB200_MODE_COST;
    const int nq = T.nq;
    auto displacement = [&]() {
        Scratch S;
        kinematics(T, x0, x0 + nq, S);
        const V3 ci = mass_center(T, S);
        kinematics(T, xf, xf + nq, S);
        const V3 cfn = mass_center(T, S);
        const V3 d = cfn - ci;
        return sqrt(dot(d, d));
    };
    switch (gr[0]) {
    case MOCOB200_GOAL_CONTROL: {
        double v = integral;
        if (ip[1]) v /= displacement();
        vals[0] = w * v;
    } break;
    case MOCOB200_GOAL_STATE_TRACKING:
    case MOCOB200_GOAL_CONTROL_TRACKING:
    case MOCOB200_GOAL_SUM_SQUARED_STATE: vals[0] = w * integral; break;
    case MOCOB200_GOAL_CONTACT_TRACKING: vals[0] = w * (integral / dp[6]); break; // MocoContactTrackingGoal.h:220-223
    case MOCOB200_GOAL_FINAL_TIME: vals[0] = w * tf; break;
    case MOCOB200_GOAL_AVERAGE_SPEED: {
        double v = dp[3] - displacement() / (tf - t0);
        if (cost) v = v * v;
        vals[0] = w * v;
    } break;
    case MOCOB200_GOAL_PERIODICITY: { // MocoPeriodicityGoal.cpp:100-127
        const int nsp = ip[0], ncp = ip[1];
        for (int k = 0; k < nsp; ++k) {
            double v = x0[ip[2 + 2 * k]] * dp[3 + k] - xf[ip[3 + 2 * k]];
            if (cost) v = v * v;
            vals[k] = w * v;
        }
        for (int k = 0; k < ncp; ++k) {
            double v = c0[ip[2 + 2 * nsp + 2 * k]] * dp[3 + nsp + k] - cf[ip[3 + 2 * nsp + 2 * k]];
            if (cost) v = v * v;
            vals[nsp + k] = w * v;
        }
    } break;
    case MOCOB200_GOAL_INITIAL_ACTIVATION: { // MocoInitialActivationGoal.cpp:41-60
        const int n = ip[0];
        for (int k = 0; k < n; ++k) {
            double v = c0[ip[1 + 2 * k]] - x0[ip[2 + 2 * k]];
            if (cost) v = v * v;
            vals[k] = w * v;
        }
    } break;
    case MOCOB200_GOAL_INITIAL_VELOCITY_EQUILIBRIUM_DGF: { // MocoInitialVelocityEquilibriumDGFGoal.cpp:36-55
        const int n = ip[0];
        Scratch S;
        kinematics(T, x0, x0 + nq, S);
        const double* auxder = d0 + (T.implicit ? nq : 0);
        for (int k = 0; k < n; ++k) {
            MuscleOut mo = muscle_force(T, ip[1 + k], x0, x0 + nq, x0 + 2 * nq, c0, auxder, S, false, true);
            double v = mo.linResidDeriv;
            if (cost) v = v * v;
            vals[k] = w * v;
        }
    } break;
    default: vals[0] = NAN; break;
    }
}

} // namespace mocob200
