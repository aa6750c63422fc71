// TEST INFRASTRUCTURE ONLY — CPU oracle: multibody kinematics / dynamics and applied forces.
// See orc_point.hpp for the formulation; reference call sites are cited per function.
// PINNED to numbers the reference produced (Simbody's own source is not under /root/reference): accelerations of the planar
// gait model along the reference's CMC run (tests/test_reference_pins.py), and the accelerations / actuator forces of the
// full-body RRA and CMC runs through the implicit residuals (tests/test_reference_dynamics_pins.py: 3e-6 ... 2.3e-4 N m).
#pragma once
#include <cstring>

#include "orc_point.hpp"

namespace orc {

constexpr int kMaxBodies = 64, kMaxDof = 64, kMaxContacts = 32;

template <class S>
struct Frames { // everything expressed in ground
    Mat3<S> R[kMaxBodies];      // body orientation
    Vec3<S> origin[kMaxBodies]; // body origin
    Vec3<S> jointPoint[kMaxBodies]; // origin of the mobilizer's M frame (where the joint axes act)
    Mat3<S> RF[kMaxBodies];     // orientation of the parent-side joint frame F
    Vec3<S> omega[kMaxBodies], vOrigin[kMaxBodies];
    Vec3<S> comOffset[kMaxBodies]; // R * com
    // per dof: unit twist of the joint (angular axis, linear axis acting at jointPoint), in ground
    Vec3<S> axA[kMaxDof], axL[kMaxDof];
    int dofBody[kMaxDof];
    // velocity-product acceleration of the joint's own motion (d/dt of the axes relative to F, times the speeds), in ground:
    // zero for the constant-axis joints, non-zero for function-based (custom) joints
    Vec3<S> biasA[kMaxBodies], biasL[kMaxBodies];
};

// A wrench accumulated about the GROUND ORIGIN.
template <class S>
struct Wrench {
    Vec3<S> f, m;
    void addForceAt(const Vec3<S>& point, const Vec3<S>& force) { f = f + force; m = m + cross(point, force); }
};

// joint axes in the F frame, acting at the M origin. Conventions of the Simbody mobilizers that
// opensim-core's joints create (SimbodyEngine/PinJoint.cpp:40-44, SliderJoint.cpp, PlanarJoint.cpp:40-44,
// WeldJoint.cpp): Pin = rotation about Z; Slider = translation along X; Planar = (rotation about Z,
// translation along X, translation along Y of F); Weld = none.
inline int jointDofs(const int* bodyRec) { return bodyRec[3]; }
inline void jointAxis(int kind, int k, double a[3], double l[3]) {
    a[0] = a[1] = a[2] = l[0] = l[1] = l[2] = 0;
    if (kind == MOCOB200_JOINT_PIN) a[2] = 1;
    else if (kind == MOCOB200_JOINT_SLIDER) l[0] = 1;
    else if (kind == MOCOB200_JOINT_PLANAR) { if (k == 0) a[2] = 1; else l[k - 1] = 1; }
}

// Scalar function of a custom joint's TransformAxis and its first derivative (Common/Constant.cpp, LinearFunction.cpp:
// SimTK::Function::Linear c0 x + c1, SimmSpline.cpp:511-685 inside MultiplierFunction.cpp), any scalar type X.
template <class X>
void axisFunction(const Model& M, const int* rec /*kind | q | spline*/, const double* d /*axis(3) c0 c1*/, const X& q, X& value, X& slope) {
    if (rec[0] == MOCOB200_FN_LINEAR) { value = X(d[3]) * q + X(d[4]); slope = X(d[3]); }
    else if (rec[0] == MOCOB200_FN_SIMMSPLINE) { value = simmSpline<X>(M, rec[2], q, false); slope = simmSpline<X>(M, rec[2], q, true); }
    else { value = X(d[4]); slope = X(0.0); }
}

// Function-based mobilizer (CustomJoint.cpp:205-273 -> SimTK::MobilizedBody::FunctionBased): X_FM(q) =
// [ R(A0, f0) R(A1, f1) R(A2, f2) , A3 f3 + A4 f4 + A5 f5 ], u = qdot. Written from that definition alone: the motion axes
// are the partial derivatives of the joint's spatial velocity (angular: successive rotation axes carried by the preceding
// rotations; linear: the translation axes), and their time derivatives come from evaluating the same expressions on dual
// numbers seeded with the joint speeds. Outputs in the F frame.
template <class S>
void customJoint(const Model& M, int cj, int q0, int nd, const S* q, const S* u, Mat3<S>& R_FM, Vec3<S>& p_FM, Vec3<S>* axA, Vec3<S>* axL,
                 Vec3<S>& biasA, Vec3<S>& biasL) {
    typedef Dual<S> X;
    const double* d = M.cjD(cj);
    const int* rec = M.cjI(cj);
    X f[6], df[6];
    for (int a = 0; a < 6; ++a) {
        const int qi = rec[3 * a + 1];
        const X arg = qi >= 0 ? X(q[qi], u[qi]) : X(0.0);
        axisFunction<X>(M, rec + 3 * a, d + 5 * a, arg, f[a], df[a]);
    }
    // rotation axes in F: W0 = A0, W1 = R0 A1, W2 = R0 R1 A2
    const Mat3<X> R0 = rotAxis<X>(f[0], d[0], d[1], d[2]), R1 = rotAxis<X>(f[1], d[5], d[6], d[7]), R2 = rotAxis<X>(f[2], d[10], d[11], d[12]);
    const Mat3<X> R01 = R0 * R1, R = R01 * R2;
    const Vec3<X> W[3] = {Vec3<X>(X(d[0]), X(d[1]), X(d[2])), R0 * Vec3<X>(X(d[5]), X(d[6]), X(d[7])), R01 * Vec3<X>(X(d[10]), X(d[11]), X(d[12]))};
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) R_FM.m[r][c] = R.m[r][c].v;
    Vec3<X> p;
    for (int a = 3; a < 6; ++a) p = p + f[a] * Vec3<X>(X(d[5 * a]), X(d[5 * a + 1]), X(d[5 * a + 2]));
    p_FM = Vec3<S>(p.x.v, p.y.v, p.z.v);
    biasA = Vec3<S>(); biasL = Vec3<S>();
    for (int k = 0; k < nd; ++k) {
        Vec3<X> a, l;
        for (int i = 0; i < 3; ++i) if (rec[3 * i + 1] == q0 + k) a = a + df[i] * W[i];
        for (int i = 3; i < 6; ++i) if (rec[3 * i + 1] == q0 + k) l = l + df[i] * Vec3<X>(X(d[5 * i]), X(d[5 * i + 1]), X(d[5 * i + 2]));
        axA[k] = Vec3<S>(a.x.v, a.y.v, a.z.v);
        axL[k] = Vec3<S>(l.x.v, l.y.v, l.z.v);
        biasA = biasA + u[q0 + k] * Vec3<S>(a.x.d, a.y.d, a.z.d);
        biasL = biasL + u[q0 + k] * Vec3<S>(l.x.d, l.y.d, l.z.d);
    }
    // What precedes is d/dt of the motion axes, exactly. The reference's mobilizer (Simbody 3.7, a dependency whose source is
    // not under /root/reference) differs from it in one case that its own records pin (tests/test_reference_dynamics_pins.py):
    // in a joint with ONE coordinate driving all of its axes (the Rajagopal knee: rotation2 and rotation3 are splines of
    // the flexion angle) the term (th1' W1) x (th2' W2) is absent from its velocity-product acceleration. With the term the
    // reference's RRA and CMC records of that model are missed by up to 0.8 N m at the hip; without it they are met to
    // 1e-4 N m. Joints whose rotation axes have a coordinate each (hips, shoulders, lumbar) meet those records with the exact
    // expression; no record covers other ways of sharing coordinates, and they keep the exact expression as well.
    if (nd == 1) {
        const int qa = rec[3 * 1 + 1], qb = rec[3 * 2 + 1];
        if (rec[3 * 1] != MOCOB200_FN_CONSTANT && rec[3 * 2] != MOCOB200_FN_CONSTANT && qa >= 0 && qb >= 0) {
            const Vec3<S> W1(W[1].x.v, W[1].y.v, W[1].z.v), W2(W[2].x.v, W[2].y.v, W[2].z.v);
            biasA = biasA - (df[1].v * u[qa] * df[2].v * u[qb]) * cross(W1, W2);
        }
    }
}

// Position and velocity kinematics, base to tip.
template <class S>
void kinematics(const Model& M, const S* q, const S* u, Frames<S>& K) {
    for (int b = 0; b < M.nb; ++b) {
        const double* d = M.bodyD(b);
        const int* bi = M.bodyI(b);
        const int parent = bi[0], kind = bi[1], q0 = bi[2];
        const Mat3<S> R_PF = Mat3<S>::fromRowMajor(d), R_BM = Mat3<S>::fromRowMajor(d + 12);
        const Vec3<S> p_PF{S(d[9]), S(d[10]), S(d[11])}, p_BM{S(d[21]), S(d[22]), S(d[23])};
        Mat3<S> R_GP; Vec3<S> oP, wP, vP;
        if (parent >= 0) { R_GP = K.R[parent]; oP = K.origin[parent]; wP = K.omega[parent]; vP = K.vOrigin[parent]; }
        const Mat3<S> R_GF = R_GP * R_PF;
        const Vec3<S> oF = oP + R_GP * p_PF;
        // joint transform X_FM(q)
        Mat3<S> R_FM; Vec3<S> p_FM;
        Vec3<S> cAxA[6], cAxL[6], cBiasA, cBiasL;
        if (kind == MOCOB200_JOINT_CUSTOM) customJoint<S>(M, M.customOf(b), q0, bi[3], q, u, R_FM, p_FM, cAxA, cAxL, cBiasA, cBiasL);
        else if (kind == MOCOB200_JOINT_PIN) R_FM = rotZ(q[q0]);
        else if (kind == MOCOB200_JOINT_SLIDER) p_FM = Vec3<S>(q[q0], S(0.0), S(0.0));
        else if (kind == MOCOB200_JOINT_PLANAR) { R_FM = rotZ(q[q0]); p_FM = Vec3<S>(q[q0 + 1], q[q0 + 2], S(0.0)); }
        const Mat3<S> R_GM = R_GF * R_FM;
        const Vec3<S> oM = oF + R_GF * p_FM;
        K.RF[b] = R_GF;
        K.R[b] = R_GM * transpose(R_BM);
        K.jointPoint[b] = oM;
        K.origin[b] = oM - K.R[b] * p_BM;
        // velocity: relative angular velocity and relative velocity of the M origin, both fixed in F
        Vec3<S> wRel, vRel;
        const int nd = jointDofs(bi);
        K.biasA[b] = kind == MOCOB200_JOINT_CUSTOM ? R_GF * cBiasA : Vec3<S>();
        K.biasL[b] = kind == MOCOB200_JOINT_CUSTOM ? R_GF * cBiasL : Vec3<S>();
        for (int k = 0; k < nd; ++k) {
            if (kind == MOCOB200_JOINT_CUSTOM) {
                K.axA[q0 + k] = R_GF * cAxA[k];
                K.axL[q0 + k] = R_GF * cAxL[k];
            } else {
                double a[3], l[3];
                jointAxis(kind, k, a, l);
                K.axA[q0 + k] = R_GF * Vec3<S>(S(a[0]), S(a[1]), S(a[2]));
                K.axL[q0 + k] = R_GF * Vec3<S>(S(l[0]), S(l[1]), S(l[2]));
            }
            K.dofBody[q0 + k] = b;
            wRel = wRel + u[q0 + k] * K.axA[q0 + k];
            vRel = vRel + u[q0 + k] * K.axL[q0 + k];
        }
        K.omega[b] = wP + wRel;
        const Vec3<S> vM = vP + cross(wP, oM - oP) + vRel;
        K.vOrigin[b] = vM + cross(K.omega[b], K.origin[b] - oM);
        K.comOffset[b] = K.R[b] * Vec3<S>(S(d[25]), S(d[26]), S(d[27]));
    }
}

template <class S> Vec3<S> stationPosition(const Frames<S>& K, int body, const Vec3<S>& local) {
    return body < 0 ? local : K.origin[body] + K.R[body] * local;
}
template <class S> Vec3<S> pointVelocity(const Frames<S>& K, int body, const Vec3<S>& pointInGround) {
    if (body < 0) return Vec3<S>();
    return K.vOrigin[body] + cross(K.omega[body], pointInGround - K.origin[body]);
}

// Inverse dynamics by Newton–Euler about the mass centres.
//   tau_k = (joint axis k) . (wrench that the joint must transmit)
// where the transmitted wrench = sum over the subtree of (m a_c, Ic alpha + w x Ic w) - applied.
// `applied[b]`: wrench applied to body b about the ground origin (may be null = none).
// withVelocity=false drops every velocity-dependent term (used to build the mass matrix).
template <class S>
void inverseDynamics(const Model& M, const Frames<S>& K, const S* u, const S* udot, const Wrench<S>* applied,
                     bool withVelocity, S* tau) {
    Vec3<S> alpha[kMaxBodies], aOrigin[kMaxBodies];
    Wrench<S> need[kMaxBodies];
    for (int b = 0; b < M.nb; ++b) {
        const double* d = M.bodyD(b);
        const int* bi = M.bodyI(b);
        const int parent = bi[0], q0 = bi[2], nd = jointDofs(bi);
        Vec3<S> alP, aP, wP, oP;
        if (parent >= 0) { alP = alpha[parent]; aP = aOrigin[parent]; wP = K.omega[parent]; oP = K.origin[parent]; }
        Vec3<S> wRel, vRel, wRelDot, vRelDot;
        for (int k = 0; k < nd; ++k) {
            if (withVelocity) { wRel = wRel + u[q0 + k] * K.axA[q0 + k]; vRel = vRel + u[q0 + k] * K.axL[q0 + k]; }
            wRelDot = wRelDot + udot[q0 + k] * K.axA[q0 + k];
            vRelDot = vRelDot + udot[q0 + k] * K.axL[q0 + k];
        }
        const Vec3<S> r = K.jointPoint[b] - oP, s = K.origin[b] - K.jointPoint[b];
        // axes are fixed in the parent: d/dt(axis) = wP x axis
        Vec3<S> al = alP + wRelDot;
        Vec3<S> aM = aP + cross(alP, r) + vRelDot;
        if (withVelocity) {
            al = al + cross(wP, wRel) + K.biasA[b];
            aM = aM + cross(wP, cross(wP, r)) + S(2.0) * cross(wP, vRel) + K.biasL[b];
        }
        Vec3<S> aB = aM + cross(al, s);
        const Vec3<S> w = K.omega[b];
        if (withVelocity) aB = aB + cross(w, cross(w, s));
        alpha[b] = al;
        aOrigin[b] = aB;
        // mass properties: inertia about the mass centre, in ground
        const S mass(d[24]);
        const Vec3<S> cLocal{S(d[25]), S(d[26]), S(d[27])};
        Mat3<S> Io;
        Io.m[0][0] = S(d[28]); Io.m[1][1] = S(d[29]); Io.m[2][2] = S(d[30]);
        Io.m[0][1] = Io.m[1][0] = S(d[31]); Io.m[0][2] = Io.m[2][0] = S(d[32]); Io.m[1][2] = Io.m[2][1] = S(d[33]);
        const Mat3<S> PA = parallelAxis(mass, cLocal);
        Mat3<S> IcLocal;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) IcLocal.m[i][j] = Io.m[i][j] - PA.m[i][j];
        const Mat3<S> Ic = K.R[b] * IcLocal * transpose(K.R[b]);
        const Vec3<S> cg = K.comOffset[b];
        Vec3<S> aC = aB + cross(al, cg);
        if (withVelocity) aC = aC + cross(w, cross(w, cg));
        const Vec3<S> F = mass * aC;
        Vec3<S> N = Ic * al;
        if (withVelocity) N = N + cross(w, Ic * w);
        Wrench<S> W;
        W.f = F;
        W.m = N + cross(K.origin[b] + cg, F);
        if (applied) { W.f = W.f - applied[b].f; W.m = W.m - applied[b].m; }
        need[b] = W;
    }
    for (int b = M.nb - 1; b >= 0; --b) {
        const int* bi = M.bodyI(b);
        const int parent = bi[0], q0 = bi[2], nd = jointDofs(bi);
        const Vec3<S> mAtJoint = need[b].m - cross(K.jointPoint[b], need[b].f);
        for (int k = 0; k < nd; ++k) tau[q0 + k] = dot(K.axA[q0 + k], mAtJoint) + dot(K.axL[q0 + k], need[b].f);
        if (parent >= 0) { need[parent].f = need[parent].f + need[b].f; need[parent].m = need[parent].m + need[b].m; }
    }
}

// Joint-space mass matrix by the composite-rigid-body method, all quantities about the ground origin.
template <class S>
void massMatrix(const Model& M, const Frames<S>& K, S* Mass /* nq x nq row-major */) {
    const int n = M.nq;
    S cm[kMaxBodies]; Vec3<S> ch[kMaxBodies]; Mat3<S> cJ[kMaxBodies]; // composite mass, first moment, inertia about ground origin
    for (int b = 0; b < M.nb; ++b) {
        const double* d = M.bodyD(b);
        const S mass(d[24]);
        const Vec3<S> cLocal{S(d[25]), S(d[26]), S(d[27])};
        Mat3<S> Io;
        Io.m[0][0] = S(d[28]); Io.m[1][1] = S(d[29]); Io.m[2][2] = S(d[30]);
        Io.m[0][1] = Io.m[1][0] = S(d[31]); Io.m[0][2] = Io.m[2][0] = S(d[32]); Io.m[1][2] = Io.m[2][1] = S(d[33]);
        const Mat3<S> PA = parallelAxis(mass, cLocal);
        Mat3<S> IcLocal;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) IcLocal.m[i][j] = Io.m[i][j] - PA.m[i][j];
        const Vec3<S> c = K.origin[b] + K.comOffset[b];
        cm[b] = mass;
        ch[b] = mass * c;
        cJ[b] = K.R[b] * IcLocal * transpose(K.R[b]) + parallelAxis(mass, c);
    }
    for (int b = M.nb - 1; b >= 0; --b) {
        const int parent = M.bodyI(b)[0];
        if (parent < 0) continue;
        cm[parent] = cm[parent] + cm[b];
        ch[parent] = ch[parent] + ch[b];
        cJ[parent] = cJ[parent] + cJ[b];
    }
    for (int i = 0; i < n * n; ++i) Mass[i] = S(0.0);
    for (int i = 0; i < n; ++i) {
        const int b = K.dofBody[i];
        // unit twist of dof i about the ground origin: angular a, linear (velocity of the point at the origin) v0
        const Vec3<S> a = K.axA[i];
        const Vec3<S> v0 = K.axL[i] - cross(a, K.jointPoint[b]);
        // momentum of the composite body moving with that twist
        const Vec3<S> lin = cm[b] * v0 + cross(a, ch[b]);
        const Vec3<S> ang = cJ[b] * a + cross(ch[b], v0);
        // project on dof i and on every dof of b's joint with index <= i, then on all ancestors' dofs
        int bb = b;
        while (bb >= 0) {
            const int* bi = M.bodyI(bb);
            const int q0 = bi[2], nd = jointDofs(bi);
            for (int k = 0; k < nd; ++k) {
                const int j = q0 + k;
                if (j > i) continue;
                const Vec3<S> aj = K.axA[j];
                const Vec3<S> v0j = K.axL[j] - cross(aj, K.jointPoint[bb]);
                const S mij = dot(aj, ang) + dot(v0j, lin);
                Mass[i * n + j] = mij;
                Mass[j * n + i] = mij;
            }
            bb = bi[0];
        }
    }
}

// Solve Mass * x = rhs by Cholesky (Mass is symmetric positive definite); Mass is overwritten.
template <class S>
void choleskySolve(int n, S* A, const S* rhs, S* x) {
    for (int j = 0; j < n; ++j) {
        S d = A[j * n + j];
        for (int k = 0; k < j; ++k) d = d - A[j * n + k] * A[j * n + k];
        d = sqrt(d);
        A[j * n + j] = d;
        for (int i = j + 1; i < n; ++i) {
            S v = A[i * n + j];
            for (int k = 0; k < j; ++k) v = v - A[i * n + k] * A[j * n + k];
            A[i * n + j] = v / d;
        }
    }
    for (int i = 0; i < n; ++i) {
        S v = rhs[i];
        for (int k = 0; k < i; ++k) v = v - A[i * n + k] * x[k];
        x[i] = v / A[i * n + i];
    }
    for (int i = n - 1; i >= 0; --i) {
        S v = x[i];
        for (int k = i + 1; k < n; ++k) v = v - A[k * n + i] * x[k];
        x[i] = v / A[i * n + i];
    }
}

} // namespace orc
