// Warp-cooperative point function (FP64, sm_100a): ONE WARP evaluates the fused CasOC point function at one
// (grid point, FD direction); all scratch lives in the warp's slice of shared memory and the 32 lanes are mapped to
// the entities of each phase:
//   kinematics        bodies of one tree level in parallel (one lane per body), level by level
//   forces            one lane per muscle (path geometry, SimmSpline, De Groote–Fregly curves) and per contact
//   gather            one lane per body / per coordinate: deterministic fixed-order sums of the point forces
//   dynamics          recursive Newton–Euler bias forces + composite-rigid-body mass matrix (origin-referenced,
//                     children pulled by their parent) + LDL^T solve (explicit) or the RNEA residual (implicit)
//   integrands        lanes over the terms of each goal, shuffle-tree reduction (fixed order)
// Same quantities as MocoCasOCProblem.h:203-356 computes through SimTK::State (see point_eval.cuh for the
// per-equation citations; the scalar curve / spline functions are shared with it). The thread-per-evaluation
// articulated-body version in point_eval.cuh is kept as an independent formulation for cross-checks.
#pragma once
#include "point_eval.cuh"

namespace mocob200 {

// Topology tables derived from mocob200_tables at create time (device pointers) + the per-warp scratch layout.
struct WarpAux {
    int nlevels;
    const int* level_start;   // [nlevels+1]: bodies are stored level by level (Simbody tree order)
    const int* child_start;   // [nb+1]
    const int* child_list;
    const int* body_pp_start; // [nb+1]: path points attached to each body
    const int* body_pp_list;
    const int* dof_body;      // [nq]
    const int* body_flags;    // [nb]: bit 0 = R_PF is the identity, bit 1 = R_BM is the identity
    const int* pp_mpp;        // [npp]: ordinal among the moving path points, or -1
    const int* mpp_pp;        // [nmpp]: path point index of each moving point
    const int* mpp_coord;     // [nmpp]
    int nmpp, npairs;
    const int* pair_i;        // mass-matrix entries (i >= j, dof j on the path from dof i to ground)
    const int* pair_j;
    // scratch layout, offsets in doubles
    int o_in, o_out, o_out2, o_R, o_p, o_w, o_v, o_cg, o_oM, o_Io, o_F, o_N, o_al, o_ac, o_cm, o_ch, o_cJ, o_a, o_l, o_tau,
        o_taub, o_lin, o_ang, o_M, o_ppos, o_prel, o_pf, o_dp, o_dpl, o_mloc, o_ref, o_mact, o_cF, o_cP;
    // shared-memory staging of the tables (per block): dbl[0, st_ndbl) (bodies, path points, muscles, contacts,
    // splines), the goal payload dbl[st_goal0, st_goal0 + st_ngoal), the whole int pool, and the topology arrays above
    // (one int pool starting at aux_base). st_doubles = size of the staged region in doubles.
    int st_ndbl, st_goal0, st_ngoal, st_nint, st_naux, st_doubles;
    const int* aux_base;
    // Finite-difference work lists (one-sided schemes). Directions are point-function input indices.
    //   p0_dirs: full evaluations (coordinates), run beside the nominal one;
    //   p1_dirs: partial evaluations that reload the nominal poses (speeds, implicit-mode accelerations);
    //   p2_dirs: partial evaluations that reload poses and velocities (time, anything unclassified);
    //   loc_*:   inputs acting through one muscle / coordinate actuator, whose columns the nominal warp writes itself
    //            (point_warp_local.cuh); dir_class[dir] == 3 marks them for k_fd1_finalize.
    int n_p0, n_p1, n_p2, n_cd, n_local, time_integrands_only;
    const int* p0_dirs; const int* p1_dirs; const int* p2_dirs;
    const int* cd_dirs; // every non-local direction (central scheme, thread-per-evaluation kernel)
    const int* cw_dirs; int n_cw; // the non-local directions that need whole evaluations (cd_dirs without p1_dirs)
    const int* loc_dir; const int* loc_kind; const int* loc_owner; const int* loc_role;
    const int* dir_class;
    int o_mlen, o_mspd, o_mout, o_Rm;
    int n_lparts; // one-sided local columns: warps per grid point, each taking the columns of 32 owners (muscles, actuators)
    int o_cjs; // per custom joint (position-dependent): 6 axis vectors in ground (18), df[6], d2f[6]
    int lockstep;
    int o_pk; // prescribed kinematics: q, u, udot at the point's time
    int npp_ext, o_warc, o_mpath; // wrap cylinders: path points + tangent-point slots; helix lengths; per-muscle current path
    int poison; // development switch (MOCOB200_POISON_SCRATCH=1): every evaluation starts from NaN-filled scratch
    int ws_doubles;
    int wpb; // warps (= evaluations in flight) per block, chosen at create from the scratch size
    // The position-dependent part of the scratch [blob_begin, blob_pos_end) and the velocity-dependent part
    // [blob_pos_end, blob_end) of the NOMINAL evaluation are saved per grid point ("blob") and reloaded by the
    // evaluations of FD directions that leave q (or q and u) untouched.
    int blob_begin, blob_pos_end, blob_end;
};

__device__ __forceinline__ V3 ldv(const double* p) { return mk(p[0], p[1], p[2]); }
__device__ __forceinline__ void stv(double* p, V3 a) { p[0] = a.x; p[1] = a.y; p[2] = a.z; }

struct WarpCtx {
    const DevTables& T;
    const WarpAux& A;
    double* ws;
    int lane;
    __device__ double* at(int off) const { return ws + off; }
};

// ---- phase 1a: pose of body b (one lane): R_GB, origin, joint point oM, joint axes (angular a, linear l acting at oM)
__device__ inline void wk_body_pose(const WarpCtx& C, int b, const double* q) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
    const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
    const int par = bi[0], kind = bi[1], q0 = bi[2];
    const int flags = A.body_flags[b];
    double RGF[9];
    V3 oF;
    if (par >= 0) {
        const double* Rp = C.at(A.o_R) + 9 * par;
        if (flags & 1) {
#pragma unroll
            for (int k = 0; k < 9; ++k) RGF[k] = Rp[k];
        } else mm3(Rp, bd, RGF);
        oF = ldv(C.at(A.o_p) + 3 * par) + rmul(Rp, ld3(bd + 9));
    } else {
#pragma unroll
        for (int k = 0; k < 9; ++k) RGF[k] = bd[k];
        oF = ld3(bd + 9);
    }
    const V3 xF = mk(RGF[0], RGF[3], RGF[6]), yF = mk(RGF[1], RGF[4], RGF[7]), zF = mk(RGF[2], RGF[5], RGF[8]);
    double RGM[9];
    V3 oM = oF;
    double* ax = C.at(A.o_a);
    double* lx = C.at(A.o_l);
    if (kind == MOCOB200_JOINT_CUSTOM) {
        // function-based joint (point_eval.cuh: custom_joint_eval): transform, motion axes, and what the velocity pass
        // needs for the velocity-product acceleration (axis vectors in ground, first and second function derivatives)
        const int cj = T.idx[T.o_body_cj + b];
        CustomJointEval E;
        custom_joint_eval(T, cj, q, E);
        mm3(RGF, E.R, RGM);
        oM = oF + rmul(RGF, E.p);
        double* cs = C.at(A.o_cjs) + 30 * cj;
        for (int i = 0; i < 6; ++i) { stv(cs + 3 * i, rmul(RGF, E.g[i])); cs[18 + i] = E.df[i]; cs[24 + i] = E.d2f[i]; }
        for (int k = 0; k < bi[3]; ++k) {
            V3 a, l;
            custom_joint_axis(E, q0 + k, a, l);
            stv(ax + 3 * (q0 + k), rmul(RGF, a));
            stv(lx + 3 * (q0 + k), rmul(RGF, l));
        }
    } else if (kind == MOCOB200_JOINT_PIN || kind == MOCOB200_JOINT_PLANAR) {
        double s, c;
        sincos(q[q0], &s, &c);
#pragma unroll
        for (int i = 0; i < 3; ++i) { // R_GM = R_GF Rz(theta)
            RGM[3 * i] = RGF[3 * i] * c + RGF[3 * i + 1] * s;
            RGM[3 * i + 1] = -RGF[3 * i] * s + RGF[3 * i + 1] * c;
            RGM[3 * i + 2] = RGF[3 * i + 2];
        }
        stv(ax + 3 * q0, zF);
        stv(lx + 3 * q0, mk(0, 0, 0));
        if (kind == MOCOB200_JOINT_PLANAR) {
            oM = oF + q[q0 + 1] * xF + q[q0 + 2] * yF;
            stv(ax + 3 * (q0 + 1), mk(0, 0, 0));
            stv(lx + 3 * (q0 + 1), xF);
            stv(ax + 3 * (q0 + 2), mk(0, 0, 0));
            stv(lx + 3 * (q0 + 2), yF);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 9; ++k) RGM[k] = RGF[k];
        if (kind == MOCOB200_JOINT_SLIDER) {
            oM = oF + q[q0] * xF;
            stv(ax + 3 * q0, mk(0, 0, 0));
            stv(lx + 3 * q0, xF);
        }
    }
    double* Rb = C.at(A.o_R) + 9 * b;
    if (flags & 2) {
#pragma unroll
        for (int k = 0; k < 9; ++k) Rb[k] = RGM[k];
    } else mmt3(RGM, bd + 12, Rb); // R_GB = R_GM R_BM^T
    stv(C.at(A.o_p) + 3 * b, oM - rmul(Rb, ld3(bd + 21)));
    stv(C.at(A.o_oM) + 3 * b, oM);
}
// ---- phase 1b: velocity and spatial acceleration of body b from the stored pose (one lane). The acceleration holds
// the velocity-product terms plus the joint accelerations udot when given; it needs no forces, so the recursive part
// of the Newton-Euler outward pass rides along here. Used unchanged by the full and the pose-reusing evaluations.
__device__ inline void wk_body_velocity(const WarpCtx& C, int b, const double* u, const double* udot) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
    const int par = bi[0], kind = bi[1], q0 = bi[2];
    V3 wP = mk(0, 0, 0), vP = mk(0, 0, 0), pP = mk(0, 0, 0), alP = mk(0, 0, 0), aP = mk(0, 0, 0);
    if (par >= 0) {
        pP = ldv(C.at(A.o_p) + 3 * par);
        wP = ldv(C.at(A.o_w) + 3 * par);
        vP = ldv(C.at(A.o_v) + 3 * par);
        alP = ldv(C.at(A.o_al) + 3 * par);
        aP = ldv(C.at(A.o_ac) + 3 * par);
    }
    V3 wrel = mk(0, 0, 0), vrel = mk(0, 0, 0), wrd = mk(0, 0, 0), vrd = mk(0, 0, 0);
    V3 bw = mk(0, 0, 0), bv = mk(0, 0, 0); // velocity-product acceleration of a function-based joint's own motion
    const double* ax = C.at(A.o_a) + 3 * q0;
    const double* lx = C.at(A.o_l) + 3 * q0;
    if (kind == MOCOB200_JOINT_CUSTOM) {
        for (int k = 0; k < bi[3]; ++k) {
            const V3 a = ldv(ax + 3 * k), l = ldv(lx + 3 * k);
            wrel = wrel + u[q0 + k] * a;
            vrel = vrel + u[q0 + k] * l;
            if (udot) { wrd = wrd + udot[q0 + k] * a; vrd = vrd + udot[q0 + k] * l; }
        }
        const int cj = T.idx[T.o_body_cj + b];
        const int* rec = T.idx + T.o_cj_i + cj * MOCOB200_CJ_ISTRIDE;
        const double* cs = C.at(A.o_cjs) + 30 * cj;
        V3 w[3];
        for (int i = 0; i < 3; ++i) {
            const double ui = rec[3 * i] != MOCOB200_FN_CONSTANT ? u[rec[3 * i + 1]] : 0.0;
            const V3 g = ldv(cs + 3 * i);
            w[i] = (cs[18 + i] * ui) * g;
            bw = bw + (cs[24 + i] * ui * ui) * g;
        }
        bw = bw + cross(w[0], w[1]) + cross(cj_one_coordinate(rec) ? w[0] : w[0] + w[1], w[2]); // point_eval.cuh: custom_joint_bias
        for (int i = 3; i < 6; ++i) {
            const double ui = rec[3 * i] != MOCOB200_FN_CONSTANT ? u[rec[3 * i + 1]] : 0.0;
            bv = bv + (cs[24 + i] * ui * ui) * ldv(cs + 3 * i);
        }
    } else if (kind == MOCOB200_JOINT_PIN || kind == MOCOB200_JOINT_PLANAR) {
        const V3 zF = ldv(ax);
        wrel = u[q0] * zF;
        if (udot) wrd = udot[q0] * zF;
        if (kind == MOCOB200_JOINT_PLANAR) {
            const V3 xF = ldv(lx + 3), yF = ldv(lx + 6);
            vrel = u[q0 + 1] * xF + u[q0 + 2] * yF;
            if (udot) vrd = udot[q0 + 1] * xF + udot[q0 + 2] * yF;
        }
    } else if (kind == MOCOB200_JOINT_SLIDER) {
        const V3 xF = ldv(lx);
        vrel = u[q0] * xF;
        if (udot) vrd = udot[q0] * xF;
    }
    const V3 oM = ldv(C.at(A.o_oM) + 3 * b), pB = ldv(C.at(A.o_p) + 3 * b);
    const V3 wB = wP + wrel;
    const V3 r = oM - pP, sv = pB - oM;
    const V3 vM = vP + cross(wP, r) + vrel;
    stv(C.at(A.o_w) + 3 * b, wB);
    stv(C.at(A.o_v) + 3 * b, vM + cross(wB, sv));
    // the joint axes are fixed in the parent: d(axis)/dt = wP x axis
    V3 al = alP + cross(wP, wrel) + wrd;
    V3 aM = aP + cross(alP, r) + cross(wP, cross(wP, r)) + 2.0 * cross(wP, vrel) + vrd;
    if (kind == MOCOB200_JOINT_CUSTOM) { al = al + bw; aM = aM + bv; }
    stv(C.at(A.o_al) + 3 * b, al);
    stv(C.at(A.o_ac) + 3 * b, aM + cross(al, sv) + cross(wB, cross(wB, sv)));
}
// non-recursive per-body quantities (all bodies in parallel): mass-centre offset and inertia about the origin, in ground
__device__ inline void wk_body_inertia(const WarpCtx& C, int b) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
    const double* Rb = C.at(A.o_R) + 9 * b;
    stv(C.at(A.o_cg) + 3 * b, rmul(Rb, ld3(bd + 25)));
    const double* in = bd + 28;
    const double I[9] = {in[0], in[3], in[4], in[3], in[1], in[5], in[4], in[5], in[2]};
    double RI[9], RIRt[9];
    mm3(Rb, I, RI);
    mmt3(RI, Rb, RIRt);
    double* Io = C.at(A.o_Io) + 6 * b;
    Io[0] = RIRt[0]; Io[1] = RIRt[4]; Io[2] = RIRt[8];
    Io[3] = 0.5 * (RIRt[1] + RIRt[3]); Io[4] = 0.5 * (RIRt[2] + RIRt[6]); Io[5] = 0.5 * (RIRt[5] + RIRt[7]);
}

__device__ __forceinline__ V3 wk_point_velocity(const WarpCtx& C, int b, V3 pointG) {
    if (b < 0) return mk(0, 0, 0);
    return ldv(C.at(C.A.o_v) + 3 * b) + cross(ldv(C.at(C.A.o_w) + 3 * b), pointG - ldv(C.at(C.A.o_p) + 3 * b));
}

// ---- phase 2a: one contact (one lane): force on the sphere body + point of application ---------------------------
__device__ inline void wk_contact(const WarpCtx& C, int ci) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const double* cd = T.dbl + T.o_contact_d + ci * MOCOB200_CONTACT_DSTRIDE;
    const int* cidx = T.idx + T.o_contact_i + ci * MOCOB200_CONTACT_ISTRIDE;
    const int bs = cidx[0], bh = cidx[1];
    const double radius = cd[3];
    const double* pr = cd + 16;
    const double diss = pr[1], us = pr[2], ud = pr[3], uv = pr[4], vt = pr[5], cf = pr[6], bd = pr[7], bv = pr[8];
    V3 nrm, oH;
    {
        const V3 xh = mk(cd[4], cd[7], cd[10]);
        const V3 ph = ld3(cd + 13);
        if (bh >= 0) {
            const double* Rh = C.at(A.o_R) + 9 * bh;
            nrm = -1.0 * rmul(Rh, xh);
            oH = ldv(C.at(A.o_p) + 3 * bh) + rmul(Rh, ph);
        } else { nrm = -1.0 * xh; oH = ph; }
    }
    const V3 centre = (bs >= 0) ? ldv(C.at(A.o_p) + 3 * bs) + rmul(C.at(A.o_R) + 9 * bs, ld3(cd)) : ld3(cd);
    const double indentation = radius - dot(centre - oH, nrm);
    const V3 cp = centre - (radius - 0.5 * indentation) * nrm;
    const V3 vrel = wk_point_velocity(C, bs, cp) - wk_point_velocity(C, bh, cp);
    const double vnormal = dot(vrel, nrm);
    const V3 vtan = vrel - vnormal * nrm;
    const double indVel = -vnormal;
    const double ind = sqrt(indentation * indentation + cf);
    const double fH = cd[DEV_CONTACT_HERTZ] * (ind * sqrt(ind)); // (4/3) k sqrt(radius k) ind^(3/2)
    const double fHs = fH * (0.5 + 0.5 * tanh(bd * indentation)) + cf;
    const double fHC = fHs * (1.0 + 1.5 * diss * indVel);
    const double fn = fHC * (0.5 + 0.5 * tanh(bv * (indVel + 2.0 / (3.0 * diss))));
    const double vslip = sqrt(dot(vtan, vtan) + cf);
    const double vr = vslip / vt;
    const double ff = fn * (fmin(vr, 1.0) * (ud + 2.0 * (us - ud) / (1.0 + vr * vr)) + uv * vslip);
    stv(C.at(A.o_cF) + 3 * ci, fn * nrm + (-ff / vslip) * vtan);
    stv(C.at(A.o_cP) + 3 * ci, cp);
}

// ---- phase 0: SimmSpline evaluations of the moving path points, one per lane: t = 6 * (moving point) + c, c < 3 the
// location components at the clamped coordinate, c >= 3 the derivative components at the unclamped coordinate
// (MovingPathPoint.cpp:216-315: the location clamps the coordinate to the spline's range, the derivatives do not)
__device__ inline void wk_moving_point_spline(const WarpCtx& C, int t, const double* q) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const int j = t / 6, c = t - 6 * j;
    const int pp = A.mpp_pp[j];
    const int* pi = T.idx + T.o_pp_i + pp * MOCOB200_PP_ISTRIDE;
    const double* pd = T.dbl + T.o_pp_d + pp * MOCOB200_PP_DSTRIDE;
    const double qv = q[pi[6]];
    if (c < 3) C.at(A.o_mloc)[3 * j + c] = simm_spline(T, pi[3 + c], fmin(fmax(qv, pd[3]), pd[4]), 0);
    else C.at(A.o_dpl)[3 * j + c - 3] = simm_spline(T, pi[c], qv, 1);
}

// ---- wrap cylinders. The two tangent points of pairing k (global number) are stations of the cylinder's body in
// slots npp + 2 k and npp + 2 k + 1 of ppos / prel / pf; the muscle's current path (GeometryPath "current_path") is kept as
// bytes in o_mpath: [count, entries...], entry >= 0 = the muscle's path point number, entry < 0 = -(2 w + 1 + which) for
// pairing w of the muscle.
__device__ __forceinline__ int wk_muscle_wrap_count(const WarpCtx& C, int m) {
    return C.T.npathwrap > 0 ? C.T.idx[C.T.o_pw_i + 2 * m + 1] : 0;
}
// global slot (into ppos / prel / pf) and body of a path entry of muscle m
__device__ __forceinline__ int wk_path_slot(const WarpCtx& C, int m, int entry, int& body, int& wrapObj) {
    const DevTables& T = C.T;
    if (entry >= 0) {
        const int pp = T.idx[T.o_mus_i + m * MOCOB200_MUS_ISTRIDE] + entry;
        body = T.idx[T.o_pp_i + pp * MOCOB200_PP_ISTRIDE];
        wrapObj = -1;
        return pp;
    }
    const int w = (-entry - 1) >> 1, which = (-entry - 1) & 1;
    const int k = T.idx[T.o_pw_i + 2 * m] + w; // global pairing
    wrapObj = T.idx[T.o_pw_i + 2 * T.nmus + k * MOCOB200_PW_ISTRIDE];
    body = T.idx[T.o_wo_i + wrapObj * MOCOB200_WO_ISTRIDE];
    return T.npp + 2 * k + which;
}
__device__ inline void wk_muscle_wraps(const WarpCtx& C, int m, unsigned active) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    signed char* mp = reinterpret_cast<signed char*>(C.at(A.o_mpath) + 4 * m);
    const int nW = T.idx[T.o_pw_i + 2 * m + 1];
    if (nW == 0) { mp[0] = 0; return; }
    for (int w = 0; w < 2 * nW; ++w) { // tangent points not on the path keep a finite position: the per-body gather multiplies it by a zero force
        const int slot = T.npp + 2 * T.idx[T.o_pw_i + 2 * m] + w;
        stv(C.at(A.o_prel) + 3 * slot, mk(0, 0, 0));
        stv(C.at(A.o_ppos) + 3 * slot, mk(0, 0, 0));
    }
    const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
    const int first = mi[0], count = mi[1];
    V3 pos[MOCOB200_MAXPPM];
    int orig[MOCOB200_MAXPPM], na = 0;
    for (int k = 0; k < count && na < MOCOB200_MAXPPM; ++k)
        if ((active >> k) & 1u) { pos[na] = ldv(C.at(A.o_ppos) + 3 * (first + k)); orig[na] = k; ++na; }
    int path[MOCOB200_MAXPPM + 2 * MOCOB200_MAXWRAP], np = na;
    for (int k = 0; k < na; ++k) path[k] = k;
    V3 wr1[MOCOB200_MAXWRAP], wr2[MOCOB200_MAXWRAP];
    double warc[MOCOB200_MAXWRAP];
    struct Frames {
        const WarpCtx& C;
        __device__ const double* R(int b) const { return C.at(C.A.o_R) + 9 * b; }
        __device__ V3 p(int b) const { return ldv(C.at(C.A.o_p) + 3 * b); }
    };
    apply_wrap_objects(T, m, Frames{C}, na, count, orig, pos, path, np, wr1, wr2, warc);
    // store: path entries (active-list indices -> path point numbers), tangent points as stations of their bodies, helix lengths
    const int pw0 = T.idx[T.o_pw_i + 2 * m];
    mp[0] = (signed char)np;
    for (int j = 0; j < np && j < 31; ++j) mp[1 + j] = (signed char)(path[j] >= 0 ? orig[path[j]] : path[j]);
    for (int j = 0; j < np; ++j) {
        if (path[j] >= 0) continue;
        const int w = (-path[j] - 1) >> 1, which = (-path[j] - 1) & 1;
        int body, wo;
        const int slot = wk_path_slot(C, m, path[j], body, wo);
        V3 r = which ? wr2[w] : wr1[w], pG = r;
        if (body >= 0) { r = rmul(C.at(A.o_R) + 9 * body, r); pG = ldv(C.at(A.o_p) + 3 * body) + r; }
        stv(C.at(A.o_prel) + 3 * slot, r);
        stv(C.at(A.o_ppos) + 3 * slot, pG);
        if (which) C.at(A.o_warc)[pw0 + w] = warc[w];
    }
}

// ---- phase 2b: path geometry of one muscle (one lane): active points, their offsets from the body origin (prel),
// ground positions (ppos) and, for moving points, dP/dq in the body frame (dpl). Position-dependent only.
// (GeometryPath.cpp:817-846; ConditionalPathPoint.cpp:160-169; MovingPathPoint.cpp:216-315)
// WRAP: this instantiation evaluates wrap cylinders (kept out of the kernels of models without them: the wrap procedure nearly
// doubles their register count and costs config 3 10 %)
template <bool WRAP>
__device__ inline void wk_muscle_geometry(const WarpCtx& C, int m, const double* q) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
    const int first = mi[0], count = mi[1];
    unsigned active = 0; // bit k: point k of this muscle is on the current path
    for (int k = 0; k < count; ++k) {
        const int pp = first + k;
        const int* pi = T.idx + T.o_pp_i + pp * MOCOB200_PP_ISTRIDE;
        const double* pd = T.dbl + T.o_pp_d + pp * MOCOB200_PP_DSTRIDE;
        const int b = pi[0], kind = pi[1];
        V3 loc = ld3(pd);
        bool on = true;
        if (kind == MOCOB200_PP_CONDITIONAL) {
            const double val = q[pi[2]];
            // an inactive point still gets a (finite) position: the per-body gather multiplies it by a zero force
            on = (val >= pd[3] - 1e-5 && val <= pd[4] + 1e-5);
        } else if (kind == MOCOB200_PP_MOVING) {
            loc = ldv(C.at(A.o_mloc) + 3 * A.pp_mpp[pp]); // evaluated by wk_moving_point_splines
        }
        V3 r = loc, pG = loc;
        if (b >= 0) {
            r = rmul(C.at(A.o_R) + 9 * b, loc);
            pG = ldv(C.at(A.o_p) + 3 * b) + r;
        }
        stv(C.at(A.o_prel) + 3 * pp, r);
        stv(C.at(A.o_ppos) + 3 * pp, pG);
        if (on) active |= 1u << k;
    }
    C.at(A.o_mact)[m] = (double)active;
    if constexpr (WRAP) { if (T.npathwrap > 0) wk_muscle_wraps(C, m, active); }
}
// ---- phase 2c: one muscle (one lane) from its stored geometry: point velocities, length, lengthening speed, muscle
// model, per-point forces (GeometryPath.cpp:848-864, 1117-1145, 306-432; Point.cpp:130-140)
__device__ inline MuscleOut wk_muscle_dynamics(const WarpCtx& C, int m, const double* u, const double* z, const double* ctl,
                                               const double* auxder, bool apply, bool wantStiffness) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
    const double* md = T.dbl + T.o_mus_d + m * MOCOB200_MUS_DSTRIDE;
    const int first = mi[0], count = mi[1], flags = mi[5];
    const unsigned active = (unsigned)C.at(A.o_mact)[m];
    const double* ppos = C.at(A.o_ppos);
    const bool wrapped = wk_muscle_wrap_count(C, m) > 0;
    const signed char* mp = wrapped ? reinterpret_cast<const signed char*>(C.at(A.o_mpath) + 4 * m) : nullptr;
    const int nent = wrapped ? mp[0] : count; // wrapped: walk the stored current path; otherwise the path points in order
    double length = 0, speed = 0;
    V3 prevP = mk(0, 0, 0), prevV = mk(0, 0, 0);
    int na = 0, prevWo = -1;
    for (int k = 0; k < nent; ++k) {
        int pp, b, wo = -1;
        if (wrapped) pp = wk_path_slot(C, m, mp[1 + k], b, wo);
        else {
            if (!((active >> k) & 1u)) continue;
            pp = first + k;
            b = T.idx[T.o_pp_i + pp * MOCOB200_PP_ISTRIDE];
        }
        const int* pi = T.idx + T.o_pp_i + (pp < T.npp ? pp : 0) * MOCOB200_PP_ISTRIDE;
        const int mc = (pp < T.npp && pi[1] == MOCOB200_PP_MOVING) ? pi[6] : -1;
        const V3 pG = ldv(ppos + 3 * pp);
        V3 vG = mk(0, 0, 0);
        if (b >= 0) vG = ldv(C.at(A.o_v) + 3 * b) + cross(ldv(C.at(A.o_w) + 3 * b), ldv(C.at(A.o_prel) + 3 * pp));
        if (mc >= 0) {
            const int j = A.pp_mpp[pp];
            const V3 dpl = ldv(C.at(A.o_dpl) + 3 * j);
            V3 dpg = dpl, vloc = u[mc] * dpl;
            if (b >= 0) { const double* Rb = C.at(A.o_R) + 9 * b; vloc = rmul(Rb, vloc); dpg = rmul(Rb, dpl); }
            vG = vG + vloc;
            stv(C.at(A.o_dp) + 3 * j, dpg);
        }
        if (na > 0) {
            const V3 r = prevP - pG, dv = prevV - vG;
            const double d = sqrt(dot(r, r));
            // between the two tangent points of one cylinder the path runs over the surface: the helix length
            length += (wo >= 0 && wo == prevWo) ? C.at(A.o_warc)[(pp - T.npp) >> 1] : d;
            speed += (d < 2.220446049250313e-16) ? sqrt(dot(dv, dv)) : dot(dv, (1.0 / d) * r);
        }
        prevP = pG;
        prevV = vG;
        prevWo = wo;
        ++na;
    }
    C.at(A.o_mlen)[m] = length;
    C.at(A.o_mspd)[m] = speed;
    const double activation = (flags & MOCOB200_MUS_FLAG_IGNORE_ACTIVATION_DYNAMICS) ? ctl[mi[2]] : z[mi[3]];
    const double ntf = (mi[4] >= 0) ? z[mi[4]] : NAN;
    const double ntfd = (mi[6] >= 0) ? auxder[mi[6]] : NAN;
    MuscleOut o = dgf_muscle(md, flags, length, speed, activation, ntf, ntfd, wantStiffness);
    if (apply) {
        // net force of the path on the body at every active point: +T dir at the start of a segment, -T dir at its
        // end; segments between points of the same body carry no force
        double* pf = C.at(A.o_pf);
        int prev = -1, pa = -1, ba = -1;
        for (int k = 0; k < nent; ++k) {
            int pb, bb, wo;
            if (wrapped) pb = wk_path_slot(C, m, mp[1 + k], bb, wo);
            else {
                if (!((active >> k) & 1u)) continue;
                pb = first + k;
                bb = T.idx[T.o_pp_i + pb * MOCOB200_PP_ISTRIDE];
            }
            if (prev >= 0) {
                if (ba != bb) {
                    V3 dir = ldv(ppos + 3 * pb) - ldv(ppos + 3 * pa);
                    const double nn = sqrt(dot(dir, dir));
                    dir = (nn < 1.9984014443252818e-14) ? (NAN * dir) : ((1.0 / nn) * dir);
                    const V3 f = o.tendonForce * dir;
                    stv(pf + 3 * pa, ldv(pf + 3 * pa) + f);
                    stv(pf + 3 * pb, ldv(pf + 3 * pb) - f);
                }
            }
            prev = k; pa = pb; ba = bb;
        }
    }
    return o;
}

} // namespace mocob200
