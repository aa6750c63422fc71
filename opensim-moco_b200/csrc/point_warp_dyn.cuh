// Warp-cooperative point function, part 2: gather of applied forces, dynamics, integrands, and the driver.
#pragma once
#include "point_warp.cuh"

namespace mocob200 {

__device__ __forceinline__ V3 sym6mul(const double* s, V3 v) { return symmul(s, v); }

// Phase stopwatch for development builds (-DMOCOB200_PROFILE_PHASES): warp 0 of block 0 records clock64() at the
// phase boundaries of its evaluation. Compiled out of the product.
#ifdef MOCOB200_PROFILE_PHASES
__device__ long long g_phase_clk[32];
__device__ long long g_item_clk[4][16384]; // [phase][item]: cycles of each fd1 work item
#define PHASE_MARK(k) do { if (lockstep) __syncthreads(); __syncwarp(); if (blockIdx.x == 0 && threadIdx.x == 0) g_phase_clk[k] = clock64(); } while (0)
#else
#define PHASE_MARK(k) do { if (lockstep) __syncthreads(); } while (0)
#endif

// ---- phase 3: applied wrench on body b about its origin (one lane per body), fixed summation order -------------
__device__ inline void wk_gather_body(const WarpCtx& C, int b) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
    const V3 pB = ldv(C.at(A.o_p) + 3 * b);
    const V3 g = ld3(T.dbl + T.o_gravity);
    const double m = bd[24];
    V3 F = m * g;                                        // gravity at the mass centre (Model.cpp:587-590)
    V3 N = cross(ldv(C.at(A.o_cg) + 3 * b), m * g);
    const double* pf = C.at(A.o_pf);
    const double* ppos = C.at(A.o_ppos);
    for (int k = A.body_pp_start[b]; k < A.body_pp_start[b + 1]; ++k) {
        const int pp = A.body_pp_list[k];
        const V3 f = ldv(pf + 3 * pp);
        F = F + f;
        N = N + cross(ldv(ppos + 3 * pp) - pB, f);
    }
    for (int f = 0; f < T.nxforce; ++f) { // ExternalForce::computeForce (Model/ExternalForce.cpp:330-355); functions of time in o_ref
        const int* xi = T.idx + T.o_xforce_i + f * MOCOB200_XFORCE_ISTRIDE;
        if (xi[0] != b) continue;
        const double* ref = C.at(A.o_ref);
        if (xi[3] >= 0) {
            V3 force = mk(ref[xi[3]], ref[xi[4]], ref[xi[5]]);
            if (xi[1] >= 0) force = rmul(C.at(A.o_R) + 9 * xi[1], force);
            V3 arm = mk(0, 0, 0);
            if (xi[6] >= 0) {
                V3 point = mk(ref[xi[6]], ref[xi[7]], ref[xi[8]]);
                if (xi[2] >= 0) point = ldv(C.at(A.o_p) + 3 * xi[2]) + rmul(C.at(A.o_R) + 9 * xi[2], point);
                arm = point - pB;
            }
            F = F + force;
            N = N + cross(arm, force);
        }
        if (xi[9] >= 0) {
            V3 torque = mk(ref[xi[9]], ref[xi[10]], ref[xi[11]]);
            if (xi[1] >= 0) torque = rmul(C.at(A.o_R) + 9 * xi[1], torque);
            N = N + torque;
        }
    }
    for (int c = 0; c < T.ncontact; ++c) {
        const int* cidx = T.idx + T.o_contact_i + c * MOCOB200_CONTACT_ISTRIDE;
        if (cidx[0] != b && cidx[1] != b) continue;
        V3 f = ldv(C.at(A.o_cF) + 3 * c);
        if (cidx[0] != b) f = -1.0 * f;
        F = F + f;
        N = N + cross(ldv(C.at(A.o_cP) + 3 * c) - pB, f);
    }
    // wrench (about the body origin) that the joint must supply = inertial - applied
    const V3 cg = ldv(C.at(A.o_cg) + 3 * b), w = ldv(C.at(A.o_w) + 3 * b), al = ldv(C.at(A.o_al) + 3 * b), aB = ldv(C.at(A.o_ac) + 3 * b);
    const double* Io = C.at(A.o_Io) + 6 * b;
    stv(C.at(A.o_F) + 3 * b, m * (aB + cross(al, cg) + cross(w, cross(w, cg))) - F);
    stv(C.at(A.o_N) + 3 * b, sym6mul(Io, al) + cross(cg, m * aB) + cross(w, sym6mul(Io, w)) - N);
}
// applied mobility force on coordinate k (one lane per coordinate)
__device__ inline void wk_gather_coord(const WarpCtx& C, int k, const double* ctl) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    double tau = 0;
    for (int j = 0; j < A.nmpp; ++j) // moving path points driven by this coordinate: (dP/dq) . (force at the point)
        if (A.mpp_coord[j] == k) tau += dot(ldv(C.at(A.o_dp) + 3 * j), ldv(C.at(A.o_pf) + 3 * A.mpp_pp[j]));
    for (int a = 0; a < T.ncact; ++a) { // CoordinateActuator.cpp:185-204
        const int* ai = T.idx + T.o_cact_i + a * MOCOB200_CACT_ISTRIDE;
        if (ai[0] == k) tau += T.dbl[T.o_cact_d + a * MOCOB200_CACT_DSTRIDE] * ctl[ai[1]];
    }
    C.at(A.o_tau)[k] = tau;
}

// ---- phase 4b: inward pass for body b (its children are final): pull the children's wrenches and composite
// inertias to this body's origin, then project the wrench on the joint axes.
__device__ inline void wk_inward(const WarpCtx& C, int b, bool composite) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
    const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
    const int q0 = bi[2], nd = bi[3];
    const V3 pB = ldv(C.at(A.o_p) + 3 * b);
    V3 F = ldv(C.at(A.o_F) + 3 * b), N = ldv(C.at(A.o_N) + 3 * b);
    double cm = bd[24];
    V3 ch = cm * ldv(C.at(A.o_cg) + 3 * b);
    double cJ[6];
    if (composite) {
#pragma unroll
        for (int k = 0; k < 6; ++k) cJ[k] = C.at(A.o_Io)[6 * b + k];
    }
    for (int k = A.child_start[b]; k < A.child_start[b + 1]; ++k) {
        const int c = A.child_list[k];
        const V3 r = ldv(C.at(A.o_p) + 3 * c) - pB;
        const V3 Fc = ldv(C.at(A.o_F) + 3 * c);
        F = F + Fc;
        N = N + ldv(C.at(A.o_N) + 3 * c) + cross(r, Fc);
        if (composite) {
            // shift the child's composite (mass mc, first moment hc, inertia Jc about its origin) by r
            const double mc = C.at(A.o_cm)[c];
            const V3 hc = ldv(C.at(A.o_ch) + 3 * c);
            const double* Jc = C.at(A.o_cJ) + 6 * c;
            const double d = 2.0 * dot(hc, r) + mc * dot(r, r);
            cJ[0] += Jc[0] + d - 2.0 * hc.x * r.x - mc * r.x * r.x;
            cJ[1] += Jc[1] + d - 2.0 * hc.y * r.y - mc * r.y * r.y;
            cJ[2] += Jc[2] + d - 2.0 * hc.z * r.z - mc * r.z * r.z;
            cJ[3] += Jc[3] - (hc.x * r.y + r.x * hc.y) - mc * r.x * r.y;
            cJ[4] += Jc[4] - (hc.x * r.z + r.x * hc.z) - mc * r.x * r.z;
            cJ[5] += Jc[5] - (hc.y * r.z + r.y * hc.z) - mc * r.y * r.z;
            ch = ch + hc + mc * r;
            cm += mc;
        }
    }
    stv(C.at(A.o_F) + 3 * b, F);
    stv(C.at(A.o_N) + 3 * b, N);
    if (composite) {
        C.at(A.o_cm)[b] = cm;
        stv(C.at(A.o_ch) + 3 * b, ch);
#pragma unroll
        for (int k = 0; k < 6; ++k) C.at(A.o_cJ)[6 * b + k] = cJ[k];
    }
    const V3 oM = ldv(C.at(A.o_oM) + 3 * b);
    const V3 NM = N + cross(pB - oM, F); // moment about the joint point
    for (int k = 0; k < nd; ++k)
        C.at(A.o_taub)[q0 + k] = dot(ldv(C.at(A.o_a) + 3 * (q0 + k)), NM) + dot(ldv(C.at(A.o_l) + 3 * (q0 + k)), F);
}

// ---- phase 4c: momentum of dof i's composite body under a unit rate of dof i (about the body origin) ------------
__device__ inline void wk_dof_momentum(const WarpCtx& C, int i) {
    const WarpAux& A = C.A;
    const int b = A.dof_body[i];
    const V3 a = ldv(C.at(A.o_a) + 3 * i);
    const V3 v0 = ldv(C.at(A.o_l) + 3 * i) + cross(a, ldv(C.at(A.o_p) + 3 * b) - ldv(C.at(A.o_oM) + 3 * b));
    const double m = C.at(A.o_cm)[b];
    const V3 h = ldv(C.at(A.o_ch) + 3 * b);
    stv(C.at(A.o_lin) + 3 * i, m * v0 + cross(a, h));
    stv(C.at(A.o_ang) + 3 * i, sym6mul(C.at(A.o_cJ) + 6 * b, a) + cross(h, v0));
}
__device__ inline void wk_mass_entry(const WarpCtx& C, int i, int j, int nq) {
    const WarpAux& A = C.A;
    const int bi = A.dof_body[i], bj = A.dof_body[j];
    const V3 aj = ldv(C.at(A.o_a) + 3 * j);
    const V3 vj = ldv(C.at(A.o_l) + 3 * j) + cross(aj, ldv(C.at(A.o_p) + 3 * bi) - ldv(C.at(A.o_oM) + 3 * bj));
    const double mij = dot(aj, ldv(C.at(A.o_ang) + 3 * i)) + dot(vj, ldv(C.at(A.o_lin) + 3 * i));
    C.at(A.o_M)[i * nq + j] = mij;
    C.at(A.o_M)[j * nq + i] = mij;
}

// ---- phase 4d: LDL^T of the mass matrix (whole warp), in place in shared memory. Right-looking: at step j the lanes
// first normalise column j (keeping the un-normalised entries in row j of the unused upper triangle), then update the
// trailing lower triangle, one (row, column) pair per lane. The factor (unit-lower L below the diagonal, D on it)
// depends on q only, so evaluations of FD directions that leave q untouched reuse the nominal factor and only run
// wk_ldlt_solve.
__device__ inline void wk_ldlt_factor(const WarpCtx& C, int n) {
    double* M = C.at(C.A.o_M);
    const int lane = C.lane;
    for (int j = 0; j < n - 1; ++j) {
        const double dj = M[j * n + j];
        for (int i = j + 1 + lane; i < n; i += 32) {
            const double a = M[i * n + j];
            M[j * n + i] = a;
            M[i * n + j] = a / dj;
        }
        __syncwarp();
        const int m = n - 1 - j;
        for (int p = lane; p < m * m; p += 32) {
            const int r = p / m, c = p - r * m;
            if (c <= r) {
                const int i = j + 1 + r, k = j + 1 + c;
                M[i * n + k] -= M[i * n + j] * M[j * n + k];
            }
        }
        __syncwarp();
    }
}
// x = M^-1 rhs from the stored factor (whole warp): L y = rhs, z = y / D, L^T x = z.
__device__ inline void wk_ldlt_solve(const WarpCtx& C, int n, const double* rhs, double* x) {
    const double* M = C.at(C.A.o_M);
    const int lane = C.lane;
    const unsigned full = 0xffffffffu;
    if (n > 32) { // one lane, plain substitution
        if (lane == 0) {
            for (int i = 0; i < n; ++i) { double v = rhs[i]; for (int k = 0; k < i; ++k) v -= M[i * n + k] * x[k]; x[i] = v; }
            for (int i = 0; i < n; ++i) x[i] /= M[i * n + i];
            for (int i = n - 1; i >= 0; --i) { double v = x[i]; for (int k = i + 1; k < n; ++k) v -= M[k * n + i] * x[k]; x[i] = v; }
        }
        __syncwarp();
        return;
    }
    const int row = lane < n ? lane : 0;
    double y = lane < n ? rhs[row] : 0.0;
    for (int j = 0; j < n - 1; ++j) { // y_j is final; eliminate it from rows i > j
        const double yj = __shfl_sync(full, y, j);
        if (lane > j && lane < n) y -= M[lane * n + j] * yj;
    }
    double xv = y / M[row * n + row];
    for (int j = n - 1; j > 0; --j) { // x_j is final; eliminate it from rows i < j
        const double xj = __shfl_sync(full, xv, j);
        if (lane < j) xv -= M[j * n + lane] * xj;
    }
    if (lane < n) x[lane] = xv;
    __syncwarp();
}

__device__ __forceinline__ double warp_sum(double v) { // fixed butterfly order => deterministic
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// The fused point function, evaluated by one warp. Inputs at ws[o_in..], outputs to ws[o_out..] (only the groups
// selected by mask are written): bit 0 = dynamics (udot|residual, zdot, aux residuals), bit 1+k = integrand k.
// reuse: 0 = compute everything; 1 = the position-dependent scratch (poses, path geometry, mass-matrix factor) is
// already in place (same q as the evaluation that produced it); 2 = the velocity-dependent scratch as well (same q, u
// and, in implicit mode, the same accelerations). The reused and the recomputed quantities come from the same device
// functions, so a reusing evaluation is bit-identical to a full one.
// time_moved: the evaluation's time differs from the one whose tracking-reference values sit in the scratch.
// lockstep: every warp of the block runs the same kind of evaluation (same mask and reuse), so they meet at a block
// barrier between phases and share their instruction fetches (the code is long and mostly straight-line).
template <int reuse, bool WRAP = false>
__device__ __forceinline__ void warp_eval_point_t(const DevTables& T, const WarpAux& A, double* ws, unsigned mask, bool time_moved,
                                                  bool lockstep) {
    const WarpCtx C{T, A, ws, (int)(threadIdx.x & 31)};
    const int lane = C.lane;
    const int nq = T.nq, nz = T.nz, nc = T.nc, nb = T.nb;
    const double* in = ws + A.o_in;
    double* out = ws + A.o_out;
    const double time = in[0];
    const int nqs = T.prescribed ? 0 : nq; // coordinate states among the inputs (none with prescribed kinematics)
    const double* states = in + 1;         // the NLP's state vector at this point (goals index into it)
    const double* q = in + 1;
    const double* u = q + nqs;
    const double* z = u + nqs;
    const double* ctl = z + nz;
    const double* mult = ctl + nc;
    const double* der = mult + T.nm;
    const double* auxder = der + ((T.implicit && !T.prescribed) ? nq : 0);
    if (T.prescribed) { // q, u, udot = the coordinates' functions of time (Components/PositionMotion.cpp:44-75)
        double* pk = ws + A.o_pk;
        if (reuse == 0 || time_moved)
            for (int k = lane; k < nq; k += 32) ppoly_eval2(T, T.idx[T.o_presc_i + k], time, pk[k], pk[nq + k], pk[2 * nq + k]);
        __syncwarp();
        q = pk; u = pk + nq; der = pk + 2 * nq;
    }
    const bool dyn = mask & 1u;
    bool need_contacts = dyn;
    {
        int gi = 0;
        for (int g2 = 0; g2 < T.ngoal; ++g2) {
            const int* gr = T.idx + T.o_goal_i + g2 * MOCOB200_GOAL_ISTRIDE;
            if (!gr[3]) continue;
            if (((mask >> (1 + gi)) & 1u) && gr[0] == MOCOB200_GOAL_CONTACT_TRACKING) need_contacts = true;
            ++gi;
        }
    }
    PHASE_MARK(0);
    // lane-parallel table look-ups first: tracking references at this time, splines of the moving path points at this q
    if (((mask >> 1) || (dyn && T.nxforce > 0)) && (reuse == 0 || time_moved))
        for (int r = lane; r < T.nppoly; r += 32) ws[A.o_ref + r] = ppoly_eval(T, r, time);
    if (dyn && reuse == 0)
        for (int t = lane; t < 6 * A.nmpp; t += 32) wk_moving_point_spline(C, t, q);
    __syncwarp();
    PHASE_MARK(11);
    if (need_contacts) {
        if (reuse < 2) {
            const double* udot = T.implicit ? der : nullptr;
            for (int L = 0; L < A.nlevels; ++L) {
                for (int b = A.level_start[L] + lane; b < A.level_start[L + 1]; b += 32) {
                    if (reuse == 0) wk_body_pose(C, b, q);
                    wk_body_velocity(C, b, u, udot);
                }
                __syncwarp();
            }
        }
        PHASE_MARK(1);
        if (reuse == 0) {
            for (int b = lane; b < nb; b += 32) wk_body_inertia(C, b);
            PHASE_MARK(2);
            if (dyn) for (int m = lane; m < T.nmus; m += 32) wk_muscle_geometry<WRAP>(C, m, q);
        }
        PHASE_MARK(3);
        for (int k = lane; k < 3 * A.npp_ext; k += 32) ws[A.o_pf + k] = 0.0;
        __syncwarp();
        const int nent = (dyn ? T.nmus : 0) + (reuse < 2 ? T.ncontact : 0);
        for (int e = lane; e < nent; e += 32) {
            if (e >= (dyn ? T.nmus : 0)) wk_contact(C, e - (dyn ? T.nmus : 0));
            else {
                const int m = e;
                const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
                const double* md = T.dbl + T.o_mus_d + m * MOCOB200_MUS_DSTRIDE;
                const MuscleOut mo = wk_muscle_dynamics(C, m, u, z, ctl, auxder, true, false);
                double* zdot = out + nq;
                double* auxres = zdot + nz;
                const double adot = mi[3] >= 0 ? activation_dot(ctl[mi[2]], z[mi[3]], md[5], md[6]) : 0.0;
                if (mi[3] >= 0) zdot[mi[3]] = adot;
                if (mi[4] >= 0) zdot[mi[4]] = mo.nfDeriv;
                if (mi[7] >= 0) auxres[mi[7]] = mo.residual;
                double* keep = ws + A.o_mout + 4 * m; // for the local Jacobian columns (point_warp_local.cuh)
                keep[0] = mo.tendonForce; keep[1] = mo.nfDeriv; keep[2] = mo.residual; keep[3] = adot;
            }
        }
        __syncwarp();
    }
    PHASE_MARK(4);
    if (dyn) {
        for (int b = lane; b < nb; b += 32) wk_gather_body(C, b);
        for (int k = lane; k < nq; k += 32) wk_gather_coord(C, k, ctl);
        __syncwarp();
        if (T.ncoupler > 0) { // constraint forces of the multipliers (few constraints: one lane)
            if (lane == 0) coupler_forces(T, q, mult, ws + A.o_tau);
            __syncwarp();
        }
        PHASE_MARK(5);
        const bool composite = !T.implicit && reuse == 0;
        for (int L = A.nlevels - 1; L >= 0; --L) {
            for (int b = A.level_start[L] + lane; b < A.level_start[L + 1]; b += 32) wk_inward(C, b, composite);
            __syncwarp();
        }
        PHASE_MARK(6);
        if (T.implicit) {
            for (int k = lane; k < nq; k += 32) out[k] = ws[A.o_tau + k] - ws[A.o_taub + k];
        } else {
            if (composite) {
                for (int k = lane; k < nq * nq; k += 32) ws[A.o_M + k] = 0.0;
                for (int i = lane; i < nq; i += 32) wk_dof_momentum(C, i);
                __syncwarp();
                for (int e = lane; e < A.npairs; e += 32) wk_mass_entry(C, A.pair_i[e], A.pair_j[e], nq);
                __syncwarp();
                PHASE_MARK(7);
                wk_ldlt_factor(C, nq);
                PHASE_MARK(8);
            }
            for (int i = lane; i < nq; i += 32) ws[A.o_tau + i] -= ws[A.o_taub + i]; // right-hand side
            __syncwarp();
            wk_ldlt_solve(C, nq, ws + A.o_tau, out);
        }
        __syncwarp();
        if (T.n_kc > 0)
            for (int k = lane; k < T.ncoupler; k += 32) coupler_errors(T, k, q, u, T.implicit ? der : out, out + nq + nz + T.n_auxres);
        __syncwarp();
    }
    PHASE_MARK(9);
    // goal integrands (MocoControlGoal.cpp:121-131, MocoStateTrackingGoal.cpp:103-117, MocoContactTrackingGoal.cpp:240-287)
    double* integ = out + nq + nz + T.n_auxres + T.n_kc;
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
            for (int k = lane; k < n; k += 32) {
                const double c = ctl[ip[3 + k]];
                val += dp[3 + k] * (expo == 2 ? c * c : abs_ipow(c, expo));
            }
        } break;
        case MOCOB200_GOAL_STATE_TRACKING:
            for (int k = lane; k < ip[0]; k += 32) {
                const double e = states[ip[1 + 2 * k]] - ws[A.o_ref + ip[2 + 2 * k]];
                val += dp[3 + k] * (e * e);
            }
            break;
        case MOCOB200_GOAL_CONTROL_TRACKING:
            for (int k = lane; k < ip[0]; k += 32) {
                const double e = ctl[ip[1 + 2 * k]] - ws[A.o_ref + ip[2 + 2 * k]];
                val += dp[3 + k] * (e * e);
            }
            break;
        case MOCOB200_GOAL_SUM_SQUARED_STATE:
            for (int k = lane; k < ip[0]; k += 32) {
                const double e = states[ip[1 + 2 * k]];
                val += dp[3 + k] * (e * e);
            }
            break;
        case MOCOB200_GOAL_CONTACT_TRACKING: {
            const int proj = ip[0], ng = ip[1];
            const V3 pv = ld3(dp + 3);
            int mygrp = -1; // lane k sums group k
            {
                const int* gp = ip + 2;
                for (int k = 0; k < ng; ++k) {
                    if (lane == k) mygrp = (int)(gp - ip);
                    gp += 1 + 2 * gp[0] + 3;
                }
            }
            if (mygrp >= 0) {
                const int* gp = ip + mygrp;
                const int ncon = gp[0];
                const double* ref = ws + A.o_ref;
                const int r0 = gp[1 + 2 * ncon], r1 = gp[2 + 2 * ncon], r2 = gp[3 + 2 * ncon];
                const V3 fr = mk(r0 >= 0 ? ref[r0] : 0.0, r1 >= 0 ? ref[r1] : 0.0, r2 >= 0 ? ref[r2] : 0.0);
                V3 fm = mk(0, 0, 0);
                for (int c = 0; c < ncon; ++c) {
                    const V3 f = ldv(ws + A.o_cF + 3 * gp[1 + 2 * c]);
                    fm = (gp[2 + 2 * c] == 0) ? fm + f : fm - f;
                }
                const V3 e3 = fm - fr;
                V3 e = e3;
                if (proj == 1) e = dot(e3, pv) * pv;
                else if (proj == 2) e = e3 - dot(e3, pv) * pv;
                val += dot(e, e);
            }
        } break;
        default: break;
        }
        val = warp_sum(val);
        if (lane == 0) integ[slot] = val;
    }
    __syncwarp();
    PHASE_MARK(10);
}
// run-time reuse level: one specialised copy of the evaluation per level
template <bool WRAP = false>
__device__ inline void warp_eval_point(const DevTables& T, const WarpAux& A, double* ws, unsigned mask, int reuse, bool time_moved,
                                       bool lockstep = false) {
    if (reuse == 0) warp_eval_point_t<0, WRAP>(T, A, ws, mask, time_moved, lockstep);
    else if (reuse == 1) warp_eval_point_t<1>(T, A, ws, mask, time_moved, lockstep);
    else warp_eval_point_t<2>(T, A, ws, mask, time_moved, lockstep);
}

} // namespace mocob200
