// Warp-cooperative point function, part 3: Jacobian columns of the inputs that act through ONE muscle or ONE
// coordinate actuator (auxiliary states, controls, auxiliary derivative variables).
//
// The multibody outputs are exactly linear in every actuator force: udot = M^-1 (tau_applied - bias) in explicit mode,
// residual = tau_applied - M udot - bias in implicit mode, and tau_applied = sum_m R_m(q) F_m + ... with R_m the
// generalized force of a unit tension in muscle m (minus its moment arms). For such an input x the finite difference
// of the point function therefore factors exactly:
//       (f(x + h) - f(x)) / h  =  M^-1 R_m (F_m(x + h) - F_m(x)) / h            (rows udot; R_m alone for residuals)
// and the rows of the muscle's own auxiliary dynamics are scalar finite differences of the muscle model. So the warp
// that evaluates the nominal point also produces these columns: R_m for all muscles (one lane each), one multi-right-
// hand-side solve with the LDL^T factor it already holds (one lane per right-hand side), then one lane per input for
// the scalar differences — instead of one full point evaluation per input. The step h and the one-sided formula are
// the caller's (CasADi's forward/backward scheme, CasOCFunction.h:38-44); only the order of the floating-point
// operations differs from differencing the whole function, and the result carries less cancellation error.
// Goal integrands are plain sums over inputs, so their rows are the differences of the one changed term.
#pragma once
#include "point_warp_dyn.cuh"

namespace mocob200 {

enum { MOCOB200_LOCAL_MUSCLE = 0, MOCOB200_LOCAL_ACTUATOR = 1 };
enum { MOCOB200_ROLE_CONTROL = 0, MOCOB200_ROLE_ACTIVATION = 1, MOCOB200_ROLE_TENDON_FORCE = 2, MOCOB200_ROLE_TENDON_FORCE_DERIV = 3 };

// R_m: generalized force of a unit tension in muscle m, from the stored path geometry (one lane). The tension pulls
// the two end points of every inter-body segment towards each other (GeometryPath.cpp:306-432); a force f at point P
// of body b loads every dof k between b and ground with a_k . ((P - oM_k) x f) + l_k . f, and a moving point adds
// (dP/dq) . f on its coordinate (MovingPathPoint.cpp:374-392).
__device__ inline void wk_unit_tension_force(const WarpCtx& C, int m, double* Rm) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const int nq = T.nq;
    const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
    const int first = mi[0], count = mi[1];
    const unsigned active = (unsigned)C.at(A.o_mact)[m];
    const double* ppos = C.at(A.o_ppos);
    for (int k = 0; k < nq; ++k) Rm[k] = 0.0;
    const bool wrapped = wk_muscle_wrap_count(C, m) > 0; // then walk the stored current path (tangent points are stations of the cylinders' bodies)
    const signed char* mp = wrapped ? reinterpret_cast<const signed char*>(C.at(A.o_mpath) + 4 * m) : nullptr;
    const int nent = wrapped ? mp[0] : count;
    const int fixedRec[MOCOB200_PP_ISTRIDE] = {-1, MOCOB200_PP_FIXED, -1, -1, -1, -1, -1, -1};
    int prev = -1, pa = -1, bodyA = -1;
    const int* ia = fixedRec;
    for (int k = 0; k < nent; ++k) {
        int pb, bodyB, wo;
        if (wrapped) pb = wk_path_slot(C, m, mp[1 + k], bodyB, wo);
        else {
            if (!((active >> k) & 1u)) continue;
            pb = first + k;
            bodyB = T.idx[T.o_pp_i + pb * MOCOB200_PP_ISTRIDE];
        }
        const int* ib = pb < T.npp ? T.idx + T.o_pp_i + pb * MOCOB200_PP_ISTRIDE : fixedRec; // a tangent point is never a moving point
        if (prev >= 0) {
            if (bodyA != bodyB) {
                V3 dir = ldv(ppos + 3 * pb) - ldv(ppos + 3 * pa);
                const double nn = sqrt(dot(dir, dir));
                dir = (nn < 1.9984014443252818e-14) ? (NAN * dir) : ((1.0 / nn) * dir);
                // the pair of opposite forces loads only the dofs between the two bodies and their common ancestor
                // (on the common part of the tree the two moments cancel exactly: dir is parallel to Pb - Pa); bodies are
                // stored parent-before-child, so the larger index is never an ancestor of the smaller
                const V3 Pa = ldv(ppos + 3 * pa), Pb = ldv(ppos + 3 * pb);
#pragma unroll 1
                for (int side = 0; side < 2; ++side) {
                    const int pp = side ? pb : pa;
                    const int* pi = side ? ib : ia;
                    if (pi[1] == MOCOB200_PP_MOVING) {
                        const V3 dpl = ldv(C.at(A.o_dpl) + 3 * A.pp_mpp[pp]);
                        const V3 dpg = pi[0] >= 0 ? rmul(C.at(A.o_R) + 9 * pi[0], dpl) : dpl;
                        Rm[pi[6]] += side ? -dot(dpg, dir) : dot(dpg, dir);
                    }
                }
                int ba = bodyA, bb = bodyB;
                while (ba != bb) {
                    const bool sideB = bb > ba;
                    const int b = sideB ? bb : ba;
                    const V3 f = sideB ? -1.0 * dir : dir;
                    const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
                    const V3 mom = cross((sideB ? Pb : Pa) - ldv(C.at(A.o_oM) + 3 * b), f);
                    for (int d = 0; d < bi[3]; ++d) {
                        const int dof = bi[2] + d;
                        Rm[dof] += dot(ldv(C.at(A.o_a) + 3 * dof), mom) + dot(ldv(C.at(A.o_l) + 3 * dof), f);
                    }
                    if (sideB) bb = bi[0]; else ba = bi[0];
                }
            }
        }
        prev = k; pa = pb; bodyA = bodyB; ia = ib;
    }
}

// x <- M^-1 x with the stored LDL^T factor, one lane, in place
__device__ inline void wk_ldlt_solve_lane(const double* M, int n, double* x) {
    for (int i = 1; i < n; ++i) { double v = x[i]; for (int k = 0; k < i; ++k) v -= M[i * n + k] * x[k]; x[i] = v; }
    for (int i = 0; i < n; ++i) x[i] /= M[i * n + i];
    for (int i = n - 2; i >= 0; --i) { double v = x[i]; for (int k = i + 1; k < n; ++k) v -= M[k * n + i] * x[k]; x[i] = v; }
}

// change of integrand `slot`'s value when input `dir` moves from v0 to v1 (all other inputs fixed): the one term of
// the goal's sum that reads it (MocoControlGoal.cpp:121-131, MocoStateTrackingGoal.cpp:103-117, ...).
__device__ inline double wk_integrand_term_delta(const DevTables& T, const WarpAux& A, const double* ws, const int* gr, int dir, double v0,
                                                 double v1) {
    const int* ip = T.idx + gr[4];
    const double* dp = T.dblG + gr[5];
    const int ns = (T.prescribed ? 0 : 2 * T.nq) + T.nz;
    const int state = (dir >= 1 && dir <= ns) ? dir - 1 : -1;               // index into the state vector
    const int control = (dir > ns && dir <= ns + T.nc) ? dir - 1 - ns : -1;
    double d = 0.0;
    switch (gr[0]) {
    case MOCOB200_GOAL_CONTROL: {
        const int expo = ip[0], n = ip[2];
        for (int k = 0; k < n; ++k)
            if (ip[3 + k] == control) {
                const double g1 = expo == 2 ? v1 * v1 : abs_ipow(v1, expo), g0 = expo == 2 ? v0 * v0 : abs_ipow(v0, expo);
                d += dp[3 + k] * g1 - dp[3 + k] * g0;
            }
    } break;
    case MOCOB200_GOAL_STATE_TRACKING:
        for (int k = 0; k < ip[0]; ++k)
            if (ip[1 + 2 * k] == state) {
                const double r = ws[A.o_ref + ip[2 + 2 * k]], e1 = v1 - r, e0 = v0 - r;
                d += dp[3 + k] * (e1 * e1) - dp[3 + k] * (e0 * e0);
            }
        break;
    case MOCOB200_GOAL_CONTROL_TRACKING:
        for (int k = 0; k < ip[0]; ++k)
            if (ip[1 + 2 * k] == control) {
                const double r = ws[A.o_ref + ip[2 + 2 * k]], e1 = v1 - r, e0 = v0 - r;
                d += dp[3 + k] * (e1 * e1) - dp[3 + k] * (e0 * e0);
            }
        break;
    case MOCOB200_GOAL_SUM_SQUARED_STATE:
        for (int k = 0; k < ip[0]; ++k)
            if (ip[1 + 2 * k] == state) d += dp[3 + k] * (v1 * v1) - dp[3 + k] * (v0 * v0);
        break;
    default: break; // contact tracking reads coordinates, speeds and time only
    }
    return d;
}

// Outputs of muscle m's own model with input `role` set to v (everything else nominal): tension, d(normalized tendon
// force)/dt, equilibrium residual, activation rate.
struct LocalMuscleEval { double F, nfDeriv, residual, actDot; };
__device__ inline LocalMuscleEval wk_local_muscle(const DevTables& T, const WarpAux& A, const double* ws, int m, int role, double v) {
    const int nq = T.nq;
    const double* in = ws + A.o_in;
    const double* z = in + 1 + (T.prescribed ? 0 : 2 * nq);
    const double* ctl = z + T.nz;
    const double* auxder = ctl + T.nc + T.nm + ((T.implicit && !T.prescribed) ? nq : 0);
    const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
    const double* md = T.dbl + T.o_mus_d + m * MOCOB200_MUS_DSTRIDE;
    const int flags = mi[5];
    double e = ctl[mi[2]], a = mi[3] >= 0 ? z[mi[3]] : NAN, ntf = mi[4] >= 0 ? z[mi[4]] : NAN, ntfd = mi[6] >= 0 ? auxder[mi[6]] : NAN;
    if (role == MOCOB200_ROLE_CONTROL) e = v;
    else if (role == MOCOB200_ROLE_ACTIVATION) a = v;
    else if (role == MOCOB200_ROLE_TENDON_FORCE) ntf = v;
    else ntfd = v;
    const bool noActDyn = flags & MOCOB200_MUS_FLAG_IGNORE_ACTIVATION_DYNAMICS;
    const MuscleOut o = dgf_muscle(md, flags, ws[A.o_mlen + m], ws[A.o_mspd + m], noActDyn ? e : a, ntf, ntfd, false);
    LocalMuscleEval r;
    r.F = o.tendonForce; r.nfDeriv = o.nfDeriv; r.residual = o.residual;
    r.actDot = mi[3] >= 0 ? activation_dot(e, a, md[5], md[6]) : 0.0;
    return r;
}

// right-hand sides of the local columns: M^-1 R_m (explicit) or R_m (implicit) per muscle, then per coordinate actuator.
// base < 0: all owners, owner r at Rm + r nq; base >= 0: the 32 owners from `base` on, owner r at Rm + (r - base) nq.
__device__ inline void wk_local_rhs(const WarpCtx& C, double* Rm, int base = -1) {
    const DevTables& T = C.T;
    const WarpAux& A = C.A;
    const int nq = T.nq;
    const int r0 = base < 0 ? 0 : base, r1 = base < 0 ? T.nmus + T.ncact : min(T.nmus + T.ncact, base + 32);
    for (int r = r0 + C.lane; r < r1; r += 32) {
        double* x = Rm + (r - r0) * nq;
        if (r < T.nmus) wk_unit_tension_force(C, r, x);
        else {
            const int a = r - T.nmus;
            const int* ai = T.idx + T.o_cact_i + a * MOCOB200_CACT_ISTRIDE;
            for (int k = 0; k < nq; ++k) x[k] = 0.0;
            x[ai[0]] = T.dbl[T.o_cact_d + a * MOCOB200_CACT_DSTRIDE];
        }
        if (!T.implicit) wk_ldlt_solve_lane(C.ws + A.o_M, nq, x);
    }
    __syncwarp();
}

// Kinematic-constraint rows of a local column (explicit mode, derivatives enforced): the acceleration-level error
// G udot + Gdot u is linear in udot, so its column is G times the udot rows just written; the position- and velocity-level
// errors do not depend on a local input.
__device__ inline void wk_local_constraint_rows(const DevTables& T, const double* q, double* col, int n_in) {
    if (T.ncoupler == 0 || !T.enforce_derivs || T.implicit) return;
    const int kc0 = T.nq + T.nz + T.n_auxres;
    for (int k = 0; k < T.ncoupler; ++k) {
        double e, d1, d2;
        coupler_function(T, k, q, e, d1, d2);
        const int* ci = T.idx + T.o_cc_i + k * MOCOB200_CC_ISTRIDE;
        col[(size_t)(kc0 + 2 * T.nm + k) * n_in] = d1 * col[(size_t)ci[1] * n_in] - col[(size_t)ci[2] * n_in];
    }
}

// CasADi's central-difference step control for one output: accumulates u = max |et / er| (finite_differences of the
// pinned fork, restated in kernels.cu fd_group); returns false when a value is not finite
__device__ __forceinline__ bool fd_central_u(double vf, double vc, double vb, double h, double reltol, double abstol, double& u) {
    if (!isfinite(vf) || !isfinite(vc) || !isfinite(vb)) return false;
    const double et = vf - 2 * vc + vb;
    const double er = reltol / h * fmax(fabs(vf - vc), fabs(vc - vb)) + abstol;
    u = fmax(u, fabs(et / er));
    return true;
}

// Central-difference version (CasADi's scheme with step refinement, one refinement loop per callback group as in the
// reference: the dynamics group, then every integrand on its own). One lane per local input.
__device__ inline void wk_local_columns_central(const DevTables& T, const WarpAux& A, double* ws, double* Rm, double h0, int h_iter,
                                                double u_aim, double reltol, double abstol, double* __restrict__ Jp,
                                                const double* __restrict__ y0, int n_in, int n_out) {
    const WarpCtx C{T, A, ws, (int)(threadIdx.x & 31)};
    const int lane = C.lane, nq = T.nq, nz = T.nz;
    const double* in = ws + A.o_in;
    wk_local_rhs(C, Rm);
    for (int t = lane; t < A.n_local; t += 32) {
        const int dir = A.loc_dir[t], kind = A.loc_kind[t], owner = A.loc_owner[t], role = A.loc_role[t];
        const double x0 = in[dir];
        double* col = Jp + dir;
        for (int o = 0; o < n_out; ++o) col[(size_t)o * n_in] = 0.0;
        const double* c = Rm + (kind == MOCOB200_LOCAL_MUSCLE ? owner : T.nmus + owner) * nq;
        // ---- dynamics group
        double h = h0;
        for (int iter = 0;; ++iter) {
            double u = 0;
            bool ok = true;
            if (kind == MOCOB200_LOCAL_ACTUATOR) {
                const double vf = x0 + h, vb = x0 - h; // force = optimal_force * control; c carries optimal_force
                for (int k = 0; k < nq; ++k) {
                    ok = fd_central_u(c[k] * vf, c[k] * x0, c[k] * vb, h, reltol, abstol, u) && ok;
                    col[(size_t)k * n_in] = (c[k] * vf - c[k] * vb) / (2 * h);
                }
            } else {
                const int* mi = T.idx + T.o_mus_i + owner * MOCOB200_MUS_ISTRIDE;
                const double* mo0 = ws + A.o_mout + 4 * owner;
                const LocalMuscleEval f = wk_local_muscle(T, A, ws, owner, role, x0 + h), b = wk_local_muscle(T, A, ws, owner, role, x0 - h);
                for (int k = 0; k < nq; ++k) {
                    // udot_k(x) = udot_k(x0) + c_k (F(x) - F(x0)): difference the tension, then map
                    ok = fd_central_u(c[k] * (f.F - mo0[0]), 0.0, c[k] * (b.F - mo0[0]), h, reltol, abstol, u) && ok;
                    col[(size_t)k * n_in] = c[k] * ((f.F - b.F) / (2 * h));
                }
                if (mi[3] >= 0) {
                    ok = fd_central_u(f.actDot, mo0[3], b.actDot, h, reltol, abstol, u) && ok;
                    col[(size_t)(nq + mi[3]) * n_in] = (f.actDot - b.actDot) / (2 * h);
                }
                if (mi[4] >= 0) {
                    ok = fd_central_u(f.nfDeriv, mo0[1], b.nfDeriv, h, reltol, abstol, u) && ok;
                    col[(size_t)(nq + mi[4]) * n_in] = (f.nfDeriv - b.nfDeriv) / (2 * h);
                }
                if (mi[7] >= 0) {
                    ok = fd_central_u(f.residual, mo0[2], b.residual, h, reltol, abstol, u) && ok;
                    col[(size_t)(nq + nz + mi[7]) * n_in] = (f.residual - b.residual) / (2 * h);
                }
            }
            if (iter == h_iter) break;
            if (!ok) h /= u_aim;
            else h *= sqrt(u_aim / fmax(1.0, u));
        }
        // ---- one group per integrand
        int slot = 0;
        for (int g = 0; g < T.ngoal; ++g) {
            const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
            if (!gr[3]) continue;
            double hg = h0, jv = 0.0;
            for (int iter = 0;; ++iter) {
                const double df = wk_integrand_term_delta(T, A, ws, gr, dir, x0, x0 + hg), db = wk_integrand_term_delta(T, A, ws, gr, dir, x0, x0 - hg);
                double u = 0;
                const bool ok = fd_central_u(df, 0.0, db, hg, reltol, abstol, u);
                jv = ok ? (df - db) / (2 * hg) : NAN;
                if (iter == h_iter) break;
                if (!ok) hg /= u_aim;
                else hg *= sqrt(u_aim / fmax(1.0, u));
            }
            col[(size_t)(nq + nz + T.n_auxres + T.n_kc + slot) * n_in] = jv;
            ++slot;
        }
        wk_local_constraint_rows(T, in + 1, col, n_in);
        for (int o = 0; o < n_out; ++o)
            if (!isfinite(y0[o])) col[(size_t)o * n_in] = NAN;
    }
    __syncwarp();
}

// The columns J[:, dir] of the "local" inputs of one grid point that belong to this warp's owners (whole warp; the scratch
// holds the reloaded blob of the nominal evaluation, y0 its outputs, Rm is scratch the evaluation no longer needs). The
// owners (muscles, then coordinate actuators) are dealt to the `nparts` warps of the grid point in chunks of 32: a warp
// forms the right-hand sides of its chunk, one owner per lane, and then the columns whose owner is in the chunk, one column
// per lane. On a full-body model (107 owners, 176 local inputs) that is one right-hand-side pass and two column passes per
// warp instead of four and six in a row.
__device__ inline void wk_local_columns(const DevTables& T, const WarpAux& A, double* ws, double* Rm, double h, bool forward,
                                        double* __restrict__ Jp, const double* __restrict__ y0, int n_in, int n_out, int part, int nparts) {
    const WarpCtx C{T, A, ws, (int)(threadIdx.x & 31)};
    const int lane = C.lane, nq = T.nq, nz = T.nz;
    const double* in = ws + A.o_in;
    const double step = forward ? h : -h;
    const int nown = T.nmus + T.ncact;
    for (int base = part * 32; base < nown; base += 32 * nparts) {
    wk_local_rhs(C, Rm, base);
    for (int t = lane; t < A.n_local; t += 32) {
        const int dir = A.loc_dir[t], kind = A.loc_kind[t], owner = A.loc_owner[t], role = A.loc_role[t];
        const int row = (kind == MOCOB200_LOCAL_MUSCLE ? owner : T.nmus + owner) - base;
        if (row < 0 || row >= 32) continue;
        const double v0 = in[dir], v1 = v0 + step;
        const double inv = forward ? 1.0 / h : -1.0 / h; // (f(x+h) - f(x)) / h or (f(x) - f(x-h)) / h
        double* col = Jp + dir;
        for (int o = 0; o < n_out; ++o) col[(size_t)o * n_in] = 0.0;
        const double* c = Rm + row * nq;
        if (kind == MOCOB200_LOCAL_ACTUATOR) {
            // force = optimal_force * control (CoordinateActuator.cpp:185-204): c already carries optimal_force
            const double dctl = (v1 - v0) * inv;
            for (int k = 0; k < nq; ++k) col[(size_t)k * n_in] = c[k] * dctl;
        } else {
            const int m = owner;
            const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
            const double* mo0 = ws + A.o_mout + 4 * m; // nominal tension, d(tendon force)/dt, residual, activation rate
            const LocalMuscleEval o1 = wk_local_muscle(T, A, ws, m, role, v1);
            const double dF = (o1.F - mo0[0]) * inv;
            for (int k = 0; k < nq; ++k) col[(size_t)k * n_in] = c[k] * dF;
            if (mi[3] >= 0) col[(size_t)(nq + mi[3]) * n_in] = (o1.actDot - mo0[3]) * inv;
            if (mi[4] >= 0) col[(size_t)(nq + mi[4]) * n_in] = (o1.nfDeriv - mo0[1]) * inv;
            if (mi[7] >= 0) col[(size_t)(nq + nz + mi[7]) * n_in] = (o1.residual - mo0[2]) * inv;
        }
        int slot = 0;
        for (int g = 0; g < T.ngoal; ++g) {
            const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
            if (!gr[3]) continue;
            col[(size_t)(nq + nz + T.n_auxres + T.n_kc + slot) * n_in] = wk_integrand_term_delta(T, A, ws, gr, dir, v0, v1) * inv;
            ++slot;
        }
        wk_local_constraint_rows(T, in + 1, col, n_in);
        // a non-finite nominal output poisons its whole row, as it does when the full function is differenced
        for (int o = 0; o < n_out; ++o)
            if (!isfinite(y0[o])) col[(size_t)o * n_in] = NAN;
    }
    __syncwarp();
    }
}

} // namespace mocob200
