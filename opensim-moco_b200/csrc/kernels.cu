// Kernels: point functions over the grid (+ CasADi-style finite differences), endpoint functions,
// and assembly of g, jac_g (CSC), f and grad_f.  FP64, sm_100a.
#include <cstdio>

#include "kernels.cuh"
#include "fd_group.cuh"

#include <cstdlib>

namespace mocob200 {

template <class SZ>
__global__ void k_points_nominal(DevTables T, DevLayout L, const double* __restrict__ x, double* __restrict__ Y, int p0, int np) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= np) return;
    const int j = p0 + t;
    double in[SZ::NIN], out[SZ::NOUT];
    gather_inputs(L, x, j, in);
    eval_point<SZ>(T, in, out, 0xffffffffu);
    for (int o = 0; o < L.n_out; ++o) Y[(size_t)j * L.n_out + o] = out[o];
}

// Muscle-tendon path length and lengthening speed of every path at every grid point: out[j][m] = (length, speed).
// What GeometryPath::getLength / getLengtheningSpeed report (GeometryPath.cpp:848-864, 1117-1145) - the quantities the
// reference's MuscleAnalysis / Actuation analyses print; one thread per (grid point, muscle).
template <class SZ>
__global__ void k_paths(DevTables T, DevLayout L, const double* __restrict__ x, double* __restrict__ out, int p0, int np) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= np * T.nmus) return;
    const int j = p0 + t / T.nmus, m = t % T.nmus;
    double in[SZ::NIN];
    gather_inputs(L, x, j, in);
    double pk[3 * SZ::NQ];
    const PointInputs P = point_inputs(T, in, pk);
    ScratchT<SZ> S;
    kinematics(T, P.q, P.u, S);
    double ls[2];
    muscle_force(T, m, P.q, P.u, P.z, P.ctl, P.auxder, S, false, false, ls);
    out[2 * ((size_t)j * T.nmus + m)] = ls[0];
    out[2 * ((size_t)j * T.nmus + m) + 1] = ls[1];
}
void launch_paths(const DevTables& T, const DevLayout& L, const double* d_x, double* d_out, int p0, int np, cudaStream_t s) {
    const int n = np * T.nmus;
    if (n <= 0) return;
    if (T.npathwrap > 0 || T.nb > SizeLarge::NB || T.nq > SizeLarge::NQ || L.n_in > SizeLarge::NIN) k_paths<SizeHuge><<<(n + 63) / 64, 64, 0, s>>>(T, L, d_x, d_out, p0, np);
    else k_paths<SizeLarge><<<(n + 63) / 64, 64, 0, s>>>(T, L, d_x, d_out, p0, np);
}

// block = one grid point; thread t < ndirs = FD direction dirs[t] (all n_in inputs when dirs == nullptr); thread ndirs =
// nominal evaluation. Columns outside the list are produced elsewhere (point_warp_local.cuh).
__global__ void __launch_bounds__(256)
k_points_fd(DevTables T, DevLayout L, FdSettings fd, const double* __restrict__ x, double* __restrict__ Y, double* __restrict__ J, int p0,
            const int* __restrict__ dirs, int ndirs) {
    extern __shared__ double sh_y0[]; // [n_out]
    const int j = p0 + blockIdx.x;
    const int tix = threadIdx.x;
    const int a = tix < ndirs ? (dirs ? dirs[tix] : tix) : (tix == ndirs ? L.n_in : L.n_in + 1);
    double in[MOCOB200_MAX_IN];
    double ya[MOCOB200_MAX_OUT], yb[MOCOB200_MAX_OUT], Jc[MOCOB200_MAX_OUT];
    const bool active = a <= L.n_in;
    if (active) gather_inputs(L, x, j, in);
    if (a == L.n_in) {
        eval_point(T, in, ya, 0xffffffffu);
        for (int o = 0; o < L.n_out; ++o) { sh_y0[o] = ya[o]; Y[(size_t)j * L.n_out + o] = ya[o]; }
    }
    const bool central = fd.scheme == MOCOB200_FD_CENTRAL;
    if (a < L.n_in && !central) {
        // forward/backward need no nominal value until the difference: evaluate before the barrier
        const double x0 = in[a];
        const double pert = (fd.scheme == MOCOB200_FD_FORWARD) ? fd.h : -fd.h;
        in[a] = x0 + pert;
        eval_point(T, in, ya, 0xffffffffu);
        in[a] = x0;
    }
    __syncthreads();
    if (a >= L.n_in) return;
    if (!central) {
        const double h = fd.h;
        for (int o = 0; o < L.n_out; ++o) Jc[o] = (fd.scheme == MOCOB200_FD_FORWARD) ? (ya[o] - sh_y0[o]) / h : (sh_y0[o] - ya[o]) / h;
    } else {
        // every reference callback refines its own step: group 0 = dynamics, group 1+k = integrand k
        const int ngroups = 1 + L.n_integrand;
        for (int gidx = 0; gidx < ngroups; ++gidx) {
            const unsigned mask = 1u << gidx;
            const int lo = gidx == 0 ? 0 : L.n_dyn_out + gidx - 1;
            const int hi = gidx == 0 ? L.n_dyn_out : lo + 1;
            fd_group([&](const double* xin, double* yout) { eval_point(T, xin, yout, mask); }, fd, in, a, sh_y0, lo, hi, yb, ya, Jc);
        }
    }
    double* Jp = J + (size_t)j * L.n_out * L.n_in;
    for (int o = 0; o < L.n_out; ++o) Jp[(size_t)o * L.n_in + a] = Jc[o];
}

// ---------------------------------------------------------------------------------------------
// d(point output row)/d(time) at grid point j: column 0 of the point's FD block, or — when the blocks of other ranks' grid points
// are not resident (sharded evaluation) — the gathered compact copy Jt[n_grid][n_out]
__device__ __forceinline__ double time_col(const DevLayout& L, const double* J, int j, int row) {
    return L.Jt ? L.Jt[(size_t)j * L.n_out + row] : J[((size_t)j * L.n_out + row) * L.n_in];
}
// integrals of the goal integrands: I_k = T * sum_j quad_j * integrand_k(j)  (CasOCTranscription.cpp:489-493)
// d_I = [I (n_integrand) | dI/dt0 | dI/dtf]
__global__ void k_integrals(DevLayout L, const double* __restrict__ x, const double* __restrict__ Y, const double* __restrict__ J, double* I, int withJ) {
    __shared__ double red[3][256];
    const int k = blockIdx.x;
    const double T = x[1] - x[0];
    double s0 = 0, s1 = 0, s2 = 0;
    const int row = L.n_dyn_out + k;
    for (int j = threadIdx.x; j < L.n_grid; j += blockDim.x) {
        const double qj = L.quad[j];
        s0 += qj * Y[(size_t)j * L.n_out + row];
        if (withJ) {
            const double dydt = time_col(L, J, j, row);
            s1 += qj * dydt * (1.0 - L.grid[j]);
            s2 += qj * dydt * L.grid[j];
        }
    }
    red[0][threadIdx.x] = s0; red[1][threadIdx.x] = s1; red[2][threadIdx.x] = s2;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a0 = 0, a1 = 0, a2 = 0;
        for (int t = 0; t < blockDim.x; ++t) { a0 += red[0][t]; a1 += red[1][t]; a2 += red[2][t]; }
        I[k] = T * a0;
        I[L.n_integrand + k] = -a0 + T * a1;
        I[2 * L.n_integrand + k] = a0 + T * a2;
    }
}

// ---------------------------------------------------------------------------------------------
// endpoint functions of all goals + their FD blocks. One block; thread a < n_ein = direction, thread n_ein = nominal.
// does the endpoint function of goal g read endpoint input a?  (columns it does not read are exactly zero under any
// finite-difference formula, so they are written without evaluating anything)
__device__ inline bool endpoint_reads(const DevTables& T, const DevLayout& L, int g, int a) {
    const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
    const int* ip = T.idx + gr[4];
    const int half = 1 + L.ns + L.nc + L.nm + L.nd;
    const bool isIntegral = a == 2 * half;
    const int side = a >= half ? 1 : 0;          // 0 = initial, 1 = final
    const int k = isIntegral ? -1 : a - side * half; // 0 = time, 1.. = states, controls, derivatives
    const bool isTime = k == 0;
    const bool isQ = k >= 1 && k <= L.nqs;
    const bool isState = k >= 1 && k <= L.ns;
    const bool isControl = k > L.ns && k <= L.ns + L.nc;
    switch (gr[0]) {
    case MOCOB200_GOAL_CONTROL: return isIntegral || (ip[1] && isQ);
    case MOCOB200_GOAL_STATE_TRACKING:
    case MOCOB200_GOAL_CONTROL_TRACKING:
    case MOCOB200_GOAL_SUM_SQUARED_STATE:
    case MOCOB200_GOAL_CONTACT_TRACKING: return isIntegral;
    case MOCOB200_GOAL_FINAL_TIME: return isTime && side == 1;
    case MOCOB200_GOAL_AVERAGE_SPEED: return isTime || isQ;
    case MOCOB200_GOAL_PERIODICITY: return isState || isControl;
    case MOCOB200_GOAL_INITIAL_ACTIVATION: return side == 0 && (isState || isControl);
    default: return !isIntegral || L.goal_integrand[g] >= 0;
    }
}

// one block per goal; thread a < n_ein = FD direction, thread n_ein = nominal evaluation.
// which: 0 = every goal, 1 = only goals without an integrand (functions of x alone), 2 = only goals with one.
__global__ void __launch_bounds__(512) k_endpoint(DevTables T, DevLayout L, FdSettings fd, const double* __restrict__ x, const double* __restrict__ I,
                           double* __restrict__ G, double* __restrict__ E, int withFd, int which) {
    extern __shared__ double sh[]; // [n_ein] nominal inputs, then [nout] nominal outputs
    const int g = blockIdx.x;
    const int a = threadIdx.x;
    const int nout = T.idx[T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE + 2];
    if (T.idx[T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE + 1] == MOCOB200_MODE_PATH_CONSTRAINT) return; // not an endpoint function (whole block)
    if (which && (which == 1) == (L.goal_integrand[g] >= 0)) return;                                  // not this launch's share
    double* sh_in = sh;
    double* sh_g0 = sh + L.n_ein;
    {   // cooperative gather of [t0, x0, c0, d0, tf, xf, cf, df, integral]
        const int half = 1 + L.ns + L.nc + L.nm + L.nd, last = L.n_grid - 1;
        for (int i = threadIdx.x; i < L.n_ein; i += blockDim.x) {
            double v;
            if (i == 2 * half) v = (L.goal_integrand[g] >= 0) ? I[L.goal_integrand[g]] : NAN; // NaN without an integrand (:495)
            else {
                const int side = i >= half ? 1 : 0, k = i - side * half, j = side ? last : 0;
                if (k == 0) v = x[side];
                else if (k <= L.ns) v = x[L.off_states + (size_t)j * L.ns + (k - 1)];
                else if (k <= L.ns + L.nc) v = x[L.off_controls + (size_t)j * L.nc + (k - 1 - L.ns)];
                else if (k <= L.ns + L.nc + L.nm) v = x[L.off_mult + (size_t)j * L.nm + (k - 1 - L.ns - L.nc)];
                else v = x[L.off_derivs + (size_t)j * L.nd + (k - 1 - L.ns - L.nc - L.nm)];
            }
            sh_in[i] = v;
        }
    }
    __syncthreads();
    double ein[2 * MOCOB200_MAX_IN + 8];
    double v0[MOCOB200_MAX_EOUT], v1[MOCOB200_MAX_EOUT], Jc[MOCOB200_MAX_EOUT];
    if (a == L.n_ein) {
        for (int i = 0; i < L.n_ein; ++i) ein[i] = sh_in[i];
        eval_endpoint(T, g, ein, L.ns, L.nc, L.nd, v0);
        for (int o = 0; o < nout; ++o) { sh_g0[o] = v0[o]; G[L.goal_eoff[g] + o] = v0[o]; }
    }
    __syncthreads();
    if (!withFd || a >= L.n_ein) return;
    const bool nanIntegral = (a == L.n_ein - 1) && L.goal_integrand[g] < 0; // structurally unused entry
    if (nanIntegral || !endpoint_reads(T, L, g, a)) {
        for (int o = 0; o < nout; ++o) {
            double z = 0.0;
            if (!nanIntegral && !isfinite(sh_g0[o])) z = NAN; // a non-finite nominal poisons every difference
            E[(size_t)(L.goal_eoff[g] + o) * L.n_ein + a] = z;
        }
        return;
    }
    for (int i = 0; i < L.n_ein; ++i) ein[i] = sh_in[i];
    fd_group([&](const double* xin, double* yout) { eval_endpoint(T, g, xin, L.ns, L.nc, L.nd, yout); }, fd, ein, a, sh_g0, 0, nout,
             v0, v1, Jc);
    for (int o = 0; o < nout; ++o) E[(size_t)(L.goal_eoff[g] + o) * L.n_ein + a] = Jc[o];
}

// ---------------------------------------------------------------------------------------------
// Hermite-Simpson midpoints with enforced constraint derivatives: qdot = u + G(q)^T gamma (CasOCTranscription.cpp:315-332 ->
// VelocityCorrection::eval CasOCFunction.cpp:286-291 -> MocoCasOCProblem.h:298-332)
__device__ __forceinline__ bool vel_corrected(const DevLayout& L, int j) { return L.nsl > 0 && L.scheme == MOCOB200_SCHEME_HERMITE_SIMPSON && (j & 1); }
__device__ inline double velcorr_value(const DevTables& T, const DevLayout& L, const double* x, int j, int s) {
    return coupler_velocity_correction(T, x + L.off_states + (size_t)j * L.ns, x + L.off_slacks + (size_t)(j / 2) * L.nsl, s);
}
// One entry of the FD block of the VelocityCorrection callback at midpoint j: d corr_s / d input a, a = 0 time, 1..2nq the
// multibody states, then the slacks; CasADi FiniteDiff on that callback (all nq outputs share the refined central step)
__device__ double velcorr_fd_entry(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* __restrict__ x, int j, int s, int a) {
    const int nmb = 2 * L.nq;
    if (a == 0) return 0.0; // the correction does not read the time: every finite difference of it is exactly zero
    double in[2 * MOCOB200_MAXQ + 16];
    const double* xs = x + L.off_states + (size_t)j * L.ns;
    const double* sl = x + L.off_slacks + (size_t)(j / 2) * L.nsl;
    for (int k = 0; k < L.nq; ++k) in[k] = xs[k];
    double* g = in + L.nq;
    for (int k = 0; k < L.nsl; ++k) g[k] = sl[k];
    double* slot = a <= L.nq ? in + (a - 1) : (a <= nmb ? nullptr : g + (a - 1 - nmb));
    if (!slot) return 0.0; // the speeds are inputs the correction does not read
    const double x0 = *slot;
    auto at = [&](int ss, double xa) { *slot = xa; const double v = coupler_velocity_correction(T, in, g, ss); *slot = x0; return v; };
    if (fd.scheme != MOCOB200_FD_CENTRAL) {
        const bool fwd = fd.scheme == MOCOB200_FD_FORWARD;
        const double vc = at(s, x0), yf = at(s, fwd ? x0 + fd.h : x0 - fd.h);
        return fwd ? (yf - vc) / fd.h : (vc - yf) / fd.h;
    }
    double h = fd.h, jv = 0;
    for (int iter = 0;; ++iter) {
        double u = 0;
        for (int ss = 0; ss < L.nq; ++ss) {
            const double vb = at(ss, x0 - h), vf = at(ss, x0 + h), vc = at(ss, x0);
            double je;
            if (!isfinite(vf) || !isfinite(vc) || !isfinite(vb)) { je = NAN; u = -1; }
            else {
                je = (vf - vb) / (2 * h);
                const double et = vf - 2 * vc + vb;
                const double er = fd.reltol / h * fmax(fabs(vf - vc), fabs(vc - vb)) + fd.abstol;
                if (u >= 0) u = fmax(u, fabs(et / er));
            }
            if (ss == s) jv = je;
        }
        if (iter == fd.h_iter) break;
        if (u < 0) h /= fd.u_aim;
        else h *= sqrt(fd.u_aim / fmax(1.0, u));
    }
    return jv;
}
__device__ __forceinline__ double xdot_val(const DevLayout& L, const double* x, const double* Y, int j, int s) {
    // L.nqs = coordinate states in the NLP (0 with prescribed kinematics: every state is then an auxiliary one)
    if (s < L.nqs) return x[L.off_states + (size_t)j * L.ns + L.nqs + s];                     // qdot = u (:313-314)
    if (s < 2 * L.nqs && L.implicit) return x[L.off_derivs + (size_t)j * L.nd + (s - L.nqs)]; // udot = w
    return Y[(size_t)j * L.n_out + L.nq + (s - 2 * L.nqs)];
}
__device__ __forceinline__ double dxdot(const DevLayout& L, const double* J, int j, int s, const ColMetaD& c) {
    if (c.kind == COL_SLACK_D) return 0.0; // enters through the velocity correction only (k_assemble_jac_velcorr)
    if (s < L.nqs) return (c.kind == COL_STATE_D && c.k == L.nqs + s) ? 1.0 : 0.0;
    if (s < 2 * L.nqs && L.implicit) return (c.kind == COL_DERIV_D && c.k == s - L.nqs) ? 1.0 : 0.0;
    return J[((size_t)j * L.n_out + L.nq + (s - 2 * L.nqs)) * L.n_in + c.in];
}
__device__ __forceinline__ double dxdot_dt(const DevLayout& L, const double* J, int j, int s) {
    if (s < L.nqs) return 0.0;
    if (s < 2 * L.nqs && L.implicit) return 0.0;
    return time_col(L, J, j, L.nq + (s - 2 * L.nqs));
}
__global__ void k_pack_time_column(DevLayout L, const double* __restrict__ J, double* __restrict__ Jt, int p0, int np) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= np * L.n_out) return;
    const size_t e = (size_t)p0 * L.n_out + t;
    Jt[e] = J[e * L.n_in];
}
void launch_pack_time_column(const DevLayout& L, const double* d_J, double* d_Jt, int p0, int np, cudaStream_t s) {
    const int n = np * L.n_out;
    if (n > 0) k_pack_time_column<<<(n + 255) / 256, 256, 0, s>>>(L, d_J, d_Jt, p0, np);
}
// defect coefficients: defect = sum_role lin_role * x_role,s + sum_role (wc_role * h) * xdot_role,s
__device__ __forceinline__ void defect_coeffs(int kind, int role, double& lin, double& wc) {
    if (kind == ROW_DEFECT_HERMITE_D) { // x_mid - (x_ip + x_i)/2 - h/8 (xdot_i - xdot_ip)
        lin = role == 1 ? 1.0 : -0.5;
        wc = role == 0 ? -1.0 / 8.0 : (role == 2 ? 1.0 / 8.0 : 0.0);
    } else if (kind == ROW_DEFECT_SIMPSON_D) { // x_ip - x_i - h/6 (xdot_ip + 4 xdot_mid + xdot_i)
        lin = role == 0 ? -1.0 : (role == 2 ? 1.0 : 0.0);
        wc = role == 1 ? -4.0 / 6.0 : -1.0 / 6.0;
    } else { // trapezoidal: x_ip - (x_i + h/2 (xdot_ip + xdot_i))
        lin = role == 0 ? -1.0 : 1.0;
        wc = -0.5;
    }
}

// ---------------------------------------------------------------------------------------------
// Path constraints (CasOC::PathConstraint::eval CasOCFunction.cpp:137-143). The one supported kind,
// MocoControlBoundConstraint (MocoControlBoundConstraint.cpp:121-148), is error_e = control_e - bound_e(time): a function
// of the mesh point's time and controls only, so it is evaluated here in the assembly kernels, not in the point kernels.
__device__ inline double pathcon_error(const DevTables& T, const int* ip, int e, double time, const double* ctl, int ck, double cv) {
    const int c = ip[2 * e];
    return (c == ck ? cv : ctl[c]) - ppoly_eval(T, ip[2 * e + 1], time);
}
// One entry of the FD block of path constraint `g` at grid point j: d error_e / d input_a, CasADi FiniteDiff on that callback
// (fd_group above: every equation of the constraint shares the refined central step). Inputs the constraint does not read
// still own a (zero-valued) structural entry of the dense block.
__device__ double pathcon_fd_entry(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* __restrict__ x, int j, int g,
                                   int e, int a) {
    const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
    const int* ip = T.idx + gr[4];
    const int neq = gr[2];
    const double t = x[0] + (x[1] - x[0]) * L.grid[j];
    const double* ctl = x + L.off_controls + (size_t)j * L.nc;
    const int ck = (a > L.ns && a <= L.ns + L.nc) ? a - 1 - L.ns : -1; // perturbed control (or none)
    const double x0 = a == 0 ? t : (ck >= 0 ? ctl[ck] : 0.0);
    auto at = [&](int ee, double xa) { return pathcon_error(T, ip, ee, a == 0 ? xa : t, ctl, ck, xa); };
    const double vc_e = pathcon_error(T, ip, e, t, ctl, -1, 0.0);
    if (fd.scheme != MOCOB200_FD_CENTRAL) {
        const bool fwd = fd.scheme == MOCOB200_FD_FORWARD;
        const double yf = at(e, fwd ? x0 + fd.h : x0 - fd.h);
        return fwd ? (yf - vc_e) / fd.h : (vc_e - yf) / fd.h;
    }
    double h = fd.h, jv = 0;
    for (int iter = 0;; ++iter) {
        double u = 0;
        for (int ee = 0; ee < neq; ++ee) {
            const double vb = at(ee, x0 - h), vf = at(ee, x0 + h), vc = pathcon_error(T, ip, ee, t, ctl, -1, 0.0);
            double je;
            if (!isfinite(vf) || !isfinite(vc) || !isfinite(vb)) { je = NAN; u = -1; }
            else {
                je = (vf - vb) / (2 * h);
                const double et = vf - 2 * vc + vb;
                const double er = fd.reltol / h * fmax(fabs(vf - vc), fabs(vc - vb)) + fd.abstol;
                if (u >= 0) u = fmax(u, fabs(et / er));
            }
            if (ee == e) jv = je;
        }
        if (iter == fd.h_iter) break;
        if (u < 0) h /= fd.u_aim;
        else h *= sqrt(fd.u_aim / fmax(1.0, u));
    }
    return jv;
}

// path-constraint rows of g and of jac_g: separate small kernels, launched only for problems that have path constraints
__global__ void k_assemble_g_path(DevTables T, DevLayout L, const double* __restrict__ x, double* __restrict__ g) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= L.n_con) return;
    const RowMetaD rm = L.rows[r];
    if (rm.kind != ROW_PATH_D) return;
    const int* gr = T.idx + T.o_goal_i + rm.pad * MOCOB200_GOAL_ISTRIDE;
    g[r] = pathcon_error(T, T.idx + gr[4], rm.b, x[0] + (x[1] - x[0]) * L.grid[rm.a], x + L.off_controls + (size_t)rm.a * L.nc, -1, 0.0);
}
__global__ void k_assemble_jac_path(DevTables T, DevLayout L, FdSettings fd, const double* __restrict__ x, double* __restrict__ vals, int k0, int k1) {
    const int k = k0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= k1) return;
    const RowMetaD rm = L.rows[L.rowidx[k]];
    if (rm.kind != ROW_PATH_D) return;
    const ColMetaD c = L.cols[L.nnzcol[k]];
    const int j = rm.a;
    if (c.kind == COL_T0_D || c.kind == COL_TF_D)
        vals[k] = pathcon_fd_entry(T, L, fd, x, j, rm.pad, rm.b, 0) * (c.kind == COL_TF_D ? L.grid[j] : 1.0 - L.grid[j]);
    else vals[k] = pathcon_fd_entry(T, L, fd, x, j, rm.pad, rm.b, c.in);
}

__global__ void k_assemble_g(DevLayout L, const double* __restrict__ x, const double* __restrict__ Y, const double* __restrict__ G,
                             double* __restrict__ g) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= L.n_con) return;
    const RowMetaD rm = L.rows[r];
    const bool hs = L.scheme == MOCOB200_SCHEME_HERMITE_SIMPSON;
    double v = 0;
    if (rm.kind == ROW_ENDPOINT_D) v = G[L.goal_eoff[rm.a] + rm.b];
    else if (rm.kind == ROW_RESIDUAL_D || rm.kind == ROW_KINEMATIC_D) v = Y[(size_t)rm.a * L.n_out + rm.b];
    else if (rm.kind == ROW_PATH_D) return; // k_assemble_g_path
    else if (rm.kind == ROW_INTERP_D) {
        const int m = rm.a, c = rm.b;
        const double* xc = x + L.off_controls;
        v = xc[(size_t)(2 * m + 1) * L.nc + c] - 0.5 * (xc[(size_t)(2 * m + 2) * L.nc + c] + xc[(size_t)(2 * m) * L.nc + c]);
    } else {
        const int m = rm.a, s = rm.b;
        const double h = (x[1] - x[0]) * (L.mesh[m + 1] - L.mesh[m]);
        const double* xs = x + L.off_states;
        if (!hs) {
            const double xi = xs[(size_t)m * L.ns + s], xp = xs[(size_t)(m + 1) * L.ns + s];
            v = xp - (xi + 0.5 * h * (xdot_val(L, x, Y, m + 1, s) + xdot_val(L, x, Y, m, s)));
        } else {
            const int i = 2 * m, mid = 2 * m + 1, ip = 2 * m + 2;
            const double xi = xs[(size_t)i * L.ns + s], xm = xs[(size_t)mid * L.ns + s], xp = xs[(size_t)ip * L.ns + s];
            const double di = xdot_val(L, x, Y, i, s), dm = xdot_val(L, x, Y, mid, s), dp = xdot_val(L, x, Y, ip, s);
            if (rm.kind == ROW_DEFECT_HERMITE_D) v = xm - 0.5 * (xp + xi) - (h / 8.0) * (di - dp);
            else v = xp - xi - (h / 6.0) * (dp + 4.0 * dm + di);
        }
    }
    g[r] = v;
}

// objective: sum of cost-mode goals (+ implicit penalties, CasOCTranscription.cpp:513-545). f[0] = total, f[1..] = terms.
__global__ void k_assemble_f(DevTables T, DevLayout L, const double* __restrict__ x, const double* __restrict__ G, double* __restrict__ f) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double total = 0;
    int term = 0;
    for (int g = 0; g < T.ngoal; ++g) {
        const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
        if (gr[1] != MOCOB200_MODE_COST) continue;
        double s = 0;
        for (int o = 0; o < gr[2]; ++o) s += G[L.goal_eoff[g] + o];
        f[1 + term++] = s;
        total += s;
    }
    const double dur = x[1] - x[0];
    if (L.minimize_mult) { // lagrange_multiplier_weight * duration * sum_j quad_j |lambda_j|^2 (CasOCTranscription.cpp:513-521)
        double acc = 0;
        for (int j = 0; j < L.n_grid; ++j) {
            double sq = 0;
            for (int k = 0; k < L.nm; ++k) { const double d = x[L.off_mult + (size_t)j * L.nm + k]; sq += d * d; }
            acc += L.quad[j] * sq;
        }
        const double s = L.mult_weight * dur * acc;
        f[1 + term++] = s;
        total += s;
    }
    for (int which = 0; which < 2; ++which) {
        const int nacc = L.nd - L.n_auxres; // acceleration variables (none in explicit mode or with prescribed kinematics)
        const bool on = which == 0 ? (L.minimize_accel && nacc > 0) : (L.minimize_auxderiv && L.n_auxres > 0);
        if (!on) continue;
        const int lo = which == 0 ? 0 : nacc, hi = which == 0 ? nacc : L.nd;
        double acc = 0;
        for (int j = 0; j < L.n_grid; ++j) {
            double sq = 0;
            for (int k = lo; k < hi; ++k) { const double d = x[L.off_derivs + (size_t)j * L.nd + k]; sq += d * d; }
            acc += L.quad[j] * sq;
        }
        const double s = (which == 0 ? L.accel_weight : L.auxderiv_weight) * dur * acc;
        f[1 + term++] = s;
        total += s;
    }
    f[0] = total;
}

// d(goal output)/d(variable of column c): endpoint block + chain through the integral
__device__ __forceinline__ double endpoint_entry(const DevLayout& L, const double* x, const double* J, const double* I, const double* E,
                                                 int g, int o, const ColMetaD& c) {
    const double* Er = E + (size_t)(L.goal_eoff[g] + o) * L.n_ein;
    const int half = 1 + L.ns + L.nc + L.nm + L.nd;
    const int slot = L.goal_integrand[g];
    const double dGdI = slot >= 0 ? Er[L.n_ein - 1] : 0.0;
    if (c.kind == COL_T0_D) return Er[0] + (slot >= 0 ? dGdI * I[L.n_integrand + slot] : 0.0);
    if (c.kind == COL_TF_D) return Er[half] + (slot >= 0 ? dGdI * I[2 * L.n_integrand + slot] : 0.0);
    if (c.kind == COL_SLACK_D) return 0.0; // slacks enter through the velocity correction only
    double v = 0;
    if (c.point == 0) v += Er[c.in];
    if (c.point == L.n_grid - 1) v += Er[half + c.in];
    if (slot >= 0) {
        const double dur = x[1] - x[0];
        v += dGdI * dur * L.quad[c.point] * J[((size_t)c.point * L.n_out + L.n_dyn_out + slot) * L.n_in + c.in];
    }
    return v;
}

// entries [k0, k1) of the CSC value array (the whole array, or the run of one block of columns when the columns are shared out
// over ranks: a column's entries need the FD blocks of that column's own grid point only, except the t0 / tf columns)
__global__ void k_assemble_jac(DevLayout L, const double* __restrict__ x, const double* __restrict__ Y, const double* __restrict__ J,
                               const double* __restrict__ I, const double* __restrict__ E, double* __restrict__ vals, int k0, int k1) {
    const int k = k0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= k1) return;
    const RowMetaD rm = L.rows[L.rowidx[k]];
    const ColMetaD c = L.cols[L.nnzcol[k]];
    const bool hs = L.scheme == MOCOB200_SCHEME_HERMITE_SIMPSON;
    const bool isTime = c.kind == COL_T0_D || c.kind == COL_TF_D;
    double v = 0;
    if (rm.kind == ROW_ENDPOINT_D) {
        v = endpoint_entry(L, x, J, I, E, rm.a, rm.b, c);
    } else if (rm.kind == ROW_RESIDUAL_D || rm.kind == ROW_KINEMATIC_D) {
        const int j = rm.a;
        if (isTime) v = time_col(L, J, j, rm.b) * (c.kind == COL_TF_D ? L.grid[j] : 1.0 - L.grid[j]);
        else v = J[((size_t)j * L.n_out + rm.b) * L.n_in + c.in];
    } else if (rm.kind == ROW_PATH_D) {
        return; // k_assemble_jac_path
    } else if (rm.kind == ROW_INTERP_D) {
        v = (c.point == 2 * rm.a + 1) ? 1.0 : -0.5;
    } else {
        const int m = rm.a, s = rm.b;
        const double dm = L.mesh[m + 1] - L.mesh[m];
        const double h = (x[1] - x[0]) * dm;
        const int base = hs ? 2 * m : m;
        if (!isTime) {
            const int role = hs ? (c.point - base) : 2 * (c.point - base);
            double lin, wc;
            defect_coeffs(rm.kind, role, lin, wc);
            v = ((c.kind == COL_STATE_D && c.k == s) ? lin : 0.0);
            if (wc != 0.0) v += wc * h * dxdot(L, J, c.point, s, c);
        } else {
            const double dhdt = (c.kind == COL_TF_D) ? dm : -dm;
            const int npts = hs ? 3 : 2;
            for (int p = 0; p < npts; ++p) {
                const int role = hs ? p : 2 * p;
                const int j = base + p;
                double lin, wc;
                defect_coeffs(rm.kind, role, lin, wc);
                if (wc == 0.0) continue;
                const double dtj = (c.kind == COL_TF_D) ? L.grid[j] : 1.0 - L.grid[j];
                v += wc * (dhdt * xdot_val(L, x, Y, j, s) + h * dxdot_dt(L, J, j, s) * dtj);
            }
        }
    }
    vals[k] = v;
}

// Velocity-correction terms of the Simpson defects of the coordinates (launched only for problems with slack variables):
// defect = x_ip - x_i - h/6 (xdot_ip + 4 xdot_mid + xdot_i) with qdot_mid = u_mid + corr(q_mid, gamma) adds -4h/6 corr.
__global__ void k_assemble_g_velcorr(DevTables T, DevLayout L, const double* __restrict__ x, double* __restrict__ g) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= L.n_intervals * L.nq) return;
    const int m = t / L.nq, s = t % L.nq;
    const double h = (x[1] - x[0]) * (L.mesh[m + 1] - L.mesh[m]);
    g[L.row_defect[m] + L.ns + s] += -(4.0 / 6.0) * h * velcorr_value(T, L, x, 2 * m + 1, s);
}
__global__ void k_assemble_jac_velcorr(DevTables T, DevLayout L, FdSettings fd, const double* __restrict__ x, double* __restrict__ vals, int k0, int k1) {
    const int k = k0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= k1) return;
    const RowMetaD rm = L.rows[L.rowidx[k]];
    if (rm.kind != ROW_DEFECT_SIMPSON_D || rm.b >= L.nq) return;
    const ColMetaD c = L.cols[L.nnzcol[k]];
    const int m = rm.a, s = rm.b, j = 2 * m + 1;
    const double dm = L.mesh[m + 1] - L.mesh[m];
    const double h = (x[1] - x[0]) * dm, wc = -4.0 / 6.0;
    if (c.kind == COL_T0_D || c.kind == COL_TF_D) { // through h (the correction itself does not read the time)
        vals[k] += wc * ((c.kind == COL_TF_D) ? dm : -dm) * velcorr_value(T, L, x, j, s);
        return;
    }
    if (c.point != j) return;
    int a = -1; // input of the VelocityCorrection callback: 0 time, 1..2nq multibody states, then slacks
    if (c.kind == COL_STATE_D && c.k < 2 * L.nq) a = 1 + c.k;
    else if (c.kind == COL_SLACK_D) a = 1 + 2 * L.nq + c.k;
    if (a < 0) return;
    vals[k] += wc * h * velcorr_fd_entry(T, L, fd, x, j, s, a);
}

__global__ void k_assemble_grad(DevTables T, DevLayout L, const double* __restrict__ x, const double* __restrict__ Y, const double* __restrict__ J,
                                const double* __restrict__ I, const double* __restrict__ E, double* __restrict__ grad) {
    const int vi = blockIdx.x * blockDim.x + threadIdx.x;
    if (vi >= L.n_var) return;
    const ColMetaD c = L.cols[vi];
    double v = 0;
    for (int g = 0; g < T.ngoal; ++g) {
        const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
        if (gr[1] != MOCOB200_MODE_COST) continue;
        for (int o = 0; o < gr[2]; ++o) v += endpoint_entry(L, x, J, I, E, g, o, c);
    }
    const double dur = x[1] - x[0];
    if (L.minimize_mult) {
        if (c.kind == COL_MULT_D) v += 2.0 * L.mult_weight * dur * L.quad[c.point] * x[vi];
        if (c.kind == COL_T0_D || c.kind == COL_TF_D) {
            double acc = 0;
            for (int j = 0; j < L.n_grid; ++j) {
                double sq = 0;
                for (int k = 0; k < L.nm; ++k) { const double d = x[L.off_mult + (size_t)j * L.nm + k]; sq += d * d; }
                acc += L.quad[j] * sq;
            }
            v += (c.kind == COL_TF_D ? 1.0 : -1.0) * L.mult_weight * acc;
        }
    }
    for (int which = 0; which < 2; ++which) {
        const int nacc = L.nd - L.n_auxres; // acceleration variables (none in explicit mode or with prescribed kinematics)
        const bool on = which == 0 ? (L.minimize_accel && nacc > 0) : (L.minimize_auxderiv && L.n_auxres > 0);
        if (!on) continue;
        const int lo = which == 0 ? 0 : nacc, hi = which == 0 ? nacc : L.nd;
        const double w = which == 0 ? L.accel_weight : L.auxderiv_weight;
        if (c.kind == COL_DERIV_D && c.k >= lo && c.k < hi) v += 2.0 * w * dur * L.quad[c.point] * x[vi];
        if (c.kind == COL_T0_D || c.kind == COL_TF_D) {
            double acc = 0;
            for (int j = 0; j < L.n_grid; ++j) {
                double sq = 0;
                for (int k = lo; k < hi; ++k) { const double d = x[L.off_derivs + (size_t)j * L.nd + k]; sq += d * d; }
                acc += L.quad[j] * sq;
            }
            v += (c.kind == COL_TF_D ? 1.0 : -1.0) * w * acc;
        }
    }
    grad[vi] = v;
}

// ---------------------------------------------------------------------------------------------
void launch_points_nominal(const DevTables& T, const DevLayout& L, const double* d_x, double* d_Y, int p0, int np, cudaStream_t s) {
    if (np <= 0) return;
    const int bs = 32;
    if (L.small_size) k_points_nominal<SizeSmall><<<(np + bs - 1) / bs, bs, 0, s>>>(T, L, d_x, d_Y, p0, np);
    else k_points_nominal<SizeLarge><<<(np + bs - 1) / bs, bs, 0, s>>>(T, L, d_x, d_Y, p0, np);
}
// Flat variant: one thread per (grid point, listed direction), nominal outputs read from Y (already evaluated). Every
// lane of every warp carries an evaluation chain, so a short direction list still fills the machine in one wave.
template <int MINB, class SZ>
__global__ void __launch_bounds__(64, MINB)
k_points_fd_flat(DevTables T, DevLayout L, FdSettings fd, const double* __restrict__ x, const double* __restrict__ Y, double* __restrict__ J,
                 int p0, int np, const int* __restrict__ dirs, int ndirs) {
    const long gid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (long)np * ndirs) return;
    const int j = p0 + (int)(gid / ndirs);
    const int a = dirs[gid % ndirs];
    double in[SZ::NIN];
    double ya[SZ::NOUT], yb[SZ::NOUT], Jc[SZ::NOUT];
    gather_inputs(L, x, j, in);
    const double* y0 = Y + (size_t)j * L.n_out;
    if (fd.scheme != MOCOB200_FD_CENTRAL) {
        fd_group([&](const double* xin, double* yout) { eval_point<SZ>(T, xin, yout, 0xffffffffu); }, fd, in, a, y0, 0, L.n_out, yb, ya, Jc);
    } else {
        const int ngroups = 1 + L.n_integrand;
        for (int gidx = 0; gidx < ngroups; ++gidx) {
            const unsigned mask = 1u << gidx;
            const int lo = gidx == 0 ? 0 : L.n_dyn_out + gidx - 1;
            const int hi = gidx == 0 ? L.n_dyn_out : lo + 1;
            fd_group([&](const double* xin, double* yout) { eval_point<SZ>(T, xin, yout, mask); }, fd, in, a, y0, lo, hi, yb, ya, Jc);
        }
    }
    double* Jp = J + (size_t)j * L.n_out * L.n_in;
    for (int o = 0; o < L.n_out; ++o) Jp[(size_t)o * L.n_in + a] = Jc[o];
}
template <int MINB, class SZ>
__global__ void __launch_bounds__(64, MINB)
k_points_fd_flat_pair(DevTables T, DevLayout L, FdSettings fd, const double* __restrict__ x, const double* __restrict__ Y,
                      double* __restrict__ J, int p0, int np, const int* __restrict__ dirs, int ndirs) {
    const long total = (long)np * ndirs * 2;
    long gid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = gid < total;
    if (!valid) gid = total - 1 - ((threadIdx.x ^ 1) & 1); // tail lanes shadow the last pair (every lane joins the shuffles)
    const int side = (int)(gid & 1);
    const long item = gid >> 1;
    const int j = p0 + (int)(item / ndirs);
    const int a = dirs[item % ndirs];
    double in[SZ::NIN];
    double yo[SZ::NOUT], Jc[SZ::NOUT];
    gather_inputs(L, x, j, in);
    const double* y0 = Y + (size_t)j * L.n_out;
    const int ngroups = 1 + L.n_integrand;
    for (int gidx = 0; gidx < ngroups; ++gidx) {
        const unsigned mask = 1u << gidx;
        const int lo = gidx == 0 ? 0 : L.n_dyn_out + gidx - 1;
        const int hi = gidx == 0 ? L.n_dyn_out : lo + 1;
        fd_group_central_pair([&](const double* xin, double* yout) { eval_point<SZ>(T, xin, yout, mask); }, fd, in, a, y0, lo, hi, yo, Jc, side);
    }
    if (!valid || side) return;
    double* Jp = J + (size_t)j * L.n_out * L.n_in;
    for (int o = 0; o < L.n_out; ++o) Jp[(size_t)o * L.n_in + a] = Jc[o];
}
void launch_points_fd_flat(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* d_x, const double* d_Y, double* d_J,
                           int p0, int np, cudaStream_t s, const int* d_dirs, int ndirs) {
    const long n = (long)np * ndirs;
    if (n <= 0) return;
    // MOCOB200_FLAT_MINBLOCKS: resident 64-thread blocks per SM the register allocation must allow (4: 255 registers,
    // 8: 128 registers); MOCOB200_FLAT_PAIR=0: central differences with both sides on one thread.
    // Measured on config 5 (profiles/r01/flat_variants_r01.txt): unpaired 4 -> 427 Jac/s, 8 -> 381; paired 4 -> 493, 8 -> 632:
    // the paired kernel has twice the threads, and only the 128-register build keeps them all resident in one wave.
    // (read per launch so that tests can switch them; steady-state launches are CUDA-graph replays)
    const char* ePair = devEnv("MOCOB200_FLAT_PAIR");
    const bool paired = fd.scheme == MOCOB200_FD_CENTRAL && (!ePair || atoi(ePair) != 0);
    const char* eMinb = devEnv("MOCOB200_FLAT_MINBLOCKS");
    const int minb = eMinb && atoi(eMinb) > 0 ? atoi(eMinb) : (paired ? 8 : 4);
    const bool small = L.small_size != 0;
    if (paired) {
        const unsigned nb = (unsigned)((2 * n + 63) / 64);
        if (small) {
            if (minb == 8) k_points_fd_flat_pair<8, SizeSmall><<<nb, 64, 0, s>>>(T, L, fd, d_x, d_Y, d_J, p0, np, d_dirs, ndirs);
            else k_points_fd_flat_pair<4, SizeSmall><<<nb, 64, 0, s>>>(T, L, fd, d_x, d_Y, d_J, p0, np, d_dirs, ndirs);
        } else {
            if (minb == 8) k_points_fd_flat_pair<8, SizeLarge><<<nb, 64, 0, s>>>(T, L, fd, d_x, d_Y, d_J, p0, np, d_dirs, ndirs);
            else k_points_fd_flat_pair<4, SizeLarge><<<nb, 64, 0, s>>>(T, L, fd, d_x, d_Y, d_J, p0, np, d_dirs, ndirs);
        }
        return;
    }
    const unsigned nb = (unsigned)((n + 63) / 64);
    if (small) {
        if (minb == 8) k_points_fd_flat<8, SizeSmall><<<nb, 64, 0, s>>>(T, L, fd, d_x, d_Y, d_J, p0, np, d_dirs, ndirs);
        else k_points_fd_flat<4, SizeSmall><<<nb, 64, 0, s>>>(T, L, fd, d_x, d_Y, d_J, p0, np, d_dirs, ndirs);
    } else {
        if (minb == 8) k_points_fd_flat<8, SizeLarge><<<nb, 64, 0, s>>>(T, L, fd, d_x, d_Y, d_J, p0, np, d_dirs, ndirs);
        else k_points_fd_flat<4, SizeLarge><<<nb, 64, 0, s>>>(T, L, fd, d_x, d_Y, d_J, p0, np, d_dirs, ndirs);
    }
}

void launch_points_fd(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* d_x, double* d_Y, double* d_J, int p0,
                      int np, cudaStream_t s, const int* d_dirs, int ndirs) {
    if (np <= 0) return;
    if (!d_dirs) ndirs = L.n_in;
    const int bs = ((ndirs + 1 + 31) / 32) * 32;
    k_points_fd<<<np, bs, L.n_out * sizeof(double), s>>>(T, L, fd, d_x, d_Y, d_J, p0, d_dirs, ndirs);
}
void launch_integrals(const DevLayout& L, const double* d_x, const double* d_Y, const double* d_J, double* d_I, bool withJ, cudaStream_t s) {
    if (L.n_integrand <= 0) return;
    k_integrals<<<L.n_integrand, 256, 0, s>>>(L, d_x, d_Y, d_J, d_I, withJ ? 1 : 0);
}
void launch_endpoint(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* d_x, const double* d_I, double* d_G,
                     double* d_E, bool withFd, cudaStream_t s, int which) {
    if (L.n_eout_total <= 0) return;
    const int bs = ((L.n_ein + 1 + 31) / 32) * 32;
    k_endpoint<<<T.ngoal, bs, (L.n_ein + MOCOB200_MAX_EOUT) * sizeof(double), s>>>(T, L, fd, d_x, d_I, d_G, d_E, withFd ? 1 : 0, which);
}
void launch_assemble_g(const DevTables& T, const DevLayout& L, const double* d_x, const double* d_Y, const double* d_G, double* d_g,
                       cudaStream_t s) {
    if (L.n_con <= 0) return;
    k_assemble_g<<<(L.n_con + 127) / 128, 128, 0, s>>>(L, d_x, d_Y, d_G, d_g);
    if (L.n_path_rows > 0) k_assemble_g_path<<<(L.n_con + 127) / 128, 128, 0, s>>>(T, L, d_x, d_g);
    if (L.nsl > 0) k_assemble_g_velcorr<<<(L.n_intervals * L.nq + 127) / 128, 128, 0, s>>>(T, L, d_x, d_g);
}
void launch_assemble_f(const DevTables& T, const DevLayout& L, const double* d_x, const double* d_G, double* d_f, cudaStream_t s) {
    k_assemble_f<<<1, 32, 0, s>>>(T, L, d_x, d_G, d_f);
}
void launch_assemble_jac(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* d_x, const double* d_Y, const double* d_J,
                         const double* d_I, const double* d_E, double* d_vals, cudaStream_t s, int k0, int k1) {
    if (k1 < 0) { k0 = 0; k1 = L.nnz; }
    const int n = k1 - k0;
    if (n <= 0) return;
    k_assemble_jac<<<(n + 255) / 256, 256, 0, s>>>(L, d_x, d_Y, d_J, d_I, d_E, d_vals, k0, k1);
    if (L.n_path_rows > 0) k_assemble_jac_path<<<(n + 255) / 256, 256, 0, s>>>(T, L, fd, d_x, d_vals, k0, k1);
    if (L.nsl > 0) k_assemble_jac_velcorr<<<(n + 255) / 256, 256, 0, s>>>(T, L, fd, d_x, d_vals, k0, k1);
}
void launch_assemble_grad(const DevTables& T, const DevLayout& L, const double* d_x, const double* d_Y, const double* d_J, const double* d_I,
                          const double* d_E, double* d_grad, cudaStream_t s) {
    k_assemble_grad<<<(L.n_var + 127) / 128, 128, 0, s>>>(T, L, d_x, d_Y, d_J, d_I, d_E, d_grad);
}

// ---------------------------------------------------------------------------------------------
// FP64 FMA saturation: 8 independent chains per thread, many resident warps per SM.
__global__ void k_fp64_peak(double* out, int iters) {
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double b = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
        a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
double measure_fp64_peak_tflops(int device) {
    cudaSetDevice(device);
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 15;
    double* d = nullptr;
    if (cudaMalloc(&d, sizeof(double) * blocks * threads) != cudaSuccess) return -1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        k_fp64_peak<<<blocks, threads>>>(d, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 8.0 * (double)iters * blocks * threads;
        if (rep > 0) best = fmax(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    return best;
}

} // namespace mocob200
