// Warp-per-evaluation kernels: point functions over the grid and their CasADi-style finite differences.
// Work item = (grid point, FD direction[, callback group]); each warp owns ws_doubles of shared memory.
#include "kernels.cuh"
#include "point_warp_dyn.cuh"

namespace mocob200 {

__device__ __forceinline__ void warp_gather_inputs(const DevLayout& L, const double* __restrict__ x, int j, double* in, int lane) {
    const double t0 = x[0], tf = x[1];
    if (lane == 0) in[0] = t0 + (tf - t0) * L.grid[j];
    const double* xs = x + L.off_states + (size_t)j * L.ns;
    for (int k = lane; k < L.ns; k += 32) in[1 + k] = xs[k];
    const double* xc = x + L.off_controls + (size_t)j * L.nc;
    for (int k = lane; k < L.nc; k += 32) in[1 + L.ns + k] = xc[k];
    const double* xd = x + L.off_derivs + (size_t)j * L.nd;
    for (int k = lane; k < L.nd; k += 32) in[1 + L.ns + L.nc + k] = xd[k];
    __syncwarp();
}

// Block prologue: copy the model tables and the topology arrays into shared memory (coalesced, all threads), and
// return table descriptors whose pointers address the staged copies. Every thread of the block must call this.
__device__ __forceinline__ double* stage_tables(const DevTables& T, const WarpAux& A, double* smem, DevTables& Ts, WarpAux& As) {
    double* sd = smem;
    const int nd = A.st_ndbl, ng = A.st_ngoal;
    int* si = (int*)(smem + ((nd + ng + 1) & ~1));
    int* sa = si + A.st_nint;
    const int tid = threadIdx.x, nt = blockDim.x;
    {
        const double2* src = (const double2*)T.dbl;
        double2* dst = (double2*)sd;
        for (int i = tid; i < nd / 2; i += nt) dst[i] = src[i];
        if ((nd & 1) && tid == 0) sd[nd - 1] = T.dbl[nd - 1];
    }
    for (int i = tid; i < ng; i += nt) sd[nd + i] = T.dbl[A.st_goal0 + i];
    for (int i = tid; i < A.st_nint; i += nt) si[i] = T.idx[i];
    for (int i = tid; i < A.st_naux; i += nt) sa[i] = A.aux_base[i];
    __syncthreads();
    Ts = T;
    Ts.dblP = T.dbl;
    Ts.dbl = sd;
    Ts.dblG = sd + nd - A.st_goal0;
    Ts.idx = si;
    As = A;
#define MOCOB200_RESTAGE(f) As.f = sa + (A.f - A.aux_base)
    MOCOB200_RESTAGE(level_start); MOCOB200_RESTAGE(child_start); MOCOB200_RESTAGE(child_list); MOCOB200_RESTAGE(body_pp_start);
    MOCOB200_RESTAGE(body_pp_list); MOCOB200_RESTAGE(dof_body); MOCOB200_RESTAGE(body_flags); MOCOB200_RESTAGE(pp_mpp);
    MOCOB200_RESTAGE(mpp_pp); MOCOB200_RESTAGE(mpp_coord); MOCOB200_RESTAGE(pair_i); MOCOB200_RESTAGE(pair_j);
#undef MOCOB200_RESTAGE
    return smem + A.st_doubles;
}

__device__ __forceinline__ void poison_scratch(const WarpAux& A, double* ws, int lane) {
    if (!A.poison) return;
    for (int i = lane; i < A.ws_doubles; i += 32) ws[i] = NAN;
    __syncwarp();
}
__device__ __forceinline__ void blob_save(const WarpAux& A, const double* ws, double* __restrict__ cfg, int j, int lane) {
    const int n = A.blob_end - A.blob_begin;
    double* dst = cfg + (size_t)j * n;
    for (int i = lane; i < n; i += 32) dst[i] = ws[A.blob_begin + i];
}
__device__ __forceinline__ void blob_load(const WarpAux& A, double* ws, const double* __restrict__ cfg, int j, int lane, int reuse) {
    const int n = A.blob_end - A.blob_begin;
    const int m = (reuse == 2 ? A.blob_end : A.blob_pos_end) - A.blob_begin;
    const double* src = cfg + (size_t)j * n;
    for (int i = lane; i < m; i += 32) ws[A.blob_begin + i] = src[i];
    __syncwarp();
}
// which part of the nominal evaluation survives a perturbation of input `dir`:
// 0 = nothing (a coordinate), 1 = poses / path geometry / mass-matrix factor (a speed, or an acceleration variable in
// implicit mode), 2 = velocities and contact forces as well (time, auxiliary states, controls, auxiliary derivatives)
__device__ __forceinline__ int reuse_level(const DevLayout& L, int dir) {
    if (dir >= 1 && dir <= L.nq) return 0;
    if (dir > L.nq && dir <= 2 * L.nq) return 1;
    if (L.implicit && dir >= 1 + L.ns + L.nc && dir < 1 + L.ns + L.nc + L.nq) return 1;
    return 2;
}

// nominal evaluation of every grid point in [p0, p0+np): Y[j][:] (+ the blob of each point when cfg != nullptr)
__global__ void k_wpoints_nominal(DevTables Tg, WarpAux Ag, DevLayout L, const double* __restrict__ x, double* __restrict__ Y,
                                  double* __restrict__ cfg, int p0, int np) {
    extern __shared__ double smem[];
    DevTables T;
    WarpAux A;
    double* scratch = stage_tables(Tg, Ag, smem, T, A);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item = blockIdx.x * (blockDim.x >> 5) + warp;
    if (item >= np) return;
    double* ws = scratch + (size_t)warp * A.ws_doubles;
    const int j = p0 + item;
    poison_scratch(A, ws, lane);
    warp_gather_inputs(L, x, j, ws + A.o_in, lane);
    warp_eval_point(T, A, ws, 0xffffffffu, 0, false);
    for (int o = lane; o < L.n_out; o += 32) Y[(size_t)j * L.n_out + o] = ws[A.o_out + o];
    if (cfg) blob_save(A, ws, cfg, j, lane);
}

// one-sided differences, two launches:
//   phase 0: item = point * (1 + nq) + k; k = 0 is the nominal evaluation (writes Y and the blob), k >= 1 perturbs
//            coordinate k (a full evaluation);
//   phase 1: item = point * (n_in - nq) + t; t = 0 perturbs time, t >= 1 perturbs input nq + t; these reload the blob.
// Perturbed evaluations write their RAW outputs into J[j][o][dir]; k_fd1_finalize turns them into differences.
__global__ void k_wpoints_fd1(DevTables Tg, WarpAux Ag, DevLayout L, FdSettings fd, const double* __restrict__ x, double* __restrict__ Y,
                              double* __restrict__ J, double* __restrict__ cfg, int p0, int np, int phase) {
    extern __shared__ double smem[];
    DevTables T;
    WarpAux A;
    double* scratch = stage_tables(Tg, Ag, smem, T, A);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long item = (long)blockIdx.x * (blockDim.x >> 5) + warp;
    const int per = phase == 0 ? 1 + L.nq : L.n_in - L.nq;
    if (item >= (long)np * per) return;
    double* ws = scratch + (size_t)warp * A.ws_doubles;
    const int j = p0 + (int)(item / per), t = (int)(item % per);
    const int dir = phase == 0 ? (t == 0 ? -1 : t) : (t == 0 ? 0 : L.nq + t);
    const int reuse = dir < 0 ? 0 : reuse_level(L, dir);
    poison_scratch(A, ws, lane);
    warp_gather_inputs(L, x, j, ws + A.o_in, lane);
    if (dir >= 0 && lane == 0) ws[A.o_in + dir] += (fd.scheme == MOCOB200_FD_FORWARD) ? fd.h : -fd.h;
    if (reuse) blob_load(A, ws, cfg, j, lane, reuse);
    __syncwarp();
    warp_eval_point(T, A, ws, 0xffffffffu, reuse, dir == 0);
    if (dir < 0) {
        for (int o = lane; o < L.n_out; o += 32) Y[(size_t)j * L.n_out + o] = ws[A.o_out + o];
        blob_save(A, ws, cfg, j, lane);
    } else {
        double* Jp = J + (size_t)j * L.n_out * L.n_in;
        for (int o = lane; o < L.n_out; o += 32) Jp[(size_t)o * L.n_in + dir] = ws[A.o_out + o];
    }
}
__global__ void k_fd1_finalize(DevLayout L, FdSettings fd, const double* __restrict__ Y, double* __restrict__ J, int p0, int np) {
    const long n = (long)np * L.n_out * L.n_in;
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const long base = (long)p0 * L.n_out * L.n_in;
    const long row = t / L.n_in; // (point - p0) * n_out + o
    const double y0 = Y[(size_t)p0 * L.n_out + row];
    const double yp = J[base + t];
    J[base + t] = (fd.scheme == MOCOB200_FD_FORWARD) ? (yp - y0) / fd.h : (y0 - yp) / fd.h;
}

// central differences with CasADi's step refinement; item = (point, dir, callback group). Needs Y and the blobs of
// the nominal evaluation.
__global__ void k_wpoints_central(DevTables Tg, WarpAux Ag, DevLayout L, FdSettings fd, const double* __restrict__ x,
                                  const double* __restrict__ Y, double* __restrict__ J, const double* __restrict__ cfg, int p0, int np) {
    extern __shared__ double smem[];
    DevTables T;
    WarpAux A;
    double* scratch = stage_tables(Tg, Ag, smem, T, A);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long item = (long)blockIdx.x * (blockDim.x >> 5) + warp;
    const int ngroups = 1 + L.n_integrand;
    const long per = (long)L.n_in * ngroups;
    if (item >= (long)np * per) return;
    double* ws = scratch + (size_t)warp * A.ws_doubles;
    const int j = p0 + (int)(item / per);
    const int rem = (int)(item % per);
    const int dir = rem / ngroups, gidx = rem % ngroups;
    const int lo = gidx == 0 ? 0 : L.n_dyn_out + gidx - 1;
    const int hi = gidx == 0 ? L.n_dyn_out : lo + 1;
    if (hi == lo) return;
    const unsigned mask = 1u << gidx;
    const int reuse = reuse_level(L, dir);
    poison_scratch(A, ws, lane);
    warp_gather_inputs(L, x, j, ws + A.o_in, lane);
    if (reuse) blob_load(A, ws, cfg, j, lane, reuse);
    const double x0 = ws[A.o_in + dir];
    const double* y0 = Y + (size_t)j * L.n_out;
    double* yb = ws + A.o_out2;
    double* yf = ws + A.o_out;
    double* Jp = J + (size_t)j * L.n_out * L.n_in;
    double h = fd.h;
    for (int iter = 0;; ++iter) {
        if (lane == 0) ws[A.o_in + dir] = x0 - h;
        __syncwarp();
        warp_eval_point(T, A, ws, mask, reuse, dir == 0);
        for (int o = lo + lane; o < hi; o += 32) yb[o] = yf[o];
        if (lane == 0) ws[A.o_in + dir] = x0 + h;
        __syncwarp();
        warp_eval_point(T, A, ws, mask, reuse, dir == 0);
        double u = 0;
        bool bad = false;
        for (int o = lo + lane; o < hi; o += 32) {
            const double vf = yf[o], vc = y0[o], vb = yb[o];
            double jv;
            if (!isfinite(vf) || !isfinite(vc) || !isfinite(vb)) { jv = NAN; bad = true; }
            else {
                jv = (vf - vb) / (2 * h);
                const double et = vf - 2 * vc + vb;
                const double er = fd.reltol / h * fmax(fabs(vf - vc), fabs(vc - vb)) + fd.abstol;
                u = fmax(u, fabs(et / er));
            }
            if (iter == fd.h_iter) Jp[(size_t)o * L.n_in + dir] = jv;
        }
        if (iter == fd.h_iter) break;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) u = fmax(u, __shfl_xor_sync(0xffffffffu, u, s));
        bad = __any_sync(0xffffffffu, bad);
        if (bad) h /= fd.u_aim;
        else h *= sqrt(fd.u_aim / fmax(1.0, u));
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------
static size_t warpSmemBytes(const WarpAux& A) { return ((size_t)A.st_doubles + (size_t)A.wpb * A.ws_doubles) * sizeof(double); }

bool warp_kernels_configure(WarpAux& A, std::string& err) {
    // choose the number of warps per block so that a block's scratch fits and several blocks share an SM
    A.wpb = 4;
    while (A.wpb > 1 && warpSmemBytes(A) > 96 * 1024) A.wpb >>= 1;
    const size_t bytes = warpSmemBytes(A);
    if (bytes > 227 * 1024) { err = "per-warp scratch exceeds shared memory"; return false; }
    // the attribute is per function, not per handle: only ever raise it
    static size_t g_maxBytes[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 8 && bytes <= g_maxBytes[dev]) return true;
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_wpoints_nominal, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_fd1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_central, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess) {
        err = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e);
        return false;
    }
    if (dev >= 0 && dev < 8) g_maxBytes[dev] = bytes;
    return true;
}

void launch_wpoints_nominal(const DevTables& T, const WarpAux& A, const DevLayout& L, const double* d_x, double* d_Y, double* d_cfg,
                            int p0, int np, cudaStream_t s) {
    if (np <= 0) return;
    k_wpoints_nominal<<<(np + A.wpb - 1) / A.wpb, A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, d_x, d_Y, d_cfg, p0, np);
}
// returns the number of kernels launched
int launch_wpoints_fd(const DevTables& T, const WarpAux& A, const DevLayout& L, const FdSettings& fd, const double* d_x, double* d_Y,
                      double* d_J, double* d_cfg, int p0, int np, cudaStream_t s, cudaEvent_t after_nominal) {
    if (np <= 0) { if (after_nominal) cudaEventRecord(after_nominal, s); return 0; }
    if (fd.scheme != MOCOB200_FD_CENTRAL) {
        const long items0 = (long)np * (1 + L.nq), items1 = (long)np * (L.n_in - L.nq);
        k_wpoints_fd1<<<(unsigned)((items0 + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np, 0);
        if (after_nominal) cudaEventRecord(after_nominal, s);
        k_wpoints_fd1<<<(unsigned)((items1 + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np, 1);
        const long n = (long)np * L.n_out * L.n_in;
        k_fd1_finalize<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(L, fd, d_Y, d_J, p0, np);
        return 3;
    }
    launch_wpoints_nominal(T, A, L, d_x, d_Y, d_cfg, p0, np, s);
    if (after_nominal) cudaEventRecord(after_nominal, s);
    const long items = (long)np * L.n_in * (1 + L.n_integrand);
    k_wpoints_central<<<(unsigned)((items + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
    return 2;
}

} // namespace mocob200

#ifdef MOCOB200_PROFILE_PHASES
extern "C" int mocob200_debug_phase_clocks(long long* out32) {
    return cudaMemcpyFromSymbol(out32, mocob200::g_phase_clk, sizeof(long long) * 32) == cudaSuccess ? 0 : 3;
}
#endif
