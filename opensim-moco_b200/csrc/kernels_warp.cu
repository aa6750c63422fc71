// Warp-per-evaluation kernels: point functions over the grid and their CasADi-style finite differences.
// Work item = (grid point, FD direction[, callback group]); each warp owns ws_doubles of shared memory.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "kernels.cuh"
#include "point_warp_local.cuh"

namespace mocob200 {

__device__ __forceinline__ void warp_gather_inputs(const DevLayout& L, const double* __restrict__ x, int j, double* in, int lane) {
    const double t0 = x[0], tf = x[1];
    if (lane == 0) in[0] = t0 + (tf - t0) * L.grid[j];
    const double* xs = x + L.off_states + (size_t)j * L.ns;
    for (int k = lane; k < L.ns; k += 32) in[1 + k] = xs[k];
    const double* xc = x + L.off_controls + (size_t)j * L.nc;
    for (int k = lane; k < L.nc; k += 32) in[1 + L.ns + k] = xc[k];
    const double* xm = x + L.off_mult + (size_t)j * L.nm;
    for (int k = lane; k < L.nm; k += 32) in[1 + L.ns + L.nc + k] = xm[k];
    const double* xd = x + L.off_derivs + (size_t)j * L.nd;
    for (int k = lane; k < L.nd; k += 32) in[1 + L.ns + L.nc + L.nm + k] = xd[k];
    __syncwarp();
}

// Copies with several independent loads in flight per thread (a plain copy loop serialises on the load latency).
template <class V>
__device__ __forceinline__ void copy_batched(V* __restrict__ dst, const V* __restrict__ src, int n, int tid, int nt) {
    int i = tid;
    for (; i + 7 * nt < n; i += 8 * nt) {
        V v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = src[i + k * nt];
#pragma unroll
        for (int k = 0; k < 8; ++k) dst[i + k * nt] = v[k];
    }
    for (; i < n; i += nt) dst[i] = src[i];
}

// Block prologue: copy the model tables and the topology arrays into shared memory (coalesced, all threads), and
// return table descriptors whose pointers address the staged copies. Every thread of the block must call this.
__device__ __forceinline__ double* stage_tables(const DevTables& T, const WarpAux& A, double* smem, DevTables& Ts, WarpAux& As) {
    double* sd = smem;
    const int nd = A.st_ndbl, ng = A.st_ngoal;
    int* si = (int*)(smem + ((nd + ng + 1) & ~1));
    int* sa = si + A.st_nint;
    const int tid = threadIdx.x, nt = blockDim.x;
    copy_batched((double2*)sd, (const double2*)T.dbl, nd / 2, tid, nt);
    if ((nd & 1) && tid == 0) sd[nd - 1] = T.dbl[nd - 1];
    copy_batched(sd + nd, T.dbl + A.st_goal0, ng, tid, nt);
    copy_batched(si, T.idx, A.st_nint, tid, nt);
    copy_batched(sa, A.aux_base, A.st_naux, tid, nt);
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
    if ((((size_t)j * n) & 1) == 0) { // 16-byte aligned blob: two doubles per load
        copy_batched((double2*)(ws + A.blob_begin), (const double2*)src, m / 2, lane, 32);
        if ((m & 1) && lane == 0) ws[A.blob_begin + m - 1] = src[m - 1];
    } else copy_batched(ws + A.blob_begin, src, m, lane, 32);
    __syncwarp();
}
// which part of the nominal evaluation survives a perturbation of input `dir`:
// 0 = nothing (a coordinate), 1 = poses / path geometry / mass-matrix factor (a speed, or an acceleration variable in
// implicit mode), 2 = velocities and contact forces as well (time, auxiliary states, controls, auxiliary derivatives)
__device__ __forceinline__ int reuse_level(const DevLayout& L, int dir) {
    if (L.prescribed) return dir == 0 ? 0 : 2; // time moves every coordinate; no coordinate, speed or acceleration inputs exist
    if (dir >= 1 && dir <= L.nq) return 0;
    if (dir > L.nq && dir <= 2 * L.nq) return 1;
    if (L.implicit && dir >= 1 + L.ns + L.nc + L.nm && dir < 1 + L.ns + L.nc + L.nm + L.nq) return 1;
    return 2;
}

// nominal evaluation of every grid point in [p0, p0+np): Y[j][:] (+ the blob of each point when cfg != nullptr)
template <bool WRAP>
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
#ifdef MOCOB200_PROFILE_PHASES
    const long long clk0 = clock64();
    struct ItemClock { long item; long long c0; int lane;
        __device__ ~ItemClock() { if (lane == 0 && item < 16384) g_item_clk[0][item] = clock64() - c0; } } itemClock{item, clk0, lane};
#endif
    poison_scratch(A, ws, lane);
    warp_gather_inputs(L, x, j, ws + A.o_in, lane);
    warp_eval_point_t<0, WRAP>(T, A, ws, 0xffffffffu, false, false);
    for (int o = lane; o < L.n_out; o += 32) Y[(size_t)j * L.n_out + o] = ws[A.o_out + o];
    if (cfg) blob_save(A, ws, cfg, j, lane);
}

// One-sided differences. Four kernels, each holding only the code of its own kind of work item (the point function is
// long, mostly straight-line code: a kernel that carries all variants runs every one of them at half speed on
// instruction fetch):
//   KIND 0  item = point * (1 + n_p0) + k; k = 0 is the nominal evaluation (writes Y and the blob), k >= 1 perturbs
//           coordinate p0_dirs[k-1] (a full evaluation);
//   KIND 1  item = point * n_p1 + t perturbs p1_dirs[t] (speeds, implicit-mode accelerations): reloads the poses;
//   KIND 2  item = point * n_p2 + t perturbs p2_dirs[t] (time, unclassified inputs): reloads poses and velocities;
//   KIND 3  item = point * n_lparts + part: the columns of the "local" inputs (point_warp_local.cuh) of 32 owners
//           (muscles, coordinate actuators) from the reloaded blob.
// KIND 1-3 run side by side once KIND 0 is done and write finished difference quotients. The coordinate evaluations of
// KIND 0 run beside the nominal one, so they leave their RAW outputs in J[j][o][dir]; the first KIND 2 item of each grid
// point (or k_fd1_finalize when there is none) turns those columns into quotients.
template <int KIND, bool WRAP = false>
__global__ void k_wpoints_fd1(DevTables Tg, WarpAux Ag, DevLayout L, FdSettings fd, const double* __restrict__ x, double* __restrict__ Y,
                              double* __restrict__ J, double* __restrict__ cfg, int p0, int np) {
    extern __shared__ double smem[];
    DevTables T;
    WarpAux A;
    double* scratch = stage_tables(Tg, Ag, smem, T, A);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long item = (long)blockIdx.x * (blockDim.x >> 5) + warp;
    const int per = KIND == 0 ? 1 + A.n_p0 : KIND == 1 ? A.n_p1 : KIND == 2 ? A.n_p2 : A.n_lparts;
    if (item >= (long)np * per) return;
    double* ws = scratch + (size_t)warp * A.ws_doubles;
    const int j = p0 + (int)(item / per), t = (int)(item % per);
    double* Jp = J + (size_t)j * L.n_out * L.n_in;
#ifdef MOCOB200_PROFILE_PHASES
    const long long clk0 = clock64();
    struct ItemClock { long item; long long c0; int lane;
        __device__ ~ItemClock() { if (lane == 0 && item < 16384) g_item_clk[KIND][item] = clock64() - c0; } } itemClock{item, clk0, lane};
#endif
    poison_scratch(A, ws, lane);
    warp_gather_inputs(L, x, j, ws + A.o_in, lane);
    if (KIND == 3) {
        blob_load(A, ws, cfg, j, lane, 2);
        wk_local_columns(T, A, ws, ws + A.o_Rm, fd.h, fd.scheme == MOCOB200_FD_FORWARD, Jp, Y + (size_t)j * L.n_out, L.n_in, L.n_out, t, per);
        return;
    }
    const int dir = KIND == 0 ? (t == 0 ? -1 : A.p0_dirs[t - 1]) : KIND == 1 ? A.p1_dirs[t] : A.p2_dirs[t];
    if (dir >= 0 && lane == 0) ws[A.o_in + dir] += (fd.scheme == MOCOB200_FD_FORWARD) ? fd.h : -fd.h;
    if (KIND > 0) blob_load(A, ws, cfg, j, lane, KIND);
    __syncwarp();
    unsigned mask = 0xffffffffu;
    if (KIND == 2 && dir == 0 && A.time_integrands_only) {
        // nothing but the tracking references depends on time: the dynamics rows keep their nominal values
        mask = 0xfffffffeu;
        for (int o = lane; o < L.n_dyn_out; o += 32) ws[A.o_out + o] = Y[(size_t)j * L.n_out + o];
    }
    warp_eval_point_t<KIND, WRAP>(T, A, ws, mask, KIND == 2 && dir == 0, KIND <= 1 && A.lockstep != 0);
    const double inv = (fd.scheme == MOCOB200_FD_FORWARD) ? 1.0 / fd.h : -1.0 / fd.h;
    if (dir < 0) {
        for (int o = lane; o < L.n_out; o += 32) Y[(size_t)j * L.n_out + o] = ws[A.o_out + o];
        blob_save(A, ws, cfg, j, lane);
    } else if (KIND == 0) {
        for (int o = lane; o < L.n_out; o += 32) Jp[(size_t)o * L.n_in + dir] = ws[A.o_out + o];
    } else {
        const double* y0 = Y + (size_t)j * L.n_out;
        for (int o = lane; o < L.n_out; o += 32) Jp[(size_t)o * L.n_in + dir] = (ws[A.o_out + o] - y0[o]) * inv;
        if (KIND == 2 && t == 0) { // finish the coordinate columns of this grid point
            for (int e = lane; e < L.n_out * A.n_p0; e += 32) {
                const int o = e / A.n_p0, c = A.p0_dirs[e - o * A.n_p0];
                double* v = Jp + (size_t)o * L.n_in + c;
                *v = (*v - y0[o]) * inv;
            }
        }
    }
}
// fallback when no KIND 2 work exists: quotients of the coordinate columns (dir_class 0) only
__global__ void k_fd1_finalize(DevLayout L, FdSettings fd, const int* __restrict__ dir_class, const double* __restrict__ Y,
                               double* __restrict__ J, int p0, int np) {
    const long n = (long)np * L.n_out * L.n_in;
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const long base = (long)p0 * L.n_out * L.n_in;
    const long row = t / L.n_in; // (point - p0) * n_out + o
    if (dir_class[t - row * L.n_in] != 0) return; // every other column is written finished
    const double y0 = Y[(size_t)p0 * L.n_out + row];
    const double yp = J[base + t];
    J[base + t] = (fd.scheme == MOCOB200_FD_FORWARD) ? (yp - y0) / fd.h : (y0 - yp) / fd.h;
}

// central differences with CasADi's step refinement; item = (point, dir, callback group). Needs Y and the blobs of
// the nominal evaluation.
template <bool WRAP>
__global__ void k_wpoints_central(DevTables Tg, WarpAux Ag, DevLayout L, FdSettings fd, const double* __restrict__ x,
                                  const double* __restrict__ Y, double* __restrict__ J, const double* __restrict__ cfg, int p0, int np) {
    extern __shared__ double smem[];
    DevTables T;
    WarpAux A;
    double* scratch = stage_tables(Tg, Ag, smem, T, A);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long item = (long)blockIdx.x * (blockDim.x >> 5) + warp;
    const int ngroups = 1 + L.n_integrand;
    const long per = (long)A.n_cd * ngroups;
    if (item >= (long)np * per) return;
    double* ws = scratch + (size_t)warp * A.ws_doubles;
    const int j = p0 + (int)(item / per);
    const int rem = (int)(item % per);
    const int dir = A.cd_dirs[rem / ngroups], gidx = rem % ngroups;
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
#pragma unroll 1
        for (int side = 0; side < 2; ++side) { // one copy of the evaluation code serves both sides
            if (lane == 0) ws[A.o_in + dir] = side ? x0 + h : x0 - h;
            __syncwarp();
            warp_eval_point<WRAP>(T, A, ws, mask, reuse, dir == 0);
            if (side == 0) for (int o = lo + lane; o < hi; o += 32) yb[o] = yf[o];
        }
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

// columns of the local inputs under the central scheme: one warp per grid point, from the reloaded blob
__global__ void k_wlocal_central(DevTables Tg, WarpAux Ag, DevLayout L, FdSettings fd, const double* __restrict__ x,
                                 const double* __restrict__ Y, double* __restrict__ J, const double* __restrict__ cfg, int p0, int np) {
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
    blob_load(A, ws, cfg, j, lane, 2);
    wk_local_columns_central(T, A, ws, ws + A.o_Rm, fd.h, fd.h_iter, fd.u_aim, fd.reltol, fd.abstol, J + (size_t)j * L.n_out * L.n_in,
                             Y + (size_t)j * L.n_out, L.n_in, L.n_out);
}

// ---------------------------------------------------------------------------------------------
static size_t smemBytesFor(const WarpAux& A, int warps) { return ((size_t)A.st_doubles + (size_t)warps * A.ws_doubles) * sizeof(double); }
static size_t warpSmemBytes(const WarpAux& A) { return smemBytesFor(A, A.wpb); }
// phase 1 of the one-sided scheme: all partial evaluations of one grid point in one block when they fit, so that the
// point's blob is read from L2 once and served to the other warps by L1
static int phase1Warps(const WarpAux& A) {
    const int per = A.n_p1;
    static const int forced = devEnv("MOCOB200_P1W") ? atoi(devEnv("MOCOB200_P1W")) : 0; // A-B tests: warps per phase-1 block
    if (forced >= 1 && forced <= 32 && smemBytesFor(A, forced) <= 220 * 1024) return forced;
    return (per >= 1 && per <= 32 && smemBytesFor(A, per) <= 220 * 1024) ? per : A.wpb;
}

bool warp_kernels_configure(WarpAux& A, std::string& err) {
    // choose the number of warps per block so that a block's scratch fits and several blocks share an SM
    A.wpb = 4;
    if (const char* e = devEnv("MOCOB200_WPB")) A.wpb = atoi(e) > 0 ? atoi(e) : 4;
    { const char* e = devEnv("MOCOB200_LOCKSTEP"); A.lockstep = (e && e[0] == '1') ? 1 : 0; }
    while (A.wpb > 1 && warpSmemBytes(A) > (devEnv("MOCOB200_WPB") ? 220 : 96) * 1024) A.wpb >>= 1;
    const size_t bytes = std::max(warpSmemBytes(A), smemBytesFor(A, phase1Warps(A)));
    if (warpSmemBytes(A) > 227 * 1024) { err = "per-warp scratch exceeds shared memory"; return false; }
    // the attribute is per function, not per handle: only ever raise it
    static size_t g_maxBytes[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 8 && bytes <= g_maxBytes[dev]) return true;
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_wpoints_nominal<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_nominal<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_fd1<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_fd1<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_central<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_fd1<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_fd1<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_fd1<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wlocal_central, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_central<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess) {
        err = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e);
        return false;
    }
    if (dev >= 0 && dev < 8) g_maxBytes[dev] = bytes;
    return true;
}

void launch_wpoints_nominal(const DevTables& T, const WarpAux& A, const DevLayout& L, const double* d_x, double* d_Y, double* d_cfg,
                            int p0, int np, cudaStream_t s) {
    if (np <= 0) return;
    static const bool timing = devEnv("MOCOB200_TIMING") != nullptr; // development stopwatch
    static cudaEvent_t ne[2] = {nullptr, nullptr};
    if (timing) {
        if (!ne[0]) { cudaEventCreate(&ne[0]); cudaEventCreate(&ne[1]); }
        else { float a = 0; if (cudaEventElapsedTime(&a, ne[0], ne[1]) == cudaSuccess) fprintf(stderr, "[mocob200 timing] nominal %.1f us (%d points)\n", 1e3 * a, np); }
        cudaEventRecord(ne[0], s);
    }
    if (T.npathwrap > 0) k_wpoints_nominal<true><<<(np + A.wpb - 1) / A.wpb, A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, d_x, d_Y, d_cfg, p0, np);
    else k_wpoints_nominal<false><<<(np + A.wpb - 1) / A.wpb, A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, d_x, d_Y, d_cfg, p0, np);
    if (timing) cudaEventRecord(ne[1], s);
}
// returns the number of kernels launched
int launch_wpoints_fd(const DevTables& T, const WarpAux& A, const DevLayout& L, const FdSettings& fd, const double* d_x, double* d_Y,
                      double* d_J, double* d_cfg, int p0, int np, cudaStream_t s, cudaEvent_t after_nominal, const SideStreams* side) {
    if (np <= 0) { if (after_nominal) cudaEventRecord(after_nominal, s); return 0; }
    if (fd.scheme != MOCOB200_FD_CENTRAL) {
        const long items0 = (long)np * (1 + A.n_p0), items1 = (long)np * A.n_p1, items2 = (long)np * A.n_p2;
        // development stopwatch (MOCOB200_TIMING=1): per-stage CUDA-event times of the previous call on stderr
        static const bool timing = devEnv("MOCOB200_TIMING") != nullptr;
        static cudaEvent_t te[4] = {nullptr, nullptr, nullptr, nullptr};
        if (timing) {
            if (!te[0]) for (auto& e : te) cudaEventCreate(&e);
            else if (cudaEventQuery(te[3]) == cudaSuccess) {
                float a = 0, b = 0, c = 0;
                cudaEventElapsedTime(&a, te[0], te[1]); cudaEventElapsedTime(&b, te[1], te[2]); cudaEventElapsedTime(&c, te[2], te[3]);
                fprintf(stderr, "[mocob200 timing] full %.1f us  partial+local %.1f us  finalize %.1f us\n", 1e3 * a, 1e3 * b, 1e3 * c);
            }
            cudaEventRecord(te[0], s);
        }
        if (T.npathwrap > 0) k_wpoints_fd1<0, true><<<(unsigned)((items0 + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
        else k_wpoints_fd1<0, false><<<(unsigned)((items0 + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
        if (after_nominal) cudaEventRecord(after_nominal, s);
        if (timing) cudaEventRecord(te[1], s);
        int launched = 1;
        // the three kinds of blob-reloading work run side by side
        const bool fork = side && side->s[0] && side->s[1];
        cudaStream_t s2 = fork ? side->s[0] : s, s3 = fork ? side->s[1] : s;
        if (fork) {
            cudaEventRecord(side->fork, s);
            cudaStreamWaitEvent(s2, side->fork, 0);
            cudaStreamWaitEvent(s3, side->fork, 0);
        }
        if (items1 > 0) {
            const int w1 = phase1Warps(A);
            k_wpoints_fd1<1><<<(unsigned)((items1 + w1 - 1) / w1), w1 * 32, smemBytesFor(A, w1), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
            ++launched;
        }
        if (items2 > 0) {
            k_wpoints_fd1<2><<<(unsigned)((items2 + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s2>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
            ++launched;
        }
        if (A.n_local > 0) {
            k_wpoints_fd1<3><<<(unsigned)(((long)np * A.n_lparts + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s3>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
            ++launched;
        }
        if (fork) {
            cudaEventRecord(side->join[0], s2);
            cudaEventRecord(side->join[1], s3);
            cudaStreamWaitEvent(s, side->join[0], 0);
            cudaStreamWaitEvent(s, side->join[1], 0);
        }
        if (timing) cudaEventRecord(te[2], s);
        if (items2 == 0) {
            const long n = (long)np * L.n_out * L.n_in;
            k_fd1_finalize<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(L, fd, A.dir_class, d_Y, d_J, p0, np);
            ++launched;
        }
        if (timing) cudaEventRecord(te[3], s);
        return launched;
    }
    launch_wpoints_nominal(T, A, L, d_x, d_Y, d_cfg, p0, np, s);
    if (after_nominal) cudaEventRecord(after_nominal, s);
    int launched = 2;
    const bool fork = A.n_local > 0 && side && side->s[0];
    if (fork) {
        cudaEventRecord(side->fork, s);
        cudaStreamWaitEvent(side->s[0], side->fork, 0);
        k_wlocal_central<<<(np + A.wpb - 1) / A.wpb, A.wpb * 32, warpSmemBytes(A), side->s[0]>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
        cudaEventRecord(side->join[0], side->s[0]);
        ++launched;
    } else if (A.n_local > 0) {
        k_wlocal_central<<<(np + A.wpb - 1) / A.wpb, A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
        ++launched;
    }
    const long items = (long)np * A.n_cd * (1 + L.n_integrand);
    if (items > 0)
        if (T.npathwrap > 0) k_wpoints_central<true><<<(unsigned)((items + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
        else k_wpoints_central<false><<<(unsigned)((items + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
    if (fork) cudaStreamWaitEvent(s, side->join[0], 0);
    return launched;
}

int launch_wlocal(const DevTables& T, const WarpAux& A, const DevLayout& L, const FdSettings& fd, const double* d_x, double* d_Y,
                  double* d_J, double* d_cfg, int p0, int np, cudaStream_t s) {
    if (np <= 0 || A.n_local <= 0) return 0;
    // d_Y and the blobs in d_cfg come from launch_wpoints_nominal
    if (fd.scheme == MOCOB200_FD_CENTRAL)
        k_wlocal_central<<<(np + A.wpb - 1) / A.wpb, A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
    else
        k_wpoints_fd1<3><<<(unsigned)(((long)np * A.n_lparts + A.wpb - 1) / A.wpb), A.wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np);
    return 1;
}

} // namespace mocob200

#ifdef MOCOB200_PROFILE_PHASES
extern "C" int mocob200_debug_item_clocks(long long* out, int phase, int n) {
    return cudaMemcpyFromSymbol(out, mocob200::g_item_clk, sizeof(long long) * n, sizeof(long long) * 16384 * phase) == cudaSuccess ? 0 : 3;
}
extern "C" int mocob200_debug_phase_clocks(long long* out32) {
    return cudaMemcpyFromSymbol(out32, mocob200::g_phase_clk, sizeof(long long) * 32) == cudaSuccess ? 0 : 3;
}
#endif
