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

// nominal evaluation of every grid point in [p0, p0+np): Y[j][:]
__global__ void k_wpoints_nominal(DevTables T, WarpAux A, DevLayout L, const double* __restrict__ x, double* __restrict__ Y, int p0, int np) {
    extern __shared__ double smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item = blockIdx.x * (blockDim.x >> 5) + warp;
    if (item >= np) return;
    double* ws = smem + (size_t)warp * A.ws_doubles;
    const int j = p0 + item;
    warp_gather_inputs(L, x, j, ws + A.o_in, lane);
    warp_eval_point(T, A, ws, 0xffffffffu);
    for (int o = lane; o < L.n_out; o += 32) Y[(size_t)j * L.n_out + o] = ws[A.o_out + o];
}

// one-sided differences: item = point * (n_in + 1) + (dir + 1); dir = -1 is the nominal evaluation.
// Writes Y (nominal) and the RAW perturbed outputs into J[j][o][dir]; k_fd1_finalize turns them into differences.
__global__ void k_wpoints_fd1(DevTables T, WarpAux A, DevLayout L, FdSettings fd, const double* __restrict__ x, double* __restrict__ Y,
                              double* __restrict__ J, int p0, int np) {
    extern __shared__ double smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long item = (long)blockIdx.x * (blockDim.x >> 5) + warp;
    const int per = L.n_in + 1;
    if (item >= (long)np * per) return;
    double* ws = smem + (size_t)warp * A.ws_doubles;
    const int j = p0 + (int)(item / per), dir = (int)(item % per) - 1;
    warp_gather_inputs(L, x, j, ws + A.o_in, lane);
    if (dir >= 0 && lane == 0) ws[A.o_in + dir] += (fd.scheme == MOCOB200_FD_FORWARD) ? fd.h : -fd.h;
    __syncwarp();
    warp_eval_point(T, A, ws, 0xffffffffu);
    if (dir < 0) {
        for (int o = lane; o < L.n_out; o += 32) Y[(size_t)j * L.n_out + o] = ws[A.o_out + o];
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

// central differences with CasADi's step refinement; item = (point, dir, callback group). Needs Y (nominal).
__global__ void k_wpoints_central(DevTables T, WarpAux A, DevLayout L, FdSettings fd, const double* __restrict__ x,
                                  const double* __restrict__ Y, double* __restrict__ J, int p0, int np) {
    extern __shared__ double smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long item = (long)blockIdx.x * (blockDim.x >> 5) + warp;
    const int ngroups = 1 + L.n_integrand;
    const long per = (long)L.n_in * ngroups;
    if (item >= (long)np * per) return;
    double* ws = smem + (size_t)warp * A.ws_doubles;
    const int j = p0 + (int)(item / per);
    const int rem = (int)(item % per);
    const int dir = rem / ngroups, gidx = rem % ngroups;
    const int lo = gidx == 0 ? 0 : L.n_dyn_out + gidx - 1;
    const int hi = gidx == 0 ? L.n_dyn_out : lo + 1;
    if (hi == lo) return;
    const unsigned mask = 1u << gidx;
    warp_gather_inputs(L, x, j, ws + A.o_in, lane);
    const double x0 = ws[A.o_in + dir];
    const double* y0 = Y + (size_t)j * L.n_out;
    double* yb = ws + A.o_out2;
    double* yf = ws + A.o_out;
    double* Jp = J + (size_t)j * L.n_out * L.n_in;
    double h = fd.h;
    for (int iter = 0;; ++iter) {
        if (lane == 0) ws[A.o_in + dir] = x0 - h;
        __syncwarp();
        warp_eval_point(T, A, ws, mask);
        for (int o = lo + lane; o < hi; o += 32) yb[o] = yf[o];
        if (lane == 0) ws[A.o_in + dir] = x0 + h;
        __syncwarp();
        warp_eval_point(T, A, ws, mask);
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
static int g_wpb = 4; // warps per block
static size_t warpSmemBytes(const WarpAux& A) { return (size_t)g_wpb * A.ws_doubles * sizeof(double); }

bool warp_kernels_configure(const WarpAux& A, std::string& err) {
    // choose the number of warps per block so that a block's scratch fits and several blocks share an SM
    g_wpb = 4;
    while (g_wpb > 1 && warpSmemBytes(A) > 96 * 1024) g_wpb >>= 1;
    const size_t bytes = warpSmemBytes(A);
    if (bytes > 227 * 1024) { err = "per-warp scratch exceeds shared memory"; return false; }
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_wpoints_nominal, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_fd1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(k_wpoints_central, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)) != cudaSuccess) {
        err = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e);
        return false;
    }
    return true;
}

void launch_wpoints_nominal(const DevTables& T, const WarpAux& A, const DevLayout& L, const double* d_x, double* d_Y, int p0, int np,
                            cudaStream_t s) {
    if (np <= 0) return;
    k_wpoints_nominal<<<(np + g_wpb - 1) / g_wpb, g_wpb * 32, warpSmemBytes(A), s>>>(T, A, L, d_x, d_Y, p0, np);
}
// returns the number of kernels launched
int launch_wpoints_fd(const DevTables& T, const WarpAux& A, const DevLayout& L, const FdSettings& fd, const double* d_x, double* d_Y,
                      double* d_J, int p0, int np, cudaStream_t s) {
    if (np <= 0) return 0;
    if (fd.scheme != MOCOB200_FD_CENTRAL) {
        const long items = (long)np * (L.n_in + 1);
        k_wpoints_fd1<<<(unsigned)((items + g_wpb - 1) / g_wpb), g_wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, p0, np);
        const long n = (long)np * L.n_out * L.n_in;
        k_fd1_finalize<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(L, fd, d_Y, d_J, p0, np);
        return 2;
    }
    launch_wpoints_nominal(T, A, L, d_x, d_Y, p0, np, s);
    const long items = (long)np * L.n_in * (1 + L.n_integrand);
    k_wpoints_central<<<(unsigned)((items + g_wpb - 1) / g_wpb), g_wpb * 32, warpSmemBytes(A), s>>>(T, A, L, fd, d_x, d_Y, d_J, p0, np);
    return 2;
}

} // namespace mocob200
