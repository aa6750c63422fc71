// Thread-per-evaluation FD kernels for the inputs that leave the coordinates untouched (speeds, implicit-mode accelerations):
// one thread per (grid point, direction[, side]); the dynamics callback group goes through eval_point_vel (point_vel.cuh),
// which reads the nominal poses / path geometry / mass-matrix factor of the grid point from the blob saved by
// k_wpoints_nominal instead of recomputing them; the integrand groups through the whole-function eval_point.
#include "fd_group.cuh"
#include "point_vel.cuh"

namespace mocob200 {

template <int MINB, class SZ>
__global__ void __launch_bounds__(64, MINB)
k_points_fd_vel_pair(DevTables T, WarpAux A, DevLayout L, FdSettings fd, const double* __restrict__ x, const double* __restrict__ Y,
                     double* __restrict__ J, const double* __restrict__ cfg, int p0, int np, const int* __restrict__ dirs, int ndirs) {
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
    const double* blobz = cfg + (size_t)j * (A.blob_end - A.blob_begin) - A.blob_begin;
    fd_group_central_pair([&](const double* xin, double* yout) { eval_point_vel<SZ>(T, A, blobz, xin, yout); }, fd, in, a, y0, 0, L.n_dyn_out,
                          yo, Jc, side);
    for (int gidx = 1; gidx <= L.n_integrand; ++gidx) {
        const int lo = L.n_dyn_out + gidx - 1;
        fd_group_central_pair([&](const double* xin, double* yout) { eval_point<SZ>(T, xin, yout, 1u << gidx); }, fd, in, a, y0, lo, lo + 1, yo, Jc,
                              side);
    }
    if (!valid || side) return;
    double* Jp = J + (size_t)j * L.n_out * L.n_in;
    for (int o = 0; o < L.n_out; ++o) Jp[(size_t)o * L.n_in + a] = Jc[o];
}

template <int MINB, class SZ>
__global__ void __launch_bounds__(64, MINB)
k_points_fd_vel(DevTables T, WarpAux A, DevLayout L, FdSettings fd, const double* __restrict__ x, const double* __restrict__ Y,
                double* __restrict__ J, const double* __restrict__ cfg, int p0, int np, const int* __restrict__ dirs, int ndirs) {
    const long gid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (long)np * ndirs) return;
    const int j = p0 + (int)(gid / ndirs);
    const int a = dirs[gid % ndirs];
    double in[SZ::NIN];
    double ya[SZ::NOUT], yb[SZ::NOUT], Jc[SZ::NOUT];
    gather_inputs(L, x, j, in);
    const double* y0 = Y + (size_t)j * L.n_out;
    const double* blobz = cfg + (size_t)j * (A.blob_end - A.blob_begin) - A.blob_begin;
    fd_group([&](const double* xin, double* yout) { eval_point_vel<SZ>(T, A, blobz, xin, yout); }, fd, in, a, y0, 0, L.n_dyn_out, yb, ya, Jc);
    for (int gidx = 1; gidx <= L.n_integrand; ++gidx) {
        const int lo = L.n_dyn_out + gidx - 1;
        fd_group([&](const double* xin, double* yout) { eval_point<SZ>(T, xin, yout, 1u << gidx); }, fd, in, a, y0, lo, lo + 1, yb, ya, Jc);
    }
    double* Jp = J + (size_t)j * L.n_out * L.n_in;
    for (int o = 0; o < L.n_out; ++o) Jp[(size_t)o * L.n_in + a] = Jc[o];
}

// d_Y and the blobs in d_cfg come from launch_wpoints_nominal; dirs = the pose-reusing directions (WarpAux::p1_dirs)
void launch_points_fd_vel(const DevTables& T, const WarpAux& A, const DevLayout& L, const FdSettings& fd, const double* d_x, const double* d_Y,
                          double* d_J, const double* d_cfg, int p0, int np, cudaStream_t s) {
    const long n = (long)np * A.n_p1;
    if (n <= 0) return;
    const bool central = fd.scheme == MOCOB200_FD_CENTRAL;
    if (central) {
        const unsigned nb = (unsigned)((2 * n + 63) / 64);
        if (L.small_size) k_points_fd_vel_pair<8, SizeSmall><<<nb, 64, 0, s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np, A.p1_dirs, A.n_p1);
        else k_points_fd_vel_pair<8, SizeLarge><<<nb, 64, 0, s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np, A.p1_dirs, A.n_p1);
    } else {
        const unsigned nb = (unsigned)((n + 63) / 64);
        if (L.small_size) k_points_fd_vel<8, SizeSmall><<<nb, 64, 0, s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np, A.p1_dirs, A.n_p1);
        else k_points_fd_vel<8, SizeLarge><<<nb, 64, 0, s>>>(T, A, L, fd, d_x, d_Y, d_J, d_cfg, p0, np, A.p1_dirs, A.n_p1);
    }
}

} // namespace mocob200
