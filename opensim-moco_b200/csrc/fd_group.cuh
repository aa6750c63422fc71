// Pieces shared by the thread-per-evaluation finite-difference kernels (kernels.cu, kernels_vel.cu).
#pragma once
#include "kernels.cuh"

namespace mocob200 {

// ---------------------------------------------------------------------------------------------
// gather the inputs of grid point j from the NLP vector x (CasOCTranscription.h:192-199 layout)
__device__ __forceinline__ void gather_inputs(const DevLayout& L, const double* __restrict__ x, int j, double* in) {
    const double t0 = x[0], tf = x[1];
    in[0] = t0 + (tf - t0) * L.grid[j]; // createTimes (CasOCTranscription.h: m_times)
    const double* xs = x + L.off_states + (size_t)j * L.ns;
    for (int k = 0; k < L.ns; ++k) in[1 + k] = xs[k];
    const double* xc = x + L.off_controls + (size_t)j * L.nc;
    for (int k = 0; k < L.nc; ++k) in[1 + L.ns + k] = xc[k];
    const double* xm = x + L.off_mult + (size_t)j * L.nm;
    for (int k = 0; k < L.nm; ++k) in[1 + L.ns + L.nc + k] = xm[k];
    const double* xd = x + L.off_derivs + (size_t)j * L.nd;
    for (int k = 0; k < L.nd; ++k) in[1 + L.ns + L.nc + L.nm + k] = xd[k];
}

// One finite-difference column of one callback group, following CasADi's FiniteDiff
// (casadi/core/finite_differences.cpp of the pinned fork; restated from its documented algorithm):
//   forward : (f(x+h v) - f(x)) / h                       h = sqrt(abstol)
//   backward: (f(x) - f(x-h v)) / h
//   central : (f(x+h v) - f(x-h v)) / 2h, h = abstol^(1/3); after each of h_iter passes the ratio
//             u = max_i |yf-2yc+yb| / (reltol/h max(|yf-yc|,|yc-yb|) + abstol) rescales h by
//             sqrt(u_aim / max(1,u)) (or h /= u_aim if an output was not finite).
// `lo,hi` select the output range of the group, `mask` the eval_point group bit.
template <class F>
__device__ inline void fd_group(F&& f, const FdSettings& fd, double* in, int a, const double* y0, int lo, int hi, double* yb,
                                double* yf, double* Jcol) {
    const double x0 = in[a];
    if (fd.scheme != MOCOB200_FD_CENTRAL) {
        const double h = fd.h;
        const double pert = (fd.scheme == MOCOB200_FD_FORWARD) ? h : -h;
        in[a] = x0 + pert;
        f(in, yf);
        in[a] = x0;
        for (int o = lo; o < hi; ++o) Jcol[o] = (fd.scheme == MOCOB200_FD_FORWARD) ? (yf[o] - y0[o]) / h : (y0[o] - yf[o]) / h;
        return;
    }
    double h = fd.h;
    for (int iter = 0;; ++iter) {
        in[a] = x0 - h;
        f(in, yb);
        in[a] = x0 + h;
        f(in, yf);
        in[a] = x0;
        double u = 0;
        for (int o = lo; o < hi; ++o) {
            const double vf = yf[o], vc = y0[o], vb = yb[o];
            if (!isfinite(vf) || !isfinite(vc) || !isfinite(vb)) { Jcol[o] = NAN; u = -1; continue; }
            Jcol[o] = (vf - vb) / (2 * h);
            const double et = vf - 2 * vc + vb;
            const double er = fd.reltol / h * fmax(fabs(vf - vc), fabs(vc - vb)) + fd.abstol;
            if (u >= 0) u = fmax(u, fabs(et / er));
        }
        if (iter == fd.h_iter) break;
        if (u < 0) h /= fd.u_aim;
        else h *= sqrt(fd.u_aim / fmax(1.0, u));
    }
}

// Central differences with the two sides of a direction on adjacent lanes (lane ^ 1): each lane evaluates one side per
// refinement pass, the pair exchanges the group's outputs by shuffle and both derive the same refined step — the
// arithmetic and its order are those of fd_group, the dependent chain per thread is half as long.
template <class F>
__device__ inline void fd_group_central_pair(F&& f, const FdSettings& fd, double* in, int a, const double* y0, int lo, int hi, double* yo,
                                             double* Jcol, int side) {
    const double x0 = in[a];
    double h = fd.h;
    for (int iter = 0;; ++iter) {
        in[a] = side ? x0 + h : x0 - h;
        f(in, yo);
        in[a] = x0;
        double u = 0;
        for (int o = lo; o < hi; ++o) {
            const double mine = yo[o];
            const double other = __shfl_xor_sync(0xffffffffu, mine, 1);
            const double vf = side ? mine : other, vb = side ? other : mine, vc = y0[o];
            if (!isfinite(vf) || !isfinite(vc) || !isfinite(vb)) { Jcol[o] = NAN; u = -1; continue; }
            Jcol[o] = (vf - vb) / (2 * h);
            const double et = vf - 2 * vc + vb;
            const double er = fd.reltol / h * fmax(fabs(vf - vc), fabs(vc - vb)) + fd.abstol;
            if (u >= 0) u = fmax(u, fabs(et / er));
        }
        if (iter == fd.h_iter) break;
        if (u < 0) h /= fd.u_aim;
        else h *= sqrt(fd.u_aim / fmax(1.0, u));
    }
}

} // namespace mocob200
