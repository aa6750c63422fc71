// Kernels of the evaluator: point functions over the grid, their finite-difference blocks,
// endpoint functions, and the assembly of g / jac_g (CSC) / grad_f.
#pragma once
#include <cstdlib>
#include <string>

#include "point_eval.cuh"
#include "point_warp.cuh"

namespace mocob200 {

// Development A-B switches (MOCOB200_FD_LOCAL, _KERNEL, _FLAT_PAIR, ...: scripts/README.md) are honoured only when
// MOCOB200_DEVELOPER_SWITCHES=1 is set as well (tests and measurement scripts do); a production process is not steered by
// stray environment variables. Settings a solver needs have ABI setters (mocob200_set_fd, _set_min_points_per_rank, ...).
inline const char* devEnv(const char* name) {
    static const bool on = [] { const char* e = std::getenv("MOCOB200_DEVELOPER_SWITCHES"); return e && e[0] == '1'; }();
    return on ? std::getenv(name) : nullptr;
}

#define MOCOB200_MAX_IN 240
#define MOCOB200_MAX_OUT 128
#define MOCOB200_MAX_EOUT 128 /* outputs of one endpoint goal */
static_assert(SizeHuge::NIN == MOCOB200_MAX_IN, "the largest size class takes every point function the library accepts");

struct FdSettings {
    int scheme;    // MOCOB200_FD_*
    double h;      // initial step
    int h_iter;    // step-size refinement iterations (central)
    double u_aim;  // target truncation/round-off ratio
    double reltol, abstol;
};

enum { ROW_ENDPOINT_D = 0, ROW_RESIDUAL_D = 1, ROW_DEFECT_TRAP_D = 2, ROW_DEFECT_HERMITE_D = 3, ROW_DEFECT_SIMPSON_D = 4, ROW_INTERP_D = 5, ROW_PATH_D = 6,
       ROW_KINEMATIC_D = 7 };
enum { COL_T0_D = 0, COL_TF_D = 1, COL_STATE_D = 2, COL_CONTROL_D = 3, COL_DERIV_D = 4, COL_MULT_D = 5, COL_SLACK_D = 6 };
struct RowMetaD { int kind, a, b, pad; };
struct ColMetaD { int kind, point, k, in; };

struct DevLayout {
    int scheme, implicit, nq, nz, ns, nc, nd, n_in, n_auxres, n_integrand, n_dyn_out, n_out;
    int prescribed, nqs; // prescribed kinematics; coordinate value states in the NLP (nq, or 0 when prescribed)
    int n_mesh, n_grid, n_intervals, n_var, n_con, nnz;
    int off_states, off_controls, off_derivs;
    int nm, nsl, n_kc, off_mult, off_slacks;   // kinematic constraints: multipliers per point, slacks per midpoint, rows per mesh point
    int minimize_mult; double mult_weight;
    int n_endpoint_rows, n_resid_rows_per_point, n_defect_rows, n_interp_rows, n_ein, n_cost_terms;
    int minimize_accel, minimize_auxderiv;
    double accel_weight, auxderiv_weight;
    const double* mesh;  // [n_mesh]
    const double* grid;  // [n_grid]
    const double* quad;  // [n_grid]
    const int* row_resid; const int* row_defect; const int* row_interp;
    const int* goal_row; const int* goal_integrand; const int* goal_eoff; // offset of goal g's block in E / G
    const RowMetaD* rows; const ColMetaD* cols;
    const int* colptr; const int* rowidx; const int* nnzcol;
    int n_eout_total; // total outputs over all goals
    int n_path_rows;  // path-constraint equations per mesh point (0: no path-constraint kernels are launched)
    const double* Jt; // sharded evaluation: the time column of every grid point's FD block, gathered ([n_grid][n_out]); else null
    int small_size;   // the model fits the SizeSmall class of the thread-per-evaluation kernels (point_eval.cuh)
};

void launch_points_nominal(const DevTables& T, const DevLayout& L, const double* d_x, double* d_Y, int p0, int np, cudaStream_t s);
// out[j][m] = (path length, lengthening speed) of muscle m at grid point j
void launch_paths(const DevTables& T, const DevLayout& L, const double* d_x, double* d_out, int p0, int np, cudaStream_t s);
// d_dirs (device, may be null = all inputs): the FD directions this launch differences
void launch_points_fd(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* d_x, double* d_Y, double* d_J,
                      int p0, int np, cudaStream_t s, const int* d_dirs = nullptr, int ndirs = 0);
// one thread per (grid point, listed direction); Y must already hold the nominal outputs
void launch_points_fd_flat(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* d_x, const double* d_Y, double* d_J,
                           int p0, int np, cudaStream_t s, const int* d_dirs, int ndirs);
// the pose-reusing directions (WarpAux::p1_dirs: speeds, implicit-mode accelerations), one thread per (point, direction[, side]);
// d_Y and the blobs in d_cfg come from launch_wpoints_nominal (kernels_vel.cu)
void launch_points_fd_vel(const DevTables& T, const WarpAux& A, const DevLayout& L, const FdSettings& fd, const double* d_x, const double* d_Y,
                          double* d_J, const double* d_cfg, int p0, int np, cudaStream_t s);
// the columns of the "local" inputs (point_warp_local.cuh) for callers that difference the other inputs with
// launch_points_fd_flat: the local-column kernel on stream s; d_Y and the blobs in d_cfg come from launch_wpoints_nominal
int launch_wlocal(const DevTables& T, const WarpAux& A, const DevLayout& L, const FdSettings& fd, const double* d_x, double* d_Y,
                  double* d_J, double* d_cfg, int p0, int np, cudaStream_t s);
// integrals[n_integrand], dIdt0[n_integrand], dIdtf[n_integrand] (derivative parts need J)
void launch_integrals(const DevLayout& L, const double* d_x, const double* d_Y, const double* d_J, double* d_I, bool withJ, cudaStream_t s);
// G[n_eout_total] values and E[n_eout_total][n_ein] FD blocks of every goal's endpoint function
// which: 0 = every goal, 1 = only the goals without an integrand (they read x alone), 2 = only the goals with one
void launch_endpoint(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* d_x, const double* d_I, double* d_G,
                     double* d_E, bool withFd, cudaStream_t s, int which = 0);
void launch_assemble_g(const DevTables& T, const DevLayout& L, const double* d_x, const double* d_Y, const double* d_G, double* d_g,
                       cudaStream_t s);
void launch_assemble_f(const DevTables& T, const DevLayout& L, const double* d_x, const double* d_G, double* d_f /*[1+n_cost_terms]*/, cudaStream_t s);
// CSC entries [k0, k1) (k1 < 0: all of them)
void launch_assemble_jac(const DevTables& T, const DevLayout& L, const FdSettings& fd, const double* d_x, const double* d_Y,
                         const double* d_J, const double* d_I, const double* d_E, double* d_vals, cudaStream_t s, int k0 = 0, int k1 = -1);
void launch_pack_time_column(const DevLayout& L, const double* d_J, double* d_Jt, int p0, int np, cudaStream_t s);
void launch_assemble_grad(const DevTables& T, const DevLayout& L, const double* d_x, const double* d_Y, const double* d_J, const double* d_I,
                          const double* d_E, double* d_grad, cudaStream_t s);
double measure_fp64_peak_tflops(int device);

// warp-per-evaluation kernels (kernels_warp.cu)
bool warp_kernels_configure(WarpAux& A, std::string& err);
void launch_wpoints_nominal(const DevTables& T, const WarpAux& A, const DevLayout& L, const double* d_x, double* d_Y, double* d_cfg,
                            int p0, int np, cudaStream_t s);
// side streams + events for kernels that run next to each other inside one evaluation (owned by the handle)
struct SideStreams {
    cudaStream_t s[2];
    cudaEvent_t fork, join[2];
};
// after_nominal (may be null) is recorded on s once Y (the nominal point outputs) is complete
int launch_wpoints_fd(const DevTables& T, const WarpAux& A, const DevLayout& L, const FdSettings& fd, const double* d_x, double* d_Y,
                      double* d_J, double* d_cfg, int p0, int np, cudaStream_t s, cudaEvent_t after_nominal, const SideStreams* side);

} // namespace mocob200
