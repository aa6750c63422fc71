// NLP layout of the transcribed problem: variable/constraint ordering and the CSC structure of
// jac_g exactly as MocoCasADiSolver hands them to IPOPT
// (CasOCIterate.h:28-45, CasOCTranscription.h:181-313, CasOCTranscription.cpp:83-446,
//  CasOCTrapezoidal.cpp, CasOCHermiteSimpson.{h,cpp}).
#pragma once
#include <vector>

#include "../../include/mocob200.h"

namespace mocob200 {

// row kinds
enum { ROW_ENDPOINT = 0, ROW_RESIDUAL = 1, ROW_DEFECT_TRAP = 2, ROW_DEFECT_HERMITE = 3, ROW_DEFECT_SIMPSON = 4, ROW_INTERP = 5, ROW_PATH = 6,
       ROW_KINEMATIC = 7 /* a = grid point of the mesh point, b = out row of the kinematic-constraint error */ };
// column kinds
enum { COL_T0 = 0, COL_TF = 1, COL_STATE = 2, COL_CONTROL = 3, COL_DERIV = 4, COL_MULT = 5,
       COL_SLACK = 6 /* point = the interior grid point the slack belongs to, k = slack index, in = -1 (not a point-function input) */ };

struct RowMeta { int kind, a, b, pad; };  // ENDPOINT: a = goal, b = output ; RESIDUAL: a = grid point, b = out row ; DEFECT_*: a = interval, b = state ; INTERP: a = interval, b = control ; PATH: a = grid point, b = equation of the path constraint, pad = its goal record
struct ColMeta { int kind, point, k, in; }; // in = index into the point-function input vector (0 = time)

struct Layout {
    int scheme, implicit;
    int nq, nz, ns, nc, nd, n_in;
    int prescribed, nqs;                       // prescribed kinematics; coordinate value states in the NLP (nq, or 0 when prescribed)
    int nm, nsl, n_kc;                         // multipliers per grid point, slacks per interior grid point, kinematic rows per mesh point
    int enforce_derivs;                        // enforce_constraint_derivatives (and kinematic constraints exist)
    int off_mult, off_slacks;
    std::vector<int> row_kin;                  // first kinematic-constraint row of each mesh point
    int n_auxres, n_integrand, n_dyn_out, n_out;
    int n_mesh, n_grid, n_intervals;
    int n_var, n_con, nnz;
    int off_states, off_controls, off_derivs;
    int n_endpoint_rows, n_resid_rows_per_point, n_defect_rows, n_interp_rows;
    int n_cost_terms;
    int n_path_rows;                           // path-constraint equations per mesh point (all path constraints)
    std::vector<int> row_path;                 // first path-constraint row of each mesh point
    std::vector<int> goal_path_off;            // offset of path constraint g inside a mesh point's path rows (or -1)
    std::vector<double> mesh, grid, quad;      // normalised times and quadrature coefficients
    std::vector<int> row_resid, row_defect, row_interp; // first row of each block (per grid point / interval)
    std::vector<int> goal_row;                 // first endpoint-constraint row of goal g (or -1)
    std::vector<int> goal_integrand;           // integrand slot of goal g (or -1)
    std::vector<RowMeta> rows;
    std::vector<ColMeta> cols;
    std::vector<int> colptr, rowidx, nnzcol;
    // endpoint function inputs: [t0, x0(ns), c0(nc), d0(nd), tf, xf, cf, df, integral]
    int n_ein;
};

// blockMask (may be null = dense callback blocks): [n_out + n_path_rows][n_in] flags (point-function outputs, then the
// path-constraint equations), non-zero where the point function's output row
// depends on its input (the per-callback Jacobian sparsity of optim_sparsity_detection = "random",
// CasOCFunction.cpp:25-105); the structure then holds only the entries CasADi's sparsity propagation would keep.
Layout buildLayout(const mocob200_tables& T, const mocob200_transcription& R, bool withStructure,
                   const unsigned char* blockMask = nullptr);
// Copies T into (dbl, idx) with every body that hangs on a WeldJoint merged into its parent (table_opt.cpp); `out` addresses the
// copies. Returns the number of bodies removed. Coordinates, states, controls and goals are untouched, so the NLP layout of T
// stays valid for the reduced tables.
int mergeWeldedBodies(const mocob200_tables& T, std::vector<double>& dbl, std::vector<int>& idx, mocob200_tables& out);
void layoutBounds(const mocob200_tables& T, const mocob200_transcription& R, const Layout& L, double* xlo, double* xhi,
                  double* glo, double* ghi);

} // namespace mocob200
