#include "layout.h"

#include <algorithm>
#include <limits>
#include <stdexcept>

namespace mocob200 {

namespace {
inline const int* goalRec(const mocob200_tables& T, int g) { return T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE; }
}

Layout buildLayout(const mocob200_tables& T, const mocob200_transcription& R, bool withStructure, const unsigned char* blockMask) {
    Layout L{};
    L.scheme = R.scheme;
    L.implicit = R.dynamics_mode == MOCOB200_DYNAMICS_IMPLICIT;
    // prescribed kinematics (PositionMotion): no coordinate value / speed states, no acceleration variables (MocoProblemRep.cpp:546-560,
    // CasOCProblem.h:485-517); the multibody residual rows stay (one per mobility)
    L.prescribed = T.prescribed ? 1 : 0;
    if (L.prescribed && !L.implicit) throw std::runtime_error("Prescribed kinematics (PositionMotion) requires implicit dynamics mode.");
    L.nq = T.nq; L.nqs = L.prescribed ? 0 : T.nq; L.nz = T.nz; L.ns = 2 * L.nqs + T.nz; L.nc = T.nc;
    L.nd = (L.implicit && !L.prescribed ? T.nq : 0) + T.nderiv_aux;
    const bool hsScheme = R.scheme == MOCOB200_SCHEME_HERMITE_SIMPSON;
    // kinematic constraints (CasOCProblem.h:473-524): one multiplier per holonomic equation; with enforce_constraint_derivatives
    // the velocity- and acceleration-level equations are rows too and every Hermite-Simpson midpoint gets one slack per equation
    L.nm = T.ncoupler;
    L.enforce_derivs = (L.nm > 0 && R.enforce_constraint_derivatives && !L.prescribed) ? 1 : 0;
    L.n_kc = L.prescribed ? 0 : (L.enforce_derivs ? 3 * L.nm : L.nm); // prescribed kinematics are assumed to obey the constraints (CasOCProblem.h:511-515)
    L.nsl = (L.enforce_derivs && hsScheme) ? L.nm : 0;
    L.n_in = 1 + L.ns + L.nc + L.nm + L.nd;
    L.n_auxres = T.nderiv_aux;
    L.goal_integrand.assign(T.ngoal, -1);
    L.goal_row.assign(T.ngoal, -1);
    L.goal_path_off.assign(T.ngoal, -1);
    L.n_integrand = 0; L.n_endpoint_rows = 0; L.n_cost_terms = 0; L.n_path_rows = 0;
    for (int g = 0; g < T.ngoal; ++g) {
        const int* gr = goalRec(T, g);
        if (gr[1] == MOCOB200_MODE_PATH_CONSTRAINT) { L.goal_path_off[g] = L.n_path_rows; L.n_path_rows += gr[2]; continue; }
        if (gr[3]) L.goal_integrand[g] = L.n_integrand++;
        if (gr[1] == MOCOB200_MODE_ENDPOINT_CONSTRAINT) { L.goal_row[g] = L.n_endpoint_rows; L.n_endpoint_rows += gr[2]; }
        else L.n_cost_terms++;
    }
    // endpoint constraints come first in goal order among themselves (createEndpointConstraintNames)
    if (L.nm && R.minimize_lagrange_multipliers) L.n_cost_terms++;
    if (L.implicit && R.minimize_accelerations) L.n_cost_terms++;
    if (L.n_auxres && R.minimize_auxiliary_derivatives) L.n_cost_terms++;
    L.n_dyn_out = L.nq + L.nz + L.n_auxres + L.n_kc;
    L.n_out = L.n_dyn_out + L.n_integrand;
    L.n_mesh = R.n_mesh;
    if (L.n_mesh < 2) throw std::runtime_error("need at least one mesh interval");
    L.n_intervals = L.n_mesh - 1;
    L.mesh.resize(L.n_mesh);
    for (int i = 0; i < L.n_mesh; ++i) L.mesh[i] = R.mesh ? R.mesh[i] : i / (double)L.n_intervals;
    const bool hs = L.scheme == MOCOB200_SCHEME_HERMITE_SIMPSON;
    L.n_grid = hs ? 2 * L.n_mesh - 1 : L.n_mesh;
    L.grid.resize(L.n_grid);
    L.quad.assign(L.n_grid, 0.0);
    for (int i = 0; i < L.n_grid; ++i) {
        if (!hs) L.grid[i] = L.mesh[i];
        else L.grid[i] = (i % 2 == 0) ? L.mesh[i / 2] : .5 * (L.mesh[i / 2] + L.mesh[i / 2 + 1]);
    }
    for (int m = 0; m < L.n_intervals; ++m) {
        const double dm = L.mesh[m + 1] - L.mesh[m];
        if (hs) {
            L.quad[2 * m] += (1.0 / 6.0) * dm;
            L.quad[2 * m + 1] += (2.0 / 3.0) * dm;
            L.quad[2 * m + 2] += (1.0 / 6.0) * dm;
        } else {
            L.quad[m] += 0.5 * dm;
            L.quad[m + 1] += 0.5 * dm;
        }
    }
    // variables: t0, tf, vec(states), vec(controls), vec(multipliers), vec(slacks at the interior points), vec(derivatives),
    // [parameters: none]  (CasOCIterate.h:28-45 enum order, CasOCTranscription.h:192-199)
    const int nInterior = L.n_grid - L.n_mesh;
    L.off_states = 2;
    L.off_controls = L.off_states + L.ns * L.n_grid;
    L.off_mult = L.off_controls + L.nc * L.n_grid;
    L.off_slacks = L.off_mult + L.nm * L.n_grid;
    L.off_derivs = L.off_slacks + L.nsl * nInterior;
    L.n_var = L.off_derivs + L.nd * L.n_grid;
    L.n_ein = 2 + 2 * (L.ns + L.nc + L.nm + L.nd) + 1;
    // constraints
    L.n_resid_rows_per_point = (L.implicit ? L.nq : 0) + L.n_auxres;
    L.n_defect_rows = (hs ? 2 : 1) * L.ns;
    L.n_interp_rows = (hs && R.interpolate_control_midpoints) ? L.nc : 0;
    L.row_resid.assign(L.n_grid, -1);
    L.row_defect.assign(L.n_intervals, -1);
    L.row_interp.assign(L.n_intervals, -1);
    L.row_path.assign(L.n_mesh, -1);
    L.row_kin.assign(L.n_mesh, -1);
    int row = L.n_endpoint_rows;
    const int ppi = hs ? 2 : 1; // grid points whose residuals are grouped with interval m
    for (int m = 0; m < L.n_intervals; ++m) {
        L.row_kin[m] = row; row += L.n_kc;         // kinematic constraints of mesh point m, then its path constraints (CasOCTranscription.h:279-283)
        L.row_path[m] = row; row += L.n_path_rows;
        for (int k = 0; k < ppi; ++k) { L.row_resid[ppi * m + k] = row; row += L.n_resid_rows_per_point; }
        L.row_defect[m] = row; row += L.n_defect_rows;
        L.row_interp[m] = row; row += L.n_interp_rows;
    }
    L.row_kin[L.n_mesh - 1] = row; row += L.n_kc;
    L.row_path[L.n_mesh - 1] = row; row += L.n_path_rows;
    L.row_resid[L.n_grid - 1] = row; row += L.n_resid_rows_per_point;
    L.n_con = row;
    L.rows.assign(L.n_con, RowMeta{0, 0, 0, 0});
    for (int g = 0; g < T.ngoal; ++g)
        if (L.goal_row[g] >= 0)
            for (int o = 0; o < goalRec(T, g)[2]; ++o) L.rows[L.goal_row[g] + o] = RowMeta{ROW_ENDPOINT, g, o, 0};
    for (int g = 0; g < T.ngoal; ++g)
        if (L.goal_path_off[g] >= 0)
            for (int im = 0; im < L.n_mesh; ++im)
                for (int e = 0; e < goalRec(T, g)[2]; ++e)
                    L.rows[L.row_path[im] + L.goal_path_off[g] + e] = RowMeta{ROW_PATH, hs ? 2 * im : im, e, g};
    for (int im = 0; im < L.n_mesh; ++im)
        for (int e = 0; e < L.n_kc; ++e) L.rows[L.row_kin[im] + e] = RowMeta{ROW_KINEMATIC, hs ? 2 * im : im, L.nq + L.nz + L.n_auxres + e, 0};
    for (int j = 0; j < L.n_grid; ++j)
        for (int o = 0; o < L.n_resid_rows_per_point; ++o) {
            // out rows: implicit multibody residual [0,nq) then aux residuals at [nq+nz, ...)
            const int outRow = (L.implicit && o < L.nq) ? o : L.nq + L.nz + (o - (L.implicit ? L.nq : 0));
            L.rows[L.row_resid[j] + o] = RowMeta{ROW_RESIDUAL, j, outRow, 0};
        }
    for (int m = 0; m < L.n_intervals; ++m) {
        for (int s = 0; s < L.ns; ++s) {
            if (hs) {
                L.rows[L.row_defect[m] + s] = RowMeta{ROW_DEFECT_HERMITE, m, s, 0};
                L.rows[L.row_defect[m] + L.ns + s] = RowMeta{ROW_DEFECT_SIMPSON, m, s, 0};
            } else
                L.rows[L.row_defect[m] + s] = RowMeta{ROW_DEFECT_TRAP, m, s, 0};
        }
        for (int c = 0; c < L.n_interp_rows; ++c) L.rows[L.row_interp[m] + c] = RowMeta{ROW_INTERP, m, c, 0};
    }
    L.cols.assign(L.n_var, ColMeta{0, 0, 0, 0});
    L.cols[0] = ColMeta{COL_T0, -1, 0, 0};
    L.cols[1] = ColMeta{COL_TF, -1, 0, 0};
    for (int j = 0; j < L.n_grid; ++j) {
        for (int k = 0; k < L.ns; ++k) L.cols[L.off_states + j * L.ns + k] = ColMeta{COL_STATE, j, k, 1 + k};
        for (int k = 0; k < L.nc; ++k) L.cols[L.off_controls + j * L.nc + k] = ColMeta{COL_CONTROL, j, k, 1 + L.ns + k};
        for (int k = 0; k < L.nm; ++k) L.cols[L.off_mult + j * L.nm + k] = ColMeta{COL_MULT, j, k, 1 + L.ns + L.nc + k};
        for (int k = 0; k < L.nd; ++k) L.cols[L.off_derivs + j * L.nd + k] = ColMeta{COL_DERIV, j, k, 1 + L.ns + L.nc + L.nm + k};
    }
    for (int i = 0; i < nInterior; ++i)
        for (int k = 0; k < L.nsl; ++k) L.cols[L.off_slacks + i * L.nsl + k] = ColMeta{COL_SLACK, 2 * i + 1, k, -1};
    if (!withStructure) { L.nnz = -1; return L; }

    // ---- CSC structure with CasADi's dense point-function blocks (optim_sparsity_detection "none")
    bool anyGoalIntegrandEC = false;
    std::vector<char> ecHasInt(T.ngoal, 0);
    for (int g = 0; g < T.ngoal; ++g)
        if (L.goal_row[g] >= 0 && goalRec(T, g)[3]) { ecHasInt[g] = 1; anyGoalIntegrandEC = true; }
    (void)anyGoalIntegrandEC;
    // does output row `o` of the point function depend on its input `in`?
    auto maskAt = [&](int o, int in) -> bool { return !blockMask || blockMask[(size_t)o * L.n_in + in] != 0; };
    // output row of residual row o of a grid point (same mapping as RowMeta above)
    auto residOutRow = [&](int o) { return (L.implicit && o < L.nq) ? o : L.nq + L.nz + (o - (L.implicit ? L.nq : 0)); };
    // does xdot_s (state s) at a point depend on input `in` of the same point?
    auto xdotDepends = [&](int s, const ColMeta& c) -> bool {
        if (s < L.nqs && L.nsl > 0 && hs && c.point % 2 == 1 && c.kind == COL_STATE && c.k < 2 * L.nqs)
            return true; // qdot_mid = u_mid + velocity correction(multibody states of the midpoint, slacks): dense callback block
        if (s < L.nqs) return c.kind == COL_STATE && c.k == L.nqs + s; // qdot = u
        if (s < 2 * L.nqs && L.implicit) return c.kind == COL_DERIV && c.k == s - L.nqs; // udot = w
        return maskAt(L.nq + (s - 2 * L.nqs), c.in); // zdot rows of the callback block: dense unless a detected mask is given
    };
    L.colptr.assign(L.n_var + 1, 0);
    L.rowidx.clear();
    L.nnzcol.clear();
    auto emit = [&](int col, int r) { L.rowidx.push_back(r); L.nnzcol.push_back(col); };
    for (int col = 0; col < L.n_var; ++col) {
        const ColMeta c = L.cols[col];
        L.colptr[col] = (int)L.rowidx.size();
        if (c.kind == COL_T0 || c.kind == COL_TF) {
            for (int r = 0; r < L.n_endpoint_rows; ++r) emit(col, r);
            for (int m = 0; m < L.n_intervals; ++m) {
                for (int e = 0; e < L.n_kc; ++e)
                    if (maskAt(L.nq + L.nz + L.n_auxres + e, 0)) emit(col, L.row_kin[m] + e);
                for (int e = 0; e < L.n_path_rows; ++e)
                    if (maskAt(L.n_out + e, 0)) emit(col, L.row_path[m] + e);
                for (int k = 0; k < ppi; ++k)
                    for (int o = 0; o < L.n_resid_rows_per_point; ++o)
                        if (maskAt(residOutRow(o), 0)) emit(col, L.row_resid[ppi * m + k] + o); // through the time input
                for (int o = 0; o < L.n_defect_rows; ++o) emit(col, L.row_defect[m] + o);      // through h, always
            }
            for (int e = 0; e < L.n_kc; ++e)
                if (maskAt(L.nq + L.nz + L.n_auxres + e, 0)) emit(col, L.row_kin[L.n_mesh - 1] + e);
            for (int e = 0; e < L.n_path_rows; ++e)
                if (maskAt(L.n_out + e, 0)) emit(col, L.row_path[L.n_mesh - 1] + e);
            for (int o = 0; o < L.n_resid_rows_per_point; ++o)
                if (maskAt(residOutRow(o), 0)) emit(col, L.row_resid[L.n_grid - 1] + o);
            continue;
        }
        if (c.kind == COL_SLACK) {
            // a slack enters through the velocity correction of its midpoint only: qdot_mid = u_mid + G^T gamma
            // (CasOCTranscription.cpp:315-332), i.e. the Simpson defects of the coordinates of its interval (the Hermite defect
            // has no xdot_mid); the VelocityCorrection callback block is dense (CasOCFunction.cpp:250-291)
            const int m = c.point / 2;
            for (int s = 0; s < L.nq; ++s) emit(col, L.row_defect[m] + L.ns + s);
            continue;
        }
        const int j = c.point;
        for (int g = 0; g < T.ngoal; ++g) {
            if (L.goal_row[g] < 0) continue;
            const bool viaIntegral = ecHasInt[g] && maskAt(L.n_dyn_out + L.goal_integrand[g], c.in);
            if (j == 0 || j == L.n_grid - 1 || viaIntegral)
                for (int o = 0; o < goalRec(T, g)[2]; ++o) emit(col, L.goal_row[g] + o);
        }
        // intervals that contain grid point j, ascending
        int ms[2], roles[2], nm = 0; // role: 0 = i (left), 1 = mid, 2 = ip (right)
        if (hs) {
            if (j % 2 == 1) { ms[nm] = j / 2; roles[nm++] = 1; }
            else {
                if (j / 2 - 1 >= 0) { ms[nm] = j / 2 - 1; roles[nm++] = 2; }
                if (j / 2 < L.n_intervals) { ms[nm] = j / 2; roles[nm++] = 0; }
            }
        } else {
            if (j - 1 >= 0) { ms[nm] = j - 1; roles[nm++] = 2; }
            if (j < L.n_intervals) { ms[nm] = j; roles[nm++] = 0; }
        }
        // the path-constraint callbacks of mesh point j: after the interval that ends at j, before the one that starts there
        bool pathDone = !(hs ? j % 2 == 0 : true) || (L.n_path_rows == 0 && L.n_kc == 0);
        auto emitPath = [&]() {
            if (pathDone) return;
            pathDone = true;
            for (int e = 0; e < L.n_kc; ++e)
                if (maskAt(L.nq + L.nz + L.n_auxres + e, c.in)) emit(col, L.row_kin[hs ? j / 2 : j] + e);
            for (int e = 0; e < L.n_path_rows; ++e)
                if (maskAt(L.n_out + e, c.in)) emit(col, L.row_path[hs ? j / 2 : j] + e);
        };
        for (int t = 0; t < nm; ++t) {
            const int m = ms[t], role = roles[t];
            if (role != 2) emitPath();
            if (role != 2) // residual rows of j sit in the block of the interval that starts at / contains j
                for (int o = 0; o < L.n_resid_rows_per_point; ++o)
                    if (maskAt(residOutRow(o), c.in)) emit(col, L.row_resid[j] + o);
            for (int s = 0; s < L.ns; ++s) {
                const bool dep = xdotDepends(s, c);
                const bool direct = c.kind == COL_STATE && c.k == s;
                if (hs) {
                    // Hermite: x_mid - (x_ip + x_i)/2 - h/8 (xdot_i - xdot_ip) ; no xdot_mid
                    if (direct || (dep && role != 1)) emit(col, L.row_defect[m] + s);
                } else {
                    if (direct || dep) emit(col, L.row_defect[m] + s);
                }
            }
            if (hs)
                for (int s = 0; s < L.ns; ++s) {
                    // Simpson: x_ip - x_i - h/6 (xdot_ip + 4 xdot_mid + xdot_i) ; no x_mid
                    const bool dep = xdotDepends(s, c);
                    const bool direct = c.kind == COL_STATE && c.k == s && role != 1;
                    if (direct || dep) emit(col, L.row_defect[m] + L.ns + s);
                }
            if (L.n_interp_rows && c.kind == COL_CONTROL) emit(col, L.row_interp[m] + c.k);
        }
        emitPath();
        if (j == L.n_grid - 1)
            for (int o = 0; o < L.n_resid_rows_per_point; ++o)
                if (maskAt(residOutRow(o), c.in)) emit(col, L.row_resid[j] + o);
    }
    L.colptr[L.n_var] = (int)L.rowidx.size();
    L.nnz = (int)L.rowidx.size();
    return L;
}

// Variable and constraint bounds (CasOCTranscription.cpp:175-250, 273-309, 429-443).
void layoutBounds(const mocob200_tables& T, const mocob200_transcription& R, const Layout& L, double* xlo, double* xhi,
                  double* glo, double* ghi) {
    const double* tb = T.dbl + T.o_time_bounds;
    const double* vb = T.dbl + T.o_bounds;
    if (xlo && xhi) {
        xlo[0] = tb[0]; xhi[0] = tb[1]; xlo[1] = tb[2]; xhi[1] = tb[3];
        for (int j = 0; j < L.n_grid; ++j) {
            const int which = j == 0 ? 2 : (j == L.n_grid - 1 ? 4 : 0);
            for (int k = 0; k < L.ns; ++k) {
                xlo[L.off_states + j * L.ns + k] = vb[6 * k + which];
                xhi[L.off_states + j * L.ns + k] = vb[6 * k + which + 1];
            }
            for (int k = 0; k < L.nc; ++k) {
                xlo[L.off_controls + j * L.nc + k] = vb[6 * (L.ns + k) + which];
                xhi[L.off_controls + j * L.nc + k] = vb[6 * (L.ns + k) + which + 1];
            }
            for (int k = 0; k < L.nm; ++k) { // MocoPhase multiplier_bounds, the same at every point (MocoProblemRep.cpp:160-200)
                xlo[L.off_mult + j * L.nm + k] = R.multiplier_bounds[0];
                xhi[L.off_mult + j * L.nm + k] = R.multiplier_bounds[1];
            }
            for (int k = 0; k < L.nd; ++k) {
                const bool accel = L.implicit && !L.prescribed && k < L.nq;
                xlo[L.off_derivs + j * L.nd + k] = accel ? R.accel_bounds[0] : R.auxderiv_bounds[0];
                xhi[L.off_derivs + j * L.nd + k] = accel ? R.accel_bounds[1] : R.auxderiv_bounds[1];
            }
        }
    }
    if (xlo && xhi)
        for (int i = 0; i < (L.n_grid - L.n_mesh) * L.nsl; ++i) { // velocity_correction_bounds (MocoCasOCProblem.cpp:183-202)
            xlo[L.off_slacks + i] = R.velocity_correction_bounds[0];
            xhi[L.off_slacks + i] = R.velocity_correction_bounds[1];
        }
    if (glo && ghi) {
        for (int r = 0; r < L.n_con; ++r) { glo[r] = 0; ghi[r] = 0; }
        for (int g = 0; g < T.ngoal; ++g) {
            const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
            const double* dp = T.dbl + gr[5];
            if (L.goal_path_off[g] >= 0) // MocoConstraintInfo bounds repeated over the mesh points (CasOCTranscription.cpp:429-432)
                for (int im = 0; im < L.n_mesh; ++im)
                    for (int e = 0; e < gr[2]; ++e) {
                        glo[L.row_path[im] + L.goal_path_off[g] + e] = dp[3 + 2 * e];
                        ghi[L.row_path[im] + L.goal_path_off[g] + e] = dp[4 + 2 * e];
                    }
            if (L.goal_row[g] < 0) continue;
            for (int o = 0; o < gr[2]; ++o) { glo[L.goal_row[g] + o] = dp[1]; ghi[L.goal_row[g] + o] = dp[2]; }
        }
    }
}

} // namespace mocob200
