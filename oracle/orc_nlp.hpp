// TEST INFRASTRUCTURE ONLY — CPU oracle: the transcribed NLP (g, f, jac_g in CSC, grad_f) built from
// the point functions the way MocoCasADiSolver + CasADi build it:
//   variables  x = [t0, tf, vec(states), vec(controls), vec(derivatives)]   (CasOCIterate.h:28-45,
//                                                                          CasOCTranscription.h:192-199)
//   constraints grouped by time (CasOCTranscription.h:219-313)
//   defects    CasOCTrapezoidal.cpp:43-58, CasOCHermiteSimpson.cpp:53-105
//   objective / endpoint constraints  CasOCTranscription.cpp:448-585
//   derivatives: finite differences of every callback (CasOCFunction.h:38-44) in CasADi's FiniteDiff
//   scheme, chained through the (linear) transcription. CasADi itself is not in the reference tree:
//   step sizes / formulas are restated from its published algorithm and are PARITY UNPINNED.
// The Jacobian is accumulated ROW-WISE into a scratch row (the way forward AD through the expression
// graph would), and its structure comes from per-row dependency sets; the CUDA side instead computes
// one CSC entry per thread from closed forms, so agreement between the two is a real check.
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

#include <stdexcept>

#include "orc_eval.hpp"

namespace orc {

struct FdOptions {
    int scheme = MOCOB200_FD_CENTRAL;
    double h = 0, reltol = 2.220446049250313e-16, abstol = 2.220446049250313e-16, u_aim = 100;
    int h_iter = 0;
    void defaults(int s) {
        scheme = s;
        const double eps = 2.220446049250313e-16;
        reltol = abstol = eps; u_aim = 100;
        if (s == MOCOB200_FD_CENTRAL) { h = std::cbrt(eps); h_iter = 1; }
        else { h = std::sqrt(eps); h_iter = 0; }
    }
};

// One directional derivative of f: R^n -> R^m along unit input `a`, CasADi FiniteDiff style
// (forward / backward: one-sided with fixed h; central: step refinement from the ratio of the
// truncation estimate yf-2yc+yb to the round-off estimate).
template <class F>
void fdColumn(F&& f, const FdOptions& fd, std::vector<double>& in, int a, const double* y0, int m, double* col) {
    const double x0 = in[a];
    std::vector<double> yf(m), yb(m);
    if (fd.scheme != MOCOB200_FD_CENTRAL) {
        const bool fwd = fd.scheme == MOCOB200_FD_FORWARD;
        in[a] = fwd ? x0 + fd.h : x0 - fd.h;
        f(in.data(), yf.data());
        in[a] = x0;
        for (int o = 0; o < m; ++o) col[o] = fwd ? (yf[o] - y0[o]) / fd.h : (y0[o] - yf[o]) / fd.h;
        return;
    }
    double h = fd.h;
    for (int iter = 0;; ++iter) {
        in[a] = x0 - h; f(in.data(), yb.data());
        in[a] = x0 + h; f(in.data(), yf.data());
        in[a] = x0;
        double u = 0;
        for (int o = 0; o < m; ++o) {
            const double vf = yf[o], vc = y0[o], vb = yb[o];
            if (!std::isfinite(vf) || !std::isfinite(vc) || !std::isfinite(vb)) { col[o] = NAN; u = -1; continue; }
            col[o] = (vf - vb) / (2 * h);
            const double errTrunc = vf - 2 * vc + vb;
            const double errRound = fd.reltol / h * std::fmax(std::fabs(vf - vc), std::fabs(vc - vb)) + fd.abstol;
            if (u >= 0) u = std::fmax(u, std::fabs(errTrunc / errRound));
        }
        if (iter == fd.h_iter) break;
        if (u < 0) h /= fd.u_aim;
        else h *= std::sqrt(fd.u_aim / std::fmax(1.0, u));
    }
}

struct Nlp {
    Model M;
    mocob200_transcription R{};
    FdOptions fd;
    bool hs = true;
    int nMesh = 0, nIntervals = 0, nGrid = 0, ns = 0, nc = 0, nd = 0, nq = 0, nIn = 0, nOut = 0;
    int nqs = 0; // coordinate value states in the NLP (nq, or 0 with prescribed kinematics)
    int offStates = 2, offControls = 0, offMult = 0, offSlacks = 0, offDerivs = 0, nVar = 0, nCon = 0, nEin = 0;
    int nm = 0, nsl = 0, nKc = 0;       // multipliers per grid point, slacks per interior point, kinematic rows per mesh point
    std::vector<int> kinRow;            // first kinematic-constraint row per mesh point
    int nResidPerPoint = 0, nEndpointRows = 0;
    std::vector<double> mesh, grid, quad;
    std::vector<int> goalRow, goalSlot, goalNout; // first endpoint row (or -1), integrand slot (or -1), outputs
    std::vector<int> residRow, defectRow, interpRow; // first row per grid point / interval
    std::vector<int> pathGoals, pathRow;              // path constraints (goal records) ; first path row per mesh point
    int nPathPerMesh = 0;
    bool interp = false;

    void init(const mocob200_tables* T, const mocob200_transcription* tr) {
        R = *tr;
        M.init(T, tr->dynamics_mode == MOCOB200_DYNAMICS_IMPLICIT, tr->enforce_constraint_derivatives != 0);
        fd.defaults(tr->fd_scheme);
        hs = tr->scheme == MOCOB200_SCHEME_HERMITE_SIMPSON;
        nMesh = tr->n_mesh; nIntervals = nMesh - 1;
        mesh.resize(nMesh);
        for (int i = 0; i < nMesh; ++i) mesh[i] = tr->mesh ? tr->mesh[i] : (double)i / nIntervals;
        // grid: HS adds interval midpoints (CasOCHermiteSimpson.h:45-66)
        grid.clear();
        for (int i = 0; i < nMesh; ++i) {
            grid.push_back(mesh[i]);
            if (hs && i + 1 < nMesh) grid.push_back(0.5 * (mesh[i] + mesh[i + 1]));
        }
        nGrid = (int)grid.size();
        quad.assign(nGrid, 0.0);
        for (int i = 0; i < nIntervals; ++i) {
            const double dm = mesh[i + 1] - mesh[i];
            if (hs) { quad[2 * i] += dm / 6.0; quad[2 * i + 1] += 2.0 * dm / 3.0; quad[2 * i + 2] += dm / 6.0; }
            else { quad[i] += 0.5 * dm; quad[i + 1] += 0.5 * dm; }
        }
        if (M.prescribed && !M.implicit) throw std::runtime_error("Prescribed kinematics (PositionMotion) requires implicit dynamics mode.");
        nq = M.nq; nqs = M.nqs; ns = 2 * M.nqs + M.nz; nc = M.nc; nd = M.nd; nIn = M.n_in; nOut = M.n_out;
        // x = [t0, tf, states, controls, multipliers, slacks (Hermite-Simpson midpoints, with enforced derivatives), derivatives]
        nm = M.nm; nKc = M.n_kc;
        nsl = (M.enforceDerivs && hs) ? nm : 0;
        offControls = offStates + ns * nGrid;
        offMult = offControls + nc * nGrid;
        offSlacks = offMult + nm * nGrid;
        offDerivs = offSlacks + nsl * (nGrid - nMesh);
        nVar = offDerivs + nd * nGrid;
        nEin = 2 * (1 + ns + nc + nm + nd) + 1;
        // rows
        goalRow.assign(T->ngoal, -1); goalSlot.assign(T->ngoal, -1); goalNout.assign(T->ngoal, 0);
        pathGoals.clear(); nPathPerMesh = 0;
        int row = 0, slot = 0;
        for (int g = 0; g < T->ngoal; ++g) {
            const int* gr = M.goalRec(g);
            goalNout[g] = gr[2];
            if (gr[1] == MOCOB200_MODE_PATH_CONSTRAINT) { pathGoals.push_back(g); nPathPerMesh += gr[2]; continue; }
            if (gr[3]) goalSlot[g] = slot++;
            if (gr[1] == MOCOB200_MODE_ENDPOINT_CONSTRAINT) { goalRow[g] = row; row += gr[2]; }
        }
        nEndpointRows = row;
        pathRow.assign(nMesh, -1);
        kinRow.assign(nMesh, -1);
        nResidPerPoint = (M.implicit ? nq : 0) + M.n_auxres;
        interp = hs && tr->interpolate_control_midpoints && nc > 0;
        residRow.assign(nGrid, -1); defectRow.assign(nIntervals, -1); interpRow.assign(nIntervals, -1);
        int igrid = 0;
        for (int im = 0; im < nMesh; ++im) {
            kinRow[im] = row; row += nKc;           // copyColumn(constraints.kinematic, imesh) comes first
            pathRow[im] = row; row += nPathPerMesh; // copyColumn(path, imesh) for every path constraint
            if (im < nIntervals) {
                while (grid[igrid] < mesh[im + 1]) { residRow[igrid] = row; row += nResidPerPoint; ++igrid; }
                defectRow[im] = row; row += (hs ? 2 : 1) * ns;
                if (interp) { interpRow[im] = row; row += nc; }
            }
        }
        residRow[nGrid - 1] = row; row += nResidPerPoint;
        nCon = row;
    }

    int meshGridIndex(int im) const { return hs ? 2 * im : im; }
    // FD block of path constraint pg at mesh point im: Jp[nout][nIn] (its own FiniteDiff function)
    void pathJacobian(const double* x, int im, int pg, double* Jp) const {
        const int nout = goalNout[pg];
        std::vector<double> in(nIn), y0(std::max(nout, 1)), col(std::max(nout, 1));
        gatherInputs(x, meshGridIndex(im), in.data());
        evalPathConstraint<double>(M, pg, in.data(), y0.data());
        auto f = [&](const double* xin, double* yout) { evalPathConstraint<double>(M, pg, xin, yout); };
        for (int a = 0; a < nIn; ++a) {
            fdColumn(f, fd, in, a, y0.data(), nout, col.data());
            for (int o = 0; o < nout; ++o) Jp[(size_t)o * nIn + a] = col[o];
        }
    }

    // ---- point functions over the trajectory
    void gatherInputs(const double* x, int j, double* in) const {
        in[0] = x[0] + (x[1] - x[0]) * grid[j];
        for (int k = 0; k < ns; ++k) in[1 + k] = x[offStates + (size_t)j * ns + k];
        for (int k = 0; k < nc; ++k) in[1 + ns + k] = x[offControls + (size_t)j * nc + k];
        for (int k = 0; k < nm; ++k) in[1 + ns + nc + k] = x[offMult + (size_t)j * nm + k];
        for (int k = 0; k < nd; ++k) in[1 + ns + nc + nm + k] = x[offDerivs + (size_t)j * nd + k];
    }
    // velocity correction at the interior (midpoint) grid point j: qdot_j = u_j + G(q_j)^T gamma_j (CasOCTranscription.cpp:315-332)
    const double* slacksAt(const double* x, int j) const { return x + offSlacks + (size_t)(j / 2) * nsl; }
    bool corrected(int j) const { return nsl > 0 && hs && (j % 2 == 1); }
    void evalPoints(const double* x, double* Y) const {
#pragma omp parallel for schedule(dynamic)
        for (int j = 0; j < nGrid; ++j) {
            std::vector<double> in(nIn);
            gatherInputs(x, j, in.data());
            evalPoint<double>(M, in.data(), Y + (size_t)j * nOut);
        }
    }
    // FD block of grid point j: J[nOut][nIn]; every reference callback (dynamics, each integrand) is its
    // own FiniteDiff function with its own step refinement.
    void pointJacobian(const double* x, int j, const double* Yj, double* J) const {
        std::vector<double> in(nIn), col(nOut), tmp(nOut);
        gatherInputs(x, j, in.data());
        const int ngroups = 1 + M.n_integrand;
        for (int a = 0; a < nIn; ++a) {
            for (int grp = 0; grp < ngroups; ++grp) {
                const int lo = grp == 0 ? 0 : M.n_dyn_out + grp - 1, hi = grp == 0 ? M.n_dyn_out : lo + 1;
                if (hi == lo) continue;
                const unsigned mask = 1u << grp;
                auto f = [&](const double* xin, double* yout) {
                    evalPoint<double>(M, xin, tmp.data(), mask);
                    for (int o = lo; o < hi; ++o) yout[o - lo] = tmp[o];
                };
                fdColumn(f, fd, in, a, Yj + lo, hi - lo, col.data());
                for (int o = lo; o < hi; ++o) J[(size_t)o * nIn + a] = col[o - lo];
            }
        }
    }
    void evalPointJacobians(const double* x, double* Y, double* J) const {
        evalPoints(x, Y);
#pragma omp parallel for schedule(dynamic)
        for (int j = 0; j < nGrid; ++j) pointJacobian(x, j, Y + (size_t)j * nOut, J + (size_t)j * nOut * nIn);
    }

    // ---- state derivatives used by the defects: qdot = u, udot = callback (explicit) or the
    // acceleration variables (implicit), zdot = callback (CasOCTranscription.cpp:313-411)
    double xdot(const double* x, const double* Y, int j, int s) const {
        if (s < nqs && corrected(j)) {
            std::vector<double> corr(nq);
            velocityCorrection<double>(M, x + offStates + (size_t)j * ns, slacksAt(x, j), corr.data());
            return x[offStates + (size_t)j * ns + nq + s] + corr[s];
        }
        if (s < nqs) return x[offStates + (size_t)j * ns + nqs + s];
        if (s < 2 * nqs && M.implicit) return x[offDerivs + (size_t)j * nd + (s - nqs)];
        return Y[(size_t)j * nOut + nq + (s - 2 * nqs)]; // auxiliary state: its zdot follows the nq multibody outputs
    }
    double integral(const double* x, const double* Y, int slot) const {
        double acc = 0;
        for (int j = 0; j < nGrid; ++j) acc += quad[j] * Y[(size_t)j * nOut + M.n_dyn_out + slot];
        return (x[1] - x[0]) * acc;
    }
    void endpointInputs(const double* x, const double* Y, int g, double* ein) const {
        const int half = 1 + ns + nc + nm + nd, last = nGrid - 1;
        ein[0] = x[0]; ein[half] = x[1];
        for (int k = 0; k < ns; ++k) { ein[1 + k] = x[offStates + k]; ein[half + 1 + k] = x[offStates + (size_t)last * ns + k]; }
        for (int k = 0; k < nc; ++k) { ein[1 + ns + k] = x[offControls + k]; ein[half + 1 + ns + k] = x[offControls + (size_t)last * nc + k]; }
        for (int k = 0; k < nm; ++k) { ein[1 + ns + nc + k] = x[offMult + k]; ein[half + 1 + ns + nc + k] = x[offMult + (size_t)last * nm + k]; }
        for (int k = 0; k < nd; ++k) { ein[1 + ns + nc + nm + k] = x[offDerivs + k]; ein[half + 1 + ns + nc + nm + k] = x[offDerivs + (size_t)last * nd + k]; }
        ein[2 * half] = goalSlot[g] >= 0 ? integral(x, Y, goalSlot[g]) : NAN;
    }

    void evalG(const double* x, double* g) const {
        std::vector<double> Y((size_t)nGrid * nOut);
        evalPoints(x, Y.data());
        assembleG(x, Y.data(), g);
    }
    void assembleG(const double* x, const double* Y, double* g) const {
        std::vector<double> ein(nEin), vals(256);
        for (int gi = 0; gi < M.T->ngoal; ++gi) {
            if (goalRow[gi] < 0) continue;
            vals.assign(std::max(256, goalNout[gi]), 0.0);
            endpointInputs(x, Y, gi, ein.data());
            evalEndpoint<double>(M, gi, ein.data(), vals.data());
            for (int o = 0; o < goalNout[gi]; ++o) g[goalRow[gi] + o] = vals[o];
        }
        for (int im = 0; im < nMesh; ++im) // kinematic-constraint errors at the mesh points (m_constraints.kinematic = out.at(3))
            for (int e = 0; e < nKc; ++e) g[kinRow[im] + e] = Y[(size_t)meshGridIndex(im) * nOut + nq + M.nz + M.n_auxres + e];
        for (int im = 0; im < nMesh; ++im) { // path constraints at the mesh points (m_meshIndices)
            std::vector<double> in(nIn);
            gatherInputs(x, meshGridIndex(im), in.data());
            int r = pathRow[im];
            for (int pg : pathGoals) { evalPathConstraint<double>(M, pg, in.data(), g + r); r += goalNout[pg]; }
        }
        for (int j = 0; j < nGrid; ++j) {
            int r = residRow[j];
            if (M.implicit) for (int k = 0; k < nq; ++k) g[r++] = Y[(size_t)j * nOut + k];
            for (int k = 0; k < M.n_auxres; ++k) g[r++] = Y[(size_t)j * nOut + nq + M.nz + k];
        }
        const double T = x[1] - x[0];
        for (int m = 0; m < nIntervals; ++m) {
            const double h = T * (mesh[m + 1] - mesh[m]);
            for (int s = 0; s < ns; ++s) {
                if (!hs) {
                    const double xi = x[offStates + (size_t)m * ns + s], xp = x[offStates + (size_t)(m + 1) * ns + s];
                    g[defectRow[m] + s] = xp - (xi + 0.5 * h * (xdot(x, Y, m + 1, s) + xdot(x, Y, m, s)));
                } else {
                    const int i = 2 * m, mid = i + 1, ip = i + 2;
                    const double xi = x[offStates + (size_t)i * ns + s], xm = x[offStates + (size_t)mid * ns + s],
                                 xp = x[offStates + (size_t)ip * ns + s];
                    const double di = xdot(x, Y, i, s), dm = xdot(x, Y, mid, s), dp = xdot(x, Y, ip, s);
                    g[defectRow[m] + s] = xm - 0.5 * (xp + xi) - (h / 8.0) * (di - dp);
                    g[defectRow[m] + ns + s] = xp - xi - (h / 6.0) * (dp + 4.0 * dm + di);
                }
            }
            if (interp)
                for (int c = 0; c < nc; ++c) {
                    const double ci = x[offControls + (size_t)(2 * m) * nc + c], cm = x[offControls + (size_t)(2 * m + 1) * nc + c],
                                 cp = x[offControls + (size_t)(2 * m + 2) * nc + c];
                    g[interpRow[m] + c] = cm - 0.5 * (cp + ci);
                }
        }
    }

    int numCostTerms() const {
        int n = 0;
        for (int g = 0; g < M.T->ngoal; ++g) if (M.goalRec(g)[1] == MOCOB200_MODE_COST) ++n;
        if (nm && R.minimize_lagrange_multipliers) ++n;
        if (M.implicit && R.minimize_accelerations) ++n;
        if (M.n_auxres && R.minimize_auxiliary_derivatives) ++n;
        return n;
    }
    // lagrange_multiplier_weight * duration * sum_j quad_j sum_k lambda_jk^2 (CasOCTranscription.cpp:513-521)
    double multiplierPenalty(const double* x) const {
        double acc = 0;
        for (int j = 0; j < nGrid; ++j) {
            double sq = 0;
            for (int k = 0; k < nm; ++k) { const double d = x[offMult + (size_t)j * nm + k]; sq += d * d; }
            acc += quad[j] * sq;
        }
        return R.lagrange_multiplier_weight * (x[1] - x[0]) * acc;
    }
    double penalty(const double* x, int lo, int hi, double w) const {
        double acc = 0;
        for (int j = 0; j < nGrid; ++j) {
            double sq = 0;
            for (int k = lo; k < hi; ++k) { const double d = x[offDerivs + (size_t)j * nd + k]; sq += d * d; }
            acc += quad[j] * sq;
        }
        return w * (x[1] - x[0]) * acc;
    }
    double evalF(const double* x, double* terms) const {
        std::vector<double> Y((size_t)nGrid * nOut);
        evalPoints(x, Y.data());
        return assembleF(x, Y.data(), terms);
    }
    double assembleF(const double* x, const double* Y, double* terms) const {
        std::vector<double> ein(nEin), vals;
        double total = 0;
        int t = 0;
        for (int gi = 0; gi < M.T->ngoal; ++gi) {
            if (M.goalRec(gi)[1] != MOCOB200_MODE_COST) continue;
            vals.assign(std::max(256, goalNout[gi]), 0.0);
            endpointInputs(x, Y, gi, ein.data());
            evalEndpoint<double>(M, gi, ein.data(), vals.data());
            double s = 0;
            for (int o = 0; o < goalNout[gi]; ++o) s += vals[o];
            if (terms) terms[t] = s;
            ++t; total += s;
        }
        if (nm && R.minimize_lagrange_multipliers) {
            const double s = multiplierPenalty(x);
            if (terms) terms[t] = s;
            ++t; total += s;
        }
        if (M.implicit && !M.prescribed && R.minimize_accelerations) {
            const double s = penalty(x, 0, nq, R.accelerations_weight);
            if (terms) terms[t] = s;
            ++t; total += s;
        }
        if (M.n_auxres && R.minimize_auxiliary_derivatives) {
            const double s = penalty(x, nd - M.n_auxres, nd, R.auxiliary_derivatives_weight);
            if (terms) terms[t] = s;
            ++t; total += s;
        }
        return total;
    }
};

} // namespace orc
