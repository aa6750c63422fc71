// TEST INFRASTRUCTURE ONLY — CPU oracle: jac_g (CSC) and grad_f by row-wise chaining of the callback
// FD blocks through the transcription, and the structural pattern CasADi would report when every
// callback has a dense Jacobian block (optim_sparsity_detection = "none", the solver default).
#pragma once
#include "orc_nlp.hpp"

namespace orc {

struct SparseRow { // scratch row with a touched list
    std::vector<double> v;
    std::vector<char> on;
    std::vector<int> touched;
    void reset(int n) { v.assign(n, 0.0); on.assign(n, 0); touched.clear(); }
    void clear() { for (int c : touched) { v[c] = 0; on[c] = 0; } touched.clear(); }
    void add(int c, double x) { if (!on[c]) { on[c] = 1; touched.push_back(c); } v[c] += x; }
    void mark(int c) { add(c, 0.0); }
};

struct JacBuilder {
    const Nlp& P;
    explicit JacBuilder(const Nlp& p) : P(p) {}

    int colOfInput(int j, int a) const { // NLP column of input a (>=1) of grid point j
        const int k = a - 1;
        if (k < P.ns) return P.offStates + j * P.ns + k;
        if (k < P.ns + P.nc) return P.offControls + j * P.nc + (k - P.ns);
        if (k < P.ns + P.nc + P.nm) return P.offMult + j * P.nm + (k - P.ns - P.nc);
        return P.offDerivs + j * P.nd + (k - P.ns - P.nc - P.nm);
    }
    // scatter  coef * d(callback output row o of point j)/d(everything)  into the row
    // (time input chains to t0 and tf: time_j = t0 + (tf - t0) grid_j)
    void addCallbackRow(SparseRow& r, const double* J, int j, int o, double coef, bool structural) const {
        const double* Jr = J ? J + ((size_t)j * P.nOut + o) * P.nIn : nullptr;
        const double gj = P.grid[j];
        if (structural) { r.mark(0); r.mark(1); for (int a = 1; a < P.nIn; ++a) r.mark(colOfInput(j, a)); return; }
        r.add(0, coef * Jr[0] * (1.0 - gj));
        r.add(1, coef * Jr[0] * gj);
        for (int a = 1; a < P.nIn; ++a) r.add(colOfInput(j, a), coef * Jr[a]);
    }
    // coef * d xdot_s(j) / d(everything)
    // FD block of the VelocityCorrection callback at midpoint j (CasOCFunction.cpp:250-291: inputs time, multibody states,
    // slacks; output the nq corrections), row s, chained into the NLP columns
    void addVelocityCorrectionRow(SparseRow& r, const double* x, int j, int s, double coef, bool structural) const {
        const int nmb = 2 * P.nq, n = 1 + nmb + P.nsl;
        const int slackCol = P.offSlacks + (j / 2) * P.nsl;
        if (structural) {
            r.mark(0); r.mark(1);
            for (int a = 0; a < nmb; ++a) r.mark(P.offStates + j * P.ns + a);
            for (int k = 0; k < P.nsl; ++k) r.mark(slackCol + k);
            return;
        }
        std::vector<double> in(n), y0(P.nq), col(P.nq);
        in[0] = x[0] + (x[1] - x[0]) * P.grid[j];
        for (int a = 0; a < nmb; ++a) in[1 + a] = x[P.offStates + (size_t)j * P.ns + a];
        for (int k = 0; k < P.nsl; ++k) in[1 + nmb + k] = x[slackCol + k];
        auto f = [&](const double* xin, double* yout) { velocityCorrection<double>(P.M, xin + 1, xin + 1 + nmb, yout); };
        f(in.data(), y0.data());
        for (int a = 0; a < n; ++a) {
            fdColumn(f, P.fd, in, a, y0.data(), P.nq, col.data());
            const double d = coef * col[s];
            if (a == 0) { r.add(0, d * (1.0 - P.grid[j])); r.add(1, d * P.grid[j]); }
            else if (a <= nmb) r.add(P.offStates + j * P.ns + (a - 1), d);
            else r.add(slackCol + (a - 1 - nmb), d);
        }
    }
    const double* xptr = nullptr;
    void addXdotRow(SparseRow& r, const double* J, int j, int s, double coef, bool structural) const {
        if (s < P.nqs && P.corrected(j)) addVelocityCorrectionRow(r, xptr, j, s, coef, structural);
        if (s < P.nqs) { if (structural) r.mark(P.offStates + j * P.ns + P.nqs + s); else r.add(P.offStates + j * P.ns + P.nqs + s, coef); return; }
        if (s < 2 * P.nqs && P.M.implicit) { const int c = P.offDerivs + j * P.nd + (s - P.nqs); if (structural) r.mark(c); else r.add(c, coef); return; }
        addCallbackRow(r, J, j, P.nq + (s - 2 * P.nqs), coef, structural);
    }
    // d(goal output o)/d(everything): endpoint FD block E[nout][nEin] + chain through the integral
    void addEndpointRow(SparseRow& r, const double* x, const double* J, const double* E, int g, int o, double coef,
                        bool structural) const {
        const int half = 1 + P.ns + P.nc + P.nm + P.nd, last = P.nGrid - 1;
        const double* Er = E ? E + (size_t)o * P.nEin : nullptr;
        const int slot = P.goalSlot[g];
        if (structural) {
            r.mark(0); r.mark(1);
            for (int a = 1; a < half; ++a) { r.mark(colOfInput(0, a)); r.mark(colOfInput(last, a)); }
            if (slot >= 0) for (int j = 0; j < P.nGrid; ++j) addCallbackRow(r, nullptr, j, P.M.n_dyn_out + slot, 0, true);
            return;
        }
        r.add(0, coef * Er[0]);
        r.add(1, coef * Er[half]);
        for (int a = 1; a < half; ++a) { r.add(colOfInput(0, a), coef * Er[a]); r.add(colOfInput(last, a), coef * Er[half + a]); }
        if (slot >= 0) {
            // I = (tf - t0) sum_j quad_j y_j
            const double dGdI = coef * Er[2 * half];
            const double T = x[1] - x[0];
            double sum = 0;
            for (int j = 0; j < P.nGrid; ++j) sum += P.quad[j] * Yptr[(size_t)j * P.nOut + P.M.n_dyn_out + slot];
            r.add(0, -dGdI * sum);
            r.add(1, dGdI * sum);
            for (int j = 0; j < P.nGrid; ++j) addCallbackRow(r, J, j, P.M.n_dyn_out + slot, dGdI * T * P.quad[j], false);
        }
    }
    const double* Yptr = nullptr;

    // endpoint FD blocks of goal g at x: E[nout][nEin]; the integral input is a NaN constant when the goal
    // has no integrand (CasOCTranscription.cpp:495) so that column is never used.
    void endpointBlock(const double* x, const double* Y, int g, std::vector<double>& E) const {
        const int nout = P.goalNout[g];
        std::vector<double> ein(P.nEin), v0(std::max(nout, 1)), col(std::max(nout, 1));
        P.endpointInputs(x, Y, g, ein.data());
        evalEndpoint<double>(P.M, g, ein.data(), v0.data());
        E.assign((size_t)nout * P.nEin, 0.0);
        auto f = [&](const double* in, double* out) { evalEndpoint<double>(P.M, g, in, out); };
        for (int a = 0; a < P.nEin; ++a) {
            if (a == P.nEin - 1 && P.goalSlot[g] < 0) continue;
            fdColumn(f, P.fd, ein, a, v0.data(), nout, col.data());
            for (int o = 0; o < nout; ++o) E[(size_t)o * P.nEin + a] = col[o];
        }
    }

    // Fill row `row` of jac_g (structural: pattern only).
    void constraintRow(SparseRow& r, int row, const double* x, const double* Y, const double* J,
                       const std::vector<std::vector<double>>* Eblocks, bool structural) const {
        const Nlp& N = P;
        // endpoint rows
        for (int g = 0; g < N.M.T->ngoal; ++g)
            if (N.goalRow[g] >= 0 && row >= N.goalRow[g] && row < N.goalRow[g] + N.goalNout[g]) {
                addEndpointRow(r, x, J, structural ? nullptr : (*Eblocks)[g].data(), g, row - N.goalRow[g], 1.0, structural);
                return;
            }
        // kinematic-constraint rows: output 3 of the multibody callback at the mesh points (CasOCTranscription.cpp:360, 398)
        if (N.nKc)
            for (int im = 0; im < N.nMesh; ++im)
                if (row >= N.kinRow[im] && row < N.kinRow[im] + N.nKc) {
                    addCallbackRow(r, J, N.meshGridIndex(im), N.nq + N.M.nz + N.M.n_auxres + (row - N.kinRow[im]), 1.0, structural);
                    return;
                }
        // path-constraint rows: each path constraint is its own callback with a dense block over the mesh point's inputs
        if (N.nPathPerMesh)
            for (int im = 0; im < N.nMesh; ++im)
                if (row >= N.pathRow[im] && row < N.pathRow[im] + N.nPathPerMesh) {
                    int k = row - N.pathRow[im];
                    const int j = N.meshGridIndex(im);
                    for (int pg : N.pathGoals) {
                        if (k >= N.goalNout[pg]) { k -= N.goalNout[pg]; continue; }
                        if (structural) { r.mark(0); r.mark(1); for (int a = 1; a < N.nIn; ++a) r.mark(colOfInput(j, a)); return; }
                        std::vector<double> Jp((size_t)N.goalNout[pg] * N.nIn);
                        N.pathJacobian(x, im, pg, Jp.data());
                        const double* Jr = Jp.data() + (size_t)k * N.nIn;
                        r.add(0, Jr[0] * (1.0 - N.grid[j]));
                        r.add(1, Jr[0] * N.grid[j]);
                        for (int a = 1; a < N.nIn; ++a) r.add(colOfInput(j, a), Jr[a]);
                        return;
                    }
                }
        // residual rows
        for (int j = 0; j < N.nGrid; ++j)
            if (row >= N.residRow[j] && row < N.residRow[j] + N.nResidPerPoint) {
                int k = row - N.residRow[j];
                int o;
                if (N.M.implicit && k < N.nq) o = k;
                else o = N.nq + N.M.nz + (k - (N.M.implicit ? N.nq : 0));
                addCallbackRow(r, J, j, o, 1.0, structural);
                return;
            }
        for (int m = 0; m < N.nIntervals; ++m) {
            const int nDef = (N.hs ? 2 : 1) * N.ns;
            if (row >= N.defectRow[m] && row < N.defectRow[m] + nDef) {
                const int k = row - N.defectRow[m];
                const int s = k % N.ns;
                const bool simpson = N.hs && k >= N.ns;
                const double dm = N.mesh[m + 1] - N.mesh[m];
                const double h = structural ? 0 : (x[1] - x[0]) * dm;
                auto X = [&](int j) { return N.offStates + j * N.ns + s; };
                // defect = lin + h * sum_p w_p xdot_p ; d = dlin + (dh) * sum + h * sum w_p dxdot_p
                int pts[3]; double w[3]; int np;
                if (!N.hs) {
                    np = 2; pts[0] = m; pts[1] = m + 1; w[0] = -0.5; w[1] = -0.5;
                    if (structural) { r.mark(X(m)); r.mark(X(m + 1)); } else { r.add(X(m), -1.0); r.add(X(m + 1), 1.0); }
                } else if (!simpson) {
                    np = 2; pts[0] = 2 * m; pts[1] = 2 * m + 2; w[0] = -1.0 / 8.0; w[1] = 1.0 / 8.0;
                    if (structural) { r.mark(X(2 * m)); r.mark(X(2 * m + 1)); r.mark(X(2 * m + 2)); }
                    else { r.add(X(2 * m), -0.5); r.add(X(2 * m + 1), 1.0); r.add(X(2 * m + 2), -0.5); }
                } else {
                    np = 3; pts[0] = 2 * m; pts[1] = 2 * m + 1; pts[2] = 2 * m + 2; w[0] = -1.0 / 6.0; w[1] = -4.0 / 6.0; w[2] = -1.0 / 6.0;
                    if (structural) { r.mark(X(2 * m)); r.mark(X(2 * m + 2)); } else { r.add(X(2 * m), -1.0); r.add(X(2 * m + 2), 1.0); }
                }
                if (structural) { r.mark(0); r.mark(1); }
                double wsum = 0;
                for (int p = 0; p < np; ++p) {
                    if (!structural) wsum += w[p] * N.xdot(x, Y, pts[p], s);
                    addXdotRow(r, J, pts[p], s, w[p] * h, structural);
                }
                if (!structural) { r.add(0, -dm * wsum); r.add(1, dm * wsum); }
                return;
            }
            if (N.interp && row >= N.interpRow[m] && row < N.interpRow[m] + N.nc) {
                const int c = row - N.interpRow[m];
                auto Cc = [&](int j) { return N.offControls + j * N.nc + c; };
                if (structural) { r.mark(Cc(2 * m)); r.mark(Cc(2 * m + 1)); r.mark(Cc(2 * m + 2)); }
                else { r.add(Cc(2 * m), -0.5); r.add(Cc(2 * m + 1), 1.0); r.add(Cc(2 * m + 2), -0.5); }
                return;
            }
        }
    }

    // CSC structure (columns ascending, rows ascending within a column)
    void structure(std::vector<int>& colptr, std::vector<int>& rowidx) {
        std::vector<std::vector<int>> byCol(P.nVar);
        SparseRow r; r.reset(P.nVar);
        for (int row = 0; row < P.nCon; ++row) {
            r.clear();
            constraintRow(r, row, nullptr, nullptr, nullptr, nullptr, true);
            for (int c : r.touched) byCol[c].push_back(row);
        }
        colptr.assign(P.nVar + 1, 0); rowidx.clear();
        for (int c = 0; c < P.nVar; ++c) {
            colptr[c] = (int)rowidx.size();
            rowidx.insert(rowidx.end(), byCol[c].begin(), byCol[c].end()); // rows were visited in ascending order
        }
        colptr[P.nVar] = (int)rowidx.size();
    }

    // values in the CSC order of structure()
    void jacobian(const double* x, const std::vector<int>& colptr, const std::vector<int>& rowidx, double* vals) {
        std::vector<double> Y((size_t)P.nGrid * P.nOut), J((size_t)P.nGrid * P.nOut * P.nIn);
        P.evalPointJacobians(x, Y.data(), J.data());
        Yptr = Y.data();
        xptr = x;
        std::vector<std::vector<double>> E(P.M.T->ngoal);
        for (int g = 0; g < P.M.T->ngoal; ++g) if (P.goalRow[g] >= 0) endpointBlock(x, Y.data(), g, E[g]);
        std::vector<int> next(colptr.begin(), colptr.end() - 1);
        SparseRow r; r.reset(P.nVar);
        for (int row = 0; row < P.nCon; ++row) {
            r.clear();
            constraintRow(r, row, x, Y.data(), J.data(), &E, false);
            // rows ascend, so every column's next free slot must be this row
            std::sort(r.touched.begin(), r.touched.end());
            for (int c : r.touched) {
                int& k = next[c];
                if (k < colptr[c + 1] && rowidx[k] == row) vals[k++] = r.v[c];
            }
        }
    }

    void gradient(const double* x, double* grad) {
        std::vector<double> Y((size_t)P.nGrid * P.nOut), J((size_t)P.nGrid * P.nOut * P.nIn);
        P.evalPointJacobians(x, Y.data(), J.data());
        Yptr = Y.data();
        SparseRow r; r.reset(P.nVar);
        std::vector<double> E;
        for (int g = 0; g < P.M.T->ngoal; ++g) {
            if (P.M.goalRec(g)[1] != MOCOB200_MODE_COST) continue;
            endpointBlock(x, Y.data(), g, E);
            for (int o = 0; o < P.goalNout[g]; ++o) addEndpointRow(r, x, J.data(), E.data(), g, o, 1.0, false);
        }
        auto pen = [&](int lo, int hi, double w) {
            const double T = x[1] - x[0];
            double acc = 0;
            for (int j = 0; j < P.nGrid; ++j) {
                double sq = 0;
                for (int k = lo; k < hi; ++k) {
                    const int c = P.offDerivs + j * P.nd + k;
                    sq += x[c] * x[c];
                    r.add(c, 2.0 * w * T * P.quad[j] * x[c]);
                }
                acc += P.quad[j] * sq;
            }
            r.add(0, -w * acc);
            r.add(1, w * acc);
        };
        if (P.nm && P.R.minimize_lagrange_multipliers) {
            const double T = x[1] - x[0], w = P.R.lagrange_multiplier_weight;
            double acc = 0;
            for (int j = 0; j < P.nGrid; ++j) {
                double sq = 0;
                for (int k = 0; k < P.nm; ++k) {
                    const int c = P.offMult + j * P.nm + k;
                    sq += x[c] * x[c];
                    r.add(c, 2.0 * w * T * P.quad[j] * x[c]);
                }
                acc += P.quad[j] * sq;
            }
            r.add(0, -w * acc);
            r.add(1, w * acc);
        }
        if (P.M.implicit && !P.M.prescribed && P.R.minimize_accelerations) pen(0, P.nq, P.R.accelerations_weight);
        if (P.M.n_auxres && P.R.minimize_auxiliary_derivatives) pen(P.nd - P.M.n_auxres, P.nd, P.R.auxiliary_derivatives_weight);
        for (int c = 0; c < P.nVar; ++c) grad[c] = r.v[c];
    }
};

} // namespace orc
