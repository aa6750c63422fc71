// Host-only half of the C ABI: example problems, save/load of flattened problems, NLP layout.
#include <cstdlib>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <limits>
#include <random>
#include <stdexcept>

#include "host_model.h"
#include "layout.h"

using namespace mocob200;

struct mocob200_problem {
    std::unique_ptr<FlatProblem> fp;
};

namespace {
void setErr(char* err, int errlen, const std::string& m) {
    if (err && errlen > 0) {
        std::snprintf(err, (size_t)errlen, "%s", m.c_str());
    }
}
const unsigned kMagic = 0x4d423230u; // "MB20"
} // namespace

extern "C" {

int mocob200_problem_create_example(const char* name, const char* data_dir, const char* options, mocob200_problem** out,
                                    char* err, int errlen) {
    if (!name || !out) return MOCOB200_ERR_ARG;
    try {
        auto p = new mocob200_problem;
        p->fp = buildExample(name, data_dir ? data_dir : ".", options ? options : "");
        *out = p;
        return MOCOB200_OK;
    } catch (const std::exception& e) {
        setErr(err, errlen, e.what());
        return MOCOB200_ERR_UNSUPPORTED;
    }
}

int mocob200_problem_save(const mocob200_problem* p, const char* path) {
    if (!p || !path) return MOCOB200_ERR_ARG;
    std::ofstream f(path, std::ios::binary);
    if (!f) return MOCOB200_ERR_IO;
    const FlatProblem& fp = *p->fp;
    unsigned magic = kMagic, ver = 2, szT = (unsigned)sizeof(mocob200_tables), szR = (unsigned)sizeof(mocob200_transcription);
    f.write((const char*)&magic, 4);
    f.write((const char*)&ver, 4);
    f.write((const char*)&szT, 4); // version 2: the struct sizes travel with the file, so later (appended) fields stay readable
    f.write((const char*)&szR, 4);
    mocob200_tables t = fp.tables;
    t.dbl = nullptr; t.idx = nullptr;
    mocob200_transcription r = fp.tr;
    r.mesh = nullptr;
    f.write((const char*)&t, sizeof t);
    f.write((const char*)&r, sizeof r);
    f.write((const char*)fp.dbl.data(), (std::streamsize)(fp.dbl.size() * sizeof(double)));
    f.write((const char*)fp.idx.data(), (std::streamsize)(fp.idx.size() * sizeof(int)));
    f.write((const char*)fp.mesh.data(), (std::streamsize)(fp.mesh.size() * sizeof(double)));
    return f ? MOCOB200_OK : MOCOB200_ERR_IO;
}

// options (may be NULL): "num_mesh_intervals=..;scheme=..;dynamics_mode=..;fd_scheme=.." override the stored transcription
int mocob200_problem_load_opts(const char* path, const char* options, mocob200_problem** out, char* err, int errlen) {
    if (!path || !out) return MOCOB200_ERR_ARG;
    try {
        std::ifstream f(path, std::ios::binary);
        if (!f) throw std::runtime_error(std::string("cannot open '") + path + "'");
        unsigned magic = 0, ver = 0;
        f.read((char*)&magic, 4);
        f.read((char*)&ver, 4);
        if (magic != kMagic || (ver != 1 && ver != 2)) throw std::runtime_error("not a mocob200 problem file");
        // version 1 files hold the structs as they were then (144 / 96 bytes); later fields are appended and zero = absent
        unsigned szT = 144, szR = 96;
        if (ver >= 2) { f.read((char*)&szT, 4); f.read((char*)&szR, 4); }
        if (szT > sizeof(mocob200_tables) || szR > sizeof(mocob200_transcription) || szT < 144 || szR < 96)
            throw std::runtime_error("problem file written by a newer library");
        auto fp = std::make_unique<FlatProblem>();
        std::memset(&fp->tables, 0, sizeof fp->tables);
        std::memset(&fp->tr, 0, sizeof fp->tr);
        f.read((char*)&fp->tables, szT);
        f.read((char*)&fp->tr, szR);
        if (szR < sizeof(mocob200_transcription)) { // solver defaults of the fields the file predates
            fp->tr.enforce_constraint_derivatives = 0;
            fp->tr.minimize_lagrange_multipliers = 0;
            fp->tr.lagrange_multiplier_weight = 1.0;
            fp->tr.velocity_correction_bounds[0] = -0.1; fp->tr.velocity_correction_bounds[1] = 0.1;
            fp->tr.multiplier_bounds[0] = -1000; fp->tr.multiplier_bounds[1] = 1000;
        }
        if (fp->tables.n_dbl < 0 || fp->tables.n_int < 0 || fp->tr.n_mesh < 2 || fp->tables.n_dbl > (1 << 28) || fp->tables.n_int > (1 << 28) ||
            fp->tr.n_mesh > (1 << 24))
            throw std::runtime_error("corrupt problem file");
        fp->dbl.resize(fp->tables.n_dbl);
        fp->idx.resize(fp->tables.n_int);
        fp->mesh.resize(fp->tr.n_mesh);
        f.read((char*)fp->dbl.data(), (std::streamsize)(fp->dbl.size() * sizeof(double)));
        f.read((char*)fp->idx.data(), (std::streamsize)(fp->idx.size() * sizeof(int)));
        f.read((char*)fp->mesh.data(), (std::streamsize)(fp->mesh.size() * sizeof(double)));
        if (!f) throw std::runtime_error("truncated problem file");
        fp->rebind();
        validateTables(fp->tables); // every offset/count inside the pools (a corrupt file must not become an out-of-bounds read)
        if (options && *options) {
            SolverSettings S;
            S.applyOptions(options); // records which keys were given; every accepted key maps onto the stored transcription
            applySettingsToTranscription(S, *fp, /*onlyKeysGiven=*/true);
        }
        fp->rebind();
        if (fp->tables.prescribed && fp->tr.dynamics_mode != MOCOB200_DYNAMICS_IMPLICIT) // MocoCasADiSolver.cpp:147-151
            throw std::runtime_error("Prescribed kinematics (PositionMotion) requires implicit dynamics mode.");
        auto p = new mocob200_problem;
        p->fp = std::move(fp);
        *out = p;
        return MOCOB200_OK;
    } catch (const std::exception& e) {
        setErr(err, errlen, e.what());
        return MOCOB200_ERR_IO;
    }
}
int mocob200_problem_load(const char* path, mocob200_problem** out, char* err, int errlen) {
    return mocob200_problem_load_opts(path, nullptr, out, err, errlen);
}

const mocob200_tables* mocob200_problem_tables(const mocob200_problem* p) { return p ? &p->fp->tables : nullptr; }
const mocob200_transcription* mocob200_problem_transcription(const mocob200_problem* p) { return p ? &p->fp->tr : nullptr; }
void mocob200_problem_destroy(mocob200_problem* p) { delete p; }

// names as '\n'-joined strings: which = 0 states, 1 controls, 2 derivatives(aux)
// ---- guesses / solutions as MocoTrajectory tables ------------------------------------------------
namespace {
// derivative variable names in NLP order: implicit-mode accelerations ("/speed" -> "/accel", MocoCasOCProblem.cpp) then the
// implicit auxiliary derivatives
std::vector<std::string> derivativeNamesOf(const FlatProblem& fp) {
    std::vector<std::string> v;
    if (fp.tr.dynamics_mode == MOCOB200_DYNAMICS_IMPLICIT && !fp.tables.prescribed)
        for (int k = 0; k < fp.tables.nq; ++k) {
            std::string n = fp.stateNames[fp.tables.nq + k];
            const size_t at = n.rfind("/speed");
            if (at != std::string::npos) n.replace(at, 6, "/accel");
            v.push_back(n);
        }
    v.insert(v.end(), fp.derivativeNames.begin(), fp.derivativeNames.end());
    return v;
}
} // namespace

// MocoTrajectory::write (MocoTrajectory.cpp:780-860): one row per grid point, columns time | states | controls |
// [multipliers: none] | derivatives | [slacks, parameters: none], metadata num_* in the header.
// index of grid point j among the interval midpoints of a Hermite-Simpson grid (-1: a mesh point)
static int hsMid(const Layout& L, int j) { return (L.n_grid != L.n_mesh && j % 2 == 1) ? j / 2 : -1; }

int mocob200_iterate_write_sto(const mocob200_problem* p, const double* x, const char* path) {
    if (!p || !x || !path) return MOCOB200_ERR_ARG;
    const FlatProblem& fp = *p->fp;
    try {
        const Layout L = buildLayout(fp.tables, fp.tr, false);
        if ((int)fp.stateNames.size() != L.ns || (int)fp.controlNames.size() != L.nc) return MOCOB200_ERR_UNSUPPORTED; // no names (loaded .mb20)
        const std::vector<std::string> dn = derivativeNamesOf(fp);
        if ((int)dn.size() != L.nd) return MOCOB200_ERR_UNSUPPORTED;
        FILE* f = std::fopen(path, "w");
        if (!f) return MOCOB200_ERR_IO;
        if ((int)fp.multiplierNames.size() != L.nm || (L.nsl && (int)fp.slackNames.size() != L.nsl)) { std::fclose(f); return MOCOB200_ERR_UNSUPPORTED; }
        // column order of MocoTrajectory::convertToTable (MocoTrajectory.cpp:793-829): states, controls, multipliers, derivatives, slacks
        std::fprintf(f, "inDegrees=no\nnum_controls=%d\nnum_derivatives=%d\nnum_multipliers=%d\nnum_parameters=0\nnum_slacks=%d\n"
                        "num_states=%d\nDataType=double\nversion=3\nOpenSimVersion=mocob200\nendheader\ntime", L.nc, L.nd, L.nm, L.nsl, L.ns);
        for (auto& n : fp.stateNames) std::fprintf(f, "\t%s", n.c_str());
        for (auto& n : fp.controlNames) std::fprintf(f, "\t%s", n.c_str());
        for (int k = 0; k < L.nm; ++k) std::fprintf(f, "\t%s", fp.multiplierNames[k].c_str());
        for (auto& n : dn) std::fprintf(f, "\t%s", n.c_str());
        for (int k = 0; k < L.nsl; ++k) std::fprintf(f, "\t%s", fp.slackNames[k].c_str());
        std::fprintf(f, "\n");
        // Slacks live on the interval midpoints only; the reference stretches those values over linspace(t0, tf) and
        // interpolates them piecewise-linearly onto every grid time (MocoCasOCProblem.h:170-185, MocoUtilities.cpp:104-141).
        const int nInt = L.n_grid - L.n_mesh;
        auto slackAt = [&](int k, int j) {
            if (nInt == 1) return j == 0 ? x[L.off_slacks + k] : std::nan("");
            const double u = L.grid[j] * (nInt - 1);
            int i = (int)std::floor(u);
            if (i >= nInt - 1) i = nInt - 2;
            const double w = u - i;
            return (1.0 - w) * x[L.off_slacks + (size_t)i * L.nsl + k] + w * x[L.off_slacks + (size_t)(i + 1) * L.nsl + k];
        };
        for (int j = 0; j < L.n_grid; ++j) {
            std::fprintf(f, "%.17g", x[0] + (x[1] - x[0]) * L.grid[j]);
            for (int k = 0; k < L.ns; ++k) std::fprintf(f, "\t%.17g", x[L.off_states + (size_t)j * L.ns + k]);
            for (int k = 0; k < L.nc; ++k) std::fprintf(f, "\t%.17g", x[L.off_controls + (size_t)j * L.nc + k]);
            for (int k = 0; k < L.nm; ++k) std::fprintf(f, "\t%.17g", x[L.off_mult + (size_t)j * L.nm + k]);
            for (int k = 0; k < L.nd; ++k) std::fprintf(f, "\t%.17g", x[L.off_derivs + (size_t)j * L.nd + k]);
            for (int k = 0; k < L.nsl; ++k) std::fprintf(f, "\t%.17g", slackAt(k, j));
            std::fprintf(f, "\n");
        }
        std::fclose(f);
        return MOCOB200_OK;
    } catch (const std::exception&) { return MOCOB200_ERR_UNSUPPORTED; }
}

// setGuessFile + the resampling of the guess onto the solver's grid (MocoTrajectory(file) .cpp:662-760, isCompatible,
// resample :610-662 with GCVSplineSet(table, {}, min(rows - 1, 5)); CasOCTranscription.cpp:595-597). Initial and final
// time come from the first and last row; every variable of the problem must have a column of its name.
int mocob200_iterate_read_sto(const mocob200_problem* p, const char* path, double* x, char* err, int errlen) {
    if (!p || !path || !x) return MOCOB200_ERR_ARG;
    const FlatProblem& fp = *p->fp;
    try {
        const Layout L = buildLayout(fp.tables, fp.tr, false);
        if ((int)fp.stateNames.size() != L.ns || (int)fp.controlNames.size() != L.nc)
            throw std::runtime_error("the problem carries no variable names (loaded from a .mb20 file)");
        const StoTable t = readSto(path);
        const int rows = (int)t.time.size();
        if (rows < 1) throw std::runtime_error("no rows in '" + std::string(path) + "'");
        std::vector<std::string> names = fp.stateNames;
        names.insert(names.end(), fp.controlNames.begin(), fp.controlNames.end());
        const std::vector<std::string> dn = derivativeNamesOf(fp);
        names.insert(names.end(), dn.begin(), dn.end());
        if ((int)fp.multiplierNames.size() != L.nm || (L.nsl && (int)fp.slackNames.size() != L.nsl))
            throw std::runtime_error("the problem carries no multiplier / slack names");
        const int kMult = L.ns + L.nc + L.nd, kSlack = kMult + L.nm; // appended after the round-1 blocks
        names.insert(names.end(), fp.multiplierNames.begin(), fp.multiplierNames.begin() + L.nm);
        names.insert(names.end(), fp.slackNames.begin(), fp.slackNames.begin() + L.nsl);
        std::vector<int> col(names.size());
        for (size_t k = 0; k < names.size(); ++k) {
            col[k] = t.col(names[k]);
            if (col[k] < 0) throw std::runtime_error("guess is not compatible with the problem: no column '" + names[k] + "'");
        }
        const double t0 = t.time.front(), tf = t.time.back();
        x[0] = t0; x[1] = tf;
        // same time grid (a solution of this very transcription): copy; otherwise interpolate with the natural spline
        bool sameGrid = rows == L.n_grid;
        for (int j = 0; sameGrid && j < rows; ++j)
            sameGrid = std::fabs(t.time[j] - (t0 + (tf - t0) * L.grid[j])) <= 1e-12 * std::fmax(1.0, std::fabs(tf));
        for (int j = 1; j < rows; ++j)
            if (!(t.time[j] > t.time[j - 1]) && !(tf == t0)) throw std::runtime_error("guess times must increase");
        std::vector<PiecewisePoly> pps;
        if (!sameGrid && rows > 1 && tf > t0) { // one factorisation for all columns
            std::vector<const std::vector<double>*> ys;
            for (size_t k = 0; k < names.size(); ++k) ys.push_back(&t.cols[col[k]]);
            pps = fitNaturalSplines(t.time, ys, std::min(rows - 1, 5));
        }
        for (size_t k = 0; k < names.size(); ++k) {
            const std::vector<double>& y = t.cols[col[k]];
            for (int j = 0; j < L.n_grid; ++j) {
                double v;
                if (sameGrid) v = y[j];
                else if (rows == 1 || !(tf > t0)) v = y[0]; // all times equal: constant (MocoTrajectory.cpp:623-641)
                else v = evalPiecewisePoly(pps[k], t0 + (tf - t0) * L.grid[j]);
                const int kk = (int)k;
                if (kk < L.ns) x[L.off_states + (size_t)j * L.ns + kk] = v;
                else if (kk < L.ns + L.nc) x[L.off_controls + (size_t)j * L.nc + (kk - L.ns)] = v;
                else if (kk < kMult) x[L.off_derivs + (size_t)j * L.nd + (kk - L.ns - L.nc)] = v;
                else if (kk < kSlack) x[L.off_mult + (size_t)j * L.nm + (kk - kMult)] = v;
                else if (hsMid(L, j) >= 0) // a slack column given on every grid time keeps its midpoint entries (CasOCTranscription.cpp:601-618)
                    x[L.off_slacks + (size_t)hsMid(L, j) * L.nsl + (kk - kSlack)] = std::isnan(v) ? 0.0 : v;
            }
        }
        return MOCOB200_OK;
    } catch (const std::exception& e) {
        setErr(err, errlen, e.what());
        return MOCOB200_ERR_IO;
    }
}

int mocob200_problem_names(const mocob200_problem* p, int which, char* buf, int buflen) {
    if (!p || !buf) return MOCOB200_ERR_ARG;
    const auto& v = which == 0 ? p->fp->stateNames : which == 1 ? p->fp->controlNames : which == 2 ? p->fp->derivativeNames
                    : which == 3 ? p->fp->multiplierNames : p->fp->slackNames;
    std::string s;
    for (auto& n : v) { s += n; s += '\n'; }
    std::snprintf(buf, (size_t)buflen, "%s", s.c_str());
    return MOCOB200_OK;
}

static void fillDims(const Layout& L, mocob200_dims* d) {
    std::memset(d, 0, sizeof *d);
    d->n_grid = L.n_grid; d->n_mesh = L.n_mesh; d->n_var = L.n_var; d->n_con = L.n_con; d->nnz_jac = L.nnz;
    d->ns = L.ns; d->nc = L.nc; d->nd = L.nd; d->n_in = L.n_in; d->n_out = L.n_out;
    d->nm = L.nm; d->nsl = L.nsl; d->n_kc = L.n_kc;
    d->n_dyn_out = L.n_dyn_out; d->n_integrand = L.n_integrand; d->n_cost_terms = L.n_cost_terms;
    d->first_interval = 0; d->n_intervals_local = L.n_intervals; d->rank = 0; d->world = 1;
}

int mocob200_layout(const mocob200_tables* T, const mocob200_transcription* R, mocob200_dims* dims) {
    if (!T || !R || !dims) return MOCOB200_ERR_ARG;
    try {
        Layout L = buildLayout(*T, *R, true);
        fillDims(L, dims);
        return MOCOB200_OK;
    } catch (const std::exception&) { return MOCOB200_ERR_UNSUPPORTED; }
}

int mocob200_layout_bounds(const mocob200_tables* T, const mocob200_transcription* R, double* xlo, double* xhi, double* glo, double* ghi) {
    if (!T || !R) return MOCOB200_ERR_ARG;
    try {
        Layout L = buildLayout(*T, *R, false);
        layoutBounds(*T, *R, L, xlo, xhi, glo, ghi);
        return MOCOB200_OK;
    } catch (const std::exception&) { return MOCOB200_ERR_UNSUPPORTED; }
}

int mocob200_layout_structure(const mocob200_tables* T, const mocob200_transcription* R, int* colptr, int* rowidx) {
    if (!T || !R || !colptr || !rowidx) return MOCOB200_ERR_ARG;
    try {
        Layout L = buildLayout(*T, *R, true);
        std::memcpy(colptr, L.colptr.data(), sizeof(int) * (L.n_var + 1));
        std::memcpy(rowidx, L.rowidx.data(), sizeof(int) * L.nnz);
        return MOCOB200_OK;
    } catch (const std::exception&) { return MOCOB200_ERR_UNSUPPORTED; }
}

int mocob200_make_iterate(const mocob200_tables* T, const mocob200_transcription* R, long seed, double spread, double* x) {
    if (!T || !R || !x) return MOCOB200_ERR_ARG;
    try {
        Layout L = buildLayout(*T, *R, false);
        std::vector<double> lo(L.n_var), hi(L.n_var);
        layoutBounds(*T, *R, L, lo.data(), hi.data(), nullptr, nullptr);
        std::mt19937_64 rng((unsigned long long)(seed < 0 ? 0 : seed));
        std::uniform_real_distribution<double> U(-1.0, 1.0);
        for (int i = 0; i < L.n_var; ++i) {
            double l = lo[i], h = hi[i], mid;
            if (!std::isinf(l) && !std::isinf(h)) mid = 0.5 * (l + h);
            else if (!std::isinf(l)) mid = l;
            else if (!std::isinf(h)) mid = h;
            else mid = 0;
            double v = mid;
            if (seed >= 0) {
                const double half = (!std::isinf(l) && !std::isinf(h)) ? 0.5 * (h - l) : 1.0;
                v = mid + spread * half * U(rng);
                if (v < l) v = l;
                if (v > h) v = h;
            }
            x[i] = v;
        }
        return MOCOB200_OK;
    } catch (const std::exception&) { return MOCOB200_ERR_UNSUPPORTED; }
}

static int g_min_points_per_rank = 96;
void mocob200_set_min_points_per_rank(int n) { g_min_points_per_rank = n > 0 ? n : 96; }

int mocob200_effective_world(int n_grid, int world) {
    // a rank is only worth its launch + exchange latency when it gets a hundred or so grid points: BASELINE config 5 (1601 points)
    // spreads over 8 GPUs, config 3 (101 points, a 0.1 ms latency-bound step) stays on one. A fixed rule, the same on every rank.
    const int min_points = g_min_points_per_rank;
    int eff = n_grid / min_points;
    if (eff < 1) eff = 1;
    return eff < world ? eff : world;
}

int mocob200_shard_range(int n_grid, int rank, int world, int* p0, int* p1) {
    if (n_grid < 0 || world < 1 || rank < 0 || rank >= world || !p0 || !p1) return MOCOB200_ERR_ARG;
    const int eff = mocob200_effective_world(n_grid, world);
    if (rank >= eff) { *p0 = *p1 = n_grid; return MOCOB200_OK; } // idle rank: small problems are not worth splitting
    *p0 = (int)((long long)n_grid * rank / eff);
    *p1 = (int)((long long)n_grid * (rank + 1) / eff);
    return MOCOB200_OK;
}

} // extern "C"
