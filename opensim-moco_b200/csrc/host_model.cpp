#include "host_model.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "xml_reader.h"

namespace mocob200 {

// ---------------------------------------------------------------- small math
Mat3 mul(const Mat3& a, const Mat3& b) {
    Mat3 c;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int k = 0; k < 3; ++k) s += a.m[3 * i + k] * b.m[3 * k + j];
            c.m[3 * i + j] = s;
        }
    return c;
}
Vec3 mul(const Mat3& a, const Vec3& x) {
    Vec3 y;
    for (int i = 0; i < 3; ++i) y.v[i] = a.m[3 * i] * x.v[0] + a.m[3 * i + 1] * x.v[1] + a.m[3 * i + 2] * x.v[2];
    return y;
}
Xform compose(const Xform& a, const Xform& b) {
    Xform c;
    c.R = mul(a.R, b.R);
    Vec3 t = mul(a.R, b.p);
    for (int i = 0; i < 3; ++i) c.p.v[i] = t.v[i] + a.p.v[i];
    return c;
}
Xform inverse(const Xform& a) {
    Xform c;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) c.R.m[3 * i + j] = a.R.m[3 * j + i];
    Vec3 t = mul(c.R, a.p);
    for (int i = 0; i < 3; ++i) c.p.v[i] = -t.v[i];
    return c;
}
Mat3 rotXYZ(double a, double b, double c) {
    // body-fixed X-Y-Z sequence: R = Rx(a) Ry(b) Rz(c)
    const double ca = std::cos(a), sa = std::sin(a), cb = std::cos(b), sb = std::sin(b), cc = std::cos(c), sc = std::sin(c);
    Mat3 Rx, Ry, Rz;
    double rx[9] = {1, 0, 0, 0, ca, -sa, 0, sa, ca};
    double ry[9] = {cb, 0, sb, 0, 1, 0, -sb, 0, cb};
    double rz[9] = {cc, -sc, 0, sc, cc, 0, 0, 0, 1};
    std::memcpy(Rx.m, rx, sizeof rx);
    std::memcpy(Ry.m, ry, sizeof ry);
    std::memcpy(Rz.m, rz, sizeof rz);
    return mul(mul(Rx, Ry), Rz);
}

int ModelSpec::bodyIndex(const std::string& n) const {
    if (n == "ground") return -1;
    for (size_t i = 0; i < bodies.size(); ++i)
        if (bodies[i].name == n) return (int)i;
    throw std::runtime_error("no body named '" + n + "'");
}

// ---------------------------------------------------------------- options
static bool boolOpt(const std::string& v) { return v == "1" || v == "true" || v == "True"; }
static Bounds boundsOpt(const std::string& v) {
    auto d = parseDoubles(v);
    if (d.size() != 2) throw std::runtime_error("expected two numbers 'lo hi', got '" + v + "'");
    return Bounds(d[0], d[1]);
}
void SolverSettings::applyOptions(const std::string& opts) {
    std::stringstream ss(opts);
    std::string kv;
    while (std::getline(ss, kv, ';')) {
        kv = trim(kv);
        if (kv.empty()) continue;
        size_t eq = kv.find('=');
        if (eq == std::string::npos) throw std::runtime_error("option without '=': " + kv);
        std::string k = trim(kv.substr(0, eq)), v = trim(kv.substr(eq + 1));
        if (k == "scheme") k = "transcription_scheme";
        if (k == "dynamics_mode") k = "multibody_dynamics_mode";
        if (k == "fd_scheme") k = "optim_finite_difference_scheme";
        if (k == "num_mesh_intervals") num_mesh_intervals = std::stoi(v);
        else if (k == "transcription_scheme") transcription_scheme = v;
        else if (k == "interpolate_control_midpoints") interpolate_control_midpoints = boolOpt(v);
        else if (k == "multibody_dynamics_mode") multibody_dynamics_mode = v;
        else if (k == "optim_finite_difference_scheme") optim_finite_difference_scheme = v;
        else if (k == "minimize_implicit_multibody_accelerations") minimize_implicit_multibody_accelerations = boolOpt(v);
        else if (k == "minimize_implicit_auxiliary_derivatives") minimize_implicit_auxiliary_derivatives = boolOpt(v);
        else if (k == "implicit_multibody_accelerations_weight") implicit_multibody_accelerations_weight = std::stod(v);
        else if (k == "implicit_auxiliary_derivatives_weight") implicit_auxiliary_derivatives_weight = std::stod(v);
        else if (k == "implicit_multibody_acceleration_bounds") implicit_multibody_acceleration_bounds = boundsOpt(v);
        else if (k == "implicit_auxiliary_derivative_bounds") implicit_auxiliary_derivative_bounds = boundsOpt(v);
        else if (k == "enforce_constraint_derivatives") enforce_constraint_derivatives = boolOpt(v);
        else if (k == "minimize_lagrange_multipliers") minimize_lagrange_multipliers = boolOpt(v);
        else if (k == "lagrange_multiplier_weight") lagrange_multiplier_weight = std::stod(v);
        else if (k == "velocity_correction_bounds") velocity_correction_bounds = boundsOpt(v);
        else if (k == "mesh") { mesh = parseDoubles(v); }
        else throw std::runtime_error("unknown solver option '" + k + "'");
        keysGiven.push_back(k);
    }
}
bool SolverSettings::given(const std::string& k) const { return std::find(keysGiven.begin(), keysGiven.end(), k) != keysGiven.end(); }

void applySettingsToTranscription(const SolverSettings& S, FlatProblem& fp, bool only) {
    mocob200_transcription& R = fp.tr;
    auto want = [&](const char* k) { return !only || S.given(k); };
    if (want("transcription_scheme")) {
        if (S.transcription_scheme == "trapezoidal") R.scheme = MOCOB200_SCHEME_TRAPEZOIDAL;
        else if (S.transcription_scheme == "hermite-simpson") R.scheme = MOCOB200_SCHEME_HERMITE_SIMPSON;
        else throw std::runtime_error("unknown transcription_scheme '" + S.transcription_scheme + "'");
    }
    if (want("multibody_dynamics_mode")) {
        if (S.multibody_dynamics_mode == "explicit") R.dynamics_mode = MOCOB200_DYNAMICS_EXPLICIT;
        else if (S.multibody_dynamics_mode == "implicit") R.dynamics_mode = MOCOB200_DYNAMICS_IMPLICIT;
        else throw std::runtime_error("unknown multibody_dynamics_mode '" + S.multibody_dynamics_mode + "'");
    }
    if (want("optim_finite_difference_scheme")) {
        if (S.optim_finite_difference_scheme == "central") R.fd_scheme = MOCOB200_FD_CENTRAL;
        else if (S.optim_finite_difference_scheme == "forward") R.fd_scheme = MOCOB200_FD_FORWARD;
        else if (S.optim_finite_difference_scheme == "backward") R.fd_scheme = MOCOB200_FD_BACKWARD;
        else throw std::runtime_error("unknown optim_finite_difference_scheme");
    }
    if (!only && S.optim_sparsity_detection != "none")
        throw std::runtime_error("optim_sparsity_detection is a property of the evaluator handle (mocob200_set_sparsity_detection)");
    if (!S.mesh.empty() && want("mesh")) {
        // validity checks of MocoCasADiSolver.cpp:192-209
        if (S.mesh.front() != 0.0 || S.mesh.back() != 1.0) throw std::runtime_error("mesh must start at 0 and end at 1");
        for (size_t i = 1; i < S.mesh.size(); ++i)
            if (!(S.mesh[i] > S.mesh[i - 1])) throw std::runtime_error("mesh must be strictly increasing");
        fp.mesh = S.mesh;
    } else if (want("num_mesh_intervals")) {
        if (S.num_mesh_intervals < 1) throw std::runtime_error("num_mesh_intervals must be >= 1");
        fp.mesh.clear();
        for (int i = 0; i <= S.num_mesh_intervals; ++i) fp.mesh.push_back(i / (double)S.num_mesh_intervals);
    }
    if (want("interpolate_control_midpoints")) R.interpolate_control_midpoints = S.interpolate_control_midpoints ? 1 : 0;
    if (want("implicit_multibody_acceleration_bounds")) {
        R.accel_bounds[0] = S.implicit_multibody_acceleration_bounds.lo;
        R.accel_bounds[1] = S.implicit_multibody_acceleration_bounds.hi;
    }
    if (want("implicit_auxiliary_derivative_bounds")) {
        R.auxderiv_bounds[0] = S.implicit_auxiliary_derivative_bounds.lo;
        R.auxderiv_bounds[1] = S.implicit_auxiliary_derivative_bounds.hi;
    }
    if (want("minimize_implicit_multibody_accelerations")) R.minimize_accelerations = S.minimize_implicit_multibody_accelerations;
    if (want("implicit_multibody_accelerations_weight")) R.accelerations_weight = S.implicit_multibody_accelerations_weight;
    if (want("minimize_implicit_auxiliary_derivatives")) R.minimize_auxiliary_derivatives = S.minimize_implicit_auxiliary_derivatives;
    if (want("implicit_auxiliary_derivatives_weight")) R.auxiliary_derivatives_weight = S.implicit_auxiliary_derivatives_weight;
    if (want("enforce_constraint_derivatives")) R.enforce_constraint_derivatives = S.enforce_constraint_derivatives;
    if (want("minimize_lagrange_multipliers")) R.minimize_lagrange_multipliers = S.minimize_lagrange_multipliers;
    if (want("lagrange_multiplier_weight")) R.lagrange_multiplier_weight = S.lagrange_multiplier_weight;
    if (want("velocity_correction_bounds")) {
        R.velocity_correction_bounds[0] = S.velocity_correction_bounds.lo;
        R.velocity_correction_bounds[1] = S.velocity_correction_bounds.hi;
    }
    if (!only) { R.multiplier_bounds[0] = -1000; R.multiplier_bounds[1] = 1000; } // MocoProblem default multiplier bounds
    fp.rebind();
}

void validateTables(const mocob200_tables& T) {
    auto bad = [](const char* what) { throw std::runtime_error(std::string("corrupt problem tables: ") + what); };
    auto dblRange = [&](int off, long n, const char* what) { if (off < 0 || n < 0 || (long)off + n > T.n_dbl) bad(what); };
    auto intRange = [&](int off, long n, const char* what) { if (off < 0 || n < 0 || (long)off + n > T.n_int) bad(what); };
    if (T.nb < 0 || T.nq < 0 || T.nz < 0 || T.nc < 0 || T.nderiv_aux < 0 || T.nmus < 0 || T.npp < 0 || T.nspline < 0 || T.ncontact < 0 ||
        T.ncact < 0 || T.ngoal < 0 || T.nppoly < 0 || T.ncustom < 0 || T.nxforce < 0 || T.ncoupler < 0) bad("negative count");
    if (T.ncoupler > 0) {
        intRange(T.o_cc_i, (long)T.ncoupler * MOCOB200_CC_ISTRIDE, "couplers");
        dblRange(T.o_cc_d, (long)T.ncoupler * MOCOB200_CC_DSTRIDE, "couplers");
        for (int k = 0; k < T.ncoupler; ++k) {
            const int* ci = T.idx + T.o_cc_i + k * MOCOB200_CC_ISTRIDE;
            if (ci[0] < 0 || ci[0] > MOCOB200_FN_SIMMSPLINE || ci[1] < 0 || ci[1] >= T.nq || ci[2] < 0 || ci[2] >= T.nq || ci[1] == ci[2] ||
                (ci[0] == MOCOB200_FN_SIMMSPLINE && (ci[3] < 0 || ci[3] >= T.nspline))) bad("coupler record");
        }
    }
    dblRange(T.o_body_d, (long)T.nb * MOCOB200_BODY_DSTRIDE, "bodies");
    intRange(T.o_body_i, (long)T.nb * MOCOB200_BODY_ISTRIDE, "bodies");
    dblRange(T.o_gravity, 3, "gravity");
    dblRange(T.o_pp_d, (long)T.npp * MOCOB200_PP_DSTRIDE, "path points");
    intRange(T.o_pp_i, (long)T.npp * MOCOB200_PP_ISTRIDE, "path points");
    dblRange(T.o_mus_d, (long)T.nmus * MOCOB200_MUS_DSTRIDE, "muscles");
    intRange(T.o_mus_i, (long)T.nmus * MOCOB200_MUS_ISTRIDE, "muscles");
    dblRange(T.o_contact_d, (long)T.ncontact * MOCOB200_CONTACT_DSTRIDE, "contacts");
    intRange(T.o_contact_i, (long)T.ncontact * MOCOB200_CONTACT_ISTRIDE, "contacts");
    dblRange(T.o_cact_d, (long)T.ncact * MOCOB200_CACT_DSTRIDE, "coordinate actuators");
    intRange(T.o_cact_i, (long)T.ncact * MOCOB200_CACT_ISTRIDE, "coordinate actuators");
    intRange(T.o_spline_i, (long)T.nspline * MOCOB200_SPLINE_ISTRIDE, "splines");
    intRange(T.o_ppoly_i, (long)T.nppoly * MOCOB200_PPOLY_ISTRIDE, "piecewise polynomials");
    intRange(T.o_goal_i, (long)T.ngoal * MOCOB200_GOAL_ISTRIDE, "goals");
    dblRange(T.o_bounds, 6L * ((T.prescribed ? 0 : 2 * T.nq) + T.nz + T.nc), "bounds");
    if (T.nwrapobj < 0 || T.npathwrap < 0) bad("negative count");
    if (T.npathwrap > 0) {
        dblRange(T.o_wo_d, (long)T.nwrapobj * MOCOB200_WO_DSTRIDE, "wrap objects");
        intRange(T.o_wo_i, (long)T.nwrapobj * MOCOB200_WO_ISTRIDE, "wrap objects");
        intRange(T.o_pw_i, 2L * T.nmus + (long)T.npathwrap * MOCOB200_PW_ISTRIDE, "path wraps");
        for (int w = 0; w < T.nwrapobj; ++w) {
            const int* wi = T.idx + T.o_wo_i + w * MOCOB200_WO_ISTRIDE;
            if (wi[0] < -1 || wi[0] >= T.nb || wi[1] < 0 || wi[1] > 1 || wi[2] < -1 || wi[2] > 1) bad("wrap object record");
            if (!(T.dbl[T.o_wo_d + w * MOCOB200_WO_DSTRIDE + 12] > 0)) bad("wrap cylinder radius");
        }
        for (int m = 0; m < T.nmus; ++m) {
            const int first = T.idx[T.o_pw_i + 2 * m], count = T.idx[T.o_pw_i + 2 * m + 1];
            if (first < 0 || count < 0 || first + count > T.npathwrap) bad("path wrap range of a muscle");
        }
        for (int k = 0; k < T.npathwrap; ++k)
            if (T.idx[T.o_pw_i + 2 * T.nmus + k * MOCOB200_PW_ISTRIDE] < 0 || T.idx[T.o_pw_i + 2 * T.nmus + k * MOCOB200_PW_ISTRIDE] >= T.nwrapobj)
                bad("path wrap object");
    }
    if (T.prescribed) {
        intRange(T.o_presc_i, T.nq, "prescribed kinematics");
        for (int k = 0; k < T.nq; ++k)
            if (T.idx[T.o_presc_i + k] < 0 || T.idx[T.o_presc_i + k] >= T.nppoly) bad("prescribed kinematics function");
    }
    dblRange(T.o_time_bounds, 4, "time bounds");
    dblRange(T.o_spline_data, 0, "spline data");
    dblRange(T.o_ppoly_data, 0, "piecewise-polynomial data");
    int nq = 0;
    for (int b = 0; b < T.nb; ++b) {
        const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
        if (bi[0] < -1 || bi[0] >= b) bad("body parent");
        if (bi[1] < 0 || bi[1] > MOCOB200_JOINT_CUSTOM || bi[3] < 0 || bi[3] > 6 || bi[2] != nq) bad("joint record");
        nq += bi[3];
    }
    if (nq != T.nq) bad("coordinate count");
    for (int s = 0; s < T.nspline; ++s) {
        const int* r = T.idx + T.o_spline_i + s * MOCOB200_SPLINE_ISTRIDE;
        if (r[1] < 2) bad("spline knots");
        dblRange(T.o_spline_data + r[0], 5L * r[1] + 1, "spline data");
        if (r[0] < 0) bad("spline data");
    }
    for (int s = 0; s < T.nppoly; ++s) {
        const int* r = T.idx + T.o_ppoly_i + s * MOCOB200_PPOLY_ISTRIDE;
        if (r[0] < 0 || r[1] < 1 || r[2] < 1) bad("piecewise polynomial record");
        dblRange(T.o_ppoly_data + r[0], (long)r[1] + 1 + (long)r[1] * r[2], "piecewise-polynomial data");
    }
    for (int p = 0; p < T.npp; ++p) {
        const int* pi = T.idx + T.o_pp_i + p * MOCOB200_PP_ISTRIDE;
        if (pi[0] < -1 || pi[0] >= T.nb || pi[1] < 0 || pi[1] > MOCOB200_PP_MOVING) bad("path point record");
        if (pi[1] == MOCOB200_PP_CONDITIONAL && (pi[2] < 0 || pi[2] >= T.nq)) bad("conditional path point coordinate");
        if (pi[1] == MOCOB200_PP_MOVING)
            for (int a = 0; a < 3; ++a)
                if (pi[3 + a] < 0 || pi[3 + a] >= T.nspline || pi[6] < 0 || pi[6] >= T.nq) bad("moving path point record");
    }
    for (int m = 0; m < T.nmus; ++m) {
        const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
        if (mi[0] < 0 || mi[1] < 2 || mi[0] + mi[1] > T.npp || mi[2] < 0 || mi[2] >= T.nc || mi[3] >= T.nz || mi[4] >= T.nz ||
            mi[6] >= T.nderiv_aux || mi[7] >= T.nderiv_aux) bad("muscle record");
    }
    for (int c = 0; c < T.ncontact; ++c) {
        const int* ci = T.idx + T.o_contact_i + c * MOCOB200_CONTACT_ISTRIDE;
        if (ci[0] < -1 || ci[0] >= T.nb || ci[1] < -1 || ci[1] >= T.nb) bad("contact record");
    }
    for (int a = 0; a < T.ncact; ++a) {
        const int* ai = T.idx + T.o_cact_i + a * MOCOB200_CACT_ISTRIDE;
        if (ai[0] < 0 || ai[0] >= T.nq || ai[1] < 0 || ai[1] >= T.nc) bad("coordinate actuator record");
    }
    for (int g = 0; g < T.ngoal; ++g) {
        const int* gr = T.idx + T.o_goal_i + g * MOCOB200_GOAL_ISTRIDE;
        if (gr[2] < 0 || gr[6] < 0 || gr[7] < 0) bad("goal record");
        intRange(gr[4], gr[6], "goal payload");
        dblRange(gr[5], gr[7], "goal payload");
    }
    if (T.ncustom > 0) {
        dblRange(T.o_cj_d, (long)T.ncustom * MOCOB200_CJ_DSTRIDE, "custom joints");
        intRange(T.o_cj_i, (long)T.ncustom * MOCOB200_CJ_ISTRIDE, "custom joints");
        intRange(T.o_body_cj, T.nb, "custom joint map");
        for (int b = 0; b < T.nb; ++b) {
            const int cj = T.idx[T.o_body_cj + b];
            const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
            if ((bi[1] == MOCOB200_JOINT_CUSTOM) != (cj >= 0) || cj >= T.ncustom) bad("custom joint map");
            if (cj < 0) continue;
            const int* ci = T.idx + T.o_cj_i + cj * MOCOB200_CJ_ISTRIDE;
            for (int a = 0; a < 6; ++a) {
                if (ci[3 * a] < 0 || ci[3 * a] > MOCOB200_FN_SIMMSPLINE) bad("custom joint function kind");
                if (ci[3 * a] != MOCOB200_FN_CONSTANT && (ci[3 * a + 1] < bi[2] || ci[3 * a + 1] >= bi[2] + bi[3])) bad("custom joint coordinate");
                if (ci[3 * a] == MOCOB200_FN_SIMMSPLINE && (ci[3 * a + 2] < 0 || ci[3 * a + 2] >= T.nspline)) bad("custom joint spline");
            }
        }
    } else {
        for (int b = 0; b < T.nb; ++b)
            if (T.idx[T.o_body_i + b * MOCOB200_BODY_ISTRIDE + 1] == MOCOB200_JOINT_CUSTOM) bad("custom joint without a record");
    }
    if (T.nxforce > 0) {
        intRange(T.o_xforce_i, (long)T.nxforce * MOCOB200_XFORCE_ISTRIDE, "external forces");
        for (int f = 0; f < T.nxforce; ++f) {
            const int* xi = T.idx + T.o_xforce_i + f * MOCOB200_XFORCE_ISTRIDE;
            if (xi[0] < 0 || xi[0] >= T.nb || xi[1] < -1 || xi[1] >= T.nb || xi[2] < -1 || xi[2] >= T.nb) bad("external force bodies");
            for (int k = 0; k < 9; ++k) if (xi[3 + k] < -1 || xi[3 + k] >= T.nppoly) bad("external force function");
        }
    }
}

// ---------------------------------------------------------------- .sto reader
int StoTable::col(const std::string& l) const {
    for (size_t i = 0; i < labels.size(); ++i)
        if (labels[i] == l) return (int)i;
    return -1;
}
StoTable readSto(const std::string& path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open '" + path + "'");
    std::string line;
    bool header = true;
    StoTable t;
    while (std::getline(f, line)) {
        if (header) {
            const std::string hl = trim(line);
            if (hl.rfind("endheader", 0) == 0) header = false;
            else if (hl.rfind("inDegrees", 0) == 0) t.inDegrees = hl.find("yes") != std::string::npos;
            continue;
        }
        std::string tl = trim(line);
        if (tl.empty()) continue;
        if (t.labels.empty() && t.time.empty() && !(std::isdigit((unsigned char)tl[0]) || tl[0] == '-' || tl[0] == '.')) {
            std::stringstream ss(tl);
            std::string tok;
            bool first = true;
            while (std::getline(ss, tok, '\t')) {
                tok = trim(tok);
                if (tok.empty()) continue;
                if (first) { first = false; continue; } // "time"
                t.labels.push_back(tok);
            }
            t.cols.resize(t.labels.size());
            continue;
        }
        std::vector<double> v = parseDoubles(tl);
        if (v.size() != t.labels.size() + 1) throw std::runtime_error("sto row width mismatch in " + path);
        t.time.push_back(v[0]);
        for (size_t i = 0; i < t.labels.size(); ++i) t.cols[i].push_back(v[i + 1]);
    }
    return t;
}

// ---------------------------------------------------------------- natural spline fit
// Interpolating natural spline of odd degree d = 2m-1 (GCVSPL with zero smoothing,
// opensim-core/OpenSim/Common/GCVSpline.cpp + gcvspl.c), as a piecewise polynomial.
static void solveDense(std::vector<double>& A, std::vector<double>& b, int n, int nrhs = 1) {
    if (nrhs != 1) { // same elimination, right-hand sides stored row-major b[n][nrhs]
        for (int c = 0; c < n; ++c) {
            int piv = c;
            double best = std::fabs(A[(size_t)c * n + c]);
            for (int r = c + 1; r < n; ++r)
                if (std::fabs(A[(size_t)r * n + c]) > best) { best = std::fabs(A[(size_t)r * n + c]); piv = r; }
            if (best == 0) throw std::runtime_error("fitNaturalSpline: singular system");
            if (piv != c) {
                for (int k = 0; k < n; ++k) std::swap(A[(size_t)piv * n + k], A[(size_t)c * n + k]);
                for (int k = 0; k < nrhs; ++k) std::swap(b[(size_t)piv * nrhs + k], b[(size_t)c * nrhs + k]);
            }
            const double inv = 1.0 / A[(size_t)c * n + c];
            for (int r = c + 1; r < n; ++r) {
                const double f = A[(size_t)r * n + c] * inv;
                if (f == 0) continue;
                for (int k = c; k < n; ++k) A[(size_t)r * n + k] -= f * A[(size_t)c * n + k];
                for (int k = 0; k < nrhs; ++k) b[(size_t)r * nrhs + k] -= f * b[(size_t)c * nrhs + k];
            }
        }
        for (int r = n - 1; r >= 0; --r)
            for (int q = 0; q < nrhs; ++q) {
                double s = b[(size_t)r * nrhs + q];
                for (int k = r + 1; k < n; ++k) s -= A[(size_t)r * n + k] * b[(size_t)k * nrhs + q];
                b[(size_t)r * nrhs + q] = s / A[(size_t)r * n + r];
            }
        return;
    }
    for (int c = 0; c < n; ++c) {
        int piv = c;
        double best = std::fabs(A[(size_t)c * n + c]);
        for (int r = c + 1; r < n; ++r)
            if (std::fabs(A[(size_t)r * n + c]) > best) { best = std::fabs(A[(size_t)r * n + c]); piv = r; }
        if (best == 0) throw std::runtime_error("fitNaturalSpline: singular system");
        if (piv != c) {
            for (int k = 0; k < n; ++k) std::swap(A[(size_t)piv * n + k], A[(size_t)c * n + k]);
            std::swap(b[piv], b[c]);
        }
        const double inv = 1.0 / A[(size_t)c * n + c];
        for (int r = c + 1; r < n; ++r) {
            const double f = A[(size_t)r * n + c] * inv;
            if (f == 0) continue;
            for (int k = c; k < n; ++k) A[(size_t)r * n + k] -= f * A[(size_t)c * n + k];
            b[r] -= f * b[c];
        }
    }
    for (int r = n - 1; r >= 0; --r) {
        double s = b[r];
        for (int k = r + 1; k < n; ++k) s -= A[(size_t)r * n + k] * b[k];
        b[r] = s / A[(size_t)r * n + r];
    }
}

// Banded LU with partial pivoting (the layout of LAPACK's dgbsv: kl sub-diagonals, ku super-diagonals, kl more for the
// fill-in of the row exchanges). AB[i][j - i + kl] holds A(i, j) for i - kl <= j <= i + ku + kl. b is [n][nrhs].
static void solveBanded(std::vector<double>& AB, std::vector<double>& b, int n, int kl, int ku, int nrhs) {
    const int w = 2 * kl + ku + 1;
    auto at = [&](int i, int j) -> double& { return AB[(size_t)i * w + (j - i + kl)]; };
    for (int c = 0; c < n; ++c) {
        const int rmax = std::min(n - 1, c + kl), jmax = std::min(n - 1, c + ku + kl);
        int piv = c;
        double best = std::fabs(at(c, c));
        for (int r = c + 1; r <= rmax; ++r)
            if (std::fabs(at(r, c)) > best) { best = std::fabs(at(r, c)); piv = r; }
        if (best == 0) throw std::runtime_error("fitNaturalSpline: singular system");
        if (piv != c) {
            for (int j = c; j <= jmax; ++j) std::swap(at(piv, j), at(c, j));
            for (int q = 0; q < nrhs; ++q) std::swap(b[(size_t)piv * nrhs + q], b[(size_t)c * nrhs + q]);
        }
        const double inv = 1.0 / at(c, c);
        for (int r = c + 1; r <= rmax; ++r) {
            const double f = at(r, c) * inv;
            if (f == 0) continue;
            for (int j = c; j <= jmax; ++j) at(r, j) -= f * at(c, j);
            for (int q = 0; q < nrhs; ++q) b[(size_t)r * nrhs + q] -= f * b[(size_t)c * nrhs + q];
        }
    }
    for (int r = n - 1; r >= 0; --r) {
        const int jmax = std::min(n - 1, r + ku + kl);
        for (int q = 0; q < nrhs; ++q) {
            double v = b[(size_t)r * nrhs + q];
            for (int j = r + 1; j <= jmax; ++j) v -= at(r, j) * b[(size_t)j * nrhs + q];
            b[(size_t)r * nrhs + q] = v / at(r, r);
        }
    }
}

PiecewisePoly fitNaturalSpline(const std::vector<double>& t, const std::vector<double>& y, int degree) {
    return fitNaturalSplines(t, {&y}, degree)[0];
}
// The same equations as below, ordered piece by piece so that the system is banded (an equation couples one piece with
// the next one at most): O(n) memory and time instead of the dense O(n^2) / O(n^3). Used for long tables only, so that the
// short ones (reference data of the examples, fixtures) keep the arithmetic their golden files were written with.
static int splineDenseMaxUnknowns() { // MOCOB200_SPLINE_DENSE_MAX=0 sends every fit through the banded solver (A-B tests)
    const char* e = std::getenv("MOCOB200_SPLINE_DENSE_MAX");
    return e ? std::atoi(e) : 2400; // 400 rows of degree 5
}
static std::vector<PiecewisePoly> fitNaturalSplinesBanded(const std::vector<double>& t, const std::vector<const std::vector<double>*>& ys, int d) {
    const int n = (int)t.size(), nrhs = (int)ys.size();
    const int m = (d + 1) / 2, np = n - 1, nc = d + 1, N = np * nc;
    struct Term { int row, col; double v; };
    std::vector<Term> terms;
    std::vector<double> b((size_t)N * nrhs, 0.0);
    auto dcoef = [](int j, int r) {
        double c = 1;
        for (int q = 0; q < r; ++q) c *= (j - q);
        return c;
    };
    int row = 0;
    for (int r = m; r <= d - 1; ++r) terms.push_back({row++, r, dcoef(r, r)}); // natural end conditions at t[0]
    for (int k = 0; k < np; ++k) {
        terms.push_back({row, k * nc, 1.0}); // interpolation at both ends of the piece
        for (int q = 0; q < nrhs; ++q) b[(size_t)row * nrhs + q] = (*ys[q])[k];
        ++row;
        for (int j = 0; j < nc; ++j) terms.push_back({row, k * nc + j, 1.0});
        for (int q = 0; q < nrhs; ++q) b[(size_t)row * nrhs + q] = (*ys[q])[k + 1];
        ++row;
        if (k + 1 < np) { // continuity of derivatives 1..d-1 with the next piece
            const double hk = t[k + 1] - t[k], hn = t[k + 2] - t[k + 1];
            for (int r = 1; r <= d - 1; ++r) {
                for (int j = r; j < nc; ++j) terms.push_back({row, k * nc + j, dcoef(j, r) / std::pow(hk, r)});
                terms.push_back({row, (k + 1) * nc + r, -dcoef(r, r) / std::pow(hn, r)});
                ++row;
            }
        } else { // natural end conditions at t[n-1]
            for (int r = m; r <= d - 1; ++r) {
                for (int j = r; j < nc; ++j) terms.push_back({row, k * nc + j, dcoef(j, r)});
                ++row;
            }
        }
    }
    if (row != N) throw std::runtime_error("fitNaturalSpline: internal row count");
    int kl = 0, ku = 0;
    for (const Term& e : terms) { kl = std::max(kl, e.row - e.col); ku = std::max(ku, e.col - e.row); }
    const int w = 2 * kl + ku + 1;
    std::vector<double> AB((size_t)N * w, 0.0);
    for (const Term& e : terms) AB[(size_t)e.row * w + (e.col - e.row + kl)] = e.v;
    solveBanded(AB, b, N, kl, ku, nrhs);
    std::vector<PiecewisePoly> out(nrhs);
    for (int q = 0; q < nrhs; ++q) {
        PiecewisePoly& pp = out[q];
        pp.order = nc;
        pp.breaks = t;
        pp.coef.resize((size_t)N);
        for (int k = 0; k < np; ++k) {
            const double hk = t[k + 1] - t[k];
            for (int j = 0; j < nc; ++j) pp.coef[(size_t)k * nc + j] = b[(size_t)(k * nc + j) * nrhs + q] / std::pow(hk, j);
        }
        pp.isSpline = true;
    }
    return out;
}
// several data columns over the same abscissae: one factorisation of the (data-independent) system
std::vector<PiecewisePoly> fitNaturalSplines(const std::vector<double>& t, const std::vector<const std::vector<double>*>& ys, int degree) {
    const int n = (int)t.size(), nrhs = (int)ys.size();
    if (nrhs < 1) return {};
    for (auto* y : ys)
        if (n < 2 || (int)y->size() != n) throw std::runtime_error("fitNaturalSpline: need >= 2 samples");
    int d = degree;
    if (d % 2 == 0) d -= 1;
    while (d > 1 && n < d + 1) d -= 2; // GCVSPL needs n >= 2m
    const int m = (d + 1) / 2, np = n - 1, nc = d + 1, N = np * nc;
    // unknowns: coefficient j of piece k in the local coordinate s = (t - t_k)/h_k in [0,1]
    if (N > splineDenseMaxUnknowns()) return fitNaturalSplinesBanded(t, ys, d);
    std::vector<double> A((size_t)N * N, 0.0), b((size_t)N * nrhs, 0.0);
    auto dcoef = [](int j, int r) { // d^r/ds^r s^j at s = 1 -> j!/(j-r)!
        double c = 1;
        for (int q = 0; q < r; ++q) c *= (j - q);
        return c;
    };
    int row = 0;
    for (int k = 0; k < np; ++k) { // interpolation at both ends of each piece
        A[(size_t)row * N + k * nc + 0] = 1;
        for (int q = 0; q < nrhs; ++q) b[(size_t)row * nrhs + q] = (*ys[q])[k];
        ++row;
        for (int j = 0; j < nc; ++j) A[(size_t)row * N + k * nc + j] = 1;
        for (int q = 0; q < nrhs; ++q) b[(size_t)row * nrhs + q] = (*ys[q])[k + 1];
        ++row;
    }
    for (int k = 0; k + 1 < np; ++k) { // continuity of derivatives 1..d-1 at interior knots
        const double hk = t[k + 1] - t[k], hn = t[k + 2] - t[k + 1];
        for (int r = 1; r <= d - 1; ++r) {
            for (int j = r; j < nc; ++j) A[(size_t)row * N + k * nc + j] = dcoef(j, r) / std::pow(hk, r);
            A[(size_t)row * N + (k + 1) * nc + r] = -dcoef(r, r) / std::pow(hn, r);
            ++row;
        }
    }
    for (int r = m; r <= d - 1; ++r) { // natural end conditions
        A[(size_t)row * N + 0 * nc + r] = dcoef(r, r);
        ++row;
        for (int j = r; j < nc; ++j) A[(size_t)row * N + (np - 1) * nc + j] = dcoef(j, r);
        ++row;
    }
    if (row != N) throw std::runtime_error("fitNaturalSpline: internal row count");
    solveDense(A, b, N, nrhs);
    std::vector<PiecewisePoly> out(nrhs);
    for (int q = 0; q < nrhs; ++q) {
        PiecewisePoly& pp = out[q];
        pp.order = nc;
        pp.breaks = t;
        pp.coef.resize((size_t)N);
        for (int k = 0; k < np; ++k) {
            const double hk = t[k + 1] - t[k];
            for (int j = 0; j < nc; ++j) pp.coef[(size_t)k * nc + j] = b[(size_t)(k * nc + j) * nrhs + q] / std::pow(hk, j);
        }
        pp.isSpline = true;
    }
    return out;
}

PiecewisePoly constantFunction(double v) {
    PiecewisePoly pp;
    pp.breaks = {0.0, 1.0};
    pp.coef = {v};
    pp.order = 1;
    return pp;
}
// y[k] + (t - x[k]) * b[k], end slopes extended beyond the first/last point (PiecewiseLinearFunction.cpp:429-463)
PiecewisePoly piecewiseLinearFunction(const std::vector<double>& x, const std::vector<double>& y) {
    if (x.size() < 2 || x.size() != y.size()) throw std::runtime_error("piecewiseLinearFunction: need >= 2 points");
    PiecewisePoly pp;
    pp.breaks = x;
    pp.order = 2;
    for (size_t k = 0; k + 1 < x.size(); ++k) {
        if (!(x[k + 1] > x[k])) throw std::runtime_error("piecewiseLinearFunction: x must increase");
        pp.coef.push_back(y[k]);
        pp.coef.push_back((y[k + 1] - y[k]) / (x[k + 1] - x[k]));
    }
    return pp;
}

double evalPiecewisePoly(const PiecewisePoly& pp, double t, int deriv) {
    const int np = (int)pp.breaks.size() - 1;
    int k = 0;
    if (t >= pp.breaks[np]) k = np - 1;
    else if (t > pp.breaks[0]) {
        int lo = 0, hi = np;
        while (hi - lo > 1) {
            int mid = (lo + hi) / 2;
            if (t >= pp.breaks[mid]) lo = mid; else hi = mid;
        }
        k = lo;
    }
    const double s = t - pp.breaks[k];
    double v = 0;
    for (int j = pp.order - 1; j >= deriv; --j) {
        double c = pp.coef[(size_t)k * pp.order + j];
        for (int q = 0; q < deriv; ++q) c *= (j - q);
        v = v * s + c;
    }
    return v;
}

// ---------------------------------------------------------------- flatten
void FlatProblem::rebind() {
    tables.dbl = dbl.data();
    tables.idx = idx.data();
    tables.n_dbl = (int)dbl.size();
    tables.n_int = (int)idx.size();
    tr.mesh = mesh.data();
    tr.n_mesh = (int)mesh.size();
}

namespace {
struct Pools {
    std::vector<double>& d;
    std::vector<int>& i;
    int dmark() const { return (int)d.size(); }
    int imark() const { return (int)i.size(); }
};
void pushX(std::vector<double>& d, const Xform& X) {
    for (int k = 0; k < 9; ++k) d.push_back(X.R.m[k]);
    for (int k = 0; k < 3; ++k) d.push_back(X.p.v[k]);
}
// Forsythe–Malcolm–Moler cubic spline (the algorithm SimmSpline::calcCoefficients uses,
// opensim-core/OpenSim/Common/SimmSpline.cpp:321-416): end conditions match the third
// derivative of the cubic through the first/last four knots.
void fmmSpline(const std::vector<double>& x, const std::vector<double>& y, std::vector<double>& b,
               std::vector<double>& c, std::vector<double>& d) {
    const int n = (int)x.size();
    const double tiny = 1e-7; // TINY_NUMBER guard of the reference (SimmMacros.h)
    b.assign(n, 0); c.assign(n, 0); d.assign(n, 0);
    auto gap = [&](int hi, int lo) { return std::max(tiny, x[hi] - x[lo]); };
    if (n < 2) return;
    if (n == 2) {
        b[0] = b[1] = (y[1] - y[0]) / gap(1, 0);
        return;
    }
    const int last = n - 1;
    // tridiagonal system: b = diagonal, d = off-diagonal, c = right-hand side
    d[0] = gap(1, 0);
    c[1] = (y[1] - y[0]) / d[0];
    for (int i = 1; i < last; ++i) {
        d[i] = gap(i + 1, i);
        b[i] = 2.0 * (d[i - 1] + d[i]);
        c[i + 1] = (y[i + 1] - y[i]) / d[i];
        c[i] = c[i + 1] - c[i];
    }
    b[0] = -d[0];
    b[last] = -d[last - 1];
    c[0] = 0; c[last] = 0;
    if (n > 3) {
        c[0] = c[2] / gap(3, 1) - c[1] / gap(2, 0);
        c[last] = c[last - 1] / gap(last, last - 2) - c[last - 2] / gap(last - 1, last - 3);
        c[0] = c[0] * d[0] * d[0] / gap(3, 0);
        c[last] = -c[last] * d[last - 1] * d[last - 1] / gap(last, last - 3);
    }
    for (int i = 1; i < n; ++i) { // forward elimination
        const double t = d[i - 1] / b[i - 1];
        b[i] -= t * d[i - 1];
        c[i] -= t * c[i - 1];
    }
    c[last] /= b[last]; // back substitution
    for (int i = last - 1; i >= 0; --i) c[i] = (c[i] - d[i] * c[i + 1]) / b[i];
    b[last] = (y[last] - y[last - 1]) / d[last - 1] + d[last - 1] * (c[last - 1] + 2.0 * c[last]);
    for (int i = 0; i < last; ++i) {
        b[i] = (y[i + 1] - y[i]) / d[i] - d[i] * (c[i + 1] + 2.0 * c[i]);
        d[i] = (c[i + 1] - c[i]) / d[i];
        c[i] *= 3.0;
    }
    c[last] *= 3.0;
    d[last] = d[last - 1];
}
} // namespace

std::unique_ptr<FlatProblem> flatten(const Problem& P, const SolverSettings& S) {
    const ModelSpec& M = P.model;
    auto fp = std::make_unique<FlatProblem>();
    std::vector<double>& D = fp->dbl;
    std::vector<int>& I = fp->idx;
    mocob200_tables& T = fp->tables;

    // ---- multibody tree in Simbody's mobilizer order: level by level from ground, joints in
    // model order within a level (MultibodyGraphMaker::growTree as driven by Model::createMultibodyTree,
    // opensim-core/OpenSim/Simulation/Model/Model.cpp:711-808).
    const int nBodies = (int)M.bodies.size();
    std::vector<int> level(nBodies, -1), order, jointOfBody(nBodies, -1);
    std::vector<char> used(M.joints.size(), 0);
    for (int lev = 1;; ++lev) {
        bool any = false;
        for (size_t j = 0; j < M.joints.size(); ++j) {
            if (used[j]) continue;
            const JointSpec& J = M.joints[j];
            const int pb = J.parent.body, cb = J.child.body;
            if (cb < 0) throw std::runtime_error("joint '" + J.name + "': reversed joints (child = ground) are unsupported");
            const int plev = pb < 0 ? 0 : level[pb];
            if (plev < 0 || plev != lev - 1) continue;
            if (level[cb] >= 0) throw std::runtime_error("joint '" + J.name + "' closes a loop (unsupported)");
            level[cb] = lev;
            jointOfBody[cb] = (int)j;
            used[j] = 1;
            order.push_back(cb);
            any = true;
        }
        if (!any) break;
    }
    if ((int)order.size() != nBodies) throw std::runtime_error("model has bodies not connected to ground by joints");
    std::vector<int> newIndex(nBodies);
    for (int k = 0; k < nBodies; ++k) newIndex[order[k]] = k;
    auto mapBody = [&](int b) { return b < 0 ? -1 : newIndex[b]; };

    // ---- splines (SimmSpline knots + cubic coefficients, pre-scaled by the MultiplierFunction scale)
    std::vector<int> splineRec; // istride records
    std::vector<double> splineData;
    auto addSpline = [&](const SplineSpec& s) {
        const int n = (int)s.x.size();
        std::vector<double> b, c, d;
        fmmSpline(s.x, s.y, b, c, d);
        const int id = (int)splineRec.size() / MOCOB200_SPLINE_ISTRIDE;
        splineRec.push_back((int)splineData.size());
        splineRec.push_back(n);
        splineRec.push_back(0);
        // layout: x[n] y[n] b[n] c[n] d[n] scale
        splineData.insert(splineData.end(), s.x.begin(), s.x.end());
        splineData.insert(splineData.end(), s.y.begin(), s.y.end());
        splineData.insert(splineData.end(), b.begin(), b.end());
        splineData.insert(splineData.end(), c.begin(), c.end());
        splineData.insert(splineData.end(), d.begin(), d.end());
        splineData.push_back(s.scale);
        return id;
    };

    T.nb = nBodies;
    T.o_body_d = (int)D.size();
    T.o_body_i = (int)I.size();
    std::map<std::string, int> coordIndex;      // coordinate name -> q index
    std::vector<std::string> coordPath;         // "/jointset/<joint>/<coord>"
    std::vector<std::array<double, 2>> coordRange;
    std::vector<double> cjD;                    // custom-joint records
    std::vector<int> cjI, bodyCj(nBodies, -1);
    int nq = 0;
    for (int k = 0; k < nBodies; ++k) {
        const BodySpec& B = M.bodies[order[k]];
        const JointSpec& J = M.joints[jointOfBody[order[k]]];
        const int expect = J.kind == MOCOB200_JOINT_WELD ? 0 : J.kind == MOCOB200_JOINT_PLANAR ? 3 : J.kind == MOCOB200_JOINT_CUSTOM ? (int)J.coords.size() : 1;
        if ((int)J.coords.size() != expect) throw std::runtime_error("joint '" + J.name + "': wrong number of coordinates");
        if (J.kind == MOCOB200_JOINT_CUSTOM) {
            // SimTK::MobilizedBody::FunctionBased as CustomJoint builds it (CustomJoint.cpp:205-273): 1..6 mobilities
            if (expect < 1 || expect > 6) throw std::runtime_error("CustomJoint '" + J.name + "' must have 1 to 6 coordinates");
            bodyCj[k] = (int)cjI.size() / MOCOB200_CJ_ISTRIDE;
            for (int a = 0; a < 6; ++a) {
                const TransformAxisSpec& A = J.axes[a];
                int local = -1;
                for (size_t c = 0; c < J.coords.size(); ++c) if (J.coords[c].name == A.coord) local = (int)c;
                if (A.fnKind != MOCOB200_FN_CONSTANT && local < 0)
                    throw std::runtime_error("CustomJoint '" + J.name + "': TransformAxis coordinate '" + A.coord + "' is not one of the joint's");
                for (int c = 0; c < 3; ++c) cjD.push_back(A.axis.v[c]);
                cjD.push_back(A.fnKind == MOCOB200_FN_LINEAR ? A.c0 : 0.0);
                cjD.push_back(A.fnKind == MOCOB200_FN_SIMMSPLINE ? 0.0 : A.c1);
                cjI.push_back(A.fnKind);
                cjI.push_back(A.fnKind == MOCOB200_FN_CONSTANT ? -1 : nq + local);
                cjI.push_back(A.fnKind == MOCOB200_FN_SIMMSPLINE ? addSpline(A.spline) : -1);
            }
        }
        pushX(D, J.parent.X); // X_PF
        pushX(D, J.child.X);  // X_BM
        // massless body with inertia -> inertia ignored (SimbodyEngine/Body.cpp:133-140)
        D.push_back(B.mass);
        for (int a = 0; a < 3; ++a) D.push_back(B.com.v[a]);
        double in[6];
        for (int a = 0; a < 6; ++a) in[a] = B.mass > 0 ? B.inertiaCom[a] : 0.0;
        const double cx = B.com.v[0], cy = B.com.v[1], cz = B.com.v[2], mm = B.mass;
        // shift from the mass centre to the body origin (Body.cpp:355-377)
        in[0] += mm * (cy * cy + cz * cz);
        in[1] += mm * (cx * cx + cz * cz);
        in[2] += mm * (cx * cx + cy * cy);
        in[3] -= mm * cx * cy;
        in[4] -= mm * cx * cz;
        in[5] -= mm * cy * cz;
        for (int a = 0; a < 6; ++a) D.push_back(in[a]);
        I.push_back(mapBody(J.parent.body));
        I.push_back(J.kind);
        I.push_back(nq);
        I.push_back(expect);
        for (auto& c : J.coords) {
            coordIndex[c.name] = nq++;
            coordPath.push_back(J.container + J.name + "/" + c.name);
            coordRange.push_back({c.lo, c.hi});
        }
    }
    T.nq = nq;
    T.o_gravity = (int)D.size();
    for (int a = 0; a < 3; ++a) D.push_back(M.gravity.v[a]);

    // ---- state / control naming (MocoUtilities.cpp:495-528, 557-587)
    std::vector<std::string>& sn = fp->stateNames;
    // prescribed kinematics: the coordinate values and speeds are not states (MocoProblemRep.cpp:546-560)
    const bool prescribed = !M.positionMotion.empty();
    std::vector<const PiecewisePoly*> motionOf(nq, nullptr);
    if (prescribed) {
        for (const auto& kv : M.positionMotion)
            for (int q = 0; q < nq; ++q) if (coordPath[q] == kv.first) motionOf[q] = &kv.second;
        for (int q = 0; q < nq; ++q)
            if (!motionOf[q]) throw std::runtime_error("PositionMotion has no function for coordinate '" + coordPath[q] + "'");
    }
    const int nqStates = prescribed ? 0 : nq;
    for (int q = 0; q < nqStates; ++q) sn.push_back(coordPath[q] + "/value");
    for (int q = 0; q < nqStates; ++q) sn.push_back(coordPath[q] + "/speed");
    std::vector<int> musAct(M.muscles.size(), -1), musTen(M.muscles.size(), -1), musDeriv(M.muscles.size(), -1);
    int nz = 0, nda = 0;
    for (size_t m = 0; m < M.muscles.size(); ++m) {
        const MuscleSpec& mu = M.muscles[m];
        if (mu.pathActuator) continue; // no states
        if (!mu.ignoreActivationDynamics) { musAct[m] = nz++; sn.push_back(mu.path + "/activation"); }
        if (!mu.ignoreTendonCompliance) {
            musTen[m] = nz++;
            sn.push_back(mu.path + "/normalized_tendon_force");
            if (mu.tendonImplicit) {
                musDeriv[m] = nda++;
                fp->derivativeNames.push_back(mu.path + "/implicitderiv_normalized_tendon_force");
            }
        }
    }
    T.nz = nz;
    T.nderiv_aux = nda;
    std::vector<std::string>& cn = fp->controlNames;
    std::vector<int> musCtl(M.muscles.size(), -1), cactCtl(M.coordActs.size(), -1);
    for (auto& ao : M.actuatorOrder) {
        if (ao.first == 0) { cactCtl[ao.second] = (int)cn.size(); cn.push_back(M.coordActs[ao.second].path); }
        else { musCtl[ao.second] = (int)cn.size(); cn.push_back(M.muscles[ao.second].path); }
    }
    T.nc = (int)cn.size();
    auto stateIndex = [&](const std::string& n) {
        for (size_t i = 0; i < sn.size(); ++i) if (sn[i] == n) return (int)i;
        throw std::runtime_error("no state named '" + n + "'");
    };
    auto controlIndex = [&](const std::string& n) {
        for (size_t i = 0; i < cn.size(); ++i) if (cn[i] == n) return (int)i;
        throw std::runtime_error("no control named '" + n + "'");
    };
    auto coordOf = [&](const std::string& nameOrPath) {
        std::string n = nameOrPath;
        size_t sl = n.rfind('/');
        if (sl != std::string::npos) n = n.substr(sl + 1);
        auto it = coordIndex.find(n);
        if (it == coordIndex.end()) throw std::runtime_error("no coordinate named '" + nameOrPath + "'");
        return it->second;
    };

    // ---- kinematic constraints: coordinate couplers (MocoCasOCProblem.cpp:140-216 names the multipliers lambda_cid<i>_p<k>)
    std::vector<int> ccI;
    std::vector<double> ccD;
    for (size_t k = 0; k < M.couplers.size(); ++k) {
        const CouplerSpec& C = M.couplers[k];
        const int qi = coordOf(C.independent), qd = coordOf(C.dependent);
        if (qi == qd) throw std::runtime_error("CoordinateCouplerConstraint '" + C.name + "': dependent == independent coordinate");
        SplineSpec sp = C.fn.spline;
        sp.scale *= C.scaleFactor;
        ccI.push_back(C.fn.fnKind);
        ccI.push_back(qi);
        ccI.push_back(qd);
        ccI.push_back(C.fn.fnKind == MOCOB200_FN_SIMMSPLINE ? addSpline(sp) : -1);
        ccD.push_back(C.fn.fnKind == MOCOB200_FN_LINEAR ? C.scaleFactor * C.fn.c0 : 0.0);
        ccD.push_back(C.fn.fnKind == MOCOB200_FN_SIMMSPLINE ? 0.0 : C.scaleFactor * C.fn.c1);
        fp->multiplierNames.push_back("lambda_cid" + std::to_string(k) + "_p0");
        fp->slackNames.push_back("gamma_cid" + std::to_string(k) + "_p0");
    }

    // ---- muscles and their path points
    T.nmus = (int)M.muscles.size();
    std::vector<double> ppD, musD;
    std::vector<int> ppI, musI;
    int npp = 0;
    for (size_t m = 0; m < M.muscles.size(); ++m) {
        const MuscleSpec& mu = M.muscles[m];
        if (mu.points.size() < 2) throw std::runtime_error("muscle '" + mu.path + "' needs >= 2 path points");
        const MuscleSpec* firstMuscle = nullptr;
        for (auto& other : M.muscles) if (!other.pathActuator) { firstMuscle = &other; break; }
        if (!mu.pathActuator && firstMuscle &&
            (std::fabs(mu.activationTimeConstant - firstMuscle->activationTimeConstant) > 0 ||
             std::fabs(mu.deactivationTimeConstant - firstMuscle->deactivationTimeConstant) > 0))
            throw std::runtime_error("muscles must share activation/deactivation time constants "
                                     "(the reference's static-local constants, DeGrooteFregly2016Muscle.cpp:194-195)");
        const int first = npp;
        for (auto& pt : mu.points) {
            for (int a = 0; a < 3; ++a) ppD.push_back(pt.location.v[a]);
            int rec[MOCOB200_PP_ISTRIDE] = {mapBody(pt.body), pt.kind, -1, -1, -1, -1, -1, -1};
            double lo = 0, hi = 0;
            if (pt.kind == MOCOB200_PP_CONDITIONAL) {
                rec[2] = coordOf(pt.condCoord);
                lo = pt.range[0];
                hi = pt.range[1];
            } else if (pt.kind == MOCOB200_PP_MOVING) {
                int cq[3] = {-1, -1, -1};
                for (int a = 0; a < 3; ++a) {
                    if (pt.coord[a].empty() || pt.fn[a].x.empty())
                        throw std::runtime_error("MovingPathPoint needs a function and a coordinate per axis");
                    rec[3 + a] = addSpline(pt.fn[a]);
                    cq[a] = coordOf(pt.coord[a]);
                }
                if (cq[1] != cq[0] || cq[2] != cq[0])
                    throw std::runtime_error("MovingPathPoint with different x/y/z coordinates is unsupported");
                rec[6] = cq[0];
                rec[7] = cq[0];
                lo = coordRange[cq[0]][0];
                hi = coordRange[cq[0]][1];
            }
            ppD.push_back(lo);
            ppD.push_back(hi);
            for (int a = 0; a < MOCOB200_PP_ISTRIDE; ++a) ppI.push_back(rec[a]);
            ++npp;
        }
        const double kT = std::log((1.0 + 0.2) / 0.2) / (1.0 + mu.tendonStrainAtOneNormForce - 1.0);
        double md[MOCOB200_MUS_DSTRIDE] = {mu.maxIsometricForce, mu.optimalFiberLength, mu.tendonSlackLength,
                                           mu.pennationAngleAtOptimal, mu.maxContractionVelocity, mu.activationTimeConstant,
                                           mu.deactivationTimeConstant, mu.fiberDamping, mu.tendonStrainAtOneNormForce,
                                           mu.passiveFiberStrainAtOneNormForce, mu.activeForceWidthScale, mu.minControl,
                                           mu.maxControl, kT};
        musD.insert(musD.end(), md, md + MOCOB200_MUS_DSTRIDE);
        int flags = 0;
        if (mu.ignoreTendonCompliance) flags |= MOCOB200_MUS_FLAG_IGNORE_TENDON_COMPLIANCE;
        if (mu.ignoreActivationDynamics) flags |= MOCOB200_MUS_FLAG_IGNORE_ACTIVATION_DYNAMICS;
        if (mu.ignorePassiveFiberForce) flags |= MOCOB200_MUS_FLAG_IGNORE_PASSIVE;
        if (mu.tendonImplicit) flags |= MOCOB200_MUS_FLAG_TENDON_IMPLICIT;
        if (mu.pathActuator) flags |= MOCOB200_MUS_FLAG_PATH_ACTUATOR | MOCOB200_MUS_FLAG_IGNORE_TENDON_COMPLIANCE | MOCOB200_MUS_FLAG_IGNORE_ACTIVATION_DYNAMICS;
        int mi[MOCOB200_MUS_ISTRIDE] = {first, (int)mu.points.size(), musCtl[m], musAct[m], musTen[m], flags,
                                        musDeriv[m], musDeriv[m], 0, 0};
        musI.insert(musI.end(), mi, mi + MOCOB200_MUS_ISTRIDE);
    }
    T.npp = npp;
    T.o_pp_d = (int)D.size(); D.insert(D.end(), ppD.begin(), ppD.end());
    T.o_pp_i = (int)I.size(); I.insert(I.end(), ppI.begin(), ppI.end());
    T.o_mus_d = (int)D.size(); D.insert(D.end(), musD.begin(), musD.end());
    T.o_mus_i = (int)I.size(); I.insert(I.end(), musI.begin(), musI.end());
    T.nspline = (int)splineRec.size() / MOCOB200_SPLINE_ISTRIDE;
    T.o_spline_data = (int)D.size(); D.insert(D.end(), splineData.begin(), splineData.end());
    T.o_spline_i = (int)I.size(); I.insert(I.end(), splineRec.begin(), splineRec.end());

    // ---- contacts
    T.ncontact = (int)M.contacts.size();
    T.o_contact_d = (int)D.size();
    T.o_contact_i = (int)I.size();
    for (auto& c : M.contacts) {
        for (int a = 0; a < 3; ++a) D.push_back(c.sphere.X.p.v[a]);
        D.push_back(c.radius);
        pushX(D, c.halfSpace.X);
        double pr[12] = {c.stiffness, c.dissipation, c.muS, c.muD, c.muV, c.vT, c.cF, c.bD, c.bV, 0, 0, 0};
        D.insert(D.end(), pr, pr + 12);
        I.push_back(mapBody(c.sphere.body));
        I.push_back(mapBody(c.halfSpace.body));
    }
    // ---- coordinate actuators
    T.ncact = (int)M.coordActs.size();
    T.o_cact_d = (int)D.size();
    T.o_cact_i = (int)I.size();
    for (size_t a = 0; a < M.coordActs.size(); ++a) {
        D.push_back(M.coordActs[a].optimalForce);
        I.push_back(coordOf(M.coordActs[a].coord));
        I.push_back(cactCtl[a]);
    }

    // ---- custom joints (before the bounds block: the warp kernels stage everything below o_bounds in shared memory)
    T.ncustom = (int)cjI.size() / MOCOB200_CJ_ISTRIDE;
    if (T.ncustom > 0) {
        T.o_cj_d = (int)D.size(); D.insert(D.end(), cjD.begin(), cjD.end());
        T.o_cj_i = (int)I.size(); I.insert(I.end(), cjI.begin(), cjI.end());
        T.o_body_cj = (int)I.size(); I.insert(I.end(), bodyCj.begin(), bodyCj.end());
    }

    // ---- kinematic constraints: the coupler records (built before the spline table was written)
    T.ncoupler = (int)M.couplers.size();
    if (T.ncoupler > 0) {
        T.o_cc_i = (int)I.size(); I.insert(I.end(), ccI.begin(), ccI.end());
        T.o_cc_d = (int)D.size(); D.insert(D.end(), ccD.begin(), ccD.end());
    }

    // ---- wrap cylinders and the (muscle, cylinder) pairings (WrapObject.cpp:97-141, PathWrap.cpp)
    {
        std::vector<int> woIndex(M.wrapObjects.size(), -1), woI, pwHead, pwRec;
        std::vector<double> woD;
        int nwo = 0;
        for (size_t w = 0; w < M.wrapObjects.size(); ++w) {
            const WrapCylinderSpec& W = M.wrapObjects[w];
            bool used = false;
            for (const MuscleSpec& mu : M.muscles) for (const auto& pw : mu.wraps) used = used || pw.object == W.name;
            if (!used) continue;
            if (!W.supported) throw std::runtime_error("wrap object '" + W.name + "': only WrapCylinder is supported (SURVEY §8 a18)");
            if (W.wrapAxis == 2 && W.wrapSign != 0) throw std::runtime_error("wrap cylinder '" + W.name + "': a z quadrant is meaningless for a cylinder");
            woIndex[w] = nwo++;
            for (int k = 0; k < 9; ++k) woD.push_back(W.X.R.m[k]);
            for (int a = 0; a < 3; ++a) woD.push_back(W.X.p.v[a]);
            woD.push_back(W.radius); woD.push_back(W.length); woD.push_back(0.0); woD.push_back(0.0);
            woI.push_back(mapBody(W.body)); woI.push_back(W.wrapAxis); woI.push_back(W.wrapSign); woI.push_back(W.active ? 1 : 0);
        }
        for (const MuscleSpec& mu : M.muscles) {
            pwHead.push_back((int)pwRec.size() / MOCOB200_PW_ISTRIDE);
            pwHead.push_back((int)mu.wraps.size());
            for (const auto& pw : mu.wraps) {
                int wo = -1;
                for (size_t w = 0; w < M.wrapObjects.size(); ++w) if (M.wrapObjects[w].name == pw.object) wo = woIndex[w];
                pwRec.push_back(wo); pwRec.push_back(pw.rangeStart); pwRec.push_back(pw.rangeEnd); pwRec.push_back(0);
            }
        }
        T.nwrapobj = nwo;
        T.npathwrap = (int)pwRec.size() / MOCOB200_PW_ISTRIDE;
        T.o_wo_d = T.o_wo_i = T.o_pw_i = 0;
        if (T.npathwrap > 0) {
            T.o_wo_d = (int)D.size(); D.insert(D.end(), woD.begin(), woD.end());
            T.o_wo_i = (int)I.size(); I.insert(I.end(), woI.begin(), woI.end());
            T.o_pw_i = (int)I.size(); I.insert(I.end(), pwHead.begin(), pwHead.end()); I.insert(I.end(), pwRec.begin(), pwRec.end());
        }
    }

    // ---- variable bounds (MocoProblemRep.cpp:300-430 ; CasOCProblem.h:603-606)
    const double inf = std::numeric_limits<double>::infinity();
    auto resolve = [&](const std::map<std::string, VariableInfo>& infos, const std::string& n, Bounds dflt) {
        VariableInfo vi;
        auto it = infos.find(n);
        if (it != infos.end()) vi = it->second;
        if (!vi.bounds.isSet()) vi.bounds = dflt;
        std::array<double, 6> o;
        o[0] = vi.bounds.lo; o[1] = vi.bounds.hi;
        o[2] = vi.initial.isSet() ? std::max(o[0], vi.initial.lo) : o[0];
        o[3] = vi.initial.isSet() ? std::min(o[1], vi.initial.hi) : o[1];
        o[4] = vi.final.isSet() ? std::max(o[0], vi.final.lo) : o[0];
        o[5] = vi.final.isSet() ? std::min(o[1], vi.final.hi) : o[1];
        return o;
    };
    for (auto& kv : P.stateInfos)
        if (std::find(sn.begin(), sn.end(), kv.first) == sn.end())
            throw std::runtime_error("state info for nonexistent state '" + kv.first + "'");
    for (auto& kv : P.controlInfos)
        if (std::find(cn.begin(), cn.end(), kv.first) == cn.end())
            throw std::runtime_error("control info for nonexistent control '" + kv.first + "'");
    std::vector<std::array<double, 6>> ctlB;
    for (size_t c = 0; c < cn.size(); ++c) {
        Bounds dflt(-inf, inf);
        for (size_t a = 0; a < M.coordActs.size(); ++a) if (cactCtl[a] == (int)c) dflt = Bounds(M.coordActs[a].minControl, M.coordActs[a].maxControl);
        for (size_t m = 0; m < M.muscles.size(); ++m) if (musCtl[m] == (int)c) dflt = Bounds(M.muscles[m].minControl, M.muscles[m].maxControl);
        ctlB.push_back(resolve(P.controlInfos, cn[c], dflt));
    }
    T.o_bounds = (int)D.size();
    for (int q = 0; q < nqStates; ++q) { auto o = resolve(P.stateInfos, sn[q], Bounds(coordRange[q][0], coordRange[q][1])); D.insert(D.end(), o.begin(), o.end()); }
    for (int q = 0; q < nqStates; ++q) { auto o = resolve(P.stateInfos, sn[nq + q], P.defaultSpeedBounds); D.insert(D.end(), o.begin(), o.end()); }
    for (size_t m = 0; m < M.muscles.size(); ++m) {
        if (musAct[m] >= 0) {
            Bounds dflt(-inf, inf);
            if (P.boundActivationFromExcitation) dflt = Bounds(ctlB[musCtl[m]][0], ctlB[musCtl[m]][1]);
            auto o = resolve(P.stateInfos, M.muscles[m].path + "/activation", dflt);
            D.insert(D.end(), o.begin(), o.end());
        }
        if (musTen[m] >= 0) { // getBoundsNormalizedTendonForce: [0, 5] (DeGrooteFregly2016Muscle.h:304-305,803-804)
            auto o = resolve(P.stateInfos, M.muscles[m].path + "/normalized_tendon_force", Bounds(0.0, 5.0));
            D.insert(D.end(), o.begin(), o.end());
        }
    }
    for (auto& o : ctlB) D.insert(D.end(), o.begin(), o.end());
    T.o_time_bounds = (int)D.size();
    D.push_back(P.timeInitial.lo); D.push_back(P.timeInitial.hi);
    D.push_back(P.timeFinal.lo); D.push_back(P.timeFinal.hi);

    // ---- goals
    std::vector<int> ppolyRec;
    std::vector<double> ppolyData;
    auto addPPoly = [&](const PiecewisePoly& pp) {
        if (pp.breaks.size() < 2) throw std::runtime_error("tracking reference spline is empty");
        const int id = (int)ppolyRec.size() / MOCOB200_PPOLY_ISTRIDE;
        ppolyRec.push_back((int)ppolyData.size());
        ppolyRec.push_back((int)pp.breaks.size() - 1);
        ppolyRec.push_back(pp.order);
        ppolyData.insert(ppolyData.end(), pp.breaks.begin(), pp.breaks.end());
        ppolyData.insert(ppolyData.end(), pp.coef.begin(), pp.coef.end());
        return id;
    };
    T.prescribed = prescribed ? 1 : 0;
    T.o_presc_i = 0;
    if (prescribed) {
        std::vector<int> ids;
        for (int q = 0; q < nq; ++q) ids.push_back(addPPoly(*motionOf[q]));
        T.o_presc_i = (int)I.size();
        I.insert(I.end(), ids.begin(), ids.end());
    }
    // ---- external forces (ExternalForce.cpp:330-355): functions of time in the piecewise-polynomial table
    {
        std::vector<int> xfI;
        for (const ExternalForceSpec& X : M.externalForces) {
            xfI.push_back(mapBody(X.appliedTo));
            xfI.push_back(mapBody(X.forceExpressedIn));
            xfI.push_back(mapBody(X.pointExpressedIn));
            for (int a = 0; a < 3; ++a) xfI.push_back(X.hasForce ? addPPoly(X.force[a]) : -1);
            for (int a = 0; a < 3; ++a) xfI.push_back(X.hasPoint ? addPPoly(X.point[a]) : -1);
            for (int a = 0; a < 3; ++a) xfI.push_back(X.hasTorque ? addPPoly(X.torque[a]) : -1);
        }
        T.nxforce = (int)M.externalForces.size();
        T.o_xforce_i = (int)I.size();
        I.insert(I.end(), xfI.begin(), xfI.end());
    }
    auto contactId = [&](const std::string& n) {
        for (size_t i = 0; i < M.contacts.size(); ++i) if (M.contacts[i].name == n) return (int)i;
        throw std::runtime_error("no contact force named '" + n + "'");
    };
    std::vector<int> goalRec, goalI;
    std::vector<double> goalD;
    for (const GoalSpec& g : P.goals) {
        int mode = g.mode;
        if (prescribed && (g.kind == MOCOB200_GOAL_AVERAGE_SPEED || g.kind == MOCOB200_GOAL_INITIAL_VELOCITY_EQUILIBRIUM_DGF ||
                           (g.kind == MOCOB200_GOAL_CONTROL && g.divideByDisplacement)))
            throw std::runtime_error("goal '" + g.name + "' evaluates the model's kinematics at the end points: not supported with prescribed kinematics");
        const bool defaultEC = g.kind == MOCOB200_GOAL_AVERAGE_SPEED || g.kind == MOCOB200_GOAL_PERIODICITY ||
                               g.kind == MOCOB200_GOAL_INITIAL_ACTIVATION;
        if (mode == MOCOB200_MODE_PATH_CONSTRAINT && g.kind != MOCOB200_PATHCON_CONTROL_BOUND)
            throw std::runtime_error("unsupported path constraint (only MocoControlBoundConstraint, SURVEY §8 a10)");
        if (mode < 0) mode = defaultEC ? MOCOB200_MODE_ENDPOINT_CONSTRAINT : MOCOB200_MODE_COST;
        const int ioff = (int)goalI.size(), doff = (int)goalD.size();
        int nout = 1, hasInt = 0;
        goalD.push_back(mode == MOCOB200_MODE_COST ? g.weight : 1.0); // m_weightToUse (MocoGoal.h:262-271)
        goalD.push_back(g.ecBounds.lo);
        goalD.push_back(g.ecBounds.hi);
        switch (g.kind) {
        case MOCOB200_GOAL_CONTROL: {
            hasInt = 1;
            std::vector<int> ci;
            std::vector<double> cw;
            for (size_t c = 0; c < cn.size(); ++c) {
                double w = 1.0;
                auto it = g.controlWeights.find(cn[c]);
                if (it != g.controlWeights.end()) w = it->second;
                if (w != 0.0) { ci.push_back((int)c); cw.push_back(w); }
            }
            for (auto& kv : g.controlWeights) controlIndex(kv.first);
            if (g.exponent < 2) throw std::runtime_error("MocoControlGoal exponent must be 2 or greater");
            goalI.push_back(g.exponent);
            goalI.push_back(g.divideByDisplacement ? 1 : 0);
            goalI.push_back((int)ci.size());
            goalI.insert(goalI.end(), ci.begin(), ci.end());
            goalD.insert(goalD.end(), cw.begin(), cw.end());
        } break;
        case MOCOB200_GOAL_STATE_TRACKING:
        case MOCOB200_GOAL_CONTROL_TRACKING:
        case MOCOB200_GOAL_SUM_SQUARED_STATE: {
            hasInt = 1;
            goalI.push_back((int)g.refs.size());
            for (auto& r : g.refs) {
                goalI.push_back(g.kind == MOCOB200_GOAL_CONTROL_TRACKING ? controlIndex(r.name) : stateIndex(r.name));
                goalI.push_back(g.kind == MOCOB200_GOAL_SUM_SQUARED_STATE ? -1 : addPPoly(r.spline));
                goalD.push_back(r.weight);
            }
        } break;
        case MOCOB200_GOAL_CONTACT_TRACKING: {
            hasInt = 1;
            goalI.push_back(g.projection);
            goalI.push_back((int)g.groups.size());
            for (int a = 0; a < 3; ++a) goalD.push_back(g.projectionVector.v[a]);
            {   // m_denominator = total mass * |gravity| (MocoContactTrackingGoal.cpp:76-81)
                double denom = 0;
                for (auto& b : M.bodies) denom += b.mass;
                const double gmag = std::sqrt(M.gravity.v[0] * M.gravity.v[0] + M.gravity.v[1] * M.gravity.v[1] +
                                              M.gravity.v[2] * M.gravity.v[2]);
                if (gmag > 1.9984014443252818e-14) denom *= gmag;
                goalD.push_back(denom);
            }
            for (auto& grp : g.groups) {
                goalI.push_back((int)grp.contacts.size());
                for (auto& cname : grp.contacts) {
                    const int cid = contactId(cname);
                    const ContactSpec& c = M.contacts[cid];
                    const int applied = M.bodyIndex(grp.appliedToBody);
                    int offset;
                    if (c.sphere.body == applied) offset = 0;           // MocoContactTrackingGoal.cpp:190-207
                    else if (c.halfSpace.body == applied) offset = 6;
                    else throw std::runtime_error("contact '" + cname + "' touches neither side of applied_to_body");
                    goalI.push_back(cid);
                    goalI.push_back(offset);
                }
                for (int a = 0; a < 3; ++a) goalI.push_back(addPPoly(grp.ref[a]));
            }
        } break;
        case MOCOB200_GOAL_FINAL_TIME: break;
        case MOCOB200_GOAL_AVERAGE_SPEED: goalD.push_back(g.desiredAverageSpeed); break;
        case MOCOB200_GOAL_PERIODICITY: {
            nout = (int)(g.statePairs.size() + g.controlPairs.size());
            goalI.push_back((int)g.statePairs.size());
            goalI.push_back((int)g.controlPairs.size());
            for (auto& pr : g.statePairs) {
                goalI.push_back(stateIndex(pr.a));
                goalI.push_back(stateIndex(pr.b.empty() ? pr.a : pr.b));
                goalD.push_back(pr.sign);
            }
            for (auto& pr : g.controlPairs) {
                goalI.push_back(controlIndex(pr.a));
                goalI.push_back(controlIndex(pr.b.empty() ? pr.a : pr.b));
                goalD.push_back(pr.sign);
            }
        } break;
        case MOCOB200_GOAL_INITIAL_ACTIVATION: {
            std::vector<int> pairs;
            for (size_t m = 0; m < M.muscles.size(); ++m)
                if (musAct[m] >= 0) { pairs.push_back(musCtl[m]); pairs.push_back(2 * nqStates + musAct[m]); }
            nout = (int)pairs.size() / 2;
            goalI.push_back(nout);
            goalI.insert(goalI.end(), pairs.begin(), pairs.end());
        } break;
        case MOCOB200_GOAL_INITIAL_VELOCITY_EQUILIBRIUM_DGF: {
            std::vector<int> ms;
            for (size_t m = 0; m < M.muscles.size(); ++m)
                if (!M.muscles[m].ignoreTendonCompliance) ms.push_back((int)m);
            nout = (int)ms.size();
            goalI.push_back(nout);
            goalI.insert(goalI.end(), ms.begin(), ms.end());
        } break;
        case MOCOB200_PATHCON_CONTROL_BOUND: { // MocoControlBoundConstraint::initializeOnModelImpl (.cpp:37-119)
            if (mode != MOCOB200_MODE_PATH_CONSTRAINT) throw std::runtime_error("MocoControlBoundConstraint is a path constraint");
            const bool hasLower = !g.lowerBound.breaks.empty(), hasUpper = !g.upperBound.breaks.empty();
            if (g.equalityWithLower && hasUpper)
                throw std::runtime_error("If equality_with_lower==true, upper bound function must not be set.");
            if (g.equalityWithLower && !hasLower)
                throw std::runtime_error("If equality_with_lower==true, lower bound function must be set.");
            auto checkTimeRange = [&](const PiecewisePoly& f) {
                if (!f.isSpline) return;
                if (f.breaks.front() > P.timeInitial.lo)
                    throw std::runtime_error("The function's minimum domain value (" + std::to_string(f.breaks.front()) +
                                             ") must be less than or equal to the minimum possible initial time (" +
                                             std::to_string(P.timeInitial.lo) + ").");
                if (f.breaks.back() < P.timeFinal.hi)
                    throw std::runtime_error("The function's maximum domain value (" + std::to_string(f.breaks.back()) +
                                             ") must be greater than or equal to the maximum possible final time (" +
                                             std::to_string(P.timeFinal.hi) + ").");
            };
            if (hasLower) checkTimeRange(g.lowerBound);
            if (hasUpper) checkTimeRange(g.upperBound);
            const int lowerId = hasLower ? addPPoly(g.lowerBound) : -1, upperId = hasUpper ? addPPoly(g.upperBound) : -1;
            nout = 0;
            if (hasLower || hasUpper)
                for (const std::string& path : g.controlPaths) {
                    const int c = controlIndex(path);
                    if (hasLower) { // 0 <= control - lower(t) <= inf   (or == 0 with equality_with_lower)
                        goalI.push_back(c); goalI.push_back(lowerId);
                        goalD.push_back(0.0); goalD.push_back(g.equalityWithLower ? 0.0 : inf);
                        ++nout;
                    }
                    if (hasUpper) { // -inf <= control - upper(t) <= 0
                        goalI.push_back(c); goalI.push_back(upperId);
                        goalD.push_back(-inf); goalD.push_back(0.0);
                        ++nout;
                    }
                }
        } break;
        default: throw std::runtime_error("unsupported goal kind");
        }
        if (mode == MOCOB200_MODE_COST && g.kind != MOCOB200_GOAL_PERIODICITY &&
            g.kind != MOCOB200_GOAL_INITIAL_ACTIVATION && g.kind != MOCOB200_GOAL_INITIAL_VELOCITY_EQUILIBRIUM_DGF)
            nout = 1;
        int rec[MOCOB200_GOAL_ISTRIDE] = {g.kind, mode, nout, hasInt, ioff, doff, (int)goalI.size() - ioff,
                                          (int)goalD.size() - doff};
        goalRec.insert(goalRec.end(), rec, rec + MOCOB200_GOAL_ISTRIDE);
    }
    T.ngoal = (int)P.goals.size();
    T.nppoly = (int)ppolyRec.size() / MOCOB200_PPOLY_ISTRIDE;
    T.o_ppoly_data = (int)D.size(); D.insert(D.end(), ppolyData.begin(), ppolyData.end());
    T.o_ppoly_i = (int)I.size(); I.insert(I.end(), ppolyRec.begin(), ppolyRec.end());
    // goal payload offsets become absolute pool offsets
    const int gI0 = (int)I.size() + (int)goalRec.size(), gD0 = (int)D.size();
    for (size_t k = 0; k < goalRec.size(); k += MOCOB200_GOAL_ISTRIDE) { goalRec[k + 4] += gI0; goalRec[k + 5] += gD0; }
    T.o_goal_i = (int)I.size();
    I.insert(I.end(), goalRec.begin(), goalRec.end());
    I.insert(I.end(), goalI.begin(), goalI.end());
    D.insert(D.end(), goalD.begin(), goalD.end());

    // ---- transcription
    std::memset(&fp->tr, 0, sizeof fp->tr);
    applySettingsToTranscription(S, *fp, false);
    fp->tr.multiplier_bounds[0] = P.multiplierBounds.lo;
    fp->tr.multiplier_bounds[1] = P.multiplierBounds.hi;
    if (P.kinematicConstraintBounds.lo != 0 || P.kinematicConstraintBounds.hi != 0)
        throw std::runtime_error("non-zero kinematic constraint bounds are unsupported");
    fp->rebind();
    return fp;
}

} // namespace mocob200
