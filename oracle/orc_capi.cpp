// TEST INFRASTRUCTURE ONLY — C entry points of the CPU oracle (liborc.so), bound by oracle/oracle_py.py.
// Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs — never by
// the product library.
#include <omp.h>

#include <cstring>
#include <memory>

#include "orc_jac.hpp"

using namespace orc;

struct orc_handle {
    Nlp nlp;
    std::vector<int> colptr, rowidx;
    bool haveStructure = false;
    void needStructure() {
        if (haveStructure) return;
        JacBuilder jb(nlp);
        jb.structure(colptr, rowidx);
        haveStructure = true;
    }
};

extern "C" {

orc_handle* orc_create(const mocob200_tables* T, const mocob200_transcription* R) {
    if (!T || !R) return nullptr;
    if (T->nb > kMaxBodies || T->nq > kMaxDof || T->ncontact > kMaxContacts) return nullptr;
    auto h = new orc_handle;
    h->nlp.init(T, R);
    return h;
}
void orc_destroy(orc_handle* h) { delete h; }
void orc_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int orc_max_threads() { return omp_get_max_threads(); }

// n_grid n_mesh n_var n_con ns nc nd n_in n_out n_dyn_out n_integrand n_cost_terms n_ein nq nz nmus ncontact
void orc_dims(const orc_handle* h, int* d) {
    const Nlp& N = h->nlp;
    int v[17] = {N.nGrid, N.nMesh, N.nVar, N.nCon, N.ns, N.nc, N.nd, N.nIn, N.nOut, N.M.n_dyn_out, N.M.n_integrand,
                 N.numCostTerms(), N.nEin, N.nq, N.M.nz, N.M.T->nmus, N.M.T->ncontact};
    std::memcpy(d, v, sizeof v);
}
void orc_set_fd(orc_handle* h, int scheme, double step, int h_iter, double u_aim) {
    h->nlp.fd.defaults(scheme);
    if (step > 0) h->nlp.fd.h = step;
    if (h_iter >= 0) h->nlp.fd.h_iter = h_iter;
    if (u_aim > 0) h->nlp.fd.u_aim = u_aim;
}
void orc_eval_point(const orc_handle* h, const double* in, double* out, unsigned mask) { evalPoint<double>(h->nlp.M, in, out, mask); }
void orc_eval_points(const orc_handle* h, const double* x, double* Y) { h->nlp.evalPoints(x, Y); }
void orc_eval_point_jacobians(const orc_handle* h, const double* x, double* Y, double* J) { h->nlp.evalPointJacobians(x, Y, J); }
void orc_eval_g(const orc_handle* h, const double* x, double* g) { h->nlp.evalG(x, g); }
double orc_eval_f(const orc_handle* h, const double* x, double* terms) { return h->nlp.evalF(x, terms); }
void orc_eval_grad_f(const orc_handle* h, const double* x, double* grad) { JacBuilder jb(h->nlp); jb.gradient(x, grad); }
int orc_nnz(orc_handle* h) { h->needStructure(); return (int)h->rowidx.size(); }
void orc_structure(orc_handle* h, int* colptr, int* rowidx) {
    h->needStructure();
    std::memcpy(colptr, h->colptr.data(), sizeof(int) * h->colptr.size());
    std::memcpy(rowidx, h->rowidx.data(), sizeof(int) * h->rowidx.size());
}
void orc_eval_jac_g(orc_handle* h, const double* x, double* vals) {
    h->needStructure();
    JacBuilder jb(h->nlp);
    jb.jacobian(x, h->colptr, h->rowidx, vals);
}
void orc_eval_endpoint(const orc_handle* h, int goal, const double* ein, double* vals) { evalEndpoint<double>(h->nlp.M, goal, ein, vals); }

// algorithmic operation counts of ONE point evaluation (SURVEY.md §8(d) convention): counts[0] = flop, [1] = libm calls
void orc_count_ops(const orc_handle* h, const double* in, unsigned mask, long long* out2) {
    const Model& M = h->nlp.M;
    std::vector<Counted> ci(M.n_in), co(M.n_out);
    for (int i = 0; i < M.n_in; ++i) ci[i] = Counted(in[i]);
    counts() = OpCounts();
    evalPoint<Counted>(M, ci.data(), co.data(), mask);
    out2[0] = counts().flop;
    out2[1] = counts().transc;
}

// ---- pieces exposed for the known-answer tests -------------------------------------------------
// DeGrooteFregly2016Muscle curves: which = 0 active f-l, 1 passive f-l, 2 f-v, 3 f-v inverse, 4 tendon force,
// 5 tendon force inverse, 6 active f-l derivative, 7 passive derivative, 8 tendon derivative
double orc_dgf_curve(int which, double x, double p0, double p1) {
    switch (which) {
    case 0: return dgf::activeForceLength<double>(x, p0);
    case 1: return dgf::passiveForce<double>(x, p0);
    case 2: return dgf::forceVelocity<double>(x);
    case 3: return dgf::forceVelocityInverse<double>(x);
    case 4: return dgf::tendonForce<double>(x, dgf::tendonKT(p0));
    case 5: return dgf::tendonForceInverse<double>(x, dgf::tendonKT(p0));
    case 6: return dgf::activeForceLengthDeriv<double>(x, p0);
    case 7: return dgf::passiveForceDeriv<double>(x, p0);
    case 8: return dgf::tendonForceDeriv<double>(x, dgf::tendonKT(p0));
    }
    return NAN;
}
double orc_activation_derivative(double e, double a, double tauA, double tauD) { return dgf::activationDerivative<double>(e, a, tauA, tauD); }
// params[14] in MOCOB200_MUS_DSTRIDE layout; out[22]
void orc_muscle(const double* params, int flags, double mtLength, double mtVelocity, double activation, double ntf,
                double ntfRate, double* out) {
    const MuscleResult<double> r = muscleModel<double>(params, flags, mtLength, mtVelocity, activation, ntf, ntfRate, true);
    const double v[22] = {r.tendonForce, r.equilibriumResidual, r.normTendonForceRate, r.linearizedResidualRate,
                          r.normFiberLength, r.cosPennation, r.normFiberVelocity, r.fiberForce, r.activeFiberForce,
                          r.passiveFiberForce, r.fiberStiffness, r.tendonStiffness, r.fiberStiffnessAlongTendon,
                          r.muscleStiffness, r.fiberVelocity, r.tendonVelocity, r.fal, r.fpe, r.fv, r.pennationAngle,
                          r.fiberLength, r.tendonLength};
    std::memcpy(out, v, sizeof v);
}
// multibody pieces: mass matrix, inverse dynamics (gravity optional), kinetic + gravitational energy, body poses
void orc_mass_matrix(const orc_handle* h, const double* q, double* Mass) {
    const Model& M = h->nlp.M;
    std::vector<double> u(M.nq, 0.0);
    static thread_local Frames<double> K;
    kinematics<double>(M, q, u.data(), K);
    massMatrix<double>(M, K, Mass);
}
void orc_inverse_dynamics(const orc_handle* h, const double* q, const double* u, const double* udot, int withGravity, double* tau) {
    const Model& M = h->nlp.M;
    static thread_local Frames<double> K;
    kinematics<double>(M, q, u, K);
    std::vector<Wrench<double>> applied(kMaxBodies);
    if (withGravity) {
        const double* g = M.T->dbl + M.T->o_gravity;
        for (int b = 0; b < M.nb; ++b)
            applied[b].addForceAt(K.origin[b] + K.comOffset[b], M.bodyD(b)[24] * Vec3<double>(g[0], g[1], g[2]));
    }
    inverseDynamics<double>(M, K, u, udot, applied.data(), true, tau);
}
// energy[0] = kinetic, [1] = gravitational potential; com[3] = whole-body mass centre
void orc_energy(const orc_handle* h, const double* q, const double* u, double* energy, double* com) {
    const Model& M = h->nlp.M;
    static thread_local Frames<double> K;
    kinematics<double>(M, q, u, K);
    const double* g = M.T->dbl + M.T->o_gravity;
    double ke = 0, pe = 0;
    for (int b = 0; b < M.nb; ++b) {
        const double* d = M.bodyD(b);
        const double m = d[24];
        const Vec3<double> cg = K.comOffset[b], c = K.origin[b] + cg;
        const Vec3<double> vc = K.vOrigin[b] + cross(K.omega[b], cg);
        Mat3<double> Io;
        Io.m[0][0] = d[28]; Io.m[1][1] = d[29]; Io.m[2][2] = d[30];
        Io.m[0][1] = Io.m[1][0] = d[31]; Io.m[0][2] = Io.m[2][0] = d[32]; Io.m[1][2] = Io.m[2][1] = d[33];
        const Mat3<double> PA = parallelAxis(m, Vec3<double>(d[25], d[26], d[27]));
        Mat3<double> IcL;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) IcL.m[i][j] = Io.m[i][j] - PA.m[i][j];
        const Vec3<double> wl = transpose(K.R[b]) * K.omega[b];
        ke += 0.5 * m * dot(vc, vc) + 0.5 * dot(wl, IcL * wl);
        pe -= m * dot(Vec3<double>(g[0], g[1], g[2]), c);
    }
    energy[0] = ke; energy[1] = pe;
    const Vec3<double> cm = massCenter<double>(M, K);
    com[0] = cm.x; com[1] = cm.y; com[2] = cm.z;
}
// path length / lengthening speed of muscle m at (q,u)
void orc_path(const orc_handle* h, int m, const double* q, const double* u, double* lengthSpeed) {
    const Model& M = h->nlp.M;
    static thread_local Frames<double> K;
    static thread_local PathState<double> P;
    kinematics<double>(M, q, u, K);
    pathGeometry<double>(M, m, q, u, K, P);
    lengthSpeed[0] = P.length; lengthSpeed[1] = P.speed;
}
// generalized forces produced by unit tension in muscle m (= minus the moment arms): tau[nq]
void orc_path_generalized_force(const orc_handle* h, int m, const double* q, double tension, double* tau) {
    const Model& M = h->nlp.M;
    std::vector<double> u(M.nq, 0.0), zero(M.nq, 0.0), mob(M.nq, 0.0), need(M.nq, 0.0);
    static thread_local Frames<double> K;
    static thread_local PathState<double> P;
    kinematics<double>(M, q, u.data(), K);
    pathGeometry<double>(M, m, q, u.data(), K, P);
    std::vector<Wrench<double>> applied(kMaxBodies);
    applyPathTension<double>(P, tension, applied.data(), mob.data());
    inverseDynamics<double>(M, K, u.data(), zero.data(), applied.data(), false, need.data());
    for (int k = 0; k < M.nq; ++k) tau[k] = mob[k] - need[k];
}
// contact force on the sphere body (ground frame) and its point of application
void orc_contact(const orc_handle* h, int c, const double* q, const double* u, double* force3, double* point3) {
    const Model& M = h->nlp.M;
    static thread_local Frames<double> K;
    kinematics<double>(M, q, u, K);
    Vec3<double> cp;
    const Vec3<double> f = contactForce<double>(M, c, K, cp);
    force3[0] = f.x; force3[1] = f.y; force3[2] = f.z;
    point3[0] = cp.x; point3[1] = cp.y; point3[2] = cp.z;
}

} // extern "C"
