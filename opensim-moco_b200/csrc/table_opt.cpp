// Device-side model reduction done once at mocob200_create: bodies attached by a WeldJoint are merged into their parent.
// A welded body adds nothing to the dynamics but a constant offset - yet every tree recursion of the kernels pays a level
// and a lane / loop trip for it (the 2D gait model has 12 mobilized bodies of which 4, the calcaneus and toes of each foot,
// are welded: 6 tree levels become 4). The reference keeps such bodies as separate mobilized bodies of zero mobility
// (WeldJoint.cpp); the merged model is the same mechanical system, so results agree to round-off (the CPU oracle keeps the
// unmerged tables, which makes every parity test a check of this reduction as well).
#include <cmath>
#include <cstring>
#include <vector>

#include "layout.h"

namespace mocob200 {
namespace {
struct M3 { double m[9]; };
M3 mulM(const double* a, const double* b) {
    M3 c;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) c.m[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
    return c;
}
M3 mulMT(const double* a, const double* b) { // a b^T
    M3 c;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) c.m[3 * i + j] = a[3 * i] * b[3 * j] + a[3 * i + 1] * b[3 * j + 1] + a[3 * i + 2] * b[3 * j + 2];
    return c;
}
void mulV(const double* R, const double* v, double* o) {
    for (int i = 0; i < 3; ++i) o[i] = R[3 * i] * v[0] + R[3 * i + 1] * v[1] + R[3 * i + 2] * v[2];
}
} // namespace

int mergeWeldedBodies(const mocob200_tables& T, std::vector<double>& dbl, std::vector<int>& idx, mocob200_tables& out) {
    dbl.assign(T.dbl, T.dbl + T.n_dbl);
    idx.assign(T.idx, T.idx + T.n_int);
    out = T;
    out.dbl = dbl.data();
    out.idx = idx.data();
    const int nb = T.nb;
    auto bodyD = [&](int b) { return dbl.data() + T.o_body_d + b * MOCOB200_BODY_DSTRIDE; };
    auto bodyI = [&](int b) { return idx.data() + T.o_body_i + b * MOCOB200_BODY_ISTRIDE; };
    // which welded bodies can go: nothing attached to them may depend on their own frame in a way a constant offset cannot express
    std::vector<char> keep(nb, 1);
    for (int b = 0; b < nb; ++b) keep[b] = !(bodyI(b)[1] == MOCOB200_JOINT_WELD && bodyI(b)[0] >= 0);
    for (int p = 0; p < T.npp; ++p) {
        const int* pi = idx.data() + T.o_pp_i + p * MOCOB200_PP_ISTRIDE;
        if (pi[1] == MOCOB200_PP_MOVING && pi[0] >= 0) keep[pi[0]] = 1; // its location functions are expressed in that frame
    }
    for (int f = 0; f < T.nxforce; ++f) {
        const int* xi = idx.data() + T.o_xforce_i + f * MOCOB200_XFORCE_ISTRIDE;
        if (xi[1] >= 0) keep[xi[1]] = 1;
        if (xi[2] >= 0) keep[xi[2]] = 1;
        if (xi[3] >= 0 && xi[6] < 0) keep[xi[0]] = 1; // applied at the body's own origin
    }
    int merged = 0;
    for (int b = 0; b < nb; ++b) merged += !keep[b];
    if (merged == 0) return 0;
    // pose of every removed body's frame in the frame of the kept body it ends up in: X_host,b
    std::vector<int> host(nb);
    std::vector<M3> Rh(nb);
    std::vector<double> ph(3 * nb, 0.0);
    const double I3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    for (int b = 0; b < nb; ++b) {
        host[b] = b;
        std::memcpy(Rh[b].m, I3, sizeof I3);
        if (keep[b]) continue;
        const double* d = bodyD(b);
        const int par = bodyI(b)[0];
        // X_par,b = X_PF X_BM^-1 : R = R_PF R_BM^T, p = p_PF - R p_BM
        const M3 R = mulMT(d, d + 12);
        double t[3], p[3];
        mulV(R.m, d + 21, t);
        for (int k = 0; k < 3; ++k) p[k] = d[9 + k] - t[k];
        // compose with the parent's own placement (the parent may itself have been merged)
        host[b] = host[par];
        Rh[b] = mulM(Rh[par].m, R.m);
        mulV(Rh[par].m, p, t);
        for (int k = 0; k < 3; ++k) ph[3 * b + k] = ph[3 * par + k] + t[k];
    }
    // mass properties into the host bodies (inertias are stored about the body origin, in the body frame)
    for (int b = 0; b < nb; ++b) {
        if (keep[b]) continue;
        const double* d = bodyD(b);
        double* hd = bodyD(host[b]);
        const double m = d[24];
        double c[3], dvec[3];
        mulV(Rh[b].m, d + 25, c);
        for (int k = 0; k < 3; ++k) dvec[k] = c[k] + ph[3 * b + k]; // the removed body's mass centre in the host frame
        // inertia about its own mass centre, in its own frame
        const double cx = d[25], cy = d[26], cz = d[27];
        double Ic[9] = {d[28] - m * (cy * cy + cz * cz), d[31] + m * cx * cy, d[32] + m * cx * cz,
                        d[31] + m * cx * cy, d[29] - m * (cx * cx + cz * cz), d[33] + m * cy * cz,
                        d[32] + m * cx * cz, d[33] + m * cy * cz, d[30] - m * (cx * cx + cy * cy)};
        const M3 RI = mulM(Rh[b].m, Ic);
        const M3 J = mulMT(RI.m, Rh[b].m); // rotated into the host frame
        const double dd = dvec[0] * dvec[0] + dvec[1] * dvec[1] + dvec[2] * dvec[2];
        const double add[6] = {J.m[0] + m * (dd - dvec[0] * dvec[0]), J.m[4] + m * (dd - dvec[1] * dvec[1]), J.m[8] + m * (dd - dvec[2] * dvec[2]),
                               J.m[1] - m * dvec[0] * dvec[1], J.m[2] - m * dvec[0] * dvec[2], J.m[5] - m * dvec[1] * dvec[2]};
        const double mh = hd[24], mt = mh + m;
        for (int k = 0; k < 6; ++k) hd[28 + k] += add[k];
        for (int k = 0; k < 3; ++k) hd[25 + k] = mt > 0 ? (mh * hd[25 + k] + m * dvec[k]) / mt : 0.0;
        hd[24] = mt;
    }
    // whatever was attached to a removed body moves to its host
    auto moveStation = [&](int b, double* loc) {
        double t[3];
        mulV(Rh[b].m, loc, t);
        for (int k = 0; k < 3; ++k) loc[k] = t[k] + ph[3 * b + k];
    };
    // the kept bodies stay level by level (a body whose parent chain shrank moves up): stable sort by the new tree level
    std::vector<int> newIndex(nb, -1), newLevel(nb, 0);
    int nb2 = 0, maxLevel = 0;
    for (int b = 0; b < nb; ++b) {
        if (!keep[b]) continue;
        ++nb2;
        const int par = bodyI(b)[0];
        newLevel[b] = par < 0 ? 0 : newLevel[host[par]] + 1;
        if (newLevel[b] > maxLevel) maxLevel = newLevel[b];
    }
    {
        int next = 0;
        for (int lev = 0; lev <= maxLevel; ++lev)
            for (int b = 0; b < nb; ++b) if (keep[b] && newLevel[b] == lev) newIndex[b] = next++;
    }
    auto remap = [&](int b) { return b < 0 ? b : newIndex[host[b]]; };
    for (int p = 0; p < T.npp; ++p) {
        int* pi = idx.data() + T.o_pp_i + p * MOCOB200_PP_ISTRIDE;
        double* pd = dbl.data() + T.o_pp_d + p * MOCOB200_PP_DSTRIDE;
        if (pi[0] >= 0 && !keep[pi[0]]) moveStation(pi[0], pd);
        pi[0] = remap(pi[0]);
    }
    for (int c = 0; c < T.ncontact; ++c) {
        int* ci = idx.data() + T.o_contact_i + c * MOCOB200_CONTACT_ISTRIDE;
        double* cd = dbl.data() + T.o_contact_d + c * MOCOB200_CONTACT_DSTRIDE;
        if (ci[0] >= 0 && !keep[ci[0]]) moveStation(ci[0], cd);
        if (ci[1] >= 0 && !keep[ci[1]]) {
            const M3 R = mulM(Rh[ci[1]].m, cd + 4);
            std::memcpy(cd + 4, R.m, sizeof R.m);
            moveStation(ci[1], cd + 13);
        }
        ci[0] = remap(ci[0]);
        ci[1] = remap(ci[1]);
    }
    for (int f = 0; f < T.nxforce; ++f) {
        int* xi = idx.data() + T.o_xforce_i + f * MOCOB200_XFORCE_ISTRIDE;
        for (int k = 0; k < 3; ++k) xi[k] = remap(xi[k]);
    }
    for (int w = 0; w < (T.npathwrap > 0 ? T.nwrapobj : 0); ++w) { // wrap cylinders: frame in the host body
        int* wi = idx.data() + T.o_wo_i + w * MOCOB200_WO_ISTRIDE;
        double* wd = dbl.data() + T.o_wo_d + w * MOCOB200_WO_DSTRIDE;
        if (wi[0] >= 0 && !keep[wi[0]]) {
            const M3 R = mulM(Rh[wi[0]].m, wd);
            std::memcpy(wd, R.m, sizeof R.m);
            moveStation(wi[0], wd + 9);
        }
        wi[0] = remap(wi[0]);
    }
    // kept bodies whose parent was removed hang on the parent's host: X_PF' = X_host,parent X_PF
    std::vector<double> newD((size_t)nb2 * MOCOB200_BODY_DSTRIDE);
    std::vector<int> newI((size_t)nb2 * MOCOB200_BODY_ISTRIDE), newCj(nb2, -1);
    for (int b = 0; b < nb; ++b) {
        if (!keep[b]) continue;
        double* d = bodyD(b);
        int* bi = bodyI(b);
        const int par = bi[0];
        if (par >= 0 && !keep[par]) {
            const M3 R = mulM(Rh[par].m, d);
            double t[3];
            mulV(Rh[par].m, d + 9, t);
            std::memcpy(d, R.m, sizeof R.m);
            for (int k = 0; k < 3; ++k) d[9 + k] = t[k] + ph[3 * par + k];
        }
        const int nbIdx = newIndex[b];
        std::memcpy(newD.data() + (size_t)nbIdx * MOCOB200_BODY_DSTRIDE, d, sizeof(double) * MOCOB200_BODY_DSTRIDE);
        int* ni = newI.data() + (size_t)nbIdx * MOCOB200_BODY_ISTRIDE;
        std::memcpy(ni, bi, sizeof(int) * MOCOB200_BODY_ISTRIDE);
        ni[0] = remap(par);
        if (T.ncustom > 0) newCj[nbIdx] = idx[T.o_body_cj + b];
    }
    std::memcpy(dbl.data() + T.o_body_d, newD.data(), newD.size() * sizeof(double));
    std::memcpy(idx.data() + T.o_body_i, newI.data(), newI.size() * sizeof(int));
    if (T.ncustom > 0) std::memcpy(idx.data() + T.o_body_cj, newCj.data(), newCj.size() * sizeof(int));
    out.nb = nb2;
    return merged;
}

} // namespace mocob200
