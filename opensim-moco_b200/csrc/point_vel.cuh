// Thread-per-evaluation point function for inputs that leave the coordinates untouched (speeds; in implicit mode the
// acceleration variables): everything that depends on q alone - body poses, joint axes, path-point positions, the active
// path points, the LDL^T factor of the mass matrix - is READ from the blob the nominal warp evaluation of the grid point
// saved (kernels_warp.cu: blob_save), shared by all such evaluations of the point and served from L1/L2; the thread keeps only
// the velocity-dependent quantities (12 doubles per body + the wrench accumulators). Compared with a whole thread
// evaluation (point_eval.cuh: articulated-body algorithm, ~6 KB of thread-local scratch) this is a recursive Newton-Euler
// pass + one substitution with the stored factor and ~2 KB of scratch.
// Same equations as point_warp.cuh / point_warp_dyn.cuh (wk_body_velocity, wk_muscle_dynamics, wk_contact, wk_gather_*,
// wk_inward, wk_ldlt_solve), one thread instead of one warp; reference: MocoCasOCProblem.h:203-297.
#pragma once
#include "point_warp_local.cuh"

namespace mocob200 {

template <class SZ>
__device__ inline void eval_point_vel(const DevTables& T, const WarpAux& A, const double* __restrict__ blobz /* blob - blob_begin */,
                                      const double* in, double* out) {
    const int nq = T.nq, nz = T.nz, nc = T.nc, nb = T.nb;
    const double* u = in + 1 + nq;
    const double* z = u + nq;
    const double* ctl = z + nz;
    const double* q = in + 1;
    const double* mult = ctl + nc;
    const double* der = mult + T.nm;
    const double* auxder = der + (T.implicit ? nq : 0);
    const double* udot = T.implicit ? der : nullptr;
    const double* R = blobz + A.o_R;
    const double* P = blobz + A.o_p;
    const double* OM = blobz + A.o_oM;
    const double* AX = blobz + A.o_a;
    const double* LX = blobz + A.o_l;
    const double* CG = blobz + A.o_cg;
    const double* IO = blobz + A.o_Io;
    const double* PPOS = blobz + A.o_ppos;
    const double* PREL = blobz + A.o_prel;
    const double* DPL = blobz + A.o_dpl;
    const double* REF = blobz + A.o_ref;
    V3 w[SZ::NB], v[SZ::NB], al[SZ::NB], ac[SZ::NB]; // spatial velocity / velocity-product (+ udot) acceleration of the body origins
    V3 Fb[SZ::NB], Nb[SZ::NB];                        // applied wrench about the body origin, then the joint wrench
    double tau[SZ::NQ];
    // ---- velocities, base to tip (wk_body_velocity)
    for (int b = 0; b < nb; ++b) {
        const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
        const int par = bi[0], kind = bi[1], q0 = bi[2], nd = bi[3];
        V3 wP = mk(0, 0, 0), vP = mk(0, 0, 0), pP = mk(0, 0, 0), alP = mk(0, 0, 0), aP = mk(0, 0, 0);
        if (par >= 0) { pP = ldv(P + 3 * par); wP = w[par]; vP = v[par]; alP = al[par]; aP = ac[par]; }
        V3 wrel = mk(0, 0, 0), vrel = mk(0, 0, 0), wrd = mk(0, 0, 0), vrd = mk(0, 0, 0);
        for (int k = 0; k < nd; ++k) {
            const V3 a = ldv(AX + 3 * (q0 + k)), l = ldv(LX + 3 * (q0 + k));
            wrel = wrel + u[q0 + k] * a;
            vrel = vrel + u[q0 + k] * l;
            if (udot) { wrd = wrd + udot[q0 + k] * a; vrd = vrd + udot[q0 + k] * l; }
        }
        V3 bw = mk(0, 0, 0), bv = mk(0, 0, 0);
        if (kind == MOCOB200_JOINT_CUSTOM) { // velocity-product acceleration of the function-based joint (point_warp.cuh)
            const int cj = T.idx[T.o_body_cj + b];
            const int* rec = T.idx + T.o_cj_i + cj * MOCOB200_CJ_ISTRIDE;
            const double* cs = blobz + A.o_cjs + 30 * cj;
            V3 ww[3];
            for (int i = 0; i < 3; ++i) {
                const double ui = rec[3 * i] != MOCOB200_FN_CONSTANT ? u[rec[3 * i + 1]] : 0.0;
                const V3 g = ldv(cs + 3 * i);
                ww[i] = (cs[18 + i] * ui) * g;
                bw = bw + (cs[24 + i] * ui * ui) * g;
            }
            bw = bw + cross(ww[0], ww[1]) + cross(cj_one_coordinate(rec) ? ww[0] : ww[0] + ww[1], ww[2]); // point_eval.cuh: custom_joint_bias
            for (int i = 3; i < 6; ++i) {
                const double ui = rec[3 * i] != MOCOB200_FN_CONSTANT ? u[rec[3 * i + 1]] : 0.0;
                bv = bv + (cs[24 + i] * ui * ui) * ldv(cs + 3 * i);
            }
        }
        const V3 oM = ldv(OM + 3 * b), pB = ldv(P + 3 * b);
        const V3 wB = wP + wrel;
        const V3 r = oM - pP, sv = pB - oM;
        w[b] = wB;
        v[b] = vP + cross(wP, r) + vrel + cross(wB, sv);
        const V3 alB = alP + cross(wP, wrel) + wrd + bw;
        const V3 aM = aP + cross(alP, r) + cross(wP, cross(wP, r)) + 2.0 * cross(wP, vrel) + vrd + bv;
        al[b] = alB;
        ac[b] = aM + cross(alB, sv) + cross(wB, cross(wB, sv));
        // applied wrench starts with gravity at the mass centre (Model.cpp:587-590)
        const double m = T.dbl[T.o_body_d + b * MOCOB200_BODY_DSTRIDE + 24];
        const V3 g = ld3(T.dbl + T.o_gravity);
        Fb[b] = m * g;
        Nb[b] = cross(ldv(CG + 3 * b), m * g);
    }
    for (int k = 0; k < nq; ++k) tau[k] = 0.0;
    // ---- muscles (wk_muscle_dynamics): speeds from the new velocities, geometry from the blob
    double* zdot = out + nq;
    double* auxres = zdot + nz;
    for (int m = 0; m < T.nmus; ++m) {
        const int* mi = T.idx + T.o_mus_i + m * MOCOB200_MUS_ISTRIDE;
        const double* md = T.dbl + T.o_mus_d + m * MOCOB200_MUS_DSTRIDE;
        const int first = mi[0], count = mi[1], flags = mi[5];
        const unsigned active = (unsigned)blobz[A.o_mact + m];
        double length = 0, speed = 0;
        V3 prevP = mk(0, 0, 0), prevV = mk(0, 0, 0);
        int na = 0;
        for (int k = 0; k < count; ++k) {
            if (!((active >> k) & 1u)) continue;
            const int pp = first + k;
            const int* pi = T.idx + T.o_pp_i + pp * MOCOB200_PP_ISTRIDE;
            const int b = pi[0];
            const V3 pG = ldv(PPOS + 3 * pp);
            V3 vG = mk(0, 0, 0);
            if (b >= 0) vG = v[b] + cross(w[b], ldv(PREL + 3 * pp));
            if (pi[1] == MOCOB200_PP_MOVING) {
                V3 vloc = u[pi[6]] * ldv(DPL + 3 * A.pp_mpp[pp]);
                if (b >= 0) vloc = rmul(R + 9 * b, vloc);
                vG = vG + vloc;
            }
            if (na > 0) {
                const V3 r = prevP - pG, dv = prevV - vG;
                const double d = sqrt(dot(r, r));
                length += d;
                speed += (d < 2.220446049250313e-16) ? sqrt(dot(dv, dv)) : dot(dv, (1.0 / d) * r);
            }
            prevP = pG;
            prevV = vG;
            ++na;
        }
        const double activation = (flags & MOCOB200_MUS_FLAG_IGNORE_ACTIVATION_DYNAMICS) ? ctl[mi[2]] : z[mi[3]];
        const double ntf = (mi[4] >= 0) ? z[mi[4]] : NAN;
        const double ntfd = (mi[6] >= 0) ? auxder[mi[6]] : NAN;
        const MuscleOut o = dgf_muscle(md, flags, length, speed, activation, ntf, ntfd, false);
        if (mi[3] >= 0) zdot[mi[3]] = activation_dot(ctl[mi[2]], z[mi[3]], md[5], md[6]);
        if (mi[4] >= 0) zdot[mi[4]] = o.nfDeriv;
        if (mi[7] >= 0) auxres[mi[7]] = o.residual;
        // equal and opposite forces on the two bodies of every inter-body segment (GeometryPath.cpp:306-432)
        int prev = -1;
        for (int k = 0; k < count; ++k) {
            if (!((active >> k) & 1u)) continue;
            if (prev >= 0) {
                const int pa = first + prev, pb = first + k;
                const int* ia = T.idx + T.o_pp_i + pa * MOCOB200_PP_ISTRIDE;
                const int* ib = T.idx + T.o_pp_i + pb * MOCOB200_PP_ISTRIDE;
                if (ia[0] != ib[0]) {
                    const V3 Pa = ldv(PPOS + 3 * pa), Pb = ldv(PPOS + 3 * pb);
                    V3 dir = Pb - Pa;
                    const double nn = sqrt(dot(dir, dir));
                    dir = (nn < 1.9984014443252818e-14) ? (NAN * dir) : ((1.0 / nn) * dir);
                    const V3 f = o.tendonForce * dir;
                    if (ia[0] >= 0) { Fb[ia[0]] = Fb[ia[0]] + f; Nb[ia[0]] = Nb[ia[0]] + cross(ldv(PREL + 3 * pa), f); }
                    if (ib[0] >= 0) { Fb[ib[0]] = Fb[ib[0]] - f; Nb[ib[0]] = Nb[ib[0]] - cross(ldv(PREL + 3 * pb), f); }
                    if (ia[1] == MOCOB200_PP_MOVING) {
                        const V3 dpl = ldv(DPL + 3 * A.pp_mpp[pa]);
                        tau[ia[6]] += dot(ia[0] >= 0 ? rmul(R + 9 * ia[0], dpl) : dpl, f);
                    }
                    if (ib[1] == MOCOB200_PP_MOVING) {
                        const V3 dpl = ldv(DPL + 3 * A.pp_mpp[pb]);
                        tau[ib[6]] -= dot(ib[0] >= 0 ? rmul(R + 9 * ib[0], dpl) : dpl, f);
                    }
                }
            }
            prev = k;
        }
    }
    // ---- contacts (wk_contact) with the new velocities
    for (int ci = 0; ci < T.ncontact; ++ci) {
        const double* cd = T.dbl + T.o_contact_d + ci * MOCOB200_CONTACT_DSTRIDE;
        const int* cidx = T.idx + T.o_contact_i + ci * MOCOB200_CONTACT_ISTRIDE;
        const int bs = cidx[0], bh = cidx[1];
        const double radius = cd[3];
        const double* pr = cd + 16;
        const double diss = pr[1], us = pr[2], ud = pr[3], uv = pr[4], vt = pr[5], cf = pr[6], bd = pr[7], bvv = pr[8];
        V3 nrm, oH;
        {
            const V3 xh = mk(cd[4], cd[7], cd[10]);
            const V3 ph = ld3(cd + 13);
            if (bh >= 0) { nrm = -1.0 * rmul(R + 9 * bh, xh); oH = ldv(P + 3 * bh) + rmul(R + 9 * bh, ph); }
            else { nrm = -1.0 * xh; oH = ph; }
        }
        const V3 centre = (bs >= 0) ? ldv(P + 3 * bs) + rmul(R + 9 * bs, ld3(cd)) : ld3(cd);
        const double indentation = radius - dot(centre - oH, nrm);
        const V3 cp = centre - (radius - 0.5 * indentation) * nrm;
        const V3 vs = bs >= 0 ? v[bs] + cross(w[bs], cp - ldv(P + 3 * bs)) : mk(0, 0, 0);
        const V3 vh = bh >= 0 ? v[bh] + cross(w[bh], cp - ldv(P + 3 * bh)) : mk(0, 0, 0);
        const V3 vrel = vs - vh;
        const double vnormal = dot(vrel, nrm);
        const V3 vtan = vrel - vnormal * nrm;
        const double indVel = -vnormal;
        const double ind = sqrt(indentation * indentation + cf);
        const double fH = cd[DEV_CONTACT_HERTZ] * (ind * sqrt(ind));
        const double fHs = fH * (0.5 + 0.5 * tanh(bd * indentation)) + cf;
        const double fHC = fHs * (1.0 + 1.5 * diss * indVel);
        const double fn = fHC * (0.5 + 0.5 * tanh(bvv * (indVel + 2.0 / (3.0 * diss))));
        const double vslip = sqrt(dot(vtan, vtan) + cf);
        const double vr = vslip / vt;
        const double ff = fn * (fmin(vr, 1.0) * (ud + 2.0 * (us - ud) / (1.0 + vr * vr)) + uv * vslip);
        const V3 f = fn * nrm + (-ff / vslip) * vtan;
        if (bs >= 0) { Fb[bs] = Fb[bs] + f; Nb[bs] = Nb[bs] + cross(cp - ldv(P + 3 * bs), f); }
        if (bh >= 0) { Fb[bh] = Fb[bh] - f; Nb[bh] = Nb[bh] - cross(cp - ldv(P + 3 * bh), f); }
    }
    // ---- external forces (functions of time, unchanged: their values sit in the blob's reference table)
    for (int f = 0; f < T.nxforce; ++f) {
        const int* xi = T.idx + T.o_xforce_i + f * MOCOB200_XFORCE_ISTRIDE;
        const int b = xi[0];
        if (xi[3] >= 0) {
            V3 force = mk(REF[xi[3]], REF[xi[4]], REF[xi[5]]);
            if (xi[1] >= 0) force = rmul(R + 9 * xi[1], force);
            V3 arm = mk(0, 0, 0);
            if (xi[6] >= 0) {
                V3 point = mk(REF[xi[6]], REF[xi[7]], REF[xi[8]]);
                if (xi[2] >= 0) point = ldv(P + 3 * xi[2]) + rmul(R + 9 * xi[2], point);
                arm = point - ldv(P + 3 * b);
            }
            Fb[b] = Fb[b] + force;
            Nb[b] = Nb[b] + cross(arm, force);
        }
        if (xi[9] >= 0) {
            V3 torque = mk(REF[xi[9]], REF[xi[10]], REF[xi[11]]);
            if (xi[1] >= 0) torque = rmul(R + 9 * xi[1], torque);
            Nb[b] = Nb[b] + torque;
        }
    }
    for (int a = 0; a < T.ncact; ++a) { // CoordinateActuator.cpp:185-204
        const int* ai = T.idx + T.o_cact_i + a * MOCOB200_CACT_ISTRIDE;
        tau[ai[0]] += T.dbl[T.o_cact_d + a * MOCOB200_CACT_DSTRIDE] * ctl[ai[1]];
    }
    coupler_forces(T, q, mult, tau);
    // ---- joint wrenches: inertial - applied, tip to base (wk_gather_body + wk_inward)
    for (int b = nb - 1; b >= 0; --b) {
        const double* bd = T.dbl + T.o_body_d + b * MOCOB200_BODY_DSTRIDE;
        const int* bi = T.idx + T.o_body_i + b * MOCOB200_BODY_ISTRIDE;
        const int par = bi[0], q0 = bi[2], nd = bi[3];
        const double m = bd[24];
        const V3 cg = ldv(CG + 3 * b), wB = w[b], alB = al[b], aB = ac[b];
        const double* Io = IO + 6 * b;
        // the children have already added their wrenches (shifted to this origin) with the opposite sign convention:
        // Fb/Nb hold applied - children's needs; the joint must supply inertial - that
        const V3 F = m * (aB + cross(alB, cg) + cross(wB, cross(wB, cg))) - Fb[b];
        const V3 N = symmul(Io, alB) + cross(cg, m * aB) + cross(wB, symmul(Io, wB)) - Nb[b];
        const V3 pB = ldv(P + 3 * b), oM = ldv(OM + 3 * b);
        const V3 NM = N + cross(pB - oM, F);
        for (int k = 0; k < nd; ++k) tau[q0 + k] -= dot(ldv(AX + 3 * (q0 + k)), NM) + dot(ldv(LX + 3 * (q0 + k)), F);
        if (par >= 0) {
            const V3 r = pB - ldv(P + 3 * par);
            Fb[par] = Fb[par] - F;
            Nb[par] = Nb[par] - (N + cross(r, F));
        }
    }
    // ---- tau now holds applied - bias: the residual (implicit) or the right-hand side of M udot = tau (explicit)
    if (T.implicit) {
        for (int k = 0; k < nq; ++k) out[k] = tau[k];
    } else {
        wk_ldlt_solve_lane(blobz + A.o_M, nq, tau);
        for (int k = 0; k < nq; ++k) out[k] = tau[k];
    }
    for (int k = 0; k < T.ncoupler; ++k) coupler_errors(T, k, q, u, T.implicit ? der : out, out + nq + nz + T.n_auxres);
}

} // namespace mocob200
