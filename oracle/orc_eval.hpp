// TEST INFRASTRUCTURE ONLY — CPU oracle: applied forces, the fused point function and the endpoint
// (goal) functions.
#pragma once
#include <stdexcept>

#include "orc_dynamics.hpp"
#include "orc_wrap.hpp"

namespace orc {

// SimTK::SmoothSphereHalfSpaceForce (Simbody 3.7; source NOT in the reference tree). Pinned against data the reference
// produced with it — example2DWalking's referenceGRF.sto along referenceCoordinates.sto: normal force 0.25 % rms of peak,
// friction loosely (the files hold no speeds; tests/test_oracle_known_answers.py) — but not to round-off: bit-level parity
// with Simbody stays UNPINNED.
// Restated from the model description cited by
// opensim-core/OpenSim/Simulation/Model/SmoothSphereHalfSpaceForce.h:28-66 (Serrancoli et al. 2019) and the
// parameter plumbing in SmoothSphereHalfSpaceForce.cpp:50-94. The half space fills +X of its frame, so the
// contact normal is the frame's -X.  Returns the force on the sphere body and the point of application.
template <class S>
Vec3<S> contactForce(const Model& M, int ci, const Frames<S>& K, Vec3<S>& applicationPoint) {
    const double* cd = M.contactD(ci);
    const int* cidx = M.contactI(ci);
    const int bs = cidx[0], bh = cidx[1];
    const double radius = cd[3];
    const double stiffness = cd[16], dissipation = cd[17], muS = cd[18], muD = cd[19], muV = cd[20], vT = cd[21],
                 cf = cd[22], bd = cd[23], bv = cd[24];
    const Mat3<S> R_BH = Mat3<S>::fromRowMajor(cd + 4);
    const Vec3<S> pH{S(cd[13]), S(cd[14]), S(cd[15])};
    Mat3<S> R_GH = R_BH;
    if (bh >= 0) R_GH = K.R[bh] * R_BH;
    const Vec3<S> planeOrigin = stationPosition(K, bh, pH);
    const Vec3<S> normal = S(-1.0) * R_GH.col(0);
    const Vec3<S> centre = stationPosition(K, bs, Vec3<S>(S(cd[0]), S(cd[1]), S(cd[2])));
    const S indentation = S(radius) - dot(centre - planeOrigin, normal);
    // contact point: bottom of the sphere moved half the indentation back towards the centre
    const Vec3<S> cp = centre - (S(radius) - S(0.5) * indentation) * normal;
    const Vec3<S> v = pointVelocity(K, bs, cp) - pointVelocity(K, bh, cp);
    const S vn = dot(v, normal);
    const Vec3<S> vt = v - vn * normal;
    const S indentationRate = -vn;
    const S k = S(0.5) * pow(S(stiffness), S(2.0 / 3.0));
    const S fHertz = S(4.0 / 3.0) * k * sqrt(S(radius) * k) * pow(sqrt(indentation * indentation + S(cf)), S(1.5));
    const S fHertzSmooth = fHertz * (S(0.5) + S(0.5) * tanh(S(bd) * indentation)) + S(cf);
    const S fHC = fHertzSmooth * (S(1.0) + S(1.5) * S(dissipation) * indentationRate);
    const S fn = fHC * (S(0.5) + S(0.5) * tanh(S(bv) * (indentationRate + S(2.0 / (3.0 * dissipation)))));
    const S vslip = sqrt(dot(vt, vt) + S(cf));
    const S vrel = vslip / S(vT);
    const S ff = fn * (fmin(vrel, S(1.0)) * (S(muD) + S(2.0 * (muS - muD)) / (S(1.0) + vrel * vrel)) + S(muV) * vslip);
    applicationPoint = cp;
    return fn * normal + (-ff / vslip) * vt;
}

// Path geometry of one muscle (GeometryPath.cpp:817-864, 1117-1145; Point.cpp:130-140;
// ConditionalPathPoint.cpp:160-169; MovingPathPoint.cpp:216-315, 374-392), with wrap cylinders
// (GeometryPath.cpp:868-1094 applyWrapObjects, :1095-1110 calcPathLengthChange, :1117-1145; orc_wrap.hpp).
template <class S>
struct PathState {
    static constexpr int kMaxPts = 40;
    int n = 0;
    Vec3<S> pos[kMaxPts], dPdq[kMaxPts]; // ground
    int body[kMaxPts], movingCoord[kMaxPts];
    S length, speed;
};
template <class S>
void pathGeometry(const Model& M, int m, const S* q, const S* u, const Frames<S>& K, PathState<S>& P) {
    const int* mi = M.musI(m);
    constexpr int kMax = PathState<S>::kMaxPts;
    // ---- the active path points, in PathPointSet order
    int nA = 0, orig[kMax], bodyA[kMax], mcA[kMax];
    Vec3<S> posA[kMax], velA[kMax], dpA[kMax];
    for (int k = 0; k < mi[1]; ++k) {
        const int* pi = M.ppI(mi[0] + k);
        const double* pd = M.ppD(mi[0] + k);
        const int b = pi[0], kind = pi[1];
        Vec3<S> loc{S(pd[0]), S(pd[1]), S(pd[2])}, vloc, dp;
        int mc = -1;
        if (kind == MOCOB200_PP_CONDITIONAL) {
            const double v = val(q[pi[2]]);
            if (!(v >= pd[3] - 1e-5 && v <= pd[4] + 1e-5)) continue; // isActive
        } else if (kind == MOCOB200_PP_MOVING) {
            mc = pi[6];
            const S qv = q[mc];
            S qc = qv; // getLocation clamps the coordinate to its range; the derivatives do not
            if (val(qv) < pd[3]) qc = S(pd[3]);
            if (val(qv) > pd[4]) qc = S(pd[4]);
            loc = Vec3<S>(simmSpline(M, pi[3], qc, false), simmSpline(M, pi[4], qc, false), simmSpline(M, pi[5], qc, false));
            dp = Vec3<S>(simmSpline(M, pi[3], qv, true), simmSpline(M, pi[4], qv, true), simmSpline(M, pi[5], qv, true));
            vloc = u[mc] * dp;
        }
        const Vec3<S> pG = stationPosition(K, b, loc);
        Vec3<S> vG = pointVelocity(K, b, pG);
        if (mc >= 0) {
            if (b >= 0) { vG = vG + K.R[b] * vloc; dp = K.R[b] * dp; }
            else vG = vG + vloc;
        }
        orig[nA] = k; bodyA[nA] = b; mcA[nA] = mc; posA[nA] = pG; velA[nA] = vG; dpA[nA] = dp;
        ++nA;
    }
    // ---- the current path: indices >= 0 are active points, -(2 w + 1) / -(2 w + 2) the two wrap points of pairing w
    int path[kMax], np = nA;
    for (int k = 0; k < nA; ++k) path[k] = k;
    const int nW = M.T->npathwrap > 0 ? M.pathWrapCount(m) : 0;
    constexpr int kMaxW = 8;
    struct WrapPts { bool on = false; wrap::V r1, r2; double arc = 0; } W[kMaxW]; // tangent points in the wrap object's BODY frame
    auto groundOf = [&](int b, const wrap::V& local) -> wrap::V { // double-precision station position
        if (b < 0) return local;
        const Vec3<S> g = stationPosition(K, b, Vec3<S>(S(local[0]), S(local[1]), S(local[2])));
        return {val(g.x), val(g.y), val(g.z)};
    };
    auto bodyFrameOf = [&](int b, const wrap::V& g) -> wrap::V { // ground point expressed in body b
        if (b < 0) return g;
        const Vec3<S> d{S(g[0]) - K.origin[b].x, S(g[1]) - K.origin[b].y, S(g[2]) - K.origin[b].z};
        const Vec3<S> l = transpose(K.R[b]) * d;
        return {val(l.x), val(l.y), val(l.z)};
    };
    auto woOfPath = [&](int e) { return e >= 0 ? -1 : M.pathWrapI(m, (-e - 1) / 2)[0]; }; // wrap object of a path entry (-1: none)
    auto groundPoint = [&](int e) -> wrap::V {
        if (e >= 0) return {val(posA[e].x), val(posA[e].y), val(posA[e].z)};
        const int w = (-e - 1) / 2;
        return groundOf(M.wrapObjI(M.pathWrapI(m, w)[0])[0], ((-e - 1) % 2 == 0) ? W[w].r1 : W[w].r2);
    };
    auto dist = [](const wrap::V& a, const wrap::V& b) { return std::sqrt(wrap::dist2(a, b)); };
    auto removeWrap = [&](int w) {
        for (int j = 0; j < np; ++j)
            if (path[j] == -(2 * w + 1)) { for (int k = j; k + 2 < np; ++k) path[k] = path[k + 2]; np -= 2; return; }
    };
    auto totalLength = [&]() {
        double len = 0;
        for (int k = 0; k + 1 < np; ++k) {
            const int a = path[k], b = path[k + 1];
            if (a < 0 && b < 0 && woOfPath(a) == woOfPath(b)) len += W[(-b - 1) / 2].arc;
            else len += dist(groundPoint(a), groundPoint(b));
        }
        return len;
    };
    if (nW > kMaxW) throw std::runtime_error("oracle: more path wraps on one muscle than kMaxW");
    if (nW > 0) {
        int order[kMaxW], result[kMaxW];
        for (int i = 0; i < nW; ++i) order[i] = i;
        const int maxIterations = nW < 2 ? 1 : 8;
        double lastLength = INFINITY;
        bool abort = false;
        for (int kk = 0; kk < maxIterations && !abort; ++kk) {
            for (int i = 0; i < nW && !abort; ++i) {
                result[i] = wrap::NoWrap;
                const int w = order[i];
                const int* pw = M.pathWrapI(m, w);
                const int* wi = M.wrapObjI(pw[0]);
                const double* wd = M.wrapObjD(pw[0]);
                bool haveBest = false;
                wrap::CylinderWrap best;
                int bestEnd = -1;
                double minChange = INFINITY;
                removeWrap(w);
                W[w].on = false;
                if (!wi[3]) continue; // inactive wrap object
                const int wrapStart = pw[1] < 1 ? 0 : pw[1] - 1, wrapEnd = pw[2] < 1 ? mi[1] - 1 : pw[2] - 1;
                int first = -1, last = -1; // active points (indices into the active list) at the ends of the range
                for (int k = 0; k < nA; ++k) if (orig[k] >= wrapStart && orig[k] <= wrapEnd) { if (first < 0) first = k; last = k; }
                if (first < 0) { abort = true; break; }
                int start = -1, end = -1;
                for (int j = 0; j < np; ++j) { if (path[j] == first) start = j; if (path[j] == last) end = j; }
                if (start < 0 || end < 0) { abort = true; break; }
                auto toWrapFrame = [&](const wrap::V& g) {
                    const wrap::V b = bodyFrameOf(wi[0], g);
                    const wrap::V d{b[0] - wd[9], b[1] - wd[10], b[2] - wd[11]};
                    return wrap::V{wd[0] * d[0] + wd[3] * d[1] + wd[6] * d[2], wd[1] * d[0] + wd[4] * d[1] + wd[7] * d[2],
                                   wd[2] * d[0] + wd[5] * d[1] + wd[8] * d[2]}; // R_BW^T d
                };
                for (int pt1 = start; pt1 < end; ++pt1) {
                    const int pt2 = pt1 + 1;
                    const int wa = woOfPath(path[pt1]), wb = woOfPath(path[pt2]);
                    if (wa >= 0 && wb >= 0 && wa == wb) continue;
                    const wrap::V g1 = groundPoint(path[pt1]), g2 = groundPoint(path[pt2]);
                    wrap::CylinderWrap cw;
                    const wrap::V a1 = toWrapFrame(g1), a2 = toWrapFrame(g2);
                    result[i] = wrap::cylinderWrapLine(wd[12], wd[13], wi[1], wi[2], a1, a2, cw);
                    if (result[i] == wrap::MandatoryWrap) { best = cw; bestEnd = pt2; haveBest = true; break; }
                    if (result[i] == wrap::Wrapped) {
                        const double change = dist(a1, cw.r1) + cw.arc + dist(a2, cw.r2) - dist(g1, g2);
                        if (change < minChange) { best = cw; bestEnd = pt2; haveBest = true; minChange = change; }
                    }
                }
                if (!haveBest) continue;
                auto toBody = [&](const wrap::V& r) { // wrap frame -> the wrap object's body frame
                    return wrap::V{wd[0] * r[0] + wd[1] * r[1] + wd[2] * r[2] + wd[9], wd[3] * r[0] + wd[4] * r[1] + wd[5] * r[2] + wd[10],
                                   wd[6] * r[0] + wd[7] * r[1] + wd[8] * r[2] + wd[11]};
                };
                W[w].on = true; W[w].r1 = toBody(best.r1); W[w].r2 = toBody(best.r2); W[w].arc = best.arc;
                if (np + 2 > kMax) throw std::runtime_error("oracle: path longer than kMaxPts");
                for (int k = np - 1; k >= bestEnd; --k) path[k + 2] = path[k];
                path[bestEnd] = -(2 * w + 1);
                path[bestEnd + 1] = -(2 * w + 2);
                np += 2;
            }
            if (abort) break;
            const double len = totalLength();
            if (std::fabs(len - lastLength) < 0.0005) break;
            lastLength = len;
            if (kk == 0 && nW > 1 && result[0] == wrap::NoWrap && result[1] == wrap::InsideRadius) {
                order[0] = 1; order[1] = 0;
                removeWrap(0);
                W[0].on = false;
            }
        }
    }
    // ---- length, lengthening speed and the point list the tension is applied through
    P.n = 0; P.length = S(0.0); P.speed = S(0.0);
    Vec3<S> prevP, prevV;
    int prevEntry = 0;
    for (int j = 0; j < np; ++j) {
        const int e = path[j];
        Vec3<S> pG, vG, dp;
        int b, mc = -1;
        if (e >= 0) { pG = posA[e]; vG = velA[e]; dp = dpA[e]; b = bodyA[e]; mc = mcA[e]; }
        else { // a wrap point: a station of the wrap object's body (PathWrapPoint)
            const int w = (-e - 1) / 2;
            const wrap::V& r = ((-e - 1) % 2 == 0) ? W[w].r1 : W[w].r2;
            b = M.wrapObjI(M.pathWrapI(m, w)[0])[0];
            pG = stationPosition(K, b, Vec3<S>(S(r[0]), S(r[1]), S(r[2])));
            vG = pointVelocity(K, b, pG);
        }
        if (P.n > 0) {
            const Vec3<S> r = prevP - pG, dv = prevV - vG;
            const S d = norm(r);
            if (e < 0 && prevEntry < 0 && woOfPath(e) == woOfPath(prevEntry)) P.length += S(W[(-e - 1) / 2].arc); // over the surface
            else P.length += d;
            if (val(d) < 2.220446049250313e-16) P.speed += norm(dv);
            else P.speed += dot(dv, (S(1.0) / d) * r);
        }
        prevP = pG; prevV = vG; prevEntry = e;
        P.pos[P.n] = pG; P.body[P.n] = b; P.movingCoord[P.n] = mc; P.dPdq[P.n] = dp;
        ++P.n;
    }
}
// GeometryPath::addInEquivalentForces (GeometryPath.cpp:306-432)
template <class S>
void applyPathTension(const PathState<S>& P, S tension, Wrench<S>* applied, S* mobilityForce) {
    for (int k = 0; k + 1 < P.n; ++k) {
        if (P.body[k] == P.body[k + 1]) continue;
        Vec3<S> dir = P.pos[k + 1] - P.pos[k];
        const S len = norm(dir);
        if (val(len) < 1.9984014443252818e-14) dir = S(NAN) * dir; // SimTK::SignificantReal
        else dir = (S(1.0) / len) * dir;
        const Vec3<S> force = tension * dir;
        if (P.body[k] >= 0) applied[P.body[k]].addForceAt(P.pos[k], force);
        if (P.body[k + 1] >= 0) applied[P.body[k + 1]].addForceAt(P.pos[k + 1], S(-1.0) * force);
        if (P.movingCoord[k] >= 0) mobilityForce[P.movingCoord[k]] += dot(P.dPdq[k], force);
        if (P.movingCoord[k + 1] >= 0) mobilityForce[P.movingCoord[k + 1]] -= dot(P.dPdq[k + 1], force);
    }
}

template <class S>
MuscleResult<S> evalMuscle(const Model& M, int m, const S* q, const S* u, const S* z, const S* ctl, const S* auxDeriv,
                           const Frames<S>& K, PathState<S>& P, bool extras) {
    const int* mi = M.musI(m);
    pathGeometry(M, m, q, u, K, P);
    const int flags = mi[5];
    if (flags & MOCOB200_MUS_FLAG_PATH_ACTUATOR) { // PathActuator::computeActuation: optimal_force x control (PathActuator.cpp:156-165)
        MuscleResult<S> r{};
        r.tendonForce = S(M.musD(m)[0]) * ctl[mi[2]];
        r.equilibriumResidual = S(0.0); r.normTendonForceRate = S(0.0); r.linearizedResidualRate = S(0.0);
        return r;
    }
    const S activation = (flags & MOCOB200_MUS_FLAG_IGNORE_ACTIVATION_DYNAMICS) ? ctl[mi[2]] : z[mi[3]];
    const S ntf = mi[4] >= 0 ? z[mi[4]] : S(NAN);
    const S ntfRate = mi[6] >= 0 ? auxDeriv[mi[6]] : S(NAN);
    return muscleModel(M.musD(m), flags, P.length, P.speed, activation, ntf, ntfRate, extras);
}

// Model::calcMassCenterPosition
template <class S>
Vec3<S> massCenter(const Model& M, const Frames<S>& K) {
    Vec3<S> acc; S total(0.0);
    for (int b = 0; b < M.nb; ++b) {
        const S m(M.bodyD(b)[24]);
        acc = acc + m * (K.origin[b] + K.comOffset[b]);
        total += m;
    }
    return (S(1.0) / total) * acc;
}

// CoordinateCouplerConstraint k (SimbodyEngine/CoordinateCouplerConstraint.cpp:40-90 CompoundFunction): position error
// scale f(q_i) - q_d and the first two derivatives of scale f at q_i.
template <class S>
void couplerFunction(const Model& M, int k, const S* q, S& err, S& d1, S& d2) {
    const int* ci = M.couplerI(k);
    const double* cd = M.couplerD(k);
    const S qi = q[ci[1]];
    S f;
    if (ci[0] == MOCOB200_FN_LINEAR) { f = S(cd[0]) * qi + S(cd[1]); d1 = S(cd[0]); d2 = S(0.0); }
    else if (ci[0] == MOCOB200_FN_SIMMSPLINE) {
        // value, slope and curvature through a dual number seeded on q_i (slope) and once more on the slope (curvature)
        const Dual<S> v = simmSpline<Dual<S>>(M, ci[3], Dual<S>(qi, S(1.0)), false);
        const Dual<S> s1 = simmSpline<Dual<S>>(M, ci[3], Dual<S>(qi, S(1.0)), true);
        f = v.v; d1 = s1.v; d2 = s1.d;
    } else { f = S(cd[1]); d1 = S(0.0); d2 = S(0.0); }
    err = f - q[ci[2]];
}
// Velocity correction of the Hermite-Simpson midpoints (MocoCasOCProblem.h:298-332; Posa et al. 2016): qdot = u + G^T gamma
// with G the constraint Jacobian: row k = d err_k / dq = (f' at the independent coordinate, -1 at the dependent one).
template <class S>
void velocityCorrection(const Model& M, const S* q, const S* gamma, S* corr) {
    for (int i = 0; i < M.nq; ++i) corr[i] = S(0.0);
    for (int k = 0; k < M.nm; ++k) {
        S e, d1, d2;
        couplerFunction(M, k, q, e, d1, d2);
        const int* ci = M.couplerI(k);
        corr[ci[1]] += d1 * gamma[k];
        corr[ci[2]] -= gamma[k];
    }
}

// The fused point function: in = [time, q, u, z, controls, multipliers, derivatives] -> out = [udot | residual, zdot,
// aux residuals, kinematic-constraint errors, integrand_0 ...]. mask bit 0 = the dynamics callback
// (MultibodySystemExplicit/Implicit::eval, CasOCFunction.cpp:232-245, 317-331 -> MocoCasOCProblem.h:203-297),
// bit 1+k = CostIntegrand k (CasOCFunction.cpp:145-151 -> MocoCasOCProblem.h:333-356).
template <class S>
void evalPoint(const Model& M, const S* in, S* out, unsigned mask = 0xffffffffu) {
    const int nq = M.nq, nz = M.nz, nc = M.nc;
    const S time = in[0];
    const int nqs = M.nqs;
    const S* states = in + 1; // the NLP's state vector at this point (goals index into it)
    const S* q = in + 1; const S* u = q + nqs; const S* z = u + nqs; const S* ctl = z + nz;
    const S* mult = ctl + nc;
    const S* der = mult + M.nm;
    const S* auxDeriv = der + ((M.implicit && !M.prescribed) ? nq : 0);
    S presc[3 * kMaxDof];
    if (M.prescribed) { // q, qdot, qdotdot of every coordinate from its function of time (Components/PositionMotion.cpp:44-75)
        for (int k = 0; k < nq; ++k) {
            const int id = M.T->idx[M.T->o_presc_i + k];
            presc[k] = piecewisePoly(M, id, time);
            presc[nq + k] = piecewisePolyDeriv(M, id, time, 1);
            presc[2 * nq + k] = piecewisePolyDeriv(M, id, time, 2);
        }
        q = presc; u = presc + nq; der = presc + 2 * nq;
    }
    const bool dyn = mask & 1u;
    bool needContact = dyn;
    for (int s = 0; s < M.n_integrand; ++s)
        if (((mask >> (1 + s)) & 1u) && M.goalRec(M.integrandGoals[s])[0] == MOCOB200_GOAL_CONTACT_TRACKING) needContact = true;
    static thread_local Frames<S> K;
    Vec3<S> contactF[kMaxContacts];
    Wrench<S> applied[kMaxBodies];
    S mobility[kMaxDof];
    if (needContact) {
        kinematics(M, q, u, K);
        for (int c = 0; c < M.T->ncontact; ++c) {
            Vec3<S> cp;
            contactF[c] = contactForce(M, c, K, cp);
            if (dyn) {
                const int* ci = M.contactI(c);
                if (ci[0] >= 0) applied[ci[0]].addForceAt(cp, contactF[c]);
                if (ci[1] >= 0) applied[ci[1]].addForceAt(cp, S(-1.0) * contactF[c]);
            }
        }
    }
    if (dyn) {
        for (int k = 0; k < nq; ++k) mobility[k] = S(0.0);
        // gravity: m g at every mass centre (Model.cpp:587-590)
        const double* g = M.T->dbl + M.T->o_gravity;
        for (int b = 0; b < M.nb; ++b) {
            const S m(M.bodyD(b)[24]);
            applied[b].addForceAt(K.origin[b] + K.comOffset[b], m * Vec3<S>(S(g[0]), S(g[1]), S(g[2])));
        }
        // ExternalForce::computeForce (Model/ExternalForce.cpp:330-355): force / torque re-expressed in ground, the point
        // taken from its expressed-in frame to ground; applied to the body at that point
        for (int f = 0; f < M.T->nxforce; ++f) {
            const int* xi = M.xforceI(f);
            const int body = xi[0];
            auto vecAt = [&](int o) { return Vec3<S>(piecewisePoly(M, xi[o], time), piecewisePoly(M, xi[o + 1], time), piecewisePoly(M, xi[o + 2], time)); };
            if (xi[3] >= 0) {
                Vec3<S> force = vecAt(3);
                if (xi[1] >= 0) force = K.R[xi[1]] * force;
                Vec3<S> point = K.origin[body];
                if (xi[6] >= 0) point = stationPosition(K, xi[2], vecAt(6));
                applied[body].addForceAt(point, force);
            }
            if (xi[9] >= 0) {
                Vec3<S> torque = vecAt(9);
                if (xi[1] >= 0) torque = K.R[xi[1]] * torque;
                applied[body].m = applied[body].m + torque;
            }
        }
        S* zdot = out + nq; S* auxres = zdot + nz;
        static thread_local PathState<S> P;
        for (int m = 0; m < M.T->nmus; ++m) {
            const int* mi = M.musI(m); const double* md = M.musD(m);
            const MuscleResult<S> r = evalMuscle(M, m, q, u, z, ctl, auxDeriv, K, P, false);
            applyPathTension(P, r.tendonForce, applied, mobility);
            if (mi[3] >= 0) zdot[mi[3]] = dgf::activationDerivative(ctl[mi[2]], z[mi[3]], md[5], md[6]);
            if (mi[4] >= 0) zdot[mi[4]] = r.normTendonForceRate;
            if (mi[7] >= 0) auxres[mi[7]] = r.equilibriumResidual;
        }
        for (int a = 0; a < M.T->ncact; ++a) { // CoordinateActuator.cpp:185-204
            const int* ai = M.T->idx + M.T->o_cact_i + a * MOCOB200_CACT_ISTRIDE;
            mobility[ai[0]] += S(M.T->dbl[M.T->o_cact_d + a * MOCOB200_CACT_DSTRIDE]) * ctl[ai[1]];
        }
        // kinematic-constraint forces from the solver's multipliers, applied like any other force: G^T (-lambda)
        // (MocoCasOCProblem.h:643-662: calcConstraintForcesFromMultipliers with the multipliers negated)
        for (int k = 0; k < M.nm; ++k) {
            S e, d1, d2;
            couplerFunction(M, k, q, e, d1, d2);
            const int* ci = M.couplerI(k);
            mobility[ci[1]] -= d1 * mult[k];
            mobility[ci[2]] += mult[k];
        }
        S tau[kMaxDof];
        if (M.implicit) {
            // residual = f_app - M udot - f_inertial (MocoCasOCProblem.h:280-285; MocoDeveloperGuide.dox:175-208)
            inverseDynamics(M, K, u, der, applied, true, tau);
            for (int k = 0; k < nq; ++k) out[k] = mobility[k] - tau[k];
        } else {
            S zero[kMaxDof];
            for (int k = 0; k < nq; ++k) zero[k] = S(0.0);
            inverseDynamics(M, K, u, zero, applied, true, tau);
            S rhs[kMaxDof];
            for (int k = 0; k < nq; ++k) rhs[k] = mobility[k] - tau[k];
            static thread_local std::vector<S> Mass;
            Mass.resize((size_t)nq * nq);
            massMatrix(M, K, Mass.data());
            choleskySolve(nq, Mass.data(), rhs, out);
        }
        // kinematic-constraint errors (MocoCasOCProblem.h:664-732): position level, then - with their derivatives enforced -
        // velocity level (G u) and acceleration level (G udot + Gdot u) with the udot of THIS model (explicit: the one just
        // solved for; implicit: the acceleration variables)
        S* kc = out + nq + nz + M.n_auxres;
        const S* udotUsed = M.implicit ? der : out;
        for (int k = 0; k < (M.n_kc > 0 ? M.nm : 0); ++k) {
            S e, d1, d2;
            couplerFunction(M, k, q, e, d1, d2);
            const int* ci = M.couplerI(k);
            kc[k] = e;
            if (M.enforceDerivs) {
                kc[M.nm + k] = d1 * u[ci[1]] - u[ci[2]];
                kc[2 * M.nm + k] = d1 * udotUsed[ci[1]] + d2 * u[ci[1]] * u[ci[1]] - udotUsed[ci[2]];
            }
        }
    }
    // integrands (MocoControlGoal.cpp:121-131, MocoStateTrackingGoal.cpp:103-117,
    // MocoControlTrackingGoal.cpp, MocoSumSquaredStateGoal.cpp, MocoContactTrackingGoal.cpp:240-287)
    S* integ = out + M.n_dyn_out;
    for (int s = 0; s < M.n_integrand; ++s) {
        if (!((mask >> (1 + s)) & 1u)) continue;
        const int* gr = M.goalRec(M.integrandGoals[s]);
        const int* ip = M.T->idx + gr[4];
        const double* dp = M.T->dbl + gr[5];
        S v(0.0);
        switch (gr[0]) {
        case MOCOB200_GOAL_CONTROL: {
            const int expo = ip[0], n = ip[2];
            for (int k = 0; k < n; ++k) {
                const S c = ctl[ip[3 + k]];
                v += S(dp[3 + k]) * (expo == 2 ? c * c : pow(fabs(c), S((double)expo)));
            }
        } break;
        case MOCOB200_GOAL_STATE_TRACKING:
            for (int k = 0; k < ip[0]; ++k) {
                const S e = states[ip[1 + 2 * k]] - piecewisePoly(M, ip[2 + 2 * k], time);
                v += S(dp[3 + k]) * (e * e);
            }
            break;
        case MOCOB200_GOAL_CONTROL_TRACKING:
            for (int k = 0; k < ip[0]; ++k) {
                const S e = ctl[ip[1 + 2 * k]] - piecewisePoly(M, ip[2 + 2 * k], time);
                v += S(dp[3 + k]) * (e * e);
            }
            break;
        case MOCOB200_GOAL_SUM_SQUARED_STATE:
            for (int k = 0; k < ip[0]; ++k) { const S e = states[ip[1 + 2 * k]]; v += S(dp[3 + k]) * (e * e); }
            break;
        case MOCOB200_GOAL_CONTACT_TRACKING: {
            const int proj = ip[0], ng = ip[1];
            const Vec3<S> pv{S(dp[3]), S(dp[4]), S(dp[5])};
            const int* gp = ip + 2;
            for (int k = 0; k < ng; ++k) {
                const int ncon = gp[0];
                Vec3<S> model;
                for (int c = 0; c < ncon; ++c) {
                    const Vec3<S> f = contactF[gp[1 + 2 * c]];
                    model = gp[2 + 2 * c] == 0 ? model + f : model - f; // record offset 0: on sphere; 6: on half space
                }
                gp += 1 + 2 * ncon;
                const Vec3<S> ref(piecewisePoly(M, gp[0], time), piecewisePoly(M, gp[1], time), piecewisePoly(M, gp[2], time));
                gp += 3;
                const Vec3<S> e3 = model - ref;
                Vec3<S> e = e3;
                if (proj == 1) e = dot(e3, pv) * pv;
                else if (proj == 2) e = e3 - dot(e3, pv) * pv;
                v += dot(e, e);
            }
        } break;
        default: break;
        }
        integ[s] = v;
    }
}

// Endpoint function of goal g (MocoGoal::calcGoal, MocoGoal.h:215-248, and each goal's calcGoalImpl):
// Path constraint g at one mesh point: in = [time, states, controls, derivatives] -> errors[n_outputs]
// (CasOC::PathConstraint::eval CasOCFunction.cpp:137-143 -> MocoCasOCProblem.h:463-484).
// Only MocoControlBoundConstraint: error = control - bound_function(time), one equation per (control, bound)
// (MocoControlBoundConstraint.cpp:121-148).
template <class S>
void evalPathConstraint(const Model& M, int g, const S* in, S* errors) {
    const int* gr = M.goalRec(g);
    const int* ip = M.T->idx + gr[4];
    const S time = in[0];
    const S* controls = in + 1 + 2 * M.nqs + M.nz;
    for (int e = 0; e < gr[2]; ++e) errors[e] = controls[ip[2 * e]] - piecewisePoly(M, ip[2 * e + 1], time);
}

// ein = [t0, x0, c0, d0, tf, xf, cf, df, integral] -> vals[n_outputs].
template <class S>
void evalEndpoint(const Model& M, int g, const S* ein, S* vals) {
    const int* gr = M.goalRec(g);
    const int* ip = M.T->idx + gr[4];
    const double* dp = M.T->dbl + gr[5];
    const int ns = 2 * M.nqs + M.nz, nc = M.nc, nd = M.nd, nq = M.nq;
    const int half = 1 + ns + nc + M.nm + nd;
    const S t0 = ein[0], tf = ein[half];
    const S* x0 = ein + 1; const S* c0 = x0 + ns; const S* d0 = c0 + nc + M.nm;
    const S* xf = ein + half + 1; const S* cf = xf + ns;
    const S integral = ein[2 * half];
    const S w(dp[0]);
    const bool cost = gr[1] == MOCOB200_MODE_COST;
    auto displacement = [&]() { // calcSystemDisplacement: |com(final) - com(initial)|
        static thread_local Frames<S> K;
        kinematics(M, x0, x0 + nq, K);
        const Vec3<S> a = massCenter(M, K);
        kinematics(M, xf, xf + nq, K);
        const Vec3<S> b = massCenter(M, K);
        return norm(b - a);
    };
    switch (gr[0]) {
    case MOCOB200_GOAL_CONTROL: {
        S v = integral;
        if (ip[1]) v = v / displacement();
        vals[0] = w * v;
    } break;
    case MOCOB200_GOAL_CONTACT_TRACKING: vals[0] = w * (integral / S(dp[6])); break; // / (total mass * |g|)
    case MOCOB200_GOAL_STATE_TRACKING:
    case MOCOB200_GOAL_CONTROL_TRACKING:
    case MOCOB200_GOAL_SUM_SQUARED_STATE: vals[0] = w * integral; break;
    case MOCOB200_GOAL_FINAL_TIME: vals[0] = w * tf; break;
    case MOCOB200_GOAL_AVERAGE_SPEED: {
        S v = S(dp[3]) - displacement() / (tf - t0);
        if (cost) v = v * v;
        vals[0] = w * v;
    } break;
    case MOCOB200_GOAL_PERIODICITY: {
        const int nsp = ip[0], ncp = ip[1];
        for (int k = 0; k < nsp; ++k) {
            S v = x0[ip[2 + 2 * k]] * S(dp[3 + k]) - xf[ip[3 + 2 * k]];
            if (cost) v = v * v;
            vals[k] = w * v;
        }
        for (int k = 0; k < ncp; ++k) {
            S v = c0[ip[2 + 2 * nsp + 2 * k]] * S(dp[3 + nsp + k]) - cf[ip[3 + 2 * nsp + 2 * k]];
            if (cost) v = v * v;
            vals[nsp + k] = w * v;
        }
    } break;
    case MOCOB200_GOAL_INITIAL_ACTIVATION:
        for (int k = 0; k < ip[0]; ++k) {
            S v = c0[ip[1 + 2 * k]] - x0[ip[2 + 2 * k]];
            if (cost) v = v * v;
            vals[k] = w * v;
        }
        break;
    case MOCOB200_GOAL_INITIAL_VELOCITY_EQUILIBRIUM_DGF: {
        static thread_local Frames<S> K;
        static thread_local PathState<S> P;
        kinematics(M, x0, x0 + nq, K);
        const S* auxDeriv = d0 + (M.implicit ? nq : 0);
        for (int k = 0; k < ip[0]; ++k) {
            const MuscleResult<S> r = evalMuscle(M, ip[1 + k], x0, x0 + nq, x0 + 2 * nq, c0, auxDeriv, K, P, true);
            S v = r.linearizedResidualRate;
            if (cost) v = v * v;
            vals[k] = w * v;
        }
    } break;
    default: vals[0] = S(NAN); break;
    }
}

} // namespace orc
