// TEST INFRASTRUCTURE ONLY — CPU oracle: a path segment wrapping over a cylinder.
// Restatement of the reference's procedure (opensim-core/OpenSim/Simulation/Wrap/WrapCylinder.cpp:176-605 wrapLine,
// :618-706 _make_spiral_path, :769-806 _adjust_tangent_point; helpers Wrap/WrapMath.cpp:73-155,195-209). The procedure is a
// heuristic, not the exact geodesic: tangent points from the planar tangent construction, slid along the axis onto the plane
// through the two path points and an "apex" point, then nudged (at most 100 times, by 1.5 x the miss) until the straight
// segments meet the helix to within 0.1 degree, the helix direction being sampled on a 2 mm discretisation. The numbers the
// reference prints (MuscleAnalysis Length, Actuation speed) are reproduced only by following the same steps.
// Everything here is in the cylinder's own frame: axis = z, centred on the origin.
#pragma once
#include <array>
#include <cmath>

namespace orc {
namespace wrap {

using V = std::array<double, 3>;
inline V operator+(const V& a, const V& b) { return {a[0] + b[0], a[1] + b[1], a[2] + b[2]}; }
inline V operator-(const V& a, const V& b) { return {a[0] - b[0], a[1] - b[1], a[2] - b[2]}; }
inline V operator*(double s, const V& a) { return {s * a[0], s * a[1], s * a[2]}; }
inline double dot(const V& a, const V& b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline V cross(const V& a, const V& b) { return {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]}; }
inline double norm(const V& a) { return std::sqrt(dot(a, a)); }
inline double dist2(const V& a, const V& b) { return dot(a - b, a - b); }
// Mtx::Normalize (Common/Mtx.h:96-103): zero vector below machine epsilon; returns the magnitude
inline double normalize(V& a) {
    const double m = norm(a);
    if (m >= 2.220446049250313e-16) a = (1.0 / m) * a;
    else a = {0, 0, 0};
    return m;
}
inline V unit(V a) { normalize(a); return a; }
inline double angleBetween(const V& a, const V& b) { return std::acos(dot((1.0 / norm(a)) * a, (1.0 / norm(b)) * b)); } // Mtx::Angle
inline int sgn(double a) { return a >= 0.0 ? 1 : -1; }                                                                 // DSIGN
inline V onAxis(const V& p) { // closest point of the z axis (WrapMath::GetClosestPointOnLineToPoint with p0 = (0,0,-1), dn = z)
    V v{p[0], p[1], p[2] + 1.0};
    const double mag = normalize(v);
    const double t = v[2] * mag;
    return {0.0, 0.0, -1.0 + t};
}

enum Action { NoWrap = 0, InsideRadius = 1, Wrapped = 2, MandatoryWrap = 3 }; // WrapObject.h:80-86

struct CylinderWrap {
    V r1, r2;          // tangent points (cylinder frame)
    double arc = 0.0;  // length of the helical part
};

// closest approach of lines p1p2 and p3p4 (WrapMath::IntersectLines): s, t as fractions of the two segments
inline bool closestApproach(const V& p1, const V& p2, const V& p3, const V& p4, V& on12, double& s, V& on34, double& t) {
    V d1 = p2 - p1, d2 = p4 - p3;
    const double m1 = normalize(d1), m2 = normalize(d2);
    const V c = cross(d1, d2);
    const double den = dot(c, c);
    if (std::fabs(den) <= 0.0000000000002) { s = t = NAN; return false; }
    const V w = p3 - p1;
    t = dot(w, cross(d1, c)) / den;
    on34 = p3 + t * d2;
    s = dot(w, cross(d2, c)) / den;
    on12 = p1 + s * d1;
    s /= m1;
    t /= m2;
    return true;
}
// WrapMath::IntersectLineSegPlane: plane n.x + d = 0
inline bool segmentPlane(const V& a, const V& b, const V& n, double d, V& hit) {
    const V ab = b - a;
    const double dp = dot(ab, n);
    if (std::fabs(dp) < 0.00001) return false;
    const double t = (-d - dot(n, a)) / dp;
    if (t < -0.00001 || t > 1.0 + 0.00001) return false;
    hit = a + t * ab;
    return true;
}

// point of the helix at fraction t (WrapCylinder.cpp:722-745): the radial unit vector uu turned about `ax` by sense*t*theta,
// plus the axial advance
inline V helixPoint(const V& r1a, const V& axial, const V& ax, const V& uu, double radius, double sense, double t, double theta) {
    const double phi = sense * t * theta, c = std::cos(phi), s = std::sin(phi);
    const V rot = c * uu + s * cross(ax, uu) + ((1.0 - c) * dot(ax, uu)) * ax; // Rodrigues
    return r1a + radius * rot + t * axial;
}
// WrapCylinder.cpp:769-806: is the segment pt -> r tangent to the helix (direction r -> w) to within 0.1 degree, measured
// against the axis? If not, slide r along the axis by 1.5 x the miss.
inline bool nudgeTangent(const V& pt, V& r, const V& w) {
    const V z{0, 0, 1};
    const V pr = unit(r - pt), rw = unit(w - r);
    const double alpha = std::acos(dot(pr, z)), omega = std::acos(dot(rw, z));
    if (!(std::fabs(alpha - omega) > 0.1 * 0.017453292519943295)) return false;
    // feet of pt and w on the line through r parallel to the axis, then where the lines pt-w and foot-foot pass closest
    auto foot = [&](const V& p) {
        V v = p - r;
        const double mag = normalize(v);
        return r + (v[2] * mag) * z;
    };
    const V pa = foot(pt), wa = foot(w);
    V i1, i2;
    double s, t;
    closestApproach(pt, w, pa, wa, i1, s, i2, t);
    r = r + 1.5 * (i2 - r);
    return true;
}

inline void spiral(double radius, const V& P1, const V& P2, bool farSide, CylinderWrap& W) {
    const double sense = farSide ? -1.0 : 1.0;
    for (int iterations = 0;; ) {
        const V r1a = onAxis(W.r1), r2a = onAxis(W.r2);
        const V axial = r2a - r1a;
        V uu = (1.0 / radius) * (W.r1 - r1a), vv = (1.0 / radius) * (W.r2 - r2a);
        double theta = angleBetween(uu, vv);
        if (farSide) theta = 2.0 * M_PI - theta;
        const double x = radius * theta, y = norm(axial);
        W.arc = std::sqrt(x * x + y * y);
        const V ax = unit(cross(uu, vv));
        int nseg = (int)(W.arc / 0.002);
        if (nseg < 1) nseg = 1;
        // the reference builds nseg surface points; only the second one (i == 1) drives the tangency check
        if (nseg > 1 && iterations < 100) {
            const double t = 1.0 / nseg;
            const V w1 = helixPoint(r1a, axial, ax, uu, radius, sense, t, theta);
            const bool a1 = nudgeTangent(P1, W.r1, w1);
            const V w2 = helixPoint(r1a, axial, ax, uu, radius, sense, 1.0 - t, theta);
            const bool a2 = nudgeTangent(P2, W.r2, w2);
            if (a1 || a2) { ++iterations; continue; }
        }
        return;
    }
}

// One segment P1 -> P2 over a cylinder of the given radius and (display) length. wrapAxis / wrapSign: the constrained half
// ("quadrant" x, -x, y, -y; sign 0 = all). Returns the WrapAction; W is valid for Wrapped / MandatoryWrap.
inline int cylinderWrapLine(double radius, double length, int wrapAxis, int wrapSign, const V& P1, const V& P2, CylinderWrap& W) {
    const double r2 = radius * radius;
    int code = Wrapped;
    const bool constrained = wrapSign != 0;
    bool farSide = false, longWrap = false;
    if (dist2(P1, onAxis(P1)) < r2 || dist2(P2, onAxis(P2)) < r2) return InsideRadius;
    V near12{0, 0, 0}, near00{0, 0, 0};
    double t12 = NAN, t00 = NAN;
    closestApproach(P1, P2, V{0, 0, -0.5 * length}, V{0, 0, 0.5 * length}, near12, t12, near00, t00);
    const bool pierces = dist2(near12, near00) < r2 && t12 > 0.0 && t12 < 1.0;
    if (!constrained) {
        if (pierces) code = MandatoryWrap;
        else return NoWrap;
    }
    const V p11 = onAxis(P1), p22 = onAxis(P2);
    const V z{0, 0, 1};
    // the two tangent points of the circle seen from P (planar construction at P's height)
    auto tangents = [&](const V& P, const V& foot, V& ta, V& tb, double& reach) {
        V vv = P - foot;
        reach = normalize(vv);
        const double along = radius * (radius / reach);
        const V pp = foot + along * vv;
        const double across = std::sqrt(r2 - along * along);
        const V uu = cross(z, vv);
        ta = pp + across * uu;
        tb = pp - across * uu;
    };
    V r1a, r1b, r2a, r2b;
    double reach1, reach2;
    tangents(P1, p11, r1a, r1b, reach1);
    tangents(P2, p22, r2a, r2b, reach2);
    if (constrained) {
        const int s1 = sgn(P1[wrapAxis]), s2 = sgn(P2[wrapAxis]);
        if (s1 == wrapSign || s2 == wrapSign) {
            if (pierces) code = MandatoryWrap;
            else if (s1 != s2 && sgn(near12[wrapAxis]) != wrapSign) code = Wrapped;
            else return NoWrap;
        }
        const double alpha = angleBetween(r1a - p11, r2b - p22), beta = angleBetween(r1b - p11, r2a - p22);
        const bool a1 = sgn(r1a[wrapAxis]) == wrapSign, b1 = sgn(r1b[wrapAxis]) == wrapSign;
        const bool a2 = sgn(r2a[wrapAxis]) == wrapSign, b2 = sgn(r2b[wrapAxis]) == wrapSign;
        // pairing (r1a, r2b) or (r1b, r2a), decided by which candidates lie on the active half
        bool useA; // true: r1 = r1a, r2 = r2b
        if (a1 == b1) { // both or neither of P1's candidates on the active half: let P2's decide
            if (a2) useA = false;
            else if (a1 || b2) useA = true;
            else { useA = alpha > beta; farSide = true; W.r1 = useA ? r1a : r1b; W.r2 = useA ? r2b : r2a; goto chosen; }
        } else useA = a1;
        W.r1 = useA ? r1a : r1b;
        W.r2 = useA ? r2b : r2a;
        farSide = useA ? (alpha > beta) : !(alpha > beta);
    chosen:
        const V sumMuscle = (W.r1 - P1) + (W.r2 - P2), sumRadial = (W.r1 - p11) + (W.r2 - p22);
        if (dot(sumRadial, sumMuscle) < 0.0) longWrap = true;
    } else {
        const V u1a = unit(r1a - p11), u1b = unit(r1b - p11), u2a = unit(r2a - p22), u2b = unit(r2b - p22);
        const double d1 = dot(u1a, u2a), d2 = dot(u1a, u2b), d3 = dot(u1b, u2a), d4 = dot(u1b, u2b);
        if (d1 > d2 && d1 > d3 && d1 > d4) { W.r1 = r1a; W.r2 = r2a; }
        else if (d2 > d3 && d2 > d4) { W.r1 = r1a; W.r2 = r2b; }
        else if (d3 > d4) { W.r1 = r1b; W.r2 = r2a; }
        else { W.r1 = r1b; W.r2 = r2b; }
    }
    // the "apex": where the bisector plane of the two path points meets the cylinder, on the near (or far) side
    const double frac = reach1 / (reach1 + reach2);
    const V mpt = P1 + frac * (P2 - P1);
    const V axispt = onAxis(mpt);
    V pn = unit(cross(unit(P1 - axispt), unit(P2 - axispt)));
    const V apexA = axispt + radius * unit(cross(z, pn)), apexB = axispt + radius * unit(cross(pn, z));
    const bool aNearer = dist2(mpt, apexA) < dist2(mpt, apexB);
    const V apex = (farSide != aNearer) ? apexA : apexB; // near side picks the nearer, far side the farther
    // slide the tangent points along the axis onto the plane through P1, P2 and the apex
    pn = unit(cross(unit(P1 - apex), unit(P2 - apex)));
    const double d = -dot(P1, pn);
    V hit;
    if (segmentPlane(W.r1 - 10.0 * z, W.r1 + 10.0 * z, pn, d, hit)) {
        const V f{0.0, 0.0, hit[2]}; // foot on the axis (the reference passes p22 as the line direction: the axis)
        if (dist2(f, p22) < dist2(p11, p22)) W.r1 = hit;
    }
    if (segmentPlane(W.r2 - 10.0 * z, W.r2 + 10.0 * z, pn, d, hit)) {
        const V f{0.0, 0.0, hit[2]};
        if (dist2(f, p11) < dist2(p22, p11)) W.r2 = hit;
    }
    const double half = 0.5 * length;
    if ((W.r1[2] < -half || W.r1[2] > half) && (W.r2[2] < -half || W.r2[2] > half)) return NoWrap;
    spiral(radius, P1, P2, longWrap, W);
    return code;
}

} // namespace wrap
} // namespace orc
