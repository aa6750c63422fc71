// One path segment over a wrap cylinder, on the device. What the reference computes is a heuristic, not the exact
// geodesic (Simulation/Wrap/WrapCylinder.cpp:176-605 wrapLine, :618-706 _make_spiral_path, :769-806 _adjust_tangent_point;
// Wrap/WrapMath.cpp:73-155,195-209): tangent points from the planar tangent construction, slid along the axis onto the plane
// through the two path points and an apex point, then nudged (at most 100 times, by 1.5 x the miss) until the straight
// segments meet the helix to within 0.1 degree, the helix direction being sampled on a 2 mm discretisation. The reference's
// printed path lengths are reproduced only by following the same steps (tests/test_wrapping.py pins them to 1e-8 m).
// Everything here is in the cylinder's own frame: axis = z, centred on the origin. One thread (or one lane) per segment.
// (included by point_eval.cuh after its 3-vector helpers)
#pragma once

namespace mocob200 {

enum { WRAP_NONE = 0, WRAP_INSIDE_RADIUS = 1, WRAP_WRAPPED = 2, WRAP_MANDATORY = 3 }; // WrapObject.h:80-86

struct CylWrap {
    V3 r1, r2;  // tangent points
    double arc; // length of the helical part
};

__device__ __forceinline__ double wr_norm(V3 a) { return sqrt(dot(a, a)); }
__device__ __forceinline__ V3 wr_unit(V3 a, double* mag = nullptr) { // Mtx::Normalize: zero below machine epsilon
    const double m = wr_norm(a);
    if (mag) *mag = m;
    return m >= 2.220446049250313e-16 ? (1.0 / m) * a : mk(0, 0, 0);
}
__device__ __forceinline__ double wr_dist2(V3 a, V3 b) { const V3 d = a - b; return dot(d, d); }
__device__ __forceinline__ double wr_angle(V3 a, V3 b) { return acos(dot((1.0 / wr_norm(a)) * a, (1.0 / wr_norm(b)) * b)); }
__device__ __forceinline__ int wr_sgn(double a) { return a >= 0.0 ? 1 : -1; }
__device__ __forceinline__ double wr_comp(V3 a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }
// closest point of the z axis, as WrapMath::GetClosestPointOnLineToPoint finds it from (0, 0, -1) along z
__device__ __forceinline__ V3 wr_on_axis(V3 p) {
    double mag;
    const V3 v = wr_unit(mk(p.x, p.y, p.z + 1.0), &mag);
    return mk(0.0, 0.0, -1.0 + v.z * mag);
}
// closest approach of lines p1p2 and p3p4 (WrapMath::IntersectLines): s, t as fractions of the two segments
__device__ inline bool wr_closest_approach(V3 p1, V3 p2, V3 p3, V3 p4, V3& on12, double& s, V3& on34, double& t) {
    double m1, m2;
    const V3 d1 = wr_unit(p2 - p1, &m1), d2 = wr_unit(p4 - p3, &m2);
    const V3 c = cross(d1, d2);
    const double den = dot(c, c);
    if (fabs(den) <= 0.0000000000002) { s = t = NAN; return false; }
    const V3 w = p3 - p1;
    t = dot(w, cross(d1, c)) / den;
    on34 = p3 + t * d2;
    s = dot(w, cross(d2, c)) / den;
    on12 = p1 + s * d1;
    s /= m1;
    t /= m2;
    return true;
}
__device__ inline bool wr_segment_plane(V3 a, V3 b, V3 n, double d, V3& hit) { // WrapMath::IntersectLineSegPlane
    const V3 ab = b - a;
    const double dp = dot(ab, n);
    if (fabs(dp) < 0.00001) return false;
    const double t = (-d - dot(n, a)) / dp;
    if (t < -0.00001 || t > 1.0 + 0.00001) return false;
    hit = a + t * ab;
    return true;
}
// point of the helix at fraction t: the radial unit vector uu turned about ax by sense t theta, plus the axial advance
__device__ inline V3 wr_helix_point(V3 r1a, V3 axial, V3 ax, V3 uu, double radius, double sense, double t, double theta) {
    double s, c;
    sincos(sense * t * theta, &s, &c);
    const V3 rot = c * uu + s * cross(ax, uu) + ((1.0 - c) * dot(ax, uu)) * ax;
    return r1a + radius * rot + t * axial;
}
// is pt -> r tangent to the helix (direction r -> w) to within 0.1 degree against the axis? if not, slide r by 1.5 x the miss
__device__ inline bool wr_nudge_tangent(V3 pt, V3& r, V3 w) {
    const V3 z = mk(0, 0, 1);
    const V3 pr = wr_unit(r - pt), rw = wr_unit(w - r);
    const double alpha = acos(pr.z), omega = acos(rw.z);
    if (!(fabs(alpha - omega) > 0.1 * 0.017453292519943295)) return false;
    double m1, m2;
    const V3 v1 = wr_unit(pt - r, &m1), v2 = wr_unit(w - r, &m2);
    const V3 pa = r + (v1.z * m1) * z, wa = r + (v2.z * m2) * z; // feet on the line through r parallel to the axis
    V3 i1, i2 = r;
    double s, t;
    wr_closest_approach(pt, w, pa, wa, i1, s, i2, t);
    r = r + 1.5 * (i2 - r);
    return true;
}
__device__ inline void wr_spiral(double radius, V3 P1, V3 P2, bool farSide, CylWrap& W) {
    const double sense = farSide ? -1.0 : 1.0;
    for (int iterations = 0;;) {
        const V3 r1a = wr_on_axis(W.r1), r2a = wr_on_axis(W.r2);
        const V3 axial = r2a - r1a;
        const V3 uu = (1.0 / radius) * (W.r1 - r1a), vv = (1.0 / radius) * (W.r2 - r2a);
        double theta = wr_angle(uu, vv);
        if (farSide) theta = 2.0 * 3.14159265358979323846 - theta;
        const double x = radius * theta, y = wr_norm(axial);
        W.arc = sqrt(x * x + y * y);
        const V3 ax = wr_unit(cross(uu, vv));
        int nseg = (int)(W.arc / 0.002);
        if (nseg < 1) nseg = 1;
        if (nseg > 1 && iterations < 100) { // only the second surface point of the reference's discretisation drives the check
            const double t = 1.0 / nseg;
            const V3 w1 = wr_helix_point(r1a, axial, ax, uu, radius, sense, t, theta);
            const bool a1 = wr_nudge_tangent(P1, W.r1, w1);
            const V3 w2 = wr_helix_point(r1a, axial, ax, uu, radius, sense, 1.0 - t, theta);
            const bool a2 = wr_nudge_tangent(P2, W.r2, w2);
            if (a1 || a2) { ++iterations; continue; }
        }
        return;
    }
}

// Segment P1 -> P2 over a cylinder of the given radius and (display) length; wrapAxis / wrapSign: the constrained half (sign 0 =
// all quadrants). Returns the WrapAction; W is valid for WRAP_WRAPPED / WRAP_MANDATORY.
__device__ inline int wr_cylinder(double radius, double length, int wrapAxis, int wrapSign, V3 P1, V3 P2, CylWrap& W) {
    const double r2 = radius * radius;
    int code = WRAP_WRAPPED;
    const bool constrained = wrapSign != 0;
    bool farSide = false, longWrap = false;
    if (wr_dist2(P1, wr_on_axis(P1)) < r2 || wr_dist2(P2, wr_on_axis(P2)) < r2) return WRAP_INSIDE_RADIUS;
    V3 near12 = mk(0, 0, 0), near00 = mk(0, 0, 0);
    double t12 = NAN, t00 = NAN;
    wr_closest_approach(P1, P2, mk(0, 0, -0.5 * length), mk(0, 0, 0.5 * length), near12, t12, near00, t00);
    const bool pierces = wr_dist2(near12, near00) < r2 && t12 > 0.0 && t12 < 1.0;
    if (!constrained) {
        if (pierces) code = WRAP_MANDATORY;
        else return WRAP_NONE;
    }
    const V3 p11 = wr_on_axis(P1), p22 = wr_on_axis(P2);
    const V3 z = mk(0, 0, 1);
    V3 r1a, r1b, r2a, r2b;
    double reach1, reach2;
    { // the two tangent points of the circle seen from each path point (planar construction at the point's height)
        V3 vv = wr_unit(P1 - p11, &reach1);
        double along = radius * (radius / reach1);
        V3 pp = p11 + along * vv;
        double across = sqrt(r2 - along * along);
        V3 uu = cross(z, vv);
        r1a = pp + across * uu;
        r1b = pp - across * uu;
        vv = wr_unit(P2 - p22, &reach2);
        along = radius * (radius / reach2);
        pp = p22 + along * vv;
        across = sqrt(r2 - along * along);
        uu = cross(z, vv);
        r2a = pp + across * uu;
        r2b = pp - across * uu;
    }
    if (constrained) {
        const int s1 = wr_sgn(wr_comp(P1, wrapAxis)), s2 = wr_sgn(wr_comp(P2, wrapAxis));
        if (s1 == wrapSign || s2 == wrapSign) {
            if (pierces) code = WRAP_MANDATORY;
            else if (s1 != s2 && wr_sgn(wr_comp(near12, wrapAxis)) != wrapSign) code = WRAP_WRAPPED;
            else return WRAP_NONE;
        }
        const double alpha = wr_angle(r1a - p11, r2b - p22), beta = wr_angle(r1b - p11, r2a - p22);
        const bool a1 = wr_sgn(wr_comp(r1a, wrapAxis)) == wrapSign, b1 = wr_sgn(wr_comp(r1b, wrapAxis)) == wrapSign;
        const bool a2 = wr_sgn(wr_comp(r2a, wrapAxis)) == wrapSign, b2 = wr_sgn(wr_comp(r2b, wrapAxis)) == wrapSign;
        bool useA, forcedFar = false; // useA: r1 = r1a, r2 = r2b; otherwise r1 = r1b, r2 = r2a
        if (a1 == b1) {
            if (a2) useA = false;
            else if (a1 || b2) useA = true;
            else { useA = alpha > beta; forcedFar = true; } // none of the four candidates on the active half
        } else useA = a1;
        W.r1 = useA ? r1a : r1b;
        W.r2 = useA ? r2b : r2a;
        farSide = forcedFar ? true : (useA ? (alpha > beta) : !(alpha > beta));
        const V3 sumMuscle = (W.r1 - P1) + (W.r2 - P2), sumRadial = (W.r1 - p11) + (W.r2 - p22);
        if (dot(sumRadial, sumMuscle) < 0.0) longWrap = true;
    } else {
        const V3 u1a = wr_unit(r1a - p11), u1b = wr_unit(r1b - p11), u2a = wr_unit(r2a - p22), u2b = wr_unit(r2b - p22);
        const double d1 = dot(u1a, u2a), d2 = dot(u1a, u2b), d3 = dot(u1b, u2a), d4 = dot(u1b, u2b);
        if (d1 > d2 && d1 > d3 && d1 > d4) { W.r1 = r1a; W.r2 = r2a; }
        else if (d2 > d3 && d2 > d4) { W.r1 = r1a; W.r2 = r2b; }
        else if (d3 > d4) { W.r1 = r1b; W.r2 = r2a; }
        else { W.r1 = r1b; W.r2 = r2b; }
    }
    // the apex: where the plane through the axis and the (weighted) middle of the segment meets the cylinder, near or far side
    const double frac = reach1 / (reach1 + reach2);
    const V3 mpt = P1 + frac * (P2 - P1);
    const V3 axispt = wr_on_axis(mpt);
    V3 pn = wr_unit(cross(wr_unit(P1 - axispt), wr_unit(P2 - axispt)));
    const V3 apexA = axispt + radius * wr_unit(cross(z, pn)), apexB = axispt + radius * wr_unit(cross(pn, z));
    const bool aNearer = wr_dist2(mpt, apexA) < wr_dist2(mpt, apexB);
    const V3 apex = (farSide != aNearer) ? apexA : apexB;
    // slide the tangent points along the axis onto the plane through P1, P2 and the apex
    pn = wr_unit(cross(wr_unit(P1 - apex), wr_unit(P2 - apex)));
    const double d = -dot(P1, pn);
    V3 hit;
    if (wr_segment_plane(W.r1 - 10.0 * z, W.r1 + 10.0 * z, pn, d, hit)) {
        const V3 f = mk(0.0, 0.0, hit.z);
        if (wr_dist2(f, p22) < wr_dist2(p11, p22)) W.r1 = hit;
    }
    if (wr_segment_plane(W.r2 - 10.0 * z, W.r2 + 10.0 * z, pn, d, hit)) {
        const V3 f = mk(0.0, 0.0, hit.z);
        if (wr_dist2(f, p11) < wr_dist2(p22, p11)) W.r2 = hit;
    }
    const double half = 0.5 * length;
    if ((W.r1.z < -half || W.r1.z > half) && (W.r2.z < -half || W.r2.z > half)) return WRAP_NONE;
    wr_spiral(radius, P1, P2, longWrap, W);
    return code;
}

} // namespace mocob200
