// TEST INFRASTRUCTURE ONLY — CPU oracle for the mocob200 evaluator.
// Nothing in the product path (opensim-moco_b200/) includes, links or calls this directory.
//
// Small fixed-size vector algebra templated on the scalar, so that the same code runs with
// `double` (the checker / CPU baseline) and with `Counted` (operation counting for the
// algorithmic-FLOP figures of DESIGN.md, convention of SURVEY.md §8(d)).
#pragma once
#include <cmath>

namespace orc {

struct OpCounts {
    long long flop = 0;    // + - * / sqrt abs min max, 1 each; every libm call also counts 1 here
    long long transc = 0;  // libm calls (exp log tanh sinh sin cos pow asin), reported separately
};
inline OpCounts& counts() {
    static thread_local OpCounts c;
    return c;
}

struct Counted {
    double v;
    Counted() : v(0) {}
    Counted(double x) : v(x) {}
    Counted(int x) : v(x) {}
};
inline Counted operator+(Counted a, Counted b) { counts().flop++; return Counted(a.v + b.v); }
inline Counted operator-(Counted a, Counted b) { counts().flop++; return Counted(a.v - b.v); }
inline Counted operator*(Counted a, Counted b) { counts().flop++; return Counted(a.v * b.v); }
inline Counted operator/(Counted a, Counted b) { counts().flop++; return Counted(a.v / b.v); }
inline Counted operator-(Counted a) { return Counted(-a.v); }
inline Counted& operator+=(Counted& a, Counted b) { a = a + b; return a; }
inline Counted& operator-=(Counted& a, Counted b) { a = a - b; return a; }
inline Counted& operator*=(Counted& a, Counted b) { a = a * b; return a; }
inline bool operator<(Counted a, Counted b) { return a.v < b.v; }
inline bool operator>(Counted a, Counted b) { return a.v > b.v; }
inline bool operator<=(Counted a, Counted b) { return a.v <= b.v; }
inline bool operator>=(Counted a, Counted b) { return a.v >= b.v; }
#define ORC_UNARY(name, isT)                                                        \
    inline Counted name(Counted a) { counts().flop++; if (isT) counts().transc++; return Counted(std::name(a.v)); }
ORC_UNARY(sqrt, false) ORC_UNARY(fabs, false) ORC_UNARY(exp, true) ORC_UNARY(log, true) ORC_UNARY(tanh, true)
ORC_UNARY(sinh, true) ORC_UNARY(sin, true) ORC_UNARY(cos, true) ORC_UNARY(asin, true)
#undef ORC_UNARY
inline Counted pow(Counted a, Counted b) { counts().flop++; counts().transc++; return Counted(std::pow(a.v, b.v)); }
inline Counted fmin(Counted a, Counted b) { counts().flop++; return Counted(std::fmin(a.v, b.v)); }
inline Counted fmax(Counted a, Counted b) { counts().flop++; return Counted(std::fmax(a.v, b.v)); }
inline bool isfinite(Counted a) { return std::isfinite(a.v); }
inline double val(Counted a) { return a.v; }
inline double val(double a) { return a; }

using std::sqrt; using std::fabs; using std::exp; using std::log; using std::tanh; using std::sinh;
using std::sin; using std::cos; using std::asin; using std::pow; using std::fmin; using std::fmax; using std::isfinite;

// First-order dual number over S: value + derivative along one direction. The custom-joint kinematics run once with
// Dual<S> (direction = the joint's speeds) so that the time derivatives of the joint's motion axes - the velocity-product
// ("H-dot u") terms of a function-based mobilizer - come from the same few lines that define its transform.
template <class S>
struct Dual {
    S v, d;
    Dual() : v(0.0), d(0.0) {}
    Dual(double x) : v(x), d(0.0) {}
    Dual(S value, S deriv) : v(value), d(deriv) {}
};
template <class S> inline Dual<S> operator+(const Dual<S>& a, const Dual<S>& b) { return {a.v + b.v, a.d + b.d}; }
template <class S> inline Dual<S> operator-(const Dual<S>& a, const Dual<S>& b) { return {a.v - b.v, a.d - b.d}; }
template <class S> inline Dual<S> operator-(const Dual<S>& a) { return {-a.v, -a.d}; }
template <class S> inline Dual<S> operator*(const Dual<S>& a, const Dual<S>& b) { return {a.v * b.v, a.v * b.d + a.d * b.v}; }
template <class S> inline Dual<S> sin(const Dual<S>& a) { return {sin(a.v), cos(a.v) * a.d}; }
template <class S> inline Dual<S> cos(const Dual<S>& a) { return {cos(a.v), -(sin(a.v) * a.d)}; }
template <class S> inline double val(const Dual<S>& a) { return val(a.v); }

template <class S>
struct Vec3 {
    S x, y, z;
    Vec3() : x(0), y(0), z(0) {}
    Vec3(S a, S b, S c) : x(a), y(b), z(c) {}
};
template <class S> inline Vec3<S> operator+(const Vec3<S>& a, const Vec3<S>& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class S> inline Vec3<S> operator-(const Vec3<S>& a, const Vec3<S>& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class S> inline Vec3<S> operator*(S s, const Vec3<S>& a) { return {s * a.x, s * a.y, s * a.z}; }
template <class S> inline S dot(const Vec3<S>& a, const Vec3<S>& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <class S> inline Vec3<S> cross(const Vec3<S>& a, const Vec3<S>& b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
template <class S> inline S norm(const Vec3<S>& a) { return sqrt(dot(a, a)); }

// 3x3 matrix stored by rows: m[r][c]
template <class S>
struct Mat3 {
    S m[3][3];
    Mat3() { for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) m[r][c] = S(r == c ? 1.0 : 0.0); }
    static Mat3 zero() { Mat3 a; for (int r = 0; r < 3; ++r) a.m[r][r] = S(0.0); return a; }
    static Mat3 fromRowMajor(const double* p) { Mat3 a; for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) a.m[r][c] = S(p[3 * r + c]); return a; }
    Vec3<S> col(int c) const { return {m[0][c], m[1][c], m[2][c]}; }
};
template <class S> inline Vec3<S> operator*(const Mat3<S>& A, const Vec3<S>& v) {
    return {A.m[0][0] * v.x + A.m[0][1] * v.y + A.m[0][2] * v.z, A.m[1][0] * v.x + A.m[1][1] * v.y + A.m[1][2] * v.z,
            A.m[2][0] * v.x + A.m[2][1] * v.y + A.m[2][2] * v.z};
}
template <class S> inline Mat3<S> operator*(const Mat3<S>& A, const Mat3<S>& B) {
    Mat3<S> C;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) C.m[r][c] = A.m[r][0] * B.m[0][c] + A.m[r][1] * B.m[1][c] + A.m[r][2] * B.m[2][c];
    return C;
}
template <class S> inline Mat3<S> transpose(const Mat3<S>& A) {
    Mat3<S> C;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C.m[r][c] = A.m[c][r];
    return C;
}
template <class S> inline Mat3<S> operator+(const Mat3<S>& A, const Mat3<S>& B) {
    Mat3<S> C;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) C.m[r][c] = A.m[r][c] + B.m[r][c];
    return C;
}
// rotation by `angle` about the unit vector (ax, ay, az): cos 1 + (1 - cos) a a^T + sin [a]x
template <class S> inline Mat3<S> rotAxis(S angle, double ax, double ay, double az) {
    const S c = cos(angle), s = sin(angle), t = S(1.0) - c;
    const double a[3] = {ax, ay, az};
    Mat3<S> R;
    for (int r = 0; r < 3; ++r)
        for (int k = 0; k < 3; ++k) R.m[r][k] = t * S(a[r] * a[k]) + (r == k ? c : S(0.0));
    R.m[0][1] = R.m[0][1] - s * S(az); R.m[0][2] = R.m[0][2] + s * S(ay);
    R.m[1][0] = R.m[1][0] + s * S(az); R.m[1][2] = R.m[1][2] - s * S(ax);
    R.m[2][0] = R.m[2][0] - s * S(ay); R.m[2][1] = R.m[2][1] + s * S(ax);
    return R;
}
template <class S> inline Mat3<S> rotZ(S angle) {
    Mat3<S> R;
    const S c = cos(angle), s = sin(angle);
    R.m[0][0] = c; R.m[0][1] = -s; R.m[1][0] = s; R.m[1][1] = c;
    return R;
}
// m (|c|^2 1 - c c^T): the parallel-axis term
template <class S> inline Mat3<S> parallelAxis(S m, const Vec3<S>& c) {
    Mat3<S> P;
    const S cc = dot(c, c);
    const S v[3] = {c.x, c.y, c.z};
    for (int r = 0; r < 3; ++r)
        for (int k = 0; k < 3; ++k) P.m[r][k] = m * ((r == k ? cc : S(0.0)) - v[r] * v[k]);
    return P;
}

} // namespace orc
