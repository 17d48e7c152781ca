// TEST INFRASTRUCTURE -- part of the CPU oracle, never linked into or called by the product path.
//
// Tiny fixed-size linear algebra used by the oracle's restatement of the OptiMo glue code.
// It restates, without Eigen, exactly the few Eigen operations the reference path uses
// (SURVEY.md §8(c)):
//   * AngleAxisd(angle, Unit{X,Y,Z}).toRotationMatrix()    -> OptiMo/joint.cpp:69-72
//   * Affine3d translate / rotate / product / inverse       -> OptiMo/joint.cpp:64-83, ikhandler.cpp:124-127
//   * Quaterniond::FromTwoVectors(a, b).toRotationMatrix()  -> OptiMo/joint.cpp:112
//   * squaredNorm / norm / isApprox                         -> OptiMo/core_rbdl.cpp:331,357,392
#ifndef ORACLE_OMATH_HPP
#define ORACLE_OMATH_HPP

#include <cmath>
#include <vector>

namespace orc {

struct V3 {
    double v[3];
    V3() : v{0.0, 0.0, 0.0} {}
    V3(double a, double b, double c) : v{a, b, c} {}
    double& operator[](int i) { return v[i]; }
    const double& operator[](int i) const { return v[i]; }
};

inline V3 operator+(const V3& a, const V3& b) { return V3(a[0] + b[0], a[1] + b[1], a[2] + b[2]); }
inline V3 operator-(const V3& a, const V3& b) { return V3(a[0] - b[0], a[1] - b[1], a[2] - b[2]); }
inline V3 operator*(double s, const V3& a) { return V3(s * a[0], s * a[1], s * a[2]); }
inline double dot(const V3& a, const V3& b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline V3 cross(const V3& a, const V3& b)
{
    return V3(a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]);
}
inline double sqnorm(const V3& a) { return dot(a, a); }
inline double norm(const V3& a) { return std::sqrt(sqnorm(a)); }
/// Eigen's normalized(): divide by the norm when the squared norm is positive, else return as is.
inline V3 normalized(const V3& a)
{
    const double z = sqnorm(a);
    if (z > 0.0) { const double n = std::sqrt(z); return V3(a[0] / n, a[1] / n, a[2] / n); }
    return a;
}

struct M3 {
    double m[3][3];
    M3() { for (auto& r : m) for (double& e : r) e = 0.0; }
    static M3 identity() { M3 r; r.m[0][0] = r.m[1][1] = r.m[2][2] = 1.0; return r; }
    double& operator()(int i, int j) { return m[i][j]; }
    const double& operator()(int i, int j) const { return m[i][j]; }
};

inline M3 operator*(const M3& a, const M3& b)
{
    M3 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            r(i, j) = a(i, 0) * b(0, j) + a(i, 1) * b(1, j) + a(i, 2) * b(2, j);
    return r;
}
inline V3 operator*(const M3& a, const V3& x)
{
    return V3(a(0, 0) * x[0] + a(0, 1) * x[1] + a(0, 2) * x[2],
              a(1, 0) * x[0] + a(1, 1) * x[1] + a(1, 2) * x[2],
              a(2, 0) * x[0] + a(2, 1) * x[1] + a(2, 2) * x[2]);
}
inline M3 transpose(const M3& a)
{
    M3 r;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r(i, j) = a(j, i);
    return r;
}

/// AngleAxisd(angle, e_axis).toRotationMatrix() for a coordinate axis (0=x,1=y,2=z), following
/// Eigen's formula: off-diagonals tmp -/+ sin*axis, diagonal (1-c)*axis_i^2 + c.
inline M3 angle_axis_unit(double angle, int axis)
{
    V3 ax; ax[axis] = 1.0;
    const double s = std::sin(angle), c = std::cos(angle);
    const V3 sin_axis = s * ax;
    const V3 cos1_axis = (1.0 - c) * ax;
    M3 r;
    double tmp;
    tmp = cos1_axis[0] * ax[1]; r(0, 1) = tmp - sin_axis[2]; r(1, 0) = tmp + sin_axis[2];
    tmp = cos1_axis[0] * ax[2]; r(0, 2) = tmp + sin_axis[1]; r(2, 0) = tmp - sin_axis[1];
    tmp = cos1_axis[1] * ax[2]; r(1, 2) = tmp - sin_axis[0]; r(2, 1) = tmp + sin_axis[0];
    for (int i = 0; i < 3; ++i) r(i, i) = cos1_axis[i] * ax[i] + c;
    return r;
}

/// Rigid/affine transform y = L x + t (Eigen::Affine3d).
struct Affine {
    M3 L; V3 t;
    Affine() : L(M3::identity()), t() {}
    void translate(const V3& d) { t = t + L * d; }   // Transform::translate: t += linear * d
    void rotate(const M3& r) { L = L * r; }           // Transform::rotate: linear *= R
};
inline Affine operator*(const Affine& a, const Affine& b)
{
    Affine r; r.L = a.L * b.L; r.t = a.L * b.t + a.t; return r;
}
inline V3 operator*(const Affine& a, const V3& x) { return a.L * x + a.t; }

/// Affine3d::inverse() in Affine mode: general 3x3 inverse of the linear part by cofactors
/// (Eigen's size-3 compute_inverse), translation = -inv * t.
inline Affine inverse(const Affine& a)
{
    const M3& m = a.L;
    auto cof = [&](int i, int j) {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        return m(i1, j1) * m(i2, j2) - m(i1, j2) * m(i2, j1);
    };
    const V3 c0(cof(0, 0), cof(1, 0), cof(2, 0));
    const double det = c0[0] * m(0, 0) + c0[1] * m(1, 0) + c0[2] * m(2, 0);
    const double invdet = 1.0 / det;
    Affine r;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r.L(i, j) = cof(j, i) * invdet;
    const V3 lt = r.L * a.t;
    r.t = V3(-lt[0], -lt[1], -lt[2]);
    return r;
}

/// Quaterniond::FromTwoVectors(a, b).toRotationMatrix()  (joint.cpp:112). The c < -1+eps branch
/// (antiparallel vectors, SVD in Eigen) is not restated; callers must avoid it.
inline M3 from_two_vectors(const V3& a, const V3& b, bool* degenerate = nullptr)
{
    const V3 v0 = normalized(a), v1 = normalized(b);
    const double c = dot(v1, v0);
    if (degenerate) *degenerate = (c < -1.0 + 1e-12);
    const V3 axis = cross(v0, v1);
    const double s = std::sqrt((1.0 + c) * 2.0);
    const double invs = 1.0 / s;
    const double x = axis[0] * invs, y = axis[1] * invs, z = axis[2] * invs, w = s * 0.5;
    const double tx = 2.0 * x, ty = 2.0 * y, tz = 2.0 * z;
    const double twx = tx * w, twy = ty * w, twz = tz * w;
    const double txx = tx * x, txy = ty * x, txz = tz * x;
    const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
    M3 r;
    r(0, 0) = 1.0 - (tyy + tzz); r(0, 1) = txy - twz;         r(0, 2) = txz + twy;
    r(1, 0) = txy + twz;         r(1, 1) = 1.0 - (txx + tzz); r(1, 2) = tyz - twx;
    r(2, 0) = txz - twy;         r(2, 1) = tyz + twx;         r(2, 2) = 1.0 - (txx + tyy);
    return r;
}

using VecN = std::vector<double>;

inline double sqnorm(const VecN& a) { double s = 0.0; for (double e : a) s += e * e; return s; }
inline double sqdist(const VecN& a, const VecN& b)
{
    double s = 0.0;
    for (size_t i = 0; i < a.size(); ++i) { const double d = a[i] - b[i]; s += d * d; }
    return s;
}
/// Eigen's DenseBase::isApprox with the default double precision 1e-12.
inline bool is_approx(const VecN& a, const VecN& b)
{
    if (a.size() != b.size()) return false;
    const double prec = 1e-12;
    return sqdist(a, b) <= prec * prec * std::fmin(sqnorm(a), sqnorm(b));
}

}  // namespace orc

#endif
