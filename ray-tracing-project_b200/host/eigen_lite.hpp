// eigen_lite.hpp - the few fixed-size Eigen types the Flyscene / BoundingBox / Tucano surface
// of the reference shows in its signatures (src/flyscene.hpp:22-186, src/BoundingBox.h:10-67),
// as plain structs with the evaluation order Eigen 3.3.7 gives them:
//   dot / squaredNorm : a0 + (a1 + a2)                 Eigen/src/Core/Redux.h:91-105
//   normalized()      : v / sqrt(squaredNorm), identity if squaredNorm <= 0   Core/Dot.h:124-134
//   cross             : (a1b2-a2b1, a2b0-a0b2, a0b1-a1b0)    Geometry/OrthoMethods.h:43-47
//   Affine3f * v      : ((c0*x + c1*y) + c2*z) + c3          Geometry/Transform.h:1372-1392
//   Affine3f::inverse : cofactor 3x3 inverse, t' = -(inv * t)  Geometry/Transform.h:1215-1226
// Host translation units are compiled with -ffp-contract=off, so every expression rounds like
// the reference's baseline x86-64 build.  They live in namespace Eigen so that code written
// against the reference's headers (lights.push_back({Eigen::Vector3f(..), ..})) compiles
// unchanged; this header must not be mixed with the real Eigen in one translation unit.
#pragma once
#include <cmath>
#include <ostream>

#include "../csrc/rtp_math.h"

namespace Eigen {

struct Vector2f {
  float v[2] = {0.f, 0.f};
  Vector2f() {}
  Vector2f(float a, float b) { v[0] = a; v[1] = b; }
  static Vector2f Zero() { return Vector2f(); }
  float &operator[](int i) { return v[i]; }
  float operator[](int i) const { return v[i]; }
  float &operator()(int i) { return v[i]; }
  float operator()(int i) const { return v[i]; }
  float x() const { return v[0]; }
  float y() const { return v[1]; }
  Vector2f operator-(const Vector2f &o) const { return Vector2f(v[0] - o.v[0], v[1] - o.v[1]); }
};

struct Vector2i {
  int v[2] = {0, 0};
  Vector2i() {}
  Vector2i(int a, int b) { v[0] = a; v[1] = b; }
  int &operator[](int i) { return v[i]; }
  int operator[](int i) const { return v[i]; }
  int &operator()(int i) { return v[i]; }
  int operator()(int i) const { return v[i]; }
};

struct Vector3f {
  float v[3] = {0.f, 0.f, 0.f};
  Vector3f() {}
  template <class A, class B, class C>
  Vector3f(A a, B b, C c) { v[0] = (float)a; v[1] = (float)b; v[2] = (float)c; }  // doubles round to float, as Eigen
  explicit Vector3f(rtp::v3 a) { v[0] = a.x; v[1] = a.y; v[2] = a.z; }
  static Vector3f Zero() { return Vector3f(); }
  static Vector3f UnitX() { return Vector3f(1.f, 0.f, 0.f); }
  static Vector3f UnitY() { return Vector3f(0.f, 1.f, 0.f); }
  static Vector3f UnitZ() { return Vector3f(0.f, 0.f, 1.f); }
  rtp::v3 r() const { return rtp::mk3(v[0], v[1], v[2]); }
  float &operator[](int i) { return v[i]; }
  float operator[](int i) const { return v[i]; }
  float &operator()(int i) { return v[i]; }
  float operator()(int i) const { return v[i]; }
  float &x() { return v[0]; }
  float &y() { return v[1]; }
  float &z() { return v[2]; }
  float x() const { return v[0]; }
  float y() const { return v[1]; }
  float z() const { return v[2]; }
  Vector3f operator+(const Vector3f &o) const { return Vector3f(rtp::vadd(r(), o.r())); }
  Vector3f operator-(const Vector3f &o) const { return Vector3f(rtp::vsub(r(), o.r())); }
  Vector3f operator-() const { return Vector3f(-v[0], -v[1], -v[2]); }
  Vector3f operator*(float s) const { return Vector3f(rtp::vmuls(r(), s)); }
  Vector3f operator/(float s) const { return Vector3f(rtp::vdivs(r(), s)); }
  Vector3f &operator+=(const Vector3f &o) { return *this = *this + o; }
  Vector3f &operator-=(const Vector3f &o) { return *this = *this - o; }
  Vector3f &operator/=(float s) { return *this = *this / s; }
  Vector3f &operator*=(float s) { return *this = *this * s; }
  float dot(const Vector3f &o) const { return rtp::vdot(r(), o.r()); }
  Vector3f cross(const Vector3f &o) const { return Vector3f(rtp::vcross(r(), o.r())); }
  float squaredNorm() const { return rtp::vdot(r(), r()); }
  float norm() const { return rtp::vnorm(r()); }
  Vector3f normalized() const { return Vector3f(rtp::vnormalized(r())); }
  void normalize() { *this = normalized(); }
  Vector3f cwiseProduct(const Vector3f &o) const { return Vector3f(v[0] * o.v[0], v[1] * o.v[1], v[2] * o.v[2]); }
  bool operator==(const Vector3f &o) const { return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2]; }
};
inline Vector3f operator*(float s, const Vector3f &a) { return a * s; }
inline std::ostream &operator<<(std::ostream &os, const Vector3f &a) { return os << a[0] << " " << a[1] << " " << a[2]; }

struct Vector4f {
  float v[4] = {0.f, 0.f, 0.f, 0.f};
  Vector4f() {}
  Vector4f(float a, float b, float c, float d) { v[0] = a; v[1] = b; v[2] = c; v[3] = d; }
  float &operator[](int i) { return v[i]; }
  float operator[](int i) const { return v[i]; }
  float &operator()(int i) { return v[i]; }
  float operator()(int i) const { return v[i]; }
  Vector3f head3() const { return Vector3f(v[0], v[1], v[2]); }
};

struct Matrix3f {
  float m[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f};  // row-major
  static Matrix3f Identity() { return Matrix3f(); }
  float &operator()(int r, int c) { return m[3 * r + c]; }
  float operator()(int r, int c) const { return m[3 * r + c]; }
  void setRow(int r, const Vector3f &a) { m[3 * r] = a[0]; m[3 * r + 1] = a[1]; m[3 * r + 2] = a[2]; }
  // coefficient-wise lazy product, 3-term redux per coefficient
  Vector3f operator*(const Vector3f &p) const {
    return Vector3f(m[0] * p[0] + (m[1] * p[1] + m[2] * p[2]), m[3] * p[0] + (m[4] * p[1] + m[5] * p[2]),
                    m[6] * p[0] + (m[7] * p[1] + m[8] * p[2]));
  }
  Matrix3f operator*(const Matrix3f &o) const {
    Matrix3f r;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) r.m[3 * i + j] = m[3 * i] * o.m[j] + (m[3 * i + 1] * o.m[3 + j] + m[3 * i + 2] * o.m[6 + j]);
    return r;
  }
  Matrix3f inverse() const {
    Matrix3f r;
    rtp::inverse3(m, 3, r.m, 3);
    return r;
  }
};

// Rotation about a unit axis: AngleAxis::toRotationMatrix (Eigen/src/Geometry/AngleAxis.h:218-242).
struct AngleAxisf {
  float angle;
  Vector3f axis;
  AngleAxisf(float a, const Vector3f &ax) : angle(a), axis(ax) {}
  Matrix3f toRotationMatrix() const {
    Matrix3f res;
    const float s = std::sin(angle), c = std::cos(angle);
    const Vector3f sin_axis = axis * s;
    const Vector3f cos1_axis = axis * (1.f - c);
    float tmp;
    tmp = cos1_axis.x() * axis.y(); res(0, 1) = tmp - sin_axis.z(); res(1, 0) = tmp + sin_axis.z();
    tmp = cos1_axis.x() * axis.z(); res(0, 2) = tmp + sin_axis.y(); res(2, 0) = tmp - sin_axis.y();
    tmp = cos1_axis.y() * axis.z(); res(1, 2) = tmp - sin_axis.x(); res(2, 1) = tmp + sin_axis.x();
    res(0, 0) = cos1_axis.x() * axis.x() + c;
    res(1, 1) = cos1_axis.y() * axis.y() + c;
    res(2, 2) = cos1_axis.z() * axis.z() + c;
    return res;
  }
  Vector3f operator*(const Vector3f &p) const { return toRotationMatrix() * p; }
};

// Only carried along in dirLights' third tuple member (never read by the ray tracer).
template <class T>
struct Quaternion {
  T w = 1, x = 0, y = 0, z = 0;
  Quaternion() {}
  Quaternion(const Matrix3f &) {}
};
typedef Quaternion<float> Quaternionf;

struct Affine3f {
  float m[12] = {1.f, 0.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 0.f, 1.f, 0.f};  // 3x4 row-major [L | t]
  static Affine3f Identity() { return Affine3f(); }
  float &operator()(int r, int c) { return m[4 * r + c]; }
  float operator()(int r, int c) const { return m[4 * r + c]; }
  const float *data() const { return m; }
  Matrix3f linear() const {
    Matrix3f r;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) r(i, j) = m[4 * i + j];
    return r;
  }
  Matrix3f rotation() const { return linear(); }  // exact only for the rigid views used here
  Vector3f translation() const { return Vector3f(m[3], m[7], m[11]); }
  void setLinear(const Matrix3f &l) {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) m[4 * i + j] = l(i, j);
  }
  // Transform::scale(s): linear *= s
  Affine3f &scale(float s) {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) m[4 * i + j] *= s;
    return *this;
  }
  // Transform::translate(v): t += linear * v
  Affine3f &translate(const Vector3f &p) {
    const Vector3f d = linear() * p;
    m[3] += d[0]; m[7] += d[1]; m[11] += d[2];
    return *this;
  }
  // Transform::rotate(R): linear = linear * R
  Affine3f &rotate(const Matrix3f &r) {
    setLinear(linear() * r);
    return *this;
  }
  Vector3f operator*(const Vector3f &p) const { return Vector3f(rtp::affine_mul(m, p.r())); }
  Vector4f operator*(const Vector4f &p) const {  // 4x4 * (x, y, z, w), last row (0 0 0 1)
    return Vector4f(((m[0] * p[0] + m[1] * p[1]) + m[2] * p[2]) + m[3] * p[3],
                    ((m[4] * p[0] + m[5] * p[1]) + m[6] * p[2]) + m[7] * p[3],
                    ((m[8] * p[0] + m[9] * p[1]) + m[10] * p[2]) + m[11] * p[3], p[3]);
  }
  // affine * affine: [L1 L2 | L1 t2 + t1]
  Affine3f operator*(const Affine3f &o) const {
    Affine3f r;
    r.setLinear(linear() * o.linear());
    const Vector3f t = linear() * o.translation();
    r.m[3] = t[0] + m[3]; r.m[7] = t[1] + m[7]; r.m[11] = t[2] + m[11];
    return r;
  }
  Affine3f inverse() const {
    Affine3f r;
    rtp::affine_inverse(m, r.m);
    return r;
  }
};

}  // namespace Eigen
