// TEST INFRASTRUCTURE (oracle shim) — not product code.
// Stand-in for the absent course-framework header `utils/vec4.h`, written from the
// uses visible in the reference: 4 doubles, 3 of them used.
// Every operation is a plain left-to-right double expression so that the CUDA path
// and the restated oracle can reproduce it bit for bit.
#ifndef SHIM_VEC4_H
#define SHIM_VEC4_H
#include <algorithm>
#include <cmath>
#include <iostream>
#include <stdexcept>
#include <vector>

struct vec4 {
  double d_[4];
  vec4() : d_{0.0, 0.0, 0.0, 0.0} {}
  explicit vec4(double s) : d_{s, s, s, s} {}
  vec4(double x, double y, double z, double w = 0.0) : d_{x, y, z, w} {}
  double &operator[](int i) { return d_[i]; }
  const double &operator[](int i) const { return d_[i]; }
  vec4 &operator+=(const vec4 &o) { d_[0] += o.d_[0]; d_[1] += o.d_[1]; d_[2] += o.d_[2]; return *this; }
  vec4 &operator-=(const vec4 &o) { d_[0] -= o.d_[0]; d_[1] -= o.d_[1]; d_[2] -= o.d_[2]; return *this; }
  vec4 &operator*=(double s) { d_[0] *= s; d_[1] *= s; d_[2] *= s; return *this; }
  vec4 &operator/=(double s) { d_[0] /= s; d_[1] /= s; d_[2] /= s; return *this; }
};

inline vec4 operator+(const vec4 &a, const vec4 &b) { return vec4(a[0] + b[0], a[1] + b[1], a[2] + b[2]); }
inline vec4 operator-(const vec4 &a, const vec4 &b) { return vec4(a[0] - b[0], a[1] - b[1], a[2] - b[2]); }
inline vec4 operator-(const vec4 &a) { return vec4(-a[0], -a[1], -a[2]); }
inline vec4 operator*(double s, const vec4 &a) { return vec4(s * a[0], s * a[1], s * a[2]); }
inline vec4 operator*(const vec4 &a, double s) { return vec4(s * a[0], s * a[1], s * a[2]); }
inline vec4 operator/(const vec4 &a, double s) { return vec4(a[0] / s, a[1] / s, a[2] / s); }
inline double dot(const vec4 &a, const vec4 &b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline vec4 cross(const vec4 &a, const vec4 &b) {
  return vec4(a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]);
}
inline double normSq(const vec4 &a) { return dot(a, a); }
inline double norm(const vec4 &a) { return std::sqrt(dot(a, a)); }
inline vec4 normalize(const vec4 &a) { return a / norm(a); }
// reflect: incoming direction v bounced off a surface with normal n  -> v - 2(n.v)n
inline vec4 reflect(const vec4 &v, const vec4 &n) { return v - (2.0 * dot(n, v)) * n; }
// mirror: direction v (pointing away from the surface) mirrored about n -> 2(n.v)n - v
inline vec4 mirror(const vec4 &v, const vec4 &n) { return (2.0 * dot(n, v)) * n - v; }
inline vec4 min(const vec4 &a, const vec4 &b) {
  return vec4(std::min(a[0], b[0]), std::min(a[1], b[1]), std::min(a[2], b[2]));
}
inline vec4 max(const vec4 &a, const vec4 &b) {
  return vec4(std::max(a[0], b[0]), std::max(a[1], b[1]), std::max(a[2], b[2]));
}
inline std::ostream &operator<<(std::ostream &os, const vec4 &a) {
  return os << a[0] << " " << a[1] << " " << a[2];
}
static_assert(sizeof(vec4) == 32, "vec4 must be 4 doubles");
#endif
