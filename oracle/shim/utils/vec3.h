// TEST INFRASTRUCTURE (oracle shim). Only det3D (myutils.cpp:33-37) touches vec3.
#ifndef SHIM_VEC3_H
#define SHIM_VEC3_H
struct vec3 {
  double d_[3];
  vec3() : d_{0.0, 0.0, 0.0} {}
  vec3(double x, double y, double z) : d_{x, y, z} {}
  double &operator[](int i) { return d_[i]; }
  const double &operator[](int i) const { return d_[i]; }
};
#endif
