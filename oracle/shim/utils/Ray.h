// TEST INFRASTRUCTURE (oracle shim). Ray{origin_, direction_}; the constructor
// normalises the direction (build-defined convention, DESIGN.md "Conventions").
#ifndef SHIM_RAY_H
#define SHIM_RAY_H
#include "utils/vec4.h"
struct Ray {
  vec4 origin_, direction_;
  Ray() {}
  Ray(const vec4 &o, const vec4 &d) : origin_(o), direction_(normalize(d)) {}
  vec4 operator()(double t) const { return origin_ + t * direction_; }
};
#endif
