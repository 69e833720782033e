// TEST INFRASTRUCTURE (oracle shim). Sphere is absent from the reference
// (SURVEY.md 8a row a10); Sphere::intersect is restated in oracle/ref_glue.cpp.
#ifndef SHIM_SPHERE_H
#define SHIM_SPHERE_H
#include "Object.h"
class Sphere : public Object {
public:
  Sphere() {}
  Sphere(const vec4 &c, double r) : center_(c), radius_(r) {}
  bool intersect(const Ray &ray, vec4 &intersection_point, vec4 &intersection_normal,
                 vec4 &intersection_diffuse, double &intersection_distance) const override;
  vec4 center_;
  double radius_ = 1.0;
};
#endif
