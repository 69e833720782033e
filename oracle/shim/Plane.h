// TEST INFRASTRUCTURE (oracle shim). Class shell for myplane.cpp:22-49.
#ifndef SHIM_PLANE_H
#define SHIM_PLANE_H
#include "Object.h"
class Plane : public Object {
public:
  Plane() {}
  Plane(const vec4 &c, const vec4 &n) : center_(c), normal_(n) {}
  bool intersect(const Ray &ray, vec4 &intersection_point, vec4 &intersection_normal,
                 vec4 &intersection_diffuse, double &intersection_distance) const override;
  vec4 center_, normal_;
};
#endif
