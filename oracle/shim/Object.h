// TEST INFRASTRUCTURE (oracle shim). Base class of analytic primitives held in
// Raytracer::objects_ (mytracer.cpp:27).
#ifndef SHIM_OBJECT_H
#define SHIM_OBJECT_H
#include "utils/Material.h"
#include "utils/Ray.h"
#include "utils/vec4.h"
class Object {
public:
  virtual ~Object() {}
  virtual bool intersect(const Ray &ray, vec4 &point, vec4 &normal, vec4 &diffuse, double &t) const = 0;
  Material material_;
};
#endif
