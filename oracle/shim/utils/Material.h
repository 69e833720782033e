// TEST INFRASTRUCTURE (oracle shim). Fields as used at mytracer_gpu.cu:381-386.
#ifndef SHIM_MATERIAL_H
#define SHIM_MATERIAL_H
#include "utils/vec4.h"
struct Material {
  vec4 ambient, diffuse, specular;
  double shininess = 0.0;
  double mirror = 0.0;
  bool shadowable = true;
};
#endif
