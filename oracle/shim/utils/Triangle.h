// TEST INFRASTRUCTURE (oracle shim). Fields as used at mybvh.cpp:226-231 and
// mytracer.cpp:226-252.
#ifndef SHIM_TRIANGLE_H
#define SHIM_TRIANGLE_H
#include "utils/vec4.h"
struct Vertex {
  vec4 position, normal;
};
struct Triangle {
  int i0 = 0, i1 = 0, i2 = 0;
  int iuv0 = 0, iuv1 = 0, iuv2 = 0;
  vec4 normal, centroid;
  int meshIdx = 0;
};
#endif
