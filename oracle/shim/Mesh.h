// TEST INFRASTRUCTURE (oracle shim). Class shell for mymesh.cpp; members named as the
// reference uses them (mymesh.cpp:25-236, mybvh.cpp:220-259, mytracer.cpp:219-287).
#ifndef SHIM_MESH_H
#define SHIM_MESH_H
#include "Object.h"
#include "utils/Image.h"
#include "utils/Triangle.h"
#include <vector>
enum Draw_mode { FLAT = 0, PHONG = 1 };
class Mesh : public Object {
public:
  Mesh() {}
  // triangle meshes are traced through the BVH, never through Object::intersect
  bool intersect(const Ray &, vec4 &, vec4 &, vec4 &, double &) const override { return false; }
  bool intersect_bounding_box(const Ray &ray) const;
  void compute_texture(int iuv0, int iuv1, int iuv2, vec4 &intersection_diffuse, const double &alpha,
                       const double &beta, const double &gamma) const;
  void compute_normals();
  bool intersect_triangle(const Triangle &triangle, const Ray &ray, vec4 &intersection_point,
                          vec4 &intersection_normal, vec4 &intersection_diffuse,
                          double &intersection_distance) const;
  std::vector<Vertex> vertices_;
  std::vector<Triangle> triangles_;
  std::vector<double> u_coordinates_, v_coordinates_;
  Image texture_;
  bool hasTexture_ = false;
  Draw_mode draw_mode_ = FLAT;
  vec4 bb_min_, bb_max_;
};
#endif
