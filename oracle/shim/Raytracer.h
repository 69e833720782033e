// TEST INFRASTRUCTURE (oracle shim). Class declaration for mytracer.cpp (CPU half):
// members as seen at mytracer.cpp:27-47,510-608.  trace / intersect_scene /
// compute_image / read_scene are absent from the reference and restated in
// oracle/ref_glue.cpp.
#ifndef SHIM_RAYTRACER_H
#define SHIM_RAYTRACER_H
#include "Mesh.h"
#include "Object.h"
#include "mybvh.h"
#include "utils/Camera.h"
#include "utils/Image.h"
#include <string>
#include <vector>
struct Light {
  vec4 position, color;
};
class Raytracer {
public:
  Raytracer() {}
  ~Raytracer();
  void init_cpu(const std::string &filename);
  void read_scene(const std::string &filename);
  void compute_image();
  vec4 trace(const Ray &ray, int depth);
  bool intersect_scene(const Ray &ray, Material &material, vec4 &point, vec4 &normal, double &t) const;
  double diffuse(const vec4 &point, const vec4 &normal, const Light &light) const;
  double reflection(const vec4 &point, const vec4 &normal, const vec4 &view, const Light &light) const;
  vec4 subtrace(const Ray &ray, const Material &material, const vec4 &point, const vec4 &normal,
                const int depth);
  vec4 lighting(const vec4 &point, const vec4 &normal, const vec4 &view, const Material &material) const;

  std::vector<Object *> objects_;
  std::vector<Mesh *> meshes_;
  std::vector<Light> lights_;
  Camera camera_;
  Image image_;
  vec4 background_, ambience_;
  int max_depth_ = 0;
  BVH bvh;
};
#endif
