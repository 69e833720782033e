// TEST INFRASTRUCTURE — glue around the reference's own CPU sources (oracle/_ref build).
//
// The reference's mybvh.cpp / mymesh.cpp / myplane.cpp / myutils.cpp / mytracer.cpp are
// compiled UNMODIFIED by path from /root/reference (oracle/Makefile).  This file restates
// only what the absent course framework held:
//   Raytracer::trace            structure follows mytracer_gpu.cu:237-312 + mytracer.cpp:546-555
//   Raytracer::intersect_scene  follows mytracer_gpu.cu:314-328 + mybvh.cpp:147-178 and the
//                               analytic-primitive conventions of myplane.cpp:22-49
//   Raytracer::compute_image    follows mytracer_gpu.cu:119-160,199-221
//   Sphere::intersect           build-defined, conventions of myplane.cpp:22-49
// plus an ID-tracking twin of BVH::intersectBVH (mybvh.cpp:147-210) and the C ABI of
// oracle_abi.h.  Never linked into the product.
#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstring>
#include <iostream>
#include <math.h>
#include <omp.h>
#include <stdexcept>
#include <string>
#include <vector>

#define private public // the ID twin walks BVH::bvhNodes_/triangles_/meshes_ (test-only)
#include "Raytracer.h"
#undef private
#include "Plane.h"
#include "Sphere.h"
#include "oracle_abi.h"

// ---------------------------------------------------------------------------------------
// restated framework pieces
// ---------------------------------------------------------------------------------------
namespace {
struct alignas(64) Counters {
  long long scene_calls = 0;   // every intersect_scene call = one ray cast
  long long reflect_rays = 0;  // trace() calls at depth >= 1 that cast a ray
};
Counters g_cnt[256];
inline Counters &cnt() { return g_cnt[omp_get_thread_num() & 255]; }
} // namespace

bool Sphere::intersect(const Ray &ray, vec4 &intersection_point, vec4 &intersection_normal,
                       vec4 &intersection_diffuse, double &intersection_distance) const {
  intersection_diffuse = material_.diffuse;
  const vec4 oc = ray.origin_ - center_;
  const double a = dot(ray.direction_, ray.direction_);
  const double b = 2.0 * dot(ray.direction_, oc);
  const double c = dot(oc, oc) - radius_ * radius_;
  const double disc = b * b - 4.0 * a * c;
  if (disc < 0.0) return false;
  const double sq = std::sqrt(disc);
  const double t1 = (-b - sq) / (2.0 * a);
  const double t2 = (-b + sq) / (2.0 * a);
  double t = t1;
  if (!(t > 1e-5)) t = t2;
  if (!(t > 1e-5)) return false;
  intersection_distance = t;
  intersection_point = ray(t);
  intersection_normal = (intersection_point - center_) / radius_;
  return true;
}

void Raytracer::read_scene(const std::string &) {
  throw std::runtime_error("read_scene: scene files are not part of the reference repo");
}

bool Raytracer::intersect_scene(const Ray &ray, Material &material, vec4 &point, vec4 &normal,
                                double &t_out) const {
  cnt().scene_calls++;
  double tmin(DBL_MAX);
  double t;
  vec4 p, n, d;
  for (const Object *o : objects_) {
    if (o->intersect(ray, p, n, d, t)) {
      if (t < tmin) {
        tmin = t;
        material = o->material_;
        material.diffuse = d;
        point = p;
        normal = n;
      }
    }
  }
  if (bvh.isInitialized()) {
    bvh.intersectBVH(ray, material, point, normal, tmin, 0);
  }
  if (tmin < DBL_MAX) t_out = tmin;
  return (tmin < DBL_MAX);
}

vec4 Raytracer::trace(const Ray &ray, int depth) {
  if (depth > max_depth_) return vec4(0.0, 0.0, 0.0);
  if (depth > 0) cnt().reflect_rays++;
  Material material;
  vec4 point, normal;
  double t;
  if (!intersect_scene(ray, material, point, normal, t)) return background_;
  vec4 color = (1.0 - material.mirror) * lighting(point, normal, -ray.direction_, material);
  color += subtrace(ray, material, point, normal, depth);
  return color;
}

void Raytracer::compute_image() {
  image_.resize(camera_.width_, camera_.height_);
#pragma omp parallel for schedule(dynamic, 1)
  for (int y = 0; y < camera_.height_; ++y) {
    for (int x = 0; x < camera_.width_; ++x) {
      vec4 color = trace(camera_.primary_ray(x, y), 0);
      image_(x, y) = min(color, vec4(1, 1, 1));
    }
  }
}

// ---------------------------------------------------------------------------------------
// ID-tracking twin of BVH::intersectBVH (same visit order, same strict '<')
// ---------------------------------------------------------------------------------------
namespace {
void bvh_twin(const BVH &bvh, const Ray &ray, int nodeIdx, double &best, int &best_id) {
  const BVH::BVHNode &node = bvh.bvhNodes_[nodeIdx];
  double dummy;
  if (!bvh.intersectAABB(ray, node.bb_min_, node.bb_max_, dummy)) return;
  if (node.triCount_ > 0) {
    double t;
    vec4 p, n, d;
    for (int i = node.firstTriIdx_; i < node.firstTriIdx_ + node.triCount_; i++) {
      Triangle *tri = bvh.triangles_.at(i);
      Mesh *mesh = bvh.meshes_.at(tri->meshIdx);
      if (mesh->intersect_triangle(*tri, ray, p, n, d, t)) {
        if (t < best) {
          best = t;
          best_id = i;
        }
      }
    }
  } else {
    const int l = node.leftChildIdx_;
    if (bvh.intersectAABB(ray, bvh.bvhNodes_[l].bb_min_, bvh.bvhNodes_[l].bb_max_, dummy))
      bvh_twin(bvh, ray, l, best, best_id);
    if (bvh.intersectAABB(ray, bvh.bvhNodes_[l + 1].bb_min_, bvh.bvhNodes_[l + 1].bb_max_, dummy))
      bvh_twin(bvh, ray, l + 1, best, best_id);
  }
}

struct Handle {
  Raytracer rt;
  long long counts[3] = {0, 0, 0};
  int ntri() const { return rt.bvh.isInitialized() ? rt.bvh.triCount : 0; }
  // nearest hit with ID, same order as Raytracer::intersect_scene
  int nearest_id(const Ray &ray, double &tbest) const {
    tbest = DBL_MAX;
    int id = -1;
    double t;
    vec4 p, n, d;
    for (size_t k = 0; k < rt.objects_.size(); ++k) {
      if (rt.objects_[k]->intersect(ray, p, n, d, t) && t < tbest) {
        tbest = t;
        id = ntri() + int(k);
      }
    }
    if (rt.bvh.isInitialized()) bvh_twin(rt.bvh, ray, 0, tbest, id);
    return id;
  }
};

Material make_material(const double *m) {
  Material mat;
  mat.ambient = vec4(m[0], m[1], m[2]);
  mat.diffuse = vec4(m[3], m[4], m[5]);
  mat.specular = vec4(m[6], m[7], m[8]);
  mat.shininess = m[9];
  mat.mirror = m[10];
  mat.shadowable = (m[11] != 0.0);
  return mat;
}
inline vec4 v3(const double *p) { return vec4(p[0], p[1], p[2]); }
} // namespace

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
extern "C" {

void *orc_new(void) { return new Handle(); }
void orc_free(void *h) { delete static_cast<Handle *>(h); }
const char *orc_kind(void) { return "reference"; }

int orc_add_mesh(void *h, int nv, const double *pos, int nt, const int *idx, const int *uvidx, int nuv,
                 const double *u, const double *v, int texW, int texH, const double *texels,
                 const double *material, int draw_mode) {
  Handle *H = static_cast<Handle *>(h);
  Mesh *m = new Mesh();
  m->vertices_.resize(nv);
  for (int i = 0; i < nv; ++i) m->vertices_[i].position = v3(pos + 3 * i);
  m->triangles_.resize(nt);
  for (int i = 0; i < nt; ++i) {
    Triangle &t = m->triangles_[i];
    t.i0 = idx[3 * i], t.i1 = idx[3 * i + 1], t.i2 = idx[3 * i + 2];
    if (uvidx) t.iuv0 = uvidx[3 * i], t.iuv1 = uvidx[3 * i + 1], t.iuv2 = uvidx[3 * i + 2];
  }
  if (nuv > 0 && u && v) {
    m->u_coordinates_.assign(u, u + nuv);
    m->v_coordinates_.assign(v, v + nuv);
  }
  if (texW > 0 && texH > 0 && texels) {
    m->texture_.resize(texW, texH);
    for (int y = 0; y < texH; ++y)
      for (int x = 0; x < texW; ++x) m->texture_(x, y) = v3(texels + 3 * (size_t(y) * texW + x));
    m->hasTexture_ = true;
  }
  m->material_ = make_material(material);
  m->draw_mode_ = draw_mode == 1 ? PHONG : FLAT;
  H->rt.meshes_.push_back(m);
  return int(H->rt.meshes_.size()) - 1;
}

int orc_add_sphere(void *h, const double *center, double radius, const double *material) {
  Handle *H = static_cast<Handle *>(h);
  Sphere *s = new Sphere(v3(center), radius);
  s->material_ = make_material(material);
  H->rt.objects_.push_back(s);
  return int(H->rt.objects_.size()) - 1;
}

int orc_add_plane(void *h, const double *center, const double *normal, const double *material) {
  Handle *H = static_cast<Handle *>(h);
  Plane *p = new Plane(v3(center), v3(normal));
  p->material_ = make_material(material);
  H->rt.objects_.push_back(p);
  return int(H->rt.objects_.size()) - 1;
}

int orc_add_light(void *h, const double *pos, const double *color) {
  Handle *H = static_cast<Handle *>(h);
  H->rt.lights_.push_back(Light{v3(pos), v3(color)});
  return int(H->rt.lights_.size()) - 1;
}

void orc_set_camera(void *h, const double *eye, const double *center, const double *up, double fovy,
                    int width, int height) {
  static_cast<Handle *>(h)->rt.camera_ = Camera(v3(eye), v3(center), v3(up), fovy, width, height);
}

void orc_get_camera(void *h, double *out) {
  const Camera &c = static_cast<Handle *>(h)->rt.camera_;
  for (int i = 0; i < 3; ++i) {
    out[i] = c.eye_[i];
    out[3 + i] = c.x_dir_[i];
    out[6 + i] = c.y_dir_[i];
    out[9 + i] = c.lower_left_[i];
  }
}

void orc_set_globals(void *h, const double *background, const double *ambience, int max_depth) {
  Handle *H = static_cast<Handle *>(h);
  H->rt.background_ = v3(background);
  H->rt.ambience_ = v3(ambience);
  H->rt.max_depth_ = max_depth;
}

void orc_build(void *h) {
  Handle *H = static_cast<Handle *>(h);
  for (Mesh *m : H->rt.meshes_) m->compute_normals(); // mymesh.cpp:103-163
  std::streambuf *old = std::cout.rdbuf(nullptr);     // silence the reference's progress prints
  H->rt.bvh.init(H->rt.meshes_);                      // mybvh.cpp:44-81
  std::cout.rdbuf(old);
}

int orc_num_triangles(void *h) { return static_cast<Handle *>(h)->ntri(); }
int orc_num_nodes(void *h) {
  Handle *H = static_cast<Handle *>(h);
  return H->rt.bvh.isInitialized() ? H->rt.bvh.nodesUsed_ : 0;
}

void orc_get_bvh(void *h, double *bbmin, double *bbmax, int *left, int *first, int *count) {
  const BVH &b = static_cast<Handle *>(h)->rt.bvh;
  for (int i = 0; i < b.nodesUsed_; ++i) {
    const BVH::BVHNode &n = b.bvhNodes_[i];
    for (int k = 0; k < 3; ++k) bbmin[3 * i + k] = n.bb_min_[k], bbmax[3 * i + k] = n.bb_max_[k];
    left[i] = n.leftChildIdx_, first[i] = n.firstTriIdx_, count[i] = n.triCount_;
  }
}

void orc_get_tri_order(void *h, int *perm) {
  const BVH &b = static_cast<Handle *>(h)->rt.bvh;
  std::vector<int> base(b.meshes_.size() + 1, 0);
  for (size_t m = 0; m < b.meshes_.size(); ++m) base[m + 1] = base[m] + int(b.meshes_[m]->triangles_.size());
  for (int i = 0; i < b.triCount; ++i) {
    const Triangle *t = b.triangles_[i];
    perm[i] = base[t->meshIdx] + int(t - b.meshes_[t->meshIdx]->triangles_.data());
  }
}

void orc_get_normals(void *h, int mesh, double *fn, double *vn) {
  const Mesh *m = static_cast<Handle *>(h)->rt.meshes_.at(mesh);
  for (size_t i = 0; i < m->triangles_.size(); ++i)
    for (int k = 0; k < 3; ++k) fn[3 * i + k] = m->triangles_[i].normal[k];
  for (size_t i = 0; i < m->vertices_.size(); ++i)
    for (int k = 0; k < 3; ++k) vn[3 * i + k] = m->vertices_[i].normal[k];
}

double orc_render(void *h, double *rgb, int *ids, double *tout, int y0, int y1, int ystep, int spp_grid,
                  int nthreads) {
  Handle *H = static_cast<Handle *>(h);
  Raytracer &rt = H->rt;
  const int W = rt.camera_.width_, Hh = rt.camera_.height_;
  if (y1 > Hh) y1 = Hh;
  if (ystep < 1) ystep = 1;
  if (nthreads > 0) omp_set_num_threads(nthreads);
  for (auto &c : g_cnt) c = Counters();
  const int nrows = (y1 - y0 + ystep - 1) / ystep;
  const int s = spp_grid < 1 ? 1 : spp_grid;
  const auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1)
  for (int r = 0; r < nrows; ++r) {
    const int y = y0 + r * ystep;
    for (int x = 0; x < W; ++x) {
      vec4 color(0.0, 0.0, 0.0);
      if (s == 1) {
        color = rt.trace(rt.camera_.primary_ray(x, y), 0);
      } else {
        for (int si = 0; si < s; si++) {
          const double xo = (si) / static_cast<double>(s) - 0.5 + 1.0 / (2.0 * s);
          for (int sj = 0; sj < s; sj++) {
            const double yo = (sj) / static_cast<double>(s) - 0.5 + 1.0 / (2.0 * s);
            color += rt.trace(rt.camera_.primary_ray(double(x) + xo, double(y) + yo), 0);
          }
        }
        color /= double(s * s);
      }
      color = min(color, vec4(1, 1, 1));
      const size_t p = size_t(y) * W + x;
      rgb[3 * p] = color[0], rgb[3 * p + 1] = color[1], rgb[3 * p + 2] = color[2];
    }
  }
  const auto t1 = std::chrono::steady_clock::now();
  long long scene = 0, refl = 0;
  for (auto &c : g_cnt) scene += c.scene_calls, refl += c.reflect_rays;
  const long long prim = (long long)nrows * W * s * s;
  H->counts[0] = prim, H->counts[1] = scene - prim - refl, H->counts[2] = refl;
  if (ids || tout) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int r = 0; r < nrows; ++r) {
      const int y = y0 + r * ystep;
      for (int x = 0; x < W; ++x) {
        double tb;
        const int id = H->nearest_id(rt.camera_.primary_ray(x, y), tb);
        const size_t p = size_t(y) * W + x;
        if (ids) ids[p] = id;
        if (tout) tout[p] = (id >= 0) ? tb : -1.0;
      }
    }
  }
  return std::chrono::duration<double>(t1 - t0).count();
}

void orc_ray_counts(void *h, long long *out3) {
  Handle *H = static_cast<Handle *>(h);
  out3[0] = H->counts[0], out3[1] = H->counts[1], out3[2] = H->counts[2];
}

void orc_trace_rays(void *h, int n, const double *origins, const double *dirs, double *rgb, int *ids,
                    double *tout) {
  Handle *H = static_cast<Handle *>(h);
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < n; ++i) {
    const Ray ray(v3(origins + 3 * i), v3(dirs + 3 * i));
    if (rgb) {
      const vec4 c = H->rt.trace(ray, 0);
      rgb[3 * i] = c[0], rgb[3 * i + 1] = c[1], rgb[3 * i + 2] = c[2];
    }
    if (ids || tout) {
      double tb;
      const int id = H->nearest_id(ray, tb);
      // the twin must agree with the unmodified reference traversal
      Material m;
      vec4 p, nn;
      double tt = -1.0;
      const bool hit = H->rt.intersect_scene(ray, m, p, nn, tt);
      if (hit != (id >= 0) || (hit && tt != tb)) throw std::runtime_error("ID twin disagrees");
      if (ids) ids[i] = id;
      if (tout) tout[i] = (id >= 0) ? tb : -1.0;
    }
  }
}

int orc_tri_test(const double *p0, const double *p1, const double *p2, const double *o, const double *d,
                 double *out4) {
  Mesh m;
  m.vertices_.resize(3);
  m.vertices_[0].position = v3(p0), m.vertices_[1].position = v3(p1), m.vertices_[2].position = v3(p2);
  Triangle t;
  t.i0 = 0, t.i1 = 1, t.i2 = 2;
  vec4 p, n, dif;
  double dist = 0.0;
  Ray ray;
  ray.origin_ = v3(o), ray.direction_ = v3(d); // NOT normalised: raw known-answer input
  const bool hit = m.intersect_triangle(t, ray, p, n, dif, dist);
  if (hit) out4[0] = dist, out4[1] = p[0], out4[2] = p[1], out4[3] = p[2];
  return hit ? 1 : 0;
}

int orc_aabb_test(const double *bbmin, const double *bbmax, const double *o, const double *d,
                  double *tmin_out) {
  BVH b;
  Ray ray;
  ray.origin_ = v3(o), ray.direction_ = v3(d);
  double tmin = 0.0;
  const bool hit = b.intersectAABB(ray, v3(bbmin), v3(bbmax), tmin);
  if (tmin_out) *tmin_out = tmin;
  return hit ? 1 : 0;
}

int orc_plane_test(const double *center, const double *normal, const double *o, const double *d,
                   double *out4) {
  Plane pl(v3(center), v3(normal));
  Ray ray;
  ray.origin_ = v3(o), ray.direction_ = v3(d);
  vec4 p, n, dif;
  double t = 0.0;
  const bool hit = pl.intersect(ray, p, n, dif, t);
  if (hit) out4[0] = t, out4[1] = p[0], out4[2] = p[1], out4[3] = p[2];
  return hit ? 1 : 0;
}

int orc_sphere_test(const double *center, double radius, const double *o, const double *d, double *out7) {
  Sphere s(v3(center), radius);
  Ray ray;
  ray.origin_ = v3(o), ray.direction_ = v3(d);
  vec4 p, n, dif;
  double t = 0.0;
  const bool hit = s.intersect(ray, p, n, dif, t);
  if (hit) {
    out7[0] = t;
    for (int k = 0; k < 3; ++k) out7[1 + k] = p[k], out7[4 + k] = n[k];
  }
  return hit ? 1 : 0;
}

void orc_lighting(void *h, const double *point, const double *normal, const double *view,
                  const double *material, double *rgb) {
  Handle *H = static_cast<Handle *>(h);
  const vec4 c = H->rt.lighting(v3(point), v3(normal), v3(view), make_material(material));
  rgb[0] = c[0], rgb[1] = c[1], rgb[2] = c[2];
}

void orc_texture(void *h, int tri, double alpha, double beta, double gamma, double *rgb) {
  const BVH &b = static_cast<Handle *>(h)->rt.bvh;
  const Triangle *t = b.triangles_.at(tri);
  const Mesh *m = b.meshes_.at(t->meshIdx);
  vec4 c = m->material_.diffuse;
  if (m->hasTexture_) m->compute_texture(t->iuv0, t->iuv1, t->iuv2, c, alpha, beta, gamma);
  rgb[0] = c[0], rgb[1] = c[1], rgb[2] = c[2];
}

} // extern "C"
