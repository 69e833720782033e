// TEST INFRASTRUCTURE — plain C++ CPU restatement of the reference's hot path ("port" oracle).
//
// Parity status: PINNED against the reference's own CPU sources.  The reference ships no tests or
// golden vectors, so this restatement is checked bit for bit against
// oracle/_ref (mybvh.cpp, mymesh.cpp, myplane.cpp, myutils.cpp, mytracer.cpp compiled unmodified)
// in tests/test_oracle.py and against the fixtures in tests/golden/ that oracle/_ref generated.
// The pieces the reference itself lacks (trace, intersect_scene, compute_image, Sphere, Camera,
// Ray, vec4) are build-defined conventions and are therefore "unpinned" by
// construction; both oracles share them.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  Compiled with -ffp-contract=off: every a*b+c below is two roundings, in
// the operand order of the cited reference lines.
#include "oracle_abi.h"
#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <omp.h>
#include <vector>

namespace {

struct V3 {
  double x = 0, y = 0, z = 0;
};
inline V3 v3(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
inline V3 v3p(const double *p) { return v3(p[0], p[1], p[2]); }
inline V3 operator+(const V3 &a, const V3 &b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(const V3 &a, const V3 &b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(const V3 &a) { return v3(-a.x, -a.y, -a.z); }
inline V3 operator*(double s, const V3 &a) { return v3(s * a.x, s * a.y, s * a.z); }
inline V3 operator/(const V3 &a, double s) { return v3(a.x / s, a.y / s, a.z / s); }
inline double dot(const V3 &a, const V3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(const V3 &a, const V3 &b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline double norm(const V3 &a) { return std::sqrt(dot(a, a)); }
inline V3 normalize(const V3 &a) { return a / norm(a); }
inline V3 reflect(const V3 &v, const V3 &n) { return v - (2.0 * dot(n, v)) * n; }
inline V3 mirror(const V3 &v, const V3 &n) { return (2.0 * dot(n, v)) * n - v; }
inline double comp(const V3 &a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }

struct Ray {
  V3 o, d;
};
inline Ray make_ray(const V3 &o, const V3 &dir) { Ray r; r.o = o; r.d = normalize(dir); return r; }
inline V3 at(const Ray &r, double t) { return r.o + t * r.d; }

struct Mat {
  V3 ka, kd, ks;
  double shininess = 0, mirror = 0;
  bool shadowable = true;
};
Mat make_mat(const double *m) {
  Mat r;
  r.ka = v3p(m); r.kd = v3p(m + 3); r.ks = v3p(m + 6);
  r.shininess = m[9]; r.mirror = m[10]; r.shadowable = m[11] != 0.0;
  return r;
}

struct MeshO {
  std::vector<V3> pos, vnormal, fnormal;
  std::vector<int> idx, uvidx;
  std::vector<double> u, v;
  int texW = 0, texH = 0;
  std::vector<V3> tex;
  Mat mat;
  int draw_mode = 0;
  int tri_base = 0;
};
struct Obj {
  int kind; // 0 sphere, 1 plane
  V3 center, normal;
  double radius;
  Mat mat;
};
struct Light {
  V3 pos, color;
};
struct TriRef {
  int mesh, local;
  V3 centroid;
};
struct Node {
  V3 mn, mx;
  int left = 0, first = 0, count = 0;
};
struct Cam {
  V3 eye, xdir, ydir, ll;
  int w = 0, h = 0;
};
struct alignas(64) Cnt {
  long long scene = 0, refl = 0;
};

// myutils.cpp:21-23,47-51
inline double det2(double a, double b, double c, double d) { return a * d - b * c; }
inline double det3(const V3 &a, const V3 &b, const V3 &c) {
  return a.x * det2(b.y, c.y, b.z, c.z) - b.x * det2(a.y, c.y, a.z, c.z) + c.x * det2(a.y, b.y, a.z, b.z);
}

struct Hit {
  Mat mat;
  V3 point, normal;
};

struct Scene {
  std::vector<MeshO> meshes;
  std::vector<Obj> objs;
  std::vector<Light> lights;
  Cam cam;
  V3 background, ambience;
  int max_depth = 0;
  std::vector<TriRef> tris;
  std::vector<Node> nodes;
  int nodes_used = 0;
  bool built = false;
  long long counts[3] = {0, 0, 0};
  Cnt cnt[256];
  Cnt &c() { return cnt[omp_get_thread_num() & 255]; }

  // ---- Mesh::compute_normals, mymesh.cpp:103-163 ---------------------------------------------
  void compute_normals(MeshO &m) {
    const double eps = 1e-12;
    const int nt = (int)m.idx.size() / 3;
    m.vnormal.assign(m.pos.size(), v3(0, 0, 0));
    m.fnormal.resize(nt);
    for (int t = 0; t < nt; ++t) {
      const V3 &p0 = m.pos[m.idx[3 * t]], &p1 = m.pos[m.idx[3 * t + 1]], &p2 = m.pos[m.idx[3 * t + 2]];
      m.fnormal[t] = normalize(cross(p1 - p0, p2 - p0));
    }
    for (int t = 0; t < nt; ++t) {
      const int i0 = m.idx[3 * t], i1 = m.idx[3 * t + 1], i2 = m.idx[3 * t + 2];
      const V3 e0 = m.pos[i1] - m.pos[i0], e1 = m.pos[i2] - m.pos[i1], e2 = m.pos[i0] - m.pos[i2];
      const double l0 = norm(e0), l1 = norm(e1), l2 = norm(e2);
      const double d0 = dot(e0, -e2), d1 = dot(e1, -e0), d2 = dot(e2, -e1);
      const double w0 = l0 * l2 + d0, w1 = l1 * l0 + d1, w2 = l2 * l1 + d2;
      if (std::fabs(w0) > eps) m.vnormal[i0] = m.vnormal[i0] + m.fnormal[t] / w0;
      if (std::fabs(w1) > eps) m.vnormal[i1] = m.vnormal[i1] + m.fnormal[t] / w1;
      if (std::fabs(w2) > eps) m.vnormal[i2] = m.vnormal[i2] + m.fnormal[t] / w2;
    }
    for (V3 &n : m.vnormal) n = normalize(n);
  }

  // ---- BVH::init / subdivide / median / inplace_partition / updateNodeBounds, mybvh.cpp:44-362 --
  void bounds(int n) {
    Node &nd = nodes[n];
    nd.mn = v3(DBL_MAX, DBL_MAX, DBL_MAX);
    nd.mx = v3(-DBL_MAX, -DBL_MAX, -DBL_MAX);
    for (int i = nd.first; i < nd.first + nd.count; ++i) {
      const MeshO &m = meshes[tris[i].mesh];
      for (int k = 0; k < 3; ++k) {
        const V3 &p = m.pos[m.idx[3 * tris[i].local + k]];
        nd.mn = v3(std::fmin(nd.mn.x, p.x), std::fmin(nd.mn.y, p.y), std::fmin(nd.mn.z, p.z));
        nd.mx = v3(std::fmax(nd.mx.x, p.x), std::fmax(nd.mx.y, p.y), std::fmax(nd.mx.z, p.z));
      }
    }
  }
  double median(int n, int axis) {
    const Node &nd = nodes[n];
    std::vector<double> a(nd.count);
    for (int i = 0; i < nd.count; ++i) a[i] = comp(tris[nd.first + i].centroid, axis);
    std::sort(a.begin(), a.end()); // order statistics only; same values as nth_element (mybvh.cpp:346-362)
    const size_t mid = a.size() / 2;
    if (a.size() % 2 == 1) return a[mid];
    return 0.5 * (a[mid - 1] + a[mid]);
  }
  void subdivide(int n, int depth) {
    if (nodes[n].count <= 2) return;                       // mybvh.cpp:270
    const int axis = depth % 3;                            // mybvh.cpp:273
    const double split = median(n, axis);
    const int first = nodes[n].first, count = nodes[n].count;
    int i = first, j = first + count - 1;
    while (i <= j) {                                       // mybvh.cpp:309-320
      if (comp(tris[i].centroid, axis) < split) ++i;
      else { std::swap(tris[i], tris[j]); --j; }
    }
    const int left_count = i - first;
    if (left_count == 0 || left_count == count) return;    // mybvh.cpp:282
    const int l = nodes_used, r = l + 1;
    nodes_used += 2;
    nodes[l].first = first; nodes[l].count = left_count;
    nodes[r].first = i;     nodes[r].count = count - left_count;
    nodes[n].left = l;
    nodes[n].count = 0;
    bounds(l);
    bounds(r);
    subdivide(l, depth + 1);
    subdivide(r, depth + 1);
  }
  void build() {
    int base = 0;
    for (MeshO &m : meshes) {
      compute_normals(m);
      m.tri_base = base;
      base += (int)m.idx.size() / 3;
    }
    tris.clear();
    for (int mi = 0; mi < (int)meshes.size(); ++mi) {      // BVH::getData, mybvh.cpp:220-237
      const MeshO &m = meshes[mi];
      for (int t = 0; t < (int)m.idx.size() / 3; ++t) {
        TriRef r;
        r.mesh = mi; r.local = t;
        r.centroid = (m.pos[m.idx[3 * t]] + m.pos[m.idx[3 * t + 1]] + m.pos[m.idx[3 * t + 2]]) / 3.0;
        tris.push_back(r);
      }
    }
    built = false;
    nodes_used = 0;
    if (meshes.empty() || tris.empty()) return;
    nodes.assign(2 * tris.size() - 1, Node());
    nodes[0].first = 0; nodes[0].count = (int)tris.size();
    nodes_used = 1;
    bounds(0);
    subdivide(0, 1);
    built = true;
  }

  // ---- BVH::intersectAABB, mybvh.cpp:99-135 --------------------------------------------------------
  static bool aabb(const Ray &r, const V3 &mn, const V3 &mx, double &tmin_out) {
    double tmin = (mn.x - r.o.x) / r.d.x, tmax = (mx.x - r.o.x) / r.d.x;
    if (tmin > tmax) std::swap(tmin, tmax);
    double tymin = (mn.y - r.o.y) / r.d.y, tymax = (mx.y - r.o.y) / r.d.y;
    if (tymin > tymax) std::swap(tymin, tymax);
    if ((tmin > tymax) || (tymin > tmax)) return false;
    tmin = std::max(tmin, tymin);
    tmax = std::min(tmax, tymax);
    double tzmin = (mn.z - r.o.z) / r.d.z, tzmax = (mx.z - r.o.z) / r.d.z;
    if (tzmin > tzmax) std::swap(tzmin, tzmax);
    if ((tmin > tzmax) || (tzmin > tmax)) return false;
    tmax = std::min(tmax, tzmax);
    tmin_out = tmin;
    return tmax > 1e-5;
  }

  // ---- Mesh::intersect_triangle + compute_texture, mymesh.cpp:70-95,176-236 ------------------------
  bool tri(const TriRef &tr, const Ray &r, V3 &point, V3 &normal, V3 &diffuse, double &t_out,
           double *bary = nullptr) const {
    const MeshO &m = meshes[tr.mesh];
    const int i0 = m.idx[3 * tr.local], i1 = m.idx[3 * tr.local + 1], i2 = m.idx[3 * tr.local + 2];
    diffuse = m.mat.kd;
    const V3 &p0 = m.pos[i0], &p1 = m.pos[i1], &p2 = m.pos[i2];
    const V3 c1 = v3(p0.x - p2.x, p0.y - p2.y, p0.z - p2.z);
    const V3 c2 = v3(p1.x - p2.x, p1.y - p2.y, p1.z - p2.z);
    const V3 c3 = v3(-r.d.x, -r.d.y, -r.d.z);
    const V3 c4 = v3(r.o.x - p2.x, r.o.y - p2.y, r.o.z - p2.z);
    const double S = det3(c1, c2, c3);
    if (std::fabs(S) < 1e-10) return false;
    const double alpha = det3(c4, c2, c3) / S;
    const double beta = det3(c1, c4, c3) / S;
    const double gamma = (1.0 - alpha - beta);
    const double t = det3(c1, c2, c4) / S;
    if (t <= 1e-5) return false;
    const bool inside = (0.0 <= alpha && alpha <= 1.0) && (0.0 <= beta && beta <= 1.0) &&
                        (0.0 <= gamma && gamma <= 1.0);
    if (!inside) return false;
    t_out = t;
    point = at(r, t);
    if (m.texW > 0) diffuse = texel(m, tr.local, alpha, beta, gamma);
    if (m.draw_mode == 0) normal = m.fnormal[tr.local];
    else normal = alpha * m.vnormal[i0] + beta * m.vnormal[i1] + gamma * m.vnormal[i2];
    if (bary) bary[0] = alpha, bary[1] = beta, bary[2] = gamma;
    return true;
  }
  static V3 texel(const MeshO &m, int local, double alpha, double beta, double gamma) {
    const int a = m.uvidx[3 * local], b = m.uvidx[3 * local + 1], c = m.uvidx[3 * local + 2];
    double u = alpha * m.u[a] + beta * m.u[b] + gamma * m.u[c];
    double v = alpha * m.v[a] + beta * m.v[b] + gamma * m.v[c];
    u = std::clamp(u, 0.0, 1.0);
    v = std::clamp(v, 0.0, 1.0);
    const unsigned W = (unsigned)m.texW, H = (unsigned)m.texH;
    const int px = (int)std::round(u * (W - 1));
    const int py = (int)std::round((1.0 - v) * (H - 1));
    return m.tex[(size_t)py * W + px];
  }
  // ---- Plane::intersect, myplane.cpp:22-49 ; Sphere::intersect (build-defined) ---------------
  static bool object_hit(const Obj &o, const Ray &r, V3 &point, V3 &normal, V3 &diffuse, double &t_out) {
    diffuse = o.mat.kd;
    if (o.kind == 1) {
      const double cosTheta = dot(o.normal, r.d);
      if (std::abs(cosTheta) < 1e-9) return false;
      const double distance = dot(o.normal, o.center);
      const double t = (distance - dot(o.normal, r.o)) / dot(o.normal, r.d);
      if (t > 1e-5) {
        point = at(r, t); t_out = t; normal = o.normal;
        return true;
      }
      return false;
    }
    const V3 oc = r.o - o.center;
    const double a = dot(r.d, r.d);
    const double b = 2.0 * dot(r.d, oc);
    const double c = dot(oc, oc) - o.radius * o.radius;
    const double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return false;
    const double sq = std::sqrt(disc);
    const double t1 = (-b - sq) / (2.0 * a), t2 = (-b + sq) / (2.0 * a);
    double t = t1;
    if (!(t > 1e-5)) t = t2;
    if (!(t > 1e-5)) return false;
    t_out = t;
    point = at(r, t);
    normal = (point - o.center) / o.radius;
    return true;
  }
  // ---- BVH::intersectBVH, mybvh.cpp:147-210 (recursive, left first, strict '<', no distance cull) --
  void bvh_hit(const Ray &r, int n, Hit &h, double &best, int &best_id) const {
    const Node &nd = nodes[n];
    double dummy;
    if (!aabb(r, nd.mn, nd.mx, dummy)) return;
    if (nd.count > 0) {
      V3 p, nn, d;
      double t;
      for (int i = nd.first; i < nd.first + nd.count; ++i) {
        if (tri(tris[i], r, p, nn, d, t) && t < best) {
          h.mat = meshes[tris[i].mesh].mat;
          h.mat.kd = d;
          h.point = p; h.normal = nn;
          best = t; best_id = i;
        }
      }
      return;
    }
    if (aabb(r, nodes[nd.left].mn, nodes[nd.left].mx, dummy)) bvh_hit(r, nd.left, h, best, best_id);
    if (aabb(r, nodes[nd.left + 1].mn, nodes[nd.left + 1].mx, dummy)) bvh_hit(r, nd.left + 1, h, best, best_id);
  }
  // ---- intersect_scene (absent from the reference; restated as in oracle/ref_glue.cpp) -----------------
  bool nearest(const Ray &r, Hit &h, double &t_out, int &id) {
    c().scene++;
    double best = DBL_MAX;
    id = -1;
    V3 p, n, d;
    double t;
    for (size_t k = 0; k < objs.size(); ++k) {
      if (object_hit(objs[k], r, p, n, d, t) && t < best) {
        best = t; id = (int)tris.size() * (built ? 1 : 0) + (int)k;
        h.mat = objs[k].mat; h.mat.kd = d; h.point = p; h.normal = n;
      }
    }
    if (built) bvh_hit(r, 0, h, best, id);
    if (best < DBL_MAX) t_out = best;
    return best < DBL_MAX;
  }
  // ---- Raytracer::lighting / diffuse / reflection, mytracer.cpp:510-534,568-608 --------------------
  V3 lighting(const V3 &point, const V3 &normal, const V3 &view, const Mat &m) {
    const double epsilon = 1e-4;
    V3 color = v3(0.0, 0.0, 0.0);
    color.x += ambience.x * m.ka.x;
    color.y += ambience.y * m.ka.y;
    color.z += ambience.z * m.ka.z;
    for (const Light &lt : lights) {
      const V3 l = normalize(lt.pos - point);
      const double diff = std::max(0.0, dot(normal, l));
      double refl = 0.0;
      if (diff > 0.0) {
        const V3 rr = normalize(mirror(l, normal));
        refl = std::max(0.0, dot(rr, view));
      }
      refl = std::pow(refl, m.shininess);
      bool shadow = false;
      if (m.shadowable) {
        const V3 ldir = normalize(lt.pos - point);
        const double ldist = norm(lt.pos - point);
        const Ray sr = make_ray(point + epsilon * ldir, ldir);
        Hit sh;
        double st;
        int sid;
        const bool hit = nearest(sr, sh, st, sid);
        shadow = hit && st < ldist && 0.0 < st;
      }
      color.x += lt.color.x * !shadow * (m.kd.x * diff + m.ks.x * refl);
      color.y += lt.color.y * !shadow * (m.kd.y * diff + m.ks.y * refl);
      color.z += lt.color.z * !shadow * (m.kd.z * diff + m.ks.z * refl);
    }
    return color;
  }
  // ---- trace + subtrace (mytracer.cpp:546-555; trace restated after mytracer_gpu.cu:237-312) -----------
  V3 trace(const Ray &r, int depth) {
    if (depth > max_depth) return v3(0.0, 0.0, 0.0);
    if (depth > 0) c().refl++;
    Hit h;
    double t;
    int id;
    if (!nearest(r, h, t, id)) return background;
    V3 color = (1.0 - h.mat.mirror) * lighting(h.point, h.normal, -r.d, h.mat);
    if (h.mat.mirror > 0.0) {
      const V3 v = reflect(r.d, h.normal);
      const Ray rr = make_ray(h.point + 1e-4 * v, v);
      color = color + h.mat.mirror * trace(rr, depth + 1);
    } else {
      color = color + v3(0, 0, 0);
    }
    return color;
  }
  Ray primary(double x, double y) const {
    return make_ray(cam.eye, cam.ll + x * cam.xdir + y * cam.ydir - cam.eye);
  }
};

inline Scene *S(void *h) { return static_cast<Scene *>(h); }
} // namespace

extern "C" {
void *orc_new(void) { return new Scene(); }
void orc_free(void *h) { delete S(h); }
const char *orc_kind(void) { return "port"; }

int orc_add_mesh(void *h, int nv, const double *pos, int nt, const int *idx, const int *uvidx, int nuv,
                 const double *u, const double *v, int texW, int texH, const double *texels,
                 const double *material, int draw_mode) {
  MeshO m;
  m.pos.resize(nv);
  for (int i = 0; i < nv; ++i) m.pos[i] = v3p(pos + 3 * i);
  m.idx.assign(idx, idx + 3 * nt);
  if (uvidx) m.uvidx.assign(uvidx, uvidx + 3 * nt); else m.uvidx.assign(3 * nt, 0);
  if (nuv > 0 && u && v) { m.u.assign(u, u + nuv); m.v.assign(v, v + nuv); }
  if (texW > 0 && texH > 0 && texels) {
    m.texW = texW; m.texH = texH;
    m.tex.resize((size_t)texW * texH);
    for (size_t i = 0; i < m.tex.size(); ++i) m.tex[i] = v3p(texels + 3 * i);
  }
  m.mat = make_mat(material);
  m.draw_mode = draw_mode == 1 ? 1 : 0;
  S(h)->meshes.push_back(std::move(m));
  return (int)S(h)->meshes.size() - 1;
}
int orc_add_sphere(void *h, const double *center, double radius, const double *material) {
  Obj o; o.kind = 0; o.center = v3p(center); o.radius = radius; o.mat = make_mat(material);
  S(h)->objs.push_back(o);
  return (int)S(h)->objs.size() - 1;
}
int orc_add_plane(void *h, const double *center, const double *normal, const double *material) {
  Obj o; o.kind = 1; o.center = v3p(center); o.normal = v3p(normal); o.radius = 0; o.mat = make_mat(material);
  S(h)->objs.push_back(o);
  return (int)S(h)->objs.size() - 1;
}
int orc_add_light(void *h, const double *pos, const double *color) {
  Light l; l.pos = v3p(pos); l.color = v3p(color);
  S(h)->lights.push_back(l);
  return (int)S(h)->lights.size() - 1;
}
void orc_set_camera(void *h, const double *eye, const double *center, const double *up, double fovy,
                    int width, int height) {
  Cam &c = S(h)->cam;
  const V3 e = v3p(eye), ce = v3p(center), u = v3p(up);
  const V3 view = normalize(ce - e);
  const double dist = norm(ce - e);
  const double img_h = 2.0 * dist * std::tan(0.5 * fovy / 180.0 * M_PI);
  const double img_w = (double(width) / double(height)) * img_h;
  const V3 right = normalize(cross(view, u));
  const V3 new_up = normalize(cross(right, view));
  c.eye = e;
  c.xdir = (img_w / double(width)) * right;
  c.ydir = (img_h / double(height)) * new_up;
  c.ll = ce - (0.5 * double(width)) * c.xdir - (0.5 * double(height)) * c.ydir;
  c.w = width; c.h = height;
}
void orc_get_camera(void *h, double *o) {
  const Cam &c = S(h)->cam;
  const V3 *v[4] = {&c.eye, &c.xdir, &c.ydir, &c.ll};
  for (int k = 0; k < 4; ++k) o[3 * k] = v[k]->x, o[3 * k + 1] = v[k]->y, o[3 * k + 2] = v[k]->z;
}
void orc_set_globals(void *h, const double *background, const double *ambience, int max_depth) {
  S(h)->background = v3p(background); S(h)->ambience = v3p(ambience); S(h)->max_depth = max_depth;
}
void orc_build(void *h) { S(h)->build(); }
int orc_num_triangles(void *h) { return S(h)->built ? (int)S(h)->tris.size() : 0; }
int orc_num_nodes(void *h) { return S(h)->built ? S(h)->nodes_used : 0; }
void orc_get_bvh(void *h, double *bbmin, double *bbmax, int *left, int *first, int *count) {
  const Scene *s = S(h);
  for (int i = 0; i < s->nodes_used; ++i) {
    const Node &n = s->nodes[i];
    bbmin[3 * i] = n.mn.x, bbmin[3 * i + 1] = n.mn.y, bbmin[3 * i + 2] = n.mn.z;
    bbmax[3 * i] = n.mx.x, bbmax[3 * i + 1] = n.mx.y, bbmax[3 * i + 2] = n.mx.z;
    left[i] = n.left, first[i] = n.first, count[i] = n.count;
  }
}
void orc_get_tri_order(void *h, int *perm) {
  const Scene *s = S(h);
  for (size_t i = 0; i < s->tris.size(); ++i) perm[i] = s->meshes[s->tris[i].mesh].tri_base + s->tris[i].local;
}
void orc_get_normals(void *h, int mesh, double *fn, double *vn) {
  const MeshO &m = S(h)->meshes.at(mesh);
  for (size_t i = 0; i < m.fnormal.size(); ++i) fn[3 * i] = m.fnormal[i].x, fn[3 * i + 1] = m.fnormal[i].y, fn[3 * i + 2] = m.fnormal[i].z;
  for (size_t i = 0; i < m.vnormal.size(); ++i) vn[3 * i] = m.vnormal[i].x, vn[3 * i + 1] = m.vnormal[i].y, vn[3 * i + 2] = m.vnormal[i].z;
}

double orc_render(void *h, double *rgb, int *ids, double *tout, int y0, int y1, int ystep, int spp_grid,
                  int nthreads) {
  Scene *s = S(h);
  const int W = s->cam.w, H = s->cam.h;
  if (y1 > H) y1 = H;
  if (ystep < 1) ystep = 1;
  if (nthreads > 0) omp_set_num_threads(nthreads);
  for (auto &c : s->cnt) c = Cnt();
  const int nrows = (y1 - y0 + ystep - 1) / ystep;
  const int g = spp_grid < 1 ? 1 : spp_grid;
  const auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1)
  for (int r = 0; r < nrows; ++r) {
    const int y = y0 + r * ystep;
    for (int x = 0; x < W; ++x) {
      V3 color = v3(0, 0, 0);
      if (g == 1) {
        color = s->trace(s->primary(x, y), 0);
      } else {
        for (int si = 0; si < g; si++) {
          const double xo = (si) / static_cast<double>(g) - 0.5 + 1.0 / (2.0 * g);
          for (int sj = 0; sj < g; sj++) {
            const double yo = (sj) / static_cast<double>(g) - 0.5 + 1.0 / (2.0 * g);
            color = color + s->trace(s->primary(double(x) + xo, double(y) + yo), 0);
          }
        }
        color = color / double(g * g);
      }
      const size_t p = (size_t)y * W + x;
      rgb[3 * p] = std::min(color.x, 1.0), rgb[3 * p + 1] = std::min(color.y, 1.0), rgb[3 * p + 2] = std::min(color.z, 1.0);
    }
  }
  const auto t1 = std::chrono::steady_clock::now();
  long long scene = 0, refl = 0;
  for (auto &c : s->cnt) scene += c.scene, refl += c.refl;
  const long long prim = (long long)nrows * W * g * g;
  s->counts[0] = prim, s->counts[1] = scene - prim - refl, s->counts[2] = refl;
  if (ids || tout) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int r = 0; r < nrows; ++r) {
      const int y = y0 + r * ystep;
      for (int x = 0; x < W; ++x) {
        Hit hh; double t = -1.0; int id;
        const bool hit = s->nearest(s->primary(x, y), hh, t, id);
        const size_t p = (size_t)y * W + x;
        if (ids) ids[p] = hit ? id : -1;
        if (tout) tout[p] = hit ? t : -1.0;
      }
    }
  }
  return std::chrono::duration<double>(t1 - t0).count();
}
void orc_ray_counts(void *h, long long *o) { for (int k = 0; k < 3; ++k) o[k] = S(h)->counts[k]; }

void orc_trace_rays(void *h, int n, const double *origins, const double *dirs, double *rgb, int *ids, double *tout) {
  Scene *s = S(h);
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < n; ++i) {
    const Ray r = make_ray(v3p(origins + 3 * i), v3p(dirs + 3 * i));
    if (rgb) { const V3 c = s->trace(r, 0); rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z; }
    if (ids || tout) {
      Hit hh; double t = -1.0; int id;
      const bool hit = s->nearest(r, hh, t, id);
      if (ids) ids[i] = hit ? id : -1;
      if (tout) tout[i] = hit ? t : -1.0;
    }
  }
}

int orc_tri_test(const double *p0, const double *p1, const double *p2, const double *o, const double *d, double *out4) {
  Scene s;
  MeshO m;
  m.pos = {v3p(p0), v3p(p1), v3p(p2)};
  m.idx = {0, 1, 2}; m.uvidx = {0, 0, 0};
  m.fnormal = {v3(0, 0, 0)};
  s.meshes.push_back(m);
  TriRef tr; tr.mesh = 0; tr.local = 0;
  Ray r; r.o = v3p(o); r.d = v3p(d);
  V3 p, n, dif; double t = 0;
  const bool hit = s.tri(tr, r, p, n, dif, t);
  if (hit) out4[0] = t, out4[1] = p.x, out4[2] = p.y, out4[3] = p.z;
  return hit;
}
int orc_aabb_test(const double *bbmin, const double *bbmax, const double *o, const double *d, double *tmin_out) {
  Ray r; r.o = v3p(o); r.d = v3p(d);
  double tmin = 0.0;
  const bool hit = Scene::aabb(r, v3p(bbmin), v3p(bbmax), tmin);
  if (tmin_out) *tmin_out = tmin;
  return hit;
}
int orc_plane_test(const double *center, const double *normal, const double *o, const double *d, double *out4) {
  Obj ob; ob.kind = 1; ob.center = v3p(center); ob.normal = v3p(normal); ob.radius = 0;
  Ray r; r.o = v3p(o); r.d = v3p(d);
  V3 p, n, dif; double t = 0;
  const bool hit = Scene::object_hit(ob, r, p, n, dif, t);
  if (hit) out4[0] = t, out4[1] = p.x, out4[2] = p.y, out4[3] = p.z;
  return hit;
}
int orc_sphere_test(const double *center, double radius, const double *o, const double *d, double *out7) {
  Obj ob; ob.kind = 0; ob.center = v3p(center); ob.radius = radius;
  Ray r; r.o = v3p(o); r.d = v3p(d);
  V3 p, n, dif; double t = 0;
  const bool hit = Scene::object_hit(ob, r, p, n, dif, t);
  if (hit) { out7[0] = t; out7[1] = p.x; out7[2] = p.y; out7[3] = p.z; out7[4] = n.x; out7[5] = n.y; out7[6] = n.z; }
  return hit;
}
void orc_lighting(void *h, const double *point, const double *normal, const double *view, const double *material, double *rgb) {
  const V3 c = S(h)->lighting(v3p(point), v3p(normal), v3p(view), make_mat(material));
  rgb[0] = c.x, rgb[1] = c.y, rgb[2] = c.z;
}
void orc_texture(void *h, int tri, double alpha, double beta, double gamma, double *rgb) {
  const Scene *s = S(h);
  const TriRef &tr = s->tris.at(tri);
  const MeshO &m = s->meshes[tr.mesh];
  V3 c = m.mat.kd;
  if (m.texW > 0) c = Scene::texel(m, tr.local, alpha, beta, gamma);
  rgb[0] = c.x, rgb[1] = c.y, rgb[2] = c.z;
}
} // extern "C"
