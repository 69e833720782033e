// rt_prep.cpp — host-side scene preparation on the caller side of the hot path: produces exactly
// what the reference's own host code hands to the kernel.
//   rt_bvh_build_median  == BVH::initSoA + subdivideSoA/medianSoA/inplace_partitionSoA/
//                           updateNodeBoundsSoA (mybvh.cpp:375-539): same topology, same
//                           node numbering, same triangle-row permutation
//   rt_compute_normals   == Mesh::compute_normals (mymesh.cpp:103-163)
//   rt_camera_look_at    pinhole camera (build-defined; DESIGN.md "Conventions")
// Compiled with -ffp-contract=off: results are bit-identical to the reference's.
#include "../../include/rt_b200.h"
#include "rt_math.cuh"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <string>
#include <vector>

namespace rtb {
void set_error(const std::string &msg); // rt_api.cpp
}

namespace {
using rtb::D3;

struct Builder {
  rt_data_soa *data;
  rt_bvh_nodes_soa *nodes;
  int nodes_used = 1;
  std::vector<int> perm;     // perm[pos] = original triangle row now at BVH position pos
  std::vector<double> cent;  // centroid of ORIGINAL row r: cent[3*r + axis]
  std::vector<double> scratch;

  // (v0 + v1 + v2) / 3.0, mybvh.cpp:493,535
  void centroids(int N) {
    cent.resize(3 * (size_t)N);
    for (int r = 0; r < N; ++r) {
      const rt_vec4 &a = data->vertexPos_[data->vertexIdx_[3 * r]];
      const rt_vec4 &b = data->vertexPos_[data->vertexIdx_[3 * r + 1]];
      const rt_vec4 &c = data->vertexPos_[data->vertexIdx_[3 * r + 2]];
      cent[3 * (size_t)r + 0] = ((a.x + b.x) + c.x) / 3.0;
      cent[3 * (size_t)r + 1] = ((a.y + b.y) + c.y) / 3.0;
      cent[3 * (size_t)r + 2] = ((a.z + b.z) + c.z) / 3.0;
    }
  }
  // mybvh.cpp:412-431
  void bounds(int n) {
    double mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, mx[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    const int f = nodes->firstTriIdx_[n], c = nodes->triCount_[n];
    for (int i = f; i < f + c; ++i) {
      const int r = perm[i];
      for (int k = 0; k < 3; ++k) {
        const rt_vec4 &v = data->vertexPos_[data->vertexIdx_[3 * r + k]];
        mn[0] = std::fmin(mn[0], v.x); mn[1] = std::fmin(mn[1], v.y); mn[2] = std::fmin(mn[2], v.z);
        mx[0] = std::fmax(mx[0], v.x); mx[1] = std::fmax(mx[1], v.y); mx[2] = std::fmax(mx[2], v.z);
      }
    }
    nodes->bb_min_[n] = rt_vec4{mn[0], mn[1], mn[2], 0.0};
    nodes->bb_max_[n] = rt_vec4{mx[0], mx[1], mx[2], 0.0};
  }
  // mybvh.cpp:346-362,521-539: median of the centroid coordinates (mean of the two middle
  // values for an even count)
  double median(int n, int axis) {
    const int f = nodes->firstTriIdx_[n], c = nodes->triCount_[n];
    scratch.resize(c);
    for (int i = 0; i < c; ++i) scratch[i] = cent[3 * (size_t)perm[f + i] + axis];
    const size_t mid = (size_t)c / 2;
    std::nth_element(scratch.begin(), scratch.begin() + mid, scratch.end());
    if (c % 2 == 1) return scratch[mid];
    const double hi = scratch[mid];
    const double lo = *std::max_element(scratch.begin(), scratch.begin() + mid);
    return 0.5 * (lo + hi);
  }
  // mybvh.cpp:438-472
  void subdivide(int n, int depth) {
    const int count = nodes->triCount_[n];
    if (count <= 2) return;
    const int axis = depth % 3;
    const double split = median(n, axis);
    const int first = nodes->firstTriIdx_[n];
    int i = first, j = first + count - 1;
    while (i <= j) { // two-pointer partition, mybvh.cpp:481-513
      if (cent[3 * (size_t)perm[i] + axis] < split) {
        i++;
      } else {
        std::swap(perm[i], perm[j]);
        j--;
      }
    }
    const int left_count = i - first;
    if (left_count == 0 || left_count == count) return;
    const int l = nodes_used, r = l + 1;
    nodes_used += 2;
    nodes->firstTriIdx_[l] = first;  nodes->triCount_[l] = left_count;
    nodes->firstTriIdx_[r] = i;      nodes->triCount_[r] = count - left_count;
    nodes->leftChildIdx_[n] = l;
    nodes->triCount_[n] = 0;
    bounds(l);
    bounds(r);
    subdivide(l, depth + 1);
    subdivide(r, depth + 1);
  }
};

template <typename T> void gather_rows(T *arr, const std::vector<int> &perm, int width) {
  if (!arr) return;
  std::vector<T> tmp(arr, arr + (size_t)perm.size() * width);
  for (size_t pos = 0; pos < perm.size(); ++pos)
    for (int k = 0; k < width; ++k) arr[pos * width + k] = tmp[(size_t)perm[pos] * width + k];
}
} // namespace

extern "C" int rt_bvh_build_median(rt_data_soa *data, rt_bvh_nodes_soa *nodes, int *nodes_used) {
  if (!data || !nodes || !nodes_used) { rtb::set_error("rt_bvh_build_median: null argument"); return RT_ERR_INVALID_ARG; }
  const int N = data->tVertexIdxCount_ / 3;
  if (N <= 0) { *nodes_used = 0; return RT_OK; } // "No mesh available", mybvh.cpp:402-404
  if (!data->vertexIdx_ || !data->vertexPos_ || !nodes->bb_min_ || !nodes->bb_max_ ||
      !nodes->leftChildIdx_ || !nodes->firstTriIdx_ || !nodes->triCount_) {
    rtb::set_error("rt_bvh_build_median: null array");
    return RT_ERR_INVALID_ARG;
  }
  for (int i = 0; i < 3 * N; ++i)
    if (data->vertexIdx_[i] < 0 || data->vertexIdx_[i] >= data->tVertexCount_) {
      rtb::set_error("rt_bvh_build_median: vertex index out of range");
      return RT_ERR_BAD_SCENE;
    }
  Builder b;
  b.data = data;
  b.nodes = nodes;
  b.perm.resize(N);
  for (int i = 0; i < N; ++i) b.perm[i] = i;
  b.centroids(N);
  for (int i = 0; i < 2 * N - 1; ++i) {
    nodes->bb_min_[i] = rt_vec4{0, 0, 0, 0};
    nodes->bb_max_[i] = rt_vec4{0, 0, 0, 0};
    nodes->leftChildIdx_[i] = 0; nodes->firstTriIdx_[i] = 0; nodes->triCount_[i] = 0;
  }
  nodes->leftChildIdx_[0] = 0; nodes->firstTriIdx_[0] = 0; nodes->triCount_[0] = N; // mybvh.cpp:393-396
  b.bounds(0);
  b.subdivide(0, 1);
  // apply the row permutation the reference performs swap by swap (mybvh.cpp:497-503)
  gather_rows(data->normals_, b.perm, 1);
  gather_rows(data->vertexIdx_, b.perm, 3);
  gather_rows(data->textureIdx_, b.perm, 3);
  *nodes_used = b.nodes_used;
  return RT_OK;
}

extern "C" int rt_compute_normals(const rt_vec4 *pos, int nv, const int *idx, int nt, rt_vec4 *fnormals,
                                  rt_vec4 *vnormals) {
  using namespace rtb;
  if (!pos || !idx || !fnormals || !vnormals || nv < 0 || nt < 0) {
    set_error("rt_compute_normals: invalid argument");
    return RT_ERR_INVALID_ARG;
  }
  const double eps = 1e-12;
  std::vector<D3> vn(nv, mk(0, 0, 0));
  auto P = [&](int i) { return mk(pos[i].x, pos[i].y, pos[i].z); };
  for (int t = 0; t < nt; ++t) {
    const int i0 = idx[3 * t], i1 = idx[3 * t + 1], i2 = idx[3 * t + 2];
    if (i0 < 0 || i1 < 0 || i2 < 0 || i0 >= nv || i1 >= nv || i2 >= nv) {
      set_error("rt_compute_normals: vertex index out of range");
      return RT_ERR_BAD_SCENE;
    }
    const D3 p0 = P(i0), p1 = P(i1), p2 = P(i2);
    // face normal, mymesh.cpp:112-117 ; cross as a.y*b.z - a.z*b.y, ...
    const D3 a = sub(p1, p0), b = sub(p2, p0);
    const D3 cr = mk(dsub(dmul(a.y, b.z), dmul(a.z, b.y)), dsub(dmul(a.z, b.x), dmul(a.x, b.z)),
                     dsub(dmul(a.x, b.y), dmul(a.y, b.x)));
    const D3 n = normalize(cr);
    fnormals[t] = rt_vec4{n.x, n.y, n.z, 0.0};
  }
  for (int t = 0; t < nt; ++t) { // angle-weighted accumulation, mymesh.cpp:120-157
    const int i0 = idx[3 * t], i1 = idx[3 * t + 1], i2 = idx[3 * t + 2];
    const D3 p0 = P(i0), p1 = P(i1), p2 = P(i2);
    const D3 n = mk(fnormals[t].x, fnormals[t].y, fnormals[t].z);
    const D3 v0 = sub(p1, p0), v1 = sub(p2, p1), v2 = sub(p0, p2);
    const double l0 = norm(v0), l1 = norm(v1), l2 = norm(v2);
    const double d0 = dot(v0, neg(v2)), d1 = dot(v1, neg(v0)), d2 = dot(v2, neg(v1));
    const double w0 = dadd(dmul(l0, l2), d0), w1 = dadd(dmul(l1, l0), d1), w2 = dadd(dmul(l2, l1), d2);
    if (std::fabs(w0) > eps) vn[i0] = add(vn[i0], divs(n, w0));
    if (std::fabs(w1) > eps) vn[i1] = add(vn[i1], divs(n, w1));
    if (std::fabs(w2) > eps) vn[i2] = add(vn[i2], divs(n, w2));
  }
  for (int i = 0; i < nv; ++i) {
    const D3 n = normalize(vn[i]);
    vnormals[i] = rt_vec4{n.x, n.y, n.z, 0.0};
  }
  return RT_OK;
}

extern "C" int rt_camera_look_at(const double eye[3], const double center[3], const double up[3],
                                 double fovy_deg, int width, int height, rt_camera *out) {
  using namespace rtb;
  if (!eye || !center || !up || !out || width <= 0 || height <= 0) {
    set_error("rt_camera_look_at: invalid argument");
    return RT_ERR_INVALID_ARG;
  }
  const D3 e = ld3(eye), c = ld3(center), u = ld3(up);
  const D3 view = normalize(sub(c, e));
  const double dist = norm(sub(c, e));
  const double img_h = dmul(dmul(2.0, dist), std::tan(dmul(ddiv(dmul(0.5, fovy_deg), 180.0), M_PI)));
  const double img_w = dmul(ddiv((double)width, (double)height), img_h);
  auto cross = [](const D3 &a, const D3 &b) {
    return mk(dsub(dmul(a.y, b.z), dmul(a.z, b.y)), dsub(dmul(a.z, b.x), dmul(a.x, b.z)),
              dsub(dmul(a.x, b.y), dmul(a.y, b.x)));
  };
  const D3 right = normalize(cross(view, u));
  const D3 new_up = normalize(cross(right, view));
  const D3 xd = scale(ddiv(img_w, (double)width), right);
  const D3 yd = scale(ddiv(img_h, (double)height), new_up);
  const D3 ll = sub(sub(c, scale(dmul(0.5, (double)width), xd)), scale(dmul(0.5, (double)height), yd));
  out->eye = rt_vec4{e.x, e.y, e.z, 0.0};
  out->x_dir = rt_vec4{xd.x, xd.y, xd.z, 0.0};
  out->y_dir = rt_vec4{yd.x, yd.y, yd.z, 0.0};
  out->lower_left = rt_vec4{ll.x, ll.y, ll.z, 0.0};
  out->width = width;
  out->height = height;
  return RT_OK;
}
