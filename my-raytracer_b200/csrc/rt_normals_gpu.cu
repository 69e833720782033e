// rt_normals_gpu.cu — Mesh::compute_normals (mymesh.cpp:103-163) on the GPU, bit-identical to the host
// mirror (rt_prep.cpp rt_compute_normals) and therefore to the reference.
//
// The reference accumulates angle-weighted face normals into the vertices in triangle order; FP64 addition
// is not associative, so a bit-identical result needs the same order per vertex.  Three passes:
//   1. one thread per triangle: face normal (normalize(cross(p1 - p0, p2 - p0))) and the three corner
//      denominators ||u|| ||v|| + u.v  (mymesh.cpp:112-146);
//   2. the 3*nt corners are sorted by vertex id with a STABLE radix sort (cub): each vertex's corners end up
//      contiguous and in ascending (triangle, corner) order — the reference's accumulation order;
//   3. one thread per vertex: adds its corners' n / w in that order (skipping |w| <= 1e-12, mymesh.cpp:149-157)
//      and normalises (a vertex no triangle uses gets 0/0 = NaN, as on the host).
// HBM-bound and tiny next to the render path; what matters is that PHONG scenes can be prepared without a
// host pass over the mesh.
#include "../../include/rt_b200.h"
#include "rt_math.cuh"
#include <cub/cub.cuh>
#include <string>

namespace rtb {
void set_error(const std::string &msg);

namespace {

#define NM_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      rtb::set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                          \
      return (e_ == cudaErrorMemoryAllocation) ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA;               \
    }                                                                                              \
  } while (0)

constexpr int kThreads = 256;
inline int blocks_for(long long n) { return (int)((n + kThreads - 1) / kThreads); }

__device__ __forceinline__ D3 ldp(const double *pos, int i) { return mk(pos[4 * (size_t)i], pos[4 * (size_t)i + 1], pos[4 * (size_t)i + 2]); }

__global__ void k_faces(const double *__restrict__ pos, const int *__restrict__ idx, int nv, int nt,
                        double *__restrict__ fn, double *__restrict__ w, int *__restrict__ key, int *__restrict__ val,
                        int *__restrict__ bad) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nt) return;
  const int i0 = idx[3 * t], i1 = idx[3 * t + 1], i2 = idx[3 * t + 2];
  if (i0 < 0 || i1 < 0 || i2 < 0 || i0 >= nv || i1 >= nv || i2 >= nv) {
    atomicExch(bad, 1);
    for (int c = 0; c < 3; ++c) key[3 * t + c] = nv, val[3 * t + c] = 3 * t + c; // parked behind every vertex
    return;
  }
  const D3 p0 = ldp(pos, i0), p1 = ldp(pos, i1), p2 = ldp(pos, i2);
  const D3 a = sub(p1, p0), b = sub(p2, p0);
  const D3 cr = mk(dsub(dmul(a.y, b.z), dmul(a.z, b.y)), dsub(dmul(a.z, b.x), dmul(a.x, b.z)),
                   dsub(dmul(a.x, b.y), dmul(a.y, b.x)));
  const D3 n = normalize(cr);
  fn[4 * (size_t)t] = n.x; fn[4 * (size_t)t + 1] = n.y; fn[4 * (size_t)t + 2] = n.z; fn[4 * (size_t)t + 3] = 0.0;
  const D3 v0 = a, v1 = sub(p2, p1), v2 = sub(p0, p2);
  const double l0 = norm(v0), l1 = norm(v1), l2 = norm(v2);
  const double d0 = dot(v0, neg(v2)), d1 = dot(v1, neg(v0)), d2 = dot(v2, neg(v1));
  w[3 * (size_t)t] = dadd(dmul(l0, l2), d0);
  w[3 * (size_t)t + 1] = dadd(dmul(l1, l0), d1);
  w[3 * (size_t)t + 2] = dadd(dmul(l2, l1), d2);
  key[3 * t] = i0; key[3 * t + 1] = i1; key[3 * t + 2] = i2;
  val[3 * t] = 3 * t; val[3 * t + 1] = 3 * t + 1; val[3 * t + 2] = 3 * t + 2;
}

// first[v] = first sorted position whose key is v (or -1)
__global__ void k_first(const int *__restrict__ key, int n, int nv, int *__restrict__ first) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const int k = key[p];
  if (k < nv && (p == 0 || key[p - 1] != k)) first[k] = p;
}

__global__ void k_vertices(const int *__restrict__ key, const int *__restrict__ val, int n, const int *__restrict__ first,
                           const double *__restrict__ fn, const double *__restrict__ w, int nv, double *__restrict__ vn) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nv) return;
  D3 acc = mk(0.0, 0.0, 0.0);
  const int f = first[v];
  if (f >= 0) {
    for (int p = f; p < n && key[p] == v; ++p) {
      const int corner = val[p];
      const double wd = w[corner];
      if (fabs(wd) > 1e-12) {
        const int t = corner / 3;
        const D3 nrm = mk(fn[4 * (size_t)t], fn[4 * (size_t)t + 1], fn[4 * (size_t)t + 2]);
        acc = add(acc, divs(nrm, wd));
      }
    }
  }
  const D3 r = normalize(acc);
  vn[4 * (size_t)v] = r.x; vn[4 * (size_t)v + 1] = r.y; vn[4 * (size_t)v + 2] = r.z; vn[4 * (size_t)v + 3] = 0.0;
}

struct Bufs {
  void *p[12] = {};
  int n = 0;
  template <typename T> cudaError_t alloc(T *&out, size_t count) {
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, count * sizeof(T) + 16);
    if (e == cudaSuccess) { p[n++] = q; out = (T *)q; }
    return e;
  }
  ~Bufs() { for (int i = 0; i < n; ++i) cudaFree(p[i]); }
};

} // namespace
} // namespace rtb

using namespace rtb;

extern "C" int rt_compute_normals_device(const rt_vec4 *pos, int nv, const int *idx, int nt, rt_vec4 *fnormals,
                                         rt_vec4 *vnormals, double *device_ms) {
  if (!pos || !idx || !fnormals || !vnormals || nv < 0 || nt < 0) {
    rtb::set_error("rt_compute_normals_device: invalid argument");
    return RT_ERR_INVALID_ARG;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    rtb::set_error("rt_compute_normals_device: no CUDA device (use rt_compute_normals on the host)");
    return RT_ERR_NO_DEVICE;
  }
  if (device_ms) *device_ms = 0.0;
  if (nv == 0 && nt == 0) return RT_OK;
  Bufs B;
  double *d_pos = nullptr, *d_fn = nullptr, *d_w = nullptr, *d_vn = nullptr;
  int *d_idx = nullptr, *d_key = nullptr, *d_val = nullptr, *d_key2 = nullptr, *d_val2 = nullptr, *d_first = nullptr, *d_bad = nullptr;
  const int nc = 3 * nt;
  NM_CUDA(B.alloc(d_pos, 4 * (size_t)nv)); NM_CUDA(B.alloc(d_idx, (size_t)nc));
  NM_CUDA(B.alloc(d_fn, 4 * (size_t)nt)); NM_CUDA(B.alloc(d_w, (size_t)nc)); NM_CUDA(B.alloc(d_vn, 4 * (size_t)nv));
  NM_CUDA(B.alloc(d_key, (size_t)nc)); NM_CUDA(B.alloc(d_val, (size_t)nc));
  NM_CUDA(B.alloc(d_key2, (size_t)nc)); NM_CUDA(B.alloc(d_val2, (size_t)nc));
  NM_CUDA(B.alloc(d_first, (size_t)nv)); NM_CUDA(B.alloc(d_bad, 1));
  NM_CUDA(cudaMemcpy(d_pos, pos, sizeof(rt_vec4) * (size_t)nv, cudaMemcpyDefault));
  NM_CUDA(cudaMemcpy(d_idx, idx, sizeof(int) * (size_t)nc, cudaMemcpyDefault));
  NM_CUDA(cudaMemset(d_bad, 0, sizeof(int)));
  NM_CUDA(cudaMemset(d_first, 0xff, sizeof(int) * (size_t)nv));
  size_t tmp_bytes = 0;
  int bits = 1;
  while ((1ll << bits) <= (long long)nv) ++bits; // keys are 0..nv
  NM_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_key, d_key2, d_val, d_val2, nc, 0, bits));
  unsigned char *d_tmp = nullptr;
  NM_CUDA(B.alloc(d_tmp, tmp_bytes));
  cudaEvent_t e0, e1;
  NM_CUDA(cudaEventCreate(&e0)); NM_CUDA(cudaEventCreate(&e1));
  NM_CUDA(cudaEventRecord(e0));
  if (nt > 0) {
    k_faces<<<blocks_for(nt), kThreads>>>(d_pos, d_idx, nv, nt, d_fn, d_w, d_key, d_val, d_bad);
    NM_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, d_key, d_key2, d_val, d_val2, nc, 0, bits));
    k_first<<<blocks_for(nc), kThreads>>>(d_key2, nc, nv, d_first);
  }
  if (nv > 0) k_vertices<<<blocks_for(nv), kThreads>>>(d_key2, d_val2, nc, d_first, d_fn, d_w, nv, d_vn);
  NM_CUDA(cudaEventRecord(e1));
  NM_CUDA(cudaEventSynchronize(e1));
  NM_CUDA(cudaGetLastError());
  float ms = 0.0f;
  NM_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  int bad = 0;
  NM_CUDA(cudaMemcpy(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost));
  if (bad) {
    rtb::set_error("rt_compute_normals_device: vertex index out of range");
    return RT_ERR_BAD_SCENE;
  }
  NM_CUDA(cudaMemcpy(fnormals, d_fn, sizeof(rt_vec4) * (size_t)nt, cudaMemcpyDefault));
  NM_CUDA(cudaMemcpy(vnormals, d_vn, sizeof(rt_vec4) * (size_t)nv, cudaMemcpyDefault));
  if (device_ms) *device_ms = ms;
  return RT_OK;
}
