// rt_bvh_gpu.cu — median-split BVH build on the GPU with the REFERENCE topology.
//
// Replaces the host build BVH::initSoA / subdivideSoA / medianSoA / inplace_partitionSoA /
// updateNodeBoundsSoA (mybvh.cpp:375-539; host mirror: rt_prep.cpp rt_bvh_build_median).  The result is
// bit-identical to the reference's: same split values, same node numbering, same triangle-row permutation
// (triangle index == BVH order is what the primitive ids of the render path are made of).
//
// The reference recursion becomes level-synchronous:
//   * every node of one depth is a contiguous range of the triangle order; one pass handles all of them;
//   * median of the centroid coordinate (axis = depth % 3, root depth 1): one segmented sort of the keys per
//     level; even counts take 0.5 * (lo + hi) of the two middle values (median_inplace, mybvh.cpp:346-362);
//   * the in-place two-pointer partition (mybvh.cpp:481-513) is sequential in the reference, but the
//     permutation it produces has a closed form (derivation in partition_dest below), so every triangle
//     computes its destination independently;
//   * nodes are created in breadth-first order and renumbered afterwards to the order the recursion
//     allocates them in (children in adjacent pairs at nodesUsed_, left subtree before right subtree,
//     mybvh.cpp:455-471): child index = base(parent), base(left) = base(parent) + 2,
//     base(right) = base(parent) + 2 + descendants(left).
#include "../../include/rt_b200.h"
#include <cfloat>
#include <cstdio>
#include <cstring>
#include <cub/cub.cuh>
#include <string>
#include <vector>

namespace rtb {
void set_error(const std::string &msg);
}

namespace {

#define BV_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      rtb::set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                          \
      return (e_ == cudaErrorMemoryAllocation) ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA;               \
    }                                                                                              \
  } while (0)

constexpr int kThreads = 256;
inline int blocks_for(long long n) { return (int)((n + kThreads - 1) / kThreads); }

// centroid of ORIGINAL row r: ((v0 + v1) + v2) / 3.0 per component (mybvh.cpp:493,535), and the
// triangle's own box (min/max of its three vertices)
__global__ void k_centroids(const double *__restrict__ vpos, const int *__restrict__ vidx, int n, double *cx, double *cy,
                            double *cz, double *tmin, double *tmax) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const double *a = vpos + 4 * (size_t)vidx[3 * r], *b = vpos + 4 * (size_t)vidx[3 * r + 1], *c = vpos + 4 * (size_t)vidx[3 * r + 2];
  cx[r] = __ddiv_rn(__dadd_rn(__dadd_rn(a[0], b[0]), c[0]), 3.0);
  cy[r] = __ddiv_rn(__dadd_rn(__dadd_rn(a[1], b[1]), c[1]), 3.0);
  cz[r] = __ddiv_rn(__dadd_rn(__dadd_rn(a[2], b[2]), c[2]), 3.0);
  for (int k = 0; k < 3; ++k) {
    tmin[3 * (size_t)r + k] = fmin(fmin(a[k], b[k]), c[k]);
    tmax[3 * (size_t)r + k] = fmax(fmax(a[k], b[k]), c[k]);
  }
}

__global__ void k_iota(int *p, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = i;
}

// position -> active segment of this level (binary search over the sorted segment starts), -1 if none
__global__ void k_segof(const int *__restrict__ seg_first, const int *__restrict__ seg_count, int nseg, int n, int *segof) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int lo = 0, hi = nseg; // last segment with first <= i
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (seg_first[mid] <= i) lo = mid + 1; else hi = mid;
  }
  const int s = lo - 1;
  segof[i] = (s >= 0 && i < seg_first[s] + seg_count[s]) ? s : -1;
}

__global__ void k_keys(const double *__restrict__ cent, const int *__restrict__ perm, int n, double *keys) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) keys[i] = cent[perm[i]];
}

// split value per segment from its sorted keys (median_inplace, mybvh.cpp:346-362)
__global__ void k_split(const double *__restrict__ sorted, const int *__restrict__ seg_first, const int *__restrict__ seg_count,
                        int nseg, double *split) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  const int f = seg_first[s], c = seg_count[s], mid = c / 2;
  if (c & 1) split[s] = sorted[f + mid];
  else split[s] = __dmul_rn(0.5, __dadd_rn(sorted[f + mid - 1], sorted[f + mid]));
}

__global__ void k_flags(const double *__restrict__ keys, const int *__restrict__ segof, const double *__restrict__ split, int n,
                        int *flag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = segof[i];
  flag[i] = (s >= 0 && keys[i] < split[s]) ? 1 : 0; // centroid[axis] < splitPos, mybvh.cpp:494
}

// ---- the reference's in-place partition, in closed form ------------------------------------------
// inplace_partitionSoA scans i upwards; an element that is not "left" (L) is swapped with a[j], j--.
// Every element is examined exactly once, either where it stands (front stream, ascending index) or
// after being pulled from the back (back stream, descending index).  With nL = #L, n = count:
//   * positions < nL: an L element stays; the k-th non-L ("hole", ascending) is replaced by the k-th L
//     element counted from the back of [nL, n);
//   * the k-th hole element lands just in front of the (k-1)-th back-L's old place: Z_{k-1} - 1 (Z_0 = n);
//   * a non-L element at x >= nL moves one place towards the front (x - 1), except a[nL] itself, which acts
//     as one more front element and lands at Z_K - 1 (K = number of holes).
// pass 1 records, per segment, the local index of the k-th hole (harr) and of the j-th back-L (zarr).
__global__ void k_roles(const int *__restrict__ flag, const int *__restrict__ pref, const int *__restrict__ segof,
                        const int *__restrict__ seg_first, const int *__restrict__ seg_count, int n, int *harr, int *zarr) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = segof[i];
  if (s < 0) return;
  const int f = seg_first[s], c = seg_count[s];
  const int nL = pref[f + c] - pref[f];
  const int x = i - f, lbefore = pref[i] - pref[f]; // # L in [0, x)
  if (x < nL) {
    if (!flag[i]) harr[f + (x - lbefore)] = x;             // k-th hole, k - 1 = # non-L in [0, x)
  } else if (flag[i]) {
    zarr[f + (nL - lbefore - 1)] = x;                      // j-th L from the back, j = # L in [x, n)
  }
}

__global__ void k_partition(const int *__restrict__ flag, const int *__restrict__ pref, const int *__restrict__ segof,
                            const int *__restrict__ seg_first, const int *__restrict__ seg_count, const int *__restrict__ harr,
                            const int *__restrict__ zarr, const int *__restrict__ perm, int n, int *perm_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = segof[i];
  if (s < 0) { perm_out[i] = perm[i]; return; }
  const int f = seg_first[s], c = seg_count[s];
  const int nL = pref[f + c] - pref[f];
  const int x = i - f, lbefore = pref[i] - pref[f];
  const int K = nL - (pref[f + nL] - pref[f]);              // holes = non-L among the first nL
  int dest;
  if (x < nL) {
    if (flag[i]) dest = x;
    else {
      const int k = x - lbefore + 1;                        // 1-based hole rank
      dest = (k == 1 ? c : zarr[f + k - 2]) - 1;            // Z_{k-1} - 1, Z_0 = n
    }
  } else if (flag[i]) {
    const int j = nL - lbefore;                             // 1-based rank from the back
    dest = harr[f + j - 1];
  } else if (x == nL) {
    dest = (K == 0 ? c : zarr[f + K - 1]) - 1;              // a[nL]: the (K+1)-th front element
  } else {
    dest = x - 1;
  }
  perm_out[f + dest] = perm[i];
}

struct NodeArrays {
  int *first, *count, *left, *parent; // breadth-first numbering
  double *bmin, *bmax;                // 3 per node
};

// children of the segments that really split (0 < nL < n); new nodes are numbered breadth-first
__global__ void k_children(const int *__restrict__ seg_first, const int *__restrict__ seg_count, const int *__restrict__ seg_node,
                           const int *__restrict__ pref, const int *__restrict__ split_rank, int nseg, int nodes_used,
                           NodeArrays na) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  const int f = seg_first[s], c = seg_count[s];
  const int nL = pref[f + c] - pref[f];
  if (nL == 0 || nL == c) return;                            // "Abort split if one side is empty", mybvh.cpp:451-452
  const int l = nodes_used + 2 * split_rank[s], p = seg_node[s];
  na.first[l] = f;          na.count[l] = nL;       na.left[l] = 0;     na.parent[l] = p;
  na.first[l + 1] = f + nL; na.count[l + 1] = c - nL; na.left[l + 1] = 0; na.parent[l + 1] = p;
  na.left[p] = l;
  na.count[p] = 0;                                           // interior <=> triCount_ == 0, mybvh.cpp:464
}

__global__ void k_split_flag(const int *__restrict__ seg_first, const int *__restrict__ seg_count, const int *__restrict__ pref,
                             int nseg, int *does_split) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  const int f = seg_first[s], c = seg_count[s];
  const int nL = pref[f + c] - pref[f];
  does_split[s] = (nL != 0 && nL != c) ? 1 : 0;
}

// next level's segments: the new children with more than 2 triangles (mybvh.cpp:441), in ascending order
__global__ void k_next_active(const NodeArrays na, int node_lo, int node_hi, int *is_active) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (node_lo + k >= node_hi) return;
  is_active[k] = na.count[node_lo + k] > 2 ? 1 : 0;
}
__global__ void k_next_segments(const NodeArrays na, int node_lo, int node_hi, const int *__restrict__ is_active,
                                const int *__restrict__ rank, int *seg_first, int *seg_count, int *seg_node) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (node_lo + k >= node_hi || !is_active[k]) return;
  const int s = rank[k];
  seg_first[s] = na.first[node_lo + k];
  seg_count[s] = na.count[node_lo + k];
  seg_node[s] = node_lo + k;
}

// order-preserving map double -> unsigned 64 for atomicMin/atomicMax
__device__ __forceinline__ unsigned long long enc(double d) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(d);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__host__ __device__ inline double dec(unsigned long long k) {
  const unsigned long long u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  double d;
  memcpy(&d, &u, 8);
  return d;
}

// position -> node created at this level (children ranges are disjoint and sorted): bounds by atomics,
// one per warp when the whole warp lies in one node (updateNodeBoundsSoA, mybvh.cpp:412-431)
__global__ void k_bounds(const int *__restrict__ perm, const double *__restrict__ tmin, const double *__restrict__ tmax,
                         const int *__restrict__ node_first, const int *__restrict__ node_count, int node_lo, int nnew, int n,
                         unsigned long long *kmin, unsigned long long *kmax) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  int node = -1;
  if (i < n) {
    int lo = 0, hi = nnew;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (node_first[node_lo + mid] <= i) lo = mid + 1; else hi = mid;
    }
    const int s = lo - 1;
    if (s >= 0 && i < node_first[node_lo + s] + node_count[node_lo + s]) node = node_lo + s;
  }
  unsigned long long mn[3], mx[3];
  if (node >= 0) {
    const int r = perm[i];
    for (int k = 0; k < 3; ++k) { mn[k] = enc(tmin[3 * (size_t)r + k]); mx[k] = enc(tmax[3 * (size_t)r + k]); }
  }
  const unsigned full = 0xffffffffu;
  const int node0 = __shfl_sync(full, node, 0);
  const bool uniform = __all_sync(full, node == node0);
  if (uniform) {
    if (node0 < 0) return;
    for (int k = 0; k < 3; ++k) {
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long a = __shfl_xor_sync(full, mn[k], o), b = __shfl_xor_sync(full, mx[k], o);
        mn[k] = a < mn[k] ? a : mn[k];
        mx[k] = b > mx[k] ? b : mx[k];
      }
    }
    if ((threadIdx.x & 31) == 0)
      for (int k = 0; k < 3; ++k) { atomicMin(kmin + 3 * (size_t)node0 + k, mn[k]); atomicMax(kmax + 3 * (size_t)node0 + k, mx[k]); }
  } else if (node >= 0) {
    for (int k = 0; k < 3; ++k) { atomicMin(kmin + 3 * (size_t)node + k, mn[k]); atomicMax(kmax + 3 * (size_t)node + k, mx[k]); }
  }
}

// ---- renumbering: breadth-first ids -> the order the reference recursion allocates nodes in ------------
__global__ void k_desc(const int *__restrict__ left, int lo, int hi, int *desc) {
  const int p = lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= hi) return;
  const int l = left[p];
  desc[p] = l ? 2 + desc[l] + desc[l + 1] : 0;
}
__global__ void k_base(const int *__restrict__ left, const int *__restrict__ desc, int lo, int hi, int *base, int *newidx) {
  const int p = lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= hi) return;
  const int l = left[p];
  if (!l) return;
  const int b = base[p];                 // nodesUsed_ when subdivideSoA(p) runs
  newidx[l] = b; newidx[l + 1] = b + 1;  // children in an adjacent pair, mybvh.cpp:456-458
  base[l] = b + 2;                       // the left subtree is subdivided first ...
  base[l + 1] = b + 2 + desc[l];         // ... and allocates all its descendants before the right one starts
}
__global__ void k_emit(const NodeArrays na, const unsigned long long *__restrict__ kmin, const unsigned long long *__restrict__ kmax,
                       const int *__restrict__ newidx, int nodes, double *out_min, double *out_max, int *out_left, int *out_first,
                       int *out_count) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= nodes) return;
  const int q = newidx[p];
  out_first[q] = na.first[p];
  out_count[q] = na.count[p];
  out_left[q] = na.left[p] ? newidx[na.left[p]] : 0;
  for (int k = 0; k < 3; ++k) {
    out_min[4 * (size_t)q + k] = dec(kmin[3 * (size_t)p + k]);
    out_max[4 * (size_t)q + k] = dec(kmax[3 * (size_t)p + k]);
  }
  out_min[4 * (size_t)q + 3] = 0.0;
  out_max[4 * (size_t)q + 3] = 0.0;
}
__global__ void k_seg_end(const int *__restrict__ f, const int *__restrict__ c, int n, int *e) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) e[s] = f[s] + c[s];
}

struct DevBuf {
  std::vector<void *> all;
  ~DevBuf() { for (void *p : all) cudaFree(p); }
  template <typename T> cudaError_t get(T *&p, size_t n) {
    cudaError_t e = cudaMalloc(&p, (n ? n : 1) * sizeof(T));
    if (e == cudaSuccess) all.push_back(p);
    return e;
  }
};

template <typename T> void gather_rows_host(T *rows, const std::vector<int> &perm, int width) {
  if (!rows) return;
  std::vector<T> tmp(rows, rows + (size_t)perm.size() * width);
  for (size_t pos = 0; pos < perm.size(); ++pos)
    for (int k = 0; k < width; ++k) rows[pos * width + k] = tmp[(size_t)perm[pos] * width + k];
}

} // namespace

extern "C" int rt_bvh_build_median_device(rt_data_soa *data, rt_bvh_nodes_soa *nodes, int *nodes_used, double *device_ms) {
  if (!data || !nodes || !nodes_used) { rtb::set_error("rt_bvh_build_median_device: null argument"); return RT_ERR_INVALID_ARG; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    rtb::set_error("rt_bvh_build_median_device: no CUDA device (use rt_bvh_build_median for the host build)");
    return RT_ERR_NO_DEVICE;
  }
  const int N = data->tVertexIdxCount_ / 3;
  if (device_ms) *device_ms = 0.0;
  if (N <= 0) { *nodes_used = 0; return RT_OK; } // "No mesh available", mybvh.cpp:402-404
  if (!data->vertexIdx_ || !data->vertexPos_ || !nodes->bb_min_ || !nodes->bb_max_ || !nodes->leftChildIdx_ ||
      !nodes->firstTriIdx_ || !nodes->triCount_) {
    rtb::set_error("rt_bvh_build_median_device: null array");
    return RT_ERR_INVALID_ARG;
  }
  const int nv = data->tVertexCount_;
  for (int i = 0; i < 3 * N; ++i)
    if (data->vertexIdx_[i] < 0 || data->vertexIdx_[i] >= nv) {
      rtb::set_error("rt_bvh_build_median_device: vertex index out of range");
      return RT_ERR_BAD_SCENE;
    }
  const int cap = 2 * N - 1;
  DevBuf B;
  double *vpos, *cent[3], *tmin, *tmax, *keys, *sorted, *split, *o_min, *o_max;
  int *vidx, *permA, *permB, *flag, *pref, *segof, *harr, *zarr, *segf[2], *segc[2], *segn[2], *sege, *dsplit, *srank, *isact, *arank;
  int *o_left, *o_first, *o_count, *desc, *base, *newidx;
  unsigned long long *kmin, *kmax;
  NodeArrays na;
  BV_CUDA(B.get(vpos, 4 * (size_t)nv)); BV_CUDA(B.get(vidx, 3 * (size_t)N));
  for (int k = 0; k < 3; ++k) BV_CUDA(B.get(cent[k], N));
  BV_CUDA(B.get(tmin, 3 * (size_t)N)); BV_CUDA(B.get(tmax, 3 * (size_t)N));
  BV_CUDA(B.get(keys, N)); BV_CUDA(B.get(sorted, N)); BV_CUDA(B.get(split, N / 2 + 2));
  BV_CUDA(B.get(permA, N)); BV_CUDA(B.get(permB, N)); BV_CUDA(B.get(flag, (size_t)N + 1)); BV_CUDA(B.get(pref, (size_t)N + 1));
  BV_CUDA(B.get(segof, N)); BV_CUDA(B.get(harr, N)); BV_CUDA(B.get(zarr, N));
  for (int k = 0; k < 2; ++k) { BV_CUDA(B.get(segf[k], N)); BV_CUDA(B.get(segc[k], N)); BV_CUDA(B.get(segn[k], N)); }
  BV_CUDA(B.get(sege, N)); BV_CUDA(B.get(dsplit, (size_t)N + 1)); BV_CUDA(B.get(srank, (size_t)N + 1));
  BV_CUDA(B.get(isact, (size_t)N + 1)); BV_CUDA(B.get(arank, (size_t)N + 1));
  BV_CUDA(B.get(na.first, cap)); BV_CUDA(B.get(na.count, cap)); BV_CUDA(B.get(na.left, cap)); BV_CUDA(B.get(na.parent, cap));
  BV_CUDA(B.get(desc, cap)); BV_CUDA(B.get(base, cap)); BV_CUDA(B.get(newidx, cap));
  BV_CUDA(B.get(kmin, 3 * (size_t)cap)); BV_CUDA(B.get(kmax, 3 * (size_t)cap));
  BV_CUDA(B.get(o_min, 4 * (size_t)cap)); BV_CUDA(B.get(o_max, 4 * (size_t)cap));
  BV_CUDA(B.get(o_left, cap)); BV_CUDA(B.get(o_first, cap)); BV_CUDA(B.get(o_count, cap));
  size_t tmp_bytes = 0, need = 0;
  {
    cub::DeviceSegmentedSort::SortKeys(nullptr, need, keys, sorted, N, N / 3 + 1, segf[0], sege);
    tmp_bytes = need;
    cub::DeviceScan::ExclusiveSum(nullptr, need, flag, pref, N + 1);
    if (need > tmp_bytes) tmp_bytes = need;
  }
  unsigned char *tmp;
  BV_CUDA(B.get(tmp, tmp_bytes));

  cudaEvent_t e0, e1;
  BV_CUDA(cudaEventCreate(&e0)); BV_CUDA(cudaEventCreate(&e1));
  BV_CUDA(cudaMemcpy(vpos, data->vertexPos_, sizeof(double) * 4 * (size_t)nv, cudaMemcpyHostToDevice));
  BV_CUDA(cudaMemcpy(vidx, data->vertexIdx_, sizeof(int) * 3 * (size_t)N, cudaMemcpyHostToDevice));
  BV_CUDA(cudaEventRecord(e0));
  k_centroids<<<blocks_for(N), kThreads>>>(vpos, vidx, N, cent[0], cent[1], cent[2], tmin, tmax);
  k_iota<<<blocks_for(N), kThreads>>>(permA, N);
  BV_CUDA(cudaMemset(kmin, 0xff, sizeof(unsigned long long) * 3 * (size_t)cap));
  BV_CUDA(cudaMemset(kmax, 0x00, sizeof(unsigned long long) * 3 * (size_t)cap));
  BV_CUDA(cudaMemset(na.left, 0, sizeof(int) * cap));
  BV_CUDA(cudaMemset(na.count, 0, sizeof(int) * cap));
  BV_CUDA(cudaMemset(na.first, 0, sizeof(int) * cap));
  {
    const int root[4] = {0, N, 0, -1};
    BV_CUDA(cudaMemcpy(na.first, &root[0], 4, cudaMemcpyHostToDevice));
    BV_CUDA(cudaMemcpy(na.count, &root[1], 4, cudaMemcpyHostToDevice));
    BV_CUDA(cudaMemcpy(na.parent, &root[3], 4, cudaMemcpyHostToDevice));
  }
  k_bounds<<<blocks_for(N), kThreads>>>(permA, tmin, tmax, na.first, na.count, 0, 1, N, kmin, kmax);
  std::vector<std::pair<int, int>> levels; // breadth-first id range of the nodes created per level
  levels.push_back({0, 1});
  int used = 1, nseg = 0, cur = 0, depth = 1;
  if (N > 2) {
    const int s0[3] = {0, N, 0};
    BV_CUDA(cudaMemcpy(segf[0], &s0[0], 4, cudaMemcpyHostToDevice));
    BV_CUDA(cudaMemcpy(segc[0], &s0[1], 4, cudaMemcpyHostToDevice));
    BV_CUDA(cudaMemcpy(segn[0], &s0[2], 4, cudaMemcpyHostToDevice));
    nseg = 1;
  }
  int *perm = permA, *perm2 = permB;
  while (nseg > 0) {
    if (depth > 4096) { rtb::set_error("rt_bvh_build_median_device: runaway depth"); return RT_ERR_LIMIT; }
    const int axis = depth % 3;                               // mybvh.cpp:444 (root depth 1: y, z, x, y, ...)
    k_segof<<<blocks_for(N), kThreads>>>(segf[cur], segc[cur], nseg, N, segof);
    k_keys<<<blocks_for(N), kThreads>>>(cent[axis], perm, N, keys);
    k_seg_end<<<blocks_for(nseg), kThreads>>>(segf[cur], segc[cur], nseg, sege);
    size_t tb = tmp_bytes;
    BV_CUDA(cub::DeviceSegmentedSort::SortKeys(tmp, tb, keys, sorted, N, nseg, segf[cur], sege));
    k_split<<<blocks_for(nseg), kThreads>>>(sorted, segf[cur], segc[cur], nseg, split);
    k_flags<<<blocks_for(N), kThreads>>>(keys, segof, split, N, flag);
    BV_CUDA(cudaMemsetAsync(flag + N, 0, sizeof(int)));
    tb = tmp_bytes;
    BV_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, flag, pref, N + 1));
    k_roles<<<blocks_for(N), kThreads>>>(flag, pref, segof, segf[cur], segc[cur], N, harr, zarr);
    k_partition<<<blocks_for(N), kThreads>>>(flag, pref, segof, segf[cur], segc[cur], harr, zarr, perm, N, perm2);
    { int *t = perm; perm = perm2; perm2 = t; }
    k_split_flag<<<blocks_for(nseg), kThreads>>>(segf[cur], segc[cur], pref, nseg, dsplit);
    BV_CUDA(cudaMemsetAsync(dsplit + nseg, 0, sizeof(int)));
    tb = tmp_bytes;
    BV_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, dsplit, srank, nseg + 1));
    int nsplit = 0;
    BV_CUDA(cudaMemcpy(&nsplit, srank + nseg, sizeof(int), cudaMemcpyDeviceToHost));
    if (nsplit == 0) break;
    if (used + 2 * nsplit > cap) { rtb::set_error("rt_bvh_build_median_device: node pool overflow"); return RT_ERR_BAD_SCENE; }
    k_children<<<blocks_for(nseg), kThreads>>>(segf[cur], segc[cur], segn[cur], pref, srank, nseg, used, na);
    const int lo = used, hi = used + 2 * nsplit;
    k_bounds<<<blocks_for(N), kThreads>>>(perm, tmin, tmax, na.first, na.count, lo, hi - lo, N, kmin, kmax);
    levels.push_back({lo, hi});
    k_next_active<<<blocks_for(hi - lo), kThreads>>>(na, lo, hi, isact);
    BV_CUDA(cudaMemsetAsync(isact + (hi - lo), 0, sizeof(int)));
    tb = tmp_bytes;
    BV_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, isact, arank, hi - lo + 1));
    int nnext = 0;
    BV_CUDA(cudaMemcpy(&nnext, arank + (hi - lo), sizeof(int), cudaMemcpyDeviceToHost));
    k_next_segments<<<blocks_for(hi - lo), kThreads>>>(na, lo, hi, isact, arank, segf[1 - cur], segc[1 - cur], segn[1 - cur]);
    used = hi;
    nseg = nnext;
    cur = 1 - cur;
    ++depth;
  }
  // renumber to the reference's allocation order and emit the five arrays
  for (int l = (int)levels.size() - 1; l >= 0; --l)
    k_desc<<<blocks_for(levels[l].second - levels[l].first), kThreads>>>(na.left, levels[l].first, levels[l].second, desc);
  {
    const int zero = 0, one = 1;
    BV_CUDA(cudaMemcpy(newidx, &zero, 4, cudaMemcpyHostToDevice));
    BV_CUDA(cudaMemcpy(base, &one, 4, cudaMemcpyHostToDevice));   // int nodesUsed_ = 1, mybvh.h:170
  }
  for (size_t l = 0; l < levels.size(); ++l)
    k_base<<<blocks_for(levels[l].second - levels[l].first), kThreads>>>(na.left, desc, levels[l].first, levels[l].second, base, newidx);
  BV_CUDA(cudaMemset(o_min, 0, sizeof(double) * 4 * (size_t)cap)); BV_CUDA(cudaMemset(o_max, 0, sizeof(double) * 4 * (size_t)cap));
  BV_CUDA(cudaMemset(o_left, 0, sizeof(int) * cap)); BV_CUDA(cudaMemset(o_first, 0, sizeof(int) * cap));
  BV_CUDA(cudaMemset(o_count, 0, sizeof(int) * cap));
  k_emit<<<blocks_for(used), kThreads>>>(na, kmin, kmax, newidx, used, o_min, o_max, o_left, o_first, o_count);
  BV_CUDA(cudaEventRecord(e1));
  BV_CUDA(cudaEventSynchronize(e1));
  BV_CUDA(cudaGetLastError());
  float ms = 0.f;
  BV_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  if (device_ms) *device_ms = ms;
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  BV_CUDA(cudaMemcpy(nodes->bb_min_, o_min, sizeof(double) * 4 * (size_t)cap, cudaMemcpyDeviceToHost));
  BV_CUDA(cudaMemcpy(nodes->bb_max_, o_max, sizeof(double) * 4 * (size_t)cap, cudaMemcpyDeviceToHost));
  BV_CUDA(cudaMemcpy(nodes->leftChildIdx_, o_left, sizeof(int) * cap, cudaMemcpyDeviceToHost));
  BV_CUDA(cudaMemcpy(nodes->firstTriIdx_, o_first, sizeof(int) * cap, cudaMemcpyDeviceToHost));
  BV_CUDA(cudaMemcpy(nodes->triCount_, o_count, sizeof(int) * cap, cudaMemcpyDeviceToHost));
  std::vector<int> hperm(N);
  BV_CUDA(cudaMemcpy(hperm.data(), perm, sizeof(int) * (size_t)N, cudaMemcpyDeviceToHost));
  // apply the row permutation the reference performs swap by swap (mybvh.cpp:497-503)
  gather_rows_host(data->normals_, hperm, 1);
  gather_rows_host(data->vertexIdx_, hperm, 3);
  gather_rows_host(data->textureIdx_, hperm, 3);
  *nodes_used = used;
  return RT_OK;
}
