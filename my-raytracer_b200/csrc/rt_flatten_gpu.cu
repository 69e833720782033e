// rt_flatten_gpu.cu — the flattener on the GPU: reference Data SoA + BVHNodes_SoA in, the HBM blob of rt_types.h out,
// without the host ever touching a per-triangle or per-node record (north star (a) "GPU-friendly flattening of the
// reference's BVH"; sits beside Raytracer::build_Data mytracer.cpp:166-296 / BVH::initSoA mybvh.cpp:375-406, whose
// arrays it consumes as they are).
//
// Same blob contract as rt_flatten.cpp (the host twin, which stays the reference for error messages and for the small
// special cases): the caller's arrays are uploaded once, then
//   1. per-triangle boxes (outward-rounded FP32) and input validation                         one thread per triangle
//   2. the SAH traversal tree on the device (rt_sah_gpu.cu), its binary NodeRec records        scan + one thread per node
//   3. its 4-wide and 8-wide twins, collapsed level by level (the greedy largest-area rule of  one thread per wide node,
//      rt_flatten.cpp collapse_wide: same structure, records numbered in level order)          one scan per level
//   4. the reference's own tree, discovered from node 0 level by level (the reference never    one thread per node,
//      records how many slots are in use, mybvh.h:57), validated like the host walk            one scan per level
//   5. TriFilter / TriExact in traversal order, TriShade in reference order, texels            one thread per record
// are written straight into (or copied device-to-device into) one allocation that rt_api.cu adopts without a copy.
// Records of a tree are numbered in level order here and in DFS pre-order by the host twin; frames do not depend on the
// numbering (hits are decided by the exact FP64 test, ties by the reference's primitive index).
//
// Anything unusual — a scene the device path does not cover (reference topology requested, a root that is a leaf,
// fewer than 3 triangles) or ANY validation failure — returns kFlattenUseHost: the host flattener then runs and
// reports the error in its own words.
#include "rt_flatten.h"
#include "rt_sah.h"
#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <string>
#include <vector>

namespace rtb {
namespace {

struct Scal {                    // scalars the kernels reduce into
  int err;                       // any validation failure
  int max_leaf, worst_stack, pad;
  unsigned long long extent;     // bits of a non-negative double: max |coordinate|
  unsigned long long box_extent; // bits of a non-negative double: max |ctr| + half
};

struct DevData {                 // device copies of the caller's arrays (only those the scene needs)
  int N, nv, num_meshes, num_uv;
  const int *vertexIdx, *vertexMeshId, *textureIdx, *drawMode, *texWidth;
  const double *vertexPos, *vertexNormals, *normals; // vec4 arrays as 4 doubles
  const double *U, *V;
};

struct RefBvh { const double *bb_min, *bb_max; const int *left, *first, *count; int cap; };

__device__ inline void atomic_max_d(unsigned long long *p, double v) {
  // warp-level maximum first: one atomic per warp
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0 && v > 0.0) atomicMax(p, (unsigned long long)__double_as_longlong(v));
}

// centre / half-extent form, outward (rt_flatten.cpp box_lohi, same arithmetic); returns max |ctr| + half
__device__ inline double box_ctr_half(const double *lo, const double *hi, float *ctr, float *half) {
  double ext = 0.0;
  for (int k = 0; k < 3; ++k) {
    const float c = __double2float_rn(__dmul_rn(0.5, __dadd_rn(lo[k], hi[k])));
    const double need = fmax(__dsub_rn(hi[k], (double)c), __dsub_rn((double)c, lo[k]));
    float h = __double2float_ru(need > 0.0 ? need : 0.0);
    h = nextafterf(h, INFINITY);
    ctr[k] = c; half[k] = h;
    ext = fmax(ext, fabs((double)c) + (double)h);
  }
  return ext;
}

// ---- 1. per-triangle boxes + validation of the index arrays ---------------------------------------------------------
__global__ void tri_box_kernel(DevData d, float *__restrict__ plo, float *__restrict__ phi, Scal *sc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double ext = 0.0;
  if (i < d.N) {
    const int v[3] = {d.vertexIdx[3 * (size_t)i], d.vertexIdx[3 * (size_t)i + 1], d.vertexIdx[3 * (size_t)i + 2]};
    bool ok = true;
    for (int k = 0; k < 3; ++k) ok = ok && v[k] >= 0 && v[k] < d.nv;
    if (ok) {
      const int mesh = d.vertexMeshId[v[0]];
      ok = mesh >= 0 && mesh < d.num_meshes;
    }
    if (!ok) { sc->err = 1; for (int k = 0; k < 3; ++k) plo[3 * (size_t)i + k] = phi[3 * (size_t)i + k] = 0.0f; }
    else {
      for (int k = 0; k < 3; ++k) {
        const double a = d.vertexPos[4 * (size_t)v[0] + k], b = d.vertexPos[4 * (size_t)v[1] + k], c = d.vertexPos[4 * (size_t)v[2] + k];
        plo[3 * (size_t)i + k] = __double2float_rd(fmin(a, fmin(b, c)));
        phi[3 * (size_t)i + k] = __double2float_ru(fmax(a, fmax(b, c)));
        ext = fmax(ext, fmax(fabs(a), fmax(fabs(b), fabs(c))));
      }
    }
  }
  atomic_max_d(&sc->extent, ext);
}

// ---- 2. binary records of the SAH tree --------------------------------------------------------------------------------
// The builder's node array has two used stretches: [0, nodes_a) and [cap_a, reserve).  `j` runs over both.
struct SahSlots {
  const SahNode *nodes;
  long long nodes_a, cap_a, total;   // total = nodes_a + (reserve - cap_a)
  __device__ long long slot_of(long long j) const { return j < nodes_a ? j : cap_a + (j - nodes_a); }
  __device__ long long index_of(long long slot) const { return slot < cap_a ? slot : nodes_a + (slot - cap_a); }
};

__global__ void sah_flag_kernel(SahSlots s, int *__restrict__ flag) {
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j < s.total) flag[j] = s.nodes[s.slot_of(j)].count == 0 ? 1 : 0;
}

__global__ void sah_emit_kernel(SahSlots s, const int *__restrict__ flag, const int *__restrict__ idx,
                                NodeRec *__restrict__ out, Scal *sc) {
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  double ext = 0.0;
  if (j < s.total && flag[j]) {
    const SahNode nd = s.nodes[s.slot_of(j)];
    const int ch[2] = {nd.left, nd.right};
    NodeRec r;
    int ref[2], cnt[2];
    for (int k = 0; k < 2; ++k) {
      const SahNode cn = s.nodes[ch[k]];
      const double lo[3] = {cn.lo[0], cn.lo[1], cn.lo[2]}, hi[3] = {cn.hi[0], cn.hi[1], cn.hi[2]};
      ext = fmax(ext, box_ctr_half(lo, hi, k ? r.c1_ctr : r.c0_ctr, k ? r.c1_half : r.c0_half));
      if (cn.count > 0) { ref[k] = cn.first; cnt[k] = cn.count; atomicMax(&sc->max_leaf, cn.count); }
      else { ref[k] = idx[s.index_of(ch[k])]; cnt[k] = 0; }
    }
    r.ref0 = ref[0]; r.cnt0 = cnt[0]; r.ref1 = ref[1]; r.cnt1 = cnt[1];
    out[idx[j]] = r;
  }
  atomic_max_d(&sc->box_extent, ext);
}

// ---- 3. W-wide twin of the SAH tree, one level of wide nodes per round ----------------------------------------------
// A wide node = a binary node's two children, then repeatedly the interior child with the largest surface area replaced
// in place by its own two children until there are W or none is interior (rt_flatten.cpp collapse_wide).
__device__ inline double sah_area(const SahNode &b) {
  const double dx = (double)b.hi[0] - b.lo[0], dy = (double)b.hi[1] - b.lo[1], dz = (double)b.hi[2] - b.lo[2];
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dy), __dmul_rn(dy, dz)), __dmul_rn(dz, dx));  // the host twin compiles without FMA contraction
}

// frontier item i = (binary node fr_node[i], stack entries above it fr_stack[i]).  Writes its child slots and how many of
// them are interior (the scan of that gives the children their record numbers).
template <int W>
__global__ void wide_expand_kernel(const SahNode *__restrict__ nodes, const int *__restrict__ fr_node, int nf,
                                   int *__restrict__ slots /* nf x W, -1 = empty */, int *__restrict__ ninterior) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nf) return;
  const SahNode root = nodes[fr_node[i]];
  int slot[W], ns = 2;
  slot[0] = root.left; slot[1] = root.right;
  while (ns < W) {
    int pick = -1;
    double best = -1.0;
    for (int k = 0; k < ns; ++k) {
      const SahNode &c = nodes[slot[k]];
      if (c.count != 0) continue;
      const double a = sah_area(c);
      if (a > best) { best = a; pick = k; }
    }
    if (pick < 0) break;
    const int l = nodes[slot[pick]].left, r = nodes[slot[pick]].right;
    for (int k = ns; k > pick + 1; --k) slot[k] = slot[k - 1];
    slot[pick] = l; slot[pick + 1] = r;
    ++ns;
  }
  int interior = 0;
  for (int k = 0; k < W; ++k) {
    slots[(size_t)i * W + k] = k < ns ? slot[k] : -1;
    if (k < ns && nodes[slot[k]].count == 0) ++interior;
  }
  ninterior[i] = interior;
}

template <int W, typename Rec>
__global__ void wide_emit_kernel(const SahNode *__restrict__ nodes, const int *__restrict__ fr_stack, int nf,
                                 const int *__restrict__ slots, const int *__restrict__ ninterior,
                                 const int *__restrict__ off, int level_base, int next_base, Rec *__restrict__ out,
                                 int *__restrict__ next_node, int *__restrict__ next_stack, Scal *sc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double ext = 0.0;
  if (i < nf) {
    Rec n;
    const int interior = ninterior[i];
    const int stack_after = fr_stack[i] + (interior > 1 ? interior - 1 : 0);
    atomicMax(&sc->worst_stack, stack_after);
    int child = off[i];
    for (int k = 0; k < W; ++k) {
      float ctr[3] = {0.0f, 0.0f, 0.0f}, half[3] = {-1e30f, -1e30f, -1e30f};
      const int s = slots[(size_t)i * W + k];
      if (s >= 0) {
        const SahNode cb = nodes[s];
        const double lo[3] = {cb.lo[0], cb.lo[1], cb.lo[2]}, hi[3] = {cb.hi[0], cb.hi[1], cb.hi[2]};
        ext = fmax(ext, box_ctr_half(lo, hi, ctr, half));
        if (cb.count > 0) { n.ref[k] = cb.first; n.cnt[k] = cb.count; }
        else {
          n.ref[k] = next_base + child; n.cnt[k] = 0;
          next_node[child] = s; next_stack[child] = stack_after;
          ++child;
        }
      } else { n.ref[k] = -1; n.cnt[k] = 0; }
      for (int a = 0; a < 3; ++a) { n.box[k >> 1][a][k & 1] = ctr[a]; n.box[k >> 1][3 + a][k & 1] = half[a]; }
    }
    out[level_base + i] = n;
  }
  atomic_max_d(&sc->box_extent, ext);
}

// ---- 4. the reference's tree, level by level from node 0 ---------------------------------------------------------------
// frontier item i = reference node fr[i] (interior).  Pass 1 validates its child pair and counts interior children.
__global__ void ref_count_kernel(RefBvh b, const int *__restrict__ fr, int nf, int *__restrict__ ninterior, Scal *sc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nf) return;
  const int l = b.left[fr[i]];
  int n = 0;
  if (l < 1 || l + 1 > b.cap - 1) sc->err = 1;            // child index out of range (rt_flatten.cpp walk)
  else n = (b.count[l] > 0 ? 0 : 1) + (b.count[l + 1] > 0 ? 0 : 1);
  ninterior[i] = n;
}

// Pass 2 writes the record (both child boxes, NodeMaxX), marks the triangles of leaf children and queues interior ones.
__global__ void ref_emit_kernel(RefBvh b, const int *__restrict__ fr, int nf, const int *__restrict__ off, int level_base,
                                int next_base, int N, NodeRec *__restrict__ out, NodeMaxX *__restrict__ maxx,
                                int *__restrict__ next, int *__restrict__ seen, Scal *sc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double ext = 0.0;
  if (i < nf) {
    const int l = b.left[fr[i]];
    if (l >= 1 && l + 1 <= b.cap - 1) {
      NodeRec r;
      int ref[2], cnt[2];
      int child = off[i];
      for (int s = 0; s < 2; ++s) {
        const int c = l + s;
        const double lo[3] = {b.bb_min[4 * (size_t)c], b.bb_min[4 * (size_t)c + 1], b.bb_min[4 * (size_t)c + 2]};
        const double hi[3] = {b.bb_max[4 * (size_t)c], b.bb_max[4 * (size_t)c + 1], b.bb_max[4 * (size_t)c + 2]};
        ext = fmax(ext, box_ctr_half(lo, hi, s ? r.c1_ctr : r.c0_ctr, s ? r.c1_half : r.c0_half));
        const int tc = b.count[c];
        if (tc > 0) {
          const int f = b.first[c];
          if (f < 0 || f > N - tc) { sc->err = 1; ref[s] = 0; cnt[s] = 1; }
          else {
            for (int t = f; t < f + tc; ++t) atomicAdd(seen + t, 1);
            ref[s] = f; cnt[s] = tc;
            atomicMax(&sc->max_leaf, tc);
          }
        } else {
          ref[s] = next_base + child; cnt[s] = 0;
          next[child] = c;
          ++child;
        }
      }
      // sibling leaves own adjacent ranges (mybvh.cpp:481-513); the kernel tests them as one range
      if (cnt[0] > 0 && cnt[1] > 0 && ref[1] != ref[0] + cnt[0]) sc->err = 1;
      r.ref0 = ref[0]; r.cnt0 = cnt[0]; r.ref1 = ref[1]; r.cnt1 = cnt[1];
      out[level_base + i] = r;
      maxx[level_base + i] = NodeMaxX{b.bb_max[4 * (size_t)l], b.bb_max[4 * (size_t)(l + 1)]};
    }
  }
  atomic_max_d(&sc->box_extent, ext);
}

__global__ void seen_check_kernel(const int *__restrict__ seen, int N, Scal *sc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N && seen[i] != 1) sc->err = 1;                  // a triangle in no leaf, or in two
}

__global__ void add_int_kernel(NodeRec *__restrict__ recs, int n, int delta) {
  // interior child references of a tree that moves `delta` records up in the NodeRec array
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (recs[i].cnt0 == 0) recs[i].ref0 += delta;
  if (recs[i].cnt1 == 0 && recs[i].c1_half[0] >= 0.0f) recs[i].ref1 += delta;
}

// ---- 5. per-triangle records ----------------------------------------------------------------------------------------------
// Thread t writes TriFilter / TriExact of traversal position t (triangle order[t]) and TriShade of reference row t.
// Same arithmetic as the host twin (rt_flatten.cpp section 2), without FMA contraction.
__global__ void tri_records_kernel(DevData d, const int *__restrict__ order, int disable_filter,
                                   TriFilter *__restrict__ filt, TriExact *__restrict__ exact, TriShade *__restrict__ shade,
                                   int32_t *__restrict__ tri_id, int32_t *__restrict__ tri_pos, Scal *sc) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= d.N) return;
  {
    const int i = order[t];
    tri_id[t] = i; tri_pos[i] = t;
    const int v0 = d.vertexIdx[3 * (size_t)i], v1 = d.vertexIdx[3 * (size_t)i + 1], v2 = d.vertexIdx[3 * (size_t)i + 2];
    TriExact e;
    for (int k = 0; k < 3; ++k) {
      e.p0[k] = d.vertexPos[4 * (size_t)v0 + k]; e.p1[k] = d.vertexPos[4 * (size_t)v1 + k]; e.p2[k] = d.vertexPos[4 * (size_t)v2 + k];
    }
    e.mesh = d.vertexMeshId[v0]; e.pad = 0;
    double e1[3], e2[3];
    for (int k = 0; k < 3; ++k) { e1[k] = __dsub_rn(e.p0[k], e.p2[k]); e2[k] = __dsub_rn(e.p1[k], e.p2[k]); }
    const double a = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(e1[0], e1[0]), __dmul_rn(e1[1], e1[1])), __dmul_rn(e1[2], e1[2])));
    const double b = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(e2[0], e2[0]), __dmul_rn(e2[1], e2[1])), __dmul_rn(e2[2], e2[2])));
    TriFilter f;
    for (int k = 0; k < 3; ++k) { f.p2[k] = __double2float_rn(e.p2[k]); f.e1[k] = __double2float_rn(e1[k]); f.e2[k] = __double2float_rn(e2[k]); }
    const double u = 5.9604644775390625e-8, tiny = 1e-30; // 2^-24
    f.k_bary = __double2float_ru(fmax(__dmul_rn(__dmul_rn(8.0, u), fmax(a, b)), tiny));
    f.k_t = __double2float_ru(fmax(__dmul_rn(__dmul_rn(__dmul_rn(8.0, u), a), b), tiny));
    f.k_det = disable_filter ? INFINITY : __double2float_ru(fmax(__dmul_rn(__dmul_rn(__dmul_rn(64.0, u), a), b), tiny));
    filt[t] = f;
    exact[t] = e;
  }
  {
    const int i = t;
    const int vs[3] = {d.vertexIdx[3 * (size_t)i], d.vertexIdx[3 * (size_t)i + 1], d.vertexIdx[3 * (size_t)i + 2]};
    const int mesh = d.vertexMeshId[vs[0]];
    TriShade s;
    for (int k = 0; k < 9; ++k) s.n[k] = 0.0;
    for (int k = 0; k < 6; ++k) s.uv[k] = 0.0;
    s.mesh = mesh; s.pad = 0;
    if (d.drawMode[mesh] == 0) {
      for (int k = 0; k < 3; ++k) s.n[k] = d.normals[4 * (size_t)i + k];
    } else {
      for (int c = 0; c < 3; ++c)
        for (int k = 0; k < 3; ++k) s.n[3 * c + k] = d.vertexNormals[4 * (size_t)vs[c] + k];
    }
    if (d.texWidth[mesh] != -1) {
      for (int c = 0; c < 3; ++c) {
        const int ti = d.textureIdx[3 * (size_t)i + c];
        if (ti < 0 || ti >= d.num_uv) { sc->err = 1; break; }
        s.uv[2 * c] = d.U[ti]; s.uv[2 * c + 1] = d.V[ti];
      }
    }
    shade[i] = s;
  }
}

__global__ void texel_kernel(const double *__restrict__ texels, size_t n, float4 *__restrict__ out) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) out[t] = make_float4((float)texels[4 * t], (float)texels[4 * t + 1], (float)texels[4 * t + 2], 1.0f);
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// every device allocation of one flatten; whatever is still listed is freed on scope exit
struct Pool {
  std::vector<void *> ptrs;
  ~Pool() { for (void *p : ptrs) cudaFree(p); }
  template <typename T> cudaError_t alloc(T *&p, size_t count) {
    void *q = nullptr;
    const cudaError_t e = cudaMalloc(&q, count ? count * sizeof(T) : 16);
    if (e == cudaSuccess) { ptrs.push_back(q); p = static_cast<T *>(q); }
    return e;
  }
  template <typename T> cudaError_t upload(T *&p, const void *src, size_t count) {   // count elements of T from host memory
    cudaError_t e = alloc(p, count);
    if (e == cudaSuccess && count) e = cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, 0);
    return e;
  }
  void release(void *p) {          // ownership leaves the pool
    for (void *&q : ptrs) if (q == p) q = nullptr;
  }
  void free_now(void *p) {
    for (void *&q : ptrs) if (q == p) { cudaFree(q); q = nullptr; }
  }
};

#define FG(call)                                                                                                       \
  do {                                                                                                                 \
    const cudaError_t e_ = (call);                                                                                     \
    if (e_ != cudaSuccess) {                                                                                           \
      err = std::string("flatten (device): ") + #call + ": " + cudaGetErrorString(e_);                                 \
      cudaGetLastError();                                                                                              \
      return e_ == cudaErrorMemoryAllocation ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA;                                     \
    }                                                                                                                  \
  } while (0)

inline unsigned blocks_of(long long n, int t = 256) { return (unsigned)((n + t - 1) / t); }

// Level-by-level collapse of the SAH tree into W-wide records (numbered in level order, root = record 0).  n_out = 0 when
// the twin is dropped (its worst-case stack need exceeds the traversal stack), as the host twin does.
template <int W, typename Rec>
int build_wide_device(Pool &pool, const SahNode *nodes, int cap_recs, Rec *&out, int &n_out, Scal *d_sc, void *d_tmp,
                      size_t tmp_bytes, std::string &err) {
  n_out = 0;
  out = nullptr;
  int *fr_node[2] = {nullptr, nullptr}, *fr_stack[2] = {nullptr, nullptr}, *slots = nullptr, *nint = nullptr, *off = nullptr;
  FG(pool.alloc(out, (size_t)cap_recs));
  for (int k = 0; k < 2; ++k) { FG(pool.alloc(fr_node[k], (size_t)cap_recs)); FG(pool.alloc(fr_stack[k], (size_t)cap_recs)); }
  FG(pool.alloc(slots, (size_t)cap_recs * W));
  FG(pool.alloc(nint, (size_t)cap_recs));
  FG(pool.alloc(off, (size_t)cap_recs));
  FG(cudaMemsetAsync(&d_sc->worst_stack, 0, sizeof(int), 0));
  FG(cudaMemsetAsync(fr_node[0], 0, sizeof(int), 0));   // the root: builder slot 0, empty stack
  FG(cudaMemsetAsync(fr_stack[0], 0, sizeof(int), 0));
  int nf = 1, level_base = 0, cur = 0;
  bool dropped = false;
  while (nf > 0) {
    wide_expand_kernel<W><<<blocks_of(nf), 256>>>(nodes, fr_node[cur], nf, slots, nint);
    size_t tb = tmp_bytes;
    FG(cub::DeviceScan::ExclusiveSum(d_tmp, tb, nint, off, nf));
    int last[2];
    FG(cudaMemcpy(&last[0], off + nf - 1, sizeof(int), cudaMemcpyDeviceToHost));
    FG(cudaMemcpy(&last[1], nint + nf - 1, sizeof(int), cudaMemcpyDeviceToHost));
    const int total = last[0] + last[1];
    const int next_base = level_base + nf;
    if ((long long)next_base + total > cap_recs) { dropped = true; break; }
    wide_emit_kernel<W, Rec><<<blocks_of(nf), 256>>>(nodes, fr_stack[cur], nf, slots, nint, off, level_base, next_base, out,
                                                      fr_node[cur ^ 1], fr_stack[cur ^ 1], d_sc);
    level_base = next_base;
    nf = total;
    cur ^= 1;
  }
  FG(cudaGetLastError());
  int worst = 0;
  FG(cudaMemcpy(&worst, &d_sc->worst_stack, sizeof(int), cudaMemcpyDeviceToHost));
  for (int k = 0; k < 2; ++k) { pool.free_now(fr_node[k]); pool.free_now(fr_stack[k]); }
  pool.free_now(slots); pool.free_now(nint); pool.free_now(off);
  if (dropped || worst > kStackMax) { pool.free_now(out); out = nullptr; return RT_OK; }
  n_out = level_base;
  return RT_OK;
}

} // namespace

int flatten_scene_device(const rt_data_soa *data, const rt_bvh_nodes_soa *bvh, const rt_object *objects, int num_objects,
                         unsigned flags, DeviceBlob &out, std::string &err) {
  out = DeviceBlob();
  // ---- what the device path covers; everything else (and every malformed input) is the host twin's -------------------
  if (!data || !bvh || (flags & RT_SCENE_REFERENCE_TOPOLOGY) || num_objects < 0 || (num_objects > 0 && !objects)) return kFlattenUseHost;
  if (data->tVertexIdxCount_ < 9 || data->tVertexIdxCount_ % 3 != 0 || data->tMeshCount_ <= 0 || data->tVertexCount_ <= 0) return kFlattenUseHost;
  const int N = data->tVertexIdxCount_ / 3, num_meshes = data->tMeshCount_, nv = data->tVertexCount_;
  if (!bvh->bb_min_ || !bvh->bb_max_ || !bvh->leftChildIdx_ || !bvh->firstTriIdx_ || !bvh->triCount_) return kFlattenUseHost;
  if (!data->vertexIdx_ || !data->vertexPos_ || !data->vertexMeshId_ || !data->normals_ || !data->vertexNormals_ ||
      !data->meshDrawMode_ || !data->meshTexWidth_ || !data->meshTexHeight_ || !data->materialAmbient_ ||
      !data->materialDiffuse_ || !data->materialSpecular_ || !data->materialShininess_ || !data->materialMirror_ ||
      !data->materialShadowable_)
    return kFlattenUseHost;
  if (bvh->triCount_[0] > 0) return kFlattenUseHost;        // the whole mesh is one leaf
  const bool timing = std::getenv("RT_FLATTEN_TIMING") != nullptr;
  auto t_last = std::chrono::steady_clock::now();
  auto mark = [&](const char *what) {
    if (!timing) return;
    cudaDeviceSynchronize();
    const auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[flatten gpu] %-28s %9.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
    t_last = now;
  };

  // ---- meshes (a handful of records: host) -------------------------------------------------------------------------------
  std::vector<MeshRec> meshes(num_meshes);
  size_t num_texels = 0;
  bool any_phong = false, any_flat = false, any_tex = false;
  for (int m = 0; m < num_meshes; ++m) {
    MeshRec &r = meshes[m];
    std::memset(&r, 0, sizeof(r));
    const rt_vec4 &ka = data->materialAmbient_[m], &kd = data->materialDiffuse_[m], &ks = data->materialSpecular_[m];
    r.ambient[0] = ka.x; r.ambient[1] = ka.y; r.ambient[2] = ka.z;
    r.diffuse[0] = kd.x; r.diffuse[1] = kd.y; r.diffuse[2] = kd.z;
    r.specular[0] = ks.x; r.specular[1] = ks.y; r.specular[2] = ks.z;
    r.shininess = data->materialShininess_[m];
    r.mirror = data->materialMirror_[m];
    r.shadowable = data->materialShadowable_[m] ? 1 : 0;
    r.draw_mode = data->meshDrawMode_[m];
    r.tex_w = data->meshTexWidth_[m];
    r.tex_h = data->meshTexHeight_[m];
    (r.draw_mode == 0 ? any_flat : any_phong) = true;
    if (r.tex_w != -1) {
      if (r.tex_w <= 0 || r.tex_h <= 0 || !data->meshTexels_ || !data->firstMeshTex_) return kFlattenUseHost;
      r.tex_offset = data->firstMeshTex_[m];
      const size_t end = (size_t)r.tex_offset + (size_t)r.tex_w * r.tex_h;
      if (end > data->tTexelCount_) return kFlattenUseHost;
      if (end > num_texels) num_texels = end;
      any_tex = true;
    }
  }
  if (any_tex && (!data->textureIdx_ || !data->textureCoordinatesU_ || !data->textureCoordinatesV_ || data->tTextCoordCount_ <= 0))
    return kFlattenUseHost;

  // ---- analytic primitives: the host twin flattens them alone (a few hundred records), merged below -------------------
  std::vector<unsigned char> mini;
  BlobHeader mh;
  std::memset(&mh, 0, sizeof(mh));
  mh.obj_root = -1; mh.obj_root4 = -1;
  if (num_objects > 0) {
    const int st = flatten_scene(nullptr, nullptr, objects, num_objects, flags, mini, err);
    if (st != RT_OK) return kFlattenUseHost;
    std::memcpy(&mh, mini.data(), sizeof(mh));
  }
  FlattenStats &stats = mutable_flatten_stats();
  stats = FlattenStats();
  mark("host part");

  // ---- upload the caller's arrays (only those this scene reads) ------------------------------------------------------------
  Pool pool;
  DevData d;
  std::memset(&d, 0, sizeof(d));
  d.N = N; d.nv = nv; d.num_meshes = num_meshes; d.num_uv = any_tex ? data->tTextCoordCount_ : 0;
  {
    int *p_i = nullptr; double *p_d = nullptr;
    FG(pool.upload(p_i, data->vertexIdx_, 3 * (size_t)N)); d.vertexIdx = p_i;
    FG(pool.upload(p_i, data->vertexMeshId_, (size_t)nv)); d.vertexMeshId = p_i;
    FG(pool.upload(p_i, data->meshDrawMode_, (size_t)num_meshes)); d.drawMode = p_i;
    FG(pool.upload(p_i, data->meshTexWidth_, (size_t)num_meshes)); d.texWidth = p_i;
    FG(pool.upload(p_d, data->vertexPos_, 4 * (size_t)nv)); d.vertexPos = p_d;
    if (any_flat) { FG(pool.upload(p_d, data->normals_, 4 * (size_t)N)); d.normals = p_d; }
    if (any_phong) { FG(pool.upload(p_d, data->vertexNormals_, 4 * (size_t)nv)); d.vertexNormals = p_d; }
    if (any_tex) {
      if (data->tTextIdxCount_ < 3 * N) return kFlattenUseHost;
      FG(pool.upload(p_i, data->textureIdx_, 3 * (size_t)N)); d.textureIdx = p_i;
      FG(pool.upload(p_d, data->textureCoordinatesU_, (size_t)d.num_uv)); d.U = p_d;
      FG(pool.upload(p_d, data->textureCoordinatesV_, (size_t)d.num_uv)); d.V = p_d;
    }
  }
  RefBvh rb;
  rb.cap = 2 * N - 1;   // slots the reference allocates (mybvh.cpp:386)
  {
    int *p_i = nullptr; double *p_d = nullptr;
    FG(pool.upload(p_d, bvh->bb_min_, 4 * (size_t)rb.cap)); rb.bb_min = p_d;
    FG(pool.upload(p_d, bvh->bb_max_, 4 * (size_t)rb.cap)); rb.bb_max = p_d;
    FG(pool.upload(p_i, bvh->leftChildIdx_, (size_t)rb.cap)); rb.left = p_i;
    FG(pool.upload(p_i, bvh->firstTriIdx_, (size_t)rb.cap)); rb.first = p_i;
    FG(pool.upload(p_i, bvh->triCount_, (size_t)rb.cap)); rb.count = p_i;
  }
  double *d_texels_in = nullptr;
  if (num_texels) FG(pool.upload(d_texels_in, data->meshTexels_, 4 * num_texels));
  Scal *d_sc = nullptr;
  FG(pool.alloc(d_sc, 1));
  FG(cudaMemsetAsync(d_sc, 0, sizeof(Scal), 0));
  mark("upload");

  // scan scratch for the largest scan of this flatten (the SAH node slots: at most 4N + 4)
  void *d_tmp = nullptr;
  size_t tmp_bytes = 0;
  {
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (int *)nullptr, (int *)nullptr, 4 * (long long)N + 8);
    unsigned char *p = nullptr;
    FG(pool.alloc(p, tmp_bytes));
    d_tmp = p;
  }

  // ---- 1. triangle boxes, 2. SAH tree on the device -----------------------------------------------------------------------
  float *d_plo = nullptr, *d_phi = nullptr;
  FG(pool.alloc(d_plo, 3 * (size_t)N));
  FG(pool.alloc(d_phi, 3 * (size_t)N));
  tri_box_kernel<<<blocks_of(N), 256>>>(d, d_plo, d_phi, d_sc);
  Scal sc;
  FG(cudaMemcpy(&sc, d_sc, sizeof(sc), cudaMemcpyDeviceToHost));
  if (sc.err) return kFlattenUseHost;
  mark("triangle boxes");
  SahDeviceTree tree;
  struct TreeGuard { SahDeviceTree &t; ~TreeGuard() { free_sah_device_tree(t); } } tree_guard{tree};
  {
    const auto tb = std::chrono::steady_clock::now();
    std::string serr;
    if (!build_sah_on_device(d_plo, d_phi, N, tree, &stats.sah_device_ms, serr)) return kFlattenUseHost;
    stats.sah_builder = 2;
    stats.sah_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tb).count();
  }
  pool.free_now(d_plo); pool.free_now(d_phi);
  if (tree.depth > kStackMax) return kFlattenUseHost;
  mark("build_sah");

  // binary records: interior nodes numbered in builder-slot order (level order on top, one block per small subtree below)
  SahSlots ss;
  ss.nodes = tree.nodes; ss.nodes_a = tree.nodes_a; ss.cap_a = (long long)tree.cap_a;
  ss.total = (long long)tree.nodes_a + ((long long)tree.reserve - (long long)tree.cap_a);
  int *d_flag = nullptr, *d_idx = nullptr;
  FG(pool.alloc(d_flag, (size_t)ss.total));
  FG(pool.alloc(d_idx, (size_t)ss.total));
  sah_flag_kernel<<<blocks_of(ss.total), 256>>>(ss, d_flag);
  {
    size_t tb = tmp_bytes;
    FG(cub::DeviceScan::ExclusiveSum(d_tmp, tb, d_flag, d_idx, (int)ss.total));
  }
  int q_base = 0;
  {
    int last[2];
    FG(cudaMemcpy(&last[0], d_idx + ss.total - 1, sizeof(int), cudaMemcpyDeviceToHost));
    FG(cudaMemcpy(&last[1], d_flag + ss.total - 1, sizeof(int), cudaMemcpyDeviceToHost));
    q_base = last[0] + last[1];
  }
  if (q_base <= 0) return kFlattenUseHost;                 // the root is a leaf: nothing to traverse
  NodeRec *d_trav = nullptr;
  FG(pool.alloc(d_trav, (size_t)q_base));
  sah_emit_kernel<<<blocks_of(ss.total), 256>>>(ss, d_flag, d_idx, d_trav, d_sc);
  FG(cudaGetLastError());
  pool.free_now(d_flag); pool.free_now(d_idx);
  mark("sah records");

  // ---- 3. wide twins -------------------------------------------------------------------------------------------------------
  Node4Rec *d_wide4 = nullptr; Node8Rec *d_wide8 = nullptr;
  int n_wide4 = 0, n_wide8 = 0;
  {
    int st = build_wide_device<4, Node4Rec>(pool, tree.nodes, q_base, d_wide4, n_wide4, d_sc, d_tmp, tmp_bytes, err);
    if (st != RT_OK) return st;
    if (n_wide4 > 0) {
      st = build_wide_device<8, Node8Rec>(pool, tree.nodes, q_base, d_wide8, n_wide8, d_sc, d_tmp, tmp_bytes, err);
      if (st != RT_OK) return st;
    }
  }
  mark("wide twins");

  // ---- 4. the reference tree ------------------------------------------------------------------------------------------------
  NodeRec *d_ref = nullptr; NodeMaxX *d_maxx = nullptr;
  int num_nodes_q = 0, ref_levels = 0;
  {
    int *fr[2] = {nullptr, nullptr}, *nint = nullptr, *off = nullptr, *seen = nullptr;
    FG(pool.alloc(d_ref, (size_t)N));
    FG(pool.alloc(d_maxx, (size_t)N));
    FG(pool.alloc(fr[0], (size_t)N)); FG(pool.alloc(fr[1], (size_t)N));
    FG(pool.alloc(nint, (size_t)N)); FG(pool.alloc(off, (size_t)N));
    FG(pool.alloc(seen, (size_t)N));
    FG(cudaMemsetAsync(seen, 0, (size_t)N * sizeof(int), 0));
    FG(cudaMemsetAsync(fr[0], 0, sizeof(int), 0));      // reference node 0
    int nf = 1, level_base = 0, cur = 0;
    while (nf > 0) {
      ++ref_levels;
      ref_count_kernel<<<blocks_of(nf), 256>>>(rb, fr[cur], nf, nint, d_sc);
      size_t tb = tmp_bytes;
      FG(cub::DeviceScan::ExclusiveSum(d_tmp, tb, nint, off, nf));
      int last[2];
      FG(cudaMemcpy(&last[0], off + nf - 1, sizeof(int), cudaMemcpyDeviceToHost));
      FG(cudaMemcpy(&last[1], nint + nf - 1, sizeof(int), cudaMemcpyDeviceToHost));
      const int total = last[0] + last[1];
      const int next_base = level_base + nf;
      // a tree over N non-empty leaves has at most N - 1 interior nodes: more means a cycle or shared children
      if ((long long)next_base + total > (long long)N - 1 || ref_levels + 1 > kStackMax) return kFlattenUseHost;
      ref_emit_kernel<<<blocks_of(nf), 256>>>(rb, fr[cur], nf, off, level_base, next_base, N, d_ref, d_maxx, fr[cur ^ 1], seen, d_sc);
      level_base = next_base;
      nf = total;
      cur ^= 1;
    }
    num_nodes_q = level_base;
    seen_check_kernel<<<blocks_of(N), 256>>>(seen, N, d_sc);
    FG(cudaGetLastError());
    pool.free_now(fr[0]); pool.free_now(fr[1]); pool.free_now(nint); pool.free_now(off); pool.free_now(seen);
  }
  FG(cudaMemcpy(&sc, d_sc, sizeof(sc), cudaMemcpyDeviceToHost));
  if (sc.err) return kFlattenUseHost;
  mark("reference tree");

  // ---- section table (the host twin's order, rt_flatten.cpp section 4) ----------------------------------------------------
  const int tri_nodes = q_base + num_nodes_q;
  const int n_wide_obj = mh.num_nodes4;                    // the object tree's 4-wide twin (its root: mh.obj_root4)
  BlobHeader h;
  std::memset(&h, 0, sizeof(h));
  h.magic = kBlobMagic; h.version = kBlobVersion;
  h.num_nodes = tri_nodes + mh.num_nodes_obj; h.num_tris = N; h.num_meshes = num_meshes; h.num_objects = num_objects;
  h.root_is_leaf = 0;
  h.max_depth = ref_levels + 1;
  if (tree.depth > h.max_depth) h.max_depth = tree.depth;
  if (mh.max_depth > h.max_depth) h.max_depth = mh.max_depth;
  h.max_leaf = sc.max_leaf;
  h.num_nodes_ref = 1 + 2 * num_nodes_q;
  h.num_texels = num_texels;
  double extent, box_extent;
  std::memcpy(&extent, &sc.extent, sizeof(double));
  std::memcpy(&box_extent, &sc.box_extent, sizeof(double));
  h.scene_extent = (float)extent;
  {
    float be = (float)box_extent;                          // round up, then the object tree's boxes
    if ((double)be < box_extent) be = std::nextafterf(be, INFINITY);
    h.box_extent = be > mh.box_extent ? be : mh.box_extent;
  }
  size_t off = align_up(sizeof(BlobHeader), kSectionAlign);
  h.off_nodes = off;   off = align_up(off + (size_t)h.num_nodes * sizeof(NodeRec), kSectionAlign);
  h.off_filter = off;  off = align_up(off + (size_t)N * sizeof(TriFilter), kSectionAlign);
  h.off_exact = off;   off = align_up(off + (size_t)N * sizeof(TriExact), kSectionAlign);
  h.off_shade = off;   off = align_up(off + (size_t)N * sizeof(TriShade), kSectionAlign);
  h.off_meshes = off;  off = align_up(off + (size_t)num_meshes * sizeof(MeshRec), kSectionAlign);
  h.off_objects = off; off = align_up(off + (size_t)num_objects * sizeof(ObjectRec), kSectionAlign);
  h.off_texels = off;  off = align_up(off + num_texels * 4 * sizeof(float), kSectionAlign);
  h.off_maxx = off;    off = align_up(off + (size_t)num_nodes_q * sizeof(NodeMaxX), kSectionAlign);
  h.off_tri_id = off;  off = align_up(off + (size_t)N * sizeof(int32_t), kSectionAlign);
  h.off_tri_pos = off; off = align_up(off + (size_t)N * sizeof(int32_t), kSectionAlign);
  h.off_objfilt = off; off = align_up(off + (size_t)num_objects * sizeof(ObjFilter), kSectionAlign);
  h.off_obj_order = off; off = align_up(off + (size_t)mh.num_obj_tree * sizeof(int32_t), kSectionAlign);
  h.off_lin_idx = off; off = align_up(off + (size_t)mh.num_lin * sizeof(int32_t), kSectionAlign);
  h.off_nodes4 = off;  off = align_up(off + ((size_t)n_wide4 + n_wide_obj) * sizeof(Node4Rec), kSectionAlign);
  h.off_nodes8 = off;  off = align_up(off + (size_t)n_wide8 * sizeof(Node8Rec), kSectionAlign);
  h.num_nodes8 = n_wide8;
  h.num_nodes4 = n_wide4 + n_wide_obj;
  h.num_nodes4_tri = n_wide4;
  h.obj_root4 = mh.obj_root4 >= 0 ? n_wide4 + mh.obj_root4 : -1;
  h.obj_root = mh.obj_root >= 0 ? tri_nodes + mh.obj_root : -1;
  h.num_nodes_obj = mh.num_nodes_obj; h.num_obj_tree = mh.num_obj_tree; h.num_lin = mh.num_lin;
  h.q_base = q_base; h.num_nodes_q = num_nodes_q; h.topology = 0;
  h.root_max_x = bvh->bb_max_[0].x;
  h.total_bytes = off;

  // ---- 5. the blob: records written in place or copied device to device ---------------------------------------------------
  unsigned char *blob = nullptr;
  FG(pool.alloc(blob, off));
  FG(cudaMemsetAsync(blob, 0, off, 0));                    // section padding is deterministic
  tri_records_kernel<<<blocks_of(N), 256>>>(d, tree.order, extent > 1e9 ? 1 : 0, reinterpret_cast<TriFilter *>(blob + h.off_filter),
                                            reinterpret_cast<TriExact *>(blob + h.off_exact),
                                            reinterpret_cast<TriShade *>(blob + h.off_shade),
                                            reinterpret_cast<int32_t *>(blob + h.off_tri_id),
                                            reinterpret_cast<int32_t *>(blob + h.off_tri_pos), d_sc);
  FG(cudaMemcpyAsync(blob + h.off_nodes, d_trav, (size_t)q_base * sizeof(NodeRec), cudaMemcpyDeviceToDevice, 0));
  if (num_nodes_q > 0) {
    add_int_kernel<<<blocks_of(num_nodes_q), 256>>>(d_ref, num_nodes_q, q_base);   // the reference tree sits behind the traversal tree
    FG(cudaMemcpyAsync(blob + h.off_nodes + (size_t)q_base * sizeof(NodeRec), d_ref, (size_t)num_nodes_q * sizeof(NodeRec),
                       cudaMemcpyDeviceToDevice, 0));
    FG(cudaMemcpyAsync(blob + h.off_maxx, d_maxx, (size_t)num_nodes_q * sizeof(NodeMaxX), cudaMemcpyDeviceToDevice, 0));
  }
  if (n_wide4 > 0) FG(cudaMemcpyAsync(blob + h.off_nodes4, d_wide4, (size_t)n_wide4 * sizeof(Node4Rec), cudaMemcpyDeviceToDevice, 0));
  if (n_wide8 > 0) FG(cudaMemcpyAsync(blob + h.off_nodes8, d_wide8, (size_t)n_wide8 * sizeof(Node8Rec), cudaMemcpyDeviceToDevice, 0));
  FG(cudaMemcpyAsync(blob + h.off_meshes, meshes.data(), (size_t)num_meshes * sizeof(MeshRec), cudaMemcpyHostToDevice, 0));
  if (num_texels) texel_kernel<<<blocks_of((long long)num_texels), 256>>>(d_texels_in, num_texels, reinterpret_cast<float4 *>(blob + h.off_texels));
  // analytic primitives from the host twin's object-only blob: interior child references move behind the triangle trees
  std::vector<NodeRec> onodes;
  std::vector<Node4Rec> owide;
  if (num_objects > 0) {
    const unsigned char *m = mini.data();
    FG(cudaMemcpyAsync(blob + h.off_objects, m + mh.off_objects, (size_t)num_objects * sizeof(ObjectRec), cudaMemcpyHostToDevice, 0));
    FG(cudaMemcpyAsync(blob + h.off_objfilt, m + mh.off_objfilt, (size_t)num_objects * sizeof(ObjFilter), cudaMemcpyHostToDevice, 0));
    if (mh.num_obj_tree) FG(cudaMemcpyAsync(blob + h.off_obj_order, m + mh.off_obj_order, (size_t)mh.num_obj_tree * sizeof(int32_t), cudaMemcpyHostToDevice, 0));
    if (mh.num_lin) FG(cudaMemcpyAsync(blob + h.off_lin_idx, m + mh.off_lin_idx, (size_t)mh.num_lin * sizeof(int32_t), cudaMemcpyHostToDevice, 0));
    if (mh.num_nodes_obj > 0) {
      onodes.resize(mh.num_nodes_obj);
      std::memcpy(onodes.data(), m + mh.off_nodes + (size_t)mh.obj_root * sizeof(NodeRec), onodes.size() * sizeof(NodeRec));
      for (NodeRec &r : onodes) {
        if (r.cnt0 == 0) r.ref0 += tri_nodes;
        if (r.cnt1 == 0 && r.c1_half[0] >= 0.0f) r.ref1 += tri_nodes;
      }
      FG(cudaMemcpyAsync(blob + h.off_nodes + (size_t)tri_nodes * sizeof(NodeRec), onodes.data(), onodes.size() * sizeof(NodeRec),
                         cudaMemcpyHostToDevice, 0));
    }
    if (n_wide_obj > 0) {
      owide.resize(n_wide_obj);
      std::memcpy(owide.data(), m + mh.off_nodes4, owide.size() * sizeof(Node4Rec));
      for (Node4Rec &r : owide)
        for (int k = 0; k < 4; ++k)
          if (r.cnt[k] == 0 && r.ref[k] >= 0) r.ref[k] += n_wide4;
      FG(cudaMemcpyAsync(blob + h.off_nodes4 + (size_t)n_wide4 * sizeof(Node4Rec), owide.data(), owide.size() * sizeof(Node4Rec),
                         cudaMemcpyHostToDevice, 0));
    }
  }
  FG(cudaMemcpyAsync(blob, &h, sizeof(h), cudaMemcpyHostToDevice, 0));
  FG(cudaGetLastError());
  FG(cudaMemcpy(&sc, d_sc, sizeof(sc), cudaMemcpyDeviceToHost));   // also orders every copy above before `mini` etc. go away
  FG(cudaDeviceSynchronize());
  if (sc.err) return kFlattenUseHost;
  mark("blob");
  pool.release(blob);
  out.d_blob = blob; out.bytes = off; out.hdr = h;
  stats.flattener = 1;
  return RT_OK;
}

} // namespace rtb
