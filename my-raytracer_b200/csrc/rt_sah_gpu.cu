// rt_sah_gpu.cu — the traversal tree's SAH build on the GPU ("GPU builder for the traversal tree the kernel
// actually walks"; sits beside BVH::initSoA mybvh.cpp:375-406, whose own topology rt_bvh_gpu.cu reproduces).
//
// Same contract as the host builder build_sah (rt_sah.h): a permutation `order` with contiguous leaf ranges and a
// binary tree of SahNode.  The frame does not depend on the tree (hits are decided by the exact FP64 test, ties by the
// reference's primitive index), so this builder need not — and does not — produce the host builder's tree; it only has
// to produce a good one, deterministically:
//
//   phase A  level-synchronous binned SAH for nodes of more than kSmall primitives.  Per level: one binning pass over
//            the primitives (20 bins x 3 axes per node: count + box; integer min/max/add atomics only, privatised in
//            shared memory while a thread block lies inside one node), one thread per node evaluates the 57 planes from
//            the bins and picks the cheapest, a scan over the left/right flags gives every primitive its stable
//            destination, the scatter also reduces the children's centroid bounds.  Children are numbered by exclusive
//            scans over the level's nodes, so node numbers and the permutation do not depend on thread timing.
//   phase B  every subtree of at most kSmall primitives is finished by ONE thread with the host builder's exact sweep
//            (sort by centroid per axis, suffix areas, leaf rule of rt_sah.cpp) into a node block reserved by a scan.
//
// Unused slots of the reserved blocks stay empty: the flattener walks the tree from node 0 and renumbers what it reaches.
#include "rt_sah.h"
#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <string>
#include <vector>

namespace rtb {
namespace {

#ifndef RT_SAHG_BINS
#define RT_SAHG_BINS 20
#endif
#ifndef RT_SAHG_SMALL
#define RT_SAHG_SMALL 16
#endif
#ifndef RT_SAHG_LEAFMAX
#define RT_SAHG_LEAFMAX 4
#endif
#ifndef RT_SAHG_TRAV
#define RT_SAHG_TRAV 1.2f
#endif
constexpr int kBinsG = RT_SAHG_BINS;
constexpr int kSmallG = RT_SAHG_SMALL;     // subtrees of at most this many primitives: one thread, exact sweep (phase B)
constexpr int kLeafMaxG = RT_SAHG_LEAFMAX; // rt_sah.cpp kLeafMax
constexpr float kTravG = RT_SAHG_TRAV;     // rt_sah.cpp kTrav
constexpr int kDepthSwitchG = 36;  // rt_sah.cpp kDepthSwitch: plain halving below, so the depth fits the traversal stack
constexpr int kBinWords = 7;       // count, lo[3], hi[3] (floats as order-preserving unsigned)
constexpr int kNodeBinWords = 3 * kBinsG * kBinWords;

struct Work {             // one node of the current level with more than kSmall primitives
  int node, first, count, depth;
  unsigned cl[3], ch[3];  // centroid bounds, order-preserving unsigned
};
struct Split {            // decision for one Work item
  int axis, plane;        // primitives whose bin on `axis` is < plane go left; axis < 0: halve by position
  int nl;
  float cl, scale;        // bin = clamp((int)((c - cl) * scale))
  float llo[3], lhi[3], rlo[3], rhi[3];
};
struct SmallRoot { int node, first, count, depth, base; };

__host__ __device__ inline unsigned f2o(float f) {
  unsigned u;
#if defined(__CUDA_ARCH__)
  u = __float_as_uint(f);
#else
  memcpy(&u, &f, 4);
#endif
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ inline float o2f(unsigned o) { return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o); }

__device__ inline float box_area(const float *lo, const float *hi) {
  const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
  if (dx < 0.0f) return 0.0f;
  return 2.0f * (dx * dy + dy * dz + dz * dx);
}
__device__ inline int bin_of(float c, float cl, float scale) {
  int k = (int)((c - cl) * scale);
  return k < 0 ? 0 : (k >= kBinsG ? kBinsG - 1 : k);
}

// centroids (same expression as rt_sah.cpp), root bounds and root centroid bounds
__global__ void init_kernel(const float *__restrict__ lo, const float *__restrict__ hi, int n, float *__restrict__ cen,
                            int *__restrict__ order, unsigned *__restrict__ root /* lo[3] hi[3] cl[3] ch[3] */) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned v[12];
  for (int k = 0; k < 3; ++k) { v[k] = 0xffffffffu; v[3 + k] = 0u; v[6 + k] = 0xffffffffu; v[9 + k] = 0u; }
  if (i < n) {
    order[i] = i;
    for (int k = 0; k < 3; ++k) {
      const float l = lo[3 * (size_t)i + k], h = hi[3 * (size_t)i + k];
      const float c = 0.5f * l + 0.5f * h;
      cen[3 * (size_t)i + k] = c;
      v[k] = f2o(l); v[3 + k] = f2o(h); v[6 + k] = f2o(c); v[9 + k] = f2o(c);
    }
  }
  for (int k = 0; k < 12; ++k) {
    const bool is_min = (k < 3) || (k >= 6 && k < 9);
    const unsigned r = is_min ? __reduce_min_sync(0xffffffffu, v[k]) : __reduce_max_sync(0xffffffffu, v[k]);
    if ((threadIdx.x & 31) == 0) { if (is_min) atomicMin(root + k, r); else atomicMax(root + k, r); }
  }
}

__global__ void fill_int_kernel(int *p, long long n, int v) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// bins of one level: empty = count 0, lo = +max, hi = -max (as ordered unsigned: lo 0xffffffff, hi 0)
__global__ void reset_bins_kernel(unsigned *bins, long long nwords) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nwords) return;
  const int w = (int)(i % kBinWords);
  bins[i] = (w == 0) ? 0u : (w <= 3 ? 0xffffffffu : 0u);
}

// Binning pass: every primitive of an active node adds itself to one bin per axis.
__global__ void __launch_bounds__(256) bin_kernel(const float *__restrict__ lo, const float *__restrict__ hi,
                                                  const float *__restrict__ cen, const int *__restrict__ order,
                                                  const int *__restrict__ slot, const Work *__restrict__ work, int n,
                                                  unsigned *__restrict__ bins) {
  __shared__ unsigned sb[kNodeBinWords];
  const int pos = blockIdx.x * blockDim.x + threadIdx.x;
  const int first_pos = blockIdx.x * blockDim.x;
  const int last_pos = min(first_pos + (int)blockDim.x, n) - 1;
  // the whole block inside one active node?  (node ranges are contiguous in position)
  const int s0 = slot[first_pos], s1 = slot[last_pos];
  const bool uniform = s0 >= 0 && s0 == s1;
  if (uniform) {
    for (int k = threadIdx.x; k < kNodeBinWords; k += blockDim.x) {
      const int w = k % kBinWords;
      sb[k] = (w == 0) ? 0u : (w <= 3 ? 0xffffffffu : 0u);
    }
    __syncthreads();
  }
  const int s = pos < n ? slot[pos] : -1;
  if (s >= 0) {
    const Work &wk = work[s];
    const int p = order[pos];
    unsigned bl[3], bh[3];
    for (int k = 0; k < 3; ++k) { bl[k] = f2o(lo[3 * (size_t)p + k]); bh[k] = f2o(hi[3 * (size_t)p + k]); }
    unsigned *dst = uniform ? sb : bins + (size_t)s * kNodeBinWords;
    for (int ax = 0; ax < 3; ++ax) {
      const float cl = o2f(wk.cl[ax]), ch = o2f(wk.ch[ax]);
      const float ext = ch - cl;
      if (!(ext > 0.0f)) continue;
      const int b = bin_of(cen[3 * (size_t)p + ax], cl, (float)kBinsG / ext);
      unsigned *q = dst + (ax * kBinsG + b) * kBinWords;
      atomicAdd(q, 1u);
      for (int k = 0; k < 3; ++k) { atomicMin(q + 1 + k, bl[k]); atomicMax(q + 4 + k, bh[k]); }
    }
  }
  if (uniform) {
    __syncthreads();
    unsigned *g = bins + (size_t)s0 * kNodeBinWords;
    for (int k = threadIdx.x; k < kNodeBinWords; k += blockDim.x) {
      const int w = k % kBinWords;
      const unsigned v = sb[k];
      if (w == 0) { if (v) atomicAdd(g + k, v); }
      else if (w <= 3) { if (v != 0xffffffffu) atomicMin(g + k, v); }
      else if (v != 0u) atomicMax(g + k, v);
    }
  }
}

// One thread per node of the level: evaluate the 3 x 15 planes from the bins (rt_sah.cpp split(), binned branch), choose the
// cheapest, write both children (numbered base + 2 i, base + 2 i + 1) and classify them: more than kSmall primitives = a
// node of the next level, otherwise the root of a phase-B subtree.
__global__ void split_kernel(const Work *__restrict__ work, int nwork, const unsigned *__restrict__ bins, SahNode *nodes,
                             int base, Split *__restrict__ split, int *__restrict__ f_large, int *__restrict__ f_small,
                             int *__restrict__ f_reserve) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nwork) return;
  const Work wk = work[i];
  const int n = wk.count;
  Split sp;
  sp.axis = -1; sp.plane = 0; sp.nl = n / 2; sp.cl = 0.0f; sp.scale = 0.0f;
  SahNode me = nodes[wk.node];
  for (int k = 0; k < 3; ++k) { sp.llo[k] = sp.rlo[k] = me.lo[k]; sp.lhi[k] = sp.rhi[k] = me.hi[k]; } // halving: parent's box
  if (wk.depth < kDepthSwitchG) {
    const float inv_area = 1.0f / fmaxf(box_area(me.lo, me.hi), FLT_MIN);
    float best = FLT_MAX;
    const unsigned *nb = bins + (size_t)i * kNodeBinWords;
    for (int ax = 0; ax < 3; ++ax) {
      const float cl = o2f(wk.cl[ax]), ch = o2f(wk.ch[ax]);
      const float ext = ch - cl;
      if (!(ext > 0.0f)) continue;
      const unsigned *b = nb + ax * kBinsG * kBinWords;
      float ra[kBinsG];
      int rc[kBinsG];
      float rlo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, rhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
      int acc = 0;
      for (int k = kBinsG - 1; k > 0; --k) {
        const unsigned *q = b + k * kBinWords;
        if (q[0]) for (int a = 0; a < 3; ++a) { rlo[a] = fminf(rlo[a], o2f(q[1 + a])); rhi[a] = fmaxf(rhi[a], o2f(q[4 + a])); }
        acc += (int)q[0];
        ra[k] = box_area(rlo, rhi);
        rc[k] = acc;
      }
      float llo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, lhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
      acc = 0;
      for (int k = 1; k < kBinsG; ++k) {
        const unsigned *q = b + (k - 1) * kBinWords;
        if (q[0]) for (int a = 0; a < 3; ++a) { llo[a] = fminf(llo[a], o2f(q[1 + a])); lhi[a] = fmaxf(lhi[a], o2f(q[4 + a])); }
        acc += (int)q[0];
        if (acc == 0 || rc[k] == 0) continue;
        const float cost = kTravG + (box_area(llo, lhi) * (float)acc + ra[k] * (float)rc[k]) * inv_area;
        if (cost < best) { best = cost; sp.axis = ax; sp.plane = k; sp.nl = acc; sp.cl = cl; sp.scale = (float)kBinsG / ext; }
      }
    }
    if (sp.axis >= 0) {
      const unsigned *b = nb + sp.axis * kBinsG * kBinWords;
      for (int a = 0; a < 3; ++a) { sp.llo[a] = sp.rlo[a] = FLT_MAX; sp.lhi[a] = sp.rhi[a] = -FLT_MAX; }
      for (int k = 0; k < kBinsG; ++k) {
        const unsigned *q = b + k * kBinWords;
        if (!q[0]) continue;
        float *lo_ = k < sp.plane ? sp.llo : sp.rlo, *hi_ = k < sp.plane ? sp.lhi : sp.rhi;
        for (int a = 0; a < 3; ++a) { lo_[a] = fminf(lo_[a], o2f(q[1 + a])); hi_[a] = fmaxf(hi_[a], o2f(q[4 + a])); }
      }
    }
  }
  split[i] = sp;
  const int cnt[2] = {sp.nl, n - sp.nl};
  const int first[2] = {wk.first, wk.first + sp.nl};
  for (int s = 0; s < 2; ++s) {
    SahNode c;
    for (int k = 0; k < 3; ++k) { c.lo[k] = s ? sp.rlo[k] : sp.llo[k]; c.hi[k] = s ? sp.rhi[k] : sp.lhi[k]; }
    c.left = c.right = -1;
    c.first = first[s]; c.count = cnt[s];
    nodes[base + 2 * i + s] = c;
    const bool large = cnt[s] > kSmallG;
    f_large[2 * i + s] = large ? 1 : 0;
    f_small[2 * i + s] = large ? 0 : 1;
    f_reserve[2 * i + s] = large ? 0 : (cnt[s] > 2 ? 2 * cnt[s] - 2 : 0);
  }
  me.left = base + 2 * i; me.right = base + 2 * i + 1; me.count = 0;
  nodes[wk.node] = me;
}

// children -> the next level's Work items / the phase-B roots (positions from the exclusive scans: deterministic)
__global__ void emit_children_kernel(const Work *__restrict__ work, int nwork, const SahNode *__restrict__ nodes, int base,
                                     const int *__restrict__ f_large, const int *__restrict__ i_large,
                                     const int *__restrict__ i_small, const int *__restrict__ o_reserve,
                                     Work *__restrict__ next, SmallRoot *__restrict__ small, int small_before,
                                     int reserve_before, int *__restrict__ child_slot) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= 2 * nwork) return;
  const int node = base + c;
  const SahNode nd = nodes[node];
  const int depth = work[c >> 1].depth + 1;
  if (f_large[c]) {
    Work w;
    w.node = node; w.first = nd.first; w.count = nd.count; w.depth = depth;
    for (int k = 0; k < 3; ++k) { w.cl[k] = 0xffffffffu; w.ch[k] = 0u; }
    next[i_large[c]] = w;
    child_slot[c] = i_large[c];
  } else {
    SmallRoot r;
    r.node = node; r.first = nd.first; r.count = nd.count; r.depth = depth; r.base = reserve_before + o_reserve[c];
    small[small_before + i_small[c]] = r;
    child_slot[c] = -1;
  }
}

// totals of the three exclusive scans (last offset + last flag) in one 12-byte read-back per level
__global__ void totals_kernel(const int *f_large, const int *i_large, const int *f_small, const int *i_small,
                              const int *f_reserve, const int *o_reserve, int last, int *tot) {
  tot[0] = i_large[last] + f_large[last];
  tot[1] = i_small[last] + f_small[last];
  tot[2] = o_reserve[last] + f_reserve[last];
}

__global__ void flag_kernel(const float *__restrict__ cen, const int *__restrict__ order, const int *__restrict__ slot,
                            const Work *__restrict__ work, const Split *__restrict__ split, int n, int *__restrict__ flags) {
  const int pos = blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= n) return;
  const int s = slot[pos];
  int left = 0;
  if (s >= 0) {
    const Split &sp = split[s];
    if (sp.axis >= 0) left = bin_of(cen[3 * (size_t)order[pos] + sp.axis], sp.cl, sp.scale) < sp.plane;
    else left = (pos - work[s].first) < sp.nl;
  }
  flags[pos] = left;
}

// stable partition of every active node's range (left primitives first), the next level's slot of every position, and the
// centroid bounds of the next level's nodes
__global__ void scatter_kernel(const float *__restrict__ cen, const int *__restrict__ order, const int *__restrict__ slot,
                               const Work *__restrict__ work, const Split *__restrict__ split, const int *__restrict__ flags,
                               const int *__restrict__ pref, const int *__restrict__ child_slot, int n,
                               int *__restrict__ order_out, int *__restrict__ slot_out, Work *__restrict__ next) {
  const int pos = blockIdx.x * blockDim.x + threadIdx.x;
  int ns = -1, prim = 0;
  if (pos < n) {
    const int s = slot[pos];
    prim = order[pos];
    if (s < 0) {
      order_out[pos] = prim;
      slot_out[pos] = -1;
    } else {
      const Work &wk = work[s];
      const int rl = pref[pos] - pref[wk.first];
      const int left = flags[pos];
      const int np = left ? wk.first + rl : wk.first + split[s].nl + ((pos - wk.first) - rl);
      ns = child_slot[2 * s + (left ? 0 : 1)];
      order_out[np] = prim;
      slot_out[np] = ns;
    }
  }
  // centroid bounds of the next level's nodes: one atomic per warp when the whole warp feeds one node
  const int ns0 = __shfl_sync(0xffffffffu, ns, 0);
  const bool same = __all_sync(0xffffffffu, ns == ns0);
  if (same) {
    if (ns0 < 0) return;
    for (int k = 0; k < 3; ++k) {
      const unsigned v = f2o(cen[3 * (size_t)prim + k]);
      const unsigned mn = __reduce_min_sync(0xffffffffu, v), mx = __reduce_max_sync(0xffffffffu, v);
      if ((threadIdx.x & 31) == 0) { atomicMin(&next[ns0].cl[k], mn); atomicMax(&next[ns0].ch[k], mx); }
    }
  } else if (ns >= 0) {
    for (int k = 0; k < 3; ++k) {
      const unsigned v = f2o(cen[3 * (size_t)prim + k]);
      atomicMin(&next[ns].cl[k], v);
      atomicMax(&next[ns].ch[k], v);
    }
  }
}

// Phase B: one thread finishes one subtree of at most kSmall primitives with the exact sweep of rt_sah.cpp (sort by
// centroid per axis, suffix areas, cost kTrav + (A_l n_l + A_r n_r) / A; leaf when n <= 2, or n <= kLeafMax and no split
// beats the leaf cost; coinciding centroids: halve while the node is large).  Nodes go into the block reserved for the root.
__global__ void small_subtree_kernel(const SmallRoot *__restrict__ roots, int nroots, const float *__restrict__ lo,
                                     const float *__restrict__ hi, const float *__restrict__ cen, int *__restrict__ order,
                                     SahNode *__restrict__ nodes, int *__restrict__ max_depth) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nroots) return;
  const SmallRoot root = roots[r];
  int ids[kSmallG], tmp[kSmallG], best_ids[kSmallG];
  float right_area[kSmallG];
  for (int i = 0; i < root.count; ++i) ids[i] = order[root.first + i];
  struct Task { int node, b, e, depth; };
  Task stack[kSmallG + 2];
  int sp = 0, next_node = root.base, deepest = root.depth;
  stack[sp++] = Task{root.node, 0, root.count, root.depth};
  while (sp > 0) {
    const Task t = stack[--sp];
    const int n = t.e - t.b;
    if (t.depth > deepest) deepest = t.depth;
    // bounds of the node (the root's came from the bins of its parent; recomputed here: tighter or equal)
    float nlo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, nhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float cl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, ch[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = t.b; i < t.e; ++i) {
      const size_t p = 3 * (size_t)ids[i];
      for (int k = 0; k < 3; ++k) {
        nlo[k] = fminf(nlo[k], lo[p + k]); nhi[k] = fmaxf(nhi[k], hi[p + k]);
        cl[k] = fminf(cl[k], cen[p + k]); ch[k] = fmaxf(ch[k], cen[p + k]);
      }
    }
    SahNode nd;
    for (int k = 0; k < 3; ++k) { nd.lo[k] = nlo[k]; nd.hi[k] = nhi[k]; }
    nd.left = nd.right = -1;
    nd.first = root.first + t.b; nd.count = n;
    int mid = -1;                                   // split position, -1 = leaf
    if (n > 2) {
      int wide = 0;
      for (int k = 1; k < 3; ++k) if (ch[k] - cl[k] > ch[wide] - cl[wide]) wide = k;
      if (!(ch[wide] - cl[wide] > 0.0f)) {
        mid = n > 2 * kLeafMaxG ? t.b + n / 2 : -1;
      } else {
        const float inv_area = 1.0f / fmaxf(box_area(nlo, nhi), FLT_MIN);
        float best = FLT_MAX;
        int best_axis = -1, best_pos = -1;
        for (int ax = 0; ax < 3; ++ax) {
          if (!(ch[ax] - cl[ax] > 0.0f)) continue;
          for (int i = 0; i < n; ++i) tmp[i] = ids[t.b + i];
          for (int i = 1; i < n; ++i) {             // insertion sort by (centroid, primitive id)
            const int x = tmp[i];
            const float cx = cen[3 * (size_t)x + ax];
            int j = i - 1;
            while (j >= 0) {
              const float cy = cen[3 * (size_t)tmp[j] + ax];
              if (cy < cx || (cy == cx && tmp[j] < x)) break;
              tmp[j + 1] = tmp[j];
              --j;
            }
            tmp[j + 1] = x;
          }
          float rlo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, rhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
          for (int i = n - 1; i > 0; --i) {
            const size_t p = 3 * (size_t)tmp[i];
            for (int k = 0; k < 3; ++k) { rlo[k] = fminf(rlo[k], lo[p + k]); rhi[k] = fmaxf(rhi[k], hi[p + k]); }
            right_area[i] = box_area(rlo, rhi);
          }
          float llo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, lhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
          bool improved = false;
          for (int i = 1; i < n; ++i) {
            const size_t p = 3 * (size_t)tmp[i - 1];
            for (int k = 0; k < 3; ++k) { llo[k] = fminf(llo[k], lo[p + k]); lhi[k] = fmaxf(lhi[k], hi[p + k]); }
            const float cost = kTravG + (box_area(llo, lhi) * (float)i + right_area[i] * (float)(n - i)) * inv_area;
            if (cost < best) { best = cost; best_axis = ax; best_pos = i; improved = true; }
          }
          if (improved) for (int i = 0; i < n; ++i) best_ids[i] = tmp[i];
        }
        if (!(best_axis < 0 || (n <= kLeafMaxG && best >= (float)n))) {
          for (int i = 0; i < n; ++i) ids[t.b + i] = best_ids[i];
          mid = t.b + best_pos;
        }
      }
    }
    if (mid >= 0) {
      nd.left = next_node; nd.right = next_node + 1; nd.count = 0;
      stack[sp++] = Task{next_node + 1, mid, t.e, t.depth + 1};
      stack[sp++] = Task{next_node, t.b, mid, t.depth + 1};
      next_node += 2;
    }
    nodes[t.node] = nd;
  }
  for (int i = 0; i < root.count; ++i) order[root.first + i] = ids[i];
  atomicMax(max_depth, deepest);
}

#define SAHG(call)                                                                                   \
  do {                                                                                               \
    cudaError_t e_ = (call);                                                                         \
    if (e_ != cudaSuccess) { err = std::string("build_sah_device: ") + #call + ": " + cudaGetErrorString(e_); goto fail; } \
  } while (0)

} // namespace

// The build on primitive boxes that already live on the device; the tree stays there (rt_flatten_gpu.cu emits the
// records from it).  Slots of t.nodes that the build never wrote hold count = -1 (0xff fill).
bool build_sah_on_device(const float *d_lo, const float *d_hi, int n, SahDeviceTree &t, double *device_ms, std::string &err) {
  t = SahDeviceTree();
  if (n <= 0) return true;
  float *d_cen = nullptr;
  int *d_order[2] = {nullptr, nullptr}, *d_slot[2] = {nullptr, nullptr}, *d_flags = nullptr, *d_pref = nullptr;
  int *d_fl = nullptr, *d_fs = nullptr, *d_fr = nullptr, *d_il = nullptr, *d_is = nullptr, *d_or = nullptr, *d_cslot = nullptr;
  unsigned *d_bins = nullptr, *d_root = nullptr;
  int *d_depth = nullptr, *d_tot = nullptr;
  Work *d_work[2] = {nullptr, nullptr};
  Split *d_split = nullptr;
  SmallRoot *d_small = nullptr;
  SahNode *d_nodes = nullptr;
  void *d_tmp = nullptr;
  size_t tmp_bytes = 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  const size_t N = (size_t)n;
  const int max_work = n / (kSmallG + 1) + 2;          // nodes of one level with more than kSmall primitives
  const size_t cap_a = 2 * N + 2;                        // phase-A nodes in front, phase-B blocks behind
  const size_t cap_nodes = cap_a + 2 * N + 2;
  const int T = 256;
  auto blocks = [&](long long m) { return (unsigned)((m + T - 1) / T); };
  bool ok = false;
  {
    size_t t1 = 0, t2 = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, t1, (int *)nullptr, (int *)nullptr, n);
    cub::DeviceScan::ExclusiveSum(nullptr, t2, (int *)nullptr, (int *)nullptr, 2 * max_work);
    tmp_bytes = t1 > t2 ? t1 : t2;
  }
  SAHG(cudaMalloc(&d_cen, 3 * N * sizeof(float)));
  for (int k = 0; k < 2; ++k) {
    SAHG(cudaMalloc(&d_order[k], N * sizeof(int)));
    SAHG(cudaMalloc(&d_slot[k], N * sizeof(int)));
    SAHG(cudaMalloc(&d_work[k], (size_t)max_work * sizeof(Work)));
  }
  SAHG(cudaMalloc(&d_flags, N * sizeof(int)));
  SAHG(cudaMalloc(&d_pref, N * sizeof(int)));
  SAHG(cudaMalloc(&d_fl, 2 * (size_t)max_work * sizeof(int)));
  SAHG(cudaMalloc(&d_fs, 2 * (size_t)max_work * sizeof(int)));
  SAHG(cudaMalloc(&d_fr, 2 * (size_t)max_work * sizeof(int)));
  SAHG(cudaMalloc(&d_il, 2 * (size_t)max_work * sizeof(int)));
  SAHG(cudaMalloc(&d_is, 2 * (size_t)max_work * sizeof(int)));
  SAHG(cudaMalloc(&d_or, 2 * (size_t)max_work * sizeof(int)));
  SAHG(cudaMalloc(&d_cslot, 2 * (size_t)max_work * sizeof(int)));
  SAHG(cudaMalloc(&d_bins, (size_t)max_work * kNodeBinWords * sizeof(unsigned)));
  SAHG(cudaMalloc(&d_root, 12 * sizeof(unsigned)));
  SAHG(cudaMalloc(&d_depth, sizeof(int)));
  SAHG(cudaMalloc(&d_tot, 4 * sizeof(int)));
  SAHG(cudaMalloc(&d_split, (size_t)max_work * sizeof(Split)));
  SAHG(cudaMalloc(&d_small, (N + 2) * sizeof(SmallRoot)));
  SAHG(cudaMalloc(&d_nodes, cap_nodes * sizeof(SahNode)));
  SAHG(cudaMemset(d_nodes, 0xff, cap_nodes * sizeof(SahNode)));
  SAHG(cudaMalloc(&d_tmp, tmp_bytes ? tmp_bytes : 16));
  SAHG(cudaEventCreate(&e0));
  SAHG(cudaEventCreate(&e1));
  SAHG(cudaEventRecord(e0));
  {
    const unsigned init[12] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    SAHG(cudaMemcpy(d_root, init, sizeof(init), cudaMemcpyHostToDevice));
    init_kernel<<<blocks(n), T>>>(d_lo, d_hi, n, d_cen, d_order[0], d_root);
    unsigned rb[12];
    SAHG(cudaMemcpy(rb, d_root, sizeof(rb), cudaMemcpyDeviceToHost));
    auto h_o2f = [](unsigned o) { unsigned u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o; float f; memcpy(&f, &u, 4); return f; };
    SahNode root;
    for (int k = 0; k < 3; ++k) { root.lo[k] = h_o2f(rb[k]); root.hi[k] = h_o2f(rb[3 + k]); }
    root.left = root.right = -1; root.first = 0; root.count = n;
    SAHG(cudaMemcpy(d_nodes, &root, sizeof(root), cudaMemcpyHostToDevice));
    int one = 1;
    SAHG(cudaMemcpy(d_depth, &one, sizeof(int), cudaMemcpyHostToDevice));
    int nwork = 0, nsmall = 0, reserve = (int)cap_a, nodes_a = 1, cur = 0;
    if (n > kSmallG) {
      Work w;
      w.node = 0; w.first = 0; w.count = n; w.depth = 1;
      for (int k = 0; k < 3; ++k) { w.cl[k] = rb[6 + k]; w.ch[k] = rb[9 + k]; }
      SAHG(cudaMemcpy(d_work[0], &w, sizeof(w), cudaMemcpyHostToDevice));
      fill_int_kernel<<<blocks(n), T>>>(d_slot[0], n, 0);
      nwork = 1;
    } else {
      SmallRoot r{0, 0, n, 1, (int)cap_a};
      SAHG(cudaMemcpy(d_small, &r, sizeof(r), cudaMemcpyHostToDevice));
      nsmall = 1;
      reserve += 2 * n;
    }
    int level = 1;
    while (nwork > 0) {
      if ((size_t)nodes_a + 2 * (size_t)nwork > cap_a) { err = "build_sah_device: node capacity exceeded"; goto fail; }
      const long long nwords = (long long)nwork * kNodeBinWords;
      reset_bins_kernel<<<blocks(nwords), T>>>(d_bins, nwords);
      bin_kernel<<<blocks(n), T>>>(d_lo, d_hi, d_cen, d_order[cur], d_slot[cur], d_work[cur], n, d_bins);
      split_kernel<<<blocks(nwork), T>>>(d_work[cur], nwork, d_bins, d_nodes, nodes_a, d_split, d_fl, d_fs, d_fr);
      SAHG(cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_fl, d_il, 2 * nwork));
      SAHG(cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_fs, d_is, 2 * nwork));
      SAHG(cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_fr, d_or, 2 * nwork));
      totals_kernel<<<1, 1>>>(d_fl, d_il, d_fs, d_is, d_fr, d_or, 2 * nwork - 1, d_tot);
      int tot[3];
      SAHG(cudaMemcpy(tot, d_tot, sizeof(tot), cudaMemcpyDeviceToHost));
      const int n_large = tot[0], n_small = tot[1], n_res = tot[2];
      if (n_large > max_work || (size_t)nsmall + n_small > N + 2 || (size_t)reserve + n_res > cap_nodes) {
        err = "build_sah_device: work list capacity exceeded"; goto fail;
      }
      emit_children_kernel<<<blocks(2 * nwork), T>>>(d_work[cur], nwork, d_nodes, nodes_a, d_fl, d_il, d_is, d_or,
                                                     d_work[cur ^ 1], d_small, nsmall, reserve, d_cslot);
      flag_kernel<<<blocks(n), T>>>(d_cen, d_order[cur], d_slot[cur], d_work[cur], d_split, n, d_flags);
      SAHG(cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_flags, d_pref, n));
      scatter_kernel<<<blocks(n), T>>>(d_cen, d_order[cur], d_slot[cur], d_work[cur], d_split, d_flags, d_pref, d_cslot, n,
                                       d_order[cur ^ 1], d_slot[cur ^ 1], d_work[cur ^ 1]);
      nodes_a += 2 * nwork;
      nsmall += n_small;
      reserve += n_res;
      nwork = n_large;
      cur ^= 1;
      ++level;
      if (level > 4096) { err = "build_sah_device: too many levels"; goto fail; }
    }
    if (nsmall > 0)
      small_subtree_kernel<<<(unsigned)((nsmall + 127) / 128), 128>>>(d_small, nsmall, d_lo, d_hi, d_cen, d_order[cur], d_nodes, d_depth);
    SAHG(cudaGetLastError());
    SAHG(cudaEventRecord(e1));
    SAHG(cudaEventSynchronize(e1));
    if (device_ms) { float ms = 0.0f; cudaEventElapsedTime(&ms, e0, e1); *device_ms = ms; }
    int dd = 0;
    SAHG(cudaMemcpy(&dd, d_depth, sizeof(int), cudaMemcpyDeviceToHost));
    t.order = d_order[cur]; d_order[cur] = nullptr;
    t.nodes = d_nodes; d_nodes = nullptr;
    t.n = n; t.nodes_a = nodes_a; t.cap_a = cap_a; t.reserve = (size_t)reserve;
    t.depth = dd > level ? dd : level;
    ok = true;
  }
fail:
  cudaFree(d_cen);
  for (int k = 0; k < 2; ++k) { cudaFree(d_order[k]); cudaFree(d_slot[k]); cudaFree(d_work[k]); }
  cudaFree(d_flags); cudaFree(d_pref); cudaFree(d_fl); cudaFree(d_fs); cudaFree(d_fr); cudaFree(d_il); cudaFree(d_is);
  cudaFree(d_or); cudaFree(d_cslot); cudaFree(d_bins); cudaFree(d_root); cudaFree(d_depth); cudaFree(d_tot); cudaFree(d_split);
  cudaFree(d_small); cudaFree(d_nodes); cudaFree(d_tmp);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  if (!ok) cudaGetLastError();
  return ok;
}

void free_sah_device_tree(SahDeviceTree &t) {
  cudaFree(t.order); cudaFree(t.nodes);
  t = SahDeviceTree();
}

// Device build of the traversal tree from host boxes.  Same contract as build_sah; `nodes` may contain unused slots (never
// referenced).  Returns false (err set) when the device build is not possible; the caller then uses the host builder.
bool build_sah_device(const float *h_lo, const float *h_hi, int n, std::vector<int> &order, std::vector<SahNode> &nodes,
                      int &depth, double *device_ms, std::string &err) {
  order.clear(); nodes.clear(); depth = 0;
  if (n <= 0) return true;
  float *d_lo = nullptr, *d_hi = nullptr;
  SahDeviceTree t;
  const size_t N = (size_t)n;
  bool ok = false;
  SAHG(cudaMalloc(&d_lo, 3 * N * sizeof(float)));
  SAHG(cudaMalloc(&d_hi, 3 * N * sizeof(float)));
  SAHG(cudaMemcpy(d_lo, h_lo, 3 * N * sizeof(float), cudaMemcpyHostToDevice));
  SAHG(cudaMemcpy(d_hi, h_hi, 3 * N * sizeof(float), cudaMemcpyHostToDevice));
  if (!build_sah_on_device(d_lo, d_hi, n, t, device_ms, err)) goto fail;
  order.resize(N);
  SAHG(cudaMemcpy(order.data(), t.order, N * sizeof(int), cudaMemcpyDeviceToHost));
  nodes.resize(t.reserve);      // phase-A nodes [0, nodes_a), phase-B blocks [cap_a, reserve)
  SAHG(cudaMemcpy(nodes.data(), t.nodes, (size_t)t.nodes_a * sizeof(SahNode), cudaMemcpyDeviceToHost));
  if (t.reserve > t.cap_a)
    SAHG(cudaMemcpy(nodes.data() + t.cap_a, t.nodes + t.cap_a, (t.reserve - t.cap_a) * sizeof(SahNode), cudaMemcpyDeviceToHost));
  depth = t.depth;
  ok = true;
fail:
  cudaFree(d_lo); cudaFree(d_hi);
  free_sah_device_tree(t);
  if (!ok) { order.clear(); nodes.clear(); depth = 0; cudaGetLastError(); }
  return ok;
}

} // namespace rtb
