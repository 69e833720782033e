// rt_sah.cpp — binary SAH BVH builder (host).  See rt_sah.h for why the product builds its own tree.
//
// Top-down: exact sweep over the three centroid orders for small nodes, 32-bin binning for large ones;
// a node becomes a leaf when it holds <= 2 primitives, or <= kLeafMax and no split beats the leaf cost.
// Below kDepthSwitch levels the builder falls back to object-median splits so that the depth stays within
// the kernel's traversal stack.  Subtrees below the top of the tree are built by worker threads into private
// node arrays that are spliced in a fixed order: the result does not depend on the thread count.
#include "rt_sah.h"
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <numeric>
#include <thread>

namespace rtb {
namespace {

#ifndef RT_SAH_LEAFMAX
#define RT_SAH_LEAFMAX 4
#endif
#ifndef RT_SAH_TRAV
#define RT_SAH_TRAV 1.2f
#endif
constexpr int kLeafMax = RT_SAH_LEAFMAX;
constexpr int kBins = 32;
constexpr int kSweepMax = 96;     // exact sweep at or below this many primitives
constexpr int kDepthSwitch = 36;  // 36 SAH levels + log2(remaining) median levels < kStackMax (64)
constexpr float kTrav = RT_SAH_TRAV;    // cost of one two-box traversal step relative to one triangle test

struct Box {
  float lo[3], hi[3];
  void reset() {
    for (int k = 0; k < 3; ++k) lo[k] = FLT_MAX, hi[k] = -FLT_MAX;
  }
  void grow(const float *l, const float *h) {
    for (int k = 0; k < 3; ++k) {
      lo[k] = l[k] < lo[k] ? l[k] : lo[k];
      hi[k] = h[k] > hi[k] ? h[k] : hi[k];
    }
  }
  void grow(const Box &b) { grow(b.lo, b.hi); }
  float area() const {
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (dx < 0.0f) return 0.0f;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
  }
};

struct Pending {
  int placeholder; // node index in the top array
  int b, e, depth;
};

// Runs fn(begin, end, chunk) over [b, e) in `parts` chunks on as many threads (results are merged by the caller
// with order-independent operations: min/max and integer sums, so the tree does not depend on the thread count).
template <typename F> void chunked(int b, int e, int parts, F fn) {
  if (parts <= 1) { fn(b, e, 0); return; }
  std::vector<std::thread> pool;
  const long long n = e - b;
  for (int t = 0; t < parts; ++t) {
    const int lo = b + (int)(n * t / parts), hi = b + (int)(n * (t + 1) / parts);
    pool.emplace_back([=]() { fn(lo, hi, t); });
  }
  for (std::thread &t : pool) t.join();
}

constexpr int kParallelMin = 400000; // nodes at least this large are binned by several threads

struct Builder {
  const float *lo, *hi;
  std::vector<float> cen; // 3 per primitive
  int *order;
  int threads = 1;

  const float *plo(int p) const { return lo + 3 * (size_t)p; }
  const float *phi(int p) const { return hi + 3 * (size_t)p; }
  float c(int p, int ax) const { return cen[3 * (size_t)p + ax]; }

  // Chooses a split of order[b, e) and partitions it; returns the split position or -1 for "make a leaf".
  int split(int b, int e, int depth, const Box &nb) {
    const int n = e - b;
    if (n <= 2) return -1;
    float cl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, ch[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    const int parts = (n >= kParallelMin) ? threads : 1;
    {
      std::vector<Box> part(parts);
      chunked(b, e, parts, [&](int lo_, int hi_, int t) {
        Box x;
        x.reset();
        for (int i = lo_; i < hi_; ++i) {
          const float *cp = &cen[3 * (size_t)order[i]];
          x.grow(cp, cp);
        }
        part[t] = x;
      });
      for (const Box &x : part)
        for (int k = 0; k < 3; ++k) {
          cl[k] = x.lo[k] < cl[k] ? x.lo[k] : cl[k];
          ch[k] = x.hi[k] > ch[k] ? x.hi[k] : ch[k];
        }
    }
    int wide = 0;
    for (int k = 1; k < 3; ++k)
      if (ch[k] - cl[k] > ch[wide] - cl[wide]) wide = k;
    if (!(ch[wide] - cl[wide] > 0.0f)) {
      // all centroids coincide: nothing to choose between; halve by position while the node is large
      return n > 2 * kLeafMax ? b + n / 2 : -1;
    }
    if (depth >= kDepthSwitch) {
      const int mid = b + n / 2;
      std::nth_element(order + b, order + mid, order + e, [&](int x, int y) {
        const float cx = c(x, wide), cy = c(y, wide);
        return cx < cy || (cx == cy && x < y);
      });
      return mid;
    }
    const float inv_area = 1.0f / std::fmax(nb.area(), FLT_MIN);
    const float leaf_cost = (float)n;
    float best = FLT_MAX;
    int best_axis = -1, best_pos = -1;
    if (n <= kSweepMax) {
      int idx[kSweepMax];
      float right_area[kSweepMax];
      int best_idx[kSweepMax];
      for (int ax = 0; ax < 3; ++ax) {
        if (!(ch[ax] - cl[ax] > 0.0f)) continue;
        for (int i = 0; i < n; ++i) idx[i] = order[b + i];
        std::sort(idx, idx + n, [&](int x, int y) {
          const float cx = c(x, ax), cy = c(y, ax);
          return cx < cy || (cx == cy && x < y);
        });
        Box r;
        r.reset();
        for (int i = n - 1; i > 0; --i) {
          r.grow(plo(idx[i]), phi(idx[i]));
          right_area[i] = r.area();
        }
        Box l;
        l.reset();
        bool improved = false;
        for (int i = 1; i < n; ++i) {
          l.grow(plo(idx[i - 1]), phi(idx[i - 1]));
          const float cost = kTrav + (l.area() * (float)i + right_area[i] * (float)(n - i)) * inv_area;
          if (cost < best) { best = cost; best_axis = ax; best_pos = i; improved = true; }
        }
        if (improved) std::copy(idx, idx + n, best_idx);
      }
      if (best_axis < 0 || (n <= kLeafMax && best >= leaf_cost)) return -1;
      std::copy(best_idx, best_idx + n, order + b);
      return b + best_pos;
    }
    // binned
    float best_scale = 0.0f, best_lo = 0.0f;
    for (int ax = 0; ax < 3; ++ax) {
      const float ext = ch[ax] - cl[ax];
      if (!(ext > 0.0f)) continue;
      const float scale = (float)kBins / ext;
      Box bb[kBins];
      int cnt[kBins];
      for (int k = 0; k < kBins; ++k) bb[k].reset(), cnt[k] = 0;
      {
        struct Bins { Box bb[kBins]; int cnt[kBins]; };
        std::vector<Bins> part(parts);
        chunked(b, e, parts, [&](int lo_, int hi_, int t) {
          Bins &q = part[t];
          for (int k = 0; k < kBins; ++k) q.bb[k].reset(), q.cnt[k] = 0;
          for (int i = lo_; i < hi_; ++i) {
            const int p = order[i];
            int k = (int)((c(p, ax) - cl[ax]) * scale);
            k = k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
            q.bb[k].grow(plo(p), phi(p));
            ++q.cnt[k];
          }
        });
        for (const Bins &q : part)
          for (int k = 0; k < kBins; ++k) {
            if (q.cnt[k]) bb[k].grow(q.bb[k]);
            cnt[k] += q.cnt[k];
          }
      }
      float ra[kBins];
      int rc[kBins];
      Box r;
      r.reset();
      int acc = 0;
      for (int k = kBins - 1; k > 0; --k) {
        r.grow(bb[k]);
        acc += cnt[k];
        ra[k] = r.area();
        rc[k] = acc;
      }
      Box l;
      l.reset();
      acc = 0;
      for (int k = 1; k < kBins; ++k) {
        l.grow(bb[k - 1]);
        acc += cnt[k - 1];
        if (acc == 0 || rc[k] == 0) continue;
        const float cost = kTrav + (l.area() * (float)acc + ra[k] * (float)rc[k]) * inv_area;
        if (cost < best) { best = cost; best_axis = ax; best_pos = k; best_scale = scale; best_lo = cl[ax]; }
      }
    }
    if (best_axis < 0) return b + n / 2;
    int *mid = std::partition(order + b, order + e, [&](int p) {
      int k = (int)((c(p, best_axis) - best_lo) * best_scale);
      k = k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
      return k < best_pos;
    });
    const int m = (int)(mid - order);
    return (m == b || m == e) ? b + n / 2 : m;
  }

  Box bounds(int b, int e) const {
    const int parts = (e - b >= kParallelMin) ? threads : 1;
    std::vector<Box> part(parts);
    chunked(b, e, parts, [&](int lo_, int hi_, int t) {
      Box x;
      x.reset();
      for (int i = lo_; i < hi_; ++i) x.grow(plo(order[i]), phi(order[i]));
      part[t] = x;
    });
    Box x;
    x.reset();
    for (const Box &q : part) x.grow(q);
    return x;
  }

  // Recursive build into `out`; subtrees of at most `grain` primitives are deferred to `pending` when it is non-null.
  int build(std::vector<SahNode> &out, int b, int e, int depth, int &max_depth, int grain, std::vector<Pending> *pending) {
    const int me = (int)out.size();
    out.emplace_back();
    if (depth > max_depth) max_depth = depth;
    const Box nb = bounds(b, e);
    {
      SahNode &n = out[me];
      for (int k = 0; k < 3; ++k) n.lo[k] = nb.lo[k], n.hi[k] = nb.hi[k];
      n.left = n.right = -1;
      n.first = b;
      n.count = e - b;
    }
    if (pending && e - b <= grain && e - b > 2) {
      pending->push_back({me, b, e, depth});
      return me;
    }
    const int mid = split(b, e, depth, nb);
    if (mid < 0) return me;
    const int l = build(out, b, mid, depth + 1, max_depth, grain, pending);
    const int r = build(out, mid, e, depth + 1, max_depth, grain, pending);
    SahNode &n = out[me];
    n.left = l; n.right = r; n.count = 0;
    return me;
  }
};

} // namespace

void build_sah(const float *lo, const float *hi, int n, std::vector<int> &order, std::vector<SahNode> &nodes,
               int &depth) {
  order.resize(n);
  std::iota(order.begin(), order.end(), 0);
  nodes.clear();
  depth = 0;
  if (n <= 0) return;
  Builder B;
  B.lo = lo; B.hi = hi; B.order = order.data();
  B.cen.resize(3 * (size_t)n);
  for (size_t i = 0; i < 3 * (size_t)n; ++i) B.cen[i] = 0.5f * lo[i] + 0.5f * hi[i];
  unsigned hw = std::thread::hardware_concurrency();
  if (hw == 0) hw = 1;
  if (hw > 32) hw = 32;
  B.threads = (int)hw;
  if (n < 200000 || hw == 1) {
    nodes.reserve(2 * (size_t)n);
    B.build(nodes, 0, n, 1, depth, 0, nullptr);
    return;
  }
  // large scenes: build the top sequentially, the deferred subtrees in parallel, then splice in order
  std::vector<Pending> pending;
  const int grain = n / (int)(8 * hw) + 1;
  B.build(nodes, 0, n, 1, depth, grain, &pending);
  std::vector<std::vector<SahNode>> sub(pending.size());
  std::vector<int> sub_depth(pending.size(), 0);
  std::atomic<size_t> next(0);
  std::vector<std::thread> pool;
  for (unsigned t = 0; t < hw; ++t)
    pool.emplace_back([&]() {
      for (;;) {
        const size_t k = next.fetch_add(1);
        if (k >= pending.size()) return;
        const Pending &p = pending[k];
        sub[k].reserve(2 * (size_t)(p.e - p.b));
        int d = p.depth;
        B.build(sub[k], p.b, p.e, p.depth, d, 0, nullptr);
        sub_depth[k] = d;
      }
    });
  for (std::thread &t : pool) t.join();
  for (size_t k = 0; k < pending.size(); ++k) {
    if (sub_depth[k] > depth) depth = sub_depth[k];
    const int base = (int)nodes.size() - 1; // sub[k][j] (j >= 1) lands at base + j; sub[k][0] replaces the placeholder
    auto fix = [&](int child) { return child < 0 ? child : base + child; };
    SahNode root = sub[k][0];
    root.left = fix(root.left); root.right = fix(root.right);
    nodes[pending[k].placeholder] = root;
    for (size_t j = 1; j < sub[k].size(); ++j) {
      SahNode x = sub[k][j];
      x.left = fix(x.left); x.right = fix(x.right);
      nodes.push_back(x);
    }
  }
}

} // namespace rtb
