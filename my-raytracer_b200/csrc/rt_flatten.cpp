// rt_flatten.cpp — "GPU-friendly flattening of the reference's BVH" (north star (a)).
//
// Input is exactly what the reference hands its kernel: struct Data (mydata.h:28-72, filled by
// Raytracer::build_Data mytracer.cpp:166-296) and BVH::BVHNodes_SoA (mybvh.h:49-55, built by
// BVH::initSoA mybvh.cpp:375-406).  The reference never records how many node slots are in use
// (mybvh.h:57), so the tree is discovered by walking from node 0.  The split topology is kept
// as is: one 64-byte NodeRec per interior node, children in the reference's left/right order,
// leaves referencing the same [firstTriIdx_, firstTriIdx_+triCount_) ranges.
#include "rt_flatten.h"
#include <chrono>
#include <mutex>
#include <thread>
#include <cstdio>
#include <cstdlib>
#include "rt_sah.h"
#include <cmath>
#include <cstring>
#include <initializer_list>
#include <limits>
#include <type_traits>

namespace rtb {
static SahDeviceFn g_sah_device = nullptr;
static thread_local FlattenStats g_stats;
void set_sah_device_builder(SahDeviceFn fn) { g_sah_device = fn; }
const FlattenStats &last_flatten_stats() { return g_stats; }
FlattenStats &mutable_flatten_stats() { return g_stats; }
namespace {

// Contiguous ranges of [0, n) on the host threads (the flattener's per-triangle passes and section copies are
// independent per element; 10 M triangles are seconds of single-thread work otherwise).
template <typename F> void parallel_ranges(size_t n, size_t min_chunk, F &&f) {
  unsigned hw = std::thread::hardware_concurrency();
  if (hw == 0) hw = 1;
  if (hw > 32) hw = 32;
  size_t parts = n / (min_chunk ? min_chunk : 1);
  if (parts > hw) parts = hw;
  if (parts <= 1) { f((size_t)0, n); return; }
  std::vector<std::thread> th;
  const size_t step = (n + parts - 1) / parts;
  for (size_t b = 0; b < n; b += step) th.emplace_back([&f, b, step, n] { f(b, b + step < n ? b + step : n); });
  for (auto &t : th) t.join();
}
void parallel_copy(void *dst, const void *src, size_t bytes) {
  parallel_ranges(bytes, (size_t)8 << 20, [&](size_t b, size_t e) {
    std::memcpy(static_cast<unsigned char *>(dst) + b, static_cast<const unsigned char *>(src) + b, e - b);
  });
}

inline float round_down(double v) {
  float f = (float)v;
  if ((double)f > v) f = std::nextafterf(f, -std::numeric_limits<float>::infinity());
  return f;
}
inline float round_up(double v) {
  float f = (float)v;
  if ((double)f < v) f = std::nextafterf(f, std::numeric_limits<float>::infinity());
  return f;
}
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

constexpr int kObjTreeMin = 8; // spheres: below this the linear per-ray loop stays

struct Walk {
  int ref_node;  // node index in the reference arrays
  int flat_idx;  // NodeRec index assigned to it
  int depth;
};

} // namespace

int flatten_scene(const rt_data_soa *data, const rt_bvh_nodes_soa *bvh, const rt_object *objects,
                  int num_objects, unsigned flags, std::vector<unsigned char> &blob, std::string &err) {
  if (num_objects < 0 || (num_objects > 0 && !objects)) {
    err = "flatten: objects pointer/count invalid";
    return RT_ERR_INVALID_ARG;
  }
  const int N = data ? data->tVertexIdxCount_ / 3 : 0;
  const int num_meshes = data ? data->tMeshCount_ : 0;
  if (data && (data->tVertexIdxCount_ < 0 || data->tVertexIdxCount_ % 3 != 0 || num_meshes < 0)) {
    err = "flatten: tVertexIdxCount_ must be a non-negative multiple of 3";
    return RT_ERR_INVALID_ARG;
  }
  if (N > 0) {
    if (!bvh || !bvh->bb_min_ || !bvh->bb_max_ || !bvh->leftChildIdx_ || !bvh->firstTriIdx_ ||
        !bvh->triCount_) {
      err = "flatten: triangles present but BVH arrays missing";
      return RT_ERR_INVALID_ARG;
    }
    if (!data->vertexIdx_ || !data->vertexPos_ || !data->vertexMeshId_ || !data->normals_ ||
        !data->vertexNormals_ || !data->meshDrawMode_ || !data->meshTexWidth_ || !data->meshTexHeight_ ||
        !data->materialAmbient_ || !data->materialDiffuse_ || !data->materialSpecular_ ||
        !data->materialShininess_ || !data->materialMirror_ || !data->materialShadowable_) {
      err = "flatten: a required Data array is null";
      return RT_ERR_INVALID_ARG;
    }
  }

  g_stats = FlattenStats();
  // RT_FLATTEN_TIMING=1: phase times on stderr (diagnostics)
  const bool timing = std::getenv("RT_FLATTEN_TIMING") != nullptr;
  auto t_last = std::chrono::steady_clock::now();
  auto mark = [&](const char *what) {
    if (!timing) return;
    const auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[flatten] %-28s %9.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
    t_last = now;
  };
  // ---- 1. walk the reference tree, number interior nodes in DFS pre-order -------------------
  std::vector<NodeRec> nodes;
  std::vector<NodeMaxX> maxx;
  int max_depth = 0, max_leaf = 0, nodes_ref = 0, root_is_leaf = 0;
  std::vector<unsigned char> tri_seen(N > 0 ? N : 0, 0);
  const int cap = 2 * N - 1; // slots the reference allocates (mybvh.cpp:386)
  double box_extent = 0.0;
  // centre / half-extent form, outward: [ctr - half, ctr + half] (as real numbers) contains the FP64 box
  auto box_lohi = [&](const double *lo, const double *hi, float *ctr, float *half) {
    for (int k = 0; k < 3; ++k) {
      const float c = (float)(0.5 * (lo[k] + hi[k]));
      const double need = std::fmax(hi[k] - (double)c, (double)c - lo[k]); // each difference: one FP64 rounding
      float h = round_up(need > 0.0 ? need : 0.0);
      h = std::nextafterf(h, std::numeric_limits<float>::infinity());        // covers that rounding
      ctr[k] = c; half[k] = h;
      const double e = std::fabs((double)c) + (double)h;
      if (e > box_extent) box_extent = e;
    }
  };
  auto box_of = [&](int r, float *ctr, float *half) {
    const rt_vec4 &a = bvh->bb_min_[r], &b = bvh->bb_max_[r];
    const double lo[3] = {a.x, a.y, a.z}, hi[3] = {b.x, b.y, b.z};
    box_lohi(lo, hi, ctr, half);
  };
  auto leaf_ok = [&](int r) -> bool {
    const int f = bvh->firstTriIdx_[r], c = bvh->triCount_[r];
    if (f < 0 || c <= 0 || f > N - c) return false;
    for (int i = f; i < f + c; ++i) {
      if (tri_seen[i]) return false;
      tri_seen[i] = 1;
    }
    if (c > max_leaf) max_leaf = c;
    return true;
  };
  if (N > 0) {
    if (bvh->triCount_[0] > 0) {
      // the whole mesh is one leaf (N <= 2 or an unsplittable root): pseudo root with an
      // empty second child so that the kernel needs no special case
      root_is_leaf = 1;
      nodes_ref = 1;
      if (!leaf_ok(0)) { err = "flatten: root leaf range invalid"; return RT_ERR_BAD_SCENE; }
      NodeRec n;
      box_of(0, n.c0_ctr, n.c0_half);
      for (int k = 0; k < 3; ++k) n.c1_ctr[k] = 0.0f, n.c1_half[k] = -1e30f;
      n.ref0 = bvh->firstTriIdx_[0]; n.cnt0 = bvh->triCount_[0];
      n.ref1 = 0; n.cnt1 = 0;
      nodes.push_back(n);
      maxx.push_back(NodeMaxX{bvh->bb_max_[0].x, -1e300});
      max_depth = 1;
    } else {
      std::vector<Walk> todo;
      nodes.emplace_back();
      maxx.emplace_back();
      todo.push_back({0, 0, 1});
      nodes_ref = 1;
      while (!todo.empty()) {
        const Walk w = todo.back();
        todo.pop_back();
        if (w.depth > max_depth) max_depth = w.depth;
        const int l = bvh->leftChildIdx_[w.ref_node];
        if (l < 1 || l + 1 > cap - 1) {
          err = "flatten: child index out of range at node " + std::to_string(w.ref_node);
          return RT_ERR_BAD_SCENE;
        }
        nodes_ref += 2;
        if (nodes_ref > cap) { err = "flatten: BVH has a cycle or more than 2N-1 nodes"; return RT_ERR_BAD_SCENE; }
        NodeRec n;
        box_of(l, n.c0_ctr, n.c0_half);
        box_of(l + 1, n.c1_ctr, n.c1_half);
        int ref[2], cnt[2];
        // right child first onto the stack so that the LEFT subtree gets the next indices
        // (DFS pre-order: a left descent walks contiguous records)
        int child_flat[2] = {-1, -1};
        for (int s = 0; s < 2; ++s) {
          const int r = l + s;
          if (bvh->triCount_[r] > 0) {
            if (!leaf_ok(r)) {
              err = "flatten: leaf range invalid/overlapping at node " + std::to_string(r);
              return RT_ERR_BAD_SCENE;
            }
            ref[s] = bvh->firstTriIdx_[r]; cnt[s] = bvh->triCount_[r];
            if (w.depth + 1 > max_depth) max_depth = w.depth + 1;
          } else {
            child_flat[s] = (int)nodes.size();
            nodes.emplace_back();
            maxx.emplace_back();
            ref[s] = child_flat[s]; cnt[s] = 0;
          }
        }
        if (child_flat[1] >= 0) todo.push_back({l + 1, child_flat[1], w.depth + 1});
        if (child_flat[0] >= 0) todo.push_back({l, child_flat[0], w.depth + 1});
        if (cnt[0] > 0 && cnt[1] > 0 && ref[1] != ref[0] + cnt[0]) {
          // BVH::inplace_partitionSoA (mybvh.cpp:481-513) always yields adjacent sibling ranges; the
          // kernel tests two sibling leaves as ONE range
          err = "flatten: sibling leaves with non-adjacent triangle ranges at node " + std::to_string(w.ref_node);
          return RT_ERR_BAD_SCENE;
        }
        n.ref0 = ref[0]; n.cnt0 = cnt[0]; n.ref1 = ref[1]; n.cnt1 = cnt[1];
        nodes[w.flat_idx] = n;
        maxx[w.flat_idx] = NodeMaxX{bvh->bb_max_[l].x, bvh->bb_max_[l + 1].x};
      }
    }
    for (int i = 0; i < N; ++i)
      if (!tri_seen[i]) { err = "flatten: triangle " + std::to_string(i) + " is in no leaf"; return RT_ERR_BAD_SCENE; }
    if (max_depth > kStackMax) {
      err = "flatten: BVH depth " + std::to_string(max_depth) + " exceeds traversal stack " + std::to_string(kStackMax);
      return RT_ERR_LIMIT;
    }
  }

  mark("reference tree walk");
  // ---- 2. per-triangle records -------------------------------------------------------------
  std::vector<TriFilter> filt(N);
  std::vector<TriExact> exact(N);
  std::vector<TriShade> shade(N);
  double extent = 0.0;
  const int nv = data ? data->tVertexCount_ : 0;
  std::mutex err_mu;
  long long err_at = -1;          // lowest failing triangle (deterministic message whatever the thread schedule)
  int err_code = RT_OK;
  auto fail_at = [&](int i, int code, const std::string &msg) {
    std::lock_guard<std::mutex> g(err_mu);
    if (err_at < 0 || i < err_at) { err_at = i; err_code = code; err = msg; }
  };
  parallel_ranges((size_t)N, 1 << 16, [&](size_t i_begin, size_t i_end) {
  double extent_local = 0.0;
  for (int i = (int)i_begin; i < (int)i_end; ++i) {
    const int v0 = data->vertexIdx_[3 * i], v1 = data->vertexIdx_[3 * i + 1], v2 = data->vertexIdx_[3 * i + 2];
    if (v0 < 0 || v1 < 0 || v2 < 0 || v0 >= nv || v1 >= nv || v2 >= nv) {
      fail_at(i, RT_ERR_BAD_SCENE, "flatten: vertex index out of range in triangle " + std::to_string(i));
      break;
    }
    const rt_vec4 &p0 = data->vertexPos_[v0], &p1 = data->vertexPos_[v1], &p2 = data->vertexPos_[v2];
    const int mesh = data->vertexMeshId_[v0]; // as mytracer_gpu.cu:492
    if (mesh < 0 || mesh >= num_meshes) {
      fail_at(i, RT_ERR_BAD_SCENE, "flatten: vertexMeshId_ out of range in triangle " + std::to_string(i));
      break;
    }
    TriExact &e = exact[i];
    e.p0[0] = p0.x; e.p0[1] = p0.y; e.p0[2] = p0.z;
    e.p1[0] = p1.x; e.p1[1] = p1.y; e.p1[2] = p1.z;
    e.p2[0] = p2.x; e.p2[1] = p2.y; e.p2[2] = p2.z;
    e.mesh = mesh; e.pad = 0;
    const double e1[3] = {p0.x - p2.x, p0.y - p2.y, p0.z - p2.z};
    const double e2[3] = {p1.x - p2.x, p1.y - p2.y, p1.z - p2.z};
    const double a = std::sqrt(e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2]);
    const double b = std::sqrt(e2[0] * e2[0] + e2[1] * e2[1] + e2[2] * e2[2]);
    for (const rt_vec4 *p : {&p0, &p1, &p2}) {
      extent_local = std::fmax(extent_local, std::fmax(std::fabs(p->x), std::fmax(std::fabs(p->y), std::fabs(p->z))));
    }
    TriFilter &f = filt[i];
    for (int k = 0; k < 3; ++k) {
      f.p2[k] = (float)e.p2[k];
      f.e1[k] = (float)e1[k];
      f.e2[k] = (float)e2[k];
    }
    const double u = 5.9604644775390625e-8; // 2^-24
    const double tiny = 1e-30;
    f.k_bary = round_up(std::fmax(8.0 * u * std::fmax(a, b), tiny));
    f.k_t = round_up(std::fmax(8.0 * u * a * b, tiny));
    f.k_det = round_up(std::fmax(64.0 * u * a * b, tiny));

    TriShade &s = shade[i];
    std::memset(&s, 0, sizeof(s));
    s.mesh = mesh;
    if (data->meshDrawMode_[mesh] == 0) {
      const rt_vec4 &n = data->normals_[i];
      s.n[0] = n.x; s.n[1] = n.y; s.n[2] = n.z;
    } else {
      const int vs[3] = {v0, v1, v2};
      for (int k = 0; k < 3; ++k) {
        const rt_vec4 &n = data->vertexNormals_[vs[k]];
        s.n[3 * k] = n.x; s.n[3 * k + 1] = n.y; s.n[3 * k + 2] = n.z;
      }
    }
    if (data->meshTexWidth_[mesh] != -1) {
      if (!data->textureIdx_ || !data->textureCoordinatesU_ || !data->textureCoordinatesV_) {
        fail_at(i, RT_ERR_INVALID_ARG, "flatten: textured mesh but texture index/coordinate arrays are null");
        break;
      }
      bool bad_uv = false;
      for (int k = 0; k < 3; ++k) {
        const int t = data->textureIdx_[3 * i + k];
        if (t < 0 || t >= data->tTextCoordCount_) { bad_uv = true; break; }
        s.uv[2 * k] = data->textureCoordinatesU_[t];
        s.uv[2 * k + 1] = data->textureCoordinatesV_[t];
      }
      if (bad_uv) {
        fail_at(i, RT_ERR_BAD_SCENE, "flatten: texture index out of range in triangle " + std::to_string(i));
        break;
      }
    }
  }
  std::lock_guard<std::mutex> g(err_mu);
  extent = std::fmax(extent, extent_local);
  });
  if (err_at >= 0) return err_code;
  mark("triangle records");
  if (extent > 1e9) {
    // FP32 products could overflow: disable the filter (everything is confirmed in FP64)
    for (TriFilter &f : filt) f.k_det = std::numeric_limits<float>::infinity();
  }

  // 4-wide twin of a binary SAH tree (rt_types.h Node4Rec): a binary node's two children, then repeatedly the interior
  // child with the largest surface area replaced IN PLACE by its own two children (so the children stay in the binary
  // tree's left-to-right order and leaf ranges ascend) until there are four or none is interior.  Records are appended
  // to `out` in DFS pre-order; returns the root's record index.  The kernel pushes at most (interior children - 1)
  // entries per visit; if the worst root-to-leaf sum does not fit the traversal stack the twin is dropped (-1).
  auto collapse_wide = [&](const std::vector<SahNode> &sn, auto &out) -> int {
    using Rec = typename std::decay_t<decltype(out)>::value_type;   // Node4Rec or Node8Rec: same fields, W children
    constexpr int W = (int)(sizeof(Rec::ref) / sizeof(int32_t));
    struct W4 { int sah_node, rec, stack_before; };
    std::vector<W4> todo4;
    const size_t base4 = out.size();
    out.emplace_back();
    todo4.push_back({0, (int)base4, 0});
    int worst_stack = 0;
    auto area = [&](int k) {
      const SahNode &b = sn[k];
      const double dx = (double)b.hi[0] - b.lo[0], dy = (double)b.hi[1] - b.lo[1], dz = (double)b.hi[2] - b.lo[2];
      return dx * dy + dy * dz + dz * dx;
    };
    while (!todo4.empty()) {
      const W4 w = todo4.back();
      todo4.pop_back();
      int slot[W], ns = 2;
      slot[0] = sn[w.sah_node].left; slot[1] = sn[w.sah_node].right;
      while (ns < W) {
        int pick = -1;
        double best_area = -1.0;
        for (int k = 0; k < ns; ++k)
          if (sn[slot[k]].count == 0 && area(slot[k]) > best_area) { best_area = area(slot[k]); pick = k; }
        if (pick < 0) break;
        const int l = sn[slot[pick]].left, r = sn[slot[pick]].right;
        for (int k = ns; k > pick + 1; --k) slot[k] = slot[k - 1];
        slot[pick] = l; slot[pick + 1] = r;
        ++ns;
      }
      Rec n;
      int child_rec[W], interior = 0;
      for (int k = 0; k < W; ++k) child_rec[k] = -1;
      for (int k = 0; k < W; ++k) {
        float ctr[3] = {0.0f, 0.0f, 0.0f}, half[3] = {-1e30f, -1e30f, -1e30f};
        if (k < ns) {
          const SahNode &cb = sn[slot[k]];
          const double lo[3] = {cb.lo[0], cb.lo[1], cb.lo[2]}, hi[3] = {cb.hi[0], cb.hi[1], cb.hi[2]};
          box_lohi(lo, hi, ctr, half);
        }
        for (int a = 0; a < 3; ++a) n.box[k >> 1][a][k & 1] = ctr[a], n.box[k >> 1][3 + a][k & 1] = half[a];
        if (k >= ns) {
          n.ref[k] = -1; n.cnt[k] = 0;
          continue;
        }
        const SahNode &cn = sn[slot[k]];
        if (cn.count > 0) { n.ref[k] = cn.first; n.cnt[k] = cn.count; }
        else {
          child_rec[k] = (int)out.size();
          out.emplace_back();
          n.ref[k] = child_rec[k]; n.cnt[k] = 0;
          ++interior;
        }
      }
      const int stack_after = w.stack_before + (interior > 1 ? interior - 1 : 0);
      if (stack_after > worst_stack) worst_stack = stack_after;
      for (int k = W - 1; k >= 0; --k)
        if (child_rec[k] >= 0) todo4.push_back({slot[k], child_rec[k], stack_after});
      out[w.rec] = n;
    }
    if (worst_stack > kStackMax || out.size() >= ((size_t)1 << 30)) {
      out.resize(base4);
      return -1;
    }
    return (int)base4;
  };

  // ---- 2b. the traversal tree (rt_sah.h): SAH over the same triangles, reference tree kept for the quirk ----
  std::vector<int> tri_id(N), tri_pos(N);
  for (int i = 0; i < N; ++i) tri_id[i] = tri_pos[i] = i;
  int q_base = 0, topology = 1;
  std::vector<Node4Rec> wide;
  std::vector<Node8Rec> wide8;
  int wide_tri = 0, obj_root4 = -1; // records of the triangle twin (root = record 0); root of the object-tree twin
  const int num_nodes_q = (int)nodes.size();
  if (!(flags & RT_SCENE_REFERENCE_TOPOLOGY) && N > 2 && !root_is_leaf) {
    std::vector<float> plo(3 * (size_t)N), phi(3 * (size_t)N);
    parallel_ranges((size_t)N, 1 << 16, [&](size_t b, size_t e_) {
      for (size_t i = b; i < e_; ++i) {
        const TriExact &e = exact[i];
        for (int k = 0; k < 3; ++k) {
          plo[3 * i + k] = round_down(std::fmin(e.p0[k], std::fmin(e.p1[k], e.p2[k])));
          phi[3 * i + k] = round_up(std::fmax(e.p0[k], std::fmax(e.p1[k], e.p2[k])));
        }
      }
    });
    std::vector<int> order;
    std::vector<SahNode> sn;
    int sah_depth = 0;
    mark("primitive boxes");
    {
      const auto tb = std::chrono::steady_clock::now();
      bool on_device = false;
      if (g_sah_device && !(flags & RT_SCENE_HOST_SAH)) {
        std::string derr;
        on_device = g_sah_device(plo.data(), phi.data(), N, order, sn, sah_depth, &g_stats.sah_device_ms, derr);
        if (!on_device && timing) std::fprintf(stderr, "[flatten] device SAH build failed (%s): host builder\n", derr.c_str());
      }
      if (!on_device) build_sah(plo.data(), phi.data(), N, order, sn, sah_depth);
      g_stats.sah_builder = on_device ? 2 : 1;
      g_stats.sah_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tb).count();
    }
    mark("build_sah");
    if (!sn.empty() && sn[0].count == 0 && sah_depth <= kStackMax) {
      // interior nodes in DFS pre-order, as for the reference tree
      std::vector<NodeRec> trav;
      std::vector<Walk> todo;
      trav.emplace_back();
      todo.push_back({0, 0, 1});
      while (!todo.empty()) {
        const Walk w = todo.back();
        todo.pop_back();
        const SahNode &sp = sn[w.ref_node];
        const int ch[2] = {sp.left, sp.right};
        NodeRec n;
        int ref[2], cnt[2], child_flat[2] = {-1, -1};
        for (int s2 = 0; s2 < 2; ++s2) {
          const SahNode &cn = sn[ch[s2]];
          const double lo[3] = {cn.lo[0], cn.lo[1], cn.lo[2]}, hi[3] = {cn.hi[0], cn.hi[1], cn.hi[2]};
          box_lohi(lo, hi, s2 ? n.c1_ctr : n.c0_ctr, s2 ? n.c1_half : n.c0_half);
          if (cn.count > 0) {
            ref[s2] = cn.first; cnt[s2] = cn.count;
            if (cn.count > max_leaf) max_leaf = cn.count;
          } else {
            child_flat[s2] = (int)trav.size();
            trav.emplace_back();
            ref[s2] = child_flat[s2]; cnt[s2] = 0;
          }
        }
        if (child_flat[1] >= 0) todo.push_back({ch[1], child_flat[1], w.depth + 1});
        if (child_flat[0] >= 0) todo.push_back({ch[0], child_flat[0], w.depth + 1});
        n.ref0 = ref[0]; n.cnt0 = cnt[0]; n.ref1 = ref[1]; n.cnt1 = cnt[1];
        trav[w.flat_idx] = n;
      }
      if (sah_depth > max_depth) max_depth = sah_depth;
      mark("sah records + reorder");
      wide_tri = collapse_wide(sn, wide) == 0 ? (int)wide.size() : 0;
      if (wide_tri > 0 && collapse_wide(sn, wide8) != 0) wide8.clear();
      mark("collapse_wide");
      q_base = (int)trav.size();
      topology = 0;
      for (NodeRec &r : nodes) { // reference tree: interior child references move behind the traversal tree
        if (r.cnt0 == 0) r.ref0 += q_base;
        if (r.cnt1 == 0 && r.c1_half[0] >= 0.0f) r.ref1 += q_base;
      }
      trav.insert(trav.end(), nodes.begin(), nodes.end());
      nodes.swap(trav);
      // triangle records into traversal order
      std::vector<TriFilter> f2(N);
      std::vector<TriExact> e2(N);
      parallel_ranges((size_t)N, 1 << 16, [&](size_t b, size_t e_) {
        for (size_t j = b; j < e_; ++j) {
          const int i = order[j];
          tri_id[j] = i; tri_pos[i] = (int)j;      // order is a permutation: every tri_pos slot is written once
          f2[j] = filt[i]; e2[j] = exact[i];
        }
      });
      filt.swap(f2);
      exact.swap(e2);
    }
  }

  // ---- 3. meshes, objects, texels ------------------------------------------------------------
  std::vector<MeshRec> meshes(num_meshes);
  size_t num_texels = 0;
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
    if (r.tex_w != -1) {
      if (r.tex_w <= 0 || r.tex_h <= 0 || !data->meshTexels_ || !data->firstMeshTex_) {
        err = "flatten: texture of mesh " + std::to_string(m) + " invalid";
        return RT_ERR_BAD_SCENE;
      }
      r.tex_offset = data->firstMeshTex_[m];
      const size_t end = (size_t)r.tex_offset + (size_t)r.tex_w * r.tex_h;
      if (end > data->tTexelCount_) {
        err = "flatten: texture of mesh " + std::to_string(m) + " exceeds tTexelCount_";
        return RT_ERR_BAD_SCENE;
      }
      if (end > num_texels) num_texels = end;
    }
  }
  std::vector<ObjectRec> objs(num_objects);
  std::vector<ObjFilter> objf(num_objects);
  for (int k = 0; k < num_objects; ++k) {
    const rt_object &o = objects[k];
    if (o.kind != RT_SPHERE && o.kind != RT_PLANE) { err = "flatten: unknown object kind"; return RT_ERR_INVALID_ARG; }
    ObjectRec &r = objs[k];
    std::memset(&r, 0, sizeof(r));
    r.kind = o.kind;
    r.shadowable = o.material.shadowable ? 1 : 0;
    r.center[0] = o.center.x; r.center[1] = o.center.y; r.center[2] = o.center.z;
    r.normal[0] = o.normal.x; r.normal[1] = o.normal.y; r.normal[2] = o.normal.z;
    r.radius = o.radius;
    r.ambient[0] = o.material.ambient.x; r.ambient[1] = o.material.ambient.y; r.ambient[2] = o.material.ambient.z;
    r.diffuse[0] = o.material.diffuse.x; r.diffuse[1] = o.material.diffuse.y; r.diffuse[2] = o.material.diffuse.z;
    r.specular[0] = o.material.specular.x; r.specular[1] = o.material.specular.y; r.specular[2] = o.material.specular.z;
    r.shininess = o.material.shininess;
    r.mirror = o.material.mirror;
    for (int a = 0; a < 3; ++a) r.cf[a] = (float)r.center[a];
    r.rf = (float)o.radius;
    r.cabs = round_up(std::fabs(r.center[0]) + std::fabs(r.center[1]) + std::fabs(r.center[2]));
    ObjFilter &q = objf[k];
    std::memset(&q, 0, sizeof(q));
    for (int a = 0; a < 3; ++a) q.c[a] = r.cf[a];
    q.r = r.rf; q.cabs = r.cabs; q.kind = r.kind;
  }

  // ---- 3b. object tree: SAH over the spheres' boxes when the linear per-ray loop would dominate ----------
  std::vector<int32_t> obj_order, lin_idx;
  int obj_root = -1, num_nodes_obj = 0;
  {
    std::vector<int> sph;
    for (int k = 0; k < num_objects; ++k)
      if (objs[k].kind == RT_SPHERE && std::isfinite(objs[k].radius) && objs[k].radius >= 0.0) sph.push_back(k);
    const int ns = (int)sph.size();
    bool built = false;
    if (ns >= kObjTreeMin && !(flags & RT_SCENE_REFERENCE_TOPOLOGY)) {
      std::vector<float> plo(3 * (size_t)ns), phi(3 * (size_t)ns);
      for (int j = 0; j < ns; ++j) {
        const ObjectRec &o = objs[sph[j]];
        for (int a = 0; a < 3; ++a) {
          // one ulp of FP64 slack on each side before the outward FP32 rounding: covers the rounding of c -+ r
          plo[3 * (size_t)j + a] = round_down(std::nextafter(o.center[a] - o.radius, -std::numeric_limits<double>::infinity()));
          phi[3 * (size_t)j + a] = round_up(std::nextafter(o.center[a] + o.radius, std::numeric_limits<double>::infinity()));
        }
      }
      std::vector<int> order;
      std::vector<SahNode> sn;
      int depth = 0;
      build_sah(plo.data(), phi.data(), ns, order, sn, depth);
      if (!sn.empty() && sn[0].count == 0 && depth <= kStackMax) {
        const int base = (int)nodes.size();
        std::vector<NodeRec> trav;
        std::vector<Walk> todo;
        trav.emplace_back();
        todo.push_back({0, 0, 1});
        while (!todo.empty()) {
          const Walk w = todo.back();
          todo.pop_back();
          const SahNode &sp = sn[w.ref_node];
          const int ch[2] = {sp.left, sp.right};
          NodeRec n;
          int ref[2], cnt[2], child_flat[2] = {-1, -1};
          for (int s2 = 0; s2 < 2; ++s2) {
            const SahNode &cn = sn[ch[s2]];
            const double lo[3] = {cn.lo[0], cn.lo[1], cn.lo[2]}, hi[3] = {cn.hi[0], cn.hi[1], cn.hi[2]};
            box_lohi(lo, hi, s2 ? n.c1_ctr : n.c0_ctr, s2 ? n.c1_half : n.c0_half);
            if (cn.count > 0) { ref[s2] = cn.first; cnt[s2] = cn.count; }
            else {
              child_flat[s2] = (int)trav.size();
              trav.emplace_back();
              ref[s2] = base + child_flat[s2]; cnt[s2] = 0;
            }
          }
          if (child_flat[1] >= 0) todo.push_back({ch[1], child_flat[1], w.depth + 1});
          if (child_flat[0] >= 0) todo.push_back({ch[0], child_flat[0], w.depth + 1});
          n.ref0 = ref[0]; n.cnt0 = cnt[0]; n.ref1 = ref[1]; n.cnt1 = cnt[1];
          trav[w.flat_idx] = n;
        }
        if (depth > max_depth) max_depth = depth;
        obj_root = base;
        num_nodes_obj = (int)trav.size();
        obj_root4 = collapse_wide(sn, wide);
        nodes.insert(nodes.end(), trav.begin(), trav.end());
        obj_order.resize(ns);
        for (int j = 0; j < ns; ++j) obj_order[j] = sph[order[j]];
        std::vector<unsigned char> in_tree(num_objects, 0);
        for (int k : sph) in_tree[k] = 1;
        for (int k = 0; k < num_objects; ++k)
          if (!in_tree[k]) lin_idx.push_back(k);
        built = true;
      }
    }
    if (!built)
      for (int k = 0; k < num_objects; ++k) lin_idx.push_back(k);
  }

  // ---- 4. assemble -----------------------------------------------------------------------------
  BlobHeader h;
  std::memset(&h, 0, sizeof(h));
  h.magic = kBlobMagic; h.version = kBlobVersion;
  h.num_nodes = (int)nodes.size(); h.num_tris = N; h.num_meshes = num_meshes; h.num_objects = num_objects;
  h.root_is_leaf = root_is_leaf; h.max_depth = max_depth; h.max_leaf = max_leaf; h.num_nodes_ref = nodes_ref;
  h.num_texels = num_texels;
  h.scene_extent = (float)extent;
  h.box_extent = round_up(box_extent);
  mark("objects");
  size_t off = align_up(sizeof(BlobHeader), kSectionAlign);
  h.off_nodes = off;   off = align_up(off + nodes.size() * sizeof(NodeRec), kSectionAlign);
  h.off_filter = off;  off = align_up(off + filt.size() * sizeof(TriFilter), kSectionAlign);
  h.off_exact = off;   off = align_up(off + exact.size() * sizeof(TriExact), kSectionAlign);
  h.off_shade = off;   off = align_up(off + shade.size() * sizeof(TriShade), kSectionAlign);
  h.off_meshes = off;  off = align_up(off + meshes.size() * sizeof(MeshRec), kSectionAlign);
  h.off_objects = off; off = align_up(off + objs.size() * sizeof(ObjectRec), kSectionAlign);
  h.off_texels = off;  off = align_up(off + num_texels * 4 * sizeof(float), kSectionAlign);
  h.off_maxx = off;    off = align_up(off + maxx.size() * sizeof(NodeMaxX), kSectionAlign);
  h.off_tri_id = off;  off = align_up(off + tri_id.size() * sizeof(int32_t), kSectionAlign);
  h.off_tri_pos = off; off = align_up(off + tri_pos.size() * sizeof(int32_t), kSectionAlign);
  h.off_objfilt = off; off = align_up(off + objf.size() * sizeof(ObjFilter), kSectionAlign);
  h.off_obj_order = off; off = align_up(off + obj_order.size() * sizeof(int32_t), kSectionAlign);
  h.off_lin_idx = off; off = align_up(off + lin_idx.size() * sizeof(int32_t), kSectionAlign);
  h.off_nodes4 = off;  off = align_up(off + wide.size() * sizeof(Node4Rec), kSectionAlign);
  h.off_nodes8 = off;  off = align_up(off + wide8.size() * sizeof(Node8Rec), kSectionAlign);
  h.num_nodes8 = (int)wide8.size();
  h.num_nodes4 = (int)wide.size();
  h.num_nodes4_tri = wide_tri; h.obj_root4 = obj_root4;
  h.obj_root = obj_root; h.num_nodes_obj = num_nodes_obj; h.num_obj_tree = (int)obj_order.size(); h.num_lin = (int)lin_idx.size();
  h.q_base = q_base; h.num_nodes_q = num_nodes_q; h.topology = topology;
  h.root_max_x = N > 0 ? bvh->bb_max_[0].x : 0.0;
  h.total_bytes = off;
  blob.clear();
  blob.resize(off);   // value-initialised (zero): section padding is deterministic
  std::memcpy(blob.data(), &h, sizeof(h));
  if (!nodes.empty()) parallel_copy(blob.data() + h.off_nodes, nodes.data(), nodes.size() * sizeof(NodeRec));
  if (!maxx.empty()) std::memcpy(blob.data() + h.off_maxx, maxx.data(), maxx.size() * sizeof(NodeMaxX));
  if (N > 0) {
    parallel_copy(blob.data() + h.off_filter, filt.data(), filt.size() * sizeof(TriFilter));
    parallel_copy(blob.data() + h.off_exact, exact.data(), exact.size() * sizeof(TriExact));
    parallel_copy(blob.data() + h.off_shade, shade.data(), shade.size() * sizeof(TriShade));
    std::memcpy(blob.data() + h.off_tri_id, tri_id.data(), tri_id.size() * sizeof(int32_t));
    std::memcpy(blob.data() + h.off_tri_pos, tri_pos.data(), tri_pos.size() * sizeof(int32_t));
  }
  if (num_meshes > 0) std::memcpy(blob.data() + h.off_meshes, meshes.data(), meshes.size() * sizeof(MeshRec));
  if (num_objects > 0) std::memcpy(blob.data() + h.off_objects, objs.data(), objs.size() * sizeof(ObjectRec));
  if (num_objects > 0) std::memcpy(blob.data() + h.off_objfilt, objf.data(), objf.size() * sizeof(ObjFilter));
  if (!obj_order.empty()) std::memcpy(blob.data() + h.off_obj_order, obj_order.data(), obj_order.size() * sizeof(int32_t));
  if (!lin_idx.empty()) std::memcpy(blob.data() + h.off_lin_idx, lin_idx.data(), lin_idx.size() * sizeof(int32_t));
  if (!wide.empty()) parallel_copy(blob.data() + h.off_nodes4, wide.data(), wide.size() * sizeof(Node4Rec));
  if (!wide8.empty()) parallel_copy(blob.data() + h.off_nodes8, wide8.data(), wide8.size() * sizeof(Node8Rec));
  float *tx = reinterpret_cast<float *>(blob.data() + h.off_texels);
  parallel_ranges(num_texels, 1 << 18, [&](size_t b, size_t e_) {
    for (size_t t = b; t < e_; ++t) {
      const rt_vec4 &c = data->meshTexels_[t];
      tx[4 * t] = (float)c.x; tx[4 * t + 1] = (float)c.y; tx[4 * t + 2] = (float)c.z; tx[4 * t + 3] = 1.0f;
    }
  });
  mark("blob assembly");
  return RT_OK;
}

int check_blob_header(const BlobHeader &h, size_t bytes, std::string &err) {
  if (h.magic != kBlobMagic || h.version != kBlobVersion) { err = "blob: bad magic/version"; return RT_ERR_BAD_SCENE; }
  if (h.total_bytes != bytes) { err = "blob: size mismatch"; return RT_ERR_BAD_SCENE; }
  const bool counts_ok = h.num_nodes >= 0 && h.num_tris >= 0 && h.num_meshes >= 0 && h.num_objects >= 0 &&
                         h.num_nodes4 >= 0 && h.num_nodes_q >= 0 && h.q_base >= 0 && h.num_nodes_obj >= 0 &&
                         h.num_obj_tree >= 0 && h.num_lin >= 0 && h.num_nodes4_tri >= 0 && h.num_texels <= bytes / 16;
  if (!counts_ok) { err = "blob: negative record count"; return RT_ERR_BAD_SCENE; }
  // every section: inside the blob, aligned for its record type, and in the flattener's order (no overlap)
  const uint64_t offs[] = {h.off_nodes, h.off_filter, h.off_exact, h.off_shade, h.off_meshes, h.off_objects, h.off_texels,
                           h.off_maxx, h.off_tri_id, h.off_tri_pos, h.off_objfilt, h.off_obj_order, h.off_lin_idx, h.off_nodes4,
                           h.off_nodes8};
  for (uint64_t o : offs)
    if (o < sizeof(BlobHeader) || o > bytes || o % 32 != 0) { err = "blob: section offset outside the blob or misaligned"; return RT_ERR_BAD_SCENE; }
  const bool ok = h.off_nodes + (size_t)h.num_nodes * sizeof(NodeRec) <= h.off_filter &&
                  h.off_filter + (size_t)h.num_tris * sizeof(TriFilter) <= h.off_exact &&
                  h.off_exact + (size_t)h.num_tris * sizeof(TriExact) <= h.off_shade &&
                  h.off_shade + (size_t)h.num_tris * sizeof(TriShade) <= h.off_meshes &&
                  h.off_meshes + (size_t)h.num_meshes * sizeof(MeshRec) <= h.off_objects &&
                  h.off_objects + (size_t)h.num_objects * sizeof(ObjectRec) <= h.off_texels &&
                  h.off_texels + h.num_texels * 16 <= h.off_maxx &&
                  h.off_maxx + (size_t)h.num_nodes_q * sizeof(NodeMaxX) <= h.off_tri_id &&
                  h.off_tri_id + (size_t)h.num_tris * sizeof(int32_t) <= h.off_tri_pos &&
                  h.off_tri_pos + (size_t)h.num_tris * sizeof(int32_t) <= h.off_objfilt &&
                  h.off_objfilt + (size_t)h.num_objects * sizeof(ObjFilter) <= h.off_obj_order &&
                  h.off_obj_order + (size_t)h.num_obj_tree * sizeof(int32_t) <= h.off_lin_idx &&
                  h.off_lin_idx + (size_t)h.num_lin * sizeof(int32_t) <= h.off_nodes4 &&
                  h.num_nodes4_tri <= h.num_nodes4 &&
                  h.obj_root4 >= -1 && h.obj_root4 < (h.num_nodes4 > 0 ? h.num_nodes4 : 1) &&
                  h.off_nodes4 + (size_t)h.num_nodes4 * sizeof(Node4Rec) <= h.off_nodes8 && h.num_nodes8 >= 0 &&
                  h.off_nodes8 + (size_t)h.num_nodes8 * sizeof(Node8Rec) <= h.total_bytes &&
                  h.num_obj_tree + h.num_lin == h.num_objects &&
                  (h.obj_root < 0 ? h.num_nodes_obj == 0 : h.obj_root == h.q_base + h.num_nodes_q) &&
                  (size_t)h.q_base + (size_t)h.num_nodes_q + (size_t)h.num_nodes_obj == (size_t)h.num_nodes;
  if (!ok) { err = "blob: section table inconsistent"; return RT_ERR_BAD_SCENE; }
  return RT_OK;
}

int check_blob(const unsigned char *blob, size_t bytes, std::string &err) {
  if (!blob || bytes < sizeof(BlobHeader)) { err = "blob: too small"; return RT_ERR_INVALID_ARG; }
  BlobHeader h;
  std::memcpy(&h, blob, sizeof(h));
  return check_blob_header(h, bytes, err);
}

SceneView view_of_blob(const unsigned char *blob) {
  BlobHeader h;
  std::memcpy(&h, blob, sizeof(h));
  SceneView v;
  v.nodes = reinterpret_cast<const NodeRec *>(blob + h.off_nodes);
  v.filt = reinterpret_cast<const TriFilter *>(blob + h.off_filter);
  v.exact = reinterpret_cast<const TriExact *>(blob + h.off_exact);
  v.shade = reinterpret_cast<const TriShade *>(blob + h.off_shade);
  v.meshes = reinterpret_cast<const MeshRec *>(blob + h.off_meshes);
  v.objects = reinterpret_cast<const ObjectRec *>(blob + h.off_objects);
  v.objfilt = reinterpret_cast<const ObjFilter *>(blob + h.off_objfilt);
  v.obj_order = reinterpret_cast<const int32_t *>(blob + h.off_obj_order);
  v.lin_idx = reinterpret_cast<const int32_t *>(blob + h.off_lin_idx);
  v.obj_root = h.obj_root; v.num_lin = h.num_lin;
  v.tri_nodes = h.q_base + h.num_nodes_q;
  v.num_nodes4 = h.num_nodes4;
  v.obj_root4 = h.obj_root4; v.num_nodes4_tri = h.num_nodes4_tri;
  v.nodes4 = h.num_nodes4 > 0 ? reinterpret_cast<const Node4Rec *>(blob + h.off_nodes4) : nullptr;
  v.nodes8 = h.num_nodes8 > 0 ? reinterpret_cast<const Node8Rec *>(blob + h.off_nodes8) : nullptr;
  v.num_nodes8 = h.num_nodes8;
  v.texels = reinterpret_cast<const float *>(blob + h.off_texels);
  v.tex_objs = nullptr;
  v.maxx = reinterpret_cast<const NodeMaxX *>(blob + h.off_maxx);
  v.root_max_x = h.root_max_x;
  v.num_nodes = h.num_nodes; v.num_tris = h.num_tris; v.num_meshes = h.num_meshes; v.num_objects = h.num_objects;
  v.tri_id = reinterpret_cast<const int32_t *>(blob + h.off_tri_id);
  v.tri_pos = reinterpret_cast<const int32_t *>(blob + h.off_tri_pos);
  v.box_extent = h.box_extent; v.q_base = h.q_base;
  return v;
}

} // namespace rtb
