// rt_core.cuh — per-lane state machine of the hot path: primary ray -> nearest-hit traversal ->
// Phong shading with any-hit shadow rays -> iterative mirror bounces.
//
// Replaces the reference's recursive CPU trace()/lighting()/subtrace() (mytracer.cpp:510-608,
// restated trace in oracle/ref_glue.cpp) and its CUDA twin trace_device/lighting_device
// (mytracer_gpu.cu:237-312,611-669).  Design points:
//   * ONE traversal loop serves primary, shadow and reflection rays, so lanes of a warp that are
//     in different path stages still execute the same traversal code together;
//   * recursion -> loop state (depth, accumulated weight), as the north star asks;
//   * shadow rays terminate on the first confirmed occluder (any-hit);
//   * bounces stop when mirror == 0 (CPU semantics, mytracer.cpp:547), unlike the reference
//     kernel that always runs max_depth bounces;
//   * HOT / COLD state split: the traversal loop keeps only FP32 ray constants and a few ints in
//     registers (struct Hot); everything FP64 — the exact ray, the best distance, the hit point,
//     normal, colours, weights (struct Cold, 29 doubles + 6 ints = 256 B per lane) — lives in
//     shared memory, one column per thread.  That halves the register count of the v1 kernel
//     (168 -> ~80) and doubles the resident warps (profiles/r01_ncu_render_kernel_v1.md).
// The same code is compiled for the device and — for tests only — for the host.
#ifndef RTB_CORE_CUH
#define RTB_CORE_CUH
#include "rt_intersect.cuh"

#ifndef RT_SKIP_BARY
#define RT_SKIP_BARY 1
#endif
#ifndef RT_CACHE_LDIR
#define RT_CACHE_LDIR 1
#endif
// Experiment ("full-lane leaf tests"): a hit leaf range of the 4-wide step is not intersected at once
// but kept in two registers; the warp intersects the postponed leaves together once RT_LEAF_DEFER_MIN lanes hold one (or
// nobody is left to wait for).  RT_LEAF_DEFER_SPEC 1: a lane holding a leaf keeps walking (speculatively, with its stale
// distance bound); 0: it waits.  Results are identical (every candidate still gets the exact test).  0 = off (default).
#ifndef RT_LEAF_DEFER
#define RT_LEAF_DEFER 0
#endif
#ifndef RT_LEAF_DEFER_MIN
#define RT_LEAF_DEFER_MIN 16
#endif
#ifndef RT_LEAF_DEFER_SPEC
#define RT_LEAF_DEFER_SPEC 1
#endif

namespace rtb {

enum Mode : int { M_IDLE = 0, M_NEAREST = 1, M_ANYHIT = 2, M_SHADE = 3 };

struct Ctx {
  SceneView sv;
  FrameParams fp;
  // lights exactly as the reference keeps them on the device: two vec4 arrays
  // (d_lightsPos / d_lightsColor, mytracer.cpp:105-118), 4 doubles per light
  const double *light_pos;
  const double *light_col;
  FrameOut out;
  const int *pixel_list; // optional: render only these pixels (y*W+x), e.g. the supersampling pass
  int list_count;
  unsigned int *cnt;     // per-thread counters (COUNT builds) or nullptr
};

// ---- cold per-lane state: shared memory columns on the device, a plain array on the host ----------
enum ColdD { // double slots
  CD_O = 0, CD_D = 3, CD_BEST_T = 6, CD_P = 7, CD_N = 10, CD_V = 13, CD_KD = 16, CD_L = 19, CD_COL = 22,
  CD_ACC = 25, CD_WEIGHT = 28, CD_LDIR = 29 /* unit vector to the light under test */,
  CD_NUM = RT_CACHE_LDIR ? 32 : 29
};
enum ColdI { CI_PX = 0, CI_PY, CI_SAMPLE, CI_DEPTH, CI_MAT, CI_LIGHT, CI_NUM };
constexpr int kColdBytesPerLane = CD_NUM * 8 + CI_NUM * 4; // 280

template <int STRIDE> struct Cold {
  double *db; // &doubles[lane]
  int *ib;    // &ints[lane]
  long long rstride; // used when STRIDE == 0 (runtime stride: global scratch columns)
  RT_HD long long st() const { return STRIDE ? (long long)STRIDE : rstride; }
  RT_HD double &d(int k) const { return db[k * st()]; }
  RT_HD int &i(int k) const { return ib[k * st()]; }
  RT_HD D3 v(int k) const { return mk(db[k * st()], db[(k + 1) * st()], db[(k + 2) * st()]); }
  RT_HD void set(int k, const D3 &a) const {
    db[k * st()] = a.x; db[(k + 1) * st()] = a.y; db[(k + 2) * st()] = a.z;
  }
};

// ---- hot per-lane state (registers) -------------------------------------------------------------------
struct Hot {
  RayF rf;
  float thi;   // FP32 upper bound of the best distance (triangle filter culling)
  float thiS;  // thi with the slab test's relative margin (box culling), this_of()
  int best_id;
  int node;    // >= 0: interior NodeRec to visit next; < 0: traversal finished
  int sp;
  int mode;
  int found;   // any-hit result
#if RT_LEAF_DEFER
  int lf_first, lf_last; // postponed leaf range of the 4-wide step (lf_last > lf_first: one is held)
#endif
  int special; // rare-path flags, tested with ONE predicate in the hot loop:
               //   S_XPAR: direction.x == 0 exactly, the reference's 0/0 slab quirk applies (rt_types.h NodeMaxX)
               //   S_OBJ : the ray is walking the tree over the analytic spheres (before the triangle tree)
};
enum Special : int { S_XPAR = 1, S_OBJ = 2 };
// Kernel configuration bits (the template parameter named COUNT throughout): K_COUNT = instrumented build,
// K_OBJ = the scene has a tree over analytic spheres.  Scenes without one run a kernel that contains none of
// that code (its mere presence in the leaf path costs 3 % on the Office scene).
// K_WIDE = ordinary rays walk the 4-wide twins of the traversal trees (Node4Rec, trav_step4): the triangle tree's
// and, with K_OBJ, the object tree's (the host picks K_WIDE only when every tree the scene has owns a twin).
// K_WIDE8 = ordinary rays walk the 8-wide twin of the triangle tree (Node8Rec, trav_step8); never together with K_OBJ.
// K_SPL = the sub-pixel samples of a pixel are traced by neighbouring LANES (rt_kernel.cuh) instead of one after the other by
// one lane; launched for 2x2 and 4x4 sample grids.
enum KernelCfg : int { K_COUNT = 1, K_OBJ = 2, K_WIDE = 4, K_WIDE8 = 8, K_SPL = 16 };

struct Mat {
  D3 ka, ks;
  double shininess, mirror;
  int shadowable;
};

#ifndef RT_LDG
#define RT_LDG 1
#endif

struct F4 {
  float x, y, z, w;
};
struct I4 {
  int x, y, z, w;
};
// 16-byte read-only loads ("vectorised __ldg", north star (a))
RT_HD F4 ld_f4(const void *p) {
#if defined(__CUDA_ARCH__)
#if RT_LDG
  const float4 v = __ldg(reinterpret_cast<const float4 *>(p));
#else
  const float4 v = *reinterpret_cast<const float4 *>(p);
#endif
  F4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w;
  return r;
#else
  return *reinterpret_cast<const F4 *>(p);
#endif
}
RT_HD I4 ld_i4(const void *p) {
#if defined(__CUDA_ARCH__)
#if RT_LDG
  const int4 v = __ldg(reinterpret_cast<const int4 *>(p));
#else
  const int4 v = *reinterpret_cast<const int4 *>(p);
#endif
  I4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w;
  return r;
#else
  return *reinterpret_cast<const I4 *>(p);
#endif
}

// 32-byte read-only load: sm_100 has 256-bit global loads (LDG.E.256), half the load instructions and L1
// requests of two 16-byte loads for a 64-byte node record
struct W8 {
  unsigned w[8];
};
#ifndef RT_LD256
#define RT_LD256 1
#endif
RT_HD W8 ld_w8(const void *p) {
  W8 r;
#if defined(__CUDA_ARCH__) && RT_LD256
  asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7])
      : "l"(p));
#elif defined(__CUDA_ARCH__)
  const uint4 a = __ldg(reinterpret_cast<const uint4 *>(p)), b = __ldg(reinterpret_cast<const uint4 *>(p) + 1);
  r.w[0] = a.x; r.w[1] = a.y; r.w[2] = a.z; r.w[3] = a.w; r.w[4] = b.x; r.w[5] = b.y; r.w[6] = b.z; r.w[7] = b.w;
#else
  memcpy(&r, p, 32);
#endif
  return r;
}
RT_HD float w2f(unsigned u) { return __int_as_float_rt((int)u); }

template <int COUNT> RT_HD void bump(const Ctx &c, int which, unsigned n = 1) {
  if (COUNT & K_COUNT) c.cnt[which] += n;
}

RT_HD void load_mat(const SceneView &sv, int mat, Mat &m) {
  if (mat >= 0) {
    const MeshRec &r = sv.meshes[mat];
    m.ka = ld3(r.ambient); m.ks = ld3(r.specular);
    m.shininess = r.shininess; m.mirror = r.mirror; m.shadowable = r.shadowable;
  } else {
    const ObjectRec &r = sv.objects[~mat];
    m.ka = ld3(r.ambient); m.ks = ld3(r.specular);
    m.shininess = r.shininess; m.mirror = r.mirror; m.shadowable = r.shadowable;
  }
}

RT_HD D3 fetch_texel(const SceneView &sv, int mesh, const MeshRec &mr, int px, int py) {
  const uint64_t local = (uint64_t)py * (uint64_t)mr.tex_w + (uint64_t)px;
#if defined(__CUDA_ARCH__)
  const float4 t = tex1Dfetch<float4>((cudaTextureObject_t)sv.tex_objs[mesh], (int)local);
  return mk((double)t.x, (double)t.y, (double)t.z);
#else
  (void)mesh;
  const float *t = sv.texels + 4 * (mr.tex_offset + local);
  return mk((double)t[0], (double)t[1], (double)t[2]);
#endif
}

// ---------------------------------------------------------------------------------------
// ray start: Ray(origin, dir) normalises dir (oracle/shim/utils/Ray.h), then the analytic
// primitives are tested in list order (restated intersect_scene), then the BVH is entered.
// ---------------------------------------------------------------------------------------
// Rays with d.x == 0 walk the reference tree with the full-featured step (trav_step<.., true>); so that the
// common step carries none of that code (predicated-off instructions still issue: it cost 11 of 70 issue slots
// per step), such a lane's node is PARKED (encoded <= -2) while the warp runs the common loop and released for a
// second loop afterwards (rt_kernel.cuh).
RT_HD int park(int node) { return -2 - node; }
RT_HD void unpark(int &node) { if (node <= -2) node = -2 - node; }

// Enter the triangle tree: ordinary rays walk the traversal tree (records from 0), rays with d.x == 0 the
// reference's own tree (records from q_base; the same tree when q_base == 0): only there does the 0/0 slab
// quirk apply.
template <int S> RT_HD void enter_triangles(const Ctx &c, Hot &H, const Cold<S> &C) {
  const SceneView &sv = c.sv;
  H.special &= S_XPAR;
  H.sp = 0;
  H.node = (sv.tri_nodes > 0) ? (H.special ? park(sv.q_base) : 0) : -1;
  // root box, tested on entry by the reference (mybvh.cpp:157): (bb_max.x - o.x)/0 = NaN rejects it
  if (H.special && C.d(CD_O) == sv.root_max_x) H.node = -1;
}

// The ray's current tree is exhausted (node < 0).  Returns true if it moved on to the next tree.
template <int COUNT, int S> RT_HD bool next_tree(const Ctx &c, Hot &H, const Cold<S> &C) {
  if (!(COUNT & K_OBJ) || !(H.special & S_OBJ) || H.mode == M_IDLE) return false;
  if (H.found) { H.special &= S_XPAR; return false; } // any-hit ray already occluded by a sphere
  enter_triangles(c, H, C);
  return H.node != -1;
}

template <int COUNT, int S>
RT_HD void start_ray(const Ctx &c, Hot &H, const Cold<S> &C, const D3 &origin, const D3 &dir, int mode, double tmax) {
  const D3 o = origin;
  const D3 d = normalize(dir);
  C.set(CD_O, o);
  C.set(CD_D, d);
  rayf_setup(H.rf, o, d, c.sv.box_extent);
  H.mode = mode;
  H.best_id = -1;
  H.found = 0;
  H.sp = 0;
#if RT_LEAF_DEFER
  H.lf_first = H.lf_last = 0;
#endif
  double best = tmax;
  float othi = thi_of(best);
  const bool exact_only = (c.fp.flags & 1) != 0;
  const SceneView &sv = c.sv;
  for (int q = 0; q < sv.num_lin; ++q) {
    const int k = sv.lin_idx[q];
    double t;
    bump<COUNT>(c, C_OBJECT);
    // FP32 pre-filter on the 32-byte ObjFilter record (a 4-wide batched variant of this loop was measured and is
    // slower: 4.6 vs 3.2 ms on the molecule scene, the loop is latency-bound at 250 k pixels)
    const char *fp = reinterpret_cast<const char *>(sv.objfilt + k);
    const F4 f0 = ld_f4(fp);
    const I4 f1 = ld_i4(fp + 16);
    const float fc[3] = {f0.x, f0.y, f0.z};
    if (f1.y == 0 && !exact_only && !sphere_filter(H.rf, fc, f0.w, __int_as_float_rt(f1.x), othi)) continue;
    const ObjectRec &ob = sv.objects[k];
    const bool hit = (ob.kind == 0) ? sphere_exact(o, d, ld3(ob.center), ob.radius, t)
                                    : plane_exact(o, d, ld3(ob.center), ld3(ob.normal), t);
    if (hit && t < best) {
      if (mode == M_ANYHIT) { H.found = 1; H.node = -1; C.d(CD_BEST_T) = best; return; }
      best = t;
      othi = thi_of(best);
      H.best_id = sv.num_tris + k;
    }
  }
  C.d(CD_BEST_T) = best;
  H.thi = thi_of(best);
  H.thiS = this_of(H.thi);
  H.special = (d.x == 0.0) ? S_XPAR : 0;
  if ((COUNT & K_OBJ) && sv.obj_root >= 0) {
    // spheres first (the reference tests objects_ before the meshes): walk their tree, then the triangles
    H.special |= S_OBJ;
    H.node = (COUNT & K_WIDE) ? sv.obj_root4 : sv.obj_root;
  } else {
    enter_triangles(c, H, C);
  }
}

// FP64 confirmation of one filter survivor + nearest / any-hit bookkeeping.  Rare (~0.7 per ray).
// j: position in traversal order (TriExact index); the reported / tie-breaking id is tri_id[j].
template <int COUNT, int S> RT_HD void confirm_tri(const Ctx &c, Hot &H, const Cold<S> &C, int j) {
  bump<COUNT>(c, C_TRI_EXACT);
  const TriExact &e = c.sv.exact[j];
  TriHit h;
  if (!tri_exact(C.v(CD_O), C.v(CD_D), ld3(e.p0), ld3(e.p1), ld3(e.p2), h)) return;
  const double best = C.d(CD_BEST_T);
  if (H.mode == M_ANYHIT) {
    // isShadow = hit && t < light_distance && 0 < t   (mytracer.cpp:599)
    if (h.t < best) { H.found = 1; H.node = -1; }
    return;
  }
  // strict '<' keeps the first hit in DFS order (mybvh.cpp:169); leaves are visited in ascending
  // triangle index there, so "lower index wins a tie" is visit-order independent.  Analytic
  // primitives are tested before the BVH, so they win ties against triangles.
  const int i = c.sv.tri_id[j];
  const bool better = (h.t < best) || (h.t == best && H.best_id >= 0 && H.best_id < c.sv.num_tris && i < H.best_id);
  if (better) {
    C.d(CD_BEST_T) = h.t;
    H.best_id = i;
    H.thi = thi_of(h.t);
    H.thiS = this_of(H.thi);
  }
}

// Spheres [first, last) of a leaf of the object tree (positions of obj_order).  Same decisions as the linear
// loop of start_ray: exact FP64 test, strict '<', and on equal t the lower object index wins — the reference
// tests objects_ in list order, so the first of two equal hits is kept.
template <int COUNT, int S>
RT_HD void obj_leaf(const Ctx &c, Hot &H, const Cold<S> &C, int first, int last, bool exact_only) {
  const SceneView &sv = c.sv;
  for (int q = first; q < last; ++q) {
    const int k = sv.obj_order[q];
    bump<COUNT>(c, C_OBJECT);
    const char *fp = reinterpret_cast<const char *>(sv.objfilt + k);
    const F4 f0 = ld_f4(fp);
    const I4 f1 = ld_i4(fp + 16);
    const float fc[3] = {f0.x, f0.y, f0.z};
    if (!exact_only && !sphere_filter(H.rf, fc, f0.w, __int_as_float_rt(f1.x), H.thi)) continue;
    const ObjectRec &ob = sv.objects[k];
    double t;
    if (!sphere_exact(C.v(CD_O), C.v(CD_D), ld3(ob.center), ob.radius, t)) continue;
    const double best = C.d(CD_BEST_T);
    const int id = sv.num_tris + k;
    if (H.mode == M_ANYHIT) {
      if (t < best) { H.found = 1; H.node = -1; return; }
      continue;
    }
    if (t < best || (t == best && H.best_id >= sv.num_tris && id < H.best_id)) {
      C.d(CD_BEST_T) = t;
      H.best_id = id;
      H.thi = thi_of(t);
      H.thiS = this_of(H.thi);
    }
  }
}

// Triangles [first, last) of one leaf — or of two sibling leaves, whose ranges are adjacent (rt_types.h).
// Leaves of the traversal tree hold positions of the triangle records; leaves of the reference tree (walked
// by d.x == 0 rays only) hold reference primitive indices, mapped through tri_pos.
template <int COUNT, bool XP, int S> RT_HD void leaf_test(const Ctx &c, Hot &H, const Cold<S> &C, int first, int last) {
  const bool exact_only = (c.fp.flags & 1) != 0;
  if ((COUNT & K_OBJ) && (H.special & S_OBJ)) { obj_leaf<COUNT>(c, H, C, first, last, exact_only); return; }
  for (int k = first; k < last; ++k) {
    bump<COUNT>(c, C_TRI_FILTER);
    bool pass = exact_only;
    int i = k;
    if (XP && H.special) i = c.sv.tri_pos[k];
    if (!exact_only) {
      const char *p = reinterpret_cast<const char *>(c.sv.filt + i);
      const F4 a = ld_f4(p), b = ld_f4(p + 16), d = ld_f4(p + 32);
      TriFilter f;
      f.p2[0] = a.x; f.p2[1] = a.y; f.p2[2] = a.z; f.k_bary = a.w;
      f.e1[0] = b.x; f.e1[1] = b.y; f.e1[2] = b.z; f.k_t = b.w;
      f.e2[0] = d.x; f.e2[1] = d.y; f.e2[2] = d.z; f.k_det = d.w;
      pass = tri_filter(H.rf, f, H.thi);
    }
    if (pass) {
      confirm_tri<COUNT>(c, H, C, i);
      if (H.found) return;
    }
  }
}

struct alignas(8) StackEntry {
  int ref;   // NodeRec index of the postponed interior child
  float tn;  // its entry distance (box_test's tn): stale entries are culled when popped
};

// One traversal step: one 64-byte node record (four 16-byte read-only loads), two conservative
// slab tests (24 FMAs), hit leaf children intersected at once (ONE loop: sibling leaf ranges are
// adjacent), nearer interior child next, the other pushed with its entry distance.
// The kernel calls this from a WARP-UNIFORM loop (while (__any_sync(node >= 0)) if (node >= 0) step):
// the vote re-converges the warp every iteration.  A per-lane `while (node >= 0)` loop lets lanes
// drift out of phase and ran 2x slower on B200 (profiles/r01_variants.md).
template <int COUNT, bool XP, int S> RT_HD void trav_step(const Ctx &c, Hot &H, const Cold<S> &C, StackEntry *stk) {
  const char *np = reinterpret_cast<const char *>(c.sv.nodes) + (size_t)(unsigned)H.node * 64u;
  const W8 lo = ld_w8(np), hi = ld_w8(np + 32);   // two 32-byte loads per 64-byte record
  I4 r; r.x = (int)hi.w[4]; r.y = (int)hi.w[5]; r.z = (int)hi.w[6]; r.w = (int)hi.w[7];
  bump<COUNT>(c, C_NODES);
  bump<COUNT>(c, C_BOX, 2);
  // lo = c0.ctr.xyz c0.half.xyz c1.ctr.xy | hi = c1.ctr.z c1.half.xyz ref0 ref1 cnt0 cnt1
  const float b0c[3] = {w2f(lo.w[0]), w2f(lo.w[1]), w2f(lo.w[2])}, b0h[3] = {w2f(lo.w[3]), w2f(lo.w[4]), w2f(lo.w[5])};
  const float b1c[3] = {w2f(lo.w[6]), w2f(lo.w[7]), w2f(hi.w[0])}, b1h[3] = {w2f(hi.w[1]), w2f(hi.w[2]), w2f(hi.w[3])};
  float tn0, tn1;
  bool h0 = box_test(H.rf, b0c, b0h, H.thiS, tn0);
  bool h1 = box_test(H.rf, b1c, b1h, H.thiS, tn1);
  if (XP && H.special == S_XPAR) { // rare (exactly axis-parallel rays in the reference tree): reproduce its NaN rejection
    const NodeMaxX mx = c.sv.maxx[H.node - c.sv.q_base];
    const double ox = C.d(CD_O);
    h0 = h0 && !(ox == mx.c0_max_x);
    h1 = h1 && !(ox == mx.c1_max_x);
  }
  const bool l0 = h0 && r.z > 0, l1 = h1 && r.w > 0;
  if (l0 || l1) {
    leaf_test<COUNT, XP>(c, H, C, l0 ? r.x : r.y, l1 ? r.y + r.w : r.x + r.z);
    if (H.found) { H.node = -1; return; }
    h0 = h0 && !l0;
    h1 = h1 && !l1;
    // a nearer hit may have been found: the surviving interior sibling is re-checked when it is
    // visited (its own children are culled against the new thiS)
  }
  int node, sp = H.sp;
  if (h0 && h1) {
    const bool far0 = tn1 < tn0;
    stk[sp].ref = far0 ? r.x : r.y;
    stk[sp].tn = fmaxf(tn0, tn1);
    ++sp;
    node = far0 ? r.y : r.x;
  } else if (h0 || h1) {
    node = h0 ? r.x : r.y;
  } else {
    node = -1;
    while (sp > 0) {
      --sp;
      const StackEntry e = stk[sp];
      if (e.tn <= H.thiS) { node = e.ref; break; } // else: culled, a nearer hit was found meanwhile
    }
  }
  H.node = node;
  H.sp = sp;
}

// Experiment (RT_PREFETCH_PUSH, out-of-cache scenes): when a child is postponed, ask L2 for its record so that the pop
// finds it on chip.  0 = off (default: a loss while the tree is L1/L2 resident, profiles/r01_variants.md).
#ifndef RT_PREFETCH_PUSH
#define RT_PREFETCH_PUSH 0
#endif
RT_HD void prefetch_rec(const void *base, int ref, unsigned bytes) {
#if defined(__CUDA_ARCH__) && RT_PREFETCH_PUSH
  const char *p = reinterpret_cast<const char *>(base) + (size_t)(unsigned)ref * bytes;
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
  if (bytes > 128u) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128));
#else
  (void)base; (void)ref; (void)bytes;
#endif
}

// One step through the 4-wide tree (K_WIDE kernels, ordinary rays only): one 128-byte Node4Rec (four 32-byte
// loads), four independent conservative slab tests (48 FMAs as 24 FFMA2, box_pair), hit leaf children intersected at
// once, the nearest hit interior child next, the others pushed with their entry distances (culled when popped).  Half the
// dependent load -> test -> branch chains of the binary step per ray, twice the independent work inside one.
RT_HD int first_bit(unsigned m) {
#if defined(__CUDA_ARCH__)
  return __ffs((int)m) - 1;
#else
  int k = 0;
  while (!((m >> k) & 1u)) ++k;
  return k;
#endif
}
// The hit leaf children [first, last) ranges of one Node4Rec (refs r*, counts n*, mask lm), ascending; neighbours own
// adjacent ranges and are merged into one loop.
template <int COUNT, int S>
RT_HD void wide_leaves(const Ctx &c, Hot &H, const Cold<S> &C, unsigned lm, int r0, int r1, int r2, int r3, int n0,
                       int n1, int n2, int n3) {
  do {
    const int k = first_bit(lm);
    lm &= lm - 1u;
    int first = k == 0 ? r0 : (k == 1 ? r1 : (k == 2 ? r2 : r3));
    int last = first + (k == 0 ? n0 : (k == 1 ? n1 : (k == 2 ? n2 : n3)));
    if (lm & (2u << k)) { // the next child is a hit leaf as well: one loop for both
      last += (k == 0 ? n1 : (k == 1 ? n2 : n3));
      lm &= ~(2u << k);
    }
    leaf_test<COUNT, false>(c, H, C, first, last);
    if (H.found) { H.node = -1; return; }
  } while (lm);
}
template <int COUNT, int S>
RT_HD void trav_step4(const Ctx &c, Hot &H, const Cold<S> &C, StackEntry *stk) {
  const char *np = reinterpret_cast<const char *>(c.sv.nodes4) + (size_t)(unsigned)H.node * 128u;
  const W8 q0 = ld_w8(np), q1 = ld_w8(np + 32), q2 = ld_w8(np + 64), q3 = ld_w8(np + 96);
  bump<COUNT>(c, C_NODES);
  bump<COUNT>(c, C_NODES4);
  // q0 q1 q2 = the two child pairs, 12 words each (rt_types.h Node4Rec) | q3 = ref[4] cnt[4]
  float w[24];
  RT_UNROLL
  for (int k = 0; k < 8; ++k) { w[k] = w2f(q0.w[k]); w[8 + k] = w2f(q1.w[k]); w[16 + k] = w2f(q2.w[k]); }
  const int r0 = (int)q3.w[0], r1 = (int)q3.w[1], r2 = (int)q3.w[2], r3 = (int)q3.w[3];
  const int n0 = (int)q3.w[4], n1 = (int)q3.w[5], n2 = (int)q3.w[6], n3 = (int)q3.w[7];
  if (COUNT & K_COUNT) bump<COUNT>(c, C_BOX, 2u + (unsigned)(r2 >= 0) + (unsigned)(r3 >= 0));
  float t0, t1, t2, t3;
  bool h0, h1, h2, h3;
  box_pair(H.rf, w, H.thiS, t0, t1, h0, h1);
  box_pair(H.rf, w + 12, H.thiS, t2, t3, h2, h3);
  unsigned lm = (h0 && n0 > 0 ? 1u : 0u) | (h1 && n1 > 0 ? 2u : 0u) | (h2 && n2 > 0 ? 4u : 0u) | (h3 && n3 > 0 ? 8u : 0u);
  if (lm) {
    h0 = h0 && !(lm & 1u); h1 = h1 && !(lm & 2u); h2 = h2 && !(lm & 4u); h3 = h3 && !(lm & 8u);
#if RT_LEAF_DEFER
    // neighbouring leaf children own adjacent ranges: a run of hit leaves is ONE range
    const unsigned low = lm & (0u - lm);
    const bool one_run = ((lm + low) & lm) == 0u;
    if (H.lf_last > H.lf_first) {               // already holding one: intersect it now (this lane only)
      leaf_test<COUNT, false>(c, H, C, H.lf_first, H.lf_last);
      H.lf_first = H.lf_last = 0;
      if (H.found) { H.node = -1; return; }
    }
    if (one_run) {
      const int kf = first_bit(lm), kl = (lm & 8u) ? 3 : ((lm & 4u) ? 2 : ((lm & 2u) ? 1 : 0));
      H.lf_first = kf == 0 ? r0 : (kf == 1 ? r1 : (kf == 2 ? r2 : r3));
      H.lf_last = (kl == 0 ? r0 + n0 : (kl == 1 ? r1 + n1 : (kl == 2 ? r2 + n2 : r3 + n3)));
    } else {
      wide_leaves<COUNT>(c, H, C, lm, r0, r1, r2, r3, n0, n1, n2, n3);
      if (H.found) return;
    }
#else
    wide_leaves<COUNT>(c, H, C, lm, r0, r1, r2, r3, n0, n1, n2, n3);
    if (H.found) return;
#endif
    // a nearer hit may have been found: surviving interior children are re-checked when visited / popped
  }
  int node, sp = H.sp;
  const float INF = 3.0e38f;
  const float u0 = h0 ? t0 : INF, u1 = h1 ? t1 : INF, u2 = h2 ? t2 : INF, u3 = h3 ? t3 : INF;
  const float m = fminf(fminf(u0, u1), fminf(u2, u3));
  if (m < INF) {
    // nearest interior child next (first of equals); the others are pushed in child order (no branches)
    const bool k0 = u0 == m, k1 = !k0 && u1 == m, k2 = !k0 && !k1 && u2 == m;
    node = k0 ? r0 : (k1 ? r1 : (k2 ? r2 : r3));
    if (h0 && !k0) { stk[sp].ref = r0; stk[sp].tn = t0; ++sp; prefetch_rec(c.sv.nodes4, r0, 128u); }
    if (h1 && !k1) { stk[sp].ref = r1; stk[sp].tn = t1; ++sp; prefetch_rec(c.sv.nodes4, r1, 128u); }
    if (h2 && !k2) { stk[sp].ref = r2; stk[sp].tn = t2; ++sp; prefetch_rec(c.sv.nodes4, r2, 128u); }
    if (h3 && (k0 || k1 || k2)) { stk[sp].ref = r3; stk[sp].tn = t3; ++sp; prefetch_rec(c.sv.nodes4, r3, 128u); }
  } else {
    node = -1;
    while (sp > 0) {
      --sp;
      const StackEntry e = stk[sp];
      if (e.tn <= H.thiS) { node = e.ref; break; }
    }
  }
  H.node = node;
  H.sp = sp;
}

#if RT_LEAF_DEFER
// Intersect the lane's postponed leaf (called warp-wide by the kernel loop when enough lanes hold one).
template <int COUNT, int S> RT_HD void flush_leaf(const Ctx &c, Hot &H, const Cold<S> &C) {
  if (H.lf_last > H.lf_first) {
    leaf_test<COUNT, false>(c, H, C, H.lf_first, H.lf_last);
    H.lf_first = H.lf_last = 0;
    if (H.found) H.node = -1;
  }
}
#endif

// One step through the 8-wide tree (K_WIDE8 kernels, ordinary rays, triangle tree only): one 256-byte Node8Rec (eight
// 32-byte loads), eight independent conservative slab tests (48 FFMA2), hit leaf children intersected at once, the
// nearest hit interior child next, the others pushed in child order with their entry distances.  A ray of the Office scene
// needs 6 of these instead of 10 four-wide steps: fewer dependent load -> test -> vote chains, more independent
// arithmetic inside each.
RT_HD int sel8(int k, int a0, int a1, int a2, int a3, int a4, int a5, int a6, int a7) {
  const int lo = (k & 2) ? ((k & 1) ? a3 : a2) : ((k & 1) ? a1 : a0);
  const int hi = (k & 2) ? ((k & 1) ? a7 : a6) : ((k & 1) ? a5 : a4);
  return (k & 4) ? hi : lo;
}
template <int COUNT, int S>
RT_HD void trav_step8(const Ctx &c, Hot &H, const Cold<S> &C, StackEntry *stk) {
  const char *np = reinterpret_cast<const char *>(c.sv.nodes8) + (size_t)(unsigned)H.node * 256u;
  const W8 q0 = ld_w8(np), q1 = ld_w8(np + 32), q2 = ld_w8(np + 64), q3 = ld_w8(np + 96), q4 = ld_w8(np + 128),
           q5 = ld_w8(np + 160), q6 = ld_w8(np + 192), q7 = ld_w8(np + 224);
  bump<COUNT>(c, C_NODES);
  bump<COUNT>(c, C_NODES8);
  // q0..q5 = the four child pairs, 12 words each | q6 = ref[8] | q7 = cnt[8]
  float w[48];
  RT_UNROLL
  for (int k = 0; k < 8; ++k) {
    w[k] = w2f(q0.w[k]); w[8 + k] = w2f(q1.w[k]); w[16 + k] = w2f(q2.w[k]);
    w[24 + k] = w2f(q3.w[k]); w[32 + k] = w2f(q4.w[k]); w[40 + k] = w2f(q5.w[k]);
  }
  const int r0 = (int)q6.w[0], r1 = (int)q6.w[1], r2 = (int)q6.w[2], r3 = (int)q6.w[3], r4 = (int)q6.w[4],
            r5 = (int)q6.w[5], r6 = (int)q6.w[6], r7 = (int)q6.w[7];
  const int n0 = (int)q7.w[0], n1 = (int)q7.w[1], n2 = (int)q7.w[2], n3 = (int)q7.w[3], n4 = (int)q7.w[4],
            n5 = (int)q7.w[5], n6 = (int)q7.w[6], n7 = (int)q7.w[7];
  if (COUNT & K_COUNT)
    bump<COUNT>(c, C_BOX, 2u + (unsigned)(r2 >= 0) + (unsigned)(r3 >= 0) + (unsigned)(r4 >= 0) + (unsigned)(r5 >= 0) +
                              (unsigned)(r6 >= 0) + (unsigned)(r7 >= 0));
  float t0, t1, t2, t3, t4, t5, t6, t7;
  bool h0, h1, h2, h3, h4, h5, h6, h7;
  box_pair(H.rf, w, H.thiS, t0, t1, h0, h1);
  box_pair(H.rf, w + 12, H.thiS, t2, t3, h2, h3);
  box_pair(H.rf, w + 24, H.thiS, t4, t5, h4, h5);
  box_pair(H.rf, w + 36, H.thiS, t6, t7, h6, h7);
  const unsigned hm = (h0 ? 1u : 0u) | (h1 ? 2u : 0u) | (h2 ? 4u : 0u) | (h3 ? 8u : 0u) | (h4 ? 16u : 0u) | (h5 ? 32u : 0u) |
                      (h6 ? 64u : 0u) | (h7 ? 128u : 0u);
  const unsigned leafm = (n0 > 0 ? 1u : 0u) | (n1 > 0 ? 2u : 0u) | (n2 > 0 ? 4u : 0u) | (n3 > 0 ? 8u : 0u) |
                         (n4 > 0 ? 16u : 0u) | (n5 > 0 ? 32u : 0u) | (n6 > 0 ? 64u : 0u) | (n7 > 0 ? 128u : 0u);
  unsigned lm = hm & leafm;
  const unsigned im = hm & ~leafm;
  while (lm) {
    // runs of neighbouring hit leaves own adjacent triangle ranges: one loop per run
    const int k = first_bit(lm);
    const unsigned run = lm >> k;
    const int len = first_bit(~run);                       // >= 1
    const int e = k + len - 1;
    lm &= ~(((1u << len) - 1u) << k);
    const int first = sel8(k, r0, r1, r2, r3, r4, r5, r6, r7);
    const int last = sel8(e, r0, r1, r2, r3, r4, r5, r6, r7) + sel8(e, n0, n1, n2, n3, n4, n5, n6, n7);
    leaf_test<COUNT, false>(c, H, C, first, last);
    if (H.found) { H.node = -1; return; }
  }
  int node, sp = H.sp;
  if (im) {
    const float INF = 3.0e38f;
    const float u0 = (im & 1u) ? t0 : INF, u1 = (im & 2u) ? t1 : INF, u2 = (im & 4u) ? t2 : INF, u3 = (im & 8u) ? t3 : INF;
    const float u4 = (im & 16u) ? t4 : INF, u5 = (im & 32u) ? t5 : INF, u6 = (im & 64u) ? t6 : INF, u7 = (im & 128u) ? t7 : INF;
    const float m = fminf(fminf(fminf(u0, u1), fminf(u2, u3)), fminf(fminf(u4, u5), fminf(u6, u7)));
    // nearest interior child next (first of equals); the others are pushed in child order
    const unsigned eq = (u0 == m ? 1u : 0u) | (u1 == m ? 2u : 0u) | (u2 == m ? 4u : 0u) | (u3 == m ? 8u : 0u) |
                        (u4 == m ? 16u : 0u) | (u5 == m ? 32u : 0u) | (u6 == m ? 64u : 0u) | (u7 == m ? 128u : 0u);
    const unsigned pickm = (eq & im) & (0u - (eq & im));   // lowest set bit
    const unsigned rest = im & ~pickm;
    node = sel8(first_bit(pickm), r0, r1, r2, r3, r4, r5, r6, r7);
    if (rest & 1u) { stk[sp].ref = r0; stk[sp].tn = t0; ++sp; prefetch_rec(c.sv.nodes8, r0, 256u); }
    if (rest & 2u) { stk[sp].ref = r1; stk[sp].tn = t1; ++sp; prefetch_rec(c.sv.nodes8, r1, 256u); }
    if (rest & 4u) { stk[sp].ref = r2; stk[sp].tn = t2; ++sp; prefetch_rec(c.sv.nodes8, r2, 256u); }
    if (rest & 8u) { stk[sp].ref = r3; stk[sp].tn = t3; ++sp; prefetch_rec(c.sv.nodes8, r3, 256u); }
    if (rest & 16u) { stk[sp].ref = r4; stk[sp].tn = t4; ++sp; prefetch_rec(c.sv.nodes8, r4, 256u); }
    if (rest & 32u) { stk[sp].ref = r5; stk[sp].tn = t5; ++sp; prefetch_rec(c.sv.nodes8, r5, 256u); }
    if (rest & 64u) { stk[sp].ref = r6; stk[sp].tn = t6; ++sp; prefetch_rec(c.sv.nodes8, r6, 256u); }
    if (rest & 128u) { stk[sp].ref = r7; stk[sp].tn = t7; ++sp; prefetch_rec(c.sv.nodes8, r7, 256u); }
  } else {
    node = -1;
    while (sp > 0) {
      --sp;
      const StackEntry e2 = stk[sp];
      if (e2.tn <= H.thiS) { node = e2.ref; break; }
    }
  }
  H.node = node;
  H.sp = sp;
}

// host/test build: run this lane's ray to completion
template <int COUNT, int S> RT_HD void traverse(const Ctx &c, Hot &H, const Cold<S> &C, StackEntry *stk) {
  do {
    if (COUNT & K_WIDE8) {
      while (H.node >= 0) trav_step8<COUNT>(c, H, C, stk);
    } else if (COUNT & K_WIDE) {
      while (H.node >= 0) trav_step4<COUNT>(c, H, C, stk); // ordinary rays; d.x == 0 rays are parked (node <= -2)
    }
    unpark(H.node);
    while (H.node >= 0) trav_step<COUNT, true>(c, H, C, stk);
  } while (next_tree<COUNT>(c, H, C));
}

// ---------------------------------------------------------------------------------------
// path logic (cold: FP64, runs once per ray, not per node)
// ---------------------------------------------------------------------------------------
// Camera::primary_ray direction (unnormalised) of the lane's pixel and sub-pixel sample (CI_PX, CI_PY, CI_SAMPLE)
template <int S> RT_HD D3 primary_dir(const Ctx &c, const Cold<S> &C) {
  const CameraRec &cam = c.fp.cam;
  double x = (double)C.i(CI_PX), y = (double)C.i(CI_PY);
  const int s = c.fp.spp_grid;
  if (s > 1) {
    // sub-pixel grid, offsets as mytracer_gpu.cu:203-207: si/s - 0.5 + 1/(2s)
    const int sample = C.i(CI_SAMPLE);
    const int si = sample / s, sj = sample % s;
    const double xo = dadd(dsub(ddiv((double)si, (double)s), 0.5), ddiv(1.0, dmul(2.0, (double)s)));
    const double yo = dadd(dsub(ddiv((double)sj, (double)s), 0.5), ddiv(1.0, dmul(2.0, (double)s)));
    x = dadd(x, xo);
    y = dadd(y, yo);
  }
  // Camera::primary_ray: Ray(eye, lower_left + x*x_dir + y*y_dir - eye)
  return sub(add(add(ld3(cam.lower_left), scale(x, ld3(cam.x_dir))), scale(y, ld3(cam.y_dir))), ld3(cam.eye));
}

// A ray the path logic wants traced next.  In the kernels for scenes with analytic primitives advance() collects the
// request of whichever branch ran and calls start_ray at ONE place: start_ray (normalisation, FP32 ray set-up, the linear
// list of analytic primitives) is the largest piece of the path logic there.
struct RayReq {
  D3 o, d;
  int mode;
  double tmax;
};

template <int COUNT, int S> RT_HD void prep_primary(const Ctx &c, const Cold<S> &C, RayReq &r) {
  r.o = ld3(c.fp.cam.eye);
  r.d = primary_dir(c, C);
  r.mode = M_NEAREST;
  r.tmax = DBL_MAX;
  C.i(CI_DEPTH) = 0;
  C.d(CD_WEIGHT) = 1.0;
  C.set(CD_COL, mk(0.0, 0.0, 0.0));
  bump<COUNT>(c, C_RAY_PRIMARY);
}

template <int COUNT, int S> RT_HD void start_primary(const Ctx &c, Hot &H, const Cold<S> &C) {
  RayReq r;
  prep_primary<COUNT>(c, C, r);
  start_ray<COUNT>(c, H, C, r.o, r.d, r.mode, r.tmax);
}

template <int S> RT_HD void write_pixel(const Ctx &c, const Cold<S> &C) {
  D3 col = C.v(CD_ACC);
  const int s2 = c.fp.spp_grid * c.fp.spp_grid;
  if (s2 > 1) col = divs(col, (double)s2);
  col = mk(dmin1(col.x), dmin1(col.y), dmin1(col.z)); // min(color, vec4(1,1,1)) mytracer_gpu.cu:155
  const size_t p = (size_t)C.i(CI_PY) * (size_t)c.fp.cam.width + (size_t)C.i(CI_PX);
  if (c.out.pixels) {
    double *q = c.out.pixels + 4 * p;
    q[0] = col.x; q[1] = col.y; q[2] = col.z; q[3] = 0.0;
  }
  if (c.out.rgba8) {
    // 8-bit quantisation (build-defined): (int)(255*clamp(c,0,1) + 0.5)
    const double r = col.x < 0.0 ? 0.0 : col.x, g = col.y < 0.0 ? 0.0 : col.y, b = col.z < 0.0 ? 0.0 : col.z;
    const unsigned ur = (unsigned)(int)dadd(dmul(255.0, r), 0.5);
    const unsigned ug = (unsigned)(int)dadd(dmul(255.0, g), 0.5);
    const unsigned ub = (unsigned)(int)dadd(dmul(255.0, b), 0.5);
    const uint32_t packed = (ur & 255u) | ((ug & 255u) << 8) | ((ub & 255u) << 16) | 0xff000000u;
    reinterpret_cast<uint32_t *>(c.out.rgba8)[p] = packed;
  }
}

// The sample's path has ended.  Returns true (request in r) when the lane goes on with the next sub-pixel sample.
template <int COUNT, bool DEFER, int S> RT_HD bool sample_done(const Ctx &c, Hot &H, const Cold<S> &C, RayReq &r) {
  if (COUNT & K_SPL) {   // the sample stays in CD_COL until the warp sums the samples of each pixel (rt_kernel.cuh)
    H.mode = M_IDLE;
    H.node = -1;
    return false;
  }
  C.set(CD_ACC, add(C.v(CD_ACC), C.v(CD_COL)));
  const int sample = C.i(CI_SAMPLE) + 1;
  C.i(CI_SAMPLE) = sample;
  if (sample < c.fp.spp_grid * c.fp.spp_grid) {
    prep_primary<COUNT>(c, C, r);
    if (DEFER) return true;
    start_ray<COUNT>(c, H, C, r.o, r.d, r.mode, r.tmax);
    return false;
  }
  write_pixel(c, C);
  H.mode = M_IDLE;
  H.node = -1;
  return false;
}

// Fill P, N, V, kd, mat for the nearest hit and start the lighting loop (L = ambient term).
template <int COUNT, int S> RT_HD void hit_setup(const Ctx &c, Hot &H, const Cold<S> &C, Mat &m) {
  const SceneView &sv = c.sv;
  const D3 o = C.v(CD_O), d = C.v(CD_D);
  const double best = C.d(CD_BEST_T);
  C.set(CD_V, neg(d));
  D3 P, N, kd;
  int mat;
  if (H.best_id < sv.num_tris) {
    const int i = H.best_id;
    const TriExact &e = sv.exact[sv.tri_pos[i]];
    TriHit h;
    h.t = best; h.alpha = 0.0; h.beta = 0.0; h.gamma = 0.0;
    const TriShade &s = sv.shade[i];
    const MeshRec &mr = sv.meshes[s.mesh];
    // barycentrics are needed for interpolated normals and texture coordinates only; the accepted test's t
    // is CD_BEST_T bit for bit (same arithmetic), so flat untextured meshes skip the FP64 recomputation
    if (!RT_SKIP_BARY || mr.draw_mode != 0 || mr.tex_w != -1)
      tri_exact(o, d, ld3(e.p0), ld3(e.p1), ld3(e.p2), h); // same arithmetic as the accepted test
    mat = s.mesh;
    P = along(o, h.t, d);                                   // ray(t), mymesh.cpp:219
    kd = ld3(mr.diffuse);
    if (mr.tex_w != -1) {
      // Mesh::compute_texture, mymesh.cpp:70-95
      double u = dadd(dadd(dmul(h.alpha, s.uv[0]), dmul(h.beta, s.uv[2])), dmul(h.gamma, s.uv[4]));
      double v = dadd(dadd(dmul(h.alpha, s.uv[1]), dmul(h.beta, s.uv[3])), dmul(h.gamma, s.uv[5]));
      u = (u < 0.0) ? 0.0 : ((1.0 < u) ? 1.0 : u);
      v = (v < 0.0) ? 0.0 : ((1.0 < v) ? 1.0 : v);
      const int tx = (int)round(dmul(u, (double)(unsigned)(mr.tex_w - 1)));
      const int ty = (int)round(dmul(dsub(1.0, v), (double)(unsigned)(mr.tex_h - 1)));
      kd = fetch_texel(sv, s.mesh, mr, tx, ty);
      bump<COUNT>(c, C_TEXEL);
    }
    if (mr.draw_mode == 0) {
      N = ld3(s.n);                                          // stored face normal, mymesh.cpp:228-229
    } else {
      // alpha*n0 + beta*n1 + gamma*n2, NOT re-normalised (mymesh.cpp:231-233)
      N = add(add(scale(h.alpha, ld3(s.n)), scale(h.beta, ld3(s.n + 3))), scale(h.gamma, ld3(s.n + 6)));
    }
  } else {
    const int k = H.best_id - sv.num_tris;
    const ObjectRec &ob = sv.objects[k];
    mat = ~k;
    P = along(o, best, d);
    kd = ld3(ob.diffuse);
    N = (ob.kind == 0) ? div3(sub(P, ld3(ob.center)), ob.radius) : ld3(ob.normal);
  }
  C.set(CD_P, P);
  C.set(CD_N, N);
  C.set(CD_KD, kd);
  C.i(CI_MAT) = mat;
  load_mat(sv, mat, m);
  // ambient term, mytracer.cpp:574-576
  C.set(CD_L, mk(dadd(0.0, dmul(c.fp.ambience[0], m.ka.x)), dadd(0.0, dmul(c.fp.ambience[1], m.ka.y)),
                 dadd(0.0, dmul(c.fp.ambience[2], m.ka.z))));
  C.i(CI_LIGHT) = 0;
  bump<COUNT>(c, C_SHADED);
}

// diffuse + specular of one unshadowed light, mytracer.cpp:510-534,581-585,603-605
// cached: CD_LDIR holds normalize(light - P) of light k (set when its shadow ray was cast)
template <int S> RT_HD void add_light(const Ctx &c, const Cold<S> &C, const Mat &m, int k, bool cached) {
  const double *lcol = c.light_col + 4 * k;
  const D3 N = C.v(CD_N);
  D3 l;
  if (RT_CACHE_LDIR && cached) {
    l = C.v(CD_LDIR);
  } else {
    const D3 lv = sub(ld3(c.light_pos + 4 * k), C.v(CD_P));
    l = normalize(lv);
  }
  const double diff = dmax0(dot(N, l));
  double refl = 0.0;
  if (diff > 0.0) {
    const D3 r = normalize(mirror(l, N));
    refl = dmax0(dot(r, C.v(CD_V)));
  }
  const double spec = pow_specular(refl, m.shininess);
  const D3 kd = C.v(CD_KD);
  D3 L = C.v(CD_L);
  L.x = dadd(L.x, dmul(lcol[0], dadd(dmul(kd.x, diff), dmul(m.ks.x, spec))));
  L.y = dadd(L.y, dmul(lcol[1], dadd(dmul(kd.y, diff), dmul(m.ks.y, spec))));
  L.z = dadd(L.z, dmul(lcol[2], dadd(dmul(kd.z, diff), dmul(m.ks.z, spec))));
  C.set(CD_L, L);
}

// One step of the path logic of a lane whose traversal has finished (H.mode != M_IDLE, H.node == -1).  Returns true
// when a ray must be started (request in r); false: the lane's state changed, look again.
// DEFER false: the ray is started right where it is requested (start_ray inlined at every site: straight-line code, the
// fastest form for the triangle-mesh kernels, profiles/r02_variants.md) and the function returns false.
template <int COUNT, bool DEFER, int S> RT_HD bool path_step(const Ctx &c, Hot &H, const Cold<S> &C, RayReq &r) {
  const SceneView &sv = c.sv;
  {
    Mat m;
    if (H.mode == M_NEAREST) {
      if (C.i(CI_DEPTH) == 0 && c.fp.spp_grid == 1) {
        const size_t p = (size_t)C.i(CI_PY) * (size_t)c.fp.cam.width + (size_t)C.i(CI_PX);
        if (c.out.hit_id) c.out.hit_id[p] = H.best_id;
        if (c.out.hit_t) c.out.hit_t[p] = (H.best_id >= 0) ? C.d(CD_BEST_T) : -1.0;
      }
      if (H.best_id < 0) {
        // miss: primary -> background; bounce -> weight * background (mytracer_gpu.cu:262,292)
        C.set(CD_COL, add(C.v(CD_COL), scale(C.d(CD_WEIGHT), ld3(c.fp.background))));
        return sample_done<COUNT, DEFER>(c, H, C, r);
      }
      hit_setup<COUNT>(c, H, C, m);
    } else { // M_ANYHIT finished: CI_LIGHT was the light under test
      load_mat(sv, C.i(CI_MAT), m);
      if (!H.found) add_light(c, C, m, C.i(CI_LIGHT), true);
      else bump<COUNT>(c, C_SHADOW_OCCLUDED);
      C.i(CI_LIGHT) = C.i(CI_LIGHT) + 1;
    }
    H.mode = M_SHADE;
    // lighting loop, mytracer.cpp:579-606
    int k = C.i(CI_LIGHT);
    while (k < c.fp.num_lights) {
      if (m.shadowable) {
        const D3 P = C.v(CD_P);
        const D3 lv = sub(ld3(c.light_pos + 4 * k), P);
        const double ldist = norm(lv);
        const D3 ldir = div3(lv, ldist);       // normalize(lv): the same sqrt, once
        bump<COUNT>(c, C_RAY_SHADOW);
        // The reference casts this shadow ray (mytracer.cpp:589-599) and the ray count includes it, but when
        // the surface faces away from the light (diff == 0, so spec = pow(0, shininess) = 0 for
        // shininess > 0, mytracer.cpp:525-533) the light adds exactly 0 whether it is occluded or not:
        // the result needs no traversal.
        if (m.shininess > 0.0 && !(0.0 < dot(C.v(CD_N), ldir))) {
          bump<COUNT>(c, C_SHADOW_TRIVIAL);
          ++k;
          continue;
        }
        if (RT_CACHE_LDIR) C.set(CD_LDIR, ldir);
        C.i(CI_LIGHT) = k;
        if (!DEFER) { start_ray<COUNT>(c, H, C, add(P, scale(1e-4, ldir)), ldir, M_ANYHIT, ldist); return false; }
        r.o = add(P, scale(1e-4, ldir)); r.d = ldir; r.mode = M_ANYHIT; r.tmax = ldist;
        return true;                      // shadow ray of light k
      }
      add_light(c, C, m, k, false);
      ++k;
    }
    C.i(CI_LIGHT) = k;
    // all lights done: colour of this bounce, then the mirror bounce (mytracer.cpp:546-555)
    const double w = C.d(CD_WEIGHT);
    C.set(CD_COL, add(C.v(CD_COL), scale(w, scale(dsub(1.0, m.mirror), C.v(CD_L)))));
    const int depth = C.i(CI_DEPTH);
    if (m.mirror > 0.0 && depth < c.fp.max_depth) {
      const D3 N = C.v(CD_N), P = C.v(CD_P);
      const D3 v = reflect(neg(C.v(CD_V)), N);
      C.d(CD_WEIGHT) = dmul(w, m.mirror);
      C.i(CI_DEPTH) = depth + 1;
      bump<COUNT>(c, C_RAY_REFLECT);
      if (!DEFER) { start_ray<COUNT>(c, H, C, add(P, scale(1e-4, v)), v, M_NEAREST, DBL_MAX); return false; }
      r.o = add(P, scale(1e-4, v)); r.d = v; r.mode = M_NEAREST; r.tmax = DBL_MAX;
      return true;
    }
    return sample_done<COUNT, DEFER>(c, H, C, r);
  }
}

// Advance a lane whose traversal has finished until it has a new ray in flight or is idle.
// Kernels whose start_ray is large - a scene with analytic primitives (their linear list is walked at every ray start), and
// the kernels without 4-wide records (scenes of analytic primitives only, A/B frames) - start rays at ONE place: half the
// code, 10 % faster on the spheres / molecule frames.  The triangle-mesh kernels keep the inlined copies (3 % faster there).
template <int COUNT, int S> RT_HD void advance(const Ctx &c, Hot &H, const Cold<S> &C) {
  constexpr bool kDefer = (COUNT & K_OBJ) != 0 || !(COUNT & K_WIDE);
  RayReq r;
  while (H.mode != M_IDLE && H.node == -1) { // -1: no ray in flight (a parked node is <= -2)
    if (path_step<COUNT, kDefer>(c, H, C, r)) start_ray<COUNT>(c, H, C, r.o, r.d, r.mode, r.tmax);
  }
}

// Begin a pixel (lane must be idle).
template <int COUNT, int S> RT_HD void begin_pixel(const Ctx &c, Hot &H, const Cold<S> &C, int px, int py) {
  C.i(CI_PX) = px;
  C.i(CI_PY) = py;
  C.i(CI_SAMPLE) = 0;
  C.set(CD_ACC, mk(0.0, 0.0, 0.0));
  start_primary<COUNT>(c, H, C);
}

// Begin ONE sub-pixel sample of a pixel (sample-per-lane mode of render_kernel).
template <int COUNT, int S> RT_HD void begin_sample(const Ctx &c, Hot &H, const Cold<S> &C, int px, int py, int sample) {
  C.i(CI_PX) = px;
  C.i(CI_PY) = py;
  C.i(CI_SAMPLE) = sample;
  start_primary<COUNT>(c, H, C);
}

// Map a work slot (owned tile j, lane l) to a pixel; false if it falls outside the image or
// the tile does not exist.  Tiles are kTileW x kTileH, numbered row-major; rank r of w owns
// tile t iff (t / chunk) % w == r.
RT_HD bool work_to_pixel(const Ctx &c, long long work, int &px, int &py) {
  const FrameParams &fp = c.fp;
  if (c.pixel_list) {
    if (work >= c.list_count) return false;
    const int p = c.pixel_list[work];
    px = p % fp.cam.width;
    py = p / fp.cam.width;
    return true;
  }
  const long long j = work / (kTileW * kTileH);
  const int l = (int)(work % (kTileW * kTileH));
  const long long chunk = fp.tile_chunk;
  const long long t = ((j / chunk) * fp.tile_world + fp.tile_rank) * chunk + (j % chunk);
  if (t >= fp.static_tiles) return false;
  const int tx = (int)(t % fp.tiles_x), ty = (int)(t / fp.tiles_x);
  px = tx * kTileW + (l % kTileW);
  py = ty * kTileH + (l / kTileW);
  return px < fp.cam.width && py < fp.cam.height;
}

// Dynamic pool: slot s of the pool covers tile static_tiles + s / 32, lane s % 32.
RT_HD bool pool_to_pixel(const Ctx &c, long long s, int &px, int &py) {
  const FrameParams &fp = c.fp;
  const long long t = fp.static_tiles + s / (kTileW * kTileH);
  const int l = (int)(s % (kTileW * kTileH));
  if (t >= (long long)fp.tiles_x * fp.tiles_y) return false;
  const int tx = (int)(t % fp.tiles_x), ty = (int)(t / fp.tiles_x);
  px = tx * kTileW + (l % kTileW);
  py = ty * kTileH + (l / kTileW);
  return px < fp.cam.width && py < fp.cam.height;
}
RT_HD long long pool_chunks(const Ctx &c) {
  const FrameParams &fp = c.fp;
  if (c.pixel_list) return 0;
  const long long dyn = (long long)fp.tiles_x * fp.tiles_y - fp.static_tiles;
  return (dyn + fp.dyn_chunk - 1) / fp.dyn_chunk;
}

// number of work slots this rank must scan
RT_HD long long work_slots(const Ctx &c) {
  const FrameParams &fp = c.fp;
  if (c.pixel_list) return c.list_count;
  const long long ntiles = fp.static_tiles;
  const long long group = (long long)fp.tile_chunk * fp.tile_world;
  const long long ngroups = (ntiles + group - 1) / group;
  return ngroups * fp.tile_chunk * (kTileW * kTileH);
}

} // namespace rtb
#endif
