// rt_wavefront.cuh — the hot path as a WAVEFRONT: a small FP32 traversal kernel and an FP64 shading kernel that
// exchange rays through path-state columns in HBM, instead of one kernel that holds both in its registers.
//
// Same functions of the reference as rt_kernel.cuh (compute_image_device / trace_device / lighting_device,
// mytracer_gpu.cu:119-160,237-312,611-669; CPU twins mytracer.cpp:510-608) and the SAME arithmetic: both
// kernels are built from the rt_core.cuh pieces (start_ray, trav_step4, confirm_tri, hit_setup, add_light,
// write_pixel), so a frame is bit-identical to the persistent megakernel's.  What changes is the schedule:
//
//   bounce 0:  TRACE(primary rays)                      -> nearest hit of every path
//              SHADE: hit set-up, ambient term; decides which lights need a shadow ray (a bit mask per path)
//                     and whether the surface spawns a mirror ray
//   bounce b:  TRACE(shadow rays of ALL lights of bounce b-1, then the mirror rays)   one launch, any-hit / nearest
//              SHADE: adds the unoccluded lights of bounce b-1 in light order (same sums as the CPU loop
//                     mytracer.cpp:579-606), then sets up the hit of bounce b
//   ...        until max_depth + 1 (shadow rays of the last hit only)
//
// A path = one work slot of this rank's tile share (tile j, lane l) — its index never changes, so path p of every
// launch is the same pixel and consecutive paths are the 8x4 pixels of one tile: a warp always traces 32 rays of
// one kind from one tile (coherent), and dead paths are skipped a warp at a time (no compaction pass).
// TRACE is a persistent grid with warp-level work fetch (one atomicAdd per 32 rays); it keeps only the FP32 ray
// constants and the traversal state in registers (the FP64 ray lives in 7 shared-memory columns for the rare exact
// confirmation), so it runs at twice the megakernel's occupancy and fits the instruction cache.  SHADE is one thread
// per path with its own register budget.
#ifndef RTB_WAVEFRONT_CUH
#define RTB_WAVEFRONT_CUH
#include "rt_core.cuh"

namespace rtb {

// path-state columns: the Cold columns of rt_core.cuh (same indices, so its functions work on them through Cold<0>)
// plus the wavefront's own
enum WfD { WD_WPREV = CD_NUM /* weight of the bounce whose lighting is pending */, WD_NUM };
enum WfI { WI_BESTID = CI_NUM, WI_MASK /* lights whose shadow ray is in flight */, WI_FLAGS, WI_NUM };
enum WfFlag : int { WF_ALIVE = 1, WF_LIGHTING = 2 /* lighting of the last hit waits for shadow rays */,
                    WF_REFLECT = 4 /* a mirror ray is in flight */ };
constexpr int kWfMaxLights = 30; // one mask bit per light

struct WfState {
  double *wd;          // WD_NUM columns of npaths doubles
  int *wi;             // WI_NUM columns of npaths ints
  unsigned char *occl; // num_lights columns of npaths bytes: any-hit result of each shadow ray
  long long npaths;
};

RT_HD Cold<0> wf_cold(const WfState &w, long long p) {
  Cold<0> G;
  G.db = w.wd + p;
  G.ib = w.wi + p;
  G.rstride = w.npaths;
  return G;
}

// Start the ray of (path p, kind) if there is one.  kind: -1 primary ray of sub-pixel sample `sample`;
// [0, L) shadow ray to that light; L the mirror ray.  T: the lane's scratch columns (CD_O, CD_D, CD_BEST_T).
// work: the work slot the path renders (wavefront: the path index itself; ray-parallel kernel: the slot its warp fetched).
template <int COUNT, int S>
RT_HD bool wf_ray_start(const Ctx &c, const WfState &w, long long p, int kind, int sample, Hot &H, const Cold<S> &T,
                        long long work = -1) {
  const Cold<0> G = wf_cold(w, p);
  if (kind < 0) {
    int px, py;
    if (!work_to_pixel(c, work >= 0 ? work : p, px, py)) { G.i(WI_FLAGS) = 0; return false; }
    G.i(CI_PX) = px; G.i(CI_PY) = py; G.i(CI_SAMPLE) = sample;
    if (sample == 0) G.set(CD_ACC, mk(0.0, 0.0, 0.0));
    G.i(CI_DEPTH) = 0;
    G.d(CD_WEIGHT) = 1.0;
    G.set(CD_COL, mk(0.0, 0.0, 0.0));
    G.i(WI_FLAGS) = WF_ALIVE;
    bump<COUNT>(c, C_RAY_PRIMARY);
    start_ray<COUNT>(c, H, T, ld3(c.fp.cam.eye), primary_dir(c, G), M_NEAREST, DBL_MAX);
    return true;
  }
  const int flags = G.i(WI_FLAGS);
  if (kind < c.fp.num_lights) {
    if (!(flags & WF_LIGHTING) || !((G.i(WI_MASK) >> kind) & 1)) return false;
    // the shadow ray of mytracer.cpp:589-599, same arithmetic as rt_core.cuh advance()
    const D3 P = G.v(CD_P);
    const D3 lv = sub(ld3(c.light_pos + 4 * kind), P);
    const double ldist = norm(lv);
    const D3 ldir = div3(lv, ldist);
    start_ray<COUNT>(c, H, T, add(P, scale(1e-4, ldir)), ldir, M_ANYHIT, ldist);
    return true;
  }
  if (!(flags & WF_REFLECT)) return false;
  const D3 N = G.v(CD_N), P = G.v(CD_P);
  const D3 v = reflect(neg(G.v(CD_V)), N);          // mytracer.cpp:546-555
  bump<COUNT>(c, C_RAY_REFLECT);
  start_ray<COUNT>(c, H, T, add(P, scale(1e-4, v)), v, M_NEAREST, DBL_MAX);
  return true;
}

template <int S>
RT_HD void wf_ray_finish(const Ctx &c, const WfState &w, long long p, int kind, const Hot &H, const Cold<S> &T) {
  if (kind >= 0 && kind < c.fp.num_lights) {
    w.occl[(size_t)kind * (size_t)w.npaths + (size_t)p] = (unsigned char)H.found;
    return;
  }
  const Cold<0> G = wf_cold(w, p);
  G.set(CD_O, T.v(CD_O));
  G.set(CD_D, T.v(CD_D));
  G.d(CD_BEST_T) = T.d(CD_BEST_T);
  G.i(WI_BESTID) = H.best_id;
}

// The path's sample is complete: accumulate it, write the pixel after the last sample (sample_done of rt_core.cuh;
// the sub-pixel samples of a pixel are separate passes over the same path columns, in sample order).
RT_HD void wf_path_done(const Ctx &c, const Cold<0> &G) {
  G.set(CD_ACC, add(G.v(CD_ACC), G.v(CD_COL)));
  if (G.i(CI_SAMPLE) + 1 == c.fp.spp_grid * c.fp.spp_grid) write_pixel(c, G);
  G.i(WI_FLAGS) = 0;
}

// One shading step of path p.  Returns true while the path still has rays in flight.
template <int COUNT> RT_HD bool wf_shade(const Ctx &c, const WfState &w, long long p) {
  const Cold<0> G = wf_cold(w, p);
  const int flags = G.i(WI_FLAGS);
  if (!(flags & WF_ALIVE)) return false;
  const SceneView &sv = c.sv;
  const int L = c.fp.num_lights;
  Mat m;
  if (flags & WF_LIGHTING) {
    // lighting loop of the previous hit, mytracer.cpp:579-606: the lights in order, each unless its shadow ray hit
    load_mat(sv, G.i(CI_MAT), m);
    const int mask = G.i(WI_MASK);
    for (int k = 0; k < L; ++k) {
      if (!((mask >> k) & 1)) continue;
      if (!w.occl[(size_t)k * (size_t)w.npaths + (size_t)p]) add_light(c, G, m, k, false);
      else bump<COUNT>(c, C_SHADOW_OCCLUDED);
    }
    G.set(CD_COL, add(G.v(CD_COL), scale(G.d(WD_WPREV), scale(dsub(1.0, m.mirror), G.v(CD_L)))));
    if (!(flags & WF_REFLECT)) { wf_path_done(c, G); return false; }
  }
  // nearest hit of the ray in flight (primary or mirror)
  Hot H;
  H.best_id = G.i(WI_BESTID);
  if (G.i(CI_DEPTH) == 0 && c.fp.spp_grid == 1) {
    const size_t px = (size_t)G.i(CI_PY) * (size_t)c.fp.cam.width + (size_t)G.i(CI_PX);
    if (c.out.hit_id) c.out.hit_id[px] = H.best_id;
    if (c.out.hit_t) c.out.hit_t[px] = (H.best_id >= 0) ? G.d(CD_BEST_T) : -1.0;
  }
  if (H.best_id < 0) {
    G.set(CD_COL, add(G.v(CD_COL), scale(G.d(CD_WEIGHT), ld3(c.fp.background))));
    wf_path_done(c, G);
    return false;
  }
  hit_setup<COUNT>(c, H, G, m);
  const double wgt = G.d(CD_WEIGHT);
  int nflags = WF_ALIVE;
  int mask = 0;
  if (m.shadowable) {
    const D3 P = G.v(CD_P), N = G.v(CD_N);
    for (int k = 0; k < L; ++k) {
      const D3 lv = sub(ld3(c.light_pos + 4 * k), P);
      const D3 ldir = div3(lv, norm(lv));
      bump<COUNT>(c, C_RAY_SHADOW);
      // a light that adds exactly zero whatever its occlusion needs no shadow ray (rt_core.cuh advance())
      if (m.shininess > 0.0 && !(0.0 < dot(N, ldir))) { bump<COUNT>(c, C_SHADOW_TRIVIAL); continue; }
      mask |= 1 << k;
    }
  } else {
    for (int k = 0; k < L; ++k) add_light(c, G, m, k, false);
  }
  if (mask) {
    G.i(WI_MASK) = mask;
    G.d(WD_WPREV) = wgt;
    nflags |= WF_LIGHTING;
  } else {
    G.set(CD_COL, add(G.v(CD_COL), scale(wgt, scale(dsub(1.0, m.mirror), G.v(CD_L)))));
  }
  const int depth = G.i(CI_DEPTH);
  if (m.mirror > 0.0 && depth < c.fp.max_depth) {
    G.d(CD_WEIGHT) = dmul(wgt, m.mirror);
    G.i(CI_DEPTH) = depth + 1;
    nflags |= WF_REFLECT;
  }
  if (nflags == WF_ALIVE) { wf_path_done(c, G); return false; }
  G.i(WI_FLAGS) = nflags;
  return true;
}

#if defined(__CUDACC__)
} // namespace rtb
#include "rt_kernel.cuh" // kBlockThreads, kFullMask
namespace rtb {
struct WfArgs {
  SceneView sv;
  FrameParams fp;
  const double *light_pos;
  const double *light_col;
  FrameOut out;
  WfState w;
  int bounce;                        // 0: primary rays; b: shadow rays of hit b-1 and the mirror rays to hit b
  int sample;                        // sub-pixel sample of this pass
  unsigned long long *task_counter;  // zeroed per launch: next batch of 32 rays
  unsigned int *pending;             // [bounce]: paths with rays in flight after SHADE(bounce); zeroed per pass
  unsigned long long *counters;      // C_COUNT entries (K_COUNT builds)
};

#ifndef RT_WF_CLAIM
#define RT_WF_CLAIM 8
#endif
#ifndef RT_WF_MIN_BLOCKS
#define RT_WF_MIN_BLOCKS 5
#endif
template <int COUNT>
__global__ void __launch_bounds__(kBlockThreads, RT_WF_MIN_BLOCKS) wf_trace_kernel(const __grid_constant__ WfArgs a) {
  if (a.bounce > 0 && a.pending[a.bounce - 1] == 0u) return; // nothing in flight: the frame finished early
  unsigned int cnt[(COUNT & K_COUNT) ? C_COUNT : 1];
  if (COUNT & K_COUNT) {
#pragma unroll
    for (int k = 0; k < C_COUNT; ++k) cnt[k] = 0;
  }
  Ctx c;
  c.sv = a.sv; c.fp = a.fp;
  c.light_pos = a.light_pos; c.light_col = a.light_col;
  c.out = a.out;
  c.pixel_list = nullptr; c.list_count = 0;
  c.cnt = cnt;
  // the exact FP64 ray of the lane (CD_O, CD_D, CD_BEST_T = columns 0..6): read by the rare FP64 confirmations only
  __shared__ double tcold[(CD_BEST_T + 1) * kBlockThreads];
  Cold<kBlockThreads> T;
  T.db = tcold + threadIdx.x;
  T.ib = nullptr;
  StackEntry stk[kStackMax];
  const unsigned lane = threadIdx.x & 31u;
  const long long nbatch = (a.w.npaths + 31) / 32;
  // After the first hits most paths of a frame may be dead (no mirror): a claim then covers RT_WF_CLAIM consecutive
  // batches, so an empty stretch of the frame costs one atomic round trip per 256 paths instead of one per 32.  While
  // more than a quarter of the paths are alive a claim is one batch: coarser claims cost 30 % on a dense stage (tail).
  const bool sparse = a.bounce > 0 && (long long)a.pending[a.bounce - 1] * 4 < a.w.npaths;
  const int per = sparse ? RT_WF_CLAIM : 1;
  const long long claims_per_kind = (nbatch + per - 1) / per;
  const long long nclaims = (a.bounce == 0 ? 1 : (long long)c.fp.num_lights + 1) * claims_per_kind;
  while (true) {
    // warp-level work fetch: one atomicAdd claims the next batches of 32 rays (one tile each, one ray kind)
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(a.task_counter, 1ull);
    t = __shfl_sync(kFullMask, t, 0);
    if ((long long)t >= nclaims) break;
    const int kind = a.bounce == 0 ? -1 : (int)((long long)t / claims_per_kind);
    const long long b0 = ((long long)t % claims_per_kind) * per;
    for (int j = 0; j < per && b0 + j < nbatch; ++j) {
    const long long p = (b0 + j) * 32 + lane;
    Hot H;
    H.mode = M_IDLE; H.node = -1; H.sp = 0; H.found = 0; H.best_id = -1; H.special = 0;
#if RT_LEAF_DEFER
    H.lf_first = H.lf_last = 0;
#endif
    const bool has = p < a.w.npaths && wf_ray_start<COUNT>(c, a.w, p, kind, a.sample, H, T);
    if (!__any_sync(kFullMask, has)) continue;
    // the traversal loops of rt_kernel.cuh, verbatim: warp-uniform step loop, d.x == 0 rays parked for the second loop
    for (;;) {
      if (COUNT & K_WIDE) {
        while (__any_sync(kFullMask, H.node >= 0)) {
          if (H.node >= 0) trav_step4<COUNT>(c, H, T, stk);
#if RT_LEAF_DEFER
          flush_leaf<COUNT>(c, H, T);     // (experiment builds: the wavefront does not postpone leaves)
#endif
        }
      } else {
        while (__any_sync(kFullMask, H.node >= 0)) {
          if (H.node >= 0) trav_step<COUNT, false>(c, H, T, stk);
        }
      }
      unpark(H.node);
      while (__any_sync(kFullMask, H.node >= 0)) {
        if (H.node >= 0) trav_step<COUNT, true>(c, H, T, stk);
      }
      if (!(COUNT & K_OBJ)) break;
      if (!__any_sync(kFullMask, next_tree<COUNT>(c, H, T))) break;
    }
    if (has) wf_ray_finish(c, a.w, p, kind, H, T);
    }
  }
  if (COUNT & K_COUNT) {
#pragma unroll
    for (int k = 0; k < C_COUNT; ++k)
      if (cnt[k]) atomicAdd(a.counters + k, (unsigned long long)cnt[k]);
  }
}

template <int COUNT>
__global__ void __launch_bounds__(kBlockThreads) wf_shade_kernel(const __grid_constant__ WfArgs a) {
  if (a.bounce > 0 && a.pending[a.bounce - 1] == 0u) return;
  unsigned int cnt[(COUNT & K_COUNT) ? C_COUNT : 1];
  if (COUNT & K_COUNT) {
#pragma unroll
    for (int k = 0; k < C_COUNT; ++k) cnt[k] = 0;
  }
  Ctx c;
  c.sv = a.sv; c.fp = a.fp;
  c.light_pos = a.light_pos; c.light_col = a.light_col;
  c.out = a.out;
  c.pixel_list = nullptr; c.list_count = 0;
  c.cnt = cnt;
  const long long p = (long long)blockIdx.x * kBlockThreads + threadIdx.x;
  const bool more = p < a.w.npaths && wf_shade<COUNT>(c, a.w, p);
  const unsigned m = __ballot_sync(kFullMask, more);
  if ((threadIdx.x & 31u) == 0 && m) atomicAdd(a.pending + a.bounce, (unsigned)__popc(m));
  if (COUNT & K_COUNT) {
#pragma unroll
    for (int k = 0; k < C_COUNT; ++k)
      if (cnt[k]) atomicAdd(a.counters + k, (unsigned long long)cnt[k]);
  }
}

// ---- the wavefront schedule INSIDE a persistent warp: rays of one path on neighbouring lanes --------------------------------
// Small frames (BASELINE configs 1 and 3: <= 350 k pixels) do not fill the machine with work, they fill it with dependent
// chains: a pixel of the rings scene is up to 6 hits x (2 shadow rays + 1 mirror ray), the persistent kernel traces them one
// after the other on one lane, a warp's 32 chains advance in lockstep (every round costs the slowest lane's traversal), and
// while the heaviest tiles run the other 2 000 warps have long finished (profiles/r02_rings_probe.log: one band of 440
// tiles takes 0.58 ms ALONE on the whole GPU).  Here a path owns G = 2, 4 or 8 neighbouring lanes: after a hit, its shadow
// rays (one lane per light) and its mirror ray are traced TOGETHER, so a path advances one bounce per round instead of
// one ray per round, and a warp holds 32 / G paths, so there are G times more units of work to spread over the machine.
// The stage functions are the wavefront's (wf_ray_start / wf_ray_finish / wf_shade: same arithmetic, same order of the
// light sums, bit-identical pixels); the path state lives in shared memory instead of HBM and the group's first lane shades.
#ifndef RT_RP_MIN_BLOCKS
#define RT_RP_MIN_BLOCKS 4
#endif
constexpr int kRpMaxPathsPerWarp = 16;                            // G = 2
constexpr int kRpPathsPerCta = kRpMaxPathsPerWarp * (kBlockThreads / 32);
template <int COUNT>
__global__ void __launch_bounds__(kBlockThreads, RT_RP_MIN_BLOCKS) rp_render_kernel(const __grid_constant__ KernelArgs a) {
  unsigned int cnt[1];
  Ctx c;
  c.sv = a.sv; c.fp = a.fp;
  c.light_pos = a.light_pos; c.light_col = a.light_col;
  c.out = a.out;
  c.pixel_list = a.pixel_list;
  c.list_count = a.list_count_ptr ? *a.list_count_ptr : a.list_count;
  c.cnt = cnt;
  __shared__ double s_wd[WD_NUM * kRpPathsPerCta];
  __shared__ int s_wi[WI_NUM * kRpPathsPerCta];
  __shared__ unsigned char s_occl[kWfMaxLights * kRpPathsPerCta];
  __shared__ double tcold[(CD_BEST_T + 1) * kBlockThreads];
  WfState w;
  w.wd = s_wd; w.wi = s_wi; w.occl = s_occl; w.npaths = kRpPathsPerCta;
  Cold<kBlockThreads> T;
  T.db = tcold + threadIdx.x;
  T.ib = nullptr;
  StackEntry stk[kStackMax];
  const unsigned lane = threadIdx.x & 31u;
  const int warp = (int)(threadIdx.x >> 5);
  const int L = c.fp.num_lights;
  const int G = L + 1 <= 2 ? 2 : (L + 1 <= 4 ? 4 : 8);           // lanes per path
  const int per_warp = 32 / G;                                    // paths per warp
  const int role = (int)lane & (G - 1);
  const long long pcol = (long long)warp * kRpMaxPathsPerWarp + ((int)lane / G);   // the path's state column
  const long long slots = work_slots(c);
  const int s2 = c.fp.spp_grid * c.fp.spp_grid;
  for (;;) {
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(a.work_counter, (unsigned long long)per_warp);
    base = __shfl_sync(kFullMask, base, 0);
    if ((long long)base >= slots) break;
    const long long work = (long long)base + ((int)lane / G);
    const bool mine = work < slots;
    for (int sample = 0; sample < s2; ++sample) {
      for (int b = 0;; ++b) {
        const int nk = b == 0 ? 1 : L + 1;                        // ray kinds of this bounce: primary | L shadow rays + mirror
        for (int k0 = 0; k0 < nk; k0 += G) {                      // (more kinds than lanes per path: several sub-rounds)
          const int kind = b == 0 ? -1 : k0 + role;
          Hot H;
          H.mode = M_IDLE; H.node = -1; H.sp = 0; H.found = 0; H.best_id = -1; H.special = 0;
#if RT_LEAF_DEFER
          H.lf_first = H.lf_last = 0;
#endif
          const bool has = mine && (b == 0 ? role == 0 : kind < nk) && wf_ray_start<COUNT>(c, w, pcol, kind, sample, H, T, work);
          if (__any_sync(kFullMask, has)) {
            for (;;) {
              if (COUNT & K_WIDE) {
                while (__any_sync(kFullMask, H.node >= 0)) {
                  if (H.node >= 0) trav_step4<COUNT>(c, H, T, stk);
#if RT_LEAF_DEFER
                  flush_leaf<COUNT>(c, H, T);
#endif
                }
              } else {
                while (__any_sync(kFullMask, H.node >= 0)) {
                  if (H.node >= 0) trav_step<COUNT, false>(c, H, T, stk);
                }
              }
              unpark(H.node);
              while (__any_sync(kFullMask, H.node >= 0)) {
                if (H.node >= 0) trav_step<COUNT, true>(c, H, T, stk);
              }
              if (!(COUNT & K_OBJ)) break;
              if (!__any_sync(kFullMask, next_tree<COUNT>(c, H, T))) break;
            }
            if (has) wf_ray_finish(c, w, pcol, kind, H, T);
          }
          __syncwarp();
        }
        // the path's first lane shades: lighting of the previous hit (shadow results of its lanes), set-up of the new one
        const bool more = mine && role == 0 && wf_shade<COUNT>(c, w, pcol);
        __syncwarp();
        if (!__any_sync(kFullMask, more)) break;
      }
    }
  }
}
#endif // __CUDACC__

} // namespace rtb
#endif
