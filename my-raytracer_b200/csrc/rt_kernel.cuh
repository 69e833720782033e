// rt_kernel.cuh — the sm_100a render kernel: persistent warps, warp-level work fetch.
//
// Replaces compute_image_device (mytracer_gpu.cu:119-160: one thread per pixel, 16x16 blocks,
// FP64 recursion-free but stack-in-local-memory traversal).  Here:
//   * the grid is sized to the machine (SM count x resident CTAs), not to the image;
//   * warp-level work fetch: when the lanes of a warp are idle, one lane claims that many work
//     slots with ONE atomicAdd and the slots are handed out by __ballot/__popc rank
//     (north star (b): "warp-level work fetch (atomics plus __shfl / __ballot)"); by default a warp
//     refills when all 32 lanes are idle, i.e. it takes one 8x4 tile at a time;
//   * work slots follow 8x4-pixel tiles so that a freshly filled warp covers a compact screen
//     patch (coherent primary rays);
//   * every lane runs the rt_core.cuh state machine; the traversal loop is shared by primary,
//     shadow and reflection rays.
#ifndef RTB_KERNEL_CUH
#define RTB_KERNEL_CUH
#include "rt_core.cuh"
#include <cuda_runtime.h>

namespace rtb {

constexpr int kBlockThreads = 128;
constexpr unsigned kFullMask = 0xffffffffu;
constexpr int kPartialRefillObjects = 32; // at or above this many analytic primitives lanes are refilled one by one

struct KernelArgs {
  SceneView sv;
  FrameParams fp;
  const double *light_pos;
  const double *light_col;
  FrameOut out;
  const int *pixel_list;
  int list_count;
  const int *list_count_ptr;        // optional: list length produced on the device (no host sync)
  unsigned long long *work_counter; // [0] local slot counter, [1] "dynamic pool is empty" flag; zeroed per launch
  unsigned long long *dyn_counter;  // cross-GPU chunk counter of this frame (may be peer memory), or nullptr
  unsigned long long *counters;     // C_COUNT entries (COUNT builds)
  double *scratch_d;                // RT_COLD == 2: CD_NUM columns of grid*block doubles
  int *scratch_i;                   //               CI_NUM columns of grid*block ints
};

#ifndef RT_COLD
#define RT_COLD 0
#endif
#ifndef RT_MIN_BLOCKS
#define RT_MIN_BLOCKS 4
#endif
#ifndef RT_SAMPLE_LANES
#define RT_SAMPLE_LANES 1   // 0: A/B, every lane traces all sub-pixel samples of its pixel one after the other (rt_api.cu)
#endif
template <int COUNT>
__global__ void __launch_bounds__(kBlockThreads, RT_MIN_BLOCKS) render_kernel(const __grid_constant__ KernelArgs a) {
  unsigned int cnt[(COUNT & K_COUNT) ? C_COUNT : 1];
  if (COUNT & K_COUNT) {
#pragma unroll
    for (int k = 0; k < C_COUNT; ++k) cnt[k] = 0;
  }
#ifdef RT_TAILPROBE
  unsigned long long t_probe0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_probe0));
#endif
  Ctx c;
  c.sv = a.sv; c.fp = a.fp;
  c.light_pos = a.light_pos; c.light_col = a.light_col;
  c.out = a.out;
  c.pixel_list = a.pixel_list;
  c.list_count = a.list_count_ptr ? *a.list_count_ptr : a.list_count;
  c.cnt = cnt;

  // cold per-lane state: one shared-memory column per thread (rt_core.cuh)
#if RT_COLD == 0   // shared-memory columns
  __shared__ double cold_d[CD_NUM * kBlockThreads];
  __shared__ int cold_i[CI_NUM * kBlockThreads];
  Cold<kBlockThreads> C;
  C.db = cold_d + threadIdx.x;
  C.ib = cold_i + threadIdx.x;
#elif RT_COLD == 1 // registers / local memory (compiler's choice)
  double cold_d[CD_NUM];
  int cold_i[CI_NUM];
  Cold<1> C;
  C.db = cold_d;
  C.ib = cold_i;
#else              // global scratch columns
  Cold<0> C;
  C.rstride = (long long)gridDim.x * kBlockThreads;
  C.db = a.scratch_d + (size_t)blockIdx.x * kBlockThreads + threadIdx.x;
  C.ib = a.scratch_i + (size_t)blockIdx.x * kBlockThreads + threadIdx.x;
#endif
  StackEntry stk[kStackMax];
  Hot H;
  H.mode = M_IDLE;
  H.node = -1;
  H.sp = 0;
  H.found = 0;
  H.best_id = -1;
  H.special = 0;
#if RT_LEAF_DEFER
  H.lf_first = H.lf_last = 0;
#endif
  const unsigned lane = threadIdx.x & 31u;
  const bool partial_refill = c.sv.num_objects >= kPartialRefillObjects;
  // Sub-pixel samples: with a 2x2 or 4x4 grid the s^2 samples of a pixel go to s^2 neighbouring LANES (a warp = 8 or 2
  // pixels) instead of being traced one after the other by one lane.  A work slot is then (pixel, sample): the unit of
  // work a warp fetches costs 1/s^2 of a tile with all its samples, so the tail of a frame (or of a GPU's share of it)
  // is that much shorter, and the rays of a warp are closer together.  The samples of a pixel are summed in sample order
  // when the whole warp has finished (same additions as the sequential loop: bit-identical pixels).
  const int s2 = c.fp.spp_grid * c.fp.spp_grid;
  // (K_SPL kernels only: the host launches them for s^2 = 4 or 16 when lanes are refilled a warp at a time.)
  const bool sample_lanes = (COUNT & K_SPL) != 0;
  const int spl = sample_lanes ? s2 : 1;     // lanes per pixel
  bool had = false;                          // this lane holds a finished sample that is not summed yet
  const long long slots = work_slots(c) * spl;
  const long long nchunks = a.dyn_counter ? pool_chunks(c) : 0;
  bool exhausted = false;   // warp-uniform: no work left anywhere
  bool local_done = false;  // warp-uniform: this rank's static share is drained
  long long wp_next = 0, wp_end = 0; // warp-private range of dynamic-pool slots (warp-uniform)

  while (true) {
    // ---- phase A: path logic (FP64, cold state) for lanes whose traversal finished -----------
    if (H.mode != M_IDLE && H.node == -1) advance<COUNT>(c, H, C);

    // ---- warp-level work fetch for idle lanes ----------------------------------------------
    const unsigned idle = __ballot_sync(kFullMask, H.mode == M_IDLE);
    // A warp takes new pixels only when ALL its lanes are idle: the 32 pixels of a tile then move through the
    // path stages together (primary rays, then the shadow rays of light 0, ...), which keeps both the traversal
    // and the FP64 path logic coherent.  Refilling single lanes as they finish keeps more lanes busy but mixes
    // ray generations in a warp and is 17 % slower (RT_REFILL_PARTIAL; profiles/r01_variants.md).
    // Exception: scenes dominated by the linear list of analytic primitives (the reference has no BVH over
    // objects_): every ray walks the same list in lockstep whatever its generation, so there is no coherence
    // to protect and full lanes win (molecule, 602 objects: 2.9 ms against 4.1 ms).
    if (sample_lanes && idle == kFullMask && __any_sync(kFullMask, had)) {
      // every lane of the warp has finished its sample: ACC = ((0 + c0) + c1) + ... in sample order, first lane writes
      const int g0 = (int)lane & ~(spl - 1);
      D3 col = mk(0.0, 0.0, 0.0);
      if (had) col = C.v(CD_COL);
      D3 acc = mk(0.0, 0.0, 0.0);
      for (int k = 0; k < spl; ++k) {
        D3 ck;
        ck.x = __shfl_sync(kFullMask, col.x, g0 + k);
        ck.y = __shfl_sync(kFullMask, col.y, g0 + k);
        ck.z = __shfl_sync(kFullMask, col.z, g0 + k);
        acc = add(acc, ck);
      }
      if (had && (int)lane == g0) {
        C.set(CD_ACC, acc);
        write_pixel(c, C);
      }
      had = false;
    }
#ifdef RT_REFILL_PARTIAL
    if (idle != 0u && !exhausted) {
#else
    if ((partial_refill ? idle != 0u : idle == kFullMask) && !exhausted) {
#endif
      const int n = __popc(idle);
      const int leader = __ffs(idle) - 1;
      const int rank_in_idle = __popc(idle & ((1u << lane) - 1u));
      if (!local_done) {
        // static share of this rank: one atomicAdd per refill, slots handed out by ballot rank
        unsigned long long base = 0;
        if ((int)lane == leader) base = atomicAdd(a.work_counter, (unsigned long long)n);
        base = __shfl_sync(kFullMask, base, leader);
        if ((long long)base >= slots) {
          local_done = true;
          exhausted = (nchunks == 0);
        } else if (H.mode == M_IDLE) {
          const long long mine = (long long)base + rank_in_idle;
          int px, py;
          if (sample_lanes) {
            if (mine < slots && work_to_pixel(c, mine / spl, px, py)) { begin_sample<COUNT>(c, H, C, px, py, (int)(mine % spl)); had = true; }
          } else if (mine < slots && work_to_pixel(c, mine, px, py)) begin_pixel<COUNT>(c, H, C, px, py);
        }
      } else {
        // cross-GPU dynamic tail: the warp claims dyn_chunk tiles at a time with ONE system-scope atomic
        // on the frame's shared counter and serves its lanes from that private range
        if (wp_next >= wp_end) {
          unsigned long long q = ~0ull;
          if ((int)lane == leader) {
            if (*(volatile unsigned long long *)(a.work_counter + 1) == 0ull) {
              q = atomicAdd_system(a.dyn_counter, 1ull);
              if ((long long)q >= nchunks) a.work_counter[1] = 1ull; // spare the other warps the remote atomic
            }
          }
          q = __shfl_sync(kFullMask, q, leader);
          if (q >= (unsigned long long)nchunks) {
            exhausted = true;
          } else {
            wp_next = (long long)q * c.fp.dyn_chunk * (kTileW * kTileH) * spl;
            wp_end = wp_next + (long long)c.fp.dyn_chunk * (kTileW * kTileH) * spl;
          }
        }
        if (!exhausted) {
          const long long mine = wp_next + rank_in_idle;
          wp_next += n;
          int px, py;
          if (sample_lanes) {
            if (mine < wp_end && pool_to_pixel(c, mine / spl, px, py)) { begin_sample<COUNT>(c, H, C, px, py, (int)(mine % spl)); had = true; }
          } else if (H.mode == M_IDLE && mine < wp_end && pool_to_pixel(c, mine, px, py)) begin_pixel<COUNT>(c, H, C, px, py);
        }
      }
    }
    const unsigned busy = __ballot_sync(kFullMask, H.mode != M_IDLE);
    if (busy == 0u) {
      if (exhausted) break;
      continue;
    }
    // ---- phase B: one traversal loop for every ray type (hot state, FP32) -------------------
    // (Two steps per vote: 13 % slower — the warp must re-converge after every step.)
    // (Leaving the loop early to refill finished lanes was measured and is a loss: lanes of one warp then
    // work on rays of different generations, the node loads stop coalescing and the shading code runs at
    // low lane counts — 10.8-17 ms instead of 7.2, profiles/r01_variants.md.)
    for (;;) {
      if (COUNT & K_WIDE8) {
        // ordinary rays walk the 8-wide tree: 6 dependent steps per ray instead of 10, eight independent box tests in each
        while (__any_sync(kFullMask, H.node >= 0)) {
          if (H.node >= 0) trav_step8<COUNT>(c, H, C, stk);
        }
      } else if (COUNT & K_WIDE) {
        // ordinary rays walk the 4-wide tree: half the dependent steps, four independent box tests in each
#if RT_LEAF_DEFER
        for (;;) {
          const unsigned pend = __ballot_sync(kFullMask, H.lf_last > H.lf_first);
          const unsigned walking = __ballot_sync(kFullMask, H.node >= 0);
          if (!(pend | walking)) break;
          // intersect the postponed leaves together: enough lanes hold one, or nobody without one is still walking
          if (pend && (__popc(pend) >= RT_LEAF_DEFER_MIN || (walking & ~pend) == 0u)) {
            flush_leaf<COUNT>(c, H, C);
            continue;
          }
#if RT_LEAF_DEFER_SPEC
          if (H.node >= 0) trav_step4<COUNT>(c, H, C, stk);
#else
          if (H.node >= 0 && !(H.lf_last > H.lf_first)) trav_step4<COUNT>(c, H, C, stk);
#endif
        }
#else
        while (__any_sync(kFullMask, H.node >= 0)) {
          if (H.node >= 0) trav_step4<COUNT>(c, H, C, stk);
        }
#endif
      } else {
        while (__any_sync(kFullMask, H.node >= 0)) {
          if (H.node >= 0) trav_step<COUNT, false>(c, H, C, stk);
        }
      }
      // rays with direction.x == 0 were parked (rt_core.cuh park()): they walk the reference tree with the step
      // that knows its NaN quirk and its primitive numbering
      unpark(H.node);
      while (__any_sync(kFullMask, H.node >= 0)) {
        if (H.node >= 0) trav_step<COUNT, true>(c, H, C, stk);
      }
      // rays that finished the tree over the analytic spheres move on to the triangle tree (rt_core.cuh)
      if (!(COUNT & K_OBJ)) break;
      if (!__any_sync(kFullMask, next_tree<COUNT>(c, H, C))) break;
    }
  }
  if (COUNT & K_COUNT) {
#pragma unroll
    for (int k = 0; k < C_COUNT; ++k)
      if (cnt[k]) atomicAdd(a.counters + k, (unsigned long long)cnt[k]);
  }
#ifdef RT_TAILPROBE
  // experiment (tools/gpu_tailprobe.py): when does each warp run out of work?  counters[0] = sum of
  // warp end times, [1] = max, [2] = min start, [3] = number of warps (globaltimer ns)
  if (!(COUNT & K_COUNT) && lane == 0) {
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    atomicAdd(a.counters + 0, t1 - t_probe0);
    atomicMax(a.counters + 1, t1);
    atomicMin(a.counters + 2, t_probe0);
    atomicAdd(a.counters + 3, 1ull);
    atomicAdd(a.counters + 4, t_probe0);
  }
#endif
}

// Contrast test of the adaptive supersampling pass (adaptive_supersampling_device,
// mytracer_gpu.cu:162-229): interior pixels whose summed squared colour distance to the 4-neighbourhood
// exceeds `threshold` are appended to `list`; a second render_kernel launch re-renders exactly those
// pixels with a subp x subp sample grid (a compaction instead of the reference's mostly-idle full grid).
__global__ void flag_contrast_kernel(const double *__restrict__ pixels, int width, int height, double threshold,
                                     int *__restrict__ list, int *__restrict__ count) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= width - 1 || y >= height - 1 || x < 1 || y < 1) return;
  const size_t p = (size_t)y * width + x;
  const D3 c = ld3(pixels + 4 * p);
  auto dist2 = [&](size_t q) {
    const D3 d = sub(c, ld3(pixels + 4 * q));
    return dot(d, d);
  };
  const double n = dadd(dadd(dadd(dist2(p + 1), dist2(p + width)), dist2(p - 1)), dist2(p - width));
  if (n > threshold) {
    const int slot = atomicAdd(count, 1);
    list[slot] = (int)p;
  }
}

} // namespace rtb
#endif
