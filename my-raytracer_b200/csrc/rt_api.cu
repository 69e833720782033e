// rt_api.cu — implementation of the C ABI declared in include/rt_b200.h.
//
// Host plumbing that the reference spreads over Raytracer::compute_image_cuda /
// prepareDeviceResources / copyLights / allocate (mytracer.cpp:68-159) and
// launch_compute_image_device (mytracer_gpu.cu:44-113), rebuilt around one resident scene blob,
// explicit streams and CUDA events, and status codes instead of print-and-continue.
// There is no CPU fallback anywhere in this file.
#include "../../include/rt_b200.h"
#include "rt_flatten.h"
#include "rt_sah.h"
#include "rt_frame.h"
#include "rt_kernel.cuh"
#include "rt_wavefront.cuh"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace rtb {
static thread_local std::string g_err;
void set_error(const std::string &m) { g_err = m; }
} // namespace rtb

using namespace rtb;

#define RT_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                               \
      return (e_ == cudaErrorMemoryAllocation) ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA;               \
    }                                                                                              \
  } while (0)

struct rt_scene {
  int device = 0;
  unsigned char *d_blob = nullptr;
  size_t bytes = 0;
  BlobHeader hdr;
  SceneView view;
  std::vector<cudaTextureObject_t> tex;
  unsigned long long *d_tex = nullptr;
  // per-frame device resources
  cudaStream_t stream = nullptr;
  // Launch ring: frames may be in flight on several streams (one work counter + event pair per launch)
  static constexpr int kRing = 8;
  cudaEvent_t ev0[kRing] = {}, ev1[kRing] = {};
  unsigned long long *d_work = nullptr; // kRing x {slot counter, pool-empty flag}
  unsigned launch_seq = 0;
  bool slot_used[kRing] = {};
  int last_slot = 0;
  unsigned long long *d_counters = nullptr;
  double *d_lights = nullptr; // pos[4*n] then col[4*n]
  int lights_cap = 0;
  // staging outputs for the host-buffer path
  uint8_t *d_rgba8 = nullptr;
  double *d_pixels = nullptr;
  int32_t *d_ids = nullptr;
  double *d_t = nullptr;
  size_t cap_rgba8 = 0, cap_pixels = 0, cap_ids = 0, cap_t = 0;
  double *d_scratch_d = nullptr; // cold per-lane state columns (RT_COLD == 2 builds)
  int *d_scratch_i = nullptr;
  int *d_list = nullptr; // supersampling pixel list
  size_t cap_list = 0;
  // wavefront pipeline (rt_wavefront.cuh): path-state columns and launch control block per ring slot
  struct WfSlot {
    double *wd = nullptr; int *wi = nullptr; unsigned char *occl = nullptr;
    unsigned long long *ctl = nullptr;   // [0, kWfCtlTasks): task counters, one per TRACE launch; then the pending[] array
    long long cap_paths = 0; int cap_lights = 0;
  } wf[kRing];
  int grid_wf[4] = {0, 0, 0, 0};         // resident CTAs x SMs of wf_trace_kernel<cfg>, cfg = (K_OBJ ? 1 : 0) | (K_WIDE ? 2 : 0)
  int wavefront_default = 0;             // RT_B200_PIPELINE=wavefront
  int grid_rp[4] = {0, 0, 0, 0};         // the same for rp_render_kernel<cfg> (rays of a path on neighbouring lanes)
  int rp_default = 0;                    // RT_B200_PIPELINE=rayparallel: every frame it covers; -1 (=persistent): never
  bool has_mirror = false;               // some mesh material reflects (paths can be deep)
  int last_pipeline = 0;                 // what the last launch ran: 0 persistent megakernel, 1 wavefront
  int grid_plain = 0, grid_count = 0;
  int grid_wide = 0, grid_wide_count = 0; // K_WIDE kernels (scenes whose trees all have 4-wide twins)
  bool wide_ok = false;
  bool wide8_ok = false, wide8_default = false;   // 8-wide twin usable; RT_B200_NODES=8
  int grid_wide8 = 0, grid_wide8_count = 0;
  int sm_count = 0, ctas_per_sm = 0; // ctas_per_sm: 0 = every resident slot (rt_scene_set_launch_width)
  size_t extra_smem = 0;
  rt_counters counters;
  rt_scene_info info;
  bool timed = false;
};

namespace {

int ensure_device(std::string &err) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    err = std::string("no CUDA device available (") + cudaGetErrorString(e) +
          "): this library has no CPU fallback";
    return RT_ERR_NO_DEVICE;
  }
  return RT_OK;
}

// Grows a staging buffer.  The zero fill is ordered on `st`, the stream every consumer of the buffer runs on
// (a cudaMemset on the legacy stream would not be ordered with a cudaStreamNonBlocking stream).
template <typename T> int ensure_cap(T *&p, size_t &cap, size_t need_bytes, cudaStream_t st) {
  if (need_bytes <= cap) return RT_OK;
  if (p) {
    RT_CUDA(cudaStreamSynchronize(st)); // nothing in flight may still use the old block
    cudaFree(p);
  }
  p = nullptr;
  cap = 0;
  RT_CUDA(cudaMalloc(&p, need_bytes));
  RT_CUDA(cudaMemsetAsync(p, 0, need_bytes, st));
  cap = need_bytes;
  return RT_OK;
}

// Payload check of an adopted blob (it may come from a broadcast buffer or from the caller): every child reference,
// leaf range and index table entry must stay inside its section, otherwise the render kernel would read out of
// bounds.  One pass over the node records and the index tables, HBM-bound, once per scene.
__global__ void validate_blob_kernel(SceneView v, int num_obj_tree, int *bad) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  bool ok = true;
  if (i < v.num_nodes) {
    const NodeRec r = v.nodes[i];
    const int limit = (v.obj_root >= 0 && i >= v.obj_root) ? num_obj_tree : v.num_tris;
    const int ref[2] = {r.ref0, r.ref1}, cnt[2] = {r.cnt0, r.cnt1};
    const bool empty[2] = {r.c0_half[0] < 0.0f, r.c1_half[0] < 0.0f};
    for (int k = 0; k < 2; ++k) {
      if (empty[k]) continue;
      if (cnt[k] < 0 || ref[k] < 0) ok = false;
      else if (cnt[k] == 0) ok = ok && ref[k] < v.num_nodes;
      else ok = ok && (long long)ref[k] + cnt[k] <= limit;
    }
  }
  if (i < v.num_nodes4) {
    const Node4Rec r = v.nodes4[i];
    const int limit = i >= v.num_nodes4_tri ? num_obj_tree : v.num_tris;
    for (int k = 0; k < 4; ++k) {
      if (r.ref[k] == -1 && r.cnt[k] == 0) continue; // empty slot
      if (r.cnt[k] < 0 || r.ref[k] < 0) ok = false;
      else if (r.cnt[k] == 0) ok = ok && r.ref[k] < v.num_nodes4;
      else ok = ok && (long long)r.ref[k] + r.cnt[k] <= limit;
    }
  }
  if (i < v.num_nodes8) {
    const Node8Rec r = v.nodes8[i];
    for (int k = 0; k < 8; ++k) {
      if (r.ref[k] == -1 && r.cnt[k] == 0) continue; // empty slot
      if (r.cnt[k] < 0 || r.ref[k] < 0) ok = false;
      else if (r.cnt[k] == 0) ok = ok && r.ref[k] < v.num_nodes8;
      else ok = ok && (long long)r.ref[k] + r.cnt[k] <= v.num_tris;
    }
  }
  if (i < v.num_tris) {
    ok = ok && (unsigned)v.tri_id[i] < (unsigned)v.num_tris && (unsigned)v.tri_pos[i] < (unsigned)v.num_tris;
    ok = ok && (unsigned)v.shade[i].mesh < (unsigned)v.num_meshes;
  }
  if (i < num_obj_tree) ok = ok && (unsigned)v.obj_order[i] < (unsigned)v.num_objects;
  if (i < v.num_lin) ok = ok && (unsigned)v.lin_idx[i] < (unsigned)v.num_objects;
  if (!ok) atomicExch(bad, 1);
}

int adopt_blob(rt_scene *s) {
  std::memset(&s->view, 0, sizeof(s->view));
  const BlobHeader &h = s->hdr;
  unsigned char *b = s->d_blob;
  s->view.nodes = reinterpret_cast<const NodeRec *>(b + h.off_nodes);
  s->view.filt = reinterpret_cast<const TriFilter *>(b + h.off_filter);
  s->view.exact = reinterpret_cast<const TriExact *>(b + h.off_exact);
  s->view.shade = reinterpret_cast<const TriShade *>(b + h.off_shade);
  s->view.meshes = reinterpret_cast<const MeshRec *>(b + h.off_meshes);
  s->view.objects = reinterpret_cast<const ObjectRec *>(b + h.off_objects);
  s->view.objfilt = reinterpret_cast<const ObjFilter *>(b + h.off_objfilt);
  s->view.obj_order = reinterpret_cast<const int32_t *>(b + h.off_obj_order);
  s->view.lin_idx = reinterpret_cast<const int32_t *>(b + h.off_lin_idx);
  s->view.obj_root = h.obj_root; s->view.num_lin = h.num_lin;
  s->view.tri_nodes = h.q_base + h.num_nodes_q;
  s->view.num_nodes4 = h.num_nodes4;
  s->view.obj_root4 = h.obj_root4; s->view.num_nodes4_tri = h.num_nodes4_tri;
  s->view.nodes4 = h.num_nodes4 > 0 ? reinterpret_cast<const Node4Rec *>(b + h.off_nodes4) : nullptr;
  s->view.nodes8 = h.num_nodes8 > 0 ? reinterpret_cast<const Node8Rec *>(b + h.off_nodes8) : nullptr;
  s->view.num_nodes8 = h.num_nodes8;
  s->view.texels = reinterpret_cast<const float *>(b + h.off_texels);
  s->view.num_nodes = h.num_nodes; s->view.num_tris = h.num_tris;
  s->view.num_meshes = h.num_meshes; s->view.num_objects = h.num_objects;
  s->view.box_extent = h.box_extent;
  s->view.maxx = reinterpret_cast<const NodeMaxX *>(b + h.off_maxx);
  s->view.tri_id = reinterpret_cast<const int32_t *>(b + h.off_tri_id);
  s->view.tri_pos = reinterpret_cast<const int32_t *>(b + h.off_tri_pos);
  s->view.q_base = h.q_base;
  s->view.root_max_x = h.root_max_x;
  {
    int *d_bad = nullptr, bad = 0;
    RT_CUDA(cudaMalloc(&d_bad, sizeof(int)));
    RT_CUDA(cudaMemset(d_bad, 0, sizeof(int)));
    long long n = h.num_nodes;
    if (h.num_nodes4 > n) n = h.num_nodes4;
    if (h.num_nodes8 > n) n = h.num_nodes8;
    if (h.num_tris > n) n = h.num_tris;
    if (h.num_objects > n) n = h.num_objects;
    if (n > 0) validate_blob_kernel<<<(unsigned)((n + 255) / 256), 256>>>(s->view, h.num_obj_tree, d_bad);
    cudaError_t e = cudaMemcpy(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost);
    cudaFree(d_bad);
    if (e != cudaSuccess) { set_error(std::string("blob validation: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    if (bad) { set_error("blob: a child reference, leaf range or index table entry lies outside its section"); return RT_ERR_BAD_SCENE; }
  }
  // texture objects over the blob's texel section ("textures go through texture objects")
  std::vector<MeshRec> meshes(h.num_meshes);
  if (h.num_meshes)
    RT_CUDA(cudaMemcpy(meshes.data(), b + h.off_meshes, sizeof(MeshRec) * h.num_meshes, cudaMemcpyDeviceToHost));
  s->tex.assign(h.num_meshes, 0);
  for (int m = 0; m < h.num_meshes; ++m) s->has_mirror = s->has_mirror || meshes[m].mirror > 0.0;
  for (int m = 0; m < h.num_meshes; ++m) {
    if (meshes[m].tex_w == -1) continue;
    cudaResourceDesc rd;
    std::memset(&rd, 0, sizeof(rd));
    rd.resType = cudaResourceTypeLinear;
    rd.res.linear.devPtr = b + h.off_texels + meshes[m].tex_offset * 16;
    rd.res.linear.desc = cudaCreateChannelDesc<float4>();
    rd.res.linear.sizeInBytes = (size_t)meshes[m].tex_w * meshes[m].tex_h * 16;
    cudaTextureDesc td;
    std::memset(&td, 0, sizeof(td));
    td.readMode = cudaReadModeElementType;
    td.filterMode = cudaFilterModePoint;
    td.addressMode[0] = cudaAddressModeClamp;
    RT_CUDA(cudaCreateTextureObject(&s->tex[m], &rd, &td, nullptr));
  }
  if (h.num_meshes) {
    RT_CUDA(cudaMalloc(&s->d_tex, sizeof(unsigned long long) * h.num_meshes));
    RT_CUDA(cudaMemcpy(s->d_tex, s->tex.data(), sizeof(unsigned long long) * h.num_meshes, cudaMemcpyHostToDevice));
  }
  s->view.tex_objs = s->d_tex;
  RT_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  for (int k = 0; k < rt_scene::kRing; ++k) {
    RT_CUDA(cudaEventCreate(&s->ev0[k]));
    RT_CUDA(cudaEventCreate(&s->ev1[k]));
  }
  RT_CUDA(cudaMalloc(&s->d_work, sizeof(unsigned long long) * 2 * rt_scene::kRing));
  RT_CUDA(cudaMalloc(&s->d_counters, sizeof(unsigned long long) * C_COUNT));
  // persistent grid: SM count x resident CTAs of each kernel variant
  cudaDeviceProp prop;
  RT_CUDA(cudaGetDeviceProperties(&prop, s->device));
  int occ0 = 0, occ1 = 0;
  // experiments only (tools/gpu_occupancy.py): RT_EXTRA_SMEM=<bytes> of unused dynamic shared memory
  // throttles the resident CTAs per SM
  s->extra_smem = 0;
  if (const char *e = std::getenv("RT_EXTRA_SMEM")) s->extra_smem = (size_t)std::atol(e);
  // the two kernel variants this scene uses: plain and instrumented, with or without the object-tree code
  const bool obj = h.obj_root >= 0;
  const void *k_plain = obj ? (const void *)render_kernel<K_OBJ> : (const void *)render_kernel<0>;
  const void *k_count = obj ? (const void *)render_kernel<K_COUNT | K_OBJ> : (const void *)render_kernel<K_COUNT>;
  if (s->extra_smem) {
    RT_CUDA(cudaFuncSetAttribute(k_plain, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->extra_smem));
    RT_CUDA(cudaFuncSetAttribute(k_count, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->extra_smem));
  }
  if (const char *e = std::getenv("RT_CARVEOUT")) { // experiments only: shared-memory carveout in percent
    RT_CUDA(cudaFuncSetAttribute(k_plain, cudaFuncAttributePreferredSharedMemoryCarveout, std::atoi(e)));
    RT_CUDA(cudaFuncSetAttribute(k_count, cudaFuncAttributePreferredSharedMemoryCarveout, std::atoi(e)));
  }
  RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ0, k_plain, kBlockThreads, s->extra_smem));
  RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ1, k_count, kBlockThreads, s->extra_smem));
  // the 4-wide tree is walked by its own kernel variants; RT_B200_WIDE=0 (or RT_FLAG_BINARY_NODES per frame)
  // keeps ordinary rays on the binary tree
  s->wide_ok = wide_usable(s->view);
  if (const char *e = std::getenv("RT_B200_WIDE")) if (std::atoi(e) == 0) s->wide_ok = false;
  if (s->wide_ok) {
    int occ2 = 0, occ3 = 0;
    const void *k_w = obj ? (const void *)render_kernel<K_WIDE | K_OBJ> : (const void *)render_kernel<K_WIDE>;
    const void *k_wc = obj ? (const void *)render_kernel<K_WIDE | K_OBJ | K_COUNT> : (const void *)render_kernel<K_WIDE | K_COUNT>;
    if (s->extra_smem) {
      RT_CUDA(cudaFuncSetAttribute(k_w, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->extra_smem));
      RT_CUDA(cudaFuncSetAttribute(k_wc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->extra_smem));
    }
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, k_w, kBlockThreads, s->extra_smem));
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ3, k_wc, kBlockThreads, s->extra_smem));
    s->grid_wide = prop.multiProcessorCount * (occ2 > 0 ? occ2 : 1);
    s->grid_wide_count = prop.multiProcessorCount * (occ3 > 0 ? occ3 : 1);
  }
  {
    // persistent grid of the wavefront TRACE kernel variants
    const void *kw[4] = {(const void *)wf_trace_kernel<0>, (const void *)wf_trace_kernel<K_OBJ>,
                         (const void *)wf_trace_kernel<K_WIDE>, (const void *)wf_trace_kernel<K_WIDE | K_OBJ>};
    for (int k = 0; k < 4; ++k) {
      int occ = 0;
      RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kw[k], kBlockThreads, 0));
      s->grid_wf[k] = prop.multiProcessorCount * (occ > 0 ? occ : 1);
    }
    if (const char *e = std::getenv("RT_B200_PIPELINE")) s->wavefront_default = std::strcmp(e, "wavefront") == 0;
    const void *kr[4] = {(const void *)rp_render_kernel<0>, (const void *)rp_render_kernel<K_OBJ>,
                         (const void *)rp_render_kernel<K_WIDE>, (const void *)rp_render_kernel<K_WIDE | K_OBJ>};
    for (int k = 0; k < 4; ++k) {
      int occ = 0;
      RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kr[k], kBlockThreads, 0));
      s->grid_rp[k] = prop.multiProcessorCount * (occ > 0 ? occ : 1);
    }
    if (const char *e = std::getenv("RT_B200_PIPELINE"))
      s->rp_default = std::strcmp(e, "rayparallel") == 0 ? 1 : (std::strcmp(e, "persistent") == 0 ? -1 : 0);
  }
  s->wide8_ok = s->wide_ok && wide8_usable(s->view);
  if (s->wide8_ok) {
    int o8 = 0, o8c = 0;
    const void *k8 = (const void *)render_kernel<K_WIDE | K_WIDE8>, *k8c = (const void *)render_kernel<K_WIDE | K_WIDE8 | K_COUNT>;
    if (s->extra_smem) {
      RT_CUDA(cudaFuncSetAttribute(k8, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->extra_smem));
      RT_CUDA(cudaFuncSetAttribute(k8c, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->extra_smem));
    }
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o8, k8, kBlockThreads, s->extra_smem));
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o8c, k8c, kBlockThreads, s->extra_smem));
    s->grid_wide8 = prop.multiProcessorCount * (o8 > 0 ? o8 : 1);
    s->grid_wide8_count = prop.multiProcessorCount * (o8c > 0 ? o8c : 1);
    if (const char *e = std::getenv("RT_B200_NODES")) s->wide8_default = std::atoi(e) == 8;
  }
  s->sm_count = prop.multiProcessorCount;
  s->grid_plain = prop.multiProcessorCount * (occ0 > 0 ? occ0 : 1);
  s->grid_count = prop.multiProcessorCount * (occ1 > 0 ? occ1 : 1);
  {
    int gmax = s->grid_plain > s->grid_count ? s->grid_plain : s->grid_count;
    if (s->grid_wide > gmax) gmax = s->grid_wide;
    if (s->grid_wide_count > gmax) gmax = s->grid_wide_count;
    if (s->grid_wide8 > gmax) gmax = s->grid_wide8;
    if (s->grid_wide8_count > gmax) gmax = s->grid_wide8_count;
    const size_t threads = (size_t)gmax * kBlockThreads;
    RT_CUDA(cudaMalloc(&s->d_scratch_d, threads * CD_NUM * sizeof(double)));
    RT_CUDA(cudaMalloc(&s->d_scratch_i, threads * CI_NUM * sizeof(int)));
  }
  std::memset(&s->info, 0, sizeof(s->info));
  s->info.num_triangles = h.num_tris; s->info.num_meshes = h.num_meshes; s->info.num_objects = h.num_objects;
  s->info.num_nodes_ref = h.num_nodes_ref; s->info.num_nodes_flat = h.num_nodes;
  s->info.max_depth_bvh = h.max_depth; s->info.max_leaf_size = h.max_leaf; s->info.blob_bytes = s->bytes;
  s->info.topology = h.topology;
  s->info.num_nodes_traversal = h.topology == 0 ? h.q_base : h.num_nodes;
  s->info.num_nodes_wide = h.num_nodes4;
  s->info.num_nodes_wide8 = h.num_nodes8;
  std::memset(&s->counters, 0, sizeof(s->counters));
  return RT_OK;
}

// Default choice of the ray-parallel kernel: triangle-mesh scenes with a reflecting material, mirror depth >= 2, and at most
// 64 k pixel samples per possible hit of a path (depth 3: 256 k, depth 5: 384 k).  Measured break-even (profiles/
// r02_rayparallel_ab2.log): Office (depth 3) wins at 130 k and 230 k pixels and loses at 360 k; rings (depth 5) wins at 88 k
// and 350 k and loses at 1.4 M; scenes with analytic primitives and scenes without mirrors lose at every size.
constexpr long long kRpAutoPixelsPerBounce = 65536;
#ifndef RT_RP_AUTO
#define RT_RP_AUTO 1
#endif
bool wavefront_ok(const rt_scene *s, const FrameParams &fp, const int *d_list, const int *d_list_count);
int launch_wavefront(rt_scene *s, const FrameParams &fp, const double *d_lpos, const double *d_lcol, const FrameOut &out,
                     cudaStream_t st);

int launch(rt_scene *s, const FrameParams &fp, const double *d_lpos, const double *d_lcol, const FrameOut &out,
           const int *d_list, int list_count, cudaStream_t st, const int *d_list_count = nullptr,
           unsigned long long *dyn_counter = nullptr) {
  if (((fp.flags & RT_FLAG_WAVEFRONT) || s->wavefront_default) && !(fp.flags & RT_FLAG_MEGAKERNEL) &&
      wavefront_ok(s, fp, d_list, d_list_count))
    return launch_wavefront(s, fp, d_lpos, d_lcol, out, st);
  s->last_pipeline = 0;
  KernelArgs a;
  a.sv = s->view; a.fp = fp;
  a.light_pos = d_lpos; a.light_col = d_lcol;
  a.out = out;
  a.pixel_list = d_list; a.list_count = list_count;
  a.list_count_ptr = d_list_count;
  const int slot = (int)(s->launch_seq++ % rt_scene::kRing);
  s->last_slot = slot;
  a.work_counter = s->d_work + 2 * slot;
  a.dyn_counter = (fp.static_tiles < fp.tiles_x * fp.tiles_y) ? dyn_counter : nullptr;
  a.counters = s->d_counters;
  a.scratch_d = s->d_scratch_d;
  a.scratch_i = s->d_scratch_i;
  const bool count = (fp.flags & RT_FLAG_COUNTERS) != 0;
  // The slot's counter and event pair may still belong to the launch kRing launches ago, possibly on another
  // stream: order this launch behind it, so any number of frames may be in flight (launches f and f + kRing
  // serialise; include/rt_b200.h).  RT_FLAG_COUNTERS launches share one counter block and are synchronous
  // (fetch_counters), so they never overlap each other.
  if (s->slot_used[slot]) RT_CUDA(cudaStreamWaitEvent(st, s->ev1[slot], 0));
  s->slot_used[slot] = true;
  RT_CUDA(cudaMemsetAsync(a.work_counter, 0, 2 * sizeof(unsigned long long), st));
  if (count) RT_CUDA(cudaMemsetAsync(s->d_counters, 0, sizeof(unsigned long long) * C_COUNT, st));
#ifdef RT_TAILPROBE
  {
    unsigned long long init[C_COUNT] = {0, 0, ~0ull};
    RT_CUDA(cudaMemcpyAsync(s->d_counters, init, sizeof(init), cudaMemcpyHostToDevice, st));
  }
#endif
  RT_CUDA(cudaEventRecord(s->ev0[slot], st));
  // frames in flight share the machine: with ctas_per_sm = k each launch takes k of the resident CTA slots of
  // every SM, so several persistent grids run side by side instead of queueing for each other's slots
  const bool wide = s->wide_ok && !(fp.flags & RT_FLAG_BINARY_NODES);
  const bool wide8 = wide && s->wide8_ok && ((fp.flags & RT_FLAG_NODES8) || s->wide8_default);
  int gc = wide ? s->grid_wide_count : s->grid_count, gp = wide ? s->grid_wide : s->grid_plain;
  if (wide8) { gc = s->grid_wide8_count; gp = s->grid_wide8; }
  if (s->ctas_per_sm > 0) {
    const int lim = s->sm_count * s->ctas_per_sm;
    gc = gc < lim ? gc : lim;
    gp = gp < lim ? gp : lim;
  }
  // Small frames are chains of dependent rays, not throughput: the rays of a path on neighbouring lanes (rp_render_kernel,
  // rt_wavefront.cuh).  Asked for by RT_FLAG_RAY_PARALLEL / RT_B200_PIPELINE=rayparallel, and by default for small frames of
  // mirror scenes (kRpAutoPixelsPerBounce); frames it does not cover (instrumented, pixel lists, dynamic tile queue, no light,
  // more than 30) stay on the persistent kernel, as does everything under RT_FLAG_MEGAKERNEL.
  {
    const long long samples = (long long)fp.cam.width * fp.cam.height * fp.spp_grid * fp.spp_grid;
    const bool covered = !count && !d_list && !d_list_count && fp.num_lights >= 1 && fp.num_lights <= kWfMaxLights &&
                         fp.static_tiles == fp.tiles_x * fp.tiles_y && !(fp.flags & (RT_FLAG_MEGAKERNEL | RT_FLAG_NODES8));
    const bool wanted = (fp.flags & RT_FLAG_RAY_PARALLEL) || s->rp_default > 0 ||
                        (s->rp_default == 0 && RT_RP_AUTO && s->hdr.num_objects == 0 && s->has_mirror && fp.max_depth >= 2 &&
                         samples <= kRpAutoPixelsPerBounce * (fp.max_depth + 1));
    if (covered && wanted) {
      const bool objt = s->hdr.obj_root >= 0;
      const int cfg = (objt ? 1 : 0) | (wide ? 2 : 0);
      int grid = s->grid_rp[cfg];
      if (s->ctas_per_sm > 0) { const int lim = s->sm_count * s->ctas_per_sm; grid = grid < lim ? grid : lim; }
      switch (cfg) {
        case 0: rp_render_kernel<0><<<grid, kBlockThreads, 0, st>>>(a); break;
        case 1: rp_render_kernel<K_OBJ><<<grid, kBlockThreads, 0, st>>>(a); break;
        case 2: rp_render_kernel<K_WIDE><<<grid, kBlockThreads, 0, st>>>(a); break;
        default: rp_render_kernel<K_WIDE | K_OBJ><<<grid, kBlockThreads, 0, st>>>(a); break;
      }
      RT_CUDA(cudaGetLastError());
      RT_CUDA(cudaEventRecord(s->ev1[slot], st));
      s->timed = true;
      s->last_pipeline = 2;
      return RT_OK;
    }
  }
  const bool obj = s->hdr.obj_root >= 0; // tree over analytic spheres: the kernel variant that can walk it
  // 2x2 / 4x4 sub-pixel grids: one sample per lane (K_SPL kernels, rt_kernel.cuh) unless the scene refills single lanes
  // (>= 32 analytic primitives), the frame is instrumented or walks the 8-wide records (A/B paths keep the sequential loop)
  const int s2 = fp.spp_grid * fp.spp_grid;
  const bool spl = RT_SAMPLE_LANES && !(fp.flags & RT_FLAG_SEQUENTIAL_SAMPLES) && (s2 == 4 || s2 == 16) && !count && !wide8 &&
                   s->hdr.num_objects < kPartialRefillObjects;
  if (spl) {
    switch ((obj ? K_OBJ : 0) | (wide ? K_WIDE : 0)) {
      case 0: render_kernel<K_SPL><<<gp, kBlockThreads, s->extra_smem, st>>>(a); break;
      case K_OBJ: render_kernel<K_SPL | K_OBJ><<<gp, kBlockThreads, s->extra_smem, st>>>(a); break;
      case K_WIDE: render_kernel<K_SPL | K_WIDE><<<gp, kBlockThreads, s->extra_smem, st>>>(a); break;
      default: render_kernel<K_SPL | K_WIDE | K_OBJ><<<gp, kBlockThreads, s->extra_smem, st>>>(a); break;
    }
  } else
  switch ((count ? K_COUNT : 0) | (obj ? K_OBJ : 0) | (wide ? K_WIDE : 0) | (wide8 ? K_WIDE8 : 0)) {
    case K_WIDE | K_WIDE8: render_kernel<K_WIDE | K_WIDE8><<<gp, kBlockThreads, s->extra_smem, st>>>(a); break;
    case K_WIDE | K_WIDE8 | K_COUNT: render_kernel<K_WIDE | K_WIDE8 | K_COUNT><<<gc, kBlockThreads, s->extra_smem, st>>>(a); break;
    case 0: render_kernel<0><<<gp, kBlockThreads, s->extra_smem, st>>>(a); break;
    case K_COUNT: render_kernel<K_COUNT><<<gc, kBlockThreads, s->extra_smem, st>>>(a); break;
    case K_OBJ: render_kernel<K_OBJ><<<gp, kBlockThreads, s->extra_smem, st>>>(a); break;
    case K_OBJ | K_COUNT: render_kernel<K_OBJ | K_COUNT><<<gc, kBlockThreads, s->extra_smem, st>>>(a); break;
    case K_WIDE: render_kernel<K_WIDE><<<gp, kBlockThreads, s->extra_smem, st>>>(a); break;
    case K_WIDE | K_COUNT: render_kernel<K_WIDE | K_COUNT><<<gc, kBlockThreads, s->extra_smem, st>>>(a); break;
    case K_WIDE | K_OBJ: render_kernel<K_WIDE | K_OBJ><<<gp, kBlockThreads, s->extra_smem, st>>>(a); break;
    default: render_kernel<K_WIDE | K_OBJ | K_COUNT><<<gc, kBlockThreads, s->extra_smem, st>>>(a); break;
  }
  RT_CUDA(cudaGetLastError());
  RT_CUDA(cudaEventRecord(s->ev1[slot], st));
  s->timed = true;
  return RT_OK;
}

// ---- the frame as a wavefront (rt_wavefront.cuh): TRACE / SHADE launches per bounce on one stream ----------------
constexpr int kWfCtlTasks = 32;     // task counters of one pass (max_depth + 2 TRACE launches)
constexpr int kWfCtlWords = 64;     // 8-byte words of a slot's control block: task counters, then pending[]
constexpr int kWfMaxDepth = kWfCtlTasks - 2;

// Can this frame run as a wavefront?  (The persistent megakernel renders everything else.)
bool wavefront_ok(const rt_scene *s, const FrameParams &fp, const int *d_list, const int *d_list_count) {
  return !d_list && !d_list_count && !(fp.flags & RT_FLAG_COUNTERS) && fp.num_lights <= kWfMaxLights &&
         fp.max_depth <= kWfMaxDepth && fp.static_tiles == fp.tiles_x * fp.tiles_y;
}

int launch_wavefront(rt_scene *s, const FrameParams &fp, const double *d_lpos, const double *d_lcol, const FrameOut &out,
                     cudaStream_t st) {
  const int slot = (int)(s->launch_seq++ % rt_scene::kRing);
  s->last_slot = slot;
  s->last_pipeline = 1;
  if (s->slot_used[slot]) RT_CUDA(cudaStreamWaitEvent(st, s->ev1[slot], 0)); // the slot's path state is free again
  s->slot_used[slot] = true;
  Ctx hc;
  hc.fp = fp; hc.pixel_list = nullptr; hc.list_count = 0;
  const long long npaths = work_slots(hc);
  rt_scene::WfSlot &ws = s->wf[slot];
  const int nl = fp.num_lights > 0 ? fp.num_lights : 1;
  if (npaths > ws.cap_paths || nl > ws.cap_lights) {
    RT_CUDA(cudaStreamSynchronize(st));
    if (ws.wd) cudaFree(ws.wd);
    if (ws.wi) cudaFree(ws.wi);
    if (ws.occl) cudaFree(ws.occl);
    ws.wd = nullptr; ws.wi = nullptr; ws.occl = nullptr; ws.cap_paths = 0; ws.cap_lights = 0;
    RT_CUDA(cudaMalloc(&ws.wd, sizeof(double) * WD_NUM * (size_t)npaths));
    RT_CUDA(cudaMalloc(&ws.wi, sizeof(int) * WI_NUM * (size_t)npaths));
    RT_CUDA(cudaMalloc(&ws.occl, (size_t)nl * (size_t)npaths));
    if (!ws.ctl) RT_CUDA(cudaMalloc(&ws.ctl, sizeof(unsigned long long) * kWfCtlWords));
    ws.cap_paths = npaths; ws.cap_lights = nl;
  }
  WfArgs a;
  a.sv = s->view; a.fp = fp;
  a.light_pos = d_lpos; a.light_col = d_lcol;
  a.out = out;
  a.w.wd = ws.wd; a.w.wi = ws.wi; a.w.occl = ws.occl; a.w.npaths = npaths;
  a.pending = reinterpret_cast<unsigned int *>(ws.ctl + kWfCtlTasks);
  a.counters = s->d_counters;
  const bool wide = s->wide_ok && !(fp.flags & RT_FLAG_BINARY_NODES);
  const bool obj = s->hdr.obj_root >= 0;
  const int cfg = (obj ? 1 : 0) | (wide ? 2 : 0);
  int grid = s->grid_wf[cfg];
  if (s->ctas_per_sm > 0) {
    // a share of the machine per frame in flight, in units of the megakernel's CTA slots (4 per SM)
    const long long lim = (long long)grid * s->ctas_per_sm / 4;
    if (lim >= s->sm_count && lim < grid) grid = (int)lim;
  }
  const unsigned sgrid = (unsigned)((npaths + kBlockThreads - 1) / kBlockThreads);
  RT_CUDA(cudaEventRecord(s->ev0[slot], st));
  const int s2 = fp.spp_grid * fp.spp_grid;
  for (int sample = 0; sample < s2; ++sample) {
    RT_CUDA(cudaMemsetAsync(ws.ctl, 0, sizeof(unsigned long long) * kWfCtlWords, st));
    a.sample = sample;
    for (int b = 0; b <= fp.max_depth + 1; ++b) {
      a.bounce = b;
      a.task_counter = ws.ctl + b;
      switch (cfg) {
        case 0: wf_trace_kernel<0><<<grid, kBlockThreads, 0, st>>>(a); wf_shade_kernel<0><<<sgrid, kBlockThreads, 0, st>>>(a); break;
        case 1: wf_trace_kernel<K_OBJ><<<grid, kBlockThreads, 0, st>>>(a); wf_shade_kernel<K_OBJ><<<sgrid, kBlockThreads, 0, st>>>(a); break;
        case 2: wf_trace_kernel<K_WIDE><<<grid, kBlockThreads, 0, st>>>(a); wf_shade_kernel<K_WIDE><<<sgrid, kBlockThreads, 0, st>>>(a); break;
        default: wf_trace_kernel<K_WIDE | K_OBJ><<<grid, kBlockThreads, 0, st>>>(a); wf_shade_kernel<K_WIDE | K_OBJ><<<sgrid, kBlockThreads, 0, st>>>(a); break;
      }
    }
  }
  RT_CUDA(cudaGetLastError());
  RT_CUDA(cudaEventRecord(s->ev1[slot], st));
  s->timed = true;
  return RT_OK;
}

int fetch_counters(rt_scene *s, cudaStream_t st) {
  unsigned long long h[C_COUNT];
  RT_CUDA(cudaMemcpyAsync(h, s->d_counters, sizeof(h), cudaMemcpyDeviceToHost, st));
  RT_CUDA(cudaStreamSynchronize(st));
  rt_counters &c = s->counters;
  c.rays_primary = h[C_RAY_PRIMARY]; c.rays_shadow = h[C_RAY_SHADOW]; c.rays_reflect = h[C_RAY_REFLECT];
  c.box_tests = h[C_BOX]; c.tri_filter = h[C_TRI_FILTER]; c.tri_exact = h[C_TRI_EXACT];
  c.object_tests = h[C_OBJECT]; c.shaded_hits = h[C_SHADED]; c.texel_fetches = h[C_TEXEL];
  c.nodes_visited = h[C_NODES];
  c.shadow_trivial = h[C_SHADOW_TRIVIAL];
  c.shadow_occluded = h[C_SHADOW_OCCLUDED];
  c.nodes4_visited = h[C_NODES4];
  c.nodes8_visited = h[C_NODES8];
  return RT_OK;
}

// ---- rt_render_device_static: the wavefront frame on scratch that is part of the loaded module ---------------------
// The wavefront kernels exchange rays through path-state columns in HBM (rt_wavefront.cuh).  launch_wavefront takes
// them from cudaMalloc; this variant takes them from ONE file-scope array, so the call allocates nothing.  The array
// is reserved on every GPU that loads the library, for the life of the process; every call on a device shares it and
// it keeps the last call's contents (path columns are fully rewritten by bounce 0 of each pass, the control block is
// cleared in-stream by the call).  Two calls that overlap on one GPU race on it.
constexpr long long kStaticMaxPaths = 65536;   // work slots of 32-pixel tiles: e.g. 256 x 256 pixels (any spp_grid)
constexpr int kStaticMaxLights = 8;
constexpr size_t kStaticOffCtl = 0;                                                           // kWfCtlWords u64
constexpr size_t kStaticOffWd = 512;                                                          // WD_NUM double columns
constexpr size_t kStaticOffWi = kStaticOffWd + sizeof(double) * WD_NUM * (size_t)kStaticMaxPaths; // WI_NUM int columns
constexpr size_t kStaticOffOccl = kStaticOffWi + sizeof(int) * WI_NUM * (size_t)kStaticMaxPaths;  // one byte column per light
constexpr size_t kStaticBytes = kStaticOffOccl + (size_t)kStaticMaxLights * (size_t)kStaticMaxPaths;
static_assert(sizeof(unsigned long long) * kWfCtlWords <= kStaticOffWd, "control block");
static_assert(kStaticOffWd % 256 == 0 && kStaticOffWi % 256 == 0 && kStaticOffOccl % 256 == 0, "column alignment");
static_assert(kStaticBytes <= (20u << 20), "static scratch is meant to stay modest (about 18.6 MB)");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

// In host code `g_static_scratch` is only a stand-in for the symbol: the address a kernel may use is per device and
// comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kStaticMaxDevices = 64;
unsigned char *g_static_ptr[kStaticMaxDevices] = {};
std::mutex g_static_mu;

int static_scratch_ptr(int device, unsigned char **out) {
  if (device < 0 || device >= kStaticMaxDevices) { set_error("rt_render_device_static: device ordinal out of range"); return RT_ERR_LIMIT; }
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[device]) {
    unsigned char *p = nullptr;
    RT_CUDA(cudaGetSymbolAddress((void **)&p, g_static_scratch));
    if (!p) { set_error("rt_render_device_static: cudaGetSymbolAddress returned a null address"); return RT_ERR_CUDA; }
    g_static_ptr[device] = p;
  }
  *out = g_static_ptr[device];
  return RT_OK;
}

} // namespace

extern "C" {

int rt_abi_version(void) { return RT_B200_ABI_VERSION; }
const char *rt_last_error(void) { return g_err.c_str(); }

int rt_init_device(int device) {
  std::string err;
  int st = ensure_device(err);
  if (st != RT_OK) { set_error(err); return st; }
  RT_CUDA(cudaSetDevice(device));
  RT_CUDA(cudaFree(nullptr));
  return RT_OK;
}

int rt_scene_create_from_blob(const void *blob, size_t bytes, int is_device, rt_scene **out) {
  if (!blob || !out) { set_error("rt_scene_create_from_blob: null argument"); return RT_ERR_INVALID_ARG; }
  std::string err;
  int st = ensure_device(err);
  if (st != RT_OK) { set_error(err); return st; }
  const auto t0 = std::chrono::steady_clock::now();
  BlobHeader h;
  if (bytes < sizeof(h)) { set_error("blob: too small"); return RT_ERR_INVALID_ARG; }
  if (is_device) RT_CUDA(cudaMemcpy(&h, blob, sizeof(h), cudaMemcpyDeviceToHost));
  else std::memcpy(&h, blob, sizeof(h));
  // section table: every offset/count inside `bytes` (the payload is checked on the device in adopt_blob)
  st = check_blob_header(h, bytes, err);
  if (st != RT_OK) { set_error(err); return st; }
  int device = 0;
  RT_CUDA(cudaGetDevice(&device));
  rt_scene *s = new rt_scene();
  s->device = device;
  s->bytes = bytes;
  s->hdr = h;
  cudaError_t e = cudaMalloc(&s->d_blob, bytes);
  if (e != cudaSuccess) { delete s; set_error("cudaMalloc(blob) failed"); return RT_ERR_OUT_OF_MEMORY; }
  e = cudaMemcpy(s->d_blob, blob, bytes, is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { cudaFree(s->d_blob); delete s; set_error("blob upload failed"); return RT_ERR_CUDA; }
  st = adopt_blob(s);
  if (st != RT_OK) { rt_scene_destroy(s); return st; }
  s->info.upload_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  *out = s;
  return RT_OK;
}

// Adopts a blob the device flattener built in HBM (no copy); frees it on failure.
static int scene_from_device_blob(DeviceBlob &db, rt_scene **out) {
  std::string err;
  int st = check_blob_header(db.hdr, db.bytes, err);
  if (st != RT_OK) { cudaFree(db.d_blob); db.d_blob = nullptr; set_error(err); return st; }
  int device = 0;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) { cudaFree(db.d_blob); db.d_blob = nullptr; set_error("cudaGetDevice failed"); return RT_ERR_CUDA; }
  rt_scene *s = new rt_scene();
  s->device = device;
  s->bytes = db.bytes;
  s->hdr = db.hdr;
  s->d_blob = db.d_blob;
  db.d_blob = nullptr;
  st = adopt_blob(s);
  if (st != RT_OK) { rt_scene_destroy(s); return st; }
  *out = s;
  return RT_OK;
}

int rt_scene_create(const rt_data_soa *data, const rt_bvh_nodes_soa *bvh, const rt_object *objects,
                    int num_objects, rt_scene **out) {
  // RT_B200_TOPOLOGY=reference: A/B switch for callers that cannot pass flags (the drop-in launcher)
  unsigned flags = 0;
  if (const char *e = std::getenv("RT_B200_TOPOLOGY"))
    if (std::strcmp(e, "reference") == 0) flags |= RT_SCENE_REFERENCE_TOPOLOGY;
  return rt_scene_create_ex(data, bvh, objects, num_objects, flags, out);
}

int rt_scene_create_ex(const rt_data_soa *data, const rt_bvh_nodes_soa *bvh, const rt_object *objects,
                       int num_objects, unsigned scene_flags, rt_scene **out) {
  if (!out) { set_error("rt_scene_create: null out"); return RT_ERR_INVALID_ARG; }
  std::string err;
  int st = ensure_device(err);
  if (st != RT_OK) { set_error(err); return st; }
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<unsigned char> blob;
  // the traversal tree is built on this GPU (rt_sah_gpu.cu) unless RT_SCENE_HOST_SAH / RT_B200_SAH=host asks for the host twin
  set_sah_device_builder(&build_sah_device);
  if (const char *e = std::getenv("RT_B200_SAH")) if (std::strcmp(e, "host") == 0) scene_flags |= RT_SCENE_HOST_SAH;
  bool host_flatten = (scene_flags & (RT_SCENE_HOST_FLATTEN | RT_SCENE_HOST_SAH | RT_SCENE_REFERENCE_TOPOLOGY)) != 0;
  if (const char *e = std::getenv("RT_B200_FLATTEN")) if (std::strcmp(e, "host") == 0) host_flatten = true;
  if (!host_flatten) {
    // the flattener on the GPU (rt_flatten_gpu.cu); whatever it declines or fails at goes to the host twin below
    DeviceBlob db;
    std::string derr;
    st = flatten_scene_device(data, bvh, objects, num_objects, scene_flags, db, derr);
    if (st == RT_OK) {
      const double flat_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      const FlattenStats fs = last_flatten_stats();
      st = scene_from_device_blob(db, out);
      if (st == RT_OK) {
        (*out)->info.flatten_ms = flat_ms;
        (*out)->info.sah_builder = fs.sah_builder;
        (*out)->info.sah_build_ms = fs.sah_ms;
        (*out)->info.sah_device_ms = fs.sah_device_ms;
        (*out)->info.flattener = 1;
        return RT_OK;
      }
    }
    if (std::getenv("RT_FLATTEN_TIMING")) std::fprintf(stderr, "[flatten gpu] declined (%d %s): host flattener\n", st, derr.c_str());
    cudaGetLastError();
  }
  st = flatten_scene(data, bvh, objects, num_objects, scene_flags, blob, err);
  if (st != RT_OK) { set_error(err); return st; }
  const double flat_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  const FlattenStats fs = last_flatten_stats();
  st = rt_scene_create_from_blob(blob.data(), blob.size(), 0, out);
  if (st == RT_OK) {
    (*out)->info.flatten_ms = flat_ms;
    (*out)->info.sah_builder = fs.sah_builder;
    (*out)->info.sah_build_ms = fs.sah_ms;
    (*out)->info.sah_device_ms = fs.sah_device_ms;
  }
  return st;
}

int rt_scene_set_launch_width(rt_scene *s, int ctas_per_sm) {
  if (!s || ctas_per_sm < 0) { set_error("rt_scene_set_launch_width: invalid argument"); return RT_ERR_INVALID_ARG; }
  s->ctas_per_sm = ctas_per_sm;
  return RT_OK;
}

int rt_scene_destroy(rt_scene *s) {
  if (!s) return RT_OK;
  for (cudaTextureObject_t t : s->tex) if (t) cudaDestroyTextureObject(t);
  auto F = [](void *p) { if (p) cudaFree(p); };
  F(s->d_blob); F(s->d_tex); F(s->d_work); F(s->d_counters); F(s->d_lights);
  F(s->d_scratch_d); F(s->d_scratch_i); F(s->d_rgba8); F(s->d_pixels); F(s->d_ids); F(s->d_t); F(s->d_list);
  for (int k = 0; k < rt_scene::kRing; ++k) {
    F(s->wf[k].wd); F(s->wf[k].wi); F(s->wf[k].occl); F(s->wf[k].ctl);
    if (s->ev0[k]) cudaEventDestroy(s->ev0[k]);
    if (s->ev1[k]) cudaEventDestroy(s->ev1[k]);
  }
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
  return RT_OK;
}

int rt_scene_get_info(const rt_scene *s, rt_scene_info *out) {
  if (!s || !out) { set_error("rt_scene_get_info: null argument"); return RT_ERR_INVALID_ARG; }
  *out = s->info;
  return RT_OK;
}

int rt_scene_blob_size(const rt_scene *s, size_t *bytes) {
  if (!s || !bytes) { set_error("rt_scene_blob_size: null argument"); return RT_ERR_INVALID_ARG; }
  *bytes = s->bytes;
  return RT_OK;
}

int rt_scene_blob_export(const rt_scene *s, void *dst, size_t bytes) {
  if (!s || !dst || bytes < s->bytes) { set_error("rt_scene_blob_export: bad argument"); return RT_ERR_INVALID_ARG; }
  RT_CUDA(cudaMemcpy(dst, s->d_blob, s->bytes, cudaMemcpyDeviceToHost));
  return RT_OK;
}

int rt_scene_blob_device_ptr(const rt_scene *s, void **dptr, size_t *bytes) {
  if (!s || !dptr || !bytes) { set_error("rt_scene_blob_device_ptr: null argument"); return RT_ERR_INVALID_ARG; }
  *dptr = s->d_blob;
  *bytes = s->bytes;
  return RT_OK;
}

int rt_render_device(rt_scene *s, const rt_camera *camera, const rt_vec4 *d_lights_pos,
                     const rt_vec4 *d_lights_color, int num_lights, const rt_render_params *params,
                     const rt_frame_out *out, void *stream) {
  if (!s) { set_error("rt_render_device: null scene"); return RT_ERR_INVALID_ARG; }
  FrameParams fp;
  std::string err;
  int st = make_frame_params(camera, num_lights, params, out, fp, err);
  if (st != RT_OK) { set_error(err); return st; }
  if (num_lights > 0 && (!d_lights_pos || !d_lights_color)) { set_error("rt_render_device: null lights"); return RT_ERR_INVALID_ARG; }
  FrameOut fo;
  fo.rgba8 = out->rgba8; fo.pixels = reinterpret_cast<double *>(out->pixels);
  fo.hit_id = out->hit_id; fo.hit_t = out->hit_t;
  cudaStream_t stx = static_cast<cudaStream_t>(stream);
  st = launch(s, fp, reinterpret_cast<const double *>(d_lights_pos), reinterpret_cast<const double *>(d_lights_color),
              fo, nullptr, 0, stx, nullptr, params->dyn_counter);
  if (st != RT_OK) return st;
#ifdef RT_TAILPROBE
  return fetch_counters(s, stx);
#endif
  if (fp.flags & RT_FLAG_COUNTERS) return fetch_counters(s, stx);
  return RT_OK;
}

// No counterpart in the reference: added for callers that must not allocate
// (CUDA graph capture, tight serving loops).  Calls no cudaMalloc / cudaFree and touches no per-scene launch state.
int rt_render_device_static(rt_scene *s, const rt_camera *camera, const rt_vec4 *d_lights_pos,
                            const rt_vec4 *d_lights_color, int num_lights, const rt_render_params *params,
                            const rt_frame_out *out, void *stream) {
  if (!s) { set_error("rt_render_device_static: null scene"); return RT_ERR_INVALID_ARG; }
  FrameParams fp;
  std::string err;
  int st = make_frame_params(camera, num_lights, params, out, fp, err);
  if (st != RT_OK) { set_error(err); return st; }
  if (num_lights > 0 && (!d_lights_pos || !d_lights_color)) { set_error("rt_render_device_static: null lights"); return RT_ERR_INVALID_ARG; }
  if ((fp.flags & RT_FLAG_COUNTERS) || fp.static_tiles != fp.tiles_x * fp.tiles_y) {
    set_error("rt_render_device_static: RT_FLAG_COUNTERS and the dynamic tile queue are not available on the static-scratch path");
    return RT_ERR_INVALID_ARG;
  }
  Ctx hc;
  hc.fp = fp; hc.pixel_list = nullptr; hc.list_count = 0;
  const long long npaths = work_slots(hc);
  if (npaths > kStaticMaxPaths) {
    set_error("rt_render_device_static: the frame needs " + std::to_string(npaths) + " path slots (32 per 8x4 tile of this rank's "
              "share), the static scratch holds " + std::to_string(kStaticMaxPaths) + " (e.g. 256x256 pixels); use rt_render_device");
    return RT_ERR_LIMIT;
  }
  if (num_lights > kStaticMaxLights) {
    set_error("rt_render_device_static: " + std::to_string(num_lights) + " lights, the static scratch holds shadow results of " +
              std::to_string(kStaticMaxLights) + "; use rt_render_device");
    return RT_ERR_LIMIT;
  }
  if (fp.max_depth > kWfMaxDepth) { set_error("rt_render_device_static: max_depth above 30"); return RT_ERR_LIMIT; }
  int cur = -1;
  RT_CUDA(cudaGetDevice(&cur));
  if (cur != s->device) { set_error("rt_render_device_static: the scene's device must be the current device"); return RT_ERR_INVALID_ARG; }
  unsigned char *base = nullptr;
  st = static_scratch_ptr(cur, &base);
  if (st != RT_OK) return st;
  cudaStream_t stx = static_cast<cudaStream_t>(stream);
  unsigned long long *ctl = reinterpret_cast<unsigned long long *>(base + kStaticOffCtl);
  WfArgs a;
  a.sv = s->view; a.fp = fp;
  a.light_pos = reinterpret_cast<const double *>(d_lights_pos); a.light_col = reinterpret_cast<const double *>(d_lights_color);
  a.out.rgba8 = out->rgba8; a.out.pixels = reinterpret_cast<double *>(out->pixels);
  a.out.hit_id = out->hit_id; a.out.hit_t = out->hit_t;
  a.w.wd = reinterpret_cast<double *>(base + kStaticOffWd);
  a.w.wi = reinterpret_cast<int *>(base + kStaticOffWi);
  a.w.occl = base + kStaticOffOccl;
  a.w.npaths = npaths;
  a.pending = reinterpret_cast<unsigned int *>(ctl + kWfCtlTasks);
  a.counters = nullptr;
  const bool wide = s->wide_ok && !(fp.flags & RT_FLAG_BINARY_NODES);
  const bool obj = s->hdr.obj_root >= 0;
  const int cfg = (obj ? 1 : 0) | (wide ? 2 : 0);
  const int grid = s->grid_wf[cfg];
  const unsigned sgrid = (unsigned)((npaths + kBlockThreads - 1) / kBlockThreads);
  const int s2 = fp.spp_grid * fp.spp_grid;
  for (int sample = 0; sample < s2; ++sample) {
    // what must start at zero: the task counters and pending[] of this pass (the array keeps the last call's contents)
    RT_CUDA(cudaMemsetAsync(ctl, 0, sizeof(unsigned long long) * kWfCtlWords, stx));
    a.sample = sample;
    for (int b = 0; b <= fp.max_depth + 1; ++b) {
      a.bounce = b;
      a.task_counter = ctl + b;
      switch (cfg) {
        case 0: wf_trace_kernel<0><<<grid, kBlockThreads, 0, stx>>>(a); wf_shade_kernel<0><<<sgrid, kBlockThreads, 0, stx>>>(a); break;
        case 1: wf_trace_kernel<K_OBJ><<<grid, kBlockThreads, 0, stx>>>(a); wf_shade_kernel<K_OBJ><<<sgrid, kBlockThreads, 0, stx>>>(a); break;
        case 2: wf_trace_kernel<K_WIDE><<<grid, kBlockThreads, 0, stx>>>(a); wf_shade_kernel<K_WIDE><<<sgrid, kBlockThreads, 0, stx>>>(a); break;
        default: wf_trace_kernel<K_WIDE | K_OBJ><<<grid, kBlockThreads, 0, stx>>>(a); wf_shade_kernel<K_WIDE | K_OBJ><<<sgrid, kBlockThreads, 0, stx>>>(a); break;
      }
    }
  }
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

int rt_render(rt_scene *s, const rt_camera *camera, const rt_vec4 *lights_pos, const rt_vec4 *lights_color,
              int num_lights, const rt_render_params *params, const rt_frame_out *out) {
  if (!s) { set_error("rt_render: null scene"); return RT_ERR_INVALID_ARG; }
  FrameParams fp;
  std::string err;
  int st = make_frame_params(camera, num_lights, params, out, fp, err);
  if (st != RT_OK) { set_error(err); return st; }
  if (num_lights > 0 && (!lights_pos || !lights_color)) { set_error("rt_render: null lights"); return RT_ERR_INVALID_ARG; }
  const size_t npx = (size_t)camera->width * camera->height;
  cudaStream_t stx = s->stream;
  // lights: host -> device every frame (copyLights, mytracer.cpp:105-118)
  if (num_lights > s->lights_cap) {
    if (s->d_lights) cudaFree(s->d_lights);
    s->d_lights = nullptr;
    RT_CUDA(cudaMalloc(&s->d_lights, sizeof(double) * 8 * num_lights));
    s->lights_cap = num_lights;
  }
  if (num_lights > 0) {
    RT_CUDA(cudaMemcpyAsync(s->d_lights, lights_pos, sizeof(double) * 4 * num_lights, cudaMemcpyHostToDevice, stx));
    RT_CUDA(cudaMemcpyAsync(s->d_lights + 4 * (size_t)s->lights_cap, lights_color, sizeof(double) * 4 * num_lights,
                            cudaMemcpyHostToDevice, stx));
  }
  FrameOut fo;
  std::memset(&fo, 0, sizeof(fo));
  const bool partial = fp.tile_world > 1 && !(fp.flags & RT_FLAG_NO_CLEAR);
  if (out->rgba8) {
    st = ensure_cap(s->d_rgba8, s->cap_rgba8, npx * 4, stx); if (st != RT_OK) return st;
    fo.rgba8 = s->d_rgba8;
    if (partial) RT_CUDA(cudaMemsetAsync(s->d_rgba8, 0, npx * 4, stx));
  }
  if (out->pixels) {
    st = ensure_cap(s->d_pixels, s->cap_pixels, npx * 32, stx); if (st != RT_OK) return st;
    fo.pixels = s->d_pixels;
    if (partial) RT_CUDA(cudaMemsetAsync(s->d_pixels, 0, npx * 32, stx));
  }
  if (out->hit_id) {
    st = ensure_cap(s->d_ids, s->cap_ids, npx * 4, stx); if (st != RT_OK) return st;
    fo.hit_id = s->d_ids;
    if (partial) RT_CUDA(cudaMemsetAsync(s->d_ids, 0xff, npx * 4, stx));
  }
  if (out->hit_t) {
    st = ensure_cap(s->d_t, s->cap_t, npx * 8, stx); if (st != RT_OK) return st;
    fo.hit_t = s->d_t;
    if (partial) RT_CUDA(cudaMemsetAsync(s->d_t, 0, npx * 8, stx));
  }
  st = launch(s, fp, s->d_lights, s->d_lights + 4 * (size_t)s->lights_cap, fo, nullptr, 0, stx, nullptr, params->dyn_counter);
  if (st != RT_OK) return st;
  if (out->rgba8) RT_CUDA(cudaMemcpyAsync(out->rgba8, s->d_rgba8, npx * 4, cudaMemcpyDeviceToHost, stx));
  if (out->pixels) RT_CUDA(cudaMemcpyAsync(out->pixels, s->d_pixels, npx * 32, cudaMemcpyDeviceToHost, stx));
  if (out->hit_id) RT_CUDA(cudaMemcpyAsync(out->hit_id, s->d_ids, npx * 4, cudaMemcpyDeviceToHost, stx));
  if (out->hit_t) RT_CUDA(cudaMemcpyAsync(out->hit_t, s->d_t, npx * 8, cudaMemcpyDeviceToHost, stx));
  if (fp.flags & RT_FLAG_COUNTERS) {
    st = fetch_counters(s, stx);
    if (st != RT_OK) return st;
  }
  RT_CUDA(cudaStreamSynchronize(stx));
  return RT_OK;
}

int rt_shared_alloc(size_t bytes, void **dptr) {
  if (!dptr || bytes == 0) { set_error("rt_shared_alloc: bad argument"); return RT_ERR_INVALID_ARG; }
  std::string err;
  int st = ensure_device(err);
  if (st != RT_OK) { set_error(err); return st; }
  RT_CUDA(cudaMalloc(dptr, bytes));   // a whole allocation of its own: cudaIpcGetMemHandle exports allocations
  RT_CUDA(cudaMemset(*dptr, 0, bytes));
  return RT_OK;
}
int rt_shared_free(void *dptr) {
  if (dptr) RT_CUDA(cudaFree(dptr));
  return RT_OK;
}
int rt_ipc_export(void *dptr, unsigned char handle[RT_IPC_HANDLE_BYTES]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "IPC handle size");
  if (!dptr || !handle) { set_error("rt_ipc_export: null argument"); return RT_ERR_INVALID_ARG; }
  cudaIpcMemHandle_t h;
  RT_CUDA(cudaIpcGetMemHandle(&h, dptr));
  std::memcpy(handle, &h, sizeof(h));
  return RT_OK;
}
int rt_ipc_open(const unsigned char handle[RT_IPC_HANDLE_BYTES], void **dptr) {
  if (!dptr || !handle) { set_error("rt_ipc_open: null argument"); return RT_ERR_INVALID_ARG; }
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof(h));
  // maps the exporting rank's allocation into this process; peer access over NVLink is enabled lazily
  RT_CUDA(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
  return RT_OK;
}
int rt_ipc_close(void *dptr) {
  if (dptr) RT_CUDA(cudaIpcCloseMemHandle(dptr));
  return RT_OK;
}

int rt_get_counters(const rt_scene *s, rt_counters *out) {
  if (!s || !out) { set_error("rt_get_counters: null argument"); return RT_ERR_INVALID_ARG; }
  *out = s->counters;
  return RT_OK;
}

int rt_last_pipeline(const rt_scene *s, int *pipeline) {
  if (!s || !pipeline) { set_error("rt_last_pipeline: null argument"); return RT_ERR_INVALID_ARG; }
  *pipeline = s->last_pipeline;
  return RT_OK;
}

int rt_last_kernel_ms(const rt_scene *s, float *ms) {
  if (!s || !ms) { set_error("rt_last_kernel_ms: null argument"); return RT_ERR_INVALID_ARG; }
  if (!s->timed) { set_error("rt_last_kernel_ms: nothing rendered yet"); return RT_ERR_INVALID_ARG; }
  RT_CUDA(cudaEventSynchronize(s->ev1[s->last_slot]));
  RT_CUDA(cudaEventElapsedTime(ms, s->ev0[s->last_slot], s->ev1[s->last_slot]));
  return RT_OK;
}

int rt_launch_compute_image(rt_vec4 *d_pixels, rt_vec4 *d_tmpPixels, rt_vec4 *h_image, int width, int height,
                            const rt_camera *camera, const rt_vec4 *d_lightsPos, const rt_vec4 *d_lightsColor,
                            int nLights, const rt_vec4 *background, const rt_vec4 *ambience, int max_depth,
                            const rt_data_soa *data, const rt_bvh_nodes_soa *bvhNodes, rt_scene **scene_cache) {
  if (!d_pixels || !d_tmpPixels || !camera || !background || !ambience || !scene_cache) {
    set_error("rt_launch_compute_image: null argument");
    return RT_ERR_INVALID_ARG;
  }
  if (width != camera->width || height != camera->height) {
    set_error("rt_launch_compute_image: width/height differ from the camera's");
    return RT_ERR_INVALID_ARG;
  }
  int st;
  if (!*scene_cache) {
    st = rt_scene_create(data, bvhNodes, nullptr, 0, scene_cache);
    if (st != RT_OK) return st;
  }
  rt_scene *s = *scene_cache;
  cudaStream_t stx = s->stream;
  const size_t npx = (size_t)width * height;
  rt_render_params p;
  std::memset(&p, 0, sizeof(p));
  p.background = *background; p.ambience = *ambience; p.max_depth = max_depth;
  p.spp_grid = 1; p.tile_world = 1; p.tile_chunk = 1;
  rt_frame_out fo;
  std::memset(&fo, 0, sizeof(fo));
  fo.pixels = d_pixels;
  // pass 1: compute_image_device (mytracer_gpu.cu:66-77)
  st = rt_render_device(s, camera, d_lightsPos, d_lightsColor, nLights, &p, &fo, stx);
  if (st != RT_OK) return st;
  // pass 2: adaptive supersampling, subp = 4, threshold = 0.02 (mytracer_gpu.cu:83-105)
  RT_CUDA(cudaMemcpyAsync(d_tmpPixels, d_pixels, npx * sizeof(rt_vec4), cudaMemcpyDeviceToDevice, stx));
  st = ensure_cap(s->d_list, s->cap_list, (npx + 1) * sizeof(int), stx);
  if (st != RT_OK) return st;
  int *d_count = s->d_list + npx;
  RT_CUDA(cudaMemsetAsync(d_count, 0, sizeof(int), stx));
  const dim3 blk(32, 8), grd((width + 31) / 32, (height + 7) / 8);
  flag_contrast_kernel<<<grd, blk, 0, stx>>>(reinterpret_cast<const double *>(d_tmpPixels), width, height, 0.02,
                                              s->d_list, d_count);
  RT_CUDA(cudaGetLastError());
  FrameParams fp;
  std::string err;
  p.spp_grid = 4;
  st = make_frame_params(camera, nLights, &p, &fo, fp, err);
  if (st != RT_OK) { set_error(err); return st; }
  FrameOut ko;
  std::memset(&ko, 0, sizeof(ko));
  ko.pixels = reinterpret_cast<double *>(d_pixels);
  st = launch(s, fp, reinterpret_cast<const double *>(d_lightsPos), reinterpret_cast<const double *>(d_lightsColor),
              ko, s->d_list, 0, stx, d_count);
  if (st != RT_OK) return st;
  // copy to host (mytracer_gpu.cu:112)
  if (h_image) RT_CUDA(cudaMemcpyAsync(h_image, d_pixels, npx * sizeof(rt_vec4), cudaMemcpyDeviceToHost, stx));
  RT_CUDA(cudaStreamSynchronize(stx));
  return RT_OK;
}

// ---- pipelined host-buffer frames (replaces the body of Raytracer::compute_image_cuda, mytracer.cpp:123-159,
// for a host that renders a sequence of frames: copyLights :105-118, the launch, the D2H copy mytracer_gpu.cu:112) ----
struct rt_frame_queue {
  rt_scene *scene = nullptr;
  int width = 0, height = 0, depth = 2, ctas_per_sm = 0, zero_copy = 0;
  size_t flush_bytes = 0;
  unsigned char *d_flush = nullptr;
  struct Slot {
    cudaStream_t compute = nullptr;
    uint8_t *d_rgba8 = nullptr;
    double *d_lights = nullptr;   // pos[4*cap] then col[4*cap]
    double *h_lights = nullptr;   // pinned staging (the caller may reuse its arrays as soon as submit returns)
    int lights_cap = 0;
    cudaEvent_t kernel_done = nullptr;
    long long ticket = -1;        // frame in this slot (-1: free)
  };
  std::vector<Slot> slots;
  cudaStream_t copy = nullptr;    // D2H engine stream: the copy of frame f overlaps the kernel of frame f + 1
  static constexpr int kTimeRing = 64;
  cudaEvent_t done[kTimeRing] = {}; // completion (after the D2H copy) of ticket t lives in done[t % kTimeRing]
  cudaEvent_t origin = nullptr;     // recorded before the first frame: zero of the completion timeline
  long long submitted = 0, retired = 0;
  double last_done_ms = 0.0;        // completion time of ticket retired - 1 on the queue's timeline
  std::vector<double> done_ms;      // [ticket % 4096]
};

namespace {
int retire_upto(rt_frame_queue *q, long long ticket) {
  while (q->retired <= ticket) {
    const long long t = q->retired;
    cudaEvent_t ev = q->done[t % rt_frame_queue::kTimeRing];
    RT_CUDA(cudaEventSynchronize(ev));
    // completion time on the queue's timeline = CUDA-event time since the first submit (frames on different
    // streams may complete out of ticket order, so each is measured against the origin, not against its predecessor)
    float ms = 0.0f;
    RT_CUDA(cudaEventElapsedTime(&ms, q->origin, ev));
    q->last_done_ms = (double)ms;
    q->done_ms[t % 4096] = q->last_done_ms;
    q->slots[t % q->depth].ticket = -1;
    q->retired = t + 1;
  }
  return RT_OK;
}
} // namespace

int rt_frame_queue_create(rt_scene *scene, int width, int height, const rt_frame_queue_opts *opts, rt_frame_queue **out) {
  if (!scene || !out || width <= 0 || height <= 0) { set_error("rt_frame_queue_create: invalid argument"); return RT_ERR_INVALID_ARG; }
  rt_frame_queue *q = new rt_frame_queue();
  q->scene = scene; q->width = width; q->height = height;
  if (opts) {
    q->depth = opts->depth > 0 ? opts->depth : 2;
    q->ctas_per_sm = opts->ctas_per_sm > 0 ? opts->ctas_per_sm : 0;
    q->flush_bytes = opts->flush_bytes;
    q->zero_copy = opts->zero_copy != 0;
  }
  // two frames of the timeline ring must never share an event while both are unretired
  if (q->depth > rt_frame_queue::kTimeRing / 2) { delete q; set_error("rt_frame_queue_create: depth > 32"); return RT_ERR_LIMIT; }
  q->done_ms.assign(4096, 0.0);
  q->slots.resize(q->depth);
  auto fail = [&](int st) { rt_frame_queue_destroy(q); return st; };
#define RT_Q(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { set_error(std::string(#call) + ": " + cudaGetErrorString(e_)); \
                        return fail(e_ == cudaErrorMemoryAllocation ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA); } } while (0)
  RT_Q(cudaStreamCreateWithFlags(&q->copy, cudaStreamNonBlocking));
  RT_Q(cudaEventCreate(&q->origin));
  for (int k = 0; k < rt_frame_queue::kTimeRing; ++k) RT_Q(cudaEventCreate(&q->done[k]));
  for (auto &sl : q->slots) {
    RT_Q(cudaStreamCreateWithFlags(&sl.compute, cudaStreamNonBlocking));
    RT_Q(cudaEventCreateWithFlags(&sl.kernel_done, cudaEventDisableTiming));
    if (!q->zero_copy) RT_Q(cudaMalloc(&sl.d_rgba8, (size_t)width * height * 4));
  }
  if (q->flush_bytes) RT_Q(cudaMalloc(&q->d_flush, q->flush_bytes));
#undef RT_Q
  *out = q;
  return RT_OK;
}

int rt_frame_submit(rt_frame_queue *q, const rt_camera *camera, const rt_vec4 *lights_pos, const rt_vec4 *lights_color,
                    int num_lights, const rt_render_params *params, uint8_t *rgba8_host, long long *ticket) {
  if (!q || !camera || !params || !rgba8_host) { set_error("rt_frame_submit: null argument"); return RT_ERR_INVALID_ARG; }
  if (camera->width != q->width || camera->height != q->height) { set_error("rt_frame_submit: camera size differs from the queue's"); return RT_ERR_INVALID_ARG; }
  if (num_lights > 0 && (!lights_pos || !lights_color)) { set_error("rt_frame_submit: null lights"); return RT_ERR_INVALID_ARG; }
  const long long t = q->submitted;
  rt_frame_queue::Slot &sl = q->slots[t % q->depth];
  int st;
  if (sl.ticket >= 0) { st = retire_upto(q, sl.ticket); if (st != RT_OK) return st; } // at most `depth` frames in flight
  if (num_lights > sl.lights_cap) {
    if (sl.d_lights) cudaFree(sl.d_lights);
    if (sl.h_lights) cudaFreeHost(sl.h_lights);
    sl.d_lights = nullptr; sl.h_lights = nullptr; sl.lights_cap = 0;
    RT_CUDA(cudaMalloc(&sl.d_lights, sizeof(double) * 8 * num_lights));
    RT_CUDA(cudaHostAlloc(&sl.h_lights, sizeof(double) * 8 * num_lights, cudaHostAllocDefault));
    sl.lights_cap = num_lights;
  }
  if (t == 0) RT_CUDA(cudaEventRecord(q->origin, sl.compute));
  if (q->d_flush) RT_CUDA(cudaMemsetAsync(q->d_flush, 1, q->flush_bytes, sl.compute)); // benchmark aid: evict L2
  double *d_lcol = sl.d_lights + 4 * (size_t)sl.lights_cap;
  if (num_lights > 0) {
    std::memcpy(sl.h_lights, lights_pos, sizeof(double) * 4 * num_lights);
    std::memcpy(sl.h_lights + 4 * (size_t)sl.lights_cap, lights_color, sizeof(double) * 4 * num_lights);
    RT_CUDA(cudaMemcpyAsync(sl.d_lights, sl.h_lights, sizeof(double) * 8 * sl.lights_cap, cudaMemcpyHostToDevice, sl.compute));
  }
  rt_frame_out fo;
  std::memset(&fo, 0, sizeof(fo));
  if (q->zero_copy) {
    // the kernel stores every finished pixel straight into the caller's pinned, mapped host frame (one frame
    // shared by all ranks' processes in the multi-GPU case): no staging buffer, no copy pass
    void *dptr = nullptr;
    RT_CUDA(cudaHostGetDevicePointer(&dptr, rgba8_host, 0));
    fo.rgba8 = static_cast<uint8_t *>(dptr);
  } else {
    fo.rgba8 = sl.d_rgba8;
  }
  const int saved = q->scene->ctas_per_sm;
  q->scene->ctas_per_sm = q->ctas_per_sm;
  st = rt_render_device(q->scene, camera, reinterpret_cast<const rt_vec4 *>(sl.d_lights), reinterpret_cast<const rt_vec4 *>(d_lcol),
                        num_lights, params, &fo, sl.compute);
  q->scene->ctas_per_sm = saved;
  if (st != RT_OK) return st;
  cudaEvent_t done = q->done[t % rt_frame_queue::kTimeRing];
  if (q->zero_copy) {
    RT_CUDA(cudaEventRecord(done, sl.compute));
  } else {
    RT_CUDA(cudaEventRecord(sl.kernel_done, sl.compute));
    RT_CUDA(cudaStreamWaitEvent(q->copy, sl.kernel_done, 0));
    RT_CUDA(cudaMemcpyAsync(rgba8_host, sl.d_rgba8, (size_t)q->width * q->height * 4, cudaMemcpyDeviceToHost, q->copy));
    RT_CUDA(cudaEventRecord(done, q->copy));
    RT_CUDA(cudaStreamWaitEvent(sl.compute, done, 0)); // the slot's next kernel must not overwrite d_rgba8 under the copy
  }
  sl.ticket = t;
  q->submitted = t + 1;
  if (ticket) *ticket = t;
  return RT_OK;
}

int rt_frame_queue_bind(rt_frame_queue *q, rt_scene *scene) {
  if (!q || !scene) { set_error("rt_frame_queue_bind: null argument"); return RT_ERR_INVALID_ARG; }
  q->scene = scene;   // frames already submitted keep the scene they were launched on
  return RT_OK;
}

int rt_frame_wait(rt_frame_queue *q, long long ticket, double *done_ms) {
  if (!q || ticket < 0 || ticket >= q->submitted) { set_error("rt_frame_wait: unknown ticket"); return RT_ERR_INVALID_ARG; }
  if (ticket + 4096 <= q->retired) { set_error("rt_frame_wait: ticket too old"); return RT_ERR_INVALID_ARG; }
  int st = retire_upto(q, ticket);
  if (st != RT_OK) return st;
  if (done_ms) *done_ms = q->done_ms[ticket % 4096];
  return RT_OK;
}

int rt_frame_queue_destroy(rt_frame_queue *q) {
  if (!q) return RT_OK;
  if (q->submitted > 0) retire_upto(q, q->submitted - 1);
  for (auto &sl : q->slots) {
    if (sl.compute) { cudaStreamSynchronize(sl.compute); cudaStreamDestroy(sl.compute); }
    if (sl.kernel_done) cudaEventDestroy(sl.kernel_done);
    if (sl.d_rgba8) cudaFree(sl.d_rgba8);
    if (sl.d_lights) cudaFree(sl.d_lights);
    if (sl.h_lights) cudaFreeHost(sl.h_lights);
  }
  if (q->copy) { cudaStreamSynchronize(q->copy); cudaStreamDestroy(q->copy); }
  for (int k = 0; k < rt_frame_queue::kTimeRing; ++k) if (q->done[k]) cudaEventDestroy(q->done[k]);
  if (q->origin) cudaEventDestroy(q->origin);
  if (q->d_flush) cudaFree(q->d_flush);
  delete q;
  return RT_OK;
}

// ---- pinned / mapped host memory for frame buffers --------------------------------------------------------
int rt_host_alloc(size_t bytes, void **p) {
  if (!p || bytes == 0) { set_error("rt_host_alloc: bad argument"); return RT_ERR_INVALID_ARG; }
  std::string err;
  int st = ensure_device(err);
  if (st != RT_OK) { set_error(err); return st; }
  RT_CUDA(cudaHostAlloc(p, bytes, cudaHostAllocPortable | cudaHostAllocMapped));
  return RT_OK;
}
int rt_host_free(void *p) {
  if (p) RT_CUDA(cudaFreeHost(p));
  return RT_OK;
}
int rt_host_register(void *p, size_t bytes, void **dptr) {
  if (!p || bytes == 0) { set_error("rt_host_register: bad argument"); return RT_ERR_INVALID_ARG; }
  std::string err;
  int st = ensure_device(err);
  if (st != RT_OK) { set_error(err); return st; }
  RT_CUDA(cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped));
  if (dptr) RT_CUDA(cudaHostGetDevicePointer(dptr, p, 0));
  return RT_OK;
}
int rt_host_unregister(void *p) {
  if (p) RT_CUDA(cudaHostUnregister(p));
  return RT_OK;
}

} // extern "C"
