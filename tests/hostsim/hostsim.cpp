// TEST INFRASTRUCTURE — host-compiled build of the kernel core (rt_core.cuh).
//
// This container has no GPU; this library lets the CPU test-suite run the SAME per-lane state
// machine, flattener and host prep code that the CUDA kernel uses, one lane at a time, and compare
// it with the oracle before GPU minutes are spent.  It is NOT a product path: nothing in
// my-raytracer_b200/ loads it, the C-ABI library never falls back to it, and bench.py never
// times it.
#include "../../my-raytracer_b200/csrc/rt_wavefront.cuh"
#include "../../my-raytracer_b200/csrc/rt_flatten.h"
#include "../../my-raytracer_b200/csrc/rt_frame.h"
#include <cstdlib>
#include <omp.h>
#include <vector>

namespace rtb {
static thread_local std::string g_err;
void set_error(const std::string &m) { g_err = m; }
} // namespace rtb

extern "C" {

const char *sim_last_error(void) { return rtb::g_err.c_str(); }

int sim_flatten(const rt_data_soa *data, const rt_bvh_nodes_soa *bvh, const rt_object *objects, int nobj,
                unsigned flags, unsigned char **blob_out, size_t *bytes) {
  std::vector<unsigned char> blob;
  std::string err;
  const int st = rtb::flatten_scene(data, bvh, objects, nobj, flags, blob, err);
  if (st != RT_OK) { rtb::set_error(err); return st; }
  *blob_out = (unsigned char *)std::malloc(blob.size());
  std::memcpy(*blob_out, blob.data(), blob.size());
  *bytes = blob.size();
  return RT_OK;
}
void sim_free(void *p) { std::free(p); }

int sim_render(const unsigned char *blob, size_t bytes, const rt_camera *cam, const rt_vec4 *lpos,
               const rt_vec4 *lcol, int nlights, const rt_render_params *params, const rt_frame_out *out,
               unsigned long long *counters, int nthreads, unsigned *cost_nodes, unsigned *cost_tris) {
  using namespace rtb;
  std::string err;
  int st = check_blob(blob, bytes, err);
  if (st != RT_OK) { set_error(err); return st; }
  Ctx c;
  st = make_frame_params(cam, nlights, params, out, c.fp, err);
  if (st != RT_OK) { set_error(err); return st; }
  c.sv = view_of_blob(blob);
  c.light_pos = reinterpret_cast<const double *>(lpos);
  c.light_col = reinterpret_cast<const double *>(lcol);
  c.pixel_list = nullptr;
  c.list_count = 0;
  c.out.rgba8 = out->rgba8;
  c.out.pixels = reinterpret_cast<double *>(out->pixels);
  c.out.hit_id = out->hit_id;
  c.out.hit_t = out->hit_t;
  const long long slots = work_slots(c);
  // dynamic tail queue: without a counter shared across processes the host build hands pool chunk q to
  // rank q % world (any assignment that covers every chunk exactly once renders the same frame)
  const long long nchunks = pool_chunks(c);
  const long long chunk_slots = (long long)c.fp.dyn_chunk * (kTileW * kTileH);
  const long long total_slots = slots + nchunks * chunk_slots;
  if (nthreads > 0) omp_set_num_threads(nthreads);
  std::vector<unsigned long long> total(C_COUNT, 0);
#pragma omp parallel
  {
    unsigned int cnt[C_COUNT] = {0};
    unsigned long long mine[C_COUNT] = {0};
    Ctx ct = c;
    ct.cnt = cnt;
    StackEntry stk[kStackMax];
    double cold_d[CD_NUM];
    int cold_i[CI_NUM];
    Cold<1> Cd;
    Cd.db = cold_d;
    Cd.ib = cold_i;
#pragma omp for schedule(dynamic, 256)
    for (long long w = 0; w < total_slots; ++w) {
      int px, py;
      if (w < slots) {
        if (!work_to_pixel(ct, w, px, py)) continue;
      } else {
        const long long s = w - slots;
        if ((s / chunk_slots) % ct.fp.tile_world != ct.fp.tile_rank) continue;
        if (!pool_to_pixel(ct, s, px, py)) continue;
      }
      Hot H;
      H.mode = M_IDLE; H.node = -1; H.sp = 0; H.found = 0; H.best_id = -1; H.special = 0;
      // the same choice of kernel variant as rt_api.cu launch(): the 4-wide twins when every tree of the scene has one
      // and the frame does not ask for the binary records
      const bool wide = wide_usable(ct.sv) && !(ct.fp.flags & RT_FLAG_BINARY_NODES);
      const bool wide8 = wide && (ct.fp.flags & RT_FLAG_NODES8) && wide8_usable(ct.sv);
      if (wide8) {
        begin_pixel<K_COUNT | K_WIDE | K_WIDE8>(ct, H, Cd, px, py);
        while (H.mode != M_IDLE) {
          if (H.node != -1) traverse<K_COUNT | K_WIDE | K_WIDE8>(ct, H, Cd, stk);
          else advance<K_COUNT | K_WIDE | K_WIDE8>(ct, H, Cd);
        }
      } else if (wide) {
        begin_pixel<K_COUNT | K_OBJ | K_WIDE>(ct, H, Cd, px, py);
        while (H.mode != M_IDLE) {
          if (H.node != -1) traverse<K_COUNT | K_OBJ | K_WIDE>(ct, H, Cd, stk);
          else advance<K_COUNT | K_OBJ | K_WIDE>(ct, H, Cd);
        }
      } else {
        begin_pixel<K_COUNT | K_OBJ>(ct, H, Cd, px, py);
        while (H.mode != M_IDLE) {
          if (H.node != -1) traverse<K_COUNT | K_OBJ>(ct, H, Cd, stk);
          else advance<K_COUNT | K_OBJ>(ct, H, Cd);
        }
      }
      if (cost_nodes) cost_nodes[(size_t)py * ct.fp.cam.width + px] = cnt[C_NODES];
      if (cost_tris) cost_tris[(size_t)py * ct.fp.cam.width + px] = cnt[C_TRI_FILTER];
      for (int k = 0; k < C_COUNT; ++k) { mine[k] += cnt[k]; cnt[k] = 0; }
    }
#pragma omp critical
    for (int k = 0; k < C_COUNT; ++k) total[k] += mine[k];
  }
  if (counters) for (int k = 0; k < C_COUNT; ++k) counters[k] = total[k];
  return RT_OK;
}

// The same frame through the WAVEFRONT schedule (rt_wavefront.cuh): the stage functions the CUDA kernels call
// (wf_ray_start / traverse / wf_ray_finish, wf_shade), one path at a time, in the launch order of rt_api.cu.
int sim_render_wavefront(const unsigned char *blob, size_t bytes, const rt_camera *cam, const rt_vec4 *lpos,
                         const rt_vec4 *lcol, int nlights, const rt_render_params *params, const rt_frame_out *out,
                         unsigned long long *counters) {
  using namespace rtb;
  std::string err;
  int st = check_blob(blob, bytes, err);
  if (st != RT_OK) { set_error(err); return st; }
  Ctx c;
  st = make_frame_params(cam, nlights, params, out, c.fp, err);
  if (st != RT_OK) { set_error(err); return st; }
  if (nlights > kWfMaxLights || c.fp.static_tiles != c.fp.tiles_x * c.fp.tiles_y) { set_error("wavefront: unsupported frame"); return RT_ERR_LIMIT; }
  c.sv = view_of_blob(blob);
  c.light_pos = reinterpret_cast<const double *>(lpos);
  c.light_col = reinterpret_cast<const double *>(lcol);
  c.pixel_list = nullptr;
  c.list_count = 0;
  c.out.rgba8 = out->rgba8;
  c.out.pixels = reinterpret_cast<double *>(out->pixels);
  c.out.hit_id = out->hit_id;
  c.out.hit_t = out->hit_t;
  unsigned int cnt[C_COUNT] = {0};
  c.cnt = cnt;
  const long long np = work_slots(c);
  std::vector<double> wd((size_t)WD_NUM * np, 0.0);
  std::vector<int> wi((size_t)WI_NUM * np, 0);
  std::vector<unsigned char> occl((size_t)(nlights > 0 ? nlights : 1) * np, 0);
  WfState w{wd.data(), wi.data(), occl.data(), np};
  const bool wide = wide_usable(c.sv) && !(c.fp.flags & RT_FLAG_BINARY_NODES);
  StackEntry stk[kStackMax];
  double tcold[CD_BEST_T + 1];
  Cold<1> T; T.db = tcold; T.ib = nullptr;
  auto trace = [&](int kind, int sample) {
    for (long long p = 0; p < np; ++p) {
      Hot H;
      H.mode = M_IDLE; H.node = -1; H.sp = 0; H.found = 0; H.best_id = -1; H.special = 0;
      bool has;
      if (wide) {
        has = wf_ray_start<K_COUNT | K_OBJ | K_WIDE>(c, w, p, kind, sample, H, T);
        if (has && H.node != -1) traverse<K_COUNT | K_OBJ | K_WIDE>(c, H, T, stk);
      } else {
        has = wf_ray_start<K_COUNT | K_OBJ>(c, w, p, kind, sample, H, T);
        if (has && H.node != -1) traverse<K_COUNT | K_OBJ>(c, H, T, stk);
      }
      if (has) wf_ray_finish(c, w, p, kind, H, T);
    }
  };
  auto shade = [&]() {
    long long more = 0;
    for (long long p = 0; p < np; ++p) more += wf_shade<K_COUNT | K_OBJ | K_WIDE>(c, w, p) ? 1 : 0;
    return more;
  };
  const int s2 = c.fp.spp_grid * c.fp.spp_grid;
  for (int sample = 0; sample < s2; ++sample) {
    trace(-1, sample);
    long long more = shade();
    for (int b = 1; b <= c.fp.max_depth + 1 && more; ++b) {
      for (int k = 0; k <= nlights; ++k) trace(k, sample);
      more = shade();
    }
    if (more) { set_error("wavefront: paths left after max_depth + 1 bounces"); return RT_ERR_LIMIT; }
  }
  if (counters) for (int k = 0; k < C_COUNT; ++k) counters[k] = cnt[k];
  return RT_OK;
}

// direct access to the primitives for property tests
int sim_tri_filter(const double *p0, const double *p1, const double *p2, const double *o, const double *d,
                   double thi) {
  using namespace rtb;
  // build the record exactly like the flattener does
  TriFilter f;
  const double e1[3] = {p0[0] - p2[0], p0[1] - p2[1], p0[2] - p2[2]};
  const double e2[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
  const double a = std::sqrt(e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2]);
  const double b = std::sqrt(e2[0] * e2[0] + e2[1] * e2[1] + e2[2] * e2[2]);
  for (int k = 0; k < 3; ++k) { f.p2[k] = (float)p2[k]; f.e1[k] = (float)e1[k]; f.e2[k] = (float)e2[k]; }
  const double u = 5.9604644775390625e-8, tiny = 1e-30;
  auto up = [](double v) { float x = (float)v; if ((double)x < v) x = nextafterf(x, INFINITY); return x; };
  f.k_bary = up(std::fmax(8.0 * u * std::fmax(a, b), tiny));
  f.k_t = up(std::fmax(8.0 * u * a * b, tiny));
  f.k_det = up(std::fmax(64.0 * u * a * b, tiny));
  RayF r;
  rayf_setup(r, ld3(o), ld3(d), 0.0f);
  return tri_filter(r, f, thi_of(thi)) ? 1 : 0;
}

int sim_tri_exact(const double *p0, const double *p1, const double *p2, const double *o, const double *d,
                  double *out4) {
  using namespace rtb;
  TriHit h;
  const bool hit = tri_exact(ld3(o), ld3(d), ld3(p0), ld3(p1), ld3(p2), h);
  if (hit) { out4[0] = h.t; out4[1] = h.alpha; out4[2] = h.beta; out4[3] = h.gamma; }
  return hit ? 1 : 0;
}

// the specular power function of the shading path (rt_math.cuh pow_specular)
double sim_pow_specular(double x, double y) { return rtb::pow_specular(x, y); }

// conservative FP32 sphere pre-filter, operands rounded exactly like the flattener does
int sim_sphere_filter(const double *center, double radius, const double *o, const double *d, double thi) {
  using namespace rtb;
  float cf[3];
  for (int k = 0; k < 3; ++k) cf[k] = (float)center[k];
  const double ca = std::fabs(center[0]) + std::fabs(center[1]) + std::fabs(center[2]);
  float cabs = (float)ca; if ((double)cabs < ca) cabs = nextafterf(cabs, INFINITY);
  RayF r;
  rayf_setup(r, ld3(o), ld3(d), 0.0f);
  return sphere_filter(r, cf, (float)radius, cabs, thi_of(thi)) ? 1 : 0;
}

// conservative FP32 box test against an FP64 box, encoded in centre/half-extent form by the flattener
// (a one-triangle-pair scene would do the same; this calls the flattener's arithmetic through a tiny BVH)
int sim_box_test(const double *bmin, const double *bmax, const double *o, const double *d, double thi,
                 double extent_scale) {
  using namespace rtb;
  float ctr[3], half[3];
  double E = 0.0;
  for (int k = 0; k < 3; ++k) {
    const float cc = (float)(0.5 * (bmin[k] + bmax[k]));
    const double need = std::fmax(bmax[k] - (double)cc, (double)cc - bmin[k]);
    float h = (float)need; if ((double)h < need) h = nextafterf(h, INFINITY);
    h = nextafterf(h, INFINITY);
    ctr[k] = cc; half[k] = h;
    E = std::fmax(E, std::fabs((double)cc) + (double)h);
  }
  E *= extent_scale; // the scene's box_extent is >= this box's
  float Ef = (float)E; if ((double)Ef < E) Ef = nextafterf(Ef, INFINITY);
  RayF r;
  rayf_setup(r, ld3(o), ld3(d), Ef);
  float tn;
  const float t = thi_of(thi);
  return box_test(r, ctr, half, this_of(t), tn) ? 1 : 0;
}

} // extern "C"
