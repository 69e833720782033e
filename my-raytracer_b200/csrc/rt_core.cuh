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
//     kernel that always runs max_depth bounces.
// The same code is compiled for the device and — for tests only — for the host.
#ifndef RTB_CORE_CUH
#define RTB_CORE_CUH
#include "rt_intersect.cuh"

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

struct Lane {
  // --- ray in flight -------------------------------------------------------------------
  D3 o, d;
  RayF rf;
  double best_t; // nearest: best distance so far; any-hit: light distance (exclusive bound)
  float thi;     // FP32 upper bound of best_t used for culling
  int best_id;
  int node;      // >= 0: interior NodeRec to visit next; < 0: traversal finished
  int sp;
  int mode;
  int found;     // any-hit result
  // --- path ------------------------------------------------------------------------------
  int px, py;
  int sample, depth;
  double weight;
  D3 col, acc;
  // --- hit being shaded ---------------------------------------------------------------
  D3 P, N, V;
  D3 kd, L;
  int mat;       // >= 0: mesh index; < 0: ~object index
  int light_k;
};

struct Mat {
  D3 ka, ks;
  double shininess, mirror;
  int shadowable;
};

template <bool COUNT> RT_HD void bump(const Ctx &c, int which, unsigned n = 1) {
  if (COUNT) c.cnt[which] += n;
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
template <bool COUNT>
RT_HD void start_ray(const Ctx &c, Lane &L, const D3 &origin, const D3 &dir, int mode, double tmax) {
  L.o = origin;
  L.d = normalize(dir);
  rayf_setup(L.rf, L.o, L.d);
  L.mode = mode;
  L.best_t = tmax;
  L.best_id = -1;
  L.found = 0;
  L.sp = 0;
  const SceneView &sv = c.sv;
  for (int k = 0; k < sv.num_objects; ++k) {
    const ObjectRec &ob = sv.objects[k];
    double t;
    bump<COUNT>(c, C_OBJECT);
    const bool hit = (ob.kind == 0) ? sphere_exact(L.o, L.d, ld3(ob.center), ob.radius, t)
                                    : plane_exact(L.o, L.d, ld3(ob.center), ld3(ob.normal), t);
    if (hit && t < L.best_t) {
      if (mode == M_ANYHIT) { L.found = 1; L.node = -1; return; }
      L.best_t = t;
      L.best_id = sv.num_tris + k;
    }
  }
  const float up = f_up(L.best_t);
  L.thi = fminf(up * (1.0f + 8.0f * kU32), FLT_MAX);
  L.node = (sv.num_nodes > 0) ? 0 : -1;
}

template <bool COUNT> RT_HD void leaf_test(const Ctx &c, Lane &L, int first, int count) {
  const SceneView &sv = c.sv;
  const bool exact_only = (c.fp.flags & 1) != 0;
  for (int i = first; i < first + count; ++i) {
    bump<COUNT>(c, C_TRI_FILTER);
    if (!exact_only) {
      const TriFilter f = sv.filt[i];
      if (!tri_filter(L.rf, f, L.thi)) continue;
    }
    bump<COUNT>(c, C_TRI_EXACT);
    const TriExact &e = sv.exact[i];
    TriHit h;
    if (!tri_exact(L.o, L.d, ld3(e.p0), ld3(e.p1), ld3(e.p2), h)) continue;
    if (L.mode == M_ANYHIT) {
      // isShadow = hit && t < light_distance && 0 < t   (mytracer.cpp:599)
      if (h.t < L.best_t) { L.found = 1; L.node = -1; return; }
    } else {
      // strict '<' keeps the first hit in DFS order (mybvh.cpp:169); leaves are visited in
      // ascending triangle index there, so "lower index wins a tie" is visit-order independent.
      // Analytic primitives are tested before the BVH, so they win ties against triangles.
      const bool better = (h.t < L.best_t) || (h.t == L.best_t && L.best_id >= 0 &&
                                               L.best_id < sv.num_tris && i < L.best_id);
      if (better) {
        L.best_t = h.t;
        L.best_id = i;
        L.thi = fminf(f_up(h.t) * (1.0f + 8.0f * kU32), FLT_MAX);
      }
    }
  }
}

// One traversal step: load one 64-byte node, test both child boxes, intersect leaf children,
// descend into the nearer interior child (push the other) or pop.
template <bool COUNT> RT_HD void trav_step(const Ctx &c, Lane &L, int *stk) {
  const NodeRec n = c.sv.nodes[L.node];
  bump<COUNT>(c, C_NODES);
  bump<COUNT>(c, C_BOX, 2);
  float tn0, tn1;
  bool h0 = box_test(L.rf, n.c0_min, n.c0_max, L.thi, tn0);
  bool h1 = box_test(L.rf, n.c1_min, n.c1_max, L.thi, tn1);
  if (h0 && n.cnt0 > 0) {
    leaf_test<COUNT>(c, L, n.ref0, n.cnt0);
    h0 = false;
    if (L.found) return;
  }
  if (h1 && n.cnt1 > 0) {
    leaf_test<COUNT>(c, L, n.ref1, n.cnt1);
    h1 = false;
    if (L.found) return;
  }
  int next;
  if (h0 && h1) {
    const bool far0 = tn1 < tn0;
    next = far0 ? n.ref1 : n.ref0;
    stk[L.sp++] = far0 ? n.ref0 : n.ref1;
  } else if (h0) {
    next = n.ref0;
  } else if (h1) {
    next = n.ref1;
  } else {
    next = (L.sp > 0) ? stk[--L.sp] : -1;
  }
  L.node = next;
}

// ---------------------------------------------------------------------------------------
// path logic
// ---------------------------------------------------------------------------------------
template <bool COUNT> RT_HD void start_primary(const Ctx &c, Lane &L) {
  const CameraRec &cam = c.fp.cam;
  double x = (double)L.px, y = (double)L.py;
  const int s = c.fp.spp_grid;
  if (s > 1) {
    // sub-pixel grid, offsets as mytracer_gpu.cu:203-207: si/s - 0.5 + 1/(2s)
    const int si = L.sample / s, sj = L.sample % s;
    const double xo = dadd(dsub(ddiv((double)si, (double)s), 0.5), ddiv(1.0, dmul(2.0, (double)s)));
    const double yo = dadd(dsub(ddiv((double)sj, (double)s), 0.5), ddiv(1.0, dmul(2.0, (double)s)));
    x = dadd(x, xo);
    y = dadd(y, yo);
  }
  // Camera::primary_ray: Ray(eye, lower_left + x*x_dir + y*y_dir - eye)
  const D3 eye = ld3(cam.eye);
  const D3 dir = sub(add(add(ld3(cam.lower_left), scale(x, ld3(cam.x_dir))), scale(y, ld3(cam.y_dir))), eye);
  L.depth = 0;
  L.weight = 1.0;
  L.col = mk(0.0, 0.0, 0.0);
  bump<COUNT>(c, C_RAY_PRIMARY);
  start_ray<COUNT>(c, L, eye, dir, M_NEAREST, DBL_MAX);
}

RT_HD void write_pixel(const Ctx &c, Lane &L) {
  D3 col = L.acc;
  const int s2 = c.fp.spp_grid * c.fp.spp_grid;
  if (s2 > 1) col = divs(col, (double)s2);
  col = mk(dmin1(col.x), dmin1(col.y), dmin1(col.z)); // min(color, vec4(1,1,1)) mytracer_gpu.cu:155
  const size_t p = (size_t)L.py * (size_t)c.fp.cam.width + (size_t)L.px;
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

template <bool COUNT> RT_HD void sample_done(const Ctx &c, Lane &L) {
  L.acc = add(L.acc, L.col);
  L.sample++;
  if (L.sample < c.fp.spp_grid * c.fp.spp_grid) {
    start_primary<COUNT>(c, L);
  } else {
    write_pixel(c, L);
    L.mode = M_IDLE;
    L.node = -1;
  }
}

// Fill P, N, kd, mat for the nearest hit and start the lighting loop (L = ambient term).
template <bool COUNT> RT_HD void hit_setup(const Ctx &c, Lane &L) {
  const SceneView &sv = c.sv;
  L.V = neg(L.d);
  if (L.best_id < sv.num_tris) {
    const int i = L.best_id;
    const TriExact &e = sv.exact[i];
    TriHit h;
    h.t = L.best_t; h.alpha = 0.0; h.beta = 0.0; h.gamma = 0.0;
    tri_exact(L.o, L.d, ld3(e.p0), ld3(e.p1), ld3(e.p2), h); // same arithmetic as the accepted test
    const TriShade &s = sv.shade[i];
    const MeshRec &mr = sv.meshes[s.mesh];
    L.mat = s.mesh;
    L.P = along(L.o, h.t, L.d);                               // ray(t), mymesh.cpp:219
    L.kd = ld3(mr.diffuse);
    if (mr.tex_w != -1) {
      // Mesh::compute_texture, mymesh.cpp:70-95
      double u = dadd(dadd(dmul(h.alpha, s.uv[0]), dmul(h.beta, s.uv[2])), dmul(h.gamma, s.uv[4]));
      double v = dadd(dadd(dmul(h.alpha, s.uv[1]), dmul(h.beta, s.uv[3])), dmul(h.gamma, s.uv[5]));
      u = (u < 0.0) ? 0.0 : ((1.0 < u) ? 1.0 : u);
      v = (v < 0.0) ? 0.0 : ((1.0 < v) ? 1.0 : v);
      const int tx = (int)round(dmul(u, (double)(unsigned)(mr.tex_w - 1)));
      const int ty = (int)round(dmul(dsub(1.0, v), (double)(unsigned)(mr.tex_h - 1)));
      L.kd = fetch_texel(sv, s.mesh, mr, tx, ty);
      bump<COUNT>(c, C_TEXEL);
    }
    if (mr.draw_mode == 0) {
      L.N = ld3(s.n);                                          // stored face normal, mymesh.cpp:228-229
    } else {
      // alpha*n0 + beta*n1 + gamma*n2, NOT re-normalised (mymesh.cpp:231-233)
      L.N = add(add(scale(h.alpha, ld3(s.n)), scale(h.beta, ld3(s.n + 3))), scale(h.gamma, ld3(s.n + 6)));
    }
  } else {
    const int k = L.best_id - sv.num_tris;
    const ObjectRec &ob = sv.objects[k];
    L.mat = ~k;
    L.P = along(L.o, L.best_t, L.d);
    L.kd = ld3(ob.diffuse);
    L.N = (ob.kind == 0) ? divs(sub(L.P, ld3(ob.center)), ob.radius) : ld3(ob.normal);
  }
  Mat m;
  load_mat(sv, L.mat, m);
  // ambient term, mytracer.cpp:574-576
  L.L = mk(dadd(0.0, dmul(c.fp.ambience[0], m.ka.x)), dadd(0.0, dmul(c.fp.ambience[1], m.ka.y)),
           dadd(0.0, dmul(c.fp.ambience[2], m.ka.z)));
  L.light_k = 0;
  bump<COUNT>(c, C_SHADED);
}

// diffuse + specular of one unshadowed light, mytracer.cpp:510-534,581-585,603-605
RT_HD void add_light(const Ctx &c, Lane &L, const Mat &m, int k) {
  const double *lcol = c.light_col + 4 * k;
  const D3 lv = sub(ld3(c.light_pos + 4 * k), L.P);
  const D3 l = normalize(lv);
  const double diff = dmax0(dot(L.N, l));
  double refl = 0.0;
  if (diff > 0.0) {
    const D3 r = normalize(mirror(l, L.N));
    refl = dmax0(dot(r, L.V));
  }
  const double spec = pow(refl, m.shininess);
  L.L.x = dadd(L.L.x, dmul(lcol[0], dadd(dmul(L.kd.x, diff), dmul(m.ks.x, spec))));
  L.L.y = dadd(L.L.y, dmul(lcol[1], dadd(dmul(L.kd.y, diff), dmul(m.ks.y, spec))));
  L.L.z = dadd(L.L.z, dmul(lcol[2], dadd(dmul(L.kd.z, diff), dmul(m.ks.z, spec))));
}

// Advance a lane whose traversal has finished until it has a new ray in flight or is idle.
template <bool COUNT> RT_HD void advance(const Ctx &c, Lane &L) {
  const SceneView &sv = c.sv;
  while (L.mode != M_IDLE && L.node < 0) {
    Mat m;
    if (L.mode == M_NEAREST) {
      if (L.depth == 0 && c.fp.spp_grid == 1) {
        const size_t p = (size_t)L.py * (size_t)c.fp.cam.width + (size_t)L.px;
        if (c.out.hit_id) c.out.hit_id[p] = L.best_id;
        if (c.out.hit_t) c.out.hit_t[p] = (L.best_id >= 0) ? L.best_t : -1.0;
      }
      if (L.best_id < 0) {
        // miss: primary -> background; bounce -> weight * background (mytracer_gpu.cu:262,292)
        L.col = add(L.col, scale(L.weight, ld3(c.fp.background)));
        sample_done<COUNT>(c, L);
        continue;
      }
      hit_setup<COUNT>(c, L);
      load_mat(sv, L.mat, m);
    } else { // M_ANYHIT finished: light_k was the light under test
      load_mat(sv, L.mat, m);
      if (!L.found) add_light(c, L, m, L.light_k);
      L.light_k++;
    }
    L.mode = M_SHADE;
    // lighting loop, mytracer.cpp:579-606
    while (L.light_k < c.fp.num_lights) {
      if (m.shadowable) {
        const D3 lv = sub(ld3(c.light_pos + 4 * L.light_k), L.P);
        const D3 ldir = normalize(lv);
        const double ldist = norm(lv);
        bump<COUNT>(c, C_RAY_SHADOW);
        start_ray<COUNT>(c, L, add(L.P, scale(1e-4, ldir)), ldir, M_ANYHIT, ldist);
        break;
      }
      add_light(c, L, m, L.light_k);
      L.light_k++;
    }
    if (L.mode == M_ANYHIT) continue; // shadow ray in flight (or already resolved: node < 0)
    // all lights done: colour of this bounce, then the mirror bounce (mytracer.cpp:546-555)
    L.col = add(L.col, scale(L.weight, scale(dsub(1.0, m.mirror), L.L)));
    if (m.mirror > 0.0 && L.depth < c.fp.max_depth) {
      const D3 v = reflect(neg(L.V), L.N);
      L.weight = dmul(L.weight, m.mirror);
      L.depth++;
      bump<COUNT>(c, C_RAY_REFLECT);
      start_ray<COUNT>(c, L, add(L.P, scale(1e-4, v)), v, M_NEAREST, DBL_MAX);
    } else {
      sample_done<COUNT>(c, L);
    }
  }
}

// Begin a pixel (lane must be idle).
template <bool COUNT> RT_HD void begin_pixel(const Ctx &c, Lane &L, int px, int py) {
  L.px = px; L.py = py;
  L.sample = 0;
  L.acc = mk(0.0, 0.0, 0.0);
  start_primary<COUNT>(c, L);
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
  const long long ntiles = (long long)fp.tiles_x * fp.tiles_y;
  if (t >= ntiles) return false;
  const int tx = (int)(t % fp.tiles_x), ty = (int)(t / fp.tiles_x);
  px = tx * kTileW + (l % kTileW);
  py = ty * kTileH + (l / kTileW);
  return px < fp.cam.width && py < fp.cam.height;
}

// number of work slots this rank must scan
RT_HD long long work_slots(const Ctx &c) {
  const FrameParams &fp = c.fp;
  if (c.pixel_list) return c.list_count;
  const long long ntiles = (long long)fp.tiles_x * fp.tiles_y;
  const long long group = (long long)fp.tile_chunk * fp.tile_world;
  const long long ngroups = (ntiles + group - 1) / group;
  return ngroups * fp.tile_chunk * (kTileW * kTileH);
}

} // namespace rtb
#endif
