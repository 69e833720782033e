// rt_intersect.cuh — intersection primitives of the hot path.
//
//   box_test         conservative FP32 slab test of one child box of a NodeRec, centre/half-extent
//                    form (replaces BVH::intersectAABB mybvh.cpp:99-135 / intersectAABB_device
//                    mytracer_gpu.cu:548-596: 9 FMAs, no divides, no selection by ray sign,
//                    culls against the current best distance)
//   tri_filter       conservative FP32 Moeller-Trumbore filter with a running error bound
//   tri_exact        FP64 confirmation == Mesh::intersect_triangle (mymesh.cpp:176-236)
//   plane_exact      == Plane::intersect (myplane.cpp:22-49)
//   sphere_exact     == Sphere::intersect as stated in DESIGN.md "Conventions"
#ifndef RTB_INTERSECT_CUH
#define RTB_INTERSECT_CUH
#include "rt_math.cuh"
#include "rt_types.h"

namespace rtb {

#ifndef RT_BOX_FOLD
#define RT_BOX_FOLD 1
#endif
// Per-ray FP32 traversal constants.
struct RayF {
  float ox, oy, oz;    // origin rounded to FP32            (triangle filter)
  float dx, dy, dz;    // direction rounded to FP32         (triangle filter)
  float ix, iy, iz;    // 1/d with |d| clamped to >= 1e-20 (no inf/NaN ever enters the slab test)
  float ax, ay, az;    // |1/d|
  float nx, ny, nz;    // -o*i - s   (near planes),  s = per-axis absolute slack, see rayf_setup
  float fx, fy, fz;    // -o*i + s   (far planes)
  float o1n;           // |ox|+|oy|+|oz|  (triangle filter error bound)
};

RT_HD float safe_rcp(float d) {
  const float ad = fabsf(d);
  const float c = ad < 1e-20f ? 1e-20f : ad;
  return copysignf(1.0f / c, d);
}

// E = SceneView::box_extent: upper bound on |ctr| + half of every child box (every axis).
RT_HD void rayf_setup(RayF &r, const D3 &o, const D3 &d, float E) {
  r.ox = (float)o.x; r.oy = (float)o.y; r.oz = (float)o.z;
  r.dx = (float)d.x; r.dy = (float)d.y; r.dz = (float)d.z;
  r.ix = safe_rcp(r.dx); r.iy = safe_rcp(r.dy); r.iz = safe_rcp(r.dz);
  r.ax = fabsf(r.ix); r.ay = fabsf(r.iy); r.az = fabsf(r.iz);
  const float px = -r.ox * r.ix, py = -r.oy * r.iy, pz = -r.oz * r.iz;
  // Absolute error of a computed slab distance  fl(c*i + fl(p -+ s -+ h|i|))  against the exact
  // ((c -+ h) - o)/d, beyond a few-ulp RELATIVE error (covered by kFarSlack and thi's margin):
  //   rounding of o to FP32 and of p = -o*i:  2u|o||i| ;  rounding of the inner FMA:
  //   u|p -+ h|i|| <= u(|o| + E)|i| ;  the outer FMA rounds relative to its result.
  // Per-axis bound s = 6u(|o| + E)|i| (+ FLT_MIN), a spatial slack of 6u(|o| + E): it stays on its own
  // axis, so a ray parallel to one slab (|i| = 1e20) keeps full culling on the other two.
  // The relative errors — i against 1/d (1.5u) and the outer FMA (u) — are relative to a slab distance of at most
  // (|o| + E)|i|, so they are absolute errors of at most 2.5u(|o| + E)|i| as well: with RT_BOX_FOLD they are folded
  // into s (7u needed in total, 8u used: a larger slack makes shadow rays re-test the surface they start on) and the box test needs no relative slack on the far distance (one FMA per box).
#if RT_BOX_FOLD
  const float k = 8.0f * kU32;
#else
  const float k = 6.0f * kU32; // 4u from the analysis above + margin
#endif
  const float sx = k * (fabsf(r.ox) + E) * r.ax + FLT_MIN;
  const float sy = k * (fabsf(r.oy) + E) * r.ay + FLT_MIN;
  const float sz = k * (fabsf(r.oz) + E) * r.az + FLT_MIN;
  r.nx = px - sx; r.ny = py - sy; r.nz = pz - sz;
  r.fx = px + sx; r.fy = py + sy; r.fz = pz + sz;
  r.o1n = fabsf(r.ox) + fabsf(r.oy) + fabsf(r.oz);
}

// relative slack on the far distance: covers the few-ulp relative rounding of every slab distance
constexpr float kFarSlack = 1.0f + 64.0f * kU32;

// Upper bounds used for culling: thi >= best t (triangle filter), thiS = thi*(1+16u) (box test).
RT_HD float thi_of(double t) { return fminf(f_up(t) * (1.0f + 8.0f * kU32), FLT_MAX); }
RT_HD float this_of(float thi) { return fminf(thi * (1.0f + 16.0f * kU32), FLT_MAX); }

// Conservative slab test of one child box in centre/half-extent form (replaces BVH::intersectAABB
// mybvh.cpp:99-135 / intersectAABB_device mytracer_gpu.cu:548-596).  Twelve FMAs, no divides and
// no selection by ray sign:
//     t_near = c*i + (n - h|i|),   t_far = c*i + (f + h|i|)        per axis
// hit iff max(tn, 0) <= min(tf, thiS) with tn = max3(t_near), tf = min3(t_far) (every rounding error, absolute
// and relative, is inside the per-axis slack s of rayf_setup; without RT_BOX_FOLD: tf*kFarSlack + FLT_MIN).
// tf < 0 (box behind the ray): tf*slack stays negative < max(tn, 0): rejected.
// tn (unclamped) orders the two children and is what the stack keeps.
RT_HD bool box_test(const RayF &r, const float *c, const float *h, float thiS, float &tn) {
  const float tnx = fmaf(c[0], r.ix, fmaf(h[0], -r.ax, r.nx));
  const float tny = fmaf(c[1], r.iy, fmaf(h[1], -r.ay, r.ny));
  const float tnz = fmaf(c[2], r.iz, fmaf(h[2], -r.az, r.nz));
  const float tfx = fmaf(c[0], r.ix, fmaf(h[0], r.ax, r.fx));
  const float tfy = fmaf(c[1], r.iy, fmaf(h[1], r.ay, r.fy));
  const float tfz = fmaf(c[2], r.iz, fmaf(h[2], r.az, r.fz));
  tn = fmax3(tnx, tny, tnz);
  const float tf = fmin3(tfx, tfy, tfz);
#if RT_BOX_FOLD
  return fmaxf(tn, 0.0f) <= fminf(tf, thiS);
#else
  return fmaxf(tn, 0.0f) <= fminf(fmaf(tf, kFarSlack, FLT_MIN), thiS);
#endif
}

// Two child boxes of a Node4Rec pair at once.  w = the pair's 12 words: cx(a) cx(b) cy(a) cy(b) cz(a) cz(b)
// hx(a) hx(b) hy(a) hy(b) hz(a) hz(b) (rt_types.h).  On sm_100 the twelve FMAs of a box pair with its sibling's: FFMA2
// (fma.rn.f32x2) multiplies a register PAIR by a broadcast scalar and adds a broadcast scalar or a pair, so the
// pair layout of the record (a 256-bit load leaves (a, b) in an aligned register pair) turns 24 FFMA into 12
// FFMA2 with no extra registers for the ray constants.  Each half is an IEEE fma: bit-identical to box_test().
#ifndef RT_FFMA2
#define RT_FFMA2 1
#endif
#if defined(__CUDA_ARCH__) && RT_FFMA2 && RT_BOX_FOLD
// (r0, r1) = (a0, a1) * s + (c0, c1)
__device__ __forceinline__ void fma2_pair(float &r0, float &r1, float a0, float a1, float s, float c0, float c1) {
  asm("{ .reg .b64 ra, rs, rc, rd; mov.b64 ra, {%2,%3}; mov.b64 rs, {%4,%4}; mov.b64 rc, {%5,%6};"
      " fma.rn.f32x2 rd, ra, rs, rc; mov.b64 {%0,%1}, rd; }"
      : "=f"(r0), "=f"(r1) : "f"(a0), "f"(a1), "f"(s), "f"(c0), "f"(c1));
}
#endif
RT_HD void box_pair(const RayF &r, const float *w, float thiS, float &tnA, float &tnB, bool &hA, bool &hB) {
#if defined(__CUDA_ARCH__) && RT_FFMA2 && RT_BOX_FOLD
  float nA, nB, fA, fB;
  float tnxA, tnxB, tnyA, tnyB, tnzA, tnzB, tfxA, tfxB, tfyA, tfyB, tfzA, tfzB;
  fma2_pair(nA, nB, w[6], w[7], -r.ax, r.nx, r.nx);  fma2_pair(tnxA, tnxB, w[0], w[1], r.ix, nA, nB);
  fma2_pair(fA, fB, w[6], w[7], r.ax, r.fx, r.fx);   fma2_pair(tfxA, tfxB, w[0], w[1], r.ix, fA, fB);
  fma2_pair(nA, nB, w[8], w[9], -r.ay, r.ny, r.ny);  fma2_pair(tnyA, tnyB, w[2], w[3], r.iy, nA, nB);
  fma2_pair(fA, fB, w[8], w[9], r.ay, r.fy, r.fy);   fma2_pair(tfyA, tfyB, w[2], w[3], r.iy, fA, fB);
  fma2_pair(nA, nB, w[10], w[11], -r.az, r.nz, r.nz); fma2_pair(tnzA, tnzB, w[4], w[5], r.iz, nA, nB);
  fma2_pair(fA, fB, w[10], w[11], r.az, r.fz, r.fz);  fma2_pair(tfzA, tfzB, w[4], w[5], r.iz, fA, fB);
  tnA = fmax3(tnxA, tnyA, tnzA);
  tnB = fmax3(tnxB, tnyB, tnzB);
  hA = fmaxf(tnA, 0.0f) <= fminf(fmin3(tfxA, tfyA, tfzA), thiS);
  hB = fmaxf(tnB, 0.0f) <= fminf(fmin3(tfxB, tfyB, tfzB), thiS);
#else
  const float cA[3] = {w[0], w[2], w[4]}, hfA[3] = {w[6], w[8], w[10]};
  const float cB[3] = {w[1], w[3], w[5]}, hfB[3] = {w[7], w[9], w[11]};
  hA = box_test(r, cA, hfA, thiS, tnA);
  hB = box_test(r, cB, hfB, thiS, tnB);
#endif
}

// FP32 conservative triangle filter.  thi = upper bound (FP32, rounded up) on the ray
// parameters still of interest.  Returns false only if the exact test certainly rejects
// the triangle or certainly yields t > thi.
RT_HD bool tri_filter(const RayF &r, const TriFilter &f, float thi) {
  const float Tx = r.ox - f.p2[0], Ty = r.oy - f.p2[1], Tz = r.oz - f.p2[2];
  // P = d x e2
  const float Px = fmaf(r.dy, f.e2[2], -r.dz * f.e2[1]);
  const float Py = fmaf(r.dz, f.e2[0], -r.dx * f.e2[2]);
  const float Pz = fmaf(r.dx, f.e2[1], -r.dy * f.e2[0]);
  const float det = fmaf(f.e1[0], Px, fmaf(f.e1[1], Py, f.e1[2] * Pz));
  const float U = fmaf(Tx, Px, fmaf(Ty, Py, Tz * Pz));
  // Q = T x e1
  const float Qx = fmaf(Ty, f.e1[2], -Tz * f.e1[1]);
  const float Qy = fmaf(Tz, f.e1[0], -Tx * f.e1[2]);
  const float Qz = fmaf(Tx, f.e1[1], -Ty * f.e1[0]);
  const float V = fmaf(r.dx, Qx, fmaf(r.dy, Qy, r.dz * Qz));
  const float W = fmaf(f.e2[0], Qx, fmaf(f.e2[1], Qy, f.e2[2] * Qz));
  // error bounds: rho = 2|o|_1 + 16|T|_1
  const float rho = fmaf(16.0f, fabsf(Tx) + fabsf(Ty) + fabsf(Tz), 2.0f * r.o1n);
  const float B = f.k_bary * rho;
  const float Bw = f.k_t * rho;
  const float Bd = f.k_det;
  const float D = fabsf(det);
  const float sg = det < 0.f ? -1.0f : 1.0f;
  const float Us = U * sg, Vs = V * sg, Ws = W * sg;
  const bool sign_unknown = D <= Bd;
  const bool inside = (Us >= -B) && (Vs >= -B) && (Us + Vs <= D + 2.0f * B + Bd);
  const bool t_lo = Ws + Bw >= 0.9e-5f * D - 1e-5f * Bd;        // exact test needs t > 1e-5
  const bool t_hi = Ws <= thi * (D + Bd) + Bw;                  // and t <= thi
  return sign_unknown || (inside && t_lo && t_hi);
}

struct TriHit {
  double t, alpha, beta, gamma;
};

// == Mesh::intersect_triangle (mymesh.cpp:176-236), geometric part.
RT_HD bool tri_exact(const D3 &o, const D3 &d, const D3 &p0, const D3 &p1, const D3 &p2, TriHit &h) {
  const D3 c1 = sub(p0, p2);
  const D3 c2 = sub(p1, p2);
  const D3 c3 = neg(d);
  const D3 c4 = sub(o, p2);
  const double S = det3(c1, c2, c3);
  if (fabs(S) < 1e-10) return false;
  // three quotients by the same determinant: one shared reciprocal (rt_math.cuh div3), bit-identical to three '/'
  const D3 q = div3(mk(det3(c4, c2, c3), det3(c1, c4, c3), det3(c1, c2, c4)), S);
  const double alpha = q.x;
  const double beta = q.y;
  const double gamma = dsub(dsub(1.0, alpha), beta);
  const double t = q.z;
  if (t <= 1e-5) return false;
  const bool inside = (0.0 <= alpha && alpha <= 1.0) && (0.0 <= beta && beta <= 1.0) &&
                      (0.0 <= gamma && gamma <= 1.0);
  if (!inside) return false;
  h.t = t; h.alpha = alpha; h.beta = beta; h.gamma = gamma;
  return true;
}

// Conservative FP32 pre-filter of Sphere::intersect: false only if the exact test certainly misses, certainly has
// both roots behind the origin, or certainly yields t > thi.  The reference tests every analytic primitive for
// every ray (linear list, no BVH over Raytracer::objects_): for the molecule scene that is 600 FP64 quadratic
// solves (sqrt + 2 divisions) per ray; the filter replaces most of them by 12 FP32 operations.
// Error bounds: delta bounds the absolute error of each component of oc = o - c computed from FP32-rounded
// operands; E bounds the error of disc = (d.oc)^2 - |oc|^2 + r^2; St the error of a root.
RT_HD bool sphere_filter(const RayF &r, const float *cf, float rf, float cabs, float thi) {
  const float ocx = r.ox - cf[0], ocy = r.oy - cf[1], ocz = r.oz - cf[2];
  const float oc2 = fmaf(ocx, ocx, fmaf(ocy, ocy, ocz * ocz));
  const float bh = fmaf(r.dx, ocx, fmaf(r.dy, ocy, r.dz * ocz));
  const float r2 = rf * rf;
  const float disc = fmaf(bh, bh, r2 - oc2);
  const float delta = 8.0f * kU32 * (r.o1n + cabs) + FLT_MIN;
  const float ocn = sqrtf(oc2) + 2.0f * delta;
  const float E = 64.0f * kU32 * (oc2 + r2) + 8.0f * ocn * delta + FLT_MIN;
  if (disc < -E) return false;
  const float sq = sqrtf(fmaxf(disc + E, 0.0f));
  const float St = 4.0f * delta + 8.0f * kU32 * ocn + 32.0f * kU32 * (fabsf(bh) + sq) + FLT_MIN;
  if (-bh + sq + St < 0.0f) return false; // both roots behind the origin (the exact test needs t > 1e-5)
  if (-bh - sq - St > thi) return false;  // even the nearer root lies beyond the best distance
  return true;
}

// == Plane::intersect (myplane.cpp:22-49)
RT_HD bool plane_exact(const D3 &o, const D3 &d, const D3 &center, const D3 &n, double &t_out) {
  const double cosTheta = dot(n, d);
  if (fabs(cosTheta) < 1e-9) return false;
  const double distance = dot(n, center);
  const double t = ddiv(dsub(distance, dot(n, o)), dot(n, d));
  if (t > 1e-5) { t_out = t; return true; }
  return false;
}

// == Sphere::intersect (absent from the reference; convention in DESIGN.md, restated
// identically in oracle/ref_glue.cpp)
RT_HD bool sphere_exact(const D3 &o, const D3 &d, const D3 &center, double radius, double &t_out) {
  const D3 oc = sub(o, center);
  const double a = dot(d, d);
  const double b = dmul(2.0, dot(d, oc));
  const double c = dsub(dot(oc, oc), dmul(radius, radius));
  const double disc = dsub(dmul(b, b), dmul(dmul(4.0, a), c));
  if (disc < 0.0) return false;
  const double sq = dsqrt(disc);
  const D3 q = div3(mk(dsub(-b, sq), dadd(-b, sq), 1.0), dmul(2.0, a)); // both roots: one shared reciprocal
  const double t1 = q.x;
  const double t2 = q.y;
  double t = t1;
  if (!(t > 1e-5)) t = t2;
  if (!(t > 1e-5)) return false;
  t_out = t;
  return true;
}

} // namespace rtb
#endif
