// rt_intersect.cuh — intersection primitives of the hot path.
//
//   box2_test        conservative FP32 slab test of the two child boxes of a NodeRec
//                    (replaces BVH::intersectAABB mybvh.cpp:99-135 / intersectAABB_device
//                    mytracer_gpu.cu:548-596: no divides, near/far planes picked by ray sign,
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

// Per-ray FP32 traversal constants.
struct RayF {
  float ox, oy, oz;    // origin rounded to FP32
  float dx, dy, dz;    // direction rounded to FP32
  float ix, iy, iz;    // 1/d with |d| clamped to >= 1e-20 (no inf/NaN ever enters the slab test)
  float lox, loy, loz; // -o*i - 4u|o*i|   (added to near-plane products)
  float hix, hiy, hiz; // -o*i + 4u|o*i|   (added to far-plane products)
  float o1n;           // |ox|+|oy|+|oz|  (triangle filter error bound)
};

RT_HD float safe_rcp(float d) {
  const float ad = fabsf(d);
  const float c = ad < 1e-20f ? 1e-20f : ad;
  return copysignf(1.0f / c, d);
}

RT_HD void rayf_setup(RayF &r, const D3 &o, const D3 &d) {
  r.ox = (float)o.x; r.oy = (float)o.y; r.oz = (float)o.z;
  r.dx = (float)d.x; r.dy = (float)d.y; r.dz = (float)d.z;
  r.ix = safe_rcp(r.dx); r.iy = safe_rcp(r.dy); r.iz = safe_rcp(r.dz);
  const float px = -r.ox * r.ix, py = -r.oy * r.iy, pz = -r.oz * r.iz;
  const float k = 4.0f * kU32;
  // slack covers the rounding of o (u|o|) and of the product o*i (u|o*i|), see DESIGN.md
  const float sx = k * fabsf(px) , sy = k * fabsf(py), sz = k * fabsf(pz);
  r.lox = px - sx - FLT_MIN; r.loy = py - sy - FLT_MIN; r.loz = pz - sz - FLT_MIN;
  r.hix = px + sx + FLT_MIN; r.hiy = py + sy + FLT_MIN; r.hiz = pz + sz + FLT_MIN;
  r.o1n = fabsf(r.ox) + fabsf(r.oy) + fabsf(r.oz);
}

// relative slack on the far distance: covers the 4u|t| rounding of every slab distance
constexpr float kFarSlack = 1.0f + 64.0f * kU32;

// Entry distance (conservative lower bound) of one child box; hit iff max(tn,0) <= min(tf', thi)
RT_HD bool box_test(const RayF &r, const float *bmin, const float *bmax, float thi, float &tn) {
  const float nx = r.ix >= 0.f ? bmin[0] : bmax[0], fx = r.ix >= 0.f ? bmax[0] : bmin[0];
  const float ny = r.iy >= 0.f ? bmin[1] : bmax[1], fy = r.iy >= 0.f ? bmax[1] : bmin[1];
  const float nz = r.iz >= 0.f ? bmin[2] : bmax[2], fz = r.iz >= 0.f ? bmax[2] : bmin[2];
  const float tnx = fmaf(nx, r.ix, r.lox), tfx = fmaf(fx, r.ix, r.hix);
  const float tny = fmaf(ny, r.iy, r.loy), tfy = fmaf(fy, r.iy, r.hiy);
  const float tnz = fmaf(nz, r.iz, r.loz), tfz = fmaf(fz, r.iz, r.hiz);
  tn = fmax3(tnx, tny, tnz);
  const float tf = fmin3(tfx, tfy, tfz);
  // tf < 0 (box behind the ray): tf*slack stays negative < max(tn,0): rejected
  const float tf2 = fminf(tf * kFarSlack + FLT_MIN, thi);
  return fmaxf(tn, 0.0f) <= tf2;
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
  const double alpha = ddiv(det3(c4, c2, c3), S);
  const double beta = ddiv(det3(c1, c4, c3), S);
  const double gamma = dsub(dsub(1.0, alpha), beta);
  const double t = ddiv(det3(c1, c2, c4), S);
  if (t <= 1e-5) return false;
  const bool inside = (0.0 <= alpha && alpha <= 1.0) && (0.0 <= beta && beta <= 1.0) &&
                      (0.0 <= gamma && gamma <= 1.0);
  if (!inside) return false;
  h.t = t; h.alpha = alpha; h.beta = beta; h.gamma = gamma;
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
  const double t1 = ddiv(dsub(-b, sq), dmul(2.0, a));
  const double t2 = ddiv(dadd(-b, sq), dmul(2.0, a));
  double t = t1;
  if (!(t > 1e-5)) t = t2;
  if (!(t > 1e-5)) return false;
  t_out = t;
  return true;
}

} // namespace rtb
#endif
