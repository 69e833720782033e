// rt_types.h — layout of the flattened scene blob in HBM (shared by the host flattener,
// the CUDA kernel and the host-compiled test build of the kernel core).
//
// One contiguous, 512-byte-sectioned blob (so it can be broadcast with a single NCCL call):
//
//   BlobHeader | NodeRec[] | TriFilter[] | TriExact[] | TriShade[] | MeshRec[] | ObjectRec[] | texels | NodeMaxX[]
//              | tri_id[] | tri_pos[] | ObjFilter[] | obj_order[] | lin_idx[] | Node4Rec[]
//
// Up to three trees share the NodeRec array (rt_sah.h); the third, over the analytic spheres of
// Raytracer::objects_ (the reference loops over them linearly per ray), is walked before the triangles: records [0, q_base) are the product's SAH traversal tree,
// records [q_base, q_base + num_nodes_q) the reference's median-split tree, walked only by rays with
// direction.x == 0 (NodeMaxX below).  With RT_SCENE_REFERENCE_TOPOLOGY there is one tree (q_base == 0) and
// every ray walks the reference topology.  TriFilter / TriExact are stored in the leaf order of the traversal
// tree; tri_id[position] is the reference's primitive index (the BVH-order row of Data, which is what the
// path reports and breaks ties with), tri_pos[] its inverse; TriShade stays in reference order.
//
// * NodeRec   64 B  (4 x float4): one record per INTERIOR node of a binary tree (the reference's:
//             mybvh.h:49-55 / mybvh.cpp:438-472) holding BOTH child boxes in FP32
//             centre / half-extent form (outward-rounded: [c-h, c+h] contains the FP64 box),
//             plus the two child references.  Same topology, half the dependent loads of the
//             reference's 76 B/node five-array layout; the centre/half form turns the slab test
//             into 9 FMAs per box with no min/max selection by ray sign (rt_intersect.cuh).
// * TriFilter 48 B  (3 x float4): FP32 Moeller-Trumbore operands + per-triangle error-bound
//             constants, BVH order.  A conservative filter: it never rejects a triangle the
//             reference's FP64 test (mymesh.cpp:176-236) accepts.
// * TriExact  80 B  (5 x 16 B): the three vertex positions in FP64, read only for filter
//             survivors; the confirmation reproduces the reference arithmetic bit for bit.
// * TriShade 128 B: normals (face or 3 vertex normals), UVs, mesh id — read once per shaded hit.
// * MeshRec  128 B: material + draw mode + texture descriptor (mydata.h:57-71).
// * ObjectRec 176 B: analytic primitives (Raytracer::objects_), FP64, + FP32 operands of the sphere pre-filter.
// * NodeMaxX  16 B: the FP64 bb_max_.x of both children, read ONLY by rays whose direction.x is exactly
//             0: the reference's slab test (mybvh.cpp:99-135) computes (bb_max.x - o.x)/0 = 0/0 = NaN when
//             o.x == bb_max.x, the NaN survives std::min and the box is REJECTED (with its subtree).
//             The kernel reproduces that one quirk so that primitive ids stay bit-exact on such rays
//             (e.g. the centre pixel column of a symmetric view over a grid line).
#ifndef RTB_TYPES_H
#define RTB_TYPES_H
#include <stddef.h>
#include <stdint.h>

namespace rtb {

constexpr uint32_t kBlobMagic = 0x42325452u; // "RT2B"
constexpr uint32_t kBlobVersion = 12;
constexpr int kStackMax = 64;       // traversal stack entries (reference: int stack[64], mytracer_gpu.cu:348)
constexpr int kTileW = 8, kTileH = 4; // one warp-sized pixel tile
constexpr size_t kSectionAlign = 512;

struct alignas(16) NodeRec {
  // c0 = left child (leftChildIdx_), c1 = right child (leftChildIdx_ + 1)
  // ctr = box centre, half = half extent >= 0 (an empty child has half = -1e30: never hit)
  float c0_ctr[3], c0_half[3];
  float c1_ctr[3], c1_half[3];
  // ref >= 0.  cnt == 0: ref is the NodeRec index of an interior child.
  //            cnt  > 0: leaf child with triangles [ref, ref + cnt) in BVH order.
  // Two sibling leaves always own adjacent ranges (ref1 == ref0 + cnt0: the in-place partition of
  // mybvh.cpp:481-513 splits one range in two); the flattener rejects anything else.
  int32_t ref0, ref1, cnt0, cnt1;
};
static_assert(sizeof(NodeRec) == 64, "NodeRec must be 64 bytes");

// 128 B (4 x 32-byte loads): one record per node of the 4-wide traversal tree, collapsed from the binary SAH
// tree (rt_flatten.cpp collapse_wide).  Four child boxes in the same outward-rounded centre / half-extent
// form.  One visit = one dependent load + four independent slab tests instead of two visits of two: half the
// dependent steps per ray and twice the instruction-level parallelism per step (the binary loop is bound by
// its dependent chain).  Children keep the left-to-right order of the binary tree, so the
// triangle ranges of leaf children are ascending (adjacent when the leaves are neighbours).
// Children are stored as two PAIRS (0,1) and (2,3) with the two boxes of a pair interleaved word by word, so that a
// 256-bit load leaves (child a, child b) of every coordinate in an aligned register pair: the operand form of the
// packed FP32 FMA of sm_100 (FFMA2, rt_intersect.cuh box_pair).
struct alignas(32) Node4Rec {
  float box[2][6][2]; // [pair][ctr.x ctr.y ctr.z half.x half.y half.z][child a, child b]; child k = 2*pair + (a: 0, b: 1)
                      // (an empty child has half = -1e30: never hit)
  int32_t ref[4];   // cnt == 0: Node4Rec index of an interior child (-1: empty slot)
  int32_t cnt[4];   // cnt  > 0: leaf child with triangle positions [ref, ref + cnt)
};
static_assert(sizeof(Node4Rec) == 128, "Node4Rec must be 128 bytes");

// 256 B (8 x 32-byte loads): one record per node of the 8-wide twin of the triangle traversal tree, collapsed from the same
// binary SAH tree as Node4Rec (rt_flatten.cpp collapse_wide): eight child boxes as four FFMA2 pairs.  One visit = one
// dependent load + eight independent slab tests: 6 instead of 10 dependent load -> test -> vote chains per ray on the
// Office scene for 1.4x the box tests (render_kernel<K_WIDE8>, rt_core.cuh trav_step8).  Triangle tree only: scenes with a
// tree over analytic spheres keep the 4-wide records.
struct alignas(32) Node8Rec {
  float box[4][6][2]; // [pair][ctr.x ctr.y ctr.z half.x half.y half.z][child a, child b]; child k = 2*pair + (a: 0, b: 1)
  int32_t ref[8];     // cnt == 0: Node8Rec index of an interior child (-1: empty slot)
  int32_t cnt[8];     // cnt  > 0: leaf child with triangle positions [ref, ref + cnt)
};
static_assert(sizeof(Node8Rec) == 256, "Node8Rec must be 256 bytes");

struct alignas(16) TriFilter {
  float p2[3], k_bary; // vertex p2 ; 8u*max(|e1|,|e2|)   (bound on the U,V numerators, per unit rho)
  float e1[3], k_t;    // p0 - p2   ; 8u*|e1||e2|         (bound on the t numerator, per unit rho)
  float e2[3], k_det;  // p1 - p2   ; 64u*|e1||e2|        (bound on the determinant)
};
static_assert(sizeof(TriFilter) == 48, "TriFilter must be 48 bytes");

struct alignas(16) TriExact {
  double p0[3], p1[3], p2[3];
  int32_t mesh, pad;
};
static_assert(sizeof(TriExact) == 80, "TriExact must be 80 bytes");

struct alignas(16) TriShade {
  double n[9];  // FLAT: n[0..2] = face normal (Data::normals_); PHONG: the 3 vertex normals
  double uv[6]; // u0 v0 u1 v1 u2 v2 (zeros when the mesh has no texture)
  int32_t mesh, pad;
};
static_assert(sizeof(TriShade) == 128, "TriShade must be 128 bytes");

struct alignas(16) MeshRec {
  double ambient[3], diffuse[3], specular[3];
  double shininess, mirror;
  int32_t shadowable, draw_mode; // draw_mode 0 FLAT, 1 PHONG (mytracer_gpu.cu:498-501)
  int32_t tex_w, tex_h;          // -1 when untextured (mytracer.cpp:273-275)
  uint64_t tex_offset;           // first texel (float4 index) inside the blob's texel section
  uint64_t pad;
};
static_assert(sizeof(MeshRec) == 128, "MeshRec must be 128 bytes");

struct alignas(16) ObjectRec {
  int32_t kind, shadowable; // kind 0 sphere, 1 plane
  double center[3];
  double normal[3];
  double radius;
  double ambient[3], diffuse[3], specular[3];
  double shininess, mirror;
  // FP32 operands of the conservative sphere pre-filter (rt_intersect.cuh sphere_filter)
  float cf[3], rf;   // centre and radius rounded to FP32
  float cabs;        // |cx| + |cy| + |cz| (error bound)
  float padf;
};
static_assert(sizeof(ObjectRec) == 176, "ObjectRec must be 176 bytes");

// 32 B per analytic primitive: everything the per-ray loop over Raytracer::objects_ touches before a sphere
// survives its FP32 pre-filter.  602 spheres = 19 KB, L1-resident next to the cold state (the 176-byte
// ObjectRec list is 106 KB and falls out of a 92 KB L1: 4.1 instead of 3.1 ms on the molecule scene).
struct alignas(16) ObjFilter {
  float c[3], r;   // centre and radius rounded to FP32
  float cabs;      // |cx| + |cy| + |cz| (error bound)
  int32_t kind;    // 0 sphere, 1 plane (planes have no filter)
  int32_t pad[2];
};
static_assert(sizeof(ObjFilter) == 32, "ObjFilter must be 32 bytes");

struct alignas(16) NodeMaxX {
  double c0_max_x, c1_max_x;
};
static_assert(sizeof(NodeMaxX) == 16, "NodeMaxX must be 16 bytes");

struct BlobHeader {
  uint32_t magic, version;
  uint64_t total_bytes;
  int32_t num_nodes, num_tris, num_meshes, num_objects;
  int32_t root_is_leaf, max_depth, max_leaf, num_nodes_ref;
  uint64_t num_texels;
  uint64_t off_nodes, off_filter, off_exact, off_shade, off_meshes, off_objects, off_texels;
  float scene_extent; // max |coordinate| over all vertices (diagnostics)
  float box_extent;   // >= |ctr| + half of every child box, every axis (slab-test error bound)
  float pad_f[2];
  uint64_t off_maxx;   // NodeMaxX[num_nodes_q], indexed by (record - q_base)
  double root_max_x;   // bb_max_.x of reference node 0 (tested on entry, mybvh.cpp:157)
  uint64_t off_tri_id, off_tri_pos; // int32[num_tris] each
  int32_t q_base;      // first record of the reference-topology tree (0: it IS the traversal tree)
  int32_t num_nodes_q; // its record count; num_nodes = total records of both trees
  int32_t topology;    // 0 SAH traversal tree + reference tree, 1 reference tree only
  int32_t num_nodes4;  // records of the 4-wide traversal tree (0: none; root = record 0)
  uint64_t off_objfilt; // ObjFilter[num_objects]
  // analytic primitives: spheres go into a third tree (records [obj_base, obj_base + num_nodes_obj)) when there
  // are enough of them; everything else (planes; all objects when there is no tree) is tested linearly per ray
  uint64_t off_obj_order; // int32[num_obj_tree]: leaf position -> object index
  uint64_t off_lin_idx;   // int32[num_lin]: objects tested linearly, in list order
  int32_t obj_root;       // record index of the object tree's root, -1: no object tree
  int32_t num_nodes_obj, num_obj_tree, num_lin;
  uint64_t off_nodes4;    // Node4Rec[num_nodes4]: the 4-wide twin of the triangle traversal tree first (root = record 0,
                          // num_nodes4_tri records), then the twin of the object tree (root = obj_root4)
  int32_t num_nodes4_tri; // 0: the triangle tree has no 4-wide twin
  int32_t obj_root4;      // -1: the object tree has none
  int32_t pad_h[2];
  uint64_t off_nodes8;    // Node8Rec[num_nodes8]: the 8-wide twin of the triangle traversal tree (root = record 0)
  int32_t num_nodes8;     // 0: none
  int32_t pad_8;
};
static_assert(sizeof(BlobHeader) % 16 == 0, "BlobHeader must keep 16-byte alignment");

// Pointers into a resident blob (device pointers in the kernel, host pointers in the test build).
struct SceneView {
  const NodeRec *nodes;
  const Node4Rec *nodes4;   // 4-wide twin of the traversal tree [0, q_base), or nullptr
  const Node8Rec *nodes8;   // 8-wide twin of the triangle traversal tree, or nullptr
  int32_t num_nodes8;
  const TriFilter *filt;
  const TriExact *exact;
  const TriShade *shade;
  const MeshRec *meshes;
  const ObjectRec *objects;
  const ObjFilter *objfilt;
  const float *texels; // 4 floats per texel (host/test build); the kernel uses tex_objs instead
  const unsigned long long *tex_objs; // cudaTextureObject_t per mesh (0 when untextured)
  const NodeMaxX *maxx;
  const int32_t *tri_id;  // traversal-order position -> reference primitive index
  const int32_t *tri_pos; // reference primitive index -> traversal-order position
  const int32_t *obj_order; // object-tree leaf position -> object index
  const int32_t *lin_idx;   // objects tested linearly at ray start
  int32_t obj_root, num_lin;
  int32_t tri_nodes;        // records of the triangle trees (0: the scene has no triangles)
  int32_t num_nodes4;       // records of the 4-wide trees (0: none)
  int32_t obj_root4;        // root of the object tree's 4-wide twin in nodes4 (-1: none)
  int32_t num_nodes4_tri;   // records of the triangle tree's twin (root = record 0; 0: none)
  double root_max_x;
  int32_t num_nodes, num_tris, num_meshes, num_objects;
  float box_extent;
  int32_t q_base;
};

// The K_WIDE kernels walk 4-wide records only (plus the reference tree for d.x == 0 rays): usable when every tree
// ordinary rays walk has a twin — the triangle tree (if the scene has one that is not the reference topology itself)
// and the object tree (if the scene has one).
inline bool wide_usable(const SceneView &v) {
  const bool tri_ok = v.tri_nodes == 0 || v.num_nodes4_tri > 0;
  const bool obj_ok = v.obj_root < 0 || v.obj_root4 >= 0;
  return v.num_nodes4 > 0 && tri_ok && obj_ok;
}

// render_kernel<K_WIDE8>: a triangle tree with an 8-wide twin and no object tree
inline bool wide8_usable(const SceneView &v) { return v.num_nodes8 > 0 && v.nodes8 != nullptr && v.obj_root < 0 && wide_usable(v); }

struct CameraRec {
  double eye[3], x_dir[3], y_dir[3], lower_left[3];
  int32_t width, height;
};

struct FrameParams {
  CameraRec cam;
  double background[3], ambience[3];
  int32_t max_depth, spp_grid;
  int32_t num_lights;
  int32_t tile_rank, tile_world, tile_chunk;
  int32_t tiles_x, tiles_y;
  int32_t flags;
  int32_t static_tiles; // tiles [0, static_tiles) are statically sharded, [static_tiles, tiles_x*tiles_y) dynamic
  int32_t dyn_chunk;    // tiles per claim from the dynamic pool
  int32_t pad;
};

struct FrameOut {
  uint8_t *rgba8;
  double *pixels; // 4 doubles per pixel
  int32_t *hit_id;
  double *hit_t;
};

// indices into the counters array (RT_FLAG_COUNTERS build)
enum Counter {
  C_RAY_PRIMARY = 0, C_RAY_SHADOW, C_RAY_REFLECT, C_BOX, C_TRI_FILTER, C_TRI_EXACT, C_OBJECT,
  C_SHADED, C_TEXEL, C_NODES, C_SHADOW_TRIVIAL, C_SHADOW_OCCLUDED, C_NODES4, C_NODES8, C_COUNT
};

} // namespace rtb
#endif
