// rt_sah.h — the product's own traversal tree: a binary surface-area-heuristic BVH over the triangles.
//
// The reference's tree (median of centroids on a cycling axis, mybvh.cpp:438-472) fixes the primitive
// numbering and one arithmetic quirk (rt_types.h NodeMaxX) — both are kept — but it is a poor
// acceleration structure: on the Office scene a random ray is expected to visit 102 interior nodes and test
// 21.5 triangles, against 24 nodes and 6.5 triangles for the tree built here.  Because every hit is decided
// by the bit-exact FP64 test and ties are broken by the reference's primitive index, the result of a
// traversal does not depend on the tree it walks; only rays with direction.x == 0 (the 0/0 quirk) must walk
// the reference's own tree, which therefore stays in the blob next to this one.
#ifndef RTB_SAH_H
#define RTB_SAH_H
#include <cstdint>
#include <string>
#include <vector>

namespace rtb {

struct SahNode {
  float lo[3], hi[3]; // contains the FP64 triangles (primitive boxes are rounded outward)
  int left, right;    // interior: child node indices; leaf: -1
  int first, count;   // leaf: positions [first, first + count) of `order`
};

// lo/hi: n primitive boxes (3 floats each).  On return `order` is a permutation of 0..n-1 whose leaf
// ranges are contiguous, `nodes[0]` is the root, `depth` the number of levels.  Deterministic for a given
// input (the worker threads only split disjoint subtrees).
void build_sah(const float *lo, const float *hi, int n, std::vector<int> &order, std::vector<SahNode> &nodes,
               int &depth);

// The same build on the GPU (rt_sah_gpu.cu; not part of host-only builds): level-synchronous binned SAH above 16
// primitives, one thread per subtree with the exact sweep below.  Same contract; `nodes` may contain unused slots.
// Returns false with `err` set if the device build could not run (the caller falls back to build_sah).
// The device-resident form: boxes in, tree left in HBM (rt_flatten_gpu.cu emits the blob's records from it).
struct SahDeviceTree {
  int *order = nullptr;       // [n] device: primitive of every position
  SahNode *nodes = nullptr;   // device: phase-A nodes [0, nodes_a), phase-B blocks [cap_a, reserve); unwritten slots: count = -1
  int n = 0, nodes_a = 0, depth = 0;
  size_t cap_a = 0, reserve = 0;
};
bool build_sah_on_device(const float *d_lo, const float *d_hi, int n, SahDeviceTree &tree, double *device_ms, std::string &err);
void free_sah_device_tree(SahDeviceTree &tree);
bool build_sah_device(const float *lo, const float *hi, int n, std::vector<int> &order, std::vector<SahNode> &nodes,
                      int &depth, double *device_ms, std::string &err);

} // namespace rtb
#endif
