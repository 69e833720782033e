// rt_flatten.h — host flattener: reference Data SoA + BVHNodes_SoA -> one HBM blob (rt_types.h).
#ifndef RTB_FLATTEN_H
#define RTB_FLATTEN_H
#include "../../include/rt_b200.h"
#include "rt_types.h"
#include <string>
#include <vector>

namespace rtb {

// Returns RT_OK or an rt_status; on failure `err` holds the message.  flags: RT_SCENE_* (rt_b200.h).
int flatten_scene(const rt_data_soa *data, const rt_bvh_nodes_soa *bvh, const rt_object *objects,
                  int num_objects, unsigned flags, std::vector<unsigned char> &blob, std::string &err);

// Device builder of the traversal tree (rt_sah.h build_sah_device), installed by the CUDA library; host-only builds
// (tests/hostsim) leave it null and flatten with the host builder.  RT_SCENE_HOST_SAH keeps the host builder.
using SahDeviceFn = bool (*)(const float *, const float *, int, std::vector<int> &, std::vector<struct SahNode> &, int &,
                             double *, std::string &);
void set_sah_device_builder(SahDeviceFn fn);
struct FlattenStats {
  int sah_builder = 0;      // 0: no SAH tree built, 1: host builder, 2: device builder
  double sah_ms = 0.0;      // wall time of the tree build (device: including its transfers)
  double sah_device_ms = 0.0;
  int flattener = 0;        // 0: host flattener (this file's twin rt_flatten.cpp), 1: device flattener (rt_flatten_gpu.cu)
};
const FlattenStats &last_flatten_stats();   // of the calling thread's last flatten_scene / flatten_scene_device
FlattenStats &mutable_flatten_stats();

// The flattener on the GPU (rt_flatten_gpu.cu; CUDA library only): same inputs, the blob is built in HBM and handed over
// as a device allocation (cudaMalloc; the caller owns it).  Returns RT_OK, or kFlattenUseHost when the scene is one the
// device path leaves to flatten_scene (reference topology, root leaf, any malformed input: the host twin words the error),
// or an rt_status of a CUDA failure (the caller may still fall back to the host twin).
struct DeviceBlob {
  unsigned char *d_blob = nullptr;
  size_t bytes = 0;
  BlobHeader hdr;
};
constexpr int kFlattenUseHost = -1;
int flatten_scene_device(const rt_data_soa *data, const rt_bvh_nodes_soa *bvh, const rt_object *objects, int num_objects,
                         unsigned flags, DeviceBlob &out, std::string &err);

// Validates a blob received from another rank (magic, version, section bounds).
int check_blob(const unsigned char *blob, size_t bytes, std::string &err);
// The same on a header already copied out of the blob (host or device resident).
int check_blob_header(const BlobHeader &h, size_t bytes, std::string &err);

// Host-side view of a blob (pointers into `blob`); tex_objs is left null.
SceneView view_of_blob(const unsigned char *blob);

} // namespace rtb
#endif
