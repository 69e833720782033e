/* rt_b200.h — C ABI of the B200-native ray-tracing hot path.
 *
 * Drop-in boundary for the reference's CUDA entry point
 *     void launch_compute_image_device(vec4* d_pixels, vec4* d_tmpPixels, vec4* d_image,
 *          const int& width, const int& height, const Camera& camera, const vec4* d_lightsPos,
 *          const vec4* d_lightsColor, const int& nLights, const vec4& background,
 *          const vec4& ambience, const int& max_depth, const Data* data,
 *          const BVH::BVHNodes_SoA* bvhNodes);            (reference mytracer_gpu.h:43-56)
 * and its companion init_device() (mytracer_gpu.h:25), called only from
 * Raytracer::compute_image_cuda() (mytracer.cpp:123-159).
 *
 * Plain C: pointers, sizes and PODs only; no C++/torch types.  Every function returns an
 * rt_status (0 = OK) instead of the reference's print-and-continue CHECK macro
 * (common/common.h:6-15); rt_last_error() gives the message of the calling thread's last
 * failure.  One rt_scene may be used by one host thread at a time.
 *
 * Layout contracts
 *   rt_vec4           == reference vec4: 4 doubles, 3 used (mydata.h uses vec4* arrays)
 *   rt_data_soa       == struct Data            (mydata.h:28-72), field for field
 *   rt_bvh_nodes_soa  == BVH::BVHNodes_SoA      (mybvh.h:49-55), field for field
 * so a maintainer can pass reinterpret_cast'ed pointers to the reference's own structs
 * (INTEGRATION.md).  Host, pinned or managed pointers are accepted for scene input.
 */
#ifndef RT_B200_H
#define RT_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_B200_ABI_VERSION 6

typedef enum rt_status {
  RT_OK = 0,
  RT_ERR_INVALID_ARG = 1,   /* null pointer, negative count, bad enum */
  RT_ERR_CUDA = 2,          /* a CUDA runtime call failed (message in rt_last_error) */
  RT_ERR_NO_DEVICE = 3,     /* no CUDA device: there is NO CPU fallback */
  RT_ERR_BAD_SCENE = 4,     /* BVH/Data arrays inconsistent (index out of range, cycle, ...) */
  RT_ERR_LIMIT = 5,         /* BVH deeper than the traversal stack, too many lights, ... */
  RT_ERR_OUT_OF_MEMORY = 6
} rt_status;

typedef struct rt_vec4 { double x, y, z, w; } rt_vec4;

/* == struct Data (mydata.h:28-72).  Triangle rows (normals_, vertexIdx_, textureIdx_) must be
 * in BVH order, i.e. after BVH::initSoA / rt_bvh_build_median permuted them. */
typedef struct rt_data_soa {
  int tMeshCount_, tVertexCount_, tVertexIdxCount_, tTextCoordCount_, tTextIdxCount_;
  size_t tTexelCount_;
  int *vertexMeshId_;          /* [tVertexCount_] */
  rt_vec4 *vertexPos_;         /* [tVertexCount_] */
  rt_vec4 *vertexNormals_;     /* [tVertexCount_] */
  rt_vec4 *normals_;           /* [tVertexIdxCount_/3] face normals */
  int *vertexIdx_;             /* [tVertexIdxCount_] global vertex ids, 3 per triangle */
  int *textureIdx_;            /* [tTextIdxCount_]  global uv ids, 3 per triangle */
  double *textureCoordinatesU_;/* [tTextCoordCount_] */
  double *textureCoordinatesV_;
  int *firstVertex_;     int *vertexCount_;        /* per mesh */
  int *firstVertexIdx_;  int *vertexIdxCount_;
  int *firstTextCoord_;  int *textCoordCount_;
  int *firstTextIdx_;    int *textIdxCount_;
  rt_vec4 *meshTexels_;        /* [tTexelCount_] all textures concatenated, row-major y*W+x */
  int *meshTexWidth_;          /* -1 when the mesh has no texture */
  int *meshTexHeight_;
  size_t *firstMeshTex_;
  int *meshDrawMode_;          /* 0 FLAT, 1 PHONG */
  rt_vec4 *materialAmbient_;
  rt_vec4 *materialDiffuse_;
  rt_vec4 *materialSpecular_;
  double *materialShininess_;
  double *materialMirror_;
  unsigned char *materialShadowable_; /* C++ bool[] */
} rt_data_soa;

/* == BVH::BVHNodes_SoA (mybvh.h:49-55).  Node 0 is the root; an interior node has
 * triCount_ == 0 and children leftChildIdx_, leftChildIdx_+1.  The reference never records
 * how many of the 2N-1 slots are used (mybvh.h:57), so consumers walk from the root. */
typedef struct rt_bvh_nodes_soa {
  rt_vec4 *bb_min_;
  rt_vec4 *bb_max_;
  int *leftChildIdx_;
  int *firstTriIdx_;
  int *triCount_;
} rt_bvh_nodes_soa;

typedef struct rt_material {
  rt_vec4 ambient, diffuse, specular;
  double shininess, mirror;
  int shadowable;
  int pad_;
} rt_material;

/* Analytic primitives of Raytracer::objects_ (mytracer.cpp:27), in list order.  They are not
 * part of struct Data; the reference's CUDA path cannot render them, the CPU
 * path can (myplane.cpp:22-49, Sphere from the course framework). */
typedef enum rt_object_kind { RT_SPHERE = 0, RT_PLANE = 1 } rt_object_kind;
typedef struct rt_object {
  int kind;           /* rt_object_kind */
  int pad_;
  rt_vec4 center;
  rt_vec4 normal;     /* planes */
  double radius;      /* spheres */
  rt_material material;
} rt_object;

/* Pinhole camera: primary_ray(x,y) = Ray(eye, lower_left + x*x_dir + y*y_dir - eye)
 * (call sites mytracer_gpu.cu:141,208-209). */
typedef struct rt_camera {
  rt_vec4 eye, x_dir, y_dir, lower_left;
  int width, height;
} rt_camera;

typedef struct rt_render_params {
  rt_vec4 background;   /* Raytracer::background_ */
  rt_vec4 ambience;     /* Raytracer::ambience_ */
  int max_depth;        /* Raytracer::max_depth_: mirror bounces per primary ray */
  int spp_grid;         /* s: s x s regular sub-pixel grid (1 = one centre ray), offsets as
                           mytracer_gpu.cu:203-207 */
  /* Image-tile sharding for multi-GPU: 8x4-pixel tiles are numbered row-major; this call
   * renders tile t iff (t / tile_chunk) % tile_world == tile_rank.  1 GPU: world 1, rank 0. */
  int tile_rank, tile_world, tile_chunk;
  int flags;            /* RT_FLAG_* */
  /* Cross-GPU dynamic tail queue (0 / NULL = off).  The LAST dyn_tiles tiles of the tile order are not
   * statically owned: once a rank has drained its static share, its warps claim them dyn_chunk tiles at
   * a time with a system-scope atomicAdd on *dyn_counter, a zero-initialised counter that every rank
   * of the frame passes (rank 0's memory, peer-mapped on the others: rt_ipc_export / rt_ipc_open). */
  int dyn_tiles, dyn_chunk;
  unsigned long long *dyn_counter;
} rt_render_params;

#define RT_FLAG_EXACT_ONLY   1 /* validation: bypass the FP32 triangle filter (every test in FP64) */
#define RT_FLAG_COUNTERS     2 /* run the instrumented kernel and fill rt_counters */
#define RT_FLAG_NO_CLEAR     4 /* keep pixels of tiles owned by other ranks untouched */
#define RT_FLAG_BINARY_NODES 8 /* A/B: ordinary rays walk the binary traversal tree instead of its 4-wide twin */
#define RT_FLAG_WAVEFRONT   16 /* render the frame as a wavefront: TRACE / SHADE kernel pairs per bounce exchanging rays
                                  through path-state columns in HBM (same arithmetic, bit-identical frame) instead of
                                  the one persistent kernel; frames the wavefront does not cover (pixel lists, dynamic
                                  tile queue, RT_FLAG_COUNTERS, > 30 lights) fall back to the persistent kernel.
                                  RT_B200_PIPELINE=wavefront in the environment selects it for every frame. */
#define RT_FLAG_NODES8      64 /* A/B: ordinary rays walk the 8-wide twin of the triangle tree (256-byte records, eight slab
                                  tests per step) instead of the 4-wide one; scenes without one (a tree over analytic
                                  spheres, tiny meshes) ignore the flag.  RT_B200_NODES=8 selects it for every frame. */
#define RT_FLAG_SEQUENTIAL_SAMPLES 128 /* A/B: with a 2x2 or 4x4 sub-pixel grid every lane traces all samples of its pixel one
                                  after the other (the round-1 schedule) instead of one sample per lane (a warp = 8 or 2 pixels,
                                  the samples of a pixel summed in sample order: bit-identical pixels, 4 / 16 x finer work
                                  units, closer rays per warp) */
#define RT_FLAG_RAY_PARALLEL 256 /* the rays of a path on neighbouring lanes (rp_render_kernel): after a hit its shadow rays (one
                                  lane per light) and its mirror ray are traced together, a path advances one bounce per
                                  round and a warp holds 8 paths instead of 32 pixels.  For frames that are chains of
                                  dependent rays rather than throughput (<= 350 k pixels, deep mirror paths); bit-identical
                                  pixels.  Chosen by default for small frames of triangle-mesh scenes with a reflecting material
                                  (mirror depth >= 2, at most 64 k pixel samples per possible hit of a path, no analytic
                                  primitives); RT_B200_PIPELINE=rayparallel selects it for every frame it covers,
                                  RT_B200_PIPELINE=persistent or RT_FLAG_MEGAKERNEL never. */
#define RT_FLAG_MEGAKERNEL  32 /* always the one persistent kernel (overrides RT_B200_PIPELINE) */

/* Output buffers; any pointer may be NULL.  All are W*H entries, index y*W + x
 * (mytracer_gpu.cu:158). */
typedef struct rt_frame_out {
  uint8_t *rgba8;    /* 4 B/pixel: (int)(255*clamp(c,0,1)+0.5), alpha 255 */
  rt_vec4 *pixels;   /* the reference's vec4 pixel buffer: rgb after min(c,1), w = 0 */
  int32_t *hit_id;   /* primary (centre) ray: BVH-order triangle row, N + object index, -1 miss */
  double *hit_t;     /* primary (centre) ray parameter, -1 on miss */
} rt_frame_out;

typedef struct rt_counters {
  unsigned long long rays_primary, rays_shadow, rays_reflect;
  unsigned long long box_tests;      /* child boxes tested (2 per visited interior node) */
  unsigned long long tri_filter;     /* FP32 filter tests */
  unsigned long long tri_exact;      /* FP64 confirmations */
  unsigned long long object_tests;   /* analytic primitive tests */
  unsigned long long shaded_hits;    /* lighting() evaluations */
  unsigned long long texel_fetches;
  unsigned long long nodes_visited;  /* interior node records loaded */
  unsigned long long shadow_trivial; /* shadow rays (counted in rays_shadow) resolved without traversal:
                                        the light's contribution is exactly zero whatever the occlusion */
  unsigned long long shadow_occluded;   /* shadow rays that found an occluder */
  unsigned long long nodes4_visited;    /* of nodes_visited: 128-byte four-child records (4-wide tree) */
  unsigned long long nodes8_visited;    /* of nodes_visited: 256-byte eight-child records (8-wide tree) */
} rt_counters;

typedef struct rt_scene_info {
  int num_triangles, num_meshes, num_objects;
  int num_nodes_ref;       /* nodes reachable from the root in the reference layout */
  int num_nodes_flat;      /* 64-byte two-child records (both trees) */
  int max_depth_bvh;
  int max_leaf_size;
  size_t blob_bytes;       /* flattened scene in HBM */
  double flatten_ms, upload_ms;
  int topology;            /* 0: SAH traversal tree + reference tree for d.x == 0 rays; 1: reference tree only */
  int num_nodes_traversal; /* records of the tree ordinary rays walk */
  int num_nodes_wide;      /* 128-byte four-child records of the 4-wide twins of the traversal trees (0: none) */
  int num_nodes_wide8;     /* 256-byte eight-child records of the triangle tree's 8-wide twin (0: none) */
  int sah_builder;         /* traversal tree built by: 0 none (reference topology / tiny mesh), 1 host threads, 2 the GPU */
  double sah_build_ms;     /* wall time of that build inside flatten_ms (GPU: including its transfers) */
  double sah_device_ms;    /* GPU builder: CUDA-event time of its kernels */
  int flattener;           /* records emitted by: 0 the host threads (rt_flatten.cpp), 1 the GPU (rt_flatten_gpu.cu) */
  int pad_;
} rt_scene_info;

typedef struct rt_scene rt_scene;

/* ---- device ------------------------------------------------------------------------ */
/* init_device() (mytracer_gpu.cu:32-39) with an explicit ordinal. */
int rt_init_device(int device);
const char *rt_last_error(void);
int rt_abi_version(void);

/* ---- scene: flatten + upload -------------------------------------------------------- */
/* Walks the reference BVH from node 0, emits 64-byte float4 two-child node records with
 * outward-rounded FP32 boxes, 48-byte FP32 triangle filter records, FP64 confirmation and
 * shading records, texture objects; uploads one contiguous blob.  objects may be NULL. */
int rt_scene_create(const rt_data_soa *data, const rt_bvh_nodes_soa *bvh, const rt_object *objects,
                    int num_objects, rt_scene **out);
/* Same with options.  By default the flattener ALSO builds its own surface-area-heuristic tree over the
 * same triangles and ordinary rays walk that one: hits are decided by the bit-exact FP64 test and ties by the
 * reference's primitive index (the BVH-order row of Data), so the frame does not depend on the tree — except
 * for rays with direction.x == 0 exactly, which keep walking the reference's tree because its slab test
 * (mybvh.cpp:99-135) rejects boxes through a 0/0 there.  RT_SCENE_REFERENCE_TOPOLOGY keeps every ray on the
 * reference's median-split topology (mybvh.cpp:438-472). */
#define RT_SCENE_REFERENCE_TOPOLOGY 1u
/* The traversal tree is built on the GPU by default (rt_sah_gpu.cu: level-synchronous binned SAH + one thread per small
 * subtree); RT_SCENE_HOST_SAH builds it with the host threads instead (the test twin; a different but equally valid
 * tree: frames are bit-identical either way). */
#define RT_SCENE_HOST_SAH 2u
/* The blob's records (triangle filter / exact / shading records, the node records of all trees, texels) are emitted on
 * the GPU from the uploaded Data / BVH arrays by default (rt_flatten_gpu.cu: 10 M triangles in under a second instead
 * of 7 s on the host threads); RT_SCENE_HOST_FLATTEN (or RT_B200_FLATTEN=host in the environment) keeps the host
 * flattener, which also handles every scene the device path declines (reference topology, a root that is a leaf,
 * malformed input: it words the error).  Frames are bit-identical either way; rt_scene_info.flattener says which ran. */
#define RT_SCENE_HOST_FLATTEN 4u
int rt_scene_create_ex(const rt_data_soa *data, const rt_bvh_nodes_soa *bvh, const rt_object *objects,
                       int num_objects, unsigned scene_flags, rt_scene **out);
int rt_scene_destroy(rt_scene *scene);
/* Persistent-grid width of the following launches: ctas_per_sm resident CTAs per SM (0 = all the kernel's
 * occupancy allows, the default).  A host that keeps D frames in flight on D streams sets occupancy / D so that
 * the D grids run side by side (each warp then renders D times more tiles of its frame, which balances short
 * multi-GPU frames) instead of queueing for each other's CTA slots. */
int rt_scene_set_launch_width(rt_scene *scene, int ctas_per_sm);
int rt_scene_get_info(const rt_scene *scene, rt_scene_info *out);

/* Multi-GPU replication: rank 0 exports the flattened blob (host copy), the caller broadcasts
 * the bytes (NCCL), every other rank adopts them. */
int rt_scene_blob_size(const rt_scene *scene, size_t *bytes);
int rt_scene_blob_export(const rt_scene *scene, void *dst_host, size_t bytes);
/* blob may be a host or a device pointer (is_device != 0: adopted with a D2D copy) */
int rt_scene_create_from_blob(const void *blob, size_t bytes, int is_device, rt_scene **out);
/* device address of the resident blob (for a device-side NCCL broadcast) */
int rt_scene_blob_device_ptr(const rt_scene *scene, void **dptr, size_t *bytes);

/* ---- render --------------------------------------------------------------------------- */
/* Raytracer::compute_image_cuda() equivalent, HOST output buffers: uploads camera/lights,
 * launches the kernel, copies the requested outputs back, synchronises. */
int rt_render(rt_scene *scene, const rt_camera *camera, const rt_vec4 *lights_pos,
              const rt_vec4 *lights_color, int num_lights, const rt_render_params *params,
              const rt_frame_out *out_host);
/* Same with DEVICE output buffers and device-resident light arrays (as the reference's
 * d_lightsPos/d_lightsColor), asynchronous on `stream` (a cudaStream_t, 0 = legacy default). */
int rt_render_device(rt_scene *scene, const rt_camera *camera, const rt_vec4 *d_lights_pos,
                     const rt_vec4 *d_lights_color, int num_lights, const rt_render_params *params,
                     const rt_frame_out *out_device, void *stream);
/* rt_render_device on memory that is part of the loaded module: allocates NOTHING (no cudaMalloc / cudaFree, no
 * allocating library routine), for callers that cannot allocate - code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it.  No counterpart in the reference.
 * Same arguments, arithmetic and (bit-identical) frame as rt_render_device; the frame runs as the wavefront
 * (RT_FLAG_WAVEFRONT schedule) with its path-state scratch in one file-scope __device__ array of 18.6 MB that every GPU
 * loading the library holds until the process exits.  LIMITS (RT_ERR_LIMIT with a message beyond them): at most
 * 65536 path slots = 32 x the 8x4-pixel tiles of this rank's share of the frame (e.g. 256 x 256 pixels; spp_grid does
 * not count, samples are passes over the same slots), at most 8 lights, max_depth <= 30.  RT_FLAG_COUNTERS and the
 * dynamic tile queue (dyn_tiles) are refused with RT_ERR_INVALID_ARG.  The scene's device must be current.  All calls of
 * a process on one GPU share the array, which keeps the last call's contents: two calls that overlap on the GPU
 * (different streams or host threads) race; order them on one stream or with events.  The first call on a
 * device looks the array's device address up (cudaGetSymbolAddress); make it before a stream capture begins.  Does not
 * update rt_last_kernel_ms / rt_last_pipeline. */
int rt_render_device_static(rt_scene *scene, const rt_camera *camera, const rt_vec4 *d_lights_pos,
                            const rt_vec4 *d_lights_color, int num_lights, const rt_render_params *params,
                            const rt_frame_out *out_device, void *stream);
/* Launches of one scene may be in flight on several streams.  Each takes its work counter and event pair from a
 * ring of 8 slots; launch f + 8 is ordered behind launch f (cudaStreamWaitEvent), so more than 8 frames in flight
 * are safe but serialise.  RT_FLAG_COUNTERS launches are synchronous (the counters are read back before the
 * call returns) and share one counter block: do not overlap them with each other. */
/* ---- pipelined host-buffer frames --------------------------------------------------------------
 * The body of Raytracer::compute_image_cuda() (mytracer.cpp:123-159: copyLights :105-118, the launch,
 * the D2H copy mytracer_gpu.cu:112) for a host that renders a SEQUENCE of frames: up to `depth` frames are in
 * flight, the D2H copy of frame f (own copy stream) overlaps the kernel of frame f + 1, and the caller's light
 * arrays are staged so they may be reused as soon as rt_frame_submit returns.
 *   rt_frame_submit   enqueues one frame; blocks only while `depth` earlier frames are still unfinished.
 *                     rgba8_host: W*H*4 bytes, pinned (rt_host_alloc / rt_host_register) for an asynchronous copy.
 *   rt_frame_wait     blocks until the frame of `ticket` (and every earlier one) is in rgba8_host; done_ms (may be
 *                     NULL) = its completion time in ms on the queue's CUDA-event timeline (0 = first submit).
 * opts.zero_copy: the render kernel stores finished pixels straight into rgba8_host (which must then be mapped
 * pinned memory: rt_host_alloc / rt_host_register) - no staging buffer and no copy pass; with tile sharding
 * (rt_render_params.tile_*) the ranks of a multi-GPU job fill ONE host frame in shared memory, each over its own
 * PCIe link.  opts.flush_bytes: benchmark aid, a memset of that many bytes ahead of every frame evicts the L2. */
typedef struct rt_frame_queue rt_frame_queue;
typedef struct rt_frame_queue_opts {
  int depth;           /* frames in flight, 1..32 (0 = 2) */
  int ctas_per_sm;     /* persistent-grid width of each frame, see rt_scene_set_launch_width (0 = all) */
  int zero_copy;
  int pad_;
  size_t flush_bytes;
} rt_frame_queue_opts;
int rt_frame_queue_create(rt_scene *scene, int width, int height, const rt_frame_queue_opts *opts,
                          rt_frame_queue **out);
int rt_frame_submit(rt_frame_queue *queue, const rt_camera *camera, const rt_vec4 *lights_pos,
                    const rt_vec4 *lights_color, int num_lights, const rt_render_params *params,
                    uint8_t *rgba8_host, long long *ticket);
int rt_frame_wait(rt_frame_queue *queue, long long ticket, double *done_ms);
/* The frames submitted from now on render `scene` (a host that switches scenes between frames; frames in flight are
 * not affected). */
int rt_frame_queue_bind(rt_frame_queue *queue, rt_scene *scene);
int rt_frame_queue_destroy(rt_frame_queue *queue);
/* Pinned, device-mapped host memory for frames: allocated here, or an existing block (e.g. a POSIX shared-memory
 * segment that every rank's process maps) registered with the CUDA driver.  dptr (may be NULL): its device
 * address, usable as rt_frame_out.rgba8 of rt_render_device. */
int rt_host_alloc(size_t bytes, void **ptr);
int rt_host_free(void *ptr);
int rt_host_register(void *ptr, size_t bytes, void **dptr);
int rt_host_unregister(void *ptr);

/* ---- peer memory for the multi-GPU frame (one process per GPU) ---------------------------- */
/* Device allocation that can be shared with the other ranks' processes (plain cudaMalloc, zero-filled):
 * rank 0 allocates the frame buffer and the dynamic-queue counters, exports 64-byte handles, the host
 * broadcasts them, every other rank opens them and passes the mapped pointers as rt_frame_out.rgba8 /
 * rt_render_params.dyn_counter: finished pixels are then stored straight into rank 0's frame buffer over
 * NVLink by the render kernel (no gather pass). */
#define RT_IPC_HANDLE_BYTES 64
int rt_shared_alloc(size_t bytes, void **dptr);
int rt_shared_free(void *dptr);
int rt_ipc_export(void *dptr, unsigned char handle[RT_IPC_HANDLE_BYTES]);
int rt_ipc_open(const unsigned char handle[RT_IPC_HANDLE_BYTES], void **dptr);
int rt_ipc_close(void *dptr);

int rt_get_counters(const rt_scene *scene, rt_counters *out);
/* CUDA-event time of the last rt_render / rt_render_device kernel on this scene (ms);
 * synchronises on the end event. */
int rt_last_kernel_ms(const rt_scene *scene, float *ms);
/* Which pipeline the last frame of this scene ran on: 0 = the persistent kernel, 2 = the ray-parallel persistent kernel
 * (RT_FLAG_RAY_PARALLEL), 1 = the wavefront (RT_FLAG_WAVEFRONT
 * asks for it; frames it does not cover fall back). */
int rt_last_pipeline(const rt_scene *scene, int *pipeline);

/* Exact-signature semantics of launch_compute_image_device (mytracer_gpu.cu:44-113): renders
 * into d_pixels (device, vec4), runs the adaptive supersampling pass (subp 4, threshold 0.02,
 * mytracer_gpu.cu:83-105,162-229) through d_tmpPixels, copies the result to h_image (host).
 * data/bvh are flattened on first use and cached in *scene_cache (may start as NULL). */
int rt_launch_compute_image(rt_vec4 *d_pixels, rt_vec4 *d_tmpPixels, rt_vec4 *h_image, int width,
                            int height, const rt_camera *camera, const rt_vec4 *d_lightsPos,
                            const rt_vec4 *d_lightsColor, int nLights, const rt_vec4 *background,
                            const rt_vec4 *ambience, int max_depth, const rt_data_soa *data,
                            const rt_bvh_nodes_soa *bvhNodes, rt_scene **scene_cache);

/* ---- host-side scene preparation (callers of the path; same results as the reference) -- */
/* BVH::initSoA (mybvh.cpp:375-539): median split on axis depth%3 (root depth 1), '<' partition,
 * pairwise child allocation; permutes data->normals_/vertexIdx_/textureIdx_ rows in place.
 * nodes: caller-allocated arrays of 2*N-1 entries.  nodes_used: out. */
int rt_bvh_build_median(rt_data_soa *data, rt_bvh_nodes_soa *nodes, int *nodes_used);
/* The same build on the GPU: level-synchronous, one segmented sort per level for the
 * medians, the reference's in-place partition in closed form, nodes renumbered to the recursion's allocation
 * order.  Same contract and bit-identical results (arrays, numbering, row permutation); host pointers in and
 * out.  device_ms (may be NULL): CUDA-event time of the device part.  RT_ERR_NO_DEVICE without a GPU. */
int rt_bvh_build_median_device(rt_data_soa *data, rt_bvh_nodes_soa *nodes, int *nodes_used, double *device_ms);
/* Mesh::compute_normals (mymesh.cpp:103-163) for one mesh: face normals and angle-weighted
 * vertex normals.  pos/vnormals: nv vec4; idx: 3*nt local vertex ids; fnormals: nt vec4. */
int rt_compute_normals(const rt_vec4 *pos, int nv, const int *idx, int nt, rt_vec4 *fnormals,
                       rt_vec4 *vnormals);
/* The same on the GPU, bit-identical (per-vertex accumulation in the reference's triangle order: stable sort
 * of the corners by vertex).  Host or managed pointers; device_ms (may be NULL) = kernel time. */
int rt_compute_normals_device(const rt_vec4 *pos, int nv, const int *idx, int nt, rt_vec4 *fnormals,
                              rt_vec4 *vnormals, double *device_ms);
/* Pinhole camera from eye/center/up/fovy(deg) (build-defined; DESIGN.md "Conventions"). */
int rt_camera_look_at(const double eye[3], const double center[3], const double up[3], double fovy_deg,
                      int width, int height, rt_camera *out);

/* ---- caller side of the boundary: scene files and images ---------------------- */
/* Scene ingestion == the course framework's read_scene/read_obj (absent; format: rt_scene_io.cpp) followed by
 * Raytracer::build_Data (mytracer.cpp:166-296), Mesh::compute_normals and BVH::initSoA: a `.sce` file with
 * OBJ meshes and PPM/PNG textures becomes exactly the arrays rt_scene_create takes.  Host-side; the flags
 * move the normals and the BVH build to the GPU (same results). */
#define RT_LOAD_DEVICE_BVH     1u
#define RT_LOAD_DEVICE_NORMALS 2u
typedef struct rt_scene_file rt_scene_file;
typedef struct rt_scene_file_view {
  rt_data_soa data;            /* pointers into the rt_scene_file; triangle rows already in BVH order */
  rt_bvh_nodes_soa bvh;
  int nodes_used;
  const rt_object *objects;
  int num_objects;
  rt_camera camera;
  const rt_vec4 *lights_pos, *lights_color;
  int num_lights;
  rt_vec4 background, ambience;
  int max_depth;
} rt_scene_file_view;
int rt_scene_file_load(const char *path, unsigned load_flags, rt_scene_file **out);
int rt_scene_file_get(const rt_scene_file *file, rt_scene_file_view *out);
int rt_scene_file_free(rt_scene_file *file);
/* Frame buffer (index y*W + x, y = 0 is the BOTTOM row: lower_left + y*y_dir) -> upright `.png` (8-bit RGBA)
 * or, for any other extension, binary PPM.  The reference writes its frames through the absent Image class
 * (outputs/ *.png). */
int rt_image_write(const char *path, const uint8_t *rgba8, int width, int height);
/* PPM (P6) or PNG (8-bit, non-interlaced) -> malloc'ed RGBA8, row 0 = top row; release with rt_image_free. */
int rt_image_read(const char *path, int *width, int *height, uint8_t **rgba8);
int rt_image_free(uint8_t *rgba8);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
