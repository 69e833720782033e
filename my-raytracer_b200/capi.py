"""ctypes mirror of include/rt_b200.h (the C-ABI boundary).

The shared library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a) as
``my-raytracer_b200/librt_b200.so``.  There is no CPU fallback: if the library is missing the
import fails loudly, and any render call without a CUDA device returns RT_ERR_NO_DEVICE.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librt_b200.so")

RT_OK = 0
RT_ERR_INVALID_ARG, RT_ERR_CUDA, RT_ERR_NO_DEVICE, RT_ERR_BAD_SCENE, RT_ERR_LIMIT, RT_ERR_OOM = 1, 2, 3, 4, 5, 6
RT_SPHERE, RT_PLANE = 0, 1
RT_FLAG_EXACT_ONLY, RT_FLAG_COUNTERS, RT_FLAG_NO_CLEAR, RT_FLAG_BINARY_NODES = 1, 2, 4, 8
RT_FLAG_WAVEFRONT, RT_FLAG_MEGAKERNEL, RT_FLAG_NODES8, RT_FLAG_SEQUENTIAL_SAMPLES, RT_FLAG_RAY_PARALLEL = 16, 32, 64, 128, 256
RT_SCENE_REFERENCE_TOPOLOGY = 1
RT_SCENE_HOST_SAH = 2
RT_SCENE_HOST_FLATTEN = 4

c_int_p = C.POINTER(C.c_int)
c_double_p = C.POINTER(C.c_double)
c_size_p = C.POINTER(C.c_size_t)
c_ubyte_p = C.POINTER(C.c_ubyte)


class Vec4(C.Structure):
    _fields_ = [("x", C.c_double), ("y", C.c_double), ("z", C.c_double), ("w", C.c_double)]

    @classmethod
    def of(cls, v):
        return cls(float(v[0]), float(v[1]), float(v[2]), 0.0)


Vec4_p = C.POINTER(Vec4)


class DataSoA(C.Structure):
    """== struct Data, reference mydata.h:28-72 (field for field)."""
    _fields_ = [
        ("tMeshCount_", C.c_int), ("tVertexCount_", C.c_int), ("tVertexIdxCount_", C.c_int),
        ("tTextCoordCount_", C.c_int), ("tTextIdxCount_", C.c_int), ("tTexelCount_", C.c_size_t),
        ("vertexMeshId_", c_int_p), ("vertexPos_", Vec4_p), ("vertexNormals_", Vec4_p),
        ("normals_", Vec4_p), ("vertexIdx_", c_int_p), ("textureIdx_", c_int_p),
        ("textureCoordinatesU_", c_double_p), ("textureCoordinatesV_", c_double_p),
        ("firstVertex_", c_int_p), ("vertexCount_", c_int_p),
        ("firstVertexIdx_", c_int_p), ("vertexIdxCount_", c_int_p),
        ("firstTextCoord_", c_int_p), ("textCoordCount_", c_int_p),
        ("firstTextIdx_", c_int_p), ("textIdxCount_", c_int_p),
        ("meshTexels_", Vec4_p), ("meshTexWidth_", c_int_p), ("meshTexHeight_", c_int_p),
        ("firstMeshTex_", c_size_p), ("meshDrawMode_", c_int_p),
        ("materialAmbient_", Vec4_p), ("materialDiffuse_", Vec4_p), ("materialSpecular_", Vec4_p),
        ("materialShininess_", c_double_p), ("materialMirror_", c_double_p),
        ("materialShadowable_", c_ubyte_p),
    ]


class BvhNodesSoA(C.Structure):
    """== BVH::BVHNodes_SoA, reference mybvh.h:49-55."""
    _fields_ = [("bb_min_", Vec4_p), ("bb_max_", Vec4_p), ("leftChildIdx_", c_int_p),
                ("firstTriIdx_", c_int_p), ("triCount_", c_int_p)]


class Material(C.Structure):
    _fields_ = [("ambient", Vec4), ("diffuse", Vec4), ("specular", Vec4), ("shininess", C.c_double),
                ("mirror", C.c_double), ("shadowable", C.c_int), ("pad_", C.c_int)]

    @classmethod
    def of(cls, m12):
        return cls(Vec4.of(m12[0:3]), Vec4.of(m12[3:6]), Vec4.of(m12[6:9]), float(m12[9]), float(m12[10]),
                   int(m12[11] != 0), 0)


class Object(C.Structure):
    _fields_ = [("kind", C.c_int), ("pad_", C.c_int), ("center", Vec4), ("normal", Vec4),
                ("radius", C.c_double), ("material", Material)]


class Camera(C.Structure):
    _fields_ = [("eye", Vec4), ("x_dir", Vec4), ("y_dir", Vec4), ("lower_left", Vec4),
                ("width", C.c_int), ("height", C.c_int)]


class RenderParams(C.Structure):
    _fields_ = [("background", Vec4), ("ambience", Vec4), ("max_depth", C.c_int), ("spp_grid", C.c_int),
                ("tile_rank", C.c_int), ("tile_world", C.c_int), ("tile_chunk", C.c_int), ("flags", C.c_int),
                ("dyn_tiles", C.c_int), ("dyn_chunk", C.c_int), ("dyn_counter", C.c_void_p)]


class FrameOut(C.Structure):
    _fields_ = [("rgba8", C.c_void_p), ("pixels", C.c_void_p), ("hit_id", C.c_void_p), ("hit_t", C.c_void_p)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_ulonglong) for n in (
        "rays_primary", "rays_shadow", "rays_reflect", "box_tests", "tri_filter", "tri_exact",
        "object_tests", "shaded_hits", "texel_fetches", "nodes_visited", "shadow_trivial",
        "shadow_occluded", "nodes4_visited", "nodes8_visited")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class SceneInfo(C.Structure):
    _fields_ = [("num_triangles", C.c_int), ("num_meshes", C.c_int), ("num_objects", C.c_int),
                ("num_nodes_ref", C.c_int), ("num_nodes_flat", C.c_int), ("max_depth_bvh", C.c_int),
                ("max_leaf_size", C.c_int), ("blob_bytes", C.c_size_t), ("flatten_ms", C.c_double),
                ("upload_ms", C.c_double), ("topology", C.c_int), ("num_nodes_traversal", C.c_int),
                ("num_nodes_wide", C.c_int), ("num_nodes_wide8", C.c_int), ("sah_builder", C.c_int),
                ("sah_build_ms", C.c_double), ("sah_device_ms", C.c_double), ("flattener", C.c_int), ("pad_", C.c_int)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class FrameQueueOpts(C.Structure):
    _fields_ = [("depth", C.c_int), ("ctas_per_sm", C.c_int), ("zero_copy", C.c_int), ("pad_", C.c_int),
                ("flush_bytes", C.c_size_t)]


class SceneFileView(C.Structure):
    _fields_ = [("data", DataSoA), ("bvh", BvhNodesSoA), ("nodes_used", C.c_int), ("objects", C.POINTER(Object)),
                ("num_objects", C.c_int), ("camera", Camera), ("lights_pos", Vec4_p), ("lights_color", Vec4_p),
                ("num_lights", C.c_int), ("background", Vec4), ("ambience", Vec4), ("max_depth", C.c_int)]


RT_LOAD_DEVICE_BVH, RT_LOAD_DEVICE_NORMALS = 1, 2

# every symbol include/rt_b200.h declares: name -> (restype, argtypes)
PROTOTYPES = {
    "rt_init_device": (C.c_int, [C.c_int]),
    "rt_last_error": (C.c_char_p, []),
    "rt_abi_version": (C.c_int, []),
    "rt_scene_create": (C.c_int, [C.POINTER(DataSoA), C.POINTER(BvhNodesSoA), C.POINTER(Object), C.c_int,
                                  C.POINTER(C.c_void_p)]),
    "rt_scene_create_ex": (C.c_int, [C.POINTER(DataSoA), C.POINTER(BvhNodesSoA), C.POINTER(Object), C.c_int, C.c_uint,
                                     C.POINTER(C.c_void_p)]),
    "rt_scene_destroy": (C.c_int, [C.c_void_p]),
    "rt_scene_set_launch_width": (C.c_int, [C.c_void_p, C.c_int]),
    "rt_scene_get_info": (C.c_int, [C.c_void_p, C.POINTER(SceneInfo)]),
    "rt_scene_blob_size": (C.c_int, [C.c_void_p, c_size_p]),
    "rt_scene_blob_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "rt_scene_create_from_blob": (C.c_int, [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_void_p)]),
    "rt_scene_blob_device_ptr": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), c_size_p]),
    "rt_render": (C.c_int, [C.c_void_p, C.POINTER(Camera), Vec4_p, Vec4_p, C.c_int, C.POINTER(RenderParams),
                            C.POINTER(FrameOut)]),
    "rt_render_device": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_void_p, C.c_void_p, C.c_int,
                                   C.POINTER(RenderParams), C.POINTER(FrameOut), C.c_void_p]),
    "rt_render_device_static": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_void_p, C.c_void_p, C.c_int,
                                   C.POINTER(RenderParams), C.POINTER(FrameOut), C.c_void_p]),
    "rt_frame_queue_create": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(FrameQueueOpts), C.POINTER(C.c_void_p)]),
    "rt_frame_submit": (C.c_int, [C.c_void_p, C.POINTER(Camera), Vec4_p, Vec4_p, C.c_int, C.POINTER(RenderParams),
                                  C.c_void_p, C.POINTER(C.c_longlong)]),
    "rt_frame_wait": (C.c_int, [C.c_void_p, C.c_longlong, c_double_p]),
    "rt_frame_queue_bind": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rt_frame_queue_destroy": (C.c_int, [C.c_void_p]),
    "rt_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "rt_host_free": (C.c_int, [C.c_void_p]),
    "rt_host_register": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "rt_host_unregister": (C.c_int, [C.c_void_p]),
    "rt_shared_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "rt_shared_free": (C.c_int, [C.c_void_p]),
    "rt_ipc_export": (C.c_int, [C.c_void_p, c_ubyte_p]),
    "rt_ipc_open": (C.c_int, [c_ubyte_p, C.POINTER(C.c_void_p)]),
    "rt_ipc_close": (C.c_int, [C.c_void_p]),
    "rt_get_counters": (C.c_int, [C.c_void_p, C.POINTER(Counters)]),
    "rt_last_kernel_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "rt_last_pipeline": (C.c_int, [C.c_void_p, c_int_p]),
    "rt_launch_compute_image": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(Camera),
                                          C.c_void_p, C.c_void_p, C.c_int, Vec4_p, Vec4_p, C.c_int,
                                          C.POINTER(DataSoA), C.POINTER(BvhNodesSoA), C.POINTER(C.c_void_p)]),
    "rt_bvh_build_median": (C.c_int, [C.POINTER(DataSoA), C.POINTER(BvhNodesSoA), c_int_p]),
    "rt_bvh_build_median_device": (C.c_int, [C.POINTER(DataSoA), C.POINTER(BvhNodesSoA), c_int_p, c_double_p]),
    "rt_compute_normals": (C.c_int, [Vec4_p, C.c_int, c_int_p, C.c_int, Vec4_p, Vec4_p]),
    "rt_compute_normals_device": (C.c_int, [Vec4_p, C.c_int, c_int_p, C.c_int, Vec4_p, Vec4_p, c_double_p]),
    "rt_scene_file_load": (C.c_int, [C.c_char_p, C.c_uint, C.POINTER(C.c_void_p)]),
    "rt_scene_file_get": (C.c_int, [C.c_void_p, C.POINTER(SceneFileView)]),
    "rt_scene_file_free": (C.c_int, [C.c_void_p]),
    "rt_image_write": (C.c_int, [C.c_char_p, C.c_void_p, C.c_int, C.c_int]),
    "rt_image_read": (C.c_int, [C.c_char_p, c_int_p, c_int_p, C.POINTER(C.c_void_p)]),
    "rt_image_free": (C.c_int, [C.c_void_p]),
    "rt_camera_look_at": (C.c_int, [c_double_p, c_double_p, c_double_p, C.c_double, C.c_int, C.c_int,
                                    C.POINTER(Camera)]),
}

_lib = None


class RtError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"rt_b200 status {status}: {msg}")
        self.status = status


def load_library(path: str | None = None):
    """Load librt_b200.so and bind every prototype.  Raises loudly when it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("RT_B200_LIB") or LIB_PATH   # RT_B200_LIB: kernel-variant experiments only
    if not os.path.exists(p):
        raise RuntimeError(
            f"{p} not found: the CUDA extension is not built (run `python -c 'import __graft_entry__ as g; "
            f"g.build()'`). There is no CPU fallback for the render path.")
    lib = C.CDLL(p)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.rt_abi_version() != 6:
        raise RuntimeError("rt_b200 ABI version mismatch")
    if path is None:
        _lib = lib
    return lib


def check(status: int, lib=None):
    if status != RT_OK:
        lib = lib or load_library()
        msg = lib.rt_last_error()
        raise RtError(status, msg.decode() if msg else "?")
