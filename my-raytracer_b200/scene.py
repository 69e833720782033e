"""Host-side scene assembly on the caller side of the C-ABI boundary.

Mirrors what the reference's host code does before it launches the kernel:

* ``HostScene``  = Raytracer::build_Data (reference mytracer.cpp:166-296: per-mesh concatenation with
  vbase / tbase / ibase offsets into the ``Data`` SoA of mydata.h) followed by BVH::initSoA
  (mybvh.cpp:375-406, through the C-ABI's ``rt_bvh_build_median``).  Vertex/face normals come from
  ``rt_compute_normals`` (== Mesh::compute_normals, mymesh.cpp:103-163).
* ``DeviceScene`` = the flattened scene resident in HBM (``rt_scene_create``) plus the render calls.

Everything numerical happens in the C-ABI library; this file only lays out numpy arrays.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import capi
from .capi import (BvhNodesSoA, Camera, Counters, DataSoA, FrameOut, Material, Object, RenderParams, SceneInfo, Vec4,
                   Vec4_p, c_double_p, c_int_p, c_size_p, c_ubyte_p, check)


def material(ambient, diffuse, specular, shininess, mirror=0.0, shadowable=True):
    """12 doubles: ambient[3] diffuse[3] specular[3] shininess mirror shadowable."""
    return np.array([*ambient, *diffuse, *specular, shininess, mirror, 1.0 if shadowable else 0.0], dtype=np.float64)


@dataclass
class MeshDesc:
    pos: np.ndarray                 # (nv, 3) float64
    idx: np.ndarray                 # (nt, 3) int32, local vertex ids
    material: np.ndarray            # (12,)
    draw_mode: int = 0              # 0 FLAT, 1 PHONG
    uvidx: np.ndarray | None = None  # (nt, 3) int32, local uv ids
    u: np.ndarray | None = None     # (nuv,)
    v: np.ndarray | None = None
    tex: np.ndarray | None = None   # (H, W, 3) float64, row 0 = top (image row y)

    def __post_init__(self):
        self.pos = np.ascontiguousarray(self.pos, dtype=np.float64).reshape(-1, 3)
        self.idx = np.ascontiguousarray(self.idx, dtype=np.int32).reshape(-1, 3)
        self.material = np.ascontiguousarray(self.material, dtype=np.float64)
        if self.uvidx is not None:
            self.uvidx = np.ascontiguousarray(self.uvidx, dtype=np.int32).reshape(-1, 3)
            self.u = np.ascontiguousarray(self.u, dtype=np.float64)
            self.v = np.ascontiguousarray(self.v, dtype=np.float64)
        if self.tex is not None:
            self.tex = np.ascontiguousarray(self.tex, dtype=np.float64)


@dataclass
class SceneDesc:
    name: str
    meshes: list = field(default_factory=list)
    objects: list = field(default_factory=list)   # ("sphere", center, radius, mat) | ("plane", center, normal, mat)
    lights: list = field(default_factory=list)    # (pos, color)
    camera: dict = field(default_factory=dict)    # eye, center, up, fovy, width, height
    background: tuple = (0.0, 0.0, 0.0)
    ambience: tuple = (0.1, 0.1, 0.1)
    max_depth: int = 3

    @property
    def num_triangles(self):
        return int(sum(m.idx.shape[0] for m in self.meshes))


def _vec4_array(a3):
    """(n,3) float64 -> (n,4) float64 contiguous (== vec4[n])."""
    a3 = np.asarray(a3, dtype=np.float64).reshape(-1, 3)
    out = np.zeros((a3.shape[0], 4), dtype=np.float64)
    out[:, :3] = a3
    return out


def _p(arr, typ):
    return arr.ctypes.data_as(typ) if arr is not None and arr.size else C.cast(None, typ)


def look_at(cam: dict, lib=None) -> Camera:
    lib = lib or capi.load_library()
    out = Camera()
    e = (C.c_double * 3)(*cam["eye"])
    c = (C.c_double * 3)(*cam["center"])
    u = (C.c_double * 3)(*cam["up"])
    check(lib.rt_camera_look_at(e, c, u, float(cam["fovy"]), int(cam["width"]), int(cam["height"]), C.byref(out)), lib)
    return out


class HostScene:
    """Data SoA (mydata.h) + BVHNodes_SoA (mybvh.h) as numpy arrays, built like the reference does."""

    def __init__(self, desc: SceneDesc, lib=None, gpu_build=False, gpu_normals=False):
        """gpu_build: BVH through rt_bvh_build_median_device; gpu_normals: normals through
        rt_compute_normals_device (same results, computed on the GPU)."""
        self.lib = lib or capi.load_library()
        self.desc = desc
        lib = self.lib
        meshes = desc.meshes
        M = len(meshes)
        nv = sum(m.pos.shape[0] for m in meshes)
        nt = sum(m.idx.shape[0] for m in meshes)
        nuv = sum((m.u.shape[0] if m.u is not None else 0) for m in meshes)
        ntex = sum((m.tex.shape[0] * m.tex.shape[1] if m.tex is not None else 0) for m in meshes)
        a = {}
        a["vertexMeshId"] = np.zeros(nv, np.int32)
        a["vertexPos"] = np.zeros((nv, 4))
        a["vertexNormals"] = np.zeros((nv, 4))
        a["normals"] = np.zeros((nt, 4))
        a["vertexIdx"] = np.zeros(3 * nt, np.int32)
        a["textureIdx"] = np.zeros(3 * nt, np.int32)
        a["U"] = np.zeros(nuv)
        a["V"] = np.zeros(nuv)
        for k in ("firstVertex", "vertexCount", "firstVertexIdx", "vertexIdxCount", "firstTextCoord", "textCoordCount",
                  "firstTextIdx", "textIdxCount", "texW", "texH", "drawMode"):
            a[k] = np.zeros(M, np.int32)
        a["firstMeshTex"] = np.zeros(M, np.uint64)
        a["texels"] = np.zeros((ntex, 4))
        a["ka"] = np.zeros((M, 4)); a["kd"] = np.zeros((M, 4)); a["ks"] = np.zeros((M, 4))
        a["shin"] = np.zeros(M); a["mirror"] = np.zeros(M); a["shadowable"] = np.zeros(M, np.uint8)
        vbase = tbase = ibase = 0
        texoff = 0
        self.mesh_tri_base = []
        for mi, m in enumerate(meshes):
            n_v, n_t = m.pos.shape[0], m.idx.shape[0]
            # Mesh::compute_normals (mymesh.cpp:103-163) through the C ABI
            pos4 = _vec4_array(m.pos)
            fn = np.zeros((n_t, 4)); vn = np.zeros((n_v, 4))
            idx = np.ascontiguousarray(m.idx.reshape(-1))
            if gpu_normals:
                check(lib.rt_compute_normals_device(_p(pos4, Vec4_p), n_v, _p(idx, c_int_p), n_t, _p(fn, Vec4_p),
                                                    _p(vn, Vec4_p), None), lib)
            else:
                check(lib.rt_compute_normals(_p(pos4, Vec4_p), n_v, _p(idx, c_int_p), n_t, _p(fn, Vec4_p), _p(vn, Vec4_p)), lib)
            # per-vertex block (mytracer.cpp:222-232)
            a["vertexPos"][vbase:vbase + n_v] = pos4
            a["vertexNormals"][vbase:vbase + n_v] = vn
            a["vertexMeshId"][vbase:vbase + n_v] = mi
            a["firstVertex"][mi], a["vertexCount"][mi] = vbase, n_v
            # texture coordinates (mytracer.cpp:234-241)
            n_uv = m.u.shape[0] if m.u is not None else 0
            if n_uv:
                a["U"][tbase:tbase + n_uv] = m.u
                a["V"][tbase:tbase + n_uv] = m.v
            a["firstTextCoord"][mi], a["textCoordCount"][mi] = tbase, n_uv
            # triangle indices with global offsets (mytracer.cpp:243-258)
            a["vertexIdx"][ibase:ibase + 3 * n_t] = vbase + idx
            uvidx = m.uvidx.reshape(-1) if m.uvidx is not None else np.zeros(3 * n_t, np.int32)
            a["textureIdx"][ibase:ibase + 3 * n_t] = tbase + uvidx
            a["normals"][ibase // 3: ibase // 3 + n_t] = fn
            a["firstVertexIdx"][mi], a["vertexIdxCount"][mi] = ibase, 3 * n_t
            a["firstTextIdx"][mi], a["textIdxCount"][mi] = ibase, 3 * n_t
            # texture (mytracer.cpp:260-276)
            if m.tex is not None:
                H, W = m.tex.shape[:2]
                a["texW"][mi], a["texH"][mi], a["firstMeshTex"][mi] = W, H, texoff
                a["texels"][texoff:texoff + W * H, :3] = m.tex.reshape(-1, 3)
                texoff += W * H
            else:
                a["texW"][mi] = a["texH"][mi] = -1
                a["firstMeshTex"][mi] = np.uint64(2**64 - 1)
            a["drawMode"][mi] = m.draw_mode
            a["ka"][mi, :3], a["kd"][mi, :3], a["ks"][mi, :3] = m.material[0:3], m.material[3:6], m.material[6:9]
            a["shin"][mi], a["mirror"][mi], a["shadowable"][mi] = m.material[9], m.material[10], int(m.material[11] != 0)
            self.mesh_tri_base.append(ibase // 3)
            vbase += n_v; ibase += 3 * n_t; tbase += n_uv
        self.arrays = a
        d = DataSoA()
        d.tMeshCount_, d.tVertexCount_, d.tVertexIdxCount_ = M, nv, 3 * nt
        d.tTextCoordCount_, d.tTextIdxCount_, d.tTexelCount_ = nuv, 3 * nt, ntex
        d.vertexMeshId_ = _p(a["vertexMeshId"], c_int_p)
        d.vertexPos_ = _p(a["vertexPos"], Vec4_p); d.vertexNormals_ = _p(a["vertexNormals"], Vec4_p)
        d.normals_ = _p(a["normals"], Vec4_p)
        d.vertexIdx_ = _p(a["vertexIdx"], c_int_p); d.textureIdx_ = _p(a["textureIdx"], c_int_p)
        d.textureCoordinatesU_ = _p(a["U"], c_double_p); d.textureCoordinatesV_ = _p(a["V"], c_double_p)
        d.firstVertex_ = _p(a["firstVertex"], c_int_p); d.vertexCount_ = _p(a["vertexCount"], c_int_p)
        d.firstVertexIdx_ = _p(a["firstVertexIdx"], c_int_p); d.vertexIdxCount_ = _p(a["vertexIdxCount"], c_int_p)
        d.firstTextCoord_ = _p(a["firstTextCoord"], c_int_p); d.textCoordCount_ = _p(a["textCoordCount"], c_int_p)
        d.firstTextIdx_ = _p(a["firstTextIdx"], c_int_p); d.textIdxCount_ = _p(a["textIdxCount"], c_int_p)
        d.meshTexels_ = _p(a["texels"], Vec4_p)
        d.meshTexWidth_ = _p(a["texW"], c_int_p); d.meshTexHeight_ = _p(a["texH"], c_int_p)
        d.firstMeshTex_ = _p(a["firstMeshTex"], c_size_p)
        d.meshDrawMode_ = _p(a["drawMode"], c_int_p)
        d.materialAmbient_ = _p(a["ka"], Vec4_p); d.materialDiffuse_ = _p(a["kd"], Vec4_p)
        d.materialSpecular_ = _p(a["ks"], Vec4_p)
        d.materialShininess_ = _p(a["shin"], c_double_p); d.materialMirror_ = _p(a["mirror"], c_double_p)
        d.materialShadowable_ = _p(a["shadowable"], c_ubyte_p)
        self.data = d
        self.num_triangles = nt
        # BVH::initSoA (mybvh.cpp:375-406): 2N-1 slots, rows of normals_/vertexIdx_/textureIdx_ permuted
        cap = max(2 * nt - 1, 1)
        b = {"bbmin": np.zeros((cap, 4)), "bbmax": np.zeros((cap, 4)), "left": np.zeros(cap, np.int32),
             "first": np.zeros(cap, np.int32), "count": np.zeros(cap, np.int32)}
        self.bvh_arrays = b
        bv = BvhNodesSoA(_p(b["bbmin"], Vec4_p), _p(b["bbmax"], Vec4_p), _p(b["left"], c_int_p),
                         _p(b["first"], c_int_p), _p(b["count"], c_int_p))
        self.bvh = bv
        self.tri_order_before = None
        used = C.c_int(0)
        if nt > 0:
            # remember the original row of every triangle through a tag carried in a copy of vertexIdx_
            orig_vidx = a["vertexIdx"].reshape(-1, 3).copy()
            self.bvh_device_ms = None
            if gpu_build:
                ms = C.c_double(0.0)
                check(lib.rt_bvh_build_median_device(C.byref(d), C.byref(bv), C.byref(used), C.byref(ms)), lib)
                self.bvh_device_ms = ms.value
            else:
                check(lib.rt_bvh_build_median(C.byref(d), C.byref(bv), C.byref(used)), lib)
            self._orig_vidx = orig_vidx
        self.nodes_used = used.value
        # analytic primitives
        objs = (Object * max(len(desc.objects), 1))()
        for k, o in enumerate(desc.objects):
            kind = o[0]
            if kind == "sphere":
                objs[k] = Object(capi.RT_SPHERE, 0, Vec4.of(o[1]), Vec4(0, 0, 0, 0), float(o[2]), Material.of(o[3]))
            elif kind == "plane":
                objs[k] = Object(capi.RT_PLANE, 0, Vec4.of(o[1]), Vec4.of(o[2]), 0.0, Material.of(o[3]))
            else:
                raise ValueError(kind)
        self.objects = objs
        self.num_objects = len(desc.objects)
        self.camera = look_at(desc.camera, lib) if desc.camera else None
        self.lights_pos = _vec4_array([l[0] for l in desc.lights]) if desc.lights else np.zeros((0, 4))
        self.lights_color = _vec4_array([l[1] for l in desc.lights]) if desc.lights else np.zeros((0, 4))

    def params(self, spp_grid=1, flags=0, tile=(0, 1, 1), max_depth=None, dyn=(0, 0, None)) -> RenderParams:
        """tile = (rank, world, chunk) static sharding; dyn = (dyn_tiles, dyn_chunk, counter device pointer):
        the cross-GPU dynamic tail queue (include/rt_b200.h)."""
        d = self.desc
        return RenderParams(Vec4.of(d.background), Vec4.of(d.ambience), d.max_depth if max_depth is None else max_depth,
                            spp_grid, tile[0], tile[1], tile[2], flags, int(dyn[0]), int(dyn[1]), dyn[2])


class DeviceScene:
    """Flattened scene resident in HBM + render entry points (host-buffer path)."""

    def __init__(self, host: HostScene = None, lib=None, blob: bytes | None = None, blob_device_ptr=None, blob_bytes=0,
                 scene_flags=0):
        """scene_flags: capi.RT_SCENE_REFERENCE_TOPOLOGY keeps every ray on the reference's median-split tree
        (default: ordinary rays walk the library's own SAH tree; same frame either way)."""
        self.lib = lib or (host.lib if host is not None else capi.load_library())
        self.host = host
        h = C.c_void_p()
        if host is not None:
            check(self.lib.rt_scene_create_ex(C.byref(host.data), C.byref(host.bvh), host.objects, host.num_objects,
                                              int(scene_flags), C.byref(h)), self.lib)
        elif blob is not None:
            buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
            check(self.lib.rt_scene_create_from_blob(buf, len(blob), 0, C.byref(h)), self.lib)
        else:
            check(self.lib.rt_scene_create_from_blob(C.c_void_p(blob_device_ptr), blob_bytes, 1, C.byref(h)), self.lib)
        self.handle = h

    def close(self):
        if self.handle:
            self.lib.rt_scene_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> dict:
        i = SceneInfo()
        check(self.lib.rt_scene_get_info(self.handle, C.byref(i)), self.lib)
        return i.as_dict()

    def export_blob(self) -> bytes:
        n = C.c_size_t()
        check(self.lib.rt_scene_blob_size(self.handle, C.byref(n)), self.lib)
        buf = (C.c_ubyte * n.value)()
        check(self.lib.rt_scene_blob_export(self.handle, buf, n.value), self.lib)
        return bytes(buf)

    def blob_device(self):
        p = C.c_void_p(); n = C.c_size_t()
        check(self.lib.rt_scene_blob_device_ptr(self.handle, C.byref(p), C.byref(n)), self.lib)
        return p.value, n.value

    def render(self, camera: Camera, lights_pos, lights_color, params: RenderParams,
               want=("rgba8", "pixels", "hit_id", "hit_t")) -> dict:
        """rt_render with freshly allocated numpy host buffers."""
        W, H = camera.width, camera.height
        out = {}
        fo = FrameOut()
        if "rgba8" in want:
            out["rgba8"] = np.zeros((H, W, 4), np.uint8); fo.rgba8 = out["rgba8"].ctypes.data
        if "pixels" in want:
            out["pixels"] = np.zeros((H, W, 4), np.float64); fo.pixels = out["pixels"].ctypes.data
        if "hit_id" in want:
            out["hit_id"] = np.full((H, W), -2, np.int32); fo.hit_id = out["hit_id"].ctypes.data
        if "hit_t" in want:
            out["hit_t"] = np.zeros((H, W), np.float64); fo.hit_t = out["hit_t"].ctypes.data
        lp = np.ascontiguousarray(lights_pos, np.float64); lc = np.ascontiguousarray(lights_color, np.float64)
        check(self.lib.rt_render(self.handle, C.byref(camera), _p(lp, Vec4_p), _p(lc, Vec4_p), lp.shape[0],
                                 C.byref(params), C.byref(fo)), self.lib)
        if params.flags & capi.RT_FLAG_COUNTERS:
            c = Counters()
            check(self.lib.rt_get_counters(self.handle, C.byref(c)), self.lib)
            out["counters"] = c.as_dict()
        return out

    def last_pipeline(self) -> str:
        """Which pipeline the last frame ran on: "megakernel" (one persistent kernel) or "wavefront"."""
        k = C.c_int()
        check(self.lib.rt_last_pipeline(self.handle, C.byref(k)), self.lib)
        return {1: "wavefront", 2: "rayparallel"}.get(k.value, "megakernel")

    def last_kernel_ms(self) -> float:
        ms = C.c_float()
        check(self.lib.rt_last_kernel_ms(self.handle, C.byref(ms)), self.lib)
        return ms.value


def quantize_rgb8(rgb):
    """The build-defined 8-bit conversion: (int)(255*clamp(c,0,1) + 0.5)."""
    c = np.clip(np.asarray(rgb, dtype=np.float64), 0.0, 1.0)
    return (255.0 * c + 0.5).astype(np.int64).astype(np.uint8)
