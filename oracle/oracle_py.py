"""TEST INFRASTRUCTURE — ctypes wrapper of the oracle libraries (oracle/oracle_abi.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this.  Two interchangeable libraries implement the ABI:

* ``reference``: oracle/_ref/libref_oracle.so — the reference's own CPU sources compiled
  unmodified from /root/reference plus the restated framework glue (ref_glue.cpp);
  ``reference_fast`` is the same at -O3 -march=x86-64-v3 for CPU-baseline timing.
* ``port``: oracle/librt_oracle.so — the plain C++ restatement (rt_oracle.cpp).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PATHS = {
    "reference": os.path.join(_HERE, "_ref", "libref_oracle.so"),
    "reference_fast": os.path.join(_HERE, "_ref", "libref_oracle_fast.so"),
    "port": os.path.join(_HERE, "librt_oracle.so"),
}
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_libs = {}


def available(kind: str) -> bool:
    return os.path.exists(PATHS[kind])


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _i(a):
    return a.ctypes.data_as(_ip) if a is not None else None


def _v(x):
    return (C.c_double * 3)(*[float(t) for t in x])


def load(kind: str = "reference"):
    if kind in _libs:
        return _libs[kind]
    lib = C.CDLL(PATHS[kind])
    lib.orc_new.restype = C.c_void_p
    lib.orc_free.argtypes = [C.c_void_p]
    lib.orc_kind.restype = C.c_char_p
    lib.orc_add_mesh.argtypes = [C.c_void_p, C.c_int, _dp, C.c_int, _ip, _ip, C.c_int, _dp, _dp, C.c_int, C.c_int, _dp,
                                 _dp, C.c_int]
    lib.orc_add_sphere.argtypes = [C.c_void_p, _dp, C.c_double, _dp]
    lib.orc_add_plane.argtypes = [C.c_void_p, _dp, _dp, _dp]
    lib.orc_add_light.argtypes = [C.c_void_p, _dp, _dp]
    lib.orc_set_camera.argtypes = [C.c_void_p, _dp, _dp, _dp, C.c_double, C.c_int, C.c_int]
    lib.orc_get_camera.argtypes = [C.c_void_p, _dp]
    lib.orc_set_globals.argtypes = [C.c_void_p, _dp, _dp, C.c_int]
    lib.orc_build.argtypes = [C.c_void_p]
    lib.orc_num_triangles.argtypes = [C.c_void_p]
    lib.orc_num_nodes.argtypes = [C.c_void_p]
    lib.orc_get_bvh.argtypes = [C.c_void_p, _dp, _dp, _ip, _ip, _ip]
    lib.orc_get_tri_order.argtypes = [C.c_void_p, _ip]
    lib.orc_get_normals.argtypes = [C.c_void_p, C.c_int, _dp, _dp]
    lib.orc_render.restype = C.c_double
    lib.orc_render.argtypes = [C.c_void_p, _dp, _ip, _dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.orc_ray_counts.argtypes = [C.c_void_p, C.POINTER(C.c_longlong)]
    lib.orc_trace_rays.argtypes = [C.c_void_p, C.c_int, _dp, _dp, _dp, _ip, _dp]
    lib.orc_tri_test.argtypes = [_dp, _dp, _dp, _dp, _dp, _dp]
    lib.orc_aabb_test.argtypes = [_dp, _dp, _dp, _dp, _dp]
    lib.orc_plane_test.argtypes = [_dp, _dp, _dp, _dp, _dp]
    lib.orc_sphere_test.argtypes = [_dp, C.c_double, _dp, _dp, _dp]
    lib.orc_lighting.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, _dp]
    lib.orc_texture.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, _dp]
    _libs[kind] = lib
    return lib


class OracleScene:
    """A scene loaded into an oracle library from a SceneDesc (duck-typed: .meshes .objects .lights
    .camera .background .ambience .max_depth)."""

    def __init__(self, desc, kind: str = "reference"):
        self.lib = load(kind)
        self.kind = kind
        self.h = C.c_void_p(self.lib.orc_new())
        self.desc = desc
        lib = self.lib
        for m in desc.meshes:
            pos = np.ascontiguousarray(m.pos, np.float64)
            idx = np.ascontiguousarray(m.idx, np.int32)
            uvidx = np.ascontiguousarray(m.uvidx, np.int32) if m.uvidx is not None else None
            u = np.ascontiguousarray(m.u, np.float64) if m.u is not None else None
            v = np.ascontiguousarray(m.v, np.float64) if m.v is not None else None
            tex = np.ascontiguousarray(m.tex, np.float64) if m.tex is not None else None
            tw, th = (tex.shape[1], tex.shape[0]) if tex is not None else (0, 0)
            mat = np.ascontiguousarray(m.material, np.float64)
            lib.orc_add_mesh(self.h, pos.shape[0], _d(pos), idx.shape[0], _i(idx), _i(uvidx), 0 if u is None else u.shape[0],
                             _d(u), _d(v), tw, th, _d(tex), _d(mat), int(m.draw_mode))
        for o in desc.objects:
            mat = np.ascontiguousarray(o[3], np.float64)
            if o[0] == "sphere":
                lib.orc_add_sphere(self.h, _v(o[1]), float(o[2]), _d(mat))
            else:
                lib.orc_add_plane(self.h, _v(o[1]), _v(o[2]), _d(mat))
        for p, c in desc.lights:
            lib.orc_add_light(self.h, _v(p), _v(c))
        cam = desc.camera
        if cam:
            lib.orc_set_camera(self.h, _v(cam["eye"]), _v(cam["center"]), _v(cam["up"]), float(cam["fovy"]),
                               int(cam["width"]), int(cam["height"]))
        lib.orc_set_globals(self.h, _v(desc.background), _v(desc.ambience), int(desc.max_depth))
        lib.orc_build(self.h)

    def __del__(self):
        try:
            if self.h:
                self.lib.orc_free(self.h)
                self.h = None
        except Exception:
            pass

    def set_camera(self, cam):
        self.lib.orc_set_camera(self.h, _v(cam["eye"]), _v(cam["center"]), _v(cam["up"]), float(cam["fovy"]),
                                int(cam["width"]), int(cam["height"]))
        self.desc.camera = cam

    def camera(self):
        out = np.zeros(12)
        self.lib.orc_get_camera(self.h, _d(out))
        return out.reshape(4, 3)

    @property
    def num_triangles(self):
        return self.lib.orc_num_triangles(self.h)

    @property
    def num_nodes(self):
        return self.lib.orc_num_nodes(self.h)

    def bvh(self):
        n = self.num_nodes
        bbmin = np.zeros((n, 3)); bbmax = np.zeros((n, 3))
        left = np.zeros(n, np.int32); first = np.zeros(n, np.int32); count = np.zeros(n, np.int32)
        if n:
            self.lib.orc_get_bvh(self.h, _d(bbmin), _d(bbmax), _i(left), _i(first), _i(count))
        return dict(bbmin=bbmin, bbmax=bbmax, left=left, first=first, count=count)

    def tri_order(self):
        perm = np.zeros(self.num_triangles, np.int32)
        if perm.size:
            self.lib.orc_get_tri_order(self.h, _i(perm))
        return perm

    def normals(self, mesh: int):
        m = self.desc.meshes[mesh]
        fn = np.zeros((m.idx.shape[0], 3)); vn = np.zeros((m.pos.shape[0], 3))
        self.lib.orc_get_normals(self.h, mesh, _d(fn), _d(vn))
        return fn, vn

    def render(self, want_ids=True, y0=0, y1=None, ystep=1, spp_grid=1, nthreads=0):
        W, H = int(self.desc.camera["width"]), int(self.desc.camera["height"])
        y1 = H if y1 is None else y1
        rgb = np.zeros((H, W, 3))
        ids = np.full((H, W), -2, np.int32) if want_ids else None
        t = np.zeros((H, W)) if want_ids else None
        secs = self.lib.orc_render(self.h, _d(rgb), _i(ids), _d(t), y0, y1, ystep, spp_grid, nthreads)
        counts = (C.c_longlong * 3)()
        self.lib.orc_ray_counts(self.h, counts)
        return dict(rgb=rgb, ids=ids, t=t, seconds=secs,
                    rays=dict(primary=counts[0], shadow=counts[1], reflect=counts[2]))

    def trace_rays(self, origins, dirs):
        o = np.ascontiguousarray(origins, np.float64); d = np.ascontiguousarray(dirs, np.float64)
        n = o.shape[0]
        rgb = np.zeros((n, 3)); ids = np.zeros(n, np.int32); t = np.zeros(n)
        self.lib.orc_trace_rays(self.h, n, _d(o), _d(d), _d(rgb), _i(ids), _d(t))
        return rgb, ids, t

    def lighting(self, point, normal, view, mat12):
        out = np.zeros(3)
        self.lib.orc_lighting(self.h, _v(point), _v(normal), _v(view), _d(np.ascontiguousarray(mat12, np.float64)), _d(out))
        return out

    def texture(self, tri, alpha, beta, gamma):
        out = np.zeros(3)
        self.lib.orc_texture(self.h, int(tri), float(alpha), float(beta), float(gamma), _d(out))
        return out


def tri_test(p0, p1, p2, o, d, kind="reference"):
    out = np.zeros(4)
    hit = load(kind).orc_tri_test(_v(p0), _v(p1), _v(p2), _v(o), _v(d), _d(out))
    return bool(hit), out


def aabb_test(bmin, bmax, o, d, kind="reference"):
    t = C.c_double(0.0)
    hit = load(kind).orc_aabb_test(_v(bmin), _v(bmax), _v(o), _v(d), C.byref(t))
    return bool(hit), t.value


def plane_test(center, normal, o, d, kind="reference"):
    out = np.zeros(4)
    hit = load(kind).orc_plane_test(_v(center), _v(normal), _v(o), _v(d), _d(out))
    return bool(hit), out


def sphere_test(center, radius, o, d, kind="reference"):
    out = np.zeros(7)
    hit = load(kind).orc_sphere_test(_v(center), float(radius), _v(o), _v(d), _d(out))
    return bool(hit), out
