"""TEST INFRASTRUCTURE — ctypes wrapper of tests/hostsim/libhostsim.so (host build of the kernel core)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None
_dp = C.POINTER(C.c_double)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(os.path.join(_HERE, "hostsim", "libhostsim.so"))
        _lib.sim_last_error.restype = C.c_char_p
        _lib.sim_render.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        for f in (_lib.sim_tri_filter, _lib.sim_box_test):
            f.argtypes = [_dp, _dp, _dp, _dp, _dp, C.c_double] if f is _lib.sim_tri_filter else [_dp, _dp, _dp, _dp, C.c_double, C.c_double]
        _lib.sim_tri_exact.argtypes = [_dp, _dp, _dp, _dp, _dp, _dp]
        _lib.sim_free.argtypes = [C.c_void_p]
    return _lib


def flatten(host, scene_flags=0):
    """HostScene -> bytes of the flattened blob (same flattener the CUDA library uses).
    scene_flags: capi.RT_SCENE_REFERENCE_TOPOLOGY keeps every ray on the reference's tree."""
    L = lib()
    blob = C.POINTER(C.c_ubyte)()
    n = C.c_size_t()
    st = L.sim_flatten(C.byref(host.data), C.byref(host.bvh), host.objects, host.num_objects,
                       C.c_uint(scene_flags), C.byref(blob), C.byref(n))
    if st != 0:
        raise RuntimeError("flatten status %d: %s" % (st, L.sim_last_error().decode()))
    out = C.string_at(blob, n.value)
    L.sim_free(blob)
    return out


def render(host, blob=None, flags=0, spp_grid=1, tile=(0, 1, 1), want_ids=True, camera=None, cost=False, dyn=(0, 0),
           wavefront=False):
    """wavefront: the frame through the wavefront schedule (rt_wavefront.cuh stage functions) instead of the
    per-lane state machine of the persistent kernel."""
    from my_raytracer_b200 import capi
    L = lib()
    blob = blob if blob is not None else flatten(host)
    buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
    cam = camera or host.camera
    W, H = cam.width, cam.height
    rgba = np.zeros((H, W, 4), np.uint8)
    px = np.zeros((H, W, 4))
    ids = np.full((H, W), -2, np.int32)
    t = np.zeros((H, W))
    fo = capi.FrameOut(rgba.ctypes.data, px.ctypes.data, ids.ctypes.data if want_ids else None,
                       t.ctypes.data if want_ids else None)
    dummy = C.c_ulonglong(0)
    p = host.params(spp_grid=spp_grid, flags=flags, tile=tile,
                    dyn=(dyn[0], dyn[1], C.addressof(dummy) if dyn[0] else None))
    cnt = (C.c_ulonglong * 16)()
    cost_n = np.zeros((H, W), np.uint32); cost_t = np.zeros((H, W), np.uint32)
    lp, lc = host.lights_pos, host.lights_color
    if wavefront:
        L.sim_render_wavefront.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                           C.c_void_p, C.c_void_p]
        st = L.sim_render_wavefront(buf, len(blob), C.byref(cam), lp.ctypes.data, lc.ctypes.data, lp.shape[0], C.byref(p),
                                    C.byref(fo), cnt)
    else:
        st = L.sim_render(buf, len(blob), C.byref(cam), lp.ctypes.data, lc.ctypes.data, lp.shape[0], C.byref(p),
                          C.byref(fo), cnt, 0, cost_n.ctypes.data if cost else None, cost_t.ctypes.data if cost else None)
    if st != 0:
        raise RuntimeError("sim_render status %d: %s" % (st, L.sim_last_error().decode()))
    names = ("rays_primary", "rays_shadow", "rays_reflect", "box_tests", "tri_filter", "tri_exact", "object_tests",
             "shaded_hits", "texel_fetches", "nodes_visited", "shadow_trivial", "shadow_occluded", "nodes4_visited",
             "nodes8_visited")
    out = dict(rgba8=rgba, pixels=px, hit_id=ids, hit_t=t, counters=dict(zip(names, [int(c) for c in cnt])))
    if cost:
        out["cost_nodes"], out["cost_tris"] = cost_n, cost_t
    return out


def _v(x):
    return (C.c_double * 3)(*[float(a) for a in x])


def tri_filter(p0, p1, p2, o, d, thi):
    return bool(lib().sim_tri_filter(_v(p0), _v(p1), _v(p2), _v(o), _v(d), float(thi)))


def tri_exact(p0, p1, p2, o, d):
    out = np.zeros(4)
    hit = lib().sim_tri_exact(_v(p0), _v(p1), _v(p2), _v(o), _v(d), out.ctypes.data_as(_dp))
    return bool(hit), out


def pow_specular(x, y):
    L = lib()
    L.sim_pow_specular.restype = C.c_double
    L.sim_pow_specular.argtypes = [C.c_double, C.c_double]
    return L.sim_pow_specular(float(x), float(y))


def sphere_filter(center, radius, o, d, thi):
    L = lib()
    L.sim_sphere_filter.argtypes = [_dp, C.c_double, _dp, _dp, C.c_double]
    return bool(L.sim_sphere_filter(_v(center), float(radius), _v(o), _v(d), float(thi)))


def box_test(bmin, bmax, o, d, thi, extent_scale=1.0):
    """extent_scale > 1 emulates a box that is small against the scene's box_extent (larger slack)."""
    return bool(lib().sim_box_test(_v(bmin), _v(bmax), _v(o), _v(d), float(thi), float(extent_scale)))
