"""Scene files on the caller side of the boundary.

* ``write_scene``  — a SceneDesc as the course framework's text scene (``.sce``) + one OBJ (+ MTL + texture image)
  per mesh: the input format of the reference's host program (format notes: csrc/rt_scene_io.cpp).
* ``FileScene``    — ``rt_scene_file_load`` result with the attributes DeviceScene needs (data, bvh, objects,
  camera, lights, params): parsing, build_Data, normals and the BVH build all happen in the C ABI library.
* ``write_image`` / ``read_image`` — ``rt_image_write`` / ``rt_image_read``.
"""
from __future__ import annotations

import ctypes as C
import os
import struct
import zlib

import numpy as np

from . import capi
from .capi import RenderParams, SceneFileView, Vec4, check


def _fmt(vals):
    return " ".join(repr(float(v)) for v in vals)


def _mat(m):
    """11 numbers + the optional shadowable flag (written only when it is not the default)."""
    return _fmt(m[:11]) + ("" if int(m[11]) else " 0")


def png_bytes(rgb_top_first: np.ndarray, filters=None, level=9) -> bytes:
    """8-bit RGB PNG with real deflate compression and (optionally) a PNG filter type per row — an encoder
    independent of the C one, used to exercise the C decoder (dynamic Huffman blocks, all five filters)."""
    h, w, _ = rgb_top_first.shape
    img = rgb_top_first.astype(np.int32)
    raw = bytearray()
    prev = np.zeros((w, 3), np.int32)
    for y in range(h):
        f = 0 if filters is None else filters[y % len(filters)]
        cur = img[y]
        a = np.vstack([np.zeros((1, 3), np.int32), cur[:-1]])
        c = np.vstack([np.zeros((1, 3), np.int32), prev[:-1]])
        if f == 0:
            out = cur
        elif f == 1:
            out = cur - a
        elif f == 2:
            out = cur - prev
        elif f == 3:
            out = cur - (a + prev) // 2
        else:
            p = a + prev - c
            pa, pb, pc = np.abs(p - a), np.abs(p - prev), np.abs(p - c)
            pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, prev, c))
            out = cur - pred
        raw.append(f)
        raw += (out & 255).astype(np.uint8).tobytes()
        prev = cur

    def chunk(t, body):
        return struct.pack(">I", len(body)) + t + body + struct.pack(">I", zlib.crc32(t + body) & 0xffffffff)
    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
            chunk(b"IDAT", zlib.compress(bytes(raw), level)) + chunk(b"IEND", b""))


def quantize_textures(desc):
    """Textures as an 8-bit image file can carry them: k / 255 (so that file and in-memory scenes are equal)."""
    for m in desc.meshes:
        if m.tex is not None:
            m.tex = np.round(np.clip(m.tex, 0.0, 1.0) * 255.0) / 255.0
    return desc


def write_scene(desc, directory: str, name: str = "scene", texture_format: str = "png") -> str:
    """Returns the path of the .sce file.  Textures must already be k/255-valued (quantize_textures)."""
    os.makedirs(directory, exist_ok=True)
    lines = [f"# {desc.name}", f"depth {int(desc.max_depth)}"]
    c = desc.camera
    lines.append("camera " + _fmt(c["eye"]) + " " + _fmt(c["center"]) + " " + _fmt(c["up"]) + " " + repr(float(c["fovy"])) +
                 f" {int(c['width'])} {int(c['height'])}")
    lines.append("background " + _fmt(desc.background))
    lines.append("ambience " + _fmt(desc.ambience))
    for pos, col in desc.lights:
        lines.append("light " + _fmt(pos) + " " + _fmt(col))
    for o in desc.objects:
        mat = _mat(o[3])
        if o[0] == "sphere":
            lines.append("sphere " + _fmt(o[1]) + " " + repr(float(o[2])) + " " + mat)
        else:
            lines.append("plane " + _fmt(o[1]) + " " + _fmt(o[2]) + " " + mat)
    for k, m in enumerate(desc.meshes):
        stem = f"{name}_mesh{k}"
        with open(os.path.join(directory, stem + ".obj"), "w") as f:
            f.write(f"# mesh {k} of {desc.name}\n")
            if m.tex is not None:
                f.write(f"mtllib {stem}.mtl\n")
            for p in m.pos:
                f.write("v " + _fmt(p) + "\n")
            if m.u is not None:
                for u, v in zip(m.u, m.v):
                    f.write("vt " + _fmt((u, v)) + "\n")
            if m.tex is not None:
                f.write("usemtl tex\n")
            for t in range(m.idx.shape[0]):
                if m.uvidx is not None:
                    f.write("f " + " ".join(f"{int(a) + 1}/{int(b) + 1}" for a, b in zip(m.idx[t], m.uvidx[t])) + "\n")
                else:
                    f.write("f " + " ".join(str(int(a) + 1) for a in m.idx[t]) + "\n")
        if m.tex is not None:
            img = np.round(m.tex * 255.0).astype(np.uint8)
            tex_name = f"{stem}.{texture_format}"
            with open(os.path.join(directory, stem + ".mtl"), "w") as f:
                f.write(f"newmtl tex\nmap_Kd {tex_name}\n")
            with open(os.path.join(directory, tex_name), "wb") as f:
                if texture_format == "png":
                    f.write(png_bytes(img, filters=[0, 1, 2, 3, 4]))
                else:
                    f.write(b"P6\n%d %d\n255\n" % (img.shape[1], img.shape[0]) + img.tobytes())
        lines.append(f"mesh {stem}.obj {'PHONG' if m.draw_mode == 1 else 'FLAT'} " + _mat(m.material))
    path = os.path.join(directory, name + ".sce")
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return path


class FileScene:
    """A scene loaded by the library from a .sce file; quacks like HostScene for DeviceScene and the renderers."""

    def __init__(self, path: str, lib=None, load_flags: int = 0):
        self.lib = lib or capi.load_library()
        h = C.c_void_p()
        check(self.lib.rt_scene_file_load(path.encode(), int(load_flags), C.byref(h)), self.lib)
        self.handle = h
        v = SceneFileView()
        check(self.lib.rt_scene_file_get(h, C.byref(v)), self.lib)
        self.view = v
        self.data, self.bvh, self.nodes_used = v.data, v.bvh, v.nodes_used
        self.objects, self.num_objects = v.objects, v.num_objects
        self.camera = v.camera
        n = v.num_lights
        self.lights_pos = np.ctypeslib.as_array(C.cast(v.lights_pos, C.POINTER(C.c_double)), (n, 4)).copy() if n else np.zeros((0, 4))
        self.lights_color = np.ctypeslib.as_array(C.cast(v.lights_color, C.POINTER(C.c_double)), (n, 4)).copy() if n else np.zeros((0, 4))
        self.num_triangles = v.data.tVertexIdxCount_ // 3

    def array(self, field: str, count: int, width: int = 1, dtype=np.float64):
        """numpy copy of one Data / BVH array (tests)."""
        holder = self.data if hasattr(self.data, field) else self.bvh
        ptr = getattr(holder, field)
        if count == 0:
            return np.zeros((0, width) if width > 1 else (0,), dtype)
        ct = {np.float64: C.c_double, np.int32: C.c_int, np.uint8: C.c_ubyte, np.uint64: C.c_size_t}[dtype]
        a = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), (count * width,)).copy()
        return a.reshape(count, width) if width > 1 else a

    def params(self, spp_grid=1, flags=0, tile=(0, 1, 1), max_depth=None, dyn=(0, 0, None)) -> RenderParams:
        v = self.view
        return RenderParams(v.background, v.ambience, v.max_depth if max_depth is None else max_depth, spp_grid,
                            tile[0], tile[1], tile[2], flags, int(dyn[0]), int(dyn[1]), dyn[2])

    def close(self):
        if self.handle:
            self.lib.rt_scene_file_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def write_image(path: str, rgba8: np.ndarray, lib=None):
    """rgba8: (H, W, 4) uint8 frame as the renderer returns it (row 0 = bottom); the file is upright."""
    lib = lib or capi.load_library()
    a = np.ascontiguousarray(rgba8, np.uint8)
    check(lib.rt_image_write(path.encode(), a.ctypes.data, a.shape[1], a.shape[0]), lib)


def read_image(path: str, lib=None) -> np.ndarray:
    """(H, W, 4) uint8, row 0 = top row of the file."""
    lib = lib or capi.load_library()
    w, h, p = C.c_int(), C.c_int(), C.c_void_p()
    check(lib.rt_image_read(path.encode(), C.byref(w), C.byref(h), C.byref(p)), lib)
    out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_ubyte)), (h.value, w.value, 4)).copy()
    lib.rt_image_free(p)
    return out
