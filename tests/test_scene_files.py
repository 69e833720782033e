"""Caller side of the boundary: .sce / OBJ / texture ingestion into the reference's Data SoA
and the image writer — host code of the C-ABI library, checked against the in-memory scene builder."""
import ctypes as C
import os
import struct
import zlib

import numpy as np
import pytest


def _file_scene(rt, lib, desc, tmp_path, fmt="png", flags=0):
    rt.scene_files.quantize_textures(desc)
    path = rt.write_scene(desc, str(tmp_path), texture_format=fmt)
    return rt.FileScene(path, lib, flags)


@pytest.mark.parametrize("scene,kw,fmt", [("mixed_small", dict(width=40, height=30), "png"),
                                          ("textured", dict(width=32, height=18, tex_size=32, slices=16, stacks=8), "ppm"),
                                          ("spheres", dict(width=20, height=20), "png")])
def test_scene_file_gives_the_same_data_soa(rt, lib, tmp_path, scene, kw, fmt):
    """write .sce/.obj/.mtl/texture -> rt_scene_file_load == HostScene(desc): every Data and BVH array, the
    objects, the camera and the lights, bit for bit (build_Data offsets, compute_normals, initSoA)."""
    desc = rt.scenes.REGISTRY[scene](**kw)
    fs = _file_scene(rt, lib, desc, tmp_path, fmt)
    host = rt.HostScene(desc, lib)
    a, d = host.arrays, host.data
    for f in ("tMeshCount_", "tVertexCount_", "tVertexIdxCount_", "tTextCoordCount_", "tTextIdxCount_", "tTexelCount_"):
        assert getattr(fs.data, f) == getattr(d, f), f
    nv, nt, M = d.tVertexCount_, d.tVertexIdxCount_ // 3, d.tMeshCount_
    eq = lambda x, y: np.array_equal(x, y, equal_nan=True)
    assert eq(fs.array("vertexPos_", nv, 4), a["vertexPos"]) and eq(fs.array("vertexNormals_", nv, 4), a["vertexNormals"])
    assert eq(fs.array("normals_", nt, 4), a["normals"])
    assert eq(fs.array("vertexIdx_", 3 * nt, 1, np.int32), a["vertexIdx"])
    assert eq(fs.array("vertexMeshId_", nv, 1, np.int32), a["vertexMeshId"])
    textured = a["texW"] != -1
    if textured.any():
        assert eq(fs.array("textureCoordinatesU_", d.tTextCoordCount_), a["U"]) and eq(fs.array("textureCoordinatesV_", d.tTextCoordCount_), a["V"])
        assert eq(fs.array("meshTexels_", d.tTexelCount_, 4)[:, :3], a["texels"][:, :3])
        vid = a["vertexIdx"].reshape(-1, 3)[:, 0]
        rows = textured[a["vertexMeshId"][vid]]                      # texture indices matter for textured meshes only
        assert eq(fs.array("textureIdx_", 3 * nt, 1, np.int32).reshape(-1, 3)[rows], a["textureIdx"].reshape(-1, 3)[rows])
    for f, k in (("meshTexWidth_", "texW"), ("meshTexHeight_", "texH"), ("meshDrawMode_", "drawMode"), ("firstVertex_", "firstVertex"),
                 ("vertexCount_", "vertexCount"), ("firstVertexIdx_", "firstVertexIdx"), ("vertexIdxCount_", "vertexIdxCount")):
        assert eq(fs.array(f, M, 1, np.int32), a[k]), f
    for f, k in (("materialAmbient_", "ka"), ("materialDiffuse_", "kd"), ("materialSpecular_", "ks")):
        assert eq(fs.array(f, M, 4), a[k])
    assert eq(fs.array("materialShininess_", M), a["shin"]) and eq(fs.array("materialMirror_", M), a["mirror"])
    if nt:
        n = host.nodes_used
        assert fs.nodes_used == n
        b = host.bvh_arrays
        assert eq(fs.array("bb_min_", n, 4), b["bbmin"][:n]) and eq(fs.array("bb_max_", n, 4), b["bbmax"][:n])
        assert eq(fs.array("leftChildIdx_", n, 1, np.int32), b["left"][:n]) and eq(fs.array("triCount_", n, 1, np.int32), b["count"][:n])
    assert fs.num_objects == host.num_objects
    assert bytes(C.string_at(fs.objects, C.sizeof(rt.capi.Object) * fs.num_objects)) == \
        bytes(C.string_at(host.objects, C.sizeof(rt.capi.Object) * host.num_objects))
    assert bytes(fs.camera) == bytes(host.camera)
    assert eq(fs.lights_pos, host.lights_pos) and eq(fs.lights_color, host.lights_color)
    assert fs.view.max_depth == desc.max_depth and fs.view.ambience.x == desc.ambience[0]


def test_scene_file_renders_like_the_in_memory_scene(rt, lib, sim, tmp_path):
    """Kernel core (host build) on the loaded scene == on the in-memory scene."""
    desc = rt.scenes.mixed_small(width=48, height=36)
    fs = _file_scene(rt, lib, desc, tmp_path)
    host = rt.HostScene(desc, lib)
    a, b = sim.render(host), sim.render(fs)
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(a[k], b[k]), k


def test_scene_file_errors(rt, lib, tmp_path):
    bad = tmp_path / "bad.sce"
    bad.write_text("camera 0 0 5 0 0 0 0 1 0 45 8 8\nteapot 1 2 3\n")
    with pytest.raises(rt.capi.RtError, match="unknown token"):
        rt.FileScene(str(bad), lib)
    bad.write_text("depth 3\nmesh missing.obj FLAT 1 1 1 1 1 1 0 0 0 1 0\ncamera 0 0 5 0 0 0 0 1 0 45 8 8\n")
    with pytest.raises(rt.capi.RtError, match="cannot open"):
        rt.FileScene(str(bad), lib)
    bad.write_text("depth 3\n# no camera\n")
    with pytest.raises(rt.capi.RtError, match="no camera"):
        rt.FileScene(str(bad), lib)
    with pytest.raises(rt.capi.RtError):
        rt.FileScene(str(tmp_path / "nope.sce"), lib)


def _decode_png_rgba(data):
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, t = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        assert zlib.crc32(t + body) & 0xffffffff == struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0]
        if t == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert (depth, ctype) == (8, 6)
        elif t == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 4 * w)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 4)


def test_image_writer_and_readers(rt, lib, tmp_path):
    rng = np.random.default_rng(3)
    frame = rng.integers(0, 256, (37, 53, 4), dtype=np.uint8)          # renderer order: row 0 = bottom
    frame[..., 3] = 255
    png, ppm = str(tmp_path / "f.png"), str(tmp_path / "f.ppm")
    rt.write_image(png, frame, lib)
    rt.write_image(ppm, frame, lib)
    # an independent decoder (zlib) sees the upright image
    assert np.array_equal(_decode_png_rgba(open(png, "rb").read()), frame[::-1])
    raw = open(ppm, "rb").read()
    assert raw.startswith(b"P6\n53 37\n255\n") and np.array_equal(np.frombuffer(raw[len(b"P6\n53 37\n255\n"):], np.uint8).reshape(37, 53, 3), frame[::-1, :, :3])
    # the library's own readers
    assert np.array_equal(rt.read_image(png, lib), frame[::-1]) and np.array_equal(rt.read_image(ppm, lib), frame[::-1])
    # PNGs from an independent encoder: real deflate (dynamic and fixed Huffman, stored), all five row filters
    smooth = (np.add.outer(np.arange(64), np.arange(48))[:, :, None] * np.array([1, 2, 3]) % 256).astype(np.uint8)
    noise = rng.integers(0, 256, (31, 17, 3), dtype=np.uint8)
    for img, level in ((smooth, 9), (smooth, 1), (noise, 6), (noise, 0)):
        p = str(tmp_path / "t.png")
        open(p, "wb").write(rt.scene_files.png_bytes(img, filters=[0, 1, 2, 3, 4], level=level))
        got = rt.read_image(p, lib)
        assert np.array_equal(got[..., :3], img) and (got[..., 3] == 255).all()
    open(str(tmp_path / "c.png"), "wb").write(b"\x89PNG\r\n\x1a\n" + b"\0" * 30)
    with pytest.raises(rt.capi.RtError):
        rt.read_image(str(tmp_path / "c.png"), lib)
