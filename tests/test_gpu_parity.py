"""GPU parity tests proper: the CUDA path, called through the C ABI, against the oracle and the golden
fixtures.  Bit-exact for primitive ids and ray parameters; 8-bit pixels within 1 LSB (north star:
>= 99.9 % of pixels; observed: identical)."""
import ast
import ctypes as C
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
SCENES = ["mixed_small", "spheres", "office", "rings", "molecule", "textured"]


def _check(out, ids, t, rgb8, rgb=None):
    assert np.array_equal(out["hit_id"], ids)
    assert np.array_equal(out["hit_t"], t)
    d = np.abs(out["rgba8"][..., :3].astype(int) - rgb8.astype(int))
    assert d.max() <= 1 and (d == 0).mean() >= 0.999
    if rgb is not None:
        assert np.abs(out["pixels"][..., :3] - rgb).max() < 1e-6      # 1e-6: FP32 texel storage


@pytest.mark.parametrize("name", SCENES)
def test_golden_scenes(rt, lib, gpu, name):
    g = np.load(os.path.join(GOLDEN, f"scene_{name}.npz"))
    host = rt.HostScene(rt.scenes.REGISTRY[name](**ast.literal_eval(str(g["kwargs"]))), lib)
    dev = rt.DeviceScene(host)
    out = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=gpu.RT_FLAG_COUNTERS))
    _check(out, g["ids"], g["t"], g["rgb8"], g["rgb"])
    c = out["counters"]
    assert [c["rays_primary"], c["rays_shadow"], c["rays_reflect"]] == list(g["rays"])
    plain = dev.render(host.camera, host.lights_pos, host.lights_color, host.params())
    exact = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=gpu.RT_FLAG_EXACT_ONLY))
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(plain[k], out[k]) and np.array_equal(exact[k], out[k])


@pytest.mark.parametrize("name,kw", [
    ("mixed_small", dict(width=333, height=201, seed=9)),          # ragged: not multiples of the 8x4 tile
    ("office", dict(width=320, height=180)),
    ("rings", dict(width=210, height=150, nu=96, nv=24)),
    ("molecule", dict(width=150, height=150, n_spheres=200)),
    ("textured", dict(width=256, height=144, tex_size=256, slices=64, stacks=32)),
    ("scatter", dict(width=256, height=144, n_objects=300)),
    ("objtree", dict(width=192, height=144)),                      # tree over analytic spheres, list-order ties
    ("objtree", dict(width=192, height=144, camera_inside=False)),
])
def test_live_oracle_scenes(rt, lib, gpu, orc, oracle_kind, name, kw):
    desc = rt.scenes.REGISTRY[name](**kw)
    host = rt.HostScene(desc, lib)
    dev = rt.DeviceScene(host)
    out = dev.render(host.camera, host.lights_pos, host.lights_color, host.params())
    ref = orc.OracleScene(desc, oracle_kind).render()
    _check(out, ref["ids"], ref["t"], rt.quantize_rgb8(ref["rgb"]), ref["rgb"])


@pytest.mark.parametrize("name,kw", [
    ("office", dict(width=480, height=270)),
    ("scatter", dict(width=256, height=144, n_objects=300)),      # has a column of d.x == 0 rays (NaN quirk path)
    ("textured", dict(width=256, height=144, tex_size=256, slices=64, stacks=32)),
    ("mixed_small", dict(width=333, height=201, seed=9)),
    ("molecule", dict(width=200, height=200, n_spheres=300)),
    ("objtree", dict(width=160, height=120, camera_inside=False)),
])
def test_frame_does_not_depend_on_the_tree(rt, lib, gpu, name, kw):
    """Library-built SAH traversal tree (default) vs every ray on the reference's median-split topology:
    bit-identical ids, distances and FP64 pixels, fewer node visits."""
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    sah = rt.DeviceScene(host)
    ref = rt.DeviceScene(host, scene_flags=gpu.RT_SCENE_REFERENCE_TOPOLOGY)
    assert ref.info()["topology"] == 1 and sah.info()["topology"] == (0 if host.num_triangles > 2 else 1)
    p = host.params(flags=gpu.RT_FLAG_COUNTERS)
    a = sah.render(host.camera, host.lights_pos, host.lights_color, p)
    b = ref.render(host.camera, host.lights_pos, host.lights_color, p)
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(a[k], b[k]), k
    for k in ("rays_primary", "rays_shadow", "rays_reflect", "shaded_hits"):
        assert a["counters"][k] == b["counters"][k]
    if name in ("molecule", "objtree"):
        assert a["counters"]["object_tests"] < b["counters"]["object_tests"]
    else:
        assert a["counters"]["nodes_visited"] < b["counters"]["nodes_visited"]


@pytest.mark.parametrize("name,kw", [
    ("office", dict(width=640, height=360)),
    ("scatter", dict(width=256, height=144, n_objects=300)),      # d.x == 0 rays stay on the reference tree
    ("textured", dict(width=256, height=144, tex_size=256, slices=64, stacks=32)),
    ("rings", dict(width=200, height=150)),
    ("molecule", dict(width=200, height=200, n_spheres=300)),     # object tree only: its 4-wide twin
    ("objtree", dict(width=192, height=144)),                      # object tree + triangles, list-order ties
])
def test_wide_tree_gives_the_same_frame_in_fewer_steps(rt, lib, gpu, name, kw):
    """The 4-wide twin of the traversal tree (render_kernel<K_WIDE>, default) against the binary tree it is
    collapsed from (RT_FLAG_BINARY_NODES), plain and instrumented kernels: bit-identical ids, distances, FP64
    pixels; the wide walk needs fewer node visits; and the FP32 filter stays a pure filter on it."""
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    dev = rt.DeviceScene(host)
    assert dev.info()["num_nodes_wide"] > 0
    args = (host.camera, host.lights_pos, host.lights_color)
    wide = dev.render(*args, host.params())
    binary = dev.render(*args, host.params(flags=gpu.RT_FLAG_BINARY_NODES))
    wide_c = dev.render(*args, host.params(flags=gpu.RT_FLAG_COUNTERS))
    binary_c = dev.render(*args, host.params(flags=gpu.RT_FLAG_COUNTERS | gpu.RT_FLAG_BINARY_NODES))
    exact = dev.render(*args, host.params(flags=gpu.RT_FLAG_EXACT_ONLY))
    for other in (binary, wide_c, binary_c, exact):
        for k in ("hit_id", "hit_t", "rgba8", "pixels"):
            assert np.array_equal(wide[k], other[k]), k
    assert wide_c["counters"]["nodes4_visited"] > 0 and binary_c["counters"]["nodes4_visited"] == 0
    ratio = 1.0 if name == "objtree" else 0.75           # objtree: a dozen spheres and a few triangles, tiny trees
    assert wide_c["counters"]["nodes_visited"] < ratio * binary_c["counters"]["nodes_visited"]
    for k in ("rays_primary", "rays_shadow", "rays_reflect", "shaded_hits", "shadow_occluded"):
        assert wide_c["counters"][k] == binary_c["counters"][k], k


def test_supersampling_grid(rt, lib, gpu):
    g = np.load(os.path.join(GOLDEN, "scene_mixed_small_spp2.npz"))
    host = rt.HostScene(rt.scenes.mixed_small(width=64, height=48), lib)
    dev = rt.DeviceScene(host)
    out = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(spp_grid=2), want=("rgba8", "pixels"))
    assert np.abs(out["pixels"][..., :3] - g["rgb"]).max() < 1e-6
    assert np.abs(out["rgba8"][..., :3].astype(int) - g["rgb8"].astype(int)).max() <= 1


def test_edge_scenes(rt, lib, gpu):
    cam = dict(eye=(0, 0, 5), center=(0, 0, 0), up=(0, 1, 0), fovy=40, width=40, height=24)
    empty = rt.HostScene(rt.SceneDesc("empty", [], [], [], cam, background=(0.25, 0.5, 0.75)), lib)
    out = rt.DeviceScene(empty).render(empty.camera, empty.lights_pos, empty.lights_color, empty.params())
    assert (out["hit_id"] == -1).all() and (out["hit_t"] == -1).all()
    assert (out["rgba8"].reshape(-1, 4) == np.array([64, 128, 191, 255], np.uint8)).all()
    pos = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0.0]])
    for ntri in (1, 2):
        d = rt.SceneDesc("tiny", [rt.MeshDesc(pos, np.array([[0, 1, 2], [0, 2, 3]], np.int32)[:ntri],
                                              rt.material((1, 1, 1), (1, 1, 1), (0, 0, 0), 1.0))], [],
                         [((0, 0, 5), (1, 1, 1))], cam, ambience=(0.5, 0.5, 0.5))
        h = rt.HostScene(d, lib)
        out = rt.DeviceScene(h).render(h.camera, h.lights_pos, h.lights_color, h.params())
        assert set(np.unique(out["hit_id"])) <= set(range(-1, ntri)) and (out["hit_id"] >= 0).any()


def test_invalid_arguments(rt, lib, gpu):
    host = rt.HostScene(rt.scenes.mixed_small(width=16, height=8), lib)
    dev = rt.DeviceScene(host)
    fo = gpu.FrameOut()
    p = host.params()
    assert lib.rt_render(None, C.byref(host.camera), None, None, 0, C.byref(p), C.byref(fo)) == gpu.RT_ERR_INVALID_ARG
    assert lib.rt_render(dev.handle, None, None, None, 0, C.byref(p), C.byref(fo)) == gpu.RT_ERR_INVALID_ARG
    bad = host.params(tile=(3, 2, 1))
    assert lib.rt_render(dev.handle, C.byref(host.camera), None, None, 0, C.byref(bad), C.byref(fo)) == gpu.RT_ERR_INVALID_ARG
    assert b"tile_rank" in lib.rt_last_error()
    with pytest.raises(gpu.RtError):
        dev.render(host.camera, host.lights_pos, host.lights_color, host.params(spp_grid=2), want=("hit_id",))


def test_blob_roundtrip_and_tile_sharding(rt, lib, gpu):
    """Multi-GPU building blocks on one GPU: a scene adopted from an exported blob renders identically,
    and the union of tile-sharded renders (any world/chunk) is the full image, bit for bit."""
    host = rt.HostScene(rt.scenes.office(width=322, height=181), lib)
    dev = rt.DeviceScene(host)
    full = dev.render(host.camera, host.lights_pos, host.lights_color, host.params())
    clone = rt.DeviceScene(lib=lib, blob=dev.export_blob())
    again = clone.render(host.camera, host.lights_pos, host.lights_color, host.params())
    for k in full:
        assert np.array_equal(full[k], again[k])
    ptr, nbytes = dev.blob_device()
    clone2 = rt.DeviceScene(lib=lib, blob_device_ptr=ptr, blob_bytes=nbytes)
    for world, chunk in ((2, 1), (4, 8), (8, 3)):
        acc = np.zeros_like(full["rgba8"]); seen = np.zeros(full["hit_id"].shape, int)
        for r in range(world):
            part = clone2.render(host.camera, host.lights_pos, host.lights_color, host.params(tile=(r, world, chunk)))
            mask = part["hit_id"] != -1 if False else (part["rgba8"][..., 3] == 255)
            seen += mask
            acc[mask] = part["rgba8"][mask]
        assert (seen == 1).all() and np.array_equal(acc, full["rgba8"])


def test_dynamic_tail_queue_and_shared_memory(rt, lib, gpu):
    """Cross-GPU dynamic tail on one GPU: several 'ranks' render one after the other into ONE shared device
    frame buffer (rt_shared_alloc, exported for IPC) and claim the last tiles from ONE shared counter.  Whoever
    runs first drains the pool; the union must be the single-rank frame bit for bit, every pixel written once."""
    import torch
    host = rt.HostScene(rt.scenes.office(width=322, height=181), lib)
    dev = rt.DeviceScene(host)
    W, H = 322, 181
    full = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(), want=("rgba8",))["rgba8"]
    fb = C.c_void_p(); ctr = C.c_void_p()
    gpu.check(lib.rt_shared_alloc(W * H * 4, C.byref(fb)), lib)
    gpu.check(lib.rt_shared_alloc(64, C.byref(ctr)), lib)
    handle = (C.c_ubyte * 64)()
    gpu.check(lib.rt_ipc_export(fb, handle), lib)                  # what rank 0 would broadcast
    assert any(handle)
    d_lpos = torch.tensor(host.lights_pos, dtype=torch.float64, device="cuda")
    d_lcol = torch.tensor(host.lights_color, dtype=torch.float64, device="cuda")
    ntiles = ((W + 7) // 8) * ((H + 3) // 4)
    from my_raytracer_b200 import dist as rdist
    fbt = rdist.wrap_device_ptr(fb.value, W * H * 4)                # torch view of the shared allocation
    for slot, (world, chunk, dyn_tiles, dyn_chunk) in enumerate(((2, 16, ntiles // 5, 8), (4, 3, ntiles, 1), (3, 5, 9, 64))):
        fbt.zero_()
        torch.cuda.synchronize()
        fo = gpu.FrameOut(fb.value, None, None, None)
        counter = ctr.value + 8 * slot                              # a fresh zero counter per frame
        order = list(range(world))[::-1] if slot % 2 else list(range(world))
        # narrower persistent grids (frames in flight side by side) render the same frame
        gpu.check(lib.rt_scene_set_launch_width(dev.handle, slot), lib)      # 0 = all slots, 1, 2 CTAs per SM
        for r in order:
            p = host.params(tile=(r, world, chunk), dyn=(dyn_tiles, dyn_chunk, counter), flags=gpu.RT_FLAG_NO_CLEAR)
            gpu.check(lib.rt_render_device(dev.handle, C.byref(host.camera), d_lpos.data_ptr(), d_lcol.data_ptr(),
                                           d_lpos.shape[0], C.byref(p), C.byref(fo), None), lib)
        torch.cuda.synchronize()
        got = fbt.cpu().numpy().reshape(H, W, 4)
        assert np.array_equal(got, full)
    gpu.check(lib.rt_scene_set_launch_width(dev.handle, 0), lib)
    assert lib.rt_scene_set_launch_width(dev.handle, -1) == gpu.RT_ERR_INVALID_ARG
    # invalid dynamic-queue arguments are rejected
    bad = host.params(dyn=(10, 4, None))
    fo = gpu.FrameOut(fb.value, None, None, None)
    assert lib.rt_render_device(dev.handle, C.byref(host.camera), d_lpos.data_ptr(), d_lcol.data_ptr(), d_lpos.shape[0],
                                C.byref(bad), C.byref(fo), None) == gpu.RT_ERR_INVALID_ARG
    del fbt
    gpu.check(lib.rt_shared_free(fb), lib)
    gpu.check(lib.rt_shared_free(ctr), lib)


def test_gpu_bvh_build_is_bit_identical_to_the_host_build(rt, lib, gpu):
    """rt_bvh_build_median_device == rt_bvh_build_median (== BVH::initSoA, pinned against the reference sources in
    tests/test_host.py): the five node arrays, the node numbering and the permuted triangle rows, on scenes with
    duplicated centroids (unsplittable leaves), tiny meshes and the Office."""
    dup = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0.0]])
    quad = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    cam = dict(eye=(0, 0, 5), center=(0, 0, 0), up=(0, 1, 0), fovy=40, width=8, height=8)
    m = rt.material((1, 1, 1), (1, 1, 1), (0, 0, 0), 1.0)
    descs = [rt.scenes.office(width=16, height=9), rt.scenes.mixed_small(width=16, height=9),
             rt.scenes.rings(width=16, height=9, nu=64, nv=16), rt.scenes.scatter(width=16, height=9, n_objects=30),
             rt.SceneDesc("dups", [rt.MeshDesc(dup, np.concatenate([quad] * 9, 0), m)], [], [((0, 0, 5), (1, 1, 1))], cam),
             rt.SceneDesc("one", [rt.MeshDesc(dup, quad[:1], m)], [], [((0, 0, 5), (1, 1, 1))], cam),
             rt.SceneDesc("three", [rt.MeshDesc(dup, np.concatenate([quad, quad[:1]], 0), m)], [], [((0, 0, 5), (1, 1, 1))], cam)]
    for d in descs:
        a = rt.HostScene(d, lib)
        b = rt.HostScene(d, lib, gpu_build=True)
        assert a.nodes_used == b.nodes_used and b.bvh_device_ms is not None
        n = a.nodes_used
        for k in ("left", "first", "count", "bbmin", "bbmax"):
            assert np.array_equal(a.bvh_arrays[k][:max(n, 1)], b.bvh_arrays[k][:max(n, 1)]), (d.name, k)
        for k in ("vertexIdx", "textureIdx", "normals"):
            assert np.array_equal(a.arrays[k], b.arrays[k], equal_nan=(k == "normals")), (d.name, k)   # degenerate faces: NaN normals


@pytest.mark.parametrize("name,kw", [("rings", dict(width=32, height=32, nu=96, nv=24)),
                                     ("textured", dict(width=32, height=32, tex_size=16, slices=64, stacks=32)),
                                     ("mixed_small", dict(width=32, height=32)),
                                     ("office", dict(width=32, height=32))])
def test_gpu_normals_are_bit_identical_to_the_host(rt, lib, gpu, name, kw):
    """rt_compute_normals_device == rt_compute_normals (== Mesh::compute_normals, mymesh.cpp:103-163): face and
    angle-weighted vertex normals, every mesh of the scene, bit for bit (degenerate faces give NaN on both)."""
    desc = rt.scenes.REGISTRY[name](**kw)
    a = rt.HostScene(desc, lib)
    b = rt.HostScene(desc, lib, gpu_normals=True)
    assert np.array_equal(a.arrays["normals"], b.arrays["normals"], equal_nan=True)
    assert np.array_equal(a.arrays["vertexNormals"], b.arrays["vertexNormals"], equal_nan=True)
    assert np.isfinite(b.arrays["vertexNormals"]).mean() > 0.99
    # error behaviour: an out-of-range vertex index is reported, not dereferenced
    pos = np.zeros((3, 4)); idx = np.array([0, 1, 7], np.int32); fn = np.zeros((1, 4)); vn = np.zeros((3, 4))
    st = lib.rt_compute_normals_device(pos.ctypes.data_as(gpu.Vec4_p), 3, idx.ctypes.data_as(gpu.c_int_p), 1,
                                       fn.ctypes.data_as(gpu.Vec4_p), vn.ctypes.data_as(gpu.Vec4_p), None)
    assert st == gpu.RT_ERR_BAD_SCENE


def test_full_size_properties(rt, lib, gpu, orc, oracle_kind):
    """BASELINE config 2 at full size (Office 34 k triangles, 1920x1080): determinism across launches,
    filter == exact-only, a checksum of the tile-sharded union, and oracle parity on a row sample."""
    desc = rt.scenes.office()
    host = rt.HostScene(desc, lib)
    dev = rt.DeviceScene(host)
    a = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(), want=("rgba8", "hit_id", "hit_t"))
    b = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=gpu.RT_FLAG_EXACT_ONLY),
                   want=("rgba8", "hit_id", "hit_t"))
    for k in a:
        assert np.array_equal(a[k], b[k])
    assert (a["hit_id"] >= 0).all()                        # camera inside the closed room
    acc = np.zeros_like(a["rgba8"])
    for r in range(4):
        part = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(tile=(r, 4, 16)), want=("rgba8",))
        acc |= part["rgba8"]
    assert np.array_equal(acc, a["rgba8"])
    rows = np.arange(7, 1080, 40)
    ref = orc.OracleScene(desc, oracle_kind).render(y0=7, ystep=40)
    assert np.array_equal(a["hit_id"][rows], ref["ids"][rows]) and np.array_equal(a["hit_t"][rows], ref["t"][rows])
    d = np.abs(a["rgba8"][rows][..., :3].astype(int) - rt.quantize_rgb8(ref["rgb"][rows]).astype(int))
    assert d.max() <= 1 and (d == 0).mean() >= 0.999


def test_drop_in_launch_with_adaptive_supersampling(rt, lib, gpu, orc, oracle_kind):
    """rt_launch_compute_image == launch_compute_image_device (mytracer_gpu.cu:44-113): pass 1, then the
    4x4 re-render of interior pixels whose 4-neighbour contrast exceeds 0.02."""
    import torch
    desc = rt.scenes.mixed_small(width=96, height=64)
    desc.objects = []                                       # the reference kernel renders meshes only
    host = rt.HostScene(desc, lib)
    W, H = 96, 64
    d_px = torch.zeros(H * W * 4, dtype=torch.float64, device="cuda")
    d_tmp = torch.zeros_like(d_px)
    h_img = np.zeros((H, W, 4))
    lp = torch.tensor(host.lights_pos, device="cuda"); lc = torch.tensor(host.lights_color, device="cuda")
    cache = C.c_void_p()
    bg, amb = gpu.Vec4.of(desc.background), gpu.Vec4.of(desc.ambience)
    gpu.check(lib.rt_launch_compute_image(d_px.data_ptr(), d_tmp.data_ptr(), h_img.ctypes.data, W, H, C.byref(host.camera),
                                          lp.data_ptr(), lc.data_ptr(), lp.shape[0], C.byref(bg), C.byref(amb),
                                          desc.max_depth, C.byref(host.data), C.byref(host.bvh), C.byref(cache)), lib)
    o = orc.OracleScene(desc, oracle_kind)
    p1 = o.render(want_ids=False)["rgb"]
    p4 = o.render(want_ids=False, spp_grid=4)["rgb"]
    c = p1
    n = np.zeros((H, W))
    n[1:-1, 1:-1] = (((c[1:-1, 1:-1] - c[1:-1, 2:]) ** 2).sum(-1) + ((c[1:-1, 1:-1] - c[2:, 1:-1]) ** 2).sum(-1) +
                     ((c[1:-1, 1:-1] - c[1:-1, :-2]) ** 2).sum(-1) + ((c[1:-1, 1:-1] - c[:-2, 1:-1]) ** 2).sum(-1))
    flagged = n > 0.02
    expect = np.where(flagged[..., None], p4, p1)
    assert 0.01 < flagged.mean() < 0.9
    assert np.abs(h_img[..., :3] - expect).max() < 1e-6
    assert np.array_equal(d_tmp.cpu().numpy().reshape(H, W, 4)[..., :3][~flagged], h_img[..., :3][~flagged])
    lib.rt_scene_destroy(cache)


def test_scene_file_to_image_on_the_gpu(rt, lib, gpu, tmp_path):
    """The whole caller-side chain through the C ABI: .sce/.obj/.png files -> rt_scene_file_load with the normals
    and the BVH built on the GPU -> rt_scene_create -> rt_render -> rt_image_write; equal to the in-memory scene
    built on the host, and the PNG on disk is the upright frame."""
    desc = rt.scenes.textured(width=192, height=108, tex_size=64, slices=32, stacks=16)
    rt.scene_files.quantize_textures(desc)
    path = rt.write_scene(desc, str(tmp_path))
    fs = rt.FileScene(path, lib, gpu.RT_LOAD_DEVICE_BVH | gpu.RT_LOAD_DEVICE_NORMALS)
    host = rt.HostScene(desc, lib)
    a = rt.DeviceScene(host).render(host.camera, host.lights_pos, host.lights_color, host.params())
    b = rt.DeviceScene(fs).render(fs.camera, fs.lights_pos, fs.lights_color, fs.params())
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(a[k], b[k]), k
    out = str(tmp_path / "frame.png")
    rt.write_image(out, b["rgba8"], lib)
    assert np.array_equal(rt.read_image(out, lib), b["rgba8"][::-1])
