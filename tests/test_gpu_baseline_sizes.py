"""GPU parity at BASELINE.json's own sizes (configs 1-5), the pipelined host-buffer C ABI (rt_frame_*), and the
payload validation of adopted blobs.  The oracle renders every 8th row of each full-size frame (seconds on the host
cores); hit ids and ray parameters must be bit-exact, 8-bit pixels within 1 LSB on >= 99.9 % of the pixels
(north star; observed: identical).  Reference semantics held: mymesh.cpp:197-215 (closed barycentric interval,
t > 1e-5), mybvh.cpp:165-176 (strict '<' keeps the first hit), mytracer.cpp:589-605 (shadow test)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rows_check(rt, out, ref, rows, ids=True):
    if ids:
        assert np.array_equal(out["hit_id"][rows], ref["ids"][rows])
        assert np.array_equal(out["hit_t"][rows], ref["t"][rows])
    d = np.abs(out["rgba8"][rows][..., :3].astype(int) - rt.quantize_rgb8(ref["rgb"][rows]).astype(int))
    assert d.max() <= 1 and (d == 0).mean() >= 0.999, (d.max(), (d == 0).mean())


@pytest.mark.parametrize("name,kw", [
    ("spheres", dict(width=500, height=500)),              # config 1: o_01_spheres at the reference default resolution
    ("office", dict(width=1920, height=1080)),             # config 2: the headline
    ("molecule", dict(width=500, height=500)),             # config 3: 600 spheres, depth 3
    ("rings", dict(width=700, height=500)),                # config 3: 65 k-triangle PHONG mesh, depth 3
    ("textured", dict(width=1920, height=1080)),           # config 4: textured PHONG mesh at 1080p
])
def test_baseline_configs_at_full_size(rt, lib, gpu, orc, oracle_kind, name, kw):
    desc = rt.scenes.REGISTRY[name](**kw)
    host = rt.HostScene(desc, lib)
    dev = rt.DeviceScene(host)
    out = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=gpu.RT_FLAG_COUNTERS),
                     want=("rgba8", "hit_id", "hit_t"))
    H = kw["height"]
    rows = np.arange(3, H, 8)
    ref = orc.OracleScene(desc, oracle_kind).render(y0=3, ystep=8)
    _rows_check(rt, out, ref, rows)
    # size-independent properties on the whole frame: the FP32 filter is only a filter, and the frame is deterministic
    exact = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(flags=gpu.RT_FLAG_EXACT_ONLY),
                       want=("rgba8", "hit_id", "hit_t"))
    for k in ("rgba8", "hit_id", "hit_t"):
        assert np.array_equal(out[k], exact[k]), k
    c = out["counters"]
    assert c["rays_primary"] == kw["width"] * H


def test_big_scene_row_sample_at_4k_4spp(rt, lib, gpu, orc, oracle_kind):
    """Config 5: 10 M triangles, 3840x2160, 2x2 sub-pixel grid (offsets mytracer_gpu.cu:203-207).  The oracle renders a
    row sample (4 spp pixels, and 1 spp ids / distances); the GPU frame must match it, and the tile-sharded union over 8
    'ranks' must equal the single frame (checksum of checksums)."""
    W, H = 3840, 2160
    desc = rt.scenes.scatter(width=W, height=H, n_objects=10000)
    host = rt.HostScene(desc, lib, gpu_build=True)
    assert host.num_triangles >= 10_000_000
    dev = rt.DeviceScene(host)
    args = (host.camera, host.lights_pos, host.lights_color)
    a4 = dev.render(*args, host.params(spp_grid=2), want=("rgba8",))
    a1 = dev.render(*args, host.params(), want=("rgba8", "hit_id", "hit_t"))
    o = orc.OracleScene(desc, oracle_kind)
    rows = np.arange(5, H, 72)                             # 30 rows
    _rows_check(rt, a4, o.render(want_ids=False, y0=5, ystep=72, spp_grid=2), rows, ids=False)
    _rows_check(rt, a1, o.render(y0=5, ystep=72), rows)
    acc = np.zeros_like(a4["rgba8"])
    for r in range(8):
        acc |= dev.render(*args, host.params(spp_grid=2, tile=(r, 8, 16)), want=("rgba8",))["rgba8"]
    assert np.array_equal(acc, a4["rgba8"])


def test_frame_queue_matches_the_synchronous_call(rt, lib, gpu):
    """rt_frame_submit / rt_frame_wait (pipelined host-buffer frames, replaces the body of
    Raytracer::compute_image_cuda mytracer.cpp:123-159): every frame of a burst equals the rt_render frame, with the D2H
    copy on its own stream, with the kernel storing straight into the mapped host frame (zero_copy), with more frames
    in flight than the library's launch ring has slots, and with the lights changing from frame to frame."""
    desc = rt.scenes.office(width=642, height=363)
    host = rt.HostScene(desc, lib)
    dev = rt.DeviceScene(host)
    W, H = 642, 363
    nb = W * H * 4
    lp = np.ascontiguousarray(host.lights_pos); lc = np.ascontiguousarray(host.lights_color)
    ref = dev.render(host.camera, lp, lc, host.params(), want=("rgba8",))["rgba8"]
    lp2 = lp.copy(); lp2[0, 1] += 0.75
    ref2 = dev.render(host.camera, lp2, lc, host.params(), want=("rgba8",))["rgba8"]
    assert not np.array_equal(ref, ref2)
    for depth, zero_copy, ctas, flush in ((2, 0, 0, 0), (4, 1, 1, 0), (12, 0, 2, 1 << 20), (1, 1, 0, 0)):
        opts = gpu.FrameQueueOpts(depth, ctas, zero_copy, 0, flush)
        q = C.c_void_p()
        gpu.check(lib.rt_frame_queue_create(dev.handle, W, H, C.byref(opts), C.byref(q)), lib)
        nframes = 2 * depth + 3
        hbuf = C.c_void_p()
        gpu.check(lib.rt_host_alloc(nb * nframes, C.byref(hbuf)), lib)
        tick = C.c_longlong()
        p = host.params()
        tickets = []
        for f in range(nframes):
            lights = lp2 if f % 3 == 1 else lp
            arr = lights.copy()
            gpu.check(lib.rt_frame_submit(q, C.byref(host.camera), arr.ctypes.data_as(gpu.Vec4_p), lc.ctypes.data_as(gpu.Vec4_p),
                                          lp.shape[0], C.byref(p), C.c_void_p(hbuf.value + f * nb), C.byref(tick)), lib)
            arr[:] = 0.0                                     # the caller's arrays may be reused once submit returns
            tickets.append(tick.value)
        assert tickets == list(range(nframes))
        done = C.c_double()
        times = []
        for f in range(nframes):
            gpu.check(lib.rt_frame_wait(q, tickets[f], C.byref(done)), lib)
            times.append(done.value)
            got = np.ctypeslib.as_array((C.c_ubyte * nb).from_address(hbuf.value + f * nb)).reshape(H, W, 4)
            assert np.array_equal(got, ref2 if f % 3 == 1 else ref), (depth, zero_copy, f)
        assert min(times) > 0.0 and max(times) == max(times[-depth:])     # one completion timeline (ms since the first submit)
        assert lib.rt_frame_wait(q, nframes, None) == gpu.RT_ERR_INVALID_ARG
        gpu.check(lib.rt_frame_queue_destroy(q), lib)
        gpu.check(lib.rt_host_free(hbuf), lib)
    # registered (not allocated) host memory works the same: this is what a shared-memory segment uses
    buf = np.zeros(nb + 4096, np.uint8)
    base = (buf.ctypes.data + 4095) & ~4095
    dptr = C.c_void_p()
    gpu.check(lib.rt_host_register(C.c_void_p(base), nb, C.byref(dptr)), lib)
    import torch
    d_lpos = torch.tensor(lp, dtype=torch.float64, device="cuda"); d_lcol = torch.tensor(lc, dtype=torch.float64, device="cuda")
    fo = gpu.FrameOut(dptr.value, None, None, None)
    gpu.check(lib.rt_render_device(dev.handle, C.byref(host.camera), d_lpos.data_ptr(), d_lcol.data_ptr(), lp.shape[0],
                                   C.byref(host.params()), C.byref(fo), None), lib)
    torch.cuda.synchronize()
    got = np.ctypeslib.as_array((C.c_ubyte * nb).from_address(base)).reshape(H, W, 4)
    assert np.array_equal(got, ref)
    gpu.check(lib.rt_host_unregister(C.c_void_p(base)), lib)


def test_many_frames_in_flight_on_many_streams(rt, lib, gpu):
    """More launches in flight than the launch ring has slots (8), on 12 streams: launch f + 8 is ordered behind
    launch f, so no two kernels ever share a work counter and every frame is complete."""
    import torch
    W, H = 322, 181
    host = rt.HostScene(rt.scenes.office(width=W, height=H), lib)
    dev = rt.DeviceScene(host)
    ref = dev.render(host.camera, host.lights_pos, host.lights_color, host.params(), want=("rgba8",))["rgba8"]
    d_lpos = torch.tensor(host.lights_pos, dtype=torch.float64, device="cuda")
    d_lcol = torch.tensor(host.lights_color, dtype=torch.float64, device="cuda")
    n = 12
    streams = [torch.cuda.Stream() for _ in range(n)]
    bufs = [torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda") for _ in range(2 * n)]
    torch.cuda.synchronize()
    for f in range(2 * n):
        fo = gpu.FrameOut(bufs[f].data_ptr(), None, None, None)
        gpu.check(lib.rt_render_device(dev.handle, C.byref(host.camera), d_lpos.data_ptr(), d_lcol.data_ptr(),
                                       d_lpos.shape[0], C.byref(host.params()), C.byref(fo),
                                       C.c_void_p(streams[f % n].cuda_stream)), lib)
    torch.cuda.synchronize()
    for f in range(2 * n):
        assert np.array_equal(bufs[f].cpu().numpy().reshape(H, W, 4), ref), f


def test_corrupt_blobs_are_rejected(rt, lib, gpu):
    """rt_scene_create_from_blob validates what it adopts: the section table against the byte count and, on the device,
    every child reference, leaf range and index-table entry — a truncated or corrupted broadcast buffer gives
    RT_ERR_BAD_SCENE instead of out-of-bounds reads in the render kernel."""
    host = rt.HostScene(rt.scenes.mixed_small(width=32, height=24), lib)
    dev = rt.DeviceScene(host)
    blob = bytearray(dev.export_blob())
    h = C.c_void_p()

    def adopt(b):
        buf = (C.c_ubyte * len(b)).from_buffer_copy(bytes(b))
        st = lib.rt_scene_create_from_blob(buf, len(b), 0, C.byref(h))
        if st == 0:
            lib.rt_scene_destroy(h)
        return st
    assert adopt(blob) == 0
    assert adopt(blob[:len(blob) - 512]) == gpu.RT_ERR_BAD_SCENE              # truncated
    from my_raytracer_b200 import blobfmt
    hdr = blobfmt.header(blob)
    # a section offset beyond the blob
    bad = bytearray(blob); blobfmt.poke_u64(bad, "off_nodes4", len(blob) + 4096)
    assert adopt(bad) == gpu.RT_ERR_BAD_SCENE
    # a child reference outside the node array (first wide record, first child)
    if hdr["num_nodes4"] > 0:
        bad = bytearray(blob)
        off = hdr["off_nodes4"] + 96                                          # ref[0]
        bad[off:off + 4] = (0x7ffffff0).to_bytes(4, "little")
        bad[off + 16:off + 20] = (0).to_bytes(4, "little")                    # cnt[0] = 0: interior
        assert adopt(bad) == gpu.RT_ERR_BAD_SCENE
    # a triangle index table entry out of range
    bad = bytearray(blob)
    off = hdr["off_tri_id"]
    bad[off:off + 4] = (hdr["num_tris"] + 5).to_bytes(4, "little")
    assert adopt(bad) == gpu.RT_ERR_BAD_SCENE


@pytest.mark.parametrize("name,kw,spp,tile", [
    ("office", dict(width=642, height=363), 1, (0, 1, 1)),
    ("office", dict(width=322, height=181), 1, (2, 3, 4)),             # a rank's tile share, pixels of other ranks untouched
    ("mixed_small", dict(width=333, height=201, seed=9), 2, (0, 1, 1)),  # 2x2 sub-pixel grid: four passes over the same paths
    ("scatter", dict(width=256, height=144, n_objects=300), 1, (0, 1, 1)),   # d.x == 0 rays
    ("textured", dict(width=256, height=144, tex_size=256, slices=64, stacks=32), 1, (0, 1, 1)),
    ("molecule", dict(width=200, height=200, n_spheres=300), 1, (0, 1, 1)),  # object tree
    ("objtree", dict(width=192, height=144), 1, (0, 1, 1)),
    ("rings", dict(width=210, height=150, nu=96, nv=24), 1, (0, 1, 1)),
])
def test_wavefront_pipeline_is_bit_identical_to_the_persistent_kernel(rt, lib, gpu, name, kw, spp, tile):
    """RT_FLAG_WAVEFRONT (TRACE / SHADE kernel pairs per bounce, rt_wavefront.cuh) against the one persistent kernel:
    the same frame bit for bit — 8-bit and FP64 pixels, primary hit ids and ray parameters — also on the binary nodes."""
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    dev = rt.DeviceScene(host)
    args = (host.camera, host.lights_pos, host.lights_color)
    want = ("rgba8", "pixels", "hit_id", "hit_t") if spp == 1 else ("rgba8", "pixels")
    mega = dev.render(*args, host.params(spp_grid=spp, tile=tile, flags=gpu.RT_FLAG_MEGAKERNEL), want=want)
    assert dev.last_pipeline() == "megakernel"
    wave = dev.render(*args, host.params(spp_grid=spp, tile=tile, flags=gpu.RT_FLAG_WAVEFRONT), want=want)
    assert dev.last_pipeline() == "wavefront"
    wave_b = dev.render(*args, host.params(spp_grid=spp, tile=tile, flags=gpu.RT_FLAG_WAVEFRONT | gpu.RT_FLAG_BINARY_NODES), want=want)
    for k in want:
        assert np.array_equal(mega[k], wave[k]), k
        assert np.array_equal(mega[k], wave_b[k]), k
    # frames the wavefront does not cover fall back to the persistent kernel
    dev.render(*args, host.params(spp_grid=spp, tile=tile, flags=gpu.RT_FLAG_WAVEFRONT | gpu.RT_FLAG_COUNTERS), want=("rgba8",))
    assert dev.last_pipeline() == "megakernel"


@pytest.mark.parametrize("name,kw", [
    ("office", dict(width=642, height=363)),
    ("rings", dict(width=210, height=150, nu=96, nv=24)),
    ("textured", dict(width=256, height=144, tex_size=256, slices=64, stacks=32)),
    ("scatter", dict(width=256, height=144, n_objects=300)),
])
def test_eight_wide_tree_on_the_gpu(rt, lib, gpu, name, kw):
    """render_kernel<K_WIDE | K_WIDE8> (RT_FLAG_NODES8) against the 4-wide default, plain and instrumented: the same frame
    bit for bit, the same rays, fewer node visits."""
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    dev = rt.DeviceScene(host)
    assert dev.info()["num_nodes_wide8"] > 0
    args = (host.camera, host.lights_pos, host.lights_color)
    w4 = dev.render(*args, host.params(flags=gpu.RT_FLAG_COUNTERS))
    w8 = dev.render(*args, host.params(flags=gpu.RT_FLAG_NODES8))
    w8c = dev.render(*args, host.params(flags=gpu.RT_FLAG_NODES8 | gpu.RT_FLAG_COUNTERS))
    for other in (w8, w8c):
        for k in ("hit_id", "hit_t", "rgba8", "pixels"):
            assert np.array_equal(w4[k], other[k]), k
    a, b = w4["counters"], w8c["counters"]
    for k in ("rays_primary", "rays_shadow", "rays_reflect", "shaded_hits", "shadow_occluded"):
        assert a[k] == b[k], k
    assert b["nodes8_visited"] > 0 and b["nodes_visited"] < 0.8 * a["nodes_visited"]


@pytest.mark.parametrize("name,kw", [
    ("office", dict(width=480, height=270)),
    ("rings", dict(width=210, height=150)),
    ("textured", dict(width=256, height=144, tex_size=256)),
    ("scatter", dict(width=256, height=144, n_objects=300)),
    ("mixed_small", dict(width=160, height=120)),
    ("spheres", dict(width=64, height=64)),                    # no triangles: no tree to build
])
def test_gpu_sah_builder_against_the_host_builder(rt, lib, gpu, name, kw):
    """The traversal tree built on the GPU (rt_sah_gpu.cu, default) against the host builder (RT_SCENE_HOST_SAH): a
    different tree, the same frame bit for bit (ids, ray parameters, FP64 pixels), comparable quality (node visits and
    triangle tests within 15 %), and a reproducible blob (two device builds are byte-identical)."""
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    dev = rt.DeviceScene(host)
    ref = rt.DeviceScene(host, scene_flags=gpu.RT_SCENE_HOST_SAH)
    tris = host.num_triangles
    assert dev.info()["sah_builder"] == (2 if tris > 2 else 0) and ref.info()["sah_builder"] == (1 if tris > 2 else 0)
    args = (host.camera, host.lights_pos, host.lights_color)
    a = dev.render(*args, host.params(flags=gpu.RT_FLAG_COUNTERS))
    b = ref.render(*args, host.params(flags=gpu.RT_FLAG_COUNTERS))
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(a[k], b[k]), k
    ca, cb = a["counters"], b["counters"]
    for k in ("rays_primary", "rays_shadow", "rays_reflect", "shaded_hits", "shadow_occluded"):
        assert ca[k] == cb[k], k
    if tris > 2:
        assert ca["nodes_visited"] < 1.15 * cb["nodes_visited"] and ca["tri_filter"] < 1.15 * cb["tri_filter"]
        again = rt.DeviceScene(host)
        assert again.export_blob() == dev.export_blob()
        # the 8-wide twin and the binary records of the GPU-built tree render the same frame too
        for fl in (gpu.RT_FLAG_NODES8, gpu.RT_FLAG_BINARY_NODES, gpu.RT_FLAG_EXACT_ONLY):
            c = dev.render(*args, host.params(flags=fl))
            for k in ("hit_id", "hit_t", "rgba8", "pixels"):
                assert np.array_equal(a[k], c[k]), (fl, k)


@pytest.mark.parametrize("name,kw", [
    ("office", dict(width=480, height=270)),
    ("rings", dict(width=210, height=150)),                    # PHONG vertex normals
    ("textured", dict(width=256, height=144, tex_size=256)),   # uv records + texels
    ("scatter", dict(width=256, height=144, n_objects=300)),   # triangle trees + a tree over analytic spheres
    ("mixed_small", dict(width=160, height=120)),              # meshes, texture, spheres and a plane tested linearly
])
def test_device_flattener_against_the_host_flattener(rt, lib, gpu, name, kw):
    """The blob emitted on the GPU (rt_flatten_gpu.cu, default) against the host flattener (RT_SCENE_HOST_FLATTEN) on the
    same GPU-built tree: per-triangle sections byte-identical, the same tree structure under a different record numbering
    (level order instead of DFS pre-order: equal record counts and equal traversal counters), the same frame bit for bit
    on every kernel variant, and a reproducible blob."""
    from my_raytracer_b200 import blobfmt
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    dev = rt.DeviceScene(host)
    ref = rt.DeviceScene(host, scene_flags=gpu.RT_SCENE_HOST_FLATTEN)
    assert dev.info()["flattener"] == 1 and ref.info()["flattener"] == 0
    assert dev.info()["sah_builder"] == 2 and ref.info()["sah_builder"] == 2
    ba, bb = dev.export_blob(), ref.export_blob()
    ha, hb = blobfmt.header(ba), blobfmt.header(bb)
    for k, v in hb.items():
        assert ha[k] == v, (k, ha[k], v)               # every count, offset and extent of the header
    n = ha["num_tris"]
    for off, size in (("off_filter", 48 * n), ("off_exact", 80 * n), ("off_shade", 128 * n), ("off_tri_id", 4 * n),
                      ("off_tri_pos", 4 * n), ("off_meshes", 128 * ha["num_meshes"]), ("off_objects", 176 * ha["num_objects"]),
                      ("off_objfilt", 32 * ha["num_objects"]), ("off_texels", 16 * ha["num_texels"]),
                      ("off_obj_order", 4 * ha["num_obj_tree"]), ("off_lin_idx", 4 * ha["num_lin"])):
        o = ha[off]
        assert ba[o:o + size] == bb[o:o + size], off
    args = (host.camera, host.lights_pos, host.lights_color)
    a = dev.render(*args, host.params(flags=gpu.RT_FLAG_COUNTERS))
    b = ref.render(*args, host.params(flags=gpu.RT_FLAG_COUNTERS))
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(a[k], b[k]), k
    assert a["counters"] == b["counters"]              # the same tree walked in the same order
    for fl in (gpu.RT_FLAG_NODES8, gpu.RT_FLAG_BINARY_NODES, gpu.RT_FLAG_EXACT_ONLY, gpu.RT_FLAG_WAVEFRONT):
        c = dev.render(*args, host.params(flags=fl))
        for k in ("hit_id", "hit_t", "rgba8", "pixels"):
            assert np.array_equal(a[k], c[k]), (fl, k)
    assert rt.DeviceScene(host).export_blob() == ba


def test_device_flattener_leaves_malformed_scenes_to_the_host_flattener(rt, lib, gpu):
    """Whatever the device flattener's validation rejects is flattened again by the host twin, which words the error:
    same status and message as with RT_SCENE_HOST_FLATTEN."""
    def attempt(mutate, flags):
        host = rt.HostScene(rt.scenes.office(width=64, height=36), lib)
        keep = mutate(host)
        h = C.c_void_p()
        st = lib.rt_scene_create_ex(C.byref(host.data), C.byref(host.bvh), host.objects, host.num_objects, flags, C.byref(h))
        msg = lib.rt_last_error()
        if st == 0:
            lib.rt_scene_destroy(h)
        del keep
        return st, msg

    def bad_child(host):
        host.bvh.leftChildIdx_[0] = 2 * host.num_triangles + 5
    def bad_vertex(host):
        host.data.vertexIdx_[7] = host.data.tVertexCount_
    def shared_leaf(host):
        # the right child of the root becomes a copy of the left one: its triangles are referenced twice, others never
        l = host.bvh.leftChildIdx_[0]
        host.bvh.leftChildIdx_[l + 1] = host.bvh.leftChildIdx_[l]
        host.bvh.triCount_[l + 1] = host.bvh.triCount_[l]
        host.bvh.firstTriIdx_[l + 1] = host.bvh.firstTriIdx_[l]
    for mutate in (bad_child, bad_vertex, shared_leaf):
        st_d, msg_d = attempt(mutate, 0)
        st_h, msg_h = attempt(mutate, gpu.RT_SCENE_HOST_FLATTEN)
        assert st_d == st_h == gpu.RT_ERR_BAD_SCENE and msg_d == msg_h, (mutate.__name__, st_d, st_h, msg_d, msg_h)


@pytest.mark.parametrize("name,kw", [
    ("office", dict(width=322, height=181)),
    ("textured", dict(width=200, height=120, tex_size=128)),
    ("scatter", dict(width=160, height=100, n_objects=20)),    # triangle trees + a tree over analytic spheres (K_OBJ kernels)
    ("mixed_small", dict(width=161, height=99)),               # ragged last tile column / row
])
def test_one_sample_per_lane_equals_the_sequential_sample_loop(rt, lib, gpu, name, kw):
    """2x2 and 4x4 sub-pixel grids: the samples of a pixel traced by neighbouring lanes and summed in sample order
    (default, K_SPL kernels) give the same FP64 pixels, bit for bit, as one lane tracing them one after the other
    (RT_FLAG_SEQUENTIAL_SAMPLES) — on the whole frame, on a rank's tile share with the dynamic tile queue, and on the
    4-wide and the binary records.  A 3x3 grid (9 lanes do not divide a warp) keeps the sequential loop."""
    import torch
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    dev = rt.DeviceScene(host)
    args = (host.camera, host.lights_pos, host.lights_color)
    for spp in (2, 4, 3):
        a = dev.render(*args, host.params(spp_grid=spp), want=("rgba8", "pixels"))
        b = dev.render(*args, host.params(spp_grid=spp, flags=gpu.RT_FLAG_SEQUENTIAL_SAMPLES), want=("rgba8", "pixels"))
        assert np.array_equal(a["pixels"], b["pixels"]) and np.array_equal(a["rgba8"], b["rgba8"]), spp
        assert a["pixels"][..., :3].max() > 0.0
        c = dev.render(*args, host.params(spp_grid=spp, flags=gpu.RT_FLAG_BINARY_NODES), want=("pixels",))
        assert np.array_equal(a["pixels"], c["pixels"]), spp
    # two ranks' shares of one frame, the last tiles from the dynamic queue (one counter, as on one GPU in the tests)
    counter = torch.zeros(1, dtype=torch.int64, device="cuda")
    W, H = host.camera.width, host.camera.height
    tiles = ((W + 7) // 8) * ((H + 3) // 4)
    whole = dev.render(*args, host.params(spp_grid=2), want=("pixels",))["pixels"]
    acc = np.zeros_like(whole)
    for rank in (0, 1):
        p = host.params(spp_grid=2, tile=(rank, 2, 4), dyn=(tiles // 3, 2, counter.data_ptr()))
        part = dev.render(*args, p, want=("pixels",))["pixels"]
        acc = np.where(part != 0.0, part, acc)
    torch.cuda.synchronize()
    assert np.array_equal(acc, whole)


@pytest.mark.parametrize("name,kw", [
    ("rings", dict(width=210, height=150)),                    # deep mirror chains on a PHONG mesh: what the kernel is for
    ("office", dict(width=322, height=181)),
    ("spheres", dict(width=128, height=96)),                   # analytic primitives only
    ("scatter", dict(width=160, height=100, n_objects=20)),    # triangle trees + a tree over analytic spheres (K_OBJ)
    ("mixed_small", dict(width=161, height=99)),
])
def test_ray_parallel_kernel_is_bit_identical_to_the_persistent_kernel(rt, lib, gpu, name, kw):
    """RT_FLAG_RAY_PARALLEL: the shadow rays and the mirror ray of a hit traced together on neighbouring lanes, the path
    state in shared memory, the wavefront's stage functions.  Same ids, ray parameters and FP64 pixels as the persistent
    kernel — with 1, 2-3, 5 and 9 lights (2, 4, 8 lanes per path; more ray kinds than lanes), a 2x2 sample grid and a
    rank's tile share."""
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    dev = rt.DeviceScene(host)
    lp0, lc0 = np.asarray(host.lights_pos, np.float64), np.asarray(host.lights_color, np.float64)
    rng = np.random.default_rng(7)

    def lights(n):
        if n <= lp0.shape[0]:
            return lp0[:n].copy(), lc0[:n].copy()
        extra = n - lp0.shape[0]
        lp = np.vstack([lp0, np.hstack([rng.uniform(-6, 6, (extra, 2)), rng.uniform(4, 9, (extra, 1)), np.zeros((extra, 1))])])
        lc = np.vstack([lc0, np.hstack([rng.uniform(0.05, 0.2, (extra, 3)), np.zeros((extra, 1))])])
        return lp, lc

    for nl, spp, tile in ((lp0.shape[0], 1, (0, 1, 1)), (1, 1, (0, 1, 1)), (5, 1, (0, 1, 1)), (9, 1, (0, 1, 1)),
                          (lp0.shape[0], 2, (0, 1, 1)), (lp0.shape[0], 1, (1, 3, 2))):
        lp, lc = lights(nl)
        want = ("rgba8", "pixels", "hit_id", "hit_t") if spp == 1 else ("rgba8", "pixels")
        a = dev.render(host.camera, lp, lc, host.params(spp_grid=spp, tile=tile, flags=gpu.RT_FLAG_MEGAKERNEL), want=want)
        assert dev.last_pipeline() == "megakernel"
        b = dev.render(host.camera, lp, lc, host.params(spp_grid=spp, tile=tile, flags=gpu.RT_FLAG_RAY_PARALLEL), want=want)
        assert dev.last_pipeline() == "rayparallel"
        for k in want:
            assert np.array_equal(a[k], b[k]), (nl, spp, tile, k)
        assert a["pixels"][..., :3].max() > 0.0
    # frames it does not cover fall back to the persistent kernel: instrumented frames, no light at all
    dev.render(host.camera, lp0, lc0, host.params(flags=gpu.RT_FLAG_RAY_PARALLEL | gpu.RT_FLAG_COUNTERS), want=("rgba8",))
    assert dev.last_pipeline() == "megakernel"
