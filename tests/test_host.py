"""Host-side logic of the product (CPU): C-ABI surface, the BVH/normal/camera mirrors against the
oracle, the flattener's structure and validation, tile sharding arithmetic, and the no-GPU error."""
import ast
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT


def test_cabi_exports_every_declared_symbol(rt, lib):
    from my_raytracer_b200 import capi
    hdr = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    declared = set(re.findall(r"^(?:int|const char \*)\s*(rt_[a-z_0-9]+)\(", hdr, re.M))
    assert declared == set(capi.PROTOTYPES), declared ^ set(capi.PROTOTYPES)
    for name in declared:
        assert hasattr(lib, name)
    assert lib.rt_abi_version() == 6


def test_struct_layouts_match_reference_headers(rt):
    """rt_data_soa / rt_bvh_nodes_soa are field-for-field mirrors of struct Data (mydata.h:28-72) and
    BVH::BVHNodes_SoA (mybvh.h:49-55): 5 ints + size_t + 27 pointers, and 5 pointers."""
    from my_raytracer_b200 import capi
    assert C.sizeof(capi.Vec4) == 32
    assert C.sizeof(capi.DataSoA) == 24 + 8 + 27 * 8
    assert C.sizeof(capi.BvhNodesSoA) == 5 * 8
    assert capi.DataSoA.vertexMeshId_.offset == 32 and capi.DataSoA.materialShadowable_.offset == 32 + 26 * 8


def test_no_gpu_fails_loudly(rt, lib):
    """Without a CUDA device every device entry point reports RT_ERR_NO_DEVICE: no CPU fallback."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from my_raytracer_b200 import capi
    assert lib.rt_init_device(0) == capi.RT_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.rt_last_error()
    host = rt.HostScene(rt.scenes.mixed_small(width=8, height=8), lib)
    with pytest.raises(capi.RtError) as e:
        rt.DeviceScene(host)
    assert e.value.status == capi.RT_ERR_NO_DEVICE


@pytest.mark.parametrize("name", ["mixed_small", "office", "rings", "textured"])
def test_bvh_build_matches_reference(rt, lib, name):
    """rt_bvh_build_median == BVH::initSoA/init: node arrays, bounds and triangle permutation."""
    g = np.load(os.path.join(GOLDEN, f"scene_{name}.npz"))
    desc = rt.scenes.REGISTRY[name](**ast.literal_eval(str(g["kwargs"])))
    host = rt.HostScene(desc, lib)
    n = host.nodes_used
    assert n == g["bvh_left"].shape[0]
    b = host.bvh_arrays
    interior = g["bvh_count"] == 0
    assert np.array_equal(b["count"][:n], g["bvh_count"])
    assert np.array_equal(b["left"][:n][interior], g["bvh_left"][interior])
    assert np.array_equal(b["first"][:n][~interior], g["bvh_first"][~interior])
    assert np.array_equal(b["bbmin"][:n, :3], g["bvh_min"]) and np.array_equal(b["bbmax"][:n, :3], g["bvh_max"])
    # the permuted vertexIdx_ rows are the original rows in the reference's BVH order
    assert np.array_equal(host.arrays["vertexIdx"].reshape(-1, 3), host._orig_vidx[g["tri_order"]])


def test_normals_and_camera_match_oracle(rt, lib, orc, oracle_kind):
    desc = rt.scenes.mixed_small(width=40, height=30)
    host = rt.HostScene(desc, lib)
    o = orc.OracleScene(desc, oracle_kind)
    order = o.tri_order()
    fn_all = np.concatenate([o.normals(m)[0] for m in range(len(desc.meshes))], 0)
    vn_all = np.concatenate([o.normals(m)[1] for m in range(len(desc.meshes))], 0)
    assert np.array_equal(host.arrays["normals"][:, :3], fn_all[order], equal_nan=True)
    assert np.array_equal(host.arrays["vertexNormals"][:, :3], vn_all, equal_nan=True)
    cam = host.camera
    mine = np.array([[v.x, v.y, v.z] for v in (cam.eye, cam.x_dir, cam.y_dir, cam.lower_left)])
    assert np.array_equal(mine, o.camera())


def test_build_data_offsets(rt, lib):
    """build_Data mirror (mytracer.cpp:166-296): global vertex/uv ids, per-mesh tables, texel layout."""
    desc = rt.scenes.mixed_small(width=8, height=8)
    host = rt.HostScene(desc, lib)
    a = host.arrays
    nv = [m.pos.shape[0] for m in desc.meshes]
    assert list(a["firstVertex"]) == list(np.cumsum([0] + nv[:-1])) and list(a["vertexCount"]) == nv
    assert list(a["vertexIdxCount"]) == [3 * m.idx.shape[0] for m in desc.meshes]
    assert list(a["texW"]) == [64, -1, -1, -1] and a["firstMeshTex"][0] == 0
    vid = a["vertexIdx"].reshape(-1, 3)
    assert ((a["vertexMeshId"][vid[:, 0]] == a["vertexMeshId"][vid[:, 1]]) &
            (a["vertexMeshId"][vid[:, 0]] == a["vertexMeshId"][vid[:, 2]])).all()
    tex = desc.meshes[0].tex
    assert np.array_equal(a["texels"][5 * 64 + 7, :3], tex[5, 7])


def _blob_sections(blob):
    import struct
    magic, ver, total = struct.unpack_from("<IIQ", blob, 0)
    ints = struct.unpack_from("<8i", blob, 16)
    ntex, = struct.unpack_from("<Q", blob, 48)
    offs = struct.unpack_from("<7Q", blob, 56)
    off_maxx, = struct.unpack_from("<Q", blob, 128)
    off_tri_id, off_tri_pos = struct.unpack_from("<2Q", blob, 144)
    q_base, num_nodes_q, topology, num_nodes4 = struct.unpack_from("<4i", blob, 160)
    off_nodes4, = struct.unpack_from("<Q", blob, 216)
    return dict(num_nodes4=num_nodes4, off_nodes4=off_nodes4, magic=magic, version=ver, total=total, num_nodes=ints[0], num_tris=ints[1], num_meshes=ints[2],
                num_objects=ints[3], root_is_leaf=ints[4], max_depth=ints[5], max_leaf=ints[6], num_nodes_ref=ints[7],
                num_texels=ntex, offs=offs, off_maxx=off_maxx, off_tri_id=off_tri_id, off_tri_pos=off_tri_pos,
                q_base=q_base, num_nodes_q=num_nodes_q, topology=topology)


def test_flatten_structure(rt, lib, sim):
    """Same split topology: every leaf range of the reference tree appears exactly once, boxes are
    outward-rounded supersets of the FP64 boxes."""
    desc = rt.scenes.office(width=16, height=9)
    host = rt.HostScene(desc, lib)
    blob = sim.flatten(host, rt.capi.RT_SCENE_REFERENCE_TOPOLOGY)
    h = _blob_sections(blob)
    n = host.nodes_used
    cnt = host.bvh_arrays["count"][:n]
    assert h["total"] == len(blob) and h["num_tris"] == 34000 and h["num_nodes_ref"] == n
    assert h["num_nodes"] == int((cnt == 0).sum()) and h["max_leaf"] == int(cnt.max())
    assert h["topology"] == 1 and h["q_base"] == 0 and h["num_nodes_q"] == h["num_nodes"]
    ident = np.arange(34000, dtype=np.int32)
    assert np.array_equal(np.frombuffer(blob, np.int32, 34000, h["off_tri_id"]), ident)
    assert np.array_equal(np.frombuffer(blob, np.int32, 34000, h["off_tri_pos"]), ident)
    nodes = np.frombuffer(blob, np.float32, h["num_nodes"] * 16, h["offs"][0]).reshape(-1, 16)
    refs = nodes[:, 12:16].view(np.int32)
    leaves = set()
    for r0, r1, c0, c1 in refs:
        if c0 > 0: leaves.add((int(r0), int(c0)))
        if c1 > 0: leaves.add((int(r1), int(c1)))
    first = host.bvh_arrays["first"][:n]
    assert leaves == {(int(f), int(c)) for f, c in zip(first[cnt > 0], cnt[cnt > 0])}
    # root record = the two children of reference node 0, conservatively rounded
    l = host.bvh_arrays["left"][0]
    mn, mx = host.bvh_arrays["bbmin"], host.bvh_arrays["bbmax"]
    # (centre / half-extent form: [c - h, c + h] evaluated in FP64 must contain the FP64 box, tightly)
    n0 = nodes[0].astype(np.float64)
    for k, ref in ((0, l), (6, l + 1)):
        c, hh = n0[k:k + 3], n0[k + 3:k + 6]
        assert (c - hh <= mn[ref, :3]).all() and (c + hh >= mx[ref, :3]).all()
        assert np.abs((c - hh) - mn[ref, :3]).max() < 1e-5 and np.abs((c + hh) - mx[ref, :3]).max() < 1e-5
    # every record of the tree: same containment against the reference boxes (walk both trees together)
    todo = [(0, 0)]
    while todo:
        ref, flat = todo.pop()
        lc = host.bvh_arrays["left"][ref]
        rec = nodes[flat].astype(np.float64)
        for s_, k in ((0, 0), (1, 6)):
            c, hh = rec[k:k + 3], rec[k + 3:k + 6]
            assert (c - hh <= mn[lc + s_, :3]).all() and (c + hh >= mx[lc + s_, :3]).all()
            if refs[flat][2 + s_] == 0:
                todo.append((lc + s_, int(refs[flat][s_])))


def test_flatten_builds_its_own_traversal_tree(rt, lib, sim):
    """Default flatten: SAH tree in records [0, q_base), the reference tree unchanged behind it; every triangle
    in exactly one SAH leaf, every child box contains the FP64 triangles below it, tri_id / tri_pos are inverse
    permutations and the filter / exact records follow the SAH leaf order."""
    desc = rt.scenes.office(width=16, height=9)
    host = rt.HostScene(desc, lib)
    ref = sim.flatten(host, rt.capi.RT_SCENE_REFERENCE_TOPOLOGY)
    blob = sim.flatten(host)
    hr, h = _blob_sections(ref), _blob_sections(blob)
    N = 34000
    assert h["topology"] == 0 and h["q_base"] > 0 and h["num_nodes_q"] == hr["num_nodes"]
    assert h["num_nodes"] == h["q_base"] + h["num_nodes_q"]
    nodes = np.frombuffer(blob, np.float32, h["num_nodes"] * 16, h["offs"][0]).reshape(-1, 16)
    refs = nodes[:, 12:16].view(np.int32)
    # the reference tree: same records, interior child references moved by q_base
    rn = np.frombuffer(ref, np.float32, hr["num_nodes"] * 16, hr["offs"][0]).reshape(-1, 16)
    rr = rn[:, 12:16].view(np.int32).copy()
    rr[:, 0] += np.where(rr[:, 2] == 0, h["q_base"], 0)
    rr[:, 1] += np.where(rr[:, 3] == 0, h["q_base"], 0)
    assert np.array_equal(nodes[h["q_base"]:, :12], rn[:, :12]) and np.array_equal(refs[h["q_base"]:], rr)
    assert blob[h["off_maxx"]:h["off_maxx"] + 16 * h["num_nodes_q"]] == ref[hr["off_maxx"]:hr["off_maxx"] + 16 * hr["num_nodes"]]
    tri_id = np.frombuffer(blob, np.int32, N, h["off_tri_id"])
    tri_pos = np.frombuffer(blob, np.int32, N, h["off_tri_pos"])
    assert np.array_equal(np.sort(tri_id), np.arange(N)) and np.array_equal(tri_pos[tri_id], np.arange(N))
    exact_r = np.frombuffer(ref, np.float64, N * 10, hr["offs"][2]).reshape(N, 10)
    exact_s = np.frombuffer(blob, np.float64, N * 10, h["offs"][2]).reshape(N, 10)
    assert np.array_equal(exact_s, exact_r[tri_id])
    filt_r = np.frombuffer(ref, np.float32, N * 12, hr["offs"][1]).reshape(N, 12)
    filt_s = np.frombuffer(blob, np.float32, N * 12, h["offs"][1]).reshape(N, 12)
    assert np.array_equal(filt_s, filt_r[tri_id])
    assert blob[h["offs"][3]:h["offs"][4]] == ref[hr["offs"][3]:hr["offs"][4]]      # TriShade stays in reference order
    verts = exact_s[:, :9].reshape(N, 3, 3)
    tmin, tmax = verts.min(1), verts.max(1)
    covered = np.zeros(N, np.int32)

    def walk(rec):
        """returns the contiguous position range below record `rec`; checks both child boxes"""
        lo_all, hi_all = N, 0
        r = nodes[rec].astype(np.float64)
        for s_, k in ((0, 0), (1, 6)):
            c, hh = r[k:k + 3], r[k + 3:k + 6]
            if refs[rec][2 + s_] > 0:
                a, b = int(refs[rec][s_]), int(refs[rec][s_] + refs[rec][2 + s_])
                covered[a:b] += 1
            else:
                a, b = walk(int(refs[rec][s_]))
            assert (c - hh <= tmin[a:b].min(0)).all() and (c + hh >= tmax[a:b].max(0)).all()
            lo_all, hi_all = min(lo_all, a), max(hi_all, b)
        if refs[rec][2] > 0 and refs[rec][3] > 0:
            assert refs[rec][1] == refs[rec][0] + refs[rec][2]      # sibling leaves own adjacent ranges
        return lo_all, hi_all

    import sys
    sys.setrecursionlimit(10000)
    assert walk(0) == (0, N) and (covered == 1).all()
    assert h["max_depth"] <= 64


def test_wide_tree_structure(rt, lib, sim):
    """The 4-wide twin of the traversal tree (Node4Rec, 128 B): children stored as two word-interleaved pairs; every
    triangle position lies in exactly one leaf; a child's box contains the FP64 vertices of everything below it; the
    children of a node own ascending, gap-free position ranges (so neighbouring leaves can be tested as one loop);
    empty slots can never be hit; and the worst-case stack need (interior children - 1 per level) fits the kernel's
    64-entry stack."""
    host = rt.HostScene(rt.scenes.office(width=16, height=9), lib)
    blob = sim.flatten(host)
    h = _blob_sections(blob)
    N = 34000
    assert h["version"] == 12 and h["num_nodes4"] > 0 and h["off_nodes4"] % 32 == 0
    assert h["off_nodes4"] + 128 * h["num_nodes4"] <= h["total"]
    assert h["num_nodes4"] < 0.6 * h["q_base"]                       # at least ~2 binary records per wide record
    rec = np.frombuffer(blob, np.float32, h["num_nodes4"] * 32, h["off_nodes4"]).reshape(-1, 32)
    box = rec[:, :24].reshape(-1, 2, 6, 2).astype(np.float64)         # [record][pair][cx cy cz hx hy hz][child a, b]
    ref = rec[:, 24:28].view(np.int32)
    cnt = rec[:, 28:32].view(np.int32)
    exact = np.frombuffer(blob, np.float64, N * 10, h["offs"][2]).reshape(N, 10)
    verts = exact[:, :9].reshape(N, 3, 3)
    tmin, tmax = verts.min(1), verts.max(1)
    covered = np.zeros(N, np.int32)
    seen = np.zeros(h["num_nodes4"], np.int32)
    worst = [0]

    def walk(r, stack):
        seen[r] += 1
        interior = int(((cnt[r] == 0) & (ref[r] >= 0)).sum())
        stack += max(interior - 1, 0)
        worst[0] = max(worst[0], stack)
        lo_all, hi_all, prev_end = N, 0, None
        for k in range(4):
            c, hh = box[r, k >> 1, 0:3, k & 1], box[r, k >> 1, 3:6, k & 1]
            if ref[r, k] < 0:                                          # empty slot: only after the used ones, never hit
                assert cnt[r, k] == 0 and (hh < -1e29).all() and k >= 2
                continue
            if cnt[r, k] > 0:
                a, b = int(ref[r, k]), int(ref[r, k] + cnt[r, k])
                assert cnt[r, k] <= h["max_leaf"]
                covered[a:b] += 1
            else:
                a, b = walk(int(ref[r, k]), stack)
            assert (c - hh <= tmin[a:b].min(0)).all() and (c + hh >= tmax[a:b].max(0)).all()
            assert prev_end is None or a == prev_end                   # ascending and adjacent
            prev_end = b
            lo_all, hi_all = min(lo_all, a), max(hi_all, b)
        return lo_all, hi_all

    import sys
    sys.setrecursionlimit(10000)
    assert walk(0, 0) == (0, N) and (covered == 1).all() and (seen == 1).all()
    assert worst[0] <= 64


def test_flatten_rejects_bad_input(rt, lib, sim):
    desc = rt.scenes.mixed_small(width=8, height=8)
    host = rt.HostScene(desc, lib)
    host.bvh_arrays["left"][0] = 0                       # cycle: root is its own child
    with pytest.raises(RuntimeError, match="out of range|cycle"):
        sim.flatten(host)
    host = rt.HostScene(desc, lib)
    host.arrays["vertexIdx"][0] = 10**6
    with pytest.raises(RuntimeError, match="vertex index"):
        sim.flatten(host)
    host = rt.HostScene(desc, lib)
    leafs = np.nonzero(host.bvh_arrays["count"][:host.nodes_used] > 0)[0]
    host.bvh_arrays["first"][leafs[3]] = host.bvh_arrays["first"][leafs[2]]   # overlapping leaves
    with pytest.raises(RuntimeError, match="leaf"):
        sim.flatten(host)


def test_flatten_edge_scenes(rt, lib, sim):
    """empty scene, analytic-only scene, 1-triangle and 2-triangle meshes (root is a leaf)."""
    empty = rt.SceneDesc("empty", [], [], [], dict(eye=(0, 0, 5), center=(0, 0, 0), up=(0, 1, 0), fovy=40, width=8, height=8))
    r = sim.render(rt.HostScene(empty, lib))
    assert (r["hit_id"] == -1).all() and (r["rgba8"][..., :3] == 0).all()
    for ntri in (1, 2):
        pos = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0.0]])
        idx = np.array([[0, 1, 2], [0, 2, 3]], np.int32)[:ntri]
        d = rt.SceneDesc("tiny", [rt.MeshDesc(pos, idx, rt.material((1, 1, 1), (1, 1, 1), (0, 0, 0), 1.0))], [],
                         [((0, 0, 5), (1, 1, 1))], dict(eye=(0, 0, 5), center=(0, 0, 0), up=(0, 1, 0), fovy=40, width=16, height=16),
                         ambience=(0.5, 0.5, 0.5))
        host = rt.HostScene(d, lib)
        assert _blob_sections(sim.flatten(host))["root_is_leaf"] == 1
        r = sim.render(host)
        assert set(np.unique(r["hit_id"])) <= {-1, 0, 1} and (r["hit_id"] >= 0).any()


def test_tile_sharding_partitions_the_image(rt, lib, sim):
    """The union over ranks of the tile-sharded renders is the full image, each pixel written once."""
    host = rt.HostScene(rt.scenes.mixed_small(width=70, height=45), lib)     # not multiples of 8x4
    blob = sim.flatten(host)
    full = sim.render(host, blob)
    for world, chunk in ((2, 1), (3, 2), (8, 4)):
        acc = np.zeros_like(full["rgba8"]); seen = np.zeros(full["hit_id"].shape, int)
        for r in range(world):
            part = sim.render(host, blob, tile=(r, world, chunk))
            mask = part["hit_id"] != -2
            seen += mask
            acc[mask] = part["rgba8"][mask]
        assert (seen == 1).all()
        assert np.array_equal(acc, full["rgba8"])
    # with a dynamic tail: the last dyn_tiles tiles are claimed in chunks instead of being owned statically
    ntiles = ((70 + 7) // 8) * ((45 + 3) // 4)
    for world, chunk, dyn_tiles, dyn_chunk in ((2, 2, 30, 4), (3, 1, ntiles, 5), (4, 3, 7, 16), (2, 2, 10 ** 6, 1)):
        acc = np.zeros_like(full["rgba8"]); seen = np.zeros(full["hit_id"].shape, int)
        for r in range(world):
            part = sim.render(host, blob, tile=(r, world, chunk), dyn=(dyn_tiles, dyn_chunk))
            mask = part["hit_id"] != -2
            seen += mask
            acc[mask] = part["rgba8"][mask]
        assert (seen == 1).all()
        assert np.array_equal(acc, full["rgba8"])


def test_blob_header_mirror_and_frame_inputs(rt, lib, sim):
    """blobfmt.BlobHeader == csrc/rt_types.h BlobHeader (256 bytes, same offsets): magic, version, byte count and an
    ascending 512-byte-aligned section table on a blob the flattener made; FrameInputs (what a rank that did not build
    the scene submits frames with) carries the same camera, lights and render parameters as the HostScene."""
    from my_raytracer_b200 import blobfmt
    from my_raytracer_b200 import dist as rdist
    desc = rt.scenes.mixed_small(width=16, height=12)
    host = rt.HostScene(desc, lib)
    blob = sim.flatten(host)
    h = blobfmt.header(blob)
    assert C.sizeof(blobfmt.BlobHeader) == 256
    assert h["magic"] == blobfmt.BLOB_MAGIC and h["version"] == blobfmt.BLOB_VERSION and h["total_bytes"] == len(blob)
    offs = [h[k] for k in ("off_nodes", "off_filter", "off_exact", "off_shade", "off_meshes", "off_objects", "off_texels",
                           "off_maxx", "off_tri_id", "off_tri_pos", "off_objfilt", "off_obj_order", "off_lin_idx", "off_nodes4", "off_nodes8")]
    assert offs == sorted(offs) and all(o % 512 == 0 and o <= len(blob) for o in offs)
    assert h["num_tris"] == host.num_triangles and h["num_objects"] == host.num_objects
    fi = rdist.FrameInputs(desc, lib)
    assert bytes(fi.camera) == bytes(host.camera)
    assert np.array_equal(fi.lights_pos, host.lights_pos) and np.array_equal(fi.lights_color, host.lights_color)
    assert bytes(fi.params(spp_grid=2, tile=(1, 4, 16))) == bytes(host.params(spp_grid=2, tile=(1, 4, 16)))
