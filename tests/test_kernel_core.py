"""The kernel core (rt_core.cuh / rt_intersect.cuh) compiled for the host, against the oracle (CPU).
Catches logic errors before GPU minutes are spent; the GPU parity tests proper are in test_gpu_parity.py."""
import ast
import os

import numpy as np
import pytest

from conftest import GOLDEN

SCENES = ["mixed_small", "spheres", "office", "rings", "molecule", "textured"]


@pytest.mark.parametrize("name", SCENES)
def test_core_matches_golden(rt, lib, sim, name):
    g = np.load(os.path.join(GOLDEN, f"scene_{name}.npz"))
    host = rt.HostScene(rt.scenes.REGISTRY[name](**ast.literal_eval(str(g["kwargs"]))), lib)
    r = sim.render(host)
    assert np.array_equal(r["hit_id"], g["ids"])                      # bit-exact primitive ids
    assert np.array_equal(r["hit_t"], g["t"])                         # bit-exact ray parameters
    d = np.abs(r["rgba8"][..., :3].astype(int) - g["rgb8"].astype(int))
    assert d.max() <= 1 and (d == 0).mean() >= 0.999
    assert np.abs(r["pixels"][..., :3] - g["rgb"]).max() < 1e-6      # FP32 texels bound the colour error
    c = r["counters"]
    assert [c["rays_primary"], c["rays_shadow"], c["rays_reflect"]] == list(g["rays"])
    assert (r["rgba8"][..., 3] == 255).all()


def test_filter_never_changes_the_result(rt, lib, sim):
    """FP32 filter + FP64 decide == FP64 everywhere (RT_FLAG_EXACT_ONLY), and the filter does filter."""
    host = rt.HostScene(rt.scenes.office(width=96, height=54), lib)
    blob = sim.flatten(host)
    a = sim.render(host, blob)
    b = sim.render(host, blob, flags=1)
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(a[k], b[k])
    assert a["counters"]["tri_exact"] < 0.5 * a["counters"]["tri_filter"]   # (0.05 on the reference topology)
    assert b["counters"]["tri_exact"] == b["counters"]["tri_filter"]


def test_supersampling_grid(rt, lib, sim):
    g = np.load(os.path.join(GOLDEN, "scene_mixed_small_spp2.npz"))
    host = rt.HostScene(rt.scenes.mixed_small(width=64, height=48), lib)
    r = sim.render(host, spp_grid=2, want_ids=False)
    assert np.abs(r["pixels"][..., :3] - g["rgb"]).max() < 1e-6
    assert np.abs(r["rgba8"][..., :3].astype(int) - g["rgb8"].astype(int)).max() <= 1


def test_tri_filter_is_conservative(sim, orc, oracle_kind):
    """Property: whenever the reference test accepts a triangle with t <= thi, the FP32 filter passes it.
    Adversarial mix: tiny far triangles, grazing rays, hits on edges/vertices, large coordinates."""
    rng = np.random.default_rng(11)
    accepted = passed_miss = misses = 0
    for i in range(6000):
        scale = 10.0 ** rng.uniform(-4, 1)
        centre = rng.uniform(-1, 1, 3) * 10.0 ** rng.uniform(-1, 2.5)
        p = centre + rng.normal(size=(3, 3)) * scale
        dist = scale * 10.0 ** rng.uniform(0, 4)
        mode = i % 4
        if mode == 0:
            w = rng.dirichlet((1, 1, 1))
        elif mode == 1:                                   # on an edge / vertex
            w = np.array([rng.random(), 0.0, 0.0]); w[1] = 1.0 - w[0]
            w = np.roll(w, rng.integers(0, 3))
        elif mode == 2:                                   # just outside an edge
            w = rng.dirichlet((1, 1, 1)); w[rng.integers(0, 3)] = -10.0 ** rng.uniform(-9, -3); w /= w.sum()
        else:
            w = rng.uniform(-0.5, 1.5, 3); w /= w.sum()
        target = w @ p
        n = np.cross(p[1] - p[0], p[2] - p[0]); n /= np.linalg.norm(n) + 1e-300
        dirn = rng.normal(size=3)
        if i % 5 == 0:                                    # grazing
            dirn = dirn - (dirn @ n) * n * (1.0 - 10.0 ** rng.uniform(-5, -1))
        dirn /= np.linalg.norm(dirn)
        o = target - dist * dirn
        hit, out = orc.tri_test(p[0], p[1], p[2], o, dirn, kind=oracle_kind)
        sh, sout = sim.tri_exact(p[0], p[1], p[2], o, dirn)
        assert sh == hit and (not hit or sout[0] == out[0])          # exact stage == reference, bit for bit
        if hit:
            accepted += 1
            assert sim.tri_filter(p[0], p[1], p[2], o, dirn, out[0]), (i, p, o, dirn)       # thi == t (tie)
            assert sim.tri_filter(p[0], p[1], p[2], o, dirn, 1e300)
        elif mode == 3 and dist < 100 * scale and np.abs(centre).max() < 100 * scale:
            # well-conditioned clear misses: here the filter must actually reject
            misses += 1
            passed_miss += sim.tri_filter(p[0], p[1], p[2], o, dirn, 1e300)
    assert accepted > 1500
    assert misses > 100 and passed_miss < 0.25 * misses   # and it is a useful filter, not a pass-through


def test_sphere_filter_is_conservative(sim, orc, oracle_kind):
    """Property: whenever the oracle's sphere test hits at parameter t, the FP32 pre-filter passes for every
    thi >= t (grazing rays, origins inside / on the sphere, far-away origins), and it rejects most clear misses."""
    rng = np.random.default_rng(21)
    hits = misses = rejected = 0
    for i in range(12000):
        s = 10.0 ** rng.uniform(-2, 2)
        c = rng.uniform(-1, 1, 3) * s * 5
        rad = s * 10.0 ** rng.uniform(-2, 0.3)
        o = c + rng.normal(size=3) * rad * 10.0 ** rng.uniform(-1, 1.5)
        if i % 5 == 0:                                   # grazing: aim at a point on the silhouette
            n = np.cross(rng.normal(size=3), o - c); n /= np.linalg.norm(n)
            tgt = c + n * rad * (1.0 + rng.choice([-1e-7, 0.0, 1e-7, -1e-3, 1e-3]))
        else:
            tgt = c + rng.normal(size=3) * rad * rng.uniform(0, 2.0)
        d = tgt - o
        if i % 7 == 0:
            d[rng.integers(0, 3)] = 0.0
        if not d.any():
            continue
        d /= np.linalg.norm(d)
        hit, out7 = orc.sphere_test(c, rad, o, d, kind=oracle_kind)
        t = out7[0]
        if hit:
            hits += 1
            assert sim.sphere_filter(c, rad, o, d, t * (1 + 1e-12)), (i, c, rad, o, d, t)
            assert sim.sphere_filter(c, rad, o, d, 1e300)
        else:
            misses += 1
            rejected += not sim.sphere_filter(c, rad, o, d, 1e300)
    assert hits > 3000 and misses > 2000 and rejected > 0.9 * misses


def test_box_test_is_conservative(sim, orc, oracle_kind):
    """Property: whenever the reference slab test accepts a box, the FP32 test on the outward-rounded box
    does too (for thi = +inf), including axis-parallel rays and flat boxes."""
    rng = np.random.default_rng(5)
    n_acc = 0
    for i in range(6000):
        s = 10.0 ** rng.uniform(-3, 2)
        mn = rng.uniform(-1, 1, 3) * s * 10
        mx = mn + rng.uniform(0, 1, 3) * s
        if i % 6 == 0:
            k = rng.integers(0, 3); mx[k] = mn[k]
        o = rng.uniform(-1, 1, 3) * s * 30
        tgt = mn + (mx - mn) * rng.uniform(-0.2, 1.2, 3)
        d = tgt - o
        if i % 5 == 0:
            d[rng.integers(0, 3)] = 0.0
        if i % 11 == 0:
            d[rng.integers(0, 3)] *= 1e-12
        if not d.any():
            continue
        d /= np.linalg.norm(d)
        hit, _ = orc.aabb_test(mn, mx, o, d, kind=oracle_kind)
        if hit:
            n_acc += 1
            assert sim.box_test(mn, mx, o, d, 1e300), (i, mn, mx, o, d)
            assert sim.box_test(mn, mx, o, d, 1e300, extent_scale=50.0), (i, mn, mx, o, d)
    assert n_acc > 1500


def test_box_test_culls_and_keeps_with_finite_tmax(sim):
    """Property with a finite best distance: a box that contains a point of the ray at parameter t0 must
    be accepted for every thi >= t0 (incl. grazing, near-parallel rays and origins on a box face), and a
    box that is clearly farther than thi must be culled (the test is a useful filter)."""
    rng = np.random.default_rng(11)
    kept = culled = far = 0
    for i in range(8000):
        s = 10.0 ** rng.uniform(-2, 2)
        mn = rng.uniform(-1, 1, 3) * s
        mx = mn + rng.uniform(0, 1, 3) * s * (10.0 ** rng.uniform(-3, 0))
        q = mn + (mx - mn) * rng.uniform(0, 1, 3)          # a point inside the box
        if i % 3 == 0:
            k = rng.integers(0, 3); q[k] = mn[k] if i % 2 else mx[k]   # on a face
        d = rng.normal(size=3)
        if i % 4 == 0:
            d[rng.integers(0, 3)] *= 10.0 ** rng.uniform(-14, -4)     # near-parallel to a slab
        if i % 7 == 0:
            d[rng.integers(0, 3)] = 0.0
        if not d.any():
            continue
        d /= np.linalg.norm(d)
        t0 = 10.0 ** rng.uniform(-3, 2) * s
        o = q - t0 * d
        # the exact point q = o + t0 d lies in the box up to FP64 rounding of o: accept at thi = t0 (1 + 1e-12)
        assert sim.box_test(mn, mx, o, d, t0 * (1 + 1e-12)), (i, mn, mx, o, d, t0)
        assert sim.box_test(mn, mx, o, d, t0 * (1 + 1e-12), extent_scale=20.0)
        kept += 1
        # entry distance is at least t0 - diag; a best distance well before that must cull the box
        diag = np.linalg.norm(mx - mn)
        if t0 > 4 * diag + 1e-3 * s:
            far += 1
            culled += not sim.box_test(mn, mx, o, d, 0.5 * (t0 - diag))
    assert kept > 7000 and far > 1000 and culled > 0.9 * far


def test_tie_break_is_visit_order_independent(rt, lib, sim, orc, oracle_kind):
    """Coplanar duplicated triangles: the lower BVH-order index must win, as in the reference's
    left-first DFS with strict '<' (mybvh.cpp:169)."""
    pos = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0.0]])
    quad = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    idx = np.concatenate([quad] * 6, 0)                   # 6 exact copies of the same two triangles
    shift = np.concatenate([pos + np.array([3.0 * k, 0, 0]) for k in range(4)], 0)
    idx2 = np.concatenate([quad + 4 * k for k in range(4)] * 3, 0)
    for p, ix in ((pos, idx), (shift, idx2)):
        desc = rt.SceneDesc("dups", [rt.MeshDesc(p, ix, rt.material((1, 1, 1), (1, 1, 1), (0, 0, 0), 1.0))], [],
                            [((0, 0, 5), (1, 1, 1))],
                            dict(eye=(p[:, 0].mean(), 0.3, 6), center=(p[:, 0].mean(), 0, 0), up=(0, 1, 0), fovy=60,
                                 width=64, height=24))
        host = rt.HostScene(desc, lib)
        r = sim.render(host)
        ref = orc.OracleScene(desc, oracle_kind).render()
        assert np.array_equal(r["hit_id"], ref["ids"]) and (ref["ids"] >= 0).sum() > 40


def test_axis_parallel_rays_reproduce_reference_nan_quirk(rt, lib, sim, orc, oracle_kind):
    """Centre pixel column of a symmetric view over a grid line: direction.x == 0 exactly and origin.x equals
    box faces, where the reference's slab test evaluates 0/0 = NaN and rejects the box whose bb_max.x equals
    origin.x (mybvh.cpp:99-135).  The ids must still be bit-exact (rt_types.h NodeMaxX)."""
    desc = rt.scenes.REGISTRY["scatter"](width=64, height=36, n_objects=40)
    host = rt.HostScene(desc, lib)
    r = sim.render(host)
    ref = orc.OracleScene(desc, oracle_kind).render()
    d = np.asarray(desc.camera["eye"], float)
    assert np.array_equal(r["hit_id"], ref["ids"]) and np.array_equal(r["hit_t"], ref["t"])
    # the quirk is actually exercised: without it the lower-index twin triangle would win on that column
    blob_ids = r["hit_id"][:, 32]
    assert (blob_ids >= 0).any()


@pytest.mark.parametrize("scene,kw", [("mixed_small", dict(width=96, height=72)),
                                      ("scatter", dict(width=64, height=36, n_objects=40)),
                                      ("rings", dict(width=70, height=50, nu=48, nv=16)),
                                      ("molecule", dict(width=90, height=90, n_spheres=150)),
                                      ("textured", dict(width=96, height=54, tex_size=64, slices=24, stacks=12))])
def test_frame_does_not_depend_on_the_tree(rt, lib, sim, scene, kw):
    """The library's own SAH traversal tree and the reference's median-split topology give bit-identical frames
    (ids, distances, FP64 pixels): every hit is decided by the exact test, ties by the reference's primitive index,
    and d.x == 0 rays keep walking the reference tree (the scatter view has a column of them)."""
    host = rt.HostScene(rt.scenes.REGISTRY[scene](**kw), lib)
    a = sim.render(host, sim.flatten(host, rt.capi.RT_SCENE_REFERENCE_TOPOLOGY))
    b = sim.render(host, sim.flatten(host))
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(a[k], b[k]), k
    if host.num_triangles:
        assert b["counters"]["nodes_visited"] < a["counters"]["nodes_visited"]
    else:       # analytic-only scene: the tree over the spheres replaces most of the per-ray object tests
        assert b["counters"]["object_tests"] < 0.5 * a["counters"]["object_tests"]


@pytest.mark.parametrize("scene,kw", [("office", dict(width=128, height=72)),
                                      ("scatter", dict(width=81, height=61, n_objects=40)),
                                      ("rings", dict(width=70, height=50, nu=48, nv=16)),
                                      ("textured", dict(width=96, height=54, tex_size=64, slices=24, stacks=12)),
                                      ("molecule", dict(width=90, height=90, n_spheres=150)),     # object tree only
                                      ("objtree", dict(width=96, height=72))])                    # object tree + triangles
def test_wide_tree_gives_the_same_frame_in_fewer_steps(rt, lib, sim, scene, kw):
    """The 4-wide twin of the traversal tree (Node4Rec, trav_step4) against the binary tree it is collapsed from
    (RT_FLAG_BINARY_NODES): bit-identical ids, distances and FP64 pixels, fewer node visits; and the FP32 filter
    stays a pure filter on the wide tree (RT_FLAG_EXACT_ONLY)."""
    host = rt.HostScene(rt.scenes.REGISTRY[scene](**kw), lib)
    blob = sim.flatten(host)
    wide = sim.render(host, blob)
    binary = sim.render(host, blob, flags=rt.capi.RT_FLAG_BINARY_NODES)
    exact = sim.render(host, blob, flags=rt.capi.RT_FLAG_EXACT_ONLY)
    assert wide["counters"]["nodes4_visited"] > 0 and binary["counters"]["nodes4_visited"] == 0
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(wide[k], binary[k]), k
        assert np.array_equal(wide[k], exact[k]), k
    ratio = 1.0 if scene == "objtree" else 0.75          # objtree: a dozen spheres and a few triangles, tiny trees
    assert wide["counters"]["nodes_visited"] < ratio * binary["counters"]["nodes_visited"]
    for k in ("rays_primary", "rays_shadow", "rays_reflect", "shaded_hits", "shadow_occluded"):
        assert wide["counters"][k] == binary["counters"][k], k


def test_pow_specular_matches_libm(sim):
    """pow_specular (own ln/exp for the Phong term, mytracer.cpp:525-533) against libm's pow on its domain: base in
    [0, 1] (+ rounding above 1), positive exponent.  Relative error <= 1e-13 wherever the result is not negligible
    (> 1e-200), exact special values, and values outside the fast domain fall back to pow() itself."""
    import math
    rng = np.random.default_rng(3)
    worst = 0.0
    for i in range(20000):
        mode = i % 5
        x = [rng.random(), 1.0 - rng.random() * 1e-3, 10.0 ** (-rng.random() * 20), 1.0 - rng.random() ** 2,
             1.0 + (rng.random() - 0.5) * 1e-15][mode]
        y = [0.3, 1, 2, 5, 10, 32, 50, 100, 128, 150, 200, 500][i % 12] * (0.5 + rng.random())
        want = math.pow(x, y)
        got = sim.pow_specular(x, y)
        if want > 1e-200:
            worst = max(worst, abs(got - want) / want)
        else:
            assert abs(got - want) <= 1e-200
    assert worst < 1e-13
    assert sim.pow_specular(0.0, 7.0) == 0.0 and sim.pow_specular(1.0, 123.0) == 1.0 and sim.pow_specular(0.5, 2.0) == 0.25
    for x, y in ((1e-310, 0.5), (0.3, 1e5), (2.5, 3.0), (0.5, -2.0), (0.0, 0.0)):      # outside the fast domain
        assert sim.pow_specular(x, y) == math.pow(x, y)


def test_object_tree_keeps_list_order_semantics(rt, lib, sim, orc, oracle_kind):
    """Spheres go into their own tree when there are >= 8 of them; the reference loops over objects_ in list order
    with a strict '<'.  Duplicated spheres (equal t: the lower index must win), a sphere around the camera (ray
    origins inside a box and inside a sphere), planes between them in the list, and triangles behind: ids,
    distances and pixels equal the oracle's and the linear loop's."""
    desc = rt.scenes.objtree()
    objs = desc.objects
    host = rt.HostScene(desc, lib)
    tree = sim.render(host, sim.flatten(host))
    lin = sim.render(host, sim.flatten(host, rt.capi.RT_SCENE_REFERENCE_TOPOLOGY))
    ref = orc.OracleScene(desc, oracle_kind).render()
    for k in ("hit_id", "hit_t", "rgba8", "pixels"):
        assert np.array_equal(tree[k], lin[k]), k
    assert np.array_equal(tree["hit_id"], ref["ids"]) and np.array_equal(tree["hit_t"], ref["t"])
    assert np.abs(tree["pixels"][..., :3] - ref["rgb"]).max() < 1e-9
    n = desc.num_triangles
    assert (tree["hit_id"] == n + len(objs) - 1).all()             # every primary ray starts inside the last sphere...
    assert tree["counters"]["nodes_visited"] > 0 and tree["counters"]["object_tests"] < lin["counters"]["object_tests"]
    # the same scene seen from outside: duplicates and shared geometry decide visible pixels
    desc = rt.scenes.objtree(camera_inside=False)
    host = rt.HostScene(desc, lib)
    tree = sim.render(host, sim.flatten(host))
    ref = orc.OracleScene(desc, oracle_kind).render()
    assert np.array_equal(tree["hit_id"], ref["ids"]) and np.array_equal(tree["hit_t"], ref["t"])
    assert np.abs(tree["pixels"][..., :3] - ref["rgb"]).max() < 1e-9
    ids = set(np.unique(tree["hit_id"]).tolist())
    assert n + 1 in ids and n + 15 not in ids and n + 5 in ids and n + 16 not in ids   # the first of two equal hits is kept


@pytest.mark.parametrize("name,kw,spp,tile", [
    ("mixed_small", dict(width=120, height=90), 1, (0, 1, 1)),
    ("mixed_small", dict(width=61, height=45), 2, (0, 1, 1)),          # sub-pixel samples = passes over the same paths
    ("office", dict(width=160, height=90), 1, (1, 3, 4)),              # a rank's tile share
    ("scatter", dict(width=96, height=54, n_objects=100), 1, (0, 1, 1)),   # d.x == 0 rays (reference tree, NaN quirk)
    ("objtree", dict(width=96, height=72), 1, (0, 1, 1)),              # tree over analytic spheres + plane + triangles
    ("textured", dict(width=96, height=54, tex_size=64, slices=32, stacks=16), 1, (0, 1, 1)),
])
def test_wavefront_schedule_renders_the_same_frame(rt, lib, sim, name, kw, spp, tile):
    """The wavefront stage functions (rt_wavefront.cuh: wf_ray_start / traverse / wf_ray_finish, wf_shade — what the
    TRACE and SHADE kernels call) in the launch order of rt_api.cu against the per-lane state machine of the persistent
    kernel: bit-identical pixels (8-bit and FP64), hit ids, ray parameters and every counter (same rays, same tests)."""
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    blob = sim.flatten(host)
    a = sim.render(host, blob, spp_grid=spp, tile=tile, want_ids=spp == 1)
    b = sim.render(host, blob, spp_grid=spp, tile=tile, want_ids=spp == 1, wavefront=True)
    for k in ("rgba8", "pixels", "hit_id", "hit_t"):
        assert np.array_equal(a[k], b[k]), k
    assert a["counters"] == b["counters"]


@pytest.mark.parametrize("name,kw", [
    ("office", dict(width=240, height=135)),
    ("rings", dict(width=140, height=100)),
    ("textured", dict(width=128, height=72, tex_size=64)),
    ("scatter", dict(width=128, height=72, n_objects=200)),           # d.x == 0 rays stay on the reference tree
    ("mixed_small", dict(width=120, height=90)),                       # analytic objects tested linearly + meshes
])
def test_eight_wide_tree_gives_the_same_frame_in_fewer_steps(rt, lib, sim, name, kw):
    """The 8-wide twin of the triangle tree (Node8Rec, trav_step8, RT_FLAG_NODES8) against the 4-wide one it shares its
    binary SAH tree with: bit-identical pixels, ids and ray parameters, the same rays, at most 70 % of the node visits."""
    from my_raytracer_b200 import blobfmt, capi
    host = rt.HostScene(rt.scenes.REGISTRY[name](**kw), lib)
    blob = sim.flatten(host)
    h = blobfmt.header(blob)
    assert 0 < h["num_nodes8"] < h["num_nodes4"] and h["off_nodes8"] % 32 == 0
    a = sim.render(host, blob)
    b = sim.render(host, blob, flags=capi.RT_FLAG_NODES8)
    for k in ("rgba8", "pixels", "hit_id", "hit_t"):
        assert np.array_equal(a[k], b[k]), k
    ca, cb = a["counters"], b["counters"]
    for k in ("rays_primary", "rays_shadow", "rays_reflect", "shaded_hits", "shadow_occluded"):
        assert ca[k] == cb[k], k
    assert cb["nodes8_visited"] > 0 and ca["nodes8_visited"] == 0
    assert cb["nodes_visited"] < 0.8 * ca["nodes_visited"]
