"""Oracle pinning (CPU): the port restatement against the golden vectors produced by the reference's
own compiled sources (oracle/_ref), and — where oracle/_ref exists — against it live, bit for bit."""
import ast
import os

import numpy as np
import pytest

from conftest import GOLDEN

SCENES = ["mixed_small", "spheres", "office", "rings", "molecule", "textured"]


def _kat():
    return np.load(os.path.join(GOLDEN, "kat_primitives.npz"))


def _kinds(orc):
    return [k for k in ("port", "reference") if orc.available(k)]


def test_port_library_present(orc):
    assert orc.available("port"), "oracle/librt_oracle.so missing: run __graft_entry__.build()"


def test_kat_triangle(orc):
    g = _kat()
    for kind in _kinds(orc):
        for row, hit, out in zip(g["tri_in"], g["tri_hit"], g["tri_out"]):
            h, o = orc.tri_test(row[0:3], row[3:6], row[6:9], row[9:12], row[12:15], kind=kind)
            assert h == bool(hit)
            if h:
                assert np.array_equal(o, out), (kind, row)   # bit-exact t and hit point
    assert g["tri_hit"].sum() > 50 and (~g["tri_hit"]).sum() > 50


def test_kat_triangle_edge_cases(orc):
    """closed [0,1] barycentric interval, t <= 1e-5 rejected, |S| < 1e-10 rejected (mymesh.cpp:197-215)."""
    p0, p1, p2 = (0, 0, 0), (1, 0, 0), (0, 1, 0)
    for kind in _kinds(orc):
        assert orc.tri_test(p0, p1, p2, (0.5, 0.0, 1.0), (0, 0, -1), kind=kind)[0]       # on an edge
        assert orc.tri_test(p0, p1, p2, (0.0, 0.0, 1.0), (0, 0, -1), kind=kind)[0]       # on a vertex
        assert not orc.tri_test(p0, p1, p2, (0.25, 0.25, 1e-5), (0, 0, -1), kind=kind)[0]  # t == 1e-5
        assert orc.tri_test(p0, p1, p2, (0.25, 0.25, 2e-5), (0, 0, -1), kind=kind)[0]
        assert not orc.tri_test(p0, p1, p2, (0.25, 0.25, -1.0), (0, 0, -1), kind=kind)[0]  # behind
        assert not orc.tri_test(p0, p1, p2, (0.25, 0.25, 1.0), (1, 0, 0), kind=kind)[0]    # parallel: S == 0
        tiny = [tuple(1e-6 * np.array(p)) for p in (p0, p1, p2)]
        assert not orc.tri_test(*tiny, (2e-7, 2e-7, 1.0), (0, 0, -1), kind=kind)[0]        # |S| = 1e-12


def test_kat_aabb(orc):
    g = _kat()
    for kind in _kinds(orc):
        for row, hit, tmin in zip(g["box_in"], g["box_hit"], g["box_tmin"]):
            h, t = orc.aabb_test(row[0:3], row[3:6], row[6:9], row[9:12], kind=kind)
            assert h == bool(hit)
            if h:
                assert t == tmin or (np.isnan(t) and np.isnan(tmin))
    # 0/0 -> NaN makes every reject comparison false: the box is accepted (mybvh.cpp:101-134)
    assert g["box_hit"][-1]


def test_kat_plane_sphere(orc):
    g = _kat()
    for kind in _kinds(orc):
        for row, hit, out in zip(g["plane_in"], g["plane_hit"], g["plane_out"]):
            h, o = orc.plane_test(row[0:3], row[3:6], row[6:9], row[9:12], kind=kind)
            assert h == bool(hit) and (not h or np.array_equal(o, out))
        for row, hit, out in zip(g["sphere_in"], g["sphere_hit"], g["sphere_out"]):
            h, o = orc.sphere_test(row[0:3], row[3], row[4:7], row[7:10], kind=kind)
            assert h == bool(hit) and (not h or np.array_equal(o, out))
    assert not g["plane_hit"][-2] and not g["plane_hit"][-1]     # parallel ray; t == 1e-5
    assert g["sphere_hit"][-2]                                   # origin inside -> far root


@pytest.mark.parametrize("name", SCENES)
def test_scene_against_golden(rt, orc, name):
    g = np.load(os.path.join(GOLDEN, f"scene_{name}.npz"))
    kw = ast.literal_eval(str(g["kwargs"]))
    desc = rt.scenes.REGISTRY[name](**kw)
    for kind in _kinds(orc):
        o = orc.OracleScene(desc, kind)
        r = o.render()
        assert np.array_equal(r["ids"], g["ids"]), kind
        assert np.array_equal(r["t"], g["t"]), kind
        assert np.array_equal(r["rgb"], g["rgb"]), kind          # bit-exact doubles
        assert [r["rays"][k] for k in ("primary", "shadow", "reflect")] == list(g["rays"])
        b = o.bvh()
        assert np.array_equal(o.tri_order(), g["tri_order"])
        assert np.array_equal(b["left"], g["bvh_left"]) and np.array_equal(b["first"], g["bvh_first"])
        assert np.array_equal(b["count"], g["bvh_count"])
        assert np.array_equal(b["bbmin"], g["bvh_min"]) and np.array_equal(b["bbmax"], g["bvh_max"])
        assert np.array_equal(o.camera(), g["camera"])


def test_supersampled_against_golden(rt, orc):
    g = np.load(os.path.join(GOLDEN, "scene_mixed_small_spp2.npz"))
    desc = rt.scenes.mixed_small(width=64, height=48)
    for kind in _kinds(orc):
        r = orc.OracleScene(desc, kind).render(want_ids=False, spp_grid=2)
        assert np.array_equal(r["rgb"], g["rgb"])


def test_port_equals_reference_live(rt, orc):
    """Where the reference sources could be compiled (this container), the port must agree bit for bit
    on arbitrary rays, normals and lighting."""
    if not orc.available("reference"):
        pytest.skip("oracle/_ref not built on this machine (no /root/reference)")
    desc = rt.scenes.mixed_small(width=32, height=24, seed=3)
    a, b = orc.OracleScene(desc, "reference"), orc.OracleScene(desc, "port")
    rng = np.random.default_rng(0)
    o = rng.uniform(-4, 4, (3000, 3)) + np.array([0, 3, 0])
    d = rng.normal(size=(3000, 3))
    ra, rb = a.trace_rays(o, d), b.trace_rays(o, d)
    for x, y in zip(ra, rb):
        assert np.array_equal(x, y)
    for m in range(len(desc.meshes)):
        for x, y in zip(a.normals(m), b.normals(m)):
            assert np.array_equal(x, y, equal_nan=True)   # NaN: normalize(0) of degenerate soup triangles
    mat = rt.material((0.1, 0.2, 0.3), (0.5, 0.6, 0.7), (0.8, 0.8, 0.8), 0.0, 0.0)   # shininess 0: pow(0,0) == 1
    la = a.lighting((0, 1, 3), (0, -1, 0), (0, 0, 1), mat)
    assert np.array_equal(la, b.lighting((0, 1, 3), (0, -1, 0), (0, 0, 1), mat))
    assert np.array_equal(a.texture(5, 0.2, 0.3, 0.5), b.texture(5, 0.2, 0.3, 0.5))


def test_phong_quirks(rt, orc, oracle_kind):
    """Appendix B items 4 and 6: back faces get diff = 0 and the specular base 0 is still raised to
    shininess (pow(0,0) == 1)."""
    desc = rt.scenes.spheres(width=8, height=8)
    o = orc.OracleScene(desc, oracle_kind)
    ns = rt.material((0.0, 0, 0), (1, 1, 1), (1, 1, 1), 0.0, 0.0, shadowable=False)
    c = o.lighting((0, 50, 0), (0, 1, 0), (0, 1, 0), ns)     # every light is behind the surface
    amb = np.array(desc.ambience) * 0.0
    assert np.allclose(c, amb + sum(np.array(l[1]) for l in desc.lights))   # kd*0 + ks*pow(0,0)=ks*1 per light
