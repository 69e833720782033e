"""Generates the committed golden fixtures from the REFERENCE oracle (oracle/_ref: the reference's own
CPU sources compiled from /root/reference).  Run in the build container only:

    python tests/golden/make_golden.py

The fixtures let the port oracle, the host test build and the CUDA path be checked on machines where
/root/reference (and therefore oracle/_ref) does not exist.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as orc  # noqa: E402
import rtpkg  # noqa: E402

rt = rtpkg.load()
KIND = "reference"
assert orc.available(KIND), "oracle/_ref is not built: run `make -C oracle ref` where /root/reference exists"


def kat_cases(seed=7, n=400):
    """Known-answer cases for single primitives, including the edge cases SURVEY 8c lists."""
    rng = np.random.default_rng(seed)
    tri, box, plane, sphere = [], [], [], []
    for i in range(n):
        p = rng.uniform(-2, 2, (3, 3))
        o = rng.uniform(-4, 4, 3)
        target = p.T @ rng.dirichlet((1, 1, 1)) if i % 3 else rng.uniform(-3, 3, 3)
        d = target - o
        d /= np.linalg.norm(d)
        tri.append((p[0], p[1], p[2], o, d))
    # edge / vertex grazes, zero direction components, t at the 1e-5 threshold, degenerate triangle
    p0, p1, p2 = np.array([0., 0, 0]), np.array([1., 0, 0]), np.array([0., 1, 0])
    for o, d in [((0.5, 0.0, 1.0), (0, 0, -1)), ((0.0, 0.0, 1.0), (0, 0, -1)), ((1.0, 0.0, 1.0), (0, 0, -1)),
                 ((0.5, 0.5, 1.0), (0, 0, -1)), ((0.25, 0.25, 1e-5), (0, 0, -1)), ((0.25, 0.25, 1.0000001e-5), (0, 0, -1)),
                 ((0.25, 0.25, 2e-5), (0, 0, -1)), ((0.25, 0.25, -1.0), (0, 0, -1)), ((0.25, 0.25, 1.0), (1, 0, 0)),
                 ((0.3, 0.3, 1.0), (0, 0, 1))]:
        tri.append((p0, p1, p2, np.array(o, float), np.array(d, float)))
    tri.append((p0, p0 * 1.0, p2, np.array([0.1, 0.1, 1.0]), np.array([0., 0, -1])))            # degenerate
    tri.append((p0 * 1e-6, p1 * 1e-6, p2 * 1e-6, np.array([2e-7, 2e-7, 1.0]), np.array([0., 0, -1])))  # |S| < 1e-10
    for i in range(n):
        mn = rng.uniform(-2, 0, 3); mx = mn + rng.uniform(0, 2, 3)
        o = rng.uniform(-4, 4, 3)
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        if i % 5 == 0:
            d[rng.integers(0, 3)] = 0.0   # axis-parallel ray: division by zero -> +-inf
            d /= np.linalg.norm(d)
        if i % 7 == 0:
            mx[rng.integers(0, 3)] = mn[rng.integers(0, 3)]  # flat box
        box.append((mn, mx, o, d))
    box.append((np.zeros(3), np.ones(3), np.array([0.0, 0.5, 0.5]), np.array([0.0, 1.0, 0.0])))   # 0/0 -> NaN accepts
    for i in range(n // 2):
        c = rng.uniform(-1, 1, 3); nrm = rng.normal(size=3); nrm /= np.linalg.norm(nrm)
        o = rng.uniform(-3, 3, 3); d = rng.normal(size=3); d /= np.linalg.norm(d)
        plane.append((c, nrm, o, d))
        sphere.append((c, float(rng.uniform(0.2, 1.5)), o, d))
    plane.append((np.zeros(3), np.array([0., 1, 0]), np.array([0., 1, 0]), np.array([1., 0, 0])))        # parallel
    plane.append((np.zeros(3), np.array([0., 1, 0]), np.array([0., 1e-5, 0]), np.array([0., -1, 0])))    # t == 1e-5
    sphere.append((np.zeros(3), 1.0, np.array([0., 0, 0.5]), np.array([0., 0, 1])))                       # origin inside
    sphere.append((np.zeros(3), 1.0, np.array([0., 1.0, -3]), np.array([0., 0, 1])))                      # tangent
    return tri, box, plane, sphere


def main():
    tri, box, plane, sphere = kat_cases()
    out = {}
    out["tri_in"] = np.array([np.concatenate(c) for c in tri])
    r = [orc.tri_test(*c, kind=KIND) for c in tri]
    out["tri_hit"] = np.array([h for h, _ in r]); out["tri_out"] = np.array([o for _, o in r])
    out["box_in"] = np.array([np.concatenate(c) for c in box])
    r = [orc.aabb_test(*c, kind=KIND) for c in box]
    out["box_hit"] = np.array([h for h, _ in r]); out["box_tmin"] = np.array([t for _, t in r])
    out["plane_in"] = np.array([np.concatenate(c) for c in plane])
    r = [orc.plane_test(*c, kind=KIND) for c in plane]
    out["plane_hit"] = np.array([h for h, _ in r]); out["plane_out"] = np.array([o for _, o in r])
    out["sphere_in"] = np.array([np.concatenate([c[0], [c[1]], c[2], c[3]]) for c in sphere])
    r = [orc.sphere_test(*c, kind=KIND) for c in sphere]
    out["sphere_hit"] = np.array([h for h, _ in r]); out["sphere_out"] = np.array([o for _, o in r])
    np.savez_compressed(os.path.join(HERE, "kat_primitives.npz"), **out)

    # small scene renders: hit ids, t, 8-bit pixels, ray counts, BVH arrays
    for name, kw in [("mixed_small", dict(width=64, height=48)), ("spheres", dict(width=64, height=64)),
                     ("office", dict(width=64, height=36)), ("rings", dict(width=56, height=40, nu=48, nv=12)),
                     ("molecule", dict(width=48, height=48, n_spheres=40)),
                     ("textured", dict(width=64, height=36, tex_size=64, slices=32, stacks=16))]:
        desc = rt.scenes.REGISTRY[name](**kw)
        o = orc.OracleScene(desc, KIND)
        r = o.render()
        b = o.bvh()
        np.savez_compressed(os.path.join(HERE, f"scene_{name}.npz"), kwargs=np.array(repr(kw)),
                            rgb8=rt.quantize_rgb8(r["rgb"]), rgb=r["rgb"], ids=r["ids"], t=r["t"],
                            rays=np.array([r["rays"]["primary"], r["rays"]["shadow"], r["rays"]["reflect"]]),
                            tri_order=o.tri_order(), bvh_left=b["left"], bvh_first=b["first"], bvh_count=b["count"],
                            bvh_min=b["bbmin"], bvh_max=b["bbmax"], camera=o.camera())
        if name == "mixed_small":
            r4 = o.render(want_ids=False, spp_grid=2)
            np.savez_compressed(os.path.join(HERE, "scene_mixed_small_spp2.npz"), rgb8=rt.quantize_rgb8(r4["rgb"]), rgb=r4["rgb"])
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
