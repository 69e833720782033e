"""Procedural scenes standing in for the course's absent assets (BASELINE.json configs).

Same named shapes, stated triangle counts, deterministic: no file I/O, numpy only.  Each
function returns a ``SceneDesc`` that both the product (HostScene -> C ABI) and the oracle
(OracleScene) consume, so the two sides always see identical doubles.
"""
from __future__ import annotations

import numpy as np

from .scene import MeshDesc, SceneDesc, material

# --------------------------------------------------------------------------------------------
# mesh builders
# --------------------------------------------------------------------------------------------


def grid_quad(origin, eu, ev, nu, nv, with_uv=False, flip=False):
    """(nu x nv) quads -> 2*nu*nv triangles spanning origin + s*eu + t*ev, s,t in [0,1].
    Face normals (cross(p1 - p0, p2 - p0), mymesh.cpp:116) point along eu x ev, or ev x eu with flip=True:
    the reference never turns a normal toward the viewer, so the winding decides which side is lit."""
    origin, eu, ev = (np.asarray(a, np.float64) for a in (origin, eu, ev))
    s = np.linspace(0.0, 1.0, nu + 1)
    t = np.linspace(0.0, 1.0, nv + 1)
    S, T = np.meshgrid(s, t, indexing="xy")                     # (nv+1, nu+1)
    pos = origin[None, :] + S.reshape(-1, 1) * eu[None, :] + T.reshape(-1, 1) * ev[None, :]
    j, i = np.meshgrid(np.arange(nv), np.arange(nu), indexing="ij")
    a = (j * (nu + 1) + i).reshape(-1)
    b = a + 1
    c = a + (nu + 1)
    d = c + 1
    idx = np.concatenate([np.stack([a, b, d], 1), np.stack([a, d, c], 1)], 0).astype(np.int32)
    if flip:
        idx = np.ascontiguousarray(idx[:, [0, 2, 1]])
    if with_uv:
        return pos, idx, S.reshape(-1).copy(), T.reshape(-1).copy()
    return pos, idx


def merge(parts):
    """[(pos, idx), ...] -> one (pos, idx)."""
    ps, is_, base = [], [], 0
    for p, i in parts:
        ps.append(p)
        is_.append(i + base)
        base += p.shape[0]
    return np.concatenate(ps, 0), np.concatenate(is_, 0).astype(np.int32)


def box(lo, hi, n=1):
    """Axis-aligned box, each face an n x n grid, outward-facing winding."""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    d = hi - lo
    ex, ey, ez = np.array([d[0], 0, 0]), np.array([0, d[1], 0]), np.array([0, 0, d[2]])
    faces = [
        grid_quad(lo, ez, ey, n, n),                      # -x
        grid_quad(lo + ex, ey, ez, n, n),                 # +x
        grid_quad(lo, ex, ez, n, n),                      # -y
        grid_quad(lo + ey, ez, ex, n, n),                 # +y
        grid_quad(lo, ey, ex, n, n),                      # -z
        grid_quad(lo + ez, ex, ey, n, n),                 # +z
    ]
    return merge(faces)


def uv_sphere(center, radius, nslices, nstacks, with_uv=False, displace=None):
    """Closed UV sphere: 2*nslices*(nstacks-1) triangles.  displace(theta, phi) -> radius factor."""
    center = np.asarray(center, np.float64)
    th = np.linspace(0.0, 2.0 * np.pi, nslices + 1)          # seam duplicated (for UVs)
    ph = np.linspace(0.0, np.pi, nstacks + 1)
    TH, PH = np.meshgrid(th, ph, indexing="xy")               # (nstacks+1, nslices+1)
    r = radius * (displace(TH, PH) if displace is not None else 1.0)
    pos = np.stack([r * np.sin(PH) * np.cos(TH), r * np.cos(PH), r * np.sin(PH) * np.sin(TH)], -1).reshape(-1, 3)
    pos = pos + center[None, :]
    tris = []
    w = nslices + 1
    for j in range(nstacks):
        for i in range(nslices):
            a, b, c, d = j * w + i, j * w + i + 1, (j + 1) * w + i, (j + 1) * w + i + 1
            if j != 0:
                tris.append((a, b, d))          # outward-facing winding
            if j != nstacks - 1:
                tris.append((a, d, c))
    idx = np.asarray(tris, np.int32)
    if with_uv:
        return pos, idx, (TH / (2.0 * np.pi)).reshape(-1).copy(), (1.0 - PH / np.pi).reshape(-1).copy()
    return pos, idx


def cylinder(base, radius, height, nseg, nstack):
    """Open-ended cylinder side along +y plus two cap fans: 2*nseg*nstack + 2*nseg triangles."""
    base = np.asarray(base, np.float64)
    th = np.linspace(0.0, 2.0 * np.pi, nseg, endpoint=False)
    ys = np.linspace(0.0, height, nstack + 1)
    ring = np.stack([radius * np.cos(th), np.zeros(nseg), radius * np.sin(th)], -1)
    pos = np.concatenate([ring + np.array([0, y, 0]) for y in ys], 0) + base[None, :]
    tris = []
    for j in range(nstack):
        for i in range(nseg):
            a, b = j * nseg + i, j * nseg + (i + 1) % nseg
            c, d = a + nseg, b + nseg
            tris += [(a, c, b), (b, c, d)]
    nb = pos.shape[0]
    pos = np.concatenate([pos, base[None, :], (base + np.array([0, height, 0]))[None, :]], 0)
    top0 = nstack * nseg
    for i in range(nseg):
        tris.append((nb, i, (i + 1) % nseg))
        tris.append((nb + 1, top0 + (i + 1) % nseg, top0 + i))
    return pos, np.asarray(tris, np.int32)


def disc_slab(center, rx, rz, thickness, nseg, nring):
    """Elliptical slab (table top): top + bottom ring grids and a rim."""
    center = np.asarray(center, np.float64)
    th = np.linspace(0.0, 2.0 * np.pi, nseg, endpoint=False)
    parts = []
    for side, y in ((1, thickness * 0.5), (-1, -thickness * 0.5)):
        rings = [np.array([[0.0, y, 0.0]])]
        for k in range(1, nring + 1):
            f = k / nring
            rings.append(np.stack([f * rx * np.cos(th), np.full(nseg, y), f * rz * np.sin(th)], -1))
        pos = np.concatenate(rings, 0)
        tris = []
        for i in range(nseg):
            a, b = 1 + i, 1 + (i + 1) % nseg
            tris.append((0, b, a) if side > 0 else (0, a, b))
        for k in range(1, nring):
            o0, o1 = 1 + (k - 1) * nseg, 1 + k * nseg
            for i in range(nseg):
                a, b = o0 + i, o0 + (i + 1) % nseg
                c, d = o1 + i, o1 + (i + 1) % nseg
                tris += ([(a, b, d), (a, d, c)] if side > 0 else [(a, d, b), (a, c, d)])
        parts.append((pos + center[None, :], np.asarray(tris, np.int32)))
    top = np.stack([rx * np.cos(th), np.full(nseg, thickness * 0.5), rz * np.sin(th)], -1)
    bot = top - np.array([0, thickness, 0])
    pos = np.concatenate([top, bot], 0) + center[None, :]
    tris = []
    for i in range(nseg):
        a, b = i, (i + 1) % nseg
        tris += [(a, b, nseg + b), (a, nseg + b, nseg + a)]
    parts.append((pos, np.asarray(tris, np.int32)))
    return merge(parts)


def torus(center, R, r, nu, nv, axis=1, with_uv=False):
    """Torus with nu x nv quads = 2*nu*nv triangles around the given axis."""
    center = np.asarray(center, np.float64)
    u = np.linspace(0.0, 2.0 * np.pi, nu, endpoint=False)
    v = np.linspace(0.0, 2.0 * np.pi, nv, endpoint=False)
    U, V = np.meshgrid(u, v, indexing="ij")
    x = (R + r * np.cos(V)) * np.cos(U)
    z = (R + r * np.cos(V)) * np.sin(U)
    y = r * np.sin(V)
    p = np.stack([x, y, z], -1).reshape(-1, 3)
    if axis == 0:
        p = p[:, [1, 0, 2]]
    elif axis == 2:
        p = p[:, [0, 2, 1]]
    tris = []
    for i in range(nu):
        for j in range(nv):
            a, b = i * nv + j, i * nv + (j + 1) % nv
            c, d = ((i + 1) % nu) * nv + j, ((i + 1) % nu) * nv + (j + 1) % nv
            tris += [(a, b, d), (a, d, c)]
    return p + center[None, :], np.asarray(tris, np.int32)


def value_noise_texture(size, seed, base=(0.5, 0.5, 0.5), amp=0.5, octaves=4):
    """Procedural RGB texture in [0,1], quantised to k/255 like an 8-bit PNG."""
    rng = np.random.default_rng(seed)
    img = np.zeros((size, size, 3))
    for o in range(octaves):
        n = 4 * 2 ** o
        g = rng.random((n + 1, n + 1, 3))
        xs = np.linspace(0, n, size, endpoint=False)
        i0 = np.floor(xs).astype(int)
        f = xs - i0
        f = f * f * (3 - 2 * f)
        a = g[i0][:, i0]; b = g[i0][:, i0 + 1]; c = g[i0 + 1][:, i0]; d = g[i0 + 1][:, i0 + 1]
        fy, fx = f[:, None, None], f[None, :, None]
        img += ((a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy) / 2 ** o
    img = img / img.max()
    img = np.asarray(base)[None, None, :] * (1 - amp) + amp * img
    return np.round(np.clip(img, 0, 1) * 255.0) / 255.0


# --------------------------------------------------------------------------------------------
# config 1 — o_01_spheres
# --------------------------------------------------------------------------------------------

def spheres(width=500, height=500):
    """3 analytic spheres on a ground plane, Phong + shadows + mirror reflections (N = 0)."""
    red = material((0.2, 0.0, 0.0), (0.8, 0.1, 0.1), (0.9, 0.9, 0.9), 100.0, 0.3)
    green = material((0.0, 0.2, 0.0), (0.1, 0.8, 0.1), (0.9, 0.9, 0.9), 200.0, 0.2)
    blue = material((0.0, 0.0, 0.2), (0.1, 0.1, 0.8), (0.9, 0.9, 0.9), 50.0, 0.3)
    ground = material((0.2, 0.2, 0.2), (0.6, 0.6, 0.6), (0.2, 0.2, 0.2), 10.0, 0.2)
    objs = [
        ("plane", (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), ground),
        ("sphere", (0.0, 3.0, -2.0), 3.0, blue),
        ("sphere", (-3.5, 1.0, 2.0), 1.0, red),
        ("sphere", (2.5, 0.5, 3.0), 0.5, green),
    ]
    lights = [((10.0, 15.0, 10.0), (0.6, 0.6, 0.6)), ((-8.0, 12.0, 6.0), (0.4, 0.4, 0.4)),
              ((0.0, 20.0, -15.0), (0.3, 0.3, 0.3))]
    cam = dict(eye=(0.0, 4.0, 14.0), center=(0.0, 2.0, 0.0), up=(0.0, 1.0, 0.0), fovy=45.0, width=width, height=height)
    return SceneDesc("o_01_spheres", [], objs, lights, cam, (0.0, 0.0, 0.0), (0.2, 0.2, 0.2), 5)


# --------------------------------------------------------------------------------------------
# config 2 — Office-class mesh scene
# --------------------------------------------------------------------------------------------

def ellipsoid(center, radii, nslices, nstacks, displace=None):
    pos, idx = uv_sphere((0, 0, 0), 1.0, nslices, nstacks, displace=displace)
    return pos * np.asarray(radii, np.float64)[None, :] + np.asarray(center, np.float64)[None, :], idx


def rot_y(pos, angle, pivot):
    c, s = np.cos(angle), np.sin(angle)
    q = pos - np.asarray(pivot, np.float64)[None, :]
    out = np.stack([c * q[:, 0] + s * q[:, 2], q[:, 1], -s * q[:, 0] + c * q[:, 2]], -1)
    return out + np.asarray(pivot, np.float64)[None, :]


def capsule_leg(base, radius, height, nseg, nstack):
    """Cylinder side with slightly conical (non-coplanar) caps."""
    pos, idx = cylinder(base, radius, height, nseg, nstack)
    pos = pos.copy()
    pos[-2, 1] -= 0.3 * radius   # bottom cap apex
    pos[-1, 1] += 0.3 * radius   # top cap apex
    return pos, idx


def wavy_sheet(origin, eu, ev, nu, nv, amp, fu, fv, flip=False):
    """Height-field sheet: grid displaced along y by a sine product (no coplanar runs)."""
    pos, idx = grid_quad(origin, eu, ev, nu, nv, flip=flip)
    s = np.linspace(0.0, 1.0, nu + 1)
    t = np.linspace(0.0, 1.0, nv + 1)
    S, T = np.meshgrid(s, t, indexing="xy")
    pos = pos.copy()
    pos[:, 1] += (amp * np.sin(fu * np.pi * S) * np.cos(fv * np.pi * T) + 0.37 * amp * np.sin(7.3 * S + 4.1 * T)).reshape(-1)
    return pos, idx


def office(width=1920, height=1080, target_tris=34000, detail=1.0):
    """Office-class scene with exactly ``target_tris`` triangles (default 34 000).

    Built like a modelled office: the big flat surfaces (room shell, cabinet and book boxes) are a
    few large triangles each, and the triangle budget sits in the curved, densely tessellated
    furniture (table top, cushions, legs, lamp, vase, plant, mugs) and a wavy carpet.  That is the
    geometry class the reference's cycling-axis median split (mybvh.cpp:273) handles well: finely
    tessellated coplanar axis-aligned grids would be unsplittable on their flat axis (mybvh.cpp:282)
    and collapse into leaves of hundreds of triangles.  ``detail`` scales tessellation (the 10x
    variant uses detail ~ 3.2 with target_tris=340000)."""
    def n(v):
        return max(3, int(round(v * detail)))
    meshes = []

    def add(parts, mat, mode=0):
        pos, idx = merge(parts) if isinstance(parts, list) else parts
        meshes.append(MeshDesc(pos, idx, mat, mode))

    wall = material((0.25, 0.24, 0.22), (0.75, 0.72, 0.66), (0.05, 0.05, 0.05), 5.0, 0.0)
    floor_m = material((0.15, 0.12, 0.10), (0.55, 0.42, 0.30), (0.3, 0.3, 0.3), 40.0, 0.15)
    glass = material((0.05, 0.08, 0.1), (0.2, 0.3, 0.4), (0.9, 0.9, 0.9), 200.0, 0.25)
    wood = material((0.12, 0.08, 0.04), (0.5, 0.33, 0.18), (0.4, 0.4, 0.4), 60.0, 0.1)
    metal = material((0.1, 0.1, 0.1), (0.35, 0.35, 0.38), (0.8, 0.8, 0.8), 120.0, 0.2)
    fabric = material((0.05, 0.08, 0.15), (0.15, 0.3, 0.6), (0.05, 0.05, 0.05), 4.0, 0.0)
    cab = material((0.15, 0.15, 0.13), (0.6, 0.6, 0.55), (0.2, 0.2, 0.2), 30.0, 0.0)
    lampm = material((0.3, 0.28, 0.2), (0.9, 0.85, 0.6), (0.5, 0.5, 0.5), 80.0, 0.0)
    leaf = material((0.05, 0.15, 0.05), (0.2, 0.6, 0.25), (0.1, 0.1, 0.1), 10.0, 0.0)
    carpet = material((0.15, 0.05, 0.05), (0.6, 0.2, 0.2), (0.02, 0.02, 0.02), 2.0, 0.0)
    X, Y, Z = 10.0, 3.0, 8.0   # room: x in [0,10], y in [0,3], z in [0,8]
    # room shell: 2 triangles per face, normals pointing inwards; floor separate (slightly mirroring)
    F = dict(flip=True)   # eu x ev points out of the room for these argument orders: flip towards the inside
    add([grid_quad((0, 0, 0), (0, 0, Z), (0, Y, 0), 1, 1, **F), grid_quad((X, 0, 0), (0, Y, 0), (0, 0, Z), 1, 1, **F),
         grid_quad((0, 0, 0), (0, Y, 0), (X, 0, 0), 1, 1, **F), grid_quad((0, 0, Z), (X, 0, 0), (0, Y, 0), 1, 1, **F),
         grid_quad((0, Y, 0), (0, 0, Z), (X, 0, 0), 1, 1, **F)], wall)
    add(grid_quad((0, 0, 0), (X, 0, 0), (0, 0, Z), 2, 2, **F), floor_m)
    # window pane on the z = 0 wall plus an exactly coplanar duplicate (equal-t tie-break exercise)
    add([grid_quad((2.0, 1.0, 0.01), (0, 1.5, 0), (6.0, 0, 0), 1, 2, **F),
         grid_quad((2.0, 1.0, 0.01), (0, 1.5, 0), (3.0, 0, 0), 1, 1, **F)], glass)
    # table: squashed ellipsoid top + 4 legs
    add(ellipsoid((5.0, 0.75, 4.2), (1.6, 0.05, 0.9), n(72), n(18)), wood, 1)
    add([capsule_leg((5.0 + sx * 1.0, 0.0, 4.2 + sz * 0.5), 0.04, 0.72, n(16), n(6))
         for sx in (-1, 1) for sz in (-1, 1)], metal)
    # chairs: cushion + back rest ellipsoids, 4 legs each
    cush, legs = [], []
    for ci, (cx, cz) in enumerate(((3.2, 4.2), (6.8, 4.2), (5.0, 2.7), (5.0, 5.7))):
        ang = 0.4 * ci + 0.15
        c0 = ellipsoid((cx, 0.46, cz), (0.27, 0.05, 0.27), n(32), n(12))
        b0 = ellipsoid((cx, 0.80, cz + 0.24), (0.26, 0.28, 0.04), n(32), n(14))
        cush += [(rot_y(c0[0], ang, (cx, 0, cz)), c0[1]), (rot_y(b0[0], ang, (cx, 0, cz)), b0[1])]
        for sx in (-1, 1):
            for sz in (-1, 1):
                l0 = capsule_leg((cx + sx * 0.2, 0.0, cz + sz * 0.2), 0.018, 0.43, n(10), n(4))
                legs.append((rot_y(l0[0], ang, (cx, 0, cz)), l0[1]))
    add(cush, fabric, 1)
    add(legs, metal)
    # cabinets, shelves, books: boxes with 12 triangles each, books slightly rotated
    boxes = [box((9.3, 0.0, 1.0 + 2.0 * i), (9.95, 2.0, 2.6 + 2.0 * i), 1) for i in range(3)]
    boxes += [box((0.05, 0.9 + 0.5 * i, 1.0), (0.45, 0.95 + 0.5 * i, 6.0), 1) for i in range(3)]
    add(boxes, cab)
    books = []
    for i in range(3):
        for j in range(24):
            z0 = 1.1 + 0.2 * j
            b = box((0.1, 0.95 + 0.5 * i, z0), (0.38, 1.25 + 0.5 * i + 0.03 * ((j * 7) % 5), z0 + 0.12 + 0.01 * (j % 3)), 1)
            books.append((rot_y(b[0], 0.03 * ((j * 5) % 7 - 3), (0.24, 0, z0)), b[1]))
    add(books, wood)
    # monitor (rounded), keyboard keys, mugs on the table
    add(ellipsoid((5.0, 1.05, 3.9), (0.42, 0.26, 0.03), n(40), n(16)), metal, 1)
    keys = [box((4.7 + 0.045 * kx, 0.805, 4.25 + 0.045 * kz), (4.74 + 0.045 * kx, 0.82 + 0.002 * ((kx + kz) % 3), 4.29 + 0.045 * kz), 1)
            for kx in range(13) for kz in range(5)]
    add(keys, metal)
    add([capsule_leg((4.2 + 0.5 * i, 0.80, 4.55 + 0.05 * i), 0.045, 0.1, n(20), n(3)) for i in range(3)], lampm)
    # globe lamp + displaced vase (PHONG) + a plant made of small ellipsoid leaves
    add(uv_sphere((5.0, 2.45, 4.2), 0.25, n(48), n(24)), lampm, 1)
    add(uv_sphere((9.6, 2.3, 1.8), 0.3, n(44), n(22),
                  displace=lambda th, ph: 1.0 + 0.25 * np.sin(3 * ph) * np.cos(2 * th)), glass, 1)
    rng = np.random.default_rng(20251112)
    leaves = []
    for i in range(int(60 * detail)):
        a = rng.uniform(0, 2 * np.pi); r = rng.uniform(0.05, 0.45); h = rng.uniform(0.5, 1.5)
        l0 = ellipsoid((0, 0, 0), (0.12, 0.015, 0.05), n(10), n(6))
        tilt = rng.uniform(-0.6, 0.6)
        q = l0[0].copy()
        q[:, 1] += tilt * q[:, 0]
        q = rot_y(q, a, (0, 0, 0)) + np.array([8.6 + r * np.cos(a), h, 7.0 + r * np.sin(a)])
        leaves.append((q, l0[1]))
    leaves.append(capsule_leg((8.6, 0.0, 7.0), 0.16, 0.45, n(20), n(4)))
    add(leaves, leaf, 1)
    # wavy carpet: fills the triangle budget exactly
    have = sum(m.idx.shape[0] for m in meshes)
    rem = target_tris - have
    if rem < 8 or rem % 2:
        raise ValueError(f"office: cannot hit {target_tris} triangles (have {have})")
    q = rem // 2
    a = max(int(np.sqrt(q * 1.3)), 1)
    b = q // a
    parts = [wavy_sheet((2.2, 0.012, 1.8), (5.6, 0, 0), (0, 0, 4.8), a, b, 0.006, 9.0, 7.0, flip=True)]
    left = q - a * b
    if left:
        parts.append(wavy_sheet((2.2, 0.012, 6.62), (5.6, 0, 0), (0, 0, 0.12), left, 1, 0.004, 5.0, 1.0, flip=True))
    add(parts, carpet)
    lights = [((5.0, 2.9, 4.0), (0.55, 0.55, 0.5)), ((2.0, 2.5, 6.5), (0.35, 0.35, 0.4)),
              ((8.0, 2.6, 1.5), (0.3, 0.3, 0.3))]
    cam = dict(eye=(1.2, 1.7, 7.4), center=(6.0, 1.0, 3.0), up=(0.0, 1.0, 0.0), fovy=60.0, width=width, height=height)
    name = "office_%dk" % round(sum(m.idx.shape[0] for m in meshes) / 1000)
    return SceneDesc(name, meshes, [], lights, cam, (0.0, 0.0, 0.0), (0.25, 0.25, 0.25), 3)


# --------------------------------------------------------------------------------------------
# config 3 — molecule (many spheres) and rings (interlocked tori), reflection depth >= 3
# --------------------------------------------------------------------------------------------

def molecule(width=500, height=500, n_spheres=600, seed=4):
    rng = np.random.default_rng(seed)
    radii = (0.25, 0.3, 0.35, 0.45)
    cols = ((0.8, 0.1, 0.1), (0.1, 0.7, 0.2), (0.15, 0.25, 0.8), (0.8, 0.8, 0.2))
    p = np.array([0.0, 2.5, 0.0])
    objs = [("plane", (0.0, 0.0, 0.0), (0.0, 1.0, 0.0),
             material((0.2, 0.2, 0.2), (0.5, 0.5, 0.5), (0.3, 0.3, 0.3), 20.0, 0.3)),
            ("plane", (0.0, 0.0, -6.0), (0.0, 0.0, 1.0),
             material((0.15, 0.15, 0.2), (0.4, 0.4, 0.5), (0.0, 0.0, 0.0), 1.0, 0.0))]
    for i in range(n_spheres):
        step = rng.normal(size=3)
        step /= np.linalg.norm(step)
        p = p + 0.55 * step
        p = np.clip(p, (-4.0, 0.6, -4.0), (4.0, 5.0, 3.0))
        k = int(rng.integers(0, 4))
        m = material(tuple(0.2 * c for c in cols[k]), cols[k], (0.8, 0.8, 0.8), 80.0, 0.15 if k == 3 else 0.0)
        objs.append(("sphere", tuple(p), radii[k], m))
    lights = [((8.0, 12.0, 10.0), (0.6, 0.6, 0.6)), ((-6.0, 9.0, 8.0), (0.4, 0.4, 0.4))]
    cam = dict(eye=(0.0, 3.5, 11.0), center=(0.0, 2.5, 0.0), up=(0.0, 1.0, 0.0), fovy=50.0, width=width, height=height)
    return SceneDesc("o_04_molecule", [], objs, lights, cam, (0.0, 0.0, 0.0), (0.2, 0.2, 0.2), 3)


def rings(width=700, height=500, nu=256, nv=64):
    """Two interlocked tori, 2 * (nu*nv*2) triangles (65 536 at the default), PHONG normals."""
    gold = material((0.2, 0.15, 0.05), (0.8, 0.6, 0.2), (0.9, 0.9, 0.9), 100.0, 0.4)
    steel = material((0.1, 0.1, 0.12), (0.5, 0.5, 0.6), (0.9, 0.9, 0.9), 150.0, 0.5)
    m1 = MeshDesc(*torus((-0.8, 0.0, 0.0), 1.6, 0.35, nu, nv, axis=1), gold, 1)
    m2 = MeshDesc(*torus((0.8, 0.0, 0.0), 1.6, 0.35, nu, nv, axis=2), steel, 1)
    lights = [((5.0, 8.0, 8.0), (0.7, 0.7, 0.7)), ((-6.0, 4.0, 6.0), (0.4, 0.4, 0.4))]
    cam = dict(eye=(0.0, 3.0, 7.0), center=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0), fovy=45.0, width=width, height=height)
    return SceneDesc("o_09_rings", [m1, m2], [], lights, cam, (0.0, 0.0, 0.0), (0.2, 0.2, 0.2), 5)


# --------------------------------------------------------------------------------------------
# config 4 — pokemon-class textured meshes with Phong-interpolated normals
# --------------------------------------------------------------------------------------------

def textured(width=1920, height=1080, tex_size=1024, slices=224, stacks=112, seed=10):
    """3 displaced UV-spheres (~50 k triangles each at the defaults) + textured ground + sky quad."""
    meshes = []
    skin = material((0.2, 0.2, 0.2), (1.0, 1.0, 1.0), (0.3, 0.3, 0.3), 30.0, 0.0)
    blobs = [((-2.2, 1.2, 0.0), 1.2, (0.9, 0.8, 0.2), 5, 3), ((0.6, 1.0, 0.8), 1.0, (0.3, 0.6, 0.9), 4, 2),
             ((2.8, 0.8, -0.4), 0.8, (0.9, 0.4, 0.3), 6, 4)]
    for bi, (c, r, col, a, b) in enumerate(blobs):
        pos, idx, u, v = uv_sphere(c, r, slices, stacks, with_uv=True,
                                   displace=lambda th, ph, a=a, b=b: 1.0 + 0.12 * np.sin(a * th) * np.sin(b * ph) ** 2)
        # push some UVs outside [0,1] to exercise the clamp (mymesh.cpp:83-84)
        u = u * 1.1 - 0.05
        tex = value_noise_texture(tex_size, seed + bi, base=col, amp=0.6)
        meshes.append(MeshDesc(pos, idx, skin, 1, idx.copy(), u, v, tex))
    # ground: gently rolling height field (no coplanar runs, see office()), sky: one quad
    pos, idx, u, v = grid_quad((-8, 0, -6), (16, 0, 0), (0, 0, 12), 64, 48, with_uv=True, flip=True)
    pos = pos.copy()
    pos[:, 1] += 0.05 * np.sin(1.7 * pos[:, 0]) * np.cos(1.3 * pos[:, 2]) + 0.02 * np.sin(5.1 * pos[:, 0] + 3.3 * pos[:, 2])
    meshes.append(MeshDesc(pos, idx, material((0.2, 0.2, 0.2), (1, 1, 1), (0.1, 0.1, 0.1), 10.0, 0.0), 0, idx.copy(),
                           u, v, value_noise_texture(tex_size, seed + 7, base=(0.3, 0.6, 0.3), amp=0.5)))
    pos, idx, u, v = grid_quad((-20, -2, -9), (40, 0, 0), (0, 22, 0), 1, 1, with_uv=True)
    meshes.append(MeshDesc(pos, idx, material((0.9, 0.9, 0.9), (0.2, 0.2, 0.2), (0, 0, 0), 1.0, 0.0, shadowable=False),
                           0, idx.copy(), u, v, value_noise_texture(tex_size // 2, seed + 9, base=(0.4, 0.6, 0.9), amp=0.3)))
    lights = [((6.0, 10.0, 8.0), (0.7, 0.7, 0.7)), ((-7.0, 6.0, 5.0), (0.3, 0.3, 0.35))]
    cam = dict(eye=(0.5, 2.4, 8.0), center=(0.3, 1.0, 0.0), up=(0.0, 1.0, 0.0), fovy=40.0, width=width, height=height)
    return SceneDesc("o_10_textured", meshes, [], lights, cam, (0.0, 0.0, 0.0), (0.35, 0.35, 0.35), 2)


# --------------------------------------------------------------------------------------------
# config 5 — scaling scene: many displaced spheres scattered in a slab
# --------------------------------------------------------------------------------------------

def scatter(width=3840, height=2160, n_objects=10000, slices=25, stacks=21, seed=5, mirror_fraction=0.1):
    """n_objects UV-spheres of 2*slices*(stacks-1) = 1000 triangles each (10 M at the defaults)."""
    rng = np.random.default_rng(seed)
    base_pos, base_idx = uv_sphere((0, 0, 0), 1.0, slices, stacks)
    nvb = base_pos.shape[0]
    centers = np.stack([rng.uniform(-60, 60, n_objects), rng.uniform(0.5, 12, n_objects),
                        rng.uniform(-60, 20, n_objects)], -1)
    radii = rng.uniform(0.25, 0.9, n_objects)
    is_mirror = rng.random(n_objects) < mirror_fraction
    cols = rng.uniform(0.2, 0.9, (n_objects, 3))
    meshes = []
    for grp, mir in ((~is_mirror, 0.0), (is_mirror, 0.3)):
        sel = np.nonzero(grp)[0]
        if sel.size == 0:
            continue
        # a handful of meshes per group (distinct materials), each many instanced-by-copy spheres
        for chunk in np.array_split(sel, 4):
            if chunk.size == 0:
                continue
            pos = (base_pos[None, :, :] * radii[chunk, None, None] + centers[chunk, None, :]).reshape(-1, 3)
            idx = (base_idx[None, :, :] + (np.arange(chunk.size) * nvb)[:, None, None]).reshape(-1, 3)
            c = cols[chunk[0]]
            meshes.append(MeshDesc(pos, idx, material(0.2 * c, c, (0.6, 0.6, 0.6), 60.0, mir), 1))
    pos, idx = grid_quad((-80, 0, -80), (160, 0, 0), (0, 0, 120), 64, 64, flip=True)
    pos = pos.copy()
    pos[:, 1] += 0.3 * np.sin(0.11 * pos[:, 0]) * np.cos(0.13 * pos[:, 2]) + 0.1 * np.sin(0.7 * pos[:, 0] + 0.5 * pos[:, 2])
    meshes.append(MeshDesc(pos, idx, material((0.2, 0.2, 0.2), (0.5, 0.5, 0.5), (0.1, 0.1, 0.1), 10.0, 0.0), 0))
    lights = [((40.0, 80.0, 60.0), (0.7, 0.7, 0.7)), ((-50.0, 60.0, 30.0), (0.4, 0.4, 0.4))]
    cam = dict(eye=(0.0, 14.0, 55.0), center=(0.0, 4.0, -10.0), up=(0.0, 1.0, 0.0), fovy=50.0, width=width, height=height)
    return SceneDesc("scatter_%dk" % (n_objects * 2 * slices * (stacks - 1) // 1000), meshes, [], lights, cam,
                     (0.1, 0.15, 0.25), (0.2, 0.2, 0.2), 3)


# --------------------------------------------------------------------------------------------
# small mixed scene for fast parity tests (meshes + analytic primitives + texture + PHONG)
# --------------------------------------------------------------------------------------------

def mixed_small(width=160, height=120, seed=1):
    rng = np.random.default_rng(seed)
    meshes = []
    pos, idx, u, v = uv_sphere((0.0, 1.0, 0.0), 1.0, 24, 12, with_uv=True)
    meshes.append(MeshDesc(pos, idx, material((0.2, 0.2, 0.2), (1, 1, 1), (0.5, 0.5, 0.5), 40.0, 0.2), 1, idx.copy(),
                           u * 1.2 - 0.1, v, value_noise_texture(64, seed, base=(0.8, 0.5, 0.3), amp=0.7)))
    meshes.append(MeshDesc(*box((-3.0, 0.0, -1.0), (-1.5, 1.5, 0.5), 3),
                           material((0.1, 0.2, 0.1), (0.2, 0.7, 0.3), (0.4, 0.4, 0.4), 20.0, 0.0), 0))
    meshes.append(MeshDesc(*torus((2.2, 0.8, 0.5), 0.7, 0.25, 24, 12), material((0.2, 0.1, 0.1), (0.8, 0.3, 0.3),
                                                                           (0.9, 0.9, 0.9), 90.0, 0.4), 1))
    # random soup with duplicated (coplanar, equal-t) triangles
    p = rng.uniform(-3, 3, (40, 3)) * np.array([1.0, 0.3, 1.0]) + np.array([0, 2.5, -2.0])
    tri = rng.integers(0, 40, (60, 3)).astype(np.int32)
    tri = np.concatenate([tri, tri[:10]], 0)
    meshes.append(MeshDesc(p, tri, material((0.1, 0.1, 0.2), (0.3, 0.3, 0.9), (0.2, 0.2, 0.2), 10.0, 0.0,
                                            shadowable=False), 0))
    objs = [("plane", (0, 0, 0), (0, 1, 0), material((0.2, 0.2, 0.2), (0.6, 0.6, 0.6), (0.2, 0.2, 0.2), 10.0, 0.3)),
            ("sphere", (0.5, 0.5, 2.5), 0.5, material((0.2, 0.2, 0.0), (0.8, 0.8, 0.1), (0.9, 0.9, 0.9), 120.0, 0.5))]
    lights = [((5.0, 8.0, 6.0), (0.6, 0.6, 0.6)), ((-4.0, 6.0, 5.0), (0.4, 0.4, 0.4))]
    cam = dict(eye=(0.5, 2.5, 8.0), center=(0.0, 1.0, 0.0), up=(0.0, 1.0, 0.0), fovy=45.0, width=width, height=height)
    return SceneDesc("mixed_small", meshes, objs, lights, cam, (0.05, 0.05, 0.1), (0.2, 0.2, 0.2), 4)


def objtree(width=96, height=72, seed=11, camera_inside=True):
    """Enough analytic spheres for the object tree (>= 8), with the cases that test the reference's list-order
    semantics: a plane in the middle of the list, an exact duplicate of object 1 at a higher index (equal t: the
    lower index must win), a sphere with the geometry of object 5 and another material, optionally a sphere around
    the camera (ray origins inside a box and inside a sphere), and a triangle mesh behind everything."""
    rng = np.random.default_rng(seed)

    def mat(c, mirror=0.0):
        return material((0.1, 0.1, 0.1), c, (0.5, 0.5, 0.5), 30.0, mirror)
    objs = []
    for k in range(14):
        ctr = tuple(rng.uniform(-2.0, 2.0, 3) * np.array([1.0, 0.6, 1.0]))
        objs.append(("sphere", ctr, float(rng.uniform(0.25, 0.7)), mat(tuple(rng.uniform(0.2, 1.0, 3)), 0.3 if k % 4 == 0 else 0.0)))
    objs.insert(3, ("plane", (0.0, -1.5, 0.0), (0.0, 1.0, 0.0), mat((0.7, 0.7, 0.7), 0.2)))
    objs.append(objs[1])
    objs.append(("sphere", objs[5][1], objs[5][2], mat((1.0, 0.0, 1.0))))
    if camera_inside:
        objs.append(("sphere", (0.0, 0.0, 6.0), 1.5, mat((0.2, 0.9, 0.9))))
    quad = grid_quad((-3.0, -1.0, -4.0), (6.0, 0.0, 0.0), (0.0, 4.0, 0.0), 6, 4)
    return SceneDesc("objtree", [MeshDesc(quad[0], quad[1], mat((0.9, 0.8, 0.3)))], objs,
                     [((3.0, 4.0, 5.0), (0.8, 0.8, 0.8)), ((-4.0, 2.0, 1.0), (0.5, 0.5, 0.6))],
                     dict(eye=(0.0, 0.0, 6.0), center=(0.0, 0.0, 0.0), up=(0, 1, 0), fovy=50, width=width, height=height),
                     ambience=(0.2, 0.2, 0.2), max_depth=3)


REGISTRY = {"objtree": objtree, "spheres": spheres, "office": office, "molecule": molecule, "rings": rings, "textured": textured,
            "scatter": scatter, "mixed_small": mixed_small}
