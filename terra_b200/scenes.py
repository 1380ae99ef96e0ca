"""Seeded procedural scenes for BASELINE.json configs C1-C5, built through the public C API only
(terra_scene_add_object -> fill triangles/properties in place -> material -> options -> commit;
reference call sequence: satellite/src/Scene.cpp:144-237).

Every builder takes a `TerraLib` (product or reference — same ABI) and returns (Scene, Camera).
Scenes keep <= 64 objects (reference realloc bug, src/Terra.c:143-146) and <= 2^23 triangles
per object (leaf index packing, src/TerraBVH.c:155).
"""
from __future__ import annotations

import numpy as np

from . import api
from .api import Scene, TerraLib, make_camera


# ------------------------------------------------------------------------------------ meshes
def quad(p0, p1, p2, p3):
    """Two triangles p0-p2-p1, p0-p3-p2 -> (2,3,3); the flat normal is cross(p3-p0, p1-p0)."""
    p0, p1, p2, p3 = (np.asarray(p, np.float32) for p in (p0, p1, p2, p3))
    return np.stack([np.stack([p0, p2, p1]), np.stack([p0, p3, p2])]).astype(np.float32)


def box(lo, hi, rot_y=0.0):
    """Axis-aligned box [lo,hi], optionally rotated about its centre's y axis. (12,3,3), outward winding."""
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    c = [(x, y, z) for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])]
    v = np.asarray(c, np.float32)     # index = 4*ix + 2*iy + iz
    faces = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    tris = np.concatenate([quad(v[a], v[d], v[c_], v[b]) for a, b, c_, d in faces])
    if rot_y:
        ctr = (lo + hi) / 2
        s, co = np.sin(rot_y), np.cos(rot_y)
        R = np.array([[co, 0, s], [0, 1, 0], [-s, 0, co]], np.float32)
        tris = ((tris - ctr) @ R.T + ctr).astype(np.float32)
    return tris


def uv_sphere(center, radius, n_lat, n_lon):
    """UV sphere with smooth normals. Triangle count = 2*n_lon*(n_lat-1). Returns (tris, normals)."""
    th = np.linspace(0, np.pi, n_lat + 1, dtype=np.float64)
    ph = np.linspace(0, 2 * np.pi, n_lon + 1, dtype=np.float64)
    T, P = np.meshgrid(th, ph, indexing="ij")
    n = np.stack([np.sin(T) * np.cos(P), np.cos(T), np.sin(T) * np.sin(P)], -1)
    tris, nrm = [], []
    for i in range(n_lat):
        a, b = n[i, :-1], n[i, 1:]
        c, d = n[i + 1, :-1], n[i + 1, 1:]
        if i != 0:
            tris.append(np.stack([a, c, b], 1))
        if i != n_lat - 1:
            tris.append(np.stack([b, c, d], 1))
    nrm = np.concatenate(tris).astype(np.float32)
    pos = (nrm * np.float32(radius) + np.asarray(center, np.float32)).astype(np.float32)
    return pos, nrm


def heightfield(nx, nz, extent, height_fn, jitter=0.25, seed=3, y0=0.0):
    """(nx*nz*2) triangle displaced grid over [0,extent]^2 with jittered vertices (unique x
    centres, so the reference's x-only sort has no ties). Returns (tris, smooth normals)."""
    rng = np.random.default_rng(seed)
    gx, gz = np.meshgrid(np.arange(nx + 1, dtype=np.float64), np.arange(nz + 1, dtype=np.float64), indexing="ij")
    gx = gx + rng.uniform(-jitter, jitter, gx.shape)
    gz = gz + rng.uniform(-jitter, jitter, gz.shape)
    x = gx * (extent / nx)
    z = gz * (extent / nz)
    y = height_fn(x, z) + y0
    P = np.stack([x, y, z], -1)
    # vertex normals from central differences
    dx = np.gradient(P, axis=0)
    dz = np.gradient(P, axis=1)
    N = np.cross(dz, dx)
    N /= np.maximum(np.linalg.norm(N, axis=-1, keepdims=True), 1e-30)
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    na, nb, nc, nd = N[:-1, :-1], N[1:, :-1], N[1:, 1:], N[:-1, 1:]
    t1 = np.stack([a, c, b], -2).reshape(-1, 3, 3)
    t2 = np.stack([a, d, c], -2).reshape(-1, 3, 3)
    n1 = np.stack([na, nc, nb], -2).reshape(-1, 3, 3)
    n2 = np.stack([na, nd, nc], -2).reshape(-1, 3, 3)
    tris = np.concatenate([t1, t2]).astype(np.float32)
    nrm = np.concatenate([n1, n2]).astype(np.float32)
    return tris, nrm


def fbm(x, z, seed=3, octaves=5, base_freq=1.0, amp=1.0):
    """Cheap sum-of-sines fBm (deterministic, no tables)."""
    rng = np.random.default_rng(seed)
    h = np.zeros_like(x)
    f, a = base_freq, amp
    for _ in range(octaves):
        px, pz, rot = rng.uniform(0, 2 * np.pi, 3)
        cx, cz = np.cos(rot), np.sin(rot)
        h += a * np.sin(f * (cx * x + cz * z) + px) * np.cos(f * (-cz * x + cx * z) + pz)
        f *= 2.03
        a *= 0.5
    return h


# ------------------------------------------------------------------------------------ C1
def cornell(lib: TerraLib, *, integrator=api.INTEGRATOR_SIMPLE, spp=16, bounces=8, jitter=0.5,
            tonemap=api.TONEMAP_NONE, materials="diffuse"):
    """Cornell box, unit scale [0,1]^3 open towards -z: 5 walls (10 tris) + short box (12) +
    tall box (12) + ceiling light quad (2) = 36 triangles. Lambert + one emissive quad.
    `materials`: 'diffuse' (C1), 'phong' (boxes become Phong), 'ggx', 'glass' (C5-like small)."""
    s = Scene(lib)
    white, red, green = (0.73, 0.73, 0.73), (0.65, 0.05, 0.05), (0.12, 0.45, 0.15)
    walls_white = np.concatenate([
        quad((0, 0, 0), (1, 0, 0), (1, 0, 1), (0, 0, 1)),        # floor
        quad((0, 1, 0), (0, 1, 1), (1, 1, 1), (1, 1, 0)),        # ceiling
        quad((0, 0, 1), (1, 0, 1), (1, 1, 1), (0, 1, 1)),        # back
    ])
    s.set_diffuse(s.add_object(walls_white), white)
    s.set_diffuse(s.add_object(quad((0, 0, 0), (0, 0, 1), (0, 1, 1), (0, 1, 0))), red)     # left
    s.set_diffuse(s.add_object(quad((1, 0, 0), (1, 1, 0), (1, 1, 1), (1, 0, 1))), green)   # right
    short = box((0.53, 0.0, 0.17), (0.83, 0.30, 0.47), rot_y=-0.29)
    tall = box((0.18, 0.0, 0.52), (0.48, 0.60, 0.82), rot_y=0.31)
    o_short, o_tall = s.add_object(short), s.add_object(tall)
    if materials == "diffuse":
        s.set_diffuse(o_short, white)
        s.set_diffuse(o_tall, white)
    elif materials == "phong":
        s.set_phong(o_short, (0.5, 0.5, 0.5), (0.4, 0.4, 0.4), 20.0)
        s.set_phong(o_tall, (0.3, 0.3, 0.6), (0.5, 0.5, 0.5), 60.0)
    elif materials == "ggx":
        s.set_ggx(o_short, (0.7, 0.6, 0.3), 0.3, 1.0)
        s.set_ggx(o_tall, (0.6, 0.6, 0.6), 0.5, 0.0)
    elif materials == "glass":
        s.set_glass(o_short, (1.0, 1.0, 1.0), 1.5)
        s.set_ggx(o_tall, (0.8, 0.8, 0.8), 0.15, 1.0)
    else:
        raise ValueError(materials)
    light = quad((0.35, 0.999, 0.35), (0.35, 0.999, 0.65), (0.65, 0.999, 0.65), (0.65, 0.999, 0.35))
    s.set_diffuse(s.add_object(light), (0.0, 0.0, 0.0), emissive=(15, 15, 15))
    s.set_options(integrator=integrator, spp=spp, bounces=bounces, jitter=jitter, tonemap=tonemap)
    s.commit()
    cam = make_camera((0.5, 0.5, -1.35), (0, 0, 1), (0, 1, 0), 45.0)
    return s, cam


# ------------------------------------------------------------------------------------ C2
def sphere_field(lib: TerraLib, *, n_spheres=125, n_lat=20, n_lon=20, extent=100.0, seed=1,
                 integrator=api.INTEGRATOR_SIMPLE, spp=256, bounces=4, jitter=0.5, ggx=True,
                 n_groups=16):
    """~100k-triangle field: n_spheres uv-spheres (2*n_lon*(n_lat-1) tris each; 125 x 760 = 95k
    + 2 floor + emissive) on a jittered grid, grouped by material into <= 64 objects. Half the
    groups Lambert, half GGX (when `ggx`), 4 emissive spheres, a large floor quad pair."""
    rng = np.random.default_rng(seed)
    s = Scene(lib)
    g = int(np.ceil(n_spheres ** (1 / 3)))
    cells = [(i, j, k) for i in range(g) for j in range(g) for k in range(g)][:n_spheres]
    cell = extent / g
    groups = [[] for _ in range(n_groups)]
    emissive = []
    for idx, (i, j, k) in enumerate(cells):
        c = (np.array([i, j, k]) + 0.5 + rng.uniform(-0.2, 0.2, 3)) * cell
        r = cell * rng.uniform(0.22, 0.32)
        pos, nrm = uv_sphere(c, r, n_lat, n_lon)
        if idx % 31 == 7 and len(emissive) < 4:
            emissive.append((pos, nrm))
        else:
            groups[idx % n_groups].append((pos, nrm))
    for gi, grp in enumerate(groups):
        if not grp:
            continue
        pos = np.concatenate([p for p, _ in grp])
        nrm = np.concatenate([n for _, n in grp])
        o = s.add_object(pos, nrm)
        alb = rng.uniform(0.2, 0.8, 3)
        if ggx and gi % 2 == 1:
            s.set_ggx(o, alb, float(rng.uniform(0.1, 0.6)), float(gi % 4 == 1))
        else:
            s.set_diffuse(o, alb)
    for pos, nrm in emissive:
        s.set_diffuse(s.add_object(pos, nrm), (0, 0, 0), emissive=(12, 11, 10))
    m = 0.15 * extent
    s.set_diffuse(s.add_object(quad((-m, -0.5, -m), (extent + m, -0.5, -m), (extent + m, -0.5, extent + m), (-m, -0.5, extent + m))),
                  (0.5, 0.5, 0.5))
    s.set_options(integrator=integrator, spp=spp, bounces=bounces, jitter=jitter)
    s.commit()
    cam = make_camera((extent * 0.5, extent * 0.62, -extent * 1.05), (0, -0.18, 1), (0, 1, 0), 45.0)
    return s, cam


# ------------------------------------------------------------------------------------ C3 / C4
def terrain(lib: TerraLib, *, n_tris=1_000_000, extent=1000.0, seed=3, n_objects=None,
            integrator=api.INTEGRATOR_SIMPLE, spp=1, bounces=4, jitter=0.5, lights=True, env_tex=None):
    """Displaced-grid terrain (fBm height, relief ~8% of extent) with ~n_tris triangles, split into
    x-strips of <= 2^23 triangles (>= 2 objects above 8.3M). Optional emissive quads
    floating above (Q2: the environment contributes nothing in parity mode)."""
    n = int(round(np.sqrt(n_tris / 2)))
    k = 2 * np.pi / extent
    tris, nrm = heightfield(n, n, extent, lambda x, z: 0.08 * extent * fbm(x * k, z * k, seed=seed, base_freq=2.0), seed=seed)
    if n_objects is None:
        n_objects = max(2 if n_tris > 4_000_000 else 1, int(np.ceil(tris.shape[0] / (1 << 23))))
    s = Scene(lib)
    rng = np.random.default_rng(seed + 100)
    order = np.argsort(tris[:, :, 0].mean(axis=1), kind="stable")
    for part in np.array_split(order, n_objects):
        part = np.sort(part)
        s.set_diffuse(s.add_object(tris[part], nrm[part]), rng.uniform(0.3, 0.7, 3))
    if lights:
        for cx, cz in ((0.3, 0.3), (0.7, 0.6), (0.5, 0.85)):
            a = 0.06 * extent
            x0, z0, y = cx * extent, cz * extent, 0.35 * extent
            q = quad((x0 - a, y, z0 - a), (x0 - a, y, z0 + a), (x0 + a, y, z0 + a), (x0 + a, y, z0 - a))
            s.set_diffuse(s.add_object(q), (0, 0, 0), emissive=(40, 38, 35))
    s.set_options(integrator=integrator, spp=spp, bounces=bounces, jitter=jitter, env=(0.4, 0.52, 1.0))
    if env_tex is not None:
        tex = s.make_texture(env_tex)
        lib.terra_attribute_init_cubemap(api.C.byref(s.options().environment_map), api.C.byref(tex))
    s.commit()
    cam = make_camera((0.5 * extent, 0.30 * extent, -0.25 * extent), (0, -0.35, 1), (0, 1, 0), 50.0)
    return s, cam


def procedural_sky(width=4096, height=2048, seed=5):
    """Float HDR lat-long sky (h,w,3)."""
    rng = np.random.default_rng(seed)
    v = (np.arange(height, dtype=np.float32) + 0.5) / height
    u = (np.arange(width, dtype=np.float32) + 0.5) / width
    V, U = np.meshgrid(v, u, indexing="ij")
    horizon = np.exp(-((V - 0.5) * 6) ** 2)
    sky = np.stack([0.3 + 0.5 * horizon, 0.45 + 0.4 * horizon, 0.9 - 0.2 * horizon], -1)
    sun = np.exp(-(((U - 0.3) * 2) ** 2 + (V - 0.25) ** 2) * 900) * 500.0
    sky = sky + sun[..., None] * np.array([1.0, 0.9, 0.7], np.float32)
    sky[V > 0.5] *= 0.25
    return (sky + rng.uniform(0, 0.01, sky.shape)).astype(np.float32)


def incoherent_rays(scene_lo, scene_hi, n, seed=7):
    """C3 ray set: origin ~U(bbox inflated 10%), direction uniform on the sphere."""
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(scene_lo, np.float64), np.asarray(scene_hi, np.float64)
    ext = hi - lo
    o = rng.uniform(lo - 0.05 * ext, hi + 0.05 * ext, (n, 3)).astype(np.float32)
    z = rng.uniform(-1, 1, n)
    ph = rng.uniform(0, 2 * np.pi, n)
    r = np.sqrt(np.maximum(0.0, 1 - z * z))
    d = np.stack([r * np.cos(ph), z, r * np.sin(ph)], -1).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    return o, d.astype(np.float32)


# ------------------------------------------------------------------------------------ C5
def glass_metal(lib: TerraLib, *, integrator=api.INTEGRATOR_SIMPLE, spp=512, bounces=32, jitter=0.5,
                n_lat=24, n_lon=32, seed=11, ggx_init=None, glass_init=None):
    """Divergence stress: glass spheres (ior 1.5) + GGX metal spheres (roughness 0.05-0.3) in a
    Lambert room with two emissive panels."""
    rng = np.random.default_rng(seed)
    s = Scene(lib)
    W = 10.0
    room = np.concatenate([
        quad((0, 0, 0), (W, 0, 0), (W, 0, W), (0, 0, W)),
        quad((0, W * 0.6, 0), (0, W * 0.6, W), (W, W * 0.6, W), (W, W * 0.6, 0)),
        quad((0, 0, W), (W, 0, W), (W, W * 0.6, W), (0, W * 0.6, W)),
        quad((0, 0, 0), (0, 0, W), (0, W * 0.6, W), (0, W * 0.6, 0)),
        quad((W, 0, 0), (W, W * 0.6, 0), (W, W * 0.6, W), (W, 0, W)),
    ])
    s.set_diffuse(s.add_object(room), (0.7, 0.7, 0.7))
    glass, metal = [], []
    for i in range(4):
        for k in range(3):
            c = ((i + 0.5) * W / 4 + rng.uniform(-0.3, 0.3), 0.0, (k + 0.8) * W / 3.6 + rng.uniform(-0.3, 0.3))
            r = rng.uniform(0.7, 1.0)
            c = (c[0], r + 0.01, c[2])
            (glass if (i + k) % 2 == 0 else metal).append(uv_sphere(c, r, n_lat, n_lon))
    o = s.add_object(np.concatenate([p for p, _ in glass]), np.concatenate([n for _, n in glass]))
    s.set_glass(o, (0.98, 0.98, 0.98), 1.5, bsdf_init=glass_init)
    for j, (p, n) in enumerate(metal):
        s.set_ggx(s.add_object(p, n), rng.uniform(0.6, 0.95, 3), float(rng.uniform(0.05, 0.3)), 1.0, bsdf_init=ggx_init)
    y = W * 0.6 - 0.002
    for x0 in (0.25 * W, 0.65 * W):
        q = quad((x0, y, 0.3 * W), (x0, y, 0.7 * W), (x0 + 0.12 * W, y, 0.7 * W), (x0 + 0.12 * W, y, 0.3 * W))
        s.set_diffuse(s.add_object(q), (0, 0, 0), emissive=(25, 25, 25))
    s.set_options(integrator=integrator, spp=spp, bounces=bounces, jitter=jitter)
    s.commit()
    cam = make_camera((W / 2, W * 0.32, -W * 0.9), (0, -0.08, 1), (0, 1, 0), 45.0)
    return s, cam
