"""Seeded inputs shared by tests/golden/make_golden.py (which runs them through the UNMODIFIED
reference to produce tests/golden/*.npz) and by the tests (which run them through the oracle and
through the CUDA path). Keeping them in one place is what makes a fixture comparable."""
from __future__ import annotations

import ctypes as C

import numpy as np

SMALL_FIELD = dict(n_spheres=8, n_lat=8, n_lon=10, extent=20.0, seed=1, ggx=False, n_groups=4)
SMALL_TERRAIN = dict(n_tris=5000, extent=100.0, seed=3, lights=True)
PRIMARY = {           # name -> (builder, kwargs, (width, height))
    "cornell": ("cornell", {}, (64, 64)),
    "field": ("sphere_field", SMALL_FIELD, (96, 54)),
    "terrain": ("terrain", SMALL_TERRAIN, (80, 45)),
}
RENDER_SEED = 41      # child does srand(seed + 1)
RENDER_CASES = {      # name -> cornell kwargs (jitter 0 => deterministic camera rays; rand() stream pinned by srand)
    "simple_diffuse": dict(integrator=0, spp=4, bounces=8, jitter=0.0, materials="diffuse"),
    "direct_diffuse": dict(integrator=1, spp=4, bounces=4, jitter=0.0, materials="diffuse"),
    "mis_diffuse": dict(integrator=2, spp=4, bounces=4, jitter=0.0, materials="diffuse"),
    "simple_phong": dict(integrator=0, spp=4, bounces=8, jitter=0.0, materials="phong"),
    "direct_phong": dict(integrator=1, spp=4, bounces=4, jitter=0.0, materials="phong"),
    "mis_phong": dict(integrator=2, spp=4, bounces=4, jitter=0.0, materials="phong"),
    # debug views, Terra.c:1140-1337 (mono / depth / normals are deterministic per pixel at jitter 0)
    "debug_mono": dict(integrator=3, spp=2, bounces=2, jitter=0.0, materials="diffuse"),
    "debug_depth": dict(integrator=4, spp=2, bounces=2, jitter=0.0, materials="diffuse"),
    "debug_normals": dict(integrator=5, spp=2, bounces=2, jitter=0.0, materials="phong"),
    "debug_mis_diffuse": dict(integrator=6, spp=4, bounces=2, jitter=0.0, materials="diffuse"),
    "debug_mis_phong": dict(integrator=6, spp=4, bounces=2, jitter=0.0, materials="phong"),
}
RENDER_SIZE = (32, 32)
# env-lit extension mode: open scenes where paths escape to the environment
ENVLIT_CASES = {
    "envlit_const": dict(integrator=0, spp=4, bounces=4, jitter=0.0, env="const"),
    "envlit_latlong": dict(integrator=0, spp=4, bounces=4, jitter=0.0, env="latlong"),
    "envlit_latlong_direct": dict(integrator=1, spp=4, bounces=3, jitter=0.0, env="latlong"),
}


def envlit_scene(scenes, api, lib, *, integrator, spp, bounces, jitter, env, commit=True):
    """Small sphere field (open to the sky) with a constant or a lat-long HDR environment map."""
    import ctypes as C
    import numpy as np
    orig = api.Scene.commit
    api.Scene.commit = lambda self: None
    try:
        s, cam = scenes.sphere_field(lib, **dict(SMALL_FIELD, integrator=integrator, spp=spp, bounces=bounces, jitter=jitter))
    finally:
        api.Scene.commit = orig
    o = s.options()
    if env == "const":
        s.const_attr(o.environment_map, (0.4, 0.52, 1.0))
    else:
        tex = s.make_texture(scenes.procedural_sky(64, 32, seed=5))
        lib.terra_attribute_init_cubemap(C.byref(o.environment_map), C.byref(tex))
    if commit:
        s.commit()
    return s, cam


def build(scenes, lib, name, **over):
    fn, kw, wh = PRIMARY[name]
    kw = dict(kw)
    kw.update(over)
    s, cam = getattr(scenes, fn)(lib, **kw)
    return s, cam, wh


def scene_bounds(scene, api):
    """Bounding box of all triangles of a Scene built through the API (reads the objects in place)."""
    lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
    objs = (api.Object * scene.n_objects).from_address(scene.objects_base)
    for o in objs:
        a = np.ctypeslib.as_array(C.cast(o.triangles, C.POINTER(C.c_float)), (o.triangles_count * 3, 3))
        lo, hi = np.minimum(lo, a.min(0)), np.maximum(hi, a.max(0))
    return lo, hi


def rays_for(scene, api, scenes, n, seed=7):
    lo, hi = scene_bounds(scene, api)
    return scenes.incoherent_rays(lo, hi, n, seed=seed)


def kat_rays():
    """Single-triangle known-answer rays: interior, exactly through an edge, exactly through a vertex,
    co-planar, parallel miss, behind the origin, axis-aligned (zero direction components -> inf
    inverse directions), negative dominant axis. Triangle = (0,0,5) (4,0,5) (0,4,5) plus a far one."""
    tri = np.array([[[0, 0, 5], [4, 0, 5], [0, 4, 5]],
                    [[-50, -50, 60], [50, -50, 60], [0, 50, 60]]], np.float32)
    o, d = [], []

    def add(origin, target=None, direction=None):
        origin = np.asarray(origin, np.float64)
        v = np.asarray(direction, np.float64) if direction is not None else np.asarray(target, np.float64) - origin
        o.append(origin)
        d.append(v / np.linalg.norm(v))

    add((1, 1, 0), direction=(0, 0, 1))            # interior, axis aligned (inv dir = inf, inf, 1)
    add((2, 2, 0), direction=(0, 0, 1))            # exactly on the hypotenuse edge -> W == 0 -> fp64 retry
    add((2, 0, 0), direction=(0, 0, 1))            # on edge a-b
    add((0, 2, 0), direction=(0, 0, 1))            # on edge a-c
    add((0, 0, 0), direction=(0, 0, 1))            # through vertex a
    add((4, 0, 0), direction=(0, 0, 1))            # through vertex b
    add((0, 4, 0), direction=(0, 0, 1))            # through vertex c
    add((1, 1, 10), direction=(0, 0, -1))          # from behind (no culling): dominant axis negative
    add((1, 1, 5), direction=(1, 0, 0))            # co-planar with the triangle
    add((1, 1, 6), direction=(0, 0, 1))            # starts past the near triangle, hits the far one
    add((1, 1, 4.9995), direction=(0, 0, 1))       # origin offset 0.001 puts the start just past? (t tiny)
    add((10, 10, 0), direction=(0, 0, 1))          # misses the near one, hits the far one
    add((100, 100, 0), direction=(0, 0, 1))        # misses both
    add((1, 1, 0), target=(1.5, 1.25, 5))          # generic
    add((-3, 7, -2), target=(1, 1, 5))             # generic oblique
    add((1, 1, 0), direction=(1, 0, 0))            # parallel to the plane, miss
    add((1, 1, 0), direction=(0, 1, 0))
    add((3, 3, 0), direction=(-1, -1, 5))          # dominant z, passes near the hypotenuse
    rng = np.random.default_rng(5)
    for _ in range(46):
        add(rng.uniform(-2, 6, 3) * (1, 1, 0) + (0, 0, rng.uniform(-3, 3)), target=rng.uniform(-1, 5, 3) * (1, 1, 0) + (0, 0, 5))
    return tri, np.asarray(o, np.float32), np.asarray(d, np.float32)


def bsdf_kat_inputs(n=256, seed=9):
    rng = np.random.default_rng(seed)
    e = rng.uniform(0, 1, (n, 3)).astype(np.float32)
    e[0] = (0, 0, 0)
    e[1] = (1, 1, 1)            # _randf() can return exactly 1.0
    nrm = rng.normal(size=(n, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    nrm[2] = (0, 1, 0)
    nrm[3] = (1, 0, 0)
    nrm[4] = (0, 0, -1)
    wo = rng.normal(size=(n, 3))
    wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    flip = (wo * nrm).sum(1) < 0
    wo[flip] = -wo[flip]          # outgoing direction on the normal's side
    return e, wo.astype(np.float32), nrm.astype(np.float32)


class no_commit:
    """Context manager: build scenes through a library's public API but skip terra_scene_commit
    (for the product, commit is the host->device crossing and needs a GPU). The oracle only needs
    the caller-owned TerraObject array and the options, both valid before commit."""

    def __init__(self, api):
        self.api = api

    def __enter__(self):
        self.orig = self.api.Scene.commit
        self.api.Scene.commit = lambda s: None
        return self

    def __exit__(self, *a):
        self.api.Scene.commit = self.orig


def build_host(scenes, api, lib, name, **over):
    with no_commit(api):
        return build(scenes, lib, name, **over)
