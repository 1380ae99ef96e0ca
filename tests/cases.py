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


# GGX / glass known-answer cases run through the reference's DEAD presets (oracle/ref_dead_presets.c) and through the
# restatements: attribute slots of include/terra_b200.h, and whether the outgoing direction is flipped to the far side
# of the normal (glass only: a path leaving the medium, total internal reflection).
DEAD_PRESET_CASES = {
    "ggx_metal": ("ggx", [(0.8, 0.6, 0.3), (0.2, 1.0, 1.5)], False),
    "ggx_plastic": ("ggx", [(0.4, 0.5, 0.6), (0.5, 0.0, 1.5)], False),
    "ggx_mixed": ("ggx", [(0.4, 0.5, 0.6), (0.35, 0.4, 1.33)], False),
    "ggx_rough_half_metal": ("ggx", [(0.9, 0.9, 0.9), (0.8, 0.5, 2.0)], False),
    "glass": ("glass", [(0.98, 0.97, 0.96), (1.5, 0, 0), (0, 0, 0)], False),
    "glass_from_inside": ("glass", [(0.9, 0.8, 0.7), (1.5, 0, 0), (0, 0, 0)], True),
    "glass_diamond_from_inside": ("glass", [(1.0, 1.0, 1.0), (2.4, 0, 0), (0, 0, 0)], True),
}


def dead_preset_inputs(flip_wo):
    e, wo, n = bsdf_kat_inputs(1024, seed=31)
    return e, (-wo if flip_wo else wo), n


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


# ---- 2-D texture-mapped material attributes (reference src/Terra.c:368-466 sampling, :1748-1752 texcoord
#      interpolation, :1804-1810 attribute dispatch): a Cornell box whose wall albedos AND emissives are textures
def _uv_quad(scenes, p0, p1, p2, p3, uv0, uv1, uv2, uv3):
    """scenes.quad's two triangles (p0-p2-p1, p0-p3-p2) with per-vertex texcoords in TEXEL units."""
    tris = scenes.quad(p0, p1, p2, p3)
    uv = np.asarray([[uv0, uv2, uv1], [uv0, uv3, uv2]], np.float32)
    return tris, uv


def texture_images():
    rng = np.random.default_rng(23)
    tex8 = rng.integers(8, 256, (7, 5, 3), dtype=np.uint8)            # h, w, c; no zero texel: every emissive lookup lights
    texf = rng.uniform(0.05, 0.9, (6, 8, 3)).astype(np.float32)
    return tex8, texf


TEXTURED_CASES = {      # name -> scene kwargs; bounces 0 + jitter 0: the image is the emissive lookup at the first hit, no random numbers
    "textured_emissive": dict(integrator=0, spp=2, bounces=0, jitter=0.0),
}
TEXTURED_SIZE = (48, 40)


def textured_cornell(scenes, api, lib, *, integrator=0, spp=2, bounces=0, jitter=0.0, emissive_scale=1.0, commit=True):
    """Five walls, each its own object and its own (texture, filter, address mode) pair bound to BOTH albedo and
    emissive; texcoords chosen so that every address mode leaves its first period without touching the texels the
    reference's mirror arithmetic indexes one past the edge (Terra.c:385-386): in a mirrored period x % w and
    y % h stay away from 0."""
    s = api.Scene(lib)
    tex8, texf = texture_images()
    if emissive_scale != 1.0:
        texf = (texf * np.float32(emissive_scale)).astype(np.float32)
    F, A = api, api
    walls = [
        # floor: u8, point, wrap — three periods in u, two in v
        (((0, 0, 0), (1, 0, 0), (1, 0, 1), (0, 0, 1)), ((0.2, 0.3), (14.6, 0.3), (14.6, 13.4), (0.2, 13.4)), tex8, F.FILTER_POINT, A.ADDRESS_WRAP),
        # ceiling: f32, bilinear, clamp — runs past the last texel
        (((0, 1, 0), (0, 1, 1), (1, 1, 1), (1, 1, 0)), ((0.4, 0.2), (0.4, 9.7), (12.3, 9.7), (12.3, 0.2)), texf, F.FILTER_BILINEAR, A.ADDRESS_CLAMP),
        # back: u8, bilinear, wrap
        (((0, 0, 1), (1, 0, 1), (1, 1, 1), (0, 1, 1)), ((0.1, 0.1), (9.3, 0.1), (9.3, 12.8), (0.1, 12.8)), tex8, F.FILTER_BILINEAR, A.ADDRESS_WRAP),
        # left: f32, point, mirror — inside the SECOND (mirrored) period of the 8x6 texture, away from its edges
        (((0, 0, 0), (0, 0, 1), (0, 1, 1), (0, 1, 0)), ((9.2, 1.2), (14.8, 1.2), (14.8, 5.8), (9.2, 5.8)), texf, F.FILTER_POINT, A.ADDRESS_MIRROR),
        # right: u8, point, clamp
        (((1, 0, 0), (1, 1, 0), (1, 1, 1), (1, 0, 1)), ((0.0, 0.0), (0.0, 8.5), (6.5, 8.5), (6.5, 0.0)), tex8, F.FILTER_POINT, A.ADDRESS_CLAMP),
    ]
    for (p, uv, img, filt, addr) in walls:
        tris, tuv = _uv_quad(scenes, *p, *uv)
        o = s.add_object(tris, texcoords=tuv)
        s.set_diffuse(o, (0.5, 0.5, 0.5))
        tex = s.make_texture(img, filter=filt, address=addr)
        lib.terra_attribute_init_texture(C.byref(o.material.attributes[0]), C.byref(tex))
        lib.terra_attribute_init_texture(C.byref(o.material.emissive), C.byref(tex))
    # a plain box in the middle so that some first hits are untextured and the walls get shadowed bounce light
    s.set_diffuse(s.add_object(scenes.box((0.35, 0.0, 0.35), (0.65, 0.45, 0.65), rot_y=0.3)), (0.7, 0.7, 0.7))
    s.set_options(integrator=integrator, spp=spp, bounces=bounces, jitter=jitter)
    if commit:
        s.commit()
    cam = scenes.make_camera((0.5, 0.5, -1.35), (0, 0, 1), (0, 1, 0), 45.0)
    return s, cam
