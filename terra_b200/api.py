"""ctypes mirror of the C ABI in include/terra_b200.h.

The binding class loads ANY library with the reference's C ABI (reference include/Terra.h:36-245): the product
(libterra_b200.so, CUDA) and — from the test harness, oracle/binding.py — the unmodified reference. That one class
drives either library *is* the drop-in property the parity tests exercise. Nothing in this package knows about the
oracle or the reference build: the harness registers what it needs through TerraLib.post_load_hooks and
Scene.missing_preset_provider.

Nothing here computes anything: it is plumbing between numpy buffers and the C ABI.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
PRODUCT_LIB = Path(os.environ.get("TERRA_B200_LIB", REPO / "terra_b200" / "libterra_b200.so"))   # override: A/B builds only

MAX_ATTR = 8


class Float2(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float)]


class Float3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]

    def __init__(self, x=0.0, y=0.0, z=0.0):
        super().__init__(float(x), float(y), float(z))

    def tuple(self):
        return (self.x, self.y, self.z)


class Float4(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float), ("w", C.c_float)]


class Float4x4(C.Structure):
    _fields_ = [("rows", Float4 * 4)]


class ShadingSurface(C.Structure):
    _fields_ = [("transform", Float4x4), ("normal", Float3), ("emissive", Float3),
                ("ior", C.c_float), ("attributes", Float3 * MAX_ATTR)]


class BSDF(C.Structure):
    _fields_ = [("sample", C.c_void_p), ("pdf", C.c_void_p), ("eval", C.c_void_p)]


class Texture(C.Structure):
    _fields_ = [("pixels", C.c_void_p), ("width", C.c_uint16), ("height", C.c_uint16),
                ("components", C.c_uint8), ("depth", C.c_uint8), ("filter", C.c_uint8),
                ("address_mode", C.c_uint8)]


class Attribute(C.Structure):
    _fields_ = [("state", C.c_void_p), ("finalize", C.c_void_p), ("eval", C.c_void_p), ("value", Float3)]


class Material(C.Structure):
    _fields_ = [("bsdf", BSDF), ("ior", C.c_float), ("emissive", Attribute),
                ("attributes", Attribute * MAX_ATTR), ("attributes_count", C.c_size_t),
                ("enable_bump_map_attr", C.c_bool), ("enable_normal_map_attr", C.c_bool)]


class Object(C.Structure):
    _fields_ = [("triangles", C.c_void_p), ("properties", C.c_void_p),
                ("triangles_count", C.c_size_t), ("material", Material)]


class SceneOptions(C.Structure):
    _fields_ = [("environment_map", Attribute), ("tonemapping_operator", C.c_int),
                ("accelerator", C.c_int), ("sampling_method", C.c_int), ("integrator", C.c_int),
                ("subpixel_jitter", C.c_float), ("samples_per_pixel", C.c_size_t),
                ("bounces", C.c_size_t), ("strata", C.c_size_t),
                ("manual_exposure", C.c_float), ("gamma", C.c_float)]


class Camera(C.Structure):
    _fields_ = [("position", Float3), ("direction", Float3), ("up", Float3), ("fov", C.c_float)]


class Framebuffer(C.Structure):
    _fields_ = [("pixels", C.c_void_p), ("results", C.c_void_p), ("width", C.c_size_t), ("height", C.c_size_t)]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("rays_traced", C.c_uint64), ("paths_traced", C.c_uint64),
                ("trace_kernel_ms", C.c_double), ("trace_kernel_launches", C.c_uint64),
                ("render_ms", C.c_double), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64), ("graph_replays", C.c_uint64),
                ("coherent_rays", C.c_uint64), ("coherent_kernel_ms", C.c_double), ("coherent_kernel_launches", C.c_uint64),
                ("coherent_nodes_visited", C.c_uint64), ("coherent_tris_tested", C.c_uint64), ("p2p_bytes", C.c_uint64)]


HIT_DTYPE = np.dtype([("t", "<f4"), ("prim", "<u4"), ("u", "<f4"), ("v", "<f4")])
RESULT_DTYPE = np.dtype([("acc", "<f4", (3,)), ("samples", "<i4")])

assert C.sizeof(ShadingSurface) == 188 and C.sizeof(Material) == 408 and C.sizeof(Object) == 432
assert C.sizeof(SceneOptions) == 96 and C.sizeof(Camera) == 40 and C.sizeof(Framebuffer) == 32
assert C.sizeof(Attribute) == 40 and C.sizeof(Texture) == 16

# enums (reference include/Terra.h:131-157)
TONEMAP_NONE, TONEMAP_LINEAR, TONEMAP_REINHARD, TONEMAP_FILMIC, TONEMAP_UNCHARTED2 = range(5)
SAMPLING_RANDOM, SAMPLING_STRATIFIED, SAMPLING_HALTON = range(3)
(INTEGRATOR_SIMPLE, INTEGRATOR_DIRECT, INTEGRATOR_DIRECT_MIS, INTEGRATOR_DEBUG_MONO,
 INTEGRATOR_DEBUG_DEPTH, INTEGRATOR_DEBUG_NORMALS, INTEGRATOR_DEBUG_MIS_WEIGHTS) = range(7)
FILTER_POINT, FILTER_BILINEAR, FILTER_TRILINEAR, FILTER_ANISOTROPIC = range(4)
ADDRESS_WRAP, ADDRESS_MIRROR, ADDRESS_CLAMP = range(3)

_vp, _sz = C.c_void_p, C.c_size_t
BVH_REFERENCE, BVH_SAH3 = 0, 1


class TerraLib:
    """One loaded Terra-ABI shared library (product or reference)."""
    post_load_hooks: list = []

    def __init__(self, path: os.PathLike | str | None = None):
        path = Path(path) if path is not None else PRODUCT_LIB
        if not path.exists():
            raise FileNotFoundError(
                f"{path} is missing — build it first (python -c 'import __graft_entry__ as g; g.build()'). "
                "There is no CPU fallback for the product library.")
        self.path = path
        self.lib = L = C.CDLL(str(path), mode=C.RTLD_GLOBAL if False else C.RTLD_LOCAL)
        self.is_product = hasattr(L, "terra_b200_device_count")
        self.is_reference = hasattr(L, "ref_bvh_build")

        older_build = bool(os.environ.get("TERRA_B200_LIB"))     # A/B runs against a previous build of the library

        def sig(name, res, *args):
            if older_build and not hasattr(L, name):
                return None                     # an entry point that build does not have yet: calling it raises
            f = getattr(L, name)
            f.restype, f.argtypes = res, list(args)
            return f

        sig("terra_scene_create", _vp)
        sig("terra_scene_add_object", C.POINTER(Object), _vp, _sz)
        sig("terra_scene_count_objects", _sz, _vp)
        sig("terra_scene_commit", None, _vp)
        sig("terra_scene_clear", None, _vp)
        sig("terra_scene_get_options", C.POINTER(SceneOptions), _vp)
        sig("terra_scene_destroy", None, _vp)
        sig("terra_framebuffer_create", C.c_bool, C.POINTER(Framebuffer), _sz, _sz)
        sig("terra_framebuffer_clear", None, C.POINTER(Framebuffer))
        sig("terra_framebuffer_destroy", None, C.POINTER(Framebuffer))
        sig("terra_texture_init", C.c_bool, C.POINTER(Texture), _sz, _sz, _sz, _vp)
        sig("terra_texture_init_hdr", C.c_bool, C.POINTER(Texture), _sz, _sz, _sz, _vp)
        sig("terra_texture_read", Float3, C.POINTER(Texture), _sz, _sz)
        sig("terra_texture_sample", Float3, _vp, _vp, _vp)
        sig("terra_texture_sample_latlong", Float3, _vp, _vp, _vp)
        sig("terra_texture_destroy", None, C.POINTER(Texture))
        sig("terra_attribute_init_constant", None, C.POINTER(Attribute), C.POINTER(Float3))
        sig("terra_attribute_init_texture", None, C.POINTER(Attribute), C.POINTER(Texture))
        sig("terra_attribute_init_cubemap", None, C.POINTER(Attribute), C.POINTER(Texture))
        sig("terra_render", None, C.POINTER(Camera), _vp, C.POINTER(Framebuffer), _sz, _sz, _sz, _sz)
        sig("terra_bsdf_diffuse_init", None, C.POINTER(BSDF))
        sig("terra_bsdf_phong_init", None, C.POINTER(BSDF))
        if self.is_product:
            sig("terra_bsdf_ggx_init", None, C.POINTER(BSDF))
            sig("terra_bsdf_glass_init", None, C.POINTER(BSDF))
            sig("terra_b200_device_count", C.c_int)
            sig("terra_b200_set_device", None, C.c_int)
            sig("terra_b200_scene_set_devices", None, _vp, C.POINTER(C.c_int), C.c_int)
            sig("terra_b200_scene_device_count", C.c_int, _vp)
            sig("terra_b200_set_fanout_min_paths", None, _vp, _sz)
            sig("terra_b200_set_fanout_chunk_paths", None, _vp, _sz)
            sig("terra_b200_probe_bandwidth", C.c_double, _sz, C.c_int)
            sig("terra_b200_set_env_lit", None, _vp, C.c_int)
            sig("terra_b200_render_async", _vp, C.POINTER(Camera), _vp, C.POINTER(Framebuffer), _sz, _sz, _sz, _sz, _sz, _vp, _vp, _vp)
            sig("terra_b200_job_poll", C.c_int, _vp)
            sig("terra_b200_job_wait", None, _vp)
            sig("terra_b200_job_cancel", None, _vp)
            sig("terra_b200_job_release", None, _vp)
            sig("terra_b200_distribution_2d_sample", None, _vp, _vp, _sz, _sz, _sz, _vp, _vp, _vp)
            sig("terra_b200_set_graph_paths", None, _vp, _sz)
            sig("terra_b200_refit", C.c_float, _vp)
            sig("terra_b200_host_alloc", _vp, _sz)
            sig("terra_b200_host_free", None, _vp)
            sig("terra_b200_commit_ms", None, _vp, C.POINTER(C.c_float))
            sig("terra_b200_lbvh_build_ms", None, C.POINTER(C.c_float), C.POINTER(C.c_float))
            sig("terra_b200_set_seed", None, _vp, C.c_uint64)
            sig("terra_b200_get_stats", None, _vp, C.POINTER(Stats))
            sig("terra_b200_reset_stats", None, _vp)
            sig("terra_b200_round_stats", None, _vp, C.POINTER(C.c_uint64))
            sig("terra_b200_incoherent_kernel", C.c_char_p, _vp)
            sig("terra_b200_set_counting", None, _vp, C.c_int)
            sig("terra_b200_bvh_nodes", _vp, _vp, C.POINTER(C.c_int))
            sig("terra_b200_bvh_stack_bound", C.c_int, _vp)
            sig("terra_b200_build_host_bvh", C.c_int, _vp, C.c_int, _vp, _vp)
            sig("terra_b200_set_bvh_mode", None, C.c_int)
            sig("terra_b200_build_host_bvh4", C.c_int, _vp, C.c_int, _vp, _vp, _vp, C.c_int, _vp)
            sig("terra_b200_trace_rays", None, _vp, _sz, _vp, _vp, _vp)
            sig("terra_b200_trace_rays_device", None, _vp, _sz, _vp, _vp, _vp, _vp)
            sig("terra_b200_trace_rays_packed_noalloc", C.c_int, _vp, _sz, _vp, _vp, _vp, _vp)
            sig("terra_b200_primary_hits", None, C.POINTER(Camera), _vp, _sz, _sz, _vp)
            sig("terra_b200_render_device", None, C.POINTER(Camera), _vp, _sz, _sz, _sz, _sz, _sz, _sz, _sz, _sz, _vp, _vp)
            sig("terra_b200_resolve_device", None, _vp, _sz, _sz, _vp, _vp, _vp)
            sig("terra_b200_bsdf_kat", None, _vp, _sz, _sz, _vp, _vp, _vp, _vp)
        for hook in TerraLib.post_load_hooks:      # test infrastructure binds the reference shim's extra entry points here
            hook(self, sig)

    def __getattr__(self, name):
        return getattr(self.lib, name)


def _ptr(a: np.ndarray) -> int:
    return a.ctypes.data


TILE_EVENT = C.CFUNCTYPE(None, _sz, _sz, _sz, _sz, _vp)      # TerraB200TileEvent


class PinnedArray:
    """numpy view of page-locked host memory from terra_b200_host_alloc (freed with the object).
    `a = PinnedArray(lib, (n, 3), np.float32); a.array[...] = ...`"""

    def __init__(self, lib: "TerraLib", shape, dtype):
        dtype = np.dtype(dtype)
        shape = (shape,) if isinstance(shape, int) else tuple(shape)
        nbytes = int(np.prod(shape)) * dtype.itemsize
        self._lib = lib
        self._p = lib.lib.terra_b200_host_alloc(max(1, nbytes))
        if not self._p:
            raise MemoryError("terra_b200_host_alloc: no CUDA device or no memory to page-lock")
        buf = (C.c_uint8 * max(1, nbytes)).from_address(self._p)
        self.array = np.frombuffer(buf, dtype=np.uint8, count=nbytes).view(dtype).reshape(shape)

    def free(self):
        if self._p:
            self.array = None
            self._lib.lib.terra_b200_host_free(self._p)
            self._p = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Scene:
    """Scene built through the public API only (reference call sequence)."""
    default_preset_overrides: dict = {}
    missing_preset_provider = None

    def __init__(self, lib: TerraLib):
        self.lib = lib
        self.h = lib.terra_scene_create()
        self.objects_base = None          # &objects[0]
        self.n_objects = 0
        self._keep = []                   # textures etc. referenced by raw pointer from attributes
        self.preset_overrides = dict(Scene.default_preset_overrides)   # name -> init function (tests plug the oracle's in)

    # -- construction -------------------------------------------------------------------
    def add_object(self, tris: np.ndarray, normals: np.ndarray | None = None,
                   texcoords: np.ndarray | None = None) -> Object:
        """tris: (n,3,3) float32 vertices; normals: (n,3,3) per-vertex; texcoords: (n,3,2)."""
        tris = np.ascontiguousarray(tris, dtype=np.float32)
        n = tris.shape[0]
        assert tris.shape == (n, 3, 3)
        if normals is None:               # flat geometric normals
            fn = np.cross(tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0])
            fn /= np.maximum(np.linalg.norm(fn, axis=1, keepdims=True), 1e-30)
            normals = np.repeat(fn[:, None, :], 3, axis=1)
        normals = np.ascontiguousarray(normals, dtype=np.float32)
        if texcoords is None:
            texcoords = np.zeros((n, 3, 2), np.float32)
        props = np.concatenate([normals.reshape(n, 9), np.asarray(texcoords, np.float32).reshape(n, 6)], axis=1)
        props = np.ascontiguousarray(props, dtype=np.float32)
        optr = self.lib.terra_scene_add_object(self.h, n)
        obj = optr.contents
        # &objects[0]: the array moves when it grows past its capacity (64, 128, ...), so follow it on every add
        # (and fill an object in before adding the next one: earlier Object references dangle after a move)
        self.objects_base = C.addressof(obj) - self.n_objects * C.sizeof(Object)
        assert obj.triangles_count == n
        C.memmove(obj.triangles, _ptr(tris), tris.nbytes)
        C.memmove(obj.properties, _ptr(props), props.nbytes)
        self.n_objects += 1
        return obj

    def const_attr(self, attr: Attribute, rgb):
        v = Float3(*rgb)
        self.lib.terra_attribute_init_constant(C.byref(attr), C.byref(v))

    def set_diffuse(self, obj: Object, albedo, emissive=(0, 0, 0)):
        self._preset_init("diffuse", self.preset_overrides.get("diffuse"))(C.byref(obj.material.bsdf))
        self.const_attr(obj.material.attributes[0], albedo)
        obj.material.attributes_count = 1
        self.const_attr(obj.material.emissive, emissive)
        obj.material.ior = 1.5

    def set_phong(self, obj: Object, albedo, specular, exponent, emissive=(0, 0, 0)):
        # slot order: reference include/TerraPresets.h:20-24, satellite/src/Scene.cpp:196-213
        self._preset_init("phong", self.preset_overrides.get("phong"))(C.byref(obj.material.bsdf))
        self.const_attr(obj.material.attributes[0], specular)
        self.const_attr(obj.material.attributes[1], albedo)
        self.const_attr(obj.material.attributes[2], (exponent, exponent, exponent))
        self.const_attr(obj.material.attributes[3], (0, 0, 0))
        obj.material.attributes_count = 4
        self.const_attr(obj.material.emissive, emissive)
        obj.material.ior = 1.5

    def _preset_init(self, name, override):
        if override is not None:
            return override
        if self.lib.is_product or name in ("diffuse", "phong"):
            return getattr(self.lib, f"terra_bsdf_{name}_init")
        # a library without live GGX / glass presets (the reference carries them only as dead code): whoever drives
        # it supplies the callbacks (tests and the bench's CPU leg plug the oracle's in, oracle/binding.py)
        if Scene.missing_preset_provider is None:
            raise RuntimeError(f"{self.lib.path.name} has no terra_bsdf_{name}_init and no provider is registered")
        return Scene.missing_preset_provider(name)

    def set_ggx(self, obj: Object, albedo, roughness, metalness, ior=1.5, emissive=(0, 0, 0), bsdf_init=None):
        self._preset_init("ggx", bsdf_init)(C.byref(obj.material.bsdf))
        self.const_attr(obj.material.attributes[0], albedo)
        self.const_attr(obj.material.attributes[1], (roughness, metalness, ior))
        obj.material.attributes_count = 2
        self.const_attr(obj.material.emissive, emissive)
        obj.material.ior = ior

    def set_glass(self, obj: Object, albedo, ior=1.5, bsdf_init=None):
        self._preset_init("glass", bsdf_init)(C.byref(obj.material.bsdf))
        self.const_attr(obj.material.attributes[0], albedo)
        self.const_attr(obj.material.attributes[1], (ior, 0, 0))
        self.const_attr(obj.material.attributes[2], (0, 0, 0))
        obj.material.attributes_count = 3
        self.const_attr(obj.material.emissive, (0, 0, 0))
        obj.material.ior = ior

    def make_texture(self, data: np.ndarray, filter=FILTER_POINT, address=ADDRESS_WRAP) -> Texture:
        """data: (h,w,c) uint8 or float32."""
        data = np.ascontiguousarray(data)
        h, w, c = data.shape
        tex = Texture()
        if data.dtype == np.uint8:
            self.lib.terra_texture_init(C.byref(tex), w, h, c, _ptr(data))
            tex.depth = 1                 # reference falls off the end without return; fields are set
        else:
            data = data.astype(np.float32)
            self.lib.terra_texture_init_hdr(C.byref(tex), w, h, c, _ptr(data))
            tex.depth = 4
        tex.filter, tex.address_mode = filter, address   # caller's job in the reference
        self._keep.append(tex)
        return tex

    def options(self) -> SceneOptions:
        return self.lib.terra_scene_get_options(self.h).contents

    def set_options(self, *, integrator=INTEGRATOR_SIMPLE, spp=16, bounces=8, jitter=0.5,
                    tonemap=TONEMAP_NONE, exposure=1.0, gamma=2.2, sampling=SAMPLING_RANDOM,
                    strata=4, env=(0, 0, 0)):
        o = self.options()
        if env is not None:
            self.const_attr(o.environment_map, env)
        o.tonemapping_operator, o.accelerator = tonemap, 0
        o.sampling_method, o.integrator = sampling, integrator
        o.subpixel_jitter, o.samples_per_pixel, o.bounces, o.strata = jitter, spp, bounces, strata
        o.manual_exposure, o.gamma = exposure, gamma

    def commit(self):
        self.lib.terra_scene_commit(self.h)

    def destroy(self):
        if self.h:
            self.lib.terra_scene_destroy(self.h)
            self.h = None

    # -- rendering ----------------------------------------------------------------------
    def render(self, cam: Camera, width: int, height: int, passes: int = 1, region=None):
        """Full-frame (or region) terra_render; returns (results structured array, pixels (h,w,3))."""
        fb = Framebuffer()
        assert self.lib.terra_framebuffer_create(C.byref(fb), width, height)
        x, y, w, h = region or (0, 0, width, height)
        for _ in range(passes):
            self.lib.terra_render(C.byref(cam), self.h, C.byref(fb), x, y, w, h)
        res = np.ctypeslib.as_array(C.cast(fb.results, C.POINTER(C.c_uint8)), (width * height * 16,)).view(RESULT_DTYPE).copy()
        pix = np.ctypeslib.as_array(C.cast(fb.pixels, C.POINTER(C.c_float)), (height, width, 3)).copy()
        self.lib.terra_framebuffer_destroy(C.byref(fb))
        return res.reshape(height, width), pix

    # -- product extensions ---------------------------------------------------------------
    def trace_rays(self, origins: np.ndarray, dirs: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        """Closest hits of a host ray batch. Page-locked arrays (`PinnedArray`) for origins / dirs / out make
        the copies overlap the traversal; plain numpy arrays give the same hits."""
        o = np.ascontiguousarray(origins, np.float32)
        d = np.ascontiguousarray(dirs, np.float32)
        hits = np.empty(o.shape[0], HIT_DTYPE) if out is None else out
        assert hits.dtype == HIT_DTYPE and hits.shape[0] == o.shape[0] and hits.flags.c_contiguous
        self.lib.terra_b200_trace_rays(self.h, o.shape[0], _ptr(o), _ptr(d), _ptr(hits))
        return hits

    def primary_hits(self, cam: Camera, width: int, height: int) -> np.ndarray:
        hits = np.empty(width * height, HIT_DTYPE)
        self.lib.terra_b200_primary_hits(C.byref(cam), self.h, width, height, _ptr(hits))
        return hits.reshape(height, width)

    def stats(self) -> Stats:
        s = Stats()
        self.lib.terra_b200_get_stats(self.h, C.byref(s))
        return s

    def build_host_bvh(self, n_tris: int, mode: int = 0):
        """Host-only BVH build (mode 0 = reference tree, 1 = sah3) + flatten.
        Returns (flat nodes (n-1,16) f32, flat tris (n,12) f32)."""
        nodes = np.zeros((n_tris * 2, 16), np.float32)
        tris = np.zeros((n_tris, 12), np.float32)
        n = self.lib.terra_b200_build_host_bvh(self.h, mode, _ptr(nodes), _ptr(tris))
        return nodes[:n].copy(), tris

    def bvh_nodes(self) -> np.ndarray:
        n = C.c_int(0)
        p = self.lib.terra_b200_bvh_nodes(self.h, C.byref(n))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), (n.value * 64,)).copy()


def make_camera(position, direction, up=(0, 1, 0), fov=45.0) -> Camera:
    return Camera(Float3(*position), Float3(*direction), Float3(*up), float(fov))


_libc = None


def srand(seed: int):
    """libc srand(): the reference (and the oracle) draw BSDF/roulette numbers from rand()."""
    global _libc
    if _libc is None:
        _libc = C.CDLL(None)
    _libc.srand(C.c_uint(seed))
