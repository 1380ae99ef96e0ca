"""TEST INFRASTRUCTURE — ctypes bindings of the checkers: oracle/libterra_oracle.so (the CPU restatement,
terra_oracle.c) and the extra entry points ref_shim.c adds to oracle/_ref/libterra_ref.so (the unmodified
reference). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this
module; nothing in the product package (terra_b200/) does, and libterra_b200.so links nothing from oracle/.

Importing it registers two hooks with the product's ABI mirror (terra_b200/api.py):
  * TerraLib.post_load_hooks: a loaded reference library gets the ref_* signatures of ref_shim.c;
  * Scene.missing_preset_provider: scenes built against a library WITHOUT live GGX / glass presets (the reference
    carries them only as dead code) get the oracle's callbacks through the function-pointer table;
and gives api.Scene the ref_* helper methods that drive the reference shim."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

from terra_b200 import api
from terra_b200.api import (BSDF, Camera, Float2, Float3, Framebuffer, HIT_DTYPE, RESULT_DTYPE, SceneOptions, ShadingSurface,
                            Scene, TerraLib, _ptr, _sz, _vp)

REPO = Path(__file__).resolve().parent.parent
ORACLE_LIB = REPO / "oracle" / "libterra_oracle.so"
REFERENCE_LIB = REPO / "oracle" / "_ref" / "libterra_ref.so"
REFERENCE_ENVLIT_LIB = REPO / "oracle" / "_ref" / "libterra_ref_envlit.so"   # reference with Terra.c:1056 restored ("patched oracle")
REFERENCE_DEAD_LIB = REPO / "oracle" / "_ref" / "libterra_ref_dead.so"       # the reference's #if 0 GGX / glass presets (ref_dead_presets.c)


def _bind_reference_shim(lib, sig):
    if not lib.is_reference:
        return
    sig("ref_bvh_build", _vp, _vp, C.c_int, C.POINTER(C.c_int))
    sig("ref_bvh_free", None, _vp)
    sig("ref_bvh_trace", None, _vp, C.c_int, _vp, _sz, _vp, _vp, _vp)
    sig("ref_bvh_trace_timed", C.c_double, _vp, C.c_int, _vp, _sz, _vp, _vp, C.POINTER(_sz))
    sig("ref_primary_hits", None, C.POINTER(Camera), _vp, _vp, _sz, _sz, _vp, _vp)
    sig("ref_raycast_surface", C.c_int, _vp, _vp, _vp, _vp, C.POINTER(ShadingSurface), _vp, C.POINTER(C.c_int))
    sig("ref_ray_count", C.c_uint64)
    sig("ref_ray_count_reset", None)
    sig("ref_render_farm", C.c_double, C.POINTER(Camera), _vp, _sz, _sz, _sz, C.c_int, C.c_uint, _sz,
        _vp, _vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), _sz)


TerraLib.post_load_hooks.append(_bind_reference_shim)


class OracleLib:
    def __init__(self, path: os.PathLike | str | None = None):
        path = Path(path) if path is not None else ORACLE_LIB
        if not path.exists():
            raise FileNotFoundError(f"{path} is missing — run `make -C oracle`")
        self.lib = L = C.CDLL(str(path))

        def sig(name, res, *args):
            f = getattr(L, name)
            f.restype, f.argtypes = res, list(args)

        sig("oracle_scene_create", _vp, _vp, C.c_int, C.POINTER(SceneOptions))
        sig("oracle_scene_destroy", None, _vp)
        sig("oracle_set_env_lit", None, _vp, C.c_int)
        sig("oracle_bvh_nodes", _vp, _vp, C.POINTER(C.c_int))
        sig("oracle_trace_rays", None, _vp, _sz, _vp, _vp, _vp)
        sig("oracle_trace_counted", None, _vp, _vp, _sz, _vp, _vp, _vp, _vp)
        sig("oracle_primary_hits", None, _vp, C.POINTER(Camera), _sz, _sz, _vp, _vp)
        sig("oracle_render", None, _vp, C.POINTER(Camera), C.POINTER(Framebuffer), _sz, _sz, _sz, _sz, C.c_uint32)
        sig("oracle_raycast_surface", C.c_int, _vp, _vp, _vp, C.POINTER(ShadingSurface), _vp, C.POINTER(C.c_int))
        sig("oracle_surface_from_normal", None, _vp, _vp, C.POINTER(ShadingSurface))
        for p in ("diffuse", "phong", "ggx", "glass"):
            sig(f"terra_oracle_bsdf_{p}_init", None, C.POINTER(BSDF))
        sig("oracle_distribution_1d_init", C.c_float, _vp, _sz, _vp)
        sig("oracle_distribution_1d_sample", C.c_float, _vp, _vp, C.c_float, _sz, C.c_float, C.POINTER(C.c_float), C.POINTER(_sz))
        sig("oracle_distribution_2d_create", _vp, _vp, _sz, _sz)
        sig("oracle_distribution_2d_destroy", None, _vp)
        sig("oracle_distribution_2d_sample", None, _vp, _sz, _vp, _vp, _vp)
        sig("oracle_env_importance", None, _vp, _sz, _sz, _vp)

    def __getattr__(self, name):
        return getattr(self.lib, name)

    def distribution_2d_sample(self, f: np.ndarray, e1: np.ndarray, e2: np.ndarray) -> np.ndarray:
        """(n,4) = s1 (row axis), s2 (column axis), pdf, 0 for a (ny,nx) function (oracle restatement)."""
        f = np.ascontiguousarray(f, np.float32)
        e1, e2 = np.ascontiguousarray(e1, np.float32), np.ascontiguousarray(e2, np.float32)
        d = self.lib.oracle_distribution_2d_create(_ptr(f), f.shape[1], f.shape[0])
        out = np.zeros((e1.shape[0], 4), np.float32)
        self.lib.oracle_distribution_2d_sample(d, e1.shape[0], _ptr(e1), _ptr(e2), _ptr(out))
        self.lib.oracle_distribution_2d_destroy(d)
        return out

    def env_importance(self, rgb: np.ndarray) -> np.ndarray:
        rgb = np.ascontiguousarray(rgb, np.float32)
        out = np.zeros(rgb.shape[:2], np.float32)
        self.lib.oracle_env_importance(_ptr(rgb), rgb.shape[1], rgb.shape[0], _ptr(out))
        return out


class _RefDistribution1D(C.Structure):          # src/TerraPrivate.h:85-90
    _fields_ = [("f", C.POINTER(C.c_float)), ("n", C.c_size_t), ("cdf", C.POINTER(C.c_float)), ("integral", C.c_float)]


class _RefDistribution2D(C.Structure):          # src/TerraPrivate.h:93-96
    _fields_ = [("marginal", _RefDistribution1D), ("conditionals", C.POINTER(_RefDistribution1D))]


def reference_distribution_2d_sample(ref: "TerraLib", f: np.ndarray, e1: np.ndarray, e2: np.ndarray):
    """The reference's own terra_distribution_2d_init / _sample (exported by libterra_ref.so, called by nothing
    in the reference). Returns ((n,4) samples like OracleLib.distribution_2d_sample, marginal cdf, conditional cdfs)."""
    L = ref.lib
    L.terra_distribution_2d_init.restype, L.terra_distribution_2d_init.argtypes = None, [C.POINTER(_RefDistribution2D), _vp, _sz, _sz]
    L.terra_distribution_2d_sample.restype = Float2
    L.terra_distribution_2d_sample.argtypes = [C.POINTER(_RefDistribution2D), C.c_float, C.c_float, C.POINTER(C.c_float)]
    f = np.ascontiguousarray(f, np.float32)
    ny, nx = f.shape
    cond = (_RefDistribution1D * ny)()
    d = _RefDistribution2D()
    d.conditionals = cond                       # the caller provides the array (Terra.c:812-814 indexes it)
    L.terra_distribution_2d_init(C.byref(d), _ptr(f), nx, ny)
    out = np.zeros((len(e1), 4), np.float32)
    pdf = C.c_float()
    for k in range(len(e1)):
        r = L.terra_distribution_2d_sample(C.byref(d), float(e1[k]), float(e2[k]), C.byref(pdf))
        out[k, 0], out[k, 1], out[k, 2] = r.x, r.y, pdf.value
    marg = np.ctypeslib.as_array(d.marginal.cdf, (ny,)).copy()
    ccdf = np.stack([np.ctypeslib.as_array(cond[i].cdf, (nx,)).copy() for i in range(ny)])
    return out, marg, ccdf


class OracleScene:
    """The oracle's view of a scene built through either Terra-ABI library (it reads the
    caller-owned TerraObject array in place)."""

    def __init__(self, oracle: OracleLib, scene: Scene):
        self.o, self.scene = oracle, scene
        self.h = oracle.oracle_scene_create(scene.objects_base, scene.n_objects, C.byref(scene.options()))

    def destroy(self):
        if self.h:
            self.o.oracle_scene_destroy(self.h)
            self.h = None

    def set_env_lit(self, on: bool = True):
        self.o.oracle_set_env_lit(self.h, 1 if on else 0)

    def bvh_nodes(self) -> np.ndarray:
        n = C.c_int(0)
        p = self.o.oracle_bvh_nodes(self.h, C.byref(n))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), (n.value * 64,)).copy()

    def trace_rays(self, origins, dirs) -> np.ndarray:
        o = np.ascontiguousarray(origins, np.float32)
        d = np.ascontiguousarray(dirs, np.float32)
        hits = np.empty(o.shape[0], HIT_DTYPE)
        self.o.oracle_trace_rays(self.h, o.shape[0], _ptr(o), _ptr(d), _ptr(hits))
        return hits

    def trace_counted(self, origins, dirs, nodes: np.ndarray | None = None):
        """Canonical near-first culled walk; `nodes` = raw 64-byte node array (default: reference tree).
        Returns (hits, nodes_fetched, triangles_tested)."""
        o = np.ascontiguousarray(origins, np.float32)
        d = np.ascontiguousarray(dirs, np.float32)
        hits = np.empty(o.shape[0], HIT_DTYPE)
        cnt = np.zeros(2, np.uint64)
        self.o.oracle_trace_counted(self.h, _ptr(nodes) if nodes is not None else None, o.shape[0], _ptr(o), _ptr(d),
                                    _ptr(hits), _ptr(cnt))
        return hits, int(cnt[0]), int(cnt[1])

    def primary_hits(self, cam: Camera, width: int, height: int):
        hits = np.empty(width * height, HIT_DTYPE)
        dirs = np.empty((width * height, 3), np.float32)
        self.o.oracle_primary_hits(self.h, C.byref(cam), width, height, _ptr(hits), _ptr(dirs))
        return hits.reshape(height, width), dirs.reshape(height, width, 3)

    def render(self, cam: Camera, width: int, height: int, pcg_seed: int = 1, region=None):
        res = np.zeros(width * height, RESULT_DTYPE)
        pix = np.zeros((height, width, 3), np.float32)
        fb = Framebuffer(_ptr(pix), _ptr(res), width, height)
        x, y, w, h = region or (0, 0, width, height)
        self.o.oracle_render(self.h, C.byref(cam), C.byref(fb), x, y, w, h, pcg_seed)
        return res.reshape(height, width), pix


# TerraBSDF callback prototypes (include/terra_b200.h; reference include/Terra.h:40-54)
BSDF_SAMPLE = C.CFUNCTYPE(Float3, C.POINTER(ShadingSurface), C.c_float, C.c_float, C.c_float, C.POINTER(Float3))
BSDF_PDF = C.CFUNCTYPE(C.c_float, C.POINTER(ShadingSurface), C.POINTER(Float3), C.POINTER(Float3))
BSDF_EVAL = C.CFUNCTYPE(Float3, C.POINTER(ShadingSurface), C.POINTER(Float3), C.POINTER(Float3))


def run_bsdf_callbacks(oracle: "OracleLib", init_fn, attributes, e123, wo, normal) -> np.ndarray:
    """Calls the host {sample, pdf, eval} callbacks that `init_fn` installs, in terra_trace's order
    (sample -> pdf -> eval on one surface record, Terra.c:1068-1075), for n input tuples.
    attributes: up to 8 rgb triples. Returns (n,8) float32 = wi.xyz, pdf, f.xyz, attributes[3].x."""
    b = BSDF()
    init_fn(C.byref(b))
    sample, pdf, ev = BSDF_SAMPLE(b.sample), BSDF_PDF(b.pdf), BSDF_EVAL(b.eval)
    attrs = np.zeros((8, 3), np.float32)
    a = np.asarray(attributes, np.float32).reshape(-1, 3)
    attrs[:a.shape[0]] = a
    n = len(e123)
    out = np.zeros((n, 8), np.float32)
    for i in range(n):
        sf = ShadingSurface()
        nrm = np.ascontiguousarray(normal[i], np.float32)
        oracle.oracle_surface_from_normal(_ptr(nrm), _ptr(attrs), C.byref(sf))
        w = Float3(*[float(x) for x in wo[i]])
        wi = sample(C.byref(sf), float(e123[i][0]), float(e123[i][1]), float(e123[i][2]), C.byref(w))
        p = pdf(C.byref(sf), C.byref(wi), C.byref(w))
        f = ev(C.byref(sf), C.byref(wi), C.byref(w))
        out[i] = (wi.x, wi.y, wi.z, p, f.x, f.y, f.z, sf.attributes[3].x)
    return out


# ---- helpers on api.Scene that drive the reference shim (lib.is_reference only)
# -- reference-shim helpers -------------------------------------------------------------
def ref_primary_hits(self, cam: Camera, width: int, height: int):
    hits = np.empty(width * height, HIT_DTYPE)
    dirs = np.empty((width * height, 3), np.float32)
    self.lib.ref_primary_hits(C.byref(cam), self.h, self.objects_base, width, height, _ptr(hits), _ptr(dirs))
    return hits.reshape(height, width), dirs.reshape(height, width, 3)

def ref_bvh(self):
    n = C.c_int(0)
    p = self.lib.ref_bvh_build(self.objects_base, self.n_objects, C.byref(n))
    return p, n.value

def ref_bvh_nodes(self) -> np.ndarray:
    p, n = self.ref_bvh()
    a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), (n * 64,)).copy()
    self.lib.ref_bvh_free(p)
    return a

def ref_trace_rays(self, origins, dirs, bvh=None) -> np.ndarray:
    o = np.ascontiguousarray(origins, np.float32)
    d = np.ascontiguousarray(dirs, np.float32)
    hits = np.empty(o.shape[0], HIT_DTYPE)
    p, n = bvh or self.ref_bvh()
    self.lib.ref_bvh_trace(p, n, self.objects_base, o.shape[0], _ptr(o), _ptr(d), _ptr(hits))
    if bvh is None:
        self.lib.ref_bvh_free(p)
    return hits

def ref_render_farm(self, cam: Camera, width, height, spp, tile=128, nproc=1, seed=1, tile_stride=1):
    res = np.zeros(width * height, RESULT_DTYPE)
    pix = np.zeros((height, width, 3), np.float32)
    rays, paths = C.c_uint64(0), C.c_uint64(0)
    secs = self.lib.ref_render_farm(C.byref(cam), self.h, width, height, tile, nproc, seed, tile_stride,
                                    _ptr(res), _ptr(pix), C.byref(rays), C.byref(paths), spp)
    return secs, rays.value, paths.value, res.reshape(height, width), pix


for _f in (ref_primary_hits, ref_bvh, ref_bvh_nodes, ref_trace_rays, ref_render_farm):
    setattr(Scene, _f.__name__, _f)

_oracle_for_presets = None


def _oracle_preset(name):
    global _oracle_for_presets
    if _oracle_for_presets is None:
        _oracle_for_presets = OracleLib()
    return getattr(_oracle_for_presets, f"terra_oracle_bsdf_{name}_init")


Scene.missing_preset_provider = staticmethod(_oracle_preset)


def harness_api():
    """terra_b200.api plus the checker bindings of this module as ONE namespace: what tests/, bench.py's CPU legs and
    smoke() use as `api` (the product package itself never sees the names added here)."""
    import types
    ns = types.ModuleType("terra_b200_harness_api")
    ns.__dict__.update({k: v for k, v in vars(api).items() if not k.startswith("__")})
    for k in ("ORACLE_LIB", "REFERENCE_LIB", "REFERENCE_ENVLIT_LIB", "REFERENCE_DEAD_LIB", "OracleLib", "OracleScene", "run_bsdf_callbacks",
              "reference_distribution_2d_sample", "BSDF_SAMPLE", "BSDF_PDF", "BSDF_EVAL"):
        setattr(ns, k, globals()[k])
    return ns
