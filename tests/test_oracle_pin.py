"""Pins the oracle (oracle/terra_oracle.c, the CPU restatement) before anything trusts it.

Two anchors, both outputs of the UNMODIFIED reference:
  * tests/golden/*.npz — committed fixtures produced by tests/golden/make_golden.py;
  * oracle/_ref/libterra_ref.so — the reference itself, when built (tests using `ref` skip otherwise).
Everything here is CPU-only. Bit-exact means equal float bit patterns (compared as uint32).
Scenes are built through the PRODUCT's public API without commit (the oracle reads the
caller-owned object array in place), so the same files also exercise the product's scene container."""
import ctypes as C
import hashlib

import numpy as np
import pytest

import cases
from conftest import golden


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def assert_hits_equal(a, b):
    assert np.array_equal(a["prim"], b["prim"])
    for f in ("t", "u", "v"):
        assert np.array_equal(bits(a[f]), bits(b[f])), f


@pytest.fixture(scope="module")
def geo():
    return golden("geometry.npz")


@pytest.fixture(scope="module", params=list(cases.PRIMARY))
def host_case(request, api, scenes, prod, oracle):
    s, cam, wh = cases.build_host(scenes, api, prod, request.param)
    osc = api.OracleScene(oracle, s)
    yield request.param, s, cam, wh, osc
    osc.destroy()
    s.destroy()


# ----------------------------------------------------------------------------- geometry vs fixtures
def test_bvh_nodes_match_golden(host_case, geo):
    """terra_bvh_create (TerraBVH.c:128-244) restated: byte-identical node array."""
    name, s, cam, wh, osc = host_case
    nodes = osc.bvh_nodes()
    assert nodes.size // 64 == int(geo[f"{name}_nodes_count"])
    assert hashlib.sha256(nodes.tobytes()).digest() == geo[f"{name}_nodes_sha256"].tobytes()
    if f"{name}_nodes" in geo.files:
        assert np.array_equal(nodes, geo[f"{name}_nodes"])


def test_primary_hits_match_golden(host_case, geo):
    """camera ray (Terra.c:1783-1799) + terra_scene_raycast (:1623-1657): id, t, u, v bit-exact."""
    name, s, cam, (w, h), osc = host_case
    hits, dirs = osc.primary_hits(cam, w, h)
    assert np.array_equal(bits(dirs), bits(geo[f"{name}_primary_dirs"]))
    assert_hits_equal(hits, geo[f"{name}_primary"])


def test_incoherent_hits_match_golden(host_case, geo, api, scenes):
    name, s, cam, wh, osc = host_case
    o, d = cases.rays_for(s, api, scenes, 4096)
    assert_hits_equal(osc.trace_rays(o, d), geo[f"{name}_rays_hits"])


def test_canonical_walk_agrees_with_reference_walk(host_case, geo, api, scenes):
    """The near-first, t-culled counter returns the reference's closest hit; ids may
    differ only on exact-t ties (first-found vs smallest id)."""
    name, s, cam, wh, osc = host_case
    o, d = cases.rays_for(s, api, scenes, 4096)
    hits, n_nodes, n_tris = osc.trace_counted(o, d)
    g = geo[f"{name}_rays_hits"]
    assert np.array_equal(bits(hits["t"]), bits(g["t"]))
    differ = hits["prim"] != g["prim"]
    assert differ.mean() < 0.01
    assert n_nodes > 0 and n_tris > 0


def test_surface_records_match_golden(host_case, geo, api, scenes):
    """terra_surface_init (Terra.c:1726-1764) + terra_f4x4_basis (TerraMath.inl:251-272)."""
    name, s, cam, wh, osc = host_case
    o, d = cases.rays_for(s, api, scenes, 4096)
    g = geo[f"{name}_surfaces"]
    objs = (api.Object * s.n_objects).from_address(s.objects_base)
    for i in range(64):
        sf = api.ShadingSurface()
        p = np.zeros(3, np.float32)
        tri = C.c_int(0)
        oi, di = np.ascontiguousarray(o[i]), np.ascontiguousarray(d[i])
        obj = osc.o.oracle_raycast_surface(osc.h, oi.ctypes.data, di.ctypes.data, C.byref(sf), p.ctypes.data, C.byref(tri))
        assert obj == int(g[i, 0])
        if obj < 0:
            continue
        assert tri.value == int(g[i, 1])
        raw = np.frombuffer(bytes(sf), np.float32)
        n_attr = min(4, int(objs[obj].material.attributes_count))
        assert np.array_equal(bits(p), bits(g[i, 2:5]))
        assert np.array_equal(bits(raw[0:12]), bits(g[i, 5:17]))          # 3 rows of the frame
        assert np.array_equal(bits(raw[16:22]), bits(g[i, 21:27]))        # normal, emissive
        assert np.array_equal(bits(raw[23:23 + 3 * n_attr]), bits(g[i, 27:27 + 3 * n_attr]))


def test_triangle_known_answers(api, prod, oracle):
    """Edge / vertex / co-planar / axis-aligned rays: the fp64 retry (TerraGeometry.c:204-208), the
    +0 sign rule (:213-227) and infinite inverse directions (Terra.c:1706-1708)."""
    tri, o, d = cases.kat_rays()
    s = api.Scene(prod)
    s.set_diffuse(s.add_object(tri), (0.5, 0.5, 0.5))
    s.set_options()
    osc = api.OracleScene(oracle, s)
    g = golden("kat_triangle.npz")["hits"]
    assert_hits_equal(osc.trace_rays(o, d), g)
    # the canonical walk must agree too (same arithmetic, different order)
    hits, _, _ = osc.trace_counted(o, d)
    assert_hits_equal(hits, g)
    assert (g["t"] >= 0).sum() >= 20 and (g["t"] < 0).sum() >= 4
    osc.destroy()
    s.destroy()
