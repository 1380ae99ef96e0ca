"""Parity tests proper: the CUDA path, called through the C ABI (ctypes -> libterra_b200.so), against
the oracle / the reference fixtures on the same seeded inputs. All marked gpu.

Bars (BASELINE.json north_star): primary closest-hit ids bit-exact outside numerical ties and t
within 1e-5 relative — here t, u, v are required BIT-identical and ids may differ only where the
other candidate has the identical t (a shared edge / coplanar overlap: with the reference tree the
rank order reproduces the reference's first-found rule, so even those must agree)."""
import ctypes as C
import os

import numpy as np
import pytest

import cases
from conftest import golden

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def assert_hits(gpu_hits, want, exact_ids):
    for f in ("t", "u", "v") if exact_ids else ("t",):
        assert np.array_equal(bits(gpu_hits[f]), bits(want[f])), f
    differ = gpu_hits["prim"] != want["prim"]
    if exact_ids:
        assert not differ.any()
    else:
        assert differ.mean() < 0.01          # exact-t ties only (t is already bit-identical everywhere)


@pytest.fixture(params=[0, 1, 2], ids=["reference_tree", "sah3_tree", "gpu_lbvh_tree"])
def bvh_mode(request, gpu):
    gpu.terra_b200_set_bvh_mode(request.param)
    yield request.param
    gpu.terra_b200_set_bvh_mode(1)


# ----------------------------------------------------------------------------- closest hit
@pytest.mark.parametrize("name", list(cases.PRIMARY))
def test_primary_hits_vs_reference_fixture(api, scenes, gpu, bvh_mode, name):
    g = golden("geometry.npz")
    s, cam, (w, h) = cases.build(scenes, gpu, name)
    hits = s.primary_hits(cam, w, h)
    assert_hits(hits, g[f"{name}_primary"], exact_ids=bvh_mode == 0)
    s.destroy()


@pytest.mark.parametrize("name", list(cases.PRIMARY))
def test_incoherent_hits_vs_reference_fixture(api, scenes, gpu, bvh_mode, name):
    g = golden("geometry.npz")
    s, cam, wh = cases.build(scenes, gpu, name)
    o, d = cases.rays_for(s, api, scenes, 4096)
    assert_hits(s.trace_rays(o, d), g[f"{name}_rays_hits"], exact_ids=bvh_mode == 0)
    s.destroy()


KERNELS = {                # TERRA_B200_INCOHERENT, TERRA_B200_STACK_SMEM (read when a scene's device state is created)
    "rounds_binary": ("v6", None),
    "rounds_binary_lean": ("v8", None),       # 4-byte stack entries, triangle-round state in shared memory (default for scenes that fit L2)
    "rounds_binary_lean_spill": ("v8", "2"),  # two shared stack rows, the rest in local memory
    "rounds_quad": ("v7", None),              # the 4-wide tree (default only for scenes that do not fit L2)
    "rounds_quad_spill": ("v7", "2"),         # two shared stack rows: almost every push goes to the local-memory part
    "per_ray_walk": ("walk", None),
}


@pytest.mark.parametrize("kernel", list(KERNELS))
@pytest.mark.parametrize("name", list(cases.PRIMARY))
def test_every_incoherent_kernel_gives_the_fixture_hits(api, scenes, gpu, monkeypatch, name, kernel):
    """Closest hits do not depend on the kernel or on the tree it walks: binary rounds, 4-wide rounds (boxes copied from
    the binary tree, up to three pushes per step, shared + local stack), per-ray walk — t, u, v bit-identical to the
    reference fixture, ids identical between kernels (equal t goes to the smaller rank whatever the visiting order)."""
    g = golden("geometry.npz")
    inc, rows = KERNELS[kernel]
    monkeypatch.setenv("TERRA_B200_INCOHERENT", inc)
    if rows:
        monkeypatch.setenv("TERRA_B200_STACK_SMEM", rows)
    s, cam, wh = cases.build(scenes, gpu, name)
    o, d = cases.rays_for(s, api, scenes, 4096)
    got = s.trace_rays(o, d)
    s.destroy()
    assert_hits(got, g[f"{name}_rays_hits"], exact_ids=False)
    monkeypatch.setenv("TERRA_B200_INCOHERENT", "v6")
    monkeypatch.delenv("TERRA_B200_STACK_SMEM", raising=False)
    s, cam, wh = cases.build(scenes, gpu, name)
    base = s.trace_rays(o, d)
    s.destroy()
    assert np.array_equal(got.view(np.uint32), base.view(np.uint32))


def test_quad_tree_kernel_renders_the_same_image(api, scenes, gpu, monkeypatch):
    """A whole frame (bounce and visibility rays through k_trace_v7, hits binned per preset) against the binary-tree
    kernel: same hits -> same paths -> the same framebuffer, bit for bit."""
    out = []
    for inc in ("v6", "v7", "v8"):
        monkeypatch.setenv("TERRA_B200_INCOHERENT", inc)
        s, cam = scenes.sphere_field(gpu, n_spheres=40, n_lat=10, n_lon=12, spp=8, bounces=4, ggx=True, integrator=api.INTEGRATOR_DIRECT)
        gpu.terra_b200_set_seed(s.h, 5)
        res, _ = s.render(cam, 160, 90)
        out.append(res["acc"].copy())
        s.destroy()
    assert out[0].sum() > 0
    assert np.array_equal(out[0].view(np.uint32), out[1].view(np.uint32))
    assert np.array_equal(out[0].view(np.uint32), out[2].view(np.uint32))


def test_triangle_known_answers(api, gpu, bvh_mode):
    """fp64 retry, +0 sign rule, infinite inverse directions, back faces, co-planar rays."""
    tri, o, d = cases.kat_rays()
    s = api.Scene(gpu)
    s.set_diffuse(s.add_object(tri), (0.5, 0.5, 0.5))
    s.set_options()
    s.commit()
    assert_hits(s.trace_rays(o, d), golden("kat_triangle.npz")["hits"], exact_ids=True)
    s.destroy()


def test_empty_and_ragged_ray_batches(api, scenes, gpu):
    s, cam, wh = cases.build(scenes, gpu, "field")
    o, d = cases.rays_for(s, api, scenes, 1000)
    full = s.trace_rays(o, d)
    assert s.trace_rays(o[:0], d[:0]).shape == (0,)
    for n in (1, 31, 33, 127, 999):          # not multiples of the warp / CTA size
        part = s.trace_rays(o[:n], d[:n])
        assert np.array_equal(part.view(np.uint32), full[:n].view(np.uint32))
    s.destroy()


def test_pipelined_host_batches_equal_one_shot(api, scenes, gpu, monkeypatch):
    """terra_b200_trace_rays moves a host batch through copy-in | trace | copy-out streams in chunks, two
    staging slots: any chunking, page-locked or pageable buffers, gives the bytes of the one-chunk call
    (ragged last chunk, more chunks than slots, a batch smaller than a chunk)."""
    s, cam, wh = cases.build(scenes, gpu, "field")
    n = 50_001
    o, d = cases.rays_for(s, api, scenes, n)
    monkeypatch.setenv("TERRA_B200_IO_CHUNK", str(1 << 20))
    want = s.trace_rays(o, d)
    po, pd = api.PinnedArray(gpu, (n, 3), np.float32), api.PinnedArray(gpu, (n, 3), np.float32)
    ph = api.PinnedArray(gpu, n, api.HIT_DTYPE)
    po.array[:] = o
    pd.array[:] = d
    for chunk in (1024, 4096, 16384, 50_001, 50_002):
        monkeypatch.setenv("TERRA_B200_IO_CHUNK", str(chunk))
        ph.array.view(np.uint8)[:] = 0xff
        got = s.trace_rays(po.array, pd.array, out=ph.array)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), chunk
        # pageable caller buffers: straight from them (driver-staged copies) ...
        monkeypatch.setenv("TERRA_B200_STAGING_MIN", "0")
        assert np.array_equal(s.trace_rays(o, d).view(np.uint32), want.view(np.uint32)), chunk
        # ... and through the library's page-locked bounce buffers (the path large pageable batches take)
        monkeypatch.setenv("TERRA_B200_STAGING_MIN", "1")
        assert np.array_equal(s.trace_rays(o, d).view(np.uint32), want.view(np.uint32)), chunk
        mixed = s.trace_rays(po.array, d, out=ph.array)                  # page-locked and pageable mixed
        assert np.array_equal(mixed.view(np.uint32), want.view(np.uint32)), chunk
        monkeypatch.delenv("TERRA_B200_STAGING_MIN")
        assert np.array_equal(s.trace_rays(po.array[:777], pd.array[:777]).view(np.uint32), want[:777].view(np.uint32))
    st = s.stats()
    assert st.h2d_bytes > 0 and st.d2h_bytes > 0
    for a in (po, pd, ph):
        a.free()
    s.destroy()


def test_two_hundred_objects_and_the_object_limit(api, scenes, gpu, oracle, bvh_mode):
    """The leaf packing gives an object 8 bits (TerraBVH.c:155): 200 objects (beyond the reference's 64
    preallocated slots, Q14) trace bit-exactly against the oracle's walk of the reference tree, object ids
    above 127 included."""
    rng = np.random.default_rng(17)
    s = api.Scene(gpu)
    for k in range(200):
        c = rng.uniform(-20, 20, 3).astype(np.float32)
        tris = (c + rng.uniform(-1.5, 1.5, (3 + k % 4, 3, 3))).astype(np.float32)
        s.set_diffuse(s.add_object(tris), (0.5, 0.5, 0.5))
    s.set_options()
    s.commit()
    assert s.n_objects == 200
    o, d = cases.rays_for(s, api, scenes, 20_000)
    got = s.trace_rays(o, d)
    osc = api.OracleScene(oracle, s)
    want = osc.trace_rays(o, d)
    osc.destroy()
    assert (want["t"] >= 0).mean() > 0.02
    assert_hits(got, want, exact_ids=(bvh_mode == 0))
    hit = got["t"] >= 0
    assert (got["prim"][hit] & 0xff).max() > 127
    s.destroy()


def test_medium_scene_vs_oracle(api, scenes, gpu, oracle):
    """100k-triangle sphere field (config C2's geometry) and a 200k-triangle terrain: 200k incoherent
    rays against the oracle's canonical walk over the same tree, plus primary hits at 320x180."""
    for fn, kw in (("sphere_field", dict(ggx=False)), ("terrain", dict(n_tris=200_000))):
        s, cam = getattr(scenes, fn)(gpu, **kw)
        osc = api.OracleScene(oracle, s)
        nodes = s.bvh_nodes()
        o, d = cases.rays_for(s, api, scenes, 200_000)
        want, n_nodes, n_tris = osc.trace_counted(o, d, nodes)
        got = s.trace_rays(o, d)
        assert np.array_equal(bits(got["t"]), bits(want["t"]))
        assert (got["prim"] != want["prim"]).mean() < 1e-3
        same = got["prim"] == want["prim"]
        assert np.array_equal(bits(got["u"][same]), bits(want["u"][same]))
        # a sample of those rays through the reference's own un-culled walk of ITS tree
        ref_walk = osc.trace_rays(o[:300], d[:300])
        assert np.array_equal(bits(got["t"][:300]), bits(ref_walk["t"]))
        # primary hits at 320x180 (camera rays: the packet kernel) against the oracle's walk of the reference tree
        w, h = 320, 180
        prim = s.primary_hits(cam, w, h)
        want_prim, _ = osc.primary_hits(cam, w, h)
        assert (want_prim["t"] >= 0).mean() > 0.3
        assert np.array_equal(bits(prim["t"]), bits(want_prim["t"]))
        assert (prim["prim"] != want_prim["prim"]).mean() < 1e-3
        same = prim["prim"] == want_prim["prim"]
        assert np.array_equal(bits(prim["u"][same]), bits(want_prim["u"][same]))
        assert np.array_equal(bits(prim["v"][same]), bits(want_prim["v"][same]))
        osc.destroy()
        s.destroy()


def test_counting_variant_matches_canonical_counter(api, scenes, gpu, oracle):
    """The work counters behind roofline.achieved: nodes fetched / triangles tested by the counting
    variant of the trace kernel stay within a few percent of the oracle's canonical near-first culled
    walk of the same tree. (The kernel for incoherent rays tests a leaf one step late and walks
    slightly larger quantised boxes, so it fetches ~2-3 % more nodes and tests ~10 % more triangles than
    the canonical walk; bench.py's roofline uses the kernel's own counts for the float-node walk.)"""
    s, cam = scenes.terrain(gpu, n_tris=200_000)
    osc = api.OracleScene(oracle, s)
    o, d = cases.rays_for(s, api, scenes, 100_000)
    _, n_nodes, n_tris = osc.trace_counted(o, d, s.bvh_nodes())
    gpu.terra_b200_set_counting(s.h, 1)
    gpu.terra_b200_reset_stats(s.h)
    s.trace_rays(o, d)
    st = s.stats()
    gpu.terra_b200_set_counting(s.h, 0)
    assert st.rays_traced == len(o)
    assert abs(st.nodes_visited - n_nodes) <= 0.05 * n_nodes
    assert abs(st.tris_tested - n_tris) <= 0.15 * n_tris
    # counting == 2: the default incoherent kernel's OWN fetches and lane states (terra_b200_round_stats). k_trace_v8
    # keeps no distance in its stack entries, so it also fetches the nodes whose entry the closest hit had culled.
    assert gpu.terra_b200_incoherent_kernel(s.h) == b"k_trace_v8"
    gpu.terra_b200_set_counting(s.h, 2)
    gpu.terra_b200_reset_stats(s.h)
    s.trace_rays(o, d)
    st2 = s.stats()
    rs = (C.c_uint64 * 10)()
    gpu.terra_b200_round_stats(s.h, rs)
    gpu.terra_b200_set_counting(s.h, 0)
    assert 0.98 * n_nodes <= st2.nodes_visited <= 1.15 * n_nodes
    assert rs[1] == st2.nodes_visited and rs[7] == st2.tris_tested and rs[9] == len(o)
    osc.destroy()
    s.destroy()
