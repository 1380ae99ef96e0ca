"""GPU BVH build (terra_b200/csrc/tb_lbvh.cu) against the reference builder's CONTRACT:
the node array must be a valid tree in the reference's 64-byte node format over exactly the scene's
triangles (reference src/TerraBVH.c:128-244, src/TerraBVH.h:13-17), and closest hits through it must
be the reference's (tests/test_gpu_parity.py runs every parity case on this tree as well)."""
import ctypes as C
import time

import numpy as np
import pytest

import cases
from conftest import golden

pytestmark = pytest.mark.gpu

NODE = np.dtype([("box", "<f4", (2, 2, 3)), ("index", "<i4", (2,)), ("type", "<i4", (2,))])


@pytest.fixture()
def lbvh(gpu):
    gpu.terra_b200_set_bvh_mode(2)
    yield gpu
    gpu.terra_b200_set_bvh_mode(1)


def check_tree(nodes, tri_of_prim):
    """Every primitive in exactly one leaf; every child box contains all triangles below it. Returns depth."""
    n = len(nodes)
    seen = {}
    depth = 0
    # iterative post-order: box of the triangles below (node, child)
    lo = np.full((n, 2, 3), np.inf); hi = np.full((n, 2, 3), -np.inf)
    order, stack = [], [(0, 1)]
    visited = np.zeros(n, bool)
    while stack:
        i, d = stack.pop()
        assert not visited[i], "a node is reachable twice"
        visited[i] = True
        depth = max(depth, d)
        order.append(i)
        for c in range(2):
            if nodes["type"][i][c] == -1:
                stack.append((int(nodes["index"][i][c]), d + 1))
    assert visited.all(), "unreachable nodes"
    for i in reversed(order):
        for c in range(2):
            if nodes["type"][i][c] == -1:
                j = int(nodes["index"][i][c])
                lo[i, c] = np.minimum(lo[j, 0], lo[j, 1]); hi[i, c] = np.maximum(hi[j, 0], hi[j, 1])
            else:
                prim = int(nodes["index"][i][c]) & 0xffffffff
                assert prim not in seen, "a triangle sits in two leaves"
                seen[prim] = True
                t = tri_of_prim(prim)
                lo[i, c] = t.min(axis=0); hi[i, c] = t.max(axis=0)
    box = nodes["box"].astype(np.float64)
    assert (box[:, :, 0, :] <= lo).all() and (box[:, :, 1, :] >= hi).all(), "a child box does not contain its triangles"
    return depth, seen


@pytest.mark.parametrize("name", list(cases.PRIMARY))
def test_lbvh_is_a_valid_reference_format_tree(api, scenes, lbvh, name):
    s, cam, _ = cases.build(scenes, lbvh, name)
    nodes = s.bvh_nodes().view(NODE)
    raw = (api.Object * s.n_objects).from_address(s.objects_base)
    objs = [np.ctypeslib.as_array(C.cast(o.triangles, C.POINTER(C.c_float)), (o.triangles_count, 3, 3)).copy() for o in raw]
    total = sum(len(o) for o in objs)
    assert len(nodes) == total - 1                   # N-1 nodes, one triangle per leaf (TerraBVH.c:134)
    depth, seen = check_tree(nodes, lambda prim: objs[prim & 0xff][prim >> 8].astype(np.float64))
    assert len(seen) == total
    assert depth == lbvh.terra_b200_bvh_stack_bound(s.h)
    s.destroy()


def test_lbvh_hits_equal_sah3_hits_on_a_medium_terrain(api, scenes, gpu):
    """200k-triangle terrain, 1M incoherent rays: t,u,v bit-identical between the GPU-built tree and the
    host SAH tree (the hit does not depend on the tree); ids may differ on exact-t ties only."""
    n_rays = 1_000_000
    res = {}
    for mode in (1, 2):
        gpu.terra_b200_set_bvh_mode(mode)
        try:
            t0 = time.time()
            s, cam = scenes.terrain(gpu, n_tris=200_000, extent=300.0, lights=False)
            commit_s = time.time() - t0
        finally:
            gpu.terra_b200_set_bvh_mode(1)
        lo = np.array([0, -30, 0], np.float32); hi = np.array([300, 30, 300], np.float32)
        o, d = scenes.incoherent_rays(lo, hi, n_rays, seed=11)
        res[mode] = (s.trace_rays(o, d), commit_s)
        if mode == 2:
            dev_ms, up_ms = C.c_float(0), C.c_float(0)
            gpu.terra_b200_lbvh_build_ms(C.byref(dev_ms), C.byref(up_ms))
            print(f"lbvh 200k tris: device {dev_ms.value:.2f} ms, with upload {up_ms.value:.2f} ms, commit {commit_s:.2f} s (sah3 commit {res[1][1]:.2f} s)")
            assert 0 < dev_ms.value < 200
        s.destroy()
    a, b = res[1][0], res[2][0]
    for f in ("t", "u", "v"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f
    assert (a["prim"] != b["prim"]).mean() < 0.01
    assert (a["t"] >= 0).mean() > 0.2


# ----------------------------------------------------------------------------- refit (the other half of row N1)
@pytest.mark.parametrize("kernel", ["auto", "v7"], ids=["default_kernels", "quad_tree_kernel"])
@pytest.mark.parametrize("mode", [1, 2], ids=["sah3_tree", "gpu_lbvh_tree"])
def test_refit_gives_the_hits_of_a_fresh_commit(api, scenes, gpu, oracle, mode, kernel, monkeypatch):
    """terra_b200_refit after editing vertices in place: the tree keeps its topology, every box is re-fitted on
    the GPU and re-quantised. Closest hits must be those of a fresh commit of the edited geometry (t bit-identical,
    ids modulo exact-t ties, u / v bit-identical where the ids agree) and of the oracle's reference walk; a refit
    with nothing moved changes nothing; renders (light areas refreshed) agree with the fresh scene's."""
    import ctypes as C
    import cases
    gpu.terra_b200_set_bvh_mode(mode)
    monkeypatch.setenv("TERRA_B200_INCOHERENT", kernel)      # v7: the 4-wide tree, whose boxes the refit re-copies (k_refit_quad)
    try:
        s, cam, wh = cases.build(scenes, gpu, "field")
        o, d = cases.rays_for(s, api, scenes, 30_000)
        before = s.trace_rays(o, d)
        assert gpu.terra_b200_refit(s.h) >= 0.0
        assert np.array_equal(s.trace_rays(o, d).view(np.uint32), before.view(np.uint32))
        objs = (api.Object * s.n_objects).from_address(s.objects_base)
        copies = []
        for k in range(s.n_objects):
            n = objs[k].triangles_count
            v = np.ctypeslib.as_array(C.cast(objs[k].triangles, C.POINTER(C.c_float)), (n, 3, 3))
            v += (0.8 * np.sin(0.9 * v[..., [2, 0, 1]] + 0.5)).astype(np.float32)       # a smooth deformation, in place
            pr = np.ctypeslib.as_array(C.cast(objs[k].properties, C.POINTER(C.c_float)), (n, 15))
            copies.append((v.copy(), pr[:, :9].reshape(n, 3, 3).copy(), pr[:, 9:].reshape(n, 3, 2).copy()))
        ms = gpu.terra_b200_refit(s.h)
        got = s.trace_rays(o, d)
        fresh = api.Scene(gpu)
        for k, (v, nrm, tex) in enumerate(copies):
            obj = fresh.add_object(v, nrm, tex)
            C.memmove(C.byref(obj.material), C.byref(objs[k].material), C.sizeof(api.Material))
        C.memmove(C.byref(fresh.options()), C.byref(s.options()), C.sizeof(api.SceneOptions))
        fresh.commit()
        want = fresh.trace_rays(o, d)
        print(f"refit of {sum(c[0].shape[0] for c in copies)} triangles: {ms:.3f} ms on the device")
        assert (bits(got["t"]) != bits(before["t"])).mean() > 0.05           # the geometry really moved
        assert np.array_equal(bits(got["t"]), bits(want["t"]))
        assert (got["prim"] != want["prim"]).mean() < 0.01
        same = got["prim"] == want["prim"]
        assert np.array_equal(bits(got["u"][same]), bits(want["u"][same])) and np.array_equal(bits(got["v"][same]), bits(want["v"][same]))
        osc = api.OracleScene(oracle, fresh)
        ref_walk = osc.trace_rays(o[:3000], d[:3000])
        osc.destroy()
        assert np.array_equal(bits(got["t"][:3000]), bits(ref_walk["t"]))
        # shading after a refit: same seed, same camera rays; the two trees may order exact-t ties differently, nothing else
        gpu.terra_b200_set_seed(s.h, 3); gpu.terra_b200_set_seed(fresh.h, 3)
        a = s.render(cam, 64, 40)[0]["acc"]
        b = fresh.render(cam, 64, 40)[0]["acc"]
        assert np.isfinite(a).all() and (a.reshape(-1, 3) != b.reshape(-1, 3)).any(axis=1).mean() < 0.02
        fresh.destroy()
        s.destroy()
    finally:
        gpu.terra_b200_set_bvh_mode(1)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)
