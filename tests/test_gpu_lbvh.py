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
