"""The C-ABI boundary and the host layer, without a GPU: every symbol include/terra_b200.h declares
is exported, struct layouts match the reference's, the scene container / framebuffer / options
behave like src/Terra.c:130-345, the host BVH equals the reference's byte for byte, and the
flattened GPU arrays round-trip to it. No compute entry point is called here."""
import ctypes as C
import hashlib
import re
import subprocess
import sys

import numpy as np
import pytest

import cases
from conftest import REPO, golden

HEADER = REPO / "include" / "terra_b200.h"


def declared_functions():
    text = HEADER.read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(terra_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(prod):
    names = declared_functions()
    assert len(names) >= 45, names
    missing = [n for n in names if not hasattr(prod.lib, n)]
    assert not missing, missing


def test_header_compiles_as_c_and_cxx(tmp_path):
    """The _Static_asserts in the header pin the x86-64 layouts measured on the reference."""
    src = tmp_path / "t.c"
    src.write_text('#include "terra_b200.h"\nint main(void){return 0;}\n')
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", f"-I{REPO / 'include'}", "-c", str(src), "-o", str(tmp_path / "t.o")], check=True)
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Werror", "-x", "c++", f"-I{REPO / 'include'}", "-c", str(src), "-o", str(tmp_path / "t2.o")], check=True)


def test_reference_caller_links_against_product(tmp_path, prod):
    """Drop-in: a caller written against the reference's call sequence (satellite/src/Scene.cpp:144-237)
    compiles against include/terra_b200.h and links against libterra_b200.so with no other library."""
    src = tmp_path / "caller.c"
    src.write_text(r'''
#include "terra_b200.h"
#include <stdio.h>
int main(void) {
    HTerraScene s = terra_scene_create();
    TerraObject* o = terra_scene_add_object(s, 2);
    TerraFloat3 a = {0.5f, 0.5f, 0.5f};
    terra_bsdf_diffuse_init(&o->material.bsdf);
    terra_attribute_init_constant(&o->material.attributes[TERRA_DIFFUSE_ALBEDO], &a);
    o->material.attributes_count = TERRA_DIFFUSE_END;
    TerraSceneOptions* opt = terra_scene_get_options(s);
    opt->samples_per_pixel = 4; opt->bounces = 2;
    TerraFramebuffer fb;
    if (!terra_framebuffer_create(&fb, 8, 8)) return 2;
    if (terra_framebuffer_create(&fb, 0, 8)) return 3;      /* Terra.c:310-312 */
    printf("%zu %d\n", terra_scene_count_objects(s), terra_b200_device_count() >= 0);
    terra_framebuffer_destroy(&fb);
    terra_scene_destroy(s);
    return 0;
}''')
    exe = tmp_path / "caller"
    libdir = REPO / "terra_b200"
    subprocess.run(["gcc", "-std=c11", f"-I{REPO / 'include'}", str(src), "-o", str(exe), f"-L{libdir}", "-lterra_b200",
                    f"-Wl,-rpath,{libdir}"], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    assert out == ["1", "1"]


def test_scene_container_and_options_double_buffer(api, prod):
    s = api.Scene(prod)
    assert prod.terra_scene_count_objects(s.h) == 0
    objs = [s.add_object(np.zeros((3, 3, 3), np.float32)) for _ in range(70)]     # past the 64 preallocated
    assert prod.terra_scene_count_objects(s.h) == 70
    o = s.options()
    o.samples_per_pixel = 7
    assert s.options().samples_per_pixel == 7       # get_options hands out the pending copy (Terra.c:251-255)
    prod.terra_scene_clear(s.h)
    assert prod.terra_scene_count_objects(s.h) == 0
    s.destroy()


def test_framebuffer_create_clear(api, prod):
    fb = api.Framebuffer()
    assert prod.terra_framebuffer_create(C.byref(fb), 5, 3)
    res = np.ctypeslib.as_array(C.cast(fb.results, C.POINTER(C.c_float)), (15 * 4,))
    pix = np.ctypeslib.as_array(C.cast(fb.pixels, C.POINTER(C.c_float)), (15 * 3,))
    assert not res.any() and not pix.any()
    res[:] = 1
    prod.terra_framebuffer_clear(C.byref(fb))
    assert not res.any()
    prod.terra_framebuffer_destroy(C.byref(fb))
    assert not prod.terra_framebuffer_create(C.byref(fb), 0, 3)


def test_no_cpu_fallback(tmp_path):
    """Without a CUDA device terra_scene_commit must fail loudly (log + abort), never fall back."""
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from terra_b200 import api, scenes\n"
            "lib = api.TerraLib()\n"
            "if lib.terra_b200_device_count() > 0: print('HAS_GPU'); sys.exit(0)\n"
            "scenes.cornell(lib)\nprint('SURVIVED')\n") % str(REPO)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    if "HAS_GPU" in r.stdout:
        pytest.skip("a CUDA device is visible")
    assert r.returncode != 0 and "SURVIVED" not in r.stdout
    assert "no CPU fallback" in r.stdout + r.stderr


# ----------------------------------------------------------------------------- host BVH + flatten
@pytest.fixture(scope="module", params=list(cases.PRIMARY))
def host_scene(request, api, scenes, prod):
    s, cam, wh = cases.build_host(scenes, api, prod, request.param)
    n_tris = sum(o.triangles_count for o in (api.Object * s.n_objects).from_address(s.objects_base))
    yield request.param, s, n_tris
    s.destroy()


def test_host_bvh_reference_mode_is_byte_identical(api, host_scene):
    """tb_bvh_build restates terra_bvh_create incl. the glibc-qsort-dependent order."""
    name, s, n_tris = host_scene
    g = golden("geometry.npz")
    flat_nodes, flat_tris = s.build_host_bvh(n_tris, api.BVH_REFERENCE)
    nodes = s.bvh_nodes()
    assert nodes.size // 64 == n_tris - 1 == int(g[f"{name}_nodes_count"])
    assert hashlib.sha256(nodes.tobytes()).digest() == g[f"{name}_nodes_sha256"].tobytes()


def _check_flat(api, s, n_tris, mode):
    flat_nodes, flat_tris = s.build_host_bvh(n_tris, mode)
    host = s.bvh_nodes().view(np.float32).reshape(-1, 16)
    hosti = host.view(np.int32)
    # boxes: same 12 floats, same order
    assert np.array_equal(flat_nodes[:, :12].view(np.uint32), host[:, :12].view(np.uint32))
    ref = flat_nodes[:, 12:14].view(np.int32)
    inner = hosti[:, 14:16] == -1
    assert np.array_equal(ref[inner], hosti[:, 12:14][inner])
    # leaves: ~ref is a rank; the rank's triangle record carries the original object | tri << 8
    ranks = ~ref[~inner]
    assert np.array_equal(np.sort(ranks), np.arange(n_tris))             # every triangle exactly once
    prim = flat_tris[:, 3].view(np.uint32)
    assert np.array_equal(prim[ranks], hosti[:, 12:14][~inner].view(np.uint32))
    objs = (api.Object * s.n_objects).from_address(s.objects_base)
    for r in np.random.default_rng(0).integers(0, n_tris, 64):
        o, t = int(prim[r]) & 0xff, int(prim[r]) >> 8
        tri = np.ctypeslib.as_array(C.cast(objs[o].triangles, C.POINTER(C.c_float)), (objs[o].triangles_count, 9))[t]
        assert np.array_equal(flat_tris[r, [0, 1, 2, 4, 5, 6, 8, 9, 10]].view(np.uint32), tri.view(np.uint32))
    return flat_nodes, flat_tris, host, hosti


def test_flatten_round_trips_reference_tree(api, host_scene):
    name, s, n_tris = host_scene
    flat_nodes, flat_tris, host, hosti = _check_flat(api, s, n_tris, api.BVH_REFERENCE)
    # rank order == the reference traversal's leaf visiting order (TerraBVH.c:266-306): simulate it
    order, stack = [], [0]
    while stack:
        n = stack.pop()
        for c in range(2):
            if hosti[n, 14 + c] == -1:
                stack.append(int(hosti[n, 12 + c]))
            else:
                order.append(int(hosti[n, 12 + c]))
    assert np.array_equal(flat_tris[:, 3].view(np.uint32), np.asarray(order, np.uint32))
    assert s.lib.terra_b200_bvh_stack_bound(s.h) < 128


def test_sah3_tree_is_a_valid_bvh(api, host_scene):
    """The default GPU tree (binned SAH over x,y,z, same 64-byte node format): N-1 nodes, every
    triangle in exactly one leaf, every child box encloses its subtree's triangles."""
    name, s, n_tris = host_scene
    flat_nodes, flat_tris, host, hosti = _check_flat(api, s, n_tris, api.BVH_SAH3)
    assert host.shape[0] == n_tris - 1
    lo = np.minimum(np.minimum(flat_tris[:, 0:3], flat_tris[:, 4:7]), flat_tris[:, 8:11])
    hi = np.maximum(np.maximum(flat_tris[:, 0:3], flat_tris[:, 4:7]), flat_tris[:, 8:11])
    ref = flat_nodes[:, 12:14].view(np.int32)
    sub_lo = np.zeros((host.shape[0], 2, 3), np.float32)
    sub_hi = np.zeros((host.shape[0], 2, 3), np.float32)
    for n in range(host.shape[0] - 1, -1, -1):      # children have larger indices than their parent
        for c in range(2):
            r = int(ref[n, c])
            if r < 0:
                sub_lo[n, c], sub_hi[n, c] = lo[~r], hi[~r]
            else:
                assert r > n
                sub_lo[n, c] = np.minimum(sub_lo[r, 0], sub_lo[r, 1])
                sub_hi[n, c] = np.maximum(sub_hi[r, 0], sub_hi[r, 1])
            box = host[n, 6 * c:6 * c + 6]
            assert (box[0:3] <= sub_lo[n, c]).all() and (box[3:6] >= sub_hi[n, c]).all()


def test_sah3_parallel_build_matches_serial(api, prod, monkeypatch):
    """The host sah3 build runs on all cores (preorder node ids, parallel binning + two-pointer partition on
    the top ranges, independent subtree tasks below, renumbering at the end); the node array must not
    depend on the thread count or on where the serial/parallel switch sits. The scene has clusters of
    identical triangles (coincident centroids -> median fallback, order-sensitive)."""
    rng = np.random.default_rng(11)
    n = 6000
    c = rng.uniform(-50, 50, (n, 1, 3)).astype(np.float32)
    tris = (c + rng.uniform(-1, 1, (n, 3, 3)).astype(np.float32)).astype(np.float32)
    tris[1000:1400] = tris[1000]                    # 400 coincident triangles
    tris[3000:3003] = tris[3000]
    s = api.Scene(prod)
    s.add_object(tris[:2500])
    s.add_object(tris[2500:])
    hashes = set()
    for threads, par_min in (("1", "65536"), ("2", "64"), ("5", "16"), ("8", "1000"), ("7", "2")):
        monkeypatch.setenv("TERRA_B200_BUILD_THREADS", threads)
        monkeypatch.setenv("TERRA_B200_BUILD_PAR_MIN", par_min)
        assert prod.lib.terra_b200_build_host_bvh(s.h, api.BVH_SAH3, None, None) == n - 1
        hashes.add(hashlib.sha256(s.bvh_nodes().tobytes()).hexdigest())
    assert len(hashes) == 1
    s.destroy()


def test_sah3_and_reference_trees_give_identical_hits(api, scenes, prod, oracle, host_scene):
    """Closest hits do not depend on the tree: canonical walk over the sah3 node array vs the
    reference walk over the reference tree (ids may differ only on exact-t ties)."""
    name, s, n_tris = host_scene
    s.build_host_bvh(n_tris, api.BVH_SAH3)
    sah_nodes = s.bvh_nodes()
    osc = api.OracleScene(oracle, s)
    o, d = cases.rays_for(s, api, scenes, 4096)
    a = osc.trace_rays(o, d)
    b, n_nodes, n_tris_tested = osc.trace_counted(o, d, sah_nodes)
    assert np.array_equal(a["t"].view(np.uint32), b["t"].view(np.uint32))
    assert (a["prim"] != b["prim"]).mean() < 0.01
    osc.destroy()


def test_quad_tree_is_the_binary_tree_regrouped(api, host_scene):
    """tb_bvh4_build (the 4-wide tree of k_trace_v7): every slot's box is the quantised box of one binary child, copied
    word for word; every leaf rank appears exactly once; every binary inner node is either a wide node or was absorbed
    into one; the stack bound covers the worst near-first walk."""
    name, s, n_tris = host_scene
    flat_nodes, _ = s.build_host_bvh(n_tris, api.BVH_SAH3)
    n_bin = n_tris - 1
    q4 = np.zeros((n_bin + 1, 16), np.uint32)
    src = np.full((n_bin + 1, 4), -2, np.int32)
    qn = np.zeros((n_bin + 1, 8), np.uint32)
    bound = C.c_int(0)
    n = s.lib.terra_b200_build_host_bvh4(s.h, api.BVH_SAH3, q4.ctypes.data, src.ctypes.data, qn.ctypes.data, n_bin + 1, C.byref(bound))
    assert 0 < n <= n_bin
    q4, src, qn = q4[:n], src[:n], qn[:n_bin]
    refs = q4[:, 12:16].view(np.int32)
    used = src >= 0
    assert used[:, :2].all()                                            # at least the two children of the binary node
    # boxes: copied from the binary child they name; empty slots are inverted boxes
    b, c = src[used] >> 1, src[used] & 1
    boxes = q4[:, :12].reshape(n, 4, 3)
    want = qn[b][np.arange(b.size)[:, None], (3 * c)[:, None] + np.arange(3)[None, :]]
    assert np.array_equal(boxes[used], want)
    assert (boxes[~used] == 0x0000ffff).all()
    # refs: leaves keep the binary ref (~rank), inner slots point at a LATER wide node (breadth-first numbering)
    bin_ref = qn[:, 6:8].view(np.int32)[b, c]
    got = refs[used]
    leaf = bin_ref < 0
    assert np.array_equal(got[leaf], bin_ref[leaf])
    assert np.array_equal(np.sort(~got[leaf]), np.arange(n_tris))       # every triangle exactly once
    rows = np.nonzero(used)[0]
    assert (got[~leaf] > rows[~leaf]).all() and (got[~leaf] < n).all()
    assert np.array_equal(np.sort(got[~leaf]), np.arange(1, n))         # every wide node but the root has exactly one parent
    # stack bound: sum of (children - 1) along the worst path, + 1
    depth = np.zeros(n, np.int64)
    for i in range(n - 1, -1, -1):
        kids = refs[i][used[i]]
        inner = kids[kids >= 0]
        depth[i] = used[i].sum() - 1 + (depth[inner].max() if inner.size else 0)
    assert bound.value == depth[0] + 1
    assert n <= (n_bin + 1) // 2 + 1 or n_tris < 8                      # a collapse: at most ~half the binary nodes remain
