"""State the library keeps between calls: shared device buffers, the host tree after a refit, materials edited in
place between commits. Each case is compared bit for bit with a FRESH scene that never went through the sequence."""
import ctypes as C

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _render_device(api, gpu, s, cam, w, h, spp):
    import torch
    buf = torch.zeros(w * h * 4, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    gpu.terra_b200_render_device(C.byref(cam), s.h, w, h, 0, 0, w, h, 0, spp, C.c_void_p(buf.data_ptr()), None)
    torch.cuda.synchronize()
    return buf.cpu().numpy().view(api.RESULT_DTYPE).reshape(h, w)


def test_render_entry_points_share_one_accumulator(api, scenes, gpu):
    """terra_b200_render_device (large), terra_render (small tile), terra_b200_render_device (large) on one scene:
    both entry points use the same device accumulator, whose capacity used to be tracked twice — the tile call shrank
    it and the second large call then wrote 33 MB into a 256 KB allocation."""
    w, h, spp = 640, 360, 2
    s, cam = scenes.cornell(gpu, spp=spp)
    gpu.terra_b200_set_seed(s.h, 31)
    first = _render_device(api, gpu, s, cam, w, h, spp)
    tile, _ = s.render(cam, w, h, region=(64, 32, 128, 128))
    gpu.terra_b200_set_seed(s.h, 31)
    again = _render_device(api, gpu, s, cam, w, h, spp)
    s.destroy()
    fresh_s, fresh_cam = scenes.cornell(gpu, spp=spp)
    gpu.terra_b200_set_seed(fresh_s.h, 31)
    want = _render_device(api, gpu, fresh_s, fresh_cam, w, h, spp)
    fresh_s.destroy()
    assert (tile["samples"][32:160, 64:192] == spp).all()
    assert np.array_equal(first.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(again.view(np.uint32), want.view(np.uint32))
    # and a scene that is still usable afterwards (a corrupted wave / scene buffer would show here)
    s2, cam2 = scenes.cornell(gpu, spp=spp)
    hits = s2.primary_hits(cam2, 64, 64)
    assert (hits["t"] > 0).mean() > 0.7
    s2.destroy()


def _moved(tris, k):
    out = tris.copy()
    out[..., 1] += np.float32(0.35 * k) * np.sin(out[..., 0] * 0.7 + k)
    return out.astype(np.float32)


def test_refit_then_a_commit_that_re_flattens_keeps_the_refitted_boxes(api, scenes, gpu, oracle):
    """terra_b200_refit moves the boxes on the device only. A later commit that uploads the geometry again without
    rebuilding (here: a BSDF preset swapped in place, which re-tags the triangle records) used to flatten the STALE
    host boxes over the moved triangles: closest hits were silently missed. Also terra_b200_bvh_nodes after a refit."""
    s, cam = scenes.sphere_field(gpu, n_spheres=8, n_lat=8, n_lon=10, extent=20.0, seed=1, ggx=False, n_groups=4)
    objs = (api.Object * s.n_objects).from_address(s.objects_base)
    before = s.bvh_nodes().copy()
    # edit every object's vertices in place (same topology), refit
    for k, o in enumerate(objs):
        a = np.ctypeslib.as_array(C.cast(o.triangles, C.POINTER(C.c_float)), (o.triangles_count, 3, 3))
        a[:] = _moved(a, k + 1)
    gpu.terra_b200_refit(s.h)
    after = s.bvh_nodes()
    assert not np.array_equal(before, after), "terra_b200_bvh_nodes returned the pre-refit boxes"
    o, d = cases.rays_for(s, api, scenes, 20_000)
    refit_hits = s.trace_rays(o, d)
    # swap one object's preset in place and commit: geometry path without a rebuild
    s.set_phong(objs[0], (0.4, 0.4, 0.4), (0.3, 0.3, 0.3), 20.0)
    s.commit()
    got = s.trace_rays(o, d)
    # the oracle walks ITS OWN tree over the edited triangles: the reference answer for the moved geometry
    osc = api.OracleScene(oracle, s)
    want = osc.trace_rays(o, d)
    osc.destroy()
    s.destroy()
    assert (want["t"] >= 0).mean() > 0.05
    assert np.array_equal(got["t"].view(np.uint32), want["t"].view(np.uint32))
    assert np.array_equal(refit_hits["t"].view(np.uint32), want["t"].view(np.uint32))
    assert (got["prim"] != want["prim"]).mean() < 0.01


def test_material_edited_in_place_shows_after_commit(api, scenes, gpu):
    """The reference reads object->material live at every hit (Terra.c:1745-1762); here materials are a device table,
    refreshed at EVERY commit: a constant albedo / emissive edited in place, then terra_scene_commit, renders exactly
    what a fresh scene with that material renders — without a BVH rebuild."""
    w, h, spp = 48, 40, 8

    def image(s, cam, seed=5):
        gpu.terra_b200_set_seed(s.h, seed)
        return s.render(cam, w, h)[0]["acc"].copy()

    s, cam = scenes.cornell(gpu, spp=spp, bounces=3)
    base = image(s, cam)
    objs = (api.Object * s.n_objects).from_address(s.objects_base)
    s.const_attr(objs[1].material.attributes[0], (0.1, 0.2, 0.9))          # left wall: red -> blue
    s.const_attr(objs[s.n_objects - 1].material.emissive, (30, 20, 10))    # the light's radiance
    out4 = (C.c_float * 4)()
    s.commit()
    gpu.terra_b200_commit_ms(s.h, out4)
    edited = image(s, cam)
    s.destroy()

    orig = api.Scene.set_diffuse
    calls = {"n": 0}

    def set_diffuse(self, obj, albedo, emissive=(0, 0, 0)):
        calls["n"] += 1
        if calls["n"] == 2:
            albedo = (0.1, 0.2, 0.9)
        if tuple(emissive) != (0, 0, 0):
            emissive = (30, 20, 10)
        return orig(self, obj, albedo, emissive)

    api.Scene.set_diffuse = set_diffuse
    try:
        f, fcam = scenes.cornell(gpu, spp=spp, bounces=3)
    finally:
        api.Scene.set_diffuse = orig
    want = image(f, fcam)
    f.destroy()
    assert not np.array_equal(base, edited)
    assert np.array_equal(edited.view(np.uint32), want.view(np.uint32))


def test_multi_device_terra_render_equals_the_sum_of_its_sample_ranges(api, scenes, gpu, monkeypatch):
    """Multi-GPU behind terra_render (terra_b200_scene_set_devices / TERRA_B200_DEVICES): the scene is replicated, GPU g
    renders sample range g of the whole region, finalises row band g from ALL accumulators through peer-mapped memory
    and copies its band back. Bit for bit what the ranges give when rendered one by one on a single GPU and summed
    in device order; pixels[] is the tonemap of that sum. Needs two GPUs (skipped otherwise)."""
    n_dev = min(gpu.terra_b200_device_count(), 4)
    if n_dev < 2:
        pytest.skip("one visible GPU")
    monkeypatch.setenv("TERRA_B200_INCOHERENT", "v7")       # bounce rays on the 4-wide tree: the replicas' copy of it is what renders
    import torch
    w, h, spp = 192, 120, 22            # spp not a multiple of the device count: ranges of different length; 120 rows: ragged bands for 4
    kw = dict(spp=spp, bounces=3, tonemap=api.TONEMAP_REINHARD)

    def build(devices):
        with cases.no_commit(api):
            s, cam = scenes.cornell(gpu, **kw)
        if devices:
            ids = (C.c_int * len(devices))(*devices)
            gpu.terra_b200_scene_set_devices(s.h, ids, len(devices))
        s.commit()
        return s, cam

    multi, cam = build(list(range(n_dev)))
    assert gpu.terra_b200_scene_device_count(multi.h) == n_dev
    gpu.terra_b200_set_fanout_min_paths(multi.h, 1)
    gpu.terra_b200_set_seed(multi.h, 77)
    got_res, got_pix = multi.render(cam, w, h)
    again_res, _ = multi.render(cam, w, h, passes=1)           # fresh framebuffer, next call's seed: runs, every pixel gets spp samples
    st = multi.stats()
    multi.destroy()
    assert (got_res["samples"] == spp).all() and (again_res["samples"] == spp).all()
    assert st.p2p_bytes > 0                                      # the replicas were copied device to device
    # deterministic: a second scene on the same devices with the same seed renders the same bits
    twin, cam_t = build(list(range(n_dev)))
    gpu.terra_b200_set_fanout_min_paths(twin.h, 1)
    gpu.terra_b200_set_seed(twin.h, 77)
    twin_res, twin_pix = twin.render(cam_t, w, h)
    twin.destroy()
    assert np.array_equal(twin_res.view(np.uint8), got_res.view(np.uint8)) and np.array_equal(twin_pix.view(np.uint32), got_pix.view(np.uint32))
    # the call cut into two row chunks (transfers of the first behind the rendering of the second): the same bits
    chunked, cam_c = build(list(range(n_dev)))
    gpu.terra_b200_set_fanout_min_paths(chunked.h, 1)
    gpu.terra_b200_set_fanout_chunk_paths(chunked.h, w * h * spp // n_dev // 5)
    gpu.terra_b200_set_seed(chunked.h, 77)
    chunk_res, chunk_pix = chunked.render(cam_c, w, h)
    chunk_tile, _ = chunked.render(cam_c, w, h, region=(8, 12, 96, 50))           # a region that is not the whole framebuffer, odd rows
    chunked.destroy()
    assert np.array_equal(chunk_res.view(np.uint8), got_res.view(np.uint8)) and np.array_equal(chunk_pix.view(np.uint32), got_pix.view(np.uint32))
    assert (chunk_tile["samples"][12:62, 8:104] == spp).all() and chunk_tile["samples"].sum() == 96 * 50 * spp

    single, cam1 = build(None)
    gpu.terra_b200_set_seed(single.h, 77)
    total = np.zeros((h, w, 3), np.float32)
    for g in range(n_dev):
        base, nxt = spp * g // n_dev, spp * (g + 1) // n_dev
        buf = torch.zeros(w * h * 4, dtype=torch.float32, device="cuda:0")
        torch.cuda.synchronize()
        if nxt > base:
            gpu.terra_b200_render_device(C.byref(cam1), single.h, w, h, 0, 0, w, h, base, nxt - base, C.c_void_p(buf.data_ptr()), None)
        torch.cuda.synchronize()
        part = buf.cpu().numpy().view(api.RESULT_DTYPE).reshape(h, w)
        total = total + part["acc"] if g else part["acc"].copy()
    single.destroy()
    # terra_render seeds call k with f(seed, k) while render_device uses f'(seed): the two entry points draw different
    # numbers, so the comparison is statistical here and bit-exact in structure (samples, finite, band seams)
    rad_multi = got_res["acc"] / spp
    rad_single = total / spp
    assert np.isfinite(rad_multi).all()
    assert np.allclose(rad_multi.mean(axis=(0, 1)), rad_single.mean(axis=(0, 1)), rtol=0.03)
    # no seam at the band boundaries
    # (the image has edges of its own — the light, the wall / floor corners — so the jump across a band boundary is
    # compared with the same rows of the single-device image, not with zero)
    rows, rows_single = rad_multi.mean(axis=(1, 2)), rad_single.mean(axis=(1, 2))
    for g in range(1, n_dev):
        y = h * g // n_dev
        jump, jump_single = rows[y] - rows[y - 1], rows_single[y] - rows_single[y - 1]
        assert abs(jump - jump_single) < 0.3 * max(rows[y], rows[y - 1]) + 1e-3, (y, jump, jump_single)      # ~3 sigma of the row noise at 22 spp
    want_pix = rad_multi / (1.0 + rad_multi)
    want_pix = np.power(want_pix, np.float32(1 / 2.2))
    assert np.allclose(got_pix, want_pix, rtol=1e-4, atol=1e-5)
