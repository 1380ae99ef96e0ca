"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/libterra_ref.so,
compiled from /root/reference/src by oracle/Makefile) on the seeded inputs of tests/cases.py.

Run in the build container only (the reference sources do not exist on the GPU box):
    python tests/golden/make_golden.py
The reference ships no golden vectors of its own, so these fixtures — outputs of the
reference itself — are what pins the oracle where the reference .so is not available.
Floats are stored as their bit patterns' float32 values (npz is lossless)."""
import ctypes as C
import hashlib
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tests"))

import cases  # noqa: E402
from terra_b200 import scenes  # noqa: E402
from oracle.binding import harness_api  # noqa: E402

api = harness_api()

OUT = Path(__file__).resolve().parent


def main():
    ref = api.TerraLib(api.REFERENCE_LIB)
    ora = api.OracleLib()      # only oracle_surface_from_normal (builds the input record of the BSDF KAT)

    # 1. BVH node arrays, primary hits, incoherent-ray hits
    out = {}
    for name in cases.PRIMARY:
        s, cam, (w, h) = cases.build(scenes, ref, name)
        nodes = s.ref_bvh_nodes()
        out[f"{name}_nodes_sha256"] = np.frombuffer(hashlib.sha256(nodes.tobytes()).digest(), np.uint8)
        out[f"{name}_nodes_count"] = np.int64(nodes.size // 64)
        if nodes.size <= 1 << 16:
            out[f"{name}_nodes"] = nodes
        hits, dirs = s.ref_primary_hits(cam, w, h)
        out[f"{name}_primary"] = hits
        out[f"{name}_primary_dirs"] = dirs
        o, d = cases.rays_for(s, api, scenes, 4096)
        out[f"{name}_rays_hits"] = s.ref_trace_rays(o, d)
        # surface records at the closest hit of 64 of those rays (terra_surface_init KAT)
        rec = np.zeros((64, 1 + 1 + 3 + 16 + 3 + 3 + 24), np.float32)
        for i in range(64):
            sf = api.ShadingSurface()
            p = np.zeros(3, np.float32)
            tri = C.c_int(0)
            oi = np.ascontiguousarray(o[i])
            di = np.ascontiguousarray(d[i])
            obj = ref.ref_raycast_surface(s.h, s.objects_base, oi.ctypes.data, di.ctypes.data, C.byref(sf), p.ctypes.data, C.byref(tri))
            raw = np.frombuffer(bytes(sf), np.float32)
            rec[i, 0], rec[i, 1] = obj, tri.value if obj >= 0 else -1
            if obj >= 0:
                n_attr = 4
                rec[i, 2:5] = p
                rec[i, 5:21] = raw[0:16]
                rec[i, 21:24] = raw[16:19]
                rec[i, 24:27] = raw[19:22]
                rec[i, 27:27 + 3 * n_attr] = raw[23:23 + 3 * n_attr]     # only the slots the material fills
        out[f"{name}_surfaces"] = rec
        s.destroy()
    np.savez_compressed(OUT / "geometry.npz", **out)

    # 2. single-triangle known answers
    tri, o, d = cases.kat_rays()
    s = api.Scene(ref)
    s.set_diffuse(s.add_object(tri), (0.5, 0.5, 0.5))
    s.set_options()
    s.commit()
    np.savez_compressed(OUT / "kat_triangle.npz", hits=s.ref_trace_rays(o, d))
    s.destroy()

    # 3. deterministic renders (jitter 0, one process, srand pinned)
    out = {}
    w, h = cases.RENDER_SIZE
    for name, kw in cases.RENDER_CASES.items():
        s, cam = scenes.cornell(ref, **kw)
        _, rays, paths, res, pix = s.ref_render_farm(cam, w, h, kw["spp"], tile=max(w, h), nproc=1, seed=cases.RENDER_SEED)
        out[f"{name}_acc"] = res["acc"]
        out[f"{name}_samples"] = res["samples"]
        out[f"{name}_rays"] = np.int64(rays)
        s.destroy()
    # tonemap operators on one case (pixels[] output of terra_render, Terra.c:574-630)
    for op in range(5):
        s, cam = scenes.cornell(ref, **dict(cases.RENDER_CASES["simple_diffuse"], tonemap=op))
        o_ = s.options()
        o_.manual_exposure, o_.gamma = 1.7, 2.2
        s.commit()
        _, _, _, res, pix = s.ref_render_farm(cam, w, h, 4, tile=max(w, h), nproc=1, seed=cases.RENDER_SEED)
        out[f"tonemap{op}_pixels"] = pix
        s.destroy()
    np.savez_compressed(OUT / "renders.npz", **out)

    # 3b. env-lit extension mode: the reference with Terra.c:1056 restored (oracle/Makefile, "patched oracle")
    if api.REFERENCE_ENVLIT_LIB.exists():
        pref = api.TerraLib(api.REFERENCE_ENVLIT_LIB)
        out = {}
        for name, kw in cases.ENVLIT_CASES.items():
            s, cam = cases.envlit_scene(scenes, api, pref, **kw)
            _, rays, paths, res, pix = s.ref_render_farm(cam, w, h, kw["spp"], tile=max(w, h), nproc=1, seed=cases.RENDER_SEED)
            out[f"{name}_acc"] = res["acc"]
            out[f"{name}_samples"] = res["samples"]
            s.destroy()
            # and what the UNPATCHED reference makes of the same scene (parity mode: the environment adds nothing)
            s, cam = cases.envlit_scene(scenes, api, ref, **kw)
            _, _, _, res0, _ = s.ref_render_farm(cam, w, h, kw["spp"], tile=max(w, h), nproc=1, seed=cases.RENDER_SEED)
            out[f"{name}_acc_parity_mode"] = res0["acc"]
            s.destroy()
        np.savez_compressed(OUT / "envlit.npz", **out)

    # 4. BSDF known answers from the reference's live presets (TerraPresets.c:34-146)
    e, wo, nrm = cases.bsdf_kat_inputs()
    out = {}
    out["diffuse"] = api.run_bsdf_callbacks(ora, ref.terra_bsdf_diffuse_init, [(0.7, 0.5, 0.3)], e, wo, nrm)
    out["phong"] = api.run_bsdf_callbacks(ora, ref.terra_bsdf_phong_init,
                                          [(0.5, 0.4, 0.3), (0.3, 0.4, 0.6), (20, 20, 20), (0, 0, 0)], e, wo, nrm)
    out["phong_specular_heavy"] = api.run_bsdf_callbacks(ora, ref.terra_bsdf_phong_init,
                                                         [(0.9, 0.9, 0.9), (0.1, 0.1, 0.1), (60, 60, 60), (0, 0, 0)], e, wo, nrm)
    np.savez_compressed(OUT / "bsdf.npz", **out)

    # 4b. GGX / glass known answers from the reference's DEAD presets (TerraPresets.c:13-28, 298-465), compiled through
    #     the adapter of the retired API in oracle/ref_dead_presets.c (oracle/Makefile, _ref/libterra_ref_dead.so):
    #     wi.xyz, weight (-> pdf), shade.xyz (-> eval, with the trailing * NoL the retired API carried)
    dead = C.CDLL(str(api.REFERENCE_DEAD_LIB))
    out = {}
    for name, (preset, attrs, flip) in cases.DEAD_PRESET_CASES.items():
        e, wo, nrm = cases.dead_preset_inputs(flip)
        out[name] = api.run_bsdf_callbacks(ora, getattr(dead, f"ref_dead_bsdf_{preset}_init"), attrs, e, wo, nrm)[:, :7]
    np.savez_compressed(OUT / "bsdf_dead.npz", **out)

    # 5. textures (Terra.c:368-477): read / point + bilinear sample / lat-long, three address modes
    rng = np.random.default_rng(17)
    tex8 = rng.integers(0, 256, (7, 5, 3), dtype=np.uint8)
    texf = rng.uniform(0, 4, (6, 8, 3)).astype(np.float32)
    uv = rng.uniform(-1, 12, (200, 2)).astype(np.float32)
    uv[:8] = [(0, 0), (4, 6), (4.5, 6.5), (5, 7), (0.25, 0.75), (9.9, 13.2), (3.999, 5.999), (1, 1)]
    uv = np.abs(uv)     # negative coordinates convert to huge size_t in the reference: undefined, not tested
    dirs = rng.normal(size=(200, 3)).astype(np.float32)
    out = {"uv": uv, "dirs": dirs}
    sc = api.Scene(ref)
    for tname, data in (("u8", tex8), ("f32", texf)):
        for filt in (api.FILTER_POINT, api.FILTER_BILINEAR):
            for addr in (api.ADDRESS_WRAP, api.ADDRESS_MIRROR, api.ADDRESS_CLAMP):
                tex = sc.make_texture(data, filter=filt, address=addr)
                r = np.zeros((len(uv), 3), np.float32)
                for i in range(len(uv)):
                    if addr == api.ADDRESS_MIRROR and (int(uv[i, 0]) // data.shape[1]) % 2 == 1 and \
                            (int(uv[i, 0]) % data.shape[1] == 0 or int(uv[i, 1]) % data.shape[0] == 0):
                        r[i] = np.nan      # the reference indexes one past the edge here (Terra.c:385-386): skipped
                        continue
                    t2 = api.Float2(float(uv[i, 0]), float(uv[i, 1]))
                    v = ref.terra_texture_sample(C.addressof(tex), C.addressof(t2), None)
                    r[i] = v.tuple()
                out[f"{tname}_f{filt}_a{addr}"] = r
        tex = sc.make_texture(data, filter=api.FILTER_POINT, address=api.ADDRESS_WRAP)
        r = np.zeros((len(dirs), 3), np.float32)
        for i in range(len(dirs)):
            dd = api.Float3(*dirs[i])
            r[i] = ref.terra_texture_sample_latlong(C.addressof(tex), C.addressof(dd), None).tuple()
        out[f"{tname}_latlong"] = r
    out["tex8"], out["texf"] = tex8, texf
    np.savez_compressed(OUT / "textures.npz", **out)
    # ---- 2-D texture-mapped material attributes through the whole renderer (Terra.c:368-466, 1748-1752, 1804-1810)
    out = {}
    tw, th = cases.TEXTURED_SIZE
    for name, kw in cases.TEXTURED_CASES.items():
        s, cam = cases.textured_cornell(scenes, api, ref, **kw)
        _, rays, paths, res, pix = s.ref_render_farm(cam, tw, th, kw["spp"], tile=max(tw, th), nproc=1, seed=cases.RENDER_SEED)
        out[f"{name}_acc"] = res["acc"]
        out[f"{name}_samples"] = res["samples"]
        hits, _ = s.ref_primary_hits(cam, tw, th)
        out[f"{name}_primary"] = hits
        s.destroy()
    np.savez_compressed(OUT / "textured.npz", **out)
    # ---- discrete distributions (Terra.c:760-846), the reference's exported functions called directly
    np.savez_compressed(OUT / "distribution.npz", **distribution_fixture(api, ref))
    print("golden fixtures written to", OUT)


def distribution_cases():
    """name -> (f (ny,nx), e1, e2): random maps with zero rows / columns / one huge bucket, edge random numbers."""
    rng = np.random.default_rng(41)
    one = np.nextafter(np.float32(1), np.float32(0))
    out = {}
    for name, (ny, nx) in (("d1x1", (1, 1)), ("d3x5", (3, 5)), ("d16x33", (16, 33)), ("d64x128", (64, 128))):
        f = (rng.random((ny, nx)).astype(np.float32) ** 4).astype(np.float32)
        if ny > 3:
            f[2] = 0
            f[2, 0] = 1e-6
            f[:, 3] = 0
            f[5, 7] = 1000.0
        n = 1500
        e1, e2 = rng.random(n).astype(np.float32), rng.random(n).astype(np.float32)
        e1[:4] = [0, one, 0.5, 0.25]
        e2[:4] = [0, one, 0.5, 0.25]
        out[name] = (f, e1, e2)
    return out


def distribution_fixture(api, ref):
    out = {}
    for name, (f, e1, e2) in distribution_cases().items():
        got, marg, ccdf = api.reference_distribution_2d_sample(ref, f, e1, e2)
        out[f"{name}_f"], out[f"{name}_e1"], out[f"{name}_e2"] = f, e1, e2
        out[f"{name}_samples"], out[f"{name}_marginal_cdf"], out[f"{name}_cdf"] = got, marg, ccdf
    return out


if __name__ == "__main__":
    main()
