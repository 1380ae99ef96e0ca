"""BSDF presets and texture lookups, CPU side: reference fixtures -> oracle -> the product's host
callbacks (the very tb_bsdf.h / tb_texture.h functions the CUDA kernels are compiled from)."""
import ctypes as C

import numpy as np
import pytest

import cases
from conftest import golden

PRESET_ATTRS = {
    "diffuse": [(0.7, 0.5, 0.3)],
    "phong": [(0.5, 0.4, 0.3), (0.3, 0.4, 0.6), (20, 20, 20), (0, 0, 0)],
    "phong_specular_heavy": [(0.9, 0.9, 0.9), (0.1, 0.1, 0.1), (60, 60, 60), (0, 0, 0)],
    "ggx_metal": [(0.8, 0.6, 0.3), (0.2, 1.0, 1.5)],
    "ggx_plastic": [(0.4, 0.5, 0.6), (0.5, 0.0, 1.5)],
    "ggx_mixed": [(0.4, 0.5, 0.6), (0.35, 0.4, 1.33)],
    "glass": [(0.98, 0.97, 0.96), (1.5, 0, 0), (0, 0, 0)],
}


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same(a, b):
    """Bit-identical, NaN == NaN (Phong's pdf is pow(negative, n) = NaN off-lobe in the reference)."""
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return np.array_equal(bits(a)[~np.isnan(a)], bits(b)[~np.isnan(b)]) and np.array_equal(np.isnan(a), np.isnan(b))


def init_of(lib, case, prefix):
    base = case.split("_")[0]
    return getattr(lib, f"{prefix}{base}_init")


@pytest.mark.parametrize("case", ["diffuse", "phong", "phong_specular_heavy"])
def test_oracle_presets_match_reference_fixture(api, oracle, case):
    e, wo, n = cases.bsdf_kat_inputs()
    out = api.run_bsdf_callbacks(oracle, init_of(oracle, case, "terra_oracle_bsdf_"), PRESET_ATTRS[case], e, wo, n)
    assert same(out, golden("bsdf.npz")[case])


@pytest.mark.parametrize("case", list(PRESET_ATTRS))
def test_product_host_presets_match_oracle(api, prod, oracle, case):
    """GGX / Glass have no live reference implementation: the oracle's restatement of
    TerraPresets.c:298-464 and the product's are written independently and must agree bit for bit (both are held
    against the reference's dead code below)."""
    e, wo, n = cases.bsdf_kat_inputs()
    a = api.run_bsdf_callbacks(oracle, init_of(oracle, case, "terra_oracle_bsdf_"), PRESET_ATTRS[case], e, wo, n)
    b = api.run_bsdf_callbacks(oracle, init_of(prod, case, "terra_bsdf_"), PRESET_ATTRS[case], e, wo, n)
    if case == "glass":     # state slot differs: oracle keeps it in attributes[2].x, column 7 reads slot 3
        a, b = a[:, :7], b[:, :7]
    assert same(a, b)


def _dead_vs_restatement(got, want, wo, n, preset):
    """got: (wi, pdf, f) of a restatement; want: (wi, weight, shade) of the reference's dead preset."""
    assert np.array_equal(bits(got[:, :3]), bits(want[:, :3])), "sampled directions differ from the reference's dead code"
    wi, pdf, f = got[:, :3].astype(np.float64), got[:, 3].astype(np.float64), got[:, 4:7].astype(np.float64)
    weight, shade = want[:, 3].astype(np.float64), want[:, 4:7].astype(np.float64)
    nol = (wi * n).sum(1)
    if preset == "glass":
        # weight: the fresnel term sample() left behind, bit for bit; shade = albedo * fresnel = eval * (n.wi), the
        # form the live integrator's f / pdf * (n.wi) turns back into albedo (eval divides by its own float32 n.wi:
        # 3e-6 relative against the float64 product here at grazing directions)
        assert np.array_equal(bits(got[:, 3]), bits(want[:, 3]))
        assert np.allclose(f * nol[:, None], shade, rtol=1e-5, atol=0)
        return
    # GGX. The retired API cached the half vector in sample(); the live restatement recomputes it as
    # normalize(wi + wo) — a few ulp apart, amplified by D's 1 - NoH^2: 6e-5 relative at most on these inputs.
    # Where sample() clamps HoV to 0 it returns wi = -wo, which has no half vector: the restatement answers pdf = 0,
    # f = 0 there and the dead code a finite weight with shade = 0 (NoL = 0) — the estimator is 0 on both sides.
    degenerate = np.linalg.norm(wi + wo, axis=1) == 0
    assert (pdf[degenerate] == 0).all() and (f[degenerate] == 0).all() and (shade[degenerate] == 0).all()
    ok = ~degenerate
    assert np.allclose(pdf[ok], weight[ok], rtol=1e-4, atol=0)
    assert np.allclose(f[ok] * np.maximum(nol[ok], 0)[:, None], shade[ok], rtol=1e-4, atol=1e-7)   # shade = eval * NoL
    assert 0.5 < ok.mean()


@pytest.mark.parametrize("case", list(cases.DEAD_PRESET_CASES))
def test_ggx_glass_restatements_match_the_reference_dead_code(api, prod, oracle, case):
    """The pin of GGX / glass: the reference's own `#if 0` presets (TerraPresets.c:13-28, 298-465), compiled through the
    adapter of the retired {sample, weight, shade} API in oracle/ref_dead_presets.c, produced tests/golden/bsdf_dead.npz;
    the oracle's restatement AND the product's host callbacks (tb_bsdf.h, what the kernels compile) must reproduce it:
    sampled directions and glass weights bit for bit, GGX weight / shade to 1e-4 relative."""
    preset, attrs, flip = cases.DEAD_PRESET_CASES[case]
    e, wo, n = cases.dead_preset_inputs(flip)
    want = golden("bsdf_dead.npz")[case]
    for lib, prefix in ((oracle, "terra_oracle_bsdf_"), (prod, "terra_bsdf_")):
        got = api.run_bsdf_callbacks(oracle, getattr(lib, f"{prefix}{preset}_init"), attrs, e, wo, n)
        _dead_vs_restatement(got, want, wo.astype(np.float64), n.astype(np.float64), preset)


def test_dead_code_fixture_is_what_the_compiled_dead_code_gives(api, oracle):
    """Where the reference was compiled (oracle/_ref): the committed fixture equals a fresh run of the dead presets."""
    if not api.REFERENCE_DEAD_LIB.exists():
        pytest.skip("oracle/_ref/libterra_ref_dead.so not built on this box")
    dead = C.CDLL(str(api.REFERENCE_DEAD_LIB))
    for case, (preset, attrs, flip) in cases.DEAD_PRESET_CASES.items():
        e, wo, n = cases.dead_preset_inputs(flip)
        got = api.run_bsdf_callbacks(oracle, getattr(dead, f"ref_dead_bsdf_{preset}_init"), attrs, e, wo, n)[:, :7]
        assert same(got, golden("bsdf_dead.npz")[case]), case


def test_ggx_energy_and_glass_weights(api, oracle):
    """No upstream result pins GGX/Glass, so check what physics gives for free: the estimator weight
    f * |n.wi| / pdf of the glass lobe is exactly the albedo (fresnel cancels), and the GGX estimator
    stays finite and mostly below 1 + tolerance for a white material (no energy explosion)."""
    e, wo, n = cases.bsdf_kat_inputs(512, seed=21)
    g = api.run_bsdf_callbacks(oracle, oracle.terra_oracle_bsdf_glass_init, [(0.9, 0.8, 0.7), (1.5, 0, 0), (0, 0, 0)], e, wo, n)
    wi, pdf, f = g[:, 0:3], g[:, 3], g[:, 4:7]
    ndl = (wi * n).sum(1)
    w = f * ndl[:, None] / pdf[:, None]
    assert np.allclose(w, np.array([0.9, 0.8, 0.7], np.float32)[None, :], rtol=1e-4)
    assert np.allclose(np.linalg.norm(wi, axis=1), 1, atol=1e-5)
    x = api.run_bsdf_callbacks(oracle, oracle.terra_oracle_bsdf_ggx_init, [(1, 1, 1), (0.4, 1.0, 1.5)], e, wo, n)
    wi, pdf, f = x[:, 0:3], np.maximum(x[:, 3], 1e-4), x[:, 4:7]
    w = f * (wi * n).sum(1)[:, None] / pdf[:, None]
    assert np.isfinite(w).all()
    assert np.median(np.abs(w)) < 1.5


# ----------------------------------------------------------------------------- textures
@pytest.fixture(scope="module")
def tex_golden():
    return golden("textures.npz")


@pytest.mark.parametrize("tname", ["u8", "f32"])
def test_texture_sampling_matches_reference_fixture(api, prod, tex_golden, tname):
    """terra_texture_read / _sample / _sample_latlong (Terra.c:368-477) of the product's host layer.
    Texel coordinates are not normalised in the reference: uv is used as a texel index."""
    g = tex_golden
    data = g["tex8"] if tname == "u8" else g["texf"]
    uv, dirs = g["uv"], g["dirs"]
    sc = api.Scene(prod)
    for filt in (api.FILTER_POINT, api.FILTER_BILINEAR):
        for addr in (api.ADDRESS_WRAP, api.ADDRESS_MIRROR, api.ADDRESS_CLAMP):
            tex = sc.make_texture(data, filter=filt, address=addr)
            want = g[f"{tname}_f{filt}_a{addr}"]
            for i in range(len(uv)):
                if np.isnan(want[i, 0]):
                    continue          # reference reads one texel past the edge there (mirror mode bug)
                t2 = api.Float2(float(uv[i, 0]), float(uv[i, 1]))
                v = prod.terra_texture_sample(C.addressof(tex), C.addressof(t2), None)
                assert same(np.array(v.tuple(), np.float32), want[i]), (filt, addr, i, uv[i], v.tuple(), want[i])
    tex = sc.make_texture(data, filter=api.FILTER_POINT, address=api.ADDRESS_WRAP)
    for i in range(len(dirs)):
        dd = api.Float3(*dirs[i])
        v = prod.terra_texture_sample_latlong(C.addressof(tex), C.addressof(dd), None)
        assert same(np.array(v.tuple(), np.float32), g[f"{tname}_latlong"][i])
    sc.destroy()


def test_texture_filters_todo_in_reference_return_black(api, prod):
    sc = api.Scene(prod)
    tex = sc.make_texture(np.full((4, 4, 3), 200, np.uint8), filter=api.FILTER_TRILINEAR)
    t2 = api.Float2(1.0, 1.0)
    assert prod.terra_texture_sample(C.addressof(tex), C.addressof(t2), None).tuple() == (0.0, 0.0, 0.0)
    sc.destroy()
