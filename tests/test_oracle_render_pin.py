"""Pins the oracle's path integrator, presets, tonemap and textures against the reference.

terra_render is deterministic at subpixel_jitter = 0 once libc's rand() stream is pinned with
srand(): camera rays do not depend on the PCG jitter numbers (Terra.c:1784-1785) and every BSDF /
roulette / light-pick number comes from rand() (Terra.c:115). The oracle draws them in the same
order, so whole images must agree bit for bit with the fixtures made by the reference."""
import ctypes as C

import numpy as np
import pytest

import cases
from conftest import golden


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def oracle_render(api, scenes, prod, oracle, kw, seed=cases.RENDER_SEED):
    w, h = cases.RENDER_SIZE
    with cases.no_commit(api):
        s, cam = scenes.cornell(prod, **kw)
    # the product's preset callbacks are host functions; the oracle calls whatever the material holds,
    # so give the scene the ORACLE's presets (pinned separately below against the product's and the reference's)
    osc = api.OracleScene(oracle, s)
    api.srand(seed + 1)
    res, pix = osc.render(cam, w, h)
    osc.destroy()
    s.destroy()
    return res, pix


@pytest.fixture(scope="module")
def renders():
    return golden("renders.npz")


@pytest.fixture()
def oracle_presets(api, oracle):
    """Scenes built in this test get the oracle's own BSDF callbacks."""
    old = dict(api.Scene.default_preset_overrides)
    api.Scene.default_preset_overrides = {p: getattr(oracle, f"terra_oracle_bsdf_{p}_init") for p in ("diffuse", "phong", "ggx", "glass")}
    yield
    api.Scene.default_preset_overrides = old


@pytest.mark.parametrize("name", list(cases.RENDER_CASES))
def test_render_bit_exact_vs_golden(api, scenes, prod, oracle, oracle_presets, renders, name):
    """terra_trace + integrators simple / direct / direct-MIS (Terra.c:1039-1097, 1339-1587) with
    Diffuse and Phong (TerraPresets.c:34-146): accumulated radiance bit-identical."""
    res, _ = oracle_render(api, scenes, prod, oracle, cases.RENDER_CASES[name])
    assert np.array_equal(res["samples"], renders[f"{name}_samples"])
    assert np.array_equal(bits(res["acc"]), bits(renders[f"{name}_acc"]))


@pytest.mark.parametrize("name", ["simple_phong", "mis_diffuse"])
def test_product_host_presets_render_bit_exact(api, scenes, prod, oracle, renders, name):
    """The host callbacks terra_bsdf_*_init of the PRODUCT install (the same tb_bsdf.h functions the
    CUDA kernels compile) reproduce the reference image when driven by the oracle's integrator."""
    res, _ = oracle_render(api, scenes, prod, oracle, cases.RENDER_CASES[name])
    assert np.array_equal(bits(res["acc"]), bits(renders[f"{name}_acc"]))


@pytest.mark.parametrize("op", range(5))
def test_tonemap_bit_exact_vs_golden(api, scenes, prod, oracle, oracle_presets, renders, op):
    """Terra.c:574-630 incl. the Filmic quirk and Uncharted2 (:1815-1828)."""
    w, h = cases.RENDER_SIZE
    with cases.no_commit(api):
        s, cam = scenes.cornell(prod, **dict(cases.RENDER_CASES["simple_diffuse"], tonemap=op))
    o = s.options()
    o.manual_exposure, o.gamma = 1.7, 2.2
    osc = api.OracleScene(oracle, s)
    api.srand(cases.RENDER_SEED + 1)
    _, pix = osc.render(cam, w, h)
    assert np.array_equal(bits(pix), bits(renders[f"tonemap{op}_pixels"]))
    osc.destroy()
    s.destroy()


def test_reference_live_vs_oracle_render(api, scenes, ref, oracle, oracle_presets):
    """Same check against the reference .so itself, on a case that is not in the fixtures."""
    kw = dict(integrator=2, spp=3, bounces=5, jitter=0.0, materials="phong")
    w, h = 24, 24
    s_ref, cam = scenes.cornell(ref, **kw)     # `ref` scenes use the reference's own presets
    _, _, _, rres, _ = s_ref.ref_render_farm(cam, w, h, 3, tile=32, nproc=1, seed=99)
    s_o, _ = scenes.cornell(ref, **kw)
    osc = api.OracleScene(oracle, s_o)
    api.srand(100)
    res, _ = osc.render(cam, w, h)
    assert np.array_equal(bits(res["acc"]), bits(rres["acc"]))
    osc.destroy(); s_o.destroy(); s_ref.destroy()


def test_oracle_presets_inside_reference_render(api, scenes, ref, oracle):
    """The oracle's Diffuse/Phong callbacks, plugged into the UNMODIFIED reference terra_render through
    the function-pointer table, reproduce the reference's own presets bit for bit. The same route is
    the oracle for GGX/Glass, which the reference only carries as dead code."""
    kw = dict(integrator=1, spp=2, bounces=4, jitter=0.0, materials="phong")
    s1, cam = scenes.cornell(ref, **kw)
    _, _, _, r1, _ = s1.ref_render_farm(cam, 24, 24, 2, tile=32, nproc=1, seed=5)
    old = dict(api.Scene.default_preset_overrides)
    api.Scene.default_preset_overrides = {p: getattr(oracle, f"terra_oracle_bsdf_{p}_init") for p in ("diffuse", "phong")}
    try:
        s2, _ = scenes.cornell(ref, **kw)
    finally:
        api.Scene.default_preset_overrides = old
    _, _, _, r2, _ = s2.ref_render_farm(cam, 24, 24, 2, tile=32, nproc=1, seed=5)
    assert np.array_equal(bits(r1["acc"]), bits(r2["acc"]))
    s1.destroy(); s2.destroy()


# ----------------------------------------------------------------------------- env-lit extension mode
def oracle_envlit_render(api, scenes, prod, oracle, kw, env_lit, seed=cases.RENDER_SEED):
    w, h = cases.RENDER_SIZE
    s, cam = cases.envlit_scene(scenes, api, prod, commit=False, **kw)
    osc = api.OracleScene(oracle, s)
    osc.set_env_lit(env_lit)
    api.srand(seed + 1)
    res, _ = osc.render(cam, w, h)
    osc.destroy(); s.destroy()
    return res


@pytest.mark.parametrize("name", list(cases.ENVLIT_CASES))
def test_envlit_oracle_bit_exact_vs_patched_reference_fixture(api, scenes, prod, oracle, oracle_presets, name):
    """the oracle's env-lit mode (miss adds throughput * environment) against the
    reference compiled with exactly Terra.c:1056 restored (tests/golden/envlit.npz, made by
    oracle/_ref/libterra_ref_envlit.so): constant and lat-long HDR environment (Terra.c:468-477),
    bit for bit; and with the mode off the oracle still equals the UNPATCHED reference."""
    g = golden("envlit.npz")
    res = oracle_envlit_render(api, scenes, prod, oracle, cases.ENVLIT_CASES[name], True)
    assert np.array_equal(bits(res["acc"]), bits(g[f"{name}_acc"]))
    assert res["acc"].sum() > g[f"{name}_acc_parity_mode"].sum()          # the environment does light the scene
    res0 = oracle_envlit_render(api, scenes, prod, oracle, cases.ENVLIT_CASES[name], False)
    assert np.array_equal(bits(res0["acc"]), bits(g[f"{name}_acc_parity_mode"]))


def test_envlit_oracle_vs_patched_reference_live(api, scenes, oracle, oracle_presets):
    if not api.REFERENCE_ENVLIT_LIB.exists():
        pytest.skip("oracle/_ref/libterra_ref_envlit.so not built here; the fixture test covers the pin")
    pref = api.TerraLib(api.REFERENCE_ENVLIT_LIB)
    kw = dict(integrator=2, spp=3, bounces=5, jitter=0.0, env="latlong")
    s, cam = cases.envlit_scene(scenes, api, pref, **kw)
    _, _, _, rres, _ = s.ref_render_farm(cam, 24, 24, 3, tile=32, nproc=1, seed=77)
    old = dict(api.Scene.default_preset_overrides)
    api.Scene.default_preset_overrides = {}
    try:
        s2, _ = cases.envlit_scene(scenes, api, pref, **kw)          # reference presets, oracle integrator
    finally:
        api.Scene.default_preset_overrides = old
    osc = api.OracleScene(oracle, s2)
    osc.set_env_lit(True)
    api.srand(78)
    res, _ = osc.render(cam, 24, 24)
    assert np.array_equal(bits(res["acc"]), bits(rres["acc"]))
    osc.destroy(); s2.destroy(); s.destroy()


@pytest.mark.parametrize("name", list(cases.TEXTURED_CASES))
def test_textured_attributes_bit_exact_vs_golden(api, scenes, prod, oracle, name):
    """2-D texture-mapped albedo / emissive (terra_texture_sample Terra.c:410-466, terra_texture_read :368-408 with
    its three address modes, texcoord interpolation :1748-1752, attribute dispatch :1804-1810) through the whole
    renderer: the oracle's integrator over the PRODUCT's host texture callbacks (tb_texture.h, the code the
    kernels compile) reproduces the fixture the unmodified reference rendered, bit for bit."""
    g = golden("textured.npz")
    w, h = cases.TEXTURED_SIZE
    s, cam = cases.textured_cornell(scenes, api, prod, commit=False, **cases.TEXTURED_CASES[name])
    osc = api.OracleScene(oracle, s)
    api.srand(cases.RENDER_SEED + 1)
    res, _ = osc.render(cam, w, h)
    osc.destroy()
    s.destroy()
    assert np.array_equal(res["samples"], g[f"{name}_samples"])
    assert np.array_equal(bits(res["acc"]), bits(g[f"{name}_acc"]))
    lit = g[f"{name}_acc"].sum(axis=-1) > 0
    assert lit.mean() > 0.5           # the textured walls fill most of the frame: the test is not vacuous
