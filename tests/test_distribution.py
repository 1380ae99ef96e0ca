"""The reference's discrete distributions (src/Terra.c:760-846; written, exported, called by nothing in the
reference): oracle restatement pinned bit-exact against the reference's own functions — the committed fixture
(outputs of libterra_ref.so, tests/golden/make_golden.py) and, where the reference was compiled, a live call —
and the CUDA path (terra_b200_distribution_2d_sample, through the C ABI) bit-exact against both."""
import sys
from pathlib import Path

import numpy as np
import pytest

from conftest import golden

sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))

CASES = ["d1x1", "d3x5", "d16x33", "d64x128"]


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("name", CASES)
def test_oracle_distribution_matches_reference_fixture(oracle, name):
    g = golden("distribution.npz")
    f, e1, e2 = g[f"{name}_f"], g[f"{name}_e1"], g[f"{name}_e2"]
    got = oracle.distribution_2d_sample(f, e1, e2)
    assert np.array_equal(bits(got), bits(g[f"{name}_samples"]))
    # the cdfs themselves: running float sum in index order, divided by the total afterwards
    ny, nx = f.shape
    cdf = np.zeros((ny, nx), np.float32)
    rows = np.zeros(ny, np.float32)
    for r in range(ny):
        rows[r] = oracle.oracle_distribution_1d_init(f[r].ctypes.data, nx, cdf[r].ctypes.data)
    marg = np.zeros(ny, np.float32)
    oracle.oracle_distribution_1d_init(rows.ctypes.data, ny, marg.ctypes.data)
    assert np.array_equal(bits(cdf), bits(g[f"{name}_cdf"]))
    assert np.array_equal(bits(marg), bits(g[f"{name}_marginal_cdf"]))


def test_oracle_distribution_matches_live_reference(api, oracle, ref):
    import make_golden
    for name, (f, e1, e2) in make_golden.distribution_cases().items():
        want, _, _ = api.reference_distribution_2d_sample(ref, f, e1, e2)
        assert np.array_equal(bits(oracle.distribution_2d_sample(f, e1, e2)), bits(want)), name
    rng = np.random.default_rng(9)
    f = rng.random((40, 70)).astype(np.float32)
    e1, e2 = rng.random(3000).astype(np.float32), rng.random(3000).astype(np.float32)
    want, _, _ = api.reference_distribution_2d_sample(ref, f, e1, e2)
    assert np.array_equal(bits(oracle.distribution_2d_sample(f, e1, e2)), bits(want))


def test_distribution_is_a_distribution(oracle):
    """Size-independent properties: bucket probabilities sum to 1, samples land in their bucket, and the
    empirical bucket frequencies follow f."""
    rng = np.random.default_rng(2)
    ny, nx = 24, 48
    f = rng.random((ny, nx)).astype(np.float32) ** 3
    f[:, 10] = 0
    n = 200_000
    e1, e2 = rng.random(n).astype(np.float32), rng.random(n).astype(np.float32)
    s = oracle.distribution_2d_sample(f, e1, e2)
    row = np.minimum((s[:, 0] * ny).astype(int), ny - 1)
    col = np.minimum((s[:, 1] * nx).astype(int), nx - 1)
    p = f.astype(np.float64) / f.sum(dtype=np.float64)
    assert np.allclose(s[:, 2], p[row, col], rtol=1e-4)
    assert (col != 10).all()
    freq = np.zeros((ny, nx))
    np.add.at(freq, (row, col), 1.0 / n)
    assert np.abs(freq - p).max() < 5 * np.sqrt(p.max() / n) + 1e-4


def test_env_importance_is_luminance_times_sin_theta(oracle):
    rng = np.random.default_rng(4)
    rgb = rng.random((16, 32, 3)).astype(np.float32) * 5
    f = oracle.env_importance(rgb)
    lum = rgb @ np.array([0.212671, 0.715160, 0.072169])
    w = np.sin(3.1416926535 * (np.arange(16) + 0.5) / 16)
    assert np.allclose(f, lum * w[:, None], rtol=1e-5)


# ----------------------------------------------------------------------------- CUDA path
@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_device_distribution_matches_reference_fixture(api, scenes, gpu, name):
    g = golden("distribution.npz")
    f, e1, e2 = g[f"{name}_f"], g[f"{name}_e1"], g[f"{name}_e2"]
    s, cam = scenes.cornell(gpu, spp=1)
    out = np.zeros((len(e1), 4), np.float32)
    gpu.terra_b200_distribution_2d_sample(s.h, f.ctypes.data, f.shape[1], f.shape[0], len(e1), e1.ctypes.data, e2.ctypes.data, out.ctypes.data)
    assert np.array_equal(bits(out), bits(g[f"{name}_samples"]))
    s.destroy()


@pytest.mark.gpu
def test_device_distribution_matches_oracle_on_an_environment_sized_map(api, scenes, gpu, oracle):
    """512 x 1024 importance map of the procedural sky (with its sun), 300k samples: the device's binary
    search returns the bucket, coordinate and pdf of the reference's linear scan, bit for bit."""
    sky = scenes.procedural_sky(1024, 512, seed=5)
    f = oracle.env_importance(np.ascontiguousarray(sky[..., :3], np.float32))
    rng = np.random.default_rng(12)
    n = 300_000
    e1, e2 = rng.random(n).astype(np.float32), rng.random(n).astype(np.float32)
    want = oracle.distribution_2d_sample(f, e1, e2)
    s, cam = scenes.cornell(gpu, spp=1)
    out = np.zeros((n, 4), np.float32)
    gpu.terra_b200_distribution_2d_sample(s.h, f.ctypes.data, f.shape[1], f.shape[0], n, e1.ctypes.data, e2.ctypes.data, out.ctypes.data)
    assert np.array_equal(bits(out), bits(want))
    s.destroy()
