import os, sys
from pathlib import Path
import numpy as np
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO)); sys.path.insert(0, str(REPO / "tests"))
import cases
from terra_b200 import scenes
from oracle.binding import harness_api
api = harness_api()
gpu = api.TerraLib()
w, h = cases.TEXTURED_SIZE
rad = lambda res: res["acc"] / np.maximum(res["samples"], 1)[..., None]
def render(spp, env=None):
    for k in ("TERRA_B200_TILE", "TERRA_B200_WARP_SAMPLES", "TERRA_B200_COHERENT", "TERRA_B200_FUSED_RAYGEN"):
        os.environ.pop(k, None)
    os.environ.update(env or {})
    s, cam = cases.textured_cornell(scenes, api, gpu, integrator=0, spp=spp, bounces=0, jitter=0.0)
    res = s.render(cam, w, h)[0]
    hits = s.primary_hits(cam, w, h)
    s.destroy()
    return rad(res), hits
base, hits = render(2)
obj = (hits["prim"] & 0xff).astype(int); obj[hits["t"] < 0] = -1
for spp, env in ((2, None), (4, None), (8, None), (32, None), (64, None), (512, None), (512, {"TERRA_B200_TILE": "0"}),
                 (512, {"TERRA_B200_WARP_SAMPLES": "1"}), (512, {"TERRA_B200_COHERENT": "v6"}), (32, {"TERRA_B200_COHERENT": "v6"})):
    g, _ = render(spp, env)
    d = np.abs(g - base).max(axis=-1)
    bad = d > 1e-4
    print(f"spp {spp} env {env}: pixels off {int(bad.sum())} by object {dict((o, int(bad[obj == o].sum())) for o in range(-1, 6))} max {d.max():.4f}")
    if bad.any():
        ys, xs = np.nonzero(bad)
        for y, x in list(zip(ys, xs))[:6]:
            print(f"      px ({x},{y}) obj {obj[y, x]} base {base[y, x]} got {g[y, x]}")
