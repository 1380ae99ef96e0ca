import os, sys
from pathlib import Path
import numpy as np
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO)); sys.path.insert(0, str(REPO / "tests"))
import cases
from terra_b200 import scenes
from oracle.binding import harness_api
api = harness_api()
ref = api.TerraLib(api.REFERENCE_LIB)
w, h = cases.TEXTURED_SIZE
rad = lambda res: res["acc"] / np.maximum(res["samples"], 1)[..., None]
g = np.load(REPO / "tests/golden/textured.npz")
fix = g["textured_emissive_acc"] / 2
print("numpy", np.__version__, "cores", os.cpu_count())
t8, tf = cases.texture_images()
print("tex8 sum", int(t8.sum()), "texf sum", float(tf.sum()))
def cpu(spp, tile, nproc, seed):
    s, cam = cases.textured_cornell(scenes, api, ref, integrator=0, spp=spp, bounces=0, jitter=0.0)
    _, _, _, res, _ = s.ref_render_farm(cam, w, h, spp, tile=tile, nproc=nproc, seed=seed)
    hits, _ = s.ref_primary_hits(cam, w, h)
    s.destroy(); return rad(res), hits
for spp, tile, nproc in ((2, 48, 1), (2, 8, 16), (512, 48, 1), (512, 8, 1), (512, 8, 16)):
    c, hits = cpu(spp, tile, nproc, 11)
    obj = (hits["prim"] & 0xff).astype(int); obj[hits["t"] < 0] = -1
    d = np.abs(c - fix).max(axis=-1)
    bad = d > 1e-4
    print(f"spp {spp} tile {tile} nproc {nproc}: pixels off {int(bad.sum())} by object {dict((o, int(bad[obj == o].sum())) for o in range(-1, 6))} max {d.max():.4f}")
    ys, xs = np.nonzero(bad)
    for y, x in list(zip(ys, xs))[:5]:
        print(f"      px ({x},{y}) obj {obj[y, x]} fixture {fix[y, x]} got {c[y, x]}")
