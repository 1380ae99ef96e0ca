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
ref = api.TerraLib(api.REFERENCE_LIB)
w, h = cases.TEXTURED_SIZE
rad = lambda res: res["acc"] / np.maximum(res["samples"], 1)[..., None]
g = np.load(REPO / "tests/golden/textured.npz")
fix = g["textured_emissive_acc"] / 2
def report(name, img, obj):
    d = np.abs(img - fix).max(axis=-1)
    bad = d > 1e-4
    print(f"{name}: pixels off {int(bad.sum())} by object {dict((o, int(bad[obj == o].sum())) for o in range(-1, 6))} max {d.max():.4f}")
    ys, xs = np.nonzero(bad)
    for y, x in list(zip(ys, xs))[:4]:
        print(f"      px ({x},{y}) obj {obj[y, x]} fixture {fix[y, x]} got {img[y, x]}")
mode = sys.argv[1]
spp = 512
def cpu():
    s, cam = cases.textured_cornell(scenes, api, ref, integrator=0, spp=spp, bounces=0, jitter=0.0)
    _, _, _, res, _ = s.ref_render_farm(cam, w, h, spp, tile=8, nproc=16, seed=11)
    hits, _ = s.ref_primary_hits(cam, w, h)
    s.destroy()
    obj = (hits["prim"] & 0xff).astype(int); obj[hits["t"] < 0] = -1
    return rad(res), obj
def dev(seed):
    s, cam = cases.textured_cornell(scenes, api, gpu, integrator=0, spp=spp, bounces=0, jitter=0.0)
    if seed is not None:
        gpu.terra_b200_set_seed(s.h, seed)
    res = s.render(cam, w, h)[0]
    s.destroy()
    return rad(res)
if mode == "cpu_first":
    A, obj = cpu(); report("cpu", A, obj)
    report("gpu after cpu, seed 7", dev(7), obj)
    report("gpu again, no seed", dev(None), obj)
else:
    G = dev(7)
    A, obj = cpu()
    report("gpu first seed 7", G, obj); report("cpu after gpu", A, obj)
    report("gpu no seed", dev(None), obj)
