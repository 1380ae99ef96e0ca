"""Diagnostic: where does the GPU image of the textured Cornell differ from the CPU renderer's?"""
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
for bounces in (0, 1, 4):
    spp = 512
    def cpu(seed):
        s, cam = cases.textured_cornell(scenes, api, ref, integrator=0, spp=spp, bounces=bounces, jitter=0.0)
        _, _, _, res, _ = s.ref_render_farm(cam, w, h, spp, tile=8, nproc=16, seed=seed)
        s.destroy(); return rad(res)
    A, B = cpu(11), cpu(23)
    out = {}
    for fused in ("1", "0"):
        os.environ["TERRA_B200_FUSED_RAYGEN"] = fused
        s, cam = cases.textured_cornell(scenes, api, gpu, integrator=0, spp=spp, bounces=bounces, jitter=0.0)
        gpu.terra_b200_set_seed(s.h, 7)
        G = rad(s.render(cam, w, h)[0])
        hits = s.primary_hits(cam, w, h)
        s.destroy()
        out[fused] = G
        obj = (hits["prim"] & 0xff).astype(int); obj[hits["t"] < 0] = -1
        rm = lambda x, y, m: float(np.sqrt(((x - y)[m] ** 2).mean())) if m.any() else 0.0
        print(f"bounces {bounces} fused {fused}: rmse GA {rm(G, A, obj >= -1):.5f} AB {rm(A, B, obj >= -1):.5f} mean G {G.mean():.5f} A {A.mean():.5f} B {B.mean():.5f}")
        for o in range(-1, 6):
            m = obj == o
            if m.any():
                print(f"    first-hit object {o}: px {m.sum()} rmse GA {rm(G, A, m):.5f} GB {rm(G, B, m):.5f} AB {rm(A, B, m):.5f}  mean G {G[m].mean():.4f} A {A[m].mean():.4f}")
