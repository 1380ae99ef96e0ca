import sys, numpy as np
sys.path.insert(0, '/root/repo')
from terra_b200 import api, scenes
prod = api.TerraLib(); ref = api.TerraLib(api.REFERENCE_LIB)
def rad(r): return r['acc'] / np.maximum(r['samples'], 1).astype(np.float32)[..., None]
w = h = 40
f = lambda x, y: float(np.sqrt(((x - y) ** 2).mean()))
for integ in (1, 0):
    for bounces in (0, 1, 2, 4):
        spp = 1024
        kw = dict(integrator=integ, bounces=bounces, jitter=0.0, spp=spp)
        sr, cam = scenes.cornell(ref, **kw)
        A = rad(sr.ref_render_farm(cam, w, h, spp, tile=8, nproc=8, seed=11)[3])
        B = rad(sr.ref_render_farm(cam, w, h, spp, tile=8, nproc=8, seed=23)[3])
        sp, cam = scenes.cornell(prod, **kw)
        prod.terra_b200_set_seed(sp.h, 7); G = rad(sp.render(cam, w, h)[0])
        prod.terra_b200_set_seed(sp.h, 8); H = rad(sp.render(cam, w, h)[0])
        print(f'integ {integ} bounces {bounces}: rmse AB {f(A,B):.5f} GH {f(G,H):.5f} GA {f(G,A):.5f} HB {f(H,B):.5f} avg-vs-avg {f((G+H)/2,(A+B)/2):.5f} mean cpu {((A+B)/2).mean():.5f} gpu {((G+H)/2).mean():.5f}')
        D = ((G + H) / 2 - (A + B) / 2).mean(-1)
        idx = np.argsort(np.abs(D).ravel())[-6:]
        print('   worst px (row, col, gpu-cpu, cpu):', [(int(i // w), int(i % w), round(float(D.ravel()[i]), 4), round(float(((A+B)/2).mean(-1).ravel()[i]), 4)) for i in idx])
        sr.destroy(); sp.destroy()
