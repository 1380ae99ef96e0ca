import sys, numpy as np
sys.path.insert(0, '/root/repo')
from terra_b200 import api, scenes
prod = api.TerraLib(); ref = api.TerraLib(api.REFERENCE_LIB)
def rad(r): return r['acc'] / np.maximum(r['samples'], 1)[..., None]
w = h = 40
for integ in (1, 2, 0):
    for bounces in (0, 1, 4):
        spp = 1024
        kw = dict(integrator=integ, bounces=bounces, jitter=0.5, spp=spp)
        sr, cam = scenes.cornell(ref, **kw)
        A = rad(sr.ref_render_farm(cam, w, h, spp, tile=8, nproc=8, seed=11)[3])
        B = rad(sr.ref_render_farm(cam, w, h, spp, tile=8, nproc=8, seed=23)[3])
        sp, cam = scenes.cornell(prod, **kw)
        prod.terra_b200_set_seed(sp.h, 7); G = rad(sp.render(cam, w, h)[0])
        prod.terra_b200_set_seed(sp.h, 8); H = rad(sp.render(cam, w, h)[0])
        f = lambda x, y: float(np.sqrt(((x - y) ** 2).mean()))
        print(f'integ {integ} bounces {bounces}: rmse AB {f(A,B):.5f} GH {f(G,H):.5f} GA {f(G,A):.5f}  mean cpu {((A+B)/2).mean():.5f} gpu {((G+H)/2).mean():.5f}')
        C_ = (A + B) / 2; K = (G + H) / 2
        rows = [(0, 8), (8, 16), (16, 24), (24, 32), (32, 40)]
        print('   row-band mean ratio gpu/cpu:', [round(float(K[a:b].mean() / C_[a:b].mean()), 4) for a, b in rows])
        d = np.abs(K - C_).sum(-1); idx = np.argsort(d.ravel())[-5:]
        print('   worst px:', [(int(i // w), int(i % w), K.reshape(-1, 3)[i].round(3).tolist(), C_.reshape(-1, 3)[i].round(3).tolist()) for i in idx])
        sr.destroy(); sp.destroy()
for mb in (8, 32, 64, 2048):
    print('probe', mb, 'MB stream GB/s', prod.terra_b200_probe_bandwidth(mb << 20, 0), 'gather GB/s', prod.terra_b200_probe_bandwidth(mb << 20, 1))
