import time, numpy as np, sys
sys.path.insert(0, '/root/repo')
from terra_b200 import api, scenes
prod = api.TerraLib(); ref = api.TerraLib(api.REFERENCE_LIB)
for mode in (0, 1):
    prod.terra_b200_set_bvh_mode(mode)
    for name, kw, wh in [('cornell', {}, (256,256)), ('sphere_field', dict(ggx=False), (480,270)), ('terrain', dict(n_tris=200000), (320,180))]:
        fn = getattr(scenes, name)
        sp, cam = fn(prod, **kw)
        sr, camr = fn(ref, **kw)
        hp = sp.primary_hits(cam, *wh)
        prod.terra_b200_set_counting(sp.h, 1); prod.terra_b200_reset_stats(sp.h); sp.primary_hits(cam, *wh); st = sp.stats(); prod.terra_b200_set_counting(sp.h, 0)
        prod.terra_b200_reset_stats(sp.h)
        t=time.time(); hp = sp.primary_hits(cam, *wh); tg=time.time()-t
        st2 = sp.stats()
        hr,_ = sr.ref_primary_hits(camr, *wh)
        n = wh[0]*wh[1]
        print('mode', mode, name, 'ids equal', (hp['prim']==hr['prim']).mean(), 't equal', (hp['t']==hr['t']).mean(), 'nodes/ray', st.nodes_visited/n, 'tris/ray', st.tris_tested/n,
              'trace ms', st2.trace_kernel_ms, 'Mrays/s', n/st2.trace_kernel_ms/1e3)
        sp.destroy(); sr.destroy()
# C3: incoherent rays vs 1M terrain
for mode in (0, 1):
    prod.terra_b200_set_bvh_mode(mode)
    t=time.time(); sp, cam = scenes.terrain(prod, n_tris=1000000, lights=False); print('terrain 1M commit s', time.time()-t, 'mode', mode)
    o, d = scenes.incoherent_rays((0,-80,0), (1000,80,1000), 4_000_000)
    sp.trace_rays(o[:1000], d[:1000])
    prod.terra_b200_set_counting(sp.h, 1); prod.terra_b200_reset_stats(sp.h); h = sp.trace_rays(o, d); st = sp.stats(); prod.terra_b200_set_counting(sp.h, 0)
    prod.terra_b200_reset_stats(sp.h)
    t=time.time(); h = sp.trace_rays(o, d); dt=time.time()-t; st2 = sp.stats()
    n=len(o)
    print('  C3 1M mode', mode, 'hit frac', (h['t']>=0).mean(), 'nodes/ray', st.nodes_visited/n, 'tris/ray', st.tris_tested/n, 'kernel ms', st2.trace_kernel_ms, 'Mrays/s kernel', n/st2.trace_kernel_ms/1e3, 'e2e Mrays/s', n/dt/1e6)
    if mode == 0:
        sr, _ = scenes.terrain(ref, n_tris=1000000, lights=False)
        t=time.time(); hr = sr.ref_trace_rays(o[:20000], d[:20000]); tr=time.time()-t
        print('  ref 20000 rays s', tr, 'ids equal', (hr['prim']==h['prim'][:20000]).mean(), 't equal', (hr['t']==h['t'][:20000]).mean())
        href = hr
    else:
        print('  vs ref ids equal', (href['prim']==h['prim'][:20000]).mean(), 't equal', (href['t']==h['t'][:20000]).mean())
    sp.destroy()
# C2 render timing, sah3
prod.terra_b200_set_bvh_mode(1)
sp, cam = scenes.sphere_field(prod, spp=8, bounces=4)
prod.terra_b200_set_seed(sp.h, 1)
sp.render(cam, 1920, 1080)
prod.terra_b200_reset_stats(sp.h)
t=time.time(); res, pix = sp.render(cam, 1920, 1080); dt=time.time()-t; st = sp.stats()
print('C2 1080p 8spp: s', dt, 'Mpaths/s', 1920*1080*8/dt/1e6, 'render_ms', st.render_ms, 'trace_ms', st.trace_kernel_ms, 'rays', st.rays_traced, 'launches', st.kernel_launches, 'mean', (res['acc']/res['samples'][...,None]).mean(axis=(0,1)))
