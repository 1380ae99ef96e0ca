import time, numpy as np, sys
sys.path.insert(0, '/root/repo')
import __graft_entry__ as g
t=time.time(); g.smoke(); print('smoke s', time.time()-t)
from terra_b200 import api, scenes
prod = api.TerraLib(); ref = api.TerraLib(api.REFERENCE_LIB)
for name, kw, wh in [('cornell', {}, (256,256)), ('sphere_field', dict(ggx=False), (480,270)), ('terrain', dict(n_tris=200000), (320,180))]:
    fn = getattr(scenes, name)
    t=time.time(); sp, cam = fn(prod, **kw); tc=time.time()-t
    sr, camr = fn(ref, **kw)
    t=time.time(); hp = sp.primary_hits(cam, *wh); tg=time.time()-t
    t=time.time(); hr,_ = sr.ref_primary_hits(camr, *wh); tr=time.time()-t
    same_id = (hp['prim']==hr['prim']); same_t = (hp['t']==hr['t'])
    print(name, 'commit s', round(tc,2), 'ids equal', same_id.mean(), 't bit-equal', same_t.mean(), 'hit frac', (hr['t']>=0).mean(), 'gpu s', round(tg,4), 'cpu s', round(tr,3))
    if not same_id.all():
        bad = np.argwhere(~same_id)[:5]; print(' mismatches', [(tuple(b), hp[tuple(b)], hr[tuple(b)]) for b in bad])
# render timing cornell
for integ in (0,1,2):
    sp, cam = scenes.cornell(prod, spp=16, integrator=integ); prod.terra_b200_set_seed(sp.h, 5)
    sp.render(cam, 256,256)
    prod.terra_b200_reset_stats(sp.h)
    t=time.time(); res,pix = sp.render(cam, 256,256); dt=time.time()-t
    st = sp.stats()
    sr, camr = scenes.cornell(ref, spp=16, integrator=integ)
    secs, rays, paths, rres, _ = sr.ref_render_farm(camr, 256,256,16, tile=64, nproc=8, seed=3)
    gm = (res['acc']/res['samples'][...,None]).mean(axis=(0,1)); cm = (rres['acc']/rres['samples'][...,None]).mean(axis=(0,1))
    print('integrator', integ, 'gpu', round(dt,4), 's', 'Mpaths/s', 256*256*16/dt/1e6, 'render_ms', st.render_ms, 'trace_ms', st.trace_kernel_ms, 'rays', st.rays_traced, 'launches', st.kernel_launches)
    print('   cpu farm', round(secs,3), 's rays', rays, 'gpu mean', gm, 'cpu mean', cm, 'rel', (gm-cm)/cm)
