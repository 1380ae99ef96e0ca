"""SHA-256 of the sah3 host node array on three scenes (CPU only): a refactor of the builder that changes
any of these changed the tree.  python scripts/bvh_hash.py [expected.json]"""
import hashlib, json, sys, time
sys.path.insert(0, ".")
from terra_b200 import api, scenes

lib = api.TerraLib(api.PRODUCT_LIB)
api.Scene.commit = lambda self: None          # host build only: no device needed
out = {}
for name, make in (("c2", lambda: scenes.sphere_field(lib, spp=4, bounces=4, ggx=True)),
                   ("terrain300k", lambda: scenes.terrain(lib, n_tris=300_000, extent=1000.0, lights=True)),
                   ("cornell", lambda: scenes.cornell(lib))):
    s = make()
    s = s[0] if isinstance(s, tuple) else s
    t0 = time.time()
    lib.lib.terra_b200_build_host_bvh(s.h, 1, None, None)
    dt = time.time() - t0
    nodes = s.bvh_nodes()
    out[name] = hashlib.sha256(nodes.tobytes()).hexdigest()
    print(name, len(nodes) // 64, out[name][:16], "%.3f s" % dt)
if len(sys.argv) > 1:
    want = json.load(open(sys.argv[1]))
    assert want == out, (want, out)
    print("identical")
