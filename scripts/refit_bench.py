"""terra_b200_refit against terra_scene_commit on the C3 terrains: wall time of both, device time of the refit
kernels, traversal rate after the refit (tree topology from the undeformed mesh) against a fresh build.
python scripts/refit_bench.py [n_tris]"""
import ctypes as C
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from terra_b200 import api, scenes

n_tris = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
extent = 1000.0 if n_tris <= 2_000_000 else 10000.0
lib = api.TerraLib()
t0 = time.time()
s, cam = scenes.terrain(lib, n_tris=n_tris, extent=extent, lights=False)
build_and_commit_s = time.time() - t0
cm = (C.c_float * 4)()
lib.terra_b200_commit_ms(s.h, cm)
lo = np.array([0, -0.09 * extent, 0], np.float32)
hi = np.array([extent, 0.09 * extent, extent], np.float32)
o, d = scenes.incoherent_rays(lo, hi, 4_000_000, seed=7)


def rate():
    s.trace_rays(o[:1000], d[:1000])
    lib.terra_b200_reset_stats(s.h)
    hits = s.trace_rays(o, d)
    st = s.stats()
    return hits, len(o) / (st.trace_kernel_ms * 1e-3) / 1e6


before, rate_before = rate()
objs = (api.Object * s.n_objects).from_address(s.objects_base)
amp = 0.01 * extent
for k in range(s.n_objects):
    n = objs[k].triangles_count
    v = np.ctypeslib.as_array(C.cast(objs[k].triangles, C.POINTER(C.c_float)), (n, 3, 3))
    v[..., 1] += (amp * np.sin(v[..., 0] * (6.0 / extent)) * np.cos(v[..., 2] * (5.0 / extent))).astype(np.float32)
t0 = time.time()
dev_ms = lib.terra_b200_refit(s.h)
refit_wall_s = time.time() - t0
after, rate_after = rate()
# a fresh commit of the deformed mesh: the yardstick for both time and hits
t0 = time.time()
fresh = api.Scene(lib)
for k in range(s.n_objects):
    n = objs[k].triangles_count
    v = np.ctypeslib.as_array(C.cast(objs[k].triangles, C.POINTER(C.c_float)), (n, 3, 3))
    fresh.set_diffuse(fresh.add_object(v), (0.5, 0.5, 0.5))
fresh.set_options()
t1 = time.time()
fresh.commit()
fresh_commit_s = time.time() - t1
want = fresh.trace_rays(o, d)
lib.terra_b200_reset_stats(fresh.h)
fresh.trace_rays(o, d)
rate_fresh = len(o) / (fresh.stats().trace_kernel_ms * 1e-3) / 1e6
same_t = bool(np.array_equal(after["t"].view(np.uint32), want["t"].view(np.uint32)))
print(json.dumps({"triangles": n_tris, "commit_ms_first": [round(float(x), 1) for x in cm], "fresh_commit_wall_s": fresh_commit_s,
                  "refit_wall_s": refit_wall_s, "refit_device_ms": dev_ms, "moved_fraction_of_hits": float((after["t"] != before["t"]).mean()),
                  "t_bit_identical_to_fresh_commit": same_t, "id_mismatch": float((after["prim"] != want["prim"]).mean()),
                  "mrays_per_s_before": rate_before, "mrays_per_s_after_refit": rate_after, "mrays_per_s_fresh_tree": rate_fresh}))
assert same_t
