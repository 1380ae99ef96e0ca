#!/usr/bin/env python
"""Lane states of k_trace_v6's rounds (terra_b200_round_stats, counting variant) on the C3 ray set and on the C2 frame's
bounce rays: the per-phase table of profiles/r2_summary.md.   python scripts/round_stats.py [--tris N]"""
import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
NAMES = ["lane_rounds", "node", "culled_pop", "second_leaf_waiting", "parked_only", "no_ray", "tri_round_lanes", "tri_tested",
         "refill_lanes_x32", "refilled"]


def report(lib, s, what, rays):
    out = (C.c_uint64 * 10)()
    lib.terra_b200_round_stats(s.h, out)
    v = dict(zip(NAMES, [int(x) for x in out]))
    st = s.stats()
    r = {"what": what, "rays": rays, "per_ray": {k: v[k] / rays for k in NAMES},
         "share_of_lane_rounds": {k: v[k] / max(v["lane_rounds"], 1) for k in NAMES[1:6]},
         "lanes_per_tri_round": v["tri_tested"] / max(v["tri_round_lanes"] / 32, 1),
         "lanes_per_refill": v["refilled"] / max(v["refill_lanes_x32"] / 32, 1),
         "warp_rounds_per_ray": v["lane_rounds"] / 32 / rays, "nodes_per_ray": st.nodes_visited / max(st.rays_traced, 1)}
    print(json.dumps(r), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--rays", type=int, default=1 << 22)
    args = ap.parse_args()
    import torch
    from terra_b200 import api, scenes
    lib = api.TerraLib()
    torch.cuda.set_device(0)
    n = args.rays
    extent = 1000.0 if args.tris <= 2_000_000 else 10000.0
    g = torch.Generator(device="cuda").manual_seed(7)
    lo = torch.tensor([0, -0.09 * extent, 0], device="cuda")
    ext = torch.tensor([extent, 0.18 * extent, extent], device="cuda")
    o4 = torch.zeros(n, 4, device="cuda"); d4 = torch.zeros(n, 4, device="cuda")
    o4[:, :3] = lo - 0.05 * ext + torch.rand(n, 3, device="cuda", generator=g) * (1.1 * ext)
    z = torch.rand(n, device="cuda", generator=g) * 2 - 1
    ph = torch.rand(n, device="cuda", generator=g) * (2 * np.pi)
    rr = torch.sqrt(torch.clamp(1 - z * z, min=0))
    d = torch.stack([rr * torch.cos(ph), z, rr * torch.sin(ph)], 1)
    d4[:, :3] = d / d.norm(dim=1, keepdim=True)
    hits = torch.empty(n, 4, device="cuda")
    s, _ = scenes.terrain(lib, n_tris=args.tris, extent=extent, lights=False)
    lib.terra_b200_set_counting(s.h, 2)
    lib.terra_b200_reset_stats(s.h)
    lib.terra_b200_trace_rays_device(s.h, n, C.c_void_p(o4.data_ptr()), C.c_void_p(d4.data_ptr()), C.c_void_p(hits.data_ptr()), None)
    torch.cuda.synchronize()
    report(lib, s, f"C3 {args.tris} triangles, uniform random rays", n)
    s.destroy()
    # C2 frame: bounce rays only (the camera rays run the packet kernel, which keeps no such counters)
    w, h, spp = 1920, 1080, 4
    s, cam = scenes.sphere_field(lib, spp=spp, bounces=4, ggx=True)
    lib.terra_b200_set_counting(s.h, 2)
    lib.terra_b200_reset_stats(s.h)
    res = torch.zeros(w * h * 4, dtype=torch.float32, device="cuda")
    lib.terra_b200_render_device(C.byref(cam), s.h, w, h, 0, 0, w, h, 0, spp, C.c_void_p(res.data_ptr()), None)
    torch.cuda.synchronize()
    st = s.stats()
    report(lib, s, "C2 frame, bounce rays", st.rays_traced - st.coherent_rays)
    s.destroy()


if __name__ == "__main__":
    main()
