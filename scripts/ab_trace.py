#!/usr/bin/env python
"""A/B of the trace kernels inside ONE process: the same seeded ray set / frame against scenes committed under
different TERRA_B200_* switches (they are read when a scene's device state is created). Prints one JSON line per
variant. Not a bench line: no CPU baseline, no roofline — the winners are re-measured with bench.py.

    python scripts/ab_trace.py c3 [--tris 1000000] [--rays 33554432]
    python scripts/ab_trace.py c2 [--spp 64]
"""
import argparse
import ctypes as C
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

C3_VARIANTS = [
    ("default (auto: v8 below 100 MB of nodes + triangles, v7 above)", dict()),
    ("v6 (binary tree, 7 CTAs, 8-byte stack entries)", dict(TERRA_B200_INCOHERENT="v6")),
    ("v7 forced (4-wide tree at any size)", dict(TERRA_B200_INCOHERENT="v7")),
    ("v8 forced (binary tree, 8 CTAs, 4-byte stack entries)", dict(TERRA_B200_INCOHERENT="v8")),
    ("sweep stack 12 rows", dict(TERRA_B200_STACK_SMEM="12")),
    ("sweep stack 24 rows", dict(TERRA_B200_STACK_SMEM="24")),
    ("sweep refetch 18", dict(TERRA_B200_REFETCH="18")),
    ("sweep refetch 26", dict(TERRA_B200_REFETCH="26")),
    ("sweep tri_min 6", dict(TERRA_B200_TRI_MIN="6")),
    ("sweep tri_min 12", dict(TERRA_B200_TRI_MIN="12")),
]
# grid over the two run-time knobs of the rounds kernels: --only grid
C3_VARIANTS += [(f"grid refetch {r} tri_min {t}", dict(TERRA_B200_REFETCH=str(r), TERRA_B200_TRI_MIN=str(t)))
                for r in (20, 22, 24, 26) for t in (6, 8, 10)]
C2_VARIANTS = [
    ("default (v8 + fused raygen)", dict()),
    ("v6 (binary tree, 7 CTAs)", dict(TERRA_B200_INCOHERENT="v6")),
    ("v7 (4-wide tree)", dict(TERRA_B200_INCOHERENT="v7")),
]
SWITCHES = ["TERRA_B200_INCOHERENT", "TERRA_B200_COHERENT", "TERRA_B200_V6_OPT", "TERRA_B200_STACK_SMEM", "TERRA_B200_TRI_MIN",
            "TERRA_B200_REFETCH", "TERRA_B200_FUSED_RAYGEN", "TERRA_B200_NODE_ORDER"]


def set_env(env):
    for k in SWITCHES:
        os.environ.pop(k, None)
    os.environ.update(env)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["c2", "c3", "c5"])
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--rays", type=int, default=1 << 25)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    import torch
    from terra_b200 import api, scenes
    lib = api.TerraLib()
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sp = C.c_void_p(stream.cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    tag = os.environ.get("TERRA_B200_LIB", "default build")

    def timed(fn):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(args.steps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream); fn(); b.record(stream)
            torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        return tot / args.steps

    if args.what == "c3":
        extent = 1000.0 if args.tris <= 2_000_000 else 10000.0
        n = args.rays
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
        want = None
        for name, env in C3_VARIANTS:
            if args.only and args.only not in name:
                continue
            set_env(env)
            s, _ = scenes.terrain(lib, n_tris=args.tris, extent=extent, lights=False)
            ms = timed(lambda: lib.terra_b200_trace_rays_device(s.h, n, C.c_void_p(o4.data_ptr()), C.c_void_p(d4.data_ptr()),
                                                                 C.c_void_p(hits.data_ptr()), sp))
            got = hits[:, 0].clone()
            same = None if want is None else bool(torch.equal(got.view(torch.int32), want.view(torch.int32)))
            if want is None:
                want = got
            # work counters on a 1M-ray sample
            lib.terra_b200_reset_stats(s.h)
            lib.terra_b200_set_counting(s.h, 2)          # the kernel's own fetches (1 = the canonical binary walk)
            m = min(n, 1 << 20)
            lib.terra_b200_trace_rays_device(s.h, m, C.c_void_p(o4.data_ptr()), C.c_void_p(d4.data_ptr()), C.c_void_p(hits.data_ptr()), sp)
            torch.cuda.synchronize()
            st = s.stats()
            lib.terra_b200_set_counting(s.h, 0)
            print(json.dumps({"ab": "c3", "build": tag, "variant": name, "tris": args.tris, "rays": n, "ms": ms,
                              "mrays_per_s": n / ms / 1e3, "hit_t_same_as_first_variant": same,
                              "hit_checksum": int(hits.view(torch.int32).to(torch.int64).sum().item()) if st.nodes_visited is not None else None,
                              "nodes_per_ray": st.nodes_visited / m, "tris_per_ray": st.tris_tested / m}), flush=True)
            s.destroy()
    else:
        w, h, spp = 1920, 1080, args.spp
        res = torch.zeros(w * h * 4, dtype=torch.float32, device="cuda")
        want = None
        for name, env in C2_VARIANTS:
            if args.only and args.only not in name:
                continue
            set_env(env)
            s, cam = (scenes.sphere_field(lib, spp=spp, bounces=4, ggx=True) if args.what == "c2"
                      else scenes.glass_metal(lib, spp=spp, bounces=32))
            lib.terra_b200_set_seed(s.h, 2024)

            def frame():
                res.zero_()
                lib.terra_b200_render_device(C.byref(cam), s.h, w, h, 0, 0, w, h, 0, spp, C.c_void_p(res.data_ptr()), sp)
            ms = timed(frame)
            st = s.stats()
            got = res.clone()
            same = None if want is None else bool(torch.equal(got.view(torch.int32), want.view(torch.int32)))
            if want is None:
                want = got
            print(json.dumps({"ab": args.what, "build": tag, "variant": name, "spp": spp, "ms": ms, "mpaths_per_s": w * h * spp / ms / 1e3,
                              "trace_ms_per_frame": st.trace_kernel_ms / (args.steps + 2), "camera_trace_ms_per_frame": st.coherent_kernel_ms / (args.steps + 2),
                              "image_same_as_first_variant": same}), flush=True)
            s.destroy()


if __name__ == "__main__":
    main()
