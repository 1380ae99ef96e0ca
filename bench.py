#!/usr/bin/env python
"""bench.py — the driver's measurement contract for the Terra hot path on B200.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload at N=1 (and, split by sample ranges, at N>1): BASELINE.json configs[1] — the procedural
~100k-triangle sphere field, mixed Lambert / GGX presets, 1920x1080, 256 spp, bounces 4. One STEP =
one full frame = 1920*1080*256 paths through the hot path (camera rays + closest hit -> shade -> ... ->
accumulate -> tonemap). Metric = Mpaths/s (1 path = 1 (pixel, sample) = one terra_trace call).

  value   device-resident: terra_b200_render_device + terra_b200_resolve_device, framebuffer in HBM
  e2e     the reference-facing call: terra_render() on a HOST TerraFramebuffer (H2D of the running
          results[], D2H of results[] + pixels[] inside the timed region)
  roofline  ONE ENTRY PER TRACE KERNEL, the one with more device time named as the dominant one:
          k_trace_packet (camera rays) and the incoherent-ray kernel the library names (k_trace_v8 on scenes that fit
          L2, k_trace_v7 beyond; bounce / visibility rays). achieved = algorithmic bytes
          (32 + 16 + 64*nodes + 36*tris per ray; visit counts from the counting variants of the
          kernels, validated against the oracle's canonical walk in tests/) x rays per launch / mean launch time
          (CUDA events the library records around every trace launch on its stream, per kind of launch).
          bound/peak: geometry that fits L2 (C1, C2, C5, C3 at 1M) is measured against the L2 streaming
          bandwidth probed on this box in this run (terra_b200_probe_bandwidth; hbm_frac beside it for
          reference); C3 at 10M and C4 against MEASURED_PEAKS.json's HBM copy bandwidth.
  cpu_baseline  the UNMODIFIED reference (oracle/_ref) on a FIXED bounded sample of the same workload, all host
          cores: a multi-process tile farm over terra_render for the render workloads, a multi-thread farm over
          terra_bvh_traverse for the traversal workloads. --impl reference runs exactly the same sample.

Other workloads (--workload c1|c3_1m|c3_10m|c4|c5) are for the profiles/ notes; the default is the contract.
Nothing here reads /root/reference. The oracle / reference .so are only ever the CPU baseline."""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent
sys.path.insert(0, str(REPO))

WORKLOADS = {
    "c1": dict(name="C1 Cornell 36 tris Lambert+emissive 256x256 16spp bounces 8", w=256, h=256, spp=16, bounces=8),
    "c2": dict(name="C2 100k-tri sphere field, mixed Lambert/GGX, 1920x1080, 256 spp, bounces 4", w=1920, h=1080, spp=256, bounces=4),
    "c4": dict(name="C4 10M-tri displaced terrain + emissive quads + HDR lat-long env, 3840x2160, 1024 spp, bounces 4", w=3840, h=2160, spp=1024, bounces=4),
    "c5": dict(name="C5 glass+metal divergence stress, 1920x1080, 512 spp, bounces 32 + RR", w=1920, h=1080, spp=512, bounces=32),
    "c3_1m": dict(name="C3 traversal: 64M incoherent rays vs 1M-triangle terrain", rays=64_000_000, tris=1_000_000),
    "c3_10m": dict(name="C3 traversal: 64M incoherent rays vs 10M-triangle terrain", rays=64_000_000, tris=10_000_000),
}


# The CPU legs (in-line cpu_baseline and --impl reference) time THE SAME fixed sample of a workload: every
# `stride`-th 128x128 tile of the frame at `spp` samples per pixel. ~10-30 s of CPU work on a 16-core host.
CPU_SAMPLE = {"c1": dict(stride=1, spp=16), "c2": dict(stride=1, spp=1), "c4": dict(stride=16, spp=1), "c5": dict(stride=1, spp=1)}
REFERENCE_WALK_MAX_TRIS = 2_000_000      # above this the reference overruns its traversal stack (see cpu_traversal_rate)
CPU_RAYS = {"c3_1m": 2_000_000, "c3_10m": 200_000}          # rays of the C3 set the CPU traverses (a 1-4 M subset)
INTEGRATORS = ["simple", "direct", "direct_mis", "debug_mono", "debug_depth", "debug_normals", "debug_mis_weights"]


def line_config(args, wl, world):
    """`config` of the JSON line: the workload and the knobs of the command line. Both arms print the same dict."""
    if "rays" in wl:
        return {"workload": wl["name"], "rays": args.rays or wl["rays"], "triangles": wl["tris"], "bvh": args.bvh,
                "l2": "flushed between timed iterations; rays+hits (3 GB) exceed L2"}
    return {"workload": wl["name"], "width": args.width or wl["w"], "height": args.height or wl["h"], "spp": args.spp or wl["spp"],
            "bounces": wl["bounces"], "integrator": INTEGRATORS[args.integrator], "split": args.split if world > 1 else "none",
            "bvh": args.bvh, "ggx_glass": "restated from the reference's dead code (TerraPresets.c:298-465) and pinned against that code compiled through an adapter of its retired API (tests/golden/bsdf_dead.npz); no live reference path renders them (SURVEY Q5)",
            "l2": "flushed between timed iterations (256 MiB write); path-state waves (up to 128M slots x 80 B) exceed L2"}


def build_scene(lib, scenes, workload, spp, integrator=0):
    if workload == "c1":
        return scenes.cornell(lib, spp=spp, bounces=8, integrator=integrator)
    if workload == "c2":
        return scenes.sphere_field(lib, spp=spp, bounces=4, ggx=True, integrator=integrator)
    if workload == "c4":
        return scenes.terrain(lib, n_tris=10_000_000, extent=10000.0, spp=spp, bounces=4, env_tex=scenes.procedural_sky(), integrator=integrator)
    if workload == "c5":
        return scenes.glass_metal(lib, spp=spp, bounces=32, integrator=integrator)
    raise ValueError(workload)


def c3_rays(extent, n, seed=7):
    """The C3 ray set on the host (numpy): origin ~U(bbox inflated 10 %), direction uniform on the sphere."""
    from terra_b200 import scenes
    lo = np.array([0, -0.09 * extent, 0], np.float32)
    hi = np.array([extent, 0.09 * extent, extent], np.float32)
    return scenes.incoherent_rays(lo, hi, n, seed=seed)


# --------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu_index = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu_index)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
            t0 = time.time()                      # nvidia-smi takes ~0.2 s to print its first line: a short timed
            while not self.rows and time.time() - t0 < 2.0:     # region must not end before the sampler is running
                time.sleep(0.02)
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------- CPU arm
def cpu_reference_rate(workload, integrator=0, seed=1, steps=1, warmup=0):
    """Times the reference's own CPU renderer on the FIXED bounded sample of `workload` (CPU_SAMPLE).

    kind "reference": oracle/_ref/libterra_ref.so = the unmodified reference sources, driven the way
    its only caller does (one terra_render per tile from every host core, satellite/src/Renderer.cpp:
    316-349) — as processes, because glibc rand() serialises threads. GGX spheres get the oracle's
    callbacks through the function-pointer table (the reference has no live GGX preset).
    kind "port": the oracle's renderer, single core, when the reference .so is absent.
    Returns (list of per-step Mpaths/s, info dict)."""
    from terra_b200 import scenes
    from oracle.binding import harness_api
    api = harness_api()
    wl = WORKLOADS[workload]
    w, h = wl["w"], wl["h"]
    cores = os.cpu_count() or 1
    tile = 128
    tx, ty = (w + tile - 1) // tile, (h + tile - 1) // tile
    stride, spp_cpu = CPU_SAMPLE[workload]["stride"], CPU_SAMPLE[workload]["spp"]
    if api.REFERENCE_LIB.exists():
        ref = api.TerraLib(api.REFERENCE_LIB)
        t0 = time.time()
        s, cam = build_scene(ref, scenes, workload, spp_cpu, integrator)
        build_s = time.time() - t0
        n_tiles = len(range(0, tx * ty, stride))
        # the 1-core figure: a sixteenth of the sample's tiles rendered by ONE process
        one_secs, _, one_paths, _, _ = s.ref_render_farm(cam, w, h, spp_cpu, tile=tile, nproc=1, seed=seed, tile_stride=stride * 16)
        one_core = one_paths / max(one_secs, 1e-6) / 1e6
        rates, rays_total, paths_total, secs_total = [], 0, 0, 0.0
        for it in range(warmup + steps):
            secs, rays, paths, _, _ = s.ref_render_farm(cam, w, h, spp_cpu, tile=tile, nproc=cores, seed=seed + it, tile_stride=stride)
            if it >= warmup:
                rates.append(paths / secs / 1e6)
                rays_total += rays; paths_total += paths; secs_total += secs
        s.destroy()
        info = {"kind": "reference", "cores": cores,
                "sample": f"every {stride}th 128x128 tile of the {w}x{h} frame ({n_tiles} tiles) at {spp_cpu} spp, "
                          f"{cores}-process tile farm over libterra_ref.so (unmodified reference, its own x-only SAH tree); "
                          f"{paths_total} paths, {rays_total} rays, {secs_total:.2f} s; reference BVH build {build_s:.2f} s excluded",
                "mrays_per_s": rays_total / secs_total / 1e6 if secs_total else None,
                "one_core_value": one_core}
        return rates, info
    # port: oracle renderer on one core, a centred crop
    prod, ora = api.TerraLib(), api.OracleLib()
    orig = api.Scene.commit
    api.Scene.commit = lambda self: None
    try:
        s, cam = build_scene(prod, scenes, workload, spp_cpu, integrator)
    finally:
        api.Scene.commit = orig
    osc = api.OracleScene(ora, s)
    cw = ch = 32
    rates = []
    for it in range(warmup + steps):
        api.srand(seed + it)
        t0 = time.time()
        osc.render(cam, w, h, region=((w - cw) // 2, (h - ch) // 2, cw, ch))
        if it >= warmup:
            rates.append(cw * ch * spp_cpu / (time.time() - t0) / 1e6)
    osc.destroy(); s.destroy()
    return rates, {"kind": "port", "cores": 1, "sample": f"centre {cw}x{ch} crop of the {w}x{h} frame at {spp_cpu} spp, oracle renderer, 1 core"}


def cpu_traversal_rate(workload, steps=1, warmup=0):
    """C3 on the CPU: the reference's own terra_bvh_create + terra_bvh_traverse (src/TerraBVH.c:128-310) over the first
    CPU_RAYS rays of the same seeded ray set, one slice per host core (threads: the traversal takes no lock and ctypes
    releases the GIL). Returns (list of per-step Mrays/s, info dict)."""
    from concurrent.futures import ThreadPoolExecutor
    from terra_b200 import scenes
    from oracle.binding import harness_api
    api = harness_api()
    wl = WORKLOADS[workload]
    cores = os.cpu_count() or 1
    n = CPU_RAYS[workload]
    extent = 1000.0 if wl["tris"] <= 2_000_000 else 10000.0
    if not api.REFERENCE_LIB.exists():
        return None, None
    ref = api.TerraLib(api.REFERENCE_LIB)
    orig = api.Scene.commit
    api.Scene.commit = lambda self: None          # terra_bvh_create is called directly below: no second build inside commit
    try:
        s, _ = scenes.terrain(ref, n_tris=wl["tris"], extent=extent, lights=False)
    finally:
        api.Scene.commit = orig
    o, d = c3_rays(extent, n)
    o, d = np.ascontiguousarray(o), np.ascontiguousarray(d)
    per = (n + cores - 1) // cores
    slices = [(k, min(n, k + per)) for k in range(0, n, per)]
    # The unmodified reference cannot walk its own tree over 10M triangles: terra_bvh_traverse keeps `int queue[64]` with
    # no overflow check (TerraBVH.c:252) and the process dies with "stack smashing detected". There the
    # oracle's restatement of the same build and the same walk (4096 entries) is timed instead: kind "port".
    port = wl["tris"] > REFERENCE_WALK_MAX_TRIS
    t0 = time.time()
    if port:
        osc = api.OracleScene(api.OracleLib(), s)           # builds the reference tree (restated terra_bvh_create)
        n_nodes = len(osc.bvh_nodes()) // 64
    else:
        bvh = s.ref_bvh()
        n_nodes = bvh[1]
    build_s = time.time() - t0

    def run(lo_hi):
        lo, hi = lo_hi
        if port:
            return int((osc.trace_rays(o[lo:hi], d[lo:hi])["t"] >= 0).sum())
        hits = C.c_size_t(0)
        ref.ref_bvh_trace_timed(bvh[0], bvh[1], s.objects_base, hi - lo, o[lo:hi].ctypes.data, d[lo:hi].ctypes.data, C.byref(hits))
        return hits.value

    rates, hits_total, secs_total = [], 0, 0.0
    with ThreadPoolExecutor(max_workers=len(slices)) as pool:
        for it in range(warmup + steps):
            t0 = time.time()
            got = list(pool.map(run, slices))
            secs = time.time() - t0
            if it >= warmup:
                rates.append(n / secs / 1e6)
                hits_total, secs_total = sum(got), secs_total + secs
    # the 1-core figure
    t0 = time.time()
    n_one = min(n, 100_000 if not port else 10_000)
    run((0, n_one))
    one_core = n_one / (time.time() - t0) / 1e6
    if port:
        osc.destroy()
    else:
        ref.ref_bvh_free(bvh[0])
    s.destroy()
    what = ("the oracle's restatement of terra_bvh_traverse (oracle/terra_oracle.c; the unmodified reference overruns its 64-entry "
            "traversal stack on this tree, TerraBVH.c:252)" if port else "terra_bvh_traverse of libterra_ref.so (unmodified reference)")
    info = {"kind": "port" if port else "reference", "cores": len(slices),
            "sample": f"first {n} rays of the seeded C3 set through {what}, the reference's own "
                      f"x-only SAH tree of {n_nodes} nodes, {len(slices)} threads; {hits_total} hits; tree build {build_s:.1f} s excluded",
            "one_core_value": one_core}
    return rates, info


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation on this box's host cores, same metric/config, same sample
    as the in-line cpu_baseline leg."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    t0 = time.time()
    if "rays" in wl:
        rates, info = cpu_traversal_rate(args.workload, steps=args.steps, warmup=args.warmup)
        if rates is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libterra_ref.so is not built on this box"}), flush=True)
            return
        metric, unit, units_per_step, scaling = "incoherent_mrays_per_s", "Mrays/s", args.rays or wl["rays"], "weak"
    else:
        rates, info = cpu_reference_rate(args.workload, integrator=args.integrator, steps=args.steps, warmup=args.warmup)
        metric, unit, scaling = "mpaths_per_s", "Mpaths/s", "strong"
        units_per_step = (args.width or wl["w"]) * (args.height or wl["h"]) * (args.spp or wl["spp"])
    value = float(np.mean(rates))
    cfg = line_config(args, wl, world)
    line = {
        "impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": units_per_step / (value * 1e6) * 1e3,
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "note": "rate measured on the bounded sample named in cpu_baseline.sample; ms_per_step is the full step extrapolated from it",
        "cpu_baseline": dict(info, value=value, unit=unit),
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.time() - t0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------- GPU arm
def measured_peaks():
    p = REPO / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def kernel_sources_sha():
    """sha256 over the CUDA sources: a committed ncu capture speaks for THIS code only if it was taken from it."""
    import hashlib
    hsh = hashlib.sha256()
    for f in sorted((REPO / "terra_b200" / "csrc").glob("*.cu*")):
        hsh.update(f.name.encode()); hsh.update(f.read_bytes())
    return hsh.hexdigest()[:16]


def profile_entry(key):
    """Entry `key` of profiles/traffic.json (numbers read off a committed `ncu --set full` capture: DRAM bytes per ray,
    issue / ALU-pipe utilisation) — only if the capture was taken from the kernel sources of this tree. A capture of
    other code is reported as stale and contributes nothing."""
    p = REPO / "profiles" / "traffic.json"
    if not p.exists():
        return None
    e = json.loads(p.read_text()).get(key)
    if not e:
        return None
    if e.get("kernel_sources_sha") != kernel_sources_sha():
        return {"stale": True, "capture": e.get("capture"), "captured_at_sources": e.get("kernel_sources_sha")}
    return e


def l2_probe(lib):
    """The north-star's denominator: L2 bandwidth measured on this box, in this run, by the library's own
    probes over a 40 MB L2-resident buffer — (i) streaming 128-bit reads, the denominator; (ii) random 64-byte record
    gathers, the access pattern of a traversal, for information."""
    return float(lib.terra_b200_probe_bandwidth(40 << 20, 0)), float(lib.terra_b200_probe_bandwidth(40 << 20, 1))


def commit_ms(lib, s):
    out = (C.c_float * 4)()
    lib.terra_b200_commit_ms(s.h, out)
    return {"bvh_build": out[0], "flatten": out[1], "quantise": out[2], "upload": out[3]}


def kernel_roofline(name, rays, launches, ms, nodes, tris, counted_rays, l2_resident, l2_stream, hbm_peak, hbm_src, profile_key):
    """One roofline entry: algorithmic bytes per ray x rays per launch / mean launch time, against L2 or HBM."""
    if not launches or not rays or not counted_rays:
        return None
    nodes_per_ray, tris_per_ray = nodes / counted_rays, tris / counted_rays
    bytes_per_ray = 32 + 16 + 64 * nodes_per_ray + 36 * tris_per_ray
    rays_per_launch, ms_per_launch = rays / launches, ms / launches
    achieved = bytes_per_ray * rays_per_launch / (ms_per_launch * 1e-3) / 1e9
    note = None
    if not l2_resident and achieved > hbm_peak:
        # Geometry larger than L2 (C3 at 10M, C4), yet the algorithmic bytes per second exceed what HBM can deliver: the
        # upper levels of the tree are served from L2 and L1 (L2 hit rate 42 % in the committed capture), so HBM is not the
        # roof of this launch. The fraction is taken against the L2 roof like the resident configs; hbm_frac and the
        # measured DRAM traffic stay beside it.
        l2_resident = True
        note = "geometry exceeds L2 but algorithmic bytes/s exceed the HBM peak (upper tree levels hit L2): L2 roof used"
    prof = profile_entry(profile_key)
    traffic = prof["dram_bytes_per_ray"] * rays_per_launch if prof and not prof.get("stale") and "dram_bytes_per_ray" in prof else None
    e = {"kernel": name, "bound": "l2" if l2_resident else "hbm", "achieved": achieved,
         "peak": l2_stream if l2_resident else hbm_peak, "unit": "GB/s",
         "frac": achieved / (l2_stream if l2_resident else hbm_peak),
         "peak_source": "L2 streaming read bandwidth probed on this box in this run (terra_b200_probe_bandwidth, 40 MB resident, 128-bit loads)"
                        if l2_resident else hbm_src,
         "hbm_frac": achieved / hbm_peak, "traffic": traffic,
         "bytes_per_ray": bytes_per_ray, "nodes_per_ray": nodes_per_ray, "tris_per_ray": tris_per_ray,
         "rays_per_launch": rays_per_launch, "ms_per_launch": ms_per_launch, "launches": launches,
         "grays_per_s": rays_per_launch / (ms_per_launch * 1e-3) / 1e9, "profile": prof}
    if note:
        e["note"] = note
    return e


def run_render_arm(args):
    import torch
    import torch.distributed as dist
    from terra_b200 import scenes
    from oracle.binding import harness_api
    api = harness_api()
    from terra_b200.multigpu import DistRenderer
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[args.workload]
    w, h = args.width or wl["w"], args.height or wl["h"]
    spp = args.spp or wl["spp"]
    paths_per_step = w * h * spp

    cpu_rates, cpu_info = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:      # before CUDA is initialised: the farm fork()s
        cpu_rates, cpu_info = cpu_reference_rate(args.workload, integrator=args.integrator)

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = api.TerraLib()
    if lib.terra_b200_device_count() <= 0:
        raise SystemExit("bench.py: no CUDA device — the product has no CPU fallback")
    lib.terra_b200_set_device(local)
    lib.terra_b200_set_bvh_mode({"reference": 0, "sah3": 1, "lbvh": 2}[args.bvh])
    t0 = time.time()
    s, cam = build_scene(lib, scenes, args.workload, spp, args.integrator)           # replicated on every GPU
    commit_s = time.time() - t0
    lib.terra_b200_set_seed(s.h, 2024)
    stream = torch.cuda.Stream()                                  # everything timed runs on ONE explicit stream (events included)
    torch.cuda.set_stream(stream)
    r = DistRenderer(lib, s, cam, w, h, mode=args.split)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")       # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        lib.terra_b200_reset_stats(s.h)
        evs = []
        for _ in range(steps):
            flush.zero_()                                             # L2 flush between timed iterations
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            fn()
            b.record(stream)
            evs.append((a, b))
        barrier()
        ms = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)                 # max over ranks
        return float(ms.item()) / steps, s.stats()

    # ---- value: device-resident frame (render on every rank -> one reduce -> resolve on rank 0)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms_dev, st = timed(lambda: r.render(spp), args.steps, args.warmup)
    clk = clocks.stop() if rank == 0 else None
    launches = st.kernel_launches
    rays_all = torch.tensor([st.rays_traced], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(rays_all)                                     # statistics only, outside every timed region
    rays_all = float(rays_all.item())

    # ---- e2e: host framebuffer in, host framebuffer out
    if world == 1:
        fb = api.Framebuffer()
        assert lib.terra_framebuffer_create(C.byref(fb), w, h)
        # warm-up >= 3 calls (the contract's W): the first call of a shape sizes buffers, the second may capture a CUDA graph
        e2e_steps = args.steps if w * h * spp <= (1 << 22) else max(1, args.steps // 2)
        ms_e2e, _ = timed(lambda: lib.terra_render(C.byref(cam), s.h, C.byref(fb), 0, 0, w, h), e2e_steps, max(1, min(3, args.warmup)))
        h2d, d2h = w * h * 16 + 40, w * h * 28
    else:
        # N > 1: the reference-facing call at N GPUs is terra_render on a scene that lives on all N devices (the
        # in-library splitter: sample ranges per GPU, peer-memory reduce-scatter fused into the finalize kernel, every
        # GPU returns its own row band over its own PCIe link). One process drives it — rank 0 — while the other ranks
        # wait at a CPU (gloo) barrier with their GPUs idle; the timed region is the call on a host framebuffer.
        cpu_group = dist.new_group(backend="gloo")
        ms_e2e = 0.0
        if rank == 0:
            orig_commit = api.Scene.commit
            api.Scene.commit = lambda self: None
            try:
                s_all, cam_all = build_scene(lib, scenes, args.workload, spp, args.integrator)
            finally:
                api.Scene.commit = orig_commit
            ids = (C.c_int * world)(*range(world))
            lib.terra_b200_scene_set_devices(s_all.h, ids, world)
            t_commit = time.perf_counter()
            s_all.commit()
            t_commit = time.perf_counter() - t_commit
            lib.terra_b200_set_seed(s_all.h, 2024)
            fb = api.Framebuffer()
            assert lib.terra_framebuffer_create(C.byref(fb), w, h)
            for _ in range(max(1, min(3, args.warmup))):
                lib.terra_render(C.byref(cam_all), s_all.h, C.byref(fb), 0, 0, w, h)
            n_e2e = max(1, args.steps // 2)
            tot = 0.0
            for _ in range(n_e2e):
                flush.zero_(); torch.cuda.synchronize()
                t0 = time.perf_counter()
                lib.terra_render(C.byref(cam_all), s_all.h, C.byref(fb), 0, 0, w, h)      # blocking: returns with the framebuffer final
                tot += (time.perf_counter() - t0) * 1e3
            ms_e2e = tot / n_e2e
            st_all = s_all.stats()
            e2e_note = {"devices": world, "p2p_bytes_at_commit": int(st_all.p2p_bytes), "render_ms_device_clock": st_all.render_ms / (n_e2e + max(1, min(3, args.warmup))),
                        "commit_s_all_devices": t_commit}
            s_all.destroy()
        dist.barrier(group=cpu_group)
        h2d, d2h = w * h * 16 + 40, w * h * 28
    # ---- small workloads are replayed from a CUDA graph, which takes no per-kernel timings: the roofline's
    #      trace-kernel times come from one extra, untimed pass launched kernel by kernel
    if st.trace_kernel_launches == 0:
        lib.terra_b200_set_graph_paths(s.h, 0)
        _, st = timed(lambda: r.render(spp), args.steps, 1)
        lib.terra_b200_set_graph_paths(s.h, 1 << 22)
    # ---- work counters for the rooflines: one small counted frame (never inside a timed region). Camera launches
    #      All launches count the canonical per-ray walk of the binary tree (k_trace<COUNT>): the work the algorithm
    #      needs, not what a kernel happens to fetch (k_trace_v8 fetches ~5 % more nodes, k_trace_v7 half as many wide ones).
    lib.terra_b200_set_counting(s.h, 1)
    lib.terra_b200_reset_stats(s.h)
    r.render(min(spp, 4 * world))
    torch.cuda.synchronize()
    stc = s.stats()
    lib.terra_b200_set_counting(s.h, 0)
    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    hbm_peak, hbm_src = measured_peaks()
    l2_stream, l2_gather = l2_probe(lib)
    l2_resident = args.workload != "c4"
    coh = kernel_roofline("k_trace_packet (camera rays)", st.coherent_rays, st.coherent_kernel_launches, st.coherent_kernel_ms,
                          stc.coherent_nodes_visited, stc.coherent_tris_tested, stc.coherent_rays,
                          l2_resident, l2_stream, hbm_peak, hbm_src, f"k_trace_packet_{args.workload}")
    inc_kernel = lib.terra_b200_incoherent_kernel(s.h).decode()
    inc = kernel_roofline(f"{inc_kernel} (bounce / visibility rays)", st.rays_traced - st.coherent_rays,
                          st.trace_kernel_launches - st.coherent_kernel_launches, st.trace_kernel_ms - st.coherent_kernel_ms,
                          stc.nodes_visited - stc.coherent_nodes_visited, stc.tris_tested - stc.coherent_tris_tested,
                          stc.rays_traced - stc.coherent_rays, l2_resident, l2_stream, hbm_peak, hbm_src, f"{inc_kernel}_{args.workload}")
    kernels = [k for k in (coh, inc) if k]
    for k in kernels:
        k["share_of_step"] = k["ms_per_launch"] * k["launches"] / args.steps / ms_dev if ms_dev else None
    if coh:
        coh["note"] = ("one node fetch serves the 32 rays of a packet; the algorithmic figure charges every ray its own canonical walk "
                       "(SURVEY 8d), so this entry says how much per-ray traffic the packet kernel REPLACES, not what it moves")
    dominant = max(kernels, key=lambda k: k["ms_per_launch"] * k["launches"]) if kernels else {}
    value = paths_per_step / (ms_dev * 1e-3) / 1e6
    e2e = paths_per_step / (ms_e2e * 1e-3) / 1e6
    line = {
        "metric": "mpaths_per_s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": line_config(args, wl, world),
        "e2e": dict({"value": e2e, "unit": "Mpaths/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e},
                    **({"path": "terra_render on one scene replicated on all GPUs (single process, peer-memory reduce-scatter)", **e2e_note}
                       if world > 1 else {"path": "terra_render"})),
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": dict(dominant, kernels=kernels, l2_probe={"stream_gbs": l2_stream, "gather_gbs": l2_gather},
                         kernel_sources_sha=kernel_sources_sha()),
        "extra": {"mrays_per_s": rays_all / args.steps / (ms_dev * 1e-3) / 1e6,
                  "rays_per_path": rays_all / args.steps / paths_per_step,
                  "trace_share_of_step": st.trace_kernel_ms / args.steps / ms_dev if ms_dev else None,
                  "scene_commit_s": commit_s, "commit_ms": commit_ms(lib, s)},
    }
    if cpu_info is not None:
        v = float(np.mean(cpu_rates))
        line["cpu_baseline"] = dict(cpu_info, value=v, unit="Mpaths/s")
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_traversal_arm(args):
    """Config C3: incoherent closest-hit Mrays/s, no shading (terra_scene_raycast minus terra_surface_init)."""
    import torch
    from terra_b200 import scenes
    from oracle.binding import harness_api
    api = harness_api()
    wl = WORKLOADS[args.workload]
    n_rays = args.rays or wl["rays"]
    cpu_rates, cpu_info = None, None
    if not args.no_cpu_baseline:
        cpu_rates, cpu_info = cpu_traversal_rate(args.workload)
    lib = api.TerraLib()
    if lib.terra_b200_device_count() <= 0:
        raise SystemExit("bench.py: no CUDA device — the product has no CPU fallback")
    torch.cuda.set_device(0)
    lib.terra_b200_set_bvh_mode({"reference": 0, "sah3": 1, "lbvh": 2}[args.bvh])
    extent = 1000.0 if wl["tris"] <= 2_000_000 else 10000.0
    t0 = time.time()
    s, cam = scenes.terrain(lib, n_tris=wl["tris"], extent=extent, lights=False)
    commit_s = time.time() - t0
    lo = np.array([0, -0.09 * extent, 0], np.float32)
    hi = np.array([extent, 0.09 * extent, extent], np.float32)
    chunk = 1 << 24
    stream_t = torch.cuda.Stream()
    torch.cuda.set_stream(stream_t)
    g = torch.Generator(device="cuda").manual_seed(7)
    o4 = torch.empty(n_rays, 4, device="cuda")
    d4 = torch.empty(n_rays, 4, device="cuda")
    ext = torch.tensor(hi - lo, device="cuda")
    lo_t = torch.tensor(lo, device="cuda")
    for a in range(0, n_rays, chunk):                                  # origin ~U(bbox inflated 10%), direction uniform on the sphere
        m = min(chunk, n_rays - a)
        o4[a:a + m, :3] = lo_t - 0.05 * ext + torch.rand(m, 3, device="cuda", generator=g) * (1.1 * ext)
        z = torch.rand(m, device="cuda", generator=g) * 2 - 1
        ph = torch.rand(m, device="cuda", generator=g) * (2 * np.pi)
        rr = torch.sqrt(torch.clamp(1 - z * z, min=0))
        d = torch.stack([rr * torch.cos(ph), z, rr * torch.sin(ph)], 1)
        d4[a:a + m, :3] = d / d.norm(dim=1, keepdim=True)
    o4[:, 3] = 0
    d4[:, 3] = 0
    hits = torch.empty(n_rays, 4, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    stream = C.c_void_p(stream_t.cuda_stream)

    def step():
        lib.terra_b200_trace_rays_device(s.h, n_rays, C.c_void_p(o4.data_ptr()), C.c_void_p(d4.data_ptr()), C.c_void_p(hits.data_ptr()), stream)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    lib.terra_b200_reset_stats(s.h)
    clocks = ClockSampler(0)
    clocks.start()
    tot = 0.0
    for _ in range(args.steps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream_t); step(); b.record(stream_t)
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    clk = clocks.stop()
    st = s.stats()
    ms = tot / args.steps
    # e2e through host buffers on a 16M-ray slice
    n_e2e = min(n_rays, 1 << 24)
    # (page-locked buffers from terra_b200_host_alloc, as the bench contract asks; best of 3 calls after one
    # warm-up call that sizes the staging buffers; the pageable figure is reported beside it)
    oh_pin, dh_pin = api.PinnedArray(lib, (n_e2e, 3), np.float32), api.PinnedArray(lib, (n_e2e, 3), np.float32)
    hh_pin = api.PinnedArray(lib, n_e2e, api.HIT_DTYPE)
    oh, dh, hh = oh_pin.array, dh_pin.array, hh_pin.array
    oh[:] = o4[:n_e2e, :3].contiguous().cpu().numpy()
    dh[:] = d4[:n_e2e, :3].contiguous().cpu().numpy()
    s.trace_rays(oh, dh, out=hh)
    e2e_s = float("inf")
    for _ in range(3):
        flush.zero_(); torch.cuda.synchronize()
        t0 = time.time()
        s.trace_rays(oh, dh, out=hh)
        e2e_s = min(e2e_s, time.time() - t0)
    if not np.array_equal(hh["t"].view(np.uint32), hits[:n_e2e, 0].contiguous().cpu().numpy().view(np.uint32)):
        raise SystemExit("bench.py: the host-buffer path disagrees with the resident path")
    # pageable caller buffers (plain numpy), output array already touched: the library stages them through its own
    # page-locked bounce buffers with host threads
    oh_page, dh_page = np.array(oh), np.array(dh)
    hh_page = np.empty(n_e2e, api.HIT_DTYPE)
    hh_page.view(np.uint8).fill(0)
    s.trace_rays(oh_page, dh_page, out=hh_page)
    e2e_pageable_s = float("inf")
    for _ in range(3):
        t0 = time.time()
        s.trace_rays(oh_page, dh_page, out=hh_page)
        e2e_pageable_s = min(e2e_pageable_s, time.time() - t0)
    if not np.array_equal(hh_page.view(np.uint32), hh.view(np.uint32)):
        raise SystemExit("bench.py: pageable and page-locked host paths disagree")
    del oh_page, dh_page, hh_page
    # work counters: the rounds kernel's own fetches over a 1M-ray sample (tests/ hold them within 5 % of the oracle's
    # canonical walk of the same tree)
    lib.terra_b200_set_counting(s.h, 1); lib.terra_b200_reset_stats(s.h)
    s.trace_rays(oh[:1_000_000], dh[:1_000_000])
    stc = s.stats(); lib.terra_b200_set_counting(s.h, 0)
    hbm_peak, hbm_src = measured_peaks()
    l2_stream, l2_gather = l2_probe(lib)
    l2_resident = wl["tris"] <= 2_000_000            # 1M triangles: 32 MB quantised nodes + 48 MB triangles < 126 MB L2
    inc_kernel = lib.terra_b200_incoherent_kernel(s.h).decode()
    roof = kernel_roofline(f"{inc_kernel} (incoherent rays)", st.rays_traced, st.trace_kernel_launches, st.trace_kernel_ms,
                           stc.nodes_visited, stc.tris_tested, stc.rays_traced, l2_resident, l2_stream, hbm_peak, hbm_src,
                           f"{inc_kernel}_{args.workload}")
    line = {
        "metric": "incoherent_mrays_per_s", "value": n_rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": line_config(args, wl, 1),
        "e2e": {"value": n_e2e / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": n_e2e * 24, "d2h_bytes_per_step": n_e2e * 16,
                "rays": n_e2e, "host_buffers": "page-locked (terra_b200_host_alloc), best of 3 calls",
                "pageable_value": n_e2e / e2e_pageable_s / 1e6},
        "gpu_launches": int(st.kernel_launches), "clocks": clk,
        "roofline": dict(roof or {}, l2_probe={"stream_gbs": l2_stream, "gather_gbs": l2_gather}, kernel_sources_sha=kernel_sources_sha(),
                         counter="k_trace<COUNT> (the canonical per-ray walk of the binary tree), 1M-ray sample"),
        "extra": {"scene_commit_s": commit_s, "commit_ms": commit_ms(lib, s)},
    }
    if cpu_info is not None:
        line["cpu_baseline"] = dict(cpu_info, value=float(np.mean(cpu_rates)), unit="Mrays/s")
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=list(WORKLOADS))
    ap.add_argument("--split", default="samples", choices=["samples", "rows"])
    ap.add_argument("--bvh", default=os.environ.get("TERRA_B200_BVH", "sah3"), choices=["reference", "sah3", "lbvh"],
                    help="tree the GPU arm walks: the reference's x-only SAH tree, the three-axis SAH tree (default), the GPU-built LBVH")
    ap.add_argument("--integrator", type=int, default=0, help="TerraIntegrator: 0 simple (the contract), 1 direct, 2 direct MIS")
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--rays", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference_arm(args)
    if args.workload.startswith("c3"):
        return run_traversal_arm(args)
    return run_render_arm(args)


if __name__ == "__main__":
    main()
