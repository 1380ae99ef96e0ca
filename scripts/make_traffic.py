#!/usr/bin/env python
"""profiles/traffic.json from the raw pages of the round's `ncu --set full` captures (gpurun_out/<tag>_*_raw.csv, made by
scripts/gpu_closing.sh) and the plain bench lines of the same commands: DRAM bytes per ray and issue figures per kernel and
workload, stamped with the sha of the kernel sources they were taken from (bench.py ignores entries of other code).
usage: python scripts/make_traffic.py <tag>"""
import csv
import json
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
import bench  # noqa: E402

tag = sys.argv[1]
G = REPO / "gpurun_out"


def rows(name):
    r = list(csv.reader((G / f"{tag}_{name}_raw.csv").read_text().splitlines()))
    hdr, units, data = r[0], r[1], r[2:]
    out = []
    for d in data:
        e = {}
        for h, u, v in zip(hdr, units, d):
            try:
                x = float(v.replace(",", ""))
            except ValueError:
                e[h] = v
                continue
            x *= {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(u, 1)
            e[h] = x
        out.append(e)
    return out


def line(name):
    return json.loads((G / f"{tag}_{name}.log").read_text().strip().splitlines()[-1])


def entry(launches, rays, capture):
    dram = sum(l["dram__bytes_read.sum"] + l["dram__bytes_write.sum"] for l in launches)
    inst = sum(l["smsp__inst_executed.sum"] for l in launches)
    t = sum(l["gpu__time_duration.sum"] for l in launches)
    w = lambda k: sum(l[k] * l["gpu__time_duration.sum"] for l in launches) / t
    return {"dram_bytes_per_ray": dram / rays, "warp_instructions_per_ray": inst / rays,
            "issue_active_frac": w("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100,
            "alu_pipe_frac": w("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active") / 100,
            "active_lanes_per_instruction": w("smsp__thread_inst_executed_per_inst_executed.ratio"),
            "l2_hit_rate": w("lts__t_sector_hit_rate.pct") / 100, "l1_hit_rate": w("l1tex__t_sector_hit_rate.pct") / 100,
            "launches_captured": len(launches), "rays_in_captured_launches": rays,
            "capture": capture, "kernel_sources_sha": bench.kernel_sources_sha()}


c2 = line("plain_c2")
paths = c2["config"]["width"] * c2["config"]["height"] * c2["config"]["spp"]
rpp = c2["extra"]["rays_per_path"]
kt = rows("k_trace")
out = {"_comment": "ncu --set full --clock-control none captures of scripts/gpu_closing.sh; per-ray figures = sums over the captured launches / "
                   "the rays those launches traced (C2: the five trace launches of one 1080p x 32 spp wave; C3: one 8M-ray launch). "
                   "bench.py multiplies dram_bytes_per_ray by its rays_per_launch and ignores entries whose kernel_sources_sha is not the tree's."}
out["k_trace_packet_c2"] = entry([l for l in kt if "k_trace_packet" in l["Kernel Name"]], paths, f"profiles/{tag}_k_trace_metrics.md (launch 0)")
# keys = "<kernel the library names for incoherent launches>_<workload>", as bench.py looks them up
out["k_trace_v8_c2"] = entry([l for l in kt if "k_trace_v8" in l["Kernel Name"]], paths * (rpp - 1), f"profiles/{tag}_k_trace_metrics.md (launches 1-4)")
out["k_trace_v8_c3_1m"] = entry(rows("c3_k_trace"), line("plain_c3")["config"]["rays"], f"profiles/{tag}_c3_k_trace_metrics.md")
out["k_trace_v7_c3_10m"] = entry(rows("c3_10m_k_trace"), line("plain_c3_10m")["config"]["rays"], f"profiles/{tag}_c3_10m_k_trace_metrics.md")
(REPO / "profiles" / "traffic.json").write_text(json.dumps(out, indent=1) + "\n")
print(json.dumps({k: {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items() if a not in ("capture",)} for k, v in out.items() if k != "_comment"}, indent=1))
