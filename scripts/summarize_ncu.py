#!/usr/bin/env python
"""Turns gpurun_out/<tag>_launches.csv (+ <tag>_<kernel>.ncu-rep) into the tracked summaries under profiles/.
usage: python scripts/summarize_ncu.py <tag> [kernel-rep-suffix ...]"""
import collections, csv, subprocess, sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
tag = sys.argv[1]
reps = sys.argv[2:] or ["k_trace"]
out = REPO / "profiles"
out.mkdir(exist_ok=True)
src = REPO / "gpurun_out" / f"{tag}_launches.csv"
if src.exists():
    lines = [l for l in open(src) if not l.startswith("==")]
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = row["Kernel Name"].split("(")[0].replace("void ", "")
        if name.startswith("k_probe_") or name.startswith("at::"):      # bench.py's own L2 probes / torch fills: not the path
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1, "nsecond": 1, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6}[row["Metric Unit"]]
        tot[name] += v; cnt[name] += 1
    s = sum(tot.values())
    with open(out / f"{tag}_launches_summary.md", "w") as f:
        f.write(f"# {tag}: ncu launch list (gpu__time_duration.sum, --clock-control none), per kernel\n\n")
        f.write("Cold-cache, serialised per-launch times: compare SHARES, not absolutes.\n\n| kernel | launches | total ms | share | avg us |\n|---|---|---|---|---|\n")
        for k, v in sorted(tot.items(), key=lambda x: -x[1]):
            f.write(f"| `{k}` | {cnt[k]} | {v / 1e6:.3f} | {v / s:.3f} | {v / cnt[k] / 1e3:.1f} |\n")
    (out / f"{tag}_launches.csv").write_text("".join(lines))
KEYS = """gpu__time_duration.sum launch__registers_per_thread launch__grid_size launch__block_size
sm__warps_active.avg.pct_of_peak_sustained_active dram__bytes_read.sum dram__bytes_write.sum lts__t_bytes.sum
lts__t_sector_hit_rate.pct l1tex__t_sector_hit_rate.pct sm__throughput.avg.pct_of_peak_sustained_elapsed
lts__throughput.avg.pct_of_peak_sustained_elapsed l1tex__throughput.avg.pct_of_peak_sustained_elapsed
gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed smsp__thread_inst_executed_per_inst_executed.ratio
smsp__inst_executed.sum smsp__issue_active.avg.pct_of_peak_sustained_active smsp__warps_eligible.avg.per_cycle_active
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_wait_per_issue_active.ratio smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio
smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio
launch__shared_mem_per_block_static launch__shared_mem_per_block_dynamic launch__occupancy_limit_registers launch__occupancy_limit_shared_mem sm__cycles_active.avg
sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active
l1tex__data_pipe_lsu_wavefronts.avg l1tex__data_pipe_lsu_wavefronts_mem_shared.avg l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg
l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum smsp__inst_executed_pipe_alu.sum smsp__inst_executed_pipe_fma.sum smsp__inst_executed_pipe_lsu.sum""".split()
for rep in reps:
    p = REPO / "gpurun_out" / f"{tag}_{rep}.ncu-rep"
    c = REPO / "gpurun_out" / f"{tag}_{rep}_raw.csv"          # made on the GPU box when the report is too large to bring back
    if c.exists():
        raw = c.read_text()
    elif p.exists():
        raw = subprocess.run(["ncu", "-i", str(p), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    else:
        continue
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(out / f"{tag}_{rep}_metrics.md", "w") as f:
        f.write(f"# {tag}: ncu --set full --clock-control none, kernel `{rep}` ({len(data)} launches captured)\n\n")
        f.write("| metric | unit | " + " | ".join(f"launch {i}" for i in range(len(data))) + " |\n|---|---|" + "---|" * len(data) + "\n")
        kn = hdr.index("Kernel Name")
        f.write("| kernel | | " + " | ".join(r[kn].split("(")[0] for r in data) + " |\n")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                f.write(f"| {k} | {units[i]} | " + " | ".join(r[i] for r in data) + " |\n")
print("wrote", sorted(x.name for x in out.iterdir()))
