#!/usr/bin/env python
"""Per-source-line share of warp instructions / stall samples from an ncu report (--import-source on).
usage: python scripts/ncu_lines.py <report.ncu-rep> [launch-skip] [top-n]"""
import csv, subprocess, sys
rep = sys.argv[1]; skip = sys.argv[2] if len(sys.argv) > 2 else "0"; topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
cur = None; out = []
for r in csv.reader(raw.splitlines()):
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if len(r) >= 2 and r[0] == "Function Name": print(r[1]); continue
    if len(r) > 10 and r[0].isdigit():
        try: out.append((cur, int(r[0]), r[1][:100], int(r[6] or 0), int(r[7] or 0), int(r[8] or 0)))
        except ValueError: pass
ti = sum(o[4] for o in out); ts = sum(o[3] for o in out)
print("warp instructions", ti, "stall samples", ts)
by = {}
for o in out:
    v = by.setdefault(o[0], [0, 0, 0]); v[0] += o[3]; v[1] += o[4]; v[2] += o[5]
for k, v in by.items(): print(f"{k:30s} samples {v[0]/ts:.3f} inst {v[1]/ti:.3f} threads/inst {v[2]/max(1,v[1]):.1f}")
for o in sorted(out, key=lambda o: -o[4])[:topn]:
    print(f"{o[0]:18s} {o[1]:4d} inst {o[4]/ti:6.3f} smp {o[3]/ts:6.3f} thr {o[5]/max(1,o[4]):5.1f} | {o[2]}")
