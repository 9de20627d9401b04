#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares from an .ncu-rep (needs -lineinfo + --import-source on).
usage: tools/ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
data = []; fname = ""
ci = cs = None
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0] == "Line No":
        ci = r.index("Instructions Executed"); cs = r.index("# Samples"); continue
    if ci is None or len(r) <= ci or r[0] == "": continue
    try: data.append((int(r[ci]), int(r[cs]), fname, r[0], r[1].strip()[:100]))
    except ValueError: pass
tot = sum(d[0] for d in data); ts = sum(d[1] for d in data)
print(f"total warp-instructions {tot:,}  stall samples {ts:,}")
for n, s, f, ln, t in sorted(data, reverse=True)[:top]:
    print(f"{100*n/tot:5.1f}% inst {100*s/max(ts,1):5.1f}% smp  {f}:{ln:>4s} {t}")
