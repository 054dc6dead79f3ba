"""Attribute the warp-stall samples and executed instructions of one ncu capture
(`--set full --import-source on`) to the phases of a kernel, by splitting its SASS stream at
the BAR.SYNC instructions.  Usage: phase_breakdown.py capture.ncu-rep"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
segs, cur = [], dict(n=0, samp=0, inst=0, ops={}, first=None)
tot_s = sum(int(r["# Samples"] or 0) for r in rows)
tot_i = sum(int(r["Instructions Executed"] or 0) for r in rows)
for r in rows:
    src = r["Source"].strip()
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    cur["n"] += 1
    cur["samp"] += int(r["# Samples"] or 0)
    cur["inst"] += int(r["Instructions Executed"] or 0)
    key = op.split(".")[0]
    cur["ops"][key] = cur["ops"].get(key, 0) + int(r["Instructions Executed"] or 0)
    if op.startswith("BAR"):
        segs.append(cur)
        cur = dict(n=0, samp=0, inst=0, ops={}, first=None)
segs.append(cur)
print(f"total samples {tot_s}, warp instructions {tot_i}")
for k, s in enumerate(segs):
    top = sorted(s["ops"].items(), key=lambda kv: -kv[1])[:6]
    print(f"seg {k:2d}: sass {s['n']:5d}  samples {100*s['samp']/max(tot_s,1):5.1f}%  inst {100*s['inst']/max(tot_i,1):5.1f}%  "
          + " ".join(f"{a}:{b/max(tot_i,1)*100:.1f}" for a, b in top))
