"""Split the SASS stream of an ncu source page at its BAR.SYNC instructions and report the share of
warp-stall samples per segment, with a fingerprint of each segment (DMMA / LDG / STG / LDS / STS /
SHFL / REDUX / MUFU counts) so that it can be matched to a phase of the kernel.
Usage: ncu -i X.ncu-rep --page source --csv | python tools/phase_split.py [min_pct]"""
import csv, sys, collections

rows = list(csv.reader(sys.stdin))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
c_src, c_smp, c_exe = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
min_pct = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
KEYS = ("DMMA", "LDG", "STG", "LDS", "STS", "SHFL", "REDUX", "MUFU", "DFMA", "DMUL", "DADD", "ATOM", "BAR", "WARPSYNC")


def new():
    return {"n": 0, "samples": 0, "exe": 0, "ops": collections.Counter(), "stalls": collections.Counter()}


segs, cur = [], new()
for r in rows[hdr_i + 1:]:
    if len(r) <= c_smp:
        continue
    src = r[c_src].strip()
    op = src.split()[0] if not src.startswith("@") else src.split()[1]
    cur["n"] += 1
    cur["samples"] += int(r[c_smp] or 0)
    cur["exe"] += int(r[c_exe] or 0)
    for key in KEYS:
        if op.startswith(key):
            cur["ops"][key] += 1
    for i, h in stall_cols:
        cur["stalls"][h] += int(r[i] or 0)
    if op.startswith("BAR"):
        segs.append(cur)
        cur = new()
segs.append(cur)
tot = sum(s["samples"] for s in segs) or 1
print(f"total samples {tot}, segments {len(segs)}")
for i, s in enumerate(segs):
    pct = 100.0 * s["samples"] / tot
    if pct < min_pct:
        continue
    top = ", ".join(f"{k[6:]} {100.0 * v / max(s['samples'], 1):.0f}%" for k, v in s["stalls"].most_common(3))
    ops = " ".join(f"{k}:{v}" for k, v in sorted(s["ops"].items()))
    print(f"seg {i:3d} {pct:5.1f}%  instrs {s['n']:5d} exe {s['exe']:>11d}  [{ops}]  stalls: {top}")
