"""Per-kernel totals (launches, ms, DRAM GB) of the LAST step in an ncu launch list written with
--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv.
Usage: launch_table.py list.csv [--detail substring]"""
import collections, csv, re, sys
rows = list(csv.DictReader(l for l in open(sys.argv[1]) if not l.startswith("==")))
byid = collections.OrderedDict()
for r in rows:
    byid.setdefault(r["ID"], {"name": r["Kernel Name"], "grid": r["Grid Size"], "block": r["Block Size"]})[r["Metric Name"]] = float(r["Metric Value"])
items = list(byid.values())
last = [i for i, v in enumerate(items) if "assemble_kernel" in v["name"]][-1]
tot = collections.OrderedDict()
for v in items[last:]:
    nm = re.sub(r"\(.*", "", v["name"]).replace("void ", "").replace("jpb::", "")
    key = nm if "factor" in nm else nm.split("<")[0]
    t = tot.setdefault(key, [0, 0.0, 0.0, 0.0])
    t[0] += 1
    t[1] += v.get("gpu__time_duration.sum", 0) / 1e6
    t[2] += v.get("dram__bytes_read.sum", 0) / 1e9
    t[3] += v.get("dram__bytes_write.sum", 0) / 1e9
    if len(sys.argv) > 3 and sys.argv[3] in nm:
        print(f"  {nm:40s} {v['grid']:>14s} {v.get('gpu__time_duration.sum', 0) / 1e3:9.1f} us")
print(f"{'kernel':44s} {'n':>4s} {'ms':>8s} {'rd GB':>7s} {'wr GB':>7s}")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    if v[1] > 0.005:
        print(f"{k:44s} {v[0]:4d} {v[1]:8.2f} {v[2]:7.1f} {v[3]:7.1f}")
print(f"total {sum(v[1] for v in tot.values()):.2f} ms")
