"""Summarise an ncu launch list (gpu__time_duration) of `bench.py --profile`: last step only."""
import csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
rows = list(csv.DictReader(lines))
idx = [i for i, r in enumerate(rows) if 'assemble_kernel' in r['Kernel Name'] or 'scale_rhs' in r['Kernel Name']]
start = [i for i, r in enumerate(rows) if 'assemble_kernel' in r['Kernel Name']][-1]
tot = {}
for r in rows[start:]:
    nm = re.sub(r'\(.*', '', r['Kernel Name']).replace('void jpb::', '').replace('jpb::', '')
    t = float(r['Metric Value']) / 1e3
    key = nm if ('factor' in nm or 'assemble_level' in nm) else nm.split('<')[0]
    tot[key] = tot.get(key, 0) + t
    if len(sys.argv) > 2 and sys.argv[2] in nm:
        print(f"{nm[-22:]:22s} {r['Grid Size']:>16s} {t:8.1f}")
print({k: round(v) for k, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v > 30})
