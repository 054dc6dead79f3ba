#!/bin/bash
# tuning sweep of the CTA width / front-size classes (development tool)
run() { echo "== $*"; env "$@" python bench.py --no-cpu --steps 3 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), {k:round(v,1) for k,v in d['phases_ms'].items()})
"; }
run JPB_NT_BIG=256
run JPB_NT_BIG=128
run JPB_NT_BIG=128 JPB_NT_MID=64
run JPB_NT_BIG=128 JPB_FRONT_MID=32
run JPB_NT_BIG=128 JPB_FRONT_MID=64
run JPB_NT_BIG=128 JPB_FRONT_BIG=96
run JPB_NT_BIG=128 JPB_NT_SMALL=128
run JPB_NT_BIG=256 JPB_FRONT_BIG=160
