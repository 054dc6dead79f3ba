#!/bin/bash
# tuning sweep (development tool): each line = one set of env overrides for bench.py
run() { echo "== $*"; env "$@" python bench.py --no-cpu --steps 3 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), round(d['e2e']['value']), {k:round(v,1) for k,v in d['phases_ms'].items()})
"; }
while read -r line; do
  [ -z "$line" ] && continue
  run $line
done
