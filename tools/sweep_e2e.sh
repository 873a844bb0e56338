#!/bin/bash
# e2e leg (host buffers) over handles x generations in flight; one line per setting
for hd in "1 2" "1 3" "1 4" "1 5" "2 2" "2 3" "2 4" "3 3"; do set -- $hd
  python bench.py --only-e2e --steps 20 --warmup 3 --sustained-s 0 --e2e-handles $1 --e2e-depth $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('handles $1 depth $2: e2e', round(d['e2e']['value']), 'bound', round(d['e2e']['host_bound_frames_per_s'] or 0), 'value', round(d['value']))"
done
