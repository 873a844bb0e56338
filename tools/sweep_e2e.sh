#!/bin/bash
# e2e leg sweep (measurements only): sub-batches x handles x generations in flight.  usage (GPU box): bash tools/sweep_e2e.sh
for sb in ${SB:-1}; do for h in ${HS:-1 2 3}; do for d in ${DS:-2 3 4}; do
  echo -n "subbatches $sb handles $h depth $d -> "
  python bench.py --only-e2e --steps 30 --warmup 3 --subbatches $sb --e2e-handles $h --e2e-depth $d 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d.get('e2e',d); print(round(e['value']))"
done; done; done
