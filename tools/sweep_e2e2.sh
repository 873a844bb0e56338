#!/bin/bash
# e2e diagnosis (measurements only)
run() { echo -n "$* -> "; python bench.py --only-e2e --steps 20 --warmup 3 "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d.get('e2e',d); print(round(e['value']))"; }
run
run --skip-match
run --subbatches 1
run --subbatches 1 --e2e-handles 1
run --frames 512
run --frames 128
run --frames 512 --e2e-handles 4
python tools/pcie_probe.py
