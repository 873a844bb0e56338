#!/bin/bash
# tools/gpu_ab.sh TAG [ENV=VAL ...] — on the GPU box: the GPU parity suite, then the resident bench line (default build) and one more
# bench line per extra "ENV=VAL" argument (A/B of runtime switches)
TAG=${1:-x}; shift
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -12 gpurun_out/${TAG}_pytest.log
run() {  # name, env...
  local name=$1; shift
  env "$@" python bench.py --steps 20 --warmup 3 --skip-secondary > gpurun_out/${TAG}_bench_${name}.json 2>gpurun_out/${TAG}_bench_${name}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_bench_${name}.json")); print("${name}", round(d["value"]), d["stage_ms_last_batch"], d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
except Exception as e:
    print("${name} failed", e, open("gpurun_out/${TAG}_bench_${name}.err").read()[-800:])
PY
}
run default X=1
i=0
for kv in "$@"; do i=$((i+1)); run "ab${i}_${kv//[^A-Za-z0-9]/_}" $kv; done
