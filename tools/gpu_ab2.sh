#!/bin/bash
# tools/gpu_ab2.sh TAG "ENV=VAL ENV2=VAL2" ... — one resident bench line per argument (each argument = one environment)
TAG=${1:-x}; shift
i=0
for kv in "" "$@"; do
  i=$((i+1))
  env $kv python bench.py --steps 20 --warmup 3 --skip-secondary > gpurun_out/${TAG}_ab${i}.json 2>gpurun_out/${TAG}_ab${i}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_ab${i}.json")); print("[${kv}]", round(d["value"]), d["stage_ms_last_batch"], d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
except Exception as e:
    print("[${kv}] failed", e, open("gpurun_out/${TAG}_ab${i}.err").read()[-600:])
PY
done
