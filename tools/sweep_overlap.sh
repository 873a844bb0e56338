#!/bin/bash
# Overlap sweep (measurements only): persistent-grid caps of FAST / describe x sub-batches x resident handles.
# usage (on the GPU box): bash tools/sweep_overlap.sh out_dir
out=${1:-gpurun_out/sweep}; mkdir -p $out
run() { # name, env..., -- bench args
  name=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" python bench.py --skip-secondary --steps 20 --warmup 3 "$@" > $out/$name.json 2>> $out/err.log
  python - "$out/$name.json" "$name" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], round(d["value"]), d["ms_per_step"], d.get("stage_ms_last_batch"))
except Exception as e: print(sys.argv[2], "failed", e)
PY
}
run base X=1 -- 
run sb1 X=1 -- --subbatches 1
run sb3 X=1 -- --subbatches 3
run sb4 X=1 -- --subbatches 4
run rh1 X=1 -- --resident-handles 1
run rh3 X=1 -- --resident-handles 3
run fast3 ORBX_FAST_PERSM=3 --
run fast2 ORBX_FAST_PERSM=2 --
run fast3_sb4 ORBX_FAST_PERSM=3 -- --subbatches 4
run fast3_rh3 ORBX_FAST_PERSM=3 -- --resident-handles 3
run desc3 ORBX_DESC_PERSM=3 --
run fast3_desc3_sb3 ORBX_FAST_PERSM=3 ORBX_DESC_PERSM=3 -- --subbatches 3
