#!/bin/bash
# On the GPU box: parity (extract tests) + two short bench runs for the base library and every liborbx_<name>.so variant.
# usage: bash tools/ab_run.sh [pytest-file] name ...
tests=${AB_TESTS:-tests/test_gpu_extract.py}
cp orb_slam3_ros2_b200/liborbx.so /tmp/base.so
for v in base "$@"; do
  if [ $v != base ]; then cp orb_slam3_ros2_b200/liborbx_$v.so orb_slam3_ros2_b200/liborbx.so; else cp /tmp/base.so orb_slam3_ros2_b200/liborbx.so; fi
  echo "== $v"
  (timeout 300 python -m pytest $tests -x -q 2>&1 | tail -1)
  for i in 1 2; do python bench.py --skip-secondary --steps 20 --warmup 3 ${AB_BENCH_ARGS} | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), round(d['ms_per_step'],4), d['stage_ms_last_batch'])"; done
done
cp /tmp/base.so orb_slam3_ros2_b200/liborbx.so
