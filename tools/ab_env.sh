#!/bin/bash
# same-box A/B of environment knobs: tools/ab_env.sh "FPB_HALO=1" "FPB_TC3_CG=1" ...  (each compared with the default, interleaved, 2 rounds)
for round in 1 2; do
  for setting in "" "$@"; do
    env $setting timeout 300 python bench.py --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('${setting:-default}', round(d['ms_per_step'],2), round(d['register_ms_p50'],2), d['clocks']['sm_mhz'])"
  done
done
