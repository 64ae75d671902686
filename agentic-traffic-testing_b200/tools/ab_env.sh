#!/bin/bash
# A/B of environment switches on the headline bench, one process per setting, device-timed numbers only:
#   bash agentic-traffic-testing_b200/tools/ab_env.sh "ATX_FUSE_QKV_ROPE=0" "ATX_QKV_SPLITS=8" ...
for cfg in "$@"; do
  env $cfg python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$cfg', 'tok/s', d['value'], 'ttft_ms', d['ttft_p50_ms'], 'decode_step_ms', d['decode_step']['ms'])"
done
