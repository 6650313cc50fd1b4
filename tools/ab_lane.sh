#!/bin/bash
# tools/ab_lane.sh "variant[:ENV=VAL]" ... : lane-kernel bench of exp/librgs_<variant>.so (GPU box)
for spec in "$@"; do
  v=${spec%%:*}; envs=${spec#*:}; [ "$envs" = "$spec" ] && envs=""
  env $envs RGS_K2_LANE=1 RGS_B200_LIB=$PWD/exp/librgs_$v.so python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$spec', '%.3e' % d['value'], 'kernel_ms %.2f' % d['roofline']['kernel_ms'])
"
done
