#!/bin/bash
# tools/ab_bench.sh v0 v1 ... : bench every exp/librgs_<v>.so with the lane and the phase kernel (GPU box)
for v in "$@"; do
  for k in lane phase; do
    if [ $k = lane ]; then export RGS_K2_LANE=1; else export RGS_K2_LANE=0; fi
    RGS_B200_LIB=$PWD/exp/librgs_$v.so python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$v $k', '%.3e' % d['value'], 'kernel_ms %.2f' % d['roofline']['kernel_ms'])
"
  done
done
