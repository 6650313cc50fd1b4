#!/bin/bash
# tools/ab_e2e.sh v1 v2 ... : device-timed and end-to-end (host buffers) bench of exp/librgs_<v>.so variants on ONE box
for v in "$@"; do
  RGS_B200_LIB=$PWD/exp/librgs_$v.so python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('$v', 'device %.4g  e2e %.4g  pass ms %s' % (d['value'], d['e2e']['value'], d['e2e']['per_rank_pass_ms']))"
done
