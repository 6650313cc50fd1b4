#!/bin/bash
# tools/ab_run.sh v1 v2 ... : device-timed bench of exp/librgs_<v>.so variants of the K2 kernel on ONE box (GPU box only)
for v in "$@"; do
  RGS_B200_LIB=$PWD/exp/librgs_$v.so python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-extras 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline())
print('$v', '%.4g game-steps/s  kernel %.2f ms' % (d['value'], d['roofline']['kernel_ms']))"
done
