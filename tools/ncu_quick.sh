#!/bin/bash
# tools/ncu_quick.sh v1 v2 ... : a handful of ncu counters of the K2 lane kernel (one launch) per exp/librgs_<v>.so A/B variant
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed_op_branch.sum,smsp__inst_executed_pipe_cbu.sum
for v in "$@"; do
  echo "== $v"
  RGS_B200_LIB=$PWD/exp/librgs_$v.so ncu --metrics $M --clock-control none -k regex:${KREGEX:-rollout_lane} -s 3 -c 1 --csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extras $BENCH_ARGS 2>/dev/null | python -c "
import sys, csv
rows = [r for r in csv.reader(l for l in sys.stdin if l.startswith('\"'))]
for r in rows[1:]:
    print('   %-95s %s %s' % (r[-3], r[-1], r[-2]))
"
done
