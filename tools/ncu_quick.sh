#!/bin/bash
# tools/ncu_quick.sh "variant[:ENV=VAL]" ... : a handful of ncu counters of the lane kernel (one launch) per A/B variant
M=gpu__time_duration.sum,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio
for spec in "$@"; do
  v=${spec%%:*}; envs=${spec#*:}; [ "$envs" = "$spec" ] && envs=""
  echo "== $spec"
  env $envs RGS_K2_LANE=1 RGS_B200_LIB=$PWD/exp/librgs_$v.so ncu --metrics $M --clock-control none -k regex:${KREGEX:-rollout_lane} -s 3 -c 1 --csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e $BENCH_ARGS 2>/dev/null | python -c "
import sys, csv
rows = [r for r in csv.reader(l for l in sys.stdin if l.startswith('\"'))]
for r in rows[1:]:
    print('   %-90s %s %s' % (r[-3], r[-1], r[-2]))
"
done
