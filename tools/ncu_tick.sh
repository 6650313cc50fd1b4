M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__cycles_active.avg,sm__cycles_elapsed.max
ncu --metrics $M --clock-control none -k regex:"rollout_lane|reset|cost|argmin" -s 20 -c 5 --csv python tools/tick.py 2>/dev/null | python -c "
import sys, csv
rows = [r for r in csv.reader(l for l in sys.stdin if l.startswith('\"'))]
for r in rows[1:]:
    print('%-40s %-70s %s %s' % (r[4][:40], r[-3], r[-1], r[-2]))
"
