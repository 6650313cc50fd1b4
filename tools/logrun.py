#!/usr/bin/env python
"""tools/logrun.py : the logging configurations of bench_extras alone (game-steps/s and trajectory GB/s at log stride 0 / 10 / 1), for A/B runs (GPU box only)."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from bench_extras import logging_runs
r = logging_runs(6545.9)
for k in ("log_stride_0", "log_stride_10", "log_stride_1", "log_stride_1_field_major"):
    v = r[k]
    print(k, "%.4g game-steps/s  %.2f ms  %.0f GB/s  loss %.3f" % (v["game_steps_per_s"], v["ms"], v["traj_write_gbs"], v.get("loss_vs_no_log", 0.0)))
