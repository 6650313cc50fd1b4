#!/usr/bin/env python
"""tools/nvn_rate.py [n_team ...] : fp32 K2 throughput of n-v-n team games alone (BASELINE configs[3]), for A/B runs (GPU box only)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from bench_extras import rollout_rate
for nt in [int(x) for x in sys.argv[1:]] or [4, 7]:
    r = rollout_rate(nt, "fast", 400_000, 200)
    print(nt, "v", nt, "%.4g game-steps/s  %.1f ms" % (r["game_steps_per_s"], r["ms"]))
