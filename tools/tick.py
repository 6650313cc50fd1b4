#!/usr/bin/env python
"""tools/tick.py : the BASELINE configs[4] sampler tick alone (device ms), for A/B runs (GPU box only)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from bench_extras import sampler_tick
print(sampler_tick())
