"""robo_game_sim_b200 -- B200-native (sm_100a) batched implementation of the GameSim physics step of
goromal/robo-game-sim, behind the reference's own Python surface.

    from robo_game_py import GameSim            # drop-in for the reference pybind11 module
    from robo_game_sim_b200 import BatchedGameSim, monte_carlo, monte_carlo_batched

The package is a thin host layer over the C ABI in include/rgs_capi.h (librgs_b200.so).  It has no
CPU implementation: importing is cheap, but the first compute call needs the built library and a GPU.
"""
from ._capi import MODE_EXACT_F64, MODE_FAST_F32, MODE_FAST_F64, EV_NAMES, launch_count  # noqa: F401
from .gamesim import BatchedGameSim, GameSim, synth_inputs  # noqa: F401
from .montecarlo import monte_carlo, monte_carlo_batched, reduce_stats, stats_to_dict  # noqa: F401
from .simstate import SimState  # noqa: F401
from . import sampler  # noqa: F401,E402  (shared-state rollout sampling, BASELINE configs[4])
