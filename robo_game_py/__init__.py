"""Drop-in replacement for the reference's pybind11 extension module `robo_game_py`
(reference src/RoboGamePy.cpp:11-25): `from robo_game_py import GameSim` keeps working for
py/run_sim.py, run_centralized_sim.py, test_dmpc.py, test_collisions.py and MonteCarlo.py."""
from robo_game_sim_b200.gamesim import BatchedGameSim, GameSim  # noqa: F401
