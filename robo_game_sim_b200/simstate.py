"""State view over the 22-double GameSim state vector -- the role of the reference's py/src/SimState.py (index
constants identical to include/SimState.h:14-23; getters return copies, as the reference's do).  Works on one state
[22] or on a batch [n, 22] (the getters slice the LAST axis), so the same controller code can be written for the
per-game and for the batched Monte Carlo loop."""
import numpy as np


class SimState:
    TAS, TBS, PK, A1, A2, B1, B2 = 0, 1, 2, 6, 10, 14, 18
    _PLAYER = {("A", 1): 6, ("A", 2): 10, ("B", 1): 14, ("B", 2): 18}

    def __init__(self, state):
        self.state = np.array(state, dtype=np.float64, copy=True)

    def get_player_state(self, team, player_id):
        try:
            o = self._PLAYER[(team, player_id)]
        except KeyError:
            print("Team or player not recognized! team = A or B, player_id = 1 or 2")   # py/src/SimState.py:30
            return None
        return self.state[..., o:o + 4].copy()

    def get_player_pos(self, team, player_id):
        return self.get_player_state(team, player_id)[..., 0:2]

    def get_player_vel(self, team, player_id):
        return self.get_player_state(team, player_id)[..., 2:4]

    def get_puck_state(self):
        return self.state[..., self.PK:self.PK + 4].copy()

    def get_puck_pos(self):
        return self.get_puck_state()[..., 0:2]

    def get_puck_vel(self):
        return self.get_puck_state()[..., 2:4]

    def get_score(self, team):
        return self.state[..., self.TAS if team == "A" else self.TBS].copy()

    def transpose(self):
        return self.state.transpose().copy()
