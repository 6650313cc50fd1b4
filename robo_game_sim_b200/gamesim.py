"""Host-side mirror of the reference's Python surface for the GameSim path.

`GameSim` is a drop-in for the pybind11 class `robo_game_py.GameSim`
(reference src/RoboGamePy.cpp:11-25): same method names, argument order, return types
(fresh float64[22] arrays) and `__repr__`; the five trailing `reset` arguments take the C++ default
values (reference src/GameSim.cpp:24-25) that four of the reference's own scripts rely on
(py/run_sim.py:43, py/test_collisions.py:26, ...).  Every call goes through the C ABI
(include/rgs_capi.h) into the sm_100a kernels; a batch of one game is latency-bound by design and
exists for API compatibility.

`BatchedGameSim` is the new entry point: N independent games resident on one B200, advanced S steps
per fused launch from device-resident action tensors (torch is used for tensor hand-off only).
"""
import ctypes as C
import os

import numpy as np

from . import _capi

_MODES = {"exact": _capi.MODE_EXACT_F64, "fast": _capi.MODE_FAST_F32, "fast64": _capi.MODE_FAST_F64,
          _capi.MODE_EXACT_F64: _capi.MODE_EXACT_F64, _capi.MODE_FAST_F32: _capi.MODE_FAST_F32,
          _capi.MODE_FAST_F64: _capi.MODE_FAST_F64}


def _vec(x, n, name):
    """pybind11/eigen.h accepts any float-convertible 1-D length-n (or (n,1)) array."""
    try:
        a = np.asarray(x, dtype=np.float64)
    except (TypeError, ValueError) as exc:
        raise TypeError(f"{name}: expected {n} floats") from exc
    if a.shape not in ((n,), (n, 1), (1, n)):
        raise TypeError(f"{name}: expected an array of shape ({n},), got {a.shape}")
    return np.ascontiguousarray(a.reshape(n))


def _flat(x):
    t = type(x)
    return t is tuple or t is list or (t is np.ndarray and x.ndim == 1)


class GameSim:
    """Drop-in for robo_game_py.GameSim (one 2v2 game, fp64 event-exact kernel K1)."""

    def __init__(self, device=0):
        self._lib = _capi.load()
        self._h = C.c_void_p()
        _capi.check(self._lib.rgs_single_create(C.byref(self._h), int(device)))
        # run() is called once per game-step by the reference's scripts: its argument and result buffers are made once
        self._act = (C.c_double * 8)()
        self._out = (C.c_double * 22)()
        self._out_np = np.frombuffer(self._out, dtype=np.float64)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.rgs_single_destroy(h)
            self._h = None

    def reset(self, dt, winning_score, x0_ball, noise, log, logname, seed=0, tau_puck=0.1, tau_player=0.5,
              player_mass=1.0, puck_mass=0.5):
        """GameSim::reset (reference src/GameSim.cpp:21-55) -> float64[22]."""
        x0 = _vec(x0_ball, 4, "x0_ball")
        if not isinstance(logname, (str, bytes)):
            raise TypeError("logname: expected str")
        name = logname.encode() if isinstance(logname, str) else logname
        out = np.empty(22, dtype=np.float64)
        _capi.check(self._lib.rgs_single_reset(
            self._h, float(dt), int(winning_score), x0.ctypes.data_as(C.POINTER(C.c_double)), float(noise),
            1 if log else 0, name, int(seed), float(tau_puck), float(tau_player), float(player_mass), float(puck_mass),
            out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def undecided(self):
        """GameSim::undecided (reference src/GameSim.cpp:57-60)."""
        flag = C.c_int32(0)
        _capi.check(self._lib.rgs_single_undecided(self._h, C.byref(flag)))
        return bool(flag.value)

    def run(self, vel_A1, vel_A2, vel_B1, vel_B2):
        """GameSim::run (reference src/GameSim.cpp:118-124): one game-step -> float64[22] copy."""
        a = self._act
        try:    # the common case: four flat length-2 sequences of Python / numpy floats
            if not (_flat(vel_A1) and _flat(vel_A2) and _flat(vel_B1) and _flat(vel_B2)):
                raise TypeError
            a[0], a[1] = vel_A1
            a[2], a[3] = vel_A2
            a[4], a[5] = vel_B1
            a[6], a[7] = vel_B2
        except (TypeError, ValueError):   # (2,1) / (1,2) arrays, other containers -- or a TypeError for the caller
            a[:] = list(np.concatenate([_vec(vel_A1, 2, "vel_A1"), _vec(vel_A2, 2, "vel_A2"), _vec(vel_B1, 2, "vel_B1"),
                                        _vec(vel_B2, 2, "vel_B2")]))
        rc = self._lib.rgs_single_run(self._h, a, self._out)
        if rc:
            _capi.check(rc)
        return self._out_np.copy()

    # --- not in the reference binding (its clock and damage only reach the log file) ---
    def time(self):
        t = C.c_double(0.0)
        _capi.check(self._lib.rgs_single_time(self._h, C.byref(t)))
        return t.value

    def damage(self):
        out = np.empty(16, dtype=np.float64)
        _capi.check(self._lib.rgs_single_damage(self._h, out.ctypes.data_as(C.POINTER(C.c_double))))
        return out.reshape(4, 4)

    def __repr__(self):
        return "GameSim Object"  # reference src/RoboGamePy.cpp:20-23


class BatchedGameSim:
    """N independent games on one GPU behind the C ABI (rgs_batch_*).

    mode: "exact" (K1, fp64, event-exact vs. the reference), "fast" (K2, fp32 throughput),
    "fast64" (K2 in fp64).  Keyword parameters are those of GameSim::reset plus the constructor
    constants (arena_x, arena_y, player_radius, puck_radius, goal_height)."""

    def __init__(self, n_games, device=None, mode="exact", players_per_team=2, first_game=0, **params):
        """params: the fields of rgs_params (dt, winning_score, noise, seed, tau_puck, tau_player, player_mass,
        puck_mass, arena_x, arena_y, player_radius, puck_radius, goal_height).  noise > 0 (process noise on the
        player accelerations, reference src/GameSim.cpp:62-68) runs in the "fast" / "fast64" modes; its streams are
        keyed by (seed, first_game + g), so a batch sharded over GPUs draws what the unsharded batch draws."""
        import torch  # tensor hand-off only

        self._torch = torch
        self._lib = _capi.load()
        if device is None:
            device = torch.cuda.current_device() if torch.cuda.is_available() else 0
        self.device = torch.device("cuda", int(device)) if not isinstance(device, torch.device) else device
        self.n_games = int(n_games)
        self.n_team = int(players_per_team)
        self.n_bodies = 2 * self.n_team + 1
        self.state_size = 2 + 4 * self.n_bodies
        p = _capi.default_params()
        if mode not in _MODES:
            raise ValueError(f"mode must be one of 'exact', 'fast', 'fast64' (got {mode!r})")
        p.mode = _MODES[mode]
        p.players_per_team = self.n_team
        for k, v in params.items():
            if not hasattr(p, k) or k in ("mode", "players_per_team"):
                raise TypeError(f"unknown parameter {k!r}")
            setattr(p, k, v)
        self.params = p
        self.mode = p.mode
        self.action_dtype = torch.float32 if p.mode == _capi.MODE_FAST_F32 else torch.float64
        self._h = C.c_void_p()
        _capi.check(self._lib.rgs_batch_create(C.byref(self._h), self.device.index or 0, self.n_games, C.byref(p)))
        if first_game:
            self.set_game_offset(first_game)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.rgs_batch_destroy(h)
            self._h = None

    # ---- helpers ----
    def _stream(self):
        return C.c_void_p(self._torch.cuda.current_stream(self.device).cuda_stream)

    def _dev_tensor(self, t, dtype, name):
        torch = self._torch
        if not isinstance(t, torch.Tensor):
            t = torch.as_tensor(np.asarray(t), dtype=dtype)
        if t.device != self.device:
            t = t.to(self.device, non_blocking=True)
        if t.dtype != dtype:
            t = t.to(dtype)
        return t.contiguous()

    def set_game_offset(self, first_game):
        """Global index of game 0 of this batch (keys the per-game process-noise streams)."""
        _capi.check(self._lib.rgs_batch_set_game_offset(self._h, int(first_game)))

    def set_params(self, **params):
        for k, v in params.items():
            if not hasattr(self.params, k) or k in ("mode", "players_per_team"):
                raise TypeError(f"unknown parameter {k!r}")
            setattr(self.params, k, v)
        _capi.check(self._lib.rgs_batch_set_params(self._h, C.byref(self.params)))

    # ---- GameSim::reset for the whole batch ----
    def reset(self, x0_ball=None, state0=None):
        """x0_ball: None (zeros), [4] (shared) or [N,4]; state0: [N, 2+4E] full-state restore."""
        torch = self._torch
        x0_ptr, stride, keep = None, 0, []
        if x0_ball is not None:
            x0 = self._dev_tensor(x0_ball, torch.float64, "x0_ball")
            if x0.numel() == 4:
                stride = 0
            elif tuple(x0.shape) == (self.n_games, 4):
                stride = 4
            else:
                raise ValueError(f"x0_ball must have shape (4,) or ({self.n_games}, 4), got {tuple(x0.shape)}")
            x0_ptr = x0.data_ptr()
            keep.append(x0)
        s0_ptr = None
        if state0 is not None:
            s0 = self._dev_tensor(state0, torch.float64, "state0")
            if tuple(s0.shape) != (self.n_games, self.state_size):
                raise ValueError(f"state0 must have shape ({self.n_games}, {self.state_size})")
            s0_ptr = s0.data_ptr()
            keep.append(s0)
        _capi.check(self._lib.rgs_batch_reset(self._h, x0_ptr, stride, s0_ptr, self._stream()))
        self._keep = keep  # borrowed until the stream op completes
        return self

    # ---- n_steps x GameSim::run for the whole batch, one fused launch ----
    def _traj_tensor(self, n_steps, log_stride, traj_layout):
        """the trajectory tensor of a run with log_stride > 0 and its device pointer; selects the layout for the call"""
        if traj_layout not in ("records", "field_major"):
            raise ValueError(f"traj_layout must be 'records' or 'field_major' (got {traj_layout!r})")
        if not log_stride:
            return None, None
        torch = self._torch
        n_rec, trec = int(n_steps) // int(log_stride), 1 + self.state_size + 4 * self.n_team * self.n_team
        shape = (self.n_games, n_rec, trec) if traj_layout == "records" else (n_rec, trec, self.n_games)
        _capi.check(self._lib.rgs_batch_set_traj_layout(self._h, 0 if traj_layout == "records" else 1))
        traj = torch.empty(shape, dtype=torch.float64, device=self.device)
        return traj, traj.data_ptr()

    def run(self, actions, n_steps=None, hold=1, log_stride=0, traj_layout="records"):
        """actions: [N, n_act, 2n, 2] per-game or [n_act, 2n, 2] shared; step k uses action
        min(k // hold, n_act-1).  Returns the trajectory tensor when log_stride > 0, else None:
        traj_layout "records" -> [N, n_steps//log_stride, 1+S+(2n)^2] (what write_logs reads), "field_major" ->
        [n_steps//log_stride, 1+S+(2n)^2, N] (coalesced stores: the layout for every-step logging of large batches;
        traj.permute(2, 0, 1) is the records view)."""
        torch = self._torch
        a = self._dev_tensor(actions, self.action_dtype, "actions")
        np2 = 2 * self.n_team
        if a.dim() == 4 and tuple(a.shape[0:1] + a.shape[2:]) == (self.n_games, np2, 2):
            n_act, stride = a.shape[1], a.shape[1] * np2 * 2
        elif a.dim() == 3 and tuple(a.shape[1:]) == (np2, 2):
            n_act, stride = a.shape[0], 0
        else:
            raise ValueError(f"actions must have shape ({self.n_games}, n_act, {np2}, 2) or (n_act, {np2}, 2), "
                             f"got {tuple(a.shape)}")
        if n_steps is None:
            n_steps = n_act * hold
        traj, traj_ptr = self._traj_tensor(n_steps, log_stride, traj_layout)
        _capi.check(self._lib.rgs_batch_run(self._h, a.data_ptr(), n_act, int(hold), stride, int(n_steps),
                                            int(log_stride), traj_ptr, self._stream()))
        self._keep = [a]
        return traj

    def run_synth(self, seed, n_steps, hold=10, u_max=10.0, first_game=0, step0=0, log_stride=0, traj_layout="records"):
        """Same, with the synthetic S0 commands generated on chip (no action tensor in HBM)."""
        torch = self._torch
        traj, traj_ptr = self._traj_tensor(n_steps, log_stride, traj_layout)
        _capi.check(self._lib.rgs_batch_run_synth(self._h, int(seed), int(first_game), int(step0), int(hold),
                                                  float(u_max), int(n_steps), int(log_stride), traj_ptr, self._stream()))
        return traj

    def run_policy(self, n_steps, log_stride=0, traj_layout="records", **policy):
        """Closed loop on the on-device surrogate of the reference's baseline team (ClassicalTeam play logic and
        targets, closed-form tracking law; `cbf=1` adds the CBF.py barrier filter): the commands are computed from
        the state inside the kernel every step.  policy: kick_velocity, gain, input_limit, safety_radius,
        barrier_gain, cbf, sweeps, kind (include/rgs_capi.h rgs_policy).  FAST modes, any team size (player 1 of a team is
        the goalie, the others field players); with process noise: 2 players per team."""
        torch = self._torch
        pol = _capi.RgsPolicy()
        self._lib.rgs_default_policy(C.byref(pol))
        for k, v in policy.items():
            if k not in dict(_capi.RgsPolicy._fields_):
                raise TypeError(f"unknown policy field {k!r}")
            setattr(pol, k, v)
        traj, traj_ptr = self._traj_tensor(n_steps, log_stride, traj_layout)
        _capi.check(self._lib.rgs_batch_run_policy(self._h, C.byref(pol), int(n_steps), int(log_stride), traj_ptr,
                                                   self._stream()))
        return traj

    # ---- read-back (device tensors) ----
    def _get(self, which):
        torch = self._torch
        np2 = 2 * self.n_team
        shapes = {"state": ((self.n_games, self.state_size), torch.float64), "time": ((self.n_games,), torch.float64),
                  "damage": ((self.n_games, np2, np2), torch.float64), "events": ((self.n_games, _capi.N_EVENTS), torch.int64)}
        shape, dt = shapes[which]
        out = torch.empty(shape, dtype=dt, device=self.device)
        ptrs = {k: (out.data_ptr() if k == which else None) for k in shapes}
        _capi.check(self._lib.rgs_batch_get(self._h, ptrs["state"], ptrs["time"], ptrs["damage"], ptrs["events"],
                                            self._stream()))
        return out

    def checkpoint(self):
        """All game records of the batch as device tensors: dict(state, time, damage, events) -- restore() takes it."""
        return {k: self._get(k) for k in ("state", "time", "damage", "events")}

    def restore(self, state=None, time=None, damage=None, events=None):
        """Overwrite the game records (rgs_batch_set): resumes a checkpoint exactly, clock, damage and event records
        included (reset(state0=...) zeroes those).  Any argument may be None = left as it is."""
        torch = self._torch
        np2 = 2 * self.n_team
        want = {"state": (state, (self.n_games, self.state_size), torch.float64), "time": (time, (self.n_games,), torch.float64),
                "damage": (damage, (self.n_games, np2, np2), torch.float64),
                "events": (events, (self.n_games, _capi.N_EVENTS), torch.int64)}
        ptrs, keep = {}, []
        for k, (x, shape, dt) in want.items():
            if x is None:
                ptrs[k] = None
                continue
            t = self._dev_tensor(x, dt, k)
            if t.numel() != int(np.prod(shape)):
                raise ValueError(f"{k} must have shape {shape}, got {tuple(t.shape)}")
            ptrs[k] = t.data_ptr()
            keep.append(t)
        _capi.check(self._lib.rgs_batch_set(self._h, ptrs["state"], ptrs["time"], ptrs["damage"], ptrs["events"],
                                            self._stream()))
        self._keep = keep
        return self

    def state(self):
        return self._get("state")

    def time(self):
        return self._get("time")

    def damage(self):
        return self._get("damage")

    def events(self):
        return self._get("events")

    def undecided(self):
        s = self.state()
        w = float(self.params.winning_score)
        return (s[:, 0] < w) & (s[:, 1] < w)

    def stats(self):
        """K4: int64[96] statistics vector on the device (sum it across GPUs with one all-reduce)."""
        torch = self._torch
        out = torch.empty(_capi.N_STATS, dtype=torch.int64, device=self.device)
        _capi.check(self._lib.rgs_batch_stats(self._h, out.data_ptr(), self._stream()))
        return out

    # ---- K3: shared-state rollout sampling ----
    def reset_shared(self, state):
        """Every game of the batch starts from the ONE state record `state` [2+4E] (BASELINE configs[4])."""
        torch = self._torch
        s0 = self._dev_tensor(state, torch.float64, "state").reshape(-1)
        if s0.numel() != self.state_size:
            raise ValueError(f"state must have {self.state_size} elements, got {s0.numel()}")
        _capi.check(self._lib.rgs_batch_reset_shared(self._h, s0.data_ptr(), self._stream()))
        self._keep = [s0]
        return self

    def terminal_cost(self, x_des, weights):
        """cost[g] = sum_{p,c} weights[p,c] * (x_g[p,c] - x_des[p,c])**2 over the 2n players of every game, and
        the arg-min.  x_des, weights: [2n,4].  Returns (cost [N] float64 device tensor, best index int, best cost)."""
        torch = self._torch
        npl = 2 * self.n_team
        xd = self._dev_tensor(x_des, torch.float64, "x_des")
        w = self._dev_tensor(weights, torch.float64, "weights")
        if xd.numel() != 4 * npl or w.numel() != 4 * npl:
            raise ValueError(f"x_des and weights must have shape ({npl}, 4)")
        cost = torch.empty(self.n_games, dtype=torch.float64, device=self.device)
        best = torch.empty(2, dtype=torch.int64, device=self.device)
        _capi.check(self._lib.rgs_batch_terminal_cost(self._h, xd.data_ptr(), w.data_ptr(), cost.data_ptr(), best.data_ptr(),
                                                      self._stream()))
        self._keep = [xd, w]
        b = best.cpu()
        return cost, int(b[0]), float(b[1:2].view(torch.float64)[0])

    # ---- the reference's binary logs for a whole batch ----
    def write_logs(self, traj, path_pattern, first_index=1, layout="reference", n_records=None, append=False):
        """One binary log per game in the reference's record format (src/GameSim.cpp:161-170, include/utils.h:66-76;
        29 doubles per record at 2v2), from the trajectory tensor a run with log_stride=1 returned.

        path_pattern: file name with one "{}" for the run index first_index + g (py/MonteCarlo.py:26 writes
                      mc_run_{i}.log with i from 1).
        layout:       "reference" reproduces the reference's damage fields byte for byte (damage.row(k) logged as 4
                      contiguous doubles of the column-major matrix); "row_major" writes the true rows
                      damage(aggressor, :) that matlab/interpret_mc_results.m:37-48 assumes.
        n_records:    [N] records per game (default: all); the reference stops logging once a game is decided, so
                      pass the games' live steps of that launch to get files of the reference's length.
        """
        torch = self._torch
        layouts = {"reference": 0, "row_major": 1}
        if layout not in layouts:
            raise ValueError("layout must be 'reference' or 'row_major'")
        np2 = 2 * self.n_team
        t = traj.detach().to("cpu", torch.float64).contiguous() if isinstance(traj, torch.Tensor) else \
            torch.from_numpy(np.ascontiguousarray(traj, dtype=np.float64))
        if t.dim() != 3 or t.shape[0] != self.n_games or t.shape[2] != 1 + self.state_size + np2 * np2:
            raise ValueError(f"traj must have shape ({self.n_games}, records, {1 + self.state_size + np2 * np2}), got {tuple(t.shape)}")
        nr_ptr, nr = None, None
        if n_records is not None:
            nr = torch.as_tensor(n_records).detach().to("cpu", torch.int64).contiguous()
            if nr.numel() != self.n_games:
                raise ValueError(f"n_records must have {self.n_games} elements")
            nr_ptr = nr.data_ptr()
        _capi.check(self._lib.rgs_write_logs(t.data_ptr(), self.n_games, int(t.shape[1]), self.n_team, nr_ptr,
                                             os.fsencode(path_pattern), int(first_index), layouts[layout], 1 if append else 0))

    # ---- host-buffer entry point (end-to-end path) ----
    def run_host(self, x0_ball, actions, n_steps, hold=1, want_state=True, want_events=True, out=None):
        """x0_ball [N,4] float64 and actions [N,n_act,2n,2] (float32 in "fast" mode, else float64)
        in HOST memory (numpy or CPU torch tensors, ideally pinned).  Resets the batch, runs n_steps
        and returns (state [N,S] f64, events [N,16] i64, stats [96] i64) as numpy arrays."""
        torch = self._torch

        def host(x, dtype):
            if isinstance(x, torch.Tensor):
                if x.is_cuda:
                    raise ValueError("run_host takes host buffers; use run() for device tensors")
                x = x.numpy()
            return np.ascontiguousarray(x, dtype=dtype)

        adt = np.float32 if self.mode == _capi.MODE_FAST_F32 else np.float64
        a = host(actions, adt)
        np2 = 2 * self.n_team
        if a.ndim != 4 or (a.shape[0], a.shape[2], a.shape[3]) != (self.n_games, np2, 2):
            raise ValueError(f"actions must have shape ({self.n_games}, n_act, {np2}, 2), got {a.shape}")
        x0 = None
        if x0_ball is not None:
            x0 = host(x0_ball, np.float64)
            if x0.shape != (self.n_games, 4):
                raise ValueError(f"x0_ball must have shape ({self.n_games}, 4)")
        out = out or {}
        st = out.get("state") if want_state else None
        ev = out.get("events") if want_events else None
        if want_state and st is None:
            st = np.empty((self.n_games, self.state_size), dtype=np.float64)
        if want_events and ev is None:
            ev = np.empty((self.n_games, _capi.N_EVENTS), dtype=np.int64)
        stats = np.zeros(_capi.N_STATS, dtype=np.int64)

        def ptr(x):
            if x is None:
                return None
            return x.data_ptr() if isinstance(x, torch.Tensor) else x.ctypes.data

        _capi.check(self._lib.rgs_batch_run_host(self._h, ptr(x0), ptr(a), a.shape[1], int(hold), int(n_steps), ptr(st),
                                                 ptr(ev), stats.ctypes.data, self._stream()))
        return st, ev, stats


def synth_inputs(n, n_act, seed=0, first_game=0, players_per_team=2, u_max=10.0, dtype=None, device=None,
                 want_x0=True, want_actions=True):
    """K5: the synthetic S0 inputs (x0_ball [n,4] f64, actions [n,n_act,2n,2]) generated on the device."""
    import torch

    lib = _capi.load()
    if device is None:
        device = torch.cuda.current_device()
    dev = torch.device("cuda", int(device)) if not isinstance(device, torch.device) else device
    dtype = dtype or torch.float64
    x0 = torch.empty((n, 4), dtype=torch.float64, device=dev) if want_x0 else None
    acts = torch.empty((n, n_act, 2 * players_per_team, 2), dtype=dtype, device=dev) if want_actions else None
    _capi.check(lib.rgs_synth_inputs(dev.index or 0, int(seed), int(first_game), int(n), int(players_per_team),
                                     int(n_act), float(u_max), x0.data_ptr() if want_x0 else None,
                                     acts.data_ptr() if want_actions else None, 32 if dtype == torch.float32 else 64,
                                     C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
    return x0, acts
