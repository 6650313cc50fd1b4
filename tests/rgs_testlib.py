"""Test-side helpers: ctypes loaders for the CHECKERS (oracle port, compiled reference, warp
emulator build of the kernel bodies) and the shared synthetic "S0" input generator.

Nothing here is imported by the product package (robo_game_sim_b200 / robo_game_py).
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
import sys  # noqa: E402

if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
ORACLE_DIR = os.path.join(ROOT, "oracle")
EMU_DIR = os.path.join(ROOT, "tests", "emu")

N_EVENTS = 16
EV = dict(GOALS_A=0, GOALS_B=1, WALL_UP=2, WALL_DOWN=3, WALL_RIGHT=4, WALL_LEFT=5, PLAYER_PLAYER=6, PLAYER_PUCK=7,
          DISCARDED=8, LOOP_ITERS=9, LIVE_STEPS=10, DECIDED_STEP=11, COUNTER_LIMIT=12, BASE_CLAMP=13, HASH=14,
          OVERLAP_FIXES=15)

MODE_EXACT_F64, MODE_FAST_F32, MODE_FAST_F64 = 64, 32, 65


class OrcParams(C.Structure):
    _fields_ = [("dt", C.c_double), ("winning_score", C.c_int32), ("players_per_team", C.c_int32),
                ("tau_puck", C.c_double), ("tau_player", C.c_double), ("player_mass", C.c_double),
                ("puck_mass", C.c_double), ("arena_x", C.c_double), ("arena_y", C.c_double),
                ("player_radius", C.c_double), ("puck_radius", C.c_double), ("goal_height", C.c_double)]


class RgsParams(C.Structure):
    _fields_ = [("dt", C.c_double), ("winning_score", C.c_int32), ("players_per_team", C.c_int32),
                ("noise", C.c_double), ("tau_puck", C.c_double), ("tau_player", C.c_double),
                ("player_mass", C.c_double), ("puck_mass", C.c_double), ("arena_x", C.c_double),
                ("arena_y", C.c_double), ("player_radius", C.c_double), ("puck_radius", C.c_double),
                ("goal_height", C.c_double), ("mode", C.c_int32), ("seed", C.c_int32)]


class RgsPolicy(C.Structure):
    _fields_ = [("kick_velocity", C.c_double), ("gain", C.c_double), ("input_limit", C.c_double),
                ("safety_radius", C.c_double), ("barrier_gain", C.c_double), ("cbf", C.c_int32), ("sweeps", C.c_int32),
                ("kind", C.c_int32), ("reserved", C.c_int32)]


def default_policy(**kw):
    """rgs_default_policy of include/rgs_capi.h."""
    d = dict(kick_velocity=4.0, gain=4.0, input_limit=10.0, safety_radius=0.4, barrier_gain=30.0, cbf=0, sweeps=8, kind=0, reserved=0)
    d.update(kw)
    return d


def _ptr(a, ct=C.c_double):
    return None if a is None else a.ctypes.data_as(C.POINTER(ct))


_cache = {}


def _make(target_dir, *args):
    subprocess.run(["make", "-C", target_dir, *args], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)


def oracle_lib():
    """Plain-C restatement (oracle/gamesim_oracle.c), built on demand with gcc."""
    if "orc" not in _cache:
        path = os.path.join(ORACLE_DIR, "liboracle_gamesim.so")
        _make(ORACLE_DIR, "port")
        lib = C.CDLL(path)
        lib.orc_default_params.argtypes = [C.POINTER(OrcParams)]
        lib.orc_reset_state.argtypes = [C.POINTER(OrcParams), C.POINTER(C.c_double), C.POINTER(C.c_double)]
        lib.orc_rollout.restype = C.c_int
        lib.orc_rollout.argtypes = [C.POINTER(OrcParams), C.c_int64, C.c_int64, C.c_int64, C.POINTER(C.c_double),
                                    C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int64,
                                    C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_double), C.c_int32]
        lib.orc_log_record.argtypes = [C.POINTER(C.c_double), C.c_double, C.POINTER(C.c_double), C.c_int,
                                       C.POINTER(C.c_double)]
        _cache["orc"] = lib
    return _cache["orc"]


def ref_lib():
    """UNMODIFIED reference GameSim compiled in place (oracle/_ref); None when neither the
    reference tree nor a prebuilt library is present."""
    if "ref" not in _cache:
        path = os.path.join(ORACLE_DIR, "_ref", "libref_gamesim.so")
        if os.path.exists("/root/reference/src/GameSim.cpp"):
            _make(ORACLE_DIR, "ref")
        if not os.path.exists(path):
            _cache["ref"] = None
            return None
        lib = C.CDLL(path)
        lib.ref_create.restype = C.c_void_p
        lib.ref_destroy.argtypes = [C.c_void_p]
        lib.ref_reset.argtypes = [C.c_void_p, C.c_double, C.c_int, C.POINTER(C.c_double), C.c_double, C.c_int,
                                  C.c_char_p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
                                  C.POINTER(C.c_double)]
        lib.ref_undecided.argtypes = [C.c_void_p]
        lib.ref_undecided.restype = C.c_int
        lib.ref_run.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        lib.ref_time.argtypes = [C.c_void_p]
        lib.ref_time.restype = C.c_double
        lib.ref_damage.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        lib.ref_flush_log.argtypes = [C.c_void_p]
        lib.ref_rollout.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int64, C.POINTER(C.c_double),
                                    C.POINTER(C.c_double)]
        lib.ref_rollout_batch.argtypes = [C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_int,
                                          C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int64, C.c_double,
                                          C.c_double, C.c_double, C.c_double, C.POINTER(C.c_double)]
        _cache["ref"] = lib
    return _cache["ref"]


def emu_lib():
    """g++ build of the kernel-body headers on the fiber warp emulator (tests/emu)."""
    if "emu" not in _cache:
        _make(EMU_DIR)
        lib = C.CDLL(os.path.join(EMU_DIR, "librgs_emu.so"))
        lib.emu_rollout.restype = C.c_int
        lib.emu_rollout.argtypes = [C.POINTER(RgsParams), C.c_int64, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                    C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_void_p, C.c_int64, C.c_int64,
                                    C.c_int64, C.c_int64, C.c_int64, C.POINTER(C.c_double), C.c_int, C.c_uint64,
                                    C.c_int64, C.c_int64, C.c_double]
        lib.emu_synth.argtypes = [C.c_uint64, C.c_int64, C.c_int64, C.c_int, C.c_int64, C.c_double,
                                  C.POINTER(C.c_double), C.POINTER(C.c_double)]
        _cache["emu"] = lib
    return _cache["emu"]


def pybind_module():
    """The pybind11 module of INTEGRATION.md section B (robo_game_sim_b200/pybind/RoboGamePyB200.cpp), built on demand
    against the in-tree librgs_b200.so and loaded WITHOUT registering it as `robo_game_py` in sys.modules (the
    repository's own ctypes package has that name)."""
    if "pyb" not in _cache:
        import glob
        import importlib.util

        d = os.path.join(ROOT, "robo_game_sim_b200", "pybind")
        _make(d)
        path = glob.glob(os.path.join(d, "robo_game_py*.so"))[0]
        spec = importlib.util.spec_from_file_location("robo_game_py", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _cache["pyb"] = mod
    return _cache["pyb"]


def default_params(n_team=2, winning_score=3, mode=MODE_EXACT_F64, **kw):
    d = dict(dt=0.05, winning_score=winning_score, players_per_team=n_team, noise=0.0, tau_puck=0.1, tau_player=0.5,
             player_mass=1.0, puck_mass=0.5, arena_x=10.0, arena_y=5.0, player_radius=0.2, puck_radius=0.175,
             goal_height=1.0, mode=mode, seed=0)
    d.update(kw)
    return d


def rgs_params(d):
    return RgsParams(**d)


def orc_params(d):
    return OrcParams(**{k: d[k] for k, _ in OrcParams._fields_})


from robo_game_sim_b200.synth_np import synth_inputs  # noqa: E402,F401  (numpy S0 generator)


def oracle_reset(pd, x0_ball):
    """[n, 2+4E] initial states for pucks x0_ball [n,4] (GameSim::reset)."""
    lib = oracle_lib()
    p = orc_params(pd)
    n = x0_ball.shape[0]
    E = 2 * pd["players_per_team"] + 1
    out = np.zeros((n, 2 + 4 * E))
    x0 = np.ascontiguousarray(x0_ball, dtype=np.float64)
    for g in range(n):
        lib.orc_reset_state(C.byref(p), _ptr(x0[g]), _ptr(out[g]))
    return out


def oracle_rollout(pd, state, actions, n_steps, hold=1, t=None, damage=None, events=None, traj=False, threads=8):
    """Oracle port over a batch.  actions [n, n_act, 2n, 2] (or [n_act, 2n, 2] shared).  Returns dict."""
    lib = oracle_lib()
    p = orc_params(pd)
    n = state.shape[0]
    nt = pd["players_per_team"]
    state = np.array(state, dtype=np.float64, order="C")
    t = np.zeros(n) if t is None else np.array(t, dtype=np.float64)
    damage = np.zeros((n, 4 * nt * nt)) if damage is None else np.array(damage, dtype=np.float64)
    events = np.zeros((n, N_EVENTS), dtype=np.int64) if events is None else np.array(events, dtype=np.int64)
    actions = np.ascontiguousarray(actions, dtype=np.float64)
    shared = actions.ndim == 3
    n_act = actions.shape[0] if shared else actions.shape[1]
    stride = 0 if shared else n_act * 4 * nt
    tr = np.zeros((n, n_steps, state.shape[1])) if traj else None
    rc = lib.orc_rollout(C.byref(p), n, n_steps, hold, _ptr(state), _ptr(t), _ptr(damage), _ptr(actions), n_act, stride,
                         _ptr(events, C.c_int64), _ptr(tr), threads)
    assert rc == 0
    return dict(state=state, t=t, damage=damage, events=events, traj=tr)


def oracle_rollout_reltime(pd, state, actions, n_steps, hold=1, threads=8):
    """The oracle stepped with its clock reset to 0 before every step (state, damage and event
    records carried over).  The reference computes base_idx = floor((t_col - t_) / dt_col_) + 1 from
    ABSOLUTE times (src/GameSim.cpp:237), so for resting contacts (time of impact == a grid time) the
    rounding noise of t_ decides between two columns; the throughput kernels use step-relative time.
    This variant is the reference arithmetic without that clock noise: it is what the K2
    formulation is compared with wherever jammed-contact event COUNTS matter."""
    st = np.array(state, dtype=np.float64, order="C")
    n = st.shape[0]
    ev = dm = None
    for k in range(n_steps):
        a = min(k // hold, actions.shape[1] - 1)
        r = oracle_rollout(pd, st, actions[:, a:a + 1], 1, 1, t=np.zeros(n), damage=dm, events=ev, threads=threads)
        st, ev, dm = r["state"], r["events"], r["damage"]
    return dict(state=st, damage=dm, events=ev)


def emu_policy_commands(pd, pol, state):
    """Commands [n, 4, 2] of the on-device surrogate policy for state records [n, 22] (emulator build, fp64)."""
    lib = emu_lib()
    p, q = rgs_params(pd), RgsPolicy(**pol)
    state = np.ascontiguousarray(state, dtype=np.float64)
    npl = 2 * pd["players_per_team"]
    u = np.zeros((state.shape[0], 2 * npl))
    for g in range(state.shape[0]):
        assert lib.emu_policy_commands(C.byref(p), C.byref(q), _ptr(state[g]), _ptr(u[g])) == 0
    return u.reshape(-1, npl, 2)


def emu_rollout(pd, state, actions, n_steps, hold=1, t=None, damage=None, events=None, log_stride=0, synth=None,
                policy=None, first_game=0):
    """The product's kernel body stepped on the CPU warp emulator (one warp, games in sequence).
    policy: a default_policy() dict = closed loop on the on-device surrogate policy (actions ignored)."""
    lib = emu_lib()
    p = rgs_params(pd)
    n = state.shape[0]
    nt = pd["players_per_team"]
    state = np.array(state, dtype=np.float64, order="C")
    t = np.zeros(n) if t is None else np.array(t, dtype=np.float64)
    damage = np.zeros((n, 4 * nt * nt)) if damage is None else np.array(damage, dtype=np.float64)
    events = np.zeros((n, N_EVENTS), dtype=np.int64) if events is None else np.array(events, dtype=np.int64)
    traj = None
    if log_stride:
        traj = np.zeros((n, n_steps // log_stride, 1 + state.shape[1] + 4 * nt * nt))
    if policy is not None:
        q = RgsPolicy(**policy)
        lib.emu_set_policy(C.byref(q))
        rc = lib.emu_rollout(C.byref(p), n, _ptr(state), _ptr(t), _ptr(damage), _ptr(events, C.c_int64), None, 1, 1,
                             0, n_steps, log_stride, _ptr(traj), 2, 0, 0, 0, 0.0)
    elif synth is not None:
        seed, first_game, step0, u_max = synth
        rc = lib.emu_rollout(C.byref(p), n, _ptr(state), _ptr(t), _ptr(damage), _ptr(events, C.c_int64), None, 1, hold,
                             0, n_steps, log_stride, _ptr(traj), 1, seed, first_game, step0, u_max)
    else:
        dt = np.float32 if pd["mode"] == MODE_FAST_F32 else np.float64
        actions = np.ascontiguousarray(actions, dtype=dt)
        shared = actions.ndim == 3
        n_act = actions.shape[0] if shared else actions.shape[1]
        stride = 0 if shared else n_act * 4 * nt
        rc = lib.emu_rollout(C.byref(p), n, _ptr(state), _ptr(t), _ptr(damage), _ptr(events, C.c_int64),
                             actions.ctypes.data_as(C.c_void_p), n_act, hold, stride, n_steps, log_stride, _ptr(traj),
                             0, 0, first_game, 0, 0.0)
    assert rc == 0
    return dict(state=state, t=t, damage=damage, events=events, traj=traj)


# ---- process-noise statistics against the fixture tests/golden/noise_stats.npz (reference run with noise > 0) ----
def noise_fixture():
    return np.load(os.path.join(ROOT, "tests", "golden", "noise_stats.npz"))


def check_noise_moments(mean, var, n, ref_mean, ref_var, n_ref, cols, var_tol, what):
    """Sample means within 4.5 combined standard errors (+1e-6 for the affine-vs-RK4 rounding), sample variances within
    var_tol of each other (the ratio of two sample variances of ~Gaussian data has sd sqrt(2/n + 2/n_ref))."""
    m, v, rm, rv = mean[..., cols], var[..., cols], ref_mean[..., cols], ref_var[..., cols]
    z = np.abs(m - rm) / (np.sqrt(v / n + rv / n_ref) + 1e-6)
    assert z.max() < 4.5, (what, "mean", float(z.max()))
    live = rv > 1e-12
    ratio = v[live] / rv[live]
    assert ratio.min() > 1 - var_tol and ratio.max() < 1 + var_tol, (what, "variance ratio", float(ratio.min()), float(ratio.max()))


def analytic_noise_var(noise, steps, dt=0.05, tau=0.5):
    """(var v, var p, cov) of the accumulated process noise of a player axis after `steps` game-steps: the per-sub-step
    increment of reference src/GameSim.cpp:62-88 (four independent stage draws) pushed through S' = M S M^T + Q."""
    h = dt / 50
    z = -h / tau
    c = [1 + z + z * z / 2 + z ** 3 / 4, 2 + z + z * z / 2, 2 + z, 1.0]
    d = [1 + z / 2 + z * z / 4, 1 + z / 2, 1.0]
    qvv = (h / 6) ** 2 * sum(x * x for x in c)
    qpp = (h * h / 6) ** 2 * sum(x * x for x in d)
    qpv = (h / 6) * (h * h / 6) * sum(a * b for a, b in zip(c, d))
    q = 1 + z / 2 + z * z / 6 + z ** 3 / 24
    R, a = 1 + z * q, h * q
    svv = spv = spp = 0.0
    for _ in range(50 * steps):
        svv, spv, spp = R * R * svv + qvv, R * (a * svv + spv) + qpv, a * a * svv + 2 * a * spv + spp + qpp
    return noise ** 2 * svv, noise ** 2 * spp, noise ** 2 * spv
