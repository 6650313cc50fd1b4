"""ctypes binding of librgs_b200.so (include/rgs_capi.h) -- the only way the Python host reaches
the CUDA kernels.  There is no CPU implementation behind it: if the library has not been built
(`python -c "import __graft_entry__ as g; g.build()"` or `make -C robo_game_sim_b200/csrc`) the
import of any compute entry point fails loudly."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RGS_B200_LIB: another build of the same library (kernel A/B measurements); never a different implementation
LIB_PATH = os.environ.get("RGS_B200_LIB") or os.path.join(_HERE, "librgs_b200.so")

RGS_OK, RGS_EINVAL, RGS_ECUDA, RGS_ENOMEM, RGS_EUNSUPPORTED = 0, -1, -2, -3, -4
MODE_EXACT_F64, MODE_FAST_F32, MODE_FAST_F64 = 64, 32, 65
N_EVENTS, N_STATS, MAX_TEAM = 16, 96, 7
NOALLOC_MAX_GAMES = 4096   # RGS_NOALLOC_MAX_GAMES: rgs_rollout_noalloc's static scratch array

EV_NAMES = ["goals_a", "goals_b", "wall_up", "wall_down", "wall_right", "wall_left", "player_player", "player_puck",
            "discarded", "loop_iters", "live_steps", "decided_step", "counter_limit", "base_clamp", "hash",
            "overlap_fixes"]
ST_GAMES, ST_WINS_A, ST_WINS_B, ST_UNDECIDED, ST_LEAD_A, ST_LEAD_B, ST_TIED = range(7)
ST_EVENTS0, ST_SCORE_HIST = 8, 32
ST_DAMAGE, ST_DAMAGE_TEAM, ST_OVERLAP_FIXES, DAMAGE_FIXED_SCALE = 64, 80, 84, float(1 << 24)


class RgsParams(C.Structure):
    _fields_ = [("dt", C.c_double), ("winning_score", C.c_int32), ("players_per_team", C.c_int32),
                ("noise", C.c_double), ("tau_puck", C.c_double), ("tau_player", C.c_double),
                ("player_mass", C.c_double), ("puck_mass", C.c_double), ("arena_x", C.c_double),
                ("arena_y", C.c_double), ("player_radius", C.c_double), ("puck_radius", C.c_double),
                ("goal_height", C.c_double), ("mode", C.c_int32), ("seed", C.c_int32)]


class RgsPolicy(C.Structure):
    """rgs_policy of include/rgs_capi.h (defaults: rgs_default_policy)."""
    _fields_ = [("kick_velocity", C.c_double), ("gain", C.c_double), ("input_limit", C.c_double),
                ("safety_radius", C.c_double), ("barrier_gain", C.c_double), ("cbf", C.c_int32), ("sweeps", C.c_int32),
                ("kind", C.c_int32), ("reserved", C.c_int32)]


# every symbol include/rgs_capi.h declares: (restype, argtypes)
_vp, _i64, _i32, _u64, _dbl = C.c_void_p, C.c_int64, C.c_int32, C.c_uint64, C.c_double
_PP = C.POINTER(RgsParams)
SIGNATURES = {
    "rgs_default_params": (None, [_PP]),
    "rgs_last_error": (C.c_char_p, []),
    "rgs_launch_count": (_i64, []),
    "rgs_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "rgs_batch_create": (C.c_int, [C.POINTER(_vp), C.c_int, _i64, _PP]),
    "rgs_batch_destroy": (C.c_int, [_vp]),
    "rgs_batch_set_params": (C.c_int, [_vp, _PP]),
    "rgs_batch_set_game_offset": (C.c_int, [_vp, _i64]),
    "rgs_batch_set_traj_layout": (C.c_int, [_vp, _i32]),
    "rgs_batch_reset": (C.c_int, [_vp, _vp, _i64, _vp, _vp]),
    "rgs_batch_run": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _i64, _i64, _vp, _vp]),
    "rgs_batch_run_synth": (C.c_int, [_vp, _u64, _i64, _i64, _i64, _dbl, _i64, _i64, _vp, _vp]),
    "rgs_default_policy": (None, [C.POINTER(RgsPolicy)]),
    "rgs_batch_run_policy": (C.c_int, [_vp, C.POINTER(RgsPolicy), _i64, _i64, _vp, _vp]),
    "rgs_batch_get": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "rgs_batch_set": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "rgs_batch_stats": (C.c_int, [_vp, _vp, _vp]),
    "rgs_batch_reset_shared": (C.c_int, [_vp, _vp, _vp]),
    "rgs_batch_terminal_cost": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "rgs_synth_inputs": (C.c_int, [C.c_int, _u64, _i64, _i64, _i32, _i64, _dbl, _vp, _vp, _i32, _vp]),
    "rgs_measure_peak": (C.c_int, [C.c_int, _i32, C.POINTER(_dbl)]),
    "rgs_write_logs": (C.c_int, [_vp, _i64, _i64, _i32, _vp, C.c_char_p, _i64, _i32, _i32]),
    "rgs_batch_run_host": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp]),
    "rgs_rollout_noalloc": (C.c_int, [C.c_int, _PP, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _vp]),
    "rgs_single_create": (C.c_int, [C.POINTER(_vp), C.c_int]),
    "rgs_single_destroy": (C.c_int, [_vp]),
    "rgs_single_reset": (C.c_int, [_vp, _dbl, _i32, C.POINTER(_dbl), _dbl, _i32, C.c_char_p, _i32, _dbl, _dbl, _dbl,
                                   _dbl, C.POINTER(_dbl)]),
    "rgs_single_undecided": (C.c_int, [_vp, C.POINTER(_i32)]),
    "rgs_single_run": (C.c_int, [_vp, C.POINTER(_dbl), C.POINTER(_dbl)]),
    "rgs_single_time": (C.c_int, [_vp, C.POINTER(_dbl)]),
    "rgs_single_damage": (C.c_int, [_vp, C.POINTER(_dbl)]),
}

_lib = None


def load():
    """Load librgs_b200.so; raises RuntimeError (never falls back) when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is not built: run `make -C robo_game_sim_b200/csrc` (nvcc, sm_100a). "
                "robo_game_sim_b200 has no CPU implementation of the GameSim step.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error():
    msg = load().rgs_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc):
    """Map a negative RGS_E* return code to the Python exception the drop-in surface promises."""
    if rc == RGS_OK:
        return
    msg = last_error()
    if rc == RGS_EINVAL:
        raise ValueError(msg)
    if rc == RGS_ENOMEM:
        raise MemoryError(msg)
    if rc == RGS_EUNSUPPORTED:
        raise NotImplementedError(msg)
    raise RuntimeError(msg)


def default_params():
    p = RgsParams()
    load().rgs_default_params(C.byref(p))
    return p


def launch_count():
    return int(load().rgs_launch_count())
