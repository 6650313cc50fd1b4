"""CPU suite, part 3: the C-ABI library loads and exports every symbol include/rgs_capi.h declares
(no compute calls without a GPU), fails loudly without a device, and the host-side logic
(argument checking of the drop-in class, sharding, configuration file, statistics naming, the
world_size-2 reduction over gloo)."""
import ctypes as C
import os
import re
import subprocess
import sys
import types

import numpy as np
import pytest

import rgs_testlib as tl

ROOT = tl.ROOT


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "rgs_capi.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rgs_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def capi():
    subprocess.run(["make", "-C", os.path.join(ROOT, "robo_game_sim_b200", "csrc"), "-j4"], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
    from robo_game_sim_b200 import _capi

    _capi.load()
    return _capi


def test_library_exports_every_declared_symbol(capi):
    declared = _header_symbols()
    assert len(declared) >= 22
    lib = C.CDLL(capi.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/rgs_capi.h but not exported"
    assert set(capi.SIGNATURES) == set(declared)      # the Python binding covers the whole header


def test_params_struct_matches_header(capi):
    p = capi.default_params()
    assert (p.dt, p.winning_score, p.players_per_team, p.noise) == (0.05, 3, 2, 0.0)
    assert (p.tau_puck, p.tau_player, p.player_mass, p.puck_mass) == (0.1, 0.5, 1.0, 0.5)
    assert (p.arena_x, p.arena_y, p.player_radius, p.puck_radius, p.goal_height) == (10.0, 5.0, 0.2, 0.175, 1.0)
    assert p.mode == capi.MODE_EXACT_F64
    assert C.sizeof(capi.RgsParams) == 8 + 4 + 4 + 8 * 10 + 4 + 4


def test_argument_errors_do_not_need_a_gpu(capi):
    lib = capi.load()
    h = C.c_void_p()
    p = capi.default_params()
    p.players_per_team = 9
    assert lib.rgs_batch_create(C.byref(h), 0, 4, C.byref(p)) == capi.RGS_EINVAL
    assert "players_per_team" in capi.last_error()
    p = capi.default_params()
    p.noise = 0.1                       # process noise: FAST modes only (statistical parity), never the exact mode
    assert lib.rgs_batch_create(C.byref(h), 0, 4, C.byref(p)) == capi.RGS_EUNSUPPORTED
    assert "FAST" in capi.last_error()
    p.noise = -1.0
    p.mode = capi.MODE_FAST_F32
    assert lib.rgs_batch_create(C.byref(h), 0, 4, C.byref(p)) == capi.RGS_EINVAL
    p = capi.default_params()
    p.mode = 7
    assert lib.rgs_batch_create(C.byref(h), 0, 4, C.byref(p)) == capi.RGS_EINVAL
    assert lib.rgs_batch_run(None, None, 1, 1, 0, 1, 0, None, None) == capi.RGS_EINVAL
    assert lib.rgs_batch_destroy(None) == capi.RGS_OK
    with pytest.raises(ValueError):
        capi.check(capi.RGS_EINVAL)


def test_no_cpu_fallback(capi):
    """Without a CUDA device the product must fail loudly, never compute on the host."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from robo_game_py import GameSim

    with pytest.raises(RuntimeError, match="no CPU path"):
        GameSim()
    h = C.c_void_p()
    p = capi.default_params()
    assert capi.load().rgs_batch_create(C.byref(h), 0, 4, C.byref(p)) == capi.RGS_ECUDA


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "robo_game_sim_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in text and "libref_gamesim" not in text and "librgs_emu" not in text, f
    for f in os.listdir(os.path.join(ROOT, "robo_game_py")):
        if f.endswith(".py"):
            text = open(os.path.join(ROOT, "robo_game_py", f)).read()
            assert "oracle" not in text


def test_gamesim_argument_conversion():
    from robo_game_sim_b200.gamesim import _vec

    assert np.array_equal(_vec([1, 2], 2, "v"), [1.0, 2.0])
    assert np.array_equal(_vec(np.zeros((2, 1)), 2, "v"), [0.0, 0.0])
    with pytest.raises(TypeError):
        _vec([1, 2, 3], 2, "v")
    with pytest.raises(TypeError):
        _vec("ab", 2, "v")


def test_shard_range_covers_all_runs():
    from robo_game_sim_b200.montecarlo import shard_range

    for n in (0, 1, 7, 8, 1000003):
        for w in (1, 2, 4, 8):
            parts = [shard_range(n, r, w) for r in range(w)]
            assert sum(c for _, c in parts) == n
            pos = 0
            for first, count in parts:
                if count:
                    assert first == pos
                pos += count


def test_configuration_txt_matches_reference_format(tmp_path):
    """13 lines, str() of each value (reference py/MonteCarlo.py:14-21)."""
    from robo_game_sim_b200.montecarlo import write_configuration

    mc = types.SimpleNamespace(num_runs=4, T=10.0, dt=0.05, winning_score=4, x0_ball=np.array([-1, 0., 0., 0.]),
                               noise_stdev=0.0, tau_puck=0.1, tau_player=0.5, player_mass=1.0, puck_mass=0.5)
    write_configuration(mc, str(tmp_path / "mc"))
    lines = open(tmp_path / "mc" / "configuration.txt").read().splitlines()
    assert lines == ["4", "10.0", "0.05", "4", "-1.0", "0.0", "0.0", "0.0", "0.0", "0.1", "0.5", "1.0", "0.5"]


def test_stats_naming():
    from robo_game_sim_b200 import _capi
    from robo_game_sim_b200.montecarlo import stats_to_dict

    s = np.zeros(96, dtype=np.int64)
    s[0], s[1], s[2], s[3] = 10, 3, 2, 5
    s[_capi.ST_EVENTS0 + 10] = 2000
    s[_capi.ST_SCORE_HIST + 5 * 1 + 2] = 7
    d = stats_to_dict(s)
    assert d["games"] == 10 and d["wins_a"] == 3 and d["live_steps"] == 2000 and d["score_hist"][1, 2] == 7


def test_numpy_s0_generator_matches_device_code_on_host():
    """robo_game_sim_b200/synth_np.py == the Philox code of csrc/rgs_device.cuh (compiled for the host)."""
    lib = tl.emu_lib()
    n, n_act, nt = 13, 7, 2
    x0 = np.zeros((n, 4))
    acts = np.zeros((n, n_act, 2 * nt, 2))
    dp = C.POINTER(C.c_double)
    lib.emu_synth(2**40 + 17, 2**33 + 5, n, nt, n_act, 10.0, x0.ctypes.data_as(dp), acts.ctypes.data_as(dp))
    gx0, gacts = tl.synth_inputs(2**40 + 17, 2**33 + 5, n, nt, n_act)
    assert np.array_equal(x0, gx0) and np.array_equal(acts, gacts)
    assert np.abs(gx0[:, 0]).max() <= 4 and np.abs(gx0[:, 1]).max() <= 2 and np.abs(gacts).max() <= 10
    # different games / seeds give different streams
    assert not np.array_equal(tl.synth_inputs(1, 0, 4, 2, 3)[1], tl.synth_inputs(2, 0, 4, 2, 3)[1])


_GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, {root!r})
import torch, torch.distributed as dist
from robo_game_sim_b200.montecarlo import reduce_stats, shard_range
rank, world = int(sys.argv[1]), int(sys.argv[2])
os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=sys.argv[3])
dist.init_process_group("gloo", rank=rank, world_size=world)
first, count = shard_range(1001, rank, world)
stats = torch.zeros(96, dtype=torch.int64)
stats[0] = count                       # games
stats[1] = sum(1 for g in range(first, first + count) if g % 3 == 0)
out = reduce_stats(stats)
assert int(out[0]) == 1001 and int(out[1]) == 334, out[:2]
dist.destroy_process_group()
print("ok")
"""


def test_stats_reduction_world_size_2_gloo(tmp_path):
    """The multi-rank path: each rank owns a slice of runs, one all-reduce(sum) of int64[96]."""
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER.format(root=ROOT))
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), str(r), "2", port], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0 and "ok" in o, o


def test_sampler_targets_follow_the_reference_play_logic():
    """Host side of the rollout sampler: DMPCTeam.execute's play decision and DMPCPlayer.get_action's
    x_des (reference py/src/DMPCTeam.py:78-88, py/src/DMPCPlayer.py:27-55), checked on hand-worked states."""
    from robo_game_sim_b200 import sampler as sp

    s = np.zeros(22)
    s[2:6] = [1.0, 0.5, 0.0, 0.0]
    assert sp.dmpc_play(s, field=-1) == sp.OFFENSE
    s[2] = 8.0                                   # field * px = 8 > 10 - 2.5
    assert sp.dmpc_play(s, field=1) == sp.DEFENSE
    s[2], s[4] = 6.0, 2.0                        # 6 > 10 - 5 and vx 2 > 1
    assert sp.dmpc_play(s, field=1) == sp.DEFENSE
    s[4] = 0.5
    assert sp.dmpc_play(s, field=1) == sp.OFFENSE
    s[2:4] = [1.0, 0.0]
    t = sp.dmpc_target(s, sp.OFFENSE, sp.ATTACKER, field=-1, goal_pos=[5.0, 0.0], home_pos=[-5.0, 0.0])
    assert np.allclose(t, [1.0, 0.0, 4.0, 0.0])                       # hit towards the goal at v_hit
    t = sp.dmpc_target(s, sp.DEFENSE, sp.ATTACKER, field=-1, goal_pos=[5.0, 0.0], home_pos=[-5.0, 0.0])
    assert np.allclose(t, [1.0, 0.0, 8.0, 0.0])                       # 2 * v_hit
    t = sp.dmpc_target(s, sp.DEFENSE, sp.DEFENDER, field=-1, goal_pos=[5.0, 0.0], home_pos=[-5.0, 0.0])
    assert np.allclose(t, [-2.0, 0.0, 0.0, 0.0])                      # half way between home and puck
    t = sp.dmpc_target(s, sp.OFFENSE, sp.DEFENDER, field=-1, goal_pos=[5.0, 0.0], home_pos=[-5.0, 0.0])
    assert np.allclose(t, [1.0, 0.0, 4.0, 0.0])                       # field * px < 0: hit
    a = sp.candidate_actions(16, 5, players=(0, 1), seed=3)
    assert a.shape == (16, 5, 4, 2) and a.dtype == np.float32
    assert np.all(a[0] == 0) and np.all(a[:, :, 2:] == 0) and np.abs(a[1:, :, :2]).max() <= 10.0 and a[1:, :, :2].std() > 3
    assert np.array_equal(a, sp.candidate_actions(16, 5, players=(0, 1), seed=3))


def test_pybind11_translation_unit_builds_and_binds():
    """INTEGRATION.md section B is a real file (robo_game_sim_b200/pybind/RoboGamePyB200.cpp): it compiles against
    pybind11 + include/rgs_capi.h, links librgs_b200.so, and exposes the reference's surface
    (src/RoboGamePy.cpp:15-24): GameSim {__init__, reset, undecided, run, __repr__}, reset with the 5 trailing
    defaults.  Without a GPU the constructor fails loudly (no CPU path)."""
    import inspect

    m = tl.pybind_module()
    assert m.__name__ == "robo_game_py"
    for name in ("reset", "undecided", "run", "__repr__", "__init__"):
        assert hasattr(m.GameSim, name)
    doc = m.GameSim.reset.__doc__
    for frag in ("seed", "= 0", "tau_puck", "= 0.1", "tau_player", "= 0.5", "player_mass", "= 1.0", "puck_mass"):
        assert frag in doc
    assert inspect.isclass(m.GameSim)
    import torch

    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU path"):
            m.GameSim()
