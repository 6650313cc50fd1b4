"""Host-side (numpy) evaluation of the synthetic "S0" input generator that the device evaluates in
K5 (rgs_synth_inputs, csrc/rgs_aux_kernels.cu) and on chip in rgs_batch_run_synth: Philox4x32-10
keyed by (seed, global game index, action index).  Input generation only -- the GameSim step itself
has no host implementation in this package."""
import numpy as np


def _philox(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(x, dtype=np.uint64) for x in np.broadcast_arrays(c0, c1, c2, c3))
    k0 = np.uint64(k0)
    k1 = np.uint64(k1)
    M0, M1, mask = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        n0 = (p1 >> np.uint64(32)) ^ c1 ^ k0
        n1 = p1 & mask
        n2 = (p0 >> np.uint64(32)) ^ c3 ^ k1
        n3 = p0 & mask
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + np.uint64(0x9E3779B9)) & mask
        k1 = (k1 + np.uint64(0xBB67AE85)) & mask
    return c0, c1, c2, c3


def _sym(bits, a):
    return (2.0 * (bits.astype(np.float64) * 2.3283064365386963e-10) - 1.0) * a


def synth_inputs(seed, first_game, n, n_team, n_act, u_max=10.0):
    """x0_ball [n,4] and actions [n,n_act,2*n_team,2] (float64), identical to rgs_synth_inputs."""
    games = np.arange(first_game, first_game + n, dtype=np.uint64)
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    o = _philox(games & np.uint64(0xFFFFFFFF), games >> np.uint64(32), np.uint64(0xFFFFFFFF), np.uint64(0xFFFFFFFF), k0, k1)
    x0 = np.stack([_sym(o[0], 4.0), _sym(o[1], 2.0), _sym(o[2], 3.0), _sym(o[3], 3.0)], axis=1)
    g = games[:, None, None]
    a = np.arange(n_act, dtype=np.uint64)[None, :, None]
    q = np.arange(n_team, dtype=np.uint64)[None, None, :]
    o = _philox(g & np.uint64(0xFFFFFFFF), g >> np.uint64(32), a & np.uint64(0xFFFFFFFF), ((a >> np.uint64(32)) << np.uint64(8)) | q, k0, k1)
    acts = np.stack([_sym(w, u_max) for w in o], axis=-1)  # [n, n_act, n_team, 4]
    return np.ascontiguousarray(x0), np.ascontiguousarray(acts.reshape(n, n_act, 2 * n_team, 2))


