// tests/emu/emu_rollout.cpp -- TEST INFRASTRUCTURE ONLY (never on the product path).
//
// Compiles the product's kernel-body headers (robo_game_sim_b200/csrc/rgs_exact.cuh, rgs_lane.cuh)
// with g++ against the fiber warp emulator (warp_emu.h) so their control flow and arithmetic can be
// checked against the oracle in this GPU-less container.  The resulting tests/emu/librgs_emu.so is
// loaded by tests/ only; robo_game_sim_b200 never loads it and has no CPU path.
#include "warp_emu.h"
//
#include "../../robo_game_sim_b200/csrc/rgs_exact.cuh"
#include "../../robo_game_sim_b200/csrc/rgs_fast.cuh"
#include "../../robo_game_sim_b200/csrc/rgs_lane.cuh"
#include "../../robo_game_sim_b200/csrc/rgs_host_params.h"

#include <memory>
#include <vector>

namespace {

template <int NT>
void run_exact(const rgs_params &p, const rgs::RunArgs &A) {
    auto W = std::make_unique<rgs::WarpSlice<double, NT>>();
    const rgs::DevParams<double> P = rgs::make_dev_params<double>(p);
    emu::run_warp([&](int lane) { rgs::rollout_exact_warp<NT>(P, A, W.get(), lane, 0, 1); });
}

template <typename T, int NT, bool kNoise = false>
void run_lane(const rgs_params &p, const rgs::RunArgs &A) {
    using S = rgs::LaneScratch<T, NT>;
    auto tab = std::make_unique<rgs::LaneTab<T>>();
    const rgs::FastParams<T> P = rgs::make_fast_params<T>(p);
    for (int ty = 0; ty < 2; ++ty)
        for (int i = 0; i < rgs::kCols; ++i) { tab->e[ty][i].g = P.tab_g[ty][i]; tab->e[ty][i].k = P.tab_k[ty][i]; }
    T fw[S::NF];
    int iw[S::NI];
    std::vector<T> scr(S::SCR_WORDS_NOISE);   // the lane's global scratch area (tracks, event counters, jam snapshot, noise)
    S s{fw, iw, 1, scr.data()};
    if (A.synth == 2) {   // closed loop on the on-device surrogate policy (with noise: 2v2 only, like the library)
        if constexpr (NT == 2 || !kNoise)
            for (int64_t g = 0; g < A.n_games; ++g) rgs::rollout_lane_game<T, NT, true, kNoise>(P, A, tab.get(), s, g);
    } else for (int64_t g = 0; g < A.n_games; ++g) rgs::rollout_lane_game<T, NT, false, kNoise>(P, A, tab.get(), s, g);
}

template <int NT>
void dispatch(const rgs_params &p, const rgs::RunArgs &A) {
    // the FAST modes run the lane body, as librgs_b200.so does
    if (p.mode == RGS_MODE_EXACT_F64) run_exact<NT>(p, A);
    else if (p.mode == RGS_MODE_FAST_F32) { if (p.noise != 0.0) run_lane<float, NT, true>(p, A); else run_lane<float, NT>(p, A); }
    else { if (p.noise != 0.0) run_lane<double, NT, true>(p, A); else run_lane<double, NT>(p, A); }
}

}  // namespace

static rgs_policy g_policy;   // used by emu_rollout calls with synth == 2
extern "C" void emu_set_policy(const rgs_policy *pol) { g_policy = *pol; }
// the policy alone: commands [2 * 2n] of one game from its state record [2 + 4E] (fp64 arithmetic)
extern "C" int emu_policy_commands(const rgs_params *p, const rgs_policy *pol, const double *state, double *u) {
    const rgs::DevParams<double> d = rgs::make_dev_params<double>(*p);
    const rgs::PolicyArgs a{pol->kick_velocity, pol->gain, pol->input_limit, pol->safety_radius, pol->barrier_gain, pol->cbf, pol->sweeps, pol->kind};
    // state record: [TAS, TBS, puck, players...]; body index: players 0..2n-1, puck 2n
    const int pk = 2 * p->players_per_team;
    auto X = [&](int e, int c) { return state[2 + 4 * (e == pk ? 0 : e + 1) + c]; };
    switch (p->players_per_team) {
    case 1: rgs::surrogate_policy<double, 1>(d, a, X, u); break;
    case 2: rgs::surrogate_policy<double, 2>(d, a, X, u); break;
    case 3: rgs::surrogate_policy<double, 3>(d, a, X, u); break;
    case 4: rgs::surrogate_policy<double, 4>(d, a, X, u); break;
    case 7: rgs::surrogate_policy<double, 7>(d, a, X, u); break;
    default: return -1;
    }
    return 0;
}

extern "C" int emu_rollout(const rgs_params *p, int64_t n_games, double *state, double *clock, double *damage,
                           long long *events, const void *actions, int64_t n_act, int64_t hold, int64_t act_game_stride,
                           int64_t n_steps, int64_t log_stride, double *traj, int synth, uint64_t seed,
                           int64_t first_game, int64_t step0, double u_max) {
    rgs::RunArgs A{};
    A.n_games = n_games;
    A.state = state;
    A.clock = clock;
    A.damage = damage;
    A.events = events;
    A.actions = actions;
    A.n_act = n_act;
    A.hold = hold;
    A.act_game_stride = act_game_stride;
    A.n_steps = n_steps;
    A.log_stride = log_stride;
    A.traj_recs = log_stride > 0 ? n_steps / log_stride : 0;
    A.traj_field_stride = 1;   // records layout
    A.traj = traj;
    A.synth = synth;
    A.seed = seed;
    A.first_game = first_game;
    A.noise_game0 = first_game;
    A.step0 = step0;
    A.u_max = u_max;
    if (synth == 2) {
        if (p->mode == RGS_MODE_EXACT_F64) return -2;
        A.pol = rgs::PolicyArgs{g_policy.kick_velocity, g_policy.gain, g_policy.input_limit, g_policy.safety_radius,
                                g_policy.barrier_gain, g_policy.cbf, g_policy.sweeps, g_policy.kind};
    }
    switch (p->players_per_team) {
    case 1: dispatch<1>(*p, A); break;
    case 2: dispatch<2>(*p, A); break;
    case 3: dispatch<3>(*p, A); break;
    case 4: dispatch<4>(*p, A); break;
    case 7: dispatch<7>(*p, A); break;
    default: return -1;
    }
    return 0;
}

// host evaluation of the synthetic S0 generators (same Philox code as the device)
extern "C" void emu_synth(uint64_t seed, int64_t first_game, int64_t n, int n_team, int64_t n_act, double u_max,
                          double *x0_ball, double *actions) {
    for (int64_t g = 0; g < n; ++g) {
        const int64_t game = first_game + g;
        if (x0_ball) {
            uint32_t o[4];
            rgs::philox4x32_10((uint32_t)game, (uint32_t)((uint64_t)game >> 32), 0xFFFFFFFFu, 0xFFFFFFFFu, (uint32_t)seed,
                               (uint32_t)(seed >> 32), o);
            x0_ball[4 * g + 0] = rgs::sym_uniform(o[0], 4.0);
            x0_ball[4 * g + 1] = rgs::sym_uniform(o[1], 2.0);
            x0_ball[4 * g + 2] = rgs::sym_uniform(o[2], 3.0);
            x0_ball[4 * g + 3] = rgs::sym_uniform(o[3], 3.0);
        }
        if (actions)
            for (int64_t a = 0; a < n_act; ++a)
                for (int c = 0; c < 4 * n_team; ++c)
                    actions[(g * n_act + a) * 4 * n_team + c] = rgs::synth_action(seed, game, a, c, u_max);
    }
}
