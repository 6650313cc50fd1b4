// oracle/ref_wrapper.cpp -- TEST INFRASTRUCTURE ONLY (never on the product path).
//
// extern "C" handle API around the UNMODIFIED reference class `GameSim`
// (/root/reference/include/GameSim.h:16-27, /root/reference/src/GameSim.cpp), which is compiled
// in place from /root/reference by oracle/Makefile against oracle/eigen_shim.  This plays the
// role of the reference's pybind11 module (reference src/RoboGamePy.cpp:11-25), which cannot be
// built here because pybind11/eigen.h needs real Eigen.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load the resulting oracle/_ref/libref_gamesim.so.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <map>
#include <random>
#include <string>
#include <vector>

// The reference exposes no getter for its clock or damage matrix (they only reach the binary
// log).  The wrapper reads them directly; every standard header the reference pulls in has
// been included above so the macro only affects the reference's own class definitions.
#define private public
#include "GameSim.h"
#undef private

namespace {
inline Eigen::Vector2d v2(const double *p) { return Eigen::Vector2d(p[0], p[1]); }
inline void copy_state(const Eigen::Matrix<double, SimState::SIZE, 1> &s, double *out) {
    for (int i = 0; i < SimState::SIZE; ++i) out[i] = s(i);
}
}  // namespace

extern "C" {

void *ref_create(void) { return new GameSim(); }
void ref_destroy(void *h) { delete static_cast<GameSim *>(h); }

// GameSim::reset (reference src/GameSim.cpp:21-55); out_state22 may be NULL.
void ref_reset(void *h, double dt, int winning_score, const double *x0_ball4, double noise, int log,
               const char *logname, int seed, double tau_puck, double tau_player, double player_mass,
               double puck_mass, double *out_state22) {
    GameSim *g = static_cast<GameSim *>(h);
    Eigen::Vector4d x0(x0_ball4[0], x0_ball4[1], x0_ball4[2], x0_ball4[3]);
    Eigen::Matrix<double, SimState::SIZE, 1> s =
        g->reset(dt, winning_score, x0, noise, log != 0, std::string(logname ? logname : ""), seed, tau_puck,
                 tau_player, player_mass, puck_mass);
    if (out_state22) copy_state(s, out_state22);
}

int ref_undecided(void *h) { return static_cast<GameSim *>(h)->undecided() ? 1 : 0; }

// GameSim::run (reference src/GameSim.cpp:118-124); actions8 = [A1x,A1y,A2x,A2y,B1x,B1y,B2x,B2y].
void ref_run(void *h, const double *actions8, double *out_state22) {
    GameSim *g = static_cast<GameSim *>(h);
    Eigen::Matrix<double, SimState::SIZE, 1> s =
        g->run(v2(actions8), v2(actions8 + 2), v2(actions8 + 4), v2(actions8 + 6));
    if (out_state22) copy_state(s, out_state22);
}

double ref_time(void *h) { return static_cast<GameSim *>(h)->t_; }

// damage(aggressor, victim), row-major into out16 (reference include/SimState.h:33-38).
void ref_damage(void *h, double *out16) {
    GameSim *g = static_cast<GameSim *>(h);
    for (int a = 0; a < 4; ++a)
        for (int v = 0; v < 4; ++v) out16[a * 4 + v] = g->state_.damage(a, v);
}

// Close the log so a reader sees every record (reference include/utils.h:61-64 only closes in
// the destructor or on the next reset).
void ref_flush_log(void *h) {
    GameSim *g = static_cast<GameSim *>(h);
    if (g->logger_.file_.is_open()) g->logger_.file_.flush();
}

// Convenience loop: n_steps calls of run() on one game, recording every returned state.
//   actions    : [n_steps, 8] doubles
//   out_states : [n_steps, 22] doubles or NULL; out_final: [22] or NULL
void ref_rollout(void *h, const double *actions, int64_t n_steps, double *out_states, double *out_final) {
    GameSim *g = static_cast<GameSim *>(h);
    Eigen::Matrix<double, SimState::SIZE, 1> s;
    for (int64_t k = 0; k < n_steps; ++k) {
        const double *a = actions + 8 * k;
        s = g->run(v2(a), v2(a + 2), v2(a + 4), v2(a + 6));
        if (out_states) copy_state(s, out_states + 22 * k);
    }
    if (out_final && n_steps > 0) copy_state(s, out_final);
}

// Batched convenience for the CPU baseline: games are stepped one after another on the
// calling thread (GameSim keeps function-local static scratch, src/GameSim.cpp:72,100, so one
// PROCESS per core is the only safe parallelism).
//   x0_ball : [n_games,4]; actions: [n_games, n_act, 8] where action index = step / hold
//   out_final : [n_games,22]
int64_t ref_rollout_batch_live(int64_t n_games, int64_t n_steps, int64_t hold, double dt, int winning_score,
                               const double *x0_ball, const double *actions, int64_t n_act, double tau_puck,
                               double tau_player, double player_mass, double puck_mass, double *out_final);
void ref_rollout_batch(int64_t n_games, int64_t n_steps, int64_t hold, double dt, int winning_score,
                       const double *x0_ball, const double *actions, int64_t n_act, double tau_puck,
                       double tau_player, double player_mass, double puck_mass, double *out_final) {
    ref_rollout_batch_live(n_games, n_steps, hold, dt, winning_score, x0_ball, actions, n_act, tau_puck, tau_player,
                           player_mass, puck_mass, out_final);
}

// Same, and returns the number of LIVE steps: run() calls made while undecided() (GameSim::run is a no-op once a
// game is decided, src/GameSim.cpp:121-123), the basis the GPU arm's game-steps/s is counted on.
int64_t ref_rollout_batch_live(int64_t n_games, int64_t n_steps, int64_t hold, double dt, int winning_score,
                               const double *x0_ball, const double *actions, int64_t n_act, double tau_puck,
                               double tau_player, double player_mass, double puck_mass, double *out_final) {
    GameSim g;
    int64_t live = 0;
    for (int64_t n = 0; n < n_games; ++n) {
        const double *x0 = x0_ball + 4 * n;
        g.reset(dt, winning_score, Eigen::Vector4d(x0[0], x0[1], x0[2], x0[3]), 0.0, false, std::string(""),
                0, tau_puck, tau_player, player_mass, puck_mass);
        Eigen::Matrix<double, SimState::SIZE, 1> s = g.state_.vector();
        for (int64_t k = 0; k < n_steps; ++k) {
            const double *a = actions + 8 * (n * n_act + k / hold);
            live += g.undecided() ? 1 : 0;
            s = g.run(v2(a), v2(a + 2), v2(a + 4), v2(a + 6));
        }
        if (out_final) copy_state(s, out_final + 22 * n);
    }
    return live;
}

}  // extern "C"
