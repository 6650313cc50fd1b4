/* oracle/gamesim_oracle.h -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Plain-C, CPU restatement of the reference GameSim step
 * (/root/reference/src/GameSim.cpp:21-623, include/GameSim.h:6-14, include/Collision.h:5-30,
 * include/SimState.h:7-59), generalised from 2v2 to n-v-n and instrumented with event counters.
 * It is pinned bit-for-bit (all 22 state doubles, clock and damage, every step) against the
 * unmodified reference compiled in place (oracle/_ref, see oracle/Makefile) by
 * tests/test_oracle.py and by the committed fixtures in tests/golden/.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product path (robo_game_sim_b200/) never does.
 */
#ifndef RGS_GAMESIM_ORACLE_H
#define RGS_GAMESIM_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_GRID 50          /* COLLISION_GRID_POINTS, reference include/GameSim.h:6 */
#define ORC_COUNTER_LIM 50   /* COLLISION_COUNTER_LIM, reference include/GameSim.h:7 */
#define ORC_MAX_TEAM 7
#define ORC_MAX_BODIES (2 * ORC_MAX_TEAM + 1)
#define ORC_N_EVENTS 16

/* per-game event record (int64[ORC_N_EVENTS]); the same layout the CUDA path reports */
enum {
    ORC_EV_GOALS_A = 0, ORC_EV_GOALS_B = 1,
    ORC_EV_WALL_UP = 2, ORC_EV_WALL_DOWN = 3, ORC_EV_WALL_RIGHT = 4, ORC_EV_WALL_LEFT = 5,
    ORC_EV_PLAYER_PLAYER = 6, ORC_EV_PLAYER_PUCK = 7,
    ORC_EV_DISCARDED = 8,      /* puck events dropped after a goal (GameSim.cpp:230-231) */
    ORC_EV_LOOP_ITERS = 9,     /* sum over steps of the do/while counter (GameSim.cpp:140-150) */
    ORC_EV_LIVE_STEPS = 10,    /* steps executed while undecided */
    ORC_EV_DECIDED_STEP = 11,  /* live-step count (>= 1) at which a team reached winning_score; 0 = undecided */
    ORC_EV_COUNTER_LIMIT = 12, /* steps that left the loop with events still queued */
    ORC_EV_BASE_CLAMP = 13,    /* base_idx > 50 (undefined behaviour in the reference) clamped */
    ORC_EV_HASH = 14,          /* FNV-1a style hash of (live step, kind, i, j[, goal]) of every event */
    ORC_EV_OVERLAP_FIXES = 15  /* correctOverlap corrections applied (GameSim.cpp:581-623) */
};

typedef struct orc_params {
    double dt;            /* 0.05 */
    int32_t winning_score;
    int32_t players_per_team; /* 2 = reference */
    double tau_puck, tau_player, player_mass, puck_mass; /* reset() arguments, GameSim.cpp:24-25 */
    double arena_x, arena_y, player_radius, puck_radius, goal_height; /* ctor constants, GameSim.cpp:5-9 */
} orc_params;

/* Fill p with the reference defaults (dt 0.05, winning score 3, 2v2, tau 0.1/0.5, mass 1.0/0.5,
 * arena 10x5, radii 0.2/0.175, goal height 1.0). */
void orc_default_params(orc_params *p);

/* GameSim::reset (GameSim.cpp:21-55): state vector [TAS,TBS,PK(4),A1..An(4 each),B1..Bn] with the
 * puck at x0_ball and players at x = -+X/4, y_k = Y/2 - (k+1/2) Y/n (reduces to GameSim.cpp:35-38
 * for n = 2).  out_state has 2 + 4*(2n+1) doubles. */
void orc_reset_state(const orc_params *p, const double *x0_ball4, double *out_state);

/* Advance n_games independent games by n_steps calls of GameSim::run (GameSim.cpp:118-171).
 *   state   [n_games, 2+4E]  in/out, reference order
 *   t       [n_games]        in/out simulator clock t_
 *   damage  [n_games, (2n)^2] in/out, damage(aggressor, victim) row-major; may be NULL
 *   actions [*, n_act, 2n, 2]; game g reads block g*act_game_stride (in doubles; 0 = shared by
 *           all games); live step k of this call uses action index min(k / hold, n_act-1)
 *   events  [n_games, ORC_N_EVENTS] in/out accumulators; may be NULL
 *   traj    [n_games, n_steps, 2+4E] state after every step; may be NULL
 *   n_threads  worker threads over games (the restatement has no shared scratch)
 * returns 0, or -1 on bad arguments. */
int orc_rollout(const orc_params *p, int64_t n_games, int64_t n_steps, int64_t hold, double *state, double *t,
                double *damage, const double *actions, int64_t n_act, int64_t act_game_stride, int64_t *events,
                double *traj, int32_t n_threads);

/* 29-double log record of one game (GameSim.cpp:161-170, include/utils.h:66-76), 2v2 only:
 * [t, TAS, TBS, puck xy, A1 xy, A2 xy, B1 xy, B2 xy, 16 damage doubles].  The reference writes
 * damage.row(k) as 4 CONTIGUOUS doubles starting at &damage(k,0) of a column-major matrix;
 * bug_compatible != 0 reproduces that, 0 writes true rows. */
void orc_log_record(const double *state22, double t, const double *damage16_rowmajor, int bug_compatible,
                    double *out29);

#ifdef __cplusplus
}
#endif
#endif /* RGS_GAMESIM_ORACLE_H */
