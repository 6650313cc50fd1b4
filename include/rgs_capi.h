/* include/rgs_capi.h -- C ABI of librgs_b200.so, the B200 (sm_100a) batched GameSim physics core.
 *
 * This is the drop-in boundary for ONE path of goromal/robo-game-sim: the GameSim step
 * (reference include/GameSim.h:16-27, src/GameSim.cpp:21-623) that the reference exposes to
 * Python through pybind11 (reference src/RoboGamePy.cpp:11-25).  Plain pointers and sizes only;
 * no torch / pybind / Eigen types cross this boundary.  Every function returns 0 on success or a
 * negative RGS_E* code; rgs_last_error() gives a thread-local message.  No exception crosses.
 *
 * Pointer naming: d_* = CUDA device pointer on the batch's device, h_* = host pointer.
 * All device work is enqueued on `stream` (a cudaStream_t passed as void*; NULL = the legacy
 * default stream) and is asynchronous unless stated otherwise.
 *
 * Layouts (game-major records: a warp owns a game, so lanes read consecutive words of ONE record):
 *   state  record  [2 + 4E] doubles, reference order [TAS, TBS, PK(px,py,vx,vy), A1.., B1..]
 *                  (reference include/SimState.h:14-23), E = 2*players_per_team + 1
 *   action record  [2n][2] commanded velocities in team order A1..An, B1..Bn
 *                  (the four Vector2d arguments of GameSim::run, reference src/GameSim.cpp:118-119)
 *   events record  int64[RGS_N_EVENTS], see enum below
 *   traj   record  [1 + (2+4E) + (2n)^2] doubles = [t, state, damage(aggressor,victim) row-major]
 */
#ifndef RGS_CAPI_H
#define RGS_CAPI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define RGS_GRID_POINTS 50 /* COLLISION_GRID_POINTS, reference include/GameSim.h:6 */
#define RGS_COUNTER_LIM 50 /* COLLISION_COUNTER_LIM, reference include/GameSim.h:7 */
#define RGS_MAX_TEAM 7
#define RGS_N_EVENTS 16
#define RGS_N_STATS 96

enum { RGS_OK = 0, RGS_EINVAL = -1, RGS_ECUDA = -2, RGS_ENOMEM = -3, RGS_EUNSUPPORTED = -4 };

/* precision / formulation of a batch */
enum {
    RGS_MODE_EXACT_F64 = 64, /* K1: reference operator order, fp64, no FMA: event-exact parity mode */
    RGS_MODE_FAST_F32 = 32,  /* K2: exact affine sub-step map, fp32, FMA: throughput mode */
    RGS_MODE_FAST_F64 = 65   /* K2 instantiated in fp64 (validation of the K2 formulation) */
};

/* events record fields (identical to the oracle's, oracle/gamesim_oracle.h) */
enum {
    RGS_EV_GOALS_A = 0, RGS_EV_GOALS_B = 1,
    RGS_EV_WALL_UP = 2, RGS_EV_WALL_DOWN = 3, RGS_EV_WALL_RIGHT = 4, RGS_EV_WALL_LEFT = 5,
    RGS_EV_PLAYER_PLAYER = 6, RGS_EV_PLAYER_PUCK = 7,
    RGS_EV_DISCARDED = 8, RGS_EV_LOOP_ITERS = 9, RGS_EV_LIVE_STEPS = 10, RGS_EV_DECIDED_STEP = 11,
    RGS_EV_COUNTER_LIMIT = 12, RGS_EV_BASE_CLAMP = 13, RGS_EV_HASH = 14, RGS_EV_OVERLAP_FIXES = 15
};

/* stats vector fields, int64[RGS_N_STATS] (K4; summed over games, all-reduced over GPUs by the caller) */
enum {
    RGS_ST_GAMES = 0, RGS_ST_WINS_A = 1, RGS_ST_WINS_B = 2, RGS_ST_UNDECIDED = 3,
    RGS_ST_LEAD_A = 4, RGS_ST_LEAD_B = 5, RGS_ST_TIED = 6, /* by score at the end, decided or not */
    RGS_ST_EVENTS0 = 8,      /* + RGS_EV_*: sums of the first 14 event fields (not the hash) */
    RGS_ST_SCORE_HIST = 32,  /* + 5*min(TAS,4) + min(TBS,4): 25-bin joint score histogram */
    /* damage (reference src/GameSim.cpp:277-282, SimState.h:33-38; read back per run by
     * matlab/interpret_mc_results.m:37-48), summed over the games in fixed point: units of 1/RGS_DAMAGE_FIXED_SCALE */
    RGS_ST_DAMAGE = 64,      /* + 4*aggressor + victim (A1,A2,B1,B2 = 0..3): 2v2 batches only, else 0 */
    RGS_ST_DAMAGE_TEAM = 80, /* + 2*(aggressor in team B) + (victim in team B): any team size */
    RGS_ST_OVERLAP_FIXES = 84 /* sum of RGS_EV_OVERLAP_FIXES */
};
#define RGS_DAMAGE_FIXED_SCALE 16777216 /* 2^24 */

/* Constructor constants (reference src/GameSim.cpp:5-9) and reset() arguments (:21-25). */
typedef struct rgs_params {
    double dt;                /* 0.05 */
    int32_t winning_score;    /* reset() argument */
    int32_t players_per_team; /* 2 = reference; up to RGS_MAX_TEAM */
    double noise;             /* process-noise stdev w_stdev_ (reset() argument, src/GameSim.cpp:30,66-67); FAST modes
                               * only (statistical, not bitwise: libstdc++'s normal stream is implementation-defined);
                               * must be 0 in RGS_MODE_EXACT_F64 */
    double tau_puck, tau_player, player_mass, puck_mass;
    double arena_x, arena_y, player_radius, puck_radius, goal_height; /* 10, 5, 0.2, 0.175, 1.0 */
    int32_t mode;             /* RGS_MODE_* */
    int32_t seed;             /* reset()'s seed (src/GameSim.cpp:41): keys the noise streams together with the GLOBAL game index */
} rgs_params;

typedef struct rgs_batch rgs_batch; /* opaque: N independent games resident on one GPU */

/* Reference defaults: GameSim::GameSim() + the default arguments of GameSim::reset. */
void rgs_default_params(rgs_params *p);

const char *rgs_last_error(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
int64_t rgs_launch_count(void);
int rgs_device_count(int *out);

/* replaces: GameSim::GameSim() (reference src/GameSim.cpp:3-17) for n_games games at once. */
int rgs_batch_create(rgs_batch **out, int device, int64_t n_games, const rgs_params *p);
int rgs_batch_destroy(rgs_batch *b);
int rgs_batch_set_params(rgs_batch *b, const rgs_params *p);

/* Global index of game 0 of this batch (default 0).  Every per-game random stream (process noise) is keyed by
 * (params.seed, game_offset + g), so a Monte Carlo batch sharded over several GPUs draws the same numbers as the
 * same runs on one GPU (py/MonteCarlo.py:28-30 seeds run i with i). */
int rgs_batch_set_game_offset(rgs_batch *b, int64_t first_game); /* n_games, team size and mode are fixed */

/* replaces: GameSim::reset (reference src/GameSim.cpp:21-55) -- K5.
 *   d_x0_ball [n_games,4] doubles (x0_stride doubles between games; 0 = one puck state shared by
 *             all games) or NULL -> zeros; players start at x=-+X/4, y_k = Y/2-(k+1/2)Y/n.
 *   d_state0  [n_games, 2+4E] doubles or NULL; when given it overrides the whole state
 *             (checkpoint restore; the reference has no equivalent).
 * Clock, damage and event records are zeroed. */
int rgs_batch_reset(rgs_batch *b, const double *d_x0_ball, int64_t x0_stride, const double *d_state0, void *stream);

/* replaces: n_steps consecutive calls of GameSim::run (reference src/GameSim.cpp:118-171) on every
 * game -- K1 / K2, ONE fused launch, state resident on chip across all n_steps.
 *   d_actions  action records of the batch's compute type (double for *_F64 modes, float for
 *              RGS_MODE_FAST_F32): game g, step k reads record
 *              d_actions + g*act_game_stride + min(k / hold, n_act-1) * 4n     (strides in elements;
 *              act_game_stride 0 = one action sequence shared by all games)
 *   d_traj     NULL, or [n_games, n_steps/log_stride] traj records: the state after every
 *              log_stride-th step (steps log_stride-1, 2*log_stride-1, ...)
 * Games whose score reached winning_score are frozen, as GameSim::run is (:121-123). */
int rgs_batch_run(rgs_batch *b, const void *d_actions, int64_t n_act, int64_t hold, int64_t act_game_stride,
                  int64_t n_steps, int64_t log_stride, double *d_traj, void *stream);

/* Layout of the d_traj tensor the run entry points fill (no reference counterpart: the reference logs one game to one
 * file, src/GameSim.cpp:161-170).  0 = records (default): [n_games, n_rec, TREC], a game's records contiguous -- the
 * layout rgs_write_logs reads.  1 = field-major (SoA): [n_rec, TREC, n_games], the same field of consecutive games
 * contiguous, so the 32 games of a warp store whole 32-byte sectors (K2: every-step logging costs 4x fewer memory
 * transactions; profiles/r02_ab_variants.md section 7).  Applies to the following rgs_batch_run* calls. */
int rgs_batch_set_traj_layout(rgs_batch *b, int32_t layout);

/* Same, with the synthetic S0 action generator evaluated on chip (no action tensor in HBM):
 * player commands ~ U(-u_max,u_max)^2 redrawn every `hold` steps from Philox4x32-10 keyed by
 * (seed, first_game + g, step0/hold + k/hold); identical to rgs_synth_actions' output. */
int rgs_batch_run_synth(rgs_batch *b, uint64_t seed, int64_t first_game, int64_t step0, int64_t hold, double u_max,
                        int64_t n_steps, int64_t log_stride, double *d_traj, void *stream);

/* Closed-loop rollout on the ON-DEVICE SURROGATE POLICY: every game-step the commands of all
 * players are computed from the state inside the kernel, so the Monte Carlo loop of py/MonteCarlo.py:35-46
 * (team_A_run_method, team_B_run_method, optional cbf_run_method, sim.run) never leaves the GPU.
 * Play selection and target states are the reference's baseline team (py/src/ClassicalTeam.py:32-75,
 * py/src/ClassicalPlayer.py:190-244); the pydrake minimum-time trajectory optimiser behind them is replaced by
 * the tracking law u = clamp(v_f + gain (p_f - p), +-input_limit).  cbf != 0 filters the commands through the
 * reference's control-barrier half-spaces (py/src/CBF.py:36-76, alpha(h) = barrier_gain h^3) with `sweeps`
 * Hildreth sweeps instead of the pydrake QP.  See robo_game_sim_b200/csrc/rgs_policy.cuh.
 * FAST modes, players_per_team == 2 (the reference's team size); RGS_EUNSUPPORTED otherwise. */
typedef struct rgs_policy {
    double kick_velocity;  /* 4     ClassicalTeam.py:12 */
    double gain;           /* 4     tracking gain [1/s] of the surrogate law */
    double input_limit;    /* 10    mc_example_1.py:24 */
    double safety_radius;  /* 0.4 = 2 * player_radius, mc_example_1.py:35 */
    double barrier_gain;   /* 30    mc_example_1.py:36 */
    int32_t cbf;           /* 0 = nominal commands, 1 = barrier filter on (mc_example_1.py:45) */
    int32_t sweeps;        /* Hildreth sweeps of the filter, e.g. 8 */
    int32_t kind;          /* 0 = ClassicalTeam play logic and targets; 1 = DMPCTeam / DMPCPlayer play logic and
                              targets x_des (py/src/DMPCTeam.py:78-93, py/src/DMPCPlayer.py:27-57; A1 / B1 attacker) */
    int32_t reserved;
} rgs_policy;
void rgs_default_policy(rgs_policy *pol);
int rgs_batch_run_policy(rgs_batch *b, const rgs_policy *pol, int64_t n_steps, int64_t log_stride, double *d_traj,
                         void *stream);

/* K3, shared-state rollout sampling (BASELINE configs[4]: 4096 candidate command sequences x 50 steps from one
 * state per controller tick).  It takes over the ROLE of the reference's per-tick trajectory optimisation
 * (DMPC.compute_control, reference py/src/DMPC.py:16-54: direct collocation over an internal linear model,
 * SNOPT), not its algorithm: candidates are rolled out through the real GameSim step instead.
 *   rgs_batch_reset_shared: every game of the batch starts from the ONE state record d_state [2+4E]
 *                           (clock, damage and event records zeroed);
 *   rgs_batch_run:          candidate g follows its own command records (act_game_stride != 0);
 *   rgs_batch_terminal_cost: d_cost[g] = sum_{p,c} w[p][c] (x_g[p][c] - x_des[p][c])^2 over the 2n players
 *                           (d_x_des, d_weights: [2n,4] doubles; the reference's terminal weight is
 *                           Omega_N_max = diag(10,10,20,20), py/src/DMPCTeam.py:56-59; weight 0 leaves a
 *                           player out), d_best[0] = arg-min (lowest index on ties), d_best[1] = the
 *                           minimum cost as double bits. */
int rgs_batch_reset_shared(rgs_batch *b, const double *d_state, void *stream);
int rgs_batch_terminal_cost(rgs_batch *b, const double *d_x_des, const double *d_weights, double *d_cost, int64_t *d_best,
                            void *stream);

/* State / clock / damage / events of all games into caller-provided DEVICE buffers (any may be NULL). */
int rgs_batch_get(rgs_batch *b, double *d_state, double *d_t, double *d_damage, int64_t *d_events, void *stream);

/* The inverse of rgs_batch_get: overwrite the records of all games from caller buffers (device or host memory; any may
 * be NULL = left as it is).  Together they checkpoint and resume a batch exactly -- clock, damage and event records
 * included, which rgs_batch_reset(d_state0) zeroes (the reference has no checkpoint; its clock and damage only reach
 * the log file, src/GameSim.cpp:161-170). */
int rgs_batch_set(rgs_batch *b, const double *d_state, const double *d_t, const double *d_damage, const int64_t *d_events,
                  void *stream);

/* K4: reduce scores + event records of the whole batch into int64[RGS_N_STATS] on the device
 * (overwritten).  Multi-GPU callers sum these vectors with one NCCL all-reduce. */
int rgs_batch_stats(rgs_batch *b, int64_t *d_stats, void *stream);

/* K5 (synthetic inputs "S0", shared by bench, tests and the CPU baseline):
 *   d_x0_ball [n,4] doubles: px~U(-4,4), py~U(-2,2), vx,vy~U(-3,3)
 *   d_actions [n, n_act, 2*players_per_team, 2] of `dtype` (32 = float, 64 = double) ~ U(-u_max,u_max)
 * keyed by (seed, first_game + g, action index); either pointer may be NULL. */
int rgs_synth_inputs(int device, uint64_t seed, int64_t first_game, int64_t n, int32_t players_per_team,
                     int64_t n_act, double u_max, double *d_x0_ball, void *d_actions, int32_t dtype, void *stream);

/* Roofline denominator measured on the spot: dependent-FMA-chain micro-kernel on every SM, fp32
 * (fp64 = 0) or fp64 (fp64 = 1); result in FMA TFLOP/s (2 flop per FMA).  Synchronous. */
int rgs_measure_peak(int device, int32_t fp64, double *out_tflops);

/* Batched log writer: one binary log file per game in the reference's record format (replaces Logger::log /
 * logVectors, reference include/utils.h:66-76, as driven by src/GameSim.cpp:161-170 with log = true; read back by
 * matlab/interpret_mc_results.m:29-50 and matlab/visualize_mc_game.m).  Host-side I/O only, no device work.
 *   h_traj       [n_games, n_rec, 1 + (2+4E) + (2n)^2] doubles in HOST memory: the trajectory tensor of
 *                rgs_batch_run / _run_synth / _run_policy with log_stride = 1, copied to the host
 *   h_n_records  [n_games] records to write per game, or NULL = n_rec (the reference stops logging once a game is
 *                decided, src/GameSim.cpp:121-123: pass the LIVE_STEPS of the launch)
 *   path_pattern file name with one "{}" that is replaced by first_index + g (py/MonteCarlo.py:26: mc_run_{i}.log)
 *   layout       0 = reference: record = t, TAS, TBS, puck x y, players x y, then the damage fields exactly as the
 *                reference writes them (damage.row(k) logged as 2n CONTIGUOUS doubles from &damage(k,0) of a
 *                column-major matrix, utils.h:75) -> byte-identical files; 1 = row_major: true rows damage(k, :)
 *   append       0 truncates the files, 1 appends (the priming step and the main run are two launches) */
int rgs_write_logs(const double *h_traj, int64_t n_games, int64_t n_rec, int32_t players_per_team,
                   const int64_t *h_n_records, const char *path_pattern, int64_t first_index, int32_t layout,
                   int32_t append);

/* Host-buffer entry point (the call MonteCarlo-style callers make; bench.py's e2e): copies
 * h_x0_ball [n_games,4] doubles and h_actions [n_games,n_act,4n] (compute type) to the device in
 * chunks overlapped with compute, runs n_steps, and returns final states [n_games,2+4E] doubles,
 * events [n_games,RGS_N_EVENTS] and stats int64[RGS_N_STATS] into host buffers (any of the three
 * may be NULL).  Synchronous: returns when the outputs are valid.  The copies and launches run on two private
 * streams; they are ordered AFTER whatever the caller has queued on `stream` (the stream of its earlier asynchronous
 * rgs_batch_run / _reset / _get calls on this batch; NULL = the legacy default stream). */
int rgs_batch_run_host(rgs_batch *b, const double *h_x0_ball, const void *h_actions, int64_t n_act, int64_t hold,
                       int64_t n_steps, double *h_state, int64_t *h_events, int64_t *h_stats, void *stream);

/* Allocation-free rollout.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops such as the K3 sampler tick,
 * harnesses that forbid it).  It is rgs_batch_run without the batch object: the fp32 throughput kernel (K2, one lane
 * per game) on n_games 2v2 games whose records the CALLER holds, with the kernel's per-lane scratch (correctOverlap
 * tracks, event counters, jam snapshot) in a file-scope static __device__ array of the library instead of a
 * cudaMalloc'ed buffer.  The call makes no cudaMalloc / cudaFree / allocating library call, on any invocation.
 *   p          mode must be RGS_MODE_FAST_F32, players_per_team 2, noise 0 (RGS_EUNSUPPORTED otherwise)
 *   n_games    0 .. RGS_NOALLOC_MAX_GAMES; more is rejected with RGS_EINVAL and a message naming the limit.  The
 *              static array is RGS_NOALLOC_MAX_GAMES x 2708 B = 11.1 MB; every GPU that loads the library holds it
 *              until the process exits
 *   d_state    [n_games, 22] doubles, in: state before the first step, out: state after the last (in place)
 *   d_clock    [n_games] doubles, in / out (the reference's t_)
 *   d_damage   [n_games, 16] doubles, in / out (accumulated onto)
 *   d_events   [n_games, RGS_N_EVENTS] int64, in / out: counters accumulated onto, the hash carried on (zero them
 *              for a fresh run)
 *   d_actions, n_act, hold, act_game_stride, n_steps: as rgs_batch_run, float records
 * Asynchronous on `stream`.  ALL calls on a device share the one array: nothing in it has to start at zero (each
 * lane re-initialises the words it reads before its game starts, so a call never sees the previous call's
 * contents), but two calls that OVERLAP on the GPU (different streams, or graph branches without a dependency) race
 * on it and corrupt each other -- order them on one stream or with an event.  The first call on a device looks the
 * array's address up (cudaGetSymbolAddress) and loads the kernel; make that call outside graph capture. */
#define RGS_NOALLOC_MAX_GAMES 4096
int rgs_rollout_noalloc(int device, const rgs_params *p, int64_t n_games, double *d_state, double *d_clock,
                        double *d_damage, int64_t *d_events, const float *d_actions, int64_t n_act, int64_t hold,
                        int64_t act_game_stride, int64_t n_steps, void *stream);

/* Batch-of-one convenience used by the drop-in class robo_game_py.GameSim
 * (replaces reference src/RoboGamePy.cpp:15-24 one to one). */
typedef struct rgs_single rgs_single;
int rgs_single_create(rgs_single **out, int device);
int rgs_single_destroy(rgs_single *s);
/* GameSim::reset: h_x0_ball[4]; logname NULL or "" with log=0; h_out_state22 receives the state. */
int rgs_single_reset(rgs_single *s, double dt, int32_t winning_score, const double *h_x0_ball, double noise,
                     int32_t log, const char *logname, int32_t seed, double tau_puck, double tau_player,
                     double player_mass, double puck_mass, double *h_out_state22);
int rgs_single_undecided(rgs_single *s, int32_t *out);
/* GameSim::run: h_actions8 = [A1x,A1y,A2x,A2y,B1x,B1y,B2x,B2y]. */
int rgs_single_run(rgs_single *s, const double *h_actions8, double *h_out_state22);
int rgs_single_time(rgs_single *s, double *out_t);
int rgs_single_damage(rgs_single *s, double *h_out16);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* RGS_CAPI_H */
