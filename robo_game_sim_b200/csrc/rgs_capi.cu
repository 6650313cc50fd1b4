// rgs_capi.cu -- the C ABI of librgs_b200.so (declared in include/rgs_capi.h).
//
// Host-side runtime of the batched GameSim path: owns the device-resident game records
// (state / clock / damage / events, game-major), validates arguments, launches the sm_100a kernels
// (K1 exact fp64, K2 fast fp32/fp64, K4 statistics, K5 reset + synthetic inputs) and implements
// the batch-of-one handle behind the drop-in Python class robo_game_py.GameSim
// (reference src/RoboGamePy.cpp:11-25).  No exception crosses this boundary and there is NO CPU
// implementation of the step anywhere in this library: without a CUDA device every entry point
// that would compute returns RGS_ECUDA.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "rgs_host_params.h"
#include "rgs_kernels.h"

namespace {

thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(cudaError_t e, const char *what) {
    return fail(RGS_ECUDA, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
}

#define RGS_CUDA(call)                                         \
    do {                                                       \
        cudaError_t e__ = (call);                              \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call);  \
    } while (0)


int check_params(const rgs_params *p) {
    if (!p) return fail(RGS_EINVAL, "params is NULL");
    if (p->players_per_team < 1 || p->players_per_team > RGS_MAX_TEAM)
        return fail(RGS_EINVAL, "players_per_team must be 1..%d (got %d)", RGS_MAX_TEAM, p->players_per_team);
    if (p->mode != RGS_MODE_EXACT_F64 && p->mode != RGS_MODE_FAST_F32 && p->mode != RGS_MODE_FAST_F64)
        return fail(RGS_EINVAL, "mode must be RGS_MODE_EXACT_F64, RGS_MODE_FAST_F32 or RGS_MODE_FAST_F64 (got %d)", p->mode);
    if (!(p->dt > 0.0)) return fail(RGS_EINVAL, "dt must be > 0");
    if (!(p->tau_puck > 0.0) || !(p->tau_player > 0.0)) return fail(RGS_EINVAL, "tau_puck and tau_player must be > 0");
    if (!(p->player_mass > 0.0) || !(p->puck_mass > 0.0)) return fail(RGS_EINVAL, "masses must be > 0");
    if (!(p->arena_x > 0.0) || !(p->arena_y > 0.0) || !(p->player_radius > 0.0) || !(p->puck_radius > 0.0))
        return fail(RGS_EINVAL, "arena and radii must be > 0");
    if (!(p->noise >= 0.0)) return fail(RGS_EINVAL, "noise must be >= 0");
    if (p->noise != 0.0 && p->mode == RGS_MODE_EXACT_F64)
        return fail(RGS_EUNSUPPORTED,
                    "process noise runs in the FAST modes (statistical parity): the reference's libstdc++ "
                    "normal_distribution stream is implementation-defined, so a bit-exact noisy mode does not exist");
    return RGS_OK;
}

}  // namespace

struct rgs_batch {
    int device = 0;
    int sm_count = 148;
    int64_t n_games = 0;
    rgs_params p{};
    int n_team = 2, E = 5, srec = 22, dmg_rec = 16;
    rgs_policy pol{};      // the on-device policy of the current rgs_batch_run_policy call
    double *state = nullptr, *clock = nullptr, *damage = nullptr;
    long long *events = nullptr, *stats = nullptr;
    unsigned long long *cost_scratch = nullptr;  // K3 arg-min partials, allocated on first use
    // K2 per-lane scratch (correctOverlap tracks, rgs_lane.cuh), one per stream a rollout can be in flight on:
    // [0] the caller's stream, [1], [2] the two private streams of rgs_batch_run_host; allocated on first use
    void *lane_scratch[3] = {nullptr, nullptr, nullptr};
    size_t lane_scratch_cap[3] = {0, 0, 0};
    size_t lane_scratch_need[2] = {0, 0};  // [noise-free, noise] kernels, queried on first use
    int64_t game_offset = 0;               // global index of game 0 (keys the process-noise streams)
    bool host_records = false;             // state / clock / damage / events live in mapped pinned HOST memory (rgs_single)
    int traj_layout = 0;                   // 0 records [N, n_rec, TREC], 1 field-major [n_rec, TREC, N] (rgs_batch_set_traj_layout)
    // staging for rgs_batch_run_host
    void *stage_actions = nullptr;
    size_t stage_actions_bytes = 0;
    double *stage_x0 = nullptr;
    cudaStream_t streams[2] = {nullptr, nullptr};
    cudaEvent_t done_ev[2] = {nullptr, nullptr};
    cudaEvent_t order_ev = nullptr;
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// one fused launch over games [g0, g1) of the batch
int run_range(rgs_batch *b, int64_t g0, int64_t g1, const void *d_actions, int64_t n_act, int64_t hold,
              int64_t act_game_stride, int64_t n_steps, int64_t log_stride, double *d_traj, int synth, uint64_t seed,
              int64_t first_game, int64_t step0, double u_max, cudaStream_t stream, int slot = 0) {
    if (g1 <= g0 || n_steps == 0) return RGS_OK;
    rgs::RunArgs A{};
    const int np = 2 * b->n_team;
    A.n_games = g1 - g0;
    A.state = b->state + g0 * b->srec;
    A.clock = b->clock + g0;
    A.damage = b->damage + g0 * b->dmg_rec;
    A.events = b->events + g0 * RGS_N_EVENTS;
    const size_t esz = b->p.mode == RGS_MODE_FAST_F32 ? sizeof(float) : sizeof(double);
    A.actions = d_actions ? static_cast<const char *>(d_actions) + (size_t)(g0 * act_game_stride) * esz : nullptr;
    A.n_act = n_act;
    A.hold = hold;
    A.act_game_stride = act_game_stride;
    A.n_steps = n_steps;
    A.log_stride = log_stride;
    A.traj_recs = log_stride > 0 ? n_steps / log_stride : 0;
    if (b->traj_layout == 0) {
        A.traj_field_stride = 1;
        A.traj = d_traj ? d_traj + g0 * A.traj_recs * (1 + b->srec + np * np) : nullptr;
    } else {   // field-major: the kernel indexes by the game's position in [g0, g1), so the base moves by g0 games
        A.traj_field_stride = b->n_games;
        A.traj = d_traj ? d_traj + g0 : nullptr;
    }
    A.synth = synth;
    A.seed = seed;
    A.first_game = first_game + g0;
    A.step0 = step0;
    A.u_max = u_max;
    if (synth == 2) {
        if (b->p.mode == RGS_MODE_EXACT_F64 || (b->n_team != 2 && b->p.noise != 0.0))
            return fail(RGS_EUNSUPPORTED, "the on-device policy runs in the FAST modes (with process noise: players_per_team == 2 only)");
        A.pol = rgs::PolicyArgs{b->pol.kick_velocity, b->pol.gain, b->pol.input_limit, b->pol.safety_radius,
                                b->pol.barrier_gain, b->pol.cbf, b->pol.sweeps, b->pol.kind};
    }
    if (b->p.mode != RGS_MODE_EXACT_F64) {
        const int nz = b->p.noise != 0.0 ? 1 : 0;
        if (b->lane_scratch_need[nz] == 0) {
            cudaError_t qe = rgs::lane_scratch_bytes(b->n_team, b->p.mode == RGS_MODE_FAST_F64, nz, b->sm_count, &b->lane_scratch_need[nz]);
            if (qe != cudaSuccess) return cuda_fail(qe, "lane scratch size query");
        }
        if (b->lane_scratch_cap[slot] < b->lane_scratch_need[nz]) {
            if (b->lane_scratch[slot]) {   // a batch switched to the (larger) noise kernels: in-flight work first
                RGS_CUDA(cudaDeviceSynchronize());
                cudaFree(b->lane_scratch[slot]);
                b->lane_scratch[slot] = nullptr;
                b->lane_scratch_cap[slot] = 0;
            }
            cudaError_t me = cudaMalloc(&b->lane_scratch[slot], b->lane_scratch_need[nz]);
            if (me != cudaSuccess) {
                cudaGetLastError();
                return fail(RGS_ENOMEM, "lane scratch of %zu bytes: %s", b->lane_scratch_need[nz], cudaGetErrorString(me));
            }
            b->lane_scratch_cap[slot] = b->lane_scratch_need[nz];
        }
        A.lane_scratch = b->lane_scratch[slot];
        A.lane_scratch_bytes = b->lane_scratch_cap[slot];
        A.noise_game0 = b->game_offset + g0;
    }
    cudaError_t e;
    switch (b->p.mode) {
    case RGS_MODE_EXACT_F64:
        e = rgs::launch_rollout_exact(b->n_team, rgs::make_dev_params<double>(b->p), A, stream, b->sm_count);
        break;
    case RGS_MODE_FAST_F32:
        e = rgs::launch_rollout_lane_f32(b->n_team, b->p, A, stream, b->sm_count);
        break;
    default:
        e = rgs::launch_rollout_lane_f64(b->n_team, b->p, A, stream, b->sm_count);
        break;
    }
    if (e != cudaSuccess) return cuda_fail(e, "rollout kernel launch");
    g_launches.fetch_add(1);
    return RGS_OK;
}

int check_run_args(rgs_batch *b, int64_t n_act, int64_t hold, int64_t n_steps, int64_t log_stride, const double *d_traj) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    if (n_steps < 0) return fail(RGS_EINVAL, "n_steps must be >= 0");
    if (hold < 1) return fail(RGS_EINVAL, "hold must be >= 1");
    if (n_act < 1) return fail(RGS_EINVAL, "n_act must be >= 1");
    if (log_stride < 0) return fail(RGS_EINVAL, "log_stride must be >= 0");
    if (d_traj && log_stride == 0) return fail(RGS_EINVAL, "d_traj given but log_stride is 0");
    return RGS_OK;
}

}  // namespace

// hooks for entry points that live in other translation units (rgs_noalloc.cu): the same argument checks, the same
// thread-local rgs_last_error() message and the same launch counter as the entry points of this file
namespace rgs {
int capi_check_params(const rgs_params *p) { return check_params(p); }
int capi_fail(int code, const char *msg) { return fail(code, "%s", msg); }
int capi_cuda_fail(cudaError_t e, const char *what) { return cuda_fail(e, what); }
void capi_count_launches(int n) { g_launches.fetch_add(n); }
}  // namespace rgs

extern "C" {

void rgs_default_params(rgs_params *p) {
    if (!p) return;
    std::memset(p, 0, sizeof(*p));
    p->dt = 0.05;
    p->winning_score = 3;
    p->players_per_team = 2;
    p->noise = 0.0;
    p->tau_puck = 0.1;
    p->tau_player = 0.5;
    p->player_mass = 1.0;
    p->puck_mass = 0.5;
    p->arena_x = 10.0;
    p->arena_y = 5.0;
    p->player_radius = 0.2;
    p->puck_radius = 0.175;
    p->goal_height = 1.0;
    p->mode = RGS_MODE_EXACT_F64;
}

const char *rgs_last_error(void) { return g_err; }
int64_t rgs_launch_count(void) { return g_launches.load(); }

int rgs_device_count(int *out) {
    if (!out) return fail(RGS_EINVAL, "out is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *out = 0;
        return cuda_fail(e, "cudaGetDeviceCount");
    }
    *out = n;
    return RGS_OK;
}

}  // extern "C"

namespace {
// host_records: the game records are allocated in mapped pinned host memory instead of device memory.  The kernels
// read and write them through the same (unified) addresses; the host reads results after a stream synchronise without
// any cudaMemcpy.  Only the batch-of-one handle behind GameSim uses it: one launch per run() call and nothing else.
int batch_create(rgs_batch **out, int device, int64_t n_games, const rgs_params *p, bool host_records) {
    if (!out) return fail(RGS_EINVAL, "out is NULL");
    *out = nullptr;
    if (n_games < 0) return fail(RGS_EINVAL, "n_games must be >= 0");
    int rc = check_params(p);
    if (rc != RGS_OK) return rc;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(RGS_ECUDA, "no CUDA device available (%s); this library has no CPU path",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= ndev) return fail(RGS_EINVAL, "device %d out of range (0..%d)", device, ndev - 1);
    DeviceGuard guard(device);
    if (!guard.ok) return fail(RGS_ECUDA, "cudaSetDevice(%d) failed", device);
    rgs_batch *b = new (std::nothrow) rgs_batch();
    if (!b) return fail(RGS_ENOMEM, "out of host memory");
    b->device = device;
    b->n_games = n_games;
    b->p = *p;
    b->n_team = p->players_per_team;
    b->E = 2 * b->n_team + 1;
    b->srec = 2 + 4 * b->E;
    b->dmg_rec = 4 * b->n_team * b->n_team;
    cudaDeviceGetAttribute(&b->sm_count, cudaDevAttrMultiProcessorCount, device);
    const size_t n = (size_t)(n_games > 0 ? n_games : 1);
    cudaError_t err = cudaSuccess;
    b->host_records = host_records;
    auto alloc = [&](void **ptr, size_t bytes) {
        return host_records ? cudaHostAlloc(ptr, bytes, cudaHostAllocMapped) : cudaMalloc(ptr, bytes);
    };
    if (err == cudaSuccess) err = alloc(reinterpret_cast<void **>(&b->state), n * b->srec * sizeof(double));
    if (err == cudaSuccess) err = alloc(reinterpret_cast<void **>(&b->clock), n * sizeof(double));
    if (err == cudaSuccess) err = alloc(reinterpret_cast<void **>(&b->damage), n * b->dmg_rec * sizeof(double));
    if (err == cudaSuccess) err = alloc(reinterpret_cast<void **>(&b->events), n * RGS_N_EVENTS * sizeof(long long));
    if (err == cudaSuccess) err = cudaMalloc(&b->stats, RGS_N_STATS * sizeof(long long));
    if (err != cudaSuccess) {
        rgs_batch_destroy(b);
        cudaGetLastError();
        return fail(err == cudaErrorMemoryAllocation ? RGS_ENOMEM : RGS_ECUDA, "device allocation for %lld games failed: %s",
                    (long long)n_games, cudaGetErrorString(err));
    }
    *out = b;
    int rc2 = rgs_batch_reset(b, nullptr, 0, nullptr, nullptr);
    if (rc2 != RGS_OK) {
        rgs_batch_destroy(b);
        *out = nullptr;
        return rc2;
    }
    return RGS_OK;
}
}  // namespace

extern "C" {

int rgs_batch_create(rgs_batch **out, int device, int64_t n_games, const rgs_params *p) {
    return batch_create(out, device, n_games, p, false);
}

int rgs_batch_destroy(rgs_batch *b) {
    if (!b) return RGS_OK;
    DeviceGuard guard(b->device);
    cudaDeviceSynchronize();
    if (b->host_records) {
        cudaFreeHost(b->state);
        cudaFreeHost(b->clock);
        cudaFreeHost(b->damage);
        cudaFreeHost(b->events);
    } else {
        cudaFree(b->state);
        cudaFree(b->clock);
        cudaFree(b->damage);
        cudaFree(b->events);
    }
    cudaFree(b->stats);
    cudaFree(b->cost_scratch);
    for (int k = 0; k < 3; ++k) cudaFree(b->lane_scratch[k]);
    cudaFree(b->stage_actions);
    cudaFree(b->stage_x0);
    for (int k = 0; k < 2; ++k) {
        if (b->streams[k]) cudaStreamDestroy(b->streams[k]);
        if (b->done_ev[k]) cudaEventDestroy(b->done_ev[k]);
    }
    if (b->order_ev) cudaEventDestroy(b->order_ev);
    delete b;
    return RGS_OK;
}

int rgs_batch_set_params(rgs_batch *b, const rgs_params *p) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    int rc = check_params(p);
    if (rc != RGS_OK) return rc;
    if (p->players_per_team != b->n_team) return fail(RGS_EINVAL, "players_per_team is fixed at create time");
    if (p->mode != b->p.mode) return fail(RGS_EINVAL, "mode is fixed at create time");
    b->p = *p;
    return RGS_OK;
}

int rgs_batch_set_game_offset(rgs_batch *b, int64_t first_game) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    b->game_offset = first_game;
    return RGS_OK;
}

int rgs_batch_reset(rgs_batch *b, const double *d_x0_ball, int64_t x0_stride, const double *d_state0, void *stream) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    if (x0_stride != 0 && x0_stride < 4) return fail(RGS_EINVAL, "x0_stride must be 0 (shared) or >= 4");
    if (b->n_games == 0) return RGS_OK;
    DeviceGuard guard(b->device);
    cudaError_t e = rgs::launch_reset(b->n_team, b->n_games, b->p, d_x0_ball, x0_stride, d_state0, b->srec, b->state, b->clock,
                                      b->damage, b->events, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "reset kernel launch");
    g_launches.fetch_add(1);
    return RGS_OK;
}

int rgs_batch_reset_shared(rgs_batch *b, const double *d_state, void *stream) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    if (!d_state) return fail(RGS_EINVAL, "d_state is NULL");
    if (b->n_games == 0) return RGS_OK;
    DeviceGuard guard(b->device);
    cudaError_t e = rgs::launch_reset(b->n_team, b->n_games, b->p, nullptr, 0, d_state, 0, b->state, b->clock, b->damage,
                                      b->events, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "reset kernel launch");
    g_launches.fetch_add(1);
    return RGS_OK;
}

int rgs_batch_terminal_cost(rgs_batch *b, const double *d_x_des, const double *d_weights, double *d_cost, int64_t *d_best,
                            void *stream) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    if (!d_x_des || !d_weights || !d_cost || !d_best) return fail(RGS_EINVAL, "d_x_des, d_weights, d_cost and d_best are required");
    if (b->n_games == 0) return fail(RGS_EINVAL, "empty batch has no minimum");
    DeviceGuard guard(b->device);
    if (!b->cost_scratch) RGS_CUDA(cudaMalloc(&b->cost_scratch, 2 * rgs::kCostBlocks * sizeof(unsigned long long)));
    cudaError_t e = rgs::launch_terminal_cost(b->n_team, b->n_games, b->state, d_x_des, d_weights, d_cost, b->cost_scratch,
                                              reinterpret_cast<long long *>(d_best), static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "terminal cost kernel launch");
    g_launches.fetch_add(2);
    return RGS_OK;
}

int rgs_batch_run(rgs_batch *b, const void *d_actions, int64_t n_act, int64_t hold, int64_t act_game_stride,
                  int64_t n_steps, int64_t log_stride, double *d_traj, void *stream) {
    int rc = check_run_args(b, n_act, hold, n_steps, log_stride, d_traj);
    if (rc != RGS_OK) return rc;
    if (b->n_games == 0) return RGS_OK;  // an empty batch has no command tensor to point at
    if (!d_actions) return fail(RGS_EINVAL, "d_actions is NULL");
    if (act_game_stride < 0) return fail(RGS_EINVAL, "act_game_stride must be >= 0");
    DeviceGuard guard(b->device);
    return run_range(b, 0, b->n_games, d_actions, n_act, hold, act_game_stride, n_steps, log_stride, d_traj, 0, 0, 0, 0, 0.0,
                     static_cast<cudaStream_t>(stream));
}

int rgs_batch_run_synth(rgs_batch *b, uint64_t seed, int64_t first_game, int64_t step0, int64_t hold, double u_max,
                        int64_t n_steps, int64_t log_stride, double *d_traj, void *stream) {
    int rc = check_run_args(b, 1, hold, n_steps, log_stride, d_traj);
    if (rc != RGS_OK) return rc;
    if (step0 < 0) return fail(RGS_EINVAL, "step0 must be >= 0");
    DeviceGuard guard(b->device);
    return run_range(b, 0, b->n_games, nullptr, 1, hold, 0, n_steps, log_stride, d_traj, 1, seed, first_game, step0, u_max,
                     static_cast<cudaStream_t>(stream));
}

void rgs_default_policy(rgs_policy *pol) {
    if (!pol) return;
    pol->kick_velocity = 4.0;
    pol->gain = 4.0;
    pol->input_limit = 10.0;
    pol->safety_radius = 0.4;
    pol->barrier_gain = 30.0;
    pol->cbf = 0;
    pol->sweeps = 8;
    pol->kind = 0;
    pol->reserved = 0;
}

int rgs_batch_set_traj_layout(rgs_batch *b, int32_t layout) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    if (layout != 0 && layout != 1) return fail(RGS_EINVAL, "trajectory layout must be 0 (records) or 1 (field-major)");
    b->traj_layout = layout;
    return RGS_OK;
}

int rgs_batch_run_policy(rgs_batch *b, const rgs_policy *pol, int64_t n_steps, int64_t log_stride, double *d_traj,
                         void *stream) {
    int rc = check_run_args(b, 1, 1, n_steps, log_stride, d_traj);
    if (rc != RGS_OK) return rc;
    if (!pol) return fail(RGS_EINVAL, "pol is NULL");
    if (pol->kind != 0 && pol->kind != 1) return fail(RGS_EINVAL, "policy: kind must be 0 (classical) or 1 (dmpc)");
    if (!(pol->input_limit > 0.0) || !(pol->gain >= 0.0)) return fail(RGS_EINVAL, "policy: input_limit must be > 0 and gain >= 0");
    if (pol->cbf && (pol->sweeps < 1 || !(pol->barrier_gain > 0.0) || !(pol->safety_radius >= 2.0 * b->p.player_radius)))
        return fail(RGS_EINVAL, "policy: the barrier filter needs sweeps >= 1, barrier_gain > 0 and safety_radius >= 2 player_radius (CBF.py:29-30)");
    DeviceGuard guard(b->device);
    b->pol = *pol;
    return run_range(b, 0, b->n_games, nullptr, 1, 1, 0, n_steps, log_stride, d_traj, 2, 0, 0, 0, 0.0,
                     static_cast<cudaStream_t>(stream));
}

int rgs_batch_get(rgs_batch *b, double *d_state, double *d_t, double *d_damage, int64_t *d_events, void *stream) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    if (b->n_games == 0) return RGS_OK;
    DeviceGuard guard(b->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const size_t n = (size_t)b->n_games;
    // cudaMemcpyDefault: the destination may be device memory or (pinned / pageable) host memory
    if (d_state) RGS_CUDA(cudaMemcpyAsync(d_state, b->state, n * b->srec * sizeof(double), cudaMemcpyDefault, s));
    if (d_t) RGS_CUDA(cudaMemcpyAsync(d_t, b->clock, n * sizeof(double), cudaMemcpyDefault, s));
    if (d_damage) RGS_CUDA(cudaMemcpyAsync(d_damage, b->damage, n * b->dmg_rec * sizeof(double), cudaMemcpyDefault, s));
    if (d_events) RGS_CUDA(cudaMemcpyAsync(d_events, b->events, n * RGS_N_EVENTS * sizeof(long long), cudaMemcpyDefault, s));
    return RGS_OK;
}

int rgs_batch_set(rgs_batch *b, const double *d_state, const double *d_t, const double *d_damage, const int64_t *d_events,
                  void *stream) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    if (b->n_games == 0) return RGS_OK;
    DeviceGuard guard(b->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const size_t n = (size_t)b->n_games;
    if (d_state) RGS_CUDA(cudaMemcpyAsync(b->state, d_state, n * b->srec * sizeof(double), cudaMemcpyDefault, s));
    if (d_t) RGS_CUDA(cudaMemcpyAsync(b->clock, d_t, n * sizeof(double), cudaMemcpyDefault, s));
    if (d_damage) RGS_CUDA(cudaMemcpyAsync(b->damage, d_damage, n * b->dmg_rec * sizeof(double), cudaMemcpyDefault, s));
    if (d_events) RGS_CUDA(cudaMemcpyAsync(b->events, d_events, n * RGS_N_EVENTS * sizeof(long long), cudaMemcpyDefault, s));
    return RGS_OK;
}

int rgs_batch_stats(rgs_batch *b, int64_t *d_stats, void *stream) {
    if (!b) return fail(RGS_EINVAL, "batch is NULL");
    if (!d_stats) return fail(RGS_EINVAL, "d_stats is NULL");
    DeviceGuard guard(b->device);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    cudaError_t e = rgs::launch_stats(b->n_team, b->n_games, b->state, b->events, b->damage, b->p.winning_score, b->stats, s, b->sm_count);
    if (e != cudaSuccess) return cuda_fail(e, "stats kernel launch");
    g_launches.fetch_add(1);
    RGS_CUDA(cudaMemcpyAsync(d_stats, b->stats, RGS_N_STATS * sizeof(long long), cudaMemcpyDefault, s));
    return RGS_OK;
}

int rgs_synth_inputs(int device, uint64_t seed, int64_t first_game, int64_t n, int32_t players_per_team, int64_t n_act,
                     double u_max, double *d_x0_ball, void *d_actions, int32_t dtype, void *stream) {
    if (n < 0 || n_act < 0) return fail(RGS_EINVAL, "n and n_act must be >= 0");
    if (players_per_team < 1 || players_per_team > RGS_MAX_TEAM) return fail(RGS_EINVAL, "players_per_team out of range");
    if (dtype != 32 && dtype != 64) return fail(RGS_EINVAL, "dtype must be 32 or 64");
    if (n == 0 || (!d_x0_ball && (!d_actions || n_act == 0))) return RGS_OK;
    DeviceGuard guard(device);
    if (!guard.ok) return fail(RGS_ECUDA, "cudaSetDevice(%d) failed", device);
    cudaError_t e = rgs::launch_synth(seed, first_game, n, players_per_team, n_act, u_max, d_x0_ball,
                                      n_act > 0 ? d_actions : nullptr, dtype, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "synth kernel launch");
    g_launches.fetch_add(1);
    return RGS_OK;
}

int rgs_measure_peak(int device, int32_t fp64, double *out_tflops) {
    if (!out_tflops) return fail(RGS_EINVAL, "out_tflops is NULL");
    DeviceGuard guard(device);
    if (!guard.ok) return fail(RGS_ECUDA, "cudaSetDevice(%d) failed", device);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    cudaError_t e = rgs::measure_fma_peak(fp64 ? 1 : 0, sms, out_tflops);
    if (e != cudaSuccess) return cuda_fail(e, "fma peak micro-kernel");
    g_launches.fetch_add(6);
    return RGS_OK;
}

int rgs_batch_run_host(rgs_batch *b, const double *h_x0_ball, const void *h_actions, int64_t n_act, int64_t hold,
                       int64_t n_steps, double *h_state, int64_t *h_events, int64_t *h_stats, void *stream) {
    int rc = check_run_args(b, n_act, hold, n_steps, 0, nullptr);
    if (rc != RGS_OK) return rc;
    if (!h_actions) return fail(RGS_EINVAL, "h_actions is NULL");
    DeviceGuard guard(b->device);
    const int64_t N = b->n_games;
    const size_t esz = b->p.mode == RGS_MODE_FAST_F32 ? sizeof(float) : sizeof(double);
    const int64_t arec = n_act * 4 * b->n_team;  // action elements per game
    const size_t need = (size_t)N * arec * esz;
    if (b->stage_actions_bytes < need) {
        cudaFree(b->stage_actions);
        b->stage_actions = nullptr;
        b->stage_actions_bytes = 0;
        cudaError_t e = cudaMalloc(&b->stage_actions, need ? need : 1);
        if (e != cudaSuccess) return fail(RGS_ENOMEM, "staging %zu action bytes: %s", need, cudaGetErrorString(e));
        b->stage_actions_bytes = need;
    }
    if (h_x0_ball && !b->stage_x0) RGS_CUDA(cudaMalloc(&b->stage_x0, (size_t)(N > 0 ? N : 1) * 4 * sizeof(double)));
    for (int k = 0; k < 2; ++k) {
        if (!b->streams[k]) RGS_CUDA(cudaStreamCreateWithFlags(&b->streams[k], cudaStreamNonBlocking));
        if (!b->done_ev[k]) RGS_CUDA(cudaEventCreateWithFlags(&b->done_ev[k], cudaEventDisableTiming));
    }
    // Work the caller queued on ITS stream for this batch (an asynchronous run / reset / get) must finish before the
    // private streams touch the same records: they are non-blocking streams and do not synchronise with it implicitly.
    if (!b->order_ev) RGS_CUDA(cudaEventCreateWithFlags(&b->order_ev, cudaEventDisableTiming));
    RGS_CUDA(cudaEventRecord(b->order_ev, static_cast<cudaStream_t>(stream)));
    RGS_CUDA(cudaStreamWaitEvent(b->streams[0], b->order_ev, 0));
    RGS_CUDA(cudaStreamWaitEvent(b->streams[1], b->order_ev, 0));
    // Game-range chunks, double-buffered over two streams: the H2D copy of chunk c+1 and the D2H
    // copy of chunk c-1 overlap the rollout of chunk c (games are independent, so a chunk is
    // simply a contiguous slice of the batch).
    // What stays exposed is the H2D copy of the FIRST chunk and the D2H copy of the LAST one, so those two ends are
    // cut finer than the middle (plan in 32nds of the batch; profiles/r02_ab_variants.md section 6).
    // Measured (tools/ab_e2e.sh, 10^6 games x 200 steps): 8 equal chunks 66.4 - 66.7 ms per pass at N = 1 and
    // 2.22e10 game-steps/s end to end on 8 GPUs; this plan 64.5 - 65.3 ms and 2.32e10 (98 % of the device-timed rate).
    // Chunk sizes are also a matter of wave quantisation on the 75 776-lane resident grid (5/32 + 10/32 loses).
#if defined(RGS_AB_PLAN) && RGS_AB_PLAN == 1
    static const int kPlan[] = {4, 4, 4, 4, 4, 4, 4, 4};
#else
    static const int kPlan[] = {2, 6, 8, 8, 6, 2};
#endif
    constexpr int kPlanLen = (int)(sizeof(kPlan) / sizeof(kPlan[0]));
    int64_t bounds[kPlanLen + 1];
    int64_t n_chunks = 1;
    bounds[0] = 0;
    bounds[1] = N;
    if (N >= 65536) {
        n_chunks = kPlanLen;
        int acc = 0;
        for (int c = 0; c < kPlanLen; ++c) {
            acc += kPlan[c];
            bounds[c + 1] = N / 32 * acc + (N % 32) * acc / 32;
        }
        bounds[kPlanLen] = N;
    }
    for (int64_t c = 0; c < n_chunks; ++c) {
        const int64_t g0 = bounds[c], g1 = bounds[c + 1];
        if (g1 <= g0) continue;
        cudaStream_t s = b->streams[c & 1];
        const int64_t ng = g1 - g0;
        RGS_CUDA(cudaMemcpyAsync(static_cast<char *>(b->stage_actions) + (size_t)g0 * arec * esz,
                                 static_cast<const char *>(h_actions) + (size_t)g0 * arec * esz, (size_t)ng * arec * esz,
                                 cudaMemcpyHostToDevice, s));
        if (h_x0_ball)
            RGS_CUDA(cudaMemcpyAsync(b->stage_x0 + g0 * 4, h_x0_ball + g0 * 4, (size_t)ng * 4 * sizeof(double),
                                     cudaMemcpyHostToDevice, s));
        cudaError_t e = rgs::launch_reset(b->n_team, ng, b->p, h_x0_ball ? b->stage_x0 + g0 * 4 : nullptr, 4, nullptr, 0,
                                          b->state + g0 * b->srec, b->clock + g0, b->damage + g0 * b->dmg_rec,
                                          b->events + g0 * RGS_N_EVENTS, s);
        if (e != cudaSuccess) return cuda_fail(e, "reset kernel launch");
        g_launches.fetch_add(1);
        rc = run_range(b, g0, g1, b->stage_actions, n_act, hold, arec, n_steps, 0, nullptr, 0, 0, 0, 0, 0.0, s, 1 + (int)(c & 1));
        if (rc != RGS_OK) return rc;
        if (h_state)
            RGS_CUDA(cudaMemcpyAsync(h_state + g0 * b->srec, b->state + g0 * b->srec, (size_t)ng * b->srec * sizeof(double),
                                     cudaMemcpyDeviceToHost, s));
        if (h_events)
            RGS_CUDA(cudaMemcpyAsync(h_events + g0 * RGS_N_EVENTS, b->events + g0 * RGS_N_EVENTS,
                                     (size_t)ng * RGS_N_EVENTS * sizeof(long long), cudaMemcpyDeviceToHost, s));
        RGS_CUDA(cudaEventRecord(b->done_ev[c & 1], s));
    }
    // statistics over the whole batch once both streams have drained into stream 0
    RGS_CUDA(cudaStreamWaitEvent(b->streams[0], b->done_ev[1], 0));
    if (h_stats) {
        rc = rgs_batch_stats(b, h_stats, b->streams[0]);
        if (rc != RGS_OK) return rc;
    }
    RGS_CUDA(cudaStreamSynchronize(b->streams[1]));
    RGS_CUDA(cudaStreamSynchronize(b->streams[0]));
    return RGS_OK;
}

// One binary log per game in the reference's on-disk format (src/GameSim.cpp:161-170, include/utils.h:66-76): per
// record t, TAS, TBS, puck x y, every player's x y, then (2n)^2 damage fields (29 doubles at 2v2).
// layout 0 ("reference"): damage.row(k) is logged by utils.h:75 as 2n CONTIGUOUS doubles from &damage(k,0) of a
// COLUMN-major matrix, i.e. linear indices k .. k+2n-1 of the column-major buffer -- not a row; kept so that files
// are byte-identical to the reference's.  layout 1 ("row_major"): the true rows, damage(aggressor, victim), which is
// what matlab/interpret_mc_results.m:37-48 assumes when it names the fields.
int rgs_write_logs(const double *h_traj, int64_t n_games, int64_t n_rec, int32_t players_per_team, const int64_t *h_n_records,
                   const char *path_pattern, int64_t first_index, int32_t layout, int32_t append) {
    if (n_games < 0 || n_rec < 0) return fail(RGS_EINVAL, "n_games and n_rec must be >= 0");
    if (n_games > 0 && n_rec > 0 && !h_traj) return fail(RGS_EINVAL, "h_traj is NULL");
    if (players_per_team < 1 || players_per_team > RGS_MAX_TEAM) return fail(RGS_EINVAL, "players_per_team out of range");
    if (layout != 0 && layout != 1) return fail(RGS_EINVAL, "layout must be 0 (reference) or 1 (row_major)");
    if (!path_pattern) return fail(RGS_EINVAL, "path_pattern is NULL");
    const std::string pat(path_pattern);
    const size_t brace = pat.find("{}");
    if (brace == std::string::npos || pat.find("{}", brace + 2) != std::string::npos)
        return fail(RGS_EINVAL, "path_pattern must contain exactly one {} (replaced by the run index)");
    const int np = 2 * players_per_team, E = np + 1, srec = 2 + 4 * E, trec = 1 + srec + np * np;
    const int frec = 3 + 2 * E + np * np;
    std::vector<double> buf;
    for (int64_t g = 0; g < n_games; ++g) {
        int64_t nr = h_n_records ? h_n_records[g] : n_rec;
        if (nr < 0) nr = 0;
        if (nr > n_rec) nr = n_rec;
        buf.resize((size_t)nr * frec);
        for (int64_t r = 0; r < nr; ++r) {
            const double *in = h_traj + ((size_t)g * n_rec + r) * trec;
            double *out = buf.data() + (size_t)r * frec;
            out[0] = in[0];
            out[1] = in[1];
            out[2] = in[2];
            for (int e = 0; e < E; ++e) {  // positions only: velocities are not logged
                out[3 + 2 * e] = in[3 + 4 * e];
                out[4 + 2 * e] = in[4 + 4 * e];
            }
            const double *dm = in + 1 + srec;  // (aggressor, victim) row-major
            double *od = out + 3 + 2 * E;
            for (int k = 0; k < np; ++k)
                for (int c = 0; c < np; ++c) {
                    if (layout == 1) {
                        od[k * np + c] = dm[k * np + c];
                    } else {
                        const int lin = k + c, row = lin % np, col = lin / np;
                        od[k * np + c] = dm[row * np + col];
                    }
                }
        }
        const std::string path = pat.substr(0, brace) + std::to_string((long long)(first_index + g)) + pat.substr(brace + 2);
        FILE *f = fopen(path.c_str(), append ? "ab" : "wb");
        if (!f) return fail(RGS_EINVAL, "cannot open %s for writing", path.c_str());
        const size_t wrote = buf.empty() ? 0 : fwrite(buf.data(), sizeof(double), buf.size(), f);
        fclose(f);
        if (wrote != buf.size()) return fail(RGS_EINVAL, "short write to %s", path.c_str());
    }
    return RGS_OK;
}

}  // extern "C"

// ---- batch-of-one handle behind robo_game_py.GameSim -----------------------------------------
struct rgs_single {
    int device = 0;
    rgs_batch *b = nullptr;        // the batch of the current game: b_exact (K1) or, with process noise, b_noise (K2 fp64)
    rgs_batch *b_exact = nullptr, *b_noise = nullptr;
    rgs_params p{};
    // Zero-copy: the game records of the handle's batches and this scratch are mapped pinned host memory, so a run()
    // call is ONE kernel launch and a stream synchronise -- the kernel fetches the eight commands and the record over
    // PCIe and posts the new state back; no cudaMemcpy on the path (54 us -> see DESIGN.md section 6, latency row).
    double *h_pin = nullptr;  // pinned, device-visible: [0..7] commands of the call, [8..11] x0_ball of reset()
    cudaStream_t stream = nullptr;
    FILE *log = nullptr;
    bool is_reset = false;
    double h_state[22];
};

extern "C" {

int rgs_single_create(rgs_single **out, int device) {
    if (!out) return fail(RGS_EINVAL, "out is NULL");
    *out = nullptr;
    rgs_single *s = new (std::nothrow) rgs_single();
    if (!s) return fail(RGS_ENOMEM, "out of host memory");
    s->device = device;
    rgs_default_params(&s->p);
    int rc = batch_create(&s->b_exact, device, 1, &s->p, true);
    if (rc != RGS_OK) { delete s; return rc; }
    s->b = s->b_exact;
    {   // the create-time reset ran on the legacy stream; this handle works on its own non-blocking stream
        DeviceGuard guard(device);
        cudaDeviceSynchronize();
    }
    DeviceGuard guard(device);
    cudaError_t e = cudaHostAlloc(reinterpret_cast<void **>(&s->h_pin), 16 * sizeof(double), cudaHostAllocMapped);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { rgs_single_destroy(s); return cuda_fail(e, "rgs_single_create"); }
    *out = s;
    return RGS_OK;
}

int rgs_single_destroy(rgs_single *s) {
    if (!s) return RGS_OK;
    {
        DeviceGuard guard(s->device);
        if (s->stream) { cudaStreamSynchronize(s->stream); cudaStreamDestroy(s->stream); }
        if (s->h_pin) cudaFreeHost(s->h_pin);
    }
    if (s->log) fclose(s->log);
    rgs_batch_destroy(s->b_exact);
    rgs_batch_destroy(s->b_noise);
    delete s;
    return RGS_OK;
}

int rgs_single_reset(rgs_single *s, double dt, int32_t winning_score, const double *h_x0_ball, double noise, int32_t log,
                     const char *logname, int32_t seed, double tau_puck, double tau_player, double player_mass,
                     double puck_mass, double *h_out_state22) {
    if (!s) return fail(RGS_EINVAL, "handle is NULL");
    if (!h_x0_ball) return fail(RGS_EINVAL, "x0_ball is NULL");
    rgs_params p = s->p;
    p.dt = dt;
    p.winning_score = winning_score;
    p.noise = noise;
    p.seed = seed;   // reng_ = std::default_random_engine(seed), src/GameSim.cpp:41
    // noise = 0 (every shipped reference script): the event-exact fp64 kernel K1.  noise > 0: the K2 formulation in
    // fp64 with Philox process noise -- statistically, not bitwise, the reference's (its normal stream is libstdc++'s)
    p.mode = noise != 0.0 ? RGS_MODE_FAST_F64 : RGS_MODE_EXACT_F64;
    if (noise != 0.0 && !s->b_noise) {
        int rcn = batch_create(&s->b_noise, s->device, 1, &p, true);
        if (rcn != RGS_OK) return rcn;
        DeviceGuard guard(s->device);
        cudaDeviceSynchronize();
    }
    s->b = noise != 0.0 ? s->b_noise : s->b_exact;
    p.tau_puck = tau_puck;
    p.tau_player = tau_player;
    p.player_mass = player_mass;
    p.puck_mass = puck_mass;
    int rc = rgs_batch_set_params(s->b, &p);
    if (rc != RGS_OK) return rc;
    s->p = p;
    DeviceGuard guard(s->device);
    std::memcpy(s->h_pin + 8, h_x0_ball, 4 * sizeof(double));
    rc = rgs_batch_reset(s->b, s->h_pin + 8, 4, nullptr, s->stream);
    if (rc != RGS_OK) return rc;
    RGS_CUDA(cudaStreamSynchronize(s->stream));
    std::memcpy(s->h_state, s->b->state, sizeof(s->h_state));   // mapped pinned memory, written by the reset kernel
    if (h_out_state22) std::memcpy(h_out_state22, s->h_state, sizeof(s->h_state));
    // Logger::open (reference include/utils.h:54-59): close, then reopen truncating; a failed open is silent
    if (s->log) { fclose(s->log); s->log = nullptr; }
    if (log && logname && logname[0]) s->log = fopen(logname, "wb");
    s->is_reset = true;
    return RGS_OK;
}

int rgs_single_undecided(rgs_single *s, int32_t *out) {
    if (!s || !out) return fail(RGS_EINVAL, "NULL argument");
    if (!s->is_reset) return fail(RGS_EINVAL, "reset() has not been called");
    *out = (s->h_state[0] < (double)s->p.winning_score) && (s->h_state[1] < (double)s->p.winning_score);
    return RGS_OK;
}

int rgs_single_run(rgs_single *s, const double *h_actions8, double *h_out_state22) {
    if (!s || !h_actions8) return fail(RGS_EINVAL, "NULL argument");
    if (!s->is_reset) return fail(RGS_EINVAL, "reset() has not been called");
    int32_t live = 0;
    rgs_single_undecided(s, &live);
    if (live) {  // GameSim::run is a no-op once decided (reference src/GameSim.cpp:121-123)
        DeviceGuard guard(s->device);
        std::memcpy(s->h_pin, h_actions8, 8 * sizeof(double));
        int rc = rgs_batch_run(s->b, s->h_pin, 1, 1, 0, 1, 0, nullptr, s->stream);
        if (rc != RGS_OK) return rc;
        RGS_CUDA(cudaStreamSynchronize(s->stream));
        std::memcpy(s->h_state, s->b->state, sizeof(s->h_state));
        if (s->log) {
            // 29-double record (reference src/GameSim.cpp:161-170).  damage.row(k) is written by
            // include/utils.h:75 as 4 CONTIGUOUS doubles from &damage(k,0) of a column-major 4x4,
            // i.e. column-major linear indices k..k+3; reproduced so files are byte-identical.
            double rec[29];
            const double *st = s->h_state, *dm = s->b->damage;  // dm: (aggressor, victim) row-major
            rec[0] = s->b->clock[0];
            rec[1] = st[0]; rec[2] = st[1];
            rec[3] = st[2]; rec[4] = st[3];
            rec[5] = st[6]; rec[6] = st[7];
            rec[7] = st[10]; rec[8] = st[11];
            rec[9] = st[14]; rec[10] = st[15];
            rec[11] = st[18]; rec[12] = st[19];
            for (int k = 0; k < 4; ++k)
                for (int c = 0; c < 4; ++c) {
                    const int lin = k + c, row = lin % 4, col = lin / 4;
                    rec[13 + 4 * k + c] = dm[row * 4 + col];
                }
            fwrite(rec, sizeof(double), 29, s->log);
            fflush(s->log);
        }
    }
    if (h_out_state22) std::memcpy(h_out_state22, s->h_state, sizeof(s->h_state));
    return RGS_OK;
}

int rgs_single_time(rgs_single *s, double *out_t) {
    if (!s || !out_t) return fail(RGS_EINVAL, "NULL argument");
    *out_t = s->b->clock[0];   // mapped pinned memory; every call of this handle ends with a stream synchronise
    return RGS_OK;
}

int rgs_single_damage(rgs_single *s, double *h_out16) {
    if (!s || !h_out16) return fail(RGS_EINVAL, "NULL argument");
    std::memcpy(h_out16, s->b->damage, 16 * sizeof(double));
    return RGS_OK;
}

}  // extern "C"
