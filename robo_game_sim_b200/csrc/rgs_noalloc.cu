// rgs_noalloc.cu -- rgs_rollout_noalloc (include/rgs_capi.h): the K2 fp32 2v2 rollout for callers that cannot
// allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart
// in the reference.
//
// The kernel is the small-batch variant of rollout_lane_kernel<float, 2, 32> (rgs_lane_launch.cuh): one-warp CTAs
// carrying a few games each, the same rollout_lane_game body.  What differs is where the per-lane global scratch
// (LaneScratch::SCR_WORDS words per lane: correctOverlap tracks, event counters, jam snapshot) lives: in the
// file-scope static __device__ array below, indexed per GAME slot (block * lanes + lane) so that
// RGS_NOALLOC_MAX_GAMES lanes' worth is all it ever needs, instead of in a cudaMalloc'ed buffer sized for the
// resident grid.  Every device that loads the library holds the array (11.1 MB) until the process exits.
//
// In host code `g_noalloc_scratch` is only the host-side stand-in of the symbol, not a device address: the address
// the kernel gets is looked up with cudaGetSymbolAddress, once per device, with that device current.
//
// All calls on a device share the array.  Nothing in it has to be zero at launch -- a lane initialises every word it
// later reads (event counters at the start of its game, tracks / checksums / jam snapshot when the event path
// first writes them; the cudaMalloc'ed scratch of the other entry points is never cleared either) -- but two calls
// that overlap on the GPU race on it.
#include <atomic>
#include <cstdio>
#include <mutex>

#include "rgs_lane.cuh"
#include "rgs_host_params.h"
#include "rgs_kernels.h"

namespace rgs {
namespace {

using ScratchLayout = LaneScratch<float, 2>;
constexpr int kNoAllocThreads = 32;   // one warp per CTA, like the small-batch path of launch_lane_nt
constexpr size_t kNoAllocWords = (size_t)RGS_NOALLOC_MAX_GAMES * ScratchLayout::SCR_WORDS;
static_assert(sizeof(float) * ScratchLayout::SCR_WORDS == 2708, "include/rgs_capi.h states 2708 bytes of scratch per game");

static __device__ __align__(256) float g_noalloc_scratch[kNoAllocWords];

__global__ void __launch_bounds__(kNoAllocThreads, 512 / kNoAllocThreads)
rollout_lane_static_scratch_kernel(const __grid_constant__ FastParams<float> P, const __grid_constant__ RunArgs A,
                                   const int lanes) {
    using S = ScratchLayout;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LaneTab<float> *tab = reinterpret_cast<LaneTab<float> *>(smem_raw);
    constexpr size_t tab_bytes = (sizeof(LaneTab<float>) + 15) & ~size_t(15);
    float *fw = reinterpret_cast<float *>(smem_raw + tab_bytes);
    int *iw = reinterpret_cast<int *>(fw + (size_t)S::NF * kNoAllocThreads);
    for (int i = threadIdx.x; i < 2 * kCols; i += kNoAllocThreads) {
        (&tab->e[0][0])[i].g = (&P.tab_g[0][0])[i];
        (&tab->e[0][0])[i].k = (&P.tab_k[0][0])[i];
    }
    __syncthreads();
    const int64_t game = (int64_t)blockIdx.x * lanes + threadIdx.x;
    if ((int)threadIdx.x >= lanes || game >= A.n_games) return;
    // one scratch slot per game: game < RGS_NOALLOC_MAX_GAMES (checked by the entry point), so the slot is inside
    // the static array whatever the grid looks like
    float *scr = static_cast<float *>(A.lane_scratch) + (size_t)game * S::SCR_WORDS;
    const S s{fw + threadIdx.x, iw + threadIdx.x, kNoAllocThreads, scr};
    rollout_lane_game<float, 2, false, false>(P, A, tab, s, game);
}

constexpr size_t noalloc_smem_bytes() {
    return ((sizeof(LaneTab<float>) + 15) & ~size_t(15)) +
           ((size_t)ScratchLayout::NF * sizeof(float) + (size_t)ScratchLayout::NI * sizeof(int)) * kNoAllocThreads;
}

constexpr int kMaxDevices = 64;
struct DeviceSetup {
    float *scratch = nullptr;   // device address of g_noalloc_scratch on this device
    int sm_count = 0;
};
DeviceSetup g_setup[kMaxDevices];
std::atomic<bool> g_ready[kMaxDevices];
std::mutex g_setup_mutex;

// once per device, with that device current: the symbol's device address, the kernel's shared-memory limit
int setup_device(int device, const DeviceSetup **out) {
    if (!g_ready[device].load(std::memory_order_acquire)) {
        std::lock_guard<std::mutex> lock(g_setup_mutex);
        if (!g_ready[device].load(std::memory_order_relaxed)) {
            DeviceSetup d;
            void *ptr = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
            if (e != cudaSuccess) return capi_cuda_fail(e, "cudaGetSymbolAddress(static lane scratch)");
            if (!ptr) return capi_fail(RGS_ECUDA, "cudaGetSymbolAddress returned a null address for the static lane scratch");
            d.scratch = static_cast<float *>(ptr);
            e = cudaFuncSetAttribute(rollout_lane_static_scratch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)noalloc_smem_bytes());
            if (e != cudaSuccess) return capi_cuda_fail(e, "cudaFuncSetAttribute(static-scratch rollout kernel)");
            e = cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device);
            if (e != cudaSuccess) return capi_cuda_fail(e, "cudaDeviceGetAttribute(multiprocessor count)");
            g_setup[device] = d;
            g_ready[device].store(true, std::memory_order_release);
        }
    }
    *out = &g_setup[device];
    return RGS_OK;
}

struct CurrentDevice {   // make `dev` current for the call, restore the caller's device afterwards
    int prev = -1;
    bool ok = true;
    explicit CurrentDevice(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; prev = -1; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~CurrentDevice() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace
}  // namespace rgs

extern "C" __attribute__((visibility("default")))
int rgs_rollout_noalloc(int device, const rgs_params *p, int64_t n_games, double *d_state, double *d_clock,
                        double *d_damage, int64_t *d_events, const float *d_actions, int64_t n_act, int64_t hold,
                        int64_t act_game_stride, int64_t n_steps, void *stream) {
    using namespace rgs;
    int rc = capi_check_params(p);
    if (rc != RGS_OK) return rc;
    if (p->mode != RGS_MODE_FAST_F32 || p->players_per_team != 2 || p->noise != 0.0)
        return capi_fail(RGS_EUNSUPPORTED, "rgs_rollout_noalloc runs RGS_MODE_FAST_F32, players_per_team == 2, noise == 0 only "
                                           "(use an rgs_batch for the other modes)");
    if (n_games < 0) return capi_fail(RGS_EINVAL, "n_games must be >= 0");
    if (n_games > RGS_NOALLOC_MAX_GAMES) {
        char msg[200];
        snprintf(msg, sizeof(msg), "rgs_rollout_noalloc: n_games = %lld exceeds RGS_NOALLOC_MAX_GAMES = %d, the size of the "
                 "static scratch array (use an rgs_batch for larger batches)", (long long)n_games, RGS_NOALLOC_MAX_GAMES);
        return capi_fail(RGS_EINVAL, msg);
    }
    if (n_steps < 0) return capi_fail(RGS_EINVAL, "n_steps must be >= 0");
    if (hold < 1) return capi_fail(RGS_EINVAL, "hold must be >= 1");
    if (n_act < 1) return capi_fail(RGS_EINVAL, "n_act must be >= 1");
    if (act_game_stride < 0) return capi_fail(RGS_EINVAL, "act_game_stride must be >= 0");
    if (device < 0 || device >= kMaxDevices) return capi_fail(RGS_EINVAL, "device out of range");
    if (n_games == 0 || n_steps == 0) return RGS_OK;
    if (!d_state || !d_clock || !d_damage || !d_events || !d_actions)
        return capi_fail(RGS_EINVAL, "d_state, d_clock, d_damage, d_events and d_actions must all be given");

    CurrentDevice cur(device);
    if (!cur.ok) return capi_fail(RGS_ECUDA, "cudaSetDevice failed");
    const DeviceSetup *dev = nullptr;
    rc = setup_device(device, &dev);
    if (rc != RGS_OK) return rc;

    RunArgs A{};
    A.n_games = n_games;
    A.state = d_state;
    A.clock = d_clock;
    A.damage = d_damage;
    A.events = reinterpret_cast<long long *>(d_events);
    A.actions = d_actions;
    A.n_act = n_act;
    A.hold = hold;
    A.act_game_stride = act_game_stride;
    A.n_steps = n_steps;
    A.lane_scratch = dev->scratch;
    A.lane_scratch_bytes = kNoAllocWords * sizeof(float);
    const FastParams<float> P = make_fast_params<float>(*p);
    // games per one-warp CTA: the small-batch rule of launch_lane_nt (aim at kSmallWarpsPerSm = 7 warps per SM)
    const int64_t warps = (int64_t)dev->sm_count * 7;
    int lanes = (int)((n_games + warps - 1) / warps);
    if (lanes < 1) lanes = 1;
    if (lanes > kNoAllocThreads) lanes = kNoAllocThreads;
    const unsigned blocks = (unsigned)((n_games + lanes - 1) / lanes);
    rollout_lane_static_scratch_kernel<<<blocks, kNoAllocThreads, noalloc_smem_bytes(), static_cast<cudaStream_t>(stream)>>>(P, A, lanes);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return capi_cuda_fail(e, "static-scratch rollout kernel launch");
    capi_count_launches(1);
    return RGS_OK;
}
