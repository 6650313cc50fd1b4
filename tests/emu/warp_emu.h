// tests/emu/warp_emu.h -- TEST INFRASTRUCTURE ONLY (never on the product path).
//
// Host stand-in for the handful of warp-level CUDA intrinsics the rollout kernels use, so that
// the SAME kernel-body headers (robo_game_sim_b200/csrc/rgs_*.cuh) can be compiled with g++ and
// stepped on the CPU in this GPU-less container.  A "warp" is 32 ucontext fibers scheduled
// round-robin on one OS thread: every collective (__shfl_sync, __ballot_sync, __syncwarp) is a
// barrier, and a barrier is simply "switch to the next lane" -- when lane 0 resumes, all 32 lanes
// have reached the same collective.  A lane that leaves the kernel body while others still wait
// at a collective (divergent collective = a bug on the GPU too) aborts with a message.
//
// Nothing under robo_game_sim_b200/ includes this file; the product has no CPU path.
#pragma once
#define RGS_HOST_EMU 1
#include <ucontext.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>

#define __host__
#define __device__
#define __forceinline__ inline
#define RGS_DEV inline
#define RGS_TRACE(...) do { if (std::getenv("RGS_EMU_TRACE")) std::fprintf(stderr, __VA_ARGS__); } while (0)

namespace emu {

constexpr int kLanes = 32;
constexpr size_t kStack = 256 * 1024;

struct Warp {
    ucontext_t ctx[kLanes];
    ucontext_t main_ctx;
    char *stacks = nullptr;
    bool done[kLanes];
    int cur = 0;
    int n_done = 0;
    uint64_t slot[2][kLanes];
    int parity[kLanes];
    std::function<void(int)> body;
};

inline Warp *&current() {
    static thread_local Warp *w = nullptr;
    return w;
}

inline int lane_id() { return current()->cur; }

inline void barrier() {
    Warp *w = current();
    const int me = w->cur;
    const int nxt = (me + 1) % kLanes;
    if (w->done[nxt]) {
        std::fprintf(stderr, "warp_emu: lane %d waits at a collective but lane %d already left the kernel\n", me, nxt);
        std::abort();
    }
    w->cur = nxt;
    swapcontext(&w->ctx[me], &w->ctx[nxt]);
}

inline void trampoline() {
    Warp *w = current();
    const int me = w->cur;
    w->body(me);
    w->done[me] = true;
    w->n_done++;
    if (w->n_done == kLanes) {
        setcontext(&w->main_ctx);
    }
    const int nxt = (me + 1) % kLanes;
    if (w->done[nxt]) {
        std::fprintf(stderr, "warp_emu: lane %d finished, lane %d finished earlier, but %d lanes still run\n", me, nxt,
                     kLanes - w->n_done);
        std::abort();
    }
    w->cur = nxt;
    setcontext(&w->ctx[nxt]);
}

// Run body(lane) for 32 lanes in lockstep-at-collectives order.
inline void run_warp(const std::function<void(int)> &body) {
    Warp w;
    w.stacks = static_cast<char *>(std::malloc(kStack * kLanes));
    w.body = body;
    for (int l = 0; l < kLanes; ++l) {
        w.done[l] = false;
        w.parity[l] = 0;
        getcontext(&w.ctx[l]);
        w.ctx[l].uc_stack.ss_sp = w.stacks + kStack * l;
        w.ctx[l].uc_stack.ss_size = kStack;
        w.ctx[l].uc_link = nullptr;
        makecontext(&w.ctx[l], reinterpret_cast<void (*)()>(trampoline), 0);
    }
    Warp *prev = current();
    current() = &w;
    w.cur = 0;
    swapcontext(&w.main_ctx, &w.ctx[0]);
    current() = prev;
    std::free(w.stacks);
}

template <typename T>
inline T exchange(T v, int src) {
    static_assert(sizeof(T) <= 8, "emu exchange: <= 8 bytes");
    Warp *w = current();
    const int me = w->cur;
    const int p = w->parity[me];
    uint64_t raw = 0;
    std::memcpy(&raw, &v, sizeof(T));
    w->slot[p][me] = raw;
    barrier();
    T out;
    std::memcpy(&out, &w->slot[p][src & 31], sizeof(T));
    w->parity[me] = p ^ 1;
    return out;
}

}  // namespace emu

// ---- the intrinsics the kernels use (full-mask, warp-uniform call sites only) ----
template <typename T>
inline T __shfl_sync(unsigned, T v, int src) { return emu::exchange(v, src); }
template <typename T>
inline T __shfl_xor_sync(unsigned, T v, int m) { return emu::exchange(v, emu::lane_id() ^ m); }
template <typename T>
inline T __shfl_down_sync(unsigned, T v, int d) {
    const int me = emu::lane_id();
    return emu::exchange(v, me + d < 32 ? me + d : me);
}
template <typename T>
inline T __shfl_up_sync(unsigned, T v, int d) {
    const int me = emu::lane_id();
    return emu::exchange(v, me - d >= 0 ? me - d : me);
}
inline unsigned __ballot_sync(unsigned, bool pred) {
    emu::Warp *w = emu::current();
    const int me = w->cur;
    const int p = w->parity[me];
    w->slot[p][me] = pred ? 1u : 0u;
    emu::barrier();
    unsigned m = 0;
    for (int l = 0; l < 32; ++l) m |= (unsigned)(w->slot[p][l] & 1u) << l;
    w->parity[me] = p ^ 1;
    return m;
}
inline bool __any_sync(unsigned mask, bool pred) { return __ballot_sync(mask, pred) != 0; }
inline bool __all_sync(unsigned mask, bool pred) { return __ballot_sync(mask, pred) == 0xffffffffu; }
inline void __syncwarp(unsigned = 0xffffffffu) { emu::barrier(); }
inline int __ffs(unsigned v) { return v ? __builtin_ctz(v) + 1 : 0; }
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
