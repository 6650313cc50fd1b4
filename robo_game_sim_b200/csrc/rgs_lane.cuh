// rgs_lane.cuh -- K2 (lane body): fused multi-step rollout in the THROUGHPUT formulation,
// ONE LANE PER GAME.
//
// Why not a warp per game here (K1 keeps that mapping): the first B200 profile of the warp-per-game
// throughput kernel (4.7e8 game-steps/s, profiles/r01_bench_first.json) showed 666 warp-instructions per game-step,
// 45 % of them in a per-column broad phase that finds nothing in 3 steps out of 4, the rest in a
// serial event path in which 31 lanes idle, and the issue slots 72 % empty on instruction-fetch
// stalls (78 KB of divergent code).  Games are independent, so the natural SIMT mapping of a huge
// Monte Carlo batch is game -> lane: every arithmetic instruction does 32 games' worth of work, the
// broad phase becomes a handful of bounding-interval tests per game-step, and only the games that
// can touch something enter the narrow phase.
//
// Formulation (same event semantics as the reference step, GameSim::updateSim, reference
// src/GameSim.cpp:126-171: overlap detected at the 50 grid columns, linear back-tracking from
// column i-1, per-owner events, earliest-first resolution with the stable-sort tie rule, prune +
// re-scan, one goal per step):
//   * time is step-relative; the 50-sub-step RK4 chain (:173-217, :70-116) is replaced by its
//     closed form (see rgs_fast.cuh): i sub-steps are the affine map v_i = u + (v0-u) R^i,
//     p_i = p0 + u i h + (v0-u) k_i, so NO grid is stored: a body's trajectory inside the step is a
//     SEGMENT {start column, state at the start column, state at the column before it} plus the
//     position offsets that correctOverlap (:581-623) added after the start;
//   * broad phase: the velocity of a segment is monotone between its end values, so the position
//     stays inside [p + min(0, vmin T), p + max(0, vmax T)]; bodies whose intervals (inflated by the
//     radii) cannot meet a wall or each other are never column-tested.  The bound is conservative,
//     so the events found are the ones the full column scan (insertCollisions :360-511) finds;
//   * narrow phase: flagged (body, wall) and (body, body) ITEMS are searched over the columns: a wall
//     item by bisection on the (convex or concave) coordinate along the wall's axis, a pair item by
//     conservative advancement (columns that the relative motion proves contact-free are skipped);
//     an item's first column with a valid time of impact is folded into its owner bodies with the
//     reference's candidate order (walls UP, DOWN, RIGHT, LEFT, then the other bodies in entity
//     order; strict '<' keeps the first);
//   * events are resolved exactly like K2's warp formulation did (rgs_fast.cuh history): advance
//     to the impact time with the RK4 polynomial, 1.01 overlap push, elastic impulse, damage,
//     advance to the next column, correctOverlap on every later column.
// correctOverlap offsets are CUMULATIVE in the reference (a column is integrated from the corrected previous
// column, :292-297, :336-340), so a body pressed against a wall or another body is corrected on every column.  The
// segment keeps the total offset (in force from the column of the last fix on), the column of the first fix (no
// offset before it) and, only when a segment has been fixed more than once, the cumulative offset of the columns in
// between in a per-lane TRACK in global memory (LaneScratch::trk; contiguous per lane, touched by such bodies only).
//
// Everything is plain scalar C++ over a per-game scratch record, so tests/emu compiles the same
// source with g++ (no warp emulation needed).  All fused multiply-adds are spelled out (fm()).
#pragma once
#include "rgs_fast.cuh"
#include "rgs_policy.cuh"

// The rollout kernels are instruction-cache-bound before they are issue-bound (profiles/): the large,
// rarely-converged narrow-phase and resolution functions are compiled once per kernel and CALLED.
#if defined(RGS_NOINLINE_BIG) && !defined(RGS_HOST_EMU)
#define RGS_DEV_BIG __device__ __noinline__
#else
#define RGS_DEV_BIG RGS_DEV
#endif

namespace rgs {

constexpr int kJamLog = 6;   // longest period of a jammed event loop that is fast-forwarded (LaneGame::JamSnap)

// per-game scratch record: word w of game slot s lives at base[w * stride + s] (bank = lane)
template <typename T, int NT>
struct LaneScratch {
    static constexpr int E = 2 * NT + 1;
    // T words
    static constexpr int SP = 0;         // [4e+c] state at the segment start column
    static constexpr int CMD = 4 * E;    // [2e+a] commanded velocity (puck: 0)
    static constexpr int PRV = 6 * E;    // [4e+c] state at the column before the segment start
    static constexpr int OFF = 10 * E;   // [2e+a] total correctOverlap offset of the segment (in force from column FL on)
    static constexpr int BB = 12 * E;    // [4e+{0..3}] bounding interval lo_x, hi_x, lo_y, hi_y of the segment
    static constexpr int QT = 16 * E;    // [e] queued / candidate event time of owner e
    static constexpr int NF = 17 * E;
    // int words
    static constexpr int CS = 0;         // [e] segment start column
    static constexpr int FL = E;         // [e] column of the last correctOverlap fix (= CS when none)
    static constexpr int QI = 2 * E;     // [e] (seq << 12) | (col << 6)... see q_pack
    static constexpr int FF = 3 * E;     // [e] column of the first correctOverlap fix (= CS when none)
    static constexpr int NI = 4 * E;
    // per-lane global scratch area, in T-sized words, contiguous per lane (the event path touches it from divergent
    // code, one lane at a time and mostly column after column, so lane-major keeps its sectors dense): the tracks; the
    // event counters of
    // this launch (RGS_EV_* order; [RGS_EV_HASH], [16] = hash lo, hi) and the per-body checksum of the track words
    // written for the segment (jam detection) as ints; then the jam snapshot (LaneGame::JamSnap).  All of it is touched
    // by the event path only -- the shared-memory record above is sized for 4 resident 128-lane CTAs per SM.
    static constexpr int TRK_WORDS = 2 * E * kCols;
    static constexpr int EVC = TRK_WORDS, TS = EVC + 17;
    static constexpr int JAM0 = TS + E;
    static constexpr int JAM_F = QT + E, JAM_I = 5 * E + 17 + 3 * kJamLog;
    static constexpr int SCR_WORDS = JAM0 + JAM_F + JAM_I;
    // process-noise kernels only: accumulated noise (dpx, dpy, dvx, dvy) since the segment start of player e at column c
    static constexpr int NZ0 = SCR_WORDS, NZ_WORDS = 4 * (E - 1) * kCols;
    static constexpr int SCR_WORDS_NOISE = SCR_WORDS + NZ_WORDS;
    T *f;
    int *i;
    int stride;
    T *trk;        // this lane's global scratch area; cumulative offset of column c, axis a of body e at trk[(e * kCols + c) * 2 + a]
    RGS_DEV T &TR(int e, int c, int a) const { return trk[(e * kCols + c) * 2 + a]; }
    RGS_DEV T &F(int w) const { return f[w * stride]; }
    RGS_DEV int &I(int w) const { return i[w * stride]; }
    RGS_DEV int &G(int w) const { return *reinterpret_cast<int *>(&trk[w]); }   // int word of the global area
};

// index of the lowest set bit of a non-zero mask
RGS_DEV int low_bit(unsigned m) {
#ifdef RGS_HOST_EMU
    return __builtin_ctz(m);
#else
    return __ffs((int)m) - 1;
#endif
}

template <typename T>
struct alignas(2 * sizeof(T)) LaneGK { T g, k; };
template <typename T>
struct LaneTab {   // block-shared coefficient table, [type][j] -> {R^j, a (1-R^j)/(1-R)}: one 64/128-bit load
    LaneGK<T> e[2][kCols];
};

// small fixed-size bitset in registers
template <int NBITS>
struct Bits {
    static constexpr int NW = (NBITS + 31) / 32;
    unsigned w[NW];
    RGS_DEV void clear() {
#pragma unroll
        for (int k = 0; k < NW; ++k) w[k] = 0u;
    }
    RGS_DEV void set(int b) {
#pragma unroll
        for (int k = 0; k < NW; ++k)
            if ((b >> 5) == k) w[k] |= 1u << (b & 31);
    }
    RGS_DEV void set_if(bool p, int b) {   // straight-line form of `if (p) set(b)`
#pragma unroll
        for (int k = 0; k < NW; ++k) w[k] |= (p && (b >> 5) == k) ? 1u << (b & 31) : 0u;
    }
    RGS_DEV bool any() const {
        unsigned a = 0u;
#pragma unroll
        for (int k = 0; k < NW; ++k) a |= w[k];
        return a != 0u;
    }
    // index of the lowest set bit (removed), -1 when empty
    RGS_DEV int pop() {
        int r = -1;
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            if (r < 0 && w[k] != 0u) {
#ifdef RGS_HOST_EMU
                const int b = __builtin_ctz(w[k]);
#else
                const int b = __ffs((int)w[k]) - 1;
#endif
                w[k] &= w[k] - 1u;
                r = 32 * k + b;
            }
        }
        return r;
    }
};

// one FMNMX instead of compare + select (+8 % on B200: the kernel is instruction-fetch-bound); differs from
// the a < b ? a : b form only in which zero it returns for (+0, -0), which no caller can observe
#if !defined(RGS_NO_FMNMX) && !defined(RGS_HOST_EMU)
RGS_DEV float tmin(float a, float b) { return fminf(a, b); }
RGS_DEV float tmax(float a, float b) { return fmaxf(a, b); }
RGS_DEV double tmin(double a, double b) { return fmin(a, b); }
RGS_DEV double tmax(double a, double b) { return fmax(a, b); }
#else
template <typename T>
RGS_DEV T tmin(T a, T b) { return a < b ? a : b; }
template <typename T>
RGS_DEV T tmax(T a, T b) { return a > b ? a : b; }
#endif

// square root used ONLY to decide how many columns a search may skip (conservative advancement keeps a 2 %
// margin, so one MUFU.SQRT is enough there; times of impact and overlaps use the IEEE sqrt)
#if !defined(RGS_HOST_EMU) && !defined(RGS_NO_APPROX_SKIP_SQRT)
RGS_DEV float skip_sqrt(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
RGS_DEV double skip_sqrt(double x) { return sqrt(x); }
#else
template <typename T>
RGS_DEV T skip_sqrt(T x) { return T(sqrt(x)); }
#endif

// v = *ptr if p (a PREDICATED load: no branch, no reconvergence region -- the track look-ups sit in every position
// evaluation of the event path and are almost never taken; as branches they cost 15 % of the kernel, profiles/r02_ab_variants.md)
#ifdef RGS_HOST_EMU
template <typename T>
inline void ld_if(bool p, const T *ptr, T &v) { if (p) v = *ptr; }
#else
RGS_DEV void ld_if(bool p, const float *ptr, float &v) {
    asm("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q ld.global.f32 %0, [%1]; }" : "+f"(v) : "l"(ptr), "r"((int)p));
}
RGS_DEV void ld_if(bool p, const double *ptr, double &v) {
    asm("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q ld.global.f64 %0, [%1]; }" : "+d"(v) : "l"(ptr), "r"((int)p));
}
#endif

// broad-phase slack: the bounding intervals are evaluated with different roundings than the
// column positions, so thresholds are inflated by a few hundred ulps of the arena size
template <typename T> struct Slack;
template <> struct Slack<float> { static constexpr float v = 2.0e-4f; };
template <> struct Slack<double> { static constexpr double v = 1.0e-9; };

template <typename T, int NT, bool kNoise = false>
struct LaneGame {
    static constexpr int E = 2 * NT + 1, PK = 2 * NT, NP = 2 * NT, NPAIR = E * (E - 1) / 2;
    using S = LaneScratch<T, NT>;
    const FastParams<T> &P;
    const LaneTab<T> *tab;
    S s;

    RGS_DEV LaneGame(const FastParams<T> &P_, const LaneTab<T> *tab_, const S &s_) : P(P_), tab(tab_), s(s_) {}

    // Event times.  The fp32 throughput mode works in step-relative time (t_ = 0 at the start of every step).  The
    // fp64 instantiation exists to validate this formulation against the oracle, so it keeps the reference's
    // ABSOLUTE clock arithmetic (T[i] = t_ + dt_col_ i :201, t_col - t_ < dt_ :501, base_idx :237, dt_t :254):
    // for contacts resting on a grid time the rounding of t_ decides between two columns.
    static constexpr bool kAbsClock = sizeof(T) == 8;
    // One inline copy of the post-event broad phase (update_bbox / flag_walls / reflag_pairs) and of fold, in rolled
    // two-trip loops, instead of one copy per body of the event: the kernels of the small teams are bound by the
    // fetch of a hot code just above the instruction cache (+3 - 5 % for 1v1 .. 4v4); for 7v7, whose code is far above it
    // anyway, the rolled form costs 7 % (profiles/r02_ab_variants.md section 9).
#ifndef RGS_AB_ONECOPY_MAXNT
#define RGS_AB_ONECOPY_MAXNT 4
#endif
    static constexpr bool kOneCopy = NT <= RGS_AB_ONECOPY_MAXNT;
    mutable T t0 = T(0);
    mutable double tclk = 0.0;   // the game clock t_ at the start of the step (fp64, t_ += dt_ like the reference)
    RGS_DEV T col_time(int c) const {
        if constexpr (kAbsClock) return t0 + P.d.dt_col * T(c);
        else return P.d.dt_col * T(c);
    }
    RGS_DEV T since_t0(T t) const {
        if constexpr (kAbsClock) return t - t0;
        else return t;
    }
    RGS_DEV bool in_step(T t) const { return since_t0(t) < P.d.dt; }

    // ---- process noise (kNoise instantiations; reference src/GameSim.cpp:62-68, see rgs_fast.cuh NoiseCoef) ----
    // Every segment of a player carries its own accumulated-noise path NZ(e, c, .) over the columns after its start
    // (the reference draws fresh noise whenever it re-integrates a body, :285-297, :329-340).  A step without
    // broad-phase flags needs only the end of the path, drawn directly from the accumulated 2-D Gaussian.
    mutable long long nz_game = 0;        // global game index
    mutable unsigned nz_step = 0, nz_epoch = 0;
    mutable bool nz_on = false;           // false while a quiet step commits from the closed form
    // Players whose noise path of the current segment exists.  Paths are generated LAZILY: a flagged step needs the
    // path of a player only if a narrow-phase item or an event evaluates it between the step's ends; every other
    // player gets the end of its path at the commit, drawn from the accumulated Gaussian like in a quiet step (the
    // same distribution: the players' noises are independent).  Generating all 2n paths for every flagged step cost
    // 50 Philox blocks per player and step.
    mutable unsigned nz_have = 0;
    RGS_DEV void ensure_path(int e) const {
        if (e != PK && !((nz_have >> e) & 1u)) {
            gen_noise_path(e, s.I(S::CS + e));   // (no event has touched e yet: its segment is the whole step, CS = 0)
            nz_have |= 1u << e;
        }
    }
    // end-of-step noise of player e from the accumulated Gaussian of the 50 sub-steps
    RGS_DEV void add_end_noise(int e) const {
        T g[4];
        normals4<T>(P.seed_lo, P.seed_hi, nz_game, nz_step, (unsigned)e, g);
        s.F(S::SP + 4 * e) += fm(P.nz_end[1], g[0], P.nz_end[2] * g[1]);
        s.F(S::SP + 4 * e + 1) += fm(P.nz_end[1], g[2], P.nz_end[2] * g[3]);
        s.F(S::SP + 4 * e + 2) += P.nz_end[0] * g[0];
        s.F(S::SP + 4 * e + 3) += P.nz_end[0] * g[2];
    }
    RGS_DEV T &NZ(int e, int c, int k) const { return s.trk[S::NZ0 + (e * kCols + c) * 4 + k]; }
    RGS_DEV void gen_noise_path(int e, int c0) const {
        T npx = T(0), npy = T(0), nvx = T(0), nvy = T(0);
        NZ(e, c0, 0) = T(0); NZ(e, c0, 1) = T(0); NZ(e, c0, 2) = T(0); NZ(e, c0, 3) = T(0);
        const T R = tab->e[0][1].g, k1 = tab->e[0][1].k;
        // One Philox block keyed by (seed, game, step, epoch, player, first column) seeds a xorshift128 stream for the
        // columns of THIS path (Marsaglia 2003; four words per column): the paths are regenerated by every event, a
        // few lanes at a time, and a Philox block per column was 41 % of the noise kernel's instructions.  Still a pure
        // function of the key, so a sharded batch draws what the unsharded batch draws.
        uint32_t xs[4];
        philox4x32_10((uint32_t)nz_game, (uint32_t)((unsigned long long)nz_game >> 32), nz_step,
                      ((nz_epoch + 1u) << 16) | ((unsigned)e << 8) | (unsigned)c0, P.seed_lo, P.seed_hi ^ 0x4E5A4E5Au, xs);
        uint32_t x = xs[0], y = xs[1], z = xs[2], w = xs[3] | 1u;
#pragma unroll 1
        for (int c = c0 + 1; c <= kGrid; ++c) {
            uint32_t o[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t t = x ^ (x << 11);
                x = y; y = z; z = w;
                w = w ^ (w >> 19) ^ t ^ (t >> 8);
                o[k] = w;
            }
            T g[4];
            box_muller(o[0], o[1], g[0], g[1]);
            box_muller(o[2], o[3], g[2], g[3]);
            npx = npx + k1 * nvx + fm(P.nz_sub[1], g[0], P.nz_sub[2] * g[1]);
            npy = npy + k1 * nvy + fm(P.nz_sub[1], g[2], P.nz_sub[2] * g[3]);
            nvx = fm(R, nvx, P.nz_sub[0] * g[0]);
            nvy = fm(R, nvy, P.nz_sub[0] * g[2]);
            NZ(e, c, 0) = npx; NZ(e, c, 1) = npy; NZ(e, c, 2) = nvx; NZ(e, c, 3) = nvy;
        }
    }
    // noise of one RK4 call of arbitrary length h (the two pieces around a time of impact, :255-256, :287-288)
    RGS_DEV void noise_kick(int e, int piece, T h, T &px, T &py, T &vx, T &vy) const {
        if (!(h > T(0))) return;
        const NoiseCoef<T> q = noise_coef<T>(h, P.neg_inv_tau[0]);
        T g[4];
        normals4<T>(P.seed_lo, P.seed_hi, nz_game, nz_step, ((nz_epoch + 1u) << 16) | ((unsigned)e << 8) | (0xF0u + (unsigned)piece), g);
        px += P.noise * fm(q.l21, g[0], q.l22 * g[1]);
        py += P.noise * fm(q.l21, g[2], q.l22 * g[3]);
        vx += P.noise * (q.l11 * g[0]);
        vy += P.noise * (q.l11 * g[2]);
    }

    // cumulative correctOverlap offset of body e at column c > CS[e]
    RGS_DEV void off_at(int e, int c, T &ox, T &oy) const {
        const bool late = c >= s.I(S::FL + e);
        ox = late ? s.F(S::OFF + 2 * e) : T(0);
        oy = late ? s.F(S::OFF + 2 * e + 1) : T(0);
        const bool mid = !late && c >= s.I(S::FF + e);
        const T *tp = &s.TR(e, c, 0);
        ld_if(mid, tp, ox);
        ld_if(mid, tp + 1, oy);
    }
    RGS_DEV T off_at1(int e, int a, int c) const {
        if (c >= s.I(S::FL + e)) return s.F(S::OFF + 2 * e + a);
        if (c < s.I(S::FF + e)) return T(0);
        return s.TR(e, c, a);
    }
    static RGS_DEV int word_bits(float x) { int b; memcpy(&b, &x, 4); return b; }
    static RGS_DEV int word_bits(double x) { long long b; memcpy(&b, &x, 8); return (int)(b ^ (b >> 32)); }
    // correctOverlap moved body e by (fx, fy) at column c > FL[e]: the offset in force so far stays the offset of
    // the columns FL[e] .. c-1 (written to the track unless this is the segment's first fix)
    RGS_DEV void record_fix(int e, int c, T fx, T fy) const {
        const int fl = s.I(S::FL + e);
        const T ox = s.F(S::OFF + 2 * e), oy = s.F(S::OFF + 2 * e + 1);
        if (fl == s.I(S::CS + e)) {
            s.I(S::FF + e) = c;
            s.G(S::TS + e) = 0;   // (the checksum only means something once the segment has been fixed twice)
        } else {
            int ts = s.G(S::TS + e);
#pragma unroll 1
            for (int k = fl; k < c; ++k) {
                s.TR(e, k, 0) = ox; s.TR(e, k, 1) = oy;
                ts = ts * 31 + (word_bits(ox) ^ (word_bits(oy) * 7) ^ k);
            }
            s.G(S::TS + e) = ts;
        }
        s.F(S::OFF + 2 * e) = ox + fx; s.F(S::OFF + 2 * e + 1) = oy + fy;
        s.I(S::FL + e) = c;
    }

    // event record of the launch: 32-bit counters (RGS_EV_* order; [RGS_EV_HASH], [16] = hash lo, hi), added to the game's
    // int64 record at the end
    RGS_DEV int &EVW(int k) const { return s.G(S::EVC + k); }
    RGS_DEV void ev_add(int k, int n = 1) const { EVW(k) += n; }
    mutable unsigned long long ev_hash = 0;   // the order-sensitive event hash of the game (RGS_EV_HASH), carried in registers
    RGS_DEV void hash_ev(long long live, int kind, int i, int j, int goal) const {
        unsigned long long h = ev_hash;
        h ^= ((unsigned long long)live << 16) | ((unsigned long long)kind << 12) | ((unsigned long long)i << 8) |
             ((unsigned long long)(j < 0 ? 15 : j) << 4) | (unsigned long long)goal;   // == hash_event() of rgs_device.cuh
        h *= 0x100000001B3ULL;
        ev_hash = h;
    }

    // state of body e at column c >= CS[e]-1 (position includes the correctOverlap offsets)
    template <bool kEnd = false>   // kEnd: c = column 50 >= FL[e] of a live segment (the commit): no look-ups, no PRV case
    RGS_DEV void eval(int e, int c, T &px, T &py, T &vx, T &vy) const {
        const int j = c - s.I(S::CS + e);
        if constexpr (!kEnd) {
            if (j < 0) {
                px = s.F(S::PRV + 4 * e); py = s.F(S::PRV + 4 * e + 1); vx = s.F(S::PRV + 4 * e + 2); vy = s.F(S::PRV + 4 * e + 3);
                return;
            }
        }
        const T spx = s.F(S::SP + 4 * e), spy = s.F(S::SP + 4 * e + 1), svx = s.F(S::SP + 4 * e + 2), svy = s.F(S::SP + 4 * e + 3);
        if (j == 0) { px = spx; py = spy; vx = svx; vy = svy; return; }
        const int ty = (e == PK);
        const T ux = s.F(S::CMD + 2 * e), uy = s.F(S::CMD + 2 * e + 1);
        const LaneGK<T> gk = tab->e[ty][j];
        const T g = gk.g, k = gk.k, tj = P.d.dt_col * T(j);
        const T wx = svx - ux, wy = svy - uy;
        T ox, oy;
        if constexpr (kEnd) { ox = s.F(S::OFF + 2 * e); oy = s.F(S::OFF + 2 * e + 1); }
        else off_at(e, c, ox, oy);
        px = fm(wx, k, fm(ux, tj, spx)) + ox;
        py = fm(wy, k, fm(uy, tj, spy)) + oy;
        vx = fm(wx, g, ux);
        vy = fm(wy, g, uy);
        if constexpr (kNoise)
            if (nz_on && ((nz_have >> e) & 1u)) { px += NZ(e, c, 0); py += NZ(e, c, 1); vx += NZ(e, c, 2); vy += NZ(e, c, 3); }
    }

    // start a new segment of body e at column c with state (px,py,vx,vy); prev = state at c-1
    RGS_DEV void start_segment(int e, int c, T px, T py, T vx, T vy, T qx, T qy, T qvx, T qvy) const {
        s.F(S::PRV + 4 * e) = qx; s.F(S::PRV + 4 * e + 1) = qy; s.F(S::PRV + 4 * e + 2) = qvx; s.F(S::PRV + 4 * e + 3) = qvy;
        s.F(S::SP + 4 * e) = px; s.F(S::SP + 4 * e + 1) = py; s.F(S::SP + 4 * e + 2) = vx; s.F(S::SP + 4 * e + 3) = vy;
        s.F(S::OFF + 2 * e) = T(0); s.F(S::OFF + 2 * e + 1) = T(0);
        s.I(S::CS + e) = c;
        s.I(S::FL + e) = c;
        s.I(S::FF + e) = c;
        if constexpr (kNoise)
            if (e != PK) { gen_noise_path(e, c); nz_have |= 1u << e; }
    }

    // bounding interval of body e over columns CS[e]..50 (monotone-velocity bound + offsets)
    RGS_DEV void update_bbox(int e) const {
        const int ty = (e == PK);
        const int cs = s.I(S::CS + e);
        const int rem = kGrid - cs;
        const T trem = P.d.dt_col * T(rem), gE = tab->e[ty][rem].g;
        const int ff = s.I(S::FF + e), fl = s.I(S::FL + e);
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const T p = s.F(S::SP + 4 * e + a), v = s.F(S::SP + 4 * e + 2 + a), u = s.F(S::CMD + 2 * e + a);
            const T vE = fm(v - u, gE, u);
            const T o1 = s.F(S::OFF + 2 * e + a);
            T olo = tmin(T(0), o1), ohi = tmax(T(0), o1);
#pragma unroll 1
            for (int k = ff; k < fl; ++k) {   // a segment fixed more than once: hull of its cumulative offsets
                const T o = s.TR(e, k, a);
                olo = tmin(olo, o); ohi = tmax(ohi, o);
            }
            const T env = (kNoise && e != PK) ? P.nz_env : T(0);
            const T lo = p + tmin(T(0), tmin(v, vE) * trem) + olo - env;
            const T hi = p + tmax(T(0), tmax(v, vE) * trem) + ohi + env;
            s.F(S::BB + 4 * e + 2 * a) = lo;
            s.F(S::BB + 4 * e + 2 * a + 1) = hi;
        }
    }

    static RGS_DEV int pair_index(int a, int b) {  // a < b
        return a * (2 * E - a - 1) / 2 + (b - a - 1);
    }
    static RGS_DEV void pair_decode(int idx, int &a, int &b) {
        a = 0;
#pragma unroll 1
        while (idx >= E - 1 - a) { idx -= E - 1 - a; ++a; }
        b = a + 1 + idx;
    }

    // broad-phase flags of body e against the four walls (bits 4e + wall)
    RGS_DEV void flag_walls(int e, Bits<4 * E> &wb) const {
        const int ty = (e == PK);
        const T sl = Slack<T>::v;
        const T lx = P.lim_x[ty] - sl, ly = P.lim_y[ty] - sl;
        wb.set_if(s.F(S::BB + 4 * e + 3) >= ly, 4 * e + WALL_UP);
        wb.set_if(s.F(S::BB + 4 * e + 2) <= -ly, 4 * e + WALL_DOWN);
        wb.set_if(s.F(S::BB + 4 * e + 1) >= lx, 4 * e + WALL_RIGHT);
        wb.set_if(s.F(S::BB + 4 * e + 0) <= -lx, 4 * e + WALL_LEFT);
    }
    // can the bounding intervals of a and b come within r_a + r_b of each other?
    RGS_DEV bool flag_pair(int a, int b) const {
        const T gx = tmax(T(0), tmax(s.F(S::BB + 4 * a) - s.F(S::BB + 4 * b + 1), s.F(S::BB + 4 * b) - s.F(S::BB + 4 * a + 1)));
        const T gy = tmax(T(0), tmax(s.F(S::BB + 4 * a + 2) - s.F(S::BB + 4 * b + 3), s.F(S::BB + 4 * b + 2) - s.F(S::BB + 4 * a + 3)));
        const T sg = P.d.sig[a == PK][b == PK] + Slack<T>::v;
        return fm(gx, gx, gy * gy) <= sg * sg;
    }

    // re-scan of body e against every other body except `skip` (its partner is flagged by its own call when
    // e < partner): flag_pair with e's interval held in registers
    RGS_DEV void reflag_pairs(int e, int skip, Bits<NPAIR> &pb) const {
        const T lx = s.F(S::BB + 4 * e), hx = s.F(S::BB + 4 * e + 1), ly = s.F(S::BB + 4 * e + 2), hy = s.F(S::BB + 4 * e + 3);
        const T sl = Slack<T>::v;
        const int te = (e == PK);
#pragma unroll 1
        for (int o = 0; o < E; ++o) {   // straight-line body (o == e gives a meaningless index that is never set)
            const T gx = tmax(T(0), tmax(lx - s.F(S::BB + 4 * o + 1), s.F(S::BB + 4 * o) - hx));
            const T gy = tmax(T(0), tmax(ly - s.F(S::BB + 4 * o + 3), s.F(S::BB + 4 * o + 2) - hy));
            const T sg = P.d.sig[te][o == PK] + sl;
            const int a = o < e ? o : e, b = o < e ? e : o;
            pb.set_if(o != e && o != skip && fm(gx, gx, gy * gy) <= sg * sg, pair_index(a, b));
        }
    }

    // ---- narrow phase: one item searched over the columns (insertCollisions :362-507, one test) ----
    // mirrored single-axis view of body a for wall wk: q = sgn * coordinate, overlap <=> q >= lim
    struct Axis {
        T sp, sv, u, w0, off_t, sgn;
        int cs, fl, ff, ty, e, ax;
    };
    RGS_DEV Axis axis_of(int a, int wk) const {
        Axis A;
        const int ax = wk >= 2 ? 0 : 1;
        const T sgn = (wk == WALL_UP || wk == WALL_RIGHT) ? T(1) : T(-1);
        A.ty = (a == PK);
        A.cs = s.I(S::CS + a);
        A.fl = s.I(S::FL + a);
        A.ff = s.I(S::FF + a);
        A.e = a; A.ax = ax; A.sgn = sgn;
        A.sp = sgn * s.F(S::SP + 4 * a + ax);
        A.sv = sgn * s.F(S::SP + 4 * a + 2 + ax);
        A.u = sgn * s.F(S::CMD + 2 * a + ax);
        A.w0 = A.sv - A.u;
        A.off_t = sgn * s.F(S::OFF + 2 * a + ax);
        return A;
    }
    // q and its velocity at column c >= cs (closed form from the segment start)
    RGS_DEV T axis_off(const Axis &A, int c) const {   // mirrored cumulative offset at column c
        T o = c >= A.fl ? A.off_t : T(0), t = T(0);
        const bool mid = c < A.fl && c >= A.ff;
        ld_if(mid, &s.TR(A.e, c, A.ax), t);
        return mid ? A.sgn * t : o;
    }
    RGS_DEV void axis_at(const Axis &A, int c, T &q, T &v) const {
        const int j = c - A.cs;
        const LaneGK<T> gk = tab->e[A.ty][j];
        q = fm(A.w0, gk.k, fm(A.u, P.d.dt_col * T(j), A.sp)) + axis_off(A, c);
        v = fm(A.w0, gk.g, A.u);
        if constexpr (kNoise)
            if (A.ty == 0) { q += A.sgn * NZ(A.e, c, A.ax); v += A.sgn * NZ(A.e, c, 2 + A.ax); }
        if (j == 0) { q = A.sp; v = A.sv; }
    }

    // first column c >= base at which body a overlaps wall wk AND the back-tracked time of impact
    // is valid (approaching at column c-1, inside this step); kGrid+1 when there is none.
    // The velocity of a segment is monotone, so q(c) is convex or concave and the first overlap
    // column is found by bisection on a monotone predicate; the reference's per-column test
    // (:374-463) then runs from that column on.
    RGS_DEV_BIG int search_wall(int a, int wk, int base, T &t_out) const {
        const Axis A = axis_of(a, wk);
        const T lim = wk >= 2 ? P.lim_x[A.ty] : P.lim_y[A.ty];
        int c0 = base;
        const bool noisy = kNoise && A.ty == 0;   // a noisy path is neither convex nor concave: plain column scan
        const bool bis = !noisy && A.fl <= base;
        if (bis) {  // no pending correctOverlap jump inside the range: q is convex / concave
            T q, v;
            axis_at(A, base, q, v);
            if (q < lim) {
                const bool falling = A.w0 > T(0);  // velocity decreases along the segment (concave q)
                int lo = base + 1, hi = kGrid + 1;
#pragma unroll 1
                for (int it = 0; it < 6; ++it) {   // straight-line body (a finished search re-evaluates a valid column)
                    const int mid = (lo + hi) >> 1;
                    const bool open = lo < hi;
                    axis_at(A, open ? mid : base, q, v);
                    const bool left = q >= lim || (falling && v <= T(0));
                    hi = (open && left) ? mid : hi;
                    lo = (open && !left) ? mid + 1 : lo;
                }
                c0 = lo;
                if (c0 > kGrid) return kGrid + 1;
            }
        }
        // columns c0-1 and c0 on the wall's axis only (the mirrored closed form equals eval() up to the sign)
        T pm, vm, p, w;
        if (c0 - 1 < A.cs) {
            const int ax = wk >= 2 ? 0 : 1;
            const T sgn = (wk == WALL_UP || wk == WALL_RIGHT) ? T(1) : T(-1);
            pm = sgn * s.F(S::PRV + 4 * a + ax);
            vm = sgn * s.F(S::PRV + 4 * a + 2 + ax);
        } else {
            axis_at(A, c0 - 1, pm, vm);
        }
        axis_at(A, c0, p, w);
        w = w - A.u;
        const T R = tab->e[A.ty][1].g, k1 = tab->e[A.ty][1].k, uh = A.u * P.d.dt_col;
        bool was_over = false;
#pragma unroll 1
        for (int c = c0; c <= kGrid; ++c) {
            const bool over = p >= lim;
            if (over && vm > T(0)) {
                const T tc = col_time(c - 1) + tmax(T(0), (lim - pm) / vm);
                if (in_step(tc)) { t_out = tc; return c; }
            }
            if constexpr (kNoise) {
                if (noisy) {
                    if (c < kGrid) {
                        T pn, vn;
                        axis_at(A, c + 1, pn, vn);
                        pm = p;
                        vm = A.u + w;
                        p = pn;
                        w = vn - A.u;
                    }
                    continue;
                }
            }
            // left the wall again on a concave arc with no pending jump: it cannot come back
            if (was_over && !over && A.fl <= c && w > T(0)) break;
            if (bis && !over && !was_over && c > c0) break;  // bisection: c0 was the only possible first overlap
            was_over = was_over || over;
            pm = p;
            vm = A.u + w;
            p = fm(w, k1, p + uh);
            w = w * R;
            if (c + 1 <= A.fl && c + 1 >= A.ff) p += axis_off(A, c + 1) - axis_off(A, c);
        }
        return kGrid + 1;
    }

    // register copy of what the position of a body at a column depends on (the column loops of
    // search_pair / march_pair evaluate positions only; same arithmetic as eval())
    struct PosSeg {
        T spx, spy, wx, wy, ux, uy, otx, oty;
        int cs, fl, ff, ty, e;
    };
    RGS_DEV PosSeg pos_seg(int e) const {
        PosSeg g;
        g.e = e;
        g.ty = (e == PK);
        g.cs = s.I(S::CS + e);
        g.fl = s.I(S::FL + e);
        g.ff = s.I(S::FF + e);
        g.spx = s.F(S::SP + 4 * e); g.spy = s.F(S::SP + 4 * e + 1);
        g.ux = s.F(S::CMD + 2 * e); g.uy = s.F(S::CMD + 2 * e + 1);
        g.wx = s.F(S::SP + 4 * e + 2) - g.ux; g.wy = s.F(S::SP + 4 * e + 3) - g.uy;
        g.otx = s.F(S::OFF + 2 * e); g.oty = s.F(S::OFF + 2 * e + 1);
        return g;
    }
    // position and velocity at column c (velocity: same arithmetic as eval())
    RGS_DEV void posvel_at(const PosSeg &g, int c, T &px, T &py, T &vx, T &vy) const {
        // c >= cs for every caller (a segment starts at the base of its event, and bases never decrease).  At c == cs the
        // closed form gives the start position exactly (k_0 = 0) and the start velocity to one rounding, which only feeds
        // the conservative column-skip bound: no special case, no branch
        const int j = c - g.cs;
        const LaneGK<T> gk = tab->e[g.ty][j];
        const T tj = P.d.dt_col * T(j);
        const bool late = c >= g.fl;
        T ox = late ? g.otx : T(0), oy = late ? g.oty : T(0);
        const bool mid = !late && c >= g.ff;
        const T *tp = &s.TR(g.e, c, 0);
        ld_if(mid, tp, ox);
        ld_if(mid, tp + 1, oy);
        px = fm(g.wx, gk.k, fm(g.ux, tj, g.spx)) + ox;
        py = fm(g.wy, gk.k, fm(g.uy, tj, g.spy)) + oy;
        vx = fm(g.wx, gk.g, g.ux);
        vy = fm(g.wy, gk.g, g.uy);
        if constexpr (kNoise)
            if (g.ty == 0) { px += NZ(g.e, c, 0); py += NZ(g.e, c, 1); vx += NZ(g.e, c, 2); vy += NZ(g.e, c, 3); }
    }
    // how far the relative velocity of two segments can drift from its value at any column: each body's velocity
    // moves monotonically from its segment-start value to its column-50 value
    RGS_DEV T drift_bound(int a, int b) const {
        const int tya = (a == PK), tyb = (b == PK);
        const T ga = tab->e[tya][kGrid - s.I(S::CS + a)].g, gb = tab->e[tyb][kGrid - s.I(S::CS + b)].g;
        T e2 = T(0);
#pragma unroll
        for (int ax = 0; ax < 2; ++ax) {
            const T wa = s.F(S::SP + 4 * a + 2 + ax) - s.F(S::CMD + 2 * a + ax), wb = s.F(S::SP + 4 * b + 2 + ax) - s.F(S::CMD + 2 * b + ax);
            // v_end - v_start = (v0 - u) (R^n - 1)
            const T ra = wa * (ga - T(1)), rb = wb * (gb - T(1));
            const T r = (ra < T(0) ? -ra : ra) + (rb < T(0) ? -rb : rb);
            e2 = fm(r, r, e2);
        }
        return skip_sqrt(e2) * T(1.001) + T(1.0e-6);
    }
    // Columns that are certainly contact-free after a column where the bodies are apart (conservative advancement).
    // Relative position d and velocity dv at that column (b minus a), |dv(t) - dv| <= eps afterwards, so
    // |d(t)| >= |d + dv t| - eps t; the earliest t with |d + dv t| = sig + eps t is the smaller positive root of
    // A t^2 + 2 B t + C.  64 = never again inside this step.
    RGS_DEV int free_columns(T dx, T dy, T dvx, T dvy, T ssq, T sig, T eps, T inv_h) const {
        const T sg = sig + Slack<T>::v;
        const T C = ssq - sg * sg;
        const T A = fm(dvx, dvx, dvy * dvy) - eps * eps;
        const T B = fm(dx, dvx, dy * dvy) - sg * eps;
        const T disc = fm(B, B, -(A * C));
        const T den = skip_sqrt(tmax(disc, T(0))) - B;
        // straight-line (selects, no early returns: every branch of the search loops is a reconvergence region)
        const bool never = disc < T(0) || !(den > T(0));
        const T cols = tmin(T(0.98) * C * inv_h / (never ? T(1) : den), T(64));
        const int n = never ? 64 : (int)cols;
        return C > T(0) ? n : 0;
    }

    // same for the pair (a, b); the time is symmetric in (a, b) as in the reference (:482-494).
    // Conservative advancement: while the bodies are apart, the relative speed bound S of the two
    // segments says how many columns can be skipped before they could touch.
    RGS_DEV_BIG int search_pair(int a, int b, int base, T &t_out) const {
        const int tya = (a == PK), tyb = (b == PK);
        const T thr = P.d.ov_le[tya][tyb], sigma = P.d.sig[tya][tyb];
        const T h = P.d.dt_col;
        // no skipping while a correctOverlap offset change of either body is still ahead
        const int fl_ab = s.I(S::FL + a) > s.I(S::FL + b) ? s.I(S::FL + a) : s.I(S::FL + b);
        const T eps = drift_bound(a, b), inv_h = T(1) / h;
        int c = base;
        const PosSeg ga = pos_seg(a), gb = pos_seg(b);
#pragma unroll 1
        while (c <= kGrid) {
            T ax, ay, avx, avy, bx, by, bvx, bvy;
            posvel_at(ga, c, ax, ay, avx, avy);
            posvel_at(gb, c, bx, by, bvx, bvy);
            const T ex = ax - bx, ey = ay - by;
            const T ssq = fm(ex, ex, ey * ey);
            if (ssq <= thr) {
                eval(a, c - 1, ax, ay, avx, avy);
                eval(b, c - 1, bx, by, bvx, bvy);
                const T dpx = bx - ax, dpy = by - ay, dvx = bvx - avx, dvy = bvy - avy;
                const T dpdv = fm(dpx, dvx, dpy * dvy), dvdv = fm(dvx, dvx, dvy * dvy), dpdp = fm(dpx, dpx, dpy * dpy);
                const T d = dpdv * dpdv - dvdv * (dpdp - sigma * sigma);
                if (dpdv < T(0) && d >= T(0)) {
                    const T tc = col_time(c - 1) + tmax(T(0), -(dpdv + T(sqrt(d))) / dvdv);
                    if (in_step(tc)) { t_out = tc; return c; }
                }
                ++c;
            } else {
                c += 1 + ((kNoise || c < fl_ab) ? 0 : free_columns(-ex, -ey, bvx - avx, bvy - avy, ssq, sigma, eps, inv_h));
            }
        }
        return kGrid + 1;
    }

    // candidate (c, t, kind, j) of owner e: keep it if it beats the owner's current candidate in the
    // reference's order: earlier column; same column: smaller t; same t: earlier in the candidate
    // order walls UP, DOWN, RIGHT, LEFT, bodies in entity order (strict '<', :389,:412,:435,:458,:495)
    RGS_DEV void fold(int e, int c, T t, int kind, int j, unsigned &fmask) const {
        const int rank = kind < PAIR ? kind : PAIR + j;
        bool better = true;
        if ((fmask >> e) & 1u) {
            const int qi = s.I(S::QI + e);
            const int bc = (qi >> 8) & 63, bk = (qi >> 4) & 15, bj = qi & 15;
            const int brank = bk < PAIR ? bk : PAIR + bj;
            const T bt = s.F(S::QT + e);
            better = c < bc || (c == bc && (t < bt || (t == bt && rank < brank)));
        }
        if (better) {
            s.F(S::QT + e) = t;
            s.I(S::QI + e) = (c << 8) | (kind << 4) | (j & 15);
            fmask |= 1u << e;
        }
    }

    RGS_DEV void run_items(Bits<4 * E> &wb, Bits<NPAIR> &pb, unsigned owners, int base, unsigned &fmask) const {
        int bit;
#pragma unroll 1
        while ((bit = wb.pop()) >= 0) {
            const int a = bit >> 2, wk = bit & 3;
            T t;
            if constexpr (kNoise) ensure_path(a);
            const int c = search_wall(a, wk, base, t);
            if (c <= kGrid) fold(a, c, t, wk, 15, fmask);
        }
#pragma unroll 1
        while ((bit = pb.pop()) >= 0) {
            int a, b;
            pair_decode(bit, a, b);
            T t;
            if constexpr (kNoise) { ensure_path(a); ensure_path(b); }
            const int c = search_pair(a, b, base, t);
            if (c <= kGrid) {
                if constexpr (kOneCopy) {
#pragma unroll 1
                    for (int k = 0; k < 2; ++k) {   // (rolled: one copy of fold)
                        const int own = k == 0 ? a : b, oth = k == 0 ? b : a;
                        if ((owners >> own) & 1u) fold(own, c, t, PAIR, oth, fmask);
                    }
                } else {
                    if ((owners >> a) & 1u) fold(a, c, t, PAIR, b, fmask);
                    if ((owners >> b) & 1u) fold(b, c, t, PAIR, a, fmask);
                }
            }
        }
    }

    // ---- event queue: <= one event per owner body (reference src/GameSim.cpp:136-150) ----
    // QI = (seq << 16) | (col << 8) | (kind << 4) | j  (j = 15 for walls)
    RGS_DEV void push(int e, unsigned &qmask, int &seq) const {
        s.I(S::QI + e) = (s.I(S::QI + e) & 0xFFFF) | (seq << 16);
        qmask |= 1u << e;
        ++seq;
    }
    // earliest time; among equal times the most recently inserted (stable insertion sort, a11)
    RGS_DEV void pop(unsigned &qmask, int &ei, int &ej, int &ekind, T &et, int &ecol) const {
        int best = -1, binfo = 0;
        T bt = T(0);
#pragma unroll 1
        for (unsigned m = qmask; m != 0u; m &= m - 1u) {  // queued owners only (usually one)
            const int e = low_bit(m);
            const T te = s.F(S::QT + e);
            const int info = s.I(S::QI + e);
            if (best < 0 || te < bt || (te == bt && (info >> 16) > (binfo >> 16))) { best = e; bt = te; binfo = info; }
        }
        ei = best;
        et = bt;
        ekind = (binfo >> 4) & 15;
        ecol = (binfo >> 8) & 255;
        ej = binfo & 15;
        if (ej == 15) ej = -1;
        qmask &= ~(1u << best);
    }
    // drop queued events sharing a body with the executed one (:347-352)
    RGS_DEV void prune(unsigned &qmask, int ei, int ej) const {
#pragma unroll 1
        for (unsigned m = qmask; m != 0u; m &= m - 1u) {
            const int e = low_bit(m);
            int cj = s.I(S::QI + e) & 15;
            if (cj == 15) cj = -1;
            const bool drop = (ei == e) || (ei == cj && cj != -1) || (ej == e) || (ej == cj && cj != -1);
            if (drop) qmask &= ~(1u << e);
        }
    }

    // correctOverlap on the columns after an impact (:285-297): visit the columns of the pair's new
    // segments where it can still overlap (conservative advancement on the relative speed bound,
    // as in search_pair), record the position offsets
    RGS_DEV void march_pair(int bi, int bj, int base) const {
        const int tyi = (bi == PK), tyj = (bj == PK);
        const T thr = P.d.ov_lt[tyi][tyj], sigma = P.d.sig[tyi][tyj];
        const T eps = drift_bound(bi, bj), inv_h = T(1) / P.d.dt_col;
        int c = base + 1;
        PosSeg gi = pos_seg(bi), gj = pos_seg(bj);
#pragma unroll 1
        while (c <= kGrid) {
            T ax, ay, bx, by, avx, avy, bvx, bvy;
            posvel_at(gi, c, ax, ay, avx, avy);
            posvel_at(gj, c, bx, by, bvx, bvy);
            const T dx = bx - ax, dy = by - ay;
            const T ssq = fm(dx, dx, dy * dy);
            if (ssq <= thr) {
                const T overlap = sigma - T(sqrt(ssq));
                const T si = P.d.fix_coef[tyi][tyj] * overlap, sj = P.d.fix_coef[tyj][tyi] * overlap;
                const T fax = -(si * dx), fay = -(si * dy), fbx = sj * dx, fby = sj * dy;
                record_fix(bi, c, fax, fay);
                record_fix(bj, c, fbx, fby);
                gi = pos_seg(bi);
                gj = pos_seg(bj);
                ev_add(RGS_EV_OVERLAP_FIXES);
                ++c;
            } else {
                c += 1 + (kNoise ? 0 : free_columns(dx, dy, bvx - avx, bvy - avy, ssq, sigma, eps, inv_h));
            }
        }
    }
    // same for a body and the wall it hit (:329-340).  q(c) is convex or concave on the new
    // segment and starts inside the limit, so it can only overlap again if its last column does.
    RGS_DEV void march_wall(int bi, int wk, int base) const {
        const Axis A = axis_of(bi, wk);
        const int ax = wk >= 2 ? 0 : 1;
        const T sgn = (wk == WALL_UP || wk == WALL_RIGHT) ? T(1) : T(-1);
        const T lim = ax == 0 ? P.lim_x[A.ty] : P.lim_y[A.ty];
        const T lim_in = ax == 0 ? P.lim_x_in[A.ty] : P.lim_y_in[A.ty];
        if constexpr (kNoise) {
            if (A.ty == 0) {   // noisy path: every column from the closed form + noise + the offsets recorded so far
                Axis B = A;
#pragma unroll 1
                for (int c = base + 1; c <= kGrid; ++c) {
                    T p, v;
                    axis_at(B, c, p, v);
                    if (p >= lim) {
                        const T f = tmin(p - P.d.wall_fix * (p - lim), lim_in) - p;
                        record_fix(bi, c, ax == 0 ? sgn * f : T(0), ax == 0 ? T(0) : sgn * f);
                        ev_add(RGS_EV_OVERLAP_FIXES);
                        if (B.fl == B.cs) B.ff = c;
                        B.fl = c;
                        B.off_t += f;
                    }
                }
                return;
            }
        }
        {
            T qe, ve;
            axis_at(A, kGrid, qe, ve);
            if (qe < lim) return;
        }
        T p = A.sp, w = A.w0;
        const T R = tab->e[A.ty][1].g, k1 = tab->e[A.ty][1].k, uh = A.u * P.d.dt_col;
#pragma unroll 1
        for (int c = base + 1; c <= kGrid; ++c) {
            p = fm(w, k1, p + uh);
            w *= R;
            if (p >= lim) {  // correctOverlap(wall) :594-623, kept strictly inside the limit
                const T pc = tmin(p - P.d.wall_fix * (p - lim), lim_in);
                const T f = pc - p;
                p = pc;
                record_fix(bi, c, ax == 0 ? sgn * f : T(0), ax == 0 ? T(0) : sgn * f);
                ev_add(RGS_EV_OVERLAP_FIXES);
            }
        }
    }

    // ---- execute one popped event (carryOutFirstCollision :237-345): advance the bodies to the time of
    // impact, resolve, start their new segments at column `base`, correctOverlap on the later columns.
    // Returns the second body to re-scan (-1 for a wall event); goal = 1 (team A) / 2 (team B) / 0.
    RGS_DEV_BIG int resolve(int ei, int ej, int ekind, T et, int ecol, int &base, int &goal, double *dmg, long long live,
                            double *dv_sq_out = nullptr) const {
        int rj = -1;
        goal = 0;
        if constexpr (kNoise) ++nz_epoch;   // fresh noise for everything this event re-integrates
        T dt_t;
        if constexpr (kAbsClock) {
            int nb = (int)floor(since_t0(et) / P.d.dt_col) + 1;  // :237
            if (nb > kGrid) { nb = kGrid; ev_add(RGS_EV_BASE_CLAMP); }
            if (nb > base) base = nb;  // event times never run backwards by more than a rounding
            dt_t = et - col_time(base - 1);  // :254 / :308
        } else {
            // The reference forms t_col = (t_ + dt_col_ (i-1)) + tau on its absolute fp64 clock (:201, :389) and takes
            // base_idx = floor((t_col - t_) / dt_col_) + 1 (:237): for a RESTING contact (tau = 0) the rounding of t_
            // decides between column i-1 and i, which sets how long a jammed event loop runs.  fp32 event times are
            // step-relative; that one decision (and dt_t, :254) is taken in fp64 exactly as the reference takes it.
            const T tau = et - col_time(ecol - 1);
            int nb;
#ifdef RGS_AB_NO64BASE
            if (false) {
#else
            if (tau == T(0)) {   // resting contact: the reference's clock rounding decides
#endif
                const double tcol = tclk + P.dt_col64 * (double)(ecol - 1);
                // (t_col - t_) / dt_col_, correctly rounded without nvcc's division expansion (whose slow path a zero
                // numerator -- a contact resting at column 1 -- would take): quotient correction with y = RN(1 / dt_col_)
                const double x = tcol - tclk, q0 = x * P.inv_dt_col64;
                const double q = x == 0.0 ? 0.0 : fm(fm(-P.dt_col64, q0, x), P.inv_dt_col64, q0);
                nb = (int)floor(q) + 1;
                if (nb > kGrid) { nb = kGrid; ev_add(RGS_EV_BASE_CLAMP); }
                if (nb > base) base = nb;
                dt_t = (T)(tcol - (tclk + P.dt_col64 * (double)(base - 1)));
            } else {
                nb = (int)floor(et / P.d.dt_col) + 1;
                if (nb > kGrid) { nb = kGrid; ev_add(RGS_EV_BASE_CLAMP); }
                if (nb > base) base = nb;
                dt_t = et - col_time(base - 1);
            }
        }
        const T h_rest = P.d.dt_col - dt_t;
        const int tyi = (ei == PK);
        T qix, qiy, qvix, qviy;
        eval(ei, base - 1, qix, qiy, qvix, qviy);
        T pix = qix, piy = qiy, vix = qvix, viy = qviy;
        const T uix = s.F(S::CMD + 2 * ei), uiy = s.F(S::CMD + 2 * ei + 1);
        advance_axis(pix, vix, uix, dt_t, P.neg_inv_tau[tyi]);
        advance_axis(piy, viy, uiy, dt_t, P.neg_inv_tau[tyi]);
        if constexpr (kNoise)
            if (!tyi) noise_kick(ei, 0, dt_t, pix, piy, vix, viy);
        if (ekind == PAIR) {  // :242-299
            const int tyj = (ej == PK);
            T qjx, qjy, qvjx, qvjy;
            eval(ej, base - 1, qjx, qjy, qvjx, qvjy);
            T pjx = qjx, pjy = qjy, vjx = qvjx, vjy = qvjy;
            const T ujx = s.F(S::CMD + 2 * ej), ujy = s.F(S::CMD + 2 * ej + 1);
            advance_axis(pjx, vjx, ujx, dt_t, P.neg_inv_tau[tyj]);
            advance_axis(pjy, vjy, ujy, dt_t, P.neg_inv_tau[tyj]);
            if constexpr (kNoise)
                if (!tyj) noise_kick(ej, 0, dt_t, pjx, pjy, vjx, vjy);
            {  // residual overlap of the constant-velocity impact-time estimate :259-265
                const T dx = pjx - pix, dy = pjy - piy;
                const T ssq = dx * dx + dy * dy;
                if (ssq <= P.d.ov_lt[tyi][tyj]) {
                    const T overlap = P.d.sig[tyi][tyj] - T(sqrt(ssq));
                    const T si = P.d.hit_coef[tyi][tyj] * overlap, sj = P.d.hit_coef[tyj][tyi] * overlap;
                    pix -= si * dx; piy -= si * dy;
                    pjx += sj * dx; pjy += sj * dy;
                }
            }
            {  // elastic impulse :268-274
                const T dvx = vix - vjx, dvy = viy - vjy, dpx = pix - pjx, dpy = piy - pjy;
                const T dot = fm(dvx, dpx, dvy * dpy), nn = fm(dpx, dpx, dpy * dpy);
                const T fi = P.d.imp_coef[tyi][tyj] * dot / nn, fj = P.d.imp_coef[tyj][tyi] * dot / nn;
                if (ei != PK && ej != PK) {  // damage :277-282
                    const double dv_sq = 0.01 * ((double)dvx * (double)dvx + (double)dvy * (double)dvy);
                    if (dv_sq_out) *dv_sq_out = dv_sq;
                    dmg[ei * NP + ej] += (double)P.d.mass[tyi] * dv_sq;
                    dmg[ej * NP + ei] += (double)P.d.mass[tyj] * dv_sq;
                    ev_add(RGS_EV_PLAYER_PLAYER);
                } else {
                    ev_add(RGS_EV_PLAYER_PUCK);
                }
                vix -= fi * dpx; viy -= fi * dpy;
                vjx += fj * dpx; vjy += fj * dpy;
            }
            advance_axis(pix, vix, uix, h_rest, P.neg_inv_tau[tyi]);
            advance_axis(piy, viy, uiy, h_rest, P.neg_inv_tau[tyi]);
            advance_axis(pjx, vjx, ujx, h_rest, P.neg_inv_tau[tyj]);
            advance_axis(pjy, vjy, ujy, h_rest, P.neg_inv_tau[tyj]);
            if constexpr (kNoise) {
                if (!tyi) noise_kick(ei, 1, h_rest, pix, piy, vix, viy);
                if (!tyj) noise_kick(ej, 1, h_rest, pjx, pjy, vjx, vjy);
            }
            {  // correctOverlap at column base :287
                const T dx = pjx - pix, dy = pjy - piy;
                const T ssq = dx * dx + dy * dy;
                if (ssq <= P.d.ov_lt[tyi][tyj]) {
                    const T overlap = P.d.sig[tyi][tyj] - T(sqrt(ssq));
                    const T si = P.d.fix_coef[tyi][tyj] * overlap, sj = P.d.fix_coef[tyj][tyi] * overlap;
                    pix -= si * dx; piy -= si * dy;
                    pjx += sj * dx; pjy += sj * dy;
                    ev_add(RGS_EV_OVERLAP_FIXES);
                }
            }
            start_segment(ei, base, pix, piy, vix, viy, qix, qiy, qvix, qviy);
            start_segment(ej, base, pjx, pjy, vjx, vjy, qjx, qjy, qvjx, qvjy);
            march_pair(ei, ej, base);
            if constexpr (!kOneCopy) update_bbox(ej);
            rj = ej;
        } else {  // wall :300-343
            switch (ekind) {  // flip only while still moving into the wall :312-318
            case WALL_UP: if (viy > T(0)) viy = -viy; break;
            case WALL_DOWN: if (viy < T(0)) viy = -viy; break;
            case WALL_RIGHT: if (vix > T(0)) vix = -vix; break;
            default: if (vix < T(0)) vix = -vix; break;
            }
            const bool is_goal = (ei == PK) && (ekind == WALL_LEFT || ekind == WALL_RIGHT) && piy < P.d.half_goal &&
                                 piy > -P.d.half_goal;  // :321-322
            if (is_goal) {
                goal = ekind == WALL_RIGHT ? 1 : 2;
                ev_add(goal == 1 ? RGS_EV_GOALS_A : RGS_EV_GOALS_B);
                pix = piy = vix = viy = T(0);  // puck reset :325
            } else {
                advance_axis(pix, vix, uix, h_rest, P.neg_inv_tau[tyi]);
                advance_axis(piy, viy, uiy, h_rest, P.neg_inv_tau[tyi]);
                if constexpr (kNoise)
                    if (!tyi) noise_kick(ei, 1, h_rest, pix, piy, vix, viy);
                {  // correctOverlap(wall) at column base :334 (folded limit, kept strictly inside)
                    const bool xa = ekind >= WALL_RIGHT;
                    const T sgn = (ekind == WALL_UP || ekind == WALL_RIGHT) ? T(1) : T(-1);
                    const T lim = xa ? P.lim_x[tyi] : P.lim_y[tyi], lim_in = xa ? P.lim_x_in[tyi] : P.lim_y_in[tyi];
                    const T q = sgn * (xa ? pix : piy);
                    if (q >= lim) {
                        const T qc = sgn * tmin(q - P.d.wall_fix * (q - lim), lim_in);
                        if (xa) pix = qc; else piy = qc;
                        ev_add(RGS_EV_OVERLAP_FIXES);
                    }
                }
            }
            ev_add(RGS_EV_WALL_UP + ekind);
            start_segment(ei, base, pix, piy, vix, viy, qix, qiy, qvix, qviy);
            if (!is_goal) march_wall(ei, ekind, base);
        }
        if constexpr (!kOneCopy) update_bbox(ei);   // (kOneCopy: the caller refreshes the bounding intervals, event_loop)
        hash_ev(live, ekind, ei, ej, goal);
        return rj;
    }

    // ---- jammed steps ------------------------------------------------------------------------------------
    // 0.03 % of the S0 steps run the event loop to its bound of 50 (:150): two bodies (and a wall) resolved at
    // the same instant over and over.  Every resolution restarts from the unchanged column base-1, so the loop
    // state becomes PERIODIC (period 2 or 3) after a few iterations.  Once the whole loop state at the top of an
    // iteration equals, bit for bit, the state a few iterations earlier, the remaining whole periods are not
    // executed: their side effects (event counters, the order-sensitive event hash, damage, push sequence
    // numbers) are replayed exactly and the loop resumes for the last partial period.  Identical results,
    // ~10 % fewer instructions overall (those steps were 12 % of them).
    // The snapshot's run-time-indexed words live in the lane's global scratch area behind the tracks (JF / JI),
    // not in local memory; only its scalars are registers.
#ifdef RGS_AB_NOJAM
    static constexpr bool kJamFF = false;
#else
    static constexpr bool kJamFF = NT <= 3 && !kNoise;
#endif   // (fresh noise every iteration: the loop state never repeats)
    struct JamSnap {
        unsigned qmask, owners, wbits, pbits, ranks;
        int base, ri, rj, scored, seq, counter, log_n;
    };
    // JF: SP, CMD, PRV, OFF, BB, and QT of the queued owners.  JI: CS, FL, QI & 0xFFFF, FF, TS [5E]; the event
    // counters [17]; per logged event kind | i << 4 | (j + 1) << 8 | goal << 12 [kJamLog] and the two halves of
    // dv_sq of a player-player impact (-1: none) [2 kJamLog]
    static constexpr int JI_EV = 5 * E, JI_LOG = 5 * E + 17, JI_DV = JI_LOG + kJamLog;
    RGS_DEV T &JF(int w) const { return s.trk[S::JAM0 + w]; }
    RGS_DEV int &JI(int w) const { return s.G(S::JAM0 + S::JAM_F + w); }
    static RGS_DEV bool same_bits(T a, T b) {
        if constexpr (sizeof(T) == 4) {
            unsigned x, y;
            memcpy(&x, &a, 4); memcpy(&y, &b, 4);
            return x == y;
        } else {
            unsigned long long x, y;
            memcpy(&x, &a, 8); memcpy(&y, &b, 8);
            return x == y;
        }
    }
    // order of the queued owners by push sequence number (the tie rule of pop() only sees this order)
    RGS_DEV unsigned queue_ranks(unsigned qmask) const {
        unsigned r = 0;
#pragma unroll 1
        for (unsigned m = qmask; m != 0u; m &= m - 1u) {
            const int e = low_bit(m);
            const int se = s.I(S::QI + e) >> 16;
            int rank = 0;
#pragma unroll 1
            for (unsigned m2 = qmask; m2 != 0u; m2 &= m2 - 1u) rank += (s.I(S::QI + low_bit(m2)) >> 16) < se;
            r |= (unsigned)rank << (3 * e);
        }
        return r;
    }
    RGS_DEV void jam_snapshot(JamSnap &J, unsigned qmask, unsigned owners, unsigned wbits, unsigned pbits, int base, int ri, int rj,
                              bool scored, int seq, int counter) const {
#pragma unroll 1
        for (int w = 0; w < S::QT; ++w) JF(w) = s.F(w);
#pragma unroll 1
        for (int e = 0; e < E; ++e) {
            JF(S::QT + e) = s.F(S::QT + e);
            JI(e) = s.I(S::CS + e);
            JI(E + e) = s.I(S::FL + e);
            JI(2 * E + e) = s.I(S::QI + e) & 0xFFFF;
            JI(3 * E + e) = s.I(S::FF + e);
            JI(4 * E + e) = s.G(S::TS + e);
        }
#pragma unroll 1
        for (int k = 0; k < 17; ++k) JI(JI_EV + k) = EVW(k);
        J.qmask = qmask; J.owners = owners; J.wbits = wbits; J.pbits = pbits; J.ranks = queue_ranks(qmask);
        J.base = base; J.ri = ri; J.rj = rj; J.scored = scored; J.seq = seq; J.counter = counter;
        J.log_n = 0;
    }
    RGS_DEV bool jam_same(const JamSnap &J, unsigned qmask, unsigned owners, unsigned wbits, unsigned pbits, int base, int ri, int rj,
                          bool scored) const {
        if (J.qmask != qmask || J.owners != owners || J.wbits != wbits || J.pbits != pbits || J.base != base || J.ri != ri ||
            J.rj != rj || J.scored != (int)scored)
            return false;
#pragma unroll 1
        for (int w = 0; w < S::QT; ++w)
            if (!same_bits(JF(w), s.F(w))) return false;
#pragma unroll 1
        for (int e = 0; e < E; ++e) {
            if (JI(e) != s.I(S::CS + e) || JI(E + e) != s.I(S::FL + e) || JI(3 * E + e) != s.I(S::FF + e)) return false;
            if (s.I(S::FL + e) != s.I(S::FF + e) && JI(4 * E + e) != s.G(S::TS + e)) return false;
            if ((qmask >> e) & 1u)
                if (!same_bits(JF(S::QT + e), s.F(S::QT + e)) || JI(2 * E + e) != (s.I(S::QI + e) & 0xFFFF)) return false;
        }
        return J.ranks == queue_ranks(qmask);
    }
    // skip the whole periods that fit before the loop bound; returns the new counter
    RGS_DEV int jam_fast_forward(const JamSnap &J, int counter, int &seq, unsigned qmask, double *dmg, long long live) const {
        const int period = counter - J.counter;
        const int m = (kCounterLim - counter) / period;
        if (m <= 0 || J.log_n != period) return counter;
        RGS_TRACE("jam fast-forward: live %lld counter %d period %d skipped %d\n", live, counter, period, m * period);
        static constexpr int kAdd[] = {RGS_EV_GOALS_A, RGS_EV_GOALS_B, RGS_EV_WALL_UP, RGS_EV_WALL_DOWN, RGS_EV_WALL_RIGHT,
                                       RGS_EV_WALL_LEFT, RGS_EV_PLAYER_PLAYER, RGS_EV_PLAYER_PUCK, RGS_EV_DISCARDED,
                                       RGS_EV_BASE_CLAMP, RGS_EV_OVERLAP_FIXES};
#pragma unroll 1
        for (int k = 0; k < (int)(sizeof(kAdd) / sizeof(int)); ++k)
            EVW(kAdd[k]) += m * (EVW(kAdd[k]) - JI(JI_EV + kAdd[k]));
#pragma unroll 1
        for (int r = 0; r < m; ++r)
#pragma unroll 1
            for (int l = 0; l < period; ++l) {
                const int c = JI(JI_LOG + l);
                const int kind = c & 15, i = (c >> 4) & 15, j = ((c >> 8) & 15) - 1, goal = (c >> 12) & 3;
                hash_ev(live, kind, i, j, goal);
                const unsigned long long db = ((unsigned long long)(unsigned)JI(JI_DV + 2 * l + 1) << 32) | (unsigned)JI(JI_DV + 2 * l);
                double dv;
                memcpy(&dv, &db, 8);
                if (dv >= 0.0) {  // damage :277-282, same operands in the same order
                    dmg[i * NP + j] += (double)P.d.mass[i == PK] * dv;
                    dmg[j * NP + i] += (double)P.d.mass[j == PK] * dv;
                }
            }
        // the sequence numbers of the queued events advance by whole periods too (their order is what matters)
        const int dseq = m * (seq - J.seq);
#pragma unroll 1
        for (unsigned q = qmask; q != 0u; q &= q - 1u) {
            const int e = low_bit(q);
            if ((s.I(S::QI + e) >> 16) >= J.seq) s.I(S::QI + e) += dseq << 16;
        }
        seq += dseq;
        return counter + m * period;
    }
    RGS_DEV void jam_log(JamSnap &J, int kind, int i, int j, int goal, double dv_sq) const {
        if (J.log_n < kJamLog) {
            JI(JI_LOG + J.log_n) = kind | (i << 4) | ((j + 1) << 8) | (goal << 12);
            unsigned long long db;
            memcpy(&db, &dv_sq, 8);
            JI(JI_DV + 2 * J.log_n) = (int)(unsigned)db;
            JI(JI_DV + 2 * J.log_n + 1) = (int)(unsigned)(db >> 32);
        }
        J.log_n++;
    }

    // ---- the event loop of one game-step (GameSim::updateSim :136-150) on the flagged items ----
    // returns the loop counter of the step
    RGS_DEV int event_loop(Bits<4 * E> &wb, Bits<NPAIR> &pb, double &sa, double &sb, double *dmg, long long live) const {
        int counter = 1;
        if (wb.any() || pb.any()) {
            unsigned qmask = 0, fmask = 0, owners = (1u << E) - 1u;
            int seq = 0, base = 1, ri = -1, rj = -1;
            bool scored = false;
            JamSnap J;
            J.counter = 0;  // no snapshot yet
#pragma unroll 1
            for (;;) {
                if constexpr (kJamFF) {
                    if (counter >= 8) {
                        if ((counter & 7) == 0) {
                            jam_snapshot(J, qmask, owners, wb.w[0], pb.w[0], base, ri, rj, scored, seq, counter);
                        } else if (J.counter != 0 && counter - J.counter <= kJamLog &&
                                   jam_same(J, qmask, owners, wb.w[0], pb.w[0], base, ri, rj, scored)) {
                            counter = jam_fast_forward(J, counter, seq, qmask, dmg, live);
                            J.counter = 0;
                        }
                    }
                }
                fmask = 0;
                run_items(wb, pb, owners, base, fmask);
                if (ri < 0) {
#pragma unroll 1
                    for (unsigned m = fmask; m != 0u; m &= m - 1u) push(low_bit(m), qmask, seq);  // ascending owners
                } else {
                    if ((fmask >> ri) & 1u) push(ri, qmask, seq);
                    if (rj >= 0 && ((fmask >> rj) & 1u)) push(rj, qmask, seq);
                }
                if (qmask == 0 || counter >= kCounterLim) break;
                int ei, ej, ekind, ecol;
                T et;
                pop(qmask, ei, ej, ekind, et, ecol);
                ri = ei; rj = -1; owners = 0;  // nothing to re-scan unless the event executes
                if ((ei == PK || ej == PK) && scored) {  // :230-231
                    ev_add(RGS_EV_DISCARDED);
                    hash_ev(live, DISCARD, ei, ej, 0);
                    if constexpr (kJamFF)
                        if (J.counter != 0) jam_log(J, DISCARD, ei, ej, 0, -1.0);
                    ++counter;
                    continue;
                }
                int goal = 0;
                double dv_sq = -1.0;
                rj = resolve(ei, ej, ekind, et, ecol, base, goal, dmg, live, &dv_sq);
                if constexpr (kJamFF)
                    if (J.counter != 0) jam_log(J, ekind, ei, ej, goal, dv_sq);
                if (goal) {
                    if (goal == 1) sa += 1.0; else sb += 1.0;
                    scored = true;
                }
                prune(qmask, ei, ej);  // :347-352
                // re-scan the bodies involved (:355-357, :147): broad phase on their new segments.  Two rolled
                // two-trip loops, so that the kernel holds ONE copy of update_bbox / flag_walls / reflag_pairs: the
                // kernel is instruction-fetch-bound and its hot code sits just above the 32 KB instruction cache
                // (about 3 % per 100 instructions of hot code, profiles/r02_ab_variants.md section 9)
                if constexpr (kOneCopy) {
                    owners = 0;
#pragma unroll 1
                    for (int k = 0; k < 2; ++k) {
                        const int e = k == 0 ? ei : rj;
                        if (e < 0) break;
                        owners |= 1u << e;
                        update_bbox(e);
                        flag_walls(e, wb);
                    }
#pragma unroll 1
                    for (int k = 0; k < 2; ++k) {
                        const int e = k == 0 ? ei : rj;
                        if (e < 0) break;
                        reflag_pairs(e, k == 0 ? -1 : ei, pb);
                    }
                } else {
                    owners = 1u << ei;
                    flag_walls(ei, wb);
                    if (rj >= 0) {
                        owners |= 1u << rj;
                        flag_walls(rj, wb);
                    }
                    reflag_pairs(ei, -1, pb);
                    if (rj >= 0) reflag_pairs(rj, ei, pb);
                }
                ++counter;
            }
            if (qmask != 0) ev_add(RGS_EV_COUNTER_LIMIT);
        }
        return counter;
    }

    // ---- one game-step; SP / CMD of the scratch record hold the state and the commands between steps ----
    // returns the event-loop counter of the step (1 when nothing happened)
    RGS_DEV int step(double t_clock, double &sa, double &sb, double *dmg, long long live) const {
        if constexpr (kAbsClock) t0 = (T)t_clock;
        tclk = t_clock;
        Bits<4 * E> wb;
        Bits<NPAIR> pb;
        unsigned need;
        step_head(wb, pb, need);
        if constexpr (kNoise) {
            nz_step = (unsigned)live;
            nz_epoch = 0;
            if (!(wb.any() || pb.any())) {
                // nothing can touch anything even with the noise envelope: only the END of every player's noise path
                // matters, drawn from the accumulated Gaussian of the 50 sub-steps (same distribution)
                nz_on = false;
                step_commit();
#pragma unroll 1
                for (int e = 0; e < NP; ++e) add_end_noise(e);
                return 1;
            }
            nz_on = true;
            nz_have = 0;
            // the paths of the bodies of the flagged items, generated here with all the lanes of the warp together; a
            // body that only gets involved later in the step gets its path then (ensure_path in run_items)
#pragma unroll 1
            for (int e = 0; e < NP; ++e)
                if ((need >> e) & 1u) ensure_path(e);
        }
        const int counter = event_loop(wb, pb, sa, sb, dmg, live);
        step_commit();
        if constexpr (kNoise) {
#pragma unroll 1
            for (int e = 0; e < NP; ++e)
                if (!((nz_have >> e) & 1u)) add_end_noise(e);   // players nothing evaluated inside the step
        }
        return counter;
    }
    // whole-step segments of all bodies and their broad-phase flags
    // need (noise kernels): the bodies of the flagged items -- their noise paths are generated up front, all lanes together
    RGS_DEV void step_head(Bits<4 * E> &wb, Bits<NPAIR> &pb, unsigned &need) const {
        wb.clear();
        pb.clear();
        need = 0u;
        {   // whole-step segments: update_bbox + flag_walls with cs = 0 and zero offsets, from registers
            const T trem = P.d.dt_col * T(kGrid), sl = Slack<T>::v;
            const T gE0 = tab->e[0][kGrid].g, gE1 = tab->e[1][kGrid].g;   // (selected, not indexed: no local-memory array)
#pragma unroll 1
            for (int e = 0; e < E; ++e) {
                const int ty = (e == PK);
                s.I(S::CS + e) = 0;
                s.I(S::FL + e) = 0;
                s.I(S::FF + e) = 0;
                T lohi[4];
#pragma unroll
                for (int a = 0; a < 2; ++a) {
                    s.F(S::OFF + 2 * e + a) = T(0);
                    const T p = s.F(S::SP + 4 * e + a), v = s.F(S::SP + 4 * e + 2 + a), u = s.F(S::CMD + 2 * e + a);
                    const T vE = fm(v - u, ty ? gE1 : gE0, u);
                    const T env = (kNoise && !ty) ? P.nz_env : T(0);
                    lohi[2 * a] = p + tmin(T(0), tmin(v, vE) * trem) - env;
                    lohi[2 * a + 1] = p + tmax(T(0), tmax(v, vE) * trem) + env;
                    s.F(S::BB + 4 * e + 2 * a) = lohi[2 * a];
                    s.F(S::BB + 4 * e + 2 * a + 1) = lohi[2 * a + 1];
                }
                const T lx = P.lim_x[ty] - sl, ly = P.lim_y[ty] - sl;
                wb.set_if(lohi[3] >= ly, 4 * e + WALL_UP);
                wb.set_if(lohi[2] <= -ly, 4 * e + WALL_DOWN);
                wb.set_if(lohi[1] >= lx, 4 * e + WALL_RIGHT);
                wb.set_if(lohi[0] <= -lx, 4 * e + WALL_LEFT);
                if constexpr (kNoise)
                    need |= (lohi[3] >= ly || lohi[2] <= -ly || lohi[1] >= lx || lohi[0] <= -lx) ? 1u << e : 0u;
            }
            // flag_pair with the first body's interval held in registers
            const T sgpp = P.d.sig[0][0] + sl, sgpk = P.d.sig[0][1] + sl;
            const T thr_pp = sgpp * sgpp, thr_pk = sgpk * sgpk;
            int pidx = 0;
#pragma unroll 1
            for (int a = 0; a < E - 1; ++a) {
                const T alx = s.F(S::BB + 4 * a), ahx = s.F(S::BB + 4 * a + 1), aly = s.F(S::BB + 4 * a + 2), ahy = s.F(S::BB + 4 * a + 3);
#pragma unroll 1
                for (int b = a + 1; b < E; ++b, ++pidx) {
                    const T gx = tmax(T(0), tmax(alx - s.F(S::BB + 4 * b + 1), s.F(S::BB + 4 * b) - ahx));
                    const T gy = tmax(T(0), tmax(aly - s.F(S::BB + 4 * b + 3), s.F(S::BB + 4 * b + 2) - ahy));
                    const bool near = fm(gx, gx, gy * gy) <= (b == PK ? thr_pk : thr_pp);
                    pb.set_if(near, pidx);
                    if constexpr (kNoise) need |= near ? (1u << a) | (1u << b) : 0u;
                }
            }
        }
    }
    // commit column 50 (:153-157)
    RGS_DEV void step_commit() const {
#pragma unroll 1
        for (int e = 0; e < E; ++e) {
            T px, py, vx, vy;
            eval<true>(e, kGrid, px, py, vx, vy);
            s.F(S::SP + 4 * e) = px; s.F(S::SP + 4 * e + 1) = py; s.F(S::SP + 4 * e + 2) = vx; s.F(S::SP + 4 * e + 3) = vy;
        }
    }

};

// trajectory record [t, state (reference order), damage] after step k
// (compiled once and CALLED: the rollout kernels are instruction-fetch-bound with a cliff at 64 KB of code,
// profiles/r02_ab_variants.md, and a non-logging launch never comes here)
#ifdef RGS_HOST_EMU
#define RGS_DEV_CALLED inline
#else
#define RGS_DEV_CALLED __device__ __noinline__
#endif
template <typename T, int NT, bool kFieldMajor>
RGS_DEV_CALLED void write_traj_lane_as(const RunArgs &A, const LaneScratch<T, NT> s, int64_t game, int64_t rec_idx, double t_clock,
                                       double sa, double sb) {
    constexpr int E = 2 * NT + 1, PK = 2 * NT, SREC = 2 + 4 * E, NP = 2 * NT, TREC = 1 + SREC + NP * NP;
    using S = LaneScratch<T, NT>;
    // record rec_idx of this game (no 64-bit divisions here); fs = distance between two fields of a record:
    // 1 in the records layout (compile-time, so the field offsets fold into the store instructions), the batch's game
    // count in the field-major layout (the 32 lanes of a warp then store 256 contiguous bytes per field)
    const int64_t fs = kFieldMajor ? A.traj_field_stride : 1;
    double *rec = kFieldMajor ? A.traj + rec_idx * TREC * fs + game : A.traj + (game * A.traj_recs + rec_idx) * TREC;
    // The damage record comes from global memory: its loads are issued FIRST, a block of 16 doubles at a time, and
    // consumed after the header and state stores, so the lane waits for one round trip per block at most (a rolled
    // load-store loop cost the every-step logging run 29 % of its throughput).  The damage record is 16-byte aligned.
    struct alignas(16) D2 { double x, y; };   // one 128-bit load
    const D2 *dsrc = reinterpret_cast<const D2 *>(A.damage + game * (NP * NP));
    double *dst = rec + (1 + SREC) * fs;
    constexpr int kPairs = NP * NP / 2, kBlock = kPairs < 8 ? kPairs : 8;
    D2 v[kBlock];
#pragma unroll
    for (int i = 0; i < kBlock; ++i) v[i] = dsrc[i];
    rec[0] = t_clock;
    rec[fs] = sa;
    rec[2 * fs] = sb;
#pragma unroll 1
    for (int slot = 0; slot < E; ++slot) {
        const int e = slot == 0 ? PK : slot - 1;
#pragma unroll
        for (int c = 0; c < 4; ++c) rec[(3 + 4 * slot + c) * fs] = (double)s.F(S::SP + 4 * e + c);
    }
#pragma unroll 1
    for (int d0 = 0;;) {
#pragma unroll
        for (int i = 0; i < kBlock; ++i)
            if (d0 + i < kPairs) { dst[2 * (d0 + i) * fs] = v[i].x; dst[(2 * (d0 + i) + 1) * fs] = v[i].y; }
        d0 += kBlock;
        if (d0 >= kPairs) break;
#pragma unroll
        for (int i = 0; i < kBlock; ++i)
            if (d0 + i < kPairs) v[i] = dsrc[d0 + i];
    }
}
template <typename T, int NT>
RGS_DEV void write_traj_lane(const RunArgs &A, const LaneScratch<T, NT> &s, int64_t game, int64_t rec_idx, double t_clock,
                             double sa, double sb) {
    if (A.traj_field_stride > 1) write_traj_lane_as<T, NT, true>(A, s, game, rec_idx, t_clock, sa, sb);
    else write_traj_lane_as<T, NT, false>(A, s, game, rec_idx, t_clock, sa, sb);
}

// all n_steps of ONE game (the body of a lane).  The game lives in the lane's scratch record (shared memory)
// for the whole launch; HBM is read once at the start and written once at the end (plus the command record every
// `hold` steps and the optional strided trajectory).
// kPolicy: the commands come from the on-device surrogate policy (rgs_policy.cuh), evaluated on the state at the
// start of EVERY game-step (the reference calls the team methods once per step, py/MonteCarlo.py:37-46); a separate
// instantiation, so that the open-loop kernels do not carry its code.
template <typename T, int NT, bool kPolicy = false, bool kNoise = false>
RGS_DEV void rollout_lane_game(const FastParams<T> &P, const RunArgs &A, const LaneTab<T> *tab, const LaneScratch<T, NT> &s,
                               int64_t game) {
    constexpr int E = 2 * NT + 1, PK = 2 * NT, SREC = 2 + 4 * E, NP = 2 * NT;
    using S = LaneScratch<T, NT>;
    const LaneGame<T, NT, kNoise> G(P, tab, s);
    if constexpr (kNoise) G.nz_game = A.noise_game0 + game;
    double *srec = A.state + game * SREC;
#pragma unroll 1
    for (int slot = 0; slot < E; ++slot) {
        const int e = slot == 0 ? PK : slot - 1;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const T v = T(srec[2 + 4 * slot + c]);
            s.F(S::SP + 4 * e + c) = v;
            s.F(S::PRV + 4 * e + c) = v;
        }
        s.F(S::CMD + 2 * e) = T(0);
        s.F(S::CMD + 2 * e + 1) = T(0);
    }
    double sa = srec[0], sb = srec[1];
    double t_clock = A.clock[game];
    // event record: the counters of this launch are 32-bit words of the scratch record (updated from divergent code
    // with run-time indices), added to the game's int64 record at the end; the hash is carried over
    long long *gev = A.events + game * RGS_N_EVENTS;
#pragma unroll 1
    for (int i = 0; i < RGS_N_EVENTS; ++i) G.EVW(i) = 0;
    {
        const unsigned long long h0 = (unsigned long long)gev[RGS_EV_HASH];
        G.ev_hash = h0;
    }
    double *dmg = A.damage + game * (NP * NP);
    long long live = gev[RGS_EV_LIVE_STEPS], loop_iters = gev[RGS_EV_LOOP_ITERS], decided = gev[RGS_EV_DECIDED_STEP];
    const T *actions = static_cast<const T *>(A.actions);
    const double wsc = (double)P.d.winning_score;
    // command record index without a division per step
    int64_t aidx = A.synth ? A.step0 / A.hold : 0;
    int64_t phase = A.synth ? A.step0 % A.hold : 0;
    bool fetch = true;
    const bool logging = A.traj != nullptr && A.log_stride > 0;
    int64_t k = 0, log_left = A.log_stride, log_rec = 0;
#pragma unroll 1
    for (; k < A.n_steps; ++k) {
        if (!(sa < wsc && sb < wsc)) break;  // undecided(), :57-60
        if constexpr (kPolicy) {
            T u[2 * NP];
            surrogate_policy<T, NT>(P.d, A.pol, [&](int e, int c) { return s.F(S::SP + 4 * e + c); }, u);
#pragma unroll
            for (int c = 0; c < 2 * NP; ++c) s.F(S::CMD + c) = u[c];
        } else if (fetch) {
            const int64_t ai = (A.synth || aidx < A.n_act) ? aidx : A.n_act - 1;
#pragma unroll 1
            for (int c = 0; c < 2 * NP; ++c)
                s.F(S::CMD + c) = A.synth ? T(synth_action(A.seed, A.first_game + game, ai, c, A.u_max))
                                          : actions[game * A.act_game_stride + ai * (2 * NP) + c];
            fetch = false;
        }
        if (++phase == A.hold) { phase = 0; ++aidx; fetch = true; }
        loop_iters += G.step(t_clock, sa, sb, dmg, live);
        t_clock += (double)P.d.dt;  // t_ += dt_ (:158), fp64 like the reference clock
        ++live;
        if (decided == 0 && !(sa < wsc && sb < wsc)) decided = live;
        if (logging && --log_left == 0) {
            log_left = A.log_stride;
            write_traj_lane<T, NT>(A, s, game, log_rec++, t_clock, sa, sb);
        }
    }
    if (logging)  // frozen steps of a decided game (GameSim::run :121-123) repeat the frozen state
        for (; k < A.n_steps; ++k)
            if (--log_left == 0) { log_left = A.log_stride; write_traj_lane<T, NT>(A, s, game, log_rec++, t_clock, sa, sb); }
    srec[0] = sa;
    srec[1] = sb;
#pragma unroll 1
    for (int slot = 0; slot < E; ++slot) {
        const int e = slot == 0 ? PK : slot - 1;
#pragma unroll
        for (int c = 0; c < 4; ++c) srec[2 + 4 * slot + c] = (double)s.F(S::SP + 4 * e + c);
    }
    A.clock[game] = t_clock;
#pragma unroll 1
    for (int i = 0; i < RGS_N_EVENTS; ++i) {
        long long v;
        if (i == RGS_EV_LIVE_STEPS) v = live;
        else if (i == RGS_EV_LOOP_ITERS) v = loop_iters;
        else if (i == RGS_EV_DECIDED_STEP) v = decided;
        else if (i == RGS_EV_HASH)
            v = (long long)G.ev_hash;
        else v = gev[i] + (long long)G.EVW(i);
        gev[i] = v;
    }
}

}  // namespace rgs
