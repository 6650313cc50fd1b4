// rgs_phase.cuh -- K2 (phased block body): the throughput formulation of rgs_lane.cuh, re-scheduled
// so that the rare, irregular event work of many games is COMPACTED across a thread block.
//
// Why: the B200 profile of the one-lane-per-game kernel (profiles/r01_ncu_lane_v2_full.md) shows 8.1
// of 32 lanes active on average.  69 % of the game-steps are contact-free, but in every warp-step a
// few of the 32 games do have a wall or pair item to search or an event to resolve, each with its
// own trip counts, so 78 % of the warp-instructions run with 1-4 lanes.  Here
//   * the game that a thread owns lives in REGISTERS; the contact-free step (bounding intervals,
//     broad-phase flags, closed-form commit of column 50) never touches shared memory;
//   * a game whose broad phase flags something takes a scratch RECORD from a block-shared pool
//     (same word-major record as rgs_lane.cuh) and posts its narrow-phase ITEMS to block-wide lists;
//   * the block then runs phases separated by __syncthreads(): SEARCH (thread = item), POP (thread =
//     owning slot: queue push / pop, end of step), RESOLVE (thread = popped event: impulse, wall
//     flip, goal, correctOverlap marches, re-flag -> new items);
//   * after kRounds rounds a game that is still mid-step (multi-event steps, jams: 0.03 % of the
//     steps run the full 50-iteration loop) simply continues in the next block-step while the other
//     slots move on: slots are not in lockstep, every slot walks its own sequence of games.
// The arithmetic, the event semantics and every device function of the narrow phase / resolution are
// the ones of rgs_lane.cuh (LaneGame): results are bit-identical to rollout_lane_game, which
// tests/test_emu_kernels.py and tests/test_gpu_parity.py check.  (The per-owner candidate fold is a
// minimum over the total order (column, time, candidate rank), so the order in which the items of
// a game are searched does not matter.)
#pragma once
#include "rgs_lane.cuh"

namespace rgs {

#ifdef RGS_HOST_EMU
// the emulator steps the threads of a block one after the other inside a phase
inline int ph_atomic_add(int *p, int v) { const int o = *p; *p = o + v; return o; }
inline bool ph_lock(int *) { return true; }
inline void ph_unlock(int *) {}
#else
RGS_DEV int ph_atomic_add(int *p, int v) { return atomicAdd(p, v); }
RGS_DEV bool ph_lock(int *l) {
    if (atomicCAS(l, 0, 1) != 0) return false;
    __threadfence_block();
    return true;
}
RGS_DEV void ph_unlock(int *l) {
    __threadfence_block();
    atomicExch(l, 0);
}
#endif

// record words beyond LaneScratch's (same word-major layout)
template <typename T, int NT>
struct PhaseRec {
    using S = LaneScratch<T, NT>;
    static constexpr int ET = S::NF;  // time of the popped event
    static constexpr int NF = S::NF + 1;
    static constexpr int QMASK = S::NI, SEQ = S::NI + 1, BASE = S::NI + 2, RI = S::NI + 3, RJ = S::NI + 4, SCORED = S::NI + 5,
                         COUNTER = S::NI + 6, FMASK = S::NI + 7, OWNERS = S::NI + 8, GAME = S::NI + 9, LIVE = S::NI + 10,
                         LOCK = S::NI + 11, EVT = S::NI + 12, GOAL = S::NI + 13, DONE = S::NI + 14;
    static constexpr int NI = S::NI + 15;
};

// block-shared state (pointers into dynamic shared memory; a plain heap block on the emulator)
template <typename T, int NT>
struct PhaseShared {
    const LaneTab<T> *tab;
    T *fw;            // [PhaseRec::NF][n_rec]
    int *iw;          // [PhaseRec::NI][n_rec]
    int n_rec;
    int *free_stack;  // [n_rec] record ids
    int *ctr;         // [8]: 0 free_top, 1 n_wall, 2 n_pair, 3 n_wall_ev, 4 n_pair_ev
    int *wall_items, *pair_items;  // [cap_items] each
    int cap_items;
    int *wall_ev, *pair_ev;        // [n_rec] each: record ids
    enum { FREE_TOP = 0, N_WALL = 1, N_PAIR = 2, N_WALL_EV = 3, N_PAIR_EV = 4 };
    RGS_DEV LaneScratch<T, NT> rec(int r) const { return LaneScratch<T, NT>{fw + r, iw + r, n_rec}; }

    static size_t bytes(int n_rec, int cap_items) {
        size_t b = (sizeof(LaneTab<T>) + 15) & ~size_t(15);
        b += (size_t)PhaseRec<T, NT>::NF * n_rec * sizeof(T);
        b += (size_t)PhaseRec<T, NT>::NI * n_rec * sizeof(int);
        b += (size_t)(n_rec + 8 + 2 * cap_items + 2 * n_rec) * sizeof(int);
        return (b + 15) & ~size_t(15);
    }
    // carve the block out of `raw` (16-byte aligned)
    RGS_DEV void carve(unsigned char *raw, int n_rec_, int cap_items_) {
        n_rec = n_rec_;
        cap_items = cap_items_;
        tab = reinterpret_cast<const LaneTab<T> *>(raw);
        raw += (sizeof(LaneTab<T>) + 15) & ~size_t(15);
        fw = reinterpret_cast<T *>(raw);
        raw += (size_t)PhaseRec<T, NT>::NF * n_rec * sizeof(T);
        iw = reinterpret_cast<int *>(raw);
        int *p = iw + (size_t)PhaseRec<T, NT>::NI * n_rec;
        free_stack = p; p += n_rec;
        ctr = p; p += 8;
        wall_items = p; p += cap_items;
        pair_items = p; p += cap_items;
        wall_ev = p; p += n_rec;
        pair_ev = p;
    }
};

// what a thread keeps in registers for the game slot it owns
template <typename T, int NT>
struct PhaseSlot {
    static constexpr int E = 2 * NT + 1, NP = 2 * NT;
    T st[4 * E];    // [4e+c] state at the step boundary, entity order (players 0..2n-1, puck 2n)
    T cmd[2 * NP];  // commanded velocities of the players
    double sa, sb, t_clock;
    long long live, loop_iters, decided;
    int64_t game, next_game, k, aidx, phase, log_left;
    int rec;        // scratch record while mid-step, else -1
    int held;       // record kept after a step ended in HEAD (the pool is only popped in HEAD and only pushed in POP)
    bool fetch;
};

template <typename T, int NT>
struct PhaseKernel {
    static constexpr int E = 2 * NT + 1, PK = 2 * NT, NP = 2 * NT, NPAIR = E * (E - 1) / 2, SREC = 2 + 4 * E;
    using S = LaneScratch<T, NT>;
    using R = PhaseRec<T, NT>;
    using Sh = PhaseShared<T, NT>;
    using Slot = PhaseSlot<T, NT>;
    using Game = LaneGame<T, NT>;

    const FastParams<T> &P;
    const RunArgs &A;
    const Sh &sh;
    RGS_DEV PhaseKernel(const FastParams<T> &P_, const RunArgs &A_, const Sh &sh_) : P(P_), A(A_), sh(sh_) {}

    // ---- record pool (allocation and release never run in the same phase) ----
    RGS_DEV int alloc() const {
        const int old = ph_atomic_add(&sh.ctr[Sh::FREE_TOP], -1);
        if (old <= 0) {
            ph_atomic_add(&sh.ctr[Sh::FREE_TOP], 1);
            return -1;
        }
        return sh.free_stack[old - 1];
    }
    RGS_DEV void release(int r) const { sh.free_stack[ph_atomic_add(&sh.ctr[Sh::FREE_TOP], 1)] = r; }

    // ---- item lists: (rec | a << 10 | (wall or b) << 14 | base << 18) ----
    RGS_DEV void fold_locked(const Game &G, int e, int c, T t, int kind, int j) const {
        int *lock = &G.s.I(R::LOCK);
        bool done = false;
        while (!done) {
            if (ph_lock(lock)) {
                unsigned fm_ = (unsigned)G.s.I(R::FMASK);
                G.fold(e, c, t, kind, j, fm_);
                G.s.I(R::FMASK) = (int)fm_;
                ph_unlock(lock);
                done = true;
            }
        }
    }
    RGS_DEV_BIG void do_wall_item(int item) const {
        const int r = item & 1023, a = (item >> 10) & 15, wk = (item >> 14) & 15, base = (item >> 18) & 63;
        const Game G(P, sh.tab, sh.rec(r));
        T t;
        const int c = G.search_wall(a, wk, base, t);
        if (c <= kGrid) fold_locked(G, a, c, t, wk, 15);
    }
    RGS_DEV_BIG void do_pair_item(int item) const {
        const int r = item & 1023, a = (item >> 10) & 15, b = (item >> 14) & 15, base = (item >> 18) & 63;
        const Game G(P, sh.tab, sh.rec(r));
        T t;
        const int c = G.search_pair(a, b, base, t);
        if (c <= kGrid) {
            const unsigned owners = (unsigned)G.s.I(R::OWNERS);
            if ((owners >> a) & 1u) fold_locked(G, a, c, t, PAIR, b);
            if ((owners >> b) & 1u) fold_locked(G, b, c, t, PAIR, a);
        }
    }
    // post the flagged items of record r; an item that does not fit is searched on the spot
    RGS_DEV void post_items(int r, Bits<4 * E> &wb, Bits<NPAIR> &pb, int base) const {
        int bit;
#pragma unroll 1
        while ((bit = wb.pop()) >= 0) {
            const int item = r | ((bit >> 2) << 10) | ((bit & 3) << 14) | (base << 18);
            const int idx = ph_atomic_add(&sh.ctr[Sh::N_WALL], 1);
            if (idx < sh.cap_items) sh.wall_items[idx] = item; else do_wall_item(item);
        }
#pragma unroll 1
        while ((bit = pb.pop()) >= 0) {
            int a, b;
            Game::pair_decode(bit, a, b);
            const int item = r | (a << 10) | (b << 14) | (base << 18);
            const int idx = ph_atomic_add(&sh.ctr[Sh::N_PAIR], 1);
            if (idx < sh.cap_items) sh.pair_items[idx] = item; else do_pair_item(item);
        }
    }

    // ---- slot <-> HBM (rollout boundaries only) ----
    RGS_DEV void load_game(Slot &ts) const {
        ts.game = ts.next_game;
        const double *srec = A.state + ts.game * SREC;
#pragma unroll
        for (int slot = 0; slot < E; ++slot) {
            const int e = slot == 0 ? PK : slot - 1;
#pragma unroll
            for (int c = 0; c < 4; ++c) ts.st[4 * e + c] = T(srec[2 + 4 * slot + c]);
        }
#pragma unroll
        for (int c = 0; c < 2 * NP; ++c) ts.cmd[c] = T(0);
        ts.sa = srec[0];
        ts.sb = srec[1];
        ts.t_clock = A.clock[ts.game];
        const long long *gev = A.events + ts.game * RGS_N_EVENTS;
        ts.live = gev[RGS_EV_LIVE_STEPS];
        ts.loop_iters = gev[RGS_EV_LOOP_ITERS];
        ts.decided = gev[RGS_EV_DECIDED_STEP];
        ts.aidx = A.synth ? A.step0 / A.hold : 0;
        ts.phase = A.synth ? A.step0 % A.hold : 0;
        ts.fetch = true;
        ts.k = 0;
        ts.log_left = A.log_stride;
    }
    RGS_DEV void write_traj(const Slot &ts, int64_t k) const {
        constexpr int TREC = 1 + SREC + NP * NP;
        double *rec = A.traj + (ts.game * (A.n_steps / A.log_stride) + ((k + 1) / A.log_stride - 1)) * TREC;
        rec[0] = ts.t_clock;
        rec[1] = ts.sa;
        rec[2] = ts.sb;
#pragma unroll
        for (int slot = 0; slot < E; ++slot) {
            const int e = slot == 0 ? PK : slot - 1;
#pragma unroll
            for (int c = 0; c < 4; ++c) rec[3 + 4 * slot + c] = (double)ts.st[4 * e + c];
        }
#pragma unroll 1
        for (int d = 0; d < NP * NP; ++d) rec[1 + SREC + d] = A.damage[ts.game * (NP * NP) + d];
    }
    RGS_DEV void store_game(Slot &ts) const {
        const bool logging = A.traj != nullptr && A.log_stride > 0;
        if (logging)  // frozen steps of a decided game (GameSim::run :121-123) repeat the frozen state
            for (; ts.k < A.n_steps; ++ts.k)
                if (--ts.log_left == 0) { ts.log_left = A.log_stride; write_traj(ts, ts.k); }
        double *srec = A.state + ts.game * SREC;
        srec[0] = ts.sa;
        srec[1] = ts.sb;
#pragma unroll
        for (int slot = 0; slot < E; ++slot) {
            const int e = slot == 0 ? PK : slot - 1;
#pragma unroll
            for (int c = 0; c < 4; ++c) srec[2 + 4 * slot + c] = (double)ts.st[4 * e + c];
        }
        A.clock[ts.game] = ts.t_clock;
        long long *gev = A.events + ts.game * RGS_N_EVENTS;
        gev[RGS_EV_LIVE_STEPS] = ts.live;
        gev[RGS_EV_LOOP_ITERS] = ts.loop_iters;
        gev[RGS_EV_DECIDED_STEP] = ts.decided;
        ts.game = -1;
    }
    RGS_DEV bool undecided(const Slot &ts) const {
        const double wsc = (double)P.d.winning_score;
        return ts.sa < wsc && ts.sb < wsc;  // :57-60
    }
    // bookkeeping at the end of a game-step whose event loop ran `counter` iterations
    RGS_DEV void finish_step(Slot &ts, int counter) const {
        ts.loop_iters += counter;
        ts.t_clock += (double)P.d.dt;  // t_ += dt_ (:158), fp64 like the reference clock
        ++ts.live;
        if (ts.decided == 0 && !undecided(ts)) ts.decided = ts.live;
        if (A.traj != nullptr && A.log_stride > 0 && --ts.log_left == 0) {
            ts.log_left = A.log_stride;
            write_traj(ts, ts.k);
        }
        ++ts.k;
        if (++ts.phase == A.hold) { ts.phase = 0; ++ts.aidx; ts.fetch = true; }
    }

    // ---- HEAD (thread = slot): start the next game-step of the slot's game.  A contact-free step
    // (no broad-phase flag) is finished here, in registers. ----
    RGS_DEV void head(Slot &ts, int64_t game_stride) const {
        if (ts.rec >= 0) {
            if (!sh.rec(ts.rec).I(R::DONE)) return;  // still inside a step
            end_step(ts, true);
        }
        for (;;) {
            if (ts.game < 0) {
                if (ts.next_game >= A.n_games) return;
                load_game(ts);
                ts.next_game += game_stride;
            }
            if (ts.k < A.n_steps && undecided(ts)) break;
            store_game(ts);
        }
        if (ts.fetch) {
            const T *actions = static_cast<const T *>(A.actions);
            const int64_t ai = (A.synth || ts.aidx < A.n_act) ? ts.aidx : A.n_act - 1;
#pragma unroll
            for (int c = 0; c < 2 * NP; ++c)
                ts.cmd[c] = A.synth ? T(synth_action(A.seed, A.first_game + ts.game, ai, c, A.u_max))
                                    : actions[ts.game * A.act_game_stride + ai * (2 * NP) + c];
            ts.fetch = false;
        }
        // bounding intervals + broad-phase flags in registers; a contact-free step ends here
        const Game G0(P, sh.tab, sh.rec(0));  // the register helpers touch no record
        T bb[4 * E];
        Bits<4 * E> wb;
        Bits<NPAIR> pb;
        G0.head_flags(ts.st, ts.cmd, bb, wb, pb);
        if (!(wb.any() || pb.any())) {
            G0.quiet_commit(ts.st, ts.cmd);
            finish_step(ts, 1);
            return;
        }
        const int r = ts.held >= 0 ? ts.held : alloc();
        if (r < 0) return;  // pool exhausted: this slot retries the step in the next block-step
        ts.held = -1;
        ts.rec = r;
        const S s = sh.rec(r);
        Game(P, sh.tab, s).spill(ts.st, ts.cmd, bb);
        s.I(R::QMASK) = 0;
        s.I(R::SEQ) = 0;
        s.I(R::BASE) = 1;
        s.I(R::RI) = -1;
        s.I(R::RJ) = -1;
        s.I(R::SCORED) = 0;
        s.I(R::COUNTER) = 1;
        s.I(R::FMASK) = 0;
        s.I(R::OWNERS) = (int)((1u << E) - 1u);
        s.I(R::GAME) = (int)ts.game;
        s.I(R::LIVE) = (int)ts.live;
        s.I(R::LOCK) = 0;
        s.I(R::GOAL) = 0;
        s.I(R::DONE) = 0;
        post_items(r, wb, pb, 1);
    }

    // ---- SEARCH (thread = item): narrow phase of every posted item ----
    RGS_DEV void search(int tid, int n_threads) const {
        const int nw = sh.ctr[Sh::N_WALL] < sh.cap_items ? sh.ctr[Sh::N_WALL] : sh.cap_items;
        const int np = sh.ctr[Sh::N_PAIR] < sh.cap_items ? sh.ctr[Sh::N_PAIR] : sh.cap_items;
#pragma unroll 1
        for (int i = tid; i < nw; i += n_threads) do_wall_item(sh.wall_items[i]);
#pragma unroll 1
        for (int i = tid; i < np; i += n_threads) do_pair_item(sh.pair_items[i]);
    }

    // ---- the queue side of the event loop (:140-150), shared by POP and RESOLVE: with nothing left to
    // search, either the step is over (queue empty or loop bound) or the earliest event is popped.
    // Returns true when an event is ready in EVT / ET, false when the step is over (DONE is set).
    RGS_DEV bool next_event(const Game &G, long long *gev, unsigned &qmask, int &counter) const {
        const S &s = G.s;
#pragma unroll 1
        for (;;) {
            if (qmask == 0 || counter >= kCounterLim) {
                if (qmask != 0) gev[RGS_EV_COUNTER_LIMIT]++;
                s.I(R::DONE) = 1;
                return false;
            }
            int ei, ej, ekind;
            T et;
            G.pop(qmask, ei, ej, ekind, et);
            s.I(R::RI) = ei;  // nothing to re-scan unless the event executes
            s.I(R::RJ) = -1;
            s.I(R::OWNERS) = 0;
            if ((ei == PK || ej == PK) && s.I(R::SCORED)) {  // :230-231
                gev[RGS_EV_DISCARDED]++;
                hash_event(gev, (long long)s.I(R::LIVE), DISCARD, ei, ej, 0);
                ++counter;
                continue;
            }
            s.I(R::EVT) = ei | ((ej + 1) << 4) | (ekind << 8);
            s.F(R::ET) = et;
            return true;
        }
    }

    // end of a step (owner thread): commit column 50 (:153-157) into the slot's registers
    RGS_DEV void end_step(Slot &ts, bool in_head) const {
        const S s = sh.rec(ts.rec);
        const Game G(P, sh.tab, s);
#pragma unroll
        for (int e = 0; e < E; ++e) G.eval(e, kGrid, ts.st[4 * e], ts.st[4 * e + 1], ts.st[4 * e + 2], ts.st[4 * e + 3]);
        const int goal = s.I(R::GOAL);
        if (goal == 1) ts.sa += 1.0;
        if (goal == 2) ts.sb += 1.0;
        const int counter = s.I(R::COUNTER);
        if (in_head) ts.held = ts.rec; else release(ts.rec);
        ts.rec = -1;
        finish_step(ts, counter);
    }

    // ---- POP (thread = owning slot): queue the candidates the searches found, then either end the
    // step or hand the earliest event to RESOLVE ----
    // first thing after the barrier that ends HEAD (the pool is popped in HEAD only)
    RGS_DEV void release_held(Slot &ts) const {
        if (ts.held >= 0) {
            release(ts.held);
            ts.held = -1;
        }
    }
    RGS_DEV void pop_phase(Slot &ts) const {
        if (ts.rec < 0) return;
        const S s = sh.rec(ts.rec);
        if (s.I(R::DONE)) return;  // ended by RESOLVE; the slot commits in its next HEAD
        const Game G(P, sh.tab, s);
        long long *gev = A.events + ts.game * RGS_N_EVENTS;
        unsigned qmask = (unsigned)s.I(R::QMASK);
        int seq = s.I(R::SEQ), counter = s.I(R::COUNTER);
        const unsigned fmask = (unsigned)s.I(R::FMASK);
        s.I(R::FMASK) = 0;
        const int ri = s.I(R::RI), rj = s.I(R::RJ);
        if (ri < 0) {
#pragma unroll 1
            for (unsigned m = fmask; m != 0u; m &= m - 1u) G.push(low_bit(m), qmask, seq);  // ascending owners
        } else {
            if ((fmask >> ri) & 1u) G.push(ri, qmask, seq);
            if (rj >= 0 && ((fmask >> rj) & 1u)) G.push(rj, qmask, seq);
        }
        const bool ev_ready = next_event(G, gev, qmask, counter);
        s.I(R::QMASK) = (int)qmask;
        s.I(R::SEQ) = seq;
        s.I(R::COUNTER) = counter;
        if (!ev_ready) {
            end_step(ts, false);
            return;
        }
        const int ekind = (s.I(R::EVT) >> 8) & 15;
        if (ekind == PAIR) sh.pair_ev[ph_atomic_add(&sh.ctr[Sh::N_PAIR_EV], 1)] = ts.rec;
        else sh.wall_ev[ph_atomic_add(&sh.ctr[Sh::N_WALL_EV], 1)] = ts.rec;
    }

    // ---- RESOLVE (thread = popped event).  When the re-scan flags nothing the same thread goes on
    // with the queue (next event, or end of step: three steps out of four with a contact end here) ----
    RGS_DEV_BIG void resolve_record(int r) const {
        const S s = sh.rec(r);
        const Game G(P, sh.tab, s);
        const int64_t game = s.I(R::GAME);
        long long *gev = A.events + game * RGS_N_EVENTS;
        double *dmg = A.damage + game * (NP * NP);
        unsigned qmask = (unsigned)s.I(R::QMASK);
        int counter = s.I(R::COUNTER), base = s.I(R::BASE);
#pragma unroll 1
        for (;;) {
            const int evt = s.I(R::EVT);
            const int ei = evt & 15, ej = ((evt >> 4) & 15) - 1, ekind = (evt >> 8) & 15;
            int goal = 0;
            const int rj = G.resolve(ei, ej, ekind, s.F(R::ET), base, goal, gev, dmg, (long long)s.I(R::LIVE));
            if (goal) {
                s.I(R::GOAL) = goal;
                s.I(R::SCORED) = 1;
            }
            G.prune(qmask, ei, ej);  // :347-352
            // re-scan the bodies involved (:355-357, :147): broad phase on their new segments
            Bits<4 * E> wb;
            Bits<NPAIR> pb;
            wb.clear();
            pb.clear();
            unsigned owners = 1u << ei;
            G.flag_walls(ei, wb);
            if (rj >= 0) {
                owners |= 1u << rj;
                G.flag_walls(rj, wb);
            }
            G.reflag_pairs(ei, -1, pb);
            if (rj >= 0) G.reflag_pairs(rj, ei, pb);
            ++counter;
            if (wb.any() || pb.any()) {
                s.I(R::RJ) = rj;
                s.I(R::OWNERS) = (int)owners;
                s.I(R::QMASK) = (int)qmask;
                s.I(R::COUNTER) = counter;
                s.I(R::BASE) = base;
                post_items(r, wb, pb, base);
                return;
            }
            if (!next_event(G, gev, qmask, counter)) break;
        }
        s.I(R::QMASK) = (int)qmask;
        s.I(R::COUNTER) = counter;
        s.I(R::BASE) = base;
    }
    RGS_DEV void resolve_phase(int tid, int n_threads) const {
        const int nw = sh.ctr[Sh::N_WALL_EV], np = sh.ctr[Sh::N_PAIR_EV];
#pragma unroll 1
        for (int i = tid; i < nw; i += n_threads) resolve_record(sh.wall_ev[i]);
#pragma unroll 1
        for (int i = tid; i < np; i += n_threads) resolve_record(sh.pair_ev[i]);
    }

    RGS_DEV bool slot_done(const Slot &ts) const { return ts.rec < 0 && ts.game < 0 && ts.next_game >= A.n_games; }
    RGS_DEV void init_slot(Slot &ts, int64_t first_game) const {
        ts.game = -1;
        ts.next_game = first_game;
        ts.rec = -1;
        ts.held = -1;
    }
};

// Block schedule (the __global__ wrapper in rgs_phase_kernel.cu and the sequential emulator driver
// in tests/emu follow it literally):
//   repeat until every slot is done:
//     HEAD | sync; release_held | for round in 0..kPhaseRounds-1: { sync; SEARCH; sync; [reset item counts] POP; sync;
//     RESOLVE; sync [reset event counts]; leave the round loop when no slot holds a record }
constexpr int kPhaseRounds = 1;

}  // namespace rgs
