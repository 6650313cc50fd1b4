/* oracle/gamesim_oracle.c -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Plain-C restatement of the reference GameSim step; see gamesim_oracle.h.  Every function cites
 * the lines of /root/reference/src/GameSim.cpp it follows.  All arithmetic is IEEE-754 double,
 * one rounding per source operator in the reference's operator order (build with
 * -ffp-contract=off, no -march: no FMA), so that states match the compiled reference bit for bit.
 *
 * Body index order is the reference's entity order [A1..An, B1..Bn, PK] (GameSim.cpp:11); the
 * state-vector order is [TAS, TBS, PK, A1..An, B1..Bn] (include/SimState.h:14-23).
 */
#include "gamesim_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

enum { K_WALL_UP = 0, K_WALL_DOWN = 1, K_WALL_RIGHT = 2, K_WALL_LEFT = 3, K_PAIR = 4, K_DISCARD = 5 };
enum { PX = 0, PY = 1, VX = 2, VY = 3 };

typedef struct {
    double t; /* absolute time of impact */
    int i;    /* owner body */
    int j;    /* other body, -1 for walls */
    int kind;
} event_t;

typedef struct {
    const orc_params *p;
    int n, E, PK;
    double dt_col;
    double score[2];
    double t;
    double x[ORC_MAX_BODIES][4];
    double u[ORC_MAX_BODIES][2];
    double *damage; /* [(2n)^2] row-major (aggressor, victim) or NULL */
    int64_t *ev;    /* [ORC_N_EVENTS] or NULL */
    int64_t scratch_ev[ORC_N_EVENTS];
    /* per-step scratch: the 51-column collision grid (GameSim.h:46) */
    double T[ORC_GRID + 1];
    double g[ORC_MAX_BODIES][ORC_GRID + 1][4];
    int scored;
    event_t q[ORC_MAX_BODIES + 1]; /* vector order: descending t, earliest at the END */
    int nq;
} game_t;

static inline double radius_of(const game_t *G, int e) { return e == G->PK ? G->p->puck_radius : G->p->player_radius; }
static inline double mass_of(const game_t *G, int e) { return e == G->PK ? G->p->puck_mass : G->p->player_mass; }

static inline void hash_event(game_t *G, int kind, int i, int j, int goal) {
    uint64_t h = (uint64_t)G->ev[ORC_EV_HASH];
    uint64_t w = ((uint64_t)G->ev[ORC_EV_LIVE_STEPS] << 16) | ((uint64_t)kind << 12) | ((uint64_t)i << 8) |
                 ((uint64_t)(j < 0 ? 15 : j) << 4) | (uint64_t)goal;
    h ^= w;
    h *= 0x100000001B3ULL;
    G->ev[ORC_EV_HASH] = (int64_t)h;
}

/* f_player / f_puck (GameSim.cpp:62-68, 90-96); process noise must be zero (its draws are
 * multiplied by w_stdev_ = 0 in every shipped script and cannot change a result). */
static inline void deriv(const game_t *G, int e, const double x[4], const double u[2], double k[4]) {
    k[PX] = x[VX];
    k[PY] = x[VY];
    if (e == G->PK) {
        k[VX] = (u[0] - 0.2 * x[VX]) / G->p->tau_puck;
        k[VY] = (u[1] - 0.2 * x[VY]) / G->p->tau_puck;
    } else {
        k[VX] = (u[0] - x[VX]) / G->p->tau_player;
        k[VY] = (u[1] - x[VY]) / G->p->tau_player;
    }
}

/* RK4_player / RK4_puck via simAgnostic (GameSim.cpp:70-88, 98-116, 549-555). */
static void rk4(const game_t *G, int e, const double x[4], const double u[2], double h, double out[4]) {
    double k1[4], k2[4], k3[4], k4[4], xs[4];
    const double h2 = h / 2.0;
    const double h6 = h / 6.0;
    deriv(G, e, x, u, k1);
    for (int c = 0; c < 4; ++c) xs[c] = x[c] + k1[c] * h2;
    deriv(G, e, xs, u, k2);
    for (int c = 0; c < 4; ++c) xs[c] = x[c] + k2[c] * h2;
    deriv(G, e, xs, u, k3);
    for (int c = 0; c < 4; ++c) xs[c] = x[c] + k3[c] * h;
    deriv(G, e, xs, u, k4);
    for (int c = 0; c < 4; ++c) xs[c] = x[c] + (((k1[c] + k2[c] * 2.0) + k3[c] * 2.0) + k4[c]) * h6;
    for (int c = 0; c < 4; ++c) out[c] = xs[c];
}

/* populateStateGrid (GameSim.cpp:173-217): bodies are integrated independently. */
static void fill_grid(game_t *G) {
    for (int i = 0; i <= ORC_GRID; ++i) G->T[i] = G->t + G->dt_col * i;
    for (int e = 0; e < G->E; ++e) {
        memcpy(G->g[e][0], G->x[e], sizeof(double) * 4);
        for (int i = 1; i <= ORC_GRID; ++i) rk4(G, e, G->g[e][i - 1], G->u[e], G->dt_col, G->g[e][i]);
    }
}

/* correctOverlap, pair form (GameSim.cpp:581-592): un-normalised separation vector. */
static void fix_pair_overlap(game_t *G, int i, int j, int col, double r_i, double r_j) {
    double *pi = G->g[i][col], *pj = G->g[j][col];
    const double dx = pj[PX] - pi[PX], dy = pj[PY] - pi[PY];
    const double overlap = r_i + r_j - sqrt(dx * dx + dy * dy);
    if (overlap > 0) {
        const double si = 1.1 * r_i / (r_i + r_j) * overlap;
        pi[PX] -= si * dx;
        pi[PY] -= si * dy;
        const double sj = 1.1 * r_j / (r_i + r_j) * overlap;
        pj[PX] += sj * dx;
        pj[PY] += sj * dy;
        G->ev[ORC_EV_OVERLAP_FIXES]++;
    }
}

/* correctOverlap, wall form (GameSim.cpp:594-623). */
static void fix_wall_overlap(game_t *G, int i, int col, double r_i, int wall) {
    double *p = G->g[i][col];
    double overlap = 0.0, cx = 0.0, cy = 0.0;
    switch (wall) {
    case K_WALL_UP:    overlap = p[PY] + r_i - G->p->arena_y / 2.0;  cx = 0.;  cy = -1.; break;
    case K_WALL_DOWN:  overlap = -p[PY] + r_i - G->p->arena_y / 2.0; cx = 0.;  cy = 1.;  break;
    case K_WALL_RIGHT: overlap = p[PX] + r_i - G->p->arena_x / 2.0;  cx = -1.; cy = 0.;  break;
    case K_WALL_LEFT:  overlap = -p[PX] + r_i - G->p->arena_x / 2.0; cx = 1.;  cy = 0.;  break;
    }
    if (overlap > 0) {
        const double s = 1.1 * overlap;
        p[PX] += s * cx;
        p[PY] += s * cy;
        G->ev[ORC_EV_OVERLAP_FIXES]++;
    }
}

/* Inner body of insertCollisions for one owner ID (GameSim.cpp:362-507): first grid column
 * i >= base whose candidate event lies inside this step. */
static void scan_body(game_t *G, int id, int base) {
    const orc_params *P = G->p;
    const double r = radius_of(G, id);
    for (int i = base; i <= ORC_GRID; ++i) {
        event_t best = {1.0e9, -1, -1, -1};
        const double *pk = G->g[id][i];
        const double *pm = G->g[id][i - 1]; /* column i-1 (i >= 1 always: base >= 1) */
        const double tm = G->T[i - 1];

        if (pk[PY] + r >= P->arena_y / 2.0) { /* WALL_UP :374-394 */
            if (pm[VY] > 0) {
                const double tau = fmax(0.0, (P->arena_y / 2.0 - r - pm[PY]) / pm[VY]);
                const double tc = tm + tau;
                if (tc < best.t) { best.t = tc; best.i = id; best.j = -1; best.kind = K_WALL_UP; }
            }
        }
        if (pk[PY] - r <= -P->arena_y / 2.0) { /* WALL_DOWN :397-417 */
            if (pm[VY] < 0) {
                const double tau = fmax(0.0, (-P->arena_y / 2.0 + r - pm[PY]) / pm[VY]);
                const double tc = tm + tau;
                if (tc < best.t) { best.t = tc; best.i = id; best.j = -1; best.kind = K_WALL_DOWN; }
            }
        }
        if (pk[PX] + r >= P->arena_x / 2.0) { /* WALL_RIGHT :420-440 */
            if (pm[VX] > 0) {
                const double tau = fmax(0.0, (P->arena_x / 2.0 - r - pm[PX]) / pm[VX]);
                const double tc = tm + tau;
                if (tc < best.t) { best.t = tc; best.i = id; best.j = -1; best.kind = K_WALL_RIGHT; }
            }
        }
        if (pk[PX] - r <= -P->arena_x / 2.0) { /* WALL_LEFT :443-463 */
            if (pm[VX] < 0) {
                const double tau = fmax(0.0, (-P->arena_x / 2.0 + r - pm[PX]) / pm[VX]);
                const double tc = tm + tau;
                if (tc < best.t) { best.t = tc; best.i = id; best.j = -1; best.kind = K_WALL_LEFT; }
            }
        }
        for (int o = 0; o < G->E; ++o) { /* INTER_AGENT :466-498, entity order */
            if (o == id) continue;
            const double ro = radius_of(G, o);
            const double *ok = G->g[o][i];
            const double ex = pk[PX] - ok[PX], ey = pk[PY] - ok[PY];
            if (sqrt(ex * ex + ey * ey) <= r + ro) {
                const double *om = G->g[o][i - 1];
                const double dpx = om[PX] - pm[PX], dpy = om[PY] - pm[PY];
                const double dvx = om[VX] - pm[VX], dvy = om[VY] - pm[VY];
                const double sigma = r + ro;
                const double dpdv = dpx * dvx + dpy * dvy;
                const double dvdv = dvx * dvx + dvy * dvy;
                const double dpdp = dpx * dpx + dpy * dpy;
                const double d = dpdv * dpdv - dvdv * (dpdp - sigma * sigma);
                if (dpdv < 0 && d >= 0) {
                    const double tau = fmax(0.0, -(dpdv + sqrt(d)) / dvdv);
                    const double tc = tm + tau;
                    if (tc < best.t) { best.t = tc; best.i = id; best.j = o; best.kind = K_PAIR; }
                }
            }
        }
        if (best.t - G->t < P->dt) { /* :501-505 */
            G->q[G->nq++] = best;
            return;
        }
    }
}

/* insertCollisions (GameSim.cpp:360-511).  std::sort on <= 16 elements is libstdc++'s insertion
 * sort, i.e. stable: equal times keep vector order, so the most recently pushed of a tie ends
 * up nearest the END and is executed first (include/Collision.h:24-30). */
static void insert_events(game_t *G, const int *checks, int n_checks, int base) {
    for (int c = 0; c < n_checks; ++c) scan_body(G, checks[c], base);
    for (int a = 1; a < G->nq; ++a) {
        event_t v = G->q[a];
        int b = a;
        while (b > 0 && v.t > G->q[b - 1].t) { G->q[b] = G->q[b - 1]; --b; }
        G->q[b] = v;
    }
}

/* carryOutFirstCollision (GameSim.cpp:219-358); returns the number of IDs to re-scan. */
static int resolve_first(game_t *G, int *base_idx, int *checks) {
    const orc_params *P = G->p;
    const event_t ev = G->q[--G->nq];

    if ((ev.i == G->PK || ev.j == G->PK) && G->scored) { /* :230-231 */
        G->ev[ORC_EV_DISCARDED]++;
        hash_event(G, K_DISCARD, ev.i, ev.j, 0);
        return 0;
    }

    int base = (int)floor((ev.t - G->t) / G->dt_col) + 1; /* :237 */
    if (base > ORC_GRID) { /* the reference would index column 51 (undefined behaviour) */
        base = ORC_GRID;
        G->ev[ORC_EV_BASE_CLAMP]++;
    }
    *base_idx = base;
    int goal = 0;

    if (ev.kind == K_PAIR) { /* :242-299 */
        const int i = ev.i, j = ev.j;
        const double m_i = mass_of(G, i), m_j = mass_of(G, j), r_i = radius_of(G, i), r_j = radius_of(G, j);
        const double dt_t = ev.t - G->T[base - 1];
        double xi[4], xj[4];
        rk4(G, i, G->g[i][base - 1], G->u[i], dt_t, xi);
        rk4(G, j, G->g[j][base - 1], G->u[j], dt_t, xj);

        const double pijx = xj[PX] - xi[PX], pijy = xj[PY] - xi[PY];
        const double overlap = r_i + r_j - sqrt(pijx * pijx + pijy * pijy);
        if (overlap > 0) { /* :261-265 */
            const double si = 1.01 * r_i / (r_i + r_j) * overlap;
            xi[PX] -= si * pijx;
            xi[PY] -= si * pijy;
            const double sj = 1.01 * r_j / (r_i + r_j) * overlap;
            xj[PX] += sj * pijx;
            xj[PY] += sj * pijy;
        }
        /* elastic impulse :268-274, evaluated left to right */
        const double p1x = xi[PX], p1y = xi[PY], v1x = xi[VX], v1y = xi[VY];
        const double p2x = xj[PX], p2y = xj[PY], v2x = xj[VX], v2y = xj[VY];
        {
            const double dvx = v1x - v2x, dvy = v1y - v2y, dpx = p1x - p2x, dpy = p1y - p2y;
            const double f = 2 * m_j / (m_i + m_j) * (dvx * dpx + dvy * dpy) / (dpx * dpx + dpy * dpy);
            xi[VX] = v1x - f * dpx;
            xi[VY] = v1y - f * dpy;
        }
        {
            const double dvx = v2x - v1x, dvy = v2y - v1y, dpx = p2x - p1x, dpy = p2y - p1y;
            const double f = 2 * m_i / (m_i + m_j) * (dvx * dpx + dvy * dpy) / (dpx * dpx + dpy * dpy);
            xj[VX] = v2x - f * dpx;
            xj[VY] = v2y - f * dpy;
        }
        if (i != G->PK && j != G->PK) { /* damage :277-282 */
            const double dvx = v1x - v2x, dvy = v1y - v2y;
            const double dv_sq = 0.01 * (dvx * dvx + dvy * dvy);
            if (G->damage) {
                const int np = 2 * G->n;
                G->damage[i * np + j] += m_i * dv_sq;
                G->damage[j * np + i] += m_j * dv_sq;
            }
            G->ev[ORC_EV_PLAYER_PLAYER]++;
        } else {
            G->ev[ORC_EV_PLAYER_PUCK]++;
        }
        /* re-simulate both bodies :285-297 */
        rk4(G, i, xi, G->u[i], G->dt_col - dt_t, G->g[i][base]);
        rk4(G, j, xj, G->u[j], G->dt_col - dt_t, G->g[j][base]);
        fix_pair_overlap(G, i, j, base, r_i, r_j);
        for (int c = base + 1; c <= ORC_GRID; ++c) {
            rk4(G, i, G->g[i][c - 1], G->u[i], G->dt_col, G->g[i][c]);
            rk4(G, j, G->g[j][c - 1], G->u[j], G->dt_col, G->g[j][c]);
            fix_pair_overlap(G, i, j, c, r_i, r_j);
        }
    } else { /* wall :300-343 */
        const int id = ev.i;
        const double r_i = radius_of(G, id);
        const double dt_t = ev.t - G->T[base - 1];
        double xi[4];
        rk4(G, id, G->g[id][base - 1], G->u[id], dt_t, xi);
        switch (ev.kind) { /* :312-318 */
        case K_WALL_UP:    if (xi[VY] > 0) xi[VY] *= -1.0; break;
        case K_WALL_DOWN:  if (xi[VY] < 0) xi[VY] *= -1.0; break;
        case K_WALL_RIGHT: if (xi[VX] > 0) xi[VX] *= -1.0; break;
        case K_WALL_LEFT:  if (xi[VX] < 0) xi[VX] *= -1.0; break;
        }
        if (id == G->PK && (ev.kind == K_WALL_LEFT || ev.kind == K_WALL_RIGHT) && xi[PY] < P->goal_height / 2.0 &&
            xi[PY] > -P->goal_height / 2.0) { /* goal :321-327 */
            if (!G->scored) {
                if (ev.kind == K_WALL_RIGHT) { G->score[0] += 1.0; G->ev[ORC_EV_GOALS_A]++; goal = 1; }
                else                         { G->score[1] += 1.0; G->ev[ORC_EV_GOALS_B]++; goal = 2; }
            }
            for (int c = 0; c < 4; ++c) G->g[id][base][c] = 0.0;
            G->scored = 1;
        } else {
            rk4(G, id, xi, G->u[id], G->dt_col - dt_t, G->g[id][base]);
            fix_wall_overlap(G, id, base, r_i, ev.kind);
        }
        for (int c = base + 1; c <= ORC_GRID; ++c) {
            rk4(G, id, G->g[id][c - 1], G->u[id], G->dt_col, G->g[id][c]);
            fix_wall_overlap(G, id, c, r_i, ev.kind);
        }
        G->ev[ORC_EV_WALL_UP + ev.kind]++;
    }
    hash_event(G, ev.kind, ev.i, ev.j, goal);

    /* drop queued events that share a body with this one :347-352 (order preserved) */
    int w = 0;
    for (int a = 0; a < G->nq; ++a) {
        const event_t *c = &G->q[a];
        const int drop = (ev.i == c->i) || (ev.i == c->j && c->j != -1) || (ev.j == c->i) ||
                         (ev.j == c->j && c->j != -1);
        if (!drop) G->q[w++] = *c;
    }
    G->nq = w;

    int n = 0;
    checks[n++] = ev.i;
    if (ev.j != -1) checks[n++] = ev.j;
    return n;
}

/* GameSim::run -> updateSim (GameSim.cpp:118-171). */
static void step_game(game_t *G, const double *act /* [2n][2] */) {
    const orc_params *P = G->p;
    if (!(G->score[0] < P->winning_score && G->score[1] < P->winning_score)) return; /* :57-60,121 */

    for (int e = 0; e < 2 * G->n; ++e) { G->u[e][0] = act[2 * e]; G->u[e][1] = act[2 * e + 1]; }
    G->u[G->PK][0] = 0.0; /* :188,214 */
    G->u[G->PK][1] = 0.0;
    G->scored = 0;
    int base = 1;
    fill_grid(G);

    int checks[ORC_MAX_BODIES];
    int n_checks, counter = 0;
    G->nq = 0;
    do { /* :140-150 */
        if (G->nq > 0) {
            n_checks = resolve_first(G, &base, checks);
        } else {
            n_checks = G->E;
            for (int e = 0; e < G->E; ++e) checks[e] = e;
        }
        insert_events(G, checks, n_checks, base);
        counter++;
    } while (G->nq > 0 && counter < ORC_COUNTER_LIM);
    if (G->nq > 0) G->ev[ORC_EV_COUNTER_LIMIT]++;
    G->ev[ORC_EV_LOOP_ITERS] += counter;

    for (int e = 0; e < G->E; ++e) memcpy(G->x[e], G->g[e][ORC_GRID], sizeof(double) * 4); /* :153-157 */
    G->t += P->dt;                                                                        /* :158 */
    G->ev[ORC_EV_LIVE_STEPS]++;
    if (G->ev[ORC_EV_DECIDED_STEP] == 0 &&
        !(G->score[0] < P->winning_score && G->score[1] < P->winning_score))
        G->ev[ORC_EV_DECIDED_STEP] = G->ev[ORC_EV_LIVE_STEPS];
}

/* ---- state vector <-> body arrays ------------------------------------------------------ */
static void load_state(game_t *G, const double *s) {
    G->score[0] = s[0];
    G->score[1] = s[1];
    memcpy(G->x[G->PK], s + 2, sizeof(double) * 4);
    for (int e = 0; e < 2 * G->n; ++e) memcpy(G->x[e], s + 6 + 4 * e, sizeof(double) * 4);
}
static void store_state(const game_t *G, double *s) {
    s[0] = G->score[0];
    s[1] = G->score[1];
    memcpy(s + 2, G->x[G->PK], sizeof(double) * 4);
    for (int e = 0; e < 2 * G->n; ++e) memcpy(s + 6 + 4 * e, G->x[e], sizeof(double) * 4);
}

void orc_default_params(orc_params *p) {
    p->dt = 0.05;
    p->winning_score = 3;
    p->players_per_team = 2;
    p->tau_puck = 0.1;
    p->tau_player = 0.5;
    p->player_mass = 1.0;
    p->puck_mass = 0.5;
    p->arena_x = 10.0;
    p->arena_y = 5.0;
    p->player_radius = 0.2;
    p->puck_radius = 0.175;
    p->goal_height = 1.0;
}

void orc_reset_state(const orc_params *p, const double *x0_ball4, double *out) {
    const int n = p->players_per_team;
    out[0] = 0.0;
    out[1] = 0.0;
    for (int c = 0; c < 4; ++c) out[2 + c] = x0_ball4 ? x0_ball4[c] : 0.0;
    for (int team = 0; team < 2; ++team) {
        for (int k = 0; k < n; ++k) {
            double *b = out + 6 + 4 * (team * n + k);
            b[PX] = team == 0 ? -p->arena_x / 4.0 : p->arena_x / 4.0; /* :35-38 */
            /* n = 2: +-arena_Y/4, bit-identical to the reference expressions */
            b[PY] = (n == 2) ? (k == 0 ? p->arena_y / 4.0 : -p->arena_y / 4.0)
                             : p->arena_y / 2.0 - (k + 0.5) * p->arena_y / n;
            b[VX] = 0.0;
            b[VY] = 0.0;
        }
    }
}

typedef struct {
    const orc_params *p;
    int64_t g0, g1, n_steps, hold, n_act, act_game_stride;
    double *state, *t, *damage, *traj;
    const double *actions;
    int64_t *events;
} job_t;

static void *run_job(void *arg) {
    job_t *J = (job_t *)arg;
    const orc_params *P = J->p;
    game_t *G = (game_t *)malloc(sizeof(game_t));
    if (!G) return NULL;
    const int n = P->players_per_team, E = 2 * n + 1, S = 2 + 4 * E, np = 2 * n;
    G->p = P;
    G->n = n;
    G->E = E;
    G->PK = 2 * n;
    G->dt_col = P->dt / ORC_GRID; /* :28 */
    for (int64_t gi = J->g0; gi < J->g1; ++gi) {
        load_state(G, J->state + gi * S);
        G->t = J->t[gi];
        G->damage = J->damage ? J->damage + gi * np * np : NULL;
        if (J->events) {
            G->ev = J->events + gi * ORC_N_EVENTS;
        } else {
            memset(G->scratch_ev, 0, sizeof(G->scratch_ev));
            G->ev = G->scratch_ev;
        }
        const double *acts = J->actions + gi * J->act_game_stride;
        for (int64_t k = 0; k < J->n_steps; ++k) {
            int64_t a = k / J->hold;
            if (a >= J->n_act) a = J->n_act - 1;
            step_game(G, acts + a * np * 2);
            if (J->traj) store_state(G, J->traj + (gi * J->n_steps + k) * S);
        }
        store_state(G, J->state + gi * S);
        J->t[gi] = G->t;
    }
    free(G);
    return NULL;
}

int orc_rollout(const orc_params *p, int64_t n_games, int64_t n_steps, int64_t hold, double *state, double *t,
                double *damage, const double *actions, int64_t n_act, int64_t act_game_stride, int64_t *events,
                double *traj, int32_t n_threads) {
    if (!p || !state || !t || !actions || n_games < 0 || n_steps < 0 || hold < 1 || n_act < 1) return -1;
    if (p->players_per_team < 1 || p->players_per_team > ORC_MAX_TEAM) return -1;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > n_games) n_threads = (int32_t)(n_games > 0 ? n_games : 1);
    job_t *jobs = (job_t *)calloc((size_t)n_threads, sizeof(job_t));
    pthread_t *th = (pthread_t *)calloc((size_t)n_threads, sizeof(pthread_t));
    if (!jobs || !th) { free(jobs); free(th); return -1; }
    for (int k = 0; k < n_threads; ++k) {
        job_t *J = &jobs[k];
        J->p = p;
        J->g0 = n_games * k / n_threads;
        J->g1 = n_games * (k + 1) / n_threads;
        J->n_steps = n_steps;
        J->hold = hold;
        J->n_act = n_act;
        J->act_game_stride = act_game_stride;
        J->state = state;
        J->t = t;
        J->damage = damage;
        J->traj = traj;
        J->actions = actions;
        J->events = events;
    }
    if (n_threads == 1) {
        run_job(&jobs[0]);
    } else {
        for (int k = 0; k < n_threads; ++k) pthread_create(&th[k], NULL, run_job, &jobs[k]);
        for (int k = 0; k < n_threads; ++k) pthread_join(th[k], NULL);
    }
    free(jobs);
    free(th);
    return 0;
}

void orc_log_record(const double *s, double t, const double *dmg, int bug_compatible, double *out) {
    out[0] = t;
    out[1] = s[0];
    out[2] = s[1];
    out[3] = s[2];  out[4] = s[3];   /* puck xy */
    out[5] = s[6];  out[6] = s[7];   /* A1 */
    out[7] = s[10]; out[8] = s[11];  /* A2 */
    out[9] = s[14]; out[10] = s[15]; /* B1 */
    out[11] = s[18]; out[12] = s[19]; /* B2 */
    for (int k = 0; k < 4; ++k) {
        for (int c = 0; c < 4; ++c) {
            if (bug_compatible) {
                /* 4 contiguous doubles from &damage(k,0) of a COLUMN-major 4x4: linear index k+c,
                 * i.e. element (row (k+c)%4, col (k+c)/4) */
                const int lin = k + c, row = lin % 4, col = lin / 4;
                out[13 + 4 * k + c] = dmg[row * 4 + col];
            } else {
                out[13 + 4 * k + c] = dmg[k * 4 + c];
            }
        }
    }
}
