/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement of the Deep-RL-Battlespace tick.
 *
 * This file is the parity oracle and the timed CPU baseline.  It is NOT part of
 * the product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product path (libbattlespace.so, CUDA)
 * never links, imports or falls back to anything in oracle/.
 *
 * Parity status: PINNED against the reference's own Python source, run unmodified
 * under oracle/stubs by oracle/ref_harness.py (fixtures in tests/golden/, made by
 * oracle/gen_golden.py; live cross-check in oracle/validate_port.py).  The
 * reference ships no tests or golden vectors of its own.
 *
 * Every function cites the reference lines it restates (paths relative to the
 * reference root).  Arithmetic is IEEE double with one rounding per operation
 * (build with -ffp-contract=off), the same libm sin/cos/atan2/sqrt CPython's
 * math module calls, and C casts where pygame 2.1.2 casts.
 *
 * Random draws are either injected by the caller (parity with the reference:
 * the draws the reference consumed are replayed) or produced by Philox4x32-10
 * with the counter layout documented in include/battlespace.h, which the CUDA
 * path implements identically (parity at scale).
 */
#define _GNU_SOURCE 1
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MAXN 8
#define MAXP (2 * MAXN)
#define SLOTS 12
#define MAXB (MAXP * SLOTS)

/* envs/sprites.py:9-10 and the PNG sizes read at sprites.py:61-62,226-227 */
#define DISP_W 1200
#define DISP_H 800
#define PLANE_W 50
#define PLANE_H 48
#define BASE_W 62
#define BASE_H 62
#define BULLET_W 6
#define BULLET_H 3

enum { W_NONE = 0, W_RED = 1, W_BLUE = 2, W_TIE = 3 };
enum { ACT_I32 = 0, ACT_F32 = 1, ACT_F64 = 2, ACT_NONE = 3 };
enum { POL_EXTERNAL = 0, POL_RANDOM = 1, POL_INSTINCT = 2, POL_INSTINCT_FF = 3 };
enum { R_HIT_BASE = 0, R_HIT_PLANE = 1, R_MISS = 2, R_DIE = 3, R_LOSE = 4 };

typedef struct { int x, y, w, h; } rect_t; /* pygame.Rect: top-left + size, ints */

typedef struct {
    rect_t rect;       /* sprites.py:312 */
    double direction;  /* sprites.py:314 */
    double dist;       /* sprites.py:317 dist_travelled */
    int owner;         /* sprites.py:310 agent_id */
} bullet_t;

typedef struct {
    rect_t plane[MAXP];
    double pdir[MAXP];
    int php[MAXP];
    uint8_t in_agents[MAXP]; /* membership of self.agents (battle_env.py:109,357) */
    uint8_t done[MAXP];      /* self.dones (battle_env.py:275,360) */
    rect_t base[2];
    int bhp[2];
    bullet_t bul[MAXB];      /* self.bullets, spawn order (battle_env.py:171,406) */
    int nb;
    double total_time;       /* battle_env.py:175 */
    int tick;                /* number of step() calls that advanced the clock */
    int env_done;            /* battle_env.py:276 */
    int winner;
    uint32_t episode;        /* Philox counter word (not in the reference) */
    double ret[2];           /* running team return of the current episode (main.py:198,203) */
} env_t;

typedef struct {
    int N, P, L, A, cont;
    double rw[5];
    double max_time;         /* battle_env.py:168 */
    int64_t E, env_offset;
    uint64_t seed;
    env_t *env;
    /* battle_env.py:102-103,169-170 counters + report sums (main.py:218-253) */
    int64_t wins_red, wins_blue, ties, total_games, env_ticks, agent_steps, ret_fx[2];
} oracle_t;

/* ------------------------------------------------------------------------- */
/* Philox4x32-10 (Salmon et al., Random123), as in include/battlespace.h      */
static void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                          uint32_t k0, uint32_t k1, uint32_t out[4]) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

enum { PURPOSE_AGENT_TICK = 0, PURPOSE_RESET_PLANE = 1, PURPOSE_RESET_BASE = 2 };

static void env_philox(const oracle_t *o, int64_t e, uint32_t episode, int tick, int agent,
                       int purpose, uint32_t w[4]) {
    uint64_t g = (uint64_t)(o->env_offset + e);
    philox4x32_10((uint32_t)g, (uint32_t)(g >> 32), episode,
                  ((uint32_t)tick << 16) | ((uint32_t)agent << 8) | (uint32_t)purpose,
                  (uint32_t)o->seed, (uint32_t)(o->seed >> 32), w);
}

static int randint_from(uint32_t w, int lo, int hi) { /* inclusive, like random.randint */
    return lo + (int)(((uint64_t)w * (uint64_t)(hi - lo + 1)) >> 32);
}

/* unit draws of one AGENT_TICK block: [shoot_u, noise_u0..2, spread_u] */
static void agent_tick_units(const uint32_t w[4], double u[5]) {
    u[1] = (double)(w[0] >> 8) * (1.0 / 16777216.0);
    u[2] = (double)(w[1] >> 8) * (1.0 / 16777216.0);
    u[3] = (double)(w[2] >> 8) * (1.0 / 16777216.0);
    u[0] = (double)(((w[0] & 0xffu) << 16) | ((w[1] & 0xffu) << 8) | (w[2] & 0xffu)) * (1.0 / 16777216.0);
    u[4] = (double)w[3] * (1.0 / 4294967296.0);
}

/* ------------------------------------------------------------------------- */
/* pygame 2.1.2 Rect pieces the hot path uses (see oracle/stubs/pygame)       */
static int rect_cx(const rect_t *r) { return r->x + (r->w >> 1); }
static int rect_cy(const rect_t *r) { return r->y + (r->h >> 1); }
static void rect_set_center(rect_t *r, double cx, double cy) {
    r->x = (int)cx - (r->w >> 1); /* C cast: truncation toward zero */
    r->y = (int)cy - (r->h >> 1);
}
static int rect_collide(const rect_t *a, const rect_t *b) {
    if (a->w == 0 || a->h == 0 || b->w == 0 || b->h == 0) return 0;
    return a->x < b->x + b->w && a->y < b->y + b->h && a->x + a->w > b->x && a->y + a->h > b->y;
}

/* Plane.forward / rotate / update screen clamp -- sprites.py:105-113,133-141 */
static void plane_clamp(rect_t *r) {
    if (r->x < 0) r->x = 0;
    if (r->x + r->w > DISP_W) r->x = DISP_W - r->w;
    if (r->y <= 0) r->y = 0;
    if (r->y + r->h >= DISP_H) r->y = DISP_H - r->h;
}

/* calc_new_xy -- sprites.py:35-42; math.radians(a) is a * (pi / 180) */
static void calc_new_xy(double ox, double oy, double speed, double time, double angle,
                        double *nx, double *ny) {
    double rad = -(angle * (M_PI / 180.0));
    double st = speed * time;
    *nx = ox + st * cos(rad);
    *ny = oy + st * sin(rad);
}

/* Plane.forward -- sprites.py:123-141 */
static void plane_forward(env_t *s, int i, double speed, double time) {
    double nx, ny;
    calc_new_xy(rect_cx(&s->plane[i]), rect_cy(&s->plane[i]), speed, time, s->pdir[i], &nx, &ny);
    rect_set_center(&s->plane[i], nx, ny);
    plane_clamp(&s->plane[i]);
}

/* Plane.rotate -- sprites.py:93-113 */
static void plane_rotate(env_t *s, int i, double angle) {
    s->pdir[i] += angle;
    while (s->pdir[i] > 360) s->pdir[i] -= 360;
    while (s->pdir[i] < 0) s->pdir[i] += 360;
    plane_clamp(&s->plane[i]);
}

/* Bullet.__init__ -- sprites.py:293-318 */
static void spawn_bullet(env_t *s, int x, int y, double angle, int owner, double u) {
    bullet_t *b = &s->bul[s->nb++];
    b->rect.w = BULLET_W; b->rect.h = BULLET_H;
    b->rect.x = 0; b->rect.y = 0;
    rect_set_center(&b->rect, x, y);
    b->direction = angle + (u * 8 - 4);
    b->dist = 0;
    b->owner = owner;
}

/* rel_angle -- battle_env.py:38-52 (Python float % is fmod plus a sign fix-up) */
static double rel_angle(int x0, int y0, double a0, int x1, int y1) {
    double dx = x0 - x1, dy = y0 - y1;
    double rads = atan2(dy, dx);
    double m = fmod(rads, 2 * M_PI);
    if (m != 0) { if (m < 0) m += 2 * M_PI; } else m = 0.0;
    double degs = m * (180.0 / M_PI);
    double rel = 180 + a0 - (360 - degs);
    if (rel < -180) rel += 360;
    if (rel > 180) rel -= 360;
    return rel;
}

/* dist -- battle_env.py:54-58 */
static double dist2d(int x1, int y1, int x0, int y0) {
    long dx = x1 - x0, dy = y1 - y0;
    return sqrt((double)(dx * dx + dy * dy));
}

static int plane_present(const env_t *s, int i) { return s->in_agents[i] && s->php[i] > 0; }

/* observe -- battle_env.py:202-244 */
static void observe(const oracle_t *o, const env_t *s, int i, float *obs) {
    const int N = o->N;
    for (int k = 0; k < o->L; ++k) obs[k] = -1.0f;
    if (!plane_present(s, i)) return;
    int team = i >= N, ot = 1 - team;
    int ax = rect_cx(&s->plane[i]), ay = rect_cy(&s->plane[i]); /* get_pos == rect.center, SURVEY 8 a9 */
    double adir = s->pdir[i];
    double diag = sqrt(pow(DISP_W, 2) + pow(DISP_H, 2));
    int bx = rect_cx(&s->base[ot]), by = rect_cy(&s->base[ot]);
    obs[0] = (float)(dist2d(ax, ay, bx, by) / diag * 2 - 1);
    obs[1] = (float)(rel_angle(ax, ay, adir, bx, by) / 360);
    for (int k = 0; k < N; ++k) {
        int j = ot * N + k;
        if (s->in_agents[j]) {
            int qx = rect_cx(&s->plane[j]), qy = rect_cy(&s->plane[j]);
            obs[2 + 3 * k] = 1.0f;
            obs[3 + 3 * k] = (float)(dist2d(ax, ay, qx, qy) / diag * 2 - 1);
            obs[4 + 3 * k] = (float)(rel_angle(ax, ay, adir, qx, qy) / 360);
        }
    }
}

static void observe_all(const oracle_t *o, const env_t *s, float *obs) {
    for (int i = 0; i < o->P; ++i) observe(o, s, i, obs + (size_t)i * o->L);
}

/* Base.reset / Plane.reset ranges -- sprites.py:74-91,238-252 */
static void reset_ranges(int team, int *xlo, int *xhi, int *ylo, int *yhi, int w, int h) {
    int xmin = w, xmax = DISP_W - w, ymin = h, ymax = DISP_H - h;
    if (team == 0) { *xlo = xmin; *xhi = xmax / 3; } else { *xlo = xmax / 3 * 2; *xhi = xmax; }
    *ylo = ymin; *yhi = ymax;
}

/* parallel_env.reset -- battle_env.py:246-279; draws[4+6N] in the reference's call order */
static void env_reset(oracle_t *o, int64_t e, const int32_t *draws) {
    env_t *s = &o->env[e];
    const int N = o->N, P = o->P;
    int32_t d[4 + 6 * MAXN];
    s->episode += 1;
    if (draws) {
        memcpy(d, draws, sizeof(int32_t) * (size_t)(4 + 6 * N));
    } else {
        uint32_t w[4];
        int xlo, xhi, ylo, yhi;
        env_philox(o, e, s->episode, 0, 0, PURPOSE_RESET_BASE, w);
        for (int t = 0; t < 2; ++t) {
            reset_ranges(t, &xlo, &xhi, &ylo, &yhi, BASE_W, BASE_H);
            d[2 * t] = randint_from(w[2 * t], xlo, xhi);
            d[2 * t + 1] = randint_from(w[2 * t + 1], ylo, yhi);
        }
        for (int i = 0; i < P; ++i) {
            int t = i >= N;
            env_philox(o, e, s->episode, 0, i, PURPOSE_RESET_PLANE, w);
            reset_ranges(t, &xlo, &xhi, &ylo, &yhi, PLANE_W, PLANE_H);
            d[4 + 3 * i] = randint_from(w[0], xlo, xhi);
            d[5 + 3 * i] = randint_from(w[1], ylo, yhi);
            d[6 + 3 * i] = t == 0 ? randint_from(w[2], 270, 450) : randint_from(w[2], 90, 270);
        }
    }
    s->winner = W_NONE;
    for (int t = 0; t < 2; ++t) {
        s->base[t].w = BASE_W; s->base[t].h = BASE_H; s->base[t].x = s->base[t].y = 0;
        s->bhp[t] = 5 * N; /* battle_env.py:91 */
        rect_set_center(&s->base[t], d[2 * t], d[2 * t + 1]);
    }
    for (int i = 0; i < P; ++i) {
        s->plane[i].w = PLANE_W; s->plane[i].h = PLANE_H; s->plane[i].x = s->plane[i].y = 0;
        s->php[i] = 4; /* battle_env.py:92 */
        rect_set_center(&s->plane[i], d[4 + 3 * i], d[5 + 3 * i]);
        int dir = d[6 + 3 * i];
        if (i < N && dir >= 360) dir -= 360; /* sprites.py:85-86 */
        s->pdir[i] = dir;
        s->in_agents[i] = 1;
        s->done[i] = 0;
    }
    s->total_time = 0; s->tick = 0; s->nb = 0; s->env_done = 0;
    s->ret[0] = s->ret[1] = 0;
}

/* tie / win -- battle_env.py:469-496 */
static void env_tie(oracle_t *o, env_t *s) {
    s->winner = W_TIE; s->env_done = 1;
    for (int i = 0; i < o->P; ++i) s->done[i] = 1;
#pragma omp atomic
    o->total_games += 1;
#pragma omp atomic
    o->ties += 1;
}
static void env_win(oracle_t *o, env_t *s, int team) {
    s->winner = team == 0 ? W_RED : W_BLUE; s->env_done = 1;
    for (int i = 0; i < o->P; ++i) s->done[i] = 1;
#pragma omp atomic
    o->total_games += 1;
    if (team == 0) {
#pragma omp atomic
        o->wins_red += 1;
    } else {
#pragma omp atomic
        o->wins_blue += 1;
    }
}

/* Bullet.update -- sprites.py:321-351.  Returns 0 none, 1 miss, 2 base, 3+k plane k of enemy team */
static int bullet_update(const oracle_t *o, env_t *s, bullet_t *b) {
    const int N = o->N;
    double nx, ny;
    calc_new_xy(rect_cx(&b->rect), rect_cy(&b->rect), 450, 0.1, b->direction, &nx, &ny);
    rect_set_center(&b->rect, nx, ny);
    b->dist += 450 * 0.1;
    if (b->dist >= 500) return 1;
    int cx = rect_cx(&b->rect), cy = rect_cy(&b->rect);
    if (cx > DISP_W || cx < 0 || cy > DISP_H || cy < 0) return 1;
    int ot = b->owner >= N ? 0 : 1;
    if (rect_collide(&b->rect, &s->base[ot])) return 2;
    for (int k = 0; k < N; ++k) { /* oteam['planes'].values(): alive planes, id order */
        int j = ot * N + k;
        if (s->in_agents[j] && rect_collide(&b->rect, &s->plane[j])) return 3 + k;
    }
    return 0;
}

/* One decoded action per agent. */
typedef struct { int disc; double c[3]; } action_t;

/*
 * parallel_env.step -- battle_env.py:281-381 with process_action (:383-424).
 * act[i] already clipped/argmax-ed; spread_u[i] = the random.random() the
 * agent's Bullet.__init__ would draw this tick (used only if it shoots).
 */
static void env_step(oracle_t *o, int64_t e, const action_t *act, const double *spread_u,
                     double *rew, int no_actions) {
    env_t *s = &o->env[e];
    const int N = o->N, P = o->P;
    for (int i = 0; i < P; ++i) rew[i] = 0;
    if (s->env_done) return;                                   /* :303-306 */
    int alive = 0;
    for (int i = 0; i < P; ++i) alive += s->in_agents[i];
    if (no_actions || alive == 0) { env_tie(o, s); return; }    /* :309-313 */
    s->total_time += 0.1;                                      /* :316 */
    s->tick += 1;
    if (s->total_time >= o->max_time) { env_tie(o, s); return; } /* :319-323 */

    for (int i = 0; i < P; ++i) {                              /* :325-329 */
        if (!s->in_agents[i]) continue;
        int ax = rect_cx(&s->plane[i]), ay = rect_cy(&s->plane[i]);
        double adir = s->pdir[i];
        if (!o->cont) {                                        /* :399-417 */
            int a = act[i].disc;
            if (a == 0) plane_forward(s, i, 215, 0.1);
            else if (a == 1) { spawn_bullet(s, ax, ay, adir, i, spread_u[i]); plane_forward(s, i, 215, 0.1); }
            else if (a == 2) { plane_rotate(s, i, 15); plane_forward(s, i, 215, 0.1); }
            else if (a == 3) { plane_rotate(s, i, -15); plane_forward(s, i, 215, 0.1); }
        } else {                                               /* :418-424 */
            double speed = ((act[i].c[0] + 1) / 2) * (275 - 200) + 200;
            plane_forward(s, i, speed, 0.1);
            plane_rotate(s, i, act[i].c[1] * 35);
            if (act[i].c[2] > 0) spawn_bullet(s, ax, ay, adir, i, spread_u[i]);
        }
    }

    int nb = s->nb, keep = 0;                                  /* :332-360 */
    for (int q = 0; q < nb; ++q) {
        bullet_t b = s->bul[q];
        int out = bullet_update(o, s, &b);
        if (out == 0) { s->bul[keep++] = b; continue; }
        if (out == 1) rew[b.owner] += o->rw[R_MISS];
        else if (out == 2) {
            int ot = b.owner >= N ? 0 : 1;
            s->bhp[ot] -= 1;
            rew[b.owner] += o->rw[R_HIT_BASE];
        } else {
            int ot = b.owner >= N ? 0 : 1, j = ot * N + (out - 3);
            s->php[j] -= 1;
            rew[b.owner] += o->rw[R_HIT_PLANE];
            if (s->php[j] <= 0) {
                s->in_agents[j] = 0;
                rew[j] += o->rw[R_DIE];
                s->done[j] = 1;
            }
        }
    }
    s->nb = keep;

    if (s->bhp[1] <= 0) {                                      /* :363-366 */
        for (int i = 0; i < N; ++i) rew[i] += o->rw[R_LOSE];
        env_win(o, s, 0);
    }
    if (s->bhp[0] <= 0) {                                      /* :369-372 */
        for (int i = N; i < P; ++i) rew[i] += o->rw[R_LOSE];
        env_win(o, s, 1);
    }
}

/* InstinctAgent.choose_action -- instinct/agent.py:10-62.  obs is the agent's own
 * float32 observation; all arithmetic in double (numpy 1.23.1 promotion). */
static void instinct_action(const oracle_t *o, const float *obs, double shoot_u,
                            const double *noise_u, action_t *out) {
    const int N = o->N;
    double diag = sqrt(pow(DISP_W, 2) + pow(DISP_H, 2));
    double dist[1 + MAXN], ang[1 + MAXN], score[1 + MAXN];
    dist[0] = ((double)obs[0] + 1) / 2 * diag;
    ang[0] = (double)obs[1] * 360;
    score[0] = dist[0] * fabs(ang[0]);
    for (int k = 0; k < N; ++k) {
        dist[1 + k] = ((double)obs[3 + 3 * k] + 1) / 2 * diag;
        ang[1 + k] = (double)obs[4 + 3 * k] * 360;
        score[1 + k] = (obs[2 + 3 * k] == 1.0f) ? dist[1 + k] * fabs(ang[1 + k]) : 1000000.0;
    }
    double mn = score[0];
    for (int k = 1; k <= N; ++k) if (score[k] < mn) mn = score[k];
    int tgt = 0;
    if (mn != score[0]) for (int k = 1; k <= N; ++k) if (score[k] == mn) { tgt = k; break; }
    double td = dist[tgt], ta = ang[tgt];
    if (o->cont) {
        double a[3] = {0, 0, 0};
        if (td < 500.0 / 3 * 2 && fabs(ta) < 20) a[2] = shoot_u < 0.6 ? 1 : -1;
        a[0] = td / diag * 2 - 1;
        if (ta > 0) { double v = -ta / 35; a[1] = v > -1 ? v : -1; }
        else { double v = -ta / 35; a[1] = v < 1 ? v : 1; }
        for (int k = 0; k < 3; ++k) {
            double v = a[k] + (-0.15 + (0.15 - -0.15) * noise_u[k]);
            out->c[k] = v < -1 ? -1 : (v > 1 ? 1 : v);
        }
        out->disc = 0;
    } else {
        if (td < 500.0 / 2 && fabs(ta) < 20) out->disc = 1;
        else if (ta > 0) out->disc = 3;
        else out->disc = 2;
    }
}

/* ------------------------------------------------------------------------- */
/* Batched API (mirrors include/battlespace.h entry for entry)               */

oracle_t *bo_create(int n_agents, int continuous, const double *rewards, int64_t n_envs,
                    int64_t env_offset, uint64_t seed) {
    if (n_agents < 1 || n_agents > MAXN || n_envs < 1) return NULL;
    oracle_t *o = (oracle_t *)calloc(1, sizeof(oracle_t));
    o->N = n_agents; o->P = 2 * n_agents; o->L = 3 * n_agents + 2;
    o->cont = continuous != 0; o->A = o->cont ? 3 : 4;
    memcpy(o->rw, rewards, sizeof(o->rw));
    o->max_time = 10 + 2 * n_agents;
    o->E = n_envs; o->env_offset = env_offset; o->seed = seed;
    o->env = (env_t *)calloc((size_t)n_envs, sizeof(env_t));
    for (int64_t e = 0; e < n_envs; ++e) { o->env[e].env_done = 1; o->env[e].winner = W_NONE; }
    return o;
}

void bo_destroy(oracle_t *o) { if (o) { free(o->env); free(o); } }

void bo_reset(oracle_t *o, const uint8_t *mask, const int32_t *draws, float *obs) {
    const int nd = 4 + 6 * o->N;
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < o->E; ++e) {
        if (!mask || mask[e]) env_reset(o, e, draws ? draws + e * nd : NULL);
        if (obs) observe_all(o, &o->env[e], obs + (size_t)e * o->P * o->L);
    }
}

static int argmax4(const float *v) { /* np.argmax: first maximum -- battle_env.py:328 */
    int b = 0;
    for (int k = 1; k < 4; ++k) if (v[k] > v[b]) b = k;
    return b;
}

static double clip1(double v) { return v < -1 ? -1 : (v > 1 ? 1 : v); } /* battle_env.py:295-297 */

static void decode_action(const oracle_t *o, const void *actions, int dtype, int64_t idx, action_t *a) {
    a->disc = 0; a->c[0] = a->c[1] = a->c[2] = 0;
    if (dtype == ACT_NONE) return;
    if (!o->cont) {
        if (dtype == ACT_I32) a->disc = ((const int32_t *)actions)[idx];
        else a->disc = argmax4((const float *)actions + idx * 4);
    } else if (dtype == ACT_F32) {
        for (int k = 0; k < 3; ++k) a->c[k] = clip1((double)((const float *)actions)[idx * 3 + k]);
    } else {
        for (int k = 0; k < 3; ++k) a->c[k] = clip1(((const double *)actions)[idx * 3 + k]);
    }
}

static void finish_tick(oracle_t *o, int64_t e, const double *rew, int was_done) {
    env_t *s = &o->env[e];
    if (was_done) return;
    double r0 = 0, r1 = 0;
    for (int i = 0; i < o->N; ++i) r0 += rew[i];
    for (int i = o->N; i < o->P; ++i) r1 += rew[i];
    s->ret[0] += r0; s->ret[1] += r1;
#pragma omp atomic
    o->env_ticks += 1;
#pragma omp atomic
    o->agent_steps += o->P;
    if (s->env_done) {
        int64_t f0 = llrint(s->ret[0] * 1048576.0), f1 = llrint(s->ret[1] * 1048576.0);
#pragma omp atomic
        o->ret_fx[0] += f0;
#pragma omp atomic
        o->ret_fx[1] += f1;
    }
}

/* bs_step: all 2N agents driven by the caller.  draws: NULL (Philox) or [E,2N,5]. */
void bo_step(oracle_t *o, const void *actions, int dtype, const double *draws, float *obs,
             double *rew, uint8_t *done, uint8_t *env_done, int8_t *winner) {
    const int P = o->P;
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < o->E; ++e) {
        env_t *s = &o->env[e];
        action_t act[MAXP]; double su[MAXP], r[MAXP];
        for (int i = 0; i < P; ++i) {
            decode_action(o, actions, dtype, e * P + i, &act[i]);
            if (draws) su[i] = draws[(e * P + i) * 5 + 4];
            else { uint32_t w[4]; double u[5]; env_philox(o, e, s->episode, s->tick, i, PURPOSE_AGENT_TICK, w); agent_tick_units(w, u); su[i] = u[4]; }
        }
        int was_done = s->env_done;
        env_step(o, e, act, su, r, dtype == ACT_NONE);
        finish_tick(o, e, r, was_done);
        if (obs) observe_all(o, s, obs + (size_t)e * P * o->L);
        for (int i = 0; i < P; ++i) { if (rew) rew[e * P + i] = r[i]; if (done) done[e * P + i] = s->done[i]; }
        if (env_done) env_done[e] = (uint8_t)s->env_done;
        if (winner) winner[e] = (int8_t)s->winner;
    }
}

/* Policy for one agent from its own observation; fills *a and the raw action row (A floats). */
static void run_policy(const oracle_t *o, int policy, const float *obs_i, const double u[5],
                       uint32_t w0, action_t *a, float *raw) {
    if (policy == POL_RANDOM) {
        if (o->cont) { for (int k = 0; k < 3; ++k) { a->c[k] = 2 * u[1 + k] - 1; if (raw) raw[k] = (float)a->c[k]; } a->disc = 0; }
        else { a->disc = (int)(w0 >> 30); if (raw) for (int k = 0; k < 4; ++k) raw[k] = k == a->disc ? 1.0f : 0.0f; }
    } else {
        instinct_action(o, obs_i, u[0], u + 1, a);
        if (o->cont) { if (policy == POL_INSTINCT_FF) a->c[2] = 1; if (raw) for (int k = 0; k < 3; ++k) raw[k] = (float)a->c[k]; }
        else if (raw) for (int k = 0; k < 4; ++k) raw[k] = k == a->disc ? 1.0f : 0.0f;
    }
}

/* bs_instinct_actions: the instinct/team.py surface.  obs [E,2N,L]; writes the chosen
 * team's actions into act_out ([E,N,3] f64 continuous, [E,N] i32 discrete). */
void bo_instinct(oracle_t *o, int team, const float *obs, const double *draws, void *act_out) {
    const int N = o->N, P = o->P, L = o->L;
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < o->E; ++e) {
        env_t *s = &o->env[e];
        for (int k = 0; k < N; ++k) {
            int i = team * N + k; double u[5]; uint32_t w[4] = {0, 0, 0, 0}; action_t a;
            if (draws) memcpy(u, draws + (e * P + i) * 5, sizeof(u));
            else { env_philox(o, e, s->episode, s->tick, i, PURPOSE_AGENT_TICK, w); agent_tick_units(w, u); }
            run_policy(o, POL_INSTINCT, obs + ((size_t)e * P + i) * L, u, w[0], &a, NULL);
            if (o->cont) memcpy((double *)act_out + (e * N + k) * 3, a.c, sizeof(a.c));
            else ((int32_t *)act_out)[e * N + k] = a.disc;
        }
    }
}

typedef struct {
    float *state, *new_state, *action, *reward; uint8_t *done;
    int64_t capacity;
} replay_t;

/*
 * bs_rollout_fused: T ticks with in-library policies, auto-reset, and one replay row per
 * env-tick for the learner team (maddpg/buffer.py:25-47 row layout, float32).
 * pol[team] in POL_*; ext_actions (POL_EXTERNAL, learner team only) is [T,E,N,A] float32.
 * Row index of (tick t, env e) = (cursor + t*E + e) mod capacity.
 */
void bo_rollout(oracle_t *o, int T, int learner, const int *pol, const float *ext_actions,
                float *rp_state, float *rp_new_state, float *rp_action, float *rp_reward,
                uint8_t *rp_done, int64_t capacity, int64_t cursor) {
    const int N = o->N, P = o->P, L = o->L, A = o->A;
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < o->E; ++e) {
        env_t *s = &o->env[e];
        float obs[MAXP * (3 * MAXN + 2)], obs2[MAXP * (3 * MAXN + 2)], raw[MAXP * 4];
        observe_all(o, s, obs);
        for (int t = 0; t < T; ++t) {
            if (s->env_done) { env_reset(o, e, NULL); observe_all(o, s, obs); }
            action_t act[MAXP]; double su[MAXP], r[MAXP];
            for (int i = 0; i < P; ++i) {
                int team = i >= N; uint32_t w[4]; double u[5];
                env_philox(o, e, s->episode, s->tick, i, PURPOSE_AGENT_TICK, w);
                agent_tick_units(w, u);
                su[i] = u[4];
                if (pol[team] == POL_EXTERNAL) {
                    int64_t idx = ((int64_t)t * o->E + e) * N + (i - team * N);
                    decode_action(o, ext_actions, ACT_F32, idx, &act[i]);
                    /* continuous: step() clips in the caller's dict (battle_env.py:295-297), so the
                     * stored action is the clipped one (main.py:195); discrete logits stay raw */
                    for (int k = 0; k < A; ++k)
                        raw[i * 4 + k] = o->cont ? (float)act[i].c[k] : ext_actions[idx * A + k];
                } else run_policy(o, pol[team], obs + (size_t)i * L, u, w[0], &act[i], raw + i * 4);
            }
            env_step(o, e, act, su, r, 0);
            finish_tick(o, e, r, 0);
            observe_all(o, s, obs2);
            if (capacity > 0) {
                int64_t row = (cursor + (int64_t)t * o->E + e) % capacity;
                for (int k = 0; k < N; ++k) {
                    int i = learner * N + k;
                    if (rp_state) memcpy(rp_state + (row * N + k) * L, obs + (size_t)i * L, sizeof(float) * L);
                    if (rp_new_state) memcpy(rp_new_state + (row * N + k) * L, obs2 + (size_t)i * L, sizeof(float) * L);
                    if (rp_action) memcpy(rp_action + (row * N + k) * A, raw + i * 4, sizeof(float) * A);
                    if (rp_reward) rp_reward[row * N + k] = (float)r[i];
                    if (rp_done) rp_done[row * N + k] = s->done[i];
                }
            }
            memcpy(obs, obs2, sizeof(float) * (size_t)P * L);
        }
    }
}

/* State export in the product's slot layout: bullets[E,2N,12,2] (x,y or -1), slot = age-1. */
void bo_get_state(const oracle_t *o, int16_t *px, int16_t *py, double *pdir, int8_t *php,
                  int16_t *base_pos, int16_t *bhp, int16_t *bullets, int32_t *tick,
                  uint8_t *env_done, int8_t *winner, uint32_t *episode) {
    const int P = o->P;
    for (int64_t e = 0; e < o->E; ++e) {
        const env_t *s = &o->env[e];
        for (int i = 0; i < P; ++i) {
            int present = plane_present(s, i);
            if (px) px[e * P + i] = present ? (int16_t)rect_cx(&s->plane[i]) : 0;
            if (py) py[e * P + i] = present ? (int16_t)rect_cy(&s->plane[i]) : 0;
            if (pdir) pdir[e * P + i] = present ? s->pdir[i] : 0;
            if (php) php[e * P + i] = present ? (int8_t)s->php[i] : 0;
        }
        for (int t = 0; t < 2; ++t) {
            if (base_pos) { base_pos[(e * 2 + t) * 2] = (int16_t)rect_cx(&s->base[t]); base_pos[(e * 2 + t) * 2 + 1] = (int16_t)rect_cy(&s->base[t]); }
            if (bhp) bhp[e * 2 + t] = (int16_t)s->bhp[t];
        }
        if (bullets) {
            int16_t *b = bullets + (size_t)e * P * SLOTS * 2;
            for (int k = 0; k < P * SLOTS * 2; ++k) b[k] = -1;
            for (int q = 0; q < s->nb; ++q) {
                int age = (int)llrint(s->bul[q].dist / 45.0);
                int16_t *d = b + ((size_t)s->bul[q].owner * SLOTS + (age - 1)) * 2;
                d[0] = (int16_t)rect_cx(&s->bul[q].rect); d[1] = (int16_t)rect_cy(&s->bul[q].rect);
            }
        }
        if (tick) tick[e] = s->tick;
        if (env_done) env_done[e] = (uint8_t)s->env_done;
        if (winner) winner[e] = (int8_t)s->winner;
        if (episode) episode[e] = s->episode;
    }
}

/* {wins_red, wins_blue, ties, total_games, sum red return * 2^20, sum blue return * 2^20, env ticks, agent steps} */
void bo_stats(const oracle_t *o, int64_t out[8]) {
    out[0] = o->wins_red; out[1] = o->wins_blue; out[2] = o->ties; out[3] = o->total_games;
    out[4] = o->ret_fx[0]; out[5] = o->ret_fx[1]; out[6] = o->env_ticks; out[7] = o->agent_steps;
}

int bo_tie_tick(int n_agents) { /* battle_env.py:316-319 accumulated in double */
    double t = 0; int k = 0;
    for (;;) { ++k; t += 0.1; if (t >= 10 + 2 * n_agents) return k; }
}
