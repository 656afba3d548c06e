// rollout2.cuh -- the fused T-tick rollout, team-per-thread mapping (round 2).
//
// Why a second mapping.  rollout_kernel (battlespace.cu) gives every PLANE a lane and an env 2N consecutive
// lanes.  Profiled at 3v3 (profiles/r2a_*): 1,501 warp-instructions per warp-tick for 5 envs = 300 per env-tick,
// issue-bound (issue slots 74 % busy), 24 of 32 lanes active on average -- per-env work is replicated on 2N
// lanes, each team's policy runs with the other team's lanes idle, and the dense bullet pass pays ~125
// instructions per 32 bullets for ballots, shared-memory atomics and compaction.
//
// Here a THREAD owns one team of one env: its N planes in registers, its team's bullets as a packed list in
// local memory (L1), in the reference's list order.  A block is 64 threads = 32 envs: warp 0 holds the red
// teams, warp 1 the blue teams, so a policy is warp-uniform and nothing is replicated more than twice.  The two
// threads of an env meet at two block barriers per tick:
//   A   policy + process_action of the own planes, spawns appended, positions published     -- battle_env.py:325-329,383-424
//   B1
//   B   the own team's bullets against the enemy base and planes, in list order.  Only the own team's bullets
//       can hurt the enemy, so this thread is the AUTHORITY for the enemy's hit points and the enemy base
//       (battle_env.py:332-360 read per team: a red bullet only ever tests blue planes); it publishes them
//   B2
//   C   own hit points / deaths arrive from the partner; rewards in the reference's accumulation order,
//       win / tie (:363-372,469-496), then the own planes' observations (:202-244) into shared memory
// The replay rows (maddpg/buffer.py:25-47) leave through shared memory: a warp's 32 consecutive ring rows are
// one contiguous block of every ring array, written with full-warp 128-bit stores (north_star: "coalesced,
// vectorised float4 HBM access") instead of one strided 32-bit store per lane and float.
//
// Arithmetic is the same device code as the lane-per-plane kernels (plane_action, instinct, observe, bullet_advance,
// rect_hit): results are bit-identical between the two mappings (tests/test_rollout_mappings.py).
#pragma once
#include "tick.cuh"

namespace bs {

// Bullets between launches, in this kernel's own layout: per block of 32 envs, slot-major and thread-minor
// ([S][64] 8-byte entries, S = 12N slots per team), so that a warp's loads and stores of slot s are coalesced.
// meta word as in the warp pools (tick.cuh) with the owner field = plane index inside the team.
// The lane-per-plane kernels keep their per-warp pools; the handle records which of the two layouts is current
// and converts on demand (pools_to_lists_kernel / lists_to_pools_kernel below).

struct RolloutArgs {
    int T, learner;
    int pol[2];
    const float *ext;     // [T,E,N,A] learner actions when pol[learner] == POL_EXTERNAL
    bs_replay rp;
    long long cursor_mod; // cursor % capacity
    long long e_mod;      // E % capacity
    long long e0, e_end;  // env range of this launch (the host entry point pipelines chunks)
};

template <int N, bool CONT>
struct R2Smem {
    static constexpr int L = 3 * N + 2, NL = N * L, NA = N * (CONT ? 3 : 4);
    float obs[2][32 * NL];        // [team][env][k][c]: the team's current observations = the rows of state / new_state
    float act[32 * NA];           // the learner's raw action rows
    uint32_t pos[2][2][N][32];    // [tick parity][team][k][env]  x | y << 16 after the move
    uint32_t rpos[2][N][32];      // the same after a reset / at kernel start (own buffer: see the hazards note)
    uint32_t hp[2][32];           // [publisher team][env]: the ENEMY's hit points, 4 bits per plane, after the bullet loop
    uint32_t died[2][32];         //   age index + 1 of the bullet that killed enemy plane j (4 bits per plane), 0 = none
    int bhp[2][32];               //   the enemy base's hit points
    double retx[2][32];           // team return of a finished game (episode log only)
#ifdef R2_SMEM_BULLETS
    uint2 blt[2][SLOTS * N][32];  // [team][slot][env]: the team's bullet list
#endif
};

// Copies the warp's staged rows (32 envs x NW floats, row-major in shared memory) to ring rows row0, row0+1, ..
// of `base`.  flat: the rows of the warp's valid envs do not wrap around the ring end.
template <int NW>
__device__ __forceinline__ void rows_out(float *base, const float *st, uint32_t row, uint32_t row0, int nvalid, bool flat, int lane) {
    if (flat) {
        float *dst = base + (size_t)row0 * NW;
        const int n = nvalid * NW;
        if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
            const float4 *s4 = reinterpret_cast<const float4 *>(st);
            float4 *d4 = reinterpret_cast<float4 *>(dst);
            const int n4 = n >> 2;
#pragma unroll
            for (int i = 0; i < (8 * NW + 31) / 32; ++i) {
                const int q = i * 32 + lane;
                if (q < n4) d4[q] = s4[q];
            }
            const int q = (n & ~3) + lane;   // fewer than 4 floats left
            if (q < n) dst[q] = st[q];
        } else {
#pragma unroll 4
            for (int q = lane; q < n; q += 32) dst[q] = st[q];
        }
    } else if (lane < nvalid) {
        for (int c = 0; c < NW; ++c) base[(size_t)row * NW + c] = st[lane * NW + c];
    }
}

// observations of own plane k -- battle_env.py:202-244 -- through the lane-per-plane code (observe<N>), so that the
// arithmetic is the same instruction sequence
template <int N>
__device__ __forceinline__ void observe_plane(const Params &p, uint32_t pxy, double dir, bool me, uint32_t ebase,
                                              const uint32_t (&exy)[N], unsigned am, float (&o)[3 * N + 2]) {
    Lane<N> s;
    s.x = (int)(pxy & 0xffffu); s.y = (int)(pxy >> 16); s.hp = me ? 1 : 0; s.dir = dir;
    s.ebx = (int)(ebase & 0xffffu); s.eby = (int)(ebase >> 16);
#pragma unroll
    for (int j = 0; j < N; ++j) s.exy[j] = exy[j];
    s.am = am;
    observe<N>(p, s, o);
}

#if defined(R2_MINB)
template <int N> struct R2Cfg { static constexpr int MINB = N <= 3 ? R2_MINB : 3; };
#elif defined(R2_SMEM_BULLETS)
template <int N> struct R2Cfg { static constexpr int MINB = N <= 3 ? 7 : (N == 4 ? 4 : 3); };
#else
template <int N> struct R2Cfg { static constexpr int MINB = N <= 3 ? 14 : (N <= 5 ? 6 : 3); };
#endif

template <int N, bool CONT>
__global__ void __launch_bounds__(64, R2Cfg<N>::MINB) rollout2_kernel(Params p, RolloutArgs ra) {
    constexpr int L = 3 * N + 2, NL = N * L, A = CONT ? 3 : 4, NA = N * A, S = SLOTS * N;
    constexpr uint32_t HP_FULL = 0x44444444u >> (4 * (8 - N));     // 4 hit points per plane, battle_env.py:92
    __shared__ __align__(16) R2Smem<N, CONT> sm;
    const int tid = threadIdx.x, lane = tid & 31, team = tid >> 5;
    const long long blk = ra.e0 / 32 + blockIdx.x;
    const long long e = blk * 32 + lane;
    const bool valid = e < ra.e_end;
    int nvalid = (int)(ra.e_end - blk * 32);
    nvalid = nvalid > 32 ? 32 : (nvalid < 0 ? 0 : nvalid);
    float *myobs = sm.obs[team] + lane * NL;

    // ---- state -> registers ------------------------------------------------------------------------------
    uint32_t pxy[N];
    double dir[N];
    uint32_t ohp = 0, ehp = 0;            // own / enemy hit points, 4 bits per plane
    uint32_t ebase = 64u | (64u << 16);   // enemy base centre
    int bhp_own = 0, bhp_enemy = 0, tick = 0, env_done = 1, winner = 0, ticks_run = 0;
    uint32_t episode = 0;
    double ret = 0;
#ifdef R2_SMEM_BULLETS
    struct BltRef { uint2 (*b)[32]; int lane; __device__ __forceinline__ uint2 &operator[](int s) const { return b[s][lane]; } };
    const BltRef blt{sm.blt[team], lane};
#else
    uint2 blt[S];
#endif
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < N; ++k) { pxy[k] = 0; dir[k] = 0; }
    if (valid) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
            const size_t a = (size_t)e * (2 * N) + team * N + k, b = (size_t)e * (2 * N) + (1 - team) * N + k;
            pxy[k] = p.plane_pos[a];
            dir[k] = p.plane_dir[a];
            const int h0 = p.plane_hp[a], h1 = p.plane_hp[b];
            ohp |= (uint32_t)(h0 > 0 ? h0 : 0) << (4 * k);
            ehp |= (uint32_t)(h1 > 0 ? h1 : 0) << (4 * k);
        }
        ebase = p.base_pos[e * 2 + (1 - team)];
        bhp_own = p.base_hp[e * 2 + team];
        bhp_enemy = p.base_hp[e * 2 + (1 - team)];
        const uint32_t m = p.meta[e];
        tick = m & 0xfff; env_done = (m >> 16) & 0xff; winner = (m >> 24) & 0xff;
        episode = p.episode[e];
        ret = p.ret[e * 2 + team];
        cnt = p.list_cnt[blk * 64 + tid];
        const uint2 *src = p.lists + (size_t)blk * S * 64 + tid;
        for (int s = 0; s < cnt; ++s) blt[s] = src[(size_t)s * 64];
    }

    const int pol = team ? ra.pol[1] : ra.pol[0];
    const bool writer = team == ra.learner && ra.rp.capacity > 0;   // warp-uniform
    uint32_t row = 0;
    if (ra.rp.capacity > 0) row = (uint32_t)((ra.cursor_mod + e % ra.rp.capacity) % ra.rp.capacity);
    const uint32_t cap32 = (uint32_t)ra.rp.capacity, emod32 = (uint32_t)ra.e_mod;

#pragma unroll 1
    for (int t = 0; t < ra.T; ++t) {
        const int buf = t & 1;
        // ---- auto-reset (the game loop's env.reset(), main.py:148) and the first observation of the launch ----
        // Hazards: the partner may still be reading pos[(t-1)&1] for its observations of tick t-1, and pos[t&1] is
        // written by phase A below, so the positions after a reset go through a buffer of their own (rpos), whose
        // previous readers have all passed B1 of an earlier tick.
        const bool need = valid && env_done;
        if (t == 0 || __any_sync(FULL, need)) {
            if (need) {   // parallel_env.reset -- battle_env.py:246-279, sprites.py:74-91,238-252
                episode += 1;
                uint32_t w[4];
                env_philox(p, e, episode, 0, 0, PURPOSE_RESET_BASE, w);
                const int rbx = randint_from(w[0], 62, 379), rby = randint_from(w[1], 62, 738);      // sprites.py:246-247
                const int bbx = randint_from(w[2], 758, 1138), bby = randint_from(w[3], 62, 738);    // sprites.py:250-251
                ebase = team == 0 ? ((uint32_t)bbx | ((uint32_t)bby << 16)) : ((uint32_t)rbx | ((uint32_t)rby << 16));
#pragma unroll
                for (int k = 0; k < N; ++k) {
                    env_philox(p, e, episode, 0, team * N + k, PURPOSE_RESET_PLANE, w);
                    int px, pd;
                    if (team == 0) { px = randint_from(w[0], 50, 383); pd = randint_from(w[2], 270, 450); if (pd >= 360) pd -= 360; }  // :82-86
                    else { px = randint_from(w[0], 766, 1150); pd = randint_from(w[2], 90, 270); }                                 // :88-91
                    const int py = randint_from(w[1], 48, 752);
                    pxy[k] = (uint32_t)px | ((uint32_t)py << 16);
                    dir[k] = (double)pd;
                }
                ohp = ehp = HP_FULL;
                bhp_own = bhp_enemy = 5 * N;                                                            // battle_env.py:91
                tick = 0; env_done = 0; winner = W_NONE; ret = 0;
                cnt = 0;                                                                                // self.bullets = [], :272
            }
#pragma unroll
            for (int k = 0; k < N; ++k) sm.rpos[team][k][lane] = pxy[k];
            __syncthreads();
            uint32_t exy[N];
            unsigned am = 0;
#pragma unroll
            for (int j = 0; j < N; ++j) { exy[j] = sm.rpos[1 - team][j][lane]; am |= (((ehp >> (4 * j)) & 15u) ? 1u : 0u) << j; }
#pragma unroll
            for (int k = 0; k < N; ++k) {
                float o[L];
                observe_plane<N>(p, pxy[k], dir[k], ((ohp >> (4 * k)) & 15u) != 0, ebase, exy, am, o);
                if (need || t == 0) {
#pragma unroll
                    for (int c = 0; c < L; ++c) myobs[k * L + c] = o[c];
                }
            }
            __syncwarp();
        }
        const uint32_t row0 = __shfl_sync(FULL, row, 0);
        const bool flat = row0 + (uint32_t)nvalid <= cap32;
        if (writer && ra.rp.state) rows_out<NL>(ra.rp.state, sm.obs[team], row, row0, nvalid, flat, lane);

        // ---- phase A: policies and process_action of the own planes ------------------------------------------
        const bool active = valid && !env_done;                              // :303-306
        const bool any_alive = (ohp | ehp) != 0;
        const bool tie_now = active && !any_alive;                           // :309-313
        const bool adv = active && any_alive;
        const int tick_before = tick;
        tick += adv ? 1 : 0;                                                 // :316
        const bool time_tie = adv && tick >= p.tie_tick;                     // :319-323
        const bool run = adv && !time_tie;
#pragma unroll
        for (int k = 0; k < N; ++k) {
            int adisc = 0;
            double a[3] = {0, 0, 0}, spread_u = 0;
            float raw[4] = {0, 0, 0, 0};
            if (valid) {
                uint32_t w[4];
                env_philox(p, e, episode, tick_before, team * N + k, PURPOSE_AGENT_TICK, w);
                spread_u = unit32(w[3]);
                if (pol == POL_EXTERNAL) {
                    const float *src = ra.ext + (((long long)t * p.E + e) * N + k) * A;
                    if (CONT) {
#pragma unroll
                        for (int c = 0; c < 3; ++c) { a[c] = clamp1((double)src[c]); raw[c] = (float)a[c]; }
                    } else {
#pragma unroll
                        for (int c = 0; c < 4; ++c) raw[c] = src[c];
                        adisc = 0;
                        float mx = raw[0];
#pragma unroll
                        for (int c = 1; c < 4; ++c) if (raw[c] > mx) { mx = raw[c]; adisc = c; }   // np.argmax: first maximum
                    }
                } else if (pol == POL_RANDOM) {
                    if (CONT) {
#pragma unroll
                        for (int c = 0; c < 3; ++c) { a[c] = sym24(w[c] >> 8); raw[c] = (float)a[c]; }
                    } else {
                        adisc = (int)(w[0] >> 30);
#pragma unroll
                        for (int c = 0; c < 4; ++c) raw[c] = c == adisc ? 1.0f : 0.0f;
                    }
                } else {
                    float o[L];
#pragma unroll
                    for (int c = 0; c < L; ++c) o[c] = myobs[k * L + c];
                    const double shoot_u = unit24(((w[0] & 0xffu) << 16) | ((w[1] & 0xffu) << 8) | (w[2] & 0xffu));
                    instinct<N, CONT>(o, shoot_u, unit24(w[0] >> 8), unit24(w[1] >> 8), unit24(w[2] >> 8), adisc, a);
                    if (CONT) {
                        if (pol == POL_INSTINCT_FF) a[2] = 1.0;
#pragma unroll
                        for (int c = 0; c < 3; ++c) raw[c] = (float)a[c];
                    } else {
#pragma unroll
                        for (int c = 0; c < 4; ++c) raw[c] = c == adisc ? 1.0f : 0.0f;
                    }
                }
            }
            if (writer) {
#pragma unroll
                for (int c = 0; c < A; ++c) sm.act[lane * NA + k * A + c] = raw[c];
            }
            int x = (int)(pxy[k] & 0xffffu), y = (int)(pxy[k] >> 16);
            uint32_t spawnP; int32_t spawnD;
            plane_action<CONT>(p, x, y, dir[k], run && ((ohp >> (4 * k)) & 15u) != 0, adisc, a, spread_u, spawnP, spawnD);
            pxy[k] = (uint32_t)x | ((uint32_t)y << 16);
            if (spawnP) { blt[cnt] = make_uint2(spawnP, pack_meta(spawnD, k, 0)); cnt += 1; }
            sm.pos[buf][team][k][lane] = pxy[k];
        }
        if (writer && ra.rp.action) {
            __syncwarp();
            rows_out<NA>(ra.rp.action, sm.act, row, row0, nvalid, flat, lane);
        }
        __syncthreads();   // B1

        // ---- phase B: the own team's bullets, in list order -- battle_env.py:332-360, sprites.py:321-351 -------
        uint32_t exy[N];
        unsigned am = 0;
#pragma unroll
        for (int j = 0; j < N; ++j) { exy[j] = sm.pos[buf][1 - team][j][lane]; am |= (((ehp >> (4 * j)) & 15u) ? 1u : 0u) << j; }
        uint32_t ev[N];
#pragma unroll
        for (int k = 0; k < N; ++k) ev[k] = 0;
        uint32_t died = 0;
        if (run) {
            const uint32_t baseT = (((ebase & 0xffffu) - 33u) | (((ebase >> 16) - 32u) << 16)) + B_BIAS;
            int out = 0;
            for (int s = 0; s < cnt; ++s) {
                const uint2 en = blt[s];
                uint32_t P = en.x, M = en.y;
                const int age = (int)(M >> 28) + 1;             // updates including this one
                const int ai = SLOTS - age;                     // age index, 0 = oldest
                const int k = (int)(M >> 23) & 7;
                unsigned code = 0;
                if (bullet_advance(P, meta_dlt(M), age)) code = 1u;                                  // miss, sprites.py:334-341
                else if (rect_hit((P & B_POS) - baseT, 67u, 64u)) { code = 2u; bhp_enemy -= 1; }   // base: no alive check, :344-346
                else {
                    const uint32_t Pp = (P & B_POS) - PLANE_OFF;
                    unsigned cm = 0;
#pragma unroll
                    for (int j = 0; j < N; ++j) cm |= (rect_hit(Pp - exy[j], 55u, 50u) ? 1u : 0u) << j;
                    cm &= am;                                   // only planes still in self.agents, battle_env.py:348-360
                    if (cm) {
                        const int j = __ffs(cm) - 1;
                        code = 3u;
                        ehp -= 1u << (4 * j);
                        if (((ehp >> (4 * j)) & 15u) == 0) { am &= ~(1u << j); died |= (uint32_t)(ai + 1) << (4 * j); }
                    }
                }
                if (code) {
#pragma unroll
                    for (int kk = 0; kk < N; ++kk) if (k == kk) ev[kk] |= code << (2 * ai);
                } else {
                    blt[out] = make_uint2(P, M + (1u << 28));
                    out += 1;
                }
            }
            cnt = out;
        }
        sm.hp[team][lane] = ehp; sm.died[team][lane] = died; sm.bhp[team][lane] = bhp_enemy;
        __syncthreads();   // B2
        ohp = sm.hp[1 - team][lane];
        const uint32_t died_own = sm.died[1 - team][lane];
        bhp_own = sm.bhp[1 - team][lane];

        // ---- phase C: rewards in the reference's accumulation order (:337-360), win / tie (:363-372,469-496) ----
        double rew[N];
#pragma unroll
        for (int k = 0; k < N; ++k) {
            double rw = 0.0;
            unsigned e2 = ev[k];
            int da = (int)((died_own >> (4 * k)) & 15u) - 1;
            // own bullet outcomes oldest first; a victim's die_punishment precedes its own same-age outcome for blue
            // (red bullets come first in the list) and follows it for red
            while (e2 != 0 || da >= 0) {
                const int ae = e2 ? ((__ffs(e2) - 1) >> 1) : 64;
                if (da >= 0 && (da < ae || (da == ae && team == 1))) { rw += p.r_die; da = -1; }
                else {
                    rw += p.r_code[(e2 >> (2 * ae)) & 3u];
                    e2 &= ~(3u << (2 * ae));
                }
            }
            rew[k] = rw;
        }
        const bool i_won = run && bhp_enemy <= 0;   // the winners' planes get lose_punishment (sic)
        const bool foe_won = run && bhp_own <= 0;
        if (i_won) {
#pragma unroll
            for (int k = 0; k < N; ++k) rew[k] += p.r_lose;
        }
        const bool red_won = team == 0 ? i_won : foe_won, blue_won = team == 0 ? foe_won : i_won;
        const bool tie = tie_now || time_tie;
        if (red_won) winner = W_RED;
        if (blue_won) winner = W_BLUE;   // both bases dead: ends as 'blue', two games counted
        if (tie) winner = W_TIE;
        if (red_won || blue_won || tie) env_done = 1;
        double rt = 0;
#pragma unroll
        for (int k = 0; k < N; ++k) rt += rew[k];
        if (active) { ret += rt; ticks_run += 1; }
        const bool ended = active && env_done;
        if (__any_sync(FULL, ended)) {   // (the same lanes in both warps of the block: the barrier below is matched)
            const long long sret = warp_sum(ended ? __double2ll_rn(ret * 1048576.0) : 0ll);
            const int nr = __popc(__ballot_sync(FULL, ended && red_won)), nb = __popc(__ballot_sync(FULL, ended && blue_won));
            const int nt = __popc(__ballot_sync(FULL, ended && tie));
#ifndef BS_NO_EPLOG
            if (p.ep_log != nullptr) {
                sm.retx[team][lane] = ret;
                __syncthreads();
                if (team == 0 && ended) {
                    const unsigned long long slot = atomicAdd(p.counters + 8, 1ull) & (unsigned long long)(p.ep_cap - 1);
                    p.ep_log[slot * 3 + 0] = (float)sm.retx[0][lane]; p.ep_log[slot * 3 + 1] = (float)sm.retx[1][lane];
                    p.ep_log[slot * 3 + 2] = tie ? 3.0f : (red_won && blue_won ? 6.0f : (red_won ? 1.0f : 2.0f));
                }
                __syncthreads();
            }
#endif
            if (lane == 0) {
                if (team == 0) {
                    if (nr) atomicAdd(p.counters + C_WINS_RED, (unsigned long long)nr);
                    if (nb) atomicAdd(p.counters + C_WINS_BLUE, (unsigned long long)nb);
                    if (nt) atomicAdd(p.counters + C_TIES, (unsigned long long)nt);
                    atomicAdd(p.counters + C_GAMES, (unsigned long long)(nr + nb + nt));
                    atomicAdd(p.counters + C_RET_RED, (unsigned long long)sret);
                } else {
                    atomicAdd(p.counters + C_RET_BLUE, (unsigned long long)sret);
                }
            }
        }

        // ---- the own planes' new observations -> shared memory; rows of the learner ----------------------------
        am = 0;
#pragma unroll
        for (int j = 0; j < N; ++j) am |= (((ehp >> (4 * j)) & 15u) ? 1u : 0u) << j;
#pragma unroll
        for (int k = 0; k < N; ++k) {
            float o[L];
            observe_plane<N>(p, pxy[k], dir[k], ((ohp >> (4 * k)) & 15u) != 0, ebase, exy, am, o);
#pragma unroll
            for (int c = 0; c < L; ++c) myobs[k * L + c] = o[c];
        }
        __syncwarp();
        if (writer) {
            if (ra.rp.new_state) rows_out<NL>(ra.rp.new_state, sm.obs[team], row, row0, nvalid, flat, lane);
            if (valid) {
#pragma unroll
                for (int k = 0; k < N; ++k) {
                    if (ra.rp.reward) ra.rp.reward[(size_t)row * N + k] = (float)rew[k];
                    if (ra.rp.done) ra.rp.done[(size_t)row * N + k] = (uint8_t)(env_done || ((ohp >> (4 * k)) & 15u) == 0);
                }
            }
        }
        row += emod32;                       // (row, emod32 < capacity <= 2^31 / N: no 32-bit overflow)
        if (row >= cap32) row -= cap32;
    }

    // ---- registers -> state ----------------------------------------------------------------------------------
    if (valid) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
            const size_t a = (size_t)e * (2 * N) + team * N + k;
            p.plane_pos[a] = pxy[k];
            p.plane_dir[a] = dir[k];
            p.plane_hp[a] = (int8_t)((ohp >> (4 * k)) & 15u);
        }
        p.base_hp[e * 2 + (1 - team)] = (int16_t)bhp_enemy;
        p.base_pos[e * 2 + (1 - team)] = ebase;
        if (team == 0) {
            p.meta[e] = (uint32_t)tick | ((uint32_t)env_done << 16) | ((uint32_t)winner << 24);
            p.episode[e] = episode;
        }
        p.ret[e * 2 + team] = ret;
        uint2 *dst = p.lists + (size_t)blk * S * 64 + tid;
        for (int s = 0; s < cnt; ++s) dst[(size_t)s * 64] = blt[s];
    }
    p.list_cnt[blk * 64 + tid] = valid ? cnt : 0;
    if (team == 0) {
        const long long tt = warp_sum((long long)ticks_run);
        if (lane == 0 && tt != 0) {
            atomicAdd(p.counters + C_TICKS, (unsigned long long)tt);
            atomicAdd(p.counters + C_AGENT_STEPS, (unsigned long long)(tt * (2 * N)));
        }
    }
}

// ---- layout conversion: per-warp pools (lane-per-plane kernels) <-> per-team lists (this kernel) -------------
// warp pools -> lists: one thread per (env, team), blocks of 64 as above.  The list must be in the reference's
// list order (oldest first, planes in team order); the order inside a pool carries no meaning, so the thread's
// entries are first placed by (age, plane) and then packed.
template <int N>
__global__ void __launch_bounds__(64) pools_to_lists_kernel(Params p) {
    constexpr int S = SLOTS * N, K = 32 / (2 * N);
    const int tid = threadIdx.x, lane = tid & 31, team = tid >> 5;
    const long long blk = blockIdx.x, e = blk * 32 + lane;
    int n = 0;
    if (e < p.E) {
        const long long gw = e / K;
        const int lo = (int)(e - gw * K) * 2 * N + team * N;   // owner lanes of this team in its warp
        uint2 tmp[S];
        for (int i = 0; i < S; ++i) tmp[i] = make_uint2(0u, 0u);
        const int cnt = p.pool_cnt[gw];
        const uint2 *src = p.pool + gw * POOL_CAP;
        for (int q = 0; q < cnt; ++q) {
            const uint2 en = src[q];
            const int owner = (int)(en.y >> 23) & 31, upd = (int)(en.y >> 28);
            if ((en.x & B_VALID) && owner >= lo && owner < lo + N && upd < SLOTS)
                tmp[(SLOTS - 1 - upd) * N + (owner - lo)] = make_uint2(en.x, (en.y & ~(31u << 23)) | ((uint32_t)(owner - lo) << 23));
        }
        uint2 *dst = p.lists + (size_t)blk * S * 64 + tid;
        for (int i = 0; i < S; ++i)
            if (tmp[i].x & B_VALID) { dst[(size_t)n * 64] = tmp[i]; n += 1; }
    }
    p.list_cnt[blk * 64 + tid] = n;
}

// lists -> warp pools: one warp per pool, lane l < 2K copies the list of (env l / 2, team l % 2) behind the others'
template <int N>
__global__ void __launch_bounds__(128) lists_to_pools_kernel(Params p) {
    constexpr int S = SLOTS * N, K = 32 / (2 * N);
    const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= p.W) return;
    const int lane = threadIdx.x & 31, sub = lane >> 1, team = lane & 1;
    const long long e = gw * K + sub;
    const bool has = sub < K && e < p.E;
    const long long blk = has ? e / 32 : 0;
    const int tid2 = team * 32 + (int)(has ? e - blk * 32 : 0);
    const int n = has ? p.list_cnt[blk * 64 + tid2] : 0;
    int off = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(FULL, off, o); if (lane >= o) off += v; }
    const int total = __shfl_sync(FULL, off, 31);
    off -= n;
    const uint2 *src = p.lists + (size_t)blk * S * 64 + tid2;
    uint2 *dst = p.pool + gw * POOL_CAP + off;
    const uint32_t lo = (uint32_t)(sub * 2 * N + team * N);
    for (int s = 0; s < n; ++s) {
        const uint2 en = src[(size_t)s * 64];
        dst[s] = make_uint2(en.x, (en.y & ~(31u << 23)) | ((lo + ((en.y >> 23) & 7u)) << 23));
    }
    if (lane == 0) p.pool_cnt[gw] = total;
}

}  // namespace bs
