// battlespace.cu -- kernels + C ABI (include/battlespace.h) of libbattlespace.so, sm_100a only.
//
// Kernels (all use the lane-per-plane mapping of tick.cuh):
//   reset_kernel     parallel_env.reset for masked envs              battle_env.py:246-279
//   step_kernel      one tick, actions from HBM and/or in-kernel policies   battle_env.py:281-381
//   rollout_kernel   T fused ticks + auto-reset + replay rows        main.py:178-213, maddpg/buffer.py:25-47
//   instinct_kernel  instinct Team.choose_actions on an obs tensor   instinct/agent.py:10-62
// There is no CPU path in this library.
#include "../../include/battlespace.h"
#include "tick.cuh"
#include "rollout2.cuh"
#include "actor.cuh"
#include "render.cuh"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

using namespace bs;

// ------------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ long long global_warp() {
    return ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
}

__device__ __forceinline__ double clip1(double v) { return clamp1(v); }  // np.clip(value, -1, 1), battle_env.py:295-297

__device__ __forceinline__ int argmax4(const float *v) {  // np.argmax: first maximum, battle_env.py:328
    int b = 0;
    float m = v[0];
#pragma unroll
    for (int c = 1; c < 4; ++c) if (v[c] > m) { m = v[c]; b = c; }
    return b;
}

template <int N>
__global__ void __launch_bounds__(BLOCK) reset_kernel(Params p, const uint8_t *mask, const int32_t *draws, float *obs) {
    constexpr int P = 2 * N, L = 3 * N + 2;
    __shared__ WarpPool pools[BLOCK / 32];
    WarpPool &w = pools[threadIdx.x >> 5];
    const long long gw = global_warp();
    Geo<N> g(gw, p.E);
    Lane<N> s;
    double ret0, ret1;
    load_lane(p, g, s, ret0, ret1);
    const bool need = g.valid() && (mask == nullptr || mask[g.e()] != 0);
    if (need) {
        reset_lane(p, g, s, draws != nullptr, draws ? draws + g.e() * (4 + 6 * N) : nullptr);
        ret0 = ret1 = 0;
    }
    const unsigned killmask = __ballot_sync(FULL, need);
    if (killmask) {  // self.bullets = [] for the envs that were reset, battle_env.py:272
        int cnt;
        pool_load(p, gw, gw < p.W, g.lane(), w, cnt);
        pool_filter(w, cnt, killmask, g.lane());
        pool_store(p, gw, gw < p.W, g.lane(), w, cnt);
    }
    gather(g, s);
    if (obs != nullptr) {
        float o[L];
        observe<N>(p, s, o);
        if (g.valid()) {
            float *dst = obs + (g.e() * P + g.i()) * L;
#pragma unroll
            for (int c = 0; c < L; ++c) dst[c] = o[c];
        }
    }
    if (need) store_lane(p, g, s, ret0, ret1, true);
}

struct StepArgs {
    const void *actions;  // layout [E, act_agents, ...]
    int dtype;
    int act_agents;       // 2N (both teams external) or N (act_team only)
    int act_team;
    int pol[2];           // per team: POL_*
    const double *draws;  // null or [E,2N,5]
    float *obs, *rew;
    uint8_t *done, *env_done;
    int8_t *winner;
    bs_replay rp;         // optional: store_transition for act_team (capacity 0 = off)
    long long cursor_mod; // cursor % capacity
    int auto_reset;       // envs that finish in this tick are reset before the call returns
    const float *obs_in;  // optional [E,2N,L]: the CURRENT observation of every env (what the previous step / reset
                          // wrote): read instead of being recomputed before the tick
};

// Decode the lane's policy into (adisc, a[3], spread_u).  o = the lane's current observation
// (only read by the instinct policies).  raw[4] = the action row a replay buffer would store.
template <int N, bool CONT>
__device__ __forceinline__ void lane_policy(const Params &p, const Geo<N> &g, const Lane<N> &s, int pol,
                                            const float (&o)[3 * N + 2], const double *draws_row,
                                            int &adisc, double (&a)[3], double &spread_u, float (&raw)[4]) {
    double u[5];
    uint32_t w[4] = {0, 0, 0, 0};
    if (draws_row != nullptr) {
#pragma unroll
        for (int c = 0; c < 5; ++c) u[c] = draws_row[c];
    } else {
        env_philox(p, g.e(), s.episode, s.tick, g.i(), PURPOSE_AGENT_TICK, w);
        agent_tick_units(w, u);
    }
    spread_u = u[4];
    if (pol == POL_RANDOM) {
        if (CONT) {
#pragma unroll
            for (int c = 0; c < 3; ++c) { a[c] = 2 * u[1 + c] - 1; raw[c] = (float)a[c]; }
            adisc = 0;
        } else {
            // Philox: the top two bits of w0; injected draws: the same two bits of noise_u0 = (w0 >> 8) / 2^24
            adisc = draws_row != nullptr ? (int)(u[1] * 4.0) : (int)(w[0] >> 30);
#pragma unroll
            for (int c = 0; c < 4; ++c) raw[c] = c == adisc ? 1.0f : 0.0f;
        }
    } else if (pol == POL_INSTINCT || pol == POL_INSTINCT_FF) {
        instinct<N, CONT>(o, u[0], u[1], u[2], u[3], adisc, a);
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

// The same for the fused rollout, where the draws are always Philox: the uniform policy's actions 2u - 1 come
// straight from the 24-bit integers (sym24) and only the units a policy consumes are built.
template <int N, bool CONT>
__device__ __forceinline__ void lane_policy_philox(const Params &p, const Geo<N> &g, const Lane<N> &s, int pol,
                                                   const float (&o)[3 * N + 2], int &adisc, double (&a)[3],
                                                   double &spread_u, float (&raw)[4]) {
    uint32_t w[4];
    env_philox(p, g.e(), s.episode, s.tick, g.i(), PURPOSE_AGENT_TICK, w);
    spread_u = unit32(w[3]);
    if (pol == POL_RANDOM) {
        if (CONT) {
#pragma unroll
            for (int c = 0; c < 3; ++c) { a[c] = sym24(w[c] >> 8); raw[c] = (float)a[c]; }
            adisc = 0;
        } else {
            adisc = (int)(w[0] >> 30);
#pragma unroll
            for (int c = 0; c < 4; ++c) raw[c] = c == adisc ? 1.0f : 0.0f;
        }
    } else if (pol == POL_INSTINCT || pol == POL_INSTINCT_FF) {
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
    else { adisc = 0; a[0] = a[1] = a[2] = 0; raw[0] = raw[1] = raw[2] = 0; }
    if (CONT) raw[3] = 0;
}

// 64-thread blocks, 12 resident (24 warps per SM at 80 registers): measured 1.5 % faster than 128 x 6 in Mode A
// (finer-grained block turnover over the 3.7 waves); 7-8 blocks of 128 and 256-thread blocks are 4 % slower.
#ifndef STEP_BLOCK
#define STEP_BLOCK 64
#endif
#ifndef STEP_MIN_BLOCKS
#define STEP_MIN_BLOCKS (768 / STEP_BLOCK)
#endif
#ifndef STEP_MIN_BLOCKS_LEAN
#define STEP_MIN_BLOCKS_LEAN (896 / STEP_BLOCK)
#endif
// LEAN = plain bs_step (Mode A): both teams' actions from the caller, no replay rows, no auto-reset, no in-kernel policy,
// no obs_in.  The general kernel needs 80 registers (24 resident warps per SM); without those paths the same tick fits
// 72 registers / 28 resident warps with no spill (r2y: 418 us at 1,048,576 envs; squeezed into 64 registers / 32 warps it
// took 422 us, into 56 or 48 registers -- 36 / 40 warps, 28 / 92 bytes of spills -- 582 / 576 us).
template <int N, bool CONT, bool LEAN>
__global__ void __launch_bounds__(STEP_BLOCK, LEAN ? STEP_MIN_BLOCKS_LEAN : STEP_MIN_BLOCKS) step_kernel(Params p, StepArgs sa) {
    constexpr int P = 2 * N, L = 3 * N + 2;
    __shared__ WarpPool pools[STEP_BLOCK / 32];
    WarpPool &w = pools[threadIdx.x >> 5];   // (the opaque-pointer form of rollout_kernel: 1.5 % slower here at 64 registers, no difference at 72)
    const long long gw = global_warp();
    Geo<N> g(gw, p.E);
    Lane<N> s;
    double ret0, ret1;
    // programmatic dependent launch (bs_rollout_actor): nothing is read before the previous grid has
    // completed; the next grid may be scheduled as soon as every block of this one has started
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;");
    int cnt;
    // every load of the prologue is issued before the first one is waited for: state, pool entries + count, and (plain
    // bs_step with float actions) the lane's action words -- one memory round trip instead of three
    float araw[4] = {0, 0, 0, 0};
    const bool early_act = LEAN && sa.dtype == ACT_F32 && sa.actions != nullptr;   // kernel-uniform
    {
        LaneRaw lr;
        uint2 pv[POOL_SPEC_N];
        lane_fetch(p, g, lr);
        pool_fetch(p, gw, gw < p.W, g.lane(), pv, cnt);
        {   // (no branch, so that the compiler keeps these loads with the others: without float actions they read the first
            //  words of the observation table, which is always there, and the result is not used)
            constexpr int A = CONT ? 3 : 4;
            const long long idx = (g.valid() ? g.e() : 0) * sa.act_agents + (sa.act_agents == P ? g.i() : g.k());
            const float *ap = early_act ? (const float *)sa.actions + idx * A : (const float *)p.obs_tab;
#pragma unroll
            for (int c = 0; c < A; ++c) araw[c] = ap[c];
        }
        lane_unpack(g, lr, s, ret0, ret1);
        pool_commit(p, gw, gw < p.W, g.lane(), pv, w, cnt);
    }
    pool_publish_base<N>(g, s, w);
    // (Measured and not kept, r2i: one prefetch.global.L2 per lane for the lines of the block that will take over this
    // block's slot -- 431.9 vs 431.2 us at 1,048,576 envs: the wait on the first loads is not the DRAM leg.)

    float o[L];
    const bool writer = !LEAN && g.valid() && sa.rp.capacity > 0 && g.team() == sa.act_team;
    long long row = 0;
    if (!LEAN && sa.rp.capacity > 0) row = (sa.cursor_mod + g.e() % sa.rp.capacity) % sa.rp.capacity;
    const bool need_obs = !LEAN && (sa.pol[0] >= POL_INSTINCT || sa.pol[1] >= POL_INSTINCT || sa.rp.capacity > 0);  // kernel-uniform
    if (need_obs && sa.obs_in != nullptr) {
        const float *src = sa.obs_in + ((g.valid() ? g.e() : 0) * P + g.i()) * L;
#pragma unroll
        for (int c = 0; c < L; ++c) o[c] = g.valid() ? src[c] : -1.0f;
    } else if (need_obs) { gather(g, s); observe<N>(p, s, o); }
    else {
#pragma unroll
        for (int c = 0; c < L; ++c) o[c] = -1.0f;
    }

    // external actions
    const int pol = LEAN ? (int)POL_EXTERNAL : (g.team() ? sa.pol[1] : sa.pol[0]);  // (a dynamic index would make the compiler copy the parameter struct to local memory)
    const bool ext = g.valid() && pol == POL_EXTERNAL && sa.dtype != ACT_NONE;
    int adisc = 0;
    double a[3] = {0, 0, 0}, spread_u = 0;
    float raw[4] = {0, 0, 0, 0};
    if (ext) {
        const long long idx = g.e() * sa.act_agents + (sa.act_agents == P ? g.i() : g.k());
        if (!CONT) {
            if (sa.dtype == ACT_I32) adisc = ((const int32_t *)sa.actions)[idx];
            else {
#pragma unroll
                for (int c = 0; c < 4; ++c) raw[c] = early_act ? araw[c] : ((const float *)sa.actions)[idx * 4 + c];
            }
        } else if (sa.dtype == ACT_F32) {
#pragma unroll
            for (int c = 0; c < 3; ++c) a[c] = (double)(early_act ? araw[c] : ((const float *)sa.actions)[idx * 3 + c]);
        } else {
#pragma unroll
            for (int c = 0; c < 3; ++c) a[c] = ((const double *)sa.actions)[idx * 3 + c];
        }
    }
    if (g.valid()) {
        const double *drow = sa.draws ? sa.draws + (g.e() * P + g.i()) * 5 : nullptr;
        if (ext) {
            if (!CONT) {
                if (sa.dtype == ACT_I32) { if (adisc >= 0 && adisc < 4) raw[adisc] = 1.0f; }
                else adisc = argmax4(raw);
            } else {
#pragma unroll
                for (int c = 0; c < 3; ++c) a[c] = clip1(a[c]);
            }
            if (drow) spread_u = drow[4];
            else {
                uint32_t w[4]; double u[5];
                env_philox(p, g.e(), s.episode, s.tick, g.i(), PURPOSE_AGENT_TICK, w);
                agent_tick_units(w, u);
                spread_u = u[4];
            }
        } else if (pol != POL_EXTERNAL) {
            lane_policy<N, CONT>(p, g, s, pol, o, drow, adisc, a, spread_u, raw);
        }
    }

    if (writer) {  // ReplayBuffer.store_transition, maddpg/buffer.py:25-47: the part known before the tick
        constexpr int A = CONT ? 3 : 4;
        if (CONT && pol == POL_EXTERNAL) {
#pragma unroll
            for (int c = 0; c < 3; ++c) raw[c] = (float)a[c];
        }
        if (sa.rp.state) {
            float *dst = sa.rp.state + (row * N + g.k()) * L;
#pragma unroll
            for (int c = 0; c < L; ++c) dst[c] = o[c];
        }
        if (sa.rp.action) {
#pragma unroll
            for (int c = 0; c < A; ++c) sa.rp.action[(row * N + g.k()) * A + c] = raw[c];
        }
    }
    double rew;
    Tally tl;
    const bool want_obs = sa.obs != nullptr || (!LEAN && sa.rp.capacity > 0);  // kernel-uniform
    {
        TickCtx c;
        tick_begin<N, CONT>(p, g, s, adisc, a, spread_u, sa.dtype == ACT_NONE, c);
        // positions are final once the planes have moved: the table gathers of the new observation are
        // issued before the bullet pass, which hides their L2 latency (42 % of this kernel's
        // long-scoreboard stalls sat on them, profiles/r1j_step_team_*)
        ObsLoads<N> q;
        gather(g, s);  // enemy positions after the move, hp before the bullet loop
        if (want_obs) observe_issue<N>(p, s, q);
        unsigned ev; int died_ai;
        bullets_dense<N, false>(g, s, w, cnt, c, ev, died_ai);
        tick_end<N>(p, g, s, c, ev, died_ai, rew, ret0, ret1, tl);
        if (want_obs) observe_finish<N>(s, q, o);
    }
    if (want_obs) {
        if (writer) {
            const long long ra_i = row * N + g.k();
            if (sa.rp.new_state) {
                float *dst = sa.rp.new_state + ra_i * L;
#pragma unroll
                for (int c = 0; c < L; ++c) dst[c] = o[c];
            }
            if (sa.rp.reward) sa.rp.reward[ra_i] = (float)rew;
            if (sa.rp.done) sa.rp.done[ra_i] = (uint8_t)(s.env_done || s.hp <= 0);
        }
    }
    if (sa.obs != nullptr) {
        if (g.valid()) {
            float *dst = sa.obs + (g.e() * P + g.i()) * L;
#pragma unroll
            for (int c = 0; c < L; ++c) dst[c] = o[c];
        }
    }
    // auto-reset (the game loop's `env.reset()` for the next game, main.py:148): rewards, dones,
    // env_done and winner below still describe the finished game; obs is the new game's first one.
    const bool fresh = !LEAN && sa.auto_reset && g.valid() && s.env_done;
    const int out_done = s.env_done, out_hp = s.hp, out_winner = s.winner;
    if (!LEAN && sa.auto_reset && __any_sync(FULL, fresh)) {
        if (fresh) { reset_lane(p, g, s, false, nullptr); ret0 = ret1 = 0; }
        pool_filter(w, cnt, __ballot_sync(FULL, fresh), g.lane());  // the finished games' bullets vanish, battle_env.py:272
        gather(g, s);
        float o2[L];
        observe<N>(p, s, o2);
        if (fresh && sa.obs != nullptr) {
            float *dst = sa.obs + (g.e() * P + g.i()) * L;
#pragma unroll
            for (int c = 0; c < L; ++c) dst[c] = o2[c];
        }
    }
    if (g.valid()) {
        const long long ai = g.e() * P + g.i();
        if (sa.rew) sa.rew[ai] = (float)rew;
        if (sa.done) sa.done[ai] = (uint8_t)(out_done || out_hp <= 0);  // self.dones, battle_env.py:275,360,478,494
        if (g.i() == 0) {
            if (sa.env_done) sa.env_done[g.e()] = (uint8_t)out_done;
            if (sa.winner) sa.winner[g.e()] = (int8_t)out_winner;
        }
    }
    store_lane(p, g, s, ret0, ret1, !LEAN && sa.auto_reset != 0);
    pool_store(p, gw, gw < p.W, g.lane(), w, cnt);
    flush_tally<N>(p, tl);
}

// 64-thread blocks x 14 resident = 28 warps per SM at 72 registers.  Round 1 ran 128 x 8 (32 warps at 64 registers; 6, 7, 9
// and 10 blocks all measured slower then); after the branch-free passes of round 2 the 8 extra registers (spills 36 -> 8
// bytes, fewer IMAD.MOV) are worth more than the 4 warps: 128 x 8 1.787e10, 128 x 7 1.816e10, 64 x 14 1.821e10,
// 32 x 28 1.806e10, 128 x 6 (80 registers) 1.776e10, 256 x 3 1.717e10 agent-steps/s (A/B on the bench command).
#ifndef ROLLOUT_BLOCK
#define ROLLOUT_BLOCK 64
#endif
#ifndef ROLLOUT_MIN_BLOCKS
#define ROLLOUT_MIN_BLOCKS (896 / ROLLOUT_BLOCK)
#endif
template <int N, bool CONT>
__global__ void __launch_bounds__(ROLLOUT_BLOCK, ROLLOUT_MIN_BLOCKS) rollout_kernel(Params p, RolloutArgs ra) {
    constexpr int L = 3 * N + 2, A = CONT ? 3 : 4;
    __shared__ WarpPool pools[ROLLOUT_BLOCK / 32];
#ifndef BS_POOL_REMAT
    // The warp's pool address in a register the compiler cannot re-derive: a shared-memory address on sm_100 carries the
    // CTA's cluster rank, so "recomputing" &pools[warp] is S2R CgaCtaId + S2R TID + 4 ALU -- and r2n's SASS did that 15
    // times, once inside every pass of the bullet loop (25 + ~10 warp-instructions per warp-tick, two S2R latencies each).
    WarpPool *wp = &pools[threadIdx.x >> 5];
    asm volatile("" : "+l"(wp));
    __builtin_assume(__isShared(wp));
    WarpPool &w = *wp;
#else
    WarpPool &w = pools[threadIdx.x >> 5];
#endif
    Geo<N> g(global_warp(), ra.e_end, ra.e0);
    Lane<N> s;
    double ret0, ret1;
    load_lane(p, g, s, ret0, ret1);
    int cnt;
    pool_load(p, ra.e0 / Geo<N>::K + global_warp(), ra.e0 + global_warp() * Geo<N>::K < ra.e_end, g.lane(), w, cnt);  // chunks start on a warp boundary
    pool_publish_base<N>(g, s, w);
    gather(g, s);
    float o[L];
    observe<N>(p, s, o);
    Tally tl;
    const bool writer = g.valid() && g.team() == ra.learner && ra.rp.capacity > 0;
    // which ring arrays this lane writes, decided once: the tick loop tests one bit instead of re-deriving `writer` and
    // null-checking five 64-bit pointers every tick
    unsigned wf = 0;
    if (writer) wf = (ra.rp.state ? 1u : 0u) | (ra.rp.action ? 2u : 0u) | (ra.rp.new_state ? 4u : 0u) | (ra.rp.reward ? 8u : 0u) | (ra.rp.done ? 16u : 0u);
    wf |= (unsigned)(g.team() ? ra.pol[1] : ra.pol[0]) << 8;   // the lane's policy rides in the same register
    asm volatile("" : "+r"(wf));   // (keep it in its register: the compiler would rather recompute it)
    // ring row of this lane's env at tick t (replay_check: capacity * N < 2^31, so 32-bit indices address every array)
    uint32_t row = 0;
    if (ra.rp.capacity > 0) row = (uint32_t)((ra.cursor_mod + g.e() % ra.rp.capacity) % ra.rp.capacity);
    const uint32_t cap32 = (uint32_t)ra.rp.capacity, emod32 = (uint32_t)ra.e_mod;

#pragma unroll 1
    for (int t = 0; t < ra.T; ++t) {
        const int pol = (int)(wf >> 8);
        // auto-reset: the game loop's env.reset() -- main.py:148
        const bool need = g.valid() && s.env_done;
        if (__any_sync(FULL, need)) {
            if (need) { reset_lane(p, g, s, false, nullptr); ret0 = ret1 = 0; pool_publish_base<N>(g, s, w); }
            pool_filter(w, cnt, __ballot_sync(FULL, need), g.lane());  // the finished games' bullets vanish, battle_env.py:272
            gather(g, s);
            float o2[L];
            observe<N>(p, s, o2);
            if (need) {
#pragma unroll
                for (int c = 0; c < L; ++c) o[c] = o2[c];
            }
        }
        const uint32_t ra_i = row * N + g.k();   // this plane's slot of the row
        if (wf & 1u) {
            float *dst = ra.rp.state + (size_t)ra_i * L;
#pragma unroll
            for (int c = 0; c < L; ++c) dst[c] = o[c];
        }
        // an in-kernel policy touches registers only: the warp's idle lanes run it too, which saves the defaults
        // (13 moves) and the divergent region that every warp pays for its two spare lanes
        int adisc;
        double a[3], spread_u;
        float raw[4];
        if (pol != POL_EXTERNAL) {
            lane_policy_philox<N, CONT>(p, g, s, pol, o, adisc, a, spread_u, raw);
        } else {
            {   // (an idle lane reads the row of the launch's first env: always there, never used)
                adisc = 0; raw[3] = 0;
                if (!CONT) a[0] = a[1] = a[2] = 0;
                const float *src = ra.ext + (((long long)t * p.E + (g.valid() ? g.e() : ra.e0)) * N + g.k()) * A;
                if (CONT) {
#pragma unroll
                    for (int c = 0; c < 3; ++c) { a[c] = clip1((double)src[c]); raw[c] = (float)a[c]; }
                } else {
#pragma unroll
                    for (int c = 0; c < 4; ++c) raw[c] = src[c];
                    adisc = argmax4(raw);
                }
                uint32_t w4[4]; double u[5];
                env_philox(p, g.e(), s.episode, s.tick, g.i(), PURPOSE_AGENT_TICK, w4);
                agent_tick_units(w4, u);
                spread_u = u[4];
            }
        }
        if (wf & 2u) {
            float *dst = ra.rp.action + (size_t)ra_i * A;
#pragma unroll
            for (int c = 0; c < A; ++c) dst[c] = raw[c];
        }
        double rew;
        {
            TickCtx c;
            tick_begin<N, CONT>(p, g, s, adisc, a, spread_u, false, c);
            unsigned ev; int died_ai;
            gather(g, s);  // enemy positions after the move, hp before the bullet loop
            bullets_dense<N, true>(g, s, w, cnt, c, ev, died_ai);
            ObsLoads<N> q;
            observe_issue<N>(p, s, q);  // positions are final; hp / alive masks are settled by then
            tick_end<N>(p, g, s, c, ev, died_ai, rew, ret0, ret1, tl);
            observe_finish<N>(s, q, o);
        }
        if (wf & 4u) {
            float *dst = ra.rp.new_state + (size_t)ra_i * L;
#pragma unroll
            for (int c = 0; c < L; ++c) dst[c] = o[c];
        }
        if (wf & 8u) ra.rp.reward[ra_i] = (float)rew;
        if (wf & 16u) ra.rp.done[ra_i] = (uint8_t)(s.env_done || s.hp <= 0);
        row += emod32;                       // (row, emod32 < capacity <= 2^31 / N: no 32-bit overflow)
        if (row >= cap32) row -= cap32;
    }
    store_lane(p, g, s, ret0, ret1, true);
    pool_store(p, ra.e0 / Geo<N>::K + global_warp(), ra.e0 + global_warp() * Geo<N>::K < ra.e_end, g.lane(), w, cnt);
    flush_tally<N>(p, tl);
}

template <int N, bool CONT>
__global__ void __launch_bounds__(BLOCK) instinct_kernel(Params p, int team, const float *obs, const double *draws,
                                                         void *out) {
    constexpr int P = 2 * N, L = 3 * N + 2;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.E * N) return;
    const long long e = t / N;
    const int k = (int)(t - e * N), i = team * N + k;
    float o[L];
    const float *src = obs + (e * P + i) * L;
#pragma unroll
    for (int c = 0; c < L; ++c) o[c] = src[c];
    double u[5];
    if (draws) {
#pragma unroll
        for (int c = 0; c < 5; ++c) u[c] = draws[(e * P + i) * 5 + c];
    } else {
        uint32_t w[4];
        const uint32_t m = p.meta[e];
        env_philox(p, e, p.episode[e], (int)(m & 0xfff), i, PURPOSE_AGENT_TICK, w);
        agent_tick_units(w, u);
    }
    int adisc; double a[3];
    instinct<N, CONT>(o, u[0], u[1], u[2], u[3], adisc, a);
    if (CONT) {
        double *dst = (double *)out + t * 3;
        dst[0] = a[0]; dst[1] = a[1]; dst[2] = a[2];
    } else {
        ((int32_t *)out)[t] = adisc;
    }
}

// ReplayBuffer.sample_buffer -- maddpg/buffer.py:49-67 (row f2).  One block row per sampled index:
// the row's N*L state words, N*A action words and N rewards / dones are gathered with consecutive
// threads on consecutive words; the per-agent copies (actor_states[k] = columns k*L.. of state_mem)
// are written from the same registers, so each ring row is read once.
struct SampleArgs {
    bs_replay rp;
    const long long *idx;
    long long B;
    int N, L, A;
    float *actor_states, *states, *actions, *rewards, *actor_new_states, *new_states;
    uint8_t *dones;
};

__global__ void __launch_bounds__(128) replay_sample_kernel(SampleArgs sa) {
    const int NL = sa.N * sa.L, NA = sa.N * sa.A;
    for (long long b = blockIdx.x; b < sa.B; b += gridDim.x) {
        long long row = sa.idx[b];
        row = row < 0 ? 0 : (row >= sa.rp.capacity ? sa.rp.capacity - 1 : row);
        for (int c = threadIdx.x; c < NL; c += blockDim.x) {
            const int k = c / sa.L, j = c - k * sa.L;
            const float v = __ldg(sa.rp.state + row * NL + c), v2 = __ldg(sa.rp.new_state + row * NL + c);
            if (sa.states) sa.states[b * NL + c] = v;
            if (sa.new_states) sa.new_states[b * NL + c] = v2;
            if (sa.actor_states) sa.actor_states[((long long)k * sa.B + b) * sa.L + j] = v;
            if (sa.actor_new_states) sa.actor_new_states[((long long)k * sa.B + b) * sa.L + j] = v2;
        }
        for (int c = threadIdx.x; c < NA; c += blockDim.x) {
            const int k = c / sa.A, j = c - k * sa.A;
            if (sa.actions) sa.actions[((long long)k * sa.B + b) * sa.A + j] = __ldg(sa.rp.action + row * NA + c);
        }
        if (threadIdx.x < sa.N) {
            if (sa.rewards) sa.rewards[b * sa.N + threadIdx.x] = __ldg(sa.rp.reward + row * sa.N + threadIdx.x);
            if (sa.dones) sa.dones[b * sa.N + threadIdx.x] = sa.rp.done[row * sa.N + threadIdx.x];
        }
    }
}

// State export / import move only the LIVE pool entries: pack copies each warp's cnt entries to off[w].. of a
// dense buffer (one warp per pool, coalesced), unpack is the inverse.  (Round 1 copied the whole W x 384-entry
// pool array: 1.2 GB for a million 5v5 envs to export a few live bullets per env.)
__global__ void __launch_bounds__(128) pool_pack_kernel(const uint2 *pool, const int *cnt, const long long *off, long long W, uint2 *dense, int unpack) {
    const long long wv = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (wv >= W) return;
    const int lane = threadIdx.x & 31, n = cnt[wv];
    uint2 *a = const_cast<uint2 *>(pool) + wv * POOL_CAP;
    uint2 *b = dense + off[wv];
    for (int q = lane; q < n; q += 32) { if (unpack) a[q] = b[q]; else b[q] = a[q]; }
}

// ------------------------------------------------------------------------------------------------
// host side: handle + C ABI
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_err;

static int fail(int code, const std::string &msg) { g_err = msg; return code; }

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(BS_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));            \
    } while (0)

struct ObsTables { double2 *tab = nullptr; size_t bytes = 0; bool persist = false; };

struct bs_handle {
    bs_config cfg;
    int N, P, L, A, K;
    bool cont;
    int64_t E, W;
    int device;
    Params p;
    ObsTables tables;
    std::vector<void *> allocs;
    // Bullets live in one of two HBM layouts: the per-warp pools of the lane-per-plane kernels (reset, step, render,
    // state export) or the per-team lists of rollout2_kernel; ensure_bullets() converts when a call needs the other one.
    int bullet_fmt = 0;              // FMT_POOLS / FMT_LISTS
    bool rollout_v2 = false;         // BS_ROLLOUT_V2=1 at bs_create: the fused rollout runs the team-per-thread kernel (rollout2.cuh;
                                     // measured slower than the lane-per-plane kernel at 3v3, DESIGN.md 7: kept for A/B and as a cross-check)
    cudaStream_t hstream = nullptr;  // used by the *_host entry points
    cudaStream_t cstream = nullptr;  // copy stream of bs_rollout_fused_host
    cudaEvent_t chunk_ev[16] = {};
    int host_chunks = 0;             // BS_HOST_CHUNKS (tuning), 0 = default
    // split-phase host rollout (bs_rollout_fused_host_begin/_end): two action staging buffers (stage slots 8, 9)
    // in rotation, so that the copies of call k+1 overlap the kernels of call k
    cudaEvent_t call_done[2] = {};   // recorded on hstream after a call's last kernel and its counter read-back
    cudaEvent_t copy_done = nullptr; // recorded on cstream after a call's last action copy
    int64_t *pinned_stats = nullptr; // [2][8] pinned host memory
    int inflight = 0, next_buf = 0;  // calls begun and not yet ended (0..2); staging buffer of the next call
    // device staging for *_host calls (grown on demand)
    void *stage[12] = {};            // slots 10, 11: offsets / dense entries of bs_get_state, bs_set_state
    size_t stage_sz[12] = {};
};

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// ---- observation tables (tick.cuh pair_obs) ------------------------------------------------------
// One copy per device, shared by every handle on it and never freed (34.6 MB).  Evaluated on the host
// with the host libm and one rounding per operation, i.e. exactly what CPython computes for
// rel_angle / dist (battle_env.py:38-58) on this machine.
static ObsTables g_tables[64];
static std::mutex g_tables_mu;

static int obs_tables_for(int device, ObsTables *out) {
    std::lock_guard<std::mutex> lk(g_tables_mu);
    if (device < 0 || device >= 64) return fail(BS_E_INVALID, "device index out of range");
    ObsTables &tb = g_tables[device];
    if (!tb.tab) {
        // record (dx, dy >= 0): atan2(dy, dx) and float32(dist/diag*2-1).  atan2(-dy, dx) is recovered on the
        // device as -atan2(dy, dx): checked here, bit for bit, against this machine's libm for every offset.
        std::vector<double2> rec((size_t)T_W * T_H);
        const volatile double diag = std::sqrt(std::pow(1200.0, 2) + std::pow(800.0, 2));  // battle_env.py:230
        for (int dy = 0; dy <= T_DY; ++dy)
            for (int dx = -T_DX; dx <= T_DX; ++dx) {
                volatile double r = std::atan2((double)dy, (double)dx);      // battle_env.py:46
                if (dy > 0) {
                    volatile double rn = std::atan2((double)-dy, (double)dx);
                    if (rn != -r) return fail(BS_E_UNSUPPORTED, "host libm atan2 is not odd in y: observation table cannot be folded");
                }
                volatile double d = std::sqrt((double)(dx * dx + dy * dy)); // battle_env.py:58
                volatile double a = d / diag;
                volatile double b = a * 2;
                const float od = (float)(b - 1);                            // stored into a float32 array, :230
                uint32_t bits; std::memcpy(&bits, &od, 4);
                double packed; const uint64_t w64 = bits; std::memcpy(&packed, &w64, 8);
                rec[(size_t)dy * T_W + (dx + T_DX)] = make_double2(r, packed);
            }
        const size_t tbytes = rec.size() * sizeof(double2);
        double2 *blk = nullptr;
        CU(cudaMalloc(&blk, tbytes));
        CU(cudaMemcpy(blk, rec.data(), tbytes, cudaMemcpyHostToDevice));
        tb.bytes = tbytes;
        // Pin the tables in L2: set aside enough persisting L2 for them (B200: 126 MB L2).
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, device));
        if (!getenv("BS_NO_L2_PERSIST") && (size_t)prop.persistingL2CacheMaxSize >= tb.bytes &&
            (size_t)prop.accessPolicyMaxWindowSize >= tb.bytes &&
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, tb.bytes + (4u << 20)) == cudaSuccess)
            tb.persist = true;
        cudaGetLastError();
        tb.tab = blk;
    }
    *out = tb;
    return BS_OK;
}

// ---- static scratch of bs_rollout_actor_static (no reference counterpart; added for callers that may not allocate) ----
// One file-scope array per loaded library image and device: every GPU that loads libbattlespace.so holds these
// BS_STATIC_SCRATCH_BYTES until the process exits.  In HOST code the name g_static_scratch is only a stand-in for the
// symbol, not a device address: the kernels get the pointer cudaGetSymbolAddress returns, resolved once per device with
// that device current.  All calls on a device share the array: calls that overlap on the GPU (two streams, two handles)
// race on it.  Nothing in it has to start at zero -- every byte a call reads was written earlier by that call (the actor
// kernel writes all E*N*A actions before the step kernel reads them; the step kernel writes env_done[0,E) before the
// next tick's actor kernel reads it as the OU reset mask; tick 0 uses the caller's noise->reset_mask).
static __device__ __align__(256) unsigned char g_static_scratch[BS_STATIC_SCRATCH_BYTES];
static unsigned char *g_static_scratch_ptr[64];
static std::mutex g_static_scratch_mu;

static int static_scratch_for(int device, unsigned char **out) {   // the caller has made `device` current
    std::lock_guard<std::mutex> lk(g_static_scratch_mu);
    if (device < 0 || device >= 64) return fail(BS_E_INVALID, "device index out of range");
    if (!g_static_scratch_ptr[device]) {
        int cur = -1;
        CU(cudaGetDevice(&cur));
        if (cur != device) return fail(BS_E_CUDA, "static scratch: the handle's device is not current");
        void *q = nullptr;
        cudaError_t ce = cudaGetSymbolAddress(&q, g_static_scratch);
        if (ce != cudaSuccess || !q)
            return fail(BS_E_CUDA, std::string("cudaGetSymbolAddress(g_static_scratch): ") + cudaGetErrorString(ce));
        g_static_scratch_ptr[device] = static_cast<unsigned char *>(q);
    }
    *out = g_static_scratch_ptr[device];
    return BS_OK;
}

template <typename T>
static int dalloc(bs_handle *h, T **ptr, size_t n) {
    void *q = nullptr;
    CU(cudaMalloc(&q, n * sizeof(T)));
    CU(cudaMemset(q, 0, n * sizeof(T)));
    h->allocs.push_back(q);
    *ptr = (T *)q;
    return BS_OK;
}

static int stage_buf(bs_handle *h, int slot, size_t bytes, void **out) {
    if (h->stage_sz[slot] < bytes) {
        if (h->stage[slot]) CU(cudaFree(h->stage[slot]));
        h->stage[slot] = nullptr; h->stage_sz[slot] = 0;
        CU(cudaMalloc(&h->stage[slot], bytes));
        h->stage_sz[slot] = bytes;
    }
    *out = h->stage[slot];
    return BS_OK;
}

static int tie_tick_for(int n_agents) {  // battle_env.py:316-319: total_time accumulates 0.1 in double
    volatile double t = 0;
    int k = 0;
    for (;;) { ++k; t = t + 0.1; if (t >= 10 + 2 * n_agents) return k; }
}

static thread_local bool g_pdl = false;  // set by bs_rollout_actor around its launch loop
static thread_local size_t g_dyn_smem = 0;  // unused dynamic shared memory of the next launch: caps the resident blocks per SM (rollout_range)

// Kernel launch with the observation tables marked persisting in L2 for this launch only (the
// caller's stream is not modified).
template <class... KArgs, class... Args>
static cudaError_t launch(const bs_handle *h, void (*kern)(KArgs...), int grid, int block, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)block); cfg.dynamicSmemBytes = g_dyn_smem; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    if (g_pdl) {  // programmatic dependent launch: this kernel may start while its predecessor drains (bs_rollout_actor)
        attr[cfg.numAttrs].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[cfg.numAttrs].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs += 1;
    }
    if (h->tables.persist) {
        cudaLaunchAttribute &a = attr[cfg.numAttrs];
        a.id = cudaLaunchAttributeAccessPolicyWindow;
        a.val.accessPolicyWindow.base_ptr = (void *)h->tables.tab;
        a.val.accessPolicyWindow.num_bytes = h->tables.bytes;
        a.val.accessPolicyWindow.hitRatio = 1.0f;
        a.val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        a.val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        cfg.attrs = attr; cfg.numAttrs += 1;
    }
    return cudaLaunchKernelEx(&cfg, kern, args...);
}

enum { FMT_POOLS = 0, FMT_LISTS = 1 };

static inline int grid_for_envs(int64_t E, int N, int block = BLOCK) {
    const int K = 32 / (2 * N);
    const int64_t warps = (E + K - 1) / K;
    return (int)((warps * 32 + block - 1) / block);
}

#define DISPATCH_N(n_, ...)                                                  \
    switch (n_) {                                                            \
        case 1: { constexpr int N_ = 1; __VA_ARGS__; } break;                \
        case 2: { constexpr int N_ = 2; __VA_ARGS__; } break;                \
        case 3: { constexpr int N_ = 3; __VA_ARGS__; } break;                \
        case 4: { constexpr int N_ = 4; __VA_ARGS__; } break;                \
        case 5: { constexpr int N_ = 5; __VA_ARGS__; } break;                \
        case 6: { constexpr int N_ = 6; __VA_ARGS__; } break;                \
        case 7: { constexpr int N_ = 7; __VA_ARGS__; } break;                \
        case 8: { constexpr int N_ = 8; __VA_ARGS__; } break;                \
        default: return fail(BS_E_INVALID, "n_agents out of range");         \
    }

// rollout2_kernel keeps both teams' observation rows of 32 envs in static shared memory: 29 KB at N = 5, over the
// 48 KB static limit from N = 7; team sizes above 5 (none of the BASELINE configs) stay on the lane-per-plane kernel.
#ifdef R2_SMEM_BULLETS   // (A/B build: the lists in shared memory instead of local memory; fits up to N = 3)
constexpr int R2_MAX_N = 3;
#define DISPATCH_R2(n_, ...)                                                 \
    switch (n_) {                                                            \
        case 1: { constexpr int N_ = 1; __VA_ARGS__; } break;                \
        case 2: { constexpr int N_ = 2; __VA_ARGS__; } break;                \
        case 3: { constexpr int N_ = 3; __VA_ARGS__; } break;                \
        default: return fail(BS_E_INVALID, "n_agents out of range");         \
    }
#else
constexpr int R2_MAX_N = 5;
#define DISPATCH_R2(n_, ...)                                                 \
    switch (n_) {                                                            \
        case 1: { constexpr int N_ = 1; __VA_ARGS__; } break;                \
        case 2: { constexpr int N_ = 2; __VA_ARGS__; } break;                \
        case 3: { constexpr int N_ = 3; __VA_ARGS__; } break;                \
        case 4: { constexpr int N_ = 4; __VA_ARGS__; } break;                \
        case 5: { constexpr int N_ = 5; __VA_ARGS__; } break;                \
        default: return fail(BS_E_INVALID, "n_agents out of range");         \
    }
#endif

// Converts the bullets to the layout the next kernel reads (a no-op when it is current); see rollout2.cuh.
static int ensure_bullets(bs_handle *h, int fmt, cudaStream_t st) {
    if (h->bullet_fmt == fmt) return BS_OK;
    if (fmt == FMT_LISTS) {
        const int grid = (int)((h->E + 31) / 32);
        DISPATCH_N(h->N, pools_to_lists_kernel<N_><<<grid, 64, 0, st>>>(h->p));
    } else {
        const int grid = (int)((h->W * 32 + 127) / 128);
        DISPATCH_N(h->N, lists_to_pools_kernel<N_><<<grid, 128, 0, st>>>(h->p));
    }
    CU(cudaGetLastError());
    h->bullet_fmt = fmt;
    return BS_OK;
}

// ---- actors (row f1) -------------------------------------------------------------------------------
static int launch_actor(bs_handle *h, const ActorArgs &aa, cudaStream_t st) {
    CU(bs_launch_actor_impl(h->N, h->cont, aa, st, g_pdl));
    return BS_OK;
}

static int actor_args(bs_handle *h, int team, const bs_actor *actors, int hidden, const float *obs, const bs_ou_noise *noise,
                      float *actions_out, ActorArgs *out) {
    if (!h || !actors || !obs || !actions_out) return fail(BS_E_INVALID, "null argument");
    if (team != 0 && team != 1) return fail(BS_E_INVALID, "team must be 0 (red) or 1 (blue)");
    if (hidden != AH) return fail(BS_E_UNSUPPORTED, "actor kernels are built for fc1_dims == fc2_dims == 64");
    ActorArgs aa{};
    for (int k = 0; k < h->N; ++k) {
        const bs_actor &a = actors[k];
        if (!a.fc1_w || !a.fc1_b || !a.ln1_g || !a.ln1_b || !a.fc2_w || !a.fc2_b || !a.ln2_g || !a.ln2_b || !a.pi_w || !a.pi_b)
            return fail(BS_E_INVALID, "null actor parameter");
        aa.net[k] = ActorNet{a.fc1_w, a.fc1_b, a.ln1_g, a.ln1_b, a.fc2_w, a.fc2_b, a.ln2_g, a.ln2_b, a.pi_w, a.pi_b};
    }
    aa.E = h->E; aa.env_offset = h->p.env_offset;
    aa.obs = obs; aa.obs_env_stride = (long long)h->P * h->L; aa.obs_agent0 = team * h->N; aa.n_team = h->N;
    aa.out = actions_out;
    if (noise) {
        if (!noise->state) return fail(BS_E_INVALID, "bs_ou_noise.state is null");
        aa.ou_state = noise->state; aa.normals = noise->normals; aa.reset_mask = noise->reset_mask;
        aa.mu = noise->mu; aa.theta = noise->theta; aa.sigma = noise->sigma; aa.scale = noise->scale;
        aa.seed_lo = (uint32_t)noise->seed; aa.seed_hi = (uint32_t)(noise->seed >> 32); aa.step = noise->step;
    }
    *out = aa;
    return BS_OK;
}

extern "C" {

int bs_version(void) { return BS_VERSION; }
const char *bs_last_error(void) { return g_err.c_str(); }

int bs_create(const bs_config *cfg, int64_t n_envs, int64_t env_offset, int device, uint64_t seed, bs_handle **out) {
    if (!cfg || !out) return fail(BS_E_INVALID, "null argument");
    if (cfg->n_agents < 1 || cfg->n_agents > BS_MAX_AGENTS) return fail(BS_E_INVALID, "n_agents must be in 1..8");
    if (n_envs < 1 || env_offset < 0) return fail(BS_E_INVALID, "n_envs must be >= 1 and env_offset >= 0");
    if (n_envs > (1ll << 30)) return fail(BS_E_INVALID, "n_envs must be <= 2^30 per handle (shard larger jobs by env_offset)");
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(BS_E_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(ce) + " (this library has no CPU path)");
    if (device < 0 || device >= ndev) return fail(BS_E_INVALID, "bad device index");
    DeviceGuard guard(device);
    bs_handle *h = new (std::nothrow) bs_handle();
    if (!h) return fail(BS_E_INVALID, "out of host memory");
    h->cfg = *cfg; h->N = cfg->n_agents; h->P = 2 * h->N; h->L = 3 * h->N + 2;
    h->cont = cfg->continuous_actions != 0; h->A = h->cont ? 3 : 4;
    h->E = n_envs; h->device = device;
    Params &p = h->p;
    p.E = n_envs; p.env_offset = env_offset;
    p.seed_lo = (uint32_t)seed; p.seed_hi = (uint32_t)(seed >> 32);
    philox_round_keys(p.seed_lo, p.seed_hi, p.rk);
    p.tie_tick = tie_tick_for(h->N);
    { const char *ex = getenv("BS_EXACT_TRIG"); p.exact_trig = ex && ex[0] == '1'; }
    p.r_hit_base = cfg->hit_base_reward; p.r_hit_plane = cfg->hit_plane_reward;
    p.r_miss = cfg->miss_punishment; p.r_die = cfg->die_punishment; p.r_lose = cfg->lose_punishment;
    p.r_code[0] = 0.0; p.r_code[1] = p.r_miss; p.r_code[2] = p.r_hit_base; p.r_code[3] = p.r_hit_plane;
    p.diag = std::sqrt(std::pow(1200.0, 2) + std::pow(800.0, 2));
    if (p.diag != DIAG) { delete h; return fail(BS_E_INVALID, "host sqrt(1200^2+800^2) differs from the device constant"); }
    const size_t EP = (size_t)n_envs * h->P;
    h->K = 32 / h->P; h->W = (n_envs + h->K - 1) / h->K;  // envs per warp, warps
    h->p.W = h->W;
    int rc;
    if ((rc = dalloc(h, &p.plane_pos, EP)) || (rc = dalloc(h, &p.plane_dir, EP)) || (rc = dalloc(h, &p.plane_hp, EP)) ||
        (rc = dalloc(h, &p.base_pos, (size_t)n_envs * 2)) || (rc = dalloc(h, &p.base_hp, (size_t)n_envs * 2)) ||
        (rc = dalloc(h, &p.pool, (size_t)h->W * POOL_CAP)) || (rc = dalloc(h, &p.pool_cnt, (size_t)h->W)) ||
        (rc = dalloc(h, &p.meta, (size_t)n_envs)) || (rc = dalloc(h, &p.episode, (size_t)n_envs)) ||
        (rc = dalloc(h, &p.ret, (size_t)n_envs * 2)) || (rc = dalloc(h, &p.counters, 9)) ||
        (rc = dalloc(h, &p.lists, (size_t)((n_envs + 31) / 32) * SLOTS * h->N * 64)) ||
        (rc = dalloc(h, &p.list_cnt, (size_t)((n_envs + 31) / 32) * 64))) {
        bs_destroy(h);
        return rc;
    }
    ObsTables tb;
    if ((rc = obs_tables_for(device, &tb))) { bs_destroy(h); return rc; }
    p.obs_tab = tb.tab; h->tables = tb;
    p.ep_log = nullptr; p.ep_cap = 0;
    // every env starts "done" (nothing spawned yet): the first reset / rollout tick spawns it
    std::vector<uint32_t> meta((size_t)n_envs, 1u << 16);
    cudaError_t ce2 = cudaMemcpy(p.meta, meta.data(), meta.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (ce2 == cudaSuccess) ce2 = cudaStreamCreateWithFlags(&h->hstream, cudaStreamNonBlocking);
    if (ce2 == cudaSuccess) ce2 = cudaStreamCreateWithFlags(&h->cstream, cudaStreamNonBlocking);
    for (auto &ev : h->chunk_ev) if (ce2 == cudaSuccess) ce2 = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    for (auto &ev : h->call_done) if (ce2 == cudaSuccess) ce2 = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    if (ce2 == cudaSuccess) ce2 = cudaEventCreateWithFlags(&h->copy_done, cudaEventDisableTiming);
    if (ce2 == cudaSuccess) ce2 = cudaMallocHost(&h->pinned_stats, 16 * sizeof(int64_t));
    if (ce2 != cudaSuccess) {   // (a failure this late must not leak the handle and its allocations)
        bs_destroy(h);
        return fail(BS_E_CUDA, std::string("bs_create: ") + cudaGetErrorString(ce2));
    }
    { const char *v2 = getenv("BS_ROLLOUT_V2"); h->rollout_v2 = v2 && v2[0] == '1'; }
    if (const char *hc = getenv("BS_HOST_CHUNKS")) h->host_chunks = atoi(hc) > 15 ? 15 : atoi(hc);
    *out = h;
    return BS_OK;
}

int bs_destroy(bs_handle *h) {
    if (!h) return BS_OK;
    DeviceGuard guard(h->device);
    cudaDeviceSynchronize();
    for (void *q : h->allocs) cudaFree(q);
    for (void *q : h->stage) if (q) cudaFree(q);
    if (h->hstream) cudaStreamDestroy(h->hstream);
    if (h->cstream) cudaStreamDestroy(h->cstream);
    for (auto ev : h->chunk_ev) if (ev) cudaEventDestroy(ev);
    for (auto ev : h->call_done) if (ev) cudaEventDestroy(ev);
    if (h->copy_done) cudaEventDestroy(h->copy_done);
    if (h->pinned_stats) cudaFreeHost(h->pinned_stats);
    delete h;
    return BS_OK;
}

int bs_reset(bs_handle *h, const uint8_t *env_mask, const int32_t *draws, float *obs, void *stream) {
    if (!h) return fail(BS_E_INVALID, "null handle");
    DeviceGuard guard(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    { int rc = ensure_bullets(h, FMT_POOLS, st); if (rc) return rc; }
    const int grid = grid_for_envs(h->E, h->N);
    DISPATCH_N(h->N, reset_kernel<N_><<<grid, BLOCK, 0, st>>>(h->p, env_mask, draws, obs));
    CU(cudaGetLastError());
    return BS_OK;
}

static int launch_step(bs_handle *h, const StepArgs &sa, cudaStream_t st) {
    { int rc = ensure_bullets(h, FMT_POOLS, st); if (rc) return rc; }
    const int grid = grid_for_envs(h->E, h->N, STEP_BLOCK);
    const bool lean = sa.pol[0] == POL_EXTERNAL && sa.pol[1] == POL_EXTERNAL && sa.rp.capacity == 0 && !sa.auto_reset &&
                      sa.obs_in == nullptr && sa.act_agents == h->P && getenv("BS_STEP_GENERAL") == nullptr;
    if (lean) {
        if (h->cont) { DISPATCH_N(h->N, CU(launch(h, step_kernel<N_, true, true>, grid, STEP_BLOCK, st, h->p, sa))); }
        else { DISPATCH_N(h->N, CU(launch(h, step_kernel<N_, false, true>, grid, STEP_BLOCK, st, h->p, sa))); }
        return BS_OK;
    }
    if (h->cont) { DISPATCH_N(h->N, CU(launch(h, step_kernel<N_, true, false>, grid, STEP_BLOCK, st, h->p, sa))); }
    else { DISPATCH_N(h->N, CU(launch(h, step_kernel<N_, false, false>, grid, STEP_BLOCK, st, h->p, sa))); }
    return BS_OK;
}

// The ring row of (tick t, env e) is (cursor + t*E + e) % capacity.  Two envs of one launch must never share
// a row (they are written by different warps): that needs capacity >= E, and, once a launch wraps the ring
// (T*E > capacity), capacity % E == 0 -- then only the SAME env comes back to a row, later ticks overwrite
// earlier ones in program order and the ring ends up holding the last capacity/E ticks of the launch.
static int replay_check(const bs_handle *h, const bs_replay *replay, int64_t T) {
    if (!replay || replay->capacity <= 0) return BS_OK;
    if (replay->capacity < h->E) return fail(BS_E_INVALID, "replay ring is smaller than n_envs: rows of different envs would collide");
    if (replay->capacity > (int64_t)0x7fffffff / h->N) return fail(BS_E_INVALID, "replay ring too large: capacity * n_agents must stay below 2^31");
    if (T * h->E > replay->capacity && replay->capacity % h->E != 0)
        return fail(BS_E_INVALID, "a launch that wraps the replay ring (T * n_envs > capacity) needs capacity % n_envs == 0");
    return BS_OK;
}

static int check_dtype(const bs_handle *h, int dt) {
    if (h->cont) { if (dt != BS_ACT_F32 && dt != BS_ACT_F64) return fail(BS_E_INVALID, "continuous env takes BS_ACT_F32 or BS_ACT_F64 actions"); }
    else if (dt != BS_ACT_I32 && dt != BS_ACT_F32) return fail(BS_E_INVALID, "discrete env takes BS_ACT_I32 or BS_ACT_F32 (logits) actions");
    return BS_OK;
}

int bs_step(bs_handle *h, const void *actions, int action_dtype, const double *draws, float *obs, float *rew,
            uint8_t *done, uint8_t *env_done, int8_t *winner, void *stream) {
    if (!h) return fail(BS_E_INVALID, "null handle");
    if (action_dtype != BS_ACT_NONE) {
        if (!actions) return fail(BS_E_INVALID, "null actions");
        int rc = check_dtype(h, action_dtype); if (rc) return rc;
    }
    DeviceGuard guard(h->device);
    StepArgs sa{actions, action_dtype, h->P, 0, {POL_EXTERNAL, POL_EXTERNAL}, draws, obs, rew, done, env_done, winner, {}, 0, 0};
    return launch_step(h, sa, (cudaStream_t)stream);
}

int bs_step_team_replay(bs_handle *h, int learner_team, const void *team_actions, int action_dtype, int other_policy,
                        const double *draws, const bs_replay *replay, int64_t cursor, int auto_reset, float *obs,
                        float *rew, uint8_t *done, uint8_t *env_done, int8_t *winner, void *stream) {
    if (!h || !team_actions) return fail(BS_E_INVALID, "null handle or actions");
    if (learner_team != 0 && learner_team != 1) return fail(BS_E_INVALID, "learner_team must be 0 (red) or 1 (blue)");
    if (other_policy < BS_POL_RANDOM || other_policy > BS_POL_INSTINCT_FF) return fail(BS_E_INVALID, "other_policy must be an in-kernel policy");
    if (cursor < 0) return fail(BS_E_INVALID, "cursor must be >= 0");
    int rc = check_dtype(h, action_dtype); if (rc) return rc;
    if ((rc = replay_check(h, replay, 1))) return rc;
    DeviceGuard guard(h->device);
    StepArgs sa{team_actions, action_dtype, h->N, learner_team, {POL_EXTERNAL, POL_EXTERNAL}, draws, obs, rew, done, env_done, winner, {}, 0, auto_reset ? 1 : 0};
    sa.pol[1 - learner_team] = other_policy;
    if (replay && replay->capacity > 0) { sa.rp = *replay; sa.cursor_mod = cursor % replay->capacity; }
    return launch_step(h, sa, (cudaStream_t)stream);
}

int bs_step_team(bs_handle *h, int learner_team, const void *team_actions, int action_dtype, int other_policy,
                 const double *draws, float *obs, float *rew, uint8_t *done, uint8_t *env_done, int8_t *winner,
                 void *stream) {
    return bs_step_team_replay(h, learner_team, team_actions, action_dtype, other_policy, draws, nullptr, 0, 0, obs, rew, done,
                               env_done, winner, stream);
}

int bs_instinct_actions(bs_handle *h, int team, const float *obs, const double *draws, void *actions_out, void *stream) {
    if (!h || !obs || !actions_out) return fail(BS_E_INVALID, "null argument");
    if (team != 0 && team != 1) return fail(BS_E_INVALID, "team must be 0 (red) or 1 (blue)");
    DeviceGuard guard(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = h->E * h->N;
    const int grid = (int)((n + BLOCK - 1) / BLOCK);
    if (h->cont) { DISPATCH_N(h->N, instinct_kernel<N_, true><<<grid, BLOCK, 0, st>>>(h->p, team, obs, draws, actions_out)); }
    else { DISPATCH_N(h->N, instinct_kernel<N_, false><<<grid, BLOCK, 0, st>>>(h->p, team, obs, draws, actions_out)); }
    CU(cudaGetLastError());
    return BS_OK;
}

static int rollout_check(bs_handle *h, int T, int learner_team, int policy_red, int policy_blue, const void *ext_actions,
                         const bs_replay *replay, int64_t cursor) {
    if (!h) return fail(BS_E_INVALID, "null handle");
    if (T < 1) return fail(BS_E_INVALID, "T must be >= 1");
    { int rc = replay_check(h, replay, T); if (rc) return rc; }
    if (learner_team != 0 && learner_team != 1) return fail(BS_E_INVALID, "learner_team must be 0 (red) or 1 (blue)");
    const int pol[2] = {policy_red, policy_blue};
    for (int t = 0; t < 2; ++t) {
        if (pol[t] < BS_POL_EXTERNAL || pol[t] > BS_POL_INSTINCT_FF) return fail(BS_E_INVALID, "unknown policy");
        if (pol[t] == BS_POL_EXTERNAL && (t != learner_team || !ext_actions))
            return fail(BS_E_INVALID, "BS_POL_EXTERNAL needs ext_actions and is only valid for learner_team");
    }
    if (cursor < 0) return fail(BS_E_INVALID, "cursor must be >= 0");
    return BS_OK;
}

// launches the fused rollout for envs [e0, e_end)
static int rollout_range(bs_handle *h, int T, int learner_team, int policy_red, int policy_blue, const float *ext_actions,
                         const bs_replay *replay, int64_t cursor, int64_t e0, int64_t e_end, cudaStream_t st) {
    RolloutArgs ra{};
    ra.T = T; ra.learner = learner_team; ra.pol[0] = policy_red; ra.pol[1] = policy_blue; ra.ext = ext_actions;
    ra.e0 = e0; ra.e_end = e_end;
    if (replay && replay->capacity > 0) {
        ra.rp = *replay;
        ra.cursor_mod = cursor % replay->capacity;
        ra.e_mod = h->E % replay->capacity;
    } else {
        std::memset(&ra.rp, 0, sizeof(ra.rp));
    }
    if (h->rollout_v2 && h->N <= R2_MAX_N) {   // team-per-thread mapping (rollout2.cuh): blocks of 32 envs, chunks start on a block boundary
        if (e0 % 32 != 0) return fail(BS_E_INVALID, "rollout chunk does not start on a 32-env boundary");
        int rc = ensure_bullets(h, FMT_LISTS, st); if (rc) return rc;
        const int grid = (int)((e_end - e0 + 31) / 32);
        if (h->cont) { DISPATCH_R2(h->N, CU(launch(h, rollout2_kernel<N_, true>, grid, 64, st, h->p, ra))); }
        else { DISPATCH_R2(h->N, CU(launch(h, rollout2_kernel<N_, false>, grid, 64, st, h->p, ra))); }
        return BS_OK;
    }
    { int rc = ensure_bullets(h, FMT_POOLS, st); if (rc) return rc; }
    const int grid = grid_for_envs(e_end - e0, h->N, ROLLOUT_BLOCK);
    struct DynScope { DynScope() { const char *v = getenv("BS_ROLLOUT_DYNSMEM"); g_dyn_smem = v ? (size_t)atoi(v) : 0; } ~DynScope() { g_dyn_smem = 0; } } dyn_scope;
    if (h->cont) { DISPATCH_N(h->N, CU(launch(h, rollout_kernel<N_, true>, grid, ROLLOUT_BLOCK, st, h->p, ra))); }
    else { DISPATCH_N(h->N, CU(launch(h, rollout_kernel<N_, false>, grid, ROLLOUT_BLOCK, st, h->p, ra))); }
    return BS_OK;
}

int bs_rollout_fused(bs_handle *h, int T, int learner_team, int policy_red, int policy_blue, const float *ext_actions,
                     const bs_replay *replay, int64_t cursor, void *stream) {
    int rc = rollout_check(h, T, learner_team, policy_red, policy_blue, ext_actions, replay, cursor); if (rc) return rc;
    DeviceGuard guard(h->device);
    return rollout_range(h, T, learner_team, policy_red, policy_blue, ext_actions, replay, cursor, 0, h->E, (cudaStream_t)stream);
}

int bs_actor_forward(bs_handle *h, int team, const bs_actor *actors, int hidden, const float *obs, const bs_ou_noise *noise,
                     float *actions_out, void *stream) {
    ActorArgs aa;
    int rc = actor_args(h, team, actors, hidden, obs, noise, actions_out, &aa); if (rc) return rc;
    DeviceGuard guard(h->device);
    return launch_actor(h, aa, (cudaStream_t)stream);
}

int bs_rollout_actor(bs_handle *h, int T, int learner_team, const bs_actor *actors, int hidden, int other_policy,
                     const bs_ou_noise *noise, const bs_replay *replay, int64_t cursor, float *obs, float *act_scratch,
                     float *rew, uint8_t *done, uint8_t *env_done, int8_t *winner, void *stream) {
    if (T < 1) return fail(BS_E_INVALID, "T must be >= 1");
    if (!env_done) return fail(BS_E_INVALID, "env_done work buffer is required");
    if (cursor < 0) return fail(BS_E_INVALID, "cursor must be >= 0");
    ActorArgs aa;
    int rc = actor_args(h, learner_team, actors, hidden, obs, noise, act_scratch, &aa); if (rc) return rc;
    if (other_policy < BS_POL_RANDOM || other_policy > BS_POL_INSTINCT_FF) return fail(BS_E_INVALID, "other_policy must be an in-kernel policy");
    if ((rc = replay_check(h, replay, T))) return rc;
    DeviceGuard guard(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    StepArgs sa{act_scratch, BS_ACT_F32, h->N, learner_team, {POL_EXTERNAL, POL_EXTERNAL}, nullptr, obs, rew, done, env_done, winner, {}, 0, 1};
    sa.obs_in = obs;  // the buffer holds every env's current observation: no need to rebuild it before the tick
    sa.pol[1 - learner_team] = other_policy;
    if (replay && replay->capacity > 0) sa.rp = *replay;
    // The two kernels of a tick are launched as programmatic dependents of each other: the actor kernel
    // stages its parameters while the step kernel drains, and each grid is scheduled as the other's blocks
    // exit instead of after a full launch gap (griddepcontrol.wait guards every dependent access).
    struct PdlScope { PdlScope() { g_pdl = getenv("BS_NO_PDL") == nullptr; } ~PdlScope() { g_pdl = false; } } pdl_scope;
    for (int t = 0; t < T; ++t) {
        if ((rc = launch_actor(h, aa, st))) return rc;
        if (sa.rp.capacity > 0) sa.cursor_mod = (cursor + (int64_t)t * h->E) % sa.rp.capacity;
        if ((rc = launch_step(h, sa, st))) return rc;
        aa.step += 1;
        if (noise) aa.reset_mask = env_done;  // envs that just finished start the next game with fresh noise
    }
    return BS_OK;
}

// bs_rollout_actor without caller work buffers and without any allocation: the action rows that travel from the
// actor kernel to the step kernel (and env_done, when the caller does not want it) live in g_static_scratch.
// Same kernels, same order as bs_rollout_actor, so the results are bit-identical (tests/test_zz_no_alloc.py).
int bs_rollout_actor_static(bs_handle *h, int T, int learner_team, const bs_actor *actors, int hidden, int other_policy,
                            const bs_ou_noise *noise, const bs_replay *replay, int64_t cursor, float *obs, float *rew,
                            uint8_t *done, uint8_t *env_done, int8_t *winner, void *stream) {
    if (!h) return fail(BS_E_INVALID, "null handle");
    if (T < 1) return fail(BS_E_INVALID, "T must be >= 1");
    if (cursor < 0) return fail(BS_E_INVALID, "cursor must be >= 0");
    const size_t act_bytes = (size_t)h->E * h->N * h->A * sizeof(float);
    const size_t done_off = (act_bytes + 255) & ~(size_t)255;
    const size_t need = done_off + (env_done ? 0 : (size_t)h->E);
    if (need > (size_t)BS_STATIC_SCRATCH_BYTES)
        return fail(BS_E_INVALID, "bs_rollout_actor_static: n_envs*N*A*4 (+ n_envs) = " + std::to_string(need) +
                                  " bytes of scratch exceed BS_STATIC_SCRATCH_BYTES = " + std::to_string((size_t)BS_STATIC_SCRATCH_BYTES) +
                                  "; use bs_rollout_actor with caller buffers or a smaller handle");
    DeviceGuard guard(h->device);
    unsigned char *scratch = nullptr;
    int rc = static_scratch_for(h->device, &scratch); if (rc) return rc;
    float *act_scratch = reinterpret_cast<float *>(scratch);
    if (!env_done) env_done = scratch + done_off;
    ActorArgs aa;
    if ((rc = actor_args(h, learner_team, actors, hidden, obs, noise, act_scratch, &aa))) return rc;
    if (other_policy < BS_POL_RANDOM || other_policy > BS_POL_INSTINCT_FF) return fail(BS_E_INVALID, "other_policy must be an in-kernel policy");
    if ((rc = replay_check(h, replay, T))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    StepArgs sa{act_scratch, BS_ACT_F32, h->N, learner_team, {POL_EXTERNAL, POL_EXTERNAL}, nullptr, obs, rew, done, env_done, winner, {}, 0, 1};
    sa.obs_in = obs;
    sa.pol[1 - learner_team] = other_policy;
    if (replay && replay->capacity > 0) sa.rp = *replay;
    struct PdlScope { PdlScope() { g_pdl = getenv("BS_NO_PDL") == nullptr; } ~PdlScope() { g_pdl = false; } } pdl_scope;
    for (int t = 0; t < T; ++t) {
        if ((rc = launch_actor(h, aa, st))) return rc;
        if (sa.rp.capacity > 0) sa.cursor_mod = (cursor + (int64_t)t * h->E) % sa.rp.capacity;
        if ((rc = launch_step(h, sa, st))) return rc;
        aa.step += 1;
        if (noise) aa.reset_mask = env_done;
    }
    return BS_OK;
}

int bs_replay_sample(const bs_replay *replay, int n_agents, int obs_len, int n_actions, const int64_t *idx, int64_t batch,
                     float *actor_states, float *states, float *actions, float *rewards, float *actor_new_states,
                     float *new_states, uint8_t *dones, void *stream) {
    if (!replay || !idx) return fail(BS_E_INVALID, "null argument");
    if (batch < 1 || replay->capacity < 1) return fail(BS_E_INVALID, "batch and capacity must be >= 1");
    if (n_agents < 1 || n_agents > BS_MAX_AGENTS || obs_len < 1 || n_actions < 1) return fail(BS_E_INVALID, "bad row shape");
    if (!replay->state || !replay->new_state || !replay->action || !replay->reward || !replay->done)
        return fail(BS_E_INVALID, "replay ring has a null array");
    cudaPointerAttributes pa;
    CU(cudaPointerGetAttributes(&pa, idx));
    if (pa.type != cudaMemoryTypeDevice && pa.type != cudaMemoryTypeManaged) return fail(BS_E_INVALID, "idx must be a device pointer");
    DeviceGuard guard(pa.device);
    SampleArgs sa{*replay, (const long long *)idx, batch, n_agents, obs_len, n_actions, actor_states, states, actions, rewards,
                  actor_new_states, new_states, dones};
    const int grid = (int)(batch < 148 * 16 ? batch : 148 * 16);
    replay_sample_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(sa);
    CU(cudaGetLastError());
    return BS_OK;
}

int bs_render(bs_handle *h, int64_t env, const bs_sprites *sprites, uint8_t *rgb, void *stream) {
    if (!h || !sprites || !rgb) return fail(BS_E_INVALID, "null argument");
    if (env < 0 || env >= h->E) return fail(BS_E_INVALID, "env index out of range");
    if (!sprites->plane[0] || !sprites->plane[1] || !sprites->base[0] || !sprites->base[1]) return fail(BS_E_INVALID, "null sprite");
    DeviceGuard guard(h->device);
    if (sprites->n_explosions < 0) return fail(BS_E_INVALID, "n_explosions must be >= 0");
    RenderArgs ra{};
    ra.env = env; ra.plane_img[0] = sprites->plane[0]; ra.plane_img[1] = sprites->plane[1];
    ra.base_img[0] = sprites->base[0]; ra.base_img[1] = sprites->base[1]; ra.rgb = rgb;
    ra.expl_img = sprites->explosion_img; ra.expl = sprites->explosions; ra.n_expl = sprites->n_explosions;
    {   // pygame.transform.rotate parameters of the env's planes, with the host libm like pygame (render.cuh PlaneRot)
        uint32_t pos[16]; double dir[16];
        cudaStream_t st = (cudaStream_t)stream;
        CU(cudaMemcpyAsync(pos, h->p.plane_pos + env * h->P, (size_t)h->P * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(dir, h->p.plane_dir + env * h->P, (size_t)h->P * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        for (int i = 0; i < h->P; ++i) {
            PlaneRot &r = ra.rot[i];
            const int w = PLANE_W, hgt = PLANE_H, cx = (int)(pos[i] & 0xffff), cy = (int)(pos[i] >> 16);
            const double angle = dir[i];
            if (std::fmod(angle, 90.0) == 0.0) {           // surf_rotate: if (!(fmod(angle, 90))) -> rotate90
                int turns = ((int)(angle / 90.0)) % 4;
                if (turns < 0) turns += 4;
                r.turns = turns; r.nx = (turns % 2) ? hgt : w; r.ny = (turns % 2) ? w : hgt;
                r.ax = r.ay = r.isin = r.icos = 0;
            } else {
                volatile double radangle = angle * .01745329251994329;
                volatile double sangle = std::sin(radangle), cangle = std::cos(radangle);
                const double x = w, y = hgt, cxr = cangle * x, cyr = cangle * y, sxr = sangle * x, syr = sangle * y;
                r.nx = (int)std::fmax(std::fmax(std::fmax(std::fabs(cxr + syr), std::fabs(cxr - syr)), std::fabs(-cxr + syr)), std::fabs(-cxr - syr));
                r.ny = (int)std::fmax(std::fmax(std::fmax(std::fabs(sxr + cyr), std::fabs(sxr - cyr)), std::fabs(-sxr + cyr)), std::fabs(-sxr - cyr));
                r.isin = (int)(sangle * 65536); r.icos = (int)(cangle * 65536);
                r.ax = ((r.nx) << 15) - (int)(cangle * ((r.nx - 1) << 15)) + ((w - r.nx) * 32768);
                r.ay = ((r.ny) << 15) - (int)(sangle * ((r.nx - 1) << 15)) + ((hgt - r.ny) * 32768);
                r.turns = -1;
            }
            r.left = cx - (r.nx >> 1); r.top = cy - (r.ny >> 1);   // Rect.center setter: x = cx - (w >> 1)
        }
    }
    const int grid = (R_W * R_H + 255) / 256;
    DISPATCH_N(h->N, render_kernel<N_><<<grid, 256, 0, (cudaStream_t)stream>>>(h->p, ra));
    CU(cudaGetLastError());
    return BS_OK;
}

int bs_episode_log(bs_handle *h, float *log, int64_t capacity) {
    if (!h) return fail(BS_E_INVALID, "null handle");
    if (log && (capacity < 1 || (capacity & (capacity - 1)) != 0)) return fail(BS_E_INVALID, "capacity must be a power of two");
    h->p.ep_log = log; h->p.ep_cap = log ? capacity : 0;
    return BS_OK;
}

int bs_episode_log_count(bs_handle *h, int64_t *count, void *stream) {
    if (!h || !count) return fail(BS_E_INVALID, "null argument");
    DeviceGuard guard(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaMemcpyAsync(count, h->p.counters + 8, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return BS_OK;
}

int bs_stats(bs_handle *h, int64_t out[8], void *stream) {
    if (!h || !out) return fail(BS_E_INVALID, "null argument");
    DeviceGuard guard(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaMemcpyAsync(out, h->p.counters, 8 * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return BS_OK;
}

int bs_stats_device(bs_handle *h, int64_t **dev_counters) {
    if (!h || !dev_counters) return fail(BS_E_INVALID, "null argument");
    *dev_counters = (int64_t *)h->p.counters;
    return BS_OK;
}

// ---- state export / import (host views; checkpoint + parity injection) ---------------------------
int bs_get_state(bs_handle *h, const bs_state_view *v) {
    if (!h || !v) return fail(BS_E_INVALID, "null argument");
    DeviceGuard guard(h->device);
    CU(cudaDeviceSynchronize());
    { int rc = ensure_bullets(h, FMT_POOLS, nullptr); if (rc) return rc; }
    CU(cudaDeviceSynchronize());
    const int64_t E = h->E; const int P = h->P; const size_t EP = (size_t)E * P;
    std::vector<uint32_t> pos(EP), bpos((size_t)E * 2), meta((size_t)E), epi((size_t)E);
    std::vector<int> pcnt((size_t)h->W);
    std::vector<double> dir(EP), ret((size_t)E * 2);
    std::vector<int8_t> hp(EP);
    std::vector<int16_t> bhp((size_t)E * 2);
    const Params &p = h->p;
    // live pool entries only: counts -> offsets (host scan) -> dense device buffer -> host
    CU(cudaMemcpy(pcnt.data(), p.pool_cnt, pcnt.size() * sizeof(int), cudaMemcpyDeviceToHost));
    std::vector<long long> poff((size_t)h->W + 1, 0);
    for (int64_t wv = 0; wv < h->W; ++wv) poff[wv + 1] = poff[wv] + (pcnt[wv] < 0 ? 0 : (pcnt[wv] > POOL_CAP ? POOL_CAP : pcnt[wv]));
    std::vector<uint2> pool((size_t)poff[h->W] + 1);
    {
        void *d_off = nullptr, *d_dense = nullptr;
        int rc = stage_buf(h, 10, poff.size() * sizeof(long long), &d_off); if (rc) return rc;
        rc = stage_buf(h, 11, pool.size() * sizeof(uint2), &d_dense); if (rc) return rc;
        CU(cudaMemcpy(d_off, poff.data(), poff.size() * sizeof(long long), cudaMemcpyHostToDevice));
        pool_pack_kernel<<<(unsigned)((h->W * 32 + 127) / 128), 128>>>(p.pool, p.pool_cnt, (const long long *)d_off, h->W, (uint2 *)d_dense, 0);
        CU(cudaGetLastError());
        CU(cudaMemcpy(pool.data(), d_dense, (size_t)poff[h->W] * sizeof(uint2), cudaMemcpyDeviceToHost));
    }
    CU(cudaMemcpy(pos.data(), p.plane_pos, EP * 4, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(dir.data(), p.plane_dir, EP * 8, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(hp.data(), p.plane_hp, EP, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(bpos.data(), p.base_pos, (size_t)E * 8, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(bhp.data(), p.base_hp, (size_t)E * 4, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(meta.data(), p.meta, (size_t)E * 4, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(epi.data(), p.episode, (size_t)E * 4, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(ret.data(), p.ret, (size_t)E * 16, cudaMemcpyDeviceToHost));
    for (size_t a = 0; a < EP; ++a) {
        if (v->plane_x) v->plane_x[a] = (int16_t)(pos[a] & 0xffff);
        if (v->plane_y) v->plane_y[a] = (int16_t)(pos[a] >> 16);
        if (v->plane_dir) v->plane_dir[a] = dir[a];
        if (v->plane_hp) v->plane_hp[a] = hp[a];
        for (int o = 0; o < SLOTS; ++o) {
            if (v->bullet_pos) { v->bullet_pos[(a * SLOTS + o) * 2] = -1; v->bullet_pos[(a * SLOTS + o) * 2 + 1] = -1; }
            if (v->bullet_vel) { int8_t *vel = v->bullet_vel + (a * SLOTS + o) * 3; vel[0] = vel[1] = vel[2] = 0; }
        }
    }
    // pool entry -> (env, plane, age): owner lane = sub * 2N + plane, env = warp * K + sub; exported index = age - 1
    for (int64_t wv = 0; wv < h->W; ++wv)
        for (int q = 0; q < (int)(poff[wv + 1] - poff[wv]); ++q) {
            const uint2 en = pool[(size_t)poff[wv] + q];
            const uint32_t ps = en.x, M = en.y;
            if (!(ps & B_VALID)) continue;
            const int owner = (M >> 23) & 31, age = (int)(M >> 28), sub = owner / P, i = owner - sub * P;
            const int64_t e = wv * h->K + sub;
            if (e >= E || age < 1 || age > SLOTS) continue;
            const size_t a = (size_t)e * P + i;
            const int o = age - 1;
            const int32_t dl = meta_dlt(M);
            if (v->bullet_pos) {
                v->bullet_pos[(a * SLOTS + o) * 2] = (int16_t)((int)(ps & 0xffff) - 64);
                v->bullet_pos[(a * SLOTS + o) * 2 + 1] = (int16_t)((int)((ps >> 16) & 0x1fff) - 64);
            }
            if (v->bullet_vel) {
                const int sx = (int16_t)(dl & 0xffff), sy = (dl - sx) >> 16;
                v->bullet_vel[(a * SLOTS + o) * 3] = (int8_t)sx;
                v->bullet_vel[(a * SLOTS + o) * 3 + 1] = (int8_t)sy;
                v->bullet_vel[(a * SLOTS + o) * 3 + 2] = (int8_t)(1 | ((ps & B_XEXACT) ? 2 : 0) | ((ps & B_YEXACT) ? 4 : 0));
            }
        }
    for (int64_t e = 0; e < E; ++e) {
        for (int t = 0; t < 2; ++t) {
            if (v->base_pos) { v->base_pos[(e * 2 + t) * 2] = (int16_t)(bpos[e * 2 + t] & 0xffff); v->base_pos[(e * 2 + t) * 2 + 1] = (int16_t)(bpos[e * 2 + t] >> 16); }
            if (v->base_hp) v->base_hp[e * 2 + t] = bhp[e * 2 + t];
            if (v->team_return) v->team_return[e * 2 + t] = ret[e * 2 + t];
        }
        if (v->tick) v->tick[e] = (int32_t)(meta[e] & 0xfff);
        if (v->env_done) v->env_done[e] = (uint8_t)((meta[e] >> 16) & 0xff);
        if (v->winner) v->winner[e] = (int8_t)((meta[e] >> 24) & 0xff);
        if (v->episode) v->episode[e] = epi[e];
    }
    return BS_OK;
}

int bs_set_state(bs_handle *h, const bs_state_view *v) {
    if (!h || !v) return fail(BS_E_INVALID, "null argument");
    if (!v->plane_x || !v->plane_y || !v->plane_dir || !v->plane_hp || !v->base_pos || !v->base_hp || !v->bullet_pos ||
        !v->bullet_vel || !v->tick || !v->env_done || !v->winner || !v->episode || !v->team_return)
        return fail(BS_E_INVALID, "bs_set_state needs every field of the view");
    DeviceGuard guard(h->device);
    CU(cudaDeviceSynchronize());
    const int64_t E = h->E; const int P = h->P; const size_t EP = (size_t)E * P;
    std::vector<uint32_t> pos(EP), bpos((size_t)E * 2), meta((size_t)E);
    std::vector<uint2> pool;          // dense: warp after warp
    std::vector<long long> poff((size_t)h->W + 1, 0);
    std::vector<int> pcnt((size_t)h->W, 0);
    // the observation tables (and the packed 11/10-bit plane words) cover on-field points only
    for (size_t a = 0; a < EP; ++a) {
        if (v->plane_x[a] < 0 || v->plane_x[a] > 1200 || v->plane_y[a] < 0 || v->plane_y[a] > 800)
            return fail(BS_E_INVALID, "bs_set_state: plane position outside the 1200x800 field");
        if (!(v->plane_dir[a] >= 0.0 && v->plane_dir[a] <= 360.0))   // Plane.rotate keeps it there, sprites.py:100-103
            return fail(BS_E_INVALID, "bs_set_state: plane direction outside [0, 360]");
    }
    for (size_t q = 0; q < (size_t)E * 2; ++q)
        if (v->base_pos[q * 2] < 0 || v->base_pos[q * 2] > 1200 || v->base_pos[q * 2 + 1] < 0 || v->base_pos[q * 2 + 1] > 800)
            return fail(BS_E_INVALID, "bs_set_state: base position outside the 1200x800 field");
    for (size_t a = 0; a < EP; ++a) pos[a] = (uint32_t)(uint16_t)v->plane_x[a] | ((uint32_t)(uint16_t)v->plane_y[a] << 16);
    // bullets -> the warp pools, oldest first (the order inside a pool carries no meaning)
    for (int64_t wv = 0; wv < h->W; poff[wv + 1] = (long long)pool.size(), ++wv)
        for (int o = SLOTS - 2; o >= 0; --o)          // age 12 never survives a tick
            for (int owner = 0; owner < h->K * P; ++owner) {
                const int sub = owner / P, i = owner - sub * P;
                const int64_t e = wv * h->K + sub;
                if (e >= E) continue;
                const size_t a = (size_t)e * P + i;
                const int8_t *vel = v->bullet_vel + (a * SLOTS + o) * 3;
                const int16_t *ps = v->bullet_pos + (a * SLOTS + o) * 2;
                if (!(vel[2] & 1)) continue;
                const uint32_t Pw = (((uint32_t)(ps[0] + 64)) | ((uint32_t)(ps[1] + 64) << 16)) | B_VALID |
                                    ((vel[2] & 2) ? B_XEXACT : 0u) | ((vel[2] & 4) ? B_YEXACT : 0u);
                const int32_t dl = (int32_t)vel[0] + (int32_t)vel[1] * 65536;
                const uint32_t M = ((uint32_t)dl & 0x7fffffu) | ((uint32_t)owner << 23) | ((uint32_t)(o + 1) << 28);
                pool.push_back(make_uint2(Pw, M)); pcnt[wv] += 1;
            }
    for (int64_t e = 0; e < E; ++e) {
        for (int t = 0; t < 2; ++t)
            bpos[e * 2 + t] = (uint32_t)(uint16_t)v->base_pos[(e * 2 + t) * 2] | ((uint32_t)(uint16_t)v->base_pos[(e * 2 + t) * 2 + 1] << 16);
        meta[e] = ((uint32_t)v->tick[e] & 0xfff) | ((uint32_t)(v->env_done[e] ? 1 : 0) << 16) | ((uint32_t)(uint8_t)v->winner[e] << 24);
    }
    const Params &p = h->p;
    CU(cudaMemcpy(p.plane_pos, pos.data(), EP * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p.plane_dir, v->plane_dir, EP * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p.plane_hp, v->plane_hp, EP, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p.base_pos, bpos.data(), (size_t)E * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p.base_hp, v->base_hp, (size_t)E * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p.pool_cnt, pcnt.data(), pcnt.size() * sizeof(int), cudaMemcpyHostToDevice));
    {
        void *d_off = nullptr, *d_dense = nullptr;
        int rc = stage_buf(h, 10, poff.size() * sizeof(long long), &d_off); if (rc) return rc;
        rc = stage_buf(h, 11, (pool.size() + 1) * sizeof(uint2), &d_dense); if (rc) return rc;
        CU(cudaMemcpy(d_off, poff.data(), poff.size() * sizeof(long long), cudaMemcpyHostToDevice));
        if (!pool.empty()) CU(cudaMemcpy(d_dense, pool.data(), pool.size() * sizeof(uint2), cudaMemcpyHostToDevice));
        pool_pack_kernel<<<(unsigned)((h->W * 32 + 127) / 128), 128>>>(p.pool, p.pool_cnt, (const long long *)d_off, h->W, (uint2 *)d_dense, 1);
        CU(cudaGetLastError());
        CU(cudaDeviceSynchronize());
    }
    h->bullet_fmt = FMT_POOLS;
    CU(cudaMemcpy(p.meta, meta.data(), (size_t)E * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p.episode, v->episode, (size_t)E * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p.ret, v->team_return, (size_t)E * 16, cudaMemcpyHostToDevice));
    return BS_OK;
}

// ---- host-buffer entry points ---------------------------------------------------------------------
static size_t action_bytes(const bs_handle *h, int dtype, int agents) {
    const size_t n = (size_t)h->E * agents;
    if (dtype == BS_ACT_I32) return n * 4;
    if (dtype == BS_ACT_F32) return n * h->A * 4;
    return n * 3 * 8;
}

#define H2D(slot, host, bytes, devptr)                                                     \
    do {                                                                                   \
        void *d_ = nullptr;                                                                \
        if (host) {                                                                        \
            int rc_ = stage_buf(h, slot, bytes, &d_); if (rc_) return rc_;                 \
            CU(cudaMemcpyAsync(d_, host, bytes, cudaMemcpyHostToDevice, h->hstream));      \
        }                                                                                  \
        devptr = decltype(devptr)(d_);                                                     \
    } while (0)
#define DOUT(slot, host, bytes, devptr)                                                    \
    do {                                                                                   \
        void *d_ = nullptr;                                                                \
        if (host) { int rc_ = stage_buf(h, slot, bytes, &d_); if (rc_) return rc_; }       \
        devptr = decltype(devptr)(d_);                                                     \
    } while (0)
#define D2H(host, devptr, bytes)                                                           \
    do { if (host) CU(cudaMemcpyAsync(host, devptr, bytes, cudaMemcpyDeviceToHost, h->hstream)); } while (0)

int bs_reset_host(bs_handle *h, const uint8_t *env_mask, const int32_t *draws, float *obs) {
    if (!h) return fail(BS_E_INVALID, "null handle");
    DeviceGuard guard(h->device);
    const uint8_t *d_mask; const int32_t *d_draws; float *d_obs;
    H2D(0, env_mask, (size_t)h->E, d_mask);
    H2D(1, draws, (size_t)h->E * (4 + 6 * h->N) * 4, d_draws);
    DOUT(2, obs, (size_t)h->E * h->P * h->L * 4, d_obs);
    int rc = bs_reset(h, d_mask, d_draws, d_obs, h->hstream); if (rc) return rc;
    D2H(obs, d_obs, (size_t)h->E * h->P * h->L * 4);
    CU(cudaStreamSynchronize(h->hstream));
    return BS_OK;
}

static int step_host_common(bs_handle *h, bool team_mode, int learner_team, const void *actions, int action_dtype,
                            int other_policy, const double *draws, float *obs, float *rew, uint8_t *done,
                            uint8_t *env_done, int8_t *winner) {
    if (!h) return fail(BS_E_INVALID, "null handle");
    int rc;
    if (action_dtype == BS_ACT_NONE && !team_mode) actions = nullptr;
    else {
        if (!actions) return fail(BS_E_INVALID, "null actions");
        rc = check_dtype(h, action_dtype); if (rc) return rc;
    }
    DeviceGuard guard(h->device);
    const size_t EP = (size_t)h->E * h->P;
    const void *d_act; const double *d_draws; float *d_obs, *d_rew; uint8_t *d_done, *d_edone; int8_t *d_win;
    H2D(0, actions, action_bytes(h, action_dtype, team_mode ? h->N : h->P), d_act);
    H2D(1, draws, EP * 5 * 8, d_draws);
    DOUT(2, obs, EP * h->L * 4, d_obs);
    DOUT(3, rew, EP * 4, d_rew);
    DOUT(4, done, EP, d_done);
    DOUT(5, env_done, (size_t)h->E, d_edone);
    DOUT(6, winner, (size_t)h->E, d_win);
    rc = team_mode ? bs_step_team(h, learner_team, d_act, action_dtype, other_policy, d_draws, d_obs, d_rew, d_done, d_edone, d_win, h->hstream)
                   : bs_step(h, d_act, action_dtype, d_draws, d_obs, d_rew, d_done, d_edone, d_win, h->hstream);
    if (rc) return rc;
    D2H(obs, d_obs, EP * h->L * 4);
    D2H(rew, d_rew, EP * 4);
    D2H(done, d_done, EP);
    D2H(env_done, d_edone, (size_t)h->E);
    D2H(winner, d_win, (size_t)h->E);
    CU(cudaStreamSynchronize(h->hstream));
    return BS_OK;
}

int bs_step_host(bs_handle *h, const void *actions, int action_dtype, const double *draws, float *obs, float *rew,
                 uint8_t *done, uint8_t *env_done, int8_t *winner) {
    return step_host_common(h, false, 0, actions, action_dtype, 0, draws, obs, rew, done, env_done, winner);
}

int bs_step_team_host(bs_handle *h, int learner_team, const void *team_actions, int action_dtype, int other_policy,
                      const double *draws, float *obs, float *rew, uint8_t *done, uint8_t *env_done, int8_t *winner) {
    return step_host_common(h, true, learner_team, team_actions, action_dtype, other_policy, draws, obs, rew, done, env_done, winner);
}

int bs_instinct_actions_host(bs_handle *h, int team, const float *obs, const double *draws, void *actions_out) {
    if (!h || !obs || !actions_out) return fail(BS_E_INVALID, "null argument");
    DeviceGuard guard(h->device);
    const size_t EP = (size_t)h->E * h->P;
    const size_t out_bytes = (size_t)h->E * h->N * (h->cont ? 24 : 4);
    const float *d_obs; const double *d_draws; void *d_out;
    H2D(2, obs, EP * h->L * 4, d_obs);
    H2D(1, draws, EP * 5 * 8, d_draws);
    DOUT(7, actions_out, out_bytes, d_out);
    int rc = bs_instinct_actions(h, team, d_obs, d_draws, d_out, h->hstream); if (rc) return rc;
    D2H(actions_out, d_out, out_bytes);
    CU(cudaStreamSynchronize(h->hstream));
    return BS_OK;
}

// The learner's actions arrive over PCIe (36 B per env-tick for 3v3), which takes longer than the
// rollout itself.  The envs are therefore cut into chunks: chunk k+1's actions are copied (strided in
// T, cudaMemcpy2DAsync on the copy stream) while chunk k's rollout runs on the compute stream.  The call
// is split in two phases so that a caller can also overlap ACROSS calls: _begin returns as soon as the host
// action buffer has been consumed (its copies are done; the kernels may still be running), _end waits for
// the oldest call in flight and hands back its counters.  Two staging buffers rotate, at most two calls are
// in flight.  bs_rollout_fused_host = _begin + _end.
int bs_rollout_fused_host_begin(bs_handle *h, int T, int learner_team, int policy_red, int policy_blue,
                                const float *ext_actions, const bs_replay *replay, int64_t cursor) {
    int rc = rollout_check(h, T, learner_team, policy_red, policy_blue, ext_actions, replay, cursor);  // (BS_POL_EXTERNAL without actions: BS_E_INVALID)
    if (rc) return rc;
    if (h->inflight >= 2) return fail(BS_E_INVALID, "two host rollouts are already in flight: call bs_rollout_fused_host_end first");
    DeviceGuard guard(h->device);
    const int buf = h->next_buf;
    if (!ext_actions) {
        rc = rollout_range(h, T, learner_team, policy_red, policy_blue, nullptr, replay, cursor, 0, h->E, h->hstream);
        if (rc) return rc;
    } else {
        const size_t row_bytes = (size_t)h->N * h->A * 4, pitch = (size_t)h->E * row_bytes;
        void *d_ext = nullptr;
        if (h->stage_sz[8 + buf] < (size_t)T * pitch) CU(cudaDeviceSynchronize());   // growing a staging buffer frees the old one
        rc = stage_buf(h, 8 + buf, (size_t)T * pitch, &d_ext); if (rc) return rc;
        // this staging buffer was last read by the call two calls ago: its kernels must have finished
        CU(cudaStreamWaitEvent(h->cstream, h->call_done[buf], 0));
        const int K = 32 / h->P;
        int chunks = h->host_chunks > 0 ? h->host_chunks : 4;   // measured best of 3,4,6,8,12 at 65,536 envs (profiles/r1m_e2e_chunks.txt)
        int64_t per = (h->E + chunks - 1) / chunks;
        per = (per + 32 * K - 1) / (32 * K) * (32 * K);    // whole warps (lane-per-plane) and whole 32-env blocks (team-per-thread) per chunk
        if (per < 4096) per = h->E;                        // small handles: one chunk
        int k = 0;
        for (int64_t e0 = 0; e0 < h->E; e0 += per, ++k) {
            const int64_t e1 = e0 + per < h->E ? e0 + per : h->E;
            CU(cudaMemcpy2DAsync((char *)d_ext + e0 * row_bytes, pitch, (const char *)ext_actions + e0 * row_bytes, pitch,
                                 (size_t)(e1 - e0) * row_bytes, (size_t)T, cudaMemcpyHostToDevice, h->cstream));
            cudaEvent_t ev = h->chunk_ev[k % 16];
            CU(cudaEventRecord(ev, h->cstream));
            CU(cudaStreamWaitEvent(h->hstream, ev, 0));
            rc = rollout_range(h, T, learner_team, policy_red, policy_blue, (const float *)d_ext, replay, cursor, e0, e1, h->hstream);
            if (rc) return rc;
        }
        CU(cudaEventRecord(h->copy_done, h->cstream));
    }
    CU(cudaMemcpyAsync(h->pinned_stats + 8 * buf, h->p.counters, 8 * sizeof(int64_t), cudaMemcpyDeviceToHost, h->hstream));
    CU(cudaEventRecord(h->call_done[buf], h->hstream));
    h->inflight += 1; h->next_buf = 1 - buf;
    if (ext_actions) CU(cudaEventSynchronize(h->copy_done));   // the caller may refill its action buffer now
    return BS_OK;
}

int bs_rollout_fused_host_end(bs_handle *h, int64_t stats_out[8]) {
    if (!h) return fail(BS_E_INVALID, "null handle");
    if (h->inflight <= 0) return fail(BS_E_INVALID, "no host rollout in flight");
    DeviceGuard guard(h->device);
    const int buf = h->inflight == 2 ? h->next_buf : 1 - h->next_buf;   // the oldest call in flight
    CU(cudaEventSynchronize(h->call_done[buf]));
    if (stats_out) std::memcpy(stats_out, h->pinned_stats + 8 * buf, 8 * sizeof(int64_t));
    h->inflight -= 1;
    return BS_OK;
}

int bs_rollout_fused_host(bs_handle *h, int T, int learner_team, int policy_red, int policy_blue,
                          const float *ext_actions, const bs_replay *replay, int64_t cursor, int64_t stats_out[8]) {
    if (h) while (h->inflight > 0) { int rc = bs_rollout_fused_host_end(h, nullptr); if (rc) return rc; }
    int rc = bs_rollout_fused_host_begin(h, T, learner_team, policy_red, policy_blue, ext_actions, replay, cursor);
    if (rc) return rc;
    return bs_rollout_fused_host_end(h, stats_out);
}

}  // extern "C"
