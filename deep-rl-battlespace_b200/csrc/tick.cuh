// tick.cuh -- device-side Battlespace tick for sm_100a.
//
// Mapping: one LANE per plane, one env per group of G = 2N consecutive lanes (red planes first),
// K = 32 / G envs per warp.  A lane keeps its own plane in registers and a replicated view of the
// enemy team (refreshed with warp shuffles); the warp's live bullets sit PACKED in a per-warp pool --
// in shared memory while a kernel runs, in HBM between launches (same layout, copied coalesced) -- so
// any lane advances any bullet and every kernel shares one bullet pass.  (History: v1 unrolled a
// 12-slot register ring: 132 KB of SASS per tick, stall_no_inst 55 %; v2 kept a per-lane ring in shared
// memory and HBM: only ~1/4 of the slots hold a bullet, the loop ran at 8 of 32 lanes; profiles/r1_*.)
// Control flow that contains a shuffle is warp-uniform; per-env conditions are predicates.
//
// Arithmetic follows the reference operation by operation in IEEE double (compiled with
// --fmad=false so that no multiply-add is contracted), see the citations on each function.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace bs {

constexpr int SLOTS = 12;
constexpr int BLOCK = 128;
constexpr unsigned FULL = 0xffffffffu;
constexpr double PI = 3.14159265358979323846;
constexpr double DEG2RAD = PI / 180.0;  // math.radians
constexpr double RAD2DEG = 180.0 / PI;  // math.degrees
constexpr double TWO_PI = 2 * PI;
constexpr double DIAG = 1442.2205101855957;  // math.sqrt(1200**2 + 800**2), battle_env.py:230 (host asserts it)

// The same doubles in constant memory.  A double literal whose low word is not zero cannot be an immediate
// operand: the compiler materialises it with two UMOV per use -- 73 warp-instructions per tick (4.6 %) in r2a.
// As constant-bank operands (`DADD R, R, c[3][..]`) they cost nothing.  (Not `const`: a const-initialised
// __constant__ would be folded back into literals.)
static __constant__ double KC_TWO_PI = TWO_PI, KC_PI = PI, KC_DEG2RAD = DEG2RAD, KC_RAD2DEG = RAD2DEG, KC_DIAG = DIAG,
                           KC_RCP_DIAG = 1.0 / DIAG, KC_RCP_360 = 1.0 / 360.0, KC_RCP_35 = 1.0 / 35.0,
                           KC_NOISE_LO = -0.15, KC_NOISE_W = 0.15 - -0.15,          // np.random.uniform(-0.15, 0.15): low + (high - low) * u
                           KC_FIRE_P = 0.6, KC_FIRE_RANGE = 500.0 / 3 * 2,          // instinct/agent.py:43-44
                           KC_TENTH = 0.1;

// Bullet words.  pos: (x+64) | (y+64) << 16 | flags; the bias keeps both fields positive while a
// bullet is one update past the field edge, so pos + dlt is ONE integer add (dlt = sx + sy*65536).
constexpr uint32_t B_VALID = 1u << 31, B_XEXACT = 1u << 30, B_YEXACT = 1u << 29, B_POS = 0x1fffffffu;
constexpr uint32_t B_BIAS = 64u | (64u << 16);

enum { W_NONE = 0, W_RED = 1, W_BLUE = 2, W_TIE = 3 };
enum { POL_EXTERNAL = 0, POL_RANDOM = 1, POL_INSTINCT = 2, POL_INSTINCT_FF = 3 };
enum { ACT_I32 = 0, ACT_F32 = 1, ACT_F64 = 2, ACT_NONE = 3 };
enum { PURPOSE_AGENT_TICK = 0, PURPOSE_RESET_PLANE = 1, PURPOSE_RESET_BASE = 2 };
enum { C_WINS_RED = 0, C_WINS_BLUE, C_TIES, C_GAMES, C_RET_RED, C_RET_BLUE, C_TICKS, C_AGENT_STEPS };

struct Params {
    long long E;          // envs in this handle
    long long W;          // warps = ceil(E / K), K = 32 / 2N envs per warp
    long long env_offset; // global index of env 0 (sharding)
    uint32_t seed_lo, seed_hi;
    uint32_t rk[20];      // Philox round keys (seed + r * Weyl constants), filled by philox_round_keys(): constant-bank
                          // operands of the round XORs instead of two integer adds per round
    int tie_tick;         // first step call whose accumulated 0.1s reach 10+2N (battle_env.py:316-319)
    int exact_trig;       // BS_EXACT_TRIG=1 at bs_create: never speculate in fp32 (cos_sin_fast); for the parity test
    double r_hit_base, r_hit_plane, r_miss, r_die, r_lose;
    double r_code[4];     // reward of a bullet event code: 0 none, 1 miss, 2 base hit, 3 plane hit (bullets_dense)
    double diag;          // == DIAG
    // SoA state (see DESIGN.md "Data layout in HBM")
    uint32_t *plane_pos;  // [E*2N]      x | y<<16
    double *plane_dir;    // [E*2N]
    int8_t *plane_hp;     // [E*2N]
    uint32_t *base_pos;   // [E*2]       x | y<<16
    int16_t *base_hp;     // [E*2]
    uint2 *pool;          // [W][POOL_CAP] W = ceil(E/K) warps: the warp's bullets, packed (x = pos word, y = meta word)
    int *pool_cnt;        // [W]         entries in use (dead entries are squeezed out by the next pass)
    uint2 *lists;         // [ceil(E/32)][12N][64]  the same bullets as per-team lists, rollout2.cuh (one of the two layouts is current)
    int *list_cnt;        // [ceil(E/32)][64]
    uint32_t *meta;       // [E]         tick | env_done<<16 | winner<<24
    uint32_t *episode;    // [E]
    double *ret;          // [E*2]
    unsigned long long *counters;  // [8] report counters + [1] episode-log cursor
    // Observation table, built once per device by the host with the host libm -- the libm CPython's
    // math module calls -- and kept resident in the 126 MB L2 (DESIGN.md "Observation table"): one
    // 16-byte record per integer offset (dx, |dy|), fetched with ONE 128-bit gather per observed pair.
    const double2 *obs_tab;  // [T_DY+1][T_W]: .x = atan2(|dy|, dx) (battle_env.py:46), low word of .y =
                             // float32(sqrt(dx^2+dy^2)/diag*2-1) (battle_env.py:54-58,230)
    // Optional per-episode score log (main.py:219-221 score_dict): ring of {red score, blue score, code};
    // code = 1 red won, 2 blue won, 3 tie, 6 both bases fell in the same tick.  counters[8] is its cursor.
    float *ep_log;        // [ep_cap][3] or nullptr
    long long ep_cap;
};

// Table geometry: every pair of on-field points (sprites.py:9-10) has dx in [-1200,1200], dy in [-800,800].
constexpr int T_DX = 1200, T_DY = 800, T_W = 2 * T_DX + 1, T_H = T_DY + 1;

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011); counter layout in include/battlespace.h
#ifndef PHILOX_INLINE
#define PHILOX_INLINE __forceinline__
#endif
static __device__ PHILOX_INLINE uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                            uint32_t k0, uint32_t k1) {
#pragma unroll   // unrolled and inlined: +4 % on the fused rollout over an out-of-line rolled loop (r1n A/B)
    for (int r = 0; r < 10; ++r) {
        uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

// the same generator with the ten round keys taken from the kernel parameters (rk[2r], rk[2r+1])
__device__ __forceinline__ uint4 philox4x32_10_rk(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, const uint32_t (&rk)[20]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned long long p0 = (unsigned long long)0xD2511F53u * c0, p1 = (unsigned long long)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ rk[2 * r], n2 = (uint32_t)(p0 >> 32) ^ c3 ^ rk[2 * r + 1];
        c0 = n0; c1 = (uint32_t)p1; c2 = n2; c3 = (uint32_t)p0;
    }
    return make_uint4(c0, c1, c2, c3);
}

inline void philox_round_keys(uint32_t seed_lo, uint32_t seed_hi, uint32_t (&rk)[20]) {
    for (int r = 0; r < 10; ++r) { rk[2 * r] = seed_lo + (uint32_t)r * 0x9E3779B9u; rk[2 * r + 1] = seed_hi + (uint32_t)r * 0xBB67AE85u; }
}

__device__ __forceinline__ void env_philox(const Params &p, long long e, uint32_t episode, int tick,
                                           int agent, int purpose, uint32_t (&w)[4]) {
    unsigned long long g = (unsigned long long)(p.env_offset + e);
    uint4 r = philox4x32_10_rk((uint32_t)g, (uint32_t)(g >> 32), episode,
                               ((uint32_t)tick << 16) | ((uint32_t)agent << 8) | (uint32_t)purpose, p.rk);
    w[0] = r.x; w[1] = r.y; w[2] = r.z; w[3] = r.w;
}

__device__ __forceinline__ int randint_from(uint32_t w, int lo, int hi) {
    return lo + (int)__umulhi(w, (uint32_t)(hi - lo + 1));
}

// k / 2^24 for a 24-bit k, exactly: place k in the top of the mantissa of a double in [1,2), subtract 1.
__device__ __forceinline__ double unit24(uint32_t k) {
    return __hiloint2double((int)(0x3ff00000u | (k >> 4)), (int)(k << 28)) - 1.0;
}
// w / 2^32, exactly, the same way
__device__ __forceinline__ double unit32(uint32_t w) {
    return __hiloint2double((int)(0x3ff00000u | (w >> 12)), (int)(w << 20)) - 1.0;
}

// unit draws of an agent-tick block: u[0]=shoot_u, u[1..3]=noise_u, u[4]=spread_u
__device__ __forceinline__ void agent_tick_units(const uint32_t (&w)[4], double (&u)[5]) {
    u[1] = unit24(w[0] >> 8);
    u[2] = unit24(w[1] >> 8);
    u[3] = unit24(w[2] >> 8);
    u[0] = unit24(((w[0] & 0xffu) << 16) | ((w[1] & 0xffu) << 8) | (w[2] & 0xffu));
    u[4] = unit32(w[3]);
}

// ---------------------------------------------------------------------------------------------
template <int N>
struct Geo {
    static constexpr int P = 2 * N;       // planes per env
    static constexpr int G = P;           // lanes per env
    static constexpr int K = 32 / G;      // envs per warp
    static constexpr int L = 3 * N + 2;   // observation length, battle_env.py:132
    // Lane constants, one byte each in two registers.  They are pure functions of threadIdx, which the compiler
    // would rather recompute than keep when registers are short -- r1o/r2a spent 75 warp-instructions per tick
    // (4.7 %) re-deriving them inside the tick loop.  The empty asm makes the packed words opaque, so a use
    // costs one byte extract.
    uint32_t pa;                          // lane | agent index i << 8 | index in team k << 16 | team << 24
    uint32_t pb;                          // first lane of: my env (gb) | my team (tb) << 8 | the enemy team (eb) << 16 | valid << 24
    int e32;                              // env index within the handle

    // the launch covers envs [e0, e_end) of the handle
    __device__ __forceinline__ Geo(long long warp_id, long long e_end, long long e0 = 0) {
        const int lane_ = threadIdx.x & 31;
        int sub = lane_ / G;
        const int i_ = lane_ - sub * G;
        const int team_ = i_ >= N;
        const int k_ = i_ - team_ * N;
        const long long e_ = e0 + warp_id * K + sub;
        const bool valid_ = sub < K && e_ < e_end;
        if (!valid_) sub = 0;  // keep shuffle sources in range for idle lanes
        const int gb_ = sub * G;
        pa = (uint32_t)lane_ | ((uint32_t)i_ << 8) | ((uint32_t)k_ << 16) | ((uint32_t)team_ << 24);
        pb = (uint32_t)gb_ | ((uint32_t)(gb_ + team_ * N) << 8) | ((uint32_t)(gb_ + (1 - team_) * N) << 16) | ((valid_ ? 1u : 0u) << 24);
        e32 = (int)e_;
#ifndef BS_GEO_RECOMPUTE   // (A/B switch: without the barrier the compiler re-derives the fields where it uses them)
        asm volatile("" : "+r"(pa), "+r"(pb), "+r"(e32));
#endif
    }
    __device__ __forceinline__ int lane() const { return (int)__byte_perm(pa, 0, 0x4440); }
    __device__ __forceinline__ int i() const { return (int)__byte_perm(pa, 0, 0x4441); }
    __device__ __forceinline__ int k() const { return (int)__byte_perm(pa, 0, 0x4442); }
    __device__ __forceinline__ int team() const { return (int)(pa >> 24); }
    __device__ __forceinline__ int gb() const { return (int)__byte_perm(pb, 0, 0x4440); }
    __device__ __forceinline__ int tb() const { return (int)__byte_perm(pb, 0, 0x4441); }
    __device__ __forceinline__ int eb() const { return (int)__byte_perm(pb, 0, 0x4442); }
    __device__ __forceinline__ bool valid() const { return (pb >> 24) != 0; }
    __device__ __forceinline__ long long e() const { return (long long)e32; }
    __device__ __forceinline__ unsigned gmask() const { return (G == 32 ? FULL : ((1u << G) - 1u)) << gb(); }
};

// Registers of one lane.
template <int N>
struct Lane {
    int x, y, hp;           // own plane: rect centre (sprites.py:84,131) and hit points
    double dir;             // own direction in degrees (sprites.py:67,99)
    int ebx, eby;           // enemy base centre
    int bhp_own, bhp_enemy; // base hit points (replicated over the group)
    int tick, env_done, winner;
    uint32_t episode;
    // replicated view of the enemy team, refreshed by gather()
    uint32_t exy[N];        // x | y << 16
    uint32_t ehp4;          // 4 bits of hp per enemy
    unsigned am;            // bit j: enemy j is in self.agents (hp > 0)
};

template <int N>
__device__ __forceinline__ void gather(const Geo<N> &g, Lane<N> &s) {
    // positions by shuffle, the alive mask from ONE ballot; the enemies' hit points are only needed when a plane may die
    // inside the bullet loop, and resolve_plane_hits fetches them itself then
    const uint32_t mine = (uint32_t)s.x | ((uint32_t)s.y << 16);
#pragma unroll
    for (int j = 0; j < N; ++j) s.exy[j] = __shfl_sync(FULL, mine, g.eb() + j);
    s.am = (__ballot_sync(FULL, s.hp > 0) >> g.eb()) & ((1u << N) - 1u);
    s.ehp4 = 0;
}

// a / b for a compile-time divisor, correctly rounded, in 3 FMA-class instructions instead of the generic
// division routine: with y = RN(1/b) and q = RN(a*y) (within 1 ulp of a/b), r = a - q*b is exact in an
// FMA and RN(q + r*y) is the correctly rounded quotient (Markstein 1990).  |a| is at most a few thousand
// here, far from over/underflow.
__device__ __forceinline__ double div_const(double a, double b, double rcp_b) {
    const double q = a * rcp_b;
    const double r = fma(-q, b, a);
    return fma(r, rcp_b, q);
}

// float32(rel / 360) -- battle_env.py:231,242: the double quotient exactly as the reference rounds it, then
// the float32 store.  (Until r1n this was rel * (1/360) with a guard that fell back to a real division
// when the product sat within 4 ulps of a float rounding midpoint.)
__device__ __forceinline__ float div360_f32(double rel) { return (float)div_const(rel, 360.0, KC_RCP_360); }

// rel_angle / 360 -- battle_env.py:38-52,231 ; dist / diag * 2 - 1 -- battle_env.py:54-58,230.
// atan2, the % 2pi, degrees() and the sqrt were evaluated by the host for every integer (dx, dy);
// what is left is the part that depends on the observer's direction.
// The look-ups depend on positions only, which are final once the planes have moved, so they are
// issued (ObsLoads) before the bullet outcome is settled and consumed after it: the L2 latency of
// the gathers hides behind tick_end instead of stalling the warp.
template <int N>
struct ObsLoads {
    double r[N + 1];   // atan2(|dy|, dx) of the pair
    float od[N + 1];
    // (dy < 0 for a pair means atan2(dy, dx) = -r[k]: atan2 is odd in y, and the host verifies that its libm honours
    // this bit for bit when it builds the table; observe_finish takes the sign from the positions)
};

__device__ __forceinline__ void pair_load(const Params &p, int dx, int dy, double &r, float &od) {
    const double2 rec = __ldg(p.obs_tab + abs(dy) * T_W + (dx + T_DX));
    r = rec.x;
    od = __int_as_float(__double2loint(rec.y));
}

// rel_angle -- battle_env.py:38-52 from r = atan2(|dy|, dx): the reference's own operations in its own
// order, one rounding each: rads %= 2*pi (for a negative angle: + 2*pi), degrees(), 360 - degs, then
// 180 + direction - that, wrapped into (-180, 180].
__device__ __forceinline__ float pair_angle(double r0, bool neg, double a0) {
    // atan2(dy, dx) = -r0 for dy < 0, and `rads %= 2*pi` then adds 2*pi: (-r0) + 2*pi == 2*pi - r0, one rounding
    // either way (r0 > 0 whenever dy != 0, so the sign test of the reference is the sign of dy).
    const double rn = KC_TWO_PI - r0;
    const double r = neg ? rn : r0;
    const double t = 360 - r * KC_RAD2DEG;
    double rel = 180 + a0 - t;
    // `if angle < -180: angle += 360` (battle_env.py:49) can never fire here: a plane's direction stays in
    // [0, 360] (Plane.rotate, sprites.py:100-103; bs_set_state rejects anything else), so 180 + a0 >= 180
    // exactly, and t <= 360 exactly (r >= 0), hence rel >= -180.  Only the upper wrap is left; it subtracts
    // 360.0 or 0.0 (only the high word differs: one select).
    rel -= __hiloint2double(rel > 180 ? 0x40768000 : 0, 0);
    return div360_f32(rel);
}

// needs enemy POSITIONS from a gather() after the move (hp may still change)
template <int N>
__device__ __forceinline__ void observe_issue(const Params &p, const Lane<N> &s, ObsLoads<N> &q) {
    // Only pairs that can still be observed are fetched: an observer or an enemy that is dead now stays
    // dead (hp / am may only shrink before observe_finish, whose selects have the last word).  The gathers
    // are this path's scarcest resource -- ~1/3 of them used to be issued for dead planes.
    const bool me = s.hp > 0;
    q.r[0] = 0; q.od[0] = 0;
    if (me) pair_load(p, s.x - s.ebx, s.y - s.eby, q.r[0], q.od[0]);
#pragma unroll
    for (int j = 0; j < N; ++j) {
        q.r[j + 1] = 0; q.od[j + 1] = 0;
        const int dy = s.y - (int)(s.exy[j] >> 16);
        if (me && ((s.am >> j) & 1u)) pair_load(p, s.x - (int)(s.exy[j] & 0xffffu), dy, q.r[j + 1], q.od[j + 1]);
    }
}

// observe -- battle_env.py:202-244, for the lane's own plane; needs the final s.hp / s.am
template <int N>
__device__ __forceinline__ void observe_finish(const Lane<N> &s, const ObsLoads<N> &q, float (&o)[3 * N + 2]) {
    const bool me = s.hp > 0;
    // the angles of all N + 1 pairs first, on every lane (an unobserved pair's record is zero: harmless), then the
    // selects: `seen ? pair_angle(..) : -1` made the compiler put a divergent region around every pair (r2l: 18
    // control-flow instructions per warp-tick here)
    float ang[N + 1];
    // the sign of dy again from the positions (still in registers: nothing moves after the planes' move) instead of
    // packing N + 1 sign bits before the bullet pass and unpacking them here
    ang[0] = pair_angle(q.r[0], s.y < s.eby, s.dir);
    const uint32_t ythr = ((uint32_t)s.y << 16) | 0xffffu;
#pragma unroll
    for (int j = 0; j < N; ++j) ang[j + 1] = pair_angle(q.r[j + 1], s.exy[j] > ythr, s.dir);   // enemy y > own y
    o[0] = me ? q.od[0] : -1.0f; o[1] = me ? ang[0] : -1.0f;
#pragma unroll
    for (int j = 0; j < N; ++j) {
        const bool seen = me && ((s.am >> j) & 1u);
        o[2 + 3 * j] = seen ? 1.0f : -1.0f; o[3 + 3 * j] = seen ? q.od[j + 1] : -1.0f;
        o[4 + 3 * j] = seen ? ang[j + 1] : -1.0f;
    }
}

template <int N>
__device__ __forceinline__ void observe(const Params &p, const Lane<N> &s, float (&o)[3 * N + 2]) {
    ObsLoads<N> q;
    observe_issue<N>(p, s, q);
    observe_finish<N>(s, q, o);
}

// min(max(v, -1), 1) for a finite v: one |v| > 1 test, then +-1 assembled from the sign bit (fmin/fmax on doubles
// cost 13 instructions each with their NaN handling; two compare-selects cost 6-7).
__device__ __forceinline__ double clamp1(double v) {
    const bool big = fabs(v) > 1.0;
    const int hi = __double2hiint(v), lo = __double2loint(v);
    return __hiloint2double(big ? (int)(((uint32_t)hi & 0x80000000u) | 0x3ff00000u) : hi, big ? 0 : lo);
}

// 2 * (k / 2^24) - 1 for a 24-bit k, exactly: (2k - 2^24) is a 25-bit integer, its conversion and the scaling
// by 2^-24 are exact, and so are the reference-order operations on k / 2^24 they replace.
__device__ __forceinline__ double sym24(uint32_t k) { return __int2double_rn((int)(2u * k) - (1 << 24)) * (1.0 / 16777216.0); }

// InstinctAgent.choose_action -- instinct/agent.py:10-62 (float32 obs decoded in double:
// numpy 1.23.1 promotion).  Continuous: a[0..2]; discrete: adisc.
template <int N, bool CONT>
__device__ __forceinline__ void instinct(const float (&o)[3 * N + 2], double shoot_u,
                                         double n0, double n1, double n2, int &adisc, double (&a)[3]) {
    double td = ((double)o[0] + 1) / 2 * KC_DIAG;
    double ta = (double)o[1] * 360;
    double best = td * fabs(ta);  // base wins ties: replaced only by a strictly smaller score
#pragma unroll
    for (int j = 0; j < N; ++j) {
        double dj = ((double)o[3 + 3 * j] + 1) / 2 * KC_DIAG;
        double aj = (double)o[4 + 3 * j] * 360;
        double sc = (o[2 + 3 * j] == 1.0f) ? dj * fabs(aj) : 1000000.0;
        if (sc < best) { best = sc; td = dj; ta = aj; }
    }
    if (CONT) {
        double a2 = 0;
        if (td < KC_FIRE_RANGE && fabs(ta) < 20) a2 = shoot_u < KC_FIRE_P ? 1.0 : -1.0;
        double a0 = div_const(td, KC_DIAG, KC_RCP_DIAG) * 2 - 1;
        // agent.py:47-50: max(v, -1) when ang > 0 (v < 0), min(v, 1) otherwise (v >= 0): one clamp covers both
        // (no NaN can occur; clamp1 keeps the sign of a zero like the reference's comparisons do)
        const double v = div_const(-ta, 35.0, KC_RCP_35), a1 = clamp1(v);
        const double t0 = a0 + (KC_NOISE_LO + KC_NOISE_W * n0);
        const double t1 = a1 + (KC_NOISE_LO + KC_NOISE_W * n1);
        const double t2 = a2 + (KC_NOISE_LO + KC_NOISE_W * n2);
        a[0] = clamp1(t0);   // np.clip, agent.py:54
        a[1] = clamp1(t1);
        a[2] = clamp1(t2);
        adisc = 0;
    } else {
        adisc = (td < 500.0 / 2 && fabs(ta) < 20) ? 1 : (ta > 0 ? 3 : 2);
        a[0] = a[1] = a[2] = 0;
    }
}

// (cos, sin)(-radians(dir)) -- calc_new_xy, sprites.py:35-42.  Out of line: two call sites per tick.
static __device__ __noinline__ double2 cos_sin_dir(double dir) {
    double rad = -(dir * KC_DEG2RAD);
    double sn, cs;
    sincos(rad, &sn, &cs);
    return make_double2(cs, sn);
}

// fp32 speculation for the two places where a cosine only feeds a truncation (plane move, bullet
// step).  (cos, sin)(-radians(dir)) with the argument reduced in double and MUFU sin/cos in float:
// absolute error <= 2^-21.2 (MUFU on [-pi, pi]) + 1.9e-7 (float(rad)) < 7e-7, i.e. < 3.2e-5 px after the
// multiplication by <= 45.  floor_guarded() accepts the float result only when it is at least FLOOR_GUARD
// (8x that bound) away from an integer -- then floor() of the exact double product is the same integer
// -- and otherwise the caller takes the exact double path (about 1 lane-call in 2,000).
constexpr float FLOOR_GUARD = 2.5e-4f;

__device__ __forceinline__ void cos_sin_fast(double dir, float &cs, float &sn) {
    // the wrap into [-pi, pi] in float: the argument then carries float(rad) 2.4e-7 + the float add 1.2e-7 + float(2 pi)
    // 1.75e-7 = 5.4e-7 instead of 1.9e-7, i.e. cos / sin within 9.6e-7 and the product with <= 45 within 4.3e-5 px:
    // FLOOR_GUARD is still 5.8x that
    float r = (float)(-(dir * KC_DEG2RAD));   // dir in [-4, 364]: rad in [-6.36, 0.07]
    if (r < -3.14159265f) r += 6.28318531f;
    cs = __cosf(r); sn = __sinf(r);
}

__device__ __forceinline__ bool floor_guarded(float v, int &fl) {
    const float f = floorf(v), fr = v - f;
    fl = (int)f;
    return fr > FLOOR_GUARD && fr < 1.0f - FLOOR_GUARD;
}

// Plane.forward + clamp -- sprites.py:123-141 (rect.center = floats: C cast, then the
// screen clamp; centre x in [25,1175], y in [24,776] for the 50x48 sprite).
// FAST (continuous mode; in discrete mode the directions are multiples of 15 degrees and hit exact
// integers all the time): x + st*cos >= 25 - 27.5, and any negative sum clamps to 25 whether it is
// truncated or floored, so trunc(x + v) == x + floor(v) for the purpose of the clamped result.
template <bool FAST>
__device__ __forceinline__ void plane_forward(int &x, int &y, double dir, double speed, bool exact) {
    const double st = speed * KC_TENTH;
    int nx, ny;
    bool done = false;
    if (FAST && !exact) {
        float cs, sn;
        cos_sin_fast(dir, cs, sn);
        const float stf = (float)st;
        int fx, fy;
        const bool okx = floor_guarded(stf * cs, fx), oky = floor_guarded(stf * sn, fy);
        done = okx && oky;
        nx = x + fx; ny = y + fy;
    }
    if (!done) {
        double2 cssn = cos_sin_dir(dir);
        nx = (int)((double)x + st * cssn.x);
        ny = (int)((double)y + st * cssn.y);
    }
    x = min(max(nx, 25), 1175);
    y = min(max(ny, 24), 776);
}

// Plane.rotate -- sprites.py:93-103
__device__ __forceinline__ double plane_rotate(double dir, double a) {
    // |a| <= 35 (a clipped turn action times max_turn, or +-15) and dir in [0, 360]: the reference's two while loops
    // (sprites.py:100-103) run at most once each; as selects they cost no divergent region
    dir += a;
    dir = dir > 360 ? dir - 360 : dir;
    dir = dir < 0 ? dir + 360 : dir;
    return dir;
}

// Report counters.  Game ends are rare (once per ~150 ticks per env), so they go straight to the
// global counters, warp-aggregated, when they happen; only the tick count lives in a register.
struct Tally {
    int ticks = 0;
};

__device__ __forceinline__ long long warp_sum(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}

// Packed rect test.  D = (bullet fields) - (target fields): both 16-bit fields must land in
// [0, rx) x [0, ry).  A negative low field borrows from the high one, but then the low test has
// already failed; a negative high field makes D huge as unsigned.  (colliderect closed forms,
// plane -28<dx<28, -26<dy<25; base -34<dx<34, -33<dy<32.)
// Both fields <= (rx - 1, ry - 1)  <=>  the per-halfword unsigned minimum leaves D unchanged: VIMNMX.U16x2 + ISETP
// instead of LOP3 + 2 ISETP (rx, ry are compile-time constants at every call site; +1.2 % on the rollout, r2w).
__device__ __forceinline__ bool rect_hit(uint32_t D, uint32_t rx, uint32_t ry) {
    return __vminu2(D, ((ry - 1u) << 16) | (rx - 1u)) == D;
}
constexpr uint32_t PLANE_OFF = B_BIAS - (27u | (25u << 16));  // pos - PLANE_OFF: fields x+27, y+25

template <int N>
__device__ __forceinline__ unsigned plane_overlaps(const Lane<N> &s, uint32_t pos) {
    const uint32_t Pp = (pos & B_POS) - PLANE_OFF;
    unsigned cm = 0;
#pragma unroll
    for (int j = 0; j < N; ++j) cm |= (rect_hit(Pp - s.exy[j], 55u, 50u) ? 1u : 0u) << j;
    return cm;
}

// ---- one env tick -- parallel_env.step, battle_env.py:281-381 -----------------------------------
// The tick is split into three parts so that the single-tick kernels (bullets in a per-lane ring)
// and the fused rollout (bullets in a dense per-warp pool) share everything but the bullet pass:
//   tick_begin   preamble + actions of the alive planes; yields the spawn words of this lane
//   bullets_*    move bullets, settle misses / base hits, resolve plane hits in list order
//   tick_end     rewards in the reference's accumulation order, base damage, win / tie, report sums
struct TickCtx {
    bool active, run, tie_now, time_tie;
    uint32_t spawnP;   // bullet pos word (0 = no spawn)
    int32_t spawnD;    // sx + sy * 65536
};

// process_action of ONE plane -- battle_env.py:383-424 -- and the bullet it may spawn (Bullet.__init__,
// sprites.py:293-318).  acts: the plane is alive and its env runs this tick.  Shared by the lane-per-plane
// kernels (tick_begin) and the team-per-thread rollout (rollout2.cuh).
//   adisc / a[3] : the decoded action (clipped / argmax-ed);  spread_u : its Bullet draw.
//   spawnP / spawnD : bullet pos word (0 = no spawn) and sx + sy * 65536.
template <bool CONT>
__device__ __forceinline__ void plane_action(const Params &p, int &x, int &y, double &dir, bool acts, int adisc,
                                             const double (&a)[3], double spread_u, uint32_t &spawnP, int32_t &spawnD) {
    const int x0 = x, y0 = y;
    const double d0 = dir;
    bool shoot = false;
    if (acts) {
        double mdir = dir, speed = 215;
        bool move = true;
        if (CONT) {
            speed = ((a[0] + 1) / 2) * (275 - 200) + 200;
        } else {
            move = adisc >= 0 && adisc <= 3;
            shoot = adisc == 1;
            if (adisc == 2) mdir = plane_rotate(dir, 15);
            if (adisc == 3) mdir = plane_rotate(dir, -15);
        }
        if (move) plane_forward<CONT>(x, y, mdir, speed, p.exact_trig != 0);
        if (CONT) { dir = plane_rotate(dir, a[1] * 35); shoot = a[2] > 0; }
        else dir = mdir;
    }
    // Bullet.__init__ -- sprites.py:293-318.  From here on the bullet is integer: its per-update
    // displacement d = 45*cos/sin is constant, and trunc(int + d) == int + floor(d) except that
    // a result in (-1,0) truncates to 0 (flagged "exact" when d itself is an integer).
    spawnP = 0; spawnD = 0;
    if (shoot) {
        const double bdir = d0 + (spread_u * 8 - 4);
        const uint32_t at = (((uint32_t)x0 | ((uint32_t)y0 << 16)) + B_BIAS) | B_VALID;
        bool done = false;
        if (CONT && !p.exact_trig) {  // fp32 speculation, see cos_sin_fast: a guarded floor is never an exact integer
            float cs, sn;
            cos_sin_fast(bdir, cs, sn);
            int fx, fy;
            const bool okx = floor_guarded(45.0f * cs, fx), oky = floor_guarded(45.0f * sn, fy);
            done = okx && oky;
            spawnP = at; spawnD = fx + fy * 65536;
        }
        if (!done) {
            double2 cssn = cos_sin_dir(bdir);
            double ddx = (450 * 0.1) * cssn.x, ddy = (450 * 0.1) * cssn.y;
            double fx = floor(ddx), fy = floor(ddy);
            spawnP = at | (fx == ddx ? B_XEXACT : 0u) | (fy == ddy ? B_YEXACT : 0u);
            spawnD = (int)fx + (int)fy * 65536;
        }
    }
}

template <int N, bool CONT>
__device__ __forceinline__ void tick_begin(const Params &p, const Geo<N> &g, Lane<N> &s, int adisc,
                                           const double (&a)[3], double spread_u, bool no_actions, TickCtx &c) {
    c.active = g.valid() && !s.env_done;                                  // :303-306
    const unsigned alive_b = __ballot_sync(FULL, c.active && s.hp > 0);
    const bool any_alive = !no_actions && (alive_b & g.gmask()) != 0;   // len(actions) == 0 also ties
    c.tie_now = c.active && !any_alive;                                 // :309-313
    const bool adv = c.active && any_alive;
    s.tick += adv ? 1 : 0;                                              // :316
    c.time_tie = adv && s.tick >= p.tie_tick;                           // :319-323
    c.run = adv && !c.time_tie;
    // ---- actions of the planes that are alive, :325-329 + process_action :383-424 -----------
    plane_action<CONT>(p, s.x, s.y, s.dir, c.run && s.hp > 0, adisc, a, spread_u, c.spawnP, c.spawnD);
}

// One Bullet.update (sprites.py:321-351) up to the plane test, on the packed words.  `age` counts
// the updates including this one.  Returns 1 = miss (range or off-screen, :334-341), 0 otherwise;
// P is advanced (with the trunc-toward-zero corner: a coordinate of -1 that came from a sum in
// (-1,0) is 0, the C cast in pygame's Rect setter).
__device__ __forceinline__ bool bullet_advance(uint32_t &P, int32_t dlt, int age) {
    P += (uint32_t)dlt;
    // a biased field of 63 is a coordinate of -1; unless the step is an exact integer it came from a sum in
    // (-1, 0), which the C cast turns into 0.  Branch-free (two compares on the packed word).
    if ((P & (0xffffu | B_XEXACT)) == 63u) P += 1u;
    if ((P & (0x1fff0000u | B_YEXACT)) == (63u << 16)) P += 0x10000u;
    const uint32_t Q = (P & B_POS) - B_BIAS;  // fields x, y (a negative one wraps to a huge unsigned value)
    const bool miss = __vminu2(Q, (800u << 16) | 1200u) != Q;   // a field beyond (1200, 800), negative ones included (VIMNMX.U16x2)
    return miss || age >= SLOTS;  // 12th update: dist_travelled >= 500, sprites.py:334-337
}

// Phase 2 of the bullet loop (ordered; only for ages where some lane of the warp has a candidate):
// resolve the N same-age bullets of each team in agent order; a plane that dies is invisible to
// later ones (battle_env.py:348-360).  cm_of(ai) = this lane's overlap mask for its bullet of age
// index ai; drop(ai) removes that bullet.  Updates ev (code 3 = plane hit), s.hp, s.ehp4, s.am.
template <int N, class CmOf, class Drop>
__device__ __forceinline__ void resolve_plane_hits(const Geo<N> &g, Lane<N> &s, unsigned cand, unsigned &ev,
                                                   int &died_ai, CmOf cm_of, Drop drop) {
    unsigned allc = __reduce_or_sync(FULL, cand);
    {   // the enemies' hit points before the bullet loop (gather() no longer carries them)
        const uint32_t myhp = (uint32_t)(s.hp > 0 ? s.hp : 0);
        s.ehp4 = 0;
#pragma unroll
        for (int j = 0; j < N; ++j) s.ehp4 |= __shfl_sync(FULL, myhp, g.eb() + j) << (4 * j);
    }
    while (allc) {
        const int ai = __ffs(allc) - 1;
        allc &= allc - 1;
        unsigned cm = 0;
        if ((cand >> ai) & 1u) cm = cm_of(ai);
        int victim = -1;
#pragma unroll 1
        for (int kk = 0; kk < N; ++kk) {
            const unsigned m = cm & s.am;
            const int c_me = m ? (__ffs(m) - 1) : -1;
            const int c_team = __shfl_sync(FULL, c_me, g.tb() + kk);   // teammate kk's victim
            const int c_foe = __shfl_sync(FULL, c_me, g.eb() + kk);    // enemy kk's victim
            if (kk == g.k()) victim = c_me;
            if (c_team >= 0) {
                s.ehp4 -= 1u << (4 * c_team);
                if (((s.ehp4 >> (4 * c_team)) & 0xfu) == 0) s.am &= ~(1u << c_team);
            }
            if (g.valid() && c_foe == g.k() && s.hp > 0) {
                s.hp -= 1;
                if (s.hp <= 0 && died_ai < 0) died_ai = ai;
            }
        }
        if (victim >= 0) { ev |= 3u << (2 * ai); drop(ai); }
    }
}

// ---- bullets in a dense per-warp pool -- battle_env.py:332-360, sprites.py:321-351 -----------------
// With one 12-slot ring per lane only ~1/4 of the slots hold a bullet, so a lane-per-owner loop runs
// at ~8 of 32 lanes (profiles/r1c_*).  The warp's live bullets are therefore kept packed, any lane
// updates any bullet, and results travel back to the owner lane as bits keyed by (owner, age) --
// which is all the ordered part of the loop needs.  Phase 1 (order free): move every bullet, settle
// misses and base hits, flag overlaps with enemy planes alive at the start of the loop; phase 2:
// resolve_plane_hits.  ev: 2 bits per age index ai (0 = oldest): 1 miss, 2 base hit, 3 plane hit.
//   pos  : as in the ring (0 = dead entry, squeezed out by the next pass)
//   meta : (sx + sy*65536) in bits 0..22 (two's complement) | owner lane << 23 | updates done << 28
constexpr int WARPS = BLOCK / 32;
constexpr int POOL_CAP = 32 * SLOTS;  // <= 11 survivors + 1 spawn per lane at the start of a pass
struct WarpPool {
    uint2 ent[POOL_CAP];         // .x = pos word, .y = meta word: the HBM layout, moved with 64-bit accesses
    uint32_t ev[32], cand[32];   // per owner lane: event codes / candidate ages of this tick
    uint32_t hits[32];           // per plane lane: bullets whose first overlapped plane (alive at the start of the loop) it is
    uint32_t lastai[32];         // per plane lane: the largest age index among those bullets (the one that lands last)
    uint32_t candt[SLOTS][32];   // per (age index, owner): pool index << 8 | overlap mask
    uint32_t plane[32];          // x | y << 16 of every lane's plane after the move (0x7fff7fff: not alive)
    uint32_t baseT[32];          // enemy-base test word | first lane of the enemy team << 27
};

__device__ __forceinline__ uint32_t pack_meta(int32_t dlt, int owner, int age) {
    return ((uint32_t)dlt & 0x7fffffu) | ((uint32_t)owner << 23) | ((uint32_t)age << 28);
}
__host__ __device__ __forceinline__ int32_t meta_dlt(uint32_t m) { return (int32_t)(m << 9) >> 9; }

template <int N>
__device__ __forceinline__ void pool_publish_base(const Geo<N> &g, const Lane<N> &s, WarpPool &w) {
    w.baseT[g.lane()] = ((((uint32_t)(s.ebx - 33)) | ((uint32_t)(s.eby - 32) << 16)) + B_BIAS) | ((uint32_t)g.eb() << 27);
}

// The bullets of an env that is reset vanish (self.bullets = [], battle_env.py:272): the callers drop them
// with pool_filter() when that happens (once per game), which keeps the per-tick loop free of that test.
// The pass body is branch-free.  The fused rollout is bound by instruction issue (+5.6 % when the branches went); the
// single-tick kernel preferred the branching form while it ran at 64 registers (424 vs 449 us per tick at 1,048,576 envs,
// r2n) and no longer does at 72 (403.9 vs 406.0 us, r2y), so the branching copy is gone.
// SINK: the three places where a lane may have nothing to publish (no spawn, no event, no hit) run without a divergent
// region, non-publishing lanes writing a harmless value to a harmless place.  +1.3 % on the issue-bound fused rollout; the latency-bound
// single-tick kernel is faster with the branches (410.0 vs 414.3 us at 1,048,576 envs, 35.2 vs 37.2 us at 65,536).
template <int N, bool SINK>
__device__ __forceinline__ void bullets_dense(const Geo<N> &g, Lane<N> &s, WarpPool &w, int &cnt, const TickCtx &c,
                                              unsigned &ev, int &died_ai) {
    const int lane = g.lane();
    const unsigned lt = (1u << lane) - 1u;
    // a plane that is not in self.agents is published far off the field: every rect test against it fails
    w.plane[lane] = s.hp > 0 ? ((uint32_t)s.x | ((uint32_t)s.y << 16)) : 0x7fff7fffu;
    const unsigned runmask = __ballot_sync(FULL, c.run);
    const unsigned sp = __ballot_sync(FULL, c.spawnP != 0);
    // (sink of the spawn store: the last slot of the pool, which is free whenever some lane does not spawn -- 32 x 11 survivors
    //  + the spawns fill it exactly; the dead word written there is beyond cnt_in.  Sink of the event atomic: the lane's own ev word,
    //  OR-ing 0 changes nothing.)
    if (SINK) *(c.spawnP ? &w.ent[cnt + __popc(sp & lt)] : &w.ent[POOL_CAP - 1]) = make_uint2(c.spawnP, pack_meta(c.spawnD, lane, 0));
    else if (c.spawnP) w.ent[cnt + __popc(sp & lt)] = make_uint2(c.spawnP, pack_meta(c.spawnD, lane, 0));
    const int cnt_in = cnt + __popc(sp);
    __syncwarp();
    // (the caller has run gather(): enemy positions after the move, hp before the bullet loop)
    int out = 0;
    {
    // Branch-free body: every lane runs the advance, the base test and the N plane tests, and the outcome is assembled
    // with selects.  The chain miss / else base / else planes cost four divergent regions (BSSY / BRA / BSYNC, ~14
    // instructions per pass) that almost every pass executed anyway, since some lane of 32 always takes each side.
#pragma unroll 1
    for (int base = 0; base < cnt_in; base += 32) {
        const int idx = base + lane;
        const uint2 en = w.ent[idx];                    // idx < 32 * ceil(cnt_in / 32) <= POOL_CAP (cnt_in <= POOL_CAP by construction)
        const uint32_t P0 = idx < cnt_in ? en.x : 0u;
        uint32_t M = en.y;
        const int owner = (M >> 23) & 31;
        const int age = (int)(M >> 28) + 1;             // updates including this one
        const int ai = SLOTS - age;                     // age index, 0 = oldest
        const bool live = (P0 & B_VALID) != 0;          // dead entries (plane hits of the last tick) are squeezed out
        const bool act = live && ((runmask >> owner) & 1u);   // the bullets of an env that did not run this tick stay put
        const uint32_t bt = w.baseT[owner];
        uint32_t P = P0;
        const bool miss = bullet_advance(P, meta_dlt(M), age);
        const bool hitb = rect_hit((P & B_POS) - (bt & 0x07ffffffu), 67u, 64u);
        const uint32_t Pp = (P & B_POS) - PLANE_OFF;
        const int eb = bt >> 27;
        unsigned cm = 0;
#pragma unroll
        for (int j = 0; j < N; ++j) cm |= (rect_hit(Pp - w.plane[eb + j], 55u, 50u) ? 1u : 0u) << j;
        const unsigned code = act ? (miss ? 1u : (hitb ? 2u : 0u)) : 0u;
        if (!act || code) cm = 0;
        if (!act) P = P0; else if (!code) M += 1u << 28;
        if (SINK) atomicOr(&w.ev[code ? owner : lane], code << (2 * ai));   // a bullet without an event ORs 0 into its lane's own word
        else if (code) atomicOr(&w.ev[owner], code << (2 * ai));
        const bool keep = live && !code;
        const unsigned kb = __ballot_sync(FULL, keep);
        const int nidx = out + __popc(kb & lt);
        if (keep) w.ent[nidx] = make_uint2(P, M);       // nidx <= idx: never ahead of a read
        if (__any_sync(FULL, cm != 0)) {                // an overlap with a plane is rare: about one bullet per warp-tick
            if (cm) {
                atomicAdd(&w.hits[eb + __ffs(cm) - 1], 1u);
                atomicMax(&w.lastai[eb + __ffs(cm) - 1], (unsigned)ai);
                atomicOr(&w.cand[owner], 1u << ai);
                w.candt[ai][owner] = ((uint32_t)nidx << 8) | cm;
            }
        }
        out += __popc(kb);
    }
    }
    cnt = out;
    __syncwarp();
    ev = w.ev[lane];
    const unsigned cand = w.cand[lane];
    const int nh = (int)w.hits[lane];
    w.ev[lane] = 0; w.cand[lane] = 0; w.hits[lane] = 0;
    died_ai = -1;
    // Phase 2.  The order of the list only matters when a plane can die inside this tick: if every plane is
    // overlapped by fewer bullets than it has hit points, self.agents does not change during the loop and
    // every candidate hits the first plane it overlaps (battle_env.py:348-360), whatever the order.
    // ... and it does not matter either when a plane takes EXACTLY as many hits as it has hit points: the bullets counted
    // in hits[] are those whose first overlapped plane it is, all of them land while it is still alive, the last one in
    // list order -- the youngest, lastai[] -- kills it, and no bullet is left over to pass through the wreck.  Only a
    // plane with MORE overlapping bullets than hit points needs the ordered pass (a few per cent of the deaths).
    int la = 0;
    if (SINK || nh) { la = (int)w.lastai[lane]; w.lastai[lane] = 0; }
    if (!__any_sync(FULL, s.hp > 0 && nh > s.hp)) {
        s.hp -= nh;                                     // (a dead plane collects no hits: it is off the field)
        if (nh && s.hp <= 0) died_ai = la;
        unsigned cc = cand;
        while (cc) {
            const int ai = __ffs(cc) - 1;
            cc &= cc - 1;
            ev |= 3u << (2 * ai);
            w.ent[w.candt[ai][lane] >> 8].x = 0;
        }
    } else {
        resolve_plane_hits<N>(g, s, cand, ev, died_ai,
            [&](int ai) { return w.candt[ai][lane] & 0xffu; },
            [&](int ai) { w.ent[w.candt[ai][lane] >> 8].x = 0; });
    }
    s.am = (__ballot_sync(FULL, s.hp > 0) >> g.eb()) & ((1u << N) - 1u);   // a plane may have died on the unordered path
}

// HBM <-> shared memory: the pool has the same layout in both, so a launch copies `cnt` packed entries
// (coalesced 8-byte words), not 24 ring words per lane.  gw = global warp index = first env / K.
// wvalid: the warp owns envs of this launch (the last block's spare warps, and in a chunked launch the
// warps past the chunk, must neither read nor write a pool).
// pool_load() in two halves (see lane_fetch): the loads of the count and of the first POOL_SPEC x 32 entries, then the rest
constexpr int POOL_SPEC_N = 4;
__device__ __forceinline__ void pool_fetch(const Params &p, long long gw, bool wvalid, int lane, uint2 (&v)[POOL_SPEC_N], int &cnt) {
    const uint2 *src = p.pool + (wvalid ? gw : 0) * POOL_CAP;
#pragma unroll
    for (int q = 0; q < POOL_SPEC_N; ++q) v[q] = src[q * 32 + lane];
    cnt = p.pool_cnt[wvalid ? gw : 0];
}
__device__ __forceinline__ void pool_commit(const Params &p, long long gw, bool wvalid, int lane, const uint2 (&v)[POOL_SPEC_N], WarpPool &w, int &cnt) {
    if (!wvalid) cnt = 0;
    const uint2 *src = p.pool + (wvalid ? gw : 0) * POOL_CAP;
#pragma unroll
    for (int q = 0; q < POOL_SPEC_N; ++q)
        if (q * 32 + lane < cnt) w.ent[q * 32 + lane] = v[q];
#pragma unroll 1
    for (int idx = POOL_SPEC_N * 32 + lane; idx < cnt; idx += 32) w.ent[idx] = src[idx];
    w.ev[lane] = 0; w.cand[lane] = 0; w.hits[lane] = 0; w.lastai[lane] = 0;
    __syncwarp();
}

__device__ __forceinline__ void pool_load(const Params &p, long long gw, bool wvalid, int lane, WarpPool &w, int &cnt) {
    // The first POOL_SPEC entries are fetched without waiting for the count (the slots always exist), so
    // that count, entries and the caller's other state loads share ONE memory round trip.
    constexpr int POOL_SPEC = 4;  // x 32 entries: covers the typical 3v3 warp (~120 live bullets)
    const uint2 *src = p.pool + (wvalid ? gw : 0) * POOL_CAP;
    uint2 v[POOL_SPEC];
#pragma unroll
    for (int q = 0; q < POOL_SPEC; ++q) v[q] = src[q * 32 + lane];
    cnt = wvalid ? p.pool_cnt[gw] : 0;
#pragma unroll
    for (int q = 0; q < POOL_SPEC; ++q)
        if (q * 32 + lane < cnt) w.ent[q * 32 + lane] = v[q];
    for (int idx = POOL_SPEC * 32 + lane; idx < cnt; idx += 32) w.ent[idx] = src[idx];
    w.ev[lane] = 0; w.cand[lane] = 0; w.hits[lane] = 0; w.lastai[lane] = 0;
    __syncwarp();
}

__device__ __forceinline__ void pool_store(const Params &p, long long gw, bool wvalid, int lane, const WarpPool &w, int cnt) {
    __syncwarp();
    if (!wvalid) return;
    uint2 *dst = p.pool + gw * POOL_CAP;
    for (int idx = lane; idx < cnt; idx += 32) dst[idx] = w.ent[idx];
    if (lane == 0) p.pool_cnt[gw] = cnt;
}

// Drops the bullets of the lanes in `killmask` (their envs were reset: self.bullets = [], battle_env.py:272)
// and squeezes dead entries out, in place.
__device__ __forceinline__ void pool_filter(WarpPool &w, int &cnt, unsigned killmask, int lane) {
    const unsigned lt = (1u << lane) - 1u;
    __syncwarp();
    int out = 0;
#pragma unroll 1
    for (int base = 0; base < cnt; base += 32) {
        const int idx = base + lane;
        uint2 en = make_uint2(0u, 0u);
        if (idx < cnt) en = w.ent[idx];
        const bool keep = (en.x & B_VALID) != 0 && !((killmask >> ((en.y >> 23) & 31)) & 1u);
        const unsigned kb = __ballot_sync(FULL, keep);
        if (keep) w.ent[out + __popc(kb & lt)] = en;    // never ahead of a read
        out += __popc(kb);
    }
    cnt = out;
    __syncwarp();
}

//   rew : this lane's reward for the tick (double, reference accumulation order).
template <int N>
__device__ __forceinline__ void tick_end(const Params &p, const Geo<N> &g, Lane<N> &s, const TickCtx &c, unsigned ev,
                                         int died_ai, double &rew, double &ret0, double &ret1, Tally &tl) {
    rew = 0.0;
    // Rewards in the reference's accumulation order (:337-360): per age, own bullet outcome; a
    // victim's die_punishment precedes its own same-age outcome for blue (red bullets come first
    // in the list) and follows it for red.  A death is rare (once per plane and game), so the common
    // loop only walks the lane's own events, oldest first; the rewards sit in a 4-entry table indexed
    // by the event code.
    if (!__any_sync(FULL, died_ai >= 0)) {
        unsigned e2 = ev;
        {   // The two oldest events without a loop: a lane rarely has more in one tick, 0.0 + a == a and a + b == b + a
            // exactly, so up to two terms need no ordering; the loop below only runs for a third one.
            const int sh1 = e2 ? ((__ffs(e2) - 1) & ~1) : 0;
            const double ra = p.r_code[(e2 >> sh1) & 3u];
            e2 &= ~(3u << sh1);
            const int sh2 = e2 ? ((__ffs(e2) - 1) & ~1) : 0;
            const double rb = p.r_code[(e2 >> sh2) & 3u];
            e2 &= ~(3u << sh2);
            rew = (rew + ra) + rb;
        }
        if (__any_sync(FULL, e2 != 0))
        while (e2 != 0) {
            const int sh = (__ffs(e2) - 1) & ~1;
            rew += p.r_code[(e2 >> sh) & 3u];
            e2 &= ~(3u << sh);
        }
    } else {
        unsigned e2 = ev;
        int da = died_ai;
        while (e2 != 0 || da >= 0) {
            const int ae = e2 ? ((__ffs(e2) - 1) >> 1) : 64;
            if (da >= 0 && (da < ae || (da == ae && g.team() == 1))) { rew += p.r_die; da = -1; }
            else {
                rew += p.r_code[(e2 >> (2 * ae)) & 3u];
                e2 &= ~(3u << (2 * ae));
            }
        }
    }
    // base damage: no alive check, hp may go negative.  code 2 = high bit only.
    const int base_hits = __popc((ev >> 1) & ~ev & 0x00555555u);
    int mine_total = 0, foe_total = 0;
#pragma unroll
    for (int kk = 0; kk < N; ++kk) {
        mine_total += __shfl_sync(FULL, base_hits, g.tb() + kk);
        foe_total += __shfl_sync(FULL, base_hits, g.eb() + kk);
    }
    s.bhp_enemy -= mine_total;
    s.bhp_own -= foe_total;

    // ---- win / tie -- :363-372, 469-496 -------------------------------------------------------
    const bool i_won = c.run && s.bhp_enemy <= 0;  // the winners' planes get lose_punishment (sic)
    const bool foe_won = c.run && s.bhp_own <= 0;
    if (i_won) rew += p.r_lose;
    const bool red_won = g.team() == 0 ? i_won : foe_won, blue_won = g.team() == 0 ? foe_won : i_won;
    const bool tie = c.tie_now || c.time_tie;
    if (red_won) s.winner = W_RED;
    if (blue_won) s.winner = W_BLUE;  // both bases dead: ends as 'blue', two games counted
    if (tie) s.winner = W_TIE;
    if (red_won || blue_won || tie) s.env_done = 1;

    // ---- report sums (main.py:193-203,218-253); the env's first lane keeps them ----------------
    double r0 = 0, r1 = 0;
    {   // each team's first lane sums its team left to right (0 + a == a exactly, so the chain starts at its own reward), then
        // red's first lane fetches blue's total: N double shuffles instead of 2N, same additions in the same order
        double t = rew;
#pragma unroll
        for (int kk = 1; kk < N; ++kk) t += __shfl_sync(FULL, rew, g.tb() + kk);
        r0 = t;
        r1 = __shfl_sync(FULL, t, g.gb() + N);
    }
    const bool first = c.active && g.i() == 0;
    if (first) { ret0 += r0; ret1 += r1; tl.ticks += 1; }
    const bool ended = first && s.env_done;
    if (__any_sync(FULL, ended)) {
        const int nr = __popc(__ballot_sync(FULL, ended && red_won)), nb = __popc(__ballot_sync(FULL, ended && blue_won));
        const int nt = __popc(__ballot_sync(FULL, ended && tie));
        const long long sr = warp_sum(ended ? __double2ll_rn(ret0 * 1048576.0) : 0ll);
        const long long sb = warp_sum(ended ? __double2ll_rn(ret1 * 1048576.0) : 0ll);
#ifndef BS_NO_EPLOG
        if (p.ep_log != nullptr && ended) {  // capacity is a power of two: no 64-bit division in the tick loop
            const unsigned long long slot = atomicAdd(p.counters + 8, 1ull) & (unsigned long long)(p.ep_cap - 1);
            p.ep_log[slot * 3 + 0] = (float)ret0; p.ep_log[slot * 3 + 1] = (float)ret1;
            p.ep_log[slot * 3 + 2] = tie ? 3.0f : (red_won && blue_won ? 6.0f : (red_won ? 1.0f : 2.0f));
        }
#endif
        if (g.lane() == 0) {
            if (nr) atomicAdd(p.counters + C_WINS_RED, (unsigned long long)nr);
            if (nb) atomicAdd(p.counters + C_WINS_BLUE, (unsigned long long)nb);
            if (nt) atomicAdd(p.counters + C_TIES, (unsigned long long)nt);
            atomicAdd(p.counters + C_GAMES, (unsigned long long)(nr + nb + nt));
            atomicAdd(p.counters + C_RET_RED, (unsigned long long)sr);
            atomicAdd(p.counters + C_RET_BLUE, (unsigned long long)sb);
        }
    }
}

// parallel_env.reset for one lane -- battle_env.py:246-279, sprites.py:74-91,238-252.
// d_* are this lane's spawn values when injected (inj=true), else Philox.
template <int N>
__device__ __forceinline__ void reset_lane(const Params &p, const Geo<N> &g, Lane<N> &s, bool inj,
                                           const int32_t *draws /* this env's [4+6N] or null */) {
    s.episode += 1;
    int rbx, rby, bbx, bby, px, py, pd;
    if (inj) {
        rbx = draws[0]; rby = draws[1]; bbx = draws[2]; bby = draws[3];
        px = draws[4 + 3 * g.i()]; py = draws[5 + 3 * g.i()]; pd = draws[6 + 3 * g.i()];
    } else {
        uint32_t w[4];
        env_philox(p, g.e(), s.episode, 0, 0, PURPOSE_RESET_BASE, w);
        rbx = randint_from(w[0], 62, 379); rby = randint_from(w[1], 62, 738);       // sprites.py:246-247
        bbx = randint_from(w[2], 758, 1138); bby = randint_from(w[3], 62, 738);     // sprites.py:250-251
        env_philox(p, g.e(), s.episode, 0, g.i(), PURPOSE_RESET_PLANE, w);
        if (g.team() == 0) { px = randint_from(w[0], 50, 383); pd = randint_from(w[2], 270, 450); }  // :82-85
        else { px = randint_from(w[0], 766, 1150); pd = randint_from(w[2], 90, 270); }             // :88-91
        py = randint_from(w[1], 48, 752);
    }
    if (g.team() == 0 && pd >= 360) pd -= 360;                                        // sprites.py:86
    s.x = px; s.y = py; s.dir = (double)pd; s.hp = 4;                               // battle_env.py:92
    s.ebx = g.team() == 0 ? bbx : rbx; s.eby = g.team() == 0 ? bby : rby;
    s.bhp_own = s.bhp_enemy = 5 * N;                                                // battle_env.py:91
    s.tick = 0; s.env_done = 0; s.winner = W_NONE;   // (the caller drops the env's bullets: kill / pool_filter)
}

// ---- state <-> registers / shared ring --------------------------------------------------------
// The same in two halves for the single-tick kernel, which lives for one tick and is bound by the latency of exactly
// these loads: lane_fetch() only ISSUES them (no divergent region: an idle lane reads env 0's words and discards
// them), so that the pool entries and the caller's actions can be requested in the same memory round trip before
// lane_unpack() waits for the first word.  (r2y: the region around load_lane() made the unpacking -- and its wait --
// precede the pool loads, and those precede the action loads: three DRAM round trips in a row.)
struct LaneRaw {
    uint32_t pos, eb, m, ep;
    double dir, r0, r1;
    int hp, bo, be;
};

template <int N>
__device__ __forceinline__ void lane_fetch(const Params &p, const Geo<N> &g, LaneRaw &r) {
    constexpr int P = 2 * N;
    const bool v = g.valid();
    const long long e = v ? g.e() : 0;
    const long long a = e * P + g.i();
    r.pos = p.plane_pos[a];
    r.dir = p.plane_dir[a];
    r.hp = p.plane_hp[a];
    r.eb = p.base_pos[e * 2 + (1 - g.team())];
    r.bo = p.base_hp[e * 2 + g.team()];
    r.be = p.base_hp[e * 2 + (1 - g.team())];
    r.m = p.meta[e];
    r.ep = p.episode[e];
    r.r0 = p.ret[e * 2]; r.r1 = p.ret[e * 2 + 1];
}

template <int N>
__device__ __forceinline__ void lane_unpack(const Geo<N> &g, const LaneRaw &r, Lane<N> &s, double &ret0, double &ret1) {
    const bool v = g.valid();
    s.x = v ? (int)(r.pos & 0xffff) : 0; s.y = v ? (int)(r.pos >> 16) : 0;
    s.dir = v ? r.dir : 0.0;
    s.hp = v ? r.hp : 0;
    s.ebx = v ? (int)(r.eb & 0xffff) : 64; s.eby = v ? (int)(r.eb >> 16) : 64;
    s.bhp_own = v ? r.bo : 0; s.bhp_enemy = v ? r.be : 0;
    s.tick = v ? (int)(r.m & 0xfff) : 0; s.env_done = v ? (int)((r.m >> 16) & 0xff) : 1; s.winner = v ? (int)((r.m >> 24) & 0xff) : 0;
    s.episode = v ? r.ep : 0u;
    const bool first = v && g.i() == 0;
    ret0 = first ? r.r0 : 0.0; ret1 = first ? r.r1 : 0.0;
}

template <int N>
__device__ __forceinline__ void load_lane(const Params &p, const Geo<N> &g, Lane<N> &s, double &ret0, double &ret1) {
    constexpr int P = 2 * N;
    s.x = s.y = 0; s.hp = 0; s.dir = 0; s.ebx = s.eby = 64; s.bhp_own = s.bhp_enemy = 0;
    s.tick = 0; s.env_done = 1; s.winner = 0; s.episode = 0; ret0 = ret1 = 0;
    if (g.valid()) {
        const long long a = g.e() * P + g.i();
        uint32_t pos = p.plane_pos[a];
        s.x = pos & 0xffff; s.y = pos >> 16;
        s.dir = p.plane_dir[a];
        s.hp = p.plane_hp[a];
        uint32_t eb = p.base_pos[g.e() * 2 + (1 - g.team())];
        s.ebx = eb & 0xffff; s.eby = eb >> 16;
        s.bhp_own = p.base_hp[g.e() * 2 + g.team()];
        s.bhp_enemy = p.base_hp[g.e() * 2 + (1 - g.team())];
        uint32_t m = p.meta[g.e()];
        s.tick = m & 0xfff; s.env_done = (m >> 16) & 0xff; s.winner = (m >> 24) & 0xff;
        s.episode = p.episode[g.e()];
        if (g.i() == 0) { ret0 = p.ret[g.e() * 2]; ret1 = p.ret[g.e() * 2 + 1]; }
    }
}

template <int N>
__device__ __forceinline__ void store_lane(const Params &p, const Geo<N> &g, const Lane<N> &s, double ret0, double ret1,
                                           bool base_pos_too) {
    constexpr int P = 2 * N;
    if (!g.valid()) return;
    const long long a = g.e() * P + g.i();
    p.plane_pos[a] = (uint32_t)s.x | ((uint32_t)s.y << 16);
    p.plane_dir[a] = s.dir;
    p.plane_hp[a] = (int8_t)s.hp;
    if (g.k() == 0) {  // first lane of each team owns its enemy's base record
        p.base_hp[g.e() * 2 + (1 - g.team())] = (int16_t)s.bhp_enemy;
        if (base_pos_too) p.base_pos[g.e() * 2 + (1 - g.team())] = (uint32_t)s.ebx | ((uint32_t)s.eby << 16);
    }
    if (g.i() == 0) {
        p.meta[g.e()] = (uint32_t)s.tick | ((uint32_t)s.env_done << 16) | ((uint32_t)s.winner << 24);
        p.episode[g.e()] = s.episode;
        p.ret[g.e() * 2] = ret0; p.ret[g.e() * 2 + 1] = ret1;
    }
}

// Launch-level flush of the tick counters: one atomic per warp per counter.
template <int N>
__device__ __forceinline__ void flush_tally(const Params &p, const Tally &tl) {
    const long long t = warp_sum((long long)tl.ticks);
    if ((threadIdx.x & 31) == 0 && t != 0) {
        atomicAdd(p.counters + C_TICKS, (unsigned long long)t);
        atomicAdd(p.counters + C_AGENT_STEPS, (unsigned long long)(t * (2 * N)));
    }
}

}  // namespace bs
