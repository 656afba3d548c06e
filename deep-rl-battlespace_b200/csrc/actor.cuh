// actor.cuh -- batched forward of the MADDPG actors, hand-written SIMT fp32 for sm_100a.
//
// Reference: ActorNetwork.forward, maddpg/networks.py:54-85 --
//     x = relu(LayerNorm(fc1(obs)));  x = relu(LayerNorm(fc2(x)));  a = tanh(pi(x))
// and NetworkedAgent.choose_action, maddpg/agent.py:25-33 -- a += OUNoise.noise(); clamp(-1, 1)
// (utils/noise.py:17-21: state += theta*(mu - state) + sigma*randn; noise = state*scale).
// Every plane of a team has its OWN actor (maddpg/team.py:17-19), so the launch is a grid of
// (row tiles) x (team slot k): block (x, k) keeps actor k's 5.4k parameters in shared memory
// (transposed: [in][out]) and walks tiles of 128 observation rows through the three layers.
// fc1/fc2 width is 64 (params.json; main.py).
//
// Why SIMT and not tcgen05: the check for this kernel is torch's fp32 forward (allow_tf32 is off by
// default) at 1e-5; K = 11..64 per layer and a LayerNorm between layers leave nothing a bf16/tf32 MMA
// could do within that tolerance without a 3-pass split.  The bound is FFMA issue: 4,992 FMA per row.
//
// fc1 and fc2 are register-tiled SIMT GEMMs: each of the 128 threads owns an 8-row x 8-column block
// of the tile's [128 x 64] output (64 independent accumulators); activations sit in shared memory as
// [k][row] and weights as [k][col], so one k-step is 4 LDS.128 (2 broadcast activation quads, 2
// weight quads, all conflict free) for 64 FMA -- FMA:LDS = 16:1.  The first version of this file ran
// one row per thread with broadcast weight loads (4:1) and stalled on shared-memory latency at 27 %
// of the FMA peak (profiles/r1h_actor_*).  LayerNorm reduces a row over the 8 threads that share it
// with three xor-shuffles (two-pass mean / variance like torch); the 64 -> A head and the noise /
// tanh / clamp epilogue run one row per thread from the staged activations.
#pragma once
#include "tick.cuh"

namespace bs {

constexpr int AH = 64;          // fc1_dims == fc2_dims == 64
constexpr int ACTOR_BLOCK = 128;
constexpr float LN_EPS = 1e-5f; // nn.LayerNorm default
enum { PURPOSE_OU_NOISE = 3 };

struct ActorNet {  // nn.Linear weight = [out, in] row-major, float32, device pointers
    const float *fc1_w, *fc1_b, *ln1_g, *ln1_b, *fc2_w, *fc2_b, *ln2_g, *ln2_b, *pi_w, *pi_b;
};

struct ActorArgs {
    ActorNet net[8];
    long long E, env_offset;
    const float *obs;            // row (e, k) at obs + e*obs_env_stride + (obs_agent0 + k)*L
    long long obs_env_stride;
    int obs_agent0, n_team;
    float *out;                  // [E, n_team, A]
    // OU noise (all optional: state == nullptr -> none)
    float *ou_state;             // [E, n_team, A]
    const float *normals;        // [E, n_team, A] injected randn draws, or nullptr -> Philox + Box-Muller
    const uint8_t *reset_mask;   // [E] != 0: OUNoise.reset() before this draw (agent.reset_noise, main.py:150)
    float mu, theta, sigma, scale;
    uint32_t seed_lo, seed_hi, step;
};

}  // namespace bs

// launches actor_kernel<3N+2, cont ? 3 : 4> (defined in actor.cu)
cudaError_t bs_launch_actor_impl(int n_agents, bool cont, const bs::ActorArgs &aa, cudaStream_t st, bool pdl = false);

#ifdef BS_ACTOR_KERNELS
namespace bs {

constexpr int TM = 128;      // rows per tile == threads per block
constexpr int WP = AH + 4;   // row stride of the transposed weights [in][out]: the transposing store at block start is then
                             // 4-way instead of 32-way bank conflicted; quads stay 16-byte aligned
constexpr int HP = TM + 4;   // row stride of the [k][row] hidden buffer: the 8 column groups of a warp then
                             // store their quads to distinct banks (4 wavefronts per STS.128, the minimum)

template <int L, int A>
struct ActorSmem {
    static constexpr int OS = 0, H = OS + L * TM, W1T = H + AH * HP, W2T = W1T + L * WP, W3T = W2T + AH * WP,
                         B1 = W3T + AH * 4, G1 = B1 + AH, E1 = G1 + AH, B2 = E1 + AH, G2 = B2 + AH, E2 = G2 + AH,
                         B3 = E2 + AH, END = B3 + 4;
    static constexpr size_t bytes() { return (size_t)END * sizeof(float); }
};

__device__ __forceinline__ void coop_copy(float *dst, const float *src, int n) {
    for (int i = threadIdx.x; i < n; i += ACTOR_BLOCK) dst[i] = __ldg(src + i);
}

__device__ __forceinline__ float4 lds4(const float *p) { return *reinterpret_cast<const float4 *>(p); }

// The 8 x 8 block is held as 8 x 4 float2 (column pairs) and updated with the packed fp32 FMA of sm_100
// (FFMA2: two IEEE fp32 FMAs per lane and instruction, one operand may be a broadcast scalar): the products are
// the same as with scalar fmaf, in the same order, but a k-step issues 32 instructions instead of 64.  The
// kernel was bound by instruction issue with FFMA at 69 % of the instructions (profiles/r2a_actor_raw.txt: FMA
// pipe 49 % busy); packed, the FMA pipe itself becomes the limiter.
// acc[r][2q..2q+1] += a[r] * w[2q..2q+1]
__device__ __forceinline__ void outer8x8(float2 (&acc)[8][4], const float4 &alo, const float4 &ahi, const float4 &wlo,
                                         const float4 &whi) {
    const float a[8] = {alo.x, alo.y, alo.z, alo.w, ahi.x, ahi.y, ahi.z, ahi.w};
    const float2 w[4] = {make_float2(wlo.x, wlo.y), make_float2(wlo.z, wlo.w), make_float2(whi.x, whi.y), make_float2(whi.z, whi.w)};
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[r][q] = __ffma2_rn(make_float2(a[r], a[r]), w[q], acc[r][q]);
}

__device__ __forceinline__ void init_bias(float2 (&acc)[8][4], const float *b, int tc) {
    const float4 lo = lds4(b + tc * 4), hi = lds4(b + 32 + tc * 4);
    const float2 w[4] = {make_float2(lo.x, lo.y), make_float2(lo.z, lo.w), make_float2(hi.x, hi.y), make_float2(hi.z, hi.w)};
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[r][q] = w[q];
}

// relu(LayerNorm(row)) for the 8 rows of the thread's block, then staged as H[col][row]
__device__ __forceinline__ void ln_relu_store(float2 (&acc)[8][4], const float *g, const float *b, float *H, int tc, int rq) {
    const float4 glo = lds4(g + tc * 4), ghi = lds4(g + 32 + tc * 4), blo = lds4(b + tc * 4), bhi = lds4(b + 32 + tc * 4);
    const float2 gg[4] = {make_float2(glo.x, glo.y), make_float2(glo.z, glo.w), make_float2(ghi.x, ghi.y), make_float2(ghi.z, ghi.w)};
    const float2 bb[4] = {make_float2(blo.x, blo.y), make_float2(blo.z, blo.w), make_float2(bhi.x, bhi.y), make_float2(bhi.z, bhi.w)};
    float mean[8], rstd[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        float s = 0.0f;   // (left to right, like the scalar version: the completed_model check sits within 2 % of its bound)
#pragma unroll
        for (int q = 0; q < 4; ++q) { s += acc[r][q].x; s += acc[r][q].y; }
        mean[r] = s;
    }
#pragma unroll
    for (int m = 1; m < 8; m <<= 1)   // the 8 rows' reductions travel together: 8 shuffles in flight per stage
#pragma unroll
        for (int r = 0; r < 8; ++r) mean[r] += __shfl_xor_sync(FULL, mean[r], m);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        mean[r] *= 1.0f / AH;
        const float2 nm = make_float2(-mean[r], -mean[r]);
        float v = 0.0f;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            acc[r][q] = __fadd2_rn(acc[r][q], nm);                                                      // acc <- x - mean
            v = fmaf(acc[r][q].x, acc[r][q].x, v); v = fmaf(acc[r][q].y, acc[r][q].y, v);
        }
        rstd[r] = v;
    }
#pragma unroll
    for (int m = 1; m < 8; m <<= 1)
#pragma unroll
        for (int r = 0; r < 8; ++r) rstd[r] += __shfl_xor_sync(FULL, rstd[r], m);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        // 1 / sqrt(var + eps): MUFU.RSQ (2 ulp) + one Newton step = the correctly rounded value to within 1 ulp in 5
        // instructions; the IEEE sqrt + division pair it replaces was 7 % of the kernel's instructions (r2e).  The
        // raw rsqrtf alone is not enough for the completed_model check (logits up to +-80 amplify the error).
        const float var = rstd[r] * (1.0f / AH) + LN_EPS;
        const float y0 = rsqrtf(var);
        const float rs = fmaf(y0, fmaf(-0.5f * var, y0 * y0, 0.5f), y0);
        const float2 rs2 = make_float2(rs, rs);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float2 y = __ffma2_rn(__fmul2_rn(acc[r][q], rs2), gg[q], bb[q]);
            acc[r][q] = make_float2(fmaxf(y.x, 0.0f), fmaxf(y.y, 0.0f));
        }
    }
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const int col = (c < 4 ? 0 : 28) + tc * 4 + c;  // columns tc*4..+3 and 32+tc*4..+3
        const int q = c >> 1;
#define ACC_AT(r) ((c & 1) ? acc[r][q].y : acc[r][q].x)
        *reinterpret_cast<float4 *>(H + col * HP + rq) = make_float4(ACC_AT(0), ACC_AT(1), ACC_AT(2), ACC_AT(3));
        *reinterpret_cast<float4 *>(H + col * HP + rq + 16) = make_float4(ACC_AT(4), ACC_AT(5), ACC_AT(6), ACC_AT(7));
#undef ACC_AT
    }
}

template <int L, int A>
__global__ void __launch_bounds__(ACTOR_BLOCK, 3) actor_kernel(ActorArgs aa) {
    using S = ActorSmem<L, A>;
    extern __shared__ float4 smem4[];
    float *sm = reinterpret_cast<float *>(smem4);
    const int tid = threadIdx.x, k = blockIdx.y;
    // Thread tile: column group tc (columns tc*4..+3 and 32+tc*4..+3) x row group: two quads of rows.  Each WARP
    // owns 32 consecutive rows of the tile (rows rq..rq+3 and rq+16..rq+19 per thread), so the activations it
    // stages between the layers are private to it and the tile loop needs no block barrier, only __syncwarp():
    // the four warps of a block drift apart and hide each other's LayerNorm / head phases (r1j: 4 __syncthreads
    // per tile at 12 resident warps left the issue slots 57 % busy).
    const int tc = tid & 7;
    const int rq = (tid & ~31) + ((tid >> 3) & 3) * 4;   // first row of the thread's first quad; second quad at rq + 16
    const ActorNet net = aa.net[k];

    // ---- actor k's parameters -> shared memory, transposed to [in][out] --------------------------
    // (coalesced global reads, all issued before the first store: one memory latency, not one per element)
    {
        constexpr int N1 = (L * AH + ACTOR_BLOCK - 1) / ACTOR_BLOCK, N2 = AH * AH / ACTOR_BLOCK;
        float v1[N1], v2[N2];
#pragma unroll
        for (int q = 0; q < N1; ++q) { const int i = tid + q * ACTOR_BLOCK; v1[q] = i < L * AH ? __ldg(net.fc1_w + i) : 0.0f; }
#pragma unroll
        for (int q = 0; q < N2; ++q) v2[q] = __ldg(net.fc2_w + tid + q * ACTOR_BLOCK);
#pragma unroll
        for (int q = 0; q < N1; ++q) { const int i = tid + q * ACTOR_BLOCK, j = i / L, c = i - j * L; if (i < L * AH) sm[S::W1T + c * WP + j] = v1[q]; }
#pragma unroll
        for (int q = 0; q < N2; ++q) { const int i = tid + q * ACTOR_BLOCK, j = i / AH, c = i - j * AH; sm[S::W2T + c * WP + j] = v2[q]; }
    }
    for (int i = tid; i < AH * 4; i += ACTOR_BLOCK) { const int c = i >> 2, a = i & 3; sm[S::W3T + i] = a < A ? __ldg(net.pi_w + a * AH + c) : 0.0f; }
    coop_copy(sm + S::B1, net.fc1_b, AH); coop_copy(sm + S::G1, net.ln1_g, AH); coop_copy(sm + S::E1, net.ln1_b, AH);
    coop_copy(sm + S::B2, net.fc2_b, AH); coop_copy(sm + S::G2, net.ln2_g, AH); coop_copy(sm + S::E2, net.ln2_b, AH);
    if (tid < 4) sm[S::B3 + tid] = tid < A ? __ldg(net.pi_b + tid) : 0.0f;
    float *Os = sm + S::OS, *H = sm + S::H;
    const float *arow = H + rq, *wcol1 = sm + S::W1T + tc * 4, *wcol2 = sm + S::W2T + tc * 4;

    const long long tiles = (aa.E + TM - 1) / TM;
    // The next tile's observation row is fetched into registers while the current tile computes.
    float onext[L];
    auto fetch = [&](long long tile) {
        const long long e = tile * TM + tid;
        const bool ok = tile < tiles && e < aa.E;
        const float *src = aa.obs + (ok ? e : 0) * aa.obs_env_stride + (long long)(aa.obs_agent0 + k) * L;
#pragma unroll
        for (int c = 0; c < L; ++c) onext[c] = ok ? __ldg(src + c) : 0.0f;
    };
    // programmatic dependent launch (bs_rollout_actor): the parameters above were staged while the
    // previous grid (the env step that writes the observations) was still draining
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;");
    fetch(blockIdx.x);
    __syncthreads();      // parameters staged (the only block barrier)
    // (Measured and not kept, r2f: 32-row units handed out warp by warp from a device counter, every block resident for
    // the whole launch -- the kernel alone 36.9 instead of 38.2 us, but the config-5 tick 65.9 instead of 61.6 us: blocks
    // that never exit leave no room for the next grid's blocks to start under programmatic dependent launch.)
    // (Measured and not kept, r2v: a static, balanced split in 32-row units -- every resident block slot gets a block and
    // the same number of units to within one, instead of 384 blocks of two tiles on 444 slots.  The kernel alone stayed
    // at 34.8 us: it is latency-bound per warp (3 warps per scheduler, issue slots 41 % busy, FMA pipe 53 %, LSU data
    // pipe 56 %), so its time is (units per warp) x (time per unit) = 2 units either way; the config-5 tick went from
    // 61.6 to 67.2 us because the step kernel's blocks no longer find SMs with a free slot early.)
    for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const long long e = tile * TM + tid;
        const bool ok = e < aa.E;
#pragma unroll
        for (int c = 0; c < L; ++c) Os[c * TM + tid] = onext[c];
        __syncwarp();     // the warp's 32 obs rows are staged; its H rows of the previous tile are free
        fetch(tile + gridDim.x);
        float2 acc[8][4];
        // ---- fc1 -----------------------------------------------------------------------------------
        init_bias(acc, sm + S::B1, tc);
#pragma unroll
        for (int c = 0; c < L; ++c)
            outer8x8(acc, lds4(Os + c * TM + rq), lds4(Os + c * TM + rq + 16), lds4(wcol1 + c * WP), lds4(wcol1 + c * WP + 32));
        ln_relu_store(acc, sm + S::G1, sm + S::E1, H, tc, rq);
        __syncwarp();
        // ---- fc2 -----------------------------------------------------------------------------------
        init_bias(acc, sm + S::B2, tc);
#pragma unroll 8
        for (int c = 0; c < AH; ++c)
            outer8x8(acc, lds4(arow + c * HP), lds4(arow + c * HP + 16), lds4(wcol2 + c * WP), lds4(wcol2 + c * WP + 32));
        __syncwarp();     // every lane has read h1 before it is overwritten by h2
        ln_relu_store(acc, sm + S::G2, sm + S::E2, H, tc, rq);
        __syncwarp();
        // ---- pi + tanh (+ OU noise, clamp): one row per thread -------------------------------------
        float2 p01 = make_float2(sm[S::B3 + 0], sm[S::B3 + 1]), p23 = make_float2(sm[S::B3 + 2], sm[S::B3 + 3]);
#pragma unroll 16
        for (int j = 0; j < AH; ++j) {
            const float x = H[j * HP + tid];
            const float4 w = lds4(sm + S::W3T + j * 4);   // (column 3 of a 3-action head is zero)
            p01 = __ffma2_rn(make_float2(x, x), make_float2(w.x, w.y), p01);
            p23 = __ffma2_rn(make_float2(x, x), make_float2(w.z, w.w), p23);
        }
        const float pi[4] = {p01.x, p01.y, p23.x, p23.y};
        if (ok) {
            const long long orow = (e * aa.n_team + k) * A;
            float nz[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            if (aa.ou_state != nullptr) {
                if (aa.normals != nullptr) {
#pragma unroll
                    for (int a = 0; a < A; ++a) nz[a] = aa.normals[orow + a];
                } else {  // Philox words -> Box-Muller pairs
                    const unsigned long long g = (unsigned long long)(aa.env_offset + e);
                    const uint4 w = philox4x32_10((uint32_t)g, (uint32_t)(g >> 32), aa.step,
                                                  ((uint32_t)k << 8) | PURPOSE_OU_NOISE, aa.seed_lo, aa.seed_hi);
                    const float u0 = ((w.x >> 8) + 1) * (1.0f / 16777216.0f), u1 = (w.y >> 8) * (1.0f / 16777216.0f);
                    const float u2 = ((w.z >> 8) + 1) * (1.0f / 16777216.0f), u3 = (w.w >> 8) * (1.0f / 16777216.0f);
                    const float r0 = sqrtf(-2.0f * logf(u0)), r1 = sqrtf(-2.0f * logf(u2));
                    float s0, c0, s1, c1;
                    sincospif(2.0f * u1, &s0, &c0); sincospif(2.0f * u3, &s1, &c1);
                    nz[0] = r0 * c0; nz[1] = r0 * s0; nz[2] = r1 * c1; nz[3] = r1 * s1;
                }
                const bool rst = aa.reset_mask != nullptr && aa.reset_mask[e] != 0;
#pragma unroll
                for (int a = 0; a < A; ++a) {  // utils/noise.py:17-21
                    const float x = rst ? aa.mu : aa.ou_state[orow + a];
                    const float st = x + (aa.theta * (aa.mu - x) + aa.sigma * nz[a]);
                    aa.ou_state[orow + a] = st;
                    nz[a] = st * aa.scale;
                }
            }
#pragma unroll
            for (int a = 0; a < A; ++a) {
                float act = tanhf(pi[a]);
                if (aa.ou_state != nullptr) act = fminf(fmaxf(act + nz[a], -1.0f), 1.0f);  // agent.py:31-32
                aa.out[orow + a] = act;
            }
        }
    }
}

}  // namespace bs
#endif
