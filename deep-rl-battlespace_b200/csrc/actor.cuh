// actor.cuh -- batched forward of the MADDPG actors, hand-written SIMT fp32 for sm_100a.
//
// Reference: ActorNetwork.forward, maddpg/networks.py:54-85 --
//     x = relu(LayerNorm(fc1(obs)));  x = relu(LayerNorm(fc2(x)));  a = tanh(pi(x))
// and NetworkedAgent.choose_action, maddpg/agent.py:25-33 -- a += OUNoise.noise(); clamp(-1, 1)
// (utils/noise.py:17-21: state += theta*(mu - state) + sigma*randn; noise = state*scale).
// Every plane of a team has its OWN actor (maddpg/team.py:17-19), so the launch is a grid of
// (row tiles) x (team slot k): block (x, k) keeps actor k's 5.4k parameters in shared memory and every
// thread carries R observation rows through the three layers on its own -- no inter-thread traffic,
// LayerNorm is a per-thread reduction over registers.  fc1/fc2 width is 64 (params.json; main.py).
//
// Why SIMT and not tcgen05: the check for this kernel is torch's fp32 forward (allow_tf32 is off by
// default) at 1e-5; K = 11..64 per layer and a LayerNorm between layers leave nothing a bf16/tf32 MMA
// could do within that tolerance without a 3-pass split.  The bound is FFMA issue: 4,992 FMA per row.
//
// Weights are read as broadcast LDS.128 (all lanes of a warp read the same address: one wavefront),
// 4 weights per load, each used by R rows: FMA:LDS = 4R:1.  The second hidden layer is produced by a
// ROLLED loop over output neurons (4 at a time) and staged per thread in shared memory (column
// [j][thread]: conflict free), because indexing a register array dynamically is not possible and
// unrolling 4,096 FMAs would not fit the instruction cache.
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

// launches actor_kernel<3N+2, cont ? 3 : 4, rows> (defined in actor.cu)
cudaError_t bs_launch_actor_impl(int n_agents, bool cont, int rows, const bs::ActorArgs &aa, cudaStream_t st);

#ifdef BS_ACTOR_KERNELS
namespace bs {

template <int L, int A>
struct ActorSmem {
    static constexpr int LP = (L + 3) & ~3;
    static constexpr int W1 = 0, W2 = W1 + AH * LP, W3 = W2 + AH * AH, B1 = W3 + A * AH, G1 = B1 + AH, E1 = G1 + AH,
                         B2 = E1 + AH, G2 = B2 + AH, E2 = G2 + AH, B3 = E2 + AH, HS = B3 + 4;
    static constexpr size_t bytes(int R) { return (size_t)(HS + AH * ACTOR_BLOCK * R) * sizeof(float); }
};

__device__ __forceinline__ void coop_copy(float *dst, const float *src, int n) {
    for (int i = threadIdx.x; i < n; i += ACTOR_BLOCK) dst[i] = __ldg(src + i);
}

template <int L, int A, int R>
__global__ void __launch_bounds__(ACTOR_BLOCK) actor_kernel(ActorArgs aa) {
    using S = ActorSmem<L, A>;
    constexpr int LP = S::LP;
    extern __shared__ float4 smem4[];
    float *sm = reinterpret_cast<float *>(smem4);
    const int tid = threadIdx.x, k = blockIdx.y;
    const ActorNet net = aa.net[k];

    // ---- actor k's parameters -> shared memory (fc1 rows padded to LP with zeros) ---------------
    for (int i = tid; i < AH * LP; i += ACTOR_BLOCK) {
        const int j = i / LP, c = i - j * LP;
        sm[S::W1 + i] = c < L ? __ldg(net.fc1_w + j * L + c) : 0.0f;
    }
    coop_copy(sm + S::W2, net.fc2_w, AH * AH);
    coop_copy(sm + S::W3, net.pi_w, A * AH);
    coop_copy(sm + S::B1, net.fc1_b, AH); coop_copy(sm + S::G1, net.ln1_g, AH); coop_copy(sm + S::E1, net.ln1_b, AH);
    coop_copy(sm + S::B2, net.fc2_b, AH); coop_copy(sm + S::G2, net.ln2_g, AH); coop_copy(sm + S::E2, net.ln2_b, AH);
    if (tid < 4) sm[S::B3 + tid] = tid < A ? __ldg(net.pi_b + tid) : 0.0f;
    __syncthreads();
    const float4 *w1 = reinterpret_cast<const float4 *>(sm + S::W1);
    const float4 *w2 = reinterpret_cast<const float4 *>(sm + S::W2);
    float *hs = sm + S::HS;  // [AH][ACTOR_BLOCK * R]
    constexpr int HST = ACTOR_BLOCK * R;

    for (long long base = (long long)blockIdx.x * HST; base < aa.E; base += (long long)gridDim.x * HST) {
        long long e[R];
        float o[R][LP];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            e[r] = base + r * ACTOR_BLOCK + tid;
            const bool ok = e[r] < aa.E;
            const float *src = aa.obs + (ok ? e[r] : 0) * aa.obs_env_stride + (long long)(aa.obs_agent0 + k) * L;
#pragma unroll
            for (int c = 0; c < LP; ++c) o[r][c] = (ok && c < L) ? __ldg(src + c) : 0.0f;
        }
        // ---- fc1 + LayerNorm + ReLU, fully in registers --------------------------------------
        float h[R][AH];
#pragma unroll
        for (int j = 0; j < AH; ++j) {
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = sm[S::B1 + j];
#pragma unroll
            for (int c4 = 0; c4 < LP / 4; ++c4) {
                const float4 w = w1[j * (LP / 4) + c4];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    acc[r] = fmaf(w.x, o[r][4 * c4 + 0], acc[r]); acc[r] = fmaf(w.y, o[r][4 * c4 + 1], acc[r]);
                    acc[r] = fmaf(w.z, o[r][4 * c4 + 2], acc[r]); acc[r] = fmaf(w.w, o[r][4 * c4 + 3], acc[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) h[r][j] = acc[r];
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            float s = 0.0f;
#pragma unroll
            for (int j = 0; j < AH; ++j) s += h[r][j];
            const float mean = s * (1.0f / AH);
            float v = 0.0f;
#pragma unroll
            for (int j = 0; j < AH; ++j) { const float d = h[r][j] - mean; v = fmaf(d, d, v); }
            const float rstd = 1.0f / sqrtf(v * (1.0f / AH) + LN_EPS);
#pragma unroll
            for (int j = 0; j < AH; ++j)
                h[r][j] = fmaxf(fmaf((h[r][j] - mean) * rstd, sm[S::G1 + j], sm[S::E1 + j]), 0.0f);
        }
        // ---- fc2: rolled over output neurons, 4 at a time; pre-norm values staged in smem ------
        float s2[R];
#pragma unroll
        for (int r = 0; r < R; ++r) s2[r] = 0.0f;
#pragma unroll 1
        for (int j = 0; j < AH; j += 4) {
            float acc[4][R];
#pragma unroll
            for (int jj = 0; jj < 4; ++jj)
#pragma unroll
                for (int r = 0; r < R; ++r) acc[jj][r] = sm[S::B2 + j + jj];
#pragma unroll
            for (int i4 = 0; i4 < AH / 4; ++i4) {
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const float4 w = w2[(j + jj) * (AH / 4) + i4];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        acc[jj][r] = fmaf(w.x, h[r][4 * i4 + 0], acc[jj][r]); acc[jj][r] = fmaf(w.y, h[r][4 * i4 + 1], acc[jj][r]);
                        acc[jj][r] = fmaf(w.z, h[r][4 * i4 + 2], acc[jj][r]); acc[jj][r] = fmaf(w.w, h[r][4 * i4 + 3], acc[jj][r]);
                    }
                }
            }
#pragma unroll
            for (int jj = 0; jj < 4; ++jj)
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    hs[(j + jj) * HST + r * ACTOR_BLOCK + tid] = acc[jj][r];
                    s2[r] += acc[jj][r];
                }
        }
        // ---- LayerNorm + ReLU + pi + tanh (+ OU noise, clamp) ---------------------------------
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float mean = s2[r] * (1.0f / AH);
            const float *col = hs + r * ACTOR_BLOCK + tid;
            float v = 0.0f;
#pragma unroll 8
            for (int j = 0; j < AH; ++j) { const float d = col[j * HST] - mean; v = fmaf(d, d, v); }
            const float rstd = 1.0f / sqrtf(v * (1.0f / AH) + LN_EPS);
            float pi[A];
#pragma unroll
            for (int a = 0; a < A; ++a) pi[a] = sm[S::B3 + a];
#pragma unroll 8
            for (int j = 0; j < AH; ++j) {
                const float x = fmaxf(fmaf((col[j * HST] - mean) * rstd, sm[S::G2 + j], sm[S::E2 + j]), 0.0f);
#pragma unroll
                for (int a = 0; a < A; ++a) pi[a] = fmaf(sm[S::W3 + a * AH + j], x, pi[a]);
            }
            if (e[r] < aa.E) {
                const long long orow = (e[r] * aa.n_team + k) * A;
                float nz[4] = {0.0f, 0.0f, 0.0f, 0.0f};
                if (aa.ou_state != nullptr) {
                    if (aa.normals != nullptr) {
#pragma unroll
                        for (int a = 0; a < A; ++a) nz[a] = aa.normals[orow + a];
                    } else {  // Philox words -> Box-Muller pairs
                        const unsigned long long g = (unsigned long long)(aa.env_offset + e[r]);
                        const uint4 w = philox4x32_10((uint32_t)g, (uint32_t)(g >> 32), aa.step,
                                                      ((uint32_t)k << 8) | PURPOSE_OU_NOISE, aa.seed_lo, aa.seed_hi);
                        const float u0 = ((w.x >> 8) + 1) * (1.0f / 16777216.0f), u1 = (w.y >> 8) * (1.0f / 16777216.0f);
                        const float u2 = ((w.z >> 8) + 1) * (1.0f / 16777216.0f), u3 = (w.w >> 8) * (1.0f / 16777216.0f);
                        const float r0 = sqrtf(-2.0f * logf(u0)), r1 = sqrtf(-2.0f * logf(u2));
                        float s0, c0, s1, c1;
                        sincospif(2.0f * u1, &s0, &c0); sincospif(2.0f * u3, &s1, &c1);
                        nz[0] = r0 * c0; nz[1] = r0 * s0; nz[2] = r1 * c1; nz[3] = r1 * s1;
                    }
                    const bool rst = aa.reset_mask != nullptr && aa.reset_mask[e[r]] != 0;
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
}

}  // namespace bs
#endif
