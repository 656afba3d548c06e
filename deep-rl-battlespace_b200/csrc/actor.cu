// actor.cu -- instantiations + launcher of the batched actor forward (actor.cuh); separate translation
// units so that they build in parallel with battlespace.cu.
#define BS_ACTOR_KERNELS
#include "actor.cuh"

using namespace bs;

template <int N, int A>
static cudaError_t launch_actor_t(const ActorArgs &aa, cudaStream_t st, bool pdl) {
    constexpr int L = 3 * N + 2;
    const size_t smem = ActorSmem<L, A>::bytes();
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_set[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(actor_kernel<L, A>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        attr_set[dev & 63] = true;
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long tiles = (aa.E + TM - 1) / TM;
    // 3 blocks are resident per SM; give every block the same number of tiles (no ragged second pass)
    long long slots = ((long long)sms * 3 + aa.n_team - 1) / aa.n_team;
    const long long passes = (tiles + slots - 1) / slots;
    long long gx = (tiles + passes - 1) / passes;
    if (gx < 1) gx = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)gx, (unsigned)aa.n_team); cfg.blockDim = dim3(ACTOR_BLOCK); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    if (pdl) {
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
    }
    return cudaLaunchKernelEx(&cfg, actor_kernel<L, A>, aa);
}

// One object per team size (make builds them in parallel): -DBS_ACTOR_N=n compiles the kernels for n
// planes per team; the dispatcher is compiled once, with -DBS_ACTOR_DISPATCH.
#ifdef BS_ACTOR_N
#define BS_CAT2(a, b) a##b
#define BS_CAT(a, b) BS_CAT2(a, b)
cudaError_t BS_CAT(bs_launch_actor_n, BS_ACTOR_N)(bool cont, const ActorArgs &aa, cudaStream_t st, bool pdl) {
    constexpr int N = BS_ACTOR_N;
    return cont ? launch_actor_t<N, 3>(aa, st, pdl) : launch_actor_t<N, 4>(aa, st, pdl);
}
#endif

#ifdef BS_ACTOR_DISPATCH
#define BS_DECL(n) cudaError_t bs_launch_actor_n##n(bool, const ActorArgs &, cudaStream_t, bool);
BS_DECL(1) BS_DECL(2) BS_DECL(3) BS_DECL(4) BS_DECL(5) BS_DECL(6) BS_DECL(7) BS_DECL(8)
cudaError_t bs_launch_actor_impl(int n_agents, bool cont, const ActorArgs &aa, cudaStream_t st, bool pdl) {
    switch (n_agents) {
        case 1: return bs_launch_actor_n1(cont, aa, st, pdl);
        case 2: return bs_launch_actor_n2(cont, aa, st, pdl);
        case 3: return bs_launch_actor_n3(cont, aa, st, pdl);
        case 4: return bs_launch_actor_n4(cont, aa, st, pdl);
        case 5: return bs_launch_actor_n5(cont, aa, st, pdl);
        case 6: return bs_launch_actor_n6(cont, aa, st, pdl);
        case 7: return bs_launch_actor_n7(cont, aa, st, pdl);
        case 8: return bs_launch_actor_n8(cont, aa, st, pdl);
    }
    return cudaErrorInvalidValue;
}
#endif
