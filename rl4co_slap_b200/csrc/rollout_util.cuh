// PTX / RNG / reward helpers shared by the two rollout kernels (am_rollout.cu: shared-memory
// resident tables, any number of starts; am_rollout_rf.cu: register-file resident tables,
// single start).
#pragma once
#include <math.h>

#include "common.cuh"

namespace rl4co {

// ---------------------------------------------------------------- PTX helpers ------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// TMA 1-D bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(void *dst, const void *src, unsigned bytes,
                                            uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

// ---------------------------------------------------------------- Philox4x32-10 ----
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0;
        key.y += W1;
    }
    return ctr;
}
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint64_t offset, uint64_t r, uint32_t t) {
    uint4 c = make_uint4((uint32_t)(offset + t), (uint32_t)((offset + t) >> 32), (uint32_t)r, (uint32_t)(r >> 32));
    uint4 o = philox4x32_10(c, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    return (float)(o.x >> 8) * (1.0f / 16777216.0f);   // [0,1)
}

// Gumbel-max sampling (register-file kernels): argmax_n (x_n + G_n), G_n = -ln(-ln u_n) i.i.d., picks n with
// probability softmax(x)_n - the Categorical(exp(logp)) draw of utils/decoding.py:381-397 without the prefix
// scan over the probabilities.  The noise depends on (seed, rollout, step, lane) only, so it is drawn before
// the step's logits exist and the selection itself is the greedy arg-max on x + G.  One Philox4x32-10 call
// gives a lane the four uniforms of its four nodes; every warp draws the same noise, so all warps agree.
__device__ __forceinline__ void gumbel4(uint64_t seed, uint64_t offset, uint64_t r, uint32_t t, int lane, float (&G)[4]) {
    const uint4 c = make_uint4((uint32_t)(offset + t), (uint32_t)((offset + t) >> 32), (uint32_t)r,
                               (uint32_t)(r >> 32) ^ ((uint32_t)(lane + 1) << 24));
    const uint4 o = philox4x32_10(c, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const uint32_t w[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        // 23 random bits: (k + 0.5) * 2^-23 is exact in fp32 for every k < 2^23 and strictly inside (0,1)
        // (with 24 bits the +0.5 is lost above 2^23 and k = 2^24 - 1 rounds to u = 1 -> G = +inf)
        const float u = ((float)(w[k] >> 9) + 0.5f) * (1.0f / 8388608.0f);
        // -ln(-ln u) = -ln2 * log2(-ln2 * log2 u)
        G[k] = -0.6931471805599453f * __log2f(-0.6931471805599453f * __log2f(u));
    }
}

// closed tour through the K storage locations of one order (slap/env.py:136-142)
__device__ __forceinline__ float slap_order_len(const float2 *locs, const int *assign,
                                                const int64_t *order, int K) {
    float2 first = locs[assign[(int)order[0]]], prev = first;
    float len = 0.f;
    for (int k = 1; k < K; ++k) {
        float2 cur = locs[assign[(int)order[k]]];
        len += leg(prev.x, prev.y, cur.x, cur.y);
        prev = cur;
    }
    return len + leg(prev.x, prev.y, first.x, first.y);
}


// host-side launchers of the register-file resident single-start kernel (am_rollout_rf.cu)
int launch_rollout_rf(const rl4co_am_rollout_args &a, cudaStream_t stream);
bool rollout_rf_supports(const rl4co_am_rollout_args &a);
// TSP / greedy / standard logit post-processing specialisation with node compaction (am_rollout_tsp.cu)
int launch_rollout_tsp_greedy(const rl4co_am_rollout_args &a, cudaStream_t stream);
bool rollout_tsp_greedy_supports(const rl4co_am_rollout_args &a);

// multistart (POMO) rollouts of tsp / cvrp on the tensor cores, starts = TMEM lanes (am_rollout_ms_tc.cu)
int launch_rollout_ms_tc(const rl4co_am_rollout_args &a, cudaStream_t stream);
bool rollout_ms_tc_supports(const rl4co_am_rollout_args &a);


// slap: constant query, incrementally maintained glimpse, two CTAs per SM (am_rollout_slap.cu)
int launch_rollout_slap(const rl4co_am_rollout_args &a, cudaStream_t stream);
bool rollout_slap_supports(const rl4co_am_rollout_args &a);

}  // namespace rl4co
