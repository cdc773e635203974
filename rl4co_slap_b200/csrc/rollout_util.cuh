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


// tsp without table 4 (args h / w_first): the first node's half of the step context, project_context.weight[:, :E] . h_first
// (context.py:123-131), formed in the rollout kernel once the first node is known.  Thread layout of the register-file
// kernels: a warp owns output dims 16 warp .. 16 warp + 15, this thread dims dim0..dim0+7 of them, and the 16 lanes
// g = lane & 15 of its half-warp split the 128 inputs, 8 each: inputs 4g..4g+3 and 64+4g..64+4g+3, so that each load
// instruction of a half-warp covers 256 contiguous bytes of one weight row.  The butterfly leaves the same sum in all 16
// lanes and lane g == 0 stores it to out16 = 16 floats of shared memory private to the warp (indexed by dim - 16 warp).
// All 32 lanes must call it; it ends with __syncwarp().
//
// Not inlined on purpose: the callers keep ~250 registers live across their decode loop and the 18 loads in flight here
// would otherwise change that loop's register allocation (spills in the step, seen with ptxas -v); as a call the cost
// stays on this once-per-rollout path.  The call leaves the function ~56 registers, not enough for 18 x 128-bit loads in
// one batch (the compiler then splits them into two dependent round trips to L2, ~2.8 k clk per rollout measured), so
// with `scratch` (4 KB of shared memory private to the warp, free at this point) the weight rows of dims 4..7 travel by
// cp.async instead and everything is in flight at once.  scratch == nullptr: registers only.
// The weight rows do not depend on the first node, only the h row does: a caller whose scratch is free from the start of
// the rollout issues first_context_prefetch() there (no registers, the 32 KB land during step 0) and passes
// prefetched = true, which leaves one L2 round trip for the other half of the rows and the h row at the end of step 0
// (measured: step 0 of a TSP100 rollout 4.6 k -> 3.7 k clk; 2.3 k with table 4.  Prefetching the other half as well, into
// unused rows of the kept logits, gained nothing more: what remains is the h row's latency and the mat-vec itself.)
__device__ __forceinline__ void first_context_prefetch(const float *w_first, float *scratch, int dim0, int g) {
    const float *wr0 = w_first + (size_t)dim0 * 128 + 4 * g;
    const uint32_t sdst = smem_u32(scratch) + (threadIdx.x & 31) * 16;
#pragma unroll
    for (int i = 4; i < 8; ++i)
#pragma unroll
        for (int half = 0; half < 2; ++half)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sdst + ((i - 4) * 2 + half) * 512),
                         "l"(wr0 + (size_t)i * 128 + 64 * half) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int kUnused = 0>   // a template only so that the header may define a non-inline function
__device__ __noinline__ void first_context_row(const float *hrow, const float *w_first, float *out16, float *scratch,
                                               bool prefetched, int dim0, int g) {
    const int lane = threadIdx.x & 31;
    const float *wr0 = w_first + (size_t)dim0 * 128 + 4 * g;
    if (scratch && !prefetched) first_context_prefetch(w_first, scratch, dim0, g);
    const float4 h0 = __ldg(reinterpret_cast<const float4 *>(hrow + 4 * g));
    const float4 h1 = __ldg(reinterpret_cast<const float4 *>(hrow + 64 + 4 * g));
    float s[8];
    auto dot8 = [&](const float4 &w0, const float4 &w1) {
        float v = h0.x * w0.x;
        v = fmaf(h0.y, w0.y, v); v = fmaf(h0.z, w0.z, v); v = fmaf(h0.w, w0.w, v);
        v = fmaf(h1.x, w1.x, v); v = fmaf(h1.y, w1.y, v); v = fmaf(h1.z, w1.z, v); v = fmaf(h1.w, w1.w, v);
        return v;
    };
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float *wr = wr0 + (size_t)i * 128;
        s[i] = dot8(__ldg(reinterpret_cast<const float4 *>(wr)), __ldg(reinterpret_cast<const float4 *>(wr + 64)));
    }
    if (scratch) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        const float4 *sw = reinterpret_cast<const float4 *>(scratch) + lane;
#pragma unroll
        for (int i = 4; i < 8; ++i) s[i] = dot8(sw[((i - 4) * 2) * 32], sw[((i - 4) * 2 + 1) * 32]);
    } else {
#pragma unroll
        for (int i = 4; i < 8; ++i) {
            const float *wr = wr0 + (size_t)i * 128;
            s[i] = dot8(__ldg(reinterpret_cast<const float4 *>(wr)), __ldg(reinterpret_cast<const float4 *>(wr + 64)));
        }
    }
#pragma unroll
    for (int m = 1; m < 16; m <<= 1)
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i] += __shfl_xor_sync(0xffffffffu, s[i], m);
    if (g == 0) {
        float4 *o = reinterpret_cast<float4 *>(out16 + (dim0 & 15));
        o[0] = make_float4(s[0], s[1], s[2], s[3]);
        o[1] = make_float4(s[4], s[5], s[6], s[7]);
    }
    __syncwarp();
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
