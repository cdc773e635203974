// K1 - TSP environment kernels (reset / step / fused replay+reward+validity).
// Reference semantics: rl4co/envs/routing/tsp/env.py:67-173, rl4co/utils/ops.py:65-101.
// HBM-bound byte/integer work: one warp per instance, 128-bit loads where the row is
// 16-byte aligned, state kept as a shared-memory bitset, one launch per call.
#include "common.cuh"

namespace rl4co {

constexpr int kWarpsPerCta = 8;

// ---------------------------------------------------------------- reset -----------
__global__ void tsp_reset_kernel(int64_t B, int N, int64_t *first_node, int64_t *current_node,
                                 int64_t *i, uint8_t *mask, float *reward, uint8_t *done,
                                 uint8_t *terminated) {
    int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t b = tid; b < B; b += stride) {
        if (first_node) first_node[b] = 0;
        if (current_node) current_node[b] = 0;
        if (i) i[b] = 0;
        if (reward) reward[b] = 0.f;
        if (done) done[b] = 0;
        if (terminated) terminated[b] = 0;
    }
    if (mask) {
        int64_t total = B * N;
        // 16-byte stores over the bulk, bytes for the tail (cudaMalloc'd rows are aligned)
        int64_t nvec = ((reinterpret_cast<uintptr_t>(mask) & 15) == 0) ? total / 16 : 0;
        uint4 ones = make_uint4(0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u);
        for (int64_t v = tid; v < nvec; v += stride) reinterpret_cast<uint4 *>(mask)[v] = ones;
        for (int64_t k = nvec * 16 + tid; k < total; k += stride) mask[k] = 1;
    }
}

// ---------------------------------------------------------------- step ------------
// Batch-global first-step test of the reference: `td["i"].all() == 0` (tsp/env.py:70)
// is true iff ANY element of i is zero.
__global__ void any_zero_kernel(int64_t B, const int64_t *i, int32_t *flag) {
    int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int found = 0;
    for (int64_t b = tid; b < B; b += stride) found |= (i[b] == 0);
    if (__any_sync(kFull, found) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

__global__ void tsp_step_kernel(int64_t B, int N, const int64_t *action, int64_t *i,
                                int64_t *first_node, int64_t *current_node, uint8_t *mask,
                                uint8_t *done, const int32_t *flag, int32_t *status) {
    int lane = threadIdx.x & 31;
    int64_t b = blockIdx.x * (int64_t)kWarpsPerCta + (threadIdx.x >> 5);
    if (b >= B) return;
    int64_t a = action[b];
    int st = 0;
    if (a < 0 || a >= N) { st = RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
    uint8_t *row = mask + b * N;
    int remaining = 0;
    for (int n = lane; n < N; n += 32) {
        uint8_t m = row[n];
        if (n == (int)a) { m = 0; row[n] = 0; }
        remaining += (m != 0);
    }
    remaining = warp_sum_int(remaining);
    if (lane == 0) {
        if (*flag) first_node[b] = a;
        current_node[b] = a;
        i[b] = i[b] + 1;
        done[b] = (remaining == 0);
        raise_status(status, st);
    }
}

// ---------------------------------------------------------------- replay ----------
// One warp per instance.  Shared memory per warp: actions as int32 [T], locs [N] float2,
// visited bitset [ceil(N/32)] words.
__global__ void __launch_bounds__(kWarpsPerCta * 32)
tsp_replay_kernel(int64_t B, int N, int T, const float *__restrict__ locs,
                  const int64_t *__restrict__ actions, int64_t *first_node,
                  int64_t *current_node, int64_t *i_out, uint8_t *mask, uint8_t *done,
                  float *reward, int check, int32_t *status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int words = (N + 31) >> 5;
    const int Tp = (T + 1) & ~1, Np = N;
    // per-warp layout: float2 locs[Np] | int act[Tp] | uint bits[words(+pad)]
    const int per_warp = Np * 8 + Tp * 4 + ((words + 3) & ~3) * 4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *base = smem_raw + (size_t)warp * per_warp;
    float2 *sl = reinterpret_cast<float2 *>(base);
    int *sa = reinterpret_cast<int *>(base + Np * 8);
    unsigned *bits = reinterpret_cast<unsigned *>(base + Np * 8 + Tp * 4);

    int64_t b = blockIdx.x * (int64_t)kWarpsPerCta + warp;
    if (b >= B) return;

    for (int w = lane; w < words; w += 32) bits[w] = 0u;

    // ---- coalesced staging: locs row = N float2 (8N bytes), actions row = T int64
    const float *lrow = locs + b * (int64_t)N * 2;
    if ((reinterpret_cast<uintptr_t>(lrow) & 15) == 0 && (N & 1) == 0) {
        const float4 *v = reinterpret_cast<const float4 *>(lrow);
        for (int k = lane; k < N / 2; k += 32) {
            float4 q = __ldg(v + k);
            sl[2 * k] = make_float2(q.x, q.y);
            sl[2 * k + 1] = make_float2(q.z, q.w);
        }
    } else {
        const float2 *v = reinterpret_cast<const float2 *>(lrow);
        for (int k = lane; k < N; k += 32) sl[k] = __ldg(v + k);
    }
    const int64_t *arow = actions + b * (int64_t)T;
    int st = 0;
    auto put = [&](int t, int64_t a) {
        if (a < 0 || a >= N) { st |= RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
        sa[t] = (int)a;
    };
    if ((reinterpret_cast<uintptr_t>(arow) & 15) == 0 && (T & 1) == 0) {
        const longlong2 *v = reinterpret_cast<const longlong2 *>(arow);
        for (int k = lane; k < T / 2; k += 32) {
            longlong2 q = __ldg(v + k);
            put(2 * k, q.x);
            put(2 * k + 1, q.y);
        }
    } else {
        for (int t = lane; t < T; t += 32) put(t, __ldg(arow + t));
    }
    __syncwarp();

    // ---- visited bitset + validity (sort(actions) == arange(T), tsp/env.py:165-173)
    int bad = 0;
    for (int t = lane; t < T; t += 32) {
        int a = sa[t];
        unsigned bit = 1u << (a & 31);
        unsigned old = atomicOr(&bits[a >> 5], bit);
        bad |= ((old & bit) != 0) | (a >= T);
    }
    // ---- tour length (utils/ops.py:93-101): legs a_t -> a_{t+1}, closed
    float len = 0.f;
    for (int t = lane; t < T; t += 32) {
        float2 p = sl[sa[t]];
        float2 q = sl[sa[t + 1 == T ? 0 : t + 1]];
        len += leg(p.x, p.y, q.x, q.y);
    }
    len = warp_sum(len);
    bad = __any_sync(kFull, bad);
    __syncwarp();

    // ---- final state after T steps (tsp/env.py:83-93)
    int remaining = 0;
    if (mask) {
        uint8_t *row = mask + b * (int64_t)N;
        for (int n = lane; n < N; n += 32) {
            uint8_t m = ((bits[n >> 5] >> (n & 31)) & 1u) ? 0 : 1;
            row[n] = m;
            remaining += m;
        }
    } else {
        for (int n = lane; n < N; n += 32) remaining += ((bits[n >> 5] >> (n & 31)) & 1u) ? 0 : 1;
    }
    remaining = warp_sum_int(remaining);
    if (lane == 0) {
        if (first_node) first_node[b] = T > 0 ? sa[0] : 0;
        if (current_node) current_node[b] = T > 0 ? sa[T - 1] : 0;
        if (i_out) i_out[b] = T;
        if (done) done[b] = (T > 0) && (remaining == 0);
        if (reward) reward[b] = -len;
        if (check && bad) st |= RL4CO_ST_INVALID_TOUR;
    }
    st = __reduce_or_sync(kFull, st);
    if (lane == 0) raise_status(status, st);
}

// ---------------------------------------------------------------- replay, persistent ----------
// N, T <= 128 (even): warps loop over instances and keep the NEXT instance's rows (locs 8N bytes, actions 8T bytes =
// at most two 16-byte registers each per lane) in flight while the current one is processed, so the HBM latency is
// hidden behind the per-instance work instead of behind occupancy alone.
// The kernel is bound by instruction issue, not by memory (round 1: 647 warp instructions per instance, 79 % issue
// utilisation, 37 % of HBM), so the per-instance work is kept minimal:
//   * the visited set is a 128-byte "seen" array in shared memory - every action stores one byte, no word selection,
//     no duplicate tracking (sort(actions) == arange(T) <=> T distinct values below T <=> popcount == T and none >= T);
//     its words, inverted, ARE the final action mask (4 nodes per 4-byte store);
//   * a lane owns four CONSECUTIVE legs: one 16-byte load of its actions, five coordinate loads;
//   * the leg lengths use sqrt.approx (<= 1 ulp): masks / states / tours are exact, the reward bar is 1e-5 relative.
__device__ __forceinline__ float leg_fast(float2 p, float2 q) {
    const float dx = __fsub_rn(q.x, p.x), dy = __fsub_rn(q.y, p.y);
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy))));
    return r;
}
__global__ void __launch_bounds__(kWarpsPerCta * 32)
tsp_replay_persistent_kernel(int64_t B, int N, int T, const float *__restrict__ locs,
                             const int64_t *__restrict__ actions, int64_t *first_node,
                             int64_t *current_node, int64_t *i_out, uint8_t *mask, uint8_t *done,
                             float *reward, int check, int32_t *status) {
    __shared__ __align__(16) float2 s_locs[kWarpsPerCta][128];
    __shared__ __align__(16) int s_act[kWarpsPerCta][136];
    __shared__ __align__(16) unsigned s_seen[kWarpsPerCta][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float2 *sl = s_locs[warp];
    int *sa = s_act[warp];
    unsigned *seen = s_seen[warp];
    uint8_t *seen8 = reinterpret_cast<uint8_t *>(seen);
    const int64_t stride = (int64_t)gridDim.x * kWarpsPerCta;
    const int nl = N >> 1, na = T >> 1;             // 16-byte pieces per row
    const bool mask4 = mask != nullptr && (N & 3) == 0 && (reinterpret_cast<uintptr_t>(mask) & 3) == 0;
    const unsigned un = (unsigned)N;

    // this lane's pieces of an instance's rows; the pointers advance by one warp stride per iteration (no 64-bit
    // multiplies in the loop), lanes without a piece never use theirs
    const bool has_l0 = lane < nl, has_l1 = lane + 32 < nl, has_a0 = lane < na, has_a1 = lane + 32 < na;
    int64_t b = blockIdx.x * (int64_t)kWarpsPerCta + warp;
    const float4 *lv = reinterpret_cast<const float4 *>(locs + b * (int64_t)N * 2) + lane;
    const longlong2 *av = reinterpret_cast<const longlong2 *>(actions + b * (int64_t)T) + lane;
    const int64_t lv_step = stride * (int64_t)nl, av_step = stride * (int64_t)na;
    float4 l0 = make_float4(0.f, 0.f, 0.f, 0.f), l1 = l0, nl0 = l0, nl1 = l0;
    longlong2 a0 = make_longlong2(0, 0), a1 = a0, na0 = a0, na1 = a0;
    auto load = [&](bool valid, float4 &x0, float4 &x1, longlong2 &y0, longlong2 &y1) {
        if (valid && has_l0) x0 = __ldg(lv);
        if (valid && has_l1) x1 = __ldg(lv + 32);
        if (valid && has_a0) y0 = __ldg(av);
        if (valid && has_a1) y1 = __ldg(av + 32);
        lv += lv_step;
        av += av_step;
    };
    for (int q = lane; q < 136; q += 32) sa[q] = 0;  // the tail beyond T is read (never used) by the 16-byte action loads
    load(b < B, l0, l1, a0, a1);
    int st = 0;
    for (; b < B; b += stride) {
        load(b + stride < B, nl0, nl1, na0, na1);   // next instance in flight
        __syncwarp();                               // the previous instance's gathers are done
        // ---- stage: coordinates as they come (two nodes per 16 bytes), the seen array cleared
        if (has_l0) reinterpret_cast<float4 *>(sl)[lane] = l0;
        if (has_l1) reinterpret_cast<float4 *>(sl)[lane + 32] = l1;
        seen[lane] = 0u;
        __syncwarp();
        // ---- actions as int (range checked), one seen byte each
        int over = 0;
        auto put = [&](int t, int64_t a) {
            const bool oob = (unsigned long long)a >= (unsigned long long)un;   // negative or >= N (torch would raise)
            st |= oob ? RL4CO_ST_ACTION_RANGE : 0;
            const int ai = oob ? (a < 0 ? 0 : N - 1) : (int)a;
            sa[t] = ai;
            seen8[ai] = 1;
            over |= ai >= T;
        };
        if (has_a0) { put(2 * lane, a0.x); put(2 * lane + 1, a0.y); }
        if (has_a1) { put(2 * lane + 64, a1.x); put(2 * lane + 65, a1.y); }
        __syncwarp();
        const unsigned word = seen[lane];           // bytes 0 / 1 of nodes 4 lane .. 4 lane + 3
        const int visited = __reduce_add_sync(kFull, __popc(word));
        // validity: sort(actions) == arange(T) (tsp/env.py:165-173)  <=>  T distinct values, all < T
        const int bad = __any_sync(kFull, over) || visited != T;
        // ---- tour length (utils/ops.py:93-101): legs a_t -> a_{t+1}, closed; lane owns t = 4 lane .. 4 lane + 3
        // (branch free: entries of `sa` beyond T are 0, legs beyond T are computed and dropped)
        const int t0 = 4 * lane;
        const int4 a4 = *reinterpret_cast<const int4 *>(sa + t0);
        const float2 home = sl[sa[0]];
        const float2 p0 = sl[a4.x], p1 = sl[a4.y], p2 = sl[a4.z], p3 = sl[a4.w], p4 = sl[sa[t0 + 4]];
        // leg k runs from node t0 + k to its successor; the successor of step T - 1 is step 0
        const float g0 = leg_fast(p0, t0 + 1 == T ? home : p1), g1 = leg_fast(p1, t0 + 2 == T ? home : p2);
        const float g2 = leg_fast(p2, t0 + 3 == T ? home : p3), g3 = leg_fast(p3, t0 + 4 == T ? home : p4);
        float len = ((t0 < T ? g0 : 0.f) + (t0 + 1 < T ? g1 : 0.f)) + ((t0 + 2 < T ? g2 : 0.f) + (t0 + 3 < T ? g3 : 0.f));
        len = warp_sum(len);
        // ---- final state after T steps (tsp/env.py:83-93)
        if (mask) {
            uint8_t *row = mask + b * (int64_t)N;
            if (mask4) {
                if (4 * lane < N) reinterpret_cast<unsigned *>(row)[lane] = word ^ 0x01010101u;   // mask = not visited
            } else {
                for (int n = lane; n < N; n += 32) row[n] = seen8[n] ? 0 : 1;
            }
        }
        if (lane == 0) {
            if (first_node) first_node[b] = T > 0 ? sa[0] : 0;
            if (current_node) current_node[b] = T > 0 ? sa[T - 1] : 0;
            if (i_out) i_out[b] = T;
            if (done) done[b] = (T > 0) && (visited == N);
            if (reward) reward[b] = -len;
            if (check && bad) st |= RL4CO_ST_INVALID_TOUR;
        }
        l0 = nl0; l1 = nl1; a0 = na0; a1 = na1;
    }
    st = __reduce_or_sync(kFull, st);
    if (lane == 0) raise_status(status, st);
}

static int grid_for(int64_t n, int per_block) {
    int64_t g = (n + per_block - 1) / per_block;
    int64_t cap = 148 * 32;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace rl4co

using namespace rl4co;

extern "C" int rl4co_tsp_reset(int64_t B, int32_t N, int64_t *first_node, int64_t *current_node,
                               int64_t *i, uint8_t *action_mask, float *reward, uint8_t *done,
                               uint8_t *terminated, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 0, "tsp_reset: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    tsp_reset_kernel<<<grid_for(B * (int64_t)((N + 15) / 16 + 1), 256), 256, 0, (cudaStream_t)stream>>>(
        B, N, first_node, current_node, i, action_mask, reward, done, terminated);
    RL4CO_CHECK_LAUNCH("tsp_reset");
    return 0;
}

extern "C" int rl4co_tsp_step(int64_t B, int32_t N, const int64_t *action, int64_t *i,
                              int64_t *first_node, int64_t *current_node, uint8_t *action_mask,
                              uint8_t *done, int32_t *flag_ws, int32_t *status,
                              rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 0, "tsp_step: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(action && i && first_node && current_node && action_mask && done && flag_ws,
                  "tsp_step: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    cudaMemsetAsync(flag_ws, 0, sizeof(int32_t), s);
    any_zero_kernel<<<grid_for(B, 256), 256, 0, s>>>(B, i, flag_ws);
    int64_t blocks = (B + kWarpsPerCta - 1) / kWarpsPerCta;
    RL4CO_REQUIRE(blocks < (1ll << 31), "tsp_step: batch too large");
    tsp_step_kernel<<<(unsigned)blocks, kWarpsPerCta * 32, 0, s>>>(
        B, N, action, i, first_node, current_node, action_mask, done, flag_ws, status);
    RL4CO_CHECK_LAUNCH("tsp_step");
    return 0;
}

extern "C" int rl4co_tsp_replay(int64_t B, int32_t N, int32_t T, const float *locs,
                                const int64_t *actions, int64_t *first_node,
                                int64_t *current_node, int64_t *i, uint8_t *action_mask,
                                uint8_t *done, float *reward, int32_t check_validity,
                                int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 0 && T >= 0, "tsp_replay: bad sizes B=%lld N=%d T=%d", (long long)B, N, T);
    if (B == 0) return 0;
    RL4CO_REQUIRE(locs && actions, "tsp_replay: null input");
    if (N <= 128 && T <= 128 && T >= 2 && (N & 1) == 0 && (T & 1) == 0 &&
        (reinterpret_cast<uintptr_t>(locs) & 15) == 0 && (reinterpret_cast<uintptr_t>(actions) & 15) == 0) {
        static int resident = 0;                            // CTAs the device holds at once (persistent grid)
        if (resident == 0) {
            int dev = 0, sm_count = 0, per_sm = 0;
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tsp_replay_persistent_kernel, kWarpsPerCta * 32, 0);
            resident = (sm_count > 0 ? sm_count : 148) * (per_sm > 0 ? per_sm : 4);
        }
        int64_t blocks = (B + kWarpsPerCta - 1) / kWarpsPerCta;
        if (blocks > resident) blocks = resident;
        tsp_replay_persistent_kernel<<<(unsigned)blocks, kWarpsPerCta * 32, 0, (cudaStream_t)stream>>>(
            B, N, T, locs, actions, first_node, current_node, i, action_mask, done, reward, check_validity, status);
        RL4CO_CHECK_LAUNCH("tsp_replay");
        return 0;
    }
    int words = (N + 31) / 32;
    int per_warp = N * 8 + ((T + 1) & ~1) * 4 + ((words + 3) & ~3) * 4;
    size_t smem = (size_t)per_warp * kWarpsPerCta;
    RL4CO_REQUIRE(smem <= 227 * 1024, "tsp_replay: N=%d T=%d needs %zu B shared memory", N, T, smem);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(tsp_replay_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int64_t blocks = (B + kWarpsPerCta - 1) / kWarpsPerCta;
    RL4CO_REQUIRE(blocks < (1ll << 31), "tsp_replay: batch too large");
    tsp_replay_kernel<<<(unsigned)blocks, kWarpsPerCta * 32, smem, (cudaStream_t)stream>>>(
        B, N, T, locs, actions, first_node, current_node, i, action_mask, done, reward,
        check_validity, status);
    RL4CO_CHECK_LAUNCH("tsp_replay");
    return 0;
}
