// K1 - SLAP (storage location assignment) environment kernels.
// Reference semantics: rl4co/envs/warehousing/slap/env.py:38-143.  The reference's _step
// clears one mask bit per instance in a python loop (:61-62) and its reward loops over the
// orders in python (:136-142); here each is one launch.
#include "common.cuh"

namespace rl4co {

constexpr int kSlapWarps = 8;

__global__ void slap_step_kernel(int64_t B, int N, int P, const int64_t *action,
                                 const float *product, int64_t pstride, int64_t *i,
                                 int32_t *assignment, uint8_t *mask, uint8_t *done,
                                 int32_t *status) {
    int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    int64_t a = action[b];
    int st = 0;
    if (a < 0 || a >= N) { st = RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
    int prod = (int)product[b * pstride];                   // product.to(torch.int), :51
    if (prod < 0 || prod >= P) { st |= RL4CO_ST_ACTION_RANGE; prod = prod < 0 ? 0 : P - 1; }
    assignment[b * (int64_t)P + prod] = (int32_t)a;         // :53
    int64_t step = i[b];
    done[b] = (step == P - 1);                              // :56 (uses i BEFORE the increment)
    mask[b * (int64_t)N + a] = 0;                           // :61-62
    i[b] = step + 1;
    if (st) raise_status(status, st);
}

// closed tour through the K storage locations of one order (utils/ops.py:93-101)
__device__ __forceinline__ float order_length(const float2 *locs, const int32_t *assign,
                                              const int64_t *order, int K, int N, int P) {
    float len = 0.f;
    int p0 = (int)order[0];
    p0 = p0 < 0 ? p0 + P : p0;                              // torch advanced indexing wraps negatives
    int l0 = assign[p0];
    l0 = l0 < 0 ? l0 + N : l0;
    float2 first = locs[l0], prev = first;
    for (int k = 1; k < K; ++k) {
        int p = (int)order[k];
        p = p < 0 ? p + P : p;
        int l = assign[p];
        l = l < 0 ? l + N : l;
        float2 cur = locs[l];
        len += leg(prev.x, prev.y, cur.x, cur.y);
        prev = cur;
    }
    len += leg(prev.x, prev.y, first.x, first.y);
    return len;
}

// one warp per instance, lanes over orders
__global__ void slap_reward_kernel(int64_t B, int N, int P, int O, int K, const float *locs,
                                   const int32_t *assignment, const int64_t *picklist,
                                   float *reward) {
    int lane = threadIdx.x & 31;
    int64_t b = blockIdx.x * (int64_t)kSlapWarps + (threadIdx.x >> 5);
    if (b >= B) return;
    const float2 *l = reinterpret_cast<const float2 *>(locs + b * (int64_t)N * 2);
    const int32_t *as = assignment + b * (int64_t)P;
    const int64_t *pl = picklist + b * (int64_t)O * K;
    float total = 0.f;
    for (int o = lane; o < O; o += 32) total += order_length(l, as, pl + (int64_t)o * K, K, N, P);
    total = warp_sum(total);
    if (lane == 0) reward[b] = -total;
}

// fused reset + P steps + reward; per-warp shared: assignment int[P]
__global__ void slap_replay_kernel(int64_t B, int N, int P, int O, int K, const float *locs,
                                   const int64_t *picklist, const int64_t *actions,
                                   int32_t *assignment, uint8_t *mask, int64_t *i_out,
                                   uint8_t *done, float *reward, int32_t *status) {
    extern __shared__ int32_t s_assign[];
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int32_t *as = s_assign + warp * P;
    int64_t b = blockIdx.x * (int64_t)kSlapWarps + warp;
    if (b >= B) return;
    int st = 0;
    if (mask) {
        uint8_t *row = mask + b * (int64_t)N;
        for (int n = lane; n < N; n += 32) row[n] = n == 0 ? 0 : 1;   // depot unavailable, :116
    }
    for (int t = lane; t < P; t += 32) {
        int64_t a = actions[b * (int64_t)P + t];
        if (a < 0 || a >= N) { st |= RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
        as[t] = (int32_t)a;                                           // product t placed at step t
    }
    __syncwarp();
    if (mask) {
        uint8_t *row = mask + b * (int64_t)N;
        for (int t = lane; t < P; t += 32) row[as[t]] = 0;
    }
    if (assignment) for (int t = lane; t < P; t += 32) assignment[b * (int64_t)P + t] = as[t];
    if (reward) {
        const float2 *l = reinterpret_cast<const float2 *>(locs + b * (int64_t)N * 2);
        const int64_t *pl = picklist + b * (int64_t)O * K;
        float total = 0.f;
        for (int o = lane; o < O; o += 32) total += order_length(l, as, pl + (int64_t)o * K, K, N, P);
        total = warp_sum(total);
        if (lane == 0) reward[b] = -total;
    }
    if (lane == 0) {
        if (i_out) i_out[b] = P;
        if (done) done[b] = P > 0;
    }
    st = __reduce_or_sync(kFull, st);
    if (lane == 0) raise_status(status, st);
}

}  // namespace rl4co

using namespace rl4co;

extern "C" int rl4co_slap_step(int64_t B, int32_t N, int32_t P, const int64_t *action,
                               const float *product, int64_t product_stride, int64_t *i,
                               int32_t *assignment, uint8_t *action_mask, uint8_t *done,
                               int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 0 && P > 0, "slap_step: bad sizes B=%lld N=%d P=%d", (long long)B, N, P);
    if (B == 0) return 0;
    RL4CO_REQUIRE(action && product && i && assignment && action_mask && done, "slap_step: null pointer");
    int64_t blocks = (B + 255) / 256;
    RL4CO_REQUIRE(blocks < (1ll << 31), "slap_step: batch too large");
    slap_step_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        B, N, P, action, product, product_stride, i, assignment, action_mask, done, status);
    RL4CO_CHECK_LAUNCH("slap_step");
    return 0;
}

extern "C" int rl4co_slap_reward(int64_t B, int32_t N, int32_t P, int32_t O, int32_t K,
                                 const float *locs, const int32_t *assignment,
                                 const int64_t *picklist, float *reward, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 0 && P > 0 && O >= 0 && K > 0, "slap_reward: bad sizes");
    if (B == 0) return 0;
    RL4CO_REQUIRE(locs && assignment && picklist && reward, "slap_reward: null pointer");
    int64_t blocks = (B + kSlapWarps - 1) / kSlapWarps;
    RL4CO_REQUIRE(blocks < (1ll << 31), "slap_reward: batch too large");
    slap_reward_kernel<<<(unsigned)blocks, kSlapWarps * 32, 0, (cudaStream_t)stream>>>(
        B, N, P, O, K, locs, assignment, picklist, reward);
    RL4CO_CHECK_LAUNCH("slap_reward");
    return 0;
}

extern "C" int rl4co_slap_replay(int64_t B, int32_t N, int32_t P, int32_t O, int32_t K,
                                 const float *locs, const int64_t *picklist,
                                 const int64_t *actions, int32_t *assignment,
                                 uint8_t *action_mask, int64_t *i, uint8_t *done, float *reward,
                                 int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 0 && P > 0 && O >= 0 && K > 0, "slap_replay: bad sizes");
    if (B == 0) return 0;
    RL4CO_REQUIRE(locs && picklist && actions, "slap_replay: null input");
    size_t smem = (size_t)P * 4 * kSlapWarps;
    RL4CO_REQUIRE(smem <= 48 * 1024, "slap_replay: P=%d too large", P);
    int64_t blocks = (B + kSlapWarps - 1) / kSlapWarps;
    RL4CO_REQUIRE(blocks < (1ll << 31), "slap_replay: batch too large");
    slap_replay_kernel<<<(unsigned)blocks, kSlapWarps * 32, smem, (cudaStream_t)stream>>>(
        B, N, P, O, K, locs, picklist, actions, assignment, action_mask, i, done, reward, status);
    RL4CO_CHECK_LAUNCH("slap_replay");
    return 0;
}
