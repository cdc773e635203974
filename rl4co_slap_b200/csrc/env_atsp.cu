// K1 - Asymmetric TSP: fused replay (reset + T x _step + reward + validity) on a cost matrix.
// Reference semantics: rl4co/envs/routing/atsp/env.py:64-171.  The state machine is the TSP one (reset and _step go
// through rl4co_tsp_reset / rl4co_tsp_step); what differs is the reward, a gather of T entries of the instance's
// [N,N] cost matrix along the closed tour (:152-162).  One warp per instance; each gathered entry costs one 32-byte
// sector, so the traffic is T sectors + the action row per instance, not the 4 N^2 bytes of the matrix.
#include "common.cuh"

namespace rl4co {

constexpr int kAtspWarps = 8;

__global__ void __launch_bounds__(kAtspWarps * 32)
atsp_replay_kernel(int64_t B, int N, int T, const float *__restrict__ cost, const int64_t *__restrict__ actions,
                   int64_t *first, int64_t *current, int64_t *i, uint8_t *mask, uint8_t *done, float *reward, int check,
                   int32_t *status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    const int per_warp = (Tp * 4 + Np + 15) & ~15;       // actions | visited bytes
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int *sa = reinterpret_cast<int *>(smem_raw + (size_t)warp * per_warp);
    uint8_t *sv = reinterpret_cast<uint8_t *>(sa + Tp);
    const int64_t b = blockIdx.x * (int64_t)kAtspWarps + warp;
    if (b >= B) return;
    for (int k = lane; k < Np; k += 32) sv[k] = 0;
    const int64_t *arow = actions + b * (int64_t)T;
    int st = 0;
    for (int t = lane; t < T; t += 32) {
        int64_t a = __ldg(arow + t);
        if (a < 0 || a >= N) { st |= RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
        sa[t] = (int)a;
    }
    __syncwarp();
    const float *crow = cost + b * (int64_t)N * N;
    float len = 0.f;
    int dup = 0;
    for (int t = lane; t < T; t += 32) {
        const int a = sa[t], nx = sa[t + 1 == T ? 0 : t + 1];
        len += __ldg(crow + (size_t)a * N + nx);                       // distance_matrix[b, src, tgt], :155-162
        const unsigned old = atomicOr(reinterpret_cast<unsigned *>(sv + (a & ~3)), 1u << (8 * (a & 3)));
        dup |= (old >> (8 * (a & 3))) & 1u;
    }
    len = warp_sum(len);
    dup = __any_sync(kFull, dup);
    __syncwarp();
    int seen = 0;
    for (int n = lane; n < N; n += 32) {
        seen += sv[n] != 0;
        if (mask) mask[b * (int64_t)N + n] = sv[n] ? 0 : 1;
    }
    seen = warp_sum_int(seen);
    if (lane == 0) {
        if (first) first[b] = T > 0 ? sa[0] : 0;
        if (current) current[b] = T > 0 ? sa[T - 1] : 0;
        if (i) i[b] = T;
        if (done) done[b] = seen == N;
        if (reward) reward[b] = -len;
        if (check && (dup || T != N || seen != N)) st |= RL4CO_ST_INVALID_TOUR;     // sort(actions) == arange(N), :164-171
    }
    st = __reduce_or_sync(kFull, st);
    if (lane == 0) raise_status(status, st);
}

}  // namespace rl4co

using namespace rl4co;

extern "C" int rl4co_atsp_replay(int64_t B, int32_t N, int32_t T, const float *cost_matrix, const int64_t *actions,
                                 int64_t *first_node, int64_t *current_node, int64_t *i, uint8_t *action_mask,
                                 uint8_t *done, float *reward, int32_t check_validity, int32_t *status,
                                 rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1 && T >= 0, "atsp_replay: bad sizes B=%lld N=%d T=%d", (long long)B, N, T);
    if (B == 0) return 0;
    RL4CO_REQUIRE(cost_matrix && actions, "atsp_replay: null input");
    const int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    const size_t smem = (size_t)((Tp * 4 + Np + 15) & ~15) * kAtspWarps;
    RL4CO_REQUIRE(smem <= 227 * 1024, "atsp_replay: N=%d T=%d needs %zu B shared memory", N, T, smem);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(atsp_replay_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int64_t blocks = (B + kAtspWarps - 1) / kAtspWarps;
    RL4CO_REQUIRE(blocks < (1ll << 31), "atsp_replay: batch too large");
    atsp_replay_kernel<<<(unsigned)blocks, kAtspWarps * 32, smem, (cudaStream_t)stream>>>(
        B, N, T, cost_matrix, actions, first_node, current_node, i, action_mask, done, reward, check_validity, status);
    RL4CO_CHECK_LAUNCH("atsp_replay");
    return 0;
}
