// K1 - Prize-collecting TSP environment kernels (action mask / step / fused replay+reward+validity).
// Reference semantics: rl4co/envs/routing/pctsp/env.py:69-200; spctsp/env.py is the same environment with the
// stochastic prize as the collected one.  Sibling environment of the routing family.
// The prize / penalty recurrences are sequential fp32 adds in the reference's order (__fadd_rn), so the depot bit of
// the mask (`cur_total_prize < 1`) is bit exact.
#include "common.cuh"

namespace rl4co {

constexpr int kPcWarps = 8;

// mask row of one instance (pctsp/env.py:141-148); all lanes of the warp participate
__device__ __forceinline__ void pctsp_mask_row(int N, int lane, float total_prize, const uint8_t *visited, uint8_t *mask) {
    int seen = 0;
    for (int n = lane + 1; n < N; n += 32) seen += visited[n] != 0;
    seen = warp_sum_int(seen);
    if (!mask) return;
    const bool closed = visited[0] != 0;
    for (int n = lane; n < N; n += 32) {
        if (n == 0) mask[0] = (total_prize < 1.0f && seen < N - 1) ? 0 : 1;
        else mask[n] = (visited[n] != 0 || closed) ? 0 : 1;
    }
}

__global__ void pctsp_mask_kernel(int64_t B, int N, const float *total_prize, const uint8_t *visited, uint8_t *mask) {
    const int lane = threadIdx.x & 31;
    const int64_t b = blockIdx.x * (int64_t)kPcWarps + (threadIdx.x >> 5);
    if (b >= B) return;
    pctsp_mask_row(N, lane, total_prize[b], visited + b * (int64_t)N, mask + b * (int64_t)N);
}

__global__ void pctsp_step_kernel(int64_t B, int N, const int64_t *action, const float *prize, const float *penalty,
                                  float *total_prize, float *total_penalty, int64_t *current, uint8_t *visited,
                                  int64_t *i, uint8_t *mask, uint8_t *done, int32_t *status) {
    const int lane = threadIdx.x & 31;
    const int64_t b = blockIdx.x * (int64_t)kPcWarps + (threadIdx.x >> 5);
    if (b >= B) return;
    int64_t a = action[b];
    int st = 0;
    if (a < 0 || a >= N) { st = RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
    uint8_t *vrow = visited + b * (int64_t)N;
    const float tp = __fadd_rn(total_prize[b], prize[b * (int64_t)N + a]);        // :73-75
    const float pn = __fadd_rn(total_penalty[b], penalty[b * (int64_t)N + a]);    // :76-78
    const int64_t step = i[b];
    __syncwarp();
    if (lane == 0) vrow[a] = 1;                                                   // :81
    __syncwarp();
    pctsp_mask_row(N, lane, tp, vrow, mask + b * (int64_t)N);
    if (lane == 0) {
        total_prize[b] = tp;
        total_penalty[b] = pn;
        current[b] = a;
        i[b] = step + 1;
        done[b] = (step > 0) && (a == 0);                                         // :84
        raise_status(status, st);
    }
}

// One warp per instance, as cvrp_replay_kernel: locs / prize / penalty / actions staged in shared memory.
__global__ void __launch_bounds__(kPcWarps * 32)
pctsp_replay_kernel(int64_t B, int N, int T, const float *__restrict__ locs, const float *__restrict__ prize,
                    const float *__restrict__ penalty, const int64_t *__restrict__ actions, int64_t *current,
                    float *total_prize, float *total_penalty, uint8_t *visited, uint8_t *mask, int64_t *i,
                    uint8_t *done, float *reward, int check, int32_t *status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    const int per_warp = N * 8 + Np * 8 + Tp * 4 + Np;   // locs | prize | penalty | actions | visited bytes
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *base = smem_raw + (size_t)warp * ((per_warp + 15) & ~15);
    float2 *sl = reinterpret_cast<float2 *>(base);
    float *sp = reinterpret_cast<float *>(base + N * 8);
    float *sq = reinterpret_cast<float *>(base + N * 8 + Np * 4);
    int *sa = reinterpret_cast<int *>(base + N * 8 + Np * 8);
    uint8_t *sv = base + N * 8 + Np * 8 + Tp * 4;

    const int64_t b = blockIdx.x * (int64_t)kPcWarps + warp;
    if (b >= B) return;
    const float2 *lrow = reinterpret_cast<const float2 *>(locs) + b * (int64_t)N;
    for (int k = lane; k < Np; k += 32) sv[k] = 0;
    float all_pen = 0.f;
    for (int k = lane; k < N; k += 32) {
        sl[k] = __ldg(lrow + k);
        sp[k] = __ldg(prize + b * (int64_t)N + k);
        sq[k] = __ldg(penalty + b * (int64_t)N + k);
        if (k > 0) all_pen += sq[k];
    }
    all_pen = warp_sum(all_pen);
    const int64_t *arow = actions + b * (int64_t)T;
    int st = 0;
    for (int t = lane; t < T; t += 32) {
        int64_t a = __ldg(arow + t);
        if (a < 0 || a >= N) { st |= RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
        sa[t] = (int)a;
    }
    __syncwarp();

    // visited, "Duplicates" (:181-186), gathered prize / saved penalty sums, tour legs depot -> a_0 ... a_{T-1} -> depot
    int dup = 0, customers = 0;
    float got = 0.f, saved = 0.f, len = 0.f;
    for (int t = lane; t <= T; t += 32) {
        const float2 p = t == 0 ? sl[0] : sl[sa[t - 1]];
        const float2 q = t == T ? sl[0] : sl[sa[t]];
        len += leg(p.x, p.y, q.x, q.y);
        if (t < T) {
            const int a = sa[t];
            const unsigned old = atomicOr(reinterpret_cast<unsigned *>(sv + (a & ~3)), 1u << (8 * (a & 3)));
            if (a != 0) { dup |= (old >> (8 * (a & 3))) & 1u; ++customers; }
            got += sp[a];
            saved += sq[a];
        }
    }
    got = warp_sum(got);
    saved = warp_sum(saved);
    len = warp_sum(len);
    customers = warp_sum_int(customers);
    dup = __any_sync(kFull, dup);
    __syncwarp();

    // env recurrences (:73-78), every lane
    // cur_total_penalty is IN/OUT: it enters holding the reset state's sum of all penalties (torch's summation order)
    float tp = 0.f, pn = total_penalty ? total_penalty[b] : all_pen;
    for (int t = 0; t < T; ++t) {
        tp = __fadd_rn(tp, sp[sa[t]]);
        pn = __fadd_rn(pn, sq[sa[t]]);
    }
    const int last = T > 0 ? sa[T - 1] : 0;
    if (visited) for (int n = lane; n < N; n += 32) visited[b * (int64_t)N + n] = sv[n];
    pctsp_mask_row(N, lane, tp, sv, mask ? mask + b * (int64_t)N : nullptr);
    if (lane == 0) {
        if (current) current[b] = last;
        if (total_prize) total_prize[b] = tp;
        if (total_penalty) total_penalty[b] = pn;
        if (i) i[b] = T;
        if (done) done[b] = (T > 1) && (last == 0);
        if (reward) reward[b] = saved - (len + all_pen);                        // :172-174
        if (check) {
            const bool enough = got >= 1.f - 1e-5f || customers == N - 1;       // :193-200
            st |= (dup ? RL4CO_ST_DUPLICATES : 0) | (enough ? 0 : RL4CO_ST_PRIZE_SHORT);
        }
    }
    st = __reduce_or_sync(kFull, st);
    if (lane == 0) raise_status(status, st);
}

}  // namespace rl4co

using namespace rl4co;

static int pc_blocks(int64_t B, unsigned *out) {
    const int64_t blocks = (B + kPcWarps - 1) / kPcWarps;
    if (blocks >= (1ll << 31)) return 1;
    *out = (unsigned)blocks;
    return 0;
}

extern "C" int rl4co_pctsp_action_mask(int64_t B, int32_t N, const float *cur_total_prize, const uint8_t *visited,
                                       uint8_t *action_mask, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1, "pctsp_action_mask: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(cur_total_prize && visited && action_mask, "pctsp_action_mask: null pointer");
    unsigned g;
    RL4CO_REQUIRE(pc_blocks(B, &g) == 0, "pctsp_action_mask: batch too large");
    pctsp_mask_kernel<<<g, kPcWarps * 32, 0, (cudaStream_t)stream>>>(B, N, cur_total_prize, visited, action_mask);
    RL4CO_CHECK_LAUNCH("pctsp_action_mask");
    return 0;
}

extern "C" int rl4co_pctsp_step(int64_t B, int32_t N, const int64_t *action, const float *real_prize,
                                const float *penalty, float *cur_total_prize, float *cur_total_penalty,
                                int64_t *current_node, uint8_t *visited, int64_t *i, uint8_t *action_mask,
                                uint8_t *done, int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1, "pctsp_step: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(action && real_prize && penalty && cur_total_prize && cur_total_penalty && current_node && visited &&
                      i && action_mask && done, "pctsp_step: null pointer");
    unsigned g;
    RL4CO_REQUIRE(pc_blocks(B, &g) == 0, "pctsp_step: batch too large");
    pctsp_step_kernel<<<g, kPcWarps * 32, 0, (cudaStream_t)stream>>>(B, N, action, real_prize, penalty, cur_total_prize,
                                                                      cur_total_penalty, current_node, visited, i,
                                                                      action_mask, done, status);
    RL4CO_CHECK_LAUNCH("pctsp_step");
    return 0;
}

extern "C" int rl4co_pctsp_replay(int64_t B, int32_t N, int32_t T, const float *locs, const float *real_prize,
                                  const float *penalty, const int64_t *actions, int64_t *current_node,
                                  float *cur_total_prize, float *cur_total_penalty, uint8_t *visited,
                                  uint8_t *action_mask, int64_t *i, uint8_t *done, float *reward,
                                  int32_t check_validity, int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1 && T >= 0, "pctsp_replay: bad sizes B=%lld N=%d T=%d", (long long)B, N, T);
    if (B == 0) return 0;
    RL4CO_REQUIRE(locs && real_prize && penalty && actions, "pctsp_replay: null input");
    const int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    const int per_warp = (N * 8 + Np * 8 + Tp * 4 + Np + 15) & ~15;
    const size_t smem = (size_t)per_warp * kPcWarps;
    RL4CO_REQUIRE(smem <= 227 * 1024, "pctsp_replay: N=%d T=%d needs %zu B shared memory", N, T, smem);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(pctsp_replay_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    unsigned g;
    RL4CO_REQUIRE(pc_blocks(B, &g) == 0, "pctsp_replay: batch too large");
    pctsp_replay_kernel<<<g, kPcWarps * 32, smem, (cudaStream_t)stream>>>(B, N, T, locs, real_prize, penalty, actions,
                                                                           current_node, cur_total_prize,
                                                                           cur_total_penalty, visited, action_mask, i,
                                                                           done, reward, check_validity, status);
    RL4CO_CHECK_LAUNCH("pctsp_replay");
    return 0;
}
