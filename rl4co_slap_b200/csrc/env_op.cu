// K1 - Orienteering Problem environment kernels (action mask / step / fused replay+reward+validity).
// Reference semantics: rl4co/envs/routing/op/env.py:74-215.  Sibling environment of the routing
// family: collect prizes, come back to the depot within a length budget.
// Distances that feed a mask comparison use leg_fma (common.cuh), the rounding sequence of torch's CPU
// vector norm, and explicit round-to-nearest adds, so `tour_length + dist > max_length` falls on the
// reference's side for borderline nodes.
#include "common.cuh"

namespace rl4co {

constexpr int kOpWarps = 8;

// mask row of one instance (op/env.py:154-170); all lanes of the warp participate.
__device__ __forceinline__ void op_mask_row(int N, int lane, const float2 *locs, const float *max_length,
                                            float tour_length, const uint8_t *visited, int current,
                                            uint8_t *mask) {
    if (!mask) return;
    const float2 pc = locs[current];
    const bool closed = visited[0] != 0;
    for (int n = lane; n < N; n += 32) {
        const float2 p = locs[n];
        const bool exceeds = __fadd_rn(tour_length, leg_fma(pc.x, pc.y, p.x, p.y)) > max_length[n];
        const bool blocked = (visited[n] != 0) | closed | exceeds;
        mask[n] = (n == 0) ? 1 : (blocked ? 0 : 1);      // the depot can always be visited (:168-169)
    }
}

__global__ void op_mask_kernel(int64_t B, int N, const float *locs, const float *max_length,
                               const float *tour_length, const uint8_t *visited, const int64_t *current,
                               uint8_t *mask) {
    const int lane = threadIdx.x & 31;
    const int64_t b = blockIdx.x * (int64_t)kOpWarps + (threadIdx.x >> 5);
    if (b >= B) return;
    int64_t c = current[b];
    c = c < 0 ? 0 : (c >= N ? N - 1 : c);
    op_mask_row(N, lane, reinterpret_cast<const float2 *>(locs) + b * (int64_t)N, max_length + b * (int64_t)N,
                tour_length[b], visited + b * (int64_t)N, (int)c, mask + b * (int64_t)N);
}

__global__ void op_step_kernel(int64_t B, int N, const int64_t *action, const float *locs, const float *prize,
                               const float *max_length, float *tour_length, float *total_prize,
                               int64_t *current, uint8_t *visited, int64_t *i, uint8_t *mask, uint8_t *done,
                               int32_t *status) {
    const int lane = threadIdx.x & 31;
    const int64_t b = blockIdx.x * (int64_t)kOpWarps + (threadIdx.x >> 5);
    if (b >= B) return;
    int64_t a = action[b];
    int st = 0;
    if (a < 0 || a >= N) { st = RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
    const float2 *lrow = reinterpret_cast<const float2 *>(locs) + b * (int64_t)N;
    uint8_t *vrow = visited + b * (int64_t)N;
    int64_t c = current[b];
    c = c < 0 ? 0 : (c >= N ? N - 1 : c);
    const float2 pp = lrow[c], pn = lrow[a];
    const float len = __fadd_rn(tour_length[b], leg_fma(pp.x, pp.y, pn.x, pn.y));      // :78-81
    const float tot = __fadd_rn(total_prize[b], prize[b * (int64_t)N + a]);            // :84-86
    const int64_t step = i[b];
    __syncwarp();
    if (lane == 0) vrow[a] = 1;                                                        // :89
    __syncwarp();
    op_mask_row(N, lane, lrow, max_length + b * (int64_t)N, len, vrow, (int)a, mask + b * (int64_t)N);
    if (lane == 0) {
        tour_length[b] = len;
        total_prize[b] = tot;
        current[b] = a;
        i[b] = step + 1;
        done[b] = (a == 0) && (step > 0);                                              // :92
        raise_status(status, st);
    }
}

// One warp per instance: locs / budgets / actions staged in shared memory, the sequential tour-length and prize
// recurrences run redundantly in every lane, visited set / duplicate check / reward lane-parallel.
__global__ void __launch_bounds__(kOpWarps * 32)
op_replay_kernel(int64_t B, int N, int T, const float *__restrict__ locs, const float *__restrict__ prize,
                 const float *__restrict__ max_length, const int64_t *__restrict__ actions, int64_t *current,
                 float *tour_length, float *total_prize, uint8_t *visited, uint8_t *mask, int64_t *i,
                 uint8_t *done, float *reward, int check, int32_t *status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    const int per_warp = N * 8 + Np * 4 + Np * 4 + Tp * 4 + Np;   // locs | budgets | prize | actions | visited bytes
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *base = smem_raw + (size_t)warp * ((per_warp + 15) & ~15);
    float2 *sl = reinterpret_cast<float2 *>(base);
    float *sm = reinterpret_cast<float *>(base + N * 8);
    float *sp = reinterpret_cast<float *>(base + N * 8 + Np * 4);
    int *sa = reinterpret_cast<int *>(base + N * 8 + Np * 8);
    uint8_t *sv = base + N * 8 + Np * 8 + Tp * 4;

    const int64_t b = blockIdx.x * (int64_t)kOpWarps + warp;
    if (b >= B) return;
    const float2 *lrow = reinterpret_cast<const float2 *>(locs) + b * (int64_t)N;
    for (int k = lane; k < Np; k += 32) sv[k] = 0;
    for (int k = lane; k < N; k += 32) {
        sl[k] = __ldg(lrow + k);
        sm[k] = __ldg(max_length + b * (int64_t)N + k);
        sp[k] = __ldg(prize + b * (int64_t)N + k);
    }
    const int64_t *arow = actions + b * (int64_t)T;
    int st = 0;
    for (int t = lane; t < T; t += 32) {
        int64_t a = __ldg(arow + t);
        if (a < 0 || a >= N) { st |= RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
        sa[t] = (int)a;
    }
    __syncwarp();

    // visited + "Duplicates" (op/env.py:194-199): every customer at most once
    int dup = 0;
    float got = 0.f;
    for (int t = lane; t < T; t += 32) {
        const int a = sa[t];
        const unsigned old = atomicOr(reinterpret_cast<unsigned *>(sv + (a & ~3)), 1u << (8 * (a & 3)));
        if (a != 0) dup |= (old >> (8 * (a & 3))) & 1u;
        got += sp[a];                                   // reward = sum of the gathered prizes (:179-182)
    }
    got = warp_sum(got);
    dup = __any_sync(kFull, dup);
    __syncwarp();

    // env recurrences (op/env.py:78-86), every lane
    float len = 0.f, tot = 0.f;
    int cur = 0;
    for (int t = 0; t < T; ++t) {
        const int a = sa[t];
        const float2 pp = sl[cur], pn = sl[a];
        len = __fadd_rn(len, leg_fma(pp.x, pp.y, pn.x, pn.y));
        tot = __fadd_rn(tot, sp[a]);
        cur = a;
    }
    // validity (:201-215): closed tour over the actions against every node's budget with the depot leg added back
    int over = 0;
    if (check) {
        float closed_len = 0.f;
        for (int t = lane; t < T; t += 32) {
            const float2 p = sl[sa[t]], q = sl[sa[t + 1 == T ? 0 : t + 1]];
            closed_len += leg(p.x, p.y, q.x, q.y);
        }
        closed_len = warp_sum(closed_len);
        for (int n = lane; n < N; n += 32) {
            const float2 p = sl[n];
            const float ml = __fadd_rn(__fadd_rn(sm[n], leg_fma(sl[0].x, sl[0].y, p.x, p.y)), 1e-6f);
            over |= !(closed_len <= __fadd_rn(ml, 1e-5f));
        }
        over = __any_sync(kFull, over);
    }

    if (visited) for (int n = lane; n < N; n += 32) visited[b * (int64_t)N + n] = sv[n];
    op_mask_row(N, lane, sl, sm, len, sv, cur, mask ? mask + b * (int64_t)N : nullptr);
    if (lane == 0) {
        if (current) current[b] = cur;
        if (tour_length) tour_length[b] = len;
        if (total_prize) total_prize[b] = tot;
        if (i) i[b] = T;
        if (done) done[b] = (T > 1) && (cur == 0);
        if (reward) reward[b] = got;
        if (check) st |= (dup ? RL4CO_ST_DUPLICATES : 0) | (over ? RL4CO_ST_OVER_LENGTH : 0);
    }
    st = __reduce_or_sync(kFull, st);
    if (lane == 0) raise_status(status, st);
}

}  // namespace rl4co

using namespace rl4co;

static int op_blocks(int64_t B, unsigned *out) {
    const int64_t blocks = (B + kOpWarps - 1) / kOpWarps;
    if (blocks >= (1ll << 31)) return 1;
    *out = (unsigned)blocks;
    return 0;
}

extern "C" int rl4co_op_action_mask(int64_t B, int32_t N, const float *locs, const float *max_length,
                                    const float *tour_length, const uint8_t *visited,
                                    const int64_t *current_node, uint8_t *action_mask, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1, "op_action_mask: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(locs && max_length && tour_length && visited && current_node && action_mask,
                  "op_action_mask: null pointer");
    unsigned g;
    RL4CO_REQUIRE(op_blocks(B, &g) == 0, "op_action_mask: batch too large");
    op_mask_kernel<<<g, kOpWarps * 32, 0, (cudaStream_t)stream>>>(B, N, locs, max_length, tour_length, visited,
                                                                   current_node, action_mask);
    RL4CO_CHECK_LAUNCH("op_action_mask");
    return 0;
}

extern "C" int rl4co_op_step(int64_t B, int32_t N, const int64_t *action, const float *locs, const float *prize,
                             const float *max_length, float *tour_length, float *current_total_prize,
                             int64_t *current_node, uint8_t *visited, int64_t *i, uint8_t *action_mask,
                             uint8_t *done, int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1, "op_step: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(action && locs && prize && max_length && tour_length && current_total_prize && current_node &&
                      visited && i && action_mask && done, "op_step: null pointer");
    unsigned g;
    RL4CO_REQUIRE(op_blocks(B, &g) == 0, "op_step: batch too large");
    op_step_kernel<<<g, kOpWarps * 32, 0, (cudaStream_t)stream>>>(B, N, action, locs, prize, max_length, tour_length,
                                                                   current_total_prize, current_node, visited, i,
                                                                   action_mask, done, status);
    RL4CO_CHECK_LAUNCH("op_step");
    return 0;
}

extern "C" int rl4co_op_replay(int64_t B, int32_t N, int32_t T, const float *locs, const float *prize,
                               const float *max_length, const int64_t *actions, int64_t *current_node,
                               float *tour_length, float *current_total_prize, uint8_t *visited,
                               uint8_t *action_mask, int64_t *i, uint8_t *done, float *reward,
                               int32_t check_validity, int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1 && T >= 0, "op_replay: bad sizes B=%lld N=%d T=%d", (long long)B, N, T);
    if (B == 0) return 0;
    RL4CO_REQUIRE(locs && prize && max_length && actions, "op_replay: null input");
    const int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    const int per_warp = (N * 8 + Np * 8 + Tp * 4 + Np + 15) & ~15;
    const size_t smem = (size_t)per_warp * kOpWarps;
    RL4CO_REQUIRE(smem <= 227 * 1024, "op_replay: N=%d T=%d needs %zu B shared memory", N, T, smem);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(op_replay_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    unsigned g;
    RL4CO_REQUIRE(op_blocks(B, &g) == 0, "op_replay: batch too large");
    op_replay_kernel<<<g, kOpWarps * 32, smem, (cudaStream_t)stream>>>(B, N, T, locs, prize, max_length, actions,
                                                                        current_node, tour_length, current_total_prize,
                                                                        visited, action_mask, i, done, reward,
                                                                        check_validity, status);
    RL4CO_CHECK_LAUNCH("op_replay");
    return 0;
}
