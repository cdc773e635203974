// K1 - Split-delivery VRP environment kernels (action mask / step / fused replay+reward+validity).
// Reference semantics: rl4co/envs/routing/sdvrp/env.py:63-164.  A customer may be visited several times; each visit
// delivers min(remaining demand, remaining capacity).  All capacity / demand arithmetic is explicit round-to-nearest
// adds and multiplies in the reference's order, so `demand_with_depot`, `used_capacity` and the masks are bit exact.
#include "common.cuh"

namespace rl4co {

constexpr int kSdWarps = 8;

// mask row (sdvrp/env.py:132-140); `rem` = demand_with_depot row (N entries, entry 0 = depot)
__device__ __forceinline__ void sdvrp_mask_row(int N, int lane, const float *rem, float used, float cap, bool at_depot,
                                               uint8_t *mask) {
    int feasible = 0;
    const bool full = used >= cap;
    for (int n = lane + 1; n < N; n += 32) {
        const bool blocked = (rem[n] == 0.f) | full;
        if (mask) mask[n] = blocked ? 0 : 1;
        feasible |= !blocked;
    }
    feasible = __any_sync(kFull, feasible);
    if (lane == 0 && mask) mask[0] = (at_depot && feasible) ? 0 : 1;
}

__global__ void sdvrp_mask_kernel(int64_t B, int N, const float *dwd, const float *used, const float *cap,
                                  const int64_t *current, uint8_t *mask) {
    const int lane = threadIdx.x & 31;
    const int64_t b = blockIdx.x * (int64_t)kSdWarps + (threadIdx.x >> 5);
    if (b >= B) return;
    sdvrp_mask_row(N, lane, dwd + b * (int64_t)N, used[b], cap[b], current[b] == 0, mask + b * (int64_t)N);
}

__global__ void sdvrp_step_kernel(int64_t B, int N, const int64_t *action, const float *cap, float *dwd, float *used,
                                  int64_t *current, uint8_t *mask, uint8_t *done, int32_t *status) {
    const int lane = threadIdx.x & 31;
    const int64_t b = blockIdx.x * (int64_t)kSdWarps + (threadIdx.x >> 5);
    if (b >= B) return;
    int64_t a = action[b];
    int st = 0;
    if (a < 0 || a >= N) { st = RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
    float *row = dwd + b * (int64_t)N;
    const float c = cap[b], u0 = used[b];
    const float delivered = fminf(row[a], __fsub_rn(c, u0));                       // :71-76
    const float u = __fmul_rn(__fadd_rn(u0, delivered), a != 0 ? 1.f : 0.f);       // :79-81
    __syncwarp();
    if (lane == 0) row[a] = __fadd_rn(row[a], -delivered);                         // scatter_add(-delivered), :84-86
    __syncwarp();
    int open = 0;
    for (int n = lane; n < N; n += 32) open |= row[n] > 0.f;                       // :89
    open = __any_sync(kFull, open);
    sdvrp_mask_row(N, lane, row, u, c, a == 0, mask + b * (int64_t)N);
    if (lane == 0) {
        used[b] = u;
        current[b] = a;
        done[b] = !open;
        raise_status(status, st);
    }
}

// One warp per instance: the delivery recurrence is sequential and runs in every lane on a shared-memory copy of the
// demands (lane 0 writes); the validity recurrence of check_solution_validity (:142-164) is the same walk with the
// depot slot holding -capacity.
__global__ void __launch_bounds__(kSdWarps * 32)
sdvrp_replay_kernel(int64_t B, int N, int T, const float *__restrict__ locs, const float *__restrict__ demand,
                    const float *__restrict__ cap_in, const int64_t *__restrict__ actions, int64_t *current,
                    float *used_out, float *dwd_out, uint8_t *mask, uint8_t *done, float *reward, int check,
                    int32_t *status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    const int per_warp = N * 8 + Np * 4 + Tp * 4;        // locs | remaining demand (depot slot 0) | actions
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *base = smem_raw + (size_t)warp * ((per_warp + 15) & ~15);
    float2 *sl = reinterpret_cast<float2 *>(base);
    float *sd = reinterpret_cast<float *>(base + N * 8);
    int *sa = reinterpret_cast<int *>(base + N * 8 + Np * 4);
    const int64_t b = blockIdx.x * (int64_t)kSdWarps + warp;
    if (b >= B) return;
    const float2 *lrow = reinterpret_cast<const float2 *>(locs) + b * (int64_t)N;
    for (int k = lane; k < N; k += 32) {
        sl[k] = __ldg(lrow + k);
        sd[k] = k == 0 ? 0.f : __ldg(demand + b * (int64_t)(N - 1) + k - 1);
    }
    const int64_t *arow = actions + b * (int64_t)T;
    int st = 0;
    for (int t = lane; t < T; t += 32) {
        int64_t a = __ldg(arow + t);
        if (a < 0 || a >= N) { st |= RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
        sa[t] = (int)a;
    }
    __syncwarp();
    const float cap = cap_in[b];
    float used = 0.f;
    int twice = 0, prev = -1, depot_seen = 0;
    for (int t = 0; t < T; ++t) {
        const int a = sa[t];
        if (check && prev == 0 && a == 0) {              // "Cannot visit depot twice if any nonzero demand"
            int open = 0;
            for (int n = lane + 1; n < N; n += 32) open |= sd[n] != 0.f;
            twice |= __any_sync(kFull, open);
        }
        const float delivered = fminf(sd[a], __fsub_rn(cap, used));
        used = __fmul_rn(__fadd_rn(used, delivered), a != 0 ? 1.f : 0.f);
        __syncwarp();
        if (lane == 0) sd[a] = __fadd_rn(sd[a], -delivered);
        __syncwarp();
        depot_seen |= a == 0;
        prev = a;
    }
    int open = 0, left = 0;
    for (int n = lane; n < N; n += 32) { open |= sd[n] > 0.f; left |= (n > 0 && sd[n] != 0.f); }
    open = __any_sync(kFull, open);
    // the reference's walk keeps -capacity in the depot slot until the first depot visit (:150,158-159), so a tour that
    // never returns to the depot fails its final `demands == 0` as well
    left = __any_sync(kFull, left) || !depot_seen;
    float len = 0.f;                                     // cvrp/env.py:151-160 (inherited): depot at both ends
    for (int t = lane; t <= T; t += 32) {
        const float2 p = t == 0 ? sl[0] : sl[sa[t - 1]];
        const float2 q = t == T ? sl[0] : sl[sa[t]];
        len += leg(p.x, p.y, q.x, q.y);
    }
    len = warp_sum(len);
    const int last = T > 0 ? sa[T - 1] : 0;
    if (dwd_out) for (int n = lane; n < N; n += 32) dwd_out[b * (int64_t)N + n] = sd[n];
    sdvrp_mask_row(N, lane, sd, used, cap, last == 0, mask ? mask + b * (int64_t)N : nullptr);
    if (lane == 0) {
        if (current) current[b] = last;
        if (used_out) used_out[b] = used;
        if (done) done[b] = (T > 0) && !open;
        if (reward) reward[b] = -len;
        if (check) st |= (twice ? RL4CO_ST_DEPOT_TWICE : 0) | (left ? RL4CO_ST_DEMAND_LEFT : 0);
    }
    st = __reduce_or_sync(kFull, st);
    if (lane == 0) raise_status(status, st);
}

}  // namespace rl4co

using namespace rl4co;

static int sd_blocks(int64_t B, unsigned *out) {
    const int64_t blocks = (B + kSdWarps - 1) / kSdWarps;
    if (blocks >= (1ll << 31)) return 1;
    *out = (unsigned)blocks;
    return 0;
}

extern "C" int rl4co_sdvrp_action_mask(int64_t B, int32_t N, const float *demand_with_depot, const float *used_capacity,
                                       const float *vehicle_capacity, const int64_t *current_node, uint8_t *action_mask,
                                       rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1, "sdvrp_action_mask: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(demand_with_depot && used_capacity && vehicle_capacity && current_node && action_mask,
                  "sdvrp_action_mask: null pointer");
    unsigned g;
    RL4CO_REQUIRE(sd_blocks(B, &g) == 0, "sdvrp_action_mask: batch too large");
    sdvrp_mask_kernel<<<g, kSdWarps * 32, 0, (cudaStream_t)stream>>>(B, N, demand_with_depot, used_capacity,
                                                                      vehicle_capacity, current_node, action_mask);
    RL4CO_CHECK_LAUNCH("sdvrp_action_mask");
    return 0;
}

extern "C" int rl4co_sdvrp_step(int64_t B, int32_t N, const int64_t *action, const float *vehicle_capacity,
                                float *demand_with_depot, float *used_capacity, int64_t *current_node,
                                uint8_t *action_mask, uint8_t *done, int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1, "sdvrp_step: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(action && vehicle_capacity && demand_with_depot && used_capacity && current_node && action_mask && done,
                  "sdvrp_step: null pointer");
    unsigned g;
    RL4CO_REQUIRE(sd_blocks(B, &g) == 0, "sdvrp_step: batch too large");
    sdvrp_step_kernel<<<g, kSdWarps * 32, 0, (cudaStream_t)stream>>>(B, N, action, vehicle_capacity, demand_with_depot,
                                                                      used_capacity, current_node, action_mask, done, status);
    RL4CO_CHECK_LAUNCH("sdvrp_step");
    return 0;
}

extern "C" int rl4co_sdvrp_replay(int64_t B, int32_t N, int32_t T, const float *locs, const float *demand,
                                  const float *vehicle_capacity, const int64_t *actions, int64_t *current_node,
                                  float *used_capacity, float *demand_with_depot, uint8_t *action_mask, uint8_t *done,
                                  float *reward, int32_t check_validity, int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1 && T >= 0, "sdvrp_replay: bad sizes B=%lld N=%d T=%d", (long long)B, N, T);
    if (B == 0) return 0;
    RL4CO_REQUIRE(locs && demand && vehicle_capacity && actions, "sdvrp_replay: null input");
    const int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    const int per_warp = (N * 8 + Np * 4 + Tp * 4 + 15) & ~15;
    const size_t smem = (size_t)per_warp * kSdWarps;
    RL4CO_REQUIRE(smem <= 227 * 1024, "sdvrp_replay: N=%d T=%d needs %zu B shared memory", N, T, smem);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(sdvrp_replay_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    unsigned g;
    RL4CO_REQUIRE(sd_blocks(B, &g) == 0, "sdvrp_replay: batch too large");
    sdvrp_replay_kernel<<<g, kSdWarps * 32, smem, (cudaStream_t)stream>>>(B, N, T, locs, demand, vehicle_capacity, actions,
                                                                           current_node, used_capacity, demand_with_depot,
                                                                           action_mask, done, reward, check_validity, status);
    RL4CO_CHECK_LAUNCH("sdvrp_replay");
    return 0;
}
