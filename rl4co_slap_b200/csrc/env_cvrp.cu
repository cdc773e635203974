// K1 - CVRP environment kernels (action mask / step / fused replay+reward+validity).
// Reference semantics: rl4co/envs/routing/cvrp/env.py:73-190.
// The capacity arithmetic uses explicit round-to-nearest add/mul intrinsics (no FMA
// contraction, no reassociation) so `used = (used + d) * (a != 0)` and the compare
// `demand + used > capacity` are bit-identical to the fp32 torch ops (SURVEY.md hard part 3).
#include "common.cuh"

namespace rl4co {

constexpr int kCvrpWarps = 8;

// mask for one instance given state; all lanes of the warp participate.
// visited: pointer to N bytes (global or shared); returns through `mask` row (may be null).
__device__ __forceinline__ void cvrp_mask_row(int N, int lane, const float *demand, float used,
                                              float cap, const uint8_t *visited, bool at_depot,
                                              uint8_t *mask) {
    int feasible = 0;
    for (int j = lane; j < N - 1; j += 32) {
        bool exceeds = __fadd_rn(demand[j], used) > cap;
        bool blocked = (visited[j + 1] != 0) | exceeds;
        if (mask) mask[j + 1] = blocked ? 0 : 1;
        feasible |= !blocked;
    }
    feasible = __any_sync(kFull, feasible);
    if (lane == 0 && mask) mask[0] = (at_depot && feasible) ? 0 : 1;
}

__global__ void cvrp_mask_kernel(int64_t B, int N, const float *demand, const float *used,
                                 const float *cap, const uint8_t *visited,
                                 const int64_t *current, uint8_t *mask) {
    int lane = threadIdx.x & 31;
    int64_t b = blockIdx.x * (int64_t)kCvrpWarps + (threadIdx.x >> 5);
    if (b >= B) return;
    cvrp_mask_row(N, lane, demand + b * (int64_t)(N - 1), used[b], cap[b], visited + b * (int64_t)N,
                  current[b] == 0, mask + b * (int64_t)N);
}

__global__ void cvrp_step_kernel(int64_t B, int N, const int64_t *action, const float *demand,
                                 const float *cap, float *used, int64_t *current,
                                 uint8_t *visited, uint8_t *mask, uint8_t *done, int32_t *status) {
    int lane = threadIdx.x & 31;
    int64_t b = blockIdx.x * (int64_t)kCvrpWarps + (threadIdx.x >> 5);
    if (b >= B) return;
    int64_t a = action[b];
    int st = 0;
    if (a < 0 || a >= N) { st = RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
    const float *drow = demand + b * (int64_t)(N - 1);
    uint8_t *vrow = visited + b * (int64_t)N;
    int idx = (int)a - 1;
    idx = idx < 0 ? 0 : (idx > N - 2 ? N - 2 : idx);          // clamp(a-1, 0, n_loc-1), :78
    float u = __fmul_rn(__fadd_rn(used[b], drow[idx]), a != 0 ? 1.f : 0.f);  // :82-84
    __syncwarp();
    if (lane == 0) vrow[a] = 1;                                // scatter(-1, cur, 1), :88
    __syncwarp();
    int cnt = 0;
    for (int n = lane; n < N; n += 32) cnt += vrow[n];         // uint8 sum, :91
    cnt = warp_sum_int(cnt);
    cvrp_mask_row(N, lane, drow, u, cap[b], vrow, a == 0, mask + b * (int64_t)N);
    if (lane == 0) {
        used[b] = u;
        current[b] = a;
        done[b] = (cnt == N);
        raise_status(status, st);
    }
}

// One warp per instance: stage locs / demand / actions in shared memory, run the (sequential)
// capacity recurrences redundantly in every lane, legs and bitset lane-parallel.
__global__ void __launch_bounds__(kCvrpWarps * 32)
cvrp_replay_kernel(int64_t B, int N, int T, const float *__restrict__ locs,
                   const float *__restrict__ demand, const float *__restrict__ cap_in,
                   const int64_t *__restrict__ actions, int64_t *current, float *used_out,
                   uint8_t *visited, uint8_t *mask, uint8_t *done, float *reward, int check,
                   int32_t *status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    const int per_warp = N * 8 + Np * 4 + Tp * 4 + Np;   // locs | demand(+depot slot) | act | visited bytes
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *base = smem_raw + (size_t)warp * ((per_warp + 15) & ~15);
    float2 *sl = reinterpret_cast<float2 *>(base);
    float *sd = reinterpret_cast<float *>(base + N * 8);           // sd[j] = demand of node j+1
    int *sa = reinterpret_cast<int *>(base + N * 8 + Np * 4);
    uint8_t *sv = base + N * 8 + Np * 4 + Tp * 4;

    int64_t b = blockIdx.x * (int64_t)kCvrpWarps + warp;
    if (b >= B) return;
    const float2 *lrow = reinterpret_cast<const float2 *>(locs + b * (int64_t)N * 2);
    for (int k = lane; k < N; k += 32) { sl[k] = __ldg(lrow + k); sv[k] = 0; }
    const float *drow = demand + b * (int64_t)(N - 1);
    for (int k = lane; k < N - 1; k += 32) sd[k] = __ldg(drow + k);
    const int64_t *arow = actions + b * (int64_t)T;
    int st = 0;
    for (int t = lane; t < T; t += 32) {
        int64_t a = __ldg(arow + t);
        if (a < 0 || a >= N) { st |= RL4CO_ST_ACTION_RANGE; a = a < 0 ? 0 : N - 1; }
        sa[t] = (int)a;
    }
    __syncwarp();
    const float cap = cap_in[b];

    // visited + duplicate customers (cvrp/env.py:164-175)
    int bad = 0;
    for (int t = lane; t < T; t += 32) {
        int a = sa[t];
        if (a != 0) {
            // byte-wise exchange through a 32-bit CAS-free path: one lane per distinct t, a
            // duplicate customer shows up as an already-set byte after the warp barrier below
            unsigned old = atomicOr(reinterpret_cast<unsigned *>(sv + (a & ~3)), 1u << (8 * (a & 3)));
            bad |= ((old >> (8 * (a & 3))) & 1u);
        } else {
            atomicOr(reinterpret_cast<unsigned *>(sv), 1u);
        }
    }
    __syncwarp();
    int cnt = 0, cust = 0;
    for (int n = lane; n < N; n += 32) { cnt += sv[n]; cust += (n > 0 && sv[n]); }
    cnt = warp_sum_int(cnt);
    cust = warp_sum_int(cust);
    bad |= (cust != N - 1);

    // env recurrence (cvrp/env.py:82-84) and validity recurrence (:178-190), every lane
    float used = 0.f, chk = 0.f;
    int over = 0;
    for (int t = 0; t < T; ++t) {
        int a = sa[t];
        int idx = a - 1;
        idx = idx < 0 ? 0 : idx;
        used = __fmul_rn(__fadd_rn(used, sd[idx]), a != 0 ? 1.f : 0.f);
        chk = __fadd_rn(chk, a == 0 ? -cap : sd[a - 1]);
        chk = chk < 0.f ? 0.f : chk;
        over |= !(chk <= __fadd_rn(cap, 1e-5f));
    }

    // reward: closed tour depot -> a_0 ... a_{T-1} -> depot (cvrp/env.py:151-160)
    float len = 0.f;
    for (int t = lane; t <= T; t += 32) {
        float2 p = t == 0 ? sl[0] : sl[sa[t - 1]];
        float2 q = t == T ? sl[0] : sl[sa[t]];
        len += leg(p.x, p.y, q.x, q.y);
    }
    len = warp_sum(len);
    bad = __any_sync(kFull, bad);

    int last = T > 0 ? sa[T - 1] : 0;
    if (visited) for (int n = lane; n < N; n += 32) visited[b * (int64_t)N + n] = sv[n];
    cvrp_mask_row(N, lane, sd, used, cap, sv, last == 0, mask ? mask + b * (int64_t)N : nullptr);
    if (lane == 0) {
        if (current) current[b] = last;
        if (used_out) used_out[b] = used;
        if (done) done[b] = (T > 0) && (cnt == N);
        if (reward) reward[b] = -len;
        if (check) st |= (bad ? RL4CO_ST_INVALID_TOUR : 0) | (over ? RL4CO_ST_OVER_CAPACITY : 0);
    }
    st = __reduce_or_sync(kFull, st);
    if (lane == 0) raise_status(status, st);
}

}  // namespace rl4co

using namespace rl4co;

static int cvrp_blocks(int64_t B, unsigned *out) {
    int64_t blocks = (B + kCvrpWarps - 1) / kCvrpWarps;
    if (blocks >= (1ll << 31)) return 1;
    *out = (unsigned)blocks;
    return 0;
}

extern "C" int rl4co_cvrp_action_mask(int64_t B, int32_t N, const float *demand,
                                      const float *used_capacity, const float *vehicle_capacity,
                                      const uint8_t *visited, const int64_t *current_node,
                                      uint8_t *action_mask, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1, "cvrp_action_mask: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(demand && used_capacity && vehicle_capacity && visited && current_node && action_mask,
                  "cvrp_action_mask: null pointer");
    unsigned g;
    RL4CO_REQUIRE(cvrp_blocks(B, &g) == 0, "cvrp_action_mask: batch too large");
    cvrp_mask_kernel<<<g, kCvrpWarps * 32, 0, (cudaStream_t)stream>>>(
        B, N, demand, used_capacity, vehicle_capacity, visited, current_node, action_mask);
    RL4CO_CHECK_LAUNCH("cvrp_action_mask");
    return 0;
}

extern "C" int rl4co_cvrp_step(int64_t B, int32_t N, const int64_t *action, const float *demand,
                               const float *vehicle_capacity, float *used_capacity,
                               int64_t *current_node, uint8_t *visited, uint8_t *action_mask,
                               uint8_t *done, int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1, "cvrp_step: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(action && demand && vehicle_capacity && used_capacity && current_node && visited &&
                      action_mask && done, "cvrp_step: null pointer");
    unsigned g;
    RL4CO_REQUIRE(cvrp_blocks(B, &g) == 0, "cvrp_step: batch too large");
    cvrp_step_kernel<<<g, kCvrpWarps * 32, 0, (cudaStream_t)stream>>>(
        B, N, action, demand, vehicle_capacity, used_capacity, current_node, visited, action_mask,
        done, status);
    RL4CO_CHECK_LAUNCH("cvrp_step");
    return 0;
}

extern "C" int rl4co_cvrp_replay(int64_t B, int32_t N, int32_t T, const float *locs,
                                 const float *demand, const float *vehicle_capacity,
                                 const int64_t *actions, int64_t *current_node,
                                 float *used_capacity, uint8_t *visited, uint8_t *action_mask,
                                 uint8_t *done, float *reward, int32_t check_validity,
                                 int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N > 1 && T >= 0, "cvrp_replay: bad sizes B=%lld N=%d T=%d", (long long)B, N, T);
    if (B == 0) return 0;
    RL4CO_REQUIRE(locs && demand && vehicle_capacity && actions, "cvrp_replay: null input");
    int Tp = (T + 3) & ~3, Np = (N + 3) & ~3;
    int per_warp = (N * 8 + Np * 4 + Tp * 4 + Np + 15) & ~15;
    size_t smem = (size_t)per_warp * kCvrpWarps;
    RL4CO_REQUIRE(smem <= 227 * 1024, "cvrp_replay: N=%d T=%d needs %zu B shared memory", N, T, smem);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(cvrp_replay_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    unsigned g;
    RL4CO_REQUIRE(cvrp_blocks(B, &g) == 0, "cvrp_replay: batch too large");
    cvrp_replay_kernel<<<g, kCvrpWarps * 32, smem, (cudaStream_t)stream>>>(
        B, N, T, locs, demand, vehicle_capacity, actions, current_node, used_capacity, visited,
        action_mask, done, reward, check_validity, status);
    RL4CO_CHECK_LAUNCH("cvrp_replay");
    return 0;
}
