// Standalone decode step on caller-provided logits, and the reward / baseline reductions.
// Reference: rl4co/utils/decoding.py:141-191,327-397 (process_logits, step, greedy, sampling);
// rl4co/models/rl/reinforce/baselines.py:57-61,186-224; rl4co/models/rl/reinforce/reinforce.py:103-106;
// rl4co/models/zoo/pomo/model.py:117-141.
#include <math.h>

#include "common.cuh"

namespace rl4co {

__device__ __forceinline__ uint4 philox_round10(uint4 ctr, uint2 key) {
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

__device__ __forceinline__ float proc_logit(float x, bool ok, float clip, float temp, bool mask_logits) {
    if (clip > 0.f) x = tanhf(x) * clip;            // decoding.py:172-173
    if (mask_logits && !ok) x = -INFINITY;          // :176-178
    return __fdiv_rn(x, temp);                      // :180
}

// one warp per row; N arbitrary (strided passes over the row, L1-resident)
__global__ void decode_step_kernel(int64_t B, int N, const float *__restrict__ logits,
                                   const uint8_t *__restrict__ mask, int mode, float temp,
                                   float clip, int mask_logits, const int64_t *action_in,
                                   uint64_t seed, uint64_t offset, int64_t *action, float *logp,
                                   float *logp_all, int32_t *status) {
    const int lane = threadIdx.x & 31;
    const int64_t b = blockIdx.x * (int64_t)(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (b >= B) return;
    const float *row = logits + b * (int64_t)N;
    const uint8_t *mrow = mask + b * (int64_t)N;
    // pass 1: max and first argmax
    float m = -INFINITY;
    int bi = 0x7fffffff;
    for (int n = lane; n < N; n += 32) {
        float x = proc_logit(row[n], mrow[n] != 0, clip, temp, mask_logits != 0);
        if (x > m || (x == m && n < bi)) { m = x; bi = n; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float om = __shfl_xor_sync(kFull, m, o);
        int oi = __shfl_xor_sync(kFull, bi, o);
        if (om > m || (om == m && oi < bi)) { m = om; bi = oi; }
    }
    // pass 2: sum exp
    float sum = 0.f;
    for (int n = lane; n < N; n += 32) {
        float x = proc_logit(row[n], mrow[n] != 0, clip, temp, mask_logits != 0);
        sum += (x == -INFINITY) ? 0.f : expf(x - m);
    }
    sum = warp_sum(sum);
    const float lse = logf(sum);
    int act = bi;
    int st = 0;
    if (mode == RL4CO_MODE_EVALUATE) {
        int64_t g = action_in[b];
        if (g < 0 || g >= N) { st |= RL4CO_ST_ACTION_RANGE; g = g < 0 ? 0 : N - 1; }
        act = (int)g;
    } else if (mode == RL4CO_MODE_SAMPLING) {
        uint4 c = make_uint4((uint32_t)offset, (uint32_t)(offset >> 32), (uint32_t)b, (uint32_t)(b >> 32));
        uint4 o = philox_round10(c, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
        float u = (float)(o.x >> 8) * (1.0f / 16777216.0f) * sum;
        // sequential inverse CDF in index order, 32 nodes per pass
        float base = 0.f;
        int pick = -1, lastf = -1;
        for (int n0 = 0; n0 < N && pick < 0; n0 += 32) {
            int n = n0 + lane;
            float e = 0.f;
            if (n < N) {
                float x = proc_logit(row[n], mrow[n] != 0, clip, temp, mask_logits != 0);
                e = (x == -INFINITY) ? 0.f : expf(x - m);
            }
            float incl = e;
#pragma unroll
            for (int o2 = 1; o2 < 32; o2 <<= 1) {
                float up = __shfl_up_sync(kFull, incl, o2);
                if (lane >= o2) incl += up;
            }
            unsigned hit = __ballot_sync(kFull, e > 0.f && base + incl > u);
            unsigned pos = __ballot_sync(kFull, e > 0.f);
            if (pos) lastf = n0 + 31 - __clz(pos);
            if (hit) pick = n0 + __ffs(hit) - 1;
            base += __shfl_sync(kFull, incl, 31);
        }
        act = pick >= 0 ? pick : (lastf >= 0 ? lastf : 0);
    }
    float xa = proc_logit(row[act], mrow[act] != 0, clip, temp, mask_logits != 0);
    float lpa = (xa - m) - lse;
    if (logp_all)
        for (int n = lane; n < N; n += 32) {
            float x = proc_logit(row[n], mrow[n] != 0, clip, temp, mask_logits != 0);
            logp_all[b * (int64_t)N + n] = (x - m) - lse;
        }
    if (lane == 0) {
        if (mask_logits && mrow[act] == 0) st |= RL4CO_ST_INFEASIBLE_ACTION;   // decoding.py:377-379
        action[b] = act;
        if (logp) logp[b] = lpa;
        raise_status(status, st);
    }
}

// ---------------------------------------------------------------- reductions -------
__global__ void reward_stats_kernel(int64_t n, const float *__restrict__ reward, double *stats) {
    __shared__ double sh[2][8];
    double s = 0.0, s2 = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double v = reward[i];
        s += v;
        s2 += v * v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(kFull, s, o);
        s2 += __shfl_xor_sync(kFull, s2, o);
    }
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sh[0][warp] = s; sh[1][warp] = s2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0, b2 = 0;
        for (int w = 0; w < (blockDim.x >> 5); ++w) { a += sh[0][w]; b2 += sh[1][w]; }
        atomicAdd(&stats[0], a);
        atomicAdd(&stats[1], b2);
        if (blockIdx.x == 0) atomicAdd(&stats[2], (double)n);
    }
}

// thread per instance; reward / ll laid out start-major [S*B]
__global__ void shared_baseline_kernel(int64_t B, int S, const float *__restrict__ reward,
                                       const float *__restrict__ ll, float *baseline,
                                       float *advantage, float *max_reward, int64_t *best_start,
                                       double *loss_partial) {
    int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    double lsum = 0.0;
    if (b < B) {
        float sum = 0.f, best = -INFINITY;
        int bi = 0;
        for (int s = 0; s < S; ++s) {
            float rr = reward[(int64_t)s * B + b];
            sum += rr;
            if (rr > best) { best = rr; bi = s; }
        }
        float bl = sum / (float)S;                               // reward.mean(dim=1), baselines.py:60-61
        if (baseline) baseline[b] = bl;
        if (max_reward) max_reward[b] = best;
        if (best_start) best_start[b] = bi;
        for (int s = 0; s < S; ++s) {
            float adv = reward[(int64_t)s * B + b] - bl;         // reinforce.py:103
            if (advantage) advantage[(int64_t)s * B + b] = adv;
            if (ll) lsum += (double)(adv * ll[(int64_t)s * B + b]);
        }
    }
    if (loss_partial) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) lsum += __shfl_xor_sync(kFull, lsum, o);
        if ((threadIdx.x & 31) == 0 && lsum != 0.0) atomicAdd(&loss_partial[0], lsum);
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&loss_partial[1], (double)B * S);
    }
}

}  // namespace rl4co

using namespace rl4co;

extern "C" int rl4co_decode_step(int64_t B, int32_t N, const float *logits, const uint8_t *mask,
                                 int32_t mode, float temperature, float tanh_clipping,
                                 int32_t mask_logits, const int64_t *action_in, uint64_t seed,
                                 uint64_t offset, int64_t *action, float *logp, float *logp_all,
                                 int32_t *status, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N >= 1, "decode_step: bad sizes");
    if (B == 0) return 0;
    RL4CO_REQUIRE(logits && mask && action, "decode_step: null pointer");
    RL4CO_REQUIRE(mode >= RL4CO_MODE_GREEDY && mode <= RL4CO_MODE_EVALUATE, "decode_step: unknown mode %d", mode);
    RL4CO_REQUIRE(mode != RL4CO_MODE_EVALUATE || action_in, "decode_step: evaluate needs action_in");
    RL4CO_REQUIRE(temperature > 0.f, "decode_step: temperature must be > 0");
    int64_t blocks = (B + 7) / 8;
    RL4CO_REQUIRE(blocks < (1ll << 31), "decode_step: batch too large");
    decode_step_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        B, N, logits, mask, mode, temperature, tanh_clipping, mask_logits, action_in, seed, offset,
        action, logp, logp_all, status);
    RL4CO_CHECK_LAUNCH("decode_step");
    return 0;
}

extern "C" int rl4co_reward_stats(int64_t n, const float *reward, double *stats, rl4co_stream_t stream) {
    RL4CO_REQUIRE(n >= 0 && reward && stats, "reward_stats: bad arguments");
    if (n == 0) return 0;
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    reward_stats_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(n, reward, stats);
    RL4CO_CHECK_LAUNCH("reward_stats");
    return 0;
}

extern "C" int rl4co_shared_baseline(int64_t B, int32_t S, const float *reward, const float *ll,
                                     float *baseline, float *advantage, float *max_reward,
                                     int64_t *best_start, double *loss_partial,
                                     rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && S >= 1 && reward, "shared_baseline: bad arguments");
    if (B == 0) return 0;
    int64_t blocks = (B + 127) / 128;
    RL4CO_REQUIRE(blocks < (1ll << 31), "shared_baseline: batch too large");
    shared_baseline_kernel<<<(unsigned)blocks, 128, 0, (cudaStream_t)stream>>>(
        B, S, reward, ll, baseline, advantage, max_reward, best_start, loss_partial);
    RL4CO_CHECK_LAUNCH("shared_baseline");
    return 0;
}
