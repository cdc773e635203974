// POMO validation / test epilogue: the steps either side of a multistart rollout.
//
//   before:  instance augmentation of the coordinates that feed the encoder
//            - dihedral-8 (rl4co/data/transforms.py:15-47): the eight symmetries of the unit square,
//            - symmetric   (transforms.py:50-89): rotation by phi about (0.5, 0.5), axes swapped for phi > 2 pi;
//   after:   best-of-starts and best-of-augmentations with the tours that achieved them
//            (rl4co/models/zoo/pomo/model.py:113-141, rl4co/utils/decoding.py:399-407).
//
// Rollout order is the reference's batchify order (utils/ops.py:11-34) applied as (n_aug, n_start):
// rollout r = (s * A + a) * B + b is start s of augmented copy a of instance b.
// All three kernels are HBM-bound elementwise / segmented-max work: coalesced along b (rewards) and along
// the tour (actions), no shared-memory staging of data, grid sized from B.
#include "common.cuh"

namespace rl4co {
namespace {

// ---------------------------------------------------------------- dihedral-8 ----------
// One thread per input point; eight float2 stores, each coalesced across the warp.
__global__ void __launch_bounds__(256)
dihedral8_kernel(int64_t n_points, const float2 *__restrict__ in, float2 *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_points) return;
    const float2 p = __ldg(in + i);
    const float x = p.x, y = p.y, fx = __fsub_rn(1.f, p.x), fy = __fsub_rn(1.f, p.y);
    // block order of transforms.py:26-36
    out[i] = make_float2(x, y);
    out[n_points + i] = make_float2(fx, y);
    out[2 * n_points + i] = make_float2(x, fy);
    out[3 * n_points + i] = make_float2(fx, fy);
    out[4 * n_points + i] = make_float2(y, x);
    out[5 * n_points + i] = make_float2(fy, x);
    out[6 * n_points + i] = make_float2(y, fx);
    out[7 * n_points + i] = make_float2(fy, fx);
}

// ---------------------------------------------------------------- symmetric -----------
// out = R(phi) (p - 0.5) + 0.5, coordinates swapped when phi > 2 pi.  Products and the sum are rounded
// separately (torch evaluates cos * x, sin * y and their difference as three ops).
__global__ void __launch_bounds__(256)
symmetric_kernel(int64_t rows, int N, const float2 *__restrict__ in, const float *__restrict__ phi,
                 float2 *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * N) return;
    const float ph = __ldg(phi + i / N);
    const float c = cosf(ph), s = sinf(ph);
    const float2 p = __ldg(in + i);
    const float x = __fsub_rn(p.x, 0.5f), y = __fsub_rn(p.y, 0.5f);
    const float xr = __fsub_rn(__fmul_rn(c, x), __fmul_rn(s, y));
    const float yr = __fadd_rn(__fmul_rn(s, x), __fmul_rn(c, y));
    const bool swap = ph > 6.283185307179586f;
    out[i] = make_float2(__fadd_rn(swap ? yr : xr, 0.5f), __fadd_rn(swap ? xr : yr, 0.5f));
}

// ---------------------------------------------------------------- best of starts / augmentations
constexpr int kBestThreads = 128;

// Phase 1: thread = instance b.  For every augmented copy a the maximum over the S starts (lowest start on
// ties, like torch.max), then the maximum over the copies of those maxima (lowest copy on ties).  Reads
// reward[(s*A + a)*B + b]: consecutive threads read consecutive addresses.
// Phase 2: the block copies the winning tours - a warp per (instance, copy) row, lanes along the tour.
__global__ void __launch_bounds__(kBestThreads)
pomo_best_kernel(int64_t B, int A, int S, int T, const float *__restrict__ reward, const int64_t *__restrict__ actions,
                 float *__restrict__ max_reward, int64_t *__restrict__ best_start,
                 int64_t *__restrict__ best_ms_actions, float *__restrict__ max_aug_reward,
                 int64_t *__restrict__ best_aug, int64_t *__restrict__ best_aug_actions) {
    extern __shared__ int s_start[];                    // [kBestThreads][A] winning start per (instance, copy)
    __shared__ int s_aug[kBestThreads];                 // winning copy per instance
    const int64_t b0 = (int64_t)blockIdx.x * kBestThreads;
    const int64_t b = b0 + threadIdx.x;
    if (b < B) {
        float best_a_val = -INFINITY;
        int best_a = 0;
        for (int a = 0; a < A; ++a) {
            float v = __ldg(reward + (int64_t)a * B + b);
            int arg = 0;
            for (int s = 1; s < S; ++s) {
                const float r = __ldg(reward + ((int64_t)s * A + a) * B + b);
                if (r > v) { v = r; arg = s; }
            }
            s_start[threadIdx.x * A + a] = arg;
            if (max_reward) max_reward[b * A + a] = v;
            if (best_start) best_start[b * A + a] = arg;
            if (a == 0 || v > best_a_val) { best_a_val = v; best_a = a; }
        }
        s_aug[threadIdx.x] = best_a;
        if (max_aug_reward) max_aug_reward[b] = best_a_val;
        if (best_aug) best_aug[b] = best_a;
    }
    __syncthreads();
    if (actions == nullptr || (best_ms_actions == nullptr && best_aug_actions == nullptr)) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_local = (int)((B - b0) < kBestThreads ? (B - b0) : kBestThreads);
    for (int item = warp; item < n_local * (A + 1); item += kBestThreads / 32) {
        const int bl = item / (A + 1), a_slot = item % (A + 1);
        const int64_t bb = b0 + bl;
        int64_t *dst;
        int a;
        if (a_slot < A) {                               // best start of copy a
            if (best_ms_actions == nullptr) continue;
            a = a_slot;
            dst = best_ms_actions + (bb * A + a) * T;
        } else {                                        // best (copy, start) overall
            if (best_aug_actions == nullptr) continue;
            a = s_aug[bl];
            dst = best_aug_actions + bb * T;
        }
        const int64_t r = ((int64_t)s_start[bl * A + a] * A + a) * B + bb;
        const int64_t *src = actions + r * T;
        for (int t = lane; t < T; t += 32) dst[t] = __ldg(src + t);
    }
}

}  // namespace
}  // namespace rl4co

using namespace rl4co;

extern "C" int rl4co_dihedral8_augment(int64_t B, int32_t N, const float *locs, float *out, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N >= 1, "dihedral8_augment: bad sizes B=%lld N=%d", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(locs && out, "dihedral8_augment: null pointer");
    RL4CO_REQUIRE(((uintptr_t)locs & 7) == 0 && ((uintptr_t)out & 7) == 0, "dihedral8_augment: pointers must be 8-byte aligned");
    const int64_t n = B * N;
    const int64_t grid = (n + 255) / 256;
    RL4CO_REQUIRE(grid < (1ll << 31), "dihedral8_augment: batch too large");
    dihedral8_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(n, reinterpret_cast<const float2 *>(locs),
                                                                       reinterpret_cast<float2 *>(out));
    RL4CO_CHECK_LAUNCH("dihedral8_augment");
    return 0;
}

extern "C" int rl4co_symmetric_augment(int64_t rows, int32_t N, const float *xy, const float *phi, float *out,
                                       rl4co_stream_t stream) {
    RL4CO_REQUIRE(rows >= 0 && N >= 1, "symmetric_augment: bad sizes rows=%lld N=%d", (long long)rows, N);
    if (rows == 0) return 0;
    RL4CO_REQUIRE(xy && phi && out, "symmetric_augment: null pointer");
    RL4CO_REQUIRE(((uintptr_t)xy & 7) == 0 && ((uintptr_t)out & 7) == 0, "symmetric_augment: pointers must be 8-byte aligned");
    const int64_t grid = (rows * N + 255) / 256;
    RL4CO_REQUIRE(grid < (1ll << 31), "symmetric_augment: batch too large");
    symmetric_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(rows, N, reinterpret_cast<const float2 *>(xy), phi,
                                                                       reinterpret_cast<float2 *>(out));
    RL4CO_CHECK_LAUNCH("symmetric_augment");
    return 0;
}

extern "C" int rl4co_pomo_best(int64_t B, int32_t A, int32_t S, int32_t T, const float *reward, const int64_t *actions,
                               float *max_reward, int64_t *best_start, int64_t *best_multistart_actions,
                               float *max_aug_reward, int64_t *best_aug, int64_t *best_aug_actions,
                               rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && A >= 1 && S >= 1 && T >= 0, "pomo_best: bad sizes B=%lld A=%d S=%d T=%d", (long long)B, A, S, T);
    if (B == 0) return 0;
    RL4CO_REQUIRE(reward != nullptr, "pomo_best: null reward");
    RL4CO_REQUIRE(actions != nullptr || (best_multistart_actions == nullptr && best_aug_actions == nullptr),
                  "pomo_best: tour outputs requested without actions");
    RL4CO_REQUIRE(A <= 64, "pomo_best: at most 64 augmented copies, got %d", A);
    const int64_t grid = (B + kBestThreads - 1) / kBestThreads;
    RL4CO_REQUIRE(grid < (1ll << 31), "pomo_best: batch too large");
    const size_t smem = (size_t)kBestThreads * A * sizeof(int);
    pomo_best_kernel<<<(unsigned)grid, kBestThreads, smem, (cudaStream_t)stream>>>(
        B, A, S, T, reward, actions, max_reward, best_start, best_multistart_actions, max_aug_reward, best_aug,
        best_aug_actions);
    RL4CO_CHECK_LAUNCH("pomo_best");
    return 0;
}
