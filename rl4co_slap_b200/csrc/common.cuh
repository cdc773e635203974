// Shared helpers for the rl4co_slap_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/rl4co_slap_b200.h"

namespace rl4co {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

// thread-local last error string (host side)
void set_error(const char *fmt, ...);

#define RL4CO_REQUIRE(cond, ...)            \
    do {                                    \
        if (!(cond)) {                      \
            rl4co::set_error(__VA_ARGS__);  \
            return 1;                       \
        }                                   \
    } while (0)

#define RL4CO_CHECK_LAUNCH(name)                                                   \
    do {                                                                           \
        cudaError_t e__ = cudaGetLastError();                                      \
        if (e__ != cudaSuccess) {                                                  \
            rl4co::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__)); \
            return 2;                                                              \
        }                                                                          \
    } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFull, v, o));
    return v;
}
__device__ __forceinline__ int warp_sum_int(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

// Euclidean leg exactly as torch's (x - y).norm(p=2, dim=-1) for a 2-vector: no FMA
// contraction so the squared terms are rounded individually like the CPU reference.
__device__ __forceinline__ float leg(float ax, float ay, float bx, float by) {
    float dx = __fsub_rn(bx, ax), dy = __fsub_rn(by, ay);
    return __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
}

// The same leg with the rounding sequence of torch's CPU vector norm for a 2-vector (measured against torch 2.11:
// sqrt(fma(dy, dy, fl(dx dx))), bit-identical on 6464 of 6464 random pairs).  Used where a distance feeds a
// COMPARISON that decides a mask bit (op/env.py:154-170), so that borderline nodes fall on the oracle's side.
__device__ __forceinline__ float leg_fma(float ax, float ay, float bx, float by) {
    float dx = __fsub_rn(bx, ax), dy = __fsub_rn(by, ay);
    return __fsqrt_rn(__fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}

__device__ __forceinline__ void raise_status(int32_t *status, int bits) {
    if (status != nullptr && bits != 0) atomicOr(status, bits);
}

}  // namespace rl4co
