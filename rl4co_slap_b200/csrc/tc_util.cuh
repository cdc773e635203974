// Shared pieces of the tcgen05 (5th-generation tensor core) kernels: TF32 hi/lo split, the 128-byte
// swizzled K-major shared-memory layout and its matrix descriptor, single-lane MMA issue, commit,
// bounded mbarrier waits.  Used by am_project_tc.cu (table projection / linear layers) and
// encoder_attention_tc.cu (Q K^T of the encoder's self-attention).
#pragma once
#include "rollout_util.cuh"

namespace rl4co {
namespace tcu {

// instruction descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptor): D = F32 (bits 4-5 = 1),
// A = B = TF32 (bits 7-9, 10-12 = 2), both K-major, N/8 at bit 17, M/16 at bit 24
__host__ __device__ constexpr uint32_t make_idesc_tf32(int m, int n) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ uint32_t to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
// byte offset of 16-byte chunk `c` of row `r` inside a [rows][128 B] image with the 128B swizzle
// (Swizzle<3,4,3>: chunk index xor row mod 8); images are 1024-byte aligned
__host__ __device__ __forceinline__ uint32_t swz(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }

// shared-memory matrix descriptor (SmemDescriptor): K-major, SWIZZLE_128B, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// descriptors travel as their low words (start address in 16-byte units | LBO); the high word is one constant.
// Everything the instruction reads must sit in uniform registers: keep the operands provably warp-uniform
// (shared-window addresses, constants, a shuffled TMEM base) or the compiler wraps every MMA in an
// elect / broadcast loop that costs more issue time than the MMA takes to execute.
__device__ __forceinline__ bool elect_one() {                  // one lane of a converged warp
    uint32_t pred;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "elect.sync _|P1, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P1;\n\t"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
constexpr uint32_t kDescHi = (uint32_t)((64ull << 32 | 1ull << 46 | 2ull << 61) >> 32);
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint32_t a_lo32, uint32_t b_lo32, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %5};\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "r"(a_lo32), "r"(b_lo32), "r"(idesc), "r"(accumulate), "r"(kDescHi)
        : "memory");
}
// the same for fp16 operands (kind::f16: A = B = F16 = format 0, K = 16 per instruction), fp32 accumulation
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n) {
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint32_t a_lo32, uint32_t b_lo32, uint32_t idesc,
                                         uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %5};\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "r"(a_lo32), "r"(b_lo32), "r"(idesc), "r"(accumulate), "r"(kDescHi)
        : "memory");
}
// the same split by truncation: hi = the top 11 significant bits of x (exactly representable in fp16 while |x| is in fp16's
// normal range), lo = fp16(x - hi) with x - hi exact - no conversion back to fp32; x = hi + lo to 21-22 significant bits
__device__ __forceinline__ void split_f16x2_trunc(float x0, float x1, uint32_t &hi, uint32_t &lo) {
    const float h0 = __uint_as_float(__float_as_uint(x0) & 0xffffe000u), h1 = __uint_as_float(__float_as_uint(x1) & 0xffffe000u);
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(h1), "f"(h0));     // first source -> upper half
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(x1 - h1), "f"(x0 - h0));
}
// two fp32 -> fp16 pairs hi, lo with x = hi + lo to 22 significant bits (lo = fp16(x - hi), exact difference)
__device__ __forceinline__ void split_f16x2(float x0, float x1, uint32_t &hi, uint32_t &lo) {
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(x1), "f"(x0));     // first source -> upper half
    float b0, b1;
    asm("{\n\t"
        ".reg .b16 l, h;\n\t"
        "mov.b32 {l, h}, %2;\n\t"
        "cvt.f32.f16 %0, l;\n\t"
        "cvt.f32.f16 %1, h;\n\t"
        "}\n" : "=f"(b0), "=f"(b1) : "r"(hi));
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(x1 - b1), "f"(x0 - b0));
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {    // arrives when all prior MMAs are done
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// bounded wait: a pipeline bug must end in a trap, not in a hung GPU
__device__ __forceinline__ void mbar_wait_bounded(uint64_t *bar, unsigned parity) {
    const uint32_t addr = smem_u32(bar);
    const long long t0 = clock64();
    while (clock64() - t0 < 4000000000ll) {                     // ~2 s
        uint32_t ok;
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n"
            : "=r"(ok)
            : "r"(addr), "r"(parity)
            : "memory");
        if (ok) return;
    }
    __trap();
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }


}  // namespace tcu
}  // namespace rl4co
