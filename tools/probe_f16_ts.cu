// probe: tcgen05.mma kind::f16 with A from TENSOR MEMORY (two fp16 packed per 32-bit column) and B from shared
// memory (K-major, 128-byte swizzled rows of 64 fp16), D fp32 in TMEM.  This is the operand arrangement of the
// multistart rollout kernel (am_rollout_ms_tc.cu): A = per-start rows (query, softmax weights, glimpse) written by
// the thread that owns the TMEM lane, B = per-instance tables resident in shared memory.
//   case 0: M128 N112 K16, B rows 0..111, K slice `ks` of a 64-wide row        (scores / logits k-step)
//   case 1: M128 N16  K16, B rows 16 j .. 16 j + 15 of a [128 x 64] tile       (P V of head j)
//   case 2: three accumulating MMAs over k-steps in two 64-wide tiles          (K loop + tile switch)
// nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probe_f16_ts tools/probe_f16_ts.cu && tools/probe_f16_ts
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
// instruction descriptor: D = F32 (bits 4-5 = 1), A = B = F16 (format 0), K-major both, N/8 at bit 17, M/16 at bit 24
__host__ __device__ constexpr uint32_t idesc_f16(int m, int n) {
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr) {     // K-major, SWIZZLE_128B, 8-row groups 1024 B apart
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc),
        "r"(idesc), "r"(acc)
        : "memory");
}

constexpr int NB = 128;                 // rows of the B tile in shared memory
constexpr int TILE = NB * 128;          // bytes per 64-wide K tile

// A [128][32] fp16 (two k-steps), B [2 tiles][128 rows][64] fp16, out [128][112] fp32
__global__ void __launch_bounds__(128) probe(const __half *A, const __half *Bm, float *out, int mode, int ks, int j) {
    extern __shared__ unsigned char raw[];
    unsigned char *smem = (unsigned char *)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    // B -> swizzled image: element (tile, n, k): byte n*128 + (((k >> 3) ^ (n & 7)) << 4) + (k & 7) * 2
    for (int e = tid; e < 2 * NB * 64; e += 128) {
        const int tile = e / (NB * 64), n = (e / 64) % NB, k = e % 64;
        *(__half *)(smem + tile * TILE + n * 128 + ((((k >> 3) ^ (n & 7))) << 4) + (k & 7) * 2) = Bm[e];
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(256u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot;
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * warp) << 16);
    // A row `tid` -> TMEM columns 0..15: column c holds fp16 elements 2c (low half) and 2c + 1 (high half)
    {
        uint32_t v[16];
        const uint32_t *row = (const uint32_t *)(A + tid * 32);
        for (int c = 0; c < 16; ++c) v[c] = row[c];
        asm volatile(
            "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
            ::"r"(lane_addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
            "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t D = tmem + 64;       // accumulator columns
    if (tid == 0) {
        const uint32_t base = smem_u32(smem);
        if (mode == 0)          // A k-step 0 x B[rows 0..111][k slice ks of tile 0]
            mma_f16_ts(D, tmem, smem_desc(base + ks * 32), idesc_f16(128, 112), 0u);
        else if (mode == 1)     // A k-step 0 x B[rows 16 j ..][k slice ks of tile 1]
            mma_f16_ts(D, tmem, smem_desc(base + TILE + j * 16 * 128 + ks * 32), idesc_f16(128, 16), 0u);
        else {                  // A k-steps 0, 1 against tile 0 slice 3 and tile 1 slice 0, accumulated; then k-step 0 again
            mma_f16_ts(D, tmem, smem_desc(base + 3 * 32), idesc_f16(128, 112), 0u);
            mma_f16_ts(D, tmem + 8, smem_desc(base + TILE), idesc_f16(128, 112), 1u);
            mma_f16_ts(D, tmem, smem_desc(base + 1 * 32), idesc_f16(128, 112), 1u);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    {
        uint32_t ok = 0;
        const long long t0 = clock64();
        while (!ok && clock64() - t0 < 2000000000ll)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                         : "=r"(ok) : "r"(smem_u32(&bar)) : "memory");
        if (!ok) __trap();
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int c = 0; c < 112; c += 16) {
        uint32_t v[16];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
              "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
            : "r"(lane_addr + 64 + c));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 16; ++i) out[tid * 112 + c + i] = __uint_as_float(v[i]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

int main() {
    const int nA = 128 * 32, nB = 2 * NB * 64;
    __half *hA = (__half *)malloc(nA * 2), *hB = (__half *)malloc(nB * 2);
    srand(7);
    for (int i = 0; i < nA; ++i) hA[i] = __float2half((rand() % 2001 - 1000) / 500.f);
    for (int i = 0; i < nB; ++i) hB[i] = __float2half((rand() % 2001 - 1000) / 700.f);
    __half *dA, *dB;
    float *dO, *hO = (float *)malloc(128 * 112 * 4);
    cudaMalloc(&dA, nA * 2); cudaMalloc(&dB, nB * 2); cudaMalloc(&dO, 128 * 112 * 4);
    cudaMemcpy(dA, hA, nA * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, nB * 2, cudaMemcpyHostToDevice);
    const int smem = 2 * TILE + 1024;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    auto Af = [&](int m, int k) { return __half2float(hA[m * 32 + k]); };
    auto Bf = [&](int tile, int n, int k) { return __half2float(hB[(tile * NB + n) * 64 + k]); };
    int bad = 0;
    for (int mode = 0; mode < 3; ++mode) {
        const int ks = mode == 2 ? 0 : 2, j = 5;
        cudaMemset(dO, 0xff, 128 * 112 * 4);
        probe<<<1, 128, smem>>>(dA, dB, dO, mode, ks, j);
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(hO, dO, 128 * 112 * 4, cudaMemcpyDeviceToHost);
        const int ncol = mode == 1 ? 16 : 112;
        double worst = 0;
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < ncol; ++n) {
                double ref = 0;
                for (int k = 0; k < 16; ++k) {
                    if (mode == 0) ref += (double)Af(m, k) * Bf(0, n, ks * 16 + k);
                    else if (mode == 1) ref += (double)Af(m, k) * Bf(1, j * 16 + n, ks * 16 + k);
                    else ref += (double)Af(m, k) * Bf(0, n, 48 + k) + (double)Af(m, 16 + k) * Bf(1, n, k) + (double)Af(m, k) * Bf(0, n, 16 + k);
                }
                const double err = fabs(ref - hO[m * 112 + n]);
                if (!(err <= 1e-4 * (1 + fabs(ref)))) { if (bad < 5) printf("  mode %d m %d n %d want %g got %g\n", mode, m, n, ref, hO[m * 112 + n]); ++bad; }
                if (err > worst) worst = err;
            }
        printf("mode %d: %s, worst |err| %.3g\n", mode, cudaGetErrorString(e), worst);
    }
    printf(bad ? "PROBE FAILED (%d mismatches)\n" : "PROBE OK\n", bad);
    return bad != 0;
}
