// K3 (tensor-core path) - projection of the encoder's node embeddings into the rollout tables
// on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM), fp32-accurate through
// a 3-term TF32 split:   a*w = a_hi*w_hi + a_hi*w_lo + a_lo*w_hi  (+ O(2^-22) dropped),
// with a_hi = rna_tf32(a), a_lo = rna_tf32(a - a_hi), products accumulated in fp32 in TMEM.
// Replaces AttentionModelDecoder._precompute_cache (rl4co/models/zoo/am/decoder.py:208-237) and
// the per-node halves of the context projection (rl4co/models/nn/env_embeddings/context.py:67-71).
//
//   tables[t][m][j] = sum_e h[m][e] * W[t*128 + j][e]        (both operands K-major)
//
// Work split: one persistent CTA per (group, table).  The CTA keeps ITS table's weights - already
// split into hi / lo and laid out in the 128-byte-swizzled K-major image the MMA reads
// (rl4co_am_pack_weights) - resident in shared memory (128 KB, one TMA bulk load), and streams
// 128-row tiles of h:
//   warps 1-8  producers: coalesced LDG of a 128 x 32 fp32 k-block, hi / lo split, swizzled
//              st.shared into a 3-stage ring, mbarrier full / empty;
//   warp 0     one elected lane issues 48 tcgen05.mma (M128 N128 K8, kind::tf32) per tile and
//              commits to the mbarriers;
//   warps 9-12 epilogue: tcgen05.ld the 128 x 128 fp32 accumulator (double buffered in TMEM,
//              2 x 128 columns) and store it to the table.
// The tables of one instance are read back by the rollout kernel row by row, so the output stays
// row-major [t][m][128].
#include "common.cuh"
#include "rollout_util.cuh"

namespace rl4co {
namespace tc {

constexpr int E = RL4CO_EMBED_DIM;       // K = 128
constexpr int BM = 128;                  // rows per tile (= TMEM lanes)
constexpr int BN = 128;                  // one table
constexpr int KB = 32;                   // fp32 per k-block = one 128-byte swizzle row
constexpr int NKB = E / KB;              // 4
constexpr int KB_BYTES = BM * KB * 4;    // 16 KB: one k-block image of 128 rows
constexpr int W_BYTES = 2 * NKB * KB_BYTES;          // 128 KB: [hi|lo][kb][128 x 128 B]
constexpr int A_STAGES = 2;
constexpr int A_STAGE_BYTES = 2 * KB_BYTES;          // hi + lo
constexpr int kProducerWarps = 8;
constexpr int kThreads = (1 + kProducerWarps + 4) * 32;
constexpr int OUT_BYTES = BM * 128;                    // staging of one 32-column block of the accumulator
constexpr int SMEM_BYTES = 1024 + W_BYTES + A_STAGES * A_STAGE_BYTES + 2 * OUT_BYTES + 256 + 512;
constexpr uint32_t kTmemCols = 256;      // two 128-column accumulators
// instruction descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptor): D = F32 (bits 4-5 = 1),
// A = B = TF32 (bits 7-9, 10-12 = 2), both K-major, N/8 at bit 17, M/16 at bit 24
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((BN >> 3) << 17) | ((BM >> 4) << 24);

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
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(kIdesc), "r"(accumulate)
        : "memory");
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

// weights -> the shared-memory image of one table: [hi|lo][kb][row j = output feature][32 k], swizzled
__global__ void __launch_bounds__(256) pack_weights_kernel(int n_tables, const float *__restrict__ W, int64_t ldw,
                                                           float *__restrict__ image) {
    const int t = blockIdx.x;
    const float *Wt = W + (size_t)t * BN * ldw;
    char *img = reinterpret_cast<char *>(image) + (size_t)t * W_BYTES;
    for (int idx = threadIdx.x; idx < BN * (E / 4); idx += blockDim.x) {
        const int j = idx / (E / 4), c4 = idx % (E / 4);        // row j, float4 chunk c4 of the row
        const float4 w = *reinterpret_cast<const float4 *>(Wt + (size_t)j * ldw + 4 * c4);
        const int kb = c4 / 8, c = c4 % 8;
        uint4 hi, lo;
        hi.x = to_tf32(w.x); hi.y = to_tf32(w.y); hi.z = to_tf32(w.z); hi.w = to_tf32(w.w);
        lo.x = to_tf32(w.x - __uint_as_float(hi.x)); lo.y = to_tf32(w.y - __uint_as_float(hi.y));
        lo.z = to_tf32(w.z - __uint_as_float(hi.z)); lo.w = to_tf32(w.w - __uint_as_float(hi.w));
        *reinterpret_cast<uint4 *>(img + (size_t)kb * KB_BYTES + swz(j, c)) = hi;
        *reinterpret_cast<uint4 *>(img + (size_t)(NKB + kb) * KB_BYTES + swz(j, c)) = lo;
    }
}

__global__ void __launch_bounds__(kThreads, 1)
am_project_tc_kernel(int64_t M, int n_tables, int n_groups, const float *__restrict__ h,
                     const float *__restrict__ image, float *__restrict__ tables, int64_t table_stride, int64_t ldo,
                     const float *__restrict__ bias, int relu, int64_t ldx, int accumulate) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char *sW = smem;                                   // 128 KB weight image
    unsigned char *sA = smem + W_BYTES;                         // A ring
    unsigned char *sO = sA + A_STAGES * A_STAGE_BYTES;          // epilogue staging, 2 x 16 KB
    uint64_t *bars = reinterpret_cast<uint64_t *>(sO + 2 * OUT_BYTES);
    uint64_t *w_full = bars, *a_full = bars + 1, *a_empty = bars + 1 + A_STAGES;
    uint64_t *acc_full = bars + 1 + 2 * A_STAGES, *acc_empty = acc_full + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_empty + 2);
    float *s_bias = reinterpret_cast<float *>(sO + 2 * OUT_BYTES + 256);   // this table's 128 bias values (zeros if none)

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int table = blockIdx.x % n_tables, group = blockIdx.x / n_tables;
    const int64_t n_tiles = (M + BM - 1) / BM;

    if (tid == 0) {
        mbar_init(w_full, 1);
        for (int s = 0; s < A_STAGES; ++s) { mbar_init(&a_full[s], kProducerWarps); mbar_init(&a_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 4); }
        fence_barrier_init();
    }
    if (tid < BN) s_bias[tid] = bias ? bias[table * BN + tid] : 0.f;
    if (warp == 0) {                                            // TMEM: 256 columns for this CTA
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================================= MMA issuer (one lane)
        if (lane == 0) {
            mbar_expect_tx(w_full, (unsigned)W_BYTES);
            const char *src = reinterpret_cast<const char *>(image) + (size_t)table * W_BYTES;
            for (int off = 0; off < W_BYTES; off += KB_BYTES) tma_load_1d(sW + off, src + off, KB_BYTES, w_full);
            mbar_wait_bounded(w_full, 0);
            // descriptors differ only in the start-address field (16-byte units, low word): one add each,
            // so that this single thread can keep the tensor pipe fed (48 MMAs of ~64 clk per tile)
            const uint64_t dA = smem_desc(smem_u32(sA)), dW = smem_desc(smem_u32(sW));
            static_assert(A_STAGES == 2 && NKB % A_STAGES == 0, "stage of k-block kb is kb & 1");
            unsigned a_phase = 0, acc_phase0 = 0, acc_phase1 = 0;
            int it = 0;
            for (int64_t tile = group; tile < n_tiles; tile += n_groups, ++it) {
                const int ab = it & 1;
                if (ab == 0) { mbar_wait_bounded(&acc_empty[0], acc_phase0 ^ 1u); acc_phase0 ^= 1u; }
                else         { mbar_wait_bounded(&acc_empty[1], acc_phase1 ^ 1u); acc_phase1 ^= 1u; }
                tc_fence_after();
                const uint32_t d = tmem_base + (uint32_t)(ab * BN);
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb) {
                    const int stage = kb & 1;
                    mbar_wait_bounded(&a_full[stage], a_phase);
                    tc_fence_after();
                    const uint64_t a_hi = dA + (uint64_t)((stage * A_STAGE_BYTES) >> 4), a_lo = a_hi + (KB_BYTES >> 4);
                    const uint64_t b_hi = dW + (uint64_t)((kb * KB_BYTES) >> 4), b_lo = b_hi + ((NKB * KB_BYTES) >> 4);
#pragma unroll
                    for (int k8 = 0; k8 < KB / 8; ++k8) {       // K = 8 tf32 = 32 bytes = 2 address units per instruction
                        umma_tf32(d, a_lo + 2 * k8, b_hi + 2 * k8, (kb | k8) != 0);
                        umma_tf32(d, a_hi + 2 * k8, b_lo + 2 * k8, 1u);
                        umma_tf32(d, a_hi + 2 * k8, b_hi + 2 * k8, 1u);
                    }
                    umma_commit(&a_empty[stage]);               // smem slot free once these MMAs have read it
                    if (stage == A_STAGES - 1) a_phase ^= 1u;
                }
                umma_commit(&acc_full[ab]);                     // accumulator complete
            }
        }
        __syncwarp();
    } else if (warp <= kProducerWarps) {
        // ================================================= producers: h tile -> hi / lo swizzled images
        const int p = tid - 32;                                 // 0..255
        const int c = p & 7, r0 = p >> 3;                       // 16-byte chunk of the k-block row, first row
        constexpr int RPT = BM / (kProducerWarps * 4);          // rows per thread and k-block (4)
        int stage = 0;
        unsigned phase = 0;
        float4 v[NKB][RPT];                                     // one tile of h: 16 LDG.128 per thread in flight
        auto load_kb = [&](int64_t tile, int kb) {
#pragma unroll
            for (int i = 0; i < RPT; ++i) {
                const int64_t m = tile * BM + r0 + 32 * i;
                v[kb][i] = (tile < n_tiles && m < M) ? __ldg(reinterpret_cast<const float4 *>(h + m * ldx + kb * KB) + c)
                                                     : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        };
#pragma unroll
        for (int kb = 0; kb < NKB; ++kb) load_kb(group, kb);
        for (int64_t tile = group; tile < n_tiles; tile += n_groups) {
#pragma unroll
            for (int kb = 0; kb < NKB; ++kb) {
                mbar_wait_bounded(&a_empty[stage], phase ^ 1u);
                unsigned char *hi_img = sA + stage * A_STAGE_BYTES, *lo_img = hi_img + KB_BYTES;
#pragma unroll
                for (int i = 0; i < RPT; ++i) {
                    const int r = r0 + 32 * i;
                    const float4 a4 = v[kb][i];
                    uint4 hi, lo;
                    hi.x = to_tf32(a4.x); hi.y = to_tf32(a4.y); hi.z = to_tf32(a4.z); hi.w = to_tf32(a4.w);
                    lo.x = to_tf32(a4.x - __uint_as_float(hi.x)); lo.y = to_tf32(a4.y - __uint_as_float(hi.y));
                    lo.z = to_tf32(a4.z - __uint_as_float(hi.z)); lo.w = to_tf32(a4.w - __uint_as_float(hi.w));
                    *reinterpret_cast<uint4 *>(hi_img + swz(r, c)) = hi;
                    *reinterpret_cast<uint4 *>(lo_img + swz(r, c)) = lo;
                }
                load_kb(tile + n_groups, kb);           // same registers, next tile: a whole tile time to land
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic writes -> async proxy (UMMA)
                __syncwarp();
                if (lane == 0) mbar_arrive(&a_full[stage]);
                if (++stage == A_STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else {
        // ================================================= epilogue: TMEM -> registers -> table rows
        const int q = warp & 3;                                 // TMEM lane quarter this warp may access
        const int row = 32 * q + lane;
        float *out = tables + (size_t)table * table_stride;
        const bool epi = bias != nullptr || relu || accumulate;
        unsigned acc_phase[2] = {0u, 0u};
        int it = 0;
        for (int64_t tile = group; tile < n_tiles; tile += n_groups, ++it) {
            const int ab = it & 1;
            mbar_wait_bounded(&acc_full[ab], acc_phase[ab]);
            acc_phase[ab] ^= 1u;
            tc_fence_after();
#pragma unroll
            for (int cb = 0; cb < BN / 32; ++cb) {
                uint32_t v[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(ab * BN + cb * 32);
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                      "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                      "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (cb == BN / 32 - 1) {                        // accumulator drained: hand it back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[ab]);
                }
                // registers (one row, 32 columns per thread) -> swizzled staging -> coalesced row stores
                unsigned char *st = sO + (cb & 1) * OUT_BYTES;
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    *reinterpret_cast<uint4 *>(st + swz(row, k)) = make_uint4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
                asm volatile("bar.sync 1, 128;" ::: "memory");     // the four epilogue warps
                const int et = tid - (1 + kProducerWarps) * 32;     // 0..127
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int idx = et + 128 * j, rr = idx >> 3, ch = idx & 7;
                    const int64_t mm = tile * BM + rr;
                    if (mm < M) {
                        float4 y = *reinterpret_cast<const float4 *>(st + swz(rr, ch));
                        float4 *dst = reinterpret_cast<float4 *>(out + mm * ldo + cb * 32 + ch * 4);
                        if (epi) {                              // linear-layer epilogue: (+ previous K chunks) + bias, ReLU
                            const float4 bb = *reinterpret_cast<const float4 *>(s_bias + cb * 32 + ch * 4);
                            y.x += bb.x; y.y += bb.y; y.z += bb.z; y.w += bb.w;
                            if (accumulate) { const float4 o = *dst; y.x += o.x; y.y += o.y; y.z += o.z; y.w += o.w; }
                            if (relu) { y.x = fmaxf(y.x, 0.f); y.y = fmaxf(y.y, 0.f); y.z = fmaxf(y.z, 0.f); y.w = fmaxf(y.w, 0.f); }
                        }
                        *dst = y;
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

}  // namespace tc

int launch_pack_weights(int n_tables, const float *W, int64_t ldw, float *image, cudaStream_t stream) {
    tc::pack_weights_kernel<<<n_tables, 256, 0, stream>>>(n_tables, W, ldw, image);
    RL4CO_CHECK_LAUNCH("am_pack_weights");
    return 0;
}

int launch_project_tc(int64_t M, int n_tables, const float *h, const float *image, float *tables,
                      int64_t table_stride, int64_t ldo, const float *bias, int relu, int64_t ldx, int accumulate,
                      cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    cudaError_t e = cudaFuncSetAttribute(tc::am_project_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         tc::SMEM_BYTES);
    if (e != cudaSuccess) {
        set_error("am_project(tc): cannot reserve %d B of shared memory: %s", tc::SMEM_BYTES, cudaGetErrorString(e));
        return 2;
    }
    const int64_t n_tiles = (M + tc::BM - 1) / tc::BM;
    int64_t groups = sm_count / n_tables;                       // persistent: one CTA per SM
    if (groups < 1) groups = 1;
    if (groups > n_tiles) groups = n_tiles;
    tc::am_project_tc_kernel<<<(unsigned)(groups * n_tables), tc::kThreads, tc::SMEM_BYTES, stream>>>(
        M, n_tables, (int)groups, h, image, tables, table_stride, ldo, bias, relu, ldx, accumulate);
    RL4CO_CHECK_LAUNCH("am_project(tc)");
    return 0;
}

}  // namespace rl4co
