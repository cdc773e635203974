// Encoder self-attention with Q K^T on the tensor cores (N <= 128 nodes, 8 heads x 16 dims, no mask,
// fp32-accurate).  Same contract as encoder_attention.cu (which stays as the SIMT cross-check): replaces
// the F.scaled_dot_product_attention call of rl4co/models/nn/attention.py:81-126 on the fused Wqkv output.
//
// The SIMT kernel spends 48 FMA per (query, key) pair: q.k twice (two-pass softmax) and p.v.  Here the
// scores S = Q K^T of one (instance, head) are ONE 128 x 128 x 16 tcgen05.mma problem (3-term TF32 split,
// fp32 accumulation in TMEM: 6 MMAs of M128 N128 K8), and the CUDA cores keep only the softmax and p.v
// (16 FMA per pair), reading the score row of "their" query straight from TMEM:
//   warps 1-4   producers: rows of Q_h, K_h, V_h (64 B each) of one (b, h) -> Q (pre-scaled by
//               log2(e)/sqrt(16)) and K as hi / lo TF32 images (K-major, 128-byte swizzled rows, the 16
//               real columns in the first 64 bytes), V as plain fp32 rows; 2 stages for Q/K, 4 for V;
//   warp 0      MMA issue (one elected lane), S into one of four 128-column TMEM buffers;
//   warps 5-16  three consumer groups of four warps (thread = query row = TMEM lane): tcgen05.ld the row,
//               max, exp2, sum, o += p V_j (V rows as broadcast LDS.128, FFMA2), normalise, store.
// A CTA walks whole instances (its 8 heads back to back), so the 64-byte column slices it reads share
// their 128-byte lines through L1 / L2.
#include <cuda_fp16.h>

#include "common.cuh"
#include "rollout_util.cuh"
#include "tc_util.cuh"

namespace rl4co {
namespace enctc {
using namespace tcu;

typedef unsigned long long u64;
constexpr int H = 8, D = 16, E = 128, MAXN = 128;
constexpr int TILE_BYTES = MAXN * 128;                  // one operand image: 128 rows x 128 B (64 B used)
constexpr int QK_STAGE_BYTES = 4 * TILE_BYTES;          // Q hi, Q lo, K hi, K lo
constexpr int QK_STAGES = 2;
constexpr int V_SLOT_BYTES = MAXN * D * 4;              // 8 KB
constexpr int S_BUFS = 4;                               // TMEM score buffers = V slots
constexpr int kProducerWarps = 4, kGroups = 3, kConsumerWarps = 4 * kGroups;
constexpr int kThreads = (1 + kProducerWarps + kConsumerWarps) * 32;       // 544
constexpr int SMEM_BYTES = 1024 + QK_STAGES * QK_STAGE_BYTES + S_BUFS * V_SLOT_BYTES + 256;
constexpr uint32_t kIdesc = make_idesc_tf32(128, 128);

__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ u64 pack2(float lo, float hi) {
    u64 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
    return d;
}
__device__ __forceinline__ void unpack2(u64 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ float ex2f(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// one 32-column slice of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
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
}

__global__ void __launch_bounds__(kThreads, 1)
encoder_attention_tc_kernel(int64_t B, int N, const float *__restrict__ qkv, float *__restrict__ out) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char *sQK = smem;                                      // [stage][Qhi | Qlo | Khi | Klo]
    unsigned char *sV = smem + QK_STAGES * QK_STAGE_BYTES;          // [slot][key][16] fp32
    uint64_t *bars = reinterpret_cast<uint64_t *>(sV + S_BUFS * V_SLOT_BYTES);
    uint64_t *in_full = bars, *in_empty = bars + QK_STAGES, *s_full = in_empty + QK_STAGES, *done = s_full + S_BUFS;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(done + S_BUFS);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // this CTA's units: instances blockIdx.x, + gridDim.x, ...; eight heads each
    const int64_t n_inst = blockIdx.x < B ? (B - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t n_units = n_inst * H;

    if (tid == 0) {
        for (int s = 0; s < QK_STAGES; ++s) { mbar_init(&in_full[s], kProducerWarps); mbar_init(&in_empty[s], 1); }
        for (int s = 0; s < S_BUFS; ++s) { mbar_init(&s_full[s], 1); mbar_init(&done[s], 4); }
        fence_barrier_init();
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================================= MMA issue
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t dQK = (uint32_t)smem_desc(smem_u32(sQK));
        for (int64_t i = 0; i < n_units; ++i) {
            const int stage = (int)(i & 1), sb = (int)(i & 3);
            mbar_wait_bounded(&in_full[stage], (unsigned)((i >> 1) & 1));
            mbar_wait_bounded(&done[sb], (unsigned)((i >> 2) & 1) ^ 1u);    // score buffer drained by its consumers
            tc_fence_after();
            const uint32_t q_hi = dQK + (uint32_t)((stage * QK_STAGE_BYTES) >> 4), q_lo = q_hi + (TILE_BYTES >> 4);
            const uint32_t k_hi = q_hi + 2 * (TILE_BYTES >> 4), k_lo = q_hi + 3 * (TILE_BYTES >> 4);
            const uint32_t d = tmem_u + (uint32_t)(sb * 128);
            if (elect_one()) {
#pragma unroll
                for (int k8 = 0; k8 < D / 8; ++k8) {            // K = 16: two instructions of K = 8 per term
                    umma_tf32(d, q_lo + 2 * k8, k_hi + 2 * k8, kIdesc, (unsigned)(k8 != 0));
                    umma_tf32(d, q_hi + 2 * k8, k_lo + 2 * k8, kIdesc, 1u);
                    umma_tf32(d, q_hi + 2 * k8, k_hi + 2 * k8, kIdesc, 1u);
                }
                umma_commit(&in_empty[stage]);
                umma_commit(&s_full[sb]);
            }
            __syncwarp();
        }
    } else if (warp <= kProducerWarps) {
        // ================================================= producers: thread = node row
        const int row = tid - 32;                               // 0..127
        const float scale = 0.25f * 1.4426950408889634f;        // 1/sqrt(16) * log2(e), folded into Q
        const uint32_t r7 = (uint32_t)(row & 7);
        if (row >= N) {                                         // padding rows of every Q / K image stay zero
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
            for (int s = 0; s < QK_STAGES; ++s)
                for (int img = 0; img < 4; ++img)
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        *reinterpret_cast<uint4 *>(sQK + s * QK_STAGE_BYTES + img * TILE_BYTES + swz(row, c)) = z;
        }
        for (int64_t i = 0; i < n_units; ++i) {
            const int stage = (int)(i & 1), sb = (int)(i & 3), h = (int)(i & 7);
            const int64_t b = blockIdx.x + (i >> 3) * gridDim.x;
            float4 q[4], k[4], v[4];
            if (row < N) {                                      // loads first: they fly while we wait for the slots
                const float4 *p = reinterpret_cast<const float4 *>(qkv + (b * N + row) * (3 * E) + h * D);
#pragma unroll
                for (int c = 0; c < 4; ++c) { q[c] = __ldg(p + c); k[c] = __ldg(p + E / 4 + c); v[c] = __ldg(p + 2 * E / 4 + c); }
            }
            mbar_wait_bounded(&in_empty[stage], (unsigned)((i >> 1) & 1) ^ 1u);
            mbar_wait_bounded(&done[sb], (unsigned)((i >> 2) & 1) ^ 1u);
            if (row < N) {
                unsigned char *st = sQK + stage * QK_STAGE_BYTES;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const uint32_t off = (uint32_t)row * 128u + (((uint32_t)c ^ r7) << 4);
                    const float4 a = make_float4(q[c].x * scale, q[c].y * scale, q[c].z * scale, q[c].w * scale);
                    uint4 hi, lo;
                    hi.x = to_tf32(a.x); hi.y = to_tf32(a.y); hi.z = to_tf32(a.z); hi.w = to_tf32(a.w);
                    lo.x = to_tf32(a.x - __uint_as_float(hi.x)); lo.y = to_tf32(a.y - __uint_as_float(hi.y));
                    lo.z = to_tf32(a.z - __uint_as_float(hi.z)); lo.w = to_tf32(a.w - __uint_as_float(hi.w));
                    *reinterpret_cast<uint4 *>(st + off) = hi;
                    *reinterpret_cast<uint4 *>(st + TILE_BYTES + off) = lo;
                    const float4 kk = k[c];
                    hi.x = to_tf32(kk.x); hi.y = to_tf32(kk.y); hi.z = to_tf32(kk.z); hi.w = to_tf32(kk.w);
                    lo.x = to_tf32(kk.x - __uint_as_float(hi.x)); lo.y = to_tf32(kk.y - __uint_as_float(hi.y));
                    lo.z = to_tf32(kk.z - __uint_as_float(hi.z)); lo.w = to_tf32(kk.w - __uint_as_float(hi.w));
                    *reinterpret_cast<uint4 *>(st + 2 * TILE_BYTES + off) = hi;
                    *reinterpret_cast<uint4 *>(st + 3 * TILE_BYTES + off) = lo;
                    *reinterpret_cast<float4 *>(sV + sb * V_SLOT_BYTES + row * (D * 4) + c * 16) = v[c];
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic writes -> async proxy (UMMA)
            __syncwarp();
            if (lane == 0) mbar_arrive(&in_full[stage]);
        }
    } else {
        // ================================================= consumers: thread = query row = TMEM lane
        const int cw = warp - (1 + kProducerWarps);             // 0..11
        const int group = cw >> 2, quarter = warp & 3;          // TMEM lanes 32 * (warp % 4) ..
        const int row = 32 * quarter + lane;
        const int full_chunks = N >> 5, tail = N & 31;
        for (int64_t i = group; i < n_units; i += kGroups) {
            const int sb = (int)(i & 3), h = (int)(i & 7);
            const int64_t b = blockIdx.x + (i >> 3) * gridDim.x;
            mbar_wait_bounded(&s_full[sb], (unsigned)((i >> 2) & 1));   // scores complete; V slot sb was filled before them
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(32 * quarter) << 16) + (uint32_t)(sb * 128);
            uint32_t sv[32];
            // pass 1: row maximum
            float m = -INFINITY;
            for (int cb = 0; cb < full_chunks; ++cb) {
                tmem_ld32(taddr + cb * 32, sv);
#pragma unroll
                for (int j = 0; j < 32; ++j) m = fmaxf(m, __uint_as_float(sv[j]));
            }
            if (tail) {
                tmem_ld32(taddr + full_chunks * 32, sv);
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < tail) m = fmaxf(m, __uint_as_float(sv[j]));
            }
            // pass 2: p = 2^(s - m), l = sum p, o = sum p V_j
            u64 o[8];
#pragma unroll
            for (int d = 0; d < 8; ++d) o[d] = 0ull;
            float l = 0.f;
            const uint32_t v_addr = smem_u32(sV + sb * V_SLOT_BYTES);
            auto key = [&](float s, int j) {
                const float p = ex2f(s - m);
                l += p;
                const u64 pp = pack2(p, p);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    u64 v0, v1;
                    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(v0), "=l"(v1) : "r"(v_addr + (uint32_t)(j * 64 + c * 16)));
                    o[2 * c] = ffma2(pp, v0, o[2 * c]);
                    o[2 * c + 1] = ffma2(pp, v1, o[2 * c + 1]);
                }
            };
            for (int cb = 0; cb < full_chunks; ++cb) {
                tmem_ld32(taddr + cb * 32, sv);
#pragma unroll
                for (int j = 0; j < 32; ++j) key(__uint_as_float(sv[j]), cb * 32 + j);
            }
            if (tail) {
                tmem_ld32(taddr + full_chunks * 32, sv);
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < tail) key(__uint_as_float(sv[j]), full_chunks * 32 + j);
            }
            tc_fence_before();                                  // TMEM buffer and V slot may be reused
            __syncwarp();
            if (lane == 0) mbar_arrive(&done[sb]);
            if (row < N) {
                const float inv = 1.f / l;
                float4 *dst = reinterpret_cast<float4 *>(out + (b * N + row) * E + h * D);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float4 y;
                    unpack2(o[2 * c], y.x, y.y);
                    unpack2(o[2 * c + 1], y.z, y.w);
                    y.x *= inv; y.y *= inv; y.z *= inv; y.w *= inv;
                    dst[c] = y;
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

}  // namespace enctc


// ============================================================================================
// Both GEMMs on the tensor cores (N <= 112): S = Q K^T as above with N = 112 key columns, then the
// consumers turn their score row into P = 2^(s - max) IN TMEM (hi / lo TF32 images written back with
// tcgen05.st: P_hi over the scores, P_lo beside them) and O = P V is a second tcgen05.mma with A = P
// read from TMEM and B = V^T (hi / lo, K-major swizzled rows of 32 keys) from shared memory:
// 14 k-steps x 3 terms of M128 N16 K8.  The CUDA cores keep only max / exp2 / sum / split and the
// final O / l.  TMEM per buffer (256 columns, two buffers): [0,112) S then P_hi, [112,224) P_lo,
// [224,240) O.  Two consumer groups of four warps, one per buffer; the MMA warp interleaves
// QK(i) with PV(i-1).
namespace enctc2 {
using namespace tcu;
using enctc::ex2f;

constexpr int H = 8, D = 16, E = 128, NK = 112;         // key columns of the score tile (multiple of 16)
constexpr int TILE_BYTES = 128 * 128;                   // Q / K operand image: 128 rows x 128 B (64 B used)
constexpr int VT_TILE = 16 * 128;                       // V^T k-block: 16 dims x 32 keys
constexpr int VT_BYTES = 4 * VT_TILE;                   // 8 KB per image (hi or lo)
constexpr int QK_STAGE_BYTES = 4 * TILE_BYTES;          // Qhi Qlo Khi Klo = 64 KB, free again once the scores are in TMEM
constexpr int QK_STAGES = 2;
constexpr int VT_SLOT_BYTES = 2 * VT_BYTES;             // VThi VTlo = 16 KB, needed until O = P V is complete
constexpr int VT_SLOTS = 4;
constexpr int kProducerWarps = 4, kConsumerWarps = 8;
constexpr int kThreads = (1 + kProducerWarps + kConsumerWarps) * 32;       // 416
constexpr int SMEM_BYTES = 1024 + QK_STAGES * QK_STAGE_BYTES + VT_SLOTS * VT_SLOT_BYTES + 256;
constexpr uint32_t kIdescS = make_idesc_tf32(128, NK);
constexpr uint32_t kIdescO = make_idesc_tf32(128, D);
constexpr int COL_PLO = NK, COL_O = 2 * NK, BUF_COLS = 256;

// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_lo32, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 db;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], db, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "r"(tmem_a), "r"(b_lo32), "r"(idesc), "r"(accumulate), "r"(kDescHi)
        : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}

__global__ void __launch_bounds__(kThreads, 1)
encoder_attention_tc2_kernel(int64_t B, int N, const float *__restrict__ qkv, float *__restrict__ out) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char *sVT = smem + QK_STAGES * QK_STAGE_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sVT + VT_SLOTS * VT_SLOT_BYTES);
    uint64_t *in_full = bars, *s_full = bars + 2, *p_full = bars + 4, *o_full = bars + 6, *buf_free = bars + 8;
    uint64_t *qk_empty = bars + 10, *vt_free = bars + 12;       // [2], [4]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 16);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t n_inst = blockIdx.x < B ? (B - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t n_units = n_inst * H;

    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(&in_full[s], kProducerWarps);     // inputs of a unit staged
            mbar_init(&s_full[s], 1);                   // scores complete (commit)
            mbar_init(&p_full[s], 4);                   // P written by the buffer's four consumer warps
            mbar_init(&o_full[s], 1);                   // O complete (commit)
            mbar_init(&buf_free[s], 4);                 // O read back: the TMEM buffer may be overwritten
            mbar_init(&qk_empty[s], 1);                 // Q / K stage read by the score MMAs (commit)
        }
        for (int s = 0; s < VT_SLOTS; ++s) mbar_init(&vt_free[s], 1);   // V^T slot read by the P V MMAs (commit)
        fence_barrier_init();
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================================= MMA issue: QK(i), then PV(i - 1)
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t d0 = (uint32_t)smem_desc(smem_u32(smem));
        for (int64_t i = 0; i <= n_units; ++i) {
            if (i < n_units) {
                const int b = (int)(i & 1);
                const unsigned ph = (unsigned)((i >> 1) & 1);
                mbar_wait_bounded(&in_full[b], ph);
                mbar_wait_bounded(&buf_free[b], ph ^ 1u);
                tc_fence_after();
                const uint32_t q_hi = d0 + (uint32_t)((b * QK_STAGE_BYTES) >> 4), q_lo = q_hi + (TILE_BYTES >> 4);
                const uint32_t k_hi = q_hi + 2 * (TILE_BYTES >> 4), k_lo = q_hi + 3 * (TILE_BYTES >> 4);
                const uint32_t d = tmem_u + (uint32_t)(b * BUF_COLS);
                if (elect_one()) {
#pragma unroll
                    for (int k8 = 0; k8 < D / 8; ++k8) {
                        umma_tf32(d, q_lo + 2 * k8, k_hi + 2 * k8, kIdescS, (unsigned)(k8 != 0));
                        umma_tf32(d, q_hi + 2 * k8, k_lo + 2 * k8, kIdescS, 1u);
                        umma_tf32(d, q_hi + 2 * k8, k_hi + 2 * k8, kIdescS, 1u);
                    }
                    umma_commit(&qk_empty[b]);
                    umma_commit(&s_full[b]);
                }
                __syncwarp();
            }
            if (i >= 1) {
                const int64_t u = i - 1;
                const int b = (int)(u & 1);
                mbar_wait_bounded(&p_full[b], (unsigned)((u >> 1) & 1));
                tc_fence_after();
                const int vs = (int)(u & 3);
                const uint32_t vt_hi = d0 + (uint32_t)((QK_STAGES * QK_STAGE_BYTES + vs * VT_SLOT_BYTES) >> 4);
                const uint32_t vt_lo = vt_hi + (VT_BYTES >> 4);
                const uint32_t p_hi = tmem_u + (uint32_t)(b * BUF_COLS), p_lo = p_hi + COL_PLO;
                const uint32_t o = tmem_u + (uint32_t)(b * BUF_COLS + COL_O);
                if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < NK / 8; ++ks) {       // 14 k-steps of 8 keys
                        const uint32_t voff = (uint32_t)(((ks >> 2) * VT_TILE + (ks & 3) * 32) >> 4);
                        umma_tf32_ts(o, p_lo + 8 * ks, vt_hi + voff, kIdescO, (unsigned)(ks != 0));
                        umma_tf32_ts(o, p_hi + 8 * ks, vt_lo + voff, kIdescO, 1u);
                        umma_tf32_ts(o, p_hi + 8 * ks, vt_hi + voff, kIdescO, 1u);
                    }
                    umma_commit(&vt_free[vs]);
                    umma_commit(&o_full[b]);
                }
                __syncwarp();
            }
        }
    } else if (warp <= kProducerWarps) {
        // ================================================= producers: thread = node row
        const int row = tid - 32;                               // 0..127
        const float scale = 0.25f * 1.4426950408889634f;        // 1/sqrt(16) * log2(e), folded into Q
        const uint32_t r7 = (uint32_t)(row & 7);
        {                                                       // zero what is never written: pad rows, all of V^T
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
            if (row >= N)
                for (int s = 0; s < QK_STAGES; ++s)
                    for (int img = 0; img < 4; ++img)
#pragma unroll
                        for (int c = 0; c < 4; ++c)
                            *reinterpret_cast<uint4 *>(smem + s * QK_STAGE_BYTES + img * TILE_BYTES + swz(row, c)) = z;
            for (int off = row * 16; off < VT_SLOTS * VT_SLOT_BYTES; off += 128 * 16)
                *reinterpret_cast<uint4 *>(sVT + off) = z;
            asm volatile("bar.sync 2, 128;" ::: "memory");      // the four producer warps: zeros before any key write
        }
        // V^T element (key j = row, dim d): k-block j >> 5, row d, 16-byte chunk ((j & 31) >> 2) ^ (d & 7), word j & 3
        const uint32_t vt_base = (uint32_t)((row >> 5) * VT_TILE + (row & 3) * 4);
        const uint32_t vt_chunk = (uint32_t)((row & 31) >> 2);
        for (int64_t i = 0; i < n_units; ++i) {
            const int st_i = (int)(i & 1), h = (int)(i & 7);
            const int64_t b = blockIdx.x + (i >> 3) * gridDim.x;
            float4 q[4], k[4], v[4];
            if (row < N) {                                      // loads first: they fly while we wait for the stage
                const float4 *p = reinterpret_cast<const float4 *>(qkv + (b * N + row) * (3 * E) + h * D);
#pragma unroll
                for (int c = 0; c < 4; ++c) { q[c] = __ldg(p + c); k[c] = __ldg(p + E / 4 + c); v[c] = __ldg(p + 2 * E / 4 + c); }
            }
            const int vs = (int)(i & 3);
            mbar_wait_bounded(&qk_empty[st_i], (unsigned)((i >> 1) & 1) ^ 1u);  // scores of unit i - 2 are in TMEM
            mbar_wait_bounded(&vt_free[vs], (unsigned)((i >> 2) & 1) ^ 1u);     // P V of unit i - 4 is complete
            if (row < N) {
                unsigned char *st = smem + st_i * QK_STAGE_BYTES;
                unsigned char *vt = sVT + vs * VT_SLOT_BYTES;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const uint32_t off = (uint32_t)row * 128u + (((uint32_t)c ^ r7) << 4);
                    const float4 a = make_float4(q[c].x * scale, q[c].y * scale, q[c].z * scale, q[c].w * scale);
                    uint4 hi, lo;
                    hi.x = to_tf32(a.x); hi.y = to_tf32(a.y); hi.z = to_tf32(a.z); hi.w = to_tf32(a.w);
                    lo.x = to_tf32(a.x - __uint_as_float(hi.x)); lo.y = to_tf32(a.y - __uint_as_float(hi.y));
                    lo.z = to_tf32(a.z - __uint_as_float(hi.z)); lo.w = to_tf32(a.w - __uint_as_float(hi.w));
                    *reinterpret_cast<uint4 *>(st + off) = hi;
                    *reinterpret_cast<uint4 *>(st + TILE_BYTES + off) = lo;
                    const float4 kk = k[c];
                    hi.x = to_tf32(kk.x); hi.y = to_tf32(kk.y); hi.z = to_tf32(kk.z); hi.w = to_tf32(kk.w);
                    lo.x = to_tf32(kk.x - __uint_as_float(hi.x)); lo.y = to_tf32(kk.y - __uint_as_float(hi.y));
                    lo.z = to_tf32(kk.z - __uint_as_float(hi.z)); lo.w = to_tf32(kk.w - __uint_as_float(hi.w));
                    *reinterpret_cast<uint4 *>(st + 2 * TILE_BYTES + off) = hi;
                    *reinterpret_cast<uint4 *>(st + 3 * TILE_BYTES + off) = lo;
                    const float vv[4] = {v[c].x, v[c].y, v[c].z, v[c].w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {               // dim d = 4 c + e of key `row`, transposed
                        const int d = 4 * c + e;
                        const uint32_t o2 = vt_base + (uint32_t)d * 128u + ((vt_chunk ^ (uint32_t)(d & 7)) << 4);
                        const uint32_t vh = to_tf32(vv[e]);
                        *reinterpret_cast<uint32_t *>(vt + o2) = vh;
                        *reinterpret_cast<uint32_t *>(vt + VT_BYTES + o2) = to_tf32(vv[e] - __uint_as_float(vh));
                    }
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic writes -> async proxy (UMMA)
            __syncwarp();
            if (lane == 0) mbar_arrive(&in_full[st_i]);
        }
    } else {
        // ================================================= consumers: thread = query row = TMEM lane
        const int cw = warp - (1 + kProducerWarps);             // 0..7
        const int buf = cw >> 2, quarter = warp & 3;
        const int row = 32 * quarter + lane;
        const uint32_t t0 = tmem_base + ((uint32_t)(32 * quarter) << 16) + (uint32_t)(buf * BUF_COLS);
        const int n_full = N >> 4, tail = N & 15;
        for (int64_t i = buf; i < n_units; i += 2) {
            const int h = (int)(i & 7);
            const int64_t b = blockIdx.x + (i >> 3) * gridDim.x;
            const unsigned ph = (unsigned)((i >> 1) & 1);
            mbar_wait_bounded(&s_full[buf], ph);
            tc_fence_after();
            // The loops over 16-column chunks stay rolled (one body for full chunks, one predicated body for
            // the chunk that holds column N - 1): unrolled per chunk, this code no longer fits the instruction
            // cache next to the other roles.  P is split by truncation (hi = p & ~0x1fff, lo = p - hi, exact);
            // p is in [0, 1], so the round-to-nearest conversion's overflow handling buys nothing here.
            uint32_t sv[16];
            float m = -INFINITY;                                // pass 1: row maximum over the N real keys
#pragma unroll 1
            for (int c = 0; c < n_full; ++c) {
                tmem_ld16(t0 + c * 16, sv);
#pragma unroll
                for (int j = 0; j < 16; ++j) m = fmaxf(m, __uint_as_float(sv[j]));
            }
            if (tail) {
                tmem_ld16(t0 + n_full * 16, sv);
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (j < tail) m = fmaxf(m, __uint_as_float(sv[j]));
            }
            float l = 0.f;                                      // pass 2: P = 2^(s - m) as hi / lo TF32, in place
            uint32_t hi[16], lo[16];
#pragma unroll 1
            for (int c = 0; c < n_full; ++c) {
                tmem_ld16(t0 + c * 16, sv);
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float p = ex2f(__uint_as_float(sv[j]) - m);
                    l += p;
                    hi[j] = __float_as_uint(p) & 0xffffe000u;
                    lo[j] = __float_as_uint(p - __uint_as_float(hi[j]));
                }
                tmem_st16(t0 + c * 16, hi);
                tmem_st16(t0 + COL_PLO + c * 16, lo);
            }
            int c_next = n_full;
            if (tail) {
                tmem_ld16(t0 + n_full * 16, sv);
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float p = j < tail ? ex2f(__uint_as_float(sv[j]) - m) : 0.f;
                    l += p;
                    hi[j] = __float_as_uint(p) & 0xffffe000u;
                    lo[j] = __float_as_uint(p - __uint_as_float(hi[j]));
                }
                tmem_st16(t0 + n_full * 16, hi);
                tmem_st16(t0 + COL_PLO + n_full * 16, lo);
                c_next = n_full + 1;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) hi[j] = 0u;
#pragma unroll 1
            for (int c = c_next; c < NK / 16; ++c) {            // key columns beyond N: P = 0
                tmem_st16(t0 + c * 16, hi);
                tmem_st16(t0 + COL_PLO + c * 16, hi);
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[buf]);
            // O = P V arrives from the tensor core; normalise and store this row's 16 outputs
            mbar_wait_bounded(&o_full[buf], ph);
            tc_fence_after();
            tmem_ld16(t0 + COL_O, sv);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&buf_free[buf]);
            if (row < N) {
                const float inv = 1.f / l;
                float4 *dst = reinterpret_cast<float4 *>(out + (b * N + row) * E + h * D);
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    dst[c] = make_float4(__uint_as_float(sv[4 * c]) * inv, __uint_as_float(sv[4 * c + 1]) * inv,
                                         __uint_as_float(sv[4 * c + 2]) * inv, __uint_as_float(sv[4 * c + 3]) * inv);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

}  // namespace enctc2


// ============================================================================================
// The two-GEMM kernel with half the tensor-memory traffic (N <= 112).  tcgen05.ld / st move 64 B/clk per SM, and in
// enctc2 one (instance, head) unit moves 128 lanes x 464 words through that port - the scores read twice (max pass, exp
// pass), P_hi and P_lo written as TF32 words, O read - i.e. ~3.7 k clk, which is what that kernel takes per unit (its
// tensor pipe is 21 % active).  Here (a) the score row is read ONCE, into 112 registers, (b) P goes back as FP16 pairs,
// hi = fp16(p), lo = fp16(p - hi) (22 significant bits), two keys per 32-bit column, scaled by 2^14 so that the residuals
// of every p that matters are normal fp16 numbers (the scale cancels in O / l), and O = P V is three kind::f16 MMAs per 16
// keys with V^T staged as fp16 hi / lo images: 128 lanes x 240 words per unit, 21 + 6 MMAs instead of 42 + 6.
// Q K^T stays 3xTF32, so q and k keep fp32's range.  V goes through fp16: |v| must stay below 65 504 (an encoder whose
// values get there has diverged), and entries below 2^-14 carry an absolute error of 3e-8 instead of a relative one.
// Roles after that (ncu source-level samples, 16 384 x 100): the four producer warps were the critical role (their
// round-to-nearest TF32 conversions with inf / nan handling, a 2-byte transposing scatter for V^T, global-load latency
// exposed once per unit), so the producers now round by integer arithmetic, double-buffer their Q / K loads and prefetch
// the CTA's next instance to L2, and V is staged by the consumer group that needs it - as an MN-major operand, which takes
// four 16-byte stores per key row instead of 32 two-byte ones.  1.60 -> 1.50 ms, and 1.56 ms with the range scale of V
// (its exchange among the group's warps: 1.69 ms with a named barrier before the softmax, 1.56 ms posted early and read
// after the softmax).  Measured and dropped: four score buffers
// with software-pipelined consumers (softmax(i + 2) before the O of unit i is read: 1.79 ms with 13 warps - 128 registers,
// the spilled loop state reloads behind the cp.async requests in flight - and 1.70 ms with two producer warps of two rows
// each, where the producers bound it again); no producer role at all, every worker group staging its own operands (1.9-
// 2.1 ms: eight warps are two per scheduler, too few to cover their own latencies); 96 of the 112 score columns resident
// and 16 re-read (1.55 ms).  What limits it is warps per scheduler x registers: the one-pass softmax wants the 112-register
// row, which caps the CTA at 13 warps.
// -> enctc4 below (the default): the two-pass rolled softmax at 72 registers, four consumer groups, 1.30 ms.
namespace enctc3 {
using namespace tcu;
using enctc::ex2f;

constexpr int H = 8, D = 16, E = 128, NK = 112;         // key columns of the score tile (multiple of 16)
constexpr int TILE_BYTES = 128 * 128;                   // Q / K operand image (TF32): 128 rows x 128 B (64 B used)
constexpr int V_HALF = 128 * 16;                        // V image (FP16, MN-major, no swizzle): [dim half][key][8 dims = 16 B]
constexpr int VT_BYTES = 2 * V_HALF;                    // 4 KB per image (hi or lo)
constexpr int QK_STAGE_BYTES = 4 * TILE_BYTES;          // Qhi Qlo Khi Klo = 64 KB, free again once the scores are in TMEM
constexpr int QK_STAGES = 2;
constexpr int VT_SLOT_BYTES = 2 * VT_BYTES;             // Vhi Vlo = 8 KB, needed until O = P V is complete
constexpr int VT_SLOTS = 2;                             // one per TMEM buffer, written by that buffer's consumer group
constexpr int kProducerWarps = 4, kConsumerWarps = 8;
constexpr int kThreads = (1 + kProducerWarps + kConsumerWarps) * 32;       // 416
constexpr int V_STAGE_BYTES = 4 * 128 * 16;               // raw V rows of one unit, [16-byte chunk][key]: 8 KB per group
constexpr int SMEM_BYTES = 1024 + QK_STAGES * QK_STAGE_BYTES + VT_SLOTS * (VT_SLOT_BYTES + V_STAGE_BYTES) + 256;
constexpr uint32_t kIdescS = make_idesc_tf32(128, NK);
constexpr uint32_t kIdescO = make_idesc_f16(128, D) | (1u << 16);      // B (= V) is MN-major: a key's dims are contiguous
// matrix descriptor of V (cute/atom/mma_traits_sm100.hpp, Major-MN, SWIZZLE_NONE: ((1,n),(8,k)):((X,SBO),(1,LBO)) in 16-byte
// units): 8 keys of a k-group 16 B apart, the next k-group LBO = 128 B on, the other 8 dims SBO = 2048 B on; version 1
constexpr uint32_t kDescLoV = 8u << 16, kDescHiV = (uint32_t)(V_HALF >> 4) | (1u << 14);
constexpr int COL_PLO = NK / 2, COL_O = NK, BUF_COLS = 256;    // P_hi [0,56) over the scores, P_lo [56,112), O [112,128)
constexpr float kPShift = 14.f;                         // P = 2^(s - max + 14): the fp16 residuals stay normal numbers

// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_lo32, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 db;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "r"(tmem_a), "r"(b_lo32), "r"(idesc), "r"(accumulate), "r"(kDescHiV)
        : "memory");
}
using enctc2::tmem_ld16;
__device__ __forceinline__ void tmem_ld32_nw(uint32_t taddr, uint32_t *v) {       // no wait: pair with tcgen05.wait::ld
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16_nw(uint32_t taddr, uint32_t *v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
// all NK score columns of this thread's lane, one wait
__device__ __forceinline__ void tmem_ld_row(uint32_t taddr, uint32_t (&v)[NK]) {
    tmem_ld32_nw(taddr, v);
    tmem_ld32_nw(taddr + 32, v + 32);
    tmem_ld32_nw(taddr + 64, v + 64);
    tmem_ld16_nw(taddr + 96, v + 96);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(v[0]),
                 "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}

__device__ __forceinline__ uint32_t rn_tf32(float x) { return (__float_as_uint(x) + 0x1000u) & 0xffffe000u; }

__global__ void __launch_bounds__(kThreads, 1)
encoder_attention_tc3_kernel(int64_t B, int N, const float *__restrict__ qkv, float *__restrict__ out) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char *sVT = smem + QK_STAGES * QK_STAGE_BYTES;
    unsigned char *sVraw = sVT + VT_SLOTS * VT_SLOT_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sVraw + VT_SLOTS * V_STAGE_BYTES);
    uint64_t *in_full = bars, *s_full = bars + 2, *p_full = bars + 4, *o_full = bars + 6, *buf_free = bars + 8;
    uint64_t *qk_empty = bars + 10, *v_rng = bars + 12;         // [2] each
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 16);
    uint32_t *s_amax = tmem_slot + 4;                           // [group][unit parity][warp of the group]: V range exchange

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t n_inst = blockIdx.x < B ? (B - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t n_units = n_inst * H;

    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(&in_full[s], kProducerWarps);     // inputs of a unit staged
            mbar_init(&s_full[s], 1);                   // scores complete (commit)
            mbar_init(&p_full[s], 4);                   // P and V^T written by the buffer's four consumer warps
            mbar_init(&o_full[s], 1);                   // O complete (commit)
            mbar_init(&buf_free[s], 4);                 // O read back: the TMEM buffer may be overwritten
            mbar_init(&qk_empty[s], 1);                 // Q / K stage read by the score MMAs (commit)
            mbar_init(&v_rng[s], 4);                    // the group's four warps have posted the range of their V rows
        }
        fence_barrier_init();
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================================= MMA issue: QK(i), then PV(i - 1)
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t d0 = (uint32_t)smem_desc(smem_u32(smem));
        const uint32_t dv0 = ((smem_u32(sVT) >> 4) & 0x3FFFu) | kDescLoV;
        for (int64_t i = 0; i <= n_units; ++i) {
            if (i < n_units) {
                const int b = (int)(i & 1);
                const unsigned ph = (unsigned)((i >> 1) & 1);
                mbar_wait_bounded(&in_full[b], ph);
                mbar_wait_bounded(&buf_free[b], ph ^ 1u);
                tc_fence_after();
                const uint32_t q_hi = d0 + (uint32_t)((b * QK_STAGE_BYTES) >> 4), q_lo = q_hi + (TILE_BYTES >> 4);
                const uint32_t k_hi = q_hi + 2 * (TILE_BYTES >> 4), k_lo = q_hi + 3 * (TILE_BYTES >> 4);
                const uint32_t d = tmem_u + (uint32_t)(b * BUF_COLS);
                if (elect_one()) {
#pragma unroll
                    for (int k8 = 0; k8 < D / 8; ++k8) {
                        umma_tf32(d, q_lo + 2 * k8, k_hi + 2 * k8, kIdescS, (unsigned)(k8 != 0));
                        umma_tf32(d, q_hi + 2 * k8, k_lo + 2 * k8, kIdescS, 1u);
                        umma_tf32(d, q_hi + 2 * k8, k_hi + 2 * k8, kIdescS, 1u);
                    }
                    umma_commit(&qk_empty[b]);
                    umma_commit(&s_full[b]);
                }
                __syncwarp();
            }
            if (i >= 1) {
                const int64_t u = i - 1;
                const int b = (int)(u & 1);
                mbar_wait_bounded(&p_full[b], (unsigned)((u >> 1) & 1));
                tc_fence_after();
                const uint32_t vt_hi = dv0 + (uint32_t)((b * VT_SLOT_BYTES) >> 4), vt_lo = vt_hi + (VT_BYTES >> 4);
                const uint32_t p_hi = tmem_u + (uint32_t)(b * BUF_COLS), p_lo = p_hi + COL_PLO;
                const uint32_t o = tmem_u + (uint32_t)(b * BUF_COLS + COL_O);
                if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < NK / 16; ++ks) {      // 7 k-steps of 16 keys (8 TMEM columns of fp16 pairs, 256 B of V)
                        umma_f16_ts(o, p_lo + 8 * ks, vt_hi + 16 * ks, kIdescO, (unsigned)(ks != 0));
                        umma_f16_ts(o, p_hi + 8 * ks, vt_lo + 16 * ks, kIdescO, 1u);
                        umma_f16_ts(o, p_hi + 8 * ks, vt_hi + 16 * ks, kIdescO, 1u);
                    }
                    umma_commit(&o_full[b]);
                }
                __syncwarp();
            }
        }
    } else if (warp <= kProducerWarps) {
        // ================================================= producers: thread = node row; Q and K of one (b, h) per unit
        // (V goes through the consumers: they have the slack, this role is the kernel's critical one)
        const int row = tid - 32;                               // 0..127
        const float scale = 0.25f * 1.4426950408889634f;        // 1/sqrt(16) * log2(e), folded into Q
        const uint32_t r7 = (uint32_t)(row & 7);
        if (row >= N) {                                         // zero what is never written: the pad rows
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
            for (int s = 0; s < QK_STAGES; ++s)
                for (int img = 0; img < 4; ++img)
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        *reinterpret_cast<uint4 *>(smem + s * QK_STAGE_BYTES + img * TILE_BYTES + swz(row, c)) = z;
        }
        auto load = [&](int64_t i, float4 (&q)[4], float4 (&k)[4]) {
            if (row < N && i < n_units) {
                const int h = (int)(i & 7);
                const int64_t b = blockIdx.x + (i >> 3) * gridDim.x;
                const float4 *p = reinterpret_cast<const float4 *>(qkv + (b * N + row) * (3 * E) + h * D);
                if (h == 0 && i + H < n_units) {                // this row of the CTA's next instance -> L2 (12 lines)
                    const char *nx = reinterpret_cast<const char *>(qkv + ((b + gridDim.x) * N + row) * (3 * E));
#pragma unroll
                    for (int c = 0; c < 12; ++c) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + 128 * c));
                }
#pragma unroll
                for (int c = 0; c < 4; ++c) { q[c] = __ldg(p + c); k[c] = __ldg(p + E / 4 + c); }
            }
        };
        // hi = x rounded to TF32's 11 significant bits by integer arithmetic (add half an ulp, clear the 13 low bits - no
        // overflow check: |x| would have to be within 2^-11 of FLT_MAX), lo = x - hi exactly, of which the tensor core
        // reads the top 11 bits: three instructions per value where cvt.rna.tf32 with its inf / nan handling costs four
        // and a second conversion for lo
        auto stage = [&](int64_t i, const float4 (&q)[4], const float4 (&k)[4]) {
            const int st_i = (int)(i & 1);
            mbar_wait_bounded(&qk_empty[st_i], (unsigned)((i >> 1) & 1) ^ 1u);  // scores of unit i - 2 are in TMEM
            if (row < N) {
                unsigned char *st = smem + st_i * QK_STAGE_BYTES;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const uint32_t off = (uint32_t)row * 128u + (((uint32_t)c ^ r7) << 4);
                    const float4 a = make_float4(q[c].x * scale, q[c].y * scale, q[c].z * scale, q[c].w * scale);
                    uint4 hi, lo;
                    hi.x = rn_tf32(a.x); hi.y = rn_tf32(a.y); hi.z = rn_tf32(a.z); hi.w = rn_tf32(a.w);
                    lo.x = __float_as_uint(a.x - __uint_as_float(hi.x)); lo.y = __float_as_uint(a.y - __uint_as_float(hi.y));
                    lo.z = __float_as_uint(a.z - __uint_as_float(hi.z)); lo.w = __float_as_uint(a.w - __uint_as_float(hi.w));
                    *reinterpret_cast<uint4 *>(st + off) = hi;
                    *reinterpret_cast<uint4 *>(st + TILE_BYTES + off) = lo;
                    const float4 kk = k[c];
                    hi.x = rn_tf32(kk.x); hi.y = rn_tf32(kk.y); hi.z = rn_tf32(kk.z); hi.w = rn_tf32(kk.w);
                    lo.x = __float_as_uint(kk.x - __uint_as_float(hi.x)); lo.y = __float_as_uint(kk.y - __uint_as_float(hi.y));
                    lo.z = __float_as_uint(kk.z - __uint_as_float(hi.z)); lo.w = __float_as_uint(kk.w - __uint_as_float(hi.w));
                    *reinterpret_cast<uint4 *>(st + 2 * TILE_BYTES + off) = hi;
                    *reinterpret_cast<uint4 *>(st + 3 * TILE_BYTES + off) = lo;
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic writes -> async proxy (UMMA)
            __syncwarp();
            if (lane == 0) mbar_arrive(&in_full[st_i]);
        };
        // two register sets: the loads of unit i + 1 are in flight while unit i is converted
        float4 qa[4], ka[4], qb[4], kb[4];
        load(0, qa, ka);
        for (int64_t i = 0; i < n_units; i += 2) {
            load(i + 1, qb, kb);
            stage(i, qa, ka);
            if (i + 1 < n_units) {
                load(i + 2, qa, ka);
                stage(i + 1, qb, kb);
            }
        }
    } else {
        // ================================================= consumers: thread = query row = TMEM lane
        const int cw = warp - (1 + kProducerWarps);             // 0..7
        const int buf = cw >> 2, quarter = warp & 3;
        const int row = 32 * quarter + lane;
        const uint32_t t0 = tmem_base + ((uint32_t)(32 * quarter) << 16) + (uint32_t)(buf * BUF_COLS);
        // The group stages the V of its own unit (thread = key row; pad keys write zeros) into the slot of its TMEM buffer,
        // as the MN-major B operand: [dim half][key][8 dims], four 16-byte stores per thread.  The slot is free once the
        // previous unit's O was seen, and the p_full arrival below publishes it together with P.
        unsigned char *vimg = sVT + buf * VT_SLOT_BYTES + row * 16;
        // the raw row travels by cp.async into a staging buffer of the group (no registers held across the softmax, whose
        // score row needs all of them): issued one unit ahead, right after P is published
        unsigned char *vraw = sVraw + buf * V_STAGE_BYTES + row * 16;
        auto fetch_v = [&](int64_t i) {
            if (row < N && i < n_units) {
                const float *p = qkv + ((blockIdx.x + (i >> 3) * gridDim.x) * N + row) * (3 * E) + 2 * E + (int)(i & 7) * D;
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(vraw + c * 2048)), "l"(p + 4 * c)
                                 : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        fetch_v(buf);
        for (int64_t i = buf; i < n_units; i += 2) {
            const int h = (int)(i & 7);
            const int64_t b = blockIdx.x + (i >> 3) * gridDim.x;
            const unsigned ph = (unsigned)((i >> 1) & 1);
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            // The V tile of the unit is scaled by a power of two so that its largest entry lands in 2^13..2^14 (as the
            // projection kernel scales its tiles): fp16's range then holds hi AND the residual lo of every entry within 2^-11
            // of the largest, whatever the magnitude of the values (without it, values below 2^-3 lose bits of lo and an
            // encoder whose values are all tiny loses them all - found by tools/soak_round2.py).  The group's 128 rows agree
            // on the scale through shared memory: every warp posts the range of its rows here and the tile is staged after
            // the softmax, when the four posts have long been made - a barrier at this point costs 0.2 ms per launch.
            uint32_t *xch = s_amax + (buf * 2 + (int)ph) * 4;
            {
                float amax = 0.f;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float4 v4 = row < N ? *reinterpret_cast<const float4 *>(vraw + c * 2048) : make_float4(0.f, 0.f, 0.f, 0.f);
                    amax = fmaxf(fmaxf(amax, fmaxf(fabsf(v4.x), fabsf(v4.y))), fmaxf(fabsf(v4.z), fabsf(v4.w)));
                }
                uint32_t ab = __float_as_uint(amax);
                asm("redux.sync.max.u32 %0, %1, 0xffffffff;" : "=r"(ab) : "r"(ab));
                if (lane == 0) {
                    xch[quarter] = ab;
                    mbar_arrive(&v_rng[buf]);
                }
            }
            mbar_wait_bounded(&s_full[buf], ph);
            tc_fence_after();
            // ONE read of the score row (112 registers); columns beyond N (K rows of zeros) -> -inf
            uint32_t sv[NK];
            tmem_ld_row(t0, sv);
#pragma unroll
            for (int c = 0; c < NK / 16; ++c)
                if (N < 16 * (c + 1)) {
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        if (16 * c + j >= N) sv[16 * c + j] = 0xff800000u;
                }
            float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
            for (int j = 0; j < NK; ++j) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(sv[j]));
            const float m = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) - kPShift;
            float l = 0.f;                                      // P = 2^(s - max + 14) as fp16 pairs hi / lo, in place
#pragma unroll
            for (int c = 0; c < NK / 16; ++c) {
                uint32_t hi[8], lo[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float p0 = ex2f(__uint_as_float(sv[16 * c + 2 * k]) - m);
                    const float p1 = ex2f(__uint_as_float(sv[16 * c + 2 * k + 1]) - m);
                    l += p0 + p1;
                    split_f16x2_trunc(p0, p1, hi[k], lo[k]);    // no F2F conversions: they share the MUFU pipe with ex2
                }
                tmem_st8(t0 + 8 * c, hi);
                tmem_st8(t0 + COL_PLO + 8 * c, lo);
            }
            float unscale;
            {                                                   // V as the MN-major fp16 operand, scaled (see above)
                mbar_wait_bounded(&v_rng[buf], ph);
                const uint4 all = *reinterpret_cast<const uint4 *>(xch);
                const uint32_t ab = max(max(all.x, all.y), max(all.z, all.w));
                int e = (int)(ab >> 23);                        // biased exponent of the largest |v| (nan / inf: 255, they propagate)
                e = e < 27 ? 27 : e;                            // all-zero or denormal tiles: any scale will do
                const float vscale = __uint_as_float((uint32_t)(267 - e) << 23);    // 2^(13 - (e - 127))
                unscale = __uint_as_float((uint32_t)(e - 13) << 23);                // 2^((e - 127) - 13)
                uint32_t vh[8], vl[8];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float4 v4 = row < N ? *reinterpret_cast<const float4 *>(vraw + c * 2048) : make_float4(0.f, 0.f, 0.f, 0.f);
                    split_f16x2_trunc(v4.x * vscale, v4.y * vscale, vh[2 * c], vl[2 * c]);
                    split_f16x2_trunc(v4.z * vscale, v4.w * vscale, vh[2 * c + 1], vl[2 * c + 1]);
                }
                *reinterpret_cast<uint4 *>(vimg) = make_uint4(vh[0], vh[1], vh[2], vh[3]);
                *reinterpret_cast<uint4 *>(vimg + V_HALF) = make_uint4(vh[4], vh[5], vh[6], vh[7]);
                *reinterpret_cast<uint4 *>(vimg + VT_BYTES) = make_uint4(vl[0], vl[1], vl[2], vl[3]);
                *reinterpret_cast<uint4 *>(vimg + VT_BYTES + V_HALF) = make_uint4(vl[4], vl[5], vl[6], vl[7]);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic writes -> async proxy (UMMA)
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[buf]);
            fetch_v(i + 2);                                     // the group's next unit: in flight during the P V
            // O = P V arrives from the tensor core; normalise and store this row's 16 outputs
            mbar_wait_bounded(&o_full[buf], ph);
            tc_fence_after();
            uint32_t ov[16];
            tmem_ld16(t0 + COL_O, ov);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&buf_free[buf]);
            if (row < N) {
                const float inv = unscale / l;
                float4 *dst = reinterpret_cast<float4 *>(out + (b * N + row) * E + h * D);
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    dst[c] = make_float4(__uint_as_float(ov[4 * c]) * inv, __uint_as_float(ov[4 * c + 1]) * inv,
                                         __uint_as_float(ov[4 * c + 2]) * inv, __uint_as_float(ov[4 * c + 3]) * inv);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

}  // namespace enctc3


// ============================================================================================
// enctc3 with FOUR consumer groups (21 warps) and the two-pass, rolled softmax of enctc2: what limited enctc3 was warps per
// scheduler x registers - its one-pass softmax holds the 112-register score row, which caps the CTA at 13 warps.  Here the
// consumers read the row twice in 16-column chunks (~50 registers), so 16 consumer warps fit (80 registers per thread) and
// four units are in flight instead of two.  P goes back IN PLACE per chunk: the 16 score columns of chunk c become 8 columns
// of fp16 hi pairs and 8 columns of lo pairs, so a buffer is 112 + 16 (O) = 128 columns and four of them fill the tensor
// memory.  Everything else as enctc3: 3xTF32 scores, kind::f16 P V, V as an MN-major range-scaled operand staged by the group.
namespace enctc4 {
using namespace tcu;
using enctc::ex2f;
using enctc2::tmem_ld16;
using enctc3::tmem_st8;
using enctc3::umma_f16_ts;
using enctc3::rn_tf32;

constexpr int H = 8, D = 16, E = 128, NK = 112;
constexpr int kGroups = 4;                              // consumer groups = TMEM buffers = V slots = units in flight
constexpr int TILE_BYTES = 128 * 128;
constexpr int V_HALF = 128 * 16, VT_BYTES = 2 * V_HALF, VT_SLOT_BYTES = 2 * VT_BYTES, V_STAGE_BYTES = 4 * 128 * 16;
constexpr int QK_STAGE_BYTES = 4 * TILE_BYTES, QK_STAGES = 2;
constexpr int kProducerWarps = 4, kConsumerWarps = 4 * kGroups;
constexpr int kThreads = (1 + kProducerWarps + kConsumerWarps) * 32;       // 672
constexpr int SMEM_BYTES = 1024 + QK_STAGES * QK_STAGE_BYTES + kGroups * (VT_SLOT_BYTES + V_STAGE_BYTES) + 512;
constexpr uint32_t kIdescS = make_idesc_tf32(128, NK);
constexpr uint32_t kIdescO = make_idesc_f16(128, D) | (1u << 16);
constexpr int BUF_COLS = 128, COL_O = NK;
constexpr float kPShift = 14.f;
#ifndef ENC4_LAG
#define ENC4_LAG 1                                      // the MMA warp issues P V of unit i - ENC4_LAG after Q K^T of unit i
                                                        // (measured per 16 384 x 100: lag 1 1.30 ms, 2 1.34, 3 1.37; unrolling
                                                        // the chunk loop by 2 / 4: 1.32 / 1.51 ms)
#endif
#ifndef ENC4_UNROLL
#define ENC4_UNROLL 1
#endif
constexpr int kChunkUnroll = ENC4_UNROLL;

__global__ void __launch_bounds__(kThreads, 1)
encoder_attention_tc4_kernel(int64_t B, int N, const float *__restrict__ qkv, float *__restrict__ out) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char *sVT = smem + QK_STAGES * QK_STAGE_BYTES;
    unsigned char *sVraw = sVT + kGroups * VT_SLOT_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sVraw + kGroups * V_STAGE_BYTES);
    uint64_t *in_full = bars, *qk_empty = bars + 2;             // [2] each
    uint64_t *s_full = bars + 4, *p_full = bars + 8, *o_full = bars + 12, *buf_free = bars + 16, *v_rng = bars + 20;   // [kGroups] each
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 24);
    uint32_t *s_amax = tmem_slot + 4;                           // [group][unit parity][warp of the group]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t n_inst = blockIdx.x < B ? (B - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t n_units = n_inst * H;

    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(&in_full[s], kProducerWarps);
            mbar_init(&qk_empty[s], 1);
        }
        for (int s = 0; s < kGroups; ++s) {
            mbar_init(&s_full[s], 1);
            mbar_init(&p_full[s], 4);
            mbar_init(&o_full[s], 1);
            mbar_init(&buf_free[s], 4);
            mbar_init(&v_rng[s], 4);
        }
        fence_barrier_init();
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================================= MMA issue: QK(i), then PV(i - ENC4_LAG)
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t d0 = (uint32_t)smem_desc(smem_u32(smem));
        const uint32_t dv0 = ((smem_u32(sVT) >> 4) & 0x3FFFu) | enctc3::kDescLoV;
        for (int64_t i = 0; i < n_units + ENC4_LAG; ++i) {
            if (i < n_units) {
                const int st = (int)(i & 1), x = (int)(i & (kGroups - 1));
                mbar_wait_bounded(&in_full[st], (unsigned)((i >> 1) & 1));
                mbar_wait_bounded(&buf_free[x], (unsigned)((i / kGroups) & 1) ^ 1u);
                tc_fence_after();
                const uint32_t q_hi = d0 + (uint32_t)((st * QK_STAGE_BYTES) >> 4), q_lo = q_hi + (TILE_BYTES >> 4);
                const uint32_t k_hi = q_hi + 2 * (TILE_BYTES >> 4), k_lo = q_hi + 3 * (TILE_BYTES >> 4);
                const uint32_t d = tmem_u + (uint32_t)(x * BUF_COLS);
                if (elect_one()) {
#pragma unroll
                    for (int k8 = 0; k8 < D / 8; ++k8) {
                        umma_tf32(d, q_lo + 2 * k8, k_hi + 2 * k8, kIdescS, (unsigned)(k8 != 0));
                        umma_tf32(d, q_hi + 2 * k8, k_lo + 2 * k8, kIdescS, 1u);
                        umma_tf32(d, q_hi + 2 * k8, k_hi + 2 * k8, kIdescS, 1u);
                    }
                    umma_commit(&qk_empty[st]);
                    umma_commit(&s_full[x]);
                }
                __syncwarp();
            }
            if (i >= ENC4_LAG) {
                const int64_t u = i - ENC4_LAG;
                const int x = (int)(u & (kGroups - 1));
                mbar_wait_bounded(&p_full[x], (unsigned)((u / kGroups) & 1));
                tc_fence_after();
                const uint32_t vt_hi = dv0 + (uint32_t)((x * VT_SLOT_BYTES) >> 4), vt_lo = vt_hi + (VT_BYTES >> 4);
                const uint32_t pb = tmem_u + (uint32_t)(x * BUF_COLS);
                const uint32_t o = pb + COL_O;
                if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < NK / 16; ++ks) {      // chunk ks: hi pairs in columns 16 ks .. + 7, lo pairs in + 8 .. + 15
                        umma_f16_ts(o, pb + 16 * ks + 8, vt_hi + 16 * ks, kIdescO, (unsigned)(ks != 0));
                        umma_f16_ts(o, pb + 16 * ks, vt_lo + 16 * ks, kIdescO, 1u);
                        umma_f16_ts(o, pb + 16 * ks, vt_hi + 16 * ks, kIdescO, 1u);
                    }
                    umma_commit(&o_full[x]);
                }
                __syncwarp();
            }
        }
    } else if (warp <= kProducerWarps) {
        // ================================================= producers: thread = node row; Q and K of one (b, h) per unit
        const int row = tid - 32;
        const float scale = 0.25f * 1.4426950408889634f;
        if (row >= N) {
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
            for (int s = 0; s < QK_STAGES; ++s)
                for (int img = 0; img < 4; ++img)
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        *reinterpret_cast<uint4 *>(smem + s * QK_STAGE_BYTES + img * TILE_BYTES + swz(row, c)) = z;
        }
        for (int64_t i = 0; i < n_units; ++i) {
            const int st_i = (int)(i & 1), h = (int)(i & 7);
            const int64_t b = blockIdx.x + (i >> 3) * gridDim.x;
            float4 q[4], k[4];
            if (row < N) {                                      // loads first: they fly while we wait for the stage
                const float4 *p = reinterpret_cast<const float4 *>(qkv + (b * N + row) * (3 * E) + h * D);
                if (h == 0 && i + H < n_units) {                // this row of the CTA's next instance -> L2 (12 lines)
                    const char *nx = reinterpret_cast<const char *>(qkv + ((b + gridDim.x) * N + row) * (3 * E));
#pragma unroll
                    for (int c = 0; c < 12; ++c) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + 128 * c));
                }
#pragma unroll
                for (int c = 0; c < 4; ++c) { q[c] = __ldg(p + c); k[c] = __ldg(p + E / 4 + c); }
            }
            mbar_wait_bounded(&qk_empty[st_i], (unsigned)((i >> 1) & 1) ^ 1u);
            if (row < N) {
                unsigned char *st = smem + st_i * QK_STAGE_BYTES;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const uint32_t off = swz(row, c);
                    const float4 a = make_float4(q[c].x * scale, q[c].y * scale, q[c].z * scale, q[c].w * scale);
                    uint4 hi, lo;
                    hi.x = rn_tf32(a.x); hi.y = rn_tf32(a.y); hi.z = rn_tf32(a.z); hi.w = rn_tf32(a.w);
                    lo.x = __float_as_uint(a.x - __uint_as_float(hi.x)); lo.y = __float_as_uint(a.y - __uint_as_float(hi.y));
                    lo.z = __float_as_uint(a.z - __uint_as_float(hi.z)); lo.w = __float_as_uint(a.w - __uint_as_float(hi.w));
                    *reinterpret_cast<uint4 *>(st + off) = hi;
                    *reinterpret_cast<uint4 *>(st + TILE_BYTES + off) = lo;
                    const float4 kk = k[c];
                    hi.x = rn_tf32(kk.x); hi.y = rn_tf32(kk.y); hi.z = rn_tf32(kk.z); hi.w = rn_tf32(kk.w);
                    lo.x = __float_as_uint(kk.x - __uint_as_float(hi.x)); lo.y = __float_as_uint(kk.y - __uint_as_float(hi.y));
                    lo.z = __float_as_uint(kk.z - __uint_as_float(hi.z)); lo.w = __float_as_uint(kk.w - __uint_as_float(hi.w));
                    *reinterpret_cast<uint4 *>(st + 2 * TILE_BYTES + off) = hi;
                    *reinterpret_cast<uint4 *>(st + 3 * TILE_BYTES + off) = lo;
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&in_full[st_i]);
        }
    } else {
        // ================================================= consumers: thread = query row = TMEM lane; group g: units g, g + 4, ...
        const int cw = warp - (1 + kProducerWarps);             // 0..15
        const int grp = cw >> 2, quarter = warp & 3;
        const int row = 32 * quarter + lane;
        const uint32_t t0 = tmem_base + ((uint32_t)(32 * quarter) << 16) + (uint32_t)(grp * BUF_COLS);
        const int n_full = N >> 4, tail = N & 15;
        unsigned char *vimg = sVT + grp * VT_SLOT_BYTES + row * 16;
        unsigned char *vraw = sVraw + grp * V_STAGE_BYTES + row * 16;
        auto fetch_v = [&](int64_t i) {
            if (row < N && i < n_units) {
                const float *p = qkv + ((blockIdx.x + (i >> 3) * gridDim.x) * N + row) * (3 * E) + 2 * E + (int)(i & 7) * D;
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(vraw + c * 2048)), "l"(p + 4 * c)
                                 : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        fetch_v(grp);
        for (int64_t i = grp; i < n_units; i += kGroups) {
            const int h = (int)(i & 7);
            const int64_t b = blockIdx.x + (i >> 3) * gridDim.x;
            const unsigned ph = (unsigned)((i / kGroups) & 1);
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            uint32_t *xch = s_amax + (grp * 2 + (int)ph) * 4;
            {                                                   // post the range of this warp's V rows (see enctc3)
                float amax = 0.f;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float4 v4 = row < N ? *reinterpret_cast<const float4 *>(vraw + c * 2048) : make_float4(0.f, 0.f, 0.f, 0.f);
                    amax = fmaxf(fmaxf(amax, fmaxf(fabsf(v4.x), fabsf(v4.y))), fmaxf(fabsf(v4.z), fabsf(v4.w)));
                }
                uint32_t ab = __float_as_uint(amax);
                asm("redux.sync.max.u32 %0, %1, 0xffffffff;" : "=r"(ab) : "r"(ab));
                if (lane == 0) {
                    xch[quarter] = ab;
                    mbar_arrive(&v_rng[grp]);
                }
            }
            mbar_wait_bounded(&s_full[grp], ph);
            tc_fence_after();
            uint32_t sv[16];
            float m = -INFINITY;                                // pass 1: row maximum over the N real keys
#pragma unroll 1
            for (int c = 0; c < n_full; ++c) {
                tmem_ld16(t0 + c * 16, sv);
#pragma unroll
                for (int j = 0; j < 16; ++j) m = fmaxf(m, __uint_as_float(sv[j]));
            }
            if (tail) {
                tmem_ld16(t0 + n_full * 16, sv);
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (j < tail) m = fmaxf(m, __uint_as_float(sv[j]));
            }
            m -= kPShift;
            float l = 0.f;                                      // pass 2: P = 2^(s - max + 14) as fp16 pairs, in place per chunk
            uint32_t hi[8], lo[8];
#pragma unroll kChunkUnroll
            for (int c = 0; c < n_full; ++c) {
                tmem_ld16(t0 + c * 16, sv);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float p0 = ex2f(__uint_as_float(sv[2 * k]) - m), p1 = ex2f(__uint_as_float(sv[2 * k + 1]) - m);
                    l += p0 + p1;
                    split_f16x2_trunc(p0, p1, hi[k], lo[k]);
                }
                tmem_st8(t0 + c * 16, hi);
                tmem_st8(t0 + c * 16 + 8, lo);
            }
            int c_next = n_full;
            if (tail) {
                tmem_ld16(t0 + n_full * 16, sv);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float p0 = 2 * k < tail ? ex2f(__uint_as_float(sv[2 * k]) - m) : 0.f;
                    const float p1 = 2 * k + 1 < tail ? ex2f(__uint_as_float(sv[2 * k + 1]) - m) : 0.f;
                    l += p0 + p1;
                    split_f16x2_trunc(p0, p1, hi[k], lo[k]);
                }
                tmem_st8(t0 + n_full * 16, hi);
                tmem_st8(t0 + n_full * 16 + 8, lo);
                c_next = n_full + 1;
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) hi[k] = 0u;
#pragma unroll 1
            for (int c = c_next; c < NK / 16; ++c) {            // key columns beyond N: P = 0
                tmem_st8(t0 + c * 16, hi);
                tmem_st8(t0 + c * 16 + 8, hi);
            }
            float unscale;
            {                                                   // V as the MN-major fp16 operand, range-scaled (see enctc3)
                mbar_wait_bounded(&v_rng[grp], ph);
                const uint4 all = *reinterpret_cast<const uint4 *>(xch);
                const uint32_t ab = max(max(all.x, all.y), max(all.z, all.w));
                int e = (int)(ab >> 23);
                e = e < 27 ? 27 : e;
                const float vscale = __uint_as_float((uint32_t)(267 - e) << 23);
                unscale = __uint_as_float((uint32_t)(e - 13) << 23);
                uint32_t vh[8], vl[8];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float4 v4 = row < N ? *reinterpret_cast<const float4 *>(vraw + c * 2048) : make_float4(0.f, 0.f, 0.f, 0.f);
                    split_f16x2_trunc(v4.x * vscale, v4.y * vscale, vh[2 * c], vl[2 * c]);
                    split_f16x2_trunc(v4.z * vscale, v4.w * vscale, vh[2 * c + 1], vl[2 * c + 1]);
                }
                *reinterpret_cast<uint4 *>(vimg) = make_uint4(vh[0], vh[1], vh[2], vh[3]);
                *reinterpret_cast<uint4 *>(vimg + V_HALF) = make_uint4(vh[4], vh[5], vh[6], vh[7]);
                *reinterpret_cast<uint4 *>(vimg + VT_BYTES) = make_uint4(vl[0], vl[1], vl[2], vl[3]);
                *reinterpret_cast<uint4 *>(vimg + VT_BYTES + V_HALF) = make_uint4(vl[4], vl[5], vl[6], vl[7]);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[grp]);
            fetch_v(i + kGroups);
            mbar_wait_bounded(&o_full[grp], ph);
            tc_fence_after();
            tmem_ld16(t0 + COL_O, sv);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&buf_free[grp]);
            if (row < N) {
                const float inv = unscale / l;
                float4 *dst = reinterpret_cast<float4 *>(out + (b * N + row) * E + h * D);
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    dst[c] = make_float4(__uint_as_float(sv[4 * c]) * inv, __uint_as_float(sv[4 * c + 1]) * inv,
                                         __uint_as_float(sv[4 * c + 2]) * inv, __uint_as_float(sv[4 * c + 3]) * inv);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

}  // namespace enctc4

int launch_encoder_attention_tc(int64_t B, int N, const float *qkv, float *out, int impl, cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    const int64_t grid = B < sm_count ? B : sm_count;
    // 0 automatic: both GEMMs on the tensor cores with P as fp16 pairs and four consumer groups (enctc4) for N <= 112, else
    // scores only (enctc)
    const int kind = impl == 0 ? (N <= enctc4::NK ? 4 : 1) : impl;
    if (kind >= 2 && N > enctc2::NK) {
        set_error("encoder_attention(tc): the two-GEMM kernels need N <= %d, got %d", enctc2::NK, N);
        return 1;
    }
    const int smem = kind == 4 ? enctc4::SMEM_BYTES : kind == 3 ? enctc3::SMEM_BYTES : (kind == 2 ? enctc2::SMEM_BYTES : enctc::SMEM_BYTES);
    const void *fn = kind == 4 ? (const void *)enctc4::encoder_attention_tc4_kernel : kind == 3 ? (const void *)enctc3::encoder_attention_tc3_kernel
                               : (kind == 2 ? (const void *)enctc2::encoder_attention_tc2_kernel
                                            : (const void *)enctc::encoder_attention_tc_kernel);
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) {
        set_error("encoder_attention(tc): cannot reserve %d B of shared memory: %s", smem, cudaGetErrorString(e));
        return 2;
    }
    if (kind == 4)
        enctc4::encoder_attention_tc4_kernel<<<(unsigned)grid, enctc4::kThreads, smem, stream>>>(B, N, qkv, out);
    else if (kind == 3)
        enctc3::encoder_attention_tc3_kernel<<<(unsigned)grid, enctc3::kThreads, smem, stream>>>(B, N, qkv, out);
    else if (kind == 2)
        enctc2::encoder_attention_tc2_kernel<<<(unsigned)grid, enctc2::kThreads, smem, stream>>>(B, N, qkv, out);
    else
        enctc::encoder_attention_tc_kernel<<<(unsigned)grid, enctc::kThreads, smem, stream>>>(B, N, qkv, out);
    RL4CO_CHECK_LAUNCH("encoder_attention(tc)");
    return 0;
}

}  // namespace rl4co

using namespace rl4co;

static int g_attention_impl = 0;    // 0 auto (enctc4 for N <= 112), 1 Q K^T only, 2 two-GEMM kernel 3xTF32, 3 / 4 two-GEMM kernels with fp16 P V

extern "C" int rl4co_set_attention_impl(int32_t impl) {
    RL4CO_REQUIRE(impl >= 0 && impl <= 4, "set_attention_impl: 0 auto, 1 scores on tensor cores, 2 both GEMMs (TF32), 3 / 4 both GEMMs (fp16 P V; 4: four consumer groups), got %d", impl);
    g_attention_impl = impl;
    return 0;
}

extern "C" int rl4co_encoder_attention_tc(int64_t B, int32_t N, const float *qkv, float *out, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N >= 1 && N <= 128, "encoder_attention_tc: B=%lld N=%d (N <= 128)", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(qkv && out, "encoder_attention_tc: null pointer");
    RL4CO_REQUIRE(((uintptr_t)qkv & 15) == 0 && ((uintptr_t)out & 15) == 0,
                  "encoder_attention_tc: pointers must be 16-byte aligned");
    return launch_encoder_attention_tc(B, N, qkv, out, g_attention_impl, (cudaStream_t)stream);
}
