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

int launch_encoder_attention_tc(int64_t B, int N, const float *qkv, float *out, cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    cudaError_t e = cudaFuncSetAttribute(enctc::encoder_attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         enctc::SMEM_BYTES);
    if (e != cudaSuccess) {
        set_error("encoder_attention(tc): cannot reserve %d B of shared memory: %s", enctc::SMEM_BYTES, cudaGetErrorString(e));
        return 2;
    }
    const int64_t grid = B < sm_count ? B : sm_count;
    enctc::encoder_attention_tc_kernel<<<(unsigned)grid, enctc::kThreads, enctc::SMEM_BYTES, stream>>>(B, N, qkv, out);
    RL4CO_CHECK_LAUNCH("encoder_attention(tc)");
    return 0;
}

}  // namespace rl4co

using namespace rl4co;

extern "C" int rl4co_encoder_attention_tc(int64_t B, int32_t N, const float *qkv, float *out, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N >= 1 && N <= 128, "encoder_attention_tc: B=%lld N=%d (N <= 128)", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(qkv && out, "encoder_attention_tc: null pointer");
    RL4CO_REQUIRE(((uintptr_t)qkv & 15) == 0 && ((uintptr_t)out & 15) == 0,
                  "encoder_attention_tc: pointers must be 16-byte aligned");
    return launch_encoder_attention_tc(B, N, qkv, out, (cudaStream_t)stream);
}
