// Encoder self-attention for the small graphs of the path (N <= 128 nodes, 8 heads x 16 dims, no mask),
// fp32.  The PyTorch encoder stays the owner of the layer; this kernel replaces its
// F.scaled_dot_product_attention call (rl4co/models/nn/attention.py:65-135), which PyTorch runs for fp32
// inputs as the sm_80 memory-efficient kernel at ~3 TFLOP/s and which was 40 % of the end-to-end time.
//
// One CTA per instance (grid-stride), warp = head.  K_h and V_h (N x 16 fp32 each) are staged in shared
// memory; a lane owns query rows lane, lane+32, ... and walks all keys twice: pass A row maximum, pass B
// exponentials (base 2, scale folded into q) and the p.V accumulation, all as FFMA2 on packed pairs with
// K / V rows read as broadcast LDS.128.  Input is the fused projection output [B, N, 3, 8, 16]
// (Wqkv), output [B, N, 128] with heads concatenated ("b h n d -> b n (h d)").
#include "rollout_util.cuh"

namespace rl4co {
namespace enc {

typedef unsigned long long u64;
constexpr int H = 8, D = 16, E = 128;

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

__global__ void __launch_bounds__(256, 2)
encoder_attention_kernel(int64_t B, int N, const float *__restrict__ qkv, float *__restrict__ out) {
    extern __shared__ __align__(16) float sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *Ks = sm + (size_t)warp * 2 * N * D, *Vs = Ks + (size_t)N * D;
    const float scale = 0.25f * 1.4426950408889634f;           // 1/sqrt(16) * log2(e)
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        const float *base = qkv + b * (int64_t)N * 3 * E;
        __syncwarp();                                           // previous instance's reads are done
        for (int idx = lane; idx < N * 4; idx += 32) {          // K_h, V_h rows: 4 float4 each
            const int n = idx >> 2, c = idx & 3;
            const float4 k = __ldg(reinterpret_cast<const float4 *>(base + (size_t)n * 3 * E + E + warp * D) + c);
            const float4 v = __ldg(reinterpret_cast<const float4 *>(base + (size_t)n * 3 * E + 2 * E + warp * D) + c);
            reinterpret_cast<float4 *>(Ks)[idx] = k;
            reinterpret_cast<float4 *>(Vs)[idx] = v;
        }
        __syncwarp();
        for (int r0 = lane; r0 < N; r0 += 64) {                 // two query rows per pass: r0 and r0 + 32
            const int r1 = r0 + 32;
            const bool has1 = r1 < N;
            u64 q0[8], q1[8];
            {
                const float4 *p0 = reinterpret_cast<const float4 *>(base + (size_t)r0 * 3 * E + warp * D);
                const float4 *p1 = reinterpret_cast<const float4 *>(base + (size_t)(has1 ? r1 : r0) * 3 * E + warp * D);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float4 a = __ldg(p0 + c), bq = __ldg(p1 + c);
                    q0[2 * c] = pack2(a.x * scale, a.y * scale); q0[2 * c + 1] = pack2(a.z * scale, a.w * scale);
                    q1[2 * c] = pack2(bq.x * scale, bq.y * scale); q1[2 * c + 1] = pack2(bq.z * scale, bq.w * scale);
                }
            }
            auto score = [&](const u64 *q, const ulonglong2 *kr) {
                u64 a = 0ull, c = 0ull;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const ulonglong2 kk = kr[i];
                    a = ffma2(q[2 * i], kk.x, a);
                    c = ffma2(q[2 * i + 1], kk.y, c);
                }
                float a0, a1, c0, c1;
                unpack2(a, a0, a1);
                unpack2(c, c0, c1);
                return (a0 + a1) + (c0 + c1);
            };
            float m0 = -INFINITY, m1 = -INFINITY;
            for (int n = 0; n < N; ++n) {                       // pass A: row maxima
                const ulonglong2 *kr = reinterpret_cast<const ulonglong2 *>(Ks + n * D);
                m0 = fmaxf(m0, score(q0, kr));
                m1 = fmaxf(m1, score(q1, kr));
            }
            u64 acc0[8], acc1[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) acc0[i] = acc1[i] = 0ull;
            float l0 = 0.f, l1 = 0.f;
            for (int n = 0; n < N; ++n) {                       // pass B: softmax numerators and p.V
                const ulonglong2 *kr = reinterpret_cast<const ulonglong2 *>(Ks + n * D);
                const ulonglong2 *vr = reinterpret_cast<const ulonglong2 *>(Vs + n * D);
                const float p0 = ex2f(score(q0, kr) - m0), p1 = ex2f(score(q1, kr) - m1);
                l0 += p0;
                l1 += p1;
                const u64 pp0 = pack2(p0, p0), pp1 = pack2(p1, p1);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const ulonglong2 vv = vr[i];
                    acc0[2 * i] = ffma2(vv.x, pp0, acc0[2 * i]); acc0[2 * i + 1] = ffma2(vv.y, pp0, acc0[2 * i + 1]);
                    acc1[2 * i] = ffma2(vv.x, pp1, acc1[2 * i]); acc1[2 * i + 1] = ffma2(vv.y, pp1, acc1[2 * i + 1]);
                }
            }
            auto store = [&](int r, const u64 *acc, float l) {
                const float inv = __fdividef(1.f, l);
                float4 *dst = reinterpret_cast<float4 *>(out + (b * (int64_t)N + r) * E + warp * D);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float x, y, z, w;
                    unpack2(acc[2 * c], x, y);
                    unpack2(acc[2 * c + 1], z, w);
                    dst[c] = make_float4(x * inv, y * inv, z * inv, w * inv);
                }
            };
            store(r0, acc0, l0);
            if (has1) store(r1, acc1, l1);
        }
    }
}

}  // namespace enc
}  // namespace rl4co

using namespace rl4co;

extern "C" int rl4co_encoder_attention(int64_t B, int32_t N, const float *qkv, float *out, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N >= 1 && N <= 128, "encoder_attention: B=%lld N=%d (N <= 128)", (long long)B, N);
    if (B == 0) return 0;
    RL4CO_REQUIRE(qkv && out, "encoder_attention: null pointer");
    RL4CO_REQUIRE(((uintptr_t)qkv & 15) == 0 && ((uintptr_t)out & 15) == 0, "encoder_attention: pointers must be 16-byte aligned");
    const size_t smem = (size_t)enc::H * 2 * N * enc::D * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(enc::encoder_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        set_error("encoder_attention: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e));
        return 2;
    }
    const int64_t grid = B < 148 * 2 ? B : 148 * 2;
    enc::encoder_attention_kernel<<<(unsigned)grid, 256, smem, (cudaStream_t)stream>>>(B, N, qkv, out);
    RL4CO_CHECK_LAUNCH("encoder_attention");
    return 0;
}
