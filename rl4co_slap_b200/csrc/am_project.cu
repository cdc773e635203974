// K3 - one-off projection of encoder node embeddings into the rollout tables, and the
// graph context.  Reference: rl4co/models/zoo/am/decoder.py:208-237 (_precompute_cache),
// rl4co/models/nn/env_embeddings/context.py:67-71 (the per-node halves of project_context).
//
// v1: fp32 SIMT register-tiled GEMM (128x128 CTA tile, 8x8 per thread, K staged in 32-wide
// slabs), exact fp32 FMA accumulation in k order so the tables match torch's fp32 Linear to
// ~1 ulp-level differences.  [the tcgen05 3xTF32 version replaces this in a later step]
#include "common.cuh"

namespace rl4co {

constexpr int PE = RL4CO_EMBED_DIM;   // K dimension = 128
constexpr int BM = 128, BN = 128, BK = 32;
constexpr int PLD = BM + 4;

__global__ void __launch_bounds__(256)
am_project_kernel(int64_t M, const float *__restrict__ h, const float *__restrict__ W,
                  float *__restrict__ tables) {
    __shared__ __align__(16) float As[BK][PLD];
    __shared__ __align__(16) float Ws[BK][PLD];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int64_t m0 = (int64_t)blockIdx.x * BM;
    const int tab = blockIdx.y;                         // one table (128 output features) per y
    const float *Wt = W + (size_t)tab * BN * PE;

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < PE; k0 += BK) {
        // stage A[m0..m0+127][k0..k0+31] and W[0..127][k0..k0+31], transposed to k-major
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            int idx = tid + it * 256;
            int row = idx >> 3, kq = idx & 7;
            float4 av = make_float4(0.f, 0.f, 0.f, 0.f);
            if (m0 + row < M) av = __ldg(reinterpret_cast<const float4 *>(h + (m0 + row) * PE + k0) + kq);
            As[kq * 4 + 0][row] = av.x; As[kq * 4 + 1][row] = av.y;
            As[kq * 4 + 2][row] = av.z; As[kq * 4 + 3][row] = av.w;
            float4 wv = __ldg(reinterpret_cast<const float4 *>(Wt + (size_t)row * PE + k0) + kq);
            Ws[kq * 4 + 0][row] = wv.x; Ws[kq * 4 + 1][row] = wv.y;
            Ws[kq * 4 + 2][row] = wv.z; Ws[kq * 4 + 3][row] = wv.w;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float4 a0 = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
            float4 a1 = *reinterpret_cast<const float4 *>(&As[k][64 + ty * 4]);
            float4 b0 = *reinterpret_cast<const float4 *>(&Ws[k][tx * 4]);
            float4 b1 = *reinterpret_cast<const float4 *>(&Ws[k][64 + tx * 4]);
            float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    float *out = tables + (size_t)tab * M * PE;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        int64_t m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m < M) {
            *reinterpret_cast<float4 *>(out + m * PE + tx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
            *reinterpret_cast<float4 *>(out + m * PE + 64 + tx * 4) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
        }
    }
}

// gctx[b] = Wfixed . mean_n h[b,n,:]   (am/decoder.py:225-228)
__global__ void __launch_bounds__(256)
am_graph_context_kernel(int64_t B, int N, const float *__restrict__ h,
                        const float *__restrict__ Wfixed, float *__restrict__ gctx) {
    extern __shared__ __align__(16) float gsm[];
    float *Wt = gsm;                         // [e][j] transposed, stride PE+1
    float *mean = gsm + PE * (PE + 1);       // [2][PE]
    float *part = mean + 2 * PE;             // [2][PE]
    const int tid = threadIdx.x;
    for (int idx = tid; idx < PE * PE; idx += 256) {
        int j = idx / PE, e = idx % PE;
        Wt[e * (PE + 1) + j] = Wfixed[idx];
    }
    __syncthreads();
    const int e = tid & (PE - 1), half = tid >> 7;
    // column sums with 16-byte loads: lane = 4 consecutive dims, warp w takes rows w, w + 8, ... (four rows in flight
    // per thread - the kernel is a pure stream over h and needs the bytes in flight, not the arithmetic)
    const int c4 = tid & 31, rg = tid >> 5;
    float4 *colsum = reinterpret_cast<float4 *>(part + 2 * PE);     // [8][32] float4
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        const float4 *hb = reinterpret_cast<const float4 *>(h + b * (int64_t)N * PE) + c4;
        float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0, s2 = s0, s3 = s0;
        int n = rg;
        for (; n + 24 < N; n += 32) {
            const float4 v0 = __ldg(hb + (size_t)n * 32), v1 = __ldg(hb + (size_t)(n + 8) * 32);
            const float4 v2 = __ldg(hb + (size_t)(n + 16) * 32), v3 = __ldg(hb + (size_t)(n + 24) * 32);
            s0.x += v0.x; s0.y += v0.y; s0.z += v0.z; s0.w += v0.w;
            s1.x += v1.x; s1.y += v1.y; s1.z += v1.z; s1.w += v1.w;
            s2.x += v2.x; s2.y += v2.y; s2.z += v2.z; s2.w += v2.w;
            s3.x += v3.x; s3.y += v3.y; s3.z += v3.z; s3.w += v3.w;
        }
        for (; n < N; n += 8) {
            const float4 v0 = __ldg(hb + (size_t)n * 32);
            s0.x += v0.x; s0.y += v0.y; s0.z += v0.z; s0.w += v0.w;
        }
        colsum[rg * 32 + c4] = make_float4((s0.x + s1.x) + (s2.x + s3.x), (s0.y + s1.y) + (s2.y + s3.y),
                                           (s0.z + s1.z) + (s2.z + s3.z), (s0.w + s1.w) + (s2.w + s3.w));
        __syncthreads();
        if (half == 0) {
            const float *cs = reinterpret_cast<const float *>(colsum);
            float m = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) m += cs[w * PE + e];
            mean[e] = m / (float)N;
        }
        __syncthreads();
        float d = 0.f;
        for (int k = half * 64; k < half * 64 + 64; ++k) d = fmaf(mean[k], Wt[k * (PE + 1) + e], d);
        part[half * PE + e] = d;
        __syncthreads();
        if (half == 0) gctx[b * PE + e] = part[e] + part[PE + e];
        __syncthreads();
    }
}

}  // namespace rl4co

namespace rl4co {
int launch_pack_weights(int n_tables, const float *W, int64_t ldw, float *image, bool f16, cudaStream_t stream);
int launch_project_tc(int64_t M, int n_tables, int n_chunks, const float *h, const float *image, float *tables,
                      int64_t table_stride, int64_t ldo, const float *bias, int relu, int64_t ldx, int accumulate,
                      const float *residual, const float *scale, const float *shift, cudaStream_t stream);
}  // namespace rl4co

using namespace rl4co;

extern "C" int64_t rl4co_am_weight_image_floats(int32_t n_tables) { return (int64_t)n_tables * 32768; }

extern "C" int rl4co_am_pack_weights(int32_t n_tables, const float *W, float *image, rl4co_stream_t stream) {
    RL4CO_REQUIRE(n_tables >= 1 && n_tables <= 16, "am_pack_weights: bad n_tables=%d", n_tables);
    RL4CO_REQUIRE(W && image, "am_pack_weights: null pointer");
    RL4CO_REQUIRE(((uintptr_t)W & 15) == 0 && ((uintptr_t)image & 15) == 0, "am_pack_weights: pointers must be 16-byte aligned");
    return launch_pack_weights(n_tables, W, 128, image, true, (cudaStream_t)stream);
}

extern "C" int rl4co_am_project_tc(int64_t M, int32_t n_tables, const float *h, const float *image,
                                   float *tables, rl4co_stream_t stream) {
    RL4CO_REQUIRE(M >= 0 && n_tables >= 1 && n_tables <= 16, "am_project_tc: bad sizes M=%lld n_tables=%d", (long long)M, n_tables);
    if (M == 0) return 0;
    RL4CO_REQUIRE(h && image && tables, "am_project_tc: null pointer");
    RL4CO_REQUIRE(((uintptr_t)h & 15) == 0 && ((uintptr_t)image & 15) == 0 && ((uintptr_t)tables & 15) == 0,
                  "am_project_tc: pointers must be 16-byte aligned");
    return launch_project_tc(M, n_tables, 1, h, image, tables, M * 128, 128, nullptr, 0, 128, 0, nullptr, nullptr, nullptr,
                             (cudaStream_t)stream);
}

extern "C" int rl4co_linear_pack(int32_t n_out, int32_t n_in, const float *W, float *image, rl4co_stream_t stream) {
    RL4CO_REQUIRE(n_out >= 128 && n_out % 128 == 0 && n_out <= 2048 && n_in >= 128 && n_in % 128 == 0 && n_in <= 2048,
                  "linear_pack: sizes must be multiples of 128 (<= 2048), got %d x %d", n_out, n_in);
    RL4CO_REQUIRE(W && image, "linear_pack: null pointer");
    RL4CO_REQUIRE(((uintptr_t)W & 15) == 0 && ((uintptr_t)image & 15) == 0, "linear_pack: pointers must be 16-byte aligned");
    const int nb = n_out / 128;
    for (int kc = 0; kc < n_in / 128; ++kc) {       // one image set per 128-wide K chunk
        int rc = launch_pack_weights(nb, W + kc * 128, n_in, image + (size_t)kc * nb * 32768, n_in == 128, (cudaStream_t)stream);
        if (rc) return rc;
    }
    return 0;
}

static int linear_tc_impl(const char *who, int64_t M, int32_t n_out, int32_t n_in, const float *x, const float *image,
                          const float *bias, int32_t relu, const float *residual, const float *scale,
                          const float *shift, float *y, rl4co_stream_t stream) {
    RL4CO_REQUIRE(M >= 0 && n_out >= 128 && n_out % 128 == 0 && n_out <= 2048 && n_in >= 128 && n_in % 128 == 0 && n_in <= 2048,
                  "%s: sizes must be multiples of 128 (<= 2048), got M=%lld %d x %d", who, (long long)M, n_out, n_in);
    if (M == 0) return 0;
    RL4CO_REQUIRE(x && image && y, "%s: null pointer", who);
    RL4CO_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)image & 15) == 0 && ((uintptr_t)y & 15) == 0 &&
                      ((uintptr_t)residual & 15) == 0,
                  "%s: pointers must be 16-byte aligned", who);
    // the K chunks of a row tile accumulate in TMEM; bias, ReLU, residual and the affine ride the single epilogue
    return launch_project_tc(M, n_out / 128, n_in / 128, x, image, y, 128, n_out, bias, relu, n_in, 0, residual, scale,
                             shift, (cudaStream_t)stream);
}

extern "C" int rl4co_linear_tc(int64_t M, int32_t n_out, int32_t n_in, const float *x, const float *image,
                               const float *bias, int32_t relu, float *y, rl4co_stream_t stream) {
    return linear_tc_impl("linear_tc", M, n_out, n_in, x, image, bias, relu, nullptr, nullptr, nullptr, y, stream);
}

extern "C" int rl4co_linear_tc_ex(int64_t M, int32_t n_out, int32_t n_in, const float *x, const float *image,
                                  const float *bias, int32_t relu, const float *residual, const float *scale,
                                  const float *shift, float *y, rl4co_stream_t stream) {
    return linear_tc_impl("linear_tc_ex", M, n_out, n_in, x, image, bias, relu, residual, scale, shift, y, stream);
}

extern "C" int rl4co_am_project(int64_t M, int32_t n_tables, const float *h, const float *W,
                                float *tables, rl4co_stream_t stream) {
    RL4CO_REQUIRE(M >= 0 && n_tables >= 1 && n_tables <= 16, "am_project: bad sizes M=%lld n_tables=%d", (long long)M, n_tables);
    if (M == 0) return 0;
    RL4CO_REQUIRE(h && W && tables, "am_project: null pointer");
    RL4CO_REQUIRE(((uintptr_t)h & 15) == 0 && ((uintptr_t)W & 15) == 0 && ((uintptr_t)tables & 15) == 0,
                  "am_project: pointers must be 16-byte aligned");
    int64_t gx = (M + BM - 1) / BM;
    RL4CO_REQUIRE(gx < (1ll << 31), "am_project: M too large");
    dim3 grid((unsigned)gx, (unsigned)n_tables);
    am_project_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(M, h, W, tables);
    RL4CO_CHECK_LAUNCH("am_project");
    return 0;
}

// The init embedding of the encoder: a Linear with 2..8 input features and 128 outputs (nn/env_embeddings/init.py:51-64,
// 111-134; examples/slap.py:11-47).  A GEMM library runs this K = 2 product as an SGEMM tile kernel at 2.6 TB/s; it is a pure
// write stream: lane = 4 output columns (weights and bias in registers), a warp writes one 512-byte row per iteration.
namespace rl4co {
constexpr int kNarrowMaxIn = 8;
__global__ void __launch_bounds__(256) linear_narrow_kernel(int64_t M, int n_in, const float *__restrict__ x,
                                                            const float *__restrict__ W, const float *__restrict__ bias,
                                                            float *__restrict__ y) {
    const int lane = threadIdx.x & 31;
    float w[4][kNarrowMaxIn], b[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        b[c] = bias ? bias[4 * lane + c] : 0.f;
#pragma unroll
        for (int k = 0; k < kNarrowMaxIn; ++k) w[c][k] = k < n_in ? W[(size_t)(4 * lane + c) * n_in + k] : 0.f;
    }
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5), me = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    for (int64_t m = me; m < M; m += warps) {
        float xv[kNarrowMaxIn];
#pragma unroll
        for (int k = 0; k < kNarrowMaxIn; ++k) xv[k] = k < n_in ? __ldg(x + m * n_in + k) : 0.f;
        float4 o;
        float *op = &o.x;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float acc = 0.f;                        // ascending k, then the bias: the order of x @ W^T + b
#pragma unroll
            for (int k = 0; k < kNarrowMaxIn; ++k)
                if (k < n_in) acc = fmaf(xv[k], w[c][k], acc);
            op[c] = acc + b[c];
        }
        *reinterpret_cast<float4 *>(y + m * 128 + 4 * lane) = o;
    }
}
}  // namespace rl4co

extern "C" int rl4co_linear_narrow(int64_t M, int32_t n_in, const float *x, const float *W, const float *bias, float *y,
                                   rl4co_stream_t stream) {
    RL4CO_REQUIRE(M >= 0 && n_in >= 1 && n_in <= kNarrowMaxIn, "linear_narrow: n_in = %d outside 1..%d", n_in, kNarrowMaxIn);
    if (M == 0) return 0;
    RL4CO_REQUIRE(x && W && y, "linear_narrow: null pointer");
    RL4CO_REQUIRE(((uintptr_t)y & 15) == 0, "linear_narrow: y must be 16-byte aligned");
    const int64_t want = (M + 7) / 8;
    const int64_t grid = want < 148 * 8 ? want : 148 * 8;
    linear_narrow_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(M, n_in, x, W, bias, y);
    RL4CO_CHECK_LAUNCH("linear_narrow");
    return 0;
}

extern "C" int rl4co_am_graph_context(int64_t B, int32_t N, const float *h, const float *Wfixed,
                                      float *gctx, rl4co_stream_t stream) {
    RL4CO_REQUIRE(B >= 0 && N >= 1, "am_graph_context: bad sizes");
    if (B == 0) return 0;
    RL4CO_REQUIRE(h && Wfixed && gctx, "am_graph_context: null pointer");
    size_t smem = (size_t)(PE * (PE + 1) + 4 * PE + 8 * PE) * sizeof(float);
    cudaFuncSetAttribute(am_graph_context_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int64_t grid = B < 148 * 3 ? B : 148 * 3;
    am_graph_context_kernel<<<(unsigned)grid, 256, smem, (cudaStream_t)stream>>>(B, N, h, Wfixed, gctx);
    RL4CO_CHECK_LAUNCH("am_graph_context");
    return 0;
}
