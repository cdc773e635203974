// probe: register-resident mat-vec inner loops (the rollout kernel's shape), 8 warps / SM.
// K[7 nodes][8 dims] fixed in registers; per iteration a fresh q[8] comes from shared memory.
//   A  FFMA,  node-major order  (acc_n += K[n][i]*q[i], i inner)   - no operand shared by neighbours
//   B  FFMA,  dim-major order   (i outer, n inner)                 - q[i] shared by 7 consecutive FFMAs
//   C  FFMA2, dim-pair-major    (q pair shared by 7 consecutive FFMA2)
//   D  FFMA2, node-major
//   E  FFMA  V-pattern: acc_d += V[n][d]*p[n]  (n outer, d inner; p[n] shared by 8 consecutive)
//   F  FFMA2 V-pattern: acc pair += V pair * (p,p)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe_fma tools/probe_fma.cu
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 2048
typedef unsigned long long u64;

__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ u64 pack(float lo, float hi) {
    u64 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
    return d;
}
__device__ __forceinline__ float2 unpack(u64 v) {
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(float *out, long long *cyc, const float *in) {
    __shared__ __align__(16) float sq[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sq[i] = in[i] + 1e-3f * i;
    float K[7][8];
#pragma unroll
    for (int n = 0; n < 7; ++n)
#pragma unroll
        for (int i = 0; i < 8; ++i) K[n][i] = in[threadIdx.x + 256 * (n * 8 + i)] + n + i;
    u64 K2[7][4];
#pragma unroll
    for (int n = 0; n < 7; ++n)
#pragma unroll
        for (int i = 0; i < 4; ++i) K2[n][i] = pack(K[n][2 * i], K[n][2 * i + 1]);
    __syncthreads();
    float sink = 0.f;
    const int half = (threadIdx.x >> 4) & 1;
    long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) {
        const float4 *qp = reinterpret_cast<const float4 *>(sq + ((it * 16 + half * 8) & 4095));
        const float4 q0 = qp[0], q1 = qp[1];
        const float q[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
        if (MODE == 0) {
            float acc[7];
#pragma unroll
            for (int n = 0; n < 7; ++n) {
                float a0 = 0.f, a1 = 0.f;
#pragma unroll
                for (int i = 0; i < 8; i += 2) { a0 = fmaf(K[n][i], q[i], a0); a1 = fmaf(K[n][i + 1], q[i + 1], a1); }
                acc[n] = a0 + a1;
            }
#pragma unroll
            for (int n = 0; n < 7; ++n) sink += acc[n];
        } else if (MODE == 1) {
            float acc[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int n = 0; n < 7; ++n) acc[n] = fmaf(K[n][i], q[i], acc[n]);
#pragma unroll
            for (int n = 0; n < 7; ++n) sink += acc[n];
        } else if (MODE == 2 || MODE == 3) {
            u64 q2[4] = {pack(q[0], q[1]), pack(q[2], q[3]), pack(q[4], q[5]), pack(q[6], q[7])};
            u64 acc[7] = {0ull, 0ull, 0ull, 0ull, 0ull, 0ull, 0ull};
            if (MODE == 2) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int n = 0; n < 7; ++n) acc[n] = ffma2(K2[n][i], q2[i], acc[n]);
            } else {
#pragma unroll
                for (int n = 0; n < 7; ++n)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[n] = ffma2(K2[n][i], q2[i], acc[n]);
            }
#pragma unroll
            for (int n = 0; n < 7; ++n) { float2 r = unpack(acc[n]); sink += r.x + r.y; }
        } else if (MODE == 4) {
            float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int n = 0; n < 7; ++n)
#pragma unroll
                for (int d = 0; d < 8; ++d) acc[d] = fmaf(K[n][d], q[n], acc[d]);
#pragma unroll
            for (int d = 0; d < 8; ++d) sink += acc[d];
        } else {
            u64 acc[4] = {0ull, 0ull, 0ull, 0ull};
#pragma unroll
            for (int n = 0; n < 7; ++n) {
                const u64 pp = pack(q[n], q[n]);
#pragma unroll
                for (int d = 0; d < 4; ++d) acc[d] = ffma2(K2[n][d], pp, acc[d]);
            }
#pragma unroll
            for (int d = 0; d < 4; ++d) { float2 r = unpack(acc[d]); sink += r.x + r.y; }
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = sink;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float *out, *in;
    long long *cyc, h[8];
    cudaMalloc(&out, 148 * 256 * 4);
    cudaMalloc(&in, 256 * 64 * 4 + 4096 * 4);
    cudaMemset(in, 0, 256 * 64 * 4 + 4096 * 4);
    cudaMalloc(&cyc, 256 * 8);
    auto rep = [&](const char *name) {
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
        double c = (double)h[0] / ITER;
        printf("%-34s %s cycles/iter %.1f  (56 MAC/thread: FMA/clk/SM %.1f)\n", name, cudaGetErrorString(e), c,
               56.0 * 256 / c);
    };
    for (int r = 0; r < 2; ++r) {
        k<0><<<sms, 256>>>(out, cyc, in); rep("A ffma  node-major");
        k<1><<<sms, 256>>>(out, cyc, in); rep("B ffma  dim-major (q shared)");
        k<2><<<sms, 256>>>(out, cyc, in); rep("C ffma2 dim-major (q shared)");
        k<3><<<sms, 256>>>(out, cyc, in); rep("D ffma2 node-major");
        k<4><<<sms, 256>>>(out, cyc, in); rep("E ffma  V-pattern (p shared)");
        k<5><<<sms, 256>>>(out, cyc, in); rep("F ffma2 V-pattern (p shared)");
    }
    return 0;
}
