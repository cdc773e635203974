// probe: SM-level throughputs the rollout kernel design depends on (B200, sm_100a)
//   1 FFMA (3-register) throughput per SM per clock, 16 warps / SM
//   2 fma.rn.f32x2 (packed) throughput
//   3 __syncthreads round trip with 512 and 256 threads, named barrier with 128 threads
//   4 LDS.128 broadcast bandwidth (all lanes same address) and distinct-address bandwidth
//   5 redux.sync / shfl latency
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe_sm tools/probe_sm.cu
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 4096

__global__ void __launch_bounds__(512, 1) k_ffma(float *out, long long *cyc, float a, float b) {
    float x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = threadIdx.x * 0.001f + i;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], a, b);
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// three distinct register operands per FFMA (the rollout kernel's shape: acc += k[i]*q[i])
__global__ void __launch_bounds__(512, 1) k_ffma3(float *out, long long *cyc, const float *in) {
    float k[16], q[16], acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 16; ++i) { k[i] = in[threadIdx.x + i * 512]; q[i] = in[threadIdx.x + i * 512 + 7]; }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i & 3] = fmaf(k[i], q[i], acc[i & 3]);
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc[0] + acc[1] + acc[2] + acc[3];
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

__global__ void __launch_bounds__(512, 1) k_ffma2(float *out, long long *cyc, const float *in) {
    unsigned long long k[8], q[8], acc[4] = {0ull, 0ull, 0ull, 0ull};
    const unsigned long long *in2 = reinterpret_cast<const unsigned long long *>(in);
#pragma unroll
    for (int i = 0; i < 8; ++i) { k[i] = in2[threadIdx.x + i * 512]; q[i] = in2[threadIdx.x + i * 512 + 3]; }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc[i & 3]) : "l"(k[i]), "l"(q[i]));
    }
    long long t1 = clock64();
    unsigned long long s = acc[0] ^ acc[1] ^ acc[2] ^ acc[3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = __uint_as_float((unsigned)(s ^ (s >> 32)));
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
__global__ void k_bar(float *out, long long *cyc) {
    __shared__ float s[512];
    s[threadIdx.x] = threadIdx.x;
    __syncthreads();
    float v = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) {
        if (MODE == 0) __syncthreads();
        else asm volatile("bar.sync %0, 128;" ::"r"(1 + (int)(threadIdx.x >> 7)));
        v += s[(threadIdx.x + it) & 511];
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = v;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// MODE 0: every lane the same float4 (broadcast); 1: four distinct float4 per warp (lane&3);
// 2: 32 distinct float4 per warp (conflict-free 512 B)
template <int MODE>
__global__ void __launch_bounds__(512, 1) k_lds(float *out, long long *cyc) {
    __shared__ float4 s[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) s[i] = make_float4(i, 1.f, 2.f, 3.f);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    int base = MODE == 0 ? 0 : MODE == 1 ? (lane & 3) : lane;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float4 v = s[(base + 32 * i + it) & 2047];
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y + acc.z + acc.w;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// dependent chain latencies on one warp
__global__ void k_lat(float *out, long long *cyc) {
    float v = threadIdx.x;
    long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) v = fmaf(v, 1.0001f, 0.5f);
    long long t1 = clock64();
    for (int it = 0; it < ITER; ++it) v += __shfl_xor_sync(0xffffffffu, v, 1);
    long long t2 = clock64();
    for (int it = 0; it < ITER; ++it) {
        float r;
        asm volatile("redux.sync.max.f32 %0, %1, %2;" : "=f"(r) : "f"(v), "r"(0xffffffffu));
        v = r + 1.f;
    }
    long long t3 = clock64();
    for (int it = 0; it < ITER; ++it) {
        float r;
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
        v = r;
    }
    long long t4 = clock64();
    for (int it = 0; it < ITER; ++it) v = tanhf(v) + 0.1f;
    long long t5 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; cyc[4] = t5 - t4; }
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float *out, *in;
    long long *cyc, h[256];
    cudaMalloc(&out, 148 * 512 * 4);
    cudaMalloc(&in, 64 * 512 * 8);
    cudaMemset(in, 0, 64 * 512 * 8);
    cudaMalloc(&cyc, 256 * 8);
    auto rep = [&](const char *name, double ops_per_thread_iter, int threads) {
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(h, cyc, 8 * 8, cudaMemcpyDeviceToHost);
        double c = (double)h[0] / ITER;
        printf("%-28s %s cycles/iter %.2f  per-SM ops/clk %.2f\n", name, cudaGetErrorString(e), c,
               ops_per_thread_iter * threads / c);
    };
    for (int r = 0; r < 2; ++r) {
        k_ffma<<<sms, 512>>>(out, cyc, 1.0001f, 0.5f); rep("ffma same-reg 16w", 16, 512);
        k_ffma3<<<sms, 512>>>(out, cyc, in); rep("ffma 3-reg 16w", 16, 512);
        k_ffma3<<<sms, 256>>>(out, cyc, in); rep("ffma 3-reg 8w", 16, 256);
        k_ffma2<<<sms, 512>>>(out, cyc, in); rep("ffma2 (f32x2) 16w [fma]", 16, 512);
        k_ffma2<<<sms, 256>>>(out, cyc, in); rep("ffma2 (f32x2) 8w [fma]", 16, 256);
        k_bar<0><<<sms, 512>>>(out, cyc); rep("__syncthreads 512 [1 op]", 1, 1);
        k_bar<0><<<sms, 256>>>(out, cyc); rep("__syncthreads 256 [1 op]", 1, 1);
        k_bar<0><<<sms, 128>>>(out, cyc); rep("__syncthreads 128 [1 op]", 1, 1);
        k_bar<1><<<sms, 512>>>(out, cyc); rep("bar.sync id,128 x4 groups", 1, 1);
        k_lds<0><<<sms, 512>>>(out, cyc); rep("lds.128 broadcast [B/clk]", 8 * 16, 512);
        k_lds<1><<<sms, 512>>>(out, cyc); rep("lds.128 4 addr [B/clk]", 8 * 16, 512);
        k_lds<2><<<sms, 512>>>(out, cyc); rep("lds.128 32 addr [B/clk]", 8 * 16, 512);
    }
    k_lat<<<1, 32>>>(out, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(h, cyc, 8 * 8, cudaMemcpyDeviceToHost);
    printf("latency (cycles): ffma %.1f shfl+add %.1f redux+add %.1f ex2 %.1f tanhf+add %.1f\n", (double)h[0] / ITER,
           (double)h[1] / ITER, (double)h[2] / ITER, (double)h[3] / ITER, (double)h[4] / ITER);
    return 0;
}
