// probe: redux.sync.max.f32 with sub-warp member masks (4 interleaved groups of 8 lanes)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float *out, int mode) {
    int lane = threadIdx.x & 31;
    float v = (float)((lane * 37) % 23) - (lane == 5 ? INFINITY : 0.f);
    unsigned mask = mode == 0 ? 0xffffffffu : (0x11111111u << (lane & 3));
    float r;
    asm volatile("redux.sync.max.f32 %0, %1, %2;" : "=f"(r) : "f"(v), "r"(mask));
    out[threadIdx.x] = r;
}
int main() {
    float *d; cudaMalloc(&d, 32 * 4);
    for (int mode = 0; mode < 2; ++mode) {
        k<<<1, 32>>>(d, mode);
        cudaError_t e = cudaDeviceSynchronize();
        float h[32]; cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
        printf("mode %d: %s :", mode, cudaGetErrorString(e));
        for (int i = 0; i < 32; ++i) printf(" %g", h[i]);
        printf("\n");
        // expected
        for (int i = 0; i < 32; ++i) {
            float m = -INFINITY;
            for (int j = 0; j < 32; ++j) if (mode == 0 || (j & 3) == (i & 3)) { float v = (float)((j * 37) % 23) - (j == 5 ? INFINITY : 0.f); m = v > m ? v : m; }
            if (m != h[i]) { printf("MISMATCH lane %d want %g got %g\n", i, m, h[i]); break; }
        }
        fflush(stdout);
    }
    return 0;
}
