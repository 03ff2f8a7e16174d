// Dependent-chain latency of FFMA2 / FFMA / IMAD on sm_100a: one warp per SM sub-partition, one chain.
#include <cstdio>
#include <cuda_runtime.h>
template <int mode>
__global__ void lat(long long* out, unsigned long long seed) {
    unsigned long long a = seed + threadIdx.x, b = 0x3f8000013f800001ull, c = 0x3f0000003f000000ull;
    float fa = 1.0f + threadIdx.x, fb = 1.0001f, fc = 0.5f;
    int ia = threadIdx.x, ib = 3, ic = 7;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 1024; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (mode == 0) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a) : "l"(b), "l"(c));
            if (mode == 1) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(fa) : "f"(fb), "f"(fc));
            if (mode == 2) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(ia) : "r"(ib), "r"(ic));
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) out[mode] = t1 - t0;
    if (a == 1 && fa == 2 && ia == 3) out[3] = 1;
}
int main() {
    long long* d; cudaMalloc(&d, 64);
    lat<0><<<1, 32>>>(d, 1); lat<1><<<1, 32>>>(d, 1); lat<2><<<1, 32>>>(d, 1); cudaDeviceSynchronize();
    long long h[4]; cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
    const char* n[3] = {"FFMA2", "FFMA", "IMAD"};
    for (int m = 0; m < 3; ++m) printf("%s dependent latency: %.2f cycles\n", n[m], h[m] / (1024.0 * 16));
    return 0;
}
