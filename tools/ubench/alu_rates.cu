// Instruction-throughput micro-benchmark for the CUDA-core ops the scaler's inner loops can be
// built from (SURVEY.md §7.3 asks for this before choosing the tap-reduction instruction mix).
// Prints lane-operations per clock per SM for each op, full-chip, 8 independent chains per thread.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o alu_rates alu_rates.cu && ./alu_rates
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;
constexpr int CHAINS = 8;

#define DEFINE_KERNEL(NAME, DECL, BODY, SINK)                                          \
__global__ void __launch_bounds__(256) NAME(unsigned* out, unsigned seed, long long* cyc) { \
    DECL;                                                                              \
    long long t0 = clock64();                                                          \
    for (int it = 0; it < ITERS; ++it) {                                               \
        _Pragma("unroll") for (int c = 0; c < CHAINS; ++c) { BODY; }                   \
    }                                                                                  \
    long long t1 = clock64();                                                          \
    unsigned acc = 0;                                                                  \
    _Pragma("unroll") for (int c = 0; c < CHAINS; ++c) { SINK; }                       \
    if (acc == 0x12345678u) out[threadIdx.x] = acc;                                    \
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;                           \
}

DEFINE_KERNEL(k_imad, int r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; int a = seed | 1; int b = seed * 3,
              asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(r[c]) : "r"(a), "r"(b)), acc += r[c])
DEFINE_KERNEL(k_ffma, float r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; float a = 1.0001f + seed; float b = 0.5f * seed,
              asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(r[c]) : "f"(a), "f"(b)), acc += __float_as_uint(r[c]))
DEFINE_KERNEL(k_ffma_rm, float r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; float a = 1.0001f + seed; float b = 0.5f * seed,
              asm volatile("fma.rm.f32 %0, %0, %1, %2;" : "+f"(r[c]) : "f"(a), "f"(b)), acc += __float_as_uint(r[c]))
DEFINE_KERNEL(k_ffma2, unsigned long long r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = (unsigned long long)(seed + c + threadIdx.x) * 0x3f80000100000001ull; unsigned long long a = 0x3f8000013f800001ull + seed; unsigned long long b = 0x3f0000003f000000ull * seed,
              asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(r[c]) : "l"(a), "l"(b)), acc += (unsigned)r[c] + (unsigned)(r[c] >> 32))
DEFINE_KERNEL(k_fadd2, unsigned long long r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = (unsigned long long)(seed + c + threadIdx.x) * 0x3f80000100000001ull; unsigned long long a = 0x3f8000013f800001ull + seed,
              asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(r[c]) : "l"(a)), acc += (unsigned)r[c] + (unsigned)(r[c] >> 32))
DEFINE_KERNEL(k_dp2a, int r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; unsigned a = seed * 77u + 5; unsigned b = seed * 3u + threadIdx.x,
              asm volatile("dp2a.lo.s32.u32 %0, %1, %2, %0;" : "+r"(r[c]) : "r"(a), "r"(b)), acc += r[c])
DEFINE_KERNEL(k_dp4a, int r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; unsigned a = seed * 77u + 5; unsigned b = seed * 3u + threadIdx.x,
              asm volatile("dp4a.s32.u32 %0, %1, %2, %0;" : "+r"(r[c]) : "r"(a), "r"(b)), acc += r[c])
DEFINE_KERNEL(k_iadd3, int r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; int a = seed | 1,
              asm volatile("add.s32 %0, %0, %1;" : "+r"(r[c]) : "r"(a)), acc += r[c])
DEFINE_KERNEL(k_lop3, unsigned r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; unsigned a = seed | 1; unsigned b = seed * 5,
              asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[c]) : "r"(a), "r"(b)), acc += r[c])
DEFINE_KERNEL(k_shf, unsigned r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; unsigned a = seed | 1; unsigned b = (seed & 7) * 8,
              asm volatile("shf.r.clamp.b32 %0, %0, %1, %2;" : "+r"(r[c]) : "r"(a), "r"(b)), acc += r[c])
DEFINE_KERNEL(k_prmt, unsigned r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; unsigned a = seed | 1; unsigned b = 0x3210 + seed,
              asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(r[c]) : "r"(a), "r"(b)), acc += r[c])
DEFINE_KERNEL(k_fmnmx, float r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; float a = 1.0001f + seed,
              asm volatile("min.f32 %0, %0, %1;" : "+f"(r[c]) : "f"(a)), acc += __float_as_uint(r[c]))
DEFINE_KERNEL(k_imnmx, int r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; int a = 100 + seed,
              asm volatile("min.s32 %0, %0, %1;" : "+r"(r[c]) : "r"(a)), acc += r[c])
DEFINE_KERNEL(k_i2ip, int r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x; int a = 100 + seed,
              asm volatile("cvt.pack.sat.u8.s32.b32 %0, %0, %1, %0;" : "+r"(r[c]) : "r"(a)), acc += r[c])
DEFINE_KERNEL(k_i2f, int r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x,
              { float f; asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(f) : "r"(r[c])); r[c] = __float_as_int(f); }, acc += r[c])
DEFINE_KERNEL(k_f2i, int r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x,
              { int f; asm volatile("cvt.rmi.s32.f32 %0, %1;" : "=r"(f) : "f"(__int_as_float(r[c]))); r[c] = f; }, acc += r[c])
DEFINE_KERNEL(k_shr, int r[CHAINS]; for (int c = 0; c < CHAINS; ++c) r[c] = seed + c + threadIdx.x,
              asm volatile("shr.s32 %0, %0, 7;" : "+r"(r[c])), acc += r[c])
// mixes: IMAD + FFMA interleaved (do the two FMA datapaths dual-issue?), FFMA2 + IMAD
DEFINE_KERNEL(k_mix_imad_ffma, int r[CHAINS]; float q[CHAINS]; for (int c = 0; c < CHAINS; ++c) { r[c] = seed + c + threadIdx.x; q[c] = r[c]; } int a = seed | 1; int b = seed * 3; float fa = 1.0001f + seed; float fb = 0.5f * seed,
              { asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(r[c]) : "r"(a), "r"(b)); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(q[c]) : "f"(fa), "f"(fb)); }, acc += r[c] + __float_as_uint(q[c]))
DEFINE_KERNEL(k_mix_ffma2_imad, int r[CHAINS]; unsigned long long q[CHAINS]; for (int c = 0; c < CHAINS; ++c) { r[c] = seed + c + threadIdx.x; q[c] = (unsigned long long)r[c] * 0x3f80000100000001ull; } int a = seed | 1; int b = seed * 3; unsigned long long fa = 0x3f8000013f800001ull + seed; unsigned long long fb = 0x3f0000003f000000ull * seed,
              { asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(r[c]) : "r"(a), "r"(b)); asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(q[c]) : "l"(fa), "l"(fb)); }, acc += r[c] + (unsigned)q[c])
DEFINE_KERNEL(k_mix_ffma2_lop3, unsigned r[CHAINS]; unsigned long long q[CHAINS]; for (int c = 0; c < CHAINS; ++c) { r[c] = seed + c + threadIdx.x; q[c] = (unsigned long long)r[c] * 0x3f80000100000001ull; } unsigned a = seed | 1; unsigned b = seed * 3; unsigned long long fa = 0x3f8000013f800001ull + seed; unsigned long long fb = 0x3f0000003f000000ull * seed,
              { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[c]) : "r"(a), "r"(b)); asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(q[c]) : "l"(fa), "l"(fb)); }, acc += r[c] + (unsigned)q[c])
DEFINE_KERNEL(k_mix_dp2a_ffma, int r[CHAINS]; float q[CHAINS]; for (int c = 0; c < CHAINS; ++c) { r[c] = seed + c + threadIdx.x; q[c] = r[c]; } unsigned a = seed * 77u + 5; unsigned b = seed * 3u + threadIdx.x; float fa = 1.0001f + seed; float fb = 0.5f * seed,
              { asm volatile("dp2a.lo.s32.u32 %0, %1, %2, %0;" : "+r"(r[c]) : "r"(a), "r"(b)); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(q[c]) : "f"(fa), "f"(fb)); }, acc += r[c] + __float_as_uint(q[c]))

// shared-memory load rates
template <int VEC>
__global__ void __launch_bounds__(256) k_lds(unsigned* out, unsigned seed, long long* cyc) {
    __shared__ __align__(16) unsigned sm[256 * 4 + 64];
    for (int i = threadIdx.x; i < 256 * 4 + 64; i += 256) sm[i] = i * seed;
    __syncthreads();
    unsigned acc = 0;
    unsigned base = (unsigned)__cvta_generic_to_shared(sm) + threadIdx.x * 4 * VEC;
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            unsigned addr = base + ((c * 16) & 63);
            if (VEC == 1) { unsigned v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr)); acc += v; }
            if (VEC == 2) { unsigned v, w; asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v), "=r"(w) : "r"(addr & ~7u)); acc += v ^ w; }
            if (VEC == 4) { unsigned v, w, x, y; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v), "=r"(w), "=r"(x), "=r"(y) : "r"(addr & ~15u)); acc += v ^ w ^ x ^ y; }
        }
    }
    long long t1 = clock64();
    if (acc == 0x12345678u) out[threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

typedef void (*kern_t)(unsigned*, unsigned, long long*);

static int run(const char* name, kern_t k, double ops_per_chain_step, int sms) {
    unsigned* out; long long* cyc;
    CK(cudaMalloc(&out, 4096)); CK(cudaMalloc(&cyc, 8));
    int blocks = sms * 8;  // 8 x 256 threads = 2048 threads = full occupancy
    k<<<blocks, 256>>>(out, 0, cyc);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0));
    k<<<blocks, 256>>>(out, 0, cyc);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    long long c; CK(cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost));
    double total = (double)blocks * 256 * ITERS * CHAINS * ops_per_chain_step;
    // per-SM rate from the in-kernel cycle count of one CTA (all 8 CTAs on an SM run concurrently)
    double per_clk_sm = (double)8 * 256 * ITERS * CHAINS * ops_per_chain_step / (double)c;
    printf("%-18s %8.3f ms  %8.2f Tlane-op/s  %7.1f lane-op/clk/SM (clock64)  est %.0f MHz\n", name, ms, total / ms * 1e-9,
           per_clk_sm, total / (ms * 1e-3) / sms / per_clk_sm * 1e-6);
    cudaFree(out); cudaFree(cyc);
    return 0;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    printf("%s  SMs=%d  clock=%d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
    int sms = p.multiProcessorCount;
#define RUN(k, ops) if (run(#k, k, ops, sms)) return 1
    for (int rep = 0; rep < 2; ++rep) {
        RUN(k_imad, 1); RUN(k_ffma, 1); RUN(k_ffma_rm, 1); RUN(k_ffma2, 1); RUN(k_fadd2, 1); RUN(k_dp2a, 1); RUN(k_dp4a, 1);
        RUN(k_iadd3, 1); RUN(k_lop3, 1); RUN(k_shf, 1); RUN(k_prmt, 1); RUN(k_fmnmx, 1); RUN(k_imnmx, 1); RUN(k_i2ip, 1);
        RUN(k_i2f, 1); RUN(k_f2i, 1); RUN(k_shr, 1);
        RUN(k_mix_imad_ffma, 2); RUN(k_mix_ffma2_imad, 2); RUN(k_mix_ffma2_lop3, 2); RUN(k_mix_dp2a_ffma, 2);
        RUN(k_lds<1>, 1); RUN(k_lds<2>, 1); RUN(k_lds<4>, 1);
    }
    return 0;
}
