// Which instructions share issue time with the packed FP32 multiply-add on sm_100a?
// For every op X this prints the issue cost (scheduler cycles per warp instruction) of X alone, of FFMA2 alone and of
// an interleaved X + FFMA2 stream; cost(X + FFMA2) < cost(X) + cost(FFMA2) means the two overlap.  The scaler's inner
// loops are budgeted with these numbers.  Also prints dependent-issue latencies.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o issue_mix issue_mix.cu && ./issue_mix
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

constexpr int ITERS = 2048;
constexpr int CH = 8;          // independent chains per op
constexpr int THREADS = 1024;  // 8 warps per scheduler, one CTA per SM

struct Regs {
    unsigned long long q[CH];   // FFMA2 chains
    float f[CH];
    unsigned u[CH];
    unsigned l[4 * CH];         // load destinations
};

enum Op { NONE, FFMA2, FFMA, FADD, IADD, IADD3, FMNMX, FMNMX3, LOP3, PRMT, I2IP, IMAD, SHF, LDS32, LDS64, LDS128, STS32, STS128, SHFL, ISETP, LEA, I2F, MOV, FMUL2, LDSU };

template <int OP>
__device__ __forceinline__ void emit(Regs &r, int c, unsigned saddr, unsigned a, unsigned b, float fa, float fb, unsigned long long qa, unsigned long long qb) {
    if (OP == FFMA2) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(r.q[c]) : "l"(qa), "l"(qb));
    if (OP == FMUL2) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(r.q[c]) : "l"(qa));
    if (OP == FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(r.f[c]) : "f"(fa), "f"(fb));
    if (OP == FADD) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(r.f[c]) : "f"(fa));
    if (OP == IADD) asm volatile("add.s32 %0, %0, %1;" : "+r"(r.u[c]) : "r"(a));
    if (OP == IADD3) asm volatile("{.reg .u32 t; add.s32 t, %0, %1; add.s32 %0, t, %2;}" : "+r"(r.u[c]) : "r"(a), "r"(b));
    if (OP == FMNMX) asm volatile("min.f32 %0, %0, %1;" : "+f"(r.f[c]) : "f"(fa));
    if (OP == FMNMX3) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(r.f[c]) : "f"(fa), "f"(fb));
    if (OP == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r.u[c]) : "r"(a), "r"(b));
    if (OP == PRMT) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(r.u[c]) : "r"(a), "r"(b));
    if (OP == I2IP) asm volatile("cvt.pack.sat.u8.s32.b32 %0, %0, %1, %0;" : "+r"(r.u[c]) : "r"(a));
    if (OP == IMAD) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(r.u[c]) : "r"(a), "r"(b));
    if (OP == SHF) asm volatile("shf.r.clamp.b32 %0, %0, %1, %2;" : "+r"(r.u[c]) : "r"(a), "r"(b));
    if (OP == LEA) asm volatile("{.reg .u32 t; shl.b32 t, %0, 2; add.s32 %0, t, %1;}" : "+r"(r.u[c]) : "r"(a));
    if (OP == I2F) { float t; asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(t) : "r"(r.u[c])); r.u[c] = __float_as_uint(t); }
    if (OP == MOV) asm volatile("mov.b32 %0, %1;" : "=r"(r.l[c]) : "r"(r.u[(c + 1) & (CH - 1)]));
    if (OP == ISETP) asm volatile("{.reg .pred p; setp.lt.s32 p, %0, %1; @p add.s32 %0, %0, 1;}" : "+r"(r.u[c]) : "r"(a));
    if (OP == LDS32) asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(r.l[c]) : "r"(saddr + 128 * c));
    if (OP == LDSU) asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(r.l[c]) : "r"((saddr & 0xffff0000u) + 128 * c + (a & 0xf00)));   // warp-uniform address
    if (OP == LDS64) asm volatile("ld.volatile.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r.l[2 * c]), "=r"(r.l[2 * c + 1]) : "r"(saddr * 2 + 256 * c));
    if (OP == LDS128) asm volatile("ld.volatile.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.l[4 * c]), "=r"(r.l[4 * c + 1]), "=r"(r.l[4 * c + 2]), "=r"(r.l[4 * c + 3]) : "r"(saddr * 4 + 512 * c));
    if (OP == STS32) asm volatile("st.volatile.shared.u32 [%0], %1;" :: "r"(saddr + 128 * c), "r"(a) : "memory");
    if (OP == STS128) asm volatile("st.volatile.shared.v4.u32 [%0], {%1,%1,%2,%2};" :: "r"(saddr * 4 + 512 * c), "r"(a), "r"(b) : "memory");
    if (OP == SHFL) asm volatile("shfl.sync.bfly.b32 %0, %0, 1, 0x1f, 0xffffffff;" : "+r"(r.u[c]));
}

// NA copies of A then NB copies of B per chain step, CH chain steps per iteration
template <int A, int NA, int B, int NB>
__global__ void __launch_bounds__(THREADS, 1) k_mix(unsigned *out, unsigned seed, long long *cyc) {
    extern __shared__ __align__(16) unsigned sm[];
    for (int i = threadIdx.x; i < 12288; i += THREADS) sm[i] = i * seed;
    __syncthreads();
    Regs r;
#pragma unroll
    for (int c = 0; c < CH; ++c) {
        r.q[c] = (unsigned long long)(seed + c + threadIdx.x) * 0x3f80000100000001ull;
        r.f[c] = seed + c + threadIdx.x;
        r.u[c] = seed + c + threadIdx.x;
    }
#pragma unroll
    for (int c = 0; c < 4 * CH; ++c) r.l[c] = 0;
    const unsigned a = seed | 1, b = seed * 3 + 0x3210;
    const float fa = 1.0001f + seed, fb = 0.5f * seed;
    const unsigned long long qa = 0x3f8000013f800001ull + seed, qb = 0x3f0000003f000000ull * seed;
    const unsigned sbase = (unsigned)__cvta_generic_to_shared(sm);
    // lane-consecutive words (the vector loads scale the lane offset by their width): conflict free
    const unsigned lane_off = (threadIdx.x & 31) * 4;
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
        const unsigned saddr_plain = sbase + lane_off + ((it & 1) << 12);   // changes every iteration: nothing to hoist
        const unsigned saddr = saddr_plain;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
#pragma unroll
            for (int k = 0; k < (NA > NB ? NA : NB); ++k) {
                if (k < NA) emit<A>(r, c, (A == LDS64 || A == LDS128 || A == STS128) ? (lane_off + ((it & 1) << 10)) + sbase / (A == LDS64 ? 2 : 4) : saddr, a, b, fa, fb, qa, qb);
                if (k < NB) emit<B>(r, c, (B == LDS64 || B == LDS128 || B == STS128) ? (lane_off + ((it & 1) << 10)) + sbase / (B == LDS64 ? 2 : 4) : saddr, a, b, fa, fb, qa, qb);
            }
        }
    }
    long long t1 = clock64();
    unsigned acc = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) acc += (unsigned)r.q[c] + (unsigned)(r.q[c] >> 32) + __float_as_uint(r.f[c]) + r.u[c];
#pragma unroll
    for (int c = 0; c < 4 * CH; ++c) acc += r.l[c];
    if (acc == 0x12345678u) out[threadIdx.x] = acc;
    // per-warp start / end stamps of the first 8 CTAs: the host takes (last end - first start) per CTA, so a warp the
    // scheduler happens to favour does not flatter the number
    if ((threadIdx.x & 31) == 0 && blockIdx.x < 8) {
        cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2] = t0;
        cyc[(blockIdx.x * 32 + (threadIdx.x >> 5)) * 2 + 1] = t1;
    }
}

// dependent-issue latency: one warp, one chain
template <int OP>
__global__ void k_lat(unsigned *out, unsigned seed, long long *cyc) {
    __shared__ unsigned sm[1024];
    for (int i = threadIdx.x; i < 1024; i += 32) sm[i] = (unsigned)__cvta_generic_to_shared(sm) + ((i * 4 + 4 * 37) & 4095);
    __syncwarp();
    unsigned long long q = (unsigned long long)(seed + threadIdx.x) * 0x3f80000100000001ull;
    const unsigned long long qa = 0x3f8000013f800001ull + seed, qb = 0x3f0000003f000000ull * seed;
    float f = seed + threadIdx.x; const float fa = 1.0001f + seed, fb = 0.5f * seed;
    unsigned u = seed + threadIdx.x; const unsigned a = seed | 1, b = seed * 3 + 0x3210;
    unsigned p = (unsigned)__cvta_generic_to_shared(sm) + threadIdx.x * 4;
    long long t0 = clock64();
#pragma unroll 16
    for (int it = 0; it < 4096; ++it) {
        if (OP == FFMA2) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(q) : "l"(qa), "l"(qb));
        if (OP == FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(fa), "f"(fb));
        if (OP == FMNMX) asm volatile("min.f32 %0, %0, %1;" : "+f"(f) : "f"(fa));
        if (OP == IADD) asm volatile("add.s32 %0, %0, %1;" : "+r"(u) : "r"(a));
        if (OP == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u) : "r"(a), "r"(b));
        if (OP == PRMT) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(u) : "r"(a), "r"(b));
        if (OP == I2IP) asm volatile("cvt.pack.sat.u8.s32.b32 %0, %0, %1, %0;" : "+r"(u) : "r"(a));
        if (OP == IMAD) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(u) : "r"(a), "r"(b));
        if (OP == LDS32) asm volatile("ld.shared.u32 %0, [%0];" : "+r"(p));
        if (OP == SHFL) asm volatile("shfl.sync.bfly.b32 %0, %0, 1, 0x1f, 0xffffffff;" : "+r"(u));
    }
    long long t1 = clock64();
    unsigned acc = (unsigned)q + (unsigned)(q >> 32) + __float_as_uint(f) + u + p;
    if (acc == 0x12345678u) out[threadIdx.x] = acc;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

typedef void (*kern_t)(unsigned *, unsigned, long long *);

static double run(kern_t k, int sms, int n_instr) {
    static unsigned *out = nullptr; static long long *cyc = nullptr;
    if (!out) { cudaMalloc(&out, 8192); cudaMalloc(&cyc, 8 * 1024); }
    cudaFuncSetAttribute(reinterpret_cast<const void *>(k), cudaFuncAttributeMaxDynamicSharedMemorySize, 49152);
    k<<<sms, THREADS, 49152>>>(out, 0, cyc);
    k<<<sms, THREADS, 49152>>>(out, 0, cyc);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError())); return -1; }
    long long c[8 * 32 * 2];
    cudaMemcpy(c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int b = 0; b < 8; ++b) {
        long long lo = c[b * 64], hi = c[b * 64 + 1];
        for (int w = 0; w < 32; ++w) { lo = c[(b * 32 + w) * 2] < lo ? c[(b * 32 + w) * 2] : lo; hi = c[(b * 32 + w) * 2 + 1] > hi ? c[(b * 32 + w) * 2 + 1] : hi; }
        avg += (hi - lo) / 8.0;
    }
    // scheduler cycles per warp instruction: 8 warps per scheduler, each issuing ITERS * CH * n_instr
    return avg / (8.0 * ITERS * CH * n_instr);
}

#define SINGLE(X, n_sass) { double c = run(k_mix<X, 1, NONE, 0>, sms, 1); printf("%-8s alone            %6.2f cycles per PTX op (%d SASS)\n", #X, c, n_sass); cost[X] = c; }
#define PAIR(A, NA, B, NB) { double c = run(k_mix<A, NA, B, NB>, sms, 1); printf("%d x %-7s + %d x %-7s  %6.2f cycles per group   sum of parts %6.2f   overlap %5.2f\n", NA, #A, NB, #B, c, NA * cost[A] + NB * cost[B], NA * cost[A] + NB * cost[B] - c); }
#define LAT(X) { static long long *cy = nullptr; static unsigned *o = nullptr; if (!cy) { cudaMalloc(&cy, 8); cudaMalloc(&o, 4096); } k_lat<X><<<1, 32>>>(o, 0, cy); k_lat<X><<<1, 32>>>(o, 0, cy); cudaDeviceSynchronize(); long long c; cudaMemcpy(&c, cy, 8, cudaMemcpyDeviceToHost); printf("latency %-8s %6.2f cycles\n", #X, c / 4096.0); }

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    printf("%s  SMs=%d\n", p.name, p.multiProcessorCount);
    int sms = p.multiProcessorCount;
    double cost[64] = {0};
    SINGLE(FFMA2, 1) SINGLE(FMUL2, 1) SINGLE(FFMA, 1) SINGLE(FADD, 1) SINGLE(IADD, 1) SINGLE(IADD3, 1) SINGLE(FMNMX, 1) SINGLE(FMNMX3, 1) SINGLE(LOP3, 1)
    SINGLE(PRMT, 1) SINGLE(I2IP, 1) SINGLE(IMAD, 1) SINGLE(SHF, 1) SINGLE(LEA, 1) SINGLE(I2F, 1) SINGLE(MOV, 1) SINGLE(ISETP, 2)
    SINGLE(LDS32, 1) SINGLE(LDSU, 1) SINGLE(LDS64, 1) SINGLE(LDS128, 1) SINGLE(STS32, 1) SINGLE(STS128, 1) SINGLE(SHFL, 1)
    printf("\n");
    PAIR(FFMA2, 1, FFMA, 1) PAIR(FFMA2, 1, FADD, 1) PAIR(FFMA2, 1, IADD, 1) PAIR(FFMA2, 1, IADD, 2) PAIR(FFMA2, 1, IADD3, 1) PAIR(FFMA2, 1, FMNMX, 1) PAIR(FFMA2, 1, FMNMX, 2)
    PAIR(FFMA2, 1, FMNMX3, 1) PAIR(FFMA2, 1, LOP3, 1) PAIR(FFMA2, 1, PRMT, 1) PAIR(FFMA2, 1, I2IP, 1) PAIR(FFMA2, 1, IMAD, 1) PAIR(FFMA2, 1, SHF, 1) PAIR(FFMA2, 1, LEA, 1)
    PAIR(FFMA2, 1, MOV, 1) PAIR(FFMA2, 1, ISETP, 1) PAIR(FFMA2, 1, LDS32, 1) PAIR(FFMA2, 1, LDS32, 2) PAIR(FFMA2, 1, LDSU, 1) PAIR(FFMA2, 1, LDS64, 1) PAIR(FFMA2, 1, LDS128, 1) PAIR(FFMA2, 2, LDS128, 1)
    PAIR(FFMA2, 4, LDS128, 1) PAIR(FFMA2, 1, STS32, 1) PAIR(FFMA2, 1, STS128, 1) PAIR(FFMA2, 1, SHFL, 1) PAIR(FFMA2, 2, SHFL, 1) PAIR(FFMA2, 2, IADD, 1) PAIR(FFMA2, 2, LOP3, 1) PAIR(FFMA2, 3, I2IP, 1)
    printf("\n");
    PAIR(FFMA, 1, IADD, 1) PAIR(FFMA, 1, LOP3, 1) PAIR(FFMA, 1, FMNMX, 1) PAIR(FFMA, 1, LDS32, 1) PAIR(FFMA, 2, LDS32, 1) PAIR(FFMA, 1, LDS128, 1) PAIR(FFMA, 1, IMAD, 1) PAIR(FFMA, 1, I2IP, 1) PAIR(FFMA, 1, PRMT, 1)
    PAIR(FFMA, 1, SHFL, 1) PAIR(FFMA, 2, IADD, 1)
    printf("\n");
    PAIR(IADD, 1, LOP3, 1) PAIR(IADD, 1, FMNMX, 1) PAIR(LOP3, 1, PRMT, 1) PAIR(LOP3, 1, I2IP, 1) PAIR(IMAD, 1, LOP3, 1) PAIR(IMAD, 1, IADD, 1) PAIR(LDS32, 1, IADD, 1) PAIR(LDS128, 1, IADD, 1) PAIR(LDS32, 1, LOP3, 1)
    PAIR(LDS32, 1, IMAD, 1) PAIR(LDS128, 1, STS128, 1) PAIR(LDS32, 1, SHFL, 1)
    printf("\n");
    LAT(FFMA2) LAT(FFMA) LAT(FMNMX) LAT(IADD) LAT(LOP3) LAT(PRMT) LAT(I2IP) LAT(IMAD) LAT(LDS32) LAT(SHFL)
    return 0;
}
