// Device helpers shared by the fused tile kernel (kernels_fused_impl.cuh) and the streaming down-scale kernel
// (kernels_stream.cu): packed FP32 arithmetic, the "integer as a subnormal bit pattern" conversions, output rules
//  and sample stores.  Everything lives in an anonymous namespace: each translation unit gets
// its own copy.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace tvs {
namespace {

enum { kSrcU8 = 0, kSrcU8Pair = 1, kSrcU16 = 2, kSrcU16Pair = 3 };
enum { kOutU8 = 0, kOutU8Dither = 1, kOutU10 = 2 };   // how the vertical sum becomes a stored sample (A.3 step 2)

// libswscale's 8x8 ordered-dither matrix for > 8-bit sources going to 8 bits (ff_dither_8x8_128)
__constant__ int kDither128i[8][8] = {
    {36, 68, 60, 92, 34, 66, 58, 90},  {100, 4, 124, 28, 98, 2, 122, 26},
    {52, 84, 44, 76, 50, 82, 42, 74},  {116, 20, 108, 12, 114, 18, 106, 10},
    {32, 64, 56, 88, 38, 70, 62, 94},  {96, 0, 120, 24, 102, 6, 126, 30},
    {48, 80, 40, 72, 54, 86, 46, 78},  {112, 16, 104, 8, 118, 22, 110, 14},
};


typedef unsigned long long f32x2;   // two floats in one 64-bit register pair (sm_100 packed FP32 ops)

__device__ __forceinline__ float fma_rm(float a, float b, float c) {
    float d;
    asm("fma.rm.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
// FFMA2: one issue slot, two fused multiply-adds (same FMA-pipe time as two FFMAs)
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
// (sat_u8(hi) << 8 | sat_u8(lo)) | upper << 16
__device__ __forceinline__ unsigned pack_sat_u8(int hi, int lo, unsigned upper) {
    unsigned d;
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi), "r"(lo), "r"(upper));
    return d;
}
__device__ __forceinline__ float byte_to_float(unsigned word, unsigned selector) {
    // selector 0x754k moves byte k of `word` under the exponent of 2^23: 0x4B0000bb is 2^23 + bb
    return __uint_as_float(__byte_perm(word, 0x4B000000u, selector)) - 8388608.0f;
}
__device__ __forceinline__ int subnormal_to_int(float v) {
    const int b = __float_as_int(v);
    return b < 0 ? -(b & 0x7fffffff) : b;
}

// Four adjacent columns of one intermediate row, as two packed pairs.
struct Quad { f32x2 lo, hi; };
__device__ __forceinline__ Quad load_quad(const float *p) {
    const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(p);
    return Quad{v.x, v.y};
}
__device__ __forceinline__ float quad_elem(const Quad &q, int k) {
    float a, b;
    unpack2(k < 2 ? q.lo : q.hi, a, b);
    return (k & 1) ? b : a;
}


// Constants of one output rule.  The float vertical sum is acc * 2^-42; the stored sample is
// clip((acc + bias) >> kShift).  q = RN(acc' * kQ) is round(acc / 2^kShift) * 2^-149, r = acc' - q * kBack the
// exact residual (scaled by 2^-42), and |r| close to 2^(kShift-1) * 2^-42 marks a sample to redo exactly.
template <int OUT> struct OutRule;
template <> struct OutRule<kOutU8>       { static constexpr int kShift = 19, kMax = 255; };
template <> struct OutRule<kOutU8Dither> { static constexpr int kShift = 19, kMax = 255; };
template <> struct OutRule<kOutU10>      { static constexpr int kShift = 17, kMax = 1023; };


// Quotient, residual and packing of one quad.
template <int OUT>
__device__ __forceinline__ float finish_quad(f32x2 a01, f32x2 a23, int dst_shift, unsigned (&word)[2], float (&r)[4]) {
    constexpr float kQ = OutRule<OUT>::kShift == 19 ? 0x1p-126f : 0x1p-124f;
    constexpr float kBack = OutRule<OUT>::kShift == 19 ? -0x1p126f : -0x1p124f;
    const f32x2 scale = pack2(kQ, kQ), back = pack2(kBack, kBack);
    const f32x2 q01 = fma2(a01, scale, 0ull), q23 = fma2(a23, scale, 0ull);
    const f32x2 r01 = fma2(q01, back, a01), r23 = fma2(q23, back, a23);
    float q0, q1, q2, q3;
    unpack2(q01, q0, q1); unpack2(q23, q2, q3);
    unpack2(r01, r[0], r[1]); unpack2(r23, r[2], r[3]);
    if (OUT == kOutU10) {
        // clamp the subnormal integer patterns to 0..1023, then pack two 16-bit samples per word
        const float top = __int_as_float(1023);
        const unsigned b0 = __float_as_uint(fminf(fmaxf(q0, 0.f), top)) << dst_shift;
        const unsigned b1 = __float_as_uint(fminf(fmaxf(q1, 0.f), top)) << dst_shift;
        const unsigned b2 = __float_as_uint(fminf(fmaxf(q2, 0.f), top)) << dst_shift;
        const unsigned b3 = __float_as_uint(fminf(fmaxf(q3, 0.f), top)) << dst_shift;
        word[0] = __byte_perm(b0, b1, 0x5410);
        word[1] = __byte_perm(b2, b3, 0x5410);
    } else {
        word[0] = pack_sat_u8(__float_as_int(q3), __float_as_int(q2), 0u);
        word[0] = pack_sat_u8(__float_as_int(q1), __float_as_int(q0), word[0]);
        word[1] = 0;
    }
    return fmaxf(fmaxf(fabsf(r[0]), fabsf(r[1])), fmaxf(fabsf(r[2]), fabsf(r[3])));
}


// Bounds-checked element stores of one quad: `xstep` bytes between samples, 1- or 2-byte samples.
template <int OUT>
__device__ __forceinline__ void store_quad_slow(uint8_t *row, int x, int dw, int xstep, const unsigned (&word)[2]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (x + k >= dw) continue;
        uint8_t *p = row + static_cast<ptrdiff_t>(x + k) * xstep;
        if (OUT == kOutU10) *reinterpret_cast<uint16_t *>(p) = static_cast<uint16_t>(word[k >> 1] >> (16 * (k & 1)));
        else *p = static_cast<uint8_t>(word[0] >> (8 * k));
    }
}
// One vector store per quad (4 bytes, or 8 for 16-bit samples); the caller has checked alignment and bounds.
template <int OUT>
__device__ __forceinline__ void store_quad_fast(uint8_t *row, int x, const unsigned (&word)[2]) {
    if (OUT == kOutU10) *reinterpret_cast<uint2 *>(row + 2 * x) = make_uint2(word[0], word[1]);
    else *reinterpret_cast<unsigned *>(row + x) = word[0];
}


__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(static_cast<unsigned>(__cvta_generic_to_shared(smem_dst))), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<unsigned>(__cvta_generic_to_shared(smem_dst))), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }


}  // namespace
}  // namespace tvs
