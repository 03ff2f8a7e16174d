// Two-pass reference-shaped kernels: a horizontal pass that writes libswscale's 15-bit intermediate to
// device scratch, a vertical pass that reads it back, and the same-size repack kernels.
//
// This path handles every shape, format and tap count (up to 256 taps) and is what the fused tile
// kernel (kernels_fused.cu) falls back to when a tile would not fit in shared memory.  It follows
// SURVEY.md A.3 step by step: hScale8To15_c / hScale16To15_c, then yuv2planeX_8_c / yuv2planeX_10_c /
// yuv2nv12cX_c, and A.4 for the unscaled converters (libswscale swscale_unscaled.c planarCopyWrapper).
#include <algorithm>

#include "tvs_internal.h"

namespace tvs {
namespace {

__constant__ uint8_t kDither128[8][8] = {
    {36, 68, 60, 92, 34, 66, 58, 90},  {100, 4, 124, 28, 98, 2, 122, 26},
    {52, 84, 44, 76, 50, 82, 42, 74},  {116, 20, 108, 12, 114, 18, 106, 10},
    {32, 64, 56, 88, 38, 70, 62, 94},  {96, 0, 120, 24, 102, 6, 126, 30},
    {48, 80, 40, 72, 54, 86, 46, 78},  {112, 16, 104, 8, 118, 22, 110, 14},
};

template <typename T>
__device__ __forceinline__ int load_sample(const uint8_t *p) {
    return static_cast<int>(*reinterpret_cast<const T *>(p));
}

// One thread per intermediate sample: mid[frame][row][x] = min((sum_j src[pos[x]+j] * coef[x][j]) >> hshift, 32767).
template <typename SrcT>
__global__ void __launch_bounds__(256)
hpass_generic_kernel(const uint8_t *__restrict__ src, ptrdiff_t pitch, ptrdiff_t frame_stride, int xstep, int src_shift,
                     int sw, int dw, const int16_t *__restrict__ coef, const int32_t *__restrict__ pos, int taps,
                     int hshift, int16_t *__restrict__ mid, size_t mid_frame_stride, int sh, int rc_coeff, int rc_offset) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = blockIdx.y;
    const int f = blockIdx.z;
    if (x >= dw) return;
    const int16_t *c = coef + static_cast<size_t>(x) * taps;
    const int first = pos[x];
    const uint8_t *p = src + f * frame_stride + r * pitch + static_cast<ptrdiff_t>(first) * xstep;
    const int n = min(taps, sw - first);   // taps past the row's end carry zero weight
    int acc = 0;
#pragma unroll 4
    for (int j = 0; j < n; ++j, p += xstep) acc += (load_sample<SrcT>(p) >> src_shift) * c[j];
    acc = min(acc >> hshift, 32767);
    if (rc_coeff) acc = min((acc * rc_coeff + rc_offset) >> 14, 32767);   // full <-> limited range (lumConvertRange / chrConvertRange)
    mid[f * mid_frame_stride + static_cast<size_t>(r) * dw + x] = static_cast<int16_t>(acc);
}

// One thread per output sample.
template <typename DstT>
__global__ void __launch_bounds__(256)
vpass_generic_kernel(const int16_t *__restrict__ mid, size_t mid_frame_stride, int sh, int dw,
                     const int16_t *__restrict__ coef, const int32_t *__restrict__ pos, int taps,
                     uint8_t *__restrict__ dst, ptrdiff_t pitch, ptrdiff_t frame_stride, int xstep, int dst_shift,
                     int rule, int dither_xoff) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    const int f = blockIdx.z;
    if (x >= dw) return;
    const int16_t *m = mid + f * mid_frame_stride + x;
    const int16_t *c = coef + static_cast<size_t>(y) * taps;
    const int first = pos[y];
    int acc = 0;
    for (int j = 0; j < taps; ++j) {
        const int sr = first + j;
        if (sr < sh) acc += static_cast<int>(m[static_cast<size_t>(sr) * dw]) * c[j];
    }
    int v;
    if (rule == kOut10) {
        v = min(max((acc + (1 << 16)) >> 17, 0), 1023) << dst_shift;
    } else {
        const int d = rule == kOut8 ? 64 : kDither128[y & 7][(x + dither_xoff) & 7];
        v = min(max((acc + (d << 12)) >> 19, 0), 255);
    }
    *reinterpret_cast<DstT *>(dst + f * frame_stride + y * pitch + static_cast<ptrdiff_t>(x) * xstep) = static_cast<DstT>(v);
}

// Four adjacent output samples per thread: one 64-bit load of the intermediate and one coefficient load per tap serve
// four multiply-adds.  The launcher uses it when rows of the intermediate are 8-byte aligned and dw is a multiple of 4.
template <typename DstT>
__global__ void __launch_bounds__(256)
vpass_quad_kernel(const int16_t *__restrict__ mid, size_t mid_frame_stride, int sh, int dw,
                  const int16_t *__restrict__ coef, const int32_t *__restrict__ pos, int taps,
                  uint8_t *__restrict__ dst, ptrdiff_t pitch, ptrdiff_t frame_stride, int xstep, int dst_shift,
                  int rule, int dither_xoff) {
    const int x = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int y = blockIdx.y;
    const int f = blockIdx.z;
    if (x >= dw) return;
    const int16_t *c = coef + static_cast<size_t>(y) * taps;
    const int first = pos[y];
    const int n = min(taps, sh - first);
    const uint2 *m = reinterpret_cast<const uint2 *>(mid + f * mid_frame_stride + static_cast<size_t>(first) * dw + x);
    const int row_step = dw >> 2;
    int a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll 4
    for (int j = 0; j < n; ++j, m += row_step) {
        const uint2 v = __ldg(m);
        const int cj = c[j];
        a0 += static_cast<int>(static_cast<short>(v.x & 0xffffu)) * cj;
        a1 += (static_cast<int>(v.x) >> 16) * cj;
        a2 += static_cast<int>(static_cast<short>(v.y & 0xffffu)) * cj;
        a3 += (static_cast<int>(v.y) >> 16) * cj;
    }
    const int acc[4] = {a0, a1, a2, a3};
    uint8_t *out = dst + f * frame_stride + y * pitch + static_cast<ptrdiff_t>(x) * xstep;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int v;
        if (rule == kOut10) {
            v = min(max((acc[k] + (1 << 16)) >> 17, 0), 1023) << dst_shift;
        } else {
            const int d = rule == kOut8 ? 64 : kDither128[y & 7][(x + k + dither_xoff) & 7];
            v = min(max((acc[k] + (d << 12)) >> 19, 0), 255);
        }
        *reinterpret_cast<DstT *>(out + static_cast<ptrdiff_t>(k) * xstep) = static_cast<DstT>(v);
    }
}

// Same-size converters.  kOpCopy moves a sample unchanged (dropping / restoring P010's padding bits);
// kOpDither2 is libswscale's 10->8 bit copy: t = (p + d[y&1][x&7]) >> 2, out = t - (t >> 8), with
// d = {1,2,1,2,..} on even rows and {3,0,3,0,..} on odd rows.  For P010 sources libswscale's scalar
// tail loop (columns >= cols & ~7) forgets to drop the padding bits; `raw_tail_from` reproduces that.
template <typename SrcT, typename DstT>
__global__ void __launch_bounds__(256)
repack_kernel(const uint8_t *__restrict__ src, ptrdiff_t spitch, ptrdiff_t sframe, int sxstep, int src_shift,
              uint8_t *__restrict__ dst, ptrdiff_t dpitch, ptrdiff_t dframe, int dxstep, int dst_shift,
              int w, int kind, int raw_tail_from, int dither_xoff) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    const int f = blockIdx.z;
    if (x >= w) return;
    unsigned v = load_sample<SrcT>(src + f * sframe + y * spitch + static_cast<ptrdiff_t>(x) * sxstep);
    if (kind == kOpDither2) {
        if (raw_tail_from < 0 || x < raw_tail_from) v >>= src_shift;
        const unsigned d = (y & 1) ? ((x & 1) ? 0u : 3u) : ((x & 1) ? 2u : 1u);
        const unsigned t = (v + d) >> 2;
        v = (t - (t >> 8)) & 0xffu;
    } else if (kind == kOpDither128) {
        const unsigned t = min((v >> src_shift) << 5, 32767u);
        v = min((t + kDither128[y & 7][(x + dither_xoff) & 7]) >> 7, 255u);
    } else {
        v = (v >> src_shift) << dst_shift;
    }
    *reinterpret_cast<DstT *>(dst + f * dframe + y * dpitch + static_cast<ptrdiff_t>(x) * dxstep) = static_cast<DstT>(v);
}

// ---- vector forms of kOpDither2 (same-size 10 -> 8 bit, e.g. a yuv420p10le / p010le source going to nv12 for the
// hardware encoder) -------------------------------------------------------------------------------------------------
// One sample, exactly as repack_kernel does it.
__device__ __forceinline__ unsigned dither2_sample(unsigned v, int x, int y) {
    const unsigned d = (y & 1) ? ((x & 1) ? 0u : 3u) : ((x & 1) ? 2u : 1u);
    const unsigned t = (v + d) >> 2;
    return (t - (t >> 8)) & 0xffu;
}
// Two 16-bit samples per word (even column in the low half).  Exact for any 16-bit input: (v + d) >> 2 is taken as
// (v >> 2) + (((v & 3) + d) >> 2), which cannot carry from one half into the other.
__device__ __forceinline__ unsigned dither2_pair(unsigned word, int shift, unsigned d2) {
    const unsigned v2 = (word >> shift) & (0xffffu >> shift) * 0x00010001u;
    const unsigned carry = (((v2 & 0x00030003u) + d2) >> 2) & 0x00010001u;
    const unsigned t2 = ((v2 >> 2) & 0x3fff3fffu) + carry;
    return (t2 - ((t2 >> 8) & 0x00ff00ffu)) & 0x00ff00ffu;   // one result byte in each half
}
// Eight consecutive samples (16 bytes, the first on an even column) -> eight bytes.
__device__ __forceinline__ uint2 dither2_oct(uint4 in, int shift, unsigned d2) {
    const unsigned a = dither2_pair(in.x, shift, d2), b = dither2_pair(in.y, shift, d2);
    const unsigned c = dither2_pair(in.z, shift, d2), d = dither2_pair(in.w, shift, d2);
    return make_uint2(__byte_perm(a, b, 0x6420), __byte_perm(c, d, 0x6420));
}

// A plane of back-to-back 16-bit samples -> a plane of back-to-back bytes, 8 samples per thread.  `raw_tail_from`
// (a multiple of 8, or -1) keeps libswscale's P010 tail quirk: columns from there on skip the shift.
__global__ void __launch_bounds__(256)
dither2_rows_kernel(const uint8_t *__restrict__ src, ptrdiff_t spitch, ptrdiff_t sframe, uint8_t *__restrict__ dst, ptrdiff_t dpitch,
                    ptrdiff_t dframe, int w, int rows, int src_shift, int raw_tail_from) {
    const int slots = (w + 7) >> 3;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = idx / slots, x = (idx - y * slots) * 8;
    if (y >= rows) return;
    const int f = blockIdx.z;
    const uint8_t *s = src + f * sframe + y * spitch + 2 * static_cast<ptrdiff_t>(x);
    uint8_t *d = dst + f * dframe + y * dpitch + x;
    const int shift = (raw_tail_from >= 0 && x >= raw_tail_from) ? 0 : src_shift;
    if (x + 8 <= w) {
        const unsigned d2 = (y & 1) ? 0x00000003u : 0x00020001u;
        *reinterpret_cast<uint2 *>(d) = dither2_oct(__ldg(reinterpret_cast<const uint4 *>(s)), shift, d2);
    } else {
        for (int k = 0; x + k < w; ++k) d[k] = static_cast<uint8_t>(dither2_sample(reinterpret_cast<const uint16_t *>(s)[k] >> shift, x + k, y));
    }
}

// Two planar 16-bit chroma planes -> one interleaved byte plane, 8 sample pairs per thread.
__global__ void __launch_bounds__(256)
dither2_interleave_kernel(const uint8_t *__restrict__ u, const uint8_t *__restrict__ v, ptrdiff_t spitch_u, ptrdiff_t spitch_v,
                          ptrdiff_t sframe_u, ptrdiff_t sframe_v, uint8_t *__restrict__ dst, ptrdiff_t dpitch, ptrdiff_t dframe, int w,
                          int rows, int src_shift) {
    const int slots = (w + 7) >> 3;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = idx / slots, x = (idx - y * slots) * 8;
    if (y >= rows) return;
    const int f = blockIdx.z;
    const uint8_t *us = u + f * sframe_u + y * spitch_u + 2 * static_cast<ptrdiff_t>(x);
    const uint8_t *vs = v + f * sframe_v + y * spitch_v + 2 * static_cast<ptrdiff_t>(x);
    uint8_t *d = dst + f * dframe + y * dpitch + 2 * x;
    if (x + 8 <= w) {
        const unsigned d2 = (y & 1) ? 0x00000003u : 0x00020001u;
        const uint2 a = dither2_oct(__ldg(reinterpret_cast<const uint4 *>(us)), src_shift, d2);
        const uint2 b = dither2_oct(__ldg(reinterpret_cast<const uint4 *>(vs)), src_shift, d2);
        uint4 o;
        o.x = __byte_perm(a.x, b.x, 0x5140); o.y = __byte_perm(a.x, b.x, 0x7362);
        o.z = __byte_perm(a.y, b.y, 0x5140); o.w = __byte_perm(a.y, b.y, 0x7362);
        *reinterpret_cast<uint4 *>(d) = o;
    } else {
        for (int k = 0; x + k < w; ++k) {
            d[2 * k] = static_cast<uint8_t>(dither2_sample(reinterpret_cast<const uint16_t *>(us)[k] >> src_shift, x + k, y));
            d[2 * k + 1] = static_cast<uint8_t>(dither2_sample(reinterpret_cast<const uint16_t *>(vs)[k] >> src_shift, x + k, y));
        }
    }
}

// kOpDither128 for the chroma of a planar 10-bit source going to NV12: 8 U + 8 V samples -> 16 interleaved bytes.
// Rows of the 8x8 dither matrix packed one byte per column.
__device__ const unsigned kDither128Packed[8][2] = {
    {0x5c3c4424u, 0x5a3a4222u}, {0x1c7c0464u, 0x1a7a0262u}, {0x4c2c5434u, 0x4a2a5232u}, {0x0c6c1474u, 0x0a6a1272u},
    {0x58384020u, 0x5e3e4626u}, {0x18780060u, 0x1e7e0666u}, {0x48285030u, 0x4e2e5636u}, {0x08681070u, 0x0e6e1676u},
};
__device__ __forceinline__ unsigned dither128_sample(unsigned v, unsigned d) {
    return min((min(v << 5, 32767u) + d) >> 7, 255u);
}
// eight 16-bit samples starting on a column that is a multiple of 8; `rot` = dither_xoff (0 or 3)
__device__ __forceinline__ uint2 dither128_oct(uint4 in, int shift, unsigned d_lo, unsigned d_hi, int rot) {
    const unsigned w[4] = {in.x, in.y, in.z, in.w};
    const unsigned long long d = (static_cast<unsigned long long>(d_hi) << 32) | d_lo;
    const unsigned long long dr = rot ? ((d >> (8 * rot)) | (d << (64 - 8 * rot))) : d;   // column k uses entry (k + rot) & 7
    unsigned out[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const unsigned v = ((w[k >> 1] >> (16 * (k & 1))) & 0xffffu) >> shift;
        out[k] = dither128_sample(v, static_cast<unsigned>(dr >> (8 * k)) & 0xffu);
    }
    return make_uint2(out[0] | (out[1] << 8) | (out[2] << 16) | (out[3] << 24), out[4] | (out[5] << 8) | (out[6] << 16) | (out[7] << 24));
}
__global__ void __launch_bounds__(256)
dither128_interleave_kernel(const uint8_t *__restrict__ u, const uint8_t *__restrict__ v, ptrdiff_t spitch_u, ptrdiff_t spitch_v,
                            ptrdiff_t sframe_u, ptrdiff_t sframe_v, uint8_t *__restrict__ dst, ptrdiff_t dpitch, ptrdiff_t dframe, int w,
                            int rows, int src_shift, int xoff_u, int xoff_v) {
    const int slots = (w + 7) >> 3;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = idx / slots, x = (idx - y * slots) * 8;
    if (y >= rows) return;
    const int f = blockIdx.z;
    const uint8_t *us = u + f * sframe_u + y * spitch_u + 2 * static_cast<ptrdiff_t>(x);
    const uint8_t *vs = v + f * sframe_v + y * spitch_v + 2 * static_cast<ptrdiff_t>(x);
    uint8_t *d = dst + f * dframe + y * dpitch + 2 * x;
    if (x + 8 <= w) {
        const unsigned d_lo = kDither128Packed[y & 7][0], d_hi = kDither128Packed[y & 7][1];
        const uint2 a = dither128_oct(__ldg(reinterpret_cast<const uint4 *>(us)), src_shift, d_lo, d_hi, xoff_u);
        const uint2 b = dither128_oct(__ldg(reinterpret_cast<const uint4 *>(vs)), src_shift, d_lo, d_hi, xoff_v);
        uint4 o;
        o.x = __byte_perm(a.x, b.x, 0x5140); o.y = __byte_perm(a.x, b.x, 0x7362);
        o.z = __byte_perm(a.y, b.y, 0x5140); o.w = __byte_perm(a.y, b.y, 0x7362);
        *reinterpret_cast<uint4 *>(d) = o;
    } else {
        for (int k = 0; x + k < w; ++k) {
            d[2 * k] = static_cast<uint8_t>(dither128_sample(reinterpret_cast<const uint16_t *>(us)[k] >> src_shift, kDither128[y & 7][(x + k + xoff_u) & 7]));
            d[2 * k + 1] = static_cast<uint8_t>(dither128_sample(reinterpret_cast<const uint16_t *>(vs)[k] >> src_shift, kDither128[y & 7][(x + k + xoff_v) & 7]));
        }
    }
}

// ---- input stage for packed RGB sources (tools/upscale_benchmark.py feeds rgb24 frames) ----------
// libswscale converts each RGB line to 16-bit Y / U / V lines before the scaler proper (input.c: rgb24ToY_c, rgb24ToUV_c,
// rgb24ToUV_half_c; RGB2YUV_SHIFT = 15, default BT.601 limited-range table).  One thread per pixel pair.
__global__ void __launch_bounds__(256)
rgb_input_kernel(const uint8_t *__restrict__ rgb, ptrdiff_t pitch, ptrdiff_t frame_stride, uint8_t *__restrict__ planes,
                 size_t planes_frame_stride, size_t u_offset, size_t c_plane_bytes, int w, int h, int half, int bgr) {
    constexpr int RY = 8414, GY = 16519, BY = 3208, RU = -4865, GU = -9528, BU = 14392, RV = 14392, GV = -12061, BV = -2332;
    const int pairs = w >> 1;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = idx / pairs, i = idx - y * pairs;
    if (y >= h) return;
    const int f = blockIdx.z;
    const uint8_t *p = rgb + f * frame_stride + y * pitch + 6 * i;
    const int ro = bgr ? 2 : 0, bo = 2 - ro;   // bgr24: the same sums with R and B swapped in memory (bgr24ToY_c, bgr24ToUV*_c)
    const int r0 = p[ro], g0 = p[1], b0 = p[bo], r1 = p[3 + ro], g1 = p[4], b1 = p[3 + bo];
    uint8_t *out = planes + f * planes_frame_stride;
    uint16_t *yrow = reinterpret_cast<uint16_t *>(out) + static_cast<size_t>(y) * w + 2 * i;
    yrow[0] = static_cast<uint16_t>((RY * r0 + GY * g0 + BY * b0 + (32 << 14) + (1 << 8)) >> 9);
    yrow[1] = static_cast<uint16_t>((RY * r1 + GY * g1 + BY * b1 + (32 << 14) + (1 << 8)) >> 9);
    uint16_t *u = reinterpret_cast<uint16_t *>(out + u_offset), *v = reinterpret_cast<uint16_t *>(out + u_offset + c_plane_bytes);
    if (half) {
        const int r = r0 + r1, g = g0 + g1, b = b0 + b1;
        u[static_cast<size_t>(y) * pairs + i] = static_cast<uint16_t>((RU * r + GU * g + BU * b + (256 << 15) + (1 << 9)) >> 10);
        v[static_cast<size_t>(y) * pairs + i] = static_cast<uint16_t>((RV * r + GV * g + BV * b + (256 << 15) + (1 << 9)) >> 10);
    } else {
        const size_t o = static_cast<size_t>(y) * w + 2 * i;
        u[o] = static_cast<uint16_t>((RU * r0 + GU * g0 + BU * b0 + (256 << 14) + (1 << 8)) >> 9);
        u[o + 1] = static_cast<uint16_t>((RU * r1 + GU * g1 + BU * b1 + (256 << 14) + (1 << 8)) >> 9);
        v[o] = static_cast<uint16_t>((RV * r0 + GV * g0 + BV * b0 + (256 << 14) + (1 << 8)) >> 9);
        v[o + 1] = static_cast<uint16_t>((RV * r1 + GV * g1 + BV * b1 + (256 << 14) + (1 << 8)) >> 9);
    }
}

// The same stage for 16-bit components (rgb48le / rgb48be; input.c: rgb48ToY_c, rgb48ToUV_c, rgb48ToUV_half_c): full 16-bit
// lines, one rounding constant, and the half-width chroma averages its two pixels (with rounding) before the sum.
__global__ void __launch_bounds__(256)
rgb48_input_kernel(const uint8_t *__restrict__ rgb, ptrdiff_t pitch, ptrdiff_t frame_stride, uint8_t *__restrict__ planes,
                   size_t planes_frame_stride, size_t u_offset, size_t c_plane_bytes, int w, int h, int half, int be, int bgr) {
    constexpr int RY = 8414, GY = 16519, BY = 3208, RU = -4865, GU = -9528, BU = 14392, RV = 14392, GV = -12061, BV = -2332;
    const int pairs = w >> 1;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = idx / pairs, i = idx - y * pairs;
    if (y >= h) return;
    const int f = blockIdx.z;
    const uint8_t *p = rgb + f * frame_stride + y * pitch + 12 * i;
    int c[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) c[k] = be ? (p[2 * k] << 8 | p[2 * k + 1]) : (p[2 * k + 1] << 8 | p[2 * k]);
    if (bgr) {   // bgr48: R and B swapped in memory
        int t = c[0]; c[0] = c[2]; c[2] = t;
        t = c[3]; c[3] = c[5]; c[5] = t;
    }
    uint8_t *out = planes + f * planes_frame_stride;
    uint16_t *yrow = reinterpret_cast<uint16_t *>(out) + static_cast<size_t>(y) * w + 2 * i;
    // (at most 28141 * 65535 + (0x2001 << 14) < 2^31: int arithmetic, as in libswscale)
    yrow[0] = static_cast<uint16_t>((RY * c[0] + GY * c[1] + BY * c[2] + (0x2001 << 14)) >> 15);
    yrow[1] = static_cast<uint16_t>((RY * c[3] + GY * c[4] + BY * c[5] + (0x2001 << 14)) >> 15);
    uint16_t *u = reinterpret_cast<uint16_t *>(out + u_offset), *v = reinterpret_cast<uint16_t *>(out + u_offset + c_plane_bytes);
    if (half) {
        const int r = (c[0] + c[3] + 1) >> 1, g = (c[1] + c[4] + 1) >> 1, b = (c[2] + c[5] + 1) >> 1;
        u[static_cast<size_t>(y) * pairs + i] = static_cast<uint16_t>((RU * r + GU * g + BU * b + (0x10001 << 14)) >> 15);
        v[static_cast<size_t>(y) * pairs + i] = static_cast<uint16_t>((RV * r + GV * g + BV * b + (0x10001 << 14)) >> 15);
    } else {
        const size_t o = static_cast<size_t>(y) * w + 2 * i;
        u[o] = static_cast<uint16_t>((RU * c[0] + GU * c[1] + BU * c[2] + (0x10001 << 14)) >> 15);
        u[o + 1] = static_cast<uint16_t>((RU * c[3] + GU * c[4] + BU * c[5] + (0x10001 << 14)) >> 15);
        v[o] = static_cast<uint16_t>((RV * c[0] + GV * c[1] + BV * c[2] + (0x10001 << 14)) >> 15);
        v[o + 1] = static_cast<uint16_t>((RV * c[3] + GV * c[4] + BV * c[5] + (0x10001 << 14)) >> 15);
    }
}

// ---- memcpy-class fast paths of the same-size converters (the DVD production shape: yuv420p -> nv12 at equal size,
// worker/tasks.py:62-63) -------------------------------------------------------------------------------------------
// Plain 2-D copy of `row_bytes` per row.  V = bytes per thread access (16, 4 or 1), chosen by the launcher from the
// alignment of both sides; the row tail that does not fill a vector is copied bytewise by the last threads.
template <int V>
__global__ void __launch_bounds__(256)
copy_rows_kernel(const uint8_t *__restrict__ src, ptrdiff_t spitch, ptrdiff_t sframe, uint8_t *__restrict__ dst, ptrdiff_t dpitch,
                 ptrdiff_t dframe, int row_bytes, int rows) {
    // rows are short (a 720-byte DVD luma row is 45 vectors), so threads are spread over (row, vector) pairs
    const int nvec = row_bytes / V, slots = nvec + (row_bytes % V ? 1 : 0);
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = idx / slots, i = idx - y * slots;
    if (y >= rows) return;
    const int f = blockIdx.z;
    const uint8_t *s = src + f * sframe + y * spitch;
    uint8_t *d = dst + f * dframe + y * dpitch;
    if (i < nvec) {
        if (V == 16) reinterpret_cast<uint4 *>(d)[i] = __ldg(reinterpret_cast<const uint4 *>(s) + i);
        else if (V == 4) reinterpret_cast<unsigned *>(d)[i] = __ldg(reinterpret_cast<const unsigned *>(s) + i);
        else d[i] = __ldg(s + i);
    } else {
        for (int k = nvec * V; k < row_bytes; ++k) d[k] = s[k];   // the bytes that do not fill a vector
    }
}

// Two planar chroma planes -> one interleaved plane (T = sample type).  Each thread reads 8 bytes of U and 8 of V and
// writes 16 interleaved bytes when everything is aligned for it; otherwise one sample pair per thread.
template <typename T, bool VEC>
__global__ void __launch_bounds__(256)
interleave_kernel(const uint8_t *__restrict__ u, const uint8_t *__restrict__ v, ptrdiff_t spitch_u, ptrdiff_t spitch_v,
                  ptrdiff_t sframe_u, ptrdiff_t sframe_v, uint8_t *__restrict__ dst, ptrdiff_t dpitch, ptrdiff_t dframe, int w,
                  int rows) {
    constexpr int PER = VEC ? 8 / static_cast<int>(sizeof(T)) : 1;   // samples of each plane per thread
    const int ngroups = w / PER, slots = ngroups + (w % PER ? 1 : 0);
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = idx / slots, i = idx - y * slots;
    if (y >= rows) return;
    const int f = blockIdx.z;
    const T *us = reinterpret_cast<const T *>(u + f * sframe_u + y * spitch_u);
    const T *vs = reinterpret_cast<const T *>(v + f * sframe_v + y * spitch_v);
    T *d = reinterpret_cast<T *>(dst + f * dframe + y * dpitch);
    if (i < ngroups) {
        if (VEC) {
            const uint2 a = __ldg(reinterpret_cast<const uint2 *>(us) + i), b = __ldg(reinterpret_cast<const uint2 *>(vs) + i);
            uint4 o;
            if (sizeof(T) == 1) {
                o.x = __byte_perm(a.x, b.x, 0x5140); o.y = __byte_perm(a.x, b.x, 0x7362);
                o.z = __byte_perm(a.y, b.y, 0x5140); o.w = __byte_perm(a.y, b.y, 0x7362);
            } else {
                o.x = __byte_perm(a.x, b.x, 0x5410); o.y = __byte_perm(a.x, b.x, 0x7632);
                o.z = __byte_perm(a.y, b.y, 0x5410); o.w = __byte_perm(a.y, b.y, 0x7632);
            }
            reinterpret_cast<uint4 *>(d)[i] = o;
        } else {
            d[2 * i] = us[i];
            d[2 * i + 1] = vs[i];
        }
    } else {
        for (int k = ngroups * PER; k < w; ++k) { d[2 * k] = us[k]; d[2 * k + 1] = vs[k]; }   // samples that do not fill a vector
    }
}

// ---- exact 2:1 "area" down-scale of 8-bit planes (the 4K -> 1080p ladder rung) --------------------------------------
// With flags=area and an exact 2:1 ratio every row of libswscale's tables is [8192, 8192] (horizontal, >> 7) and
// [2048, 2048] (vertical, +2^18 >> 19) on sample pairs 2x, 2x+1 / rows 2y, 2y+1, so the whole chain collapses to
//     out = (p00 + p01 + p10 + p11 + 2) >> 2
// exactly: the 15-bit intermediate (p0 + p1) * 64 never reaches the 32767 clamp and the final value never leaves
// 0..255.  PAIR = false: one component per byte (planar planes, luma).  PAIR = true: interleaved UV bytes, the two
// components are filtered independently and stay interleaved.  Each thread produces 8 output bytes from 2 x 16 source
// bytes with 16-bit SIMD-in-a-register adds; VEC = false is the scalar form for unaligned rows.
template <bool PAIR, bool VEC>
__global__ void __launch_bounds__(256)
box2_kernel(const uint8_t *__restrict__ src, ptrdiff_t spitch, ptrdiff_t sframe, uint8_t *__restrict__ dst, ptrdiff_t dpitch,
            ptrdiff_t dframe, int out_bytes, int rows) {
    const int per = VEC ? 8 : (PAIR ? 2 : 1);                 // output bytes per thread
    const int slots = (out_bytes + per - 1) / per;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = idx / slots, i = idx - y * slots;
    if (y >= rows) return;
    const int f = blockIdx.z;
    const uint8_t *s0 = src + f * sframe + static_cast<ptrdiff_t>(2 * y) * spitch, *s1 = s0 + spitch;
    uint8_t *d = dst + f * dframe + y * dpitch;
    if (VEC && 8 * i + 8 <= out_bytes) {
        const uint4 a = __ldg(reinterpret_cast<const uint4 *>(s0) + i), b = __ldg(reinterpret_cast<const uint4 *>(s1) + i);
        const unsigned wa[4] = {a.x, a.y, a.z, a.w}, wb[4] = {b.x, b.y, b.z, b.w};
        unsigned o[4];   // per source word: two results in the low bytes of its 16-bit halves
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            // planar: [b0+b1, b2+b3]; interleaved: [U0, U1] and [V0, V1] in the halves of two masked words
            const unsigned even = (wa[k] & 0x00ff00ffu) + (wb[k] & 0x00ff00ffu);             // bytes 0 and 2 of both rows
            const unsigned odd = ((wa[k] >> 8) & 0x00ff00ffu) + ((wb[k] >> 8) & 0x00ff00ffu); // bytes 1 and 3
            if (!PAIR) {
                o[k] = ((even + odd + 0x00020002u) >> 2) & 0x00ff00ffu;                        // out0 | out1 << 16
            } else {
                const unsigned u = ((even & 0xffffu) + (even >> 16) + 2u) >> 2, v = ((odd & 0xffffu) + (odd >> 16) + 2u) >> 2;
                o[k] = u | (v << 8);                                                            // U | V << 8
            }
        }
        uint2 out;
        if (!PAIR) {
            out.x = __byte_perm(o[0], o[1], 0x6420);   // bytes 0,2 of o[0], then 0,2 of o[1]
            out.y = __byte_perm(o[2], o[3], 0x6420);
        } else {
            out.x = o[0] | (o[1] << 16);
            out.y = o[2] | (o[3] << 16);
        }
        reinterpret_cast<uint2 *>(d)[i] = out;
    } else {
        for (int k = per * i; k < min(per * i + per, out_bytes); ++k) {
            const int sx = PAIR ? 4 * (k >> 1) + (k & 1) : 2 * k, step = PAIR ? 2 : 1;
            d[k] = static_cast<uint8_t>((s0[sx] + s0[sx + step] + s1[sx] + s1[sx + step] + 2) >> 2);
        }
    }
}

inline dim3 plane_grid(int w, int h, int n_frames) { return dim3((w + 255) / 256, h, n_frames); }

}  // namespace

int launch_hpass_generic(const PlaneOp &op, const SampleView &src, const DeviceTable &h, int16_t *mid,
                         size_t mid_frame_stride, int n_frames, cudaStream_t stream) {
    const dim3 grid = plane_grid(op.dw, op.sh, n_frames);
    int16_t *m = mid + op.mid_offset;
    if (op.src_bits == 8)
        hpass_generic_kernel<uint8_t><<<grid, 256, 0, stream>>>(src.base, src.pitch, src.frame_stride, src.xstep,
            op.src_shift, op.sw, op.dw, h.coef, h.pos, h.taps, op.hshift, m, mid_frame_stride, op.sh, op.rc_coeff, op.rc_offset);
    else
        hpass_generic_kernel<uint16_t><<<grid, 256, 0, stream>>>(src.base, src.pitch, src.frame_stride, src.xstep,
            op.src_shift, op.sw, op.dw, h.coef, h.pos, h.taps, op.hshift, m, mid_frame_stride, op.sh, op.rc_coeff, op.rc_offset);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_vpass_generic(const PlaneOp &op, const SampleView &dst, const DeviceTable &v, const int16_t *mid,
                         size_t mid_frame_stride, int n_frames, cudaStream_t stream) {
    const int16_t *m = mid + op.mid_offset;
    if ((op.dw & 3) == 0 && (mid_frame_stride & 3) == 0 && (reinterpret_cast<uintptr_t>(m) & 7) == 0) {
        const int threads = op.dw / 4 >= 256 ? 256 : 64;
        const dim3 g((op.dw / 4 + threads - 1) / threads, op.dh, n_frames);
        if (op.dst_bits == 8)
            vpass_quad_kernel<uint8_t><<<g, threads, 0, stream>>>(m, mid_frame_stride, op.sh, op.dw, v.coef, v.pos, v.taps,
                dst.base, dst.pitch, dst.frame_stride, dst.xstep, op.dst_shift, op.rule, op.dither_xoff);
        else
            vpass_quad_kernel<uint16_t><<<g, threads, 0, stream>>>(m, mid_frame_stride, op.sh, op.dw, v.coef, v.pos, v.taps,
                dst.base, dst.pitch, dst.frame_stride, dst.xstep, op.dst_shift, op.rule, op.dither_xoff);
        return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
    }
    const dim3 grid = plane_grid(op.dw, op.dh, n_frames);
    if (op.dst_bits == 8)
        vpass_generic_kernel<uint8_t><<<grid, 256, 0, stream>>>(m, mid_frame_stride, op.sh, op.dw, v.coef, v.pos, v.taps,
            dst.base, dst.pitch, dst.frame_stride, dst.xstep, op.dst_shift, op.rule, op.dither_xoff);
    else
        vpass_generic_kernel<uint16_t><<<grid, 256, 0, stream>>>(m, mid_frame_stride, op.sh, op.dw, v.coef, v.pos, v.taps,
            dst.base, dst.pitch, dst.frame_stride, dst.xstep, op.dst_shift, op.rule, op.dither_xoff);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_repack(const PlaneOp &op, const SampleView &src, const SampleView &dst, int n_frames, cudaStream_t stream) {
    const dim3 grid = plane_grid(op.dw, op.dh, n_frames);
#define TVS_REPACK(S, D)                                                                                          \
    repack_kernel<S, D><<<grid, 256, 0, stream>>>(src.base, src.pitch, src.frame_stride, src.xstep, op.src_shift, \
        dst.base, dst.pitch, dst.frame_stride, dst.xstep, op.dst_shift, op.dw, op.kind, op.raw_tail_from, op.dither_xoff)
    if (op.src_bits == 8 && op.dst_bits == 8) TVS_REPACK(uint8_t, uint8_t);
    else if (op.src_bits == 8) TVS_REPACK(uint8_t, uint16_t);
    else if (op.dst_bits == 8) TVS_REPACK(uint16_t, uint8_t);
    else TVS_REPACK(uint16_t, uint16_t);
#undef TVS_REPACK
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

namespace {
inline bool aligned_for(const void *p, ptrdiff_t pitch, ptrdiff_t frame, int n_frames, int a) {
    return ((reinterpret_cast<uintptr_t>(p) | static_cast<uintptr_t>(pitch) | (n_frames > 1 ? static_cast<uintptr_t>(frame) : 0u)) & (a - 1)) == 0;
}
}  // namespace

// libswscale's same-size p010le -> p010le copy (see copy_plane in oracle/swscale_port.c): the stored words are copied, but
// the first word of every group of four and the row's last (words % 4) words lose their six padding bits.  One thread per
// group of four words; a no-op conversion nobody should need fast.
__global__ void __launch_bounds__(256)
p010_copy_kernel(const uint8_t *__restrict__ src, ptrdiff_t spitch, ptrdiff_t sframe, uint8_t *__restrict__ dst, ptrdiff_t dpitch,
                 ptrdiff_t dframe, int words, int rows) {
    const int groups = (words + 3) >> 2;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = idx / groups, g = idx - y * groups;
    if (y >= rows) return;
    const uint16_t *s = reinterpret_cast<const uint16_t *>(src + blockIdx.z * sframe + y * spitch);
    uint16_t *d = reinterpret_cast<uint16_t *>(dst + blockIdx.z * dframe + y * dpitch);
    for (int k = 0; k < 4; ++k) {
        const int i = 4 * g + k;
        if (i >= words) break;
        uint16_t v = s[i];
        if (k == 0 || i >= (words & ~3)) v &= 0xffc0;
        d[i] = v;
    }
}

int launch_p010_copy(const SampleView &src, const SampleView &dst, int words, int rows, int n_frames, cudaStream_t stream) {
    const unsigned total = static_cast<unsigned>((words + 3) / 4) * rows;
    const dim3 grid((total + 255) / 256, 1, n_frames);
    p010_copy_kernel<<<grid, 256, 0, stream>>>(src.base, src.pitch, src.frame_stride, dst.base, dst.pitch, dst.frame_stride, words, rows);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_copy_rows(const SampleView &src_in, const SampleView &dst_in, int row_bytes, int rows, int n_frames, cudaStream_t stream) {
    SampleView src = src_in, dst = dst_in;
    // Tightly packed planes (what a rawvideo pipe delivers) are one long row: 854-wide frames then copy with 16-byte vectors
    // although their rows are only 2-byte aligned.
    if (rows > 1 && src.pitch == row_bytes && dst.pitch == row_bytes && static_cast<long long>(row_bytes) * rows < (1LL << 30)) {
        row_bytes *= rows; rows = 1; src.pitch = dst.pitch = 0;
    }
    const int v = aligned_for(src.base, src.pitch, src.frame_stride, n_frames, 16) && aligned_for(dst.base, dst.pitch, dst.frame_stride, n_frames, 16) ? 16
                : aligned_for(src.base, src.pitch, src.frame_stride, n_frames, 4) && aligned_for(dst.base, dst.pitch, dst.frame_stride, n_frames, 4) ? 4 : 1;
    const int slots = row_bytes / v + (row_bytes % v ? 1 : 0);
    const dim3 grid((static_cast<unsigned>(slots) * rows + 255) / 256, 1, n_frames);
    if (v == 16) copy_rows_kernel<16><<<grid, 256, 0, stream>>>(src.base, src.pitch, src.frame_stride, dst.base, dst.pitch, dst.frame_stride, row_bytes, rows);
    else if (v == 4) copy_rows_kernel<4><<<grid, 256, 0, stream>>>(src.base, src.pitch, src.frame_stride, dst.base, dst.pitch, dst.frame_stride, row_bytes, rows);
    else copy_rows_kernel<1><<<grid, 256, 0, stream>>>(src.base, src.pitch, src.frame_stride, dst.base, dst.pitch, dst.frame_stride, row_bytes, rows);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_interleave(const SampleView &u_in, const SampleView &v_in, const SampleView &dst_in, int sample_bytes, int w, int rows, int n_frames,
                      cudaStream_t stream) {
    SampleView u = u_in, v = v_in, dst = dst_in;
    if (rows > 1 && u.pitch == static_cast<ptrdiff_t>(w) * sample_bytes && v.pitch == u.pitch && dst.pitch == 2 * u.pitch &&
        static_cast<long long>(w) * rows < (1LL << 29)) {   // tightly packed planes: one long row (see launch_copy_rows)
        w *= rows; rows = 1; u.pitch = v.pitch = dst.pitch = 0;
    }
    const bool vec = aligned_for(u.base, u.pitch, u.frame_stride, n_frames, 8) && aligned_for(v.base, v.pitch, v.frame_stride, n_frames, 8) &&
                     aligned_for(dst.base, dst.pitch, dst.frame_stride, n_frames, 16);
    const int per = vec ? 8 / sample_bytes : 1;
    const int slots = w / per + (w % per ? 1 : 0);
    const dim3 grid((static_cast<unsigned>(slots) * rows + 255) / 256, 1, n_frames);
#define TVS_ILV(T, V) interleave_kernel<T, V><<<grid, 256, 0, stream>>>(u.base, v.base, u.pitch, v.pitch, u.frame_stride, v.frame_stride, \
                                                                      dst.base, dst.pitch, dst.frame_stride, w, rows)
    if (sample_bytes == 1) { if (vec) TVS_ILV(uint8_t, true); else TVS_ILV(uint8_t, false); }
    else { if (vec) TVS_ILV(uint16_t, true); else TVS_ILV(uint16_t, false); }
#undef TVS_ILV
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_dither2_rows(const PlaneOp &op, const SampleView &src, const SampleView &dst, int n_frames, cudaStream_t stream) {
    if (!aligned_for(src.base, src.pitch, src.frame_stride, n_frames, 16) || !aligned_for(dst.base, dst.pitch, dst.frame_stride, n_frames, 8) ||
        (op.raw_tail_from >= 0 && (op.raw_tail_from & 7)))
        return 0;   // not this kernel's case: the caller falls back to the per-sample kernel
    const int slots = (op.dw + 7) / 8;
    const dim3 grid((static_cast<unsigned>(slots) * op.dh + 255) / 256, 1, n_frames);
    dither2_rows_kernel<<<grid, 256, 0, stream>>>(src.base, src.pitch, src.frame_stride, dst.base, dst.pitch, dst.frame_stride, op.dw, op.dh,
                                                  op.src_shift, op.raw_tail_from);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_dither2_interleave(const PlaneOp &op, const SampleView &u, const SampleView &v, const SampleView &dst, int n_frames,
                              cudaStream_t stream) {
    if (!aligned_for(u.base, u.pitch, u.frame_stride, n_frames, 16) || !aligned_for(v.base, v.pitch, v.frame_stride, n_frames, 16) ||
        !aligned_for(dst.base, dst.pitch, dst.frame_stride, n_frames, 16))
        return 0;
    const int slots = (op.dw + 7) / 8;
    const dim3 grid((static_cast<unsigned>(slots) * op.dh + 255) / 256, 1, n_frames);
    dither2_interleave_kernel<<<grid, 256, 0, stream>>>(u.base, v.base, u.pitch, v.pitch, u.frame_stride, v.frame_stride, dst.base, dst.pitch,
                                                        dst.frame_stride, op.dw, op.dh, op.src_shift);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_dither128_interleave(const PlaneOp &op, const SampleView &u, const SampleView &v, const SampleView &dst, int n_frames,
                                cudaStream_t stream) {
    if (!aligned_for(u.base, u.pitch, u.frame_stride, n_frames, 16) || !aligned_for(v.base, v.pitch, v.frame_stride, n_frames, 16) ||
        !aligned_for(dst.base, dst.pitch, dst.frame_stride, n_frames, 16))
        return 0;
    const int slots = (op.dw + 7) / 8;
    const dim3 grid((static_cast<unsigned>(slots) * op.dh + 255) / 256, 1, n_frames);
    dither128_interleave_kernel<<<grid, 256, 0, stream>>>(u.base, v.base, u.pitch, v.pitch, u.frame_stride, v.frame_stride, dst.base,
                                                          dst.pitch, dst.frame_stride, op.dw, op.dh, op.src_shift, 0, 3);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_rgb_input(const uint8_t *rgb, ptrdiff_t pitch, ptrdiff_t frame_stride, uint8_t *planes, size_t planes_frame_stride,
                     size_t u_offset_bytes, size_t c_plane_bytes, int w, int h, bool half, bool bgr, int wide, int n_frames, cudaStream_t stream) {
    const unsigned total = static_cast<unsigned>(w / 2) * h;
    const dim3 grid((total + 255) / 256, 1, n_frames);
    if (wide)   // 1 = rgb48le, 2 = rgb48be
        rgb48_input_kernel<<<grid, 256, 0, stream>>>(rgb, pitch, frame_stride, planes, planes_frame_stride, u_offset_bytes, c_plane_bytes, w, h,
                                                     half ? 1 : 0, wide == 2 ? 1 : 0, bgr ? 1 : 0);
    else
        rgb_input_kernel<<<grid, 256, 0, stream>>>(rgb, pitch, frame_stride, planes, planes_frame_stride, u_offset_bytes, c_plane_bytes, w, h,
                                                   half ? 1 : 0, bgr ? 1 : 0);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_box2(const SampleView &src, const SampleView &dst, bool pair, int out_bytes, int rows, int n_frames, cudaStream_t stream) {
    const bool vec = aligned_for(src.base, src.pitch, src.frame_stride, n_frames, 16) && aligned_for(dst.base, dst.pitch, dst.frame_stride, n_frames, 8);
    const int per = vec ? 8 : (pair ? 2 : 1);
    const int slots = (out_bytes + per - 1) / per;
    const dim3 grid((static_cast<unsigned>(slots) * rows + 255) / 256, 1, n_frames);
#define TVS_BOX(P, V) box2_kernel<P, V><<<grid, 256, 0, stream>>>(src.base, src.pitch, src.frame_stride, dst.base, dst.pitch, dst.frame_stride, out_bytes, rows)
    if (pair) { if (vec) TVS_BOX(true, true); else TVS_BOX(true, false); }
    else { if (vec) TVS_BOX(false, true); else TVS_BOX(false, false); }
#undef TVS_BOX
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tvs
