// Streaming ("line buffer") kernel for vertical down-scales.
//
// The tile kernel (kernels_fused_impl.cuh) pays for every tile with a vertical halo, a register-ring warm-up and
// per-tile bookkeeping.  That is cheap when a tile's source footprint is small (up-scales) and expensive when the
// source is the bulk of the data: a 1080p -> 720p tile only fits 16 output rows in shared memory.  Here a CTA owns
// a 128-column strip of one plane of one frame and walks down the whole height instead:
//
//   for each block of KB source rows:   cp.async the raw rows (double buffered, one block ahead)
//                                       convert to floats                          -> S      (own chunks, no barrier)
//                                       horizontal pass of those rows              -> M ring (barrier before and after)
//                                       vertical pass of every output row whose window is now complete -> HBM
//
// The intermediate lives in a shared-memory ring of RM rows (a power of two >= the live window + KB), so no source
// row is ever filtered twice, the lane's horizontal coefficients stay in registers for the whole strip, and shared
// memory per CTA is a few tens of KB (5+ CTAs per SM).  Vertical windows of consecutive output rows barely overlap
// when down-scaling, so the vertical pass simply reads its VT rows from the ring (no register ring to maintain).
// Arithmetic, exactness argument and output rules are those of the tile kernel (scale_device.cuh).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "scale_device.cuh"
#include "tvs_internal.h"

namespace tvs {
namespace {

constexpr int kStripW = 128;              // output columns per CTA: lane * 4 + k
constexpr int kSThreads = 128;
constexpr int kSWarps = kSThreads / 32;
constexpr int kMaxStrips = 96;
constexpr int kMaxStreamOps = 3;
constexpr int kStreamCtasPerSM = 5;
constexpr size_t kStreamSmemBudget = (227 * 1024) / kStreamCtasPerSM - 1536;

struct StreamOpParams {
    const uint8_t *src; ptrdiff_t src_pitch, src_fstride; int src_xstep;
    uint8_t *dst; ptrdiff_t dst_pitch, dst_fstride; int dst_xstep;
    int sw, sh, dw, dh;
    int strips, strip_begin;
    const float *hcoef;    // [dw][HT]             c * 2^-(23+hshift) in tap order (zero padded to HT)
    const int *hpos;       // [dw]                 first source column of each output column
    const short *hcoef_i;  // [dw][h_taps]         the integer coefficients (raw-byte horizontal pass)
    int h_taps;            // real horizontal taps (row length of hcoef_i)
    int pf_step_rows, pf_step_chunks;   // kSThreads split into whole staged rows + chunks (prefetch loop stride)
    const int *vpos;       // [dh]
    const float *vpair;    // [dh][VT][2]          (g, g) * 2^107 in tap order
    const int *vint;       // [dh][VT]             the same coefficients as integers (exact path)
    const int *block_yend; // [n_blocks]           output rows complete once source block b has been filtered
    int kb, n_blocks;      // source rows per block
    int rm_mask;           // rows of the intermediate ring - 1 (power of two)
    int s_pitch, s_chunks, s_chunks_log2;
    int s_split;           // 16-bit sources at 2:1: a staged row keeps its even samples first and its odd samples from
                           // this float offset on, so that lanes two samples apart read neighbouring words (0 = natural order)
    unsigned chunk_magic, word_magic;   // ceil(2^32 / s_chunks), ceil(2^32 / (4 * s_chunks)): exact division of small indices
    const uint8_t *src_lo; size_t src_extent;
    float v_thr, v_init;
    int src_mode; float src_scale, h_addend;
    int dst_shift, dither_xoff;
    short strip_c0[kMaxStrips], strip_nc[kMaxStrips];
};

struct StreamLaunchParams {
    StreamOpParams op[kMaxStreamOps];
    int n_ops;
    int m_floats, s_floats, raw_words;   // shared-memory layout (element counts), maxima over the planes of the launch
};

template <int VT, int OUT>
__device__ __noinline__ uint2 exact_fixup_ring(const float *m_col, int mask, int first, const int *coef, float v_thr, float r0,
                                               float r1, float r2, float r3, int4 dither, int dst_shift, uint2 packed) {
    unsigned word[2] = {packed.x, packed.y};
    const float r[4] = {r0, r1, r2, r3};
    const int d[4] = {dither.x, dither.y, dither.z, dither.w};
    int a[4] = {0, 0, 0, 0};
    for (int j = 0; j < VT; ++j) {
        const float4 t = *reinterpret_cast<const float4 *>(m_col + ((first + j) & mask) * kStripW);
        const int c = __ldg(coef + j);
        a[0] += subnormal_to_int(t.x) * c; a[1] += subnormal_to_int(t.y) * c;
        a[2] += subnormal_to_int(t.z) * c; a[3] += subnormal_to_int(t.w) * c;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (fabsf(r[k]) >= v_thr) {
            const int bias = OUT == kOutU8 ? (64 << 12) : OUT == kOutU8Dither ? (d[k] << 12) : (1 << 16);
            const unsigned v = static_cast<unsigned>(min(max((a[k] + bias) >> OutRule<OUT>::kShift, 0), OutRule<OUT>::kMax));
            if (OUT == kOutU10) {
                const int sh = 16 * (k & 1);
                word[k >> 1] = (word[k >> 1] & ~(0xffffu << sh)) | ((v << dst_shift) << sh);
            } else {
                word[0] = (word[0] & ~(0xffu << (8 * k))) | (v << (8 * k));
            }
        }
    }
    return make_uint2(word[0], word[1]);
}

// cp.async the raw bytes of source rows [r_b, r_b + kb) of the strip's footprint (see prefetch_tile_source of the tile
// kernel for the chunk / alignment rules; identical here).
__device__ __forceinline__ void prefetch_block(const StreamOpParams &op, const uint8_t *strip_base, const uint8_t *lo, const uint8_t *hi,
                                               int r_b, int rows, int last_byte, unsigned *raw, int tid) {
    const int s_chunks = op.s_chunks;
    const ptrdiff_t pitch = op.src_pitch;
    const uint8_t *base = strip_base + static_cast<ptrdiff_t>(r_b) * pitch;
    const unsigned mis0 = static_cast<unsigned>(reinterpret_cast<uintptr_t>(base)) & 15u;
    const unsigned pmod = static_cast<unsigned>(pitch) & 15u;
    // (row, chunk) pairs are spread over all threads (a row is rarely a power-of-two number of chunks); the pair of
    // the next index is derived from the previous one, no division in the loop
    const int total = rows * s_chunks;
    const int step_rows = op.pf_step_rows, step_chunks = op.pf_step_chunks;   // kSThreads = step_rows * s_chunks + step_chunks
    int row = static_cast<int>(__umulhi(static_cast<unsigned>(tid), op.chunk_magic));
    int ch = tid - row * s_chunks;
    const int pitch32 = static_cast<int>(pitch);   // a block spans far less than 2 GB
    // every chunk of an interior block lies inside the plane's allocation: no per-chunk range checks
    const bool interior = base - 16 >= lo && base + static_cast<ptrdiff_t>(rows - 1) * pitch + 16 * s_chunks <= hi;
    for (int idx = tid; idx < total; idx += kSThreads) {
        const unsigned mis = (mis0 + row * pmod) & 15u;
        uint4 *dst = reinterpret_cast<uint4 *>(raw) + idx;
        const int first = 16 * ch - static_cast<int>(mis);
        const uint8_t *g = base + (row * pitch32 + first);
        if (first > last_byte) {
            *dst = make_uint4(0u, 0u, 0u, 0u);
        } else if (interior || (g >= lo && g + 16 <= hi)) {
            cp_async16(dst, g);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                unsigned *d = reinterpret_cast<unsigned *>(dst) + k;
                if (first + 4 * k <= last_byte && first + 4 * k + 3 >= 0) cp_async4(d, g + 4 * k);
                else *d = 0u;
            }
        }
        row += step_rows;
        ch += step_chunks;
        if (ch >= s_chunks) { ch -= s_chunks; ++row; }
    }
}

// One 16-bit sample (byte pair `selector` of `word`) -> (v * scale - 512) as a float, in two instructions: PRMT puts
// the sample under the exponent of 2^23 and one FMA removes 2^23 * scale + 512 (`bias`); every step is exact.
__device__ __forceinline__ float wide_to_float(unsigned word, unsigned selector, float scale, float bias) {
    return fmaf(__uint_as_float(__byte_perm(word, 0x4B000000u, selector)), scale, bias);
}

// Raw staged bytes -> float samples of S (the caller has made the copied chunks visible with a barrier).
// 8-bit sources: consecutive threads take consecutive 4-byte words, so a warp's 16-byte stores land on consecutive
// banks.  16-bit sources: a thread takes a whole 16-byte chunk (8 samples, or 4 of an interleaved pair plane) and
// steps its (row, chunk) position without divisions.
__device__ __forceinline__ void convert_block(const StreamOpParams &op, const unsigned *raw, float *S, int rows, unsigned mis0, unsigned pmod,
                                              int tid) {
    const int s_pitch = op.s_pitch;
    const int src_mode = op.src_mode;
    if (src_mode >= kSrcU16) {
        const int s_chunks = op.s_chunks;
        const float sscale = op.src_scale, bias = -(8388608.0f * sscale + 512.0f);
        const int total = rows * s_chunks;
        const int step_rows = op.pf_step_rows, step_chunks = op.pf_step_chunks;
        int row = static_cast<int>(__umulhi(static_cast<unsigned>(tid), op.chunk_magic));
        int ch = tid - row * s_chunks;
        // P010 keeps its 10 bits in the top of the word: whatever a decoder left in the six bits below is dropped first, as
        // libswscale's integer shift drops it (sscale < 1 marks that layout)
        const unsigned keep = sscale < 1.0f ? 0xFFC0FFC0u : 0xFFFFFFFFu;
        for (int idx = tid; idx < total; idx += kSThreads) {
            uint4 v = reinterpret_cast<const uint4 *>(raw)[idx];
            v.x &= keep; v.y &= keep; v.z &= keep; v.w &= keep;
            float *srow = S + row * s_pitch;
            if (src_mode == kSrcU16 && op.s_split) {   // eight samples per chunk, even and odd positions apart
                reinterpret_cast<float4 *>(srow)[ch] = make_float4(wide_to_float(v.x, 0x7410u, sscale, bias), wide_to_float(v.y, 0x7410u, sscale, bias),
                                                                  wide_to_float(v.z, 0x7410u, sscale, bias), wide_to_float(v.w, 0x7410u, sscale, bias));
                reinterpret_cast<float4 *>(srow + op.s_split)[ch] = make_float4(wide_to_float(v.x, 0x7432u, sscale, bias), wide_to_float(v.y, 0x7432u, sscale, bias),
                                                                               wide_to_float(v.z, 0x7432u, sscale, bias), wide_to_float(v.w, 0x7432u, sscale, bias));
            } else if (src_mode == kSrcU16) {          // eight samples per chunk
                reinterpret_cast<float4 *>(srow)[2 * ch] = make_float4(wide_to_float(v.x, 0x7410u, sscale, bias), wide_to_float(v.x, 0x7432u, sscale, bias),
                                                                      wide_to_float(v.y, 0x7410u, sscale, bias), wide_to_float(v.y, 0x7432u, sscale, bias));
                reinterpret_cast<float4 *>(srow)[2 * ch + 1] = make_float4(wide_to_float(v.z, 0x7410u, sscale, bias), wide_to_float(v.z, 0x7432u, sscale, bias),
                                                                          wide_to_float(v.w, 0x7410u, sscale, bias), wide_to_float(v.w, 0x7432u, sscale, bias));
            } else if (op.s_split) {        // interleaved 16-bit chroma, even and odd sample positions apart
                const unsigned sel = ((mis0 + row * pmod) & 2u) ? 0x7432u : 0x7410u;
                reinterpret_cast<float2 *>(srow)[ch] = make_float2(wide_to_float(v.x, sel, sscale, bias), wide_to_float(v.z, sel, sscale, bias));
                reinterpret_cast<float2 *>(srow + op.s_split)[ch] = make_float2(wide_to_float(v.y, sel, sscale, bias), wide_to_float(v.w, sel, sscale, bias));
            } else {                        // interleaved 16-bit chroma: one sample of this plane per word
                const unsigned sel = ((mis0 + row * pmod) & 2u) ? 0x7432u : 0x7410u;
                reinterpret_cast<float4 *>(srow)[ch] = make_float4(wide_to_float(v.x, sel, sscale, bias), wide_to_float(v.y, sel, sscale, bias),
                                                                  wide_to_float(v.z, sel, sscale, bias), wide_to_float(v.w, sel, sscale, bias));
            }
            row += step_rows;
            ch += step_chunks;
            if (ch >= s_chunks) { ch -= s_chunks; ++row; }
        }
        return;
    }
    const int words = op.s_chunks * 4;
    const unsigned magic = op.word_magic;
    for (int idx = tid; idx < rows * words; idx += kSThreads) {
        const int row = static_cast<int>(__umulhi(static_cast<unsigned>(idx), magic));
        const int w = idx - row * words;
        const unsigned v = raw[idx];
        float *srow = S + row * s_pitch;
        if (src_mode == kSrcU8) {              // four samples per word
            reinterpret_cast<float4 *>(srow)[w] =
                make_float4(byte_to_float(v, 0x7540u), byte_to_float(v, 0x7541u), byte_to_float(v, 0x7542u), byte_to_float(v, 0x7543u));
        } else {                               // interleaved 8-bit chroma: this plane is every other byte
            const unsigned p = (mis0 + row * pmod) & 1u;
            reinterpret_cast<float2 *>(srow)[w] = make_float2(byte_to_float(v, 0x7540u | p), byte_to_float(v, 0x7542u | p));
        }
    }
}

__device__ __forceinline__ int dp2a_lo(unsigned pixels, int coef_pair, int acc) {
    int d;   // acc + s16(coef_pair.lo) * u8(pixels.b0) + s16(coef_pair.hi) * u8(pixels.b1)
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(coef_pair), "r"(pixels), "r"(acc));
    return d;
}
__device__ __forceinline__ int dp2a_hi(unsigned pixels, int coef_pair, int acc) {
    int d;   // the same on bytes 2 and 3
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(coef_pair), "r"(pixels), "r"(acc));
    return d;
}

// NW = 0: the horizontal pass runs on float samples (any source).  NW > 0: 8-bit sources whose rows are 4-byte
// aligned relative to each other skip the float conversion altogether: the pass reads the raw staged words and does
// two taps per IDP.2A, with each column's 16-bit coefficients pre-placed at the byte positions of its window inside NW
// consecutive words (the placement depends only on the column and on the strip's alignment, not on the row).
// Resident CTAs the register budget is set for (measured, one B200): the raw-byte instances gain 6-11 % at 6 CTAs
// (80 registers) over 4 (128), the float instances are best at 5 (96 registers).
constexpr int stream_min_ctas(int nw) { return nw > 0 ? 6 : 5; }
template <int HT, int VT, int OUT, int NW>
__global__ void __launch_bounds__(kSThreads, stream_min_ctas(NW))
stream_scale_kernel(const __grid_constant__ StreamLaunchParams P) {
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    int sx = blockIdx.x, oi = 0;
    if (P.n_ops > 1 && sx >= P.op[1].strip_begin) oi = 1;
    if (P.n_ops > 2 && sx >= P.op[2].strip_begin) oi = 2;
    const StreamOpParams &op = P.op[oi];
    sx -= op.strip_begin;
    const int f = blockIdx.y;

    float *M = smem;                                                // [rm_mask + 1][128] intermediate ring, t * 2^-149
    float *S = M + P.m_floats;                                      // [kb][s_pitch]      float samples of the current block
    unsigned *RAW = reinterpret_cast<unsigned *>(S + P.s_floats);   // 2 x [kb][s_chunks] raw 16-byte chunks

    const int dw = op.dw, dh = op.dh, sh = op.sh, kb = op.kb, mask = op.rm_mask;
    const int x0 = sx * kStripW;
    const int c0 = op.strip_c0[sx], ncols = op.strip_nc[sx];
    const int xs = op.src_xstep;
    const int sample_bytes = op.src_mode >= kSrcU16 ? 2 : 1;
    const int last_byte = (min(ncols, op.sw - c0) - 1) * xs + sample_bytes - 1;
    const uint8_t *strip_base = op.src + f * op.src_fstride + static_cast<ptrdiff_t>(c0) * xs;   // row 0 of the footprint
    const uint8_t *lo = op.src_lo + f * op.src_fstride, *hi = lo + op.src_extent;
    const ptrdiff_t src_pitch = op.src_pitch;
    const unsigned pmod = static_cast<unsigned>(src_pitch) & 15u;
    const int raw_words = P.raw_words;

    // Horizontal pass: a lane owns columns lane, lane + 32, lane + 64, lane + 96 of the strip.  When down-scaling,
    // neighbouring output columns read windows that are a whole source step apart, so consecutive lanes touch
    // consecutive shared-memory words (no bank conflicts) and every multiply-add is a real tap.  The coefficients of
    // the four columns stay in registers for the whole strip.
    float ck[NW ? 1 : 4][NW ? 1 : HT];
    int cp[NW ? 4 : 1][NW ? 2 * NW : 1];   // raw mode: (c[p], c[p+1]) pairs by byte position inside the NW words
    int hstart[4];                       // float mode: first sample of the window; raw mode: its first word
    const int src_mode = op.src_mode;
    if (NW == 0) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = min(x0 + lane + 32 * k, dw - 1);
            hstart[k] = __ldg(op.hpos + x) - c0;
#pragma unroll
            for (int j = 0; j < HT; ++j) ck[k][j] = __ldg(op.hcoef + static_cast<size_t>(x) * HT + j);
        }
    } else {
        // rows are a multiple of 4 bytes apart, so the byte phase of a window is the same in every row
        const int a = static_cast<int>(reinterpret_cast<uintptr_t>(strip_base) & 3u);
        const int taps = op.h_taps;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = min(x0 + lane + 32 * k, dw - 1);
            const int byte0 = a + (__ldg(op.hpos + x) - c0) * xs;      // position of tap 0 relative to the aligned row start
            hstart[k] = byte0 & ~3;   // byte offset of the window's first word
            const int phase = byte0 & 3;
            const short *ci = op.hcoef_i + static_cast<size_t>(x) * taps;
#pragma unroll
            for (int h = 0; h < 2 * NW; ++h) {
                // byte positions 2h and 2h + 1 of the window's words hold taps (2h - phase) / xs and (2h + 1 - phase) / xs
                int lo16 = 0, hi16 = 0;
                const int p0 = 2 * h - phase, p1 = p0 + 1;
                const int xsh = xs - 1;   // the byte step of a sample is 1 or 2
                if (p0 >= 0 && (p0 & xsh) == 0 && (p0 >> xsh) < taps) lo16 = __ldg(ci + (p0 >> xsh));
                if (p1 >= 0 && (p1 & xsh) == 0 && (p1 >> xsh) < taps) hi16 = __ldg(ci + (p1 >> xsh));
                cp[k][h] = (lo16 & 0xffff) | (hi16 << 16);
            }
        }
    }
    const float h_addend = op.h_addend, t_max = __int_as_float(32767);

    // rows the vertical pass may touch before they are produced (tap padding) must read as finite values
    for (int i = tid; i < (mask + 1) * (kStripW / 4); i += kSThreads) reinterpret_cast<float4 *>(M)[i] = make_float4(0.f, 0.f, 0.f, 0.f);

    // vertical-pass constants
    const int xq = x0 + 4 * lane;
    const int dst_xstep = op.dst_xstep, dst_shift = op.dst_shift;
    const ptrdiff_t dst_pitch = op.dst_pitch;
    uint8_t *dst_base = op.dst + f * op.dst_fstride;
    constexpr int kSample = OUT == kOutU10 ? 2 : 1, kAlign = 4 * kSample - 1;
    const bool fast_store = dst_xstep == kSample && (dst_pitch & kAlign) == 0 && x0 + kStripW <= dw &&
                            ((reinterpret_cast<uintptr_t>(dst_base) + static_cast<uintptr_t>(x0) * kSample) & kAlign) == 0;
    const float v_thr = op.v_thr, v_init = op.v_init;
    const int dcol = (xq + op.dither_xoff) & 7;
    const float *m_col = M + 4 * lane;

    prefetch_block(op, strip_base, lo, hi, 0, min(kb, sh), last_byte, RAW, tid);
    cp_async_commit();
    int y_next = 0;
    for (int b = 0; b < op.n_blocks; ++b) {
        const int r_b = b * kb, rows = min(kb, sh - r_b);
        const int y_end = __ldg(op.block_yend + b);   // needed after the horizontal pass; asked for now
        cp_async_wait_all();
        if (b + 1 < op.n_blocks) {
            prefetch_block(op, strip_base, lo, hi, r_b + kb, min(kb, sh - r_b - kb), last_byte, RAW + ((b + 1) & 1) * raw_words, tid);
            cp_async_commit();
        }
        __syncthreads();   // the block's raw rows are visible to everyone; S is free (previous horizontal pass is over)
        const unsigned mis0 = static_cast<unsigned>(reinterpret_cast<uintptr_t>(strip_base + static_cast<ptrdiff_t>(r_b) * src_pitch)) & 15u;
        if (NW == 0) {
            convert_block(op, RAW + (b & 1) * raw_words, S, rows, mis0, pmod, tid);
            __syncthreads();   // S holds the block
        }

        // ---- horizontal pass of the block's rows into the ring ----
        if (NW == 0) {
            const int split = op.s_split;
            for (int i = warp; i < rows; i += kSWarps) {
                const unsigned mis = (mis0 + i * pmod) & 15u;
                const int mis_samples = src_mode == kSrcU8 ? mis : src_mode == kSrcU16Pair ? (mis >> 2) : (mis >> 1);
                const float *srow = S + i * op.s_pitch;
                float *mrow = M + ((r_b + i) & mask) * kStripW + lane;
                // taps in pairs: one FFMA2 does two of them (the sums are exact integers * 2^-(23+hshift), so the order of
                // the additions does not matter).  `pe` / `po` address the even / odd taps, STEP floats apart.
                auto column = [&](const float *pe, const float *po, auto step_tag, int k) {
                    constexpr int STEP = decltype(step_tag)::value;
                    f32x2 acc2 = 0ull;
#pragma unroll
                    for (int j = 0; j + 1 < HT; j += 2) acc2 = fma2(pack2(pe[(j >> 1) * STEP], po[(j >> 1) * STEP]), pack2(ck[k][j], ck[k][j + 1]), acc2);
                    float even, odd;
                    unpack2(acc2, even, odd);
                    float acc = even + odd;
                    if (HT & 1) acc = fmaf(pe[(HT >> 1) * STEP], ck[k][HT - 1], acc);
                    return acc;
                };
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int h0 = hstart[k] + mis_samples;
                    float acc;
                    if (split) {
                        // split rows: the window's even-position samples in one half of the row, the odd ones in the other
                        const int par = h0 & 1;
                        acc = column(srow + (par ? split : 0) + (h0 >> 1), srow + (par ? 0 : split) + (h0 >> 1) + par,
                                     std::integral_constant<int, 1>{}, k);
                    } else {
                        acc = column(srow + h0, srow + h0 + 1, std::integral_constant<int, 2>{}, k);
                    }
                    mrow[32 * k] = fminf(fma_rm(acc, 0x1p-126f, h_addend), t_max);   // floor(sum >> hshift), clamped above
                }
            }
        } else {
            const char *rblock = reinterpret_cast<const char *>(RAW + (b & 1) * raw_words);
            const int row_bytes = op.s_chunks * 16;
            for (int i = warp; i < rows; i += kSWarps) {
                const unsigned mis = (mis0 + i * pmod) & 15u;        // a multiple of 4 apart from row to row
                const char *rrow = rblock + i * row_bytes + (mis & ~3u);
                unsigned *mrow = reinterpret_cast<unsigned *>(M) + ((r_b + i) & mask) * kStripW + lane;
                unsigned v[4][NW ? NW : 1];
#pragma unroll
                for (int k = 0; k < 4; ++k)
#pragma unroll
                    for (int w = 0; w < NW; ++w) v[k][w] = *reinterpret_cast<const unsigned *>(rrow + hstart[k] + 4 * w);
                // the four columns' dot products advance together: four independent dependency chains
                int acc[4] = {0, 0, 0, 0};
#pragma unroll
                for (int w = 0; w < NW; ++w) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) acc[k] = dp2a_lo(v[k][w], cp[k][2 * w], acc[k]);
#pragma unroll
                    for (int k = 0; k < 4; ++k) acc[k] = dp2a_hi(v[k][w], cp[k][2 * w + 1], acc[k]);
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int t = min(acc[k] >> 7, 32767);
                    mrow[32 * k] = static_cast<unsigned>(abs(t)) | (static_cast<unsigned>(t) & 0x80000000u);   // sign-magnitude = t * 2^-149
                }
            }
        }
        __syncthreads();   // the ring holds every row up to r_b + rows

        // ---- vertical pass of the output rows whose windows are now complete ----
        for (int y = y_next + warp; y < y_end; y += kSWarps) {
            const int first = __ldg(op.vpos + y);
            const f32x2 *gp = reinterpret_cast<const f32x2 *>(op.vpair) + static_cast<size_t>(y) * VT;
            f32x2 a01 = pack2(v_init, v_init), a23 = a01;
            int4 dither = make_int4(0, 0, 0, 0);
            if (OUT == kOutU8Dither) {
                const int *tab = kDither128i[y & 7];
                dither = make_int4(tab[dcol], tab[(dcol + 1) & 7], tab[(dcol + 2) & 7], tab[(dcol + 3) & 7]);
                a01 = pack2(fmaf(static_cast<float>(dither.x - 64), 0x1p-30f, v_init), fmaf(static_cast<float>(dither.y - 64), 0x1p-30f, v_init));
                a23 = pack2(fmaf(static_cast<float>(dither.z - 64), 0x1p-30f, v_init), fmaf(static_cast<float>(dither.w - 64), 0x1p-30f, v_init));
            }
            // even and odd taps accumulate separately (two dependency chains per pair instead of one) when there are
            // enough taps for that to matter; the extra addition is counted in the error bound (stream_plan_create)
            f32x2 b01 = 0ull, b23 = 0ull;
            // this row's (g, g) pairs: two per 128-bit load when the row starts on a 16-byte boundary
            f32x2 gg[VT];
            if constexpr (VT % 2 == 0) {
#pragma unroll
                for (int j = 0; j < VT; j += 2) {
                    const ulonglong2 c2 = __ldg(reinterpret_cast<const ulonglong2 *>(gp + j));
                    gg[j] = c2.x; gg[j + 1] = c2.y;
                }
            } else {
#pragma unroll
                for (int j = 0; j < VT; ++j) gg[j] = __ldg(gp + j);
            }
            // The window's rows sit at consecutive ring slots unless it wraps: then every address is one base register
            // plus a constant.
            const int slot0 = first & mask;
            if (slot0 + VT <= mask + 1) {
                const float *w0 = m_col + slot0 * kStripW;
#pragma unroll
                for (int j = 0; j < VT; ++j) {
                    const Quad q = load_quad(w0 + j * kStripW);
                    if (VT >= 4 && (j & 1)) { b01 = fma2(q.lo, gg[j], b01); b23 = fma2(q.hi, gg[j], b23); }
                    else { a01 = fma2(q.lo, gg[j], a01); a23 = fma2(q.hi, gg[j], a23); }
                }
            } else {
#pragma unroll
                for (int j = 0; j < VT; ++j) {
                    const Quad q = load_quad(m_col + ((first + j) & mask) * kStripW);
                    if (VT >= 4 && (j & 1)) { b01 = fma2(q.lo, gg[j], b01); b23 = fma2(q.hi, gg[j], b23); }
                    else { a01 = fma2(q.lo, gg[j], a01); a23 = fma2(q.hi, gg[j], a23); }
                }
            }
            if (VT >= 4) {
                const f32x2 one = pack2(1.0f, 1.0f);
                a01 = fma2(b01, one, a01);
                a23 = fma2(b23, one, a23);
            }
            unsigned word[2];
            float r[4];
            const float worst = finish_quad<OUT>(a01, a23, dst_shift, word, r);
            if (worst >= v_thr) {
                const uint2 fx = exact_fixup_ring<VT, OUT>(m_col, mask, first, op.vint + static_cast<size_t>(y) * VT, v_thr, r[0], r[1], r[2],
                                                           r[3], dither, dst_shift, make_uint2(word[0], word[1]));
                word[0] = fx.x; word[1] = fx.y;
            }
            uint8_t *drow = dst_base + static_cast<ptrdiff_t>(y) * dst_pitch;
            if (fast_store) store_quad_fast<OUT>(drow, xq, word);
            else store_quad_slow<OUT>(drow, xq, dw, dst_xstep, word);
        }
        y_next = y_end;
    }
    (void)dh;
}

using StreamFn = void (*)(const StreamLaunchParams);
// horizontal taps, vertical taps, raw-byte words (0 = float horizontal pass); one kernel per output rule (raw: 8-bit out only)
struct StreamInstance { int ht, vt, nw; StreamFn fn[3]; };
#define TVS_SINST(H, V) {H, V, 0, {stream_scale_kernel<H, V, kOutU8, 0>, stream_scale_kernel<H, V, kOutU8Dither, 0>, stream_scale_kernel<H, V, kOutU10, 0>}}
#define TVS_SRAW(H, V, W) {H, V, W, {stream_scale_kernel<H, V, kOutU8, W>, nullptr, nullptr}}
const StreamInstance kStreamInstances[] = {
    TVS_SINST(2, 2), TVS_SINST(4, 2), TVS_SINST(4, 3),  TVS_SINST(4, 4),  TVS_SINST(6, 4),  TVS_SINST(6, 6),   TVS_SINST(8, 6),
    TVS_SINST(8, 8), TVS_SINST(9, 9), TVS_SINST(12, 9), TVS_SINST(12, 12), TVS_SINST(16, 12),
    // raw-byte horizontal pass: NW words cover phase (<= 3) + the window's bytes
    TVS_SRAW(2, 2, 2), TVS_SRAW(4, 4, 2), TVS_SRAW(4, 4, 3), TVS_SRAW(6, 6, 3), TVS_SRAW(6, 6, 4), TVS_SRAW(9, 9, 3), TVS_SRAW(9, 9, 5),
    TVS_SRAW(12, 12, 4), TVS_SRAW(12, 12, 7),
};
#undef TVS_SINST
#undef TVS_SRAW

// test / profiling aid: TVS_STREAM_FLOAT=1 keeps the float horizontal pass even where the raw-byte one applies
#ifdef TVS_PROFILING
const bool g_stream_force_float = [] { const char *e = getenv("TVS_STREAM_FLOAT"); return e && e[0] == '1'; }();
#else
constexpr bool g_stream_force_float = false;   // production builds read no environment switches
#endif

template <typename T>
size_t append_blob(std::vector<uint8_t> &blob, const std::vector<T> &v) {
    const size_t off = (blob.size() + 255) & ~static_cast<size_t>(255);
    blob.resize(off + v.size() * sizeof(T));
    memcpy(blob.data() + off, v.data(), v.size() * sizeof(T));
    return off;
}

}  // namespace

struct StreamOp {
    int op_index = 0, instance = -1, raw_instance = -1, out_mode = 0, src_mode = 0, h_taps = 0;
    int strips = 0, vtaps = 0, kb = 0, n_blocks = 0, rm_rows = 0;
    int s_pitch = 0, s_chunks = 0, s_chunks_log2 = 0, s_split = 0;
    float v_thr = 0, v_init = 0, src_scale = 1.f, h_addend = 0.f;
    size_t hdense_off = 0, hint_off = 0, hbase_off = 0, vpos_off = 0, vpair_off = 0, vint_off = 0, yend_off = 0;
    std::vector<short> strip_c0, strip_nc;
    size_t smem_bytes = 0;
};

struct StreamPlan {
    std::vector<StreamOp> ops;
    uint8_t *blob = nullptr;
    size_t blob_bytes = 0;
};

namespace {
size_t stream_smem_bytes(int rm_rows, int kb, int s_pitch, int s_chunks) {
    return (static_cast<size_t>(rm_rows) * kStripW + static_cast<size_t>(kb) * s_pitch + 2 * static_cast<size_t>(kb) * s_chunks * 4) * 4;
}
}  // namespace

StreamPlan *stream_plan_create(const Plan &plan, const AxisTable host_tables[4], int device, std::string &err) {
    (void)device;
    std::vector<uint8_t> blob;
    auto sp = new StreamPlan;
    auto bail = [&](const char *why) { err = why; delete sp; return static_cast<StreamPlan *>(nullptr); };
    for (size_t oi = 0; oi < plan.ops.size(); ++oi) {
        const PlaneOp &op = plan.ops[oi];
        if (op.kind != kOpScale) continue;
        if (op.sh <= op.dh) return bail("not a vertical down-scale");
        const AxisTable &h = host_tables[op.htable], &v = host_tables[op.vtable];
        const bool wide_src = op.src_bits > 8;
        const int sample_bytes = wide_src ? 2 : 1;
        if (op.src_xstep != sample_bytes && op.src_xstep != 2 * sample_bytes) return bail("unexpected sample spacing");
        StreamOp so;
        so.op_index = static_cast<int>(oi);
        so.src_mode = wide_src ? (op.src_xstep == 2 ? kSrcU16 : kSrcU16Pair) : (op.src_xstep == 1 ? kSrcU8 : kSrcU8Pair);
        so.out_mode = op.rule == kOut10 ? kOutU10 : op.rule == kOut8Dither ? kOutU8Dither : kOutU8;

        // exactness preconditions of the float horizontal sum (see kernels_fused_impl.cuh)
        long long worst_abs = 0;
        for (int x = 0; x < op.dw; ++x) {
            long long s = 0, a = 0;
            for (int j = 0; j < h.taps; ++j) { const int c = h.coef[static_cast<size_t>(x) * h.taps + j]; s += c; a += std::abs(c); }
            if (s != (1 << 14)) return bail("horizontal row does not sum to one");
            worst_abs = std::max(worst_abs, a);
        }
        if ((wide_src ? 512 : 255) * worst_abs >= (1LL << 24)) return bail("horizontal sum not exact in fp32");
        int inst = -1;
        const int n_inst = static_cast<int>(sizeof(kStreamInstances) / sizeof(kStreamInstances[0]));
        for (int i = 0; i < n_inst; ++i) {
            if (kStreamInstances[i].nw != 0 || kStreamInstances[i].ht < h.taps || kStreamInstances[i].vt < v.taps) continue;
            if (inst < 0 || kStreamInstances[i].ht + kStreamInstances[i].vt < kStreamInstances[inst].ht + kStreamInstances[inst].vt) inst = i;
        }
        if (inst < 0) return bail("no streaming kernel instance for this filter size");
        // raw-byte horizontal pass (8-bit in, plain 8-bit out): the words a window can touch at the worst byte phase
        so.h_taps = h.taps;
        bool fits16 = true;
        for (int c : h.coef) fits16 = fits16 && c >= -32768 && c <= 32767;
        if (!wide_src && so.out_mode == kOutU8 && fits16) {
            const int need_words = (3 + (h.taps - 1) * op.src_xstep + 1 + 3) / 4;
            for (int i = 0; i < n_inst; ++i) {
                const StreamInstance &k = kStreamInstances[i];
                if (k.nw < need_words || k.vt < v.taps || k.ht < h.taps) continue;
                if (so.raw_instance < 0 || k.vt + k.nw < kStreamInstances[so.raw_instance].vt + kStreamInstances[so.raw_instance].nw) so.raw_instance = i;
            }
            // both variants share one vertical layout
            if (so.raw_instance >= 0 && kStreamInstances[so.raw_instance].vt != kStreamInstances[inst].vt) {
                int both = -1;
                for (int i = 0; i < n_inst; ++i) {
                    const StreamInstance &k = kStreamInstances[i];
                    if (k.nw != 0 || k.ht < h.taps || k.vt != kStreamInstances[so.raw_instance].vt) continue;
                    if (both < 0 || k.ht < kStreamInstances[both].ht) both = i;
                }
                if (both >= 0) inst = both; else so.raw_instance = -1;
            }
        }
        const int HT = kStreamInstances[inst].ht, VT = kStreamInstances[inst].vt;
        so.instance = inst;
        so.vtaps = v.taps;
        so.strips = (op.dw + kStripW - 1) / kStripW;
        if (so.strips > kMaxStrips) return bail("plane too wide for the strip table");

        // float vertical sum: residual threshold / accumulator start (tvs_internal.h)
        const VerticalFloatRule vrule = vertical_float_rule_split(v, VT, so.out_mode == kOutU10, so.out_mode == kOutU8Dither);
        if (!vrule.ok) return bail("vertical error bound too large");
        so.v_thr = vrule.thr;
        so.v_init = vrule.init;
        so.src_scale = op.src_shift ? std::ldexp(1.0f, -op.src_shift) : 1.0f;
        so.h_addend = wide_src ? std::ldexp(16384.0f, -149) : 0.0f;

        std::vector<float> hcoef(static_cast<size_t>(op.dw) * HT, 0.0f);
        for (int x = 0; x < op.dw; ++x)
            for (int j = 0; j < h.taps; ++j)
                hcoef[static_cast<size_t>(x) * HT + j] = std::ldexp(static_cast<float>(h.coef[static_cast<size_t>(x) * h.taps + j]), -(23 + op.hshift));
        int ncols_max = 0;
        for (int sx = 0; sx < so.strips; ++sx) {
            const int xa = sx * kStripW, xb = std::min(xa + kStripW, op.dw) - 1;
            so.strip_c0.push_back(static_cast<short>(h.pos[xa]));
            so.strip_nc.push_back(static_cast<short>(h.pos[xb] + HT - h.pos[xa]));
            ncols_max = std::max(ncols_max, h.pos[xb] + HT - h.pos[xa]);
        }
        so.s_chunks = std::max(2, (15 + (ncols_max - 1) * op.src_xstep + sample_bytes - 1) / 16 + 1);   // >= 2: the reciprocal fits 32 bits
        if (so.raw_instance >= 0) so.s_chunks += 2;   // the raw pass reads whole words, up to NW of them past a window's start
        so.s_chunks_log2 = 0;
        while ((1 << so.s_chunks_log2) < so.s_chunks) ++so.s_chunks_log2;
        so.s_pitch = so.s_chunks * (16 / op.src_xstep);
        if ((so.s_pitch & 31) == 0) so.s_pitch += 4;
        if (so.src_mode == kSrcU16 || so.src_mode == kSrcU16Pair) {
            // windows exactly two samples apart (2:1): keep even and odd samples of a staged row apart, so that lanes read
            // consecutive words instead of every other one (2-way bank conflicts otherwise)
            int twos = 0;
            for (int x = 1; x < op.dw; ++x) twos += h.pos[x] - h.pos[x - 1] == 2;
            if (op.dw > 8 && twos >= (op.dw - 1) * 9 / 10) so.s_split = so.s_chunks * (so.src_mode == kSrcU16 ? 4 : 2);
        }

        // vertical tables in tap order, (g, g) pairs and integers
        std::vector<float> vpair(static_cast<size_t>(op.dh) * VT * 2, 0.0f);
        std::vector<int> vint(static_cast<size_t>(op.dh) * VT, 0);
        for (int y = 0; y < op.dh; ++y)
            for (int j = 0; j < v.taps; ++j) {
                const int c = v.coef[static_cast<size_t>(y) * v.taps + j];
                vpair[(static_cast<size_t>(y) * VT + j) * 2] = vpair[(static_cast<size_t>(y) * VT + j) * 2 + 1] = std::ldexp(static_cast<float>(c), 107);
                vint[static_cast<size_t>(y) * VT + j] = c;
            }

        // block size: as many source rows per step as shared memory allows (fewer barriers), and the ring that holds
        // the live window plus one block
        for (int kb = 32;; kb /= 2) {
            const int n_blocks = (op.sh + kb - 1) / kb;
            std::vector<int> yend(n_blocks);
            int y = 0, live = 0;
            for (int b = 0; b < n_blocks; ++b) {
                const int hr_end = std::min((b + 1) * kb, op.sh);
                const int ystart = y;
                while (y < op.dh && (std::min(v.pos[y] + v.taps, op.sh) <= hr_end)) ++y;
                if (b == n_blocks - 1) y = op.dh;
                yend[b] = y;
                // rows alive while block b is being written: from the first row still needed to the end of the block
                const int first_needed = ystart < op.dh ? v.pos[ystart] : hr_end;
                live = std::max(live, b * kb + kb - std::min(first_needed, b * kb));
            }
            int rm = 1;
            while (rm < live + VT) rm *= 2;   // + VT: tap padding may index a little past the window
            so.kb = kb; so.n_blocks = n_blocks; so.rm_rows = rm;
            so.smem_bytes = stream_smem_bytes(rm, kb, so.s_pitch, so.s_chunks);
            if (so.smem_bytes <= kStreamSmemBudget || kb == 4) {
                so.yend_off = append_blob(blob, yend);
                break;
            }
        }
        if (so.smem_bytes > 200 * 1024) return bail("strip does not fit in shared memory");
        so.hdense_off = append_blob(blob, hcoef);
        if (so.raw_instance >= 0) so.hint_off = append_blob(blob, std::vector<short>(h.coef.begin(), h.coef.end()));
        so.hbase_off = append_blob(blob, h.pos);
        so.vpos_off = append_blob(blob, v.pos);
        so.vpair_off = append_blob(blob, vpair);
        so.vint_off = append_blob(blob, vint);
        sp->ops.push_back(so);
    }
    if (sp->ops.empty()) return bail("nothing to scale");
    if (cudaMalloc(&sp->blob, blob.size()) != cudaSuccess ||
        cudaMemcpy(sp->blob, blob.data(), blob.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
        cudaGetLastError();
        if (sp->blob) cudaFree(sp->blob);
        return bail("device allocation for the streaming tables failed");
    }
    sp->blob_bytes = blob.size();
    for (const StreamOp &so : sp->ops) {
        cudaFuncSetAttribute(reinterpret_cast<const void *>(kStreamInstances[so.instance].fn[so.out_mode]),
                             cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (so.raw_instance >= 0)
            cudaFuncSetAttribute(reinterpret_cast<const void *>(kStreamInstances[so.raw_instance].fn[so.out_mode]),
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    return sp;
}

void stream_plan_destroy(StreamPlan *p) {
    if (!p) return;
    if (p->blob) cudaFree(p->blob);
    delete p;
}

int stream_launch(const StreamPlan *p, const Plan &plan, const SampleView src_views[], const SampleView dst_views[], int n_frames,
                  cudaStream_t stream) {
    int launched = 0;
    std::vector<bool> done(p->ops.size(), false);
    // the raw-byte variant needs source rows a multiple of 4 bytes apart (a window's byte phase is then row-independent)
    std::vector<int> chosen(p->ops.size());
    for (size_t i = 0; i < p->ops.size(); ++i) {
        const StreamOp &so = p->ops[i];
        const bool raw_ok = so.raw_instance >= 0 && (src_views[so.op_index].pitch & 3) == 0 && !g_stream_force_float;
        chosen[i] = raw_ok ? so.raw_instance : so.instance;
    }
    for (size_t i = 0; i < p->ops.size(); ++i) {
        if (done[i]) continue;
        const bool raw = kStreamInstances[chosen[i]].nw != 0;
        StreamLaunchParams lp{};
        int strips = 0, m_floats = 0, s_floats = 0, raw_words = 0;
        for (size_t j = i; j < p->ops.size() && lp.n_ops < kMaxStreamOps; ++j) {
            const StreamOp &so = p->ops[j];
            if (done[j] || chosen[j] != chosen[i] || so.out_mode != p->ops[i].out_mode) continue;
            done[j] = true;
            const PlaneOp &op = plan.ops[so.op_index];
            const SampleView &s = src_views[so.op_index], &d = dst_views[so.op_index];
            StreamOpParams &q = lp.op[lp.n_ops++];
            q.src = s.base; q.src_pitch = s.pitch; q.src_fstride = s.frame_stride; q.src_xstep = s.xstep;
            q.dst = d.base; q.dst_pitch = d.pitch; q.dst_fstride = d.frame_stride; q.dst_xstep = d.xstep;
            q.sw = op.sw; q.sh = op.sh; q.dw = op.dw; q.dh = op.dh;
            q.strips = so.strips; q.strip_begin = strips;
            q.hcoef = reinterpret_cast<const float *>(p->blob + so.hdense_off);
            q.hpos = reinterpret_cast<const int *>(p->blob + so.hbase_off);
            q.hcoef_i = reinterpret_cast<const short *>(p->blob + so.hint_off);
            q.h_taps = so.h_taps;
            q.pf_step_rows = kSThreads / so.s_chunks; q.pf_step_chunks = kSThreads % so.s_chunks;
            q.vpos = reinterpret_cast<const int *>(p->blob + so.vpos_off);
            q.vpair = reinterpret_cast<const float *>(p->blob + so.vpair_off);
            q.vint = reinterpret_cast<const int *>(p->blob + so.vint_off);
            q.block_yend = reinterpret_cast<const int *>(p->blob + so.yend_off);
            q.kb = so.kb; q.n_blocks = so.n_blocks; q.rm_mask = so.rm_rows - 1;
            q.s_pitch = so.s_pitch; q.s_chunks = so.s_chunks; q.s_chunks_log2 = so.s_chunks_log2; q.s_split = so.s_split;
            q.chunk_magic = static_cast<unsigned>((0x100000000ull + so.s_chunks - 1) / so.s_chunks);
            q.word_magic = static_cast<unsigned>((0x100000000ull + 4 * so.s_chunks - 1) / (4 * so.s_chunks));
            q.src_lo = s.base - op.src_offset;
            q.src_extent = static_cast<size_t>(op.sh - 1) * static_cast<size_t>(s.pitch) + static_cast<size_t>(op.sw) * op.src_xstep;
            q.v_thr = so.v_thr; q.v_init = so.v_init; q.src_mode = so.src_mode; q.src_scale = so.src_scale; q.h_addend = so.h_addend;
            q.dst_shift = op.dst_shift; q.dither_xoff = op.dither_xoff;
            std::copy(so.strip_c0.begin(), so.strip_c0.end(), q.strip_c0);
            std::copy(so.strip_nc.begin(), so.strip_nc.end(), q.strip_nc);
            strips += so.strips;
            m_floats = std::max(m_floats, so.rm_rows * kStripW);
            if (!raw) s_floats = std::max(s_floats, so.kb * so.s_pitch);
            raw_words = std::max(raw_words, so.kb * so.s_chunks * 4);
        }
        lp.m_floats = m_floats; lp.s_floats = s_floats; lp.raw_words = raw_words;
        const size_t smem = static_cast<size_t>(m_floats + s_floats + 2 * raw_words) * 4;
        if (smem > 200 * 1024) return -2;
        const dim3 grid(strips, n_frames);
        {
            // room for the CTAs the register budget allows, no more: the rest of the 256 KB stays L1 for the tables
            const size_t want = static_cast<size_t>(stream_min_ctas(kStreamInstances[chosen[i]].nw)) * (smem + 1024);
            const int pct = static_cast<int>((want * 100 + (228 << 10) - 1) / (228 << 10));
            cudaFuncSetAttribute(reinterpret_cast<const void *>(kStreamInstances[chosen[i]].fn[p->ops[i].out_mode]),
                                 cudaFuncAttributePreferredSharedMemoryCarveout, std::min(100, std::max(pct, 15)));
        }
        kStreamInstances[chosen[i]].fn[p->ops[i].out_mode]<<<grid, kSThreads, smem, stream>>>(lp);
        if (cudaPeekAtLastError() != cudaSuccess) return -1;
        ++launched;
    }
    return launched;
}

}  // namespace tvs
