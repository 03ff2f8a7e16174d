// Slab kernel: one autonomous warp per 128-column slab.  Runs the up-scales of BASELINE.json (480p -> 1080p, 1080p -> 4K),
// the worker's 480p -> 1080p NV12 target, 16-bit planar down-scales (two columns per lane: 4K 10-bit -> 1080p) and, through
// the split-sample (WIDE) instances, the 15- and 16-bit planes that the packed-RGB input stages produce.
//
// A CTA is ONE warp.  It owns a *slab* - 128 output columns of one plane of one frame, lane L the columns 4L .. 4L+3 - and
// walks it from its first output row to its last, source row by source row:
//
//   every 4 source rows   cp.async the slab's raw rows, three blocks ahead                     -> RAW (warp private)
//                         bytes -> exact floats, one 32-bit word per lane                        -> S   (warp private)
//   every source row      horizontal pass of the lane's 4 columns (dense coefficients in registers for the whole slab);
//                         the 4 results go straight into the lane's vertical register window (the last VT rows)
//   then                  every output row whose window ends at that source row: VT x 4 FMAs from the ring, rounding,
//                         packing, one 4-byte store per lane (a full 128-byte line per warp)             -> HBM
//
// Because a lane filters horizontally exactly the columns it later filters vertically, the intermediate never leaves the
// register file: no shared-memory intermediate, no barrier of any kind after the warp's own __syncwarp around the
// conversion, no tile halo (every source row is filtered once per slab), no ring warm-up per tile.  Warps are
// completely independent, so an SM simply runs as many as fit (about 20 at 96 registers) and they cover each other's
// latencies.  Two earlier forms of this kernel split the passes between a producer and a consumer warp meeting in a
// shared-memory ring; their hand-over (mbarrier try_wait loops, or named barriers of which an SM only has 64) cost
// more than it saved - profiles/r02_slab_history.md keeps the measurements.
//
// Instruction budget (profiles/r02_issue_mix_raw.txt): on sm_100a a packed FFMA2 holds the issue port for 2.19 cycles
// and does not overlap integer work, while scalar FFMA issues in 1.02 and lets IADD / LOP3 slip in beside it, so both
// passes use scalar FFMA.  Arithmetic, exactness argument and output rules are those of DESIGN.md section 4 (the H sum
// is an exact integer, floored by one FMA.RM onto the subnormal grid; the V sum carries a bounded error and samples whose
// residual is within that bound of a rounding boundary are recomputed in int32 from the register ring).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "scale_device.cuh"
#include "tvs_internal.h"

namespace tvs {
namespace {

constexpr int kSlabThreads = 32;   // one warp
constexpr int kBlockRows = 4;      // source rows staged and converted together
constexpr int kRawBlocks = 4;      // raw staging ring in blocks: the warp prefetches kRawBlocks - 1 blocks ahead
constexpr int kMaxSlabOps = 3;
// A fifth source mode, of this kernel only: 16-bit samples of up to 15 significant bits (the planes libswscale's RGB input
// stage makes).  Sample x coefficient no longer fits a float exactly, so each sample is staged as two floats, its low
// byte and its high byte, and the horizontal sum runs once over each (see the WIDE horizontal pass).
constexpr int kSrcU16Split = 4;
// The floor-by-addition constant M of the WIDE pass, in units of the scaled coefficients (2^-(23 + hshift - 8)).  M plus the
// quotient floor(B / 256) must stay inside [2^23, 2^24), where a float's unit is 1, and the high-byte sum A started at -M must
// stay below 2^24 in magnitude: 1.5 * 2^23 for the RGB24 stage's 15-bit samples (high byte <= 127, shift 13), 2^23 + 2^15 for
// RGB48's 16-bit samples (high byte <= 255, shift 15); slab_plan_create checks both against the table's largest row.
template <int WIDE> struct WideMagic { static constexpr float value = WIDE == 15 ? 0x1.01p-7f : 0x1.8p-5f; };
#ifndef TVS_SLAB_MIN_CTAS
#define TVS_SLAB_MIN_CTAS 21
#endif

struct SlabOpParams {
    const uint8_t *src; ptrdiff_t src_pitch, src_fstride; int src_xstep;
    uint8_t *dst; ptrdiff_t dst_pitch, dst_fstride; int dst_xstep;
    int sw, sh, dw, dh;
    int n_groups;          // groups of C adjacent output columns (one group per lane)
    int n_slabs;           // ceil(dw / (32 * C))
    int chunk_rows;        // output rows per work item
    int n_chunks;          // ceil(dh / chunk_rows)
    int item_begin;        // first item of this plane inside a frame's item list
    const float *hdense;   // [n_groups][C][SPAN]  c * 2^-(23+hshift), zero outside a column's own window
    const int *hbase;      // [n_groups]           first source column of the group's span
    const int *vpos;       // [dh + 1]             first source row of each output row (one padding entry)
    int vclass;            // which SlabVClass of the launch holds this plane's vertical tables
    const uint8_t *src_lo; // first byte of the stored source plane (frame 0) and its extent: bounds for 16-byte reads
    size_t src_extent;
    float v_thr, v_init;
    float h_addend;        // subnormal offset of the H result: 16384 * 2^-149 when samples were centred on 512, else 0
    int src_mode;          // kSrcU8 / kSrcU8Pair / kSrcU16 / kSrcU16Pair
    int src_shift;         // P010: the 10 bits sit 6 places up in the word
    int dst_shift, dither_xoff;
    const unsigned char *win_aligned;   // [n_slabs] every lane's window starts a multiple of 4 samples after the slab's first one
};

// Vertical tables of one plane geometry (luma, or the two chroma planes), held in the kernel parameters: the constant
// bank serves them through the uniform datapath, so the coefficients of an output row cost no vector registers, no
// global loads and no latency to speak of.  Up-scale tables repeat (480 -> 1080 has 9 distinct rows plus a few edge
// rows), so the rows are stored once and indexed per output row.
// All classes of a launch share one pool of 32-bit words inside the parameters (kernel parameters may be 32,764 bytes):
//   coef  [n_distinct][VT] floats   g * 2^107 in tap order
//   idx   [dh] uint16               coefficient row of each output row
//   yend  [sh + VT + 1] uint16      yend[r] = output rows whose VT-row window ends at or before source row r
// 480 -> 1080 lanczos needs 11 KB of it, 1080 -> 2160 10 KB; a shape whose tables do not fit runs on the tile kernel.
constexpr int kSlabPoolWords = 7800;
struct SlabVClass { int coef_off, idx_off, yend_off; };   // word offsets into the pool

struct SlabLaunchParams {
    SlabOpParams op[kMaxSlabOps];
    SlabVClass vc[2];
    unsigned pool[kSlabPoolWords];
    int n_ops;
    int items_per_frame, n_items;
    int raw_pitch;         // bytes per staged raw row (multiple of 16); a converted row holds raw_pitch / sample spacing floats
    unsigned chunk_magic;  // ceil(2^32 / (raw_pitch / 16)): exact division of small indices by the chunks per row
};

template <int N>
__device__ __forceinline__ void cp_async_wait_pending() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// One work item: output rows [ya, yb) of slab `slab` of plane `oi` of frame `f`; source rows [r_begin, r_end) feed it
// (rows at or past the plane's height are zero rows the vertical window may touch).
struct SlabItem { int oi, f, slab, ya, yb, r_begin, r_end; };

__device__ __forceinline__ SlabItem decode_item(const SlabLaunchParams &P, int item, int vt) {
    SlabItem it;
    it.f = item / P.items_per_frame;
    int rem = item - it.f * P.items_per_frame;
    it.oi = 0;
    if (P.n_ops > 1 && rem >= P.op[1].item_begin) it.oi = 1;
    if (P.n_ops > 2 && rem >= P.op[2].item_begin) it.oi = 2;
    const SlabOpParams &op = P.op[it.oi];
    rem -= op.item_begin;
    const int chunk = rem / op.n_slabs;            // slabs of one row chunk are consecutive items: neighbours in L2
    it.slab = rem - chunk * op.n_slabs;
    it.ya = chunk * op.chunk_rows;
    it.yb = min(it.ya + op.chunk_rows, op.dh);
    it.r_begin = __ldg(op.vpos + it.ya);
    it.r_end = __ldg(op.vpos + it.yb - 1) + vt;
    return it;
}

// Everything needed to stage one block of source rows, fixed for the life of an item.  Offsets are 32-bit and relative
// to the first byte of the frame's stored plane (a plane the slab kernel accepts is smaller than 1 GiB).
struct SlabStage {
    const uint8_t *lo;        // the stored plane of this frame: nothing outside [lo, lo + extent) is read
    int extent;
    int foot;                 // offset of the first byte of the footprint of the item's first source row
    int pitch;
    int rows_real;            // source rows of the item that exist
    int last_byte;            // offset of the last needed byte inside a footprint row
    int s_chunks;             // 16-byte chunks per staged row
    unsigned chunk_magic;
};

// A chunk at the edge of the plane: 4-byte words that each contain at least one needed byte, zeros elsewhere.  Rare, and
// kept out of line so that the staging loop stays short (the hot code of this kernel has to fit the instruction cache).
__device__ __noinline__ void slab_prefetch_edge(uint4 *dst, const uint8_t *g, int first, int last_byte) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        unsigned *d = reinterpret_cast<unsigned *>(dst) + k;
        if (first + 4 * k <= last_byte && first + 4 * k + 3 >= 0) cp_async4(d, g + 4 * k);
        else *d = 0u;
    }
}

// Start the asynchronous copy of raw source rows [row0, row0 + kBlockRows) of the slab's footprint (rows counted from
// the item's first source row) into `raw_block` ([kBlockRows][s_chunks] 16-byte chunks).  Rows are read as aligned
// chunks starting at the chunk that holds the row's first sample, so byte i of a footprint row lands at position
// mis + i of the staged row, mis (0..15) being the misalignment of the row's first byte; any pitch works.  A chunk that
// would reach outside the plane is fetched as 4-byte words that each contain at least one needed byte; chunks and words
// past the end of the row, and rows below the plane, are zeroed.
template <int C>
__device__ __forceinline__ void slab_prefetch_block(const SlabStage &st, int row0, uint4 *raw_block, int lane) {
    const int total = kBlockRows * st.s_chunks;
    const unsigned lo_mis = static_cast<unsigned>(reinterpret_cast<uintptr_t>(st.lo)) & 15u;
#pragma unroll 1
    for (int idx = lane; idx < total; idx += 32) {
        const int rr = static_cast<int>(__umulhi(static_cast<unsigned>(idx), st.chunk_magic));
        const int ch = idx - rr * st.s_chunks;
        uint4 *dst = raw_block + idx;
        const int row = row0 + rr;
        const int rowo = st.foot + row * st.pitch;                                      // offset of the row's first needed byte
        const int first = 16 * ch - static_cast<int>((lo_mis + static_cast<unsigned>(rowo)) & 15u);   // chunk start, relative to it
        const int go = rowo + first;                                                    // chunk start, relative to the plane
        if (row >= st.rows_real || first > st.last_byte) {
            *dst = make_uint4(0u, 0u, 0u, 0u);
        } else if (go >= 0 && go + 16 <= st.extent) {
            cp_async16(dst, st.lo + go);
        } else if (C == 2) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                unsigned *d = reinterpret_cast<unsigned *>(dst) + k;
                if (first + 4 * k <= st.last_byte && first + 4 * k + 3 >= 0) cp_async4(d, st.lo + go + 4 * k);
                else *d = 0u;
            }
        } else {
            slab_prefetch_edge(dst, st.lo + go, first, st.last_byte);
        }
    }
}

// Raw words of one staged block -> exact float samples.  The converted block mirrors the raw block: raw word i becomes
// floats [i * n, (i + 1) * n) with n = 4 / (bytes between samples), so sample j of a footprint row sits at float index
// (mis >> log2(bytes between samples)) + j of its converted row whatever the alignment.
__device__ __forceinline__ void slab_convert_block_impl(const unsigned *raw, float *S, int words_per_row, int src_mode, int src_shift,
                                                        unsigned mis0, unsigned pmod, int row0, int lane) {
    const int total = kBlockRows * words_per_row;
    // `S` may sit a few floats off its natural alignment (see the aligned-window path of the horizontal pass): vector
    // stores only when the destination of every word is aligned for them
    const unsigned s_mis = static_cast<unsigned>(reinterpret_cast<uintptr_t>(S)) & 15u;
    if (src_mode == kSrcU8) {                       // 4 samples per word
        if (s_mis == 0) {
#pragma unroll 2
            for (int i = lane; i < total; i += 32) {
                const unsigned v = raw[i];
                reinterpret_cast<float4 *>(S)[i] = make_float4(byte_to_float(v, 0x7540u), byte_to_float(v, 0x7541u),
                                                               byte_to_float(v, 0x7542u), byte_to_float(v, 0x7543u));
            }
        } else {
#pragma unroll 2
            for (int i = lane; i < total; i += 32) {
                const unsigned v = raw[i];
                float *o = S + 4 * i;
                o[0] = byte_to_float(v, 0x7540u); o[1] = byte_to_float(v, 0x7541u);
                o[2] = byte_to_float(v, 0x7542u); o[3] = byte_to_float(v, 0x7543u);
            }
        }
    } else if (src_mode == kSrcU16) {               // two 16-bit samples per word, centred on 512
        const unsigned keep = 0xffffu >> src_shift;
        const unsigned mask = keep | (keep << 16);  // P010: integer shift first, as libswscale does; planar words pass unchanged
        if ((s_mis & 7u) == 0) {
#pragma unroll 2
            for (int i = lane; i < total; i += 32) {
                const unsigned v = (raw[i] >> src_shift) & mask;
                reinterpret_cast<float2 *>(S)[i] = make_float2(byte_to_float(v, 0x7410u) - 512.0f, byte_to_float(v, 0x7432u) - 512.0f);
            }
        } else {
#pragma unroll 2
            for (int i = lane; i < total; i += 32) {
                const unsigned v = (raw[i] >> src_shift) & mask;
                S[2 * i] = byte_to_float(v, 0x7410u) - 512.0f;
                S[2 * i + 1] = byte_to_float(v, 0x7432u) - 512.0f;
            }
        }
    } else {
        // interleaved chroma: which byte / half of a word belongs to this plane depends on the row's alignment
#pragma unroll 1
        for (int rr = 0; rr < kBlockRows; ++rr) {
            const unsigned mis = (mis0 + static_cast<unsigned>(row0 + rr) * pmod) & 15u;
            const unsigned *rw = raw + rr * words_per_row;
            if (src_mode == kSrcU8Pair) {
                const unsigned p = mis & 1u;
                float2 *out = reinterpret_cast<float2 *>(S) + rr * words_per_row;
                for (int w = lane; w < words_per_row; w += 32) {
                    const unsigned v = rw[w];
                    out[w] = make_float2(byte_to_float(v, 0x7540u | p), byte_to_float(v, 0x7542u | p));
                }
            } else {
                const unsigned sel = (mis & 2u) ? 0x7432u : 0x7410u;
                const unsigned keep = 0xffffu >> src_shift;
                float *out = S + rr * words_per_row;
                for (int w = lane; w < words_per_row; w += 32) {
                    const unsigned v = (rw[w] >> src_shift) & (keep | (keep << 16));
                    out[w] = byte_to_float(v, sel) - 512.0f;
                }
            }
        }
    }
}

// kSrcU16Split (WIDE instances only; kept apart so that the other instances' code is untouched by it): two 16-bit samples
// per word -> low bytes in S, high bytes in the block behind it.
__device__ __forceinline__ void slab_convert_block_split(const unsigned *raw, float *S, int words_per_row, int lane) {
    const int total = kBlockRows * words_per_row;
    const unsigned s_mis = static_cast<unsigned>(reinterpret_cast<uintptr_t>(S)) & 15u;
    float *hi = S + kBlockRows * (2 * words_per_row);
    if ((s_mis & 7u) == 0) {
#pragma unroll 2
        for (int i = lane; i < total; i += 32) {
            const unsigned v = raw[i];
            reinterpret_cast<float2 *>(S)[i] = make_float2(byte_to_float(v, 0x7540u), byte_to_float(v, 0x7542u));
            reinterpret_cast<float2 *>(hi)[i] = make_float2(byte_to_float(v, 0x7541u), byte_to_float(v, 0x7543u));
        }
    } else {
#pragma unroll 2
        for (int i = lane; i < total; i += 32) {
            const unsigned v = raw[i];
            S[2 * i] = byte_to_float(v, 0x7540u); S[2 * i + 1] = byte_to_float(v, 0x7542u);
            hi[2 * i] = byte_to_float(v, 0x7541u); hi[2 * i + 1] = byte_to_float(v, 0x7543u);
        }
    }
}

__device__ __noinline__ void slab_convert_block_general(const unsigned *raw, float *S, int words_per_row, int src_mode, int src_shift,
                                                        unsigned mis0, unsigned pmod, int row0, int lane) {
    slab_convert_block_impl(raw, S, words_per_row, src_mode, src_shift, mis0, pmod, row0, lane);
}

// What is inline here is hot code, and the hot code of this kernel has to stay small (profiles/r02_slab_history.md).
// The 4-column instances (up-scales, mostly of 8-bit material) keep only the conversion of 8-bit planar sources into a
// naturally aligned block inline and call the rest; the 2-column instances (down-scales, where 10-bit 4K sources are
// the heavy case and a call in the block loop costs registers) inline everything.
template <int C>
__device__ __forceinline__ void slab_convert_block(const unsigned *raw, float *S, int words_per_row, int src_mode, int src_shift,
                                                   unsigned mis0, unsigned pmod, int row0, int lane) {
    if (C == 2) {
        slab_convert_block_impl(raw, S, words_per_row, src_mode, src_shift, mis0, pmod, row0, lane);
    } else if (src_mode == kSrcU8 && (static_cast<unsigned>(reinterpret_cast<uintptr_t>(S)) & 15u) == 0) {
        const int total = kBlockRows * words_per_row;
#pragma unroll 2
        for (int i = lane; i < total; i += 32) {
            const unsigned v = raw[i];
            reinterpret_cast<float4 *>(S)[i] = make_float4(byte_to_float(v, 0x7540u), byte_to_float(v, 0x7541u),
                                                           byte_to_float(v, 0x7542u), byte_to_float(v, 0x7543u));
        }
    } else {
        slab_convert_block_general(raw, S, words_per_row, src_mode, src_shift, mis0, pmod, row0, lane);
    }
}

// The register ring, passed by value to the (rare, out-of-line) exact path.
template <int VT, int C> struct SlabRing { float w[VT][C]; };

// Exact redo of the samples of one quad whose float residual sits inside the error band of a rounding boundary: the
// intermediates are the bit patterns held in the ring, the coefficients the integer copy of the row's record.
template <int VT, int C, int OUT>
__device__ __noinline__ uint2 slab_fixup(SlabRing<VT, C> ring, float4 g0, float4 g1, float4 g2, float v_thr, float4 rr, int4 dither,
                                         int dst_shift, uint2 packed) {
    unsigned word[2] = {packed.x, packed.y};
    const float r[4] = {rr.x, rr.y, rr.z, rr.w};
    const int d[4] = {dither.x, dither.y, dither.z, dither.w};
    int a[4] = {0, 0, 0, 0};
    const float gf[12] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w, g2.x, g2.y, g2.z, g2.w};
#pragma unroll
    for (int s = 0; s < VT; ++s) {
        const int c = static_cast<int>(gf[s] * 0x1p-107f);   // the table holds c * 2^107, exactly
#pragma unroll
        for (int k = 0; k < C; ++k) a[k] += subnormal_to_int(ring.w[s][k]) * c;
    }
#pragma unroll
    for (int k = 0; k < C; ++k) {
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

// Quotient, residual and packing of C float sums (scalar form of finish_quad).  8-bit samples land in the low C bytes
// of word[0]; 10-bit samples two per word.
template <int C, int OUT>
__device__ __forceinline__ float slab_finish(const float (&a)[C], int dst_shift, unsigned (&word)[2], float (&r)[4]) {
    constexpr float kQ = OutRule<OUT>::kShift == 19 ? 0x1p-126f : 0x1p-124f;
    constexpr float kBack = OutRule<OUT>::kShift == 19 ? -0x1p126f : -0x1p124f;
    float q[4] = {0.f, 0.f, 0.f, 0.f};
    r[0] = r[1] = r[2] = r[3] = 0.f;
#pragma unroll
    for (int c = 0; c < C; ++c) q[c] = a[c] * kQ;                    // round(acc / 2^shift) as a subnormal bit pattern
#pragma unroll
    for (int c = 0; c < C; ++c) r[c] = fmaf(q[c], kBack, a[c]);     // exact residual
    if (OUT == kOutU10) {
        const float top = __int_as_float(1023);
        const unsigned b0 = __float_as_uint(fminf(fmaxf(q[0], 0.f), top)) << dst_shift;
        const unsigned b1 = __float_as_uint(fminf(fmaxf(q[1], 0.f), top)) << dst_shift;
        word[0] = __byte_perm(b0, b1, 0x5410);
        word[1] = 0;
        if (C == 4) {
            const unsigned b2 = __float_as_uint(fminf(fmaxf(q[2], 0.f), top)) << dst_shift;
            const unsigned b3 = __float_as_uint(fminf(fmaxf(q[3], 0.f), top)) << dst_shift;
            word[1] = __byte_perm(b2, b3, 0x5410);
        }
    } else {
        word[0] = 0;
        if (C == 4) word[0] = pack_sat_u8(__float_as_int(q[3]), __float_as_int(q[2]), 0u);
        word[0] = pack_sat_u8(__float_as_int(q[1]), __float_as_int(q[0]), word[0]);
        word[1] = 0;
    }
    if (C == 4) return fmaxf(fmaxf(fabsf(r[0]), fabsf(r[1])), fmaxf(fabsf(r[2]), fabsf(r[3])));
    return fmaxf(fabsf(r[0]), fabsf(r[1]));
}

// C samples of one lane: one vector store (the caller has checked alignment and bounds), or bounds-checked element stores
// with `xstep` bytes between samples.
template <int C, int OUT>
__device__ __forceinline__ void slab_store_fast(uint8_t *p, const unsigned (&word)[2]) {
    if (OUT == kOutU10) {
        if (C == 4) *reinterpret_cast<uint2 *>(p) = make_uint2(word[0], word[1]);
        else *reinterpret_cast<unsigned *>(p) = word[0];
    } else {
        if (C == 4) *reinterpret_cast<unsigned *>(p) = word[0];
        else *reinterpret_cast<unsigned short *>(p) = static_cast<unsigned short>(word[0]);
    }
}
template <int C, int OUT>
__device__ __forceinline__ void slab_store_slow(uint8_t *row, int x, int dw, int xstep, const unsigned (&word)[2]) {
#pragma unroll
    for (int k = 0; k < C; ++k) {
        if (x + k >= dw) continue;
        uint8_t *p = row + static_cast<ptrdiff_t>(x + k) * xstep;
        if (OUT == kOutU10) *reinterpret_cast<uint16_t *>(p) = static_cast<uint16_t>(word[k >> 1] >> (16 * (k & 1)));
        else *p = static_cast<uint8_t>(word[0] >> (8 * k));
    }
}

// FAST: the slab is full width, samples are back to back and every row of it starts on a store-sized boundary: one
// vector store per lane and row.  Otherwise bounds-checked element stores (partial slabs, interleaved chroma, odd pitches).
// WIDE: the source is staged as low and high bytes (kSrcU16Split).  With coefficients c scaled by 2^-28 the pass forms
// A = sum(high * c) and B = sum(low * c), both exact, and needs floor((256 A + B) >> 13) = floor((A + floor(B / 256)) / 32):
// floor(B / 256) comes from one FMA rounded down onto 1.5 * 2^23 (where floats are integers), the A chain starts at minus
// that constant so that the two add up exactly, and the last division is the usual multiplication rounded down onto the
// subnormal grid.
template <int C, int SPAN, int VT, int OUT, bool FAST, int WIDE>
__device__ __forceinline__ void slab_run_item(const SlabLaunchParams &P, const SlabItem &it, float *S, uint4 *RAW, int lane) {
    const SlabOpParams &op = P.op[it.oi];
    const float *v_coef = reinterpret_cast<const float *>(P.pool + P.vc[op.vclass].coef_off);
    const unsigned short *v_idx = reinterpret_cast<const unsigned short *>(P.pool + P.vc[op.vclass].idx_off);
    const unsigned short *v_yend = reinterpret_cast<const unsigned short *>(P.pool + P.vc[op.vclass].yend_off) + it.r_begin;

    // ---- horizontal side: this lane's 4 output columns, dense coefficients over their common source span ----
    constexpr int kSlabW = 32 * C;   // output columns per slab: lane * C + k
    const int g_first = it.slab * 32;
    const int grp = min(g_first + lane, op.n_groups - 1);
    float cd[C][SPAN];
    {
        static_assert((C * SPAN) % 4 == 0, "a lane's coefficients are read as 16-byte pieces");
        const float4 *src = reinterpret_cast<const float4 *>(op.hdense + static_cast<size_t>(grp) * (C * SPAN));
        float *flat = &cd[0][0];
#pragma unroll
        for (int i = 0; i < C * SPAN / 4; ++i) {
            const float4 v = __ldg(src + i);
            flat[4 * i] = v.x; flat[4 * i + 1] = v.y; flat[4 * i + 2] = v.z; flat[4 * i + 3] = v.w;
        }
    }
    const int xs = op.src_xstep;
    const int mshift = op.src_mode == kSrcU8 ? 0 : op.src_mode == kSrcU16Pair ? 2 : 1;
    const int rows_total = it.r_end - it.r_begin;                       // source rows the vertical windows will touch
    const int n_blocks = (rows_total + kBlockRows - 1) / kBlockRows;
    const int s_pitch = P.raw_pitch >> mshift;                          // floats per converted row
    const int words_per_row = P.raw_pitch >> 2;
    const int src_mode = op.src_mode, src_shift = op.src_shift;
    SlabStage st;
    int hb;
    {
        const int c0 = __ldg(op.hbase + g_first);
        const int g_last = min(g_first + 31, op.n_groups - 1);
        const int ncols = min(__ldg(op.hbase + g_last) + SPAN, op.sw) - c0;   // samples past the plane's width carry zero weight
        hb = __ldg(op.hbase + grp) - c0;
        st.last_byte = (ncols - 1) * xs + (src_mode >= kSrcU16 ? 1 : 0);
        st.lo = op.src_lo + it.f * op.src_fstride;
        st.extent = static_cast<int>(op.src_extent);
        st.pitch = static_cast<int>(op.src_pitch);
        st.foot = static_cast<int>(op.src - op.src_lo) + it.r_begin * st.pitch + c0 * xs;
        st.rows_real = min(it.r_end, op.sh) - it.r_begin;
        st.s_chunks = P.raw_pitch >> 4;
        st.chunk_magic = P.chunk_magic;
    }
    const int rows_real = st.rows_real;
    const unsigned mis0 = (static_cast<unsigned>(reinterpret_cast<uintptr_t>(st.lo)) + static_cast<unsigned>(st.foot)) & 15u;
    const unsigned pmod = static_cast<unsigned>(st.pitch) & 15u;
    const float h_addend = op.h_addend;
    const float t_max = __int_as_float(32767);
    const int block_chunks = kBlockRows * st.s_chunks;
    // Aligned windows: when the lanes' windows start a multiple of 4 samples apart (exact 2:1, 4:1 ... down-scales, where
    // scalar loads 4 words apart would be 4-way bank conflicts) and every row has the same alignment, the converted block is
    // placed a few floats off so that each window starts on a 16-byte boundary and is read with 128-bit loads.
    const bool aligned_win = C == 2 && pmod == 0 && (src_mode == kSrcU8 || src_mode == (WIDE ? kSrcU16Split : kSrcU16)) && __ldg(op.win_aligned + it.slab) != 0;
    if (aligned_win) S += (4 - static_cast<int>((mis0 >> mshift) & 3u)) & 3;
    const float *s_lane = S + hb;

    // ---- vertical side ----
    const int x = it.slab * kSlabW + C * lane;
    const int dw = op.dw, dst_xstep = op.dst_xstep, dst_shift = op.dst_shift;
    const int dst_pitch = static_cast<int>(op.dst_pitch);
    uint8_t *drow = op.dst + it.f * op.dst_fstride + static_cast<ptrdiff_t>(it.ya) * dst_pitch;
    if (FAST) drow += static_cast<ptrdiff_t>(x) * (OUT == kOutU10 ? 2 : 1);    // this lane's first sample of the row
    const float v_thr = op.v_thr, v_init = op.v_init;
    const int dcol = (x + op.dither_xoff) & 7;
    SlabRing<VT, C> ring;
#pragma unroll
    for (int s = 0; s < VT; ++s) {
#pragma unroll
        for (int c = 0; c < C; ++c) ring.w[s][c] = 0.f;
    }
    int y = it.ya;
    const int yb = it.yb;

    unsigned mis = mis0;
    int row = 0;                    // next source row to filter

    // One source row: the ring moves up by one row (ring.w[0] is the oldest row, ring.w[VT - 1] the newest: one code
    // copy serves every row, at the price of 4 * (VT - 1) register moves), the horizontal pass fills the newest row,
    // then every output row whose window ends here is produced.
    auto row_body = [&](const float *sp) {
#pragma unroll
        for (int s = 0; s + 1 < VT; ++s) {
#pragma unroll
            for (int c = 0; c < C; ++c) ring.w[s][c] = ring.w[s + 1][c];
        }
        float (&o)[C] = ring.w[VT - 1];
        if (row < rows_real) {
            const float *win = sp + (mis >> mshift);
            float px[(SPAN + 3) & ~3];
            if (C == 2 && aligned_win) {
#pragma unroll
                for (int i = 0; i < (SPAN + 3) / 4; ++i) {
                    const float4 v = reinterpret_cast<const float4 *>(win)[i];
                    px[4 * i] = v.x; px[4 * i + 1] = v.y; px[4 * i + 2] = v.z; px[4 * i + 3] = v.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < SPAN; ++i) px[i] = win[i];
            }
            float t[C];
#pragma unroll
            for (int c = 0; c < C; ++c) t[c] = px[0] * cd[c][0];
#pragma unroll
            for (int i = 1; i < SPAN; ++i) {
#pragma unroll
                for (int c = 0; c < C; ++c) t[c] = fmaf(px[i], cd[c][i], t[c]);
            }
            if constexpr (WIDE != 0) {
                constexpr float kWideMagic = WideMagic<WIDE>::value;
                // t holds B (low bytes); now A over the high bytes, started at minus the magic constant
                const int hi_off = kBlockRows * s_pitch;   // the high bytes of a converted block sit this far behind the low bytes
                if (C == 2 && aligned_win) {
#pragma unroll
                    for (int i = 0; i < (SPAN + 3) / 4; ++i) {
                        const float4 v = reinterpret_cast<const float4 *>(win + hi_off)[i];
                        px[4 * i] = v.x; px[4 * i + 1] = v.y; px[4 * i + 2] = v.z; px[4 * i + 3] = v.w;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < SPAN; ++i) px[i] = win[hi_off + i];
                }
                float ta[C];
#pragma unroll
                for (int c = 0; c < C; ++c) ta[c] = fmaf(px[0], cd[c][0], -kWideMagic);
#pragma unroll
                for (int i = 1; i < SPAN; ++i) {
#pragma unroll
                    for (int c = 0; c < C; ++c) ta[c] = fmaf(px[i], cd[c][i], ta[c]);
                }
#pragma unroll
                for (int c = 0; c < C; ++c) t[c] = __fmaf_rd(t[c], 0x1p-8f, kWideMagic) + ta[c];
            }
            // floor(sum >> hshift) lands on the subnormal grid (one multiplication rounded toward minus infinity; for
            // samples centred on 512 the offset is an integer on that grid, so adding it afterwards is exact), clamped
            // above like libswscale's int16 store
#pragma unroll
            for (int c = 0; c < C; ++c) o[c] = fminf(__fmul_rd(t[c], 0x1p-126f) + h_addend, t_max);
        } else {
#pragma unroll
            for (int c = 0; c < C; ++c) o[c] = 0.f;   // a row below the plane: zero weight in every window that reaches it
        }
        mis = (mis + pmod) & 15u;
        ++row;
        const int y_stop = min(static_cast<int>(v_yend[row - 1]), yb);   // rows complete now that this source row is in
#pragma unroll 1
        while (y < y_stop) {
            const float *gp = v_coef + static_cast<int>(v_idx[y]) * VT;
            float g[12] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int s = 0; s < VT; ++s) g[s] = gp[s];
            float a[C];
            int4 dither = make_int4(0, 0, 0, 0);
            if (OUT == kOutU8Dither) {
                const int *tab = kDither128i[y & 7];
                dither = make_int4(tab[dcol], tab[(dcol + 1) & 7], tab[(dcol + 2) & 7], tab[(dcol + 3) & 7]);
                const int dv[4] = {dither.x, dither.y, dither.z, dither.w};
                // (d - 64) * 4096 * 2^-42: the rounding quotient already adds 64 << 12
#pragma unroll
                for (int c = 0; c < C; ++c) a[c] = fmaf(static_cast<float>(dv[c] - 64), 0x1p-30f, v_init);
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c) a[c] = v_init;
            }
#pragma unroll
            for (int s = 0; s < VT; ++s) {
#pragma unroll
                for (int c = 0; c < C; ++c) a[c] = fmaf(ring.w[s][c], g[s], a[c]);
            }
            unsigned word[2];
            float r[4];
            const float worst = slab_finish<C, OUT>(a, dst_shift, word, r);
            if (worst >= v_thr) {
                const uint2 fx = slab_fixup<VT, C, OUT>(ring, make_float4(g[0], g[1], g[2], g[3]), make_float4(g[4], g[5], g[6], g[7]),
                                                        make_float4(g[8], g[9], g[10], g[11]), v_thr, make_float4(r[0], r[1], r[2], r[3]),
                                                        dither, dst_shift, make_uint2(word[0], word[1]));
                word[0] = fx.x; word[1] = fx.y;
            }
            if (FAST) slab_store_fast<C, OUT>(drow, word);
            else slab_store_slow<C, OUT>(drow, x, dw, dst_xstep, word);
            drow += dst_pitch;
            ++y;
        }
    };

    // prologue: the first kRawBlocks - 1 blocks are on their way before anything is converted
#pragma unroll 1
    for (int b = 0; b < kRawBlocks - 1; ++b) {
        if (b < n_blocks) slab_prefetch_block<C>(st, b * kBlockRows, RAW + b * block_chunks, lane);
        cp_async_commit();
    }
#pragma unroll 1
    for (int b = 0; b < n_blocks; ++b) {
        cp_async_wait_pending<kRawBlocks - 2>();   // this lane's copies of block b have landed ...
        __syncwarp();                              // ... and so have everybody else's; the previous block's S reads are done
        if constexpr (WIDE != 0)
            slab_convert_block_split(reinterpret_cast<const unsigned *>(RAW + (b % kRawBlocks) * block_chunks), S, words_per_row, lane);
        else
            slab_convert_block<C>(reinterpret_cast<const unsigned *>(RAW + (b % kRawBlocks) * block_chunks), S, words_per_row, src_mode,
                               src_shift, mis0, pmod, b * kBlockRows, lane);
        __syncwarp();                              // S is complete; the raw block is free again
        if (b + kRawBlocks - 1 < n_blocks)
            slab_prefetch_block<C>(st, (b + kRawBlocks - 1) * kBlockRows, RAW + ((b + kRawBlocks - 1) % kRawBlocks) * block_chunks, lane);
        cp_async_commit();

        const int rows_here = min(kBlockRows, rows_total - row);
        const float *sp = s_lane;
#pragma unroll 1
        for (int rr = 0; rr < rows_here; ++rr, sp += s_pitch) {
            row_body(sp);
        }
    }
    cp_async_wait_pending<0>();
}

// Resident warps per SM the register budget is set for: 21 (80 registers) for short filters, 18 (96 registers) once the
// horizontal coefficients and the vertical window together need more (measured: profiles/r02_slab_history.md).
template <int C, int SPAN, int VT> struct SlabBudget { static constexpr int kMinCtas = SPAN + VT >= 26 ? 16 : C * (SPAN + VT) >= 52 ? 18 : TVS_SLAB_MIN_CTAS; };

template <int C, int SPAN, int VT, int OUT, int WIDE>
__global__ void __launch_bounds__(kSlabThreads, SlabBudget<C, SPAN, VT>::kMinCtas)
slab_scale_kernel(const __grid_constant__ SlabLaunchParams P) {
    constexpr int kSlabW = 32 * C;
    extern __shared__ __align__(16) float smem[];
    float *S = smem;                                                       // [kBlockRows][raw_pitch] floats at most, + a zero tail
    const int s_floats = kBlockRows * P.raw_pitch + 32;
    uint4 *RAW = reinterpret_cast<uint4 *>(S + s_floats);                  // [kRawBlocks][kBlockRows][raw_pitch / 16] chunks
    const int lane = threadIdx.x;
    // converted rows are read a little past their last written sample (dense windows of the slab's last columns run
    // into the next row, the last row's into the tail): those floats must be finite
    for (int i = lane; i < s_floats; i += kSlabThreads) S[i] = 0.f;
    __syncwarp();

#pragma unroll 1
    for (int item = blockIdx.x; item < P.n_items; item += gridDim.x) {
        const SlabItem it = decode_item(P, item, VT);
        const SlabOpParams &op = P.op[it.oi];
        constexpr int kSample = OUT == kOutU10 ? 2 : 1, kAlign = C * kSample - 1;
        const uintptr_t d0 = reinterpret_cast<uintptr_t>(op.dst + it.f * op.dst_fstride) + static_cast<uintptr_t>(it.slab * kSlabW) * kSample;
        const bool fast = op.dst_xstep == kSample && (op.dst_pitch & kAlign) == 0 && (d0 & kAlign) == 0 && (it.slab + 1) * kSlabW <= op.dw;
        if (fast) slab_run_item<C, SPAN, VT, OUT, true, WIDE>(P, it, S, RAW, lane);
        else slab_run_item<C, SPAN, VT, OUT, false, WIDE>(P, it, S, RAW, lane);
        __syncwarp();
    }
}

using SlabKernelFn = void (*)(const SlabLaunchParams);
// `cols` adjacent output columns per lane (a slab is 32 * cols columns wide): 4 wherever the dense horizontal window of
// four neighbours stays short (up-scales, mild down-scales), 2 for stronger down-scales, whose windows are long and far apart.
// one kernel per output rule; fn_wide13: RGB24 / BGR24 planes (never dithered), fn_wide15: RGB48 planes (8-bit output is dithered)
struct SlabInstance { int cols, span, vt; SlabKernelFn fn[3], fn_wide13[3], fn_wide15[3]; };
#define TVS_SLAB_INST(C, S, V) {C, S, V, {slab_scale_kernel<C, S, V, kOutU8, 0>, slab_scale_kernel<C, S, V, kOutU8Dither, 0>, slab_scale_kernel<C, S, V, kOutU10, 0>}, \
                                {slab_scale_kernel<C, S, V, kOutU8, 13>, nullptr, slab_scale_kernel<C, S, V, kOutU10, 13>}, \
                                {nullptr, slab_scale_kernel<C, S, V, kOutU8Dither, 15>, slab_scale_kernel<C, S, V, kOutU10, 15>}}
const SlabInstance kSlabInstances[] = {
    TVS_SLAB_INST(4, 4, 2), TVS_SLAB_INST(4, 6, 4), TVS_SLAB_INST(4, 8, 4), TVS_SLAB_INST(4, 8, 6), TVS_SLAB_INST(4, 12, 6), TVS_SLAB_INST(4, 12, 8),
    TVS_SLAB_INST(2, 6, 4), TVS_SLAB_INST(2, 8, 6), TVS_SLAB_INST(2, 10, 8), TVS_SLAB_INST(2, 12, 10), TVS_SLAB_INST(2, 14, 12),
};
#undef TVS_SLAB_INST

template <typename T>
size_t slab_append(std::vector<uint8_t> &blob, const std::vector<T> &v) {
    const size_t off = (blob.size() + 255) & ~static_cast<size_t>(255);
    blob.resize(off + v.size() * sizeof(T));
    memcpy(blob.data() + off, v.data(), v.size() * sizeof(T));
    return off;
}

}  // namespace

struct SlabOp {
    int op_index = 0, instance = -1, out_mode = 0, src_mode = 0;
    int n_groups = 0, n_slabs = 0;
    int s_pitch = 0, raw_pitch = 0;
    float v_thr = 0, v_init = 0, h_addend = 0;
    size_t win_aligned_off = 0;
    size_t hdense_off = 0, hbase_off = 0, vpos_off = 0;
    int vtable = 0;                         // which vertical table of the context (2 luma, 3 chroma): planes that share it share a class
    std::vector<float> vcoef;               // [n_distinct][VT] distinct coefficient rows, c * 2^107 in tap order
    std::vector<unsigned short> vidx;       // [dh]             coefficient row of each output row
    std::vector<unsigned short> vyend;      // [sh + VT + 1]    see SlabLaunchParams
    size_t pool_words() const { return vcoef.size() + (vidx.size() + 1) / 2 + (vyend.size() + 1) / 2; }
};

struct SlabPlan {
    std::vector<SlabOp> ops;
    uint8_t *blob = nullptr;
    size_t blob_bytes = 0;
    int device = 0, sm_count = 148;
};

// The float vertical sum of this kernel runs in tap order, one scalar FMA per tap: follow that order row by row (each FMA rounds by at most half an ulp of its own partial sum, bounded through
// |t| <= 32767) instead of charging every tap half an ulp of the largest possible sum.
static VerticalFloatRule slab_vertical_rule(const AxisTable &v, int vt_inst, bool out10, bool dither) {
    const VerticalFloatRule coarse = vertical_float_rule(v, vt_inst, out10, dither);
    if (!coarse.ok || coarse.thr > 1.0e38f) return coarse;   // exact tables keep their tie-free treatment
    auto half_ulp = [](double bound) { return bound < 1.0 ? 0.0 : 0.5 * std::ldexp(1.0, static_cast<int>(std::floor(std::log2(bound))) + 1 - 24); };
    const double init_abs = dither ? 64.0 * 4096.0 : 0.0;
    double e_max = 0.0;
    std::vector<int> by_slot(vt_inst);
    for (int y = 0; y < v.length; ++y) {
        std::fill(by_slot.begin(), by_slot.end(), 0);
        for (int j = 0; j < v.taps; ++j) by_slot[j] = std::abs(static_cast<int>(v.coef[static_cast<size_t>(y) * v.taps + j]));
        double run = init_abs, e = 0.0;
        for (int s = 0; s < vt_inst; ++s) {
            if (!by_slot[s]) continue;   // a zero coefficient leaves the accumulator unchanged, exactly
            run += 32767.0 * by_slot[s];
            e += half_ulp(run * (1.0 + 1e-6));
        }
        e_max = std::max(e_max, e);
    }
    const double half_step = out10 ? 65536.0 : 262144.0;
    if (e_max > half_step / 8) return {false, 0.f, 0.f};
    return {true, static_cast<float>(std::ldexp(half_step - e_max - 1.0, -42)), coarse.init};
}

SlabPlan *slab_plan_create(const Plan &plan, const AxisTable host_tables[4], int device, std::string &err) {
    std::vector<uint8_t> blob;
    auto sp = new SlabPlan;
    sp->device = device;
    cudaDeviceGetAttribute(&sp->sm_count, cudaDevAttrMultiProcessorCount, device);
    for (size_t oi = 0; oi < plan.ops.size(); ++oi) {
        const PlaneOp &op = plan.ops[oi];
        if (op.kind != kOpScale) continue;
        const AxisTable &h = host_tables[op.htable], &v = host_tables[op.vtable];
        const bool wide_src = op.src_bits > 8;
        const int sample_bytes = wide_src ? 2 : 1;
        if (op.src_xstep != sample_bytes && op.src_xstep != 2 * sample_bytes) { err = "unexpected sample spacing"; delete sp; return nullptr; }
        // the RGB input stages' planes: 15-bit samples shifted by 13 (rgb24 / bgr24), 16-bit samples shifted by 15 (rgb48)
        const bool split = op.src_bits == 16 && op.src_xstep == 2 &&
                           ((op.hshift == 13 && op.rule != kOut8Dither) || (op.hshift == 15 && op.rule != kOut8));
        if (op.src_bits > 10 && !split) { err = "samples wider than 10 bits"; delete sp; return nullptr; }
        if (static_cast<long long>(op.sh) * op.sw * op.src_xstep >= (1LL << 30)) { err = "plane too large for 32-bit offsets"; delete sp; return nullptr; }
        const int src_mode = split ? kSrcU16Split : wide_src ? (op.src_xstep == 2 ? kSrcU16 : kSrcU16Pair) : (op.src_xstep == 1 ? kSrcU8 : kSrcU8Pair);
        const int out_mode = op.rule == kOut10 ? kOutU10 : op.rule == kOut8Dither ? kOutU8Dither : kOutU8;
        for (int y = 1; y < op.dh; ++y)
            if (v.pos[y] < v.pos[y - 1]) { err = "vertical positions go backwards"; delete sp; return nullptr; }
        // float exactness of the horizontal sum (see the tile kernel)
        long long worst_abs = 0;
        for (int x = 0; x < op.dw; ++x) {
            long long s = 0, a = 0;
            for (int j = 0; j < h.taps; ++j) { const int c = h.coef[static_cast<size_t>(x) * h.taps + j]; s += c; a += std::abs(c); }
            if (s != (1 << 14)) { err = "horizontal row does not sum to one"; delete sp; return nullptr; }
            worst_abs = std::max(worst_abs, a);
        }
        if (split) {
            // low bytes: |B| <= 255 * worst_abs, and M + floor(B / 256) must stay inside [2^23, 2^24); high bytes (<= hi_max) on
            // top of the start value -M must stay below 2^24 (see WideMagic)
            const long long hi_max = op.hshift == 15 ? 255 : 127, magic = op.hshift == 15 ? (1LL << 23) + (1LL << 15) : (3LL << 22);
            const long long bq = 255 * worst_abs / 256 + 1;
            if (255 * worst_abs >= (1LL << 24) || magic - bq < (1LL << 23) || magic + bq >= (1LL << 24) ||
                hi_max * worst_abs + magic + bq >= (1LL << 24)) {
                err = "horizontal sum not exact in fp32"; delete sp; return nullptr;
            }
        } else if ((wide_src ? 512 : 255) * worst_abs >= (1LL << 24)) { err = "horizontal sum not exact in fp32"; delete sp; return nullptr; }
        // smallest instance whose dense window covers every group of `cols` neighbouring columns and whose vertical
        // window covers the taps; four columns per lane are preferred (less staging and bookkeeping per sample)
        int inst = -1;
        for (size_t i = 0; i < sizeof(kSlabInstances) / sizeof(kSlabInstances[0]); ++i) {
            const int cols = kSlabInstances[i].cols;
            if (kSlabInstances[i].vt < v.taps) continue;
            int span = 0;
            for (int xa = 0; xa < op.dw; xa += cols) span = std::max(span, h.pos[std::min(xa + cols, op.dw) - 1] + h.taps - h.pos[xa]);
            if (kSlabInstances[i].span < span) continue;
            auto cost = [](const SlabInstance &k) { return (k.cols == 4 ? 0 : 1000) + k.cols * (k.span + k.vt); };
            if (inst < 0 || cost(kSlabInstances[i]) < cost(kSlabInstances[inst])) inst = static_cast<int>(i);
        }
        if (inst < 0) { err = "no slab kernel instance for this filter size"; delete sp; return nullptr; }
        const int C = kSlabInstances[inst].cols, SPAN = kSlabInstances[inst].span, VT = kSlabInstances[inst].vt;
        const int n_groups = (op.dw + C - 1) / C;
        const VerticalFloatRule vrule = slab_vertical_rule(v, VT, out_mode == kOutU10, out_mode == kOutU8Dither);
        if (!vrule.ok) { err = "vertical error bound too large"; delete sp; return nullptr; }

        SlabOp so;
        so.op_index = static_cast<int>(oi);
        so.instance = inst;
        so.out_mode = out_mode;
        so.src_mode = src_mode;
        so.n_groups = n_groups;
        so.n_slabs = (op.dw + 32 * C - 1) / (32 * C);
        so.v_thr = vrule.thr;
        so.v_init = vrule.init;
        so.h_addend = wide_src && !split ? std::ldexp(16384.0f, -149) : 0.0f;   // (512 << 14) >> 9, as a subnormal
        const int coef_shift = split ? op.hshift - 8 : op.hshift;   // split samples: the high byte already carries a factor 256

        std::vector<float> hdense(static_cast<size_t>(n_groups) * C * SPAN, 0.0f);
        std::vector<int> hbase(n_groups);
        for (int g = 0; g < n_groups; ++g) {
            const int xa = g * C;
            hbase[g] = h.pos[xa];
            for (int k = 0; k < C && xa + k < op.dw; ++k) {
                const int x = xa + k;
                for (int j = 0; j < h.taps; ++j) {
                    const int c = h.coef[static_cast<size_t>(x) * h.taps + j];
                    // scaled so that one FMA by 2^-126 lands floor(sum >> hshift) on the subnormal grid
                    if (c) hdense[(static_cast<size_t>(g) * C + k) * SPAN + (h.pos[x] - hbase[g] + j)] = std::ldexp(static_cast<float>(c), -(23 + coef_shift));
                }
            }
        }
        std::vector<unsigned char> win_aligned(so.n_slabs, 1);
        for (int g = 0; g < n_groups; ++g)
            if ((hbase[g] - hbase[g / 32 * 32]) % 4 != 0) win_aligned[g / 32] = 0;
        // widest slab footprint -> staged row sizes (worst alignment)
        int ncols_max = 0;
        for (int s = 0; s < so.n_slabs; ++s) {
            const int ga = s * 32, gb = std::min(ga + 31, n_groups - 1);
            ncols_max = std::max(ncols_max, std::min(hbase[gb] + SPAN, op.sw) - hbase[ga]);
        }
        // at least two chunks: the kernel divides by the chunk count with a 32-bit reciprocal, and ceil(2^32 / 1) does not fit
        // (found by tools/gpu_fuzz.py on one-column chroma planes, whose rows all landed in row 0 of the block)
        const int s_chunks = std::max(2, (15 + (ncols_max - 1) * op.src_xstep + sample_bytes - 1) / 16 + 1);
        so.raw_pitch = 16 * s_chunks;
        // a converted row holds (16 / spacing) samples per chunk, and a lane's window may start up to SPAN short of its end
        so.s_pitch = ((s_chunks * (16 / op.src_xstep) + SPAN + 3) & ~3) + 4;

        std::vector<int> vpos(op.dh + 1, v.pos[op.dh - 1]);
        for (int y = 0; y < op.dh; ++y) vpos[y] = v.pos[y];
        // vertical tables for the kernel parameters: distinct coefficient rows, a row index per output row, and per
        // source row the number of output rows whose VT-row window [pos, pos + VT) ends at or before it
        if (op.dh > 65535) { err = "plane taller than the slab kernel's tables"; delete sp; return nullptr; }
        so.vtable = op.vtable;
        so.vidx.resize(op.dh);
        for (int y = 0; y < op.dh; ++y) {
            float row[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};   // VT <= 12
            for (int j = 0; j < v.taps; ++j) row[j] = std::ldexp(static_cast<float>(v.coef[static_cast<size_t>(y) * v.taps + j]), 107);
            int id = -1;
            for (size_t k = 0; k + VT <= so.vcoef.size() && id < 0; k += VT)
                if (memcmp(&so.vcoef[k], row, VT * sizeof(float)) == 0) id = static_cast<int>(k / VT);
            if (id < 0) {
                id = static_cast<int>(so.vcoef.size() / VT);
                so.vcoef.insert(so.vcoef.end(), row, row + VT);
                if (so.vcoef.size() > static_cast<size_t>(kSlabPoolWords)) { err = "vertical tables do not fit in the kernel parameters"; delete sp; return nullptr; }
            }
            so.vidx[y] = static_cast<unsigned short>(id);
        }
        so.vyend.assign(op.sh + VT + 1, static_cast<unsigned short>(op.dh));
        {
            int y = 0;
            for (int r = 0; r < op.sh + VT + 1; ++r) {
                while (y < op.dh && v.pos[y] + VT - 1 <= r) ++y;
                so.vyend[r] = static_cast<unsigned short>(y);
            }
        }
        so.hdense_off = slab_append(blob, hdense);
        so.win_aligned_off = slab_append(blob, win_aligned);
        so.hbase_off = slab_append(blob, hbase);
        so.vpos_off = slab_append(blob, vpos);
        sp->ops.push_back(so);
    }
    if (sp->ops.empty()) { err = "nothing to scale"; delete sp; return nullptr; }
    {
        // the vertical tables of all planes (one set per distinct table) must fit in the parameter pool together
        size_t words = 0;
        std::vector<int> seen;
        for (const SlabOp &so : sp->ops) {
            if (std::find(seen.begin(), seen.end(), so.vtable) != seen.end()) continue;
            seen.push_back(so.vtable);
            words += so.pool_words();
        }
        if (seen.size() > 2 || words > static_cast<size_t>(kSlabPoolWords)) { err = "vertical tables do not fit in the kernel parameters"; delete sp; return nullptr; }
    }
    if (cudaMalloc(&sp->blob, blob.size()) != cudaSuccess ||
        cudaMemcpy(sp->blob, blob.data(), blob.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
        err = "device allocation for the slab tables failed";
        cudaGetLastError();
        if (sp->blob) cudaFree(sp->blob);
        delete sp;
        return nullptr;
    }
    sp->blob_bytes = blob.size();
    return sp;
}

void slab_plan_destroy(SlabPlan *p) {
    if (!p) return;
    if (p->blob) cudaFree(p->blob);
    delete p;
}

int slab_launch(const SlabPlan *p, const Plan &plan, const SampleView src_views[], const SampleView dst_views[], int n_frames,
                cudaStream_t stream) {
    int launched = 0;
    std::vector<bool> done(p->ops.size(), false);
    for (size_t i = 0; i < p->ops.size(); ++i) {
        if (done[i]) continue;
        // one launch covers every plane that runs on the same kernel instance
        SlabLaunchParams lp{};
        std::vector<size_t> members;
        for (size_t j = i; j < p->ops.size() && members.size() < kMaxSlabOps; ++j)
            if (!done[j] && p->ops[j].instance == p->ops[i].instance && p->ops[j].out_mode == p->ops[i].out_mode &&
                (p->ops[j].src_mode == kSrcU16Split) == (p->ops[i].src_mode == kSrcU16Split)) { done[j] = true; members.push_back(j); }
        // Row chunks: every item should cost about the same (all slabs are 128 columns wide, so cost ~ output rows) and
        // there should be several items per resident CTA; a chunk is never shorter than 64 rows (ring warm-up, tables).
        int min_dh = 1 << 30;
        for (size_t j : members) min_dh = std::min(min_dh, plan.ops[p->ops[j].op_index].dh);
        int chunk_rows = std::max(64, std::min(min_dh, 1024));
        const SlabInstance &inst = kSlabInstances[p->ops[i].instance];
        const SlabKernelFn fn = p->ops[i].src_mode != kSrcU16Split ? inst.fn[p->ops[i].out_mode]
                              : plan.ops[p->ops[i].op_index].hshift == 15 ? inst.fn_wide15[p->ops[i].out_mode] : inst.fn_wide13[p->ops[i].out_mode];
        if (!fn) return -1;
        size_t smem = 0;
        {
            int raw_pitch = 0;
            for (size_t j : members) raw_pitch = std::max(raw_pitch, p->ops[j].raw_pitch);
            smem = (static_cast<size_t>(kBlockRows) * raw_pitch + 32) * 4 + static_cast<size_t>(kRawBlocks) * kBlockRows * raw_pitch;
        }
        if (smem > 48 * 1024) cudaFuncSetAttribute(reinterpret_cast<const void *>(fn), cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        // Some twenty one-warp CTAs share an SM.  Ask for a shared-memory carve-out that holds them (the default sizes it
        // for far fewer) but no more: the rest of the 256 KB stays L1, which serves the vertical tables and the source rows.
        {
            const size_t want = 24 * (smem + 1024);
            int pct = static_cast<int>((want * 100 + (228 << 10) - 1) / (228 << 10));
            cudaFuncSetAttribute(reinterpret_cast<const void *>(fn), cudaFuncAttributePreferredSharedMemoryCarveout, std::min(100, std::max(pct, 15)));
        }
        int resident = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, fn, kSlabThreads, smem) != cudaSuccess || resident < 1) return -1;
        const long long slots = static_cast<long long>(resident) * p->sm_count;
        for (;;) {
            long long items = 0;
            for (size_t j : members) {
                const PlaneOp &op = plan.ops[p->ops[j].op_index];
                items += static_cast<long long>(p->ops[j].n_slabs) * ((op.dh + chunk_rows - 1) / chunk_rows);
            }
            if (items * n_frames >= 4 * slots || chunk_rows <= 64) break;
            chunk_rows = std::max(64, chunk_rows / 2);
        }
        int items = 0;
        int n_classes = 0, class_vtable[2] = {-1, -1};
        size_t pool_used = 0;
        for (size_t j : members) {
            const SlabOp &so = p->ops[j];
            const PlaneOp &op = plan.ops[so.op_index];
            const SampleView &s = src_views[so.op_index], &d = dst_views[so.op_index];
            SlabOpParams &q = lp.op[lp.n_ops++];
            q.src = s.base; q.src_pitch = s.pitch; q.src_fstride = s.frame_stride; q.src_xstep = s.xstep;
            q.dst = d.base; q.dst_pitch = d.pitch; q.dst_fstride = d.frame_stride; q.dst_xstep = d.xstep;
            q.sw = op.sw; q.sh = op.sh; q.dw = op.dw; q.dh = op.dh;
            q.n_groups = so.n_groups; q.n_slabs = so.n_slabs;
            const int n_chunks = (op.dh + chunk_rows - 1) / chunk_rows;
            q.chunk_rows = (op.dh + n_chunks - 1) / n_chunks;     // equal chunks
            q.n_chunks = (op.dh + q.chunk_rows - 1) / q.chunk_rows;
            q.item_begin = items;
            items += q.n_slabs * q.n_chunks;
            q.hdense = reinterpret_cast<const float *>(p->blob + so.hdense_off);
            q.hbase = reinterpret_cast<const int *>(p->blob + so.hbase_off);
            q.vpos = reinterpret_cast<const int *>(p->blob + so.vpos_off);
            // planes with the same vertical table (the two chroma planes) share a class
            int cls = -1;
            for (int k = 0; k < n_classes; ++k) if (class_vtable[k] == so.vtable) cls = k;
            if (cls < 0) {
                if (n_classes == 2) return -1;
                cls = n_classes++;
                class_vtable[cls] = so.vtable;
                SlabVClass &vc = lp.vc[cls];
                if (pool_used + so.pool_words() > static_cast<size_t>(kSlabPoolWords)) return -1;
                vc.coef_off = static_cast<int>(pool_used);
                memcpy(lp.pool + pool_used, so.vcoef.data(), so.vcoef.size() * sizeof(float));
                pool_used += so.vcoef.size();
                vc.idx_off = static_cast<int>(pool_used);
                memcpy(lp.pool + pool_used, so.vidx.data(), so.vidx.size() * sizeof(unsigned short));
                pool_used += (so.vidx.size() + 1) / 2;
                vc.yend_off = static_cast<int>(pool_used);
                memcpy(lp.pool + pool_used, so.vyend.data(), so.vyend.size() * sizeof(unsigned short));
                pool_used += (so.vyend.size() + 1) / 2;
            }
            q.vclass = cls;
            q.src_lo = s.base - op.src_offset;
            q.src_extent = static_cast<size_t>(op.sh - 1) * static_cast<size_t>(s.pitch) + static_cast<size_t>(op.sw) * op.src_xstep;
            q.v_thr = so.v_thr; q.v_init = so.v_init; q.h_addend = so.h_addend;
            q.src_mode = so.src_mode; q.src_shift = op.src_shift;
            q.dst_shift = op.dst_shift; q.dither_xoff = op.dither_xoff;
            q.win_aligned = p->blob + so.win_aligned_off;
            lp.raw_pitch = std::max(lp.raw_pitch, so.raw_pitch);   // every plane of the launch stages the launch's chunk count per row
        }
        lp.chunk_magic = static_cast<unsigned>((0x100000000ULL + (lp.raw_pitch >> 4) - 1) / (lp.raw_pitch >> 4));
        lp.items_per_frame = items;
        const long long total = static_cast<long long>(items) * n_frames;
        if (total > 0x7fffffffLL) return -1;
        lp.n_items = static_cast<int>(total);
        const int grid = static_cast<int>(std::min<long long>(total, slots));
        fn<<<grid, kSlabThreads, smem, stream>>>(lp);
        if (cudaPeekAtLastError() != cudaSuccess) return -1;
        ++launched;
    }
    return launched;
}

size_t slab_plan_device_bytes(const SlabPlan *p) { return p ? p->blob_bytes : 0; }

}  // namespace tvs
