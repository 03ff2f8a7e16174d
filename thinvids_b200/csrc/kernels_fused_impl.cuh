// Fused tile kernel: stage the source footprint of one output tile in shared memory, run libswscale's
// horizontal pass into a shared-memory intermediate, then the vertical pass straight to global memory.
// HBM sees each source byte once (plus the tile halo, which L2 absorbs) and each destination byte once.
//
// Arithmetic is bit-exact with SURVEY.md A.3 but runs on the FP32 FMA pipe, which on sm_100a issues twice
// as many multiply-adds per clock as IMAD/IDP (profiles/r01_alu_rates.md):
//   * H pass: pixels are exact small floats and coefficients are c * 2^-30, so the 6..24-term sum is an exact
//     integer * 2^-30 (|sum| < 2^24).  One FMA with round-toward-minus-infinity by 2^-126 then yields
//     floor(sum / 128) * 2^-149: a SUBNORMAL float whose bit pattern is the integer t itself (sign-magnitude
//     for negative t).  FMNMX against the pattern 32767 is the upper clamp.  Subnormal operands run at full
//     rate on this chip (profiles/r01_denorm_probe.txt).
//   * V pass: acc' = sum t_d * (g * 2^107) = acc * 2^-42 accumulates with at most `e` absolute rounding error
//     (|acc| < 2^28).  q = RN(acc' * 2^-126) is round(acc / 2^19) * 2^-149, again an integer bit pattern, which
//     cvt.pack.sat.u8.s32 clamps and packs.  The residual r = acc' - q * 2^126 is exact; when |r| is within `e`
//     of the rounding boundary (about 0.02 % of samples) the sample is recomputed in int32, so the result is
//     exact by construction, not by luck.  RN ties are inside that band.
// Tile geometry: 256 output columns (two halves of 128; a lane owns 4 adjacent columns of each half) by `th`
// output rows, 256 threads.  In the H pass warps 0-3 / 4-7 take the left / right half and a lane keeps the dense
// coefficients of its 4 columns over their common source span in registers for the life of the tile.  In the
// V pass each warp walks down a strip of th/8 rows with a register ring of VT intermediate rows, producing
// 8 samples per lane per row; row coefficients and start rows come from a shared-memory copy.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "scale_device.cuh"
#include "tvs_internal.h"

namespace tvs {
namespace {

// This file is compiled twice (kernels_fused_w256.cu, kernels_fused_w128.cu) with TVS_HALVES = 2 and 1: 256-column
// tiles with 2 CTAs of 256 threads per SM (best for up-scales: 8 samples per lane per row in the vertical pass) and
// 128-column tiles with 4 CTAs of 128 threads per SM (best for down-scales, whose source footprint is what fills
// shared memory).  TVS_FUSED_FN gives the entry points of each build their own names.
#ifndef TVS_HALVES
#error "include this file from kernels_fused_w256.cu / kernels_fused_w128.cu"
#endif
#ifndef TVS_CTAS_PER_SM
#define TVS_CTAS_PER_SM (TVS_HALVES == 2 ? 2 : 4)
#endif
#define TVS_CAT2(a, b) a##b
#define TVS_CAT(a, b) TVS_CAT2(a, b)
#define TVS_FUSED_FN(name) TVS_CAT(name, TVS_FUSED_SUFFIX)
constexpr int kHalves = TVS_HALVES;        // 128-column halves per tile: 4 warps each in the H pass
constexpr int kCtasPerSM = TVS_CTAS_PER_SM;
constexpr int kHalfW = 128;               // columns per half tile: lane * 4 + k
constexpr int kTileW = kHalves * kHalfW;
constexpr int kThreads = 128 * kHalves;
constexpr int kWarps = kThreads / 32;
constexpr int kThreadsLog2 = kHalves == 2 ? 8 : 7;   // the staging loops give each row 2^k <= kThreads threads
constexpr int kGroup = 4;                 // adjacent output columns per lane
constexpr int kGroupsPerHalf = kHalfW / kGroup;
constexpr int kMaxFusedOps = 3;
constexpr int kMaxTilesX = 64, kMaxTilesY = 160;   // tile descriptors travel in the kernel parameters
constexpr size_t kSmemBudget = (227 * 1024) / kCtasPerSM - 1536;   // shared memory per CTA so that kCtasPerSM fit on an SM

struct FusedOpParams {
    const uint8_t *src; ptrdiff_t src_pitch, src_fstride; int src_xstep;
    uint8_t *dst; ptrdiff_t dst_pitch, dst_fstride; int dst_xstep;
    int sw, sh, dw, dh;
    int th, tiles_x, tiles_y, tile_begin;
    int n_groups;
    const float *hdense;   // [n_groups][4][SPAN]  c * 2^-30, zero outside a column's own window
    const int *hbase;      // [n_groups]           first source column of the group's span
    const int *vpos;       // [dh]
    const float *vring;    // [rows][VT][2]        (g, g) * 2^107, indexed by ring slot = source_row % VT
    const int *vring_i;    // [dh][VT]             the same coefficients as integers (exact path)
    int vtaps;
    int s_pitch;           // floats per staged source row
    int s_chunks;          // 16-byte source chunks staged per row
    int s_chunks_log2;     // log2 of the power of two >= s_chunks (index split in the staging loops)
    const uint8_t *src_lo; // first byte of the stored source plane (frame 0) and its extent: bounds for 16-byte reads
    size_t src_extent;
    int m_rows;            // rows of the shared-memory tiles
    float v_thr;           // |residual| at or above this needs the exact path
    float v_init;          // start value of the vertical accumulators (see vertical_float_rule)
    int v_mono;            // consecutive output rows start at most one source row apart (no vertical shrink)
    int src_mode;          // kSrcU8 / kSrcU8Pair / kSrcU16 / kSrcU16Pair: sample width and spacing in the source plane
    float src_scale;       // 16-bit sources: 1, or 2^-6 for P010 (its 10 bits sit at the top of the word)
    float h_addend;        // subnormal offset of the H result: 16384 * 2^-149 when samples were centred on 512, else 0
    int dst_shift;         // 10-bit destinations: left shift of the stored value (P010: 6)
    int dither_xoff;       // ordered dither: column offset of this plane (3 for V, else 0)
    // per tile column: first source column and column count of the footprint; per tile row: source row range
    short tile_c0[kMaxTilesX], tile_nc[kMaxTilesX], tile_r0[kMaxTilesY], tile_r1[kMaxTilesY];
};

struct FusedLaunchParams {
    FusedOpParams op[kMaxFusedOps];
    int n_ops;
    int tiles_per_frame, total_tiles;
    // shared-memory layout, sized for the largest plane of the launch (element counts)
    int m_floats, s_floats, raw_words, cv_pairs, vp_ints;
    int debug_skip;   // profiling aid (TVS_DEBUG_SKIP): bit 0 skips the conversion, bit 1 the H pass, bit 2 the V pass
};

// Pull intermediate rows [next_row, need_end) into the register rings (left and right half).  Ring slot =
// row % VT, so the slot a row lands in is a compile-time constant inside each case; the switch enters at the
// current slot and falls through, which keeps the rings in registers (no dynamic indexing) at the cost of one
// uniform branch per call.  Rows past the tile's last real row are zero-filled in shared memory, so the loads
// need no guard.
template <int VT, bool DUAL>
__device__ __forceinline__ void load_ring_rows(Quad (&wa)[VT], Quad (&wb)[VT], int &slot, int &next_row,
                                               int need_end, const float *&mp) {
#define TVS_RING_CASE(s)                                                                           \
    case s:                                                                                        \
        if constexpr (s < VT) {                                                                    \
            wa[s] = load_quad(mp);                                                                 \
            if (DUAL) wb[s] = load_quad(mp + kHalfW);                                              \
            mp += kTileW;                                                                          \
            if (++next_row >= need_end) { slot = (s + 1 == VT) ? 0 : s + 1; return; }              \
        }
    for (;;) {
        switch (slot) {
            TVS_RING_CASE(0) TVS_RING_CASE(1) TVS_RING_CASE(2) TVS_RING_CASE(3) TVS_RING_CASE(4) TVS_RING_CASE(5)
            TVS_RING_CASE(6) TVS_RING_CASE(7) TVS_RING_CASE(8) TVS_RING_CASE(9) TVS_RING_CASE(10) TVS_RING_CASE(11)
        }
        slot = 0;
    }
#undef TVS_RING_CASE
}

// Exact redo of the samples of one quad whose float residual sits inside the error band of a rounding boundary.
// Rare (about one warp-row in fifty), so it re-reads its VT intermediate rows from shared memory instead of
// keeping the register ring alive across a call.  `m_first` points at this quad's columns in row `first`;
// `dither` are the four ordered-dither values (kOutU8Dither only).
template <int VT, int OUT>
__device__ __noinline__ uint2 exact_fixup(const float *m_first, int first, const int *exact_coef, float v_thr,
                                          float r0, float r1, float r2, float r3, int4 dither, int dst_shift,
                                          uint2 packed) {
    unsigned word[2] = {packed.x, packed.y};
    const float r[4] = {r0, r1, r2, r3};
    const int d[4] = {dither.x, dither.y, dither.z, dither.w};
    int slot = first % VT;
    int a[4] = {0, 0, 0, 0};
    for (int j = 0; j < VT; ++j) {
        const float4 t = *reinterpret_cast<const float4 *>(m_first + j * kTileW);
        const int c = __ldg(exact_coef + slot);
        a[0] += subnormal_to_int(t.x) * c; a[1] += subnormal_to_int(t.y) * c;
        a[2] += subnormal_to_int(t.z) * c; a[3] += subnormal_to_int(t.w) * c;
        slot = slot + 1 == VT ? 0 : slot + 1;
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

// Per-row extras of the output rule: the accumulator start values (dither) for the lane's four columns.
struct RowExtras {
    f32x2 init01, init23;   // (d - 64) * 4096 * 2^-42 per column, or zero
    int4 dither;            // the integer dither values, for the exact path
};

// One output row, 4 + 4 columns (left and right half of the tile): float sums, rounded quotients, packed samples.
// `gg` holds each ring coefficient twice, (g, g), so one FFMA2 serves two columns.  The four accumulator chains
// are independent, which is what hides the FMA latency at 4 warps per scheduler.
template <int VT, bool RIGHT, int OUT>
__device__ __forceinline__ void vertical_row(const Quad (&wa)[VT], const Quad (&wb)[VT], const f32x2 (&gg)[VT], float v_thr,
                                             const int *exact_coef, const float *m_first, int first, const RowExtras &ex,
                                             int dst_shift, unsigned (&word_a)[2], unsigned (&word_b)[2]) {
    f32x2 a01 = ex.init01, a23 = ex.init23, b01 = ex.init01, b23 = ex.init23;   // the halves are 128 = 0 mod 8 columns apart
#pragma unroll
    for (int s = 0; s < VT; ++s) {
        a01 = fma2(wa[s].lo, gg[s], a01);
        a23 = fma2(wa[s].hi, gg[s], a23);
        if (RIGHT) {
            b01 = fma2(wb[s].lo, gg[s], b01);
            b23 = fma2(wb[s].hi, gg[s], b23);
        }
    }
    float ra[4], rb[4];
    const float worst_a = finish_quad<OUT>(a01, a23, dst_shift, word_a, ra);
    const float worst_b = RIGHT ? finish_quad<OUT>(b01, b23, dst_shift, word_b, rb) : 0.0f;
    if (fmaxf(worst_a, worst_b) >= v_thr) {   // one branch on the common path; each half is redone only if it asked
        if (worst_a >= v_thr) {
            const uint2 fa = exact_fixup<VT, OUT>(m_first, first, exact_coef, v_thr, ra[0], ra[1], ra[2], ra[3], ex.dither, dst_shift,
                                                  make_uint2(word_a[0], word_a[1]));
            word_a[0] = fa.x; word_a[1] = fa.y;
        }
        if (RIGHT && worst_b >= v_thr) {
            const uint2 fb = exact_fixup<VT, OUT>(m_first + kHalfW, first, exact_coef, v_thr, rb[0], rb[1], rb[2], rb[3], ex.dither,
                                                  dst_shift, make_uint2(word_b[0], word_b[1]));
            word_b[0] = fb.x; word_b[1] = fb.y;
        }
    }
}

// Vertical pass of one warp over its strip of output rows [ya, yb).
// MODE kStoreBoth: the tile is full width and every destination row is aligned for vector stores: each quad is
// one store and nothing is predicated.  kStoreLeft: the same for a tile whose right half is empty.
// kStoreGeneric: anything else (partial quads, interleaved chroma, unaligned rows): bounds-checked element stores.
enum { kStoreBoth = 0, kStoreLeft = 1, kStoreGeneric = 2 };

template <int VT>
__device__ __forceinline__ void load_row_coefs(f32x2 (&gg)[VT], const f32x2 *cv) {
    if constexpr (VT % 2 == 0) {
#pragma unroll
        for (int s = 0; s < VT; s += 2) {
            const ulonglong2 c2 = *reinterpret_cast<const ulonglong2 *>(cv + s);
            gg[s] = c2.x; gg[s + 1] = c2.y;
        }
    } else {
#pragma unroll
        for (int s = 0; s < VT; ++s) gg[s] = cv[s];
    }
}

// DUAL = false keeps a single ring (one half of the tile per call): used when VT is too large for two rings to
// fit in registers; `M` and `xa` then address the half being processed and MODE kStoreBoth is not used.
template <int VT, int MODE, int OUT, bool DUAL>
__device__ __forceinline__ void vertical_strip(const float *M, const f32x2 *cv, const int *vp, const int *exact, int r0,
                                               int ya, int yb, uint8_t *drow, ptrdiff_t dst_pitch, int dst_xstep, int dw,
                                               int xa, bool right, float v_thr, int lane, int dst_shift, int dither_xoff,
                                               float v_init) {
    const int xb = xa + kHalfW;
    const float *m_col = M + 4 * lane;
    Quad wa[VT], wb[VT];
#pragma unroll
    for (int s = 0; s < VT; ++s) wa[s] = wb[s] = Quad{0ull, 0ull};
    int next_row = vp[0];
    int slot = next_row % VT;
    const float *mp = m_col + (next_row - r0) * kTileW;
    int y = ya;
    const int dcol = (xa + dither_xoff) & 7;   // dither column of the quad's first sample (4 * (lane & 1) + offset)

    // one output row; `gg` are this row's coefficients, `first` its first source row
    auto emit = [&](const f32x2 (&gg)[VT], int first) {
        if (next_row < first) {   // steps larger than the ring: skip rows nobody needs
            next_row = first; slot = first % VT; mp = m_col + (first - r0) * kTileW;
        }
        if (next_row < first + VT) load_ring_rows<VT, DUAL>(wa, wb, slot, next_row, first + VT, mp);
        const float *m_first = m_col + (first - r0) * kTileW;
        RowExtras ex;
        ex.init01 = ex.init23 = pack2(v_init, v_init);
        ex.dither = make_int4(0, 0, 0, 0);
        if (OUT == kOutU8Dither) {
            const int *drow_tab = kDither128i[y & 7];
            ex.dither = make_int4(drow_tab[dcol], drow_tab[(dcol + 1) & 7], drow_tab[(dcol + 2) & 7], drow_tab[(dcol + 3) & 7]);
            // (d - 64) * 4096 * 2^-42: the rounding quotient below already adds 64 << 12
            ex.init01 = pack2(fmaf(static_cast<float>(ex.dither.x - 64), 0x1p-30f, v_init), fmaf(static_cast<float>(ex.dither.y - 64), 0x1p-30f, v_init));
            ex.init23 = pack2(fmaf(static_cast<float>(ex.dither.z - 64), 0x1p-30f, v_init), fmaf(static_cast<float>(ex.dither.w - 64), 0x1p-30f, v_init));
        }
        unsigned word_a[2], word_b[2] = {0, 0};
        if (MODE == kStoreBoth && DUAL) {
            vertical_row<VT, true, OUT>(wa, wb, gg, v_thr, exact, m_first, first, ex, dst_shift, word_a, word_b);
            store_quad_fast<OUT>(drow, xa, word_a);
            store_quad_fast<OUT>(drow, xb, word_b);
        } else if (MODE == kStoreLeft) {
            vertical_row<VT, false, OUT>(wa, wb, gg, v_thr, exact, m_first, first, ex, dst_shift, word_a, word_b);
            store_quad_fast<OUT>(drow, xa, word_a);
        } else if (DUAL && right) {
            vertical_row<VT, true, OUT>(wa, wb, gg, v_thr, exact, m_first, first, ex, dst_shift, word_a, word_b);
            store_quad_slow<OUT>(drow, xa, dw, dst_xstep, word_a);
            store_quad_slow<OUT>(drow, xb, dw, dst_xstep, word_b);
        } else {
            vertical_row<VT, false, OUT>(wa, wb, gg, v_thr, exact, m_first, first, ex, dst_shift, word_a, word_b);
            store_quad_slow<OUT>(drow, xa, dw, dst_xstep, word_a);
        }
        drow += dst_pitch;
        exact += VT;
        ++y;
    };

    // Rows are taken two at a time so that each row's start position and coefficients are requested one row
    // ahead of their use without register-to-register copies (both tables carry a padding row).
    f32x2 g0[VT], g1[VT];
    load_row_coefs<VT>(g0, cv);
    int f0 = vp[0];
    while (y + 1 < yb) {
        load_row_coefs<VT>(g1, cv + VT);
        const int f1 = vp[1];
        emit(g0, f0);
        cv += 2 * VT; vp += 2;
        load_row_coefs<VT>(g0, cv);
        f0 = vp[0];
        emit(g1, f1);
    }
    if (y < yb) emit(g0, f0);
}

// The same strip for planes that are not shrunk vertically: consecutive output rows then start at the same source
// row or the next one, so the walk can be driven by SOURCE rows.  Each step loads one intermediate row into its ring
// slot -- a compile-time constant inside the unrolled step, no dispatch per output row -- and then emits every
// output row whose window that row completed.  A row's coefficients and start position are requested right after
// the previous row's multiply-adds, under the latency of its rounding / packing / store.
template <int VT, int MODE, int OUT, bool DUAL>
__device__ __forceinline__ void vertical_strip_mono(const float *M, const f32x2 *cv, const int *vp, const int *exact, int r0,
                                                    int ya, int yb, uint8_t *drow, ptrdiff_t dst_pitch, int dst_xstep, int dw,
                                                    int xa, bool right, float v_thr, int lane, int dst_shift, int dither_xoff,
                                                    float v_init) {
    constexpr bool kRight = DUAL && MODE == kStoreBoth;
    const int xb = xa + kHalfW;
    const float *m_col = M + 4 * lane;
    Quad wa[VT], wb[VT];
#pragma unroll
    for (int s = 0; s < VT; ++s) wa[s] = wb[s] = Quad{0ull, 0ull};
    int next_row = vp[0];
    int slot = next_row % VT;
    const float *mp = m_col + (next_row - r0) * kTileW;
    int y = ya;
    const int dcol = (xa + dither_xoff) & 7;
    // this lane's first sample of the current destination row (vector-store modes); the right half is a constant
    // offset away, and the exact-path table is addressed from the row number only when it is needed
    uint8_t *dlane = drow + static_cast<ptrdiff_t>(xa) * (OUT == kOutU10 ? 2 : 1);
    const bool both = kRight || (DUAL && MODE == kStoreGeneric && right);
    f32x2 gg[VT];
    load_row_coefs<VT>(gg, cv);
    int first = next_row;

    // every output row whose last source row is now in the ring; true once the strip is finished
    auto emit_ready = [&]() -> bool {
        while (first + VT <= next_row) {
            RowExtras ex;
            ex.init01 = ex.init23 = pack2(v_init, v_init);
            ex.dither = make_int4(0, 0, 0, 0);
            if (OUT == kOutU8Dither) {
                const int *drow_tab = kDither128i[y & 7];
                ex.dither = make_int4(drow_tab[dcol], drow_tab[(dcol + 1) & 7], drow_tab[(dcol + 2) & 7], drow_tab[(dcol + 3) & 7]);
                ex.init01 = pack2(fmaf(static_cast<float>(ex.dither.x - 64), 0x1p-30f, v_init), fmaf(static_cast<float>(ex.dither.y - 64), 0x1p-30f, v_init));
                ex.init23 = pack2(fmaf(static_cast<float>(ex.dither.z - 64), 0x1p-30f, v_init), fmaf(static_cast<float>(ex.dither.w - 64), 0x1p-30f, v_init));
            }
            f32x2 a01 = ex.init01, a23 = ex.init23, b01 = ex.init01, b23 = ex.init23;
#pragma unroll
            for (int s = 0; s < VT; ++s) {
                a01 = fma2(wa[s].lo, gg[s], a01);
                a23 = fma2(wa[s].hi, gg[s], a23);
                if (DUAL && (kRight || MODE == kStoreGeneric)) {
                    b01 = fma2(wb[s].lo, gg[s], b01);
                    b23 = fma2(wb[s].hi, gg[s], b23);
                }
            }
            // the next row's tables (both carry a padding row past the strip)
            cv += VT; ++vp;
            load_row_coefs<VT>(gg, cv);
            const int first_next = vp[0];

            unsigned word_a[2], word_b[2] = {0, 0};
            float ra[4], rb[4];
            const float worst_a = finish_quad<OUT>(a01, a23, dst_shift, word_a, ra);
            const float worst_b = both ? finish_quad<OUT>(b01, b23, dst_shift, word_b, rb) : 0.0f;
            if (fmaxf(worst_a, worst_b) >= v_thr) {
                const float *m_first = m_col + (first - r0) * kTileW;
                if (worst_a >= v_thr) {
                    const uint2 fa = exact_fixup<VT, OUT>(m_first, first, exact + (y - ya) * VT, v_thr, ra[0], ra[1], ra[2], ra[3], ex.dither, dst_shift,
                                                          make_uint2(word_a[0], word_a[1]));
                    word_a[0] = fa.x; word_a[1] = fa.y;
                }
                if (both && worst_b >= v_thr) {
                    const uint2 fb = exact_fixup<VT, OUT>(m_first + kHalfW, first, exact + (y - ya) * VT, v_thr, rb[0], rb[1], rb[2], rb[3], ex.dither,
                                                          dst_shift, make_uint2(word_b[0], word_b[1]));
                    word_b[0] = fb.x; word_b[1] = fb.y;
                }
            }
            if (MODE == kStoreGeneric) {
                store_quad_slow<OUT>(drow, xa, dw, dst_xstep, word_a);
                if (both) store_quad_slow<OUT>(drow, xb, dw, dst_xstep, word_b);
                drow += dst_pitch;
            } else {
                store_quad_fast<OUT>(dlane, 0, word_a);
                if (kRight) store_quad_fast<OUT>(dlane, kHalfW, word_b);
                dlane += dst_pitch;
            }
            first = first_next;
            if (++y >= yb) return true;
        }
        return false;
    };

#define TVS_MONO_STEP(s)                                                   \
    case s:                                                                \
        if constexpr (s < VT) {                                            \
            wa[s] = load_quad(mp);                                         \
            if (DUAL) wb[s] = load_quad(mp + kHalfW);                      \
            mp += kTileW;                                                  \
            ++next_row;                                                    \
            if (emit_ready()) return;                                      \
        }
    for (;;) {
        switch (slot) {
            TVS_MONO_STEP(0) TVS_MONO_STEP(1) TVS_MONO_STEP(2) TVS_MONO_STEP(3) TVS_MONO_STEP(4) TVS_MONO_STEP(5)
            TVS_MONO_STEP(6) TVS_MONO_STEP(7) TVS_MONO_STEP(8) TVS_MONO_STEP(9) TVS_MONO_STEP(10) TVS_MONO_STEP(11)
        }
        slot = 0;
    }
#undef TVS_MONO_STEP
}

struct TileCoord { int oi, f, tx, ty; };

__device__ __forceinline__ TileCoord decode_tile(const FusedLaunchParams &P, int tile) {
    TileCoord c;
    c.f = tile / P.tiles_per_frame;
    int rem = tile - c.f * P.tiles_per_frame;
    c.oi = 0;
    if (P.n_ops > 1 && rem >= P.op[1].tile_begin) c.oi = 1;
    if (P.n_ops > 2 && rem >= P.op[2].tile_begin) c.oi = 2;
    rem -= P.op[c.oi].tile_begin;
    const int tiles_y = P.op[c.oi].tiles_y;     // tiles of one column are consecutive: they share coefficients and halo rows
    c.tx = rem / tiles_y;
    c.ty = rem - c.tx * tiles_y;
    return c;
}

// The tile after `c` in the launch's tile order (rows of a tile column, columns of a plane, planes of a frame,
// frames): no divisions, unlike decode_tile.
__device__ __forceinline__ TileCoord next_tile(const FusedLaunchParams &P, TileCoord c) {
    if (++c.ty < P.op[c.oi].tiles_y) return c;
    c.ty = 0;
    if (++c.tx < P.op[c.oi].tiles_x) return c;
    c.tx = 0;
    if (++c.oi < P.n_ops) return c;
    c.oi = 0;
    ++c.f;
    return c;
}

// First byte of the tile's source footprint (sample c0 of row r0 of frame f).
__device__ __forceinline__ const uint8_t *tile_src_base(const FusedOpParams &op, const TileCoord &c) {
    return op.src + c.f * op.src_fstride + static_cast<ptrdiff_t>(op.tile_r0[c.ty]) * op.src_pitch +
           static_cast<ptrdiff_t>(op.tile_c0[c.tx]) * op.src_xstep;
}

// Start the asynchronous copy of the raw source footprint of tile `c`.  Rows are read as aligned 16-byte chunks
// starting at the chunk that holds the row's first sample, so byte i of the footprint row lands at position
// mis + i of the staged row, mis (0..15) being the misalignment of the row's first byte; any pitch works.  A chunk
// that would reach outside the plane (first / last bytes of the allocation) is fetched as 4-byte words that each
// contain at least one needed byte; chunks and words past the end of the row are zeroed.
__device__ __forceinline__ void prefetch_tile_source(const FusedLaunchParams &P, const TileCoord &c, unsigned *raw, int tid) {
    const FusedOpParams &op = P.op[c.oi];
    const int c0 = op.tile_c0[c.tx];
    const int nrows = op.tile_r1[c.ty] - op.tile_r0[c.ty];
    const int xs = op.src_xstep;
    const int sample_bytes = op.src_mode >= kSrcU16 ? 2 : 1;
    // offset of the last needed byte of a footprint row (samples past the plane's width carry zero weight)
    const int last_byte = (min(static_cast<int>(op.tile_nc[c.tx]), op.sw - c0) - 1) * xs + sample_bytes - 1;
    const int clog = min(op.s_chunks_log2, kThreadsLog2), cmask = (1 << clog) - 1, s_chunks = op.s_chunks;
    const uint8_t *src_base = tile_src_base(op, c);
    const uint8_t *lo = op.src_lo + c.f * op.src_fstride, *hi = lo + op.src_extent;
    const unsigned mis0 = static_cast<unsigned>(reinterpret_cast<uintptr_t>(src_base)) & 15u;
    const unsigned pmod = static_cast<unsigned>(op.src_pitch) & 15u;
    const ptrdiff_t src_pitch = op.src_pitch;
    const int rstep = kThreads >> clog;
    for (int ch = tid & cmask; ch < s_chunks; ch += 1 << clog) {   // one pass unless a row is longer than 256 chunks
        int row = tid >> clog;
        uint4 *dst = reinterpret_cast<uint4 *>(raw) + row * s_chunks + ch;
        const uint8_t *src = src_base + row * src_pitch + 16 * ch;
        unsigned mis = (mis0 + row * pmod) & 15u;
        const unsigned mstep = (rstep * pmod) & 15u;
        for (; row < nrows; row += rstep, dst += rstep * s_chunks, src += rstep * src_pitch, mis = (mis + mstep) & 15u) {
            const int first = 16 * ch - static_cast<int>(mis);      // footprint-row offset of the chunk's first byte
            const uint8_t *g = src - mis;
            if (first > last_byte) {
                *dst = make_uint4(0u, 0u, 0u, 0u);
            } else if (g >= lo && g + 16 <= hi) {
                cp_async16(dst, g);
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    unsigned *d = reinterpret_cast<unsigned *>(dst) + k;
                    if (first + 4 * k <= last_byte && first + 4 * k + 3 >= 0) cp_async4(d, g + 4 * k);
                    else *d = 0u;
                }
            }
        }
    }
}

template <int VT>
__device__ __forceinline__ void prefetch_tile_tables(const FusedLaunchParams &P, const TileCoord &c, f32x2 *cv, int *vp, int tid) {
    const FusedOpParams &op = P.op[c.oi];
    const int y0 = c.ty * op.th;
    const char *cvg = reinterpret_cast<const char *>(op.vring) + static_cast<size_t>(y0) * VT * sizeof(f32x2);
    for (int i = tid; i < op.th * VT / 2; i += kThreads) cp_async16(reinterpret_cast<char *>(cv) + 16 * i, cvg + 16 * i);
    const char *vpg = reinterpret_cast<const char *>(op.vpos + y0);
    for (int i = tid; i < op.th / 4; i += kThreads) cp_async16(reinterpret_cast<char *>(vp) + 16 * i, vpg + 16 * i);
}

// Horizontal coefficients of the 64 column groups of tile column `tx` (2 halves x 32 lanes), laid out so that a
// warp's 128-bit reads are conflict free: CD[(half * N4 + i) * 32 + lane] as 16-byte units, N4 = 4 * SPAN / 4.
template <int SPAN>
__device__ __forceinline__ void prefetch_hcoef(const FusedOpParams &op, int tx, float4 *cd, int *hb, int tid) {
    constexpr int N4 = kGroup * SPAN / 4;
    const int g0 = tx * kHalves * kGroupsPerHalf;
    for (int i = tid; i < kHalves * kGroupsPerHalf * N4; i += kThreads) {
        const int grp = i / N4, part = i - grp * N4;     // consecutive threads read consecutive 16-byte pieces
        const int g = min(g0 + grp, op.n_groups - 1);
        const int half = grp >> 5, ln = grp & 31;
        cp_async16(cd + (half * N4 + part) * 32 + ln, op.hdense + static_cast<size_t>(g) * (kGroup * SPAN) + 4 * part);
    }
    if (tid < kHalves * kGroupsPerHalf) hb[tid] = __ldg(op.hbase + min(g0 + tid, op.n_groups - 1)) - op.tile_c0[tx];
}

// Persistent kernel: each CTA walks a contiguous range of tiles.  While tile i is filtered, the source bytes and
// row tables of tile i+1 stream into the other half of a double buffer (cp.async), so no phase waits on HBM.
template <int SPAN, int VT, int OUT>
__global__ void __launch_bounds__(kThreads, kCtasPerSM)
fused_scale_kernel(const __grid_constant__ FusedLaunchParams P) {
    static_assert(SPAN % 2 == 0, "the packed horizontal pass pairs up source samples");
    constexpr int N4 = kGroup * SPAN / 4;
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int half = warp >> 2;

    float *M = smem;                                             // [m_rows + VT][256] intermediate t * 2^-149
    float *S = M + P.m_floats;                                   // [m_rows][s_pitch]  source samples as floats
    unsigned *RAW = reinterpret_cast<unsigned *>(S + P.s_floats);   // 2 x [m_rows][s_chunks] raw 16-byte source chunks
    f32x2 *CVB = reinterpret_cast<f32x2 *>(RAW + 2 * P.raw_words);  // 2 x [th + 1][VT] ring-ordered (g, g) pairs
    int *VPB = reinterpret_cast<int *>(CVB + 2 * P.cv_pairs);       // 2 x [th + 1]  first source row per output row
    float4 *CD = reinterpret_cast<float4 *>(VPB + 2 * P.vp_ints);   // [2][N4][32]   horizontal coefficients of the column
    int *HB = reinterpret_cast<int *>(CD + kHalves * N4 * 32);            // [64]          span start of each column group

    // contiguous share of the tile list
    const int per = P.total_tiles / gridDim.x, extra = P.total_tiles % gridDim.x;
    const int t_begin = blockIdx.x * per + min(static_cast<int>(blockIdx.x), extra);
    const int t_end = t_begin + per + (static_cast<int>(blockIdx.x) < extra ? 1 : 0);
    if (t_begin >= t_end) return;

    int buf = 0;
    TileCoord cur = decode_tile(P, t_begin);
    prefetch_tile_source(P, cur, RAW, tid);
    prefetch_tile_tables<VT>(P, cur, CVB, VPB, tid);
    prefetch_hcoef<SPAN>(P.op[cur.oi], cur.tx, CD, HB, tid);
    cp_async_commit();

    for (int tile = t_begin; tile < t_end; ++tile, buf ^= 1) {
        const FusedOpParams &op = P.op[cur.oi];
        // No barrier here: a thread converts exactly the source words it prefetched itself, so its own
        // cp.async completion is enough, and the conversion only writes S, which nobody reads during the vertical
        // pass.  Warps that finish a tile early therefore run ahead into the next tile's conversion.
        cp_async_wait_all();
        TileCoord nxt = cur;
        const bool more = tile + 1 < t_end;
        if (more) {
            nxt = next_tile(P, cur);
            prefetch_tile_source(P, nxt, RAW + (buf ^ 1) * P.raw_words, tid);
            cp_async_commit();
        }

        const int th = op.th, dw = op.dw;
        const int x0 = cur.tx * kTileW;
        const int y0 = cur.ty * th, y1 = min(y0 + th, op.dh);
        const int r0 = op.tile_r0[cur.ty];
        const int nrows = op.tile_r1[cur.ty] - r0;
        const int s_pitch = op.s_pitch, xs = op.src_xstep;
        const unsigned mis0 = static_cast<unsigned>(reinterpret_cast<uintptr_t>(tile_src_base(op, cur))) & 15u;
        const unsigned pmod = static_cast<unsigned>(op.src_pitch) & 15u;
        const unsigned *raw = RAW + buf * P.raw_words;

        // ---- raw words -> float samples ----
        if (!(P.debug_skip & 1)) {
            const int clog = min(op.s_chunks_log2, kThreadsLog2), cmask = (1 << clog) - 1, s_chunks = op.s_chunks;
            const int rstep = kThreads >> clog;
            const int src_mode = op.src_mode;
            const float sscale = op.src_scale;
            const unsigned keep16 = sscale < 1.0f ? 0xFFC0FFC0u : 0xFFFFFFFFu;   // P010: drop whatever sits in the six padding bits, as libswscale's shift does
            // a thread converts the chunks it prefetched: byte p of a staged row is sample p / spacing of S
            for (int ch = tid & cmask; ch < s_chunks; ch += 1 << clog) {
                int row = tid >> clog;
                const uint4 *rw = reinterpret_cast<const uint4 *>(raw) + row * s_chunks + ch;
                if (src_mode == kSrcU8) {              // 16 samples per chunk
                    float4 *out = reinterpret_cast<float4 *>(S + row * s_pitch) + 4 * ch;
                    for (; row < nrows; row += rstep, rw += rstep * s_chunks, out += rstep * (s_pitch / 4)) {
                        const uint4 v = *rw;
                        const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            out[k] = make_float4(byte_to_float(w[k], 0x7540u), byte_to_float(w[k], 0x7541u),
                                                 byte_to_float(w[k], 0x7542u), byte_to_float(w[k], 0x7543u));
                    }
                } else if (src_mode == kSrcU8Pair) {   // interleaved 8-bit chroma: this plane is every other byte
                    float4 *out = reinterpret_cast<float4 *>(S + row * s_pitch) + 2 * ch;
                    for (; row < nrows; row += rstep, rw += rstep * s_chunks, out += rstep * (s_pitch / 4)) {
                        const unsigned p = (mis0 + row * pmod) & 1u;
                        const uint4 v = *rw;
                        out[0] = make_float4(byte_to_float(v.x, 0x7540u | p), byte_to_float(v.x, 0x7542u | p),
                                             byte_to_float(v.y, 0x7540u | p), byte_to_float(v.y, 0x7542u | p));
                        out[1] = make_float4(byte_to_float(v.z, 0x7540u | p), byte_to_float(v.z, 0x7542u | p),
                                             byte_to_float(v.w, 0x7540u | p), byte_to_float(v.w, 0x7542u | p));
                    }
                } else if (src_mode == kSrcU16) {      // eight 16-bit samples per chunk, centred on 512
                    float4 *out = reinterpret_cast<float4 *>(S + row * s_pitch) + 2 * ch;
                    for (; row < nrows; row += rstep, rw += rstep * s_chunks, out += rstep * (s_pitch / 4)) {
                        uint4 v = *rw;
                        v.x &= keep16; v.y &= keep16; v.z &= keep16; v.w &= keep16;
                        out[0] = make_float4(fmaf(byte_to_float(v.x, 0x7410u), sscale, -512.0f), fmaf(byte_to_float(v.x, 0x7432u), sscale, -512.0f),
                                             fmaf(byte_to_float(v.y, 0x7410u), sscale, -512.0f), fmaf(byte_to_float(v.y, 0x7432u), sscale, -512.0f));
                        out[1] = make_float4(fmaf(byte_to_float(v.z, 0x7410u), sscale, -512.0f), fmaf(byte_to_float(v.z, 0x7432u), sscale, -512.0f),
                                             fmaf(byte_to_float(v.w, 0x7410u), sscale, -512.0f), fmaf(byte_to_float(v.w, 0x7432u), sscale, -512.0f));
                    }
                } else {                               // interleaved 16-bit chroma: one sample of this plane per 4 bytes
                    float4 *out = reinterpret_cast<float4 *>(S + row * s_pitch) + ch;
                    for (; row < nrows; row += rstep, rw += rstep * s_chunks, out += rstep * (s_pitch / 4)) {
                        const unsigned sel = ((mis0 + row * pmod) & 2u) ? 0x7432u : 0x7410u;
                        uint4 v = *rw;
                        v.x &= keep16; v.y &= keep16; v.z &= keep16; v.w &= keep16;
                        *out = make_float4(fmaf(byte_to_float(v.x, sel), sscale, -512.0f), fmaf(byte_to_float(v.y, sel), sscale, -512.0f),
                                           fmaf(byte_to_float(v.z, sel), sscale, -512.0f), fmaf(byte_to_float(v.w, sel), sscale, -512.0f));
                    }
                }
            }
        }
        __syncthreads();   // (A) S is complete, the tables copied a tile ago are visible, and the previous tile's M is free

        // the other halves of the table buffers were last read by the previous tile's vertical pass: refill them now
        if (more) {
            prefetch_tile_tables<VT>(P, nxt, CVB + (buf ^ 1) * P.cv_pairs, VPB + (buf ^ 1) * P.vp_ints, tid);
            cp_async_commit();
        }
        // rows the vertical ring may touch beyond the last real row read as zero
        for (int i = tid; i < VT * (kTileW / 4); i += kThreads)
            reinterpret_cast<float4 *>(M + nrows * kTileW)[i] = make_float4(0.f, 0.f, 0.f, 0.f);

        // ---- horizontal pass: S -> M ----------------------------------------------------------------
        if (x0 + half * kHalfW < dw && !(P.debug_skip & 2)) {
            f32x2 cd[kGroup][SPAN / 2];      // this lane's dense coefficients, (c[i], c[i+1]) pairs
            {
                f32x2 *flat = &cd[0][0];
#pragma unroll
                for (int i = 0; i < N4; ++i) {
                    const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(CD + (half * N4 + i) * 32 + lane);
                    flat[2 * i] = v.x; flat[2 * i + 1] = v.y;
                }
            }
            const int hb = HB[half * kGroupsPerHalf + lane];
            const int src_mode_h = op.src_mode;
            const float h_addend = op.h_addend;
            const float t_max = __int_as_float(32767);
            float *mcol = M + half * kHalfW + 4 * lane;
            for (int row = warp & 3; row < nrows; row += 4) {
                const unsigned mis = (mis0 + row * pmod) & 15u;
                // position of the row's first sample inside the staged row (see prefetch_tile_source)
                const float *sp = S + row * s_pitch + hb + (src_mode_h == kSrcU8 ? mis : src_mode_h == kSrcU16Pair ? (mis >> 2) : (mis >> 1));
                f32x2 px[SPAN / 2];
#pragma unroll
                for (int i = 0; i < SPAN / 2; ++i) px[i] = pack2(sp[2 * i], sp[2 * i + 1]);
                float4 o;
                float *ov = reinterpret_cast<float *>(&o);
#pragma unroll
                for (int k = 0; k < kGroup; ++k) {
                    f32x2 acc = 0ull;
#pragma unroll
                    for (int i = 0; i < SPAN / 2; ++i) acc = fma2(px[i], cd[k][i], acc);
                    float even, odd;
                    unpack2(acc, even, odd);
                    ov[k] = fminf(fma_rm(even + odd, 0x1p-126f, h_addend), t_max);   // floor(sum >> hshift), clamped above
                }
                *reinterpret_cast<float4 *>(mcol + row * kTileW) = o;
            }
        }
        __syncthreads();   // (B) M is complete

        // the next tile starts a new column: its horizontal coefficients can stream in under the vertical pass
        if (more && (nxt.oi != cur.oi || nxt.tx != cur.tx)) {
            prefetch_hcoef<SPAN>(P.op[nxt.oi], nxt.tx, CD, HB, tid);
            cp_async_commit();
        }

        // ---- vertical pass: M -> destination --------------------------------------------------------
        const int strip = th / kWarps;
        const int ya = y0 + warp * strip, yb = min(ya + strip, y1);
        if (ya < yb && !(P.debug_skip & 4)) {
            const int dst_xstep = op.dst_xstep;
            const ptrdiff_t dst_pitch = op.dst_pitch;
            uint8_t *drow = op.dst + cur.f * op.dst_fstride + static_cast<ptrdiff_t>(ya) * dst_pitch;
            // vector stores need samples back to back and every row of this tile column on a store-sized boundary
            constexpr int kSample = OUT == kOutU10 ? 2 : 1, kAlign = 4 * kSample - 1;
            const bool words = dst_xstep == kSample && (dst_pitch & kAlign) == 0 &&
                               ((reinterpret_cast<uintptr_t>(drow) + static_cast<uintptr_t>(x0) * kSample) & kAlign) == 0;
            const int *exact = op.vring_i + static_cast<size_t>(ya) * VT;
            const f32x2 *cv = CVB + buf * P.cv_pairs + (ya - y0) * VT;
            const int *vp = VPB + buf * P.vp_ints + (ya - y0);
            const int xa = x0 + 4 * lane;
            // the source-row-driven walk pays off for short filters; with more taps its unrolled steps outgrow the
            // instruction cache's benefit (measured: +4 % at 4 taps, -1 % at 6)
            const bool mono = VT <= 4 && op.v_mono != 0;
            constexpr bool kDual = kHalves == 2 && VT <= 8;   // two register rings of VT rows x 4 columns only fit up to 8 taps
            if constexpr (kDual) {
                if (words && x0 + kTileW <= dw)
                    { if (mono) vertical_strip_mono<VT, kStoreBoth, OUT, true>(M, cv, vp, exact, r0, ya, yb, drow, dst_pitch, kSample, dw, xa, true,
                                                              op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); else vertical_strip<VT, kStoreBoth, OUT, true>(M, cv, vp, exact, r0, ya, yb, drow, dst_pitch, kSample, dw, xa, true,
                                                              op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); }
                else if (words && x0 + kHalfW == dw)
                    { if (mono) vertical_strip_mono<VT, kStoreLeft, OUT, true>(M, cv, vp, exact, r0, ya, yb, drow, dst_pitch, kSample, dw, xa, false,
                                                              op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); else vertical_strip<VT, kStoreLeft, OUT, true>(M, cv, vp, exact, r0, ya, yb, drow, dst_pitch, kSample, dw, xa, false,
                                                              op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); }
                else
                    { if (mono) vertical_strip_mono<VT, kStoreGeneric, OUT, true>(M, cv, vp, exact, r0, ya, yb, drow, dst_pitch, dst_xstep, dw, xa,
                                                                 x0 + kHalfW < dw, op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); else vertical_strip<VT, kStoreGeneric, OUT, true>(M, cv, vp, exact, r0, ya, yb, drow, dst_pitch, dst_xstep, dw, xa,
                                                                 x0 + kHalfW < dw, op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); }
            } else {
                for (int h = 0; h < kHalves && x0 + h * kHalfW < dw; ++h) {
                    const int xh = xa + h * kHalfW;
                    if (words && x0 + (h + 1) * kHalfW <= dw)
                        { if (mono) vertical_strip_mono<VT, kStoreLeft, OUT, false>(M + h * kHalfW, cv, vp, exact, r0, ya, yb, drow, dst_pitch, kSample, dw,
                                                                   xh, false, op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); else vertical_strip<VT, kStoreLeft, OUT, false>(M + h * kHalfW, cv, vp, exact, r0, ya, yb, drow, dst_pitch, kSample, dw,
                                                                   xh, false, op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); }
                    else
                        { if (mono) vertical_strip_mono<VT, kStoreGeneric, OUT, false>(M + h * kHalfW, cv, vp, exact, r0, ya, yb, drow, dst_pitch, dst_xstep,
                                                                      dw, xh, false, op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); else vertical_strip<VT, kStoreGeneric, OUT, false>(M + h * kHalfW, cv, vp, exact, r0, ya, yb, drow, dst_pitch, dst_xstep,
                                                                      dw, xh, false, op.v_thr, lane, op.dst_shift, op.dither_xoff, op.v_init); }
                }
            }
        }
        cur = nxt;
    }
}

using KernelFn = void (*)(const FusedLaunchParams);
struct Instance { int span, vt; KernelFn fn[3]; };   // one kernel per output rule
#define TVS_INST(S, V) {S, V, {fused_scale_kernel<S, V, kOutU8>, fused_scale_kernel<S, V, kOutU8Dither>, fused_scale_kernel<S, V, kOutU10>}}
const Instance kInstances[] = {
    TVS_INST(4, 1),  TVS_INST(8, 1),  TVS_INST(4, 2),  TVS_INST(8, 2),  TVS_INST(6, 4),  TVS_INST(8, 4),
    TVS_INST(8, 6),  TVS_INST(12, 3), TVS_INST(12, 4), TVS_INST(12, 6), TVS_INST(16, 6), TVS_INST(16, 9),
    TVS_INST(20, 12), TVS_INST(24, 12),
};
#undef TVS_INST

}  // namespace

struct FusedOp {
    int op_index = 0;
    int instance = -1;
    int th = 0, tiles_x = 0, tiles_y = 0, n_groups = 0, s_pitch = 0, s_chunks = 0, s_chunks_log2 = 0, m_rows = 0;
    size_t smem_bytes = 0;
    float v_thr = 0, v_init = 0;
    int v_mono = 0;
    // offsets (in bytes) into the device blob
    size_t hdense_off = 0, hbase_off = 0, vpos_off = 0, vring_off = 0, vring_i_off = 0;
    int vtaps = 0;
    int src_mode = 0, out_mode = 0;
    float src_scale = 1.f, h_addend = 0.f;
    std::vector<short> tile_c0, tile_nc, tile_r0, tile_r1;
};

struct FusedPlan {
    std::vector<FusedOp> ops;
    uint8_t *blob = nullptr;
    size_t blob_bytes = 0;
    int device = 0;
    int sm_count = 148;
    bool prefer_two_pass = false;   // wide filters: the dense horizontal pass wastes more than the scratch round trip costs
};

namespace {

template <typename T>
size_t append(std::vector<uint8_t> &blob, const std::vector<T> &v) {
    size_t off = (blob.size() + 255) & ~static_cast<size_t>(255);
    blob.resize(off + v.size() * sizeof(T));
    memcpy(blob.data() + off, v.data(), v.size() * sizeof(T));
    return off;
}

struct SmemLayout { int m_floats, s_floats, raw_words, cv_pairs, vp_ints; };
SmemLayout smem_layout(int m_rows, int s_pitch, int s_chunks, int th, int vt) {
    SmemLayout l;
    l.m_floats = (m_rows + vt) * kTileW;
    l.s_floats = m_rows * s_pitch;
    l.raw_words = m_rows * s_chunks * 4;
    l.cv_pairs = ((th + 1) * vt + 1) & ~1;
    l.vp_ints = (th + 1 + 3) & ~3;
    return l;
}
size_t layout_bytes(const SmemLayout &l, int span) {
    const size_t hcoef = static_cast<size_t>(kHalves * kGroupsPerHalf) * (kGroup * span + 1) * 4;   // CD + HB
    return static_cast<size_t>(l.m_floats + l.s_floats + 2 * l.raw_words + 2 * l.vp_ints) * 4 + static_cast<size_t>(2 * l.cv_pairs) * 8 + hcoef;
}
size_t fused_smem_bytes(int m_rows, int s_pitch, int s_chunks, int th, int vt, int span) {
    return layout_bytes(smem_layout(m_rows, s_pitch, s_chunks, th, vt), span);
}

// Worst-case absolute rounding error of the float vertical sum, in units of the integer sum.
double vertical_error_bound(const AxisTable &v, int vt_inst) {
    long long worst = 0;
    for (int y = 0; y < v.length; ++y) {
        long long s = 0;
        for (int j = 0; j < v.taps; ++j) s += std::abs(static_cast<int>(v.coef[static_cast<size_t>(y) * v.taps + j]));
        worst = std::max(worst, s);
    }
    const double bound = 32767.0 * static_cast<double>(worst) + 1.0;
    const double ulp = std::ldexp(1.0, static_cast<int>(std::floor(std::log2(bound))) - 23);
    return 0.5 * ulp * vt_inst;
}

}  // namespace

FusedPlan *TVS_FUSED_FN(fused_plan_create)(const Plan &plan, const AxisTable host_tables[4], int device, std::string &err) {
    std::vector<uint8_t> blob;
    auto fp = new FusedPlan;
    fp->device = device;
    cudaDeviceGetAttribute(&fp->sm_count, cudaDevAttrMultiProcessorCount, device);
    for (size_t oi = 0; oi < plan.ops.size(); ++oi) {
        const PlaneOp &op = plan.ops[oi];
        if (op.kind != kOpScale) continue;
        const AxisTable &h = host_tables[op.htable], &v = host_tables[op.vtable];
        const bool wide_src = op.src_bits > 8;
        const int sample_bytes = wide_src ? 2 : 1;
        const int src_mode = wide_src ? (op.src_xstep == 2 ? kSrcU16 : kSrcU16Pair) : (op.src_xstep == 1 ? kSrcU8 : kSrcU8Pair);
        if (op.src_xstep != sample_bytes && op.src_xstep != 2 * sample_bytes) { err = "unexpected sample spacing"; delete fp; return nullptr; }
        const int out_mode = op.rule == kOut10 ? kOutU10 : op.rule == kOut8Dither ? kOutU8Dither : kOutU8;

        // float exactness of the horizontal sum: rows must sum to 2^14 (so that centring 10-bit samples on 512 is
        // a constant offset) and max|sample| * sum|c| must stay below 2^24
        const int n_groups = (op.dw + kGroup - 1) / kGroup;
        int span = 0;
        long long worst_abs = 0;
        for (int x = 0; x < op.dw; ++x) {
            long long s = 0, a = 0;
            for (int j = 0; j < h.taps; ++j) { const int c = h.coef[static_cast<size_t>(x) * h.taps + j]; s += c; a += std::abs(c); }
            if (s != (1 << 14)) { err = "horizontal row does not sum to one"; delete fp; return nullptr; }
            worst_abs = std::max(worst_abs, a);
        }
        if ((wide_src ? 512 : 255) * worst_abs >= (1LL << 24)) { err = "horizontal sum not exact in fp32"; delete fp; return nullptr; }
        for (int g = 0; g < n_groups; ++g) {
            const int xa = g * kGroup, xb = std::min(xa + kGroup, op.dw) - 1;
            span = std::max(span, h.pos[xb] + h.taps - h.pos[xa]);
        }
        int inst = -1;
        for (size_t i = 0; i < sizeof(kInstances) / sizeof(kInstances[0]); ++i) {
            if (kInstances[i].span < span || kInstances[i].vt < v.taps) continue;
            if (inst < 0 || kInstances[i].span + kInstances[i].vt < kInstances[inst].span + kInstances[inst].vt) inst = static_cast<int>(i);
        }
        if (inst < 0) { err = "no kernel instance for this filter size"; delete fp; return nullptr; }
        const int SPAN = kInstances[inst].span, VT = kInstances[inst].vt;
        const VerticalFloatRule vrule = vertical_float_rule(v, VT, out_mode == kOutU10, out_mode == kOutU8Dither);
        if (!vrule.ok) { err = "vertical error bound too large"; delete fp; return nullptr; }

        FusedOp fo;
        fo.op_index = static_cast<int>(oi);
        fo.instance = inst;
        fo.n_groups = n_groups;
        fo.vtaps = v.taps;
        fo.tiles_x = (op.dw + kTileW - 1) / kTileW;
        fo.v_thr = vrule.thr;
        fo.v_init = vrule.init;
        fo.src_mode = src_mode;
        fo.out_mode = out_mode;
        fo.src_scale = op.src_shift ? std::ldexp(1.0f, -op.src_shift) : 1.0f;
        fo.h_addend = wide_src ? std::ldexp(16384.0f, -149) : 0.0f;   // (512 << 14) >> 9, as a subnormal

        // dense horizontal coefficients and span starts
        std::vector<float> hdense(static_cast<size_t>(n_groups) * kGroup * SPAN, 0.0f);
        std::vector<int> hbase(n_groups);
        for (int g = 0; g < n_groups; ++g) {
            const int xa = g * kGroup;
            hbase[g] = h.pos[xa];
            for (int k = 0; k < kGroup && xa + k < op.dw; ++k) {
                const int x = xa + k;
                for (int j = 0; j < h.taps; ++j) {
                    const int c = h.coef[static_cast<size_t>(x) * h.taps + j];
                    // scaled so that one FMA by 2^-126 lands floor(sum >> hshift) on the subnormal grid
                    if (c) hdense[(static_cast<size_t>(g) * kGroup + k) * SPAN + (h.pos[x] - hbase[g] + j)] =
                               std::ldexp(static_cast<float>(c), -(23 + op.hshift));
                }
            }
        }
        int ncols_max = 0;
        for (int tx = 0; tx < fo.tiles_x; ++tx) {
            const int ga = tx * kHalves * kGroupsPerHalf, gb = std::min(ga + kHalves * kGroupsPerHalf - 1, n_groups - 1);
            ncols_max = std::max(ncols_max, hbase[gb] + SPAN - hbase[ga]);
        }
        // 16-byte chunks per staged row for the worst alignment, and the float pitch that holds their samples
        fo.s_chunks = (15 + (ncols_max - 1) * op.src_xstep + sample_bytes - 1) / 16 + 1;
        fo.s_chunks_log2 = 0;
        while ((1 << fo.s_chunks_log2) < fo.s_chunks) ++fo.s_chunks_log2;
        fo.s_pitch = fo.s_chunks * (16 / op.src_xstep);            // samples per chunk = 16 / spacing (a multiple of 4)
        if ((fo.s_pitch & 31) == 0) fo.s_pitch += 4;                // keep consecutive rows off the same banks

        // tile height: as tall as shared memory allows (less vertical halo), a multiple of the warp count
        auto rows_for = [&](int th, int &tiles_y) {
            tiles_y = (op.dh + th - 1) / th;
            int m = 0;
            for (int ty = 0; ty < tiles_y; ++ty) {
                const int ya = ty * th, yb = std::min(ya + th, op.dh);
                m = std::max(m, std::min(v.pos[yb - 1] + v.taps, op.sh) - v.pos[ya]);
            }
            return m;
        };
        int th = 128;
        while (th > kWarps && th - kWarps >= op.dh) th -= kWarps;
        for (;; th -= kWarps) {
            fo.m_rows = rows_for(th, fo.tiles_y);
            fo.smem_bytes = fused_smem_bytes(fo.m_rows, fo.s_pitch, fo.s_chunks, th, VT, SPAN);
            if (fo.smem_bytes <= kSmemBudget || th == kWarps) break;
        }
        if (fo.smem_bytes > 226 * 1024) { err = "tile does not fit in shared memory"; delete fp; return nullptr; }
        fo.th = th;
        if (fo.tiles_x > kMaxTilesX || fo.tiles_y > kMaxTilesY) { err = "too many tiles for the descriptor table"; delete fp; return nullptr; }

        // vertical coefficients indexed by ring slot, and start rows; both padded to whole tiles so the kernel
        // can copy a tile's worth without bounds checks
        const int padded_rows = fo.tiles_y * th;
        std::vector<float> vring(static_cast<size_t>(padded_rows) * VT * 2, 0.0f);   // (g, g) pairs
        std::vector<int> vring_i(static_cast<size_t>(padded_rows) * VT, 0);
        std::vector<int> vpos_padded(padded_rows, v.pos[op.dh - 1]);
        fo.v_mono = 1;
        for (int y = 1; y < op.dh; ++y)
            if (v.pos[y] < v.pos[y - 1] || v.pos[y] > v.pos[y - 1] + 1) fo.v_mono = 0;
        for (int y = 0; y < op.dh; ++y) {
            vpos_padded[y] = v.pos[y];
            for (int j = 0; j < v.taps; ++j) {
                const int c = v.coef[static_cast<size_t>(y) * v.taps + j];
                const int slot = (v.pos[y] + j) % VT;
                vring[(static_cast<size_t>(y) * VT + slot) * 2] = vring[(static_cast<size_t>(y) * VT + slot) * 2 + 1] =
                    std::ldexp(static_cast<float>(c), 107);
                vring_i[static_cast<size_t>(y) * VT + slot] = c;
            }
        }
        for (int tx = 0; tx < fo.tiles_x; ++tx) {
            const int ga = tx * kHalves * kGroupsPerHalf, gb = std::min(ga + kHalves * kGroupsPerHalf - 1, n_groups - 1);
            fo.tile_c0.push_back(static_cast<short>(hbase[ga]));
            fo.tile_nc.push_back(static_cast<short>(hbase[gb] + SPAN - hbase[ga]));
        }
        for (int ty = 0; ty < fo.tiles_y; ++ty) {
            const int ya = ty * th, yb = std::min(ya + th, op.dh);
            fo.tile_r0.push_back(static_cast<short>(v.pos[ya]));
            fo.tile_r1.push_back(static_cast<short>(std::min(v.pos[yb - 1] + v.taps, op.sh)));
        }

        fo.hdense_off = append(blob, hdense);
        fo.hbase_off = append(blob, hbase);
        fo.vpos_off = append(blob, vpos_padded);
        fo.vring_off = append(blob, vring);
        fo.vring_i_off = append(blob, vring_i);
        if (SPAN >= 16 || VT > 8) fp->prefer_two_pass = true;   // measured on C4: 7.6K (fused) vs 12.0K frames/s (two-pass)
        fp->ops.push_back(fo);
    }
    if (fp->ops.empty()) { err = "nothing to scale"; delete fp; return nullptr; }
    // planes that share a kernel instance run in one launch whose shared-memory layout is the component-wise
    // maximum over them; make sure that still fits
    for (const FusedOp &a : fp->ops) {
        SmemLayout lay{0, 0, 0, 0, 0};
        for (const FusedOp &b : fp->ops) {
            if (b.instance != a.instance || b.out_mode != a.out_mode) continue;
            const SmemLayout l = smem_layout(b.m_rows, b.s_pitch, b.s_chunks, b.th, kInstances[b.instance].vt);
            lay.m_floats = std::max(lay.m_floats, l.m_floats); lay.s_floats = std::max(lay.s_floats, l.s_floats);
            lay.raw_words = std::max(lay.raw_words, l.raw_words); lay.cv_pairs = std::max(lay.cv_pairs, l.cv_pairs);
            lay.vp_ints = std::max(lay.vp_ints, l.vp_ints);
        }
        if (layout_bytes(lay, kInstances[a.instance].span) > 226 * 1024) { err = "combined tile layout does not fit in shared memory"; delete fp; return nullptr; }
    }
    if (cudaMalloc(&fp->blob, blob.size()) != cudaSuccess ||
        cudaMemcpy(fp->blob, blob.data(), blob.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
        err = "device allocation for the fused tables failed";
        cudaGetLastError();
        if (fp->blob) cudaFree(fp->blob);
        delete fp;
        return nullptr;
    }
    fp->blob_bytes = blob.size();
    for (const FusedOp &fo : fp->ops) {
        cudaFuncSetAttribute(reinterpret_cast<const void *>(kInstances[fo.instance].fn[fo.out_mode]),
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
    }
    return fp;
}

void TVS_FUSED_FN(fused_plan_destroy)(FusedPlan *p) {
    if (!p) return;
    if (p->blob) cudaFree(p->blob);
    delete p;
}

size_t TVS_FUSED_FN(fused_plan_device_bytes)(const FusedPlan *p) { return p ? p->blob_bytes : 0; }

bool TVS_FUSED_FN(fused_plan_prefers_two_pass)(const FusedPlan *p) { return p && p->prefer_two_pass; }

int TVS_FUSED_FN(fused_launch)(const FusedPlan *p, const Plan &plan, const SampleView src_views[], const SampleView dst_views[],
                 int n_frames, cudaStream_t stream) {
    int launched = 0;
    std::vector<bool> done(p->ops.size(), false);
    for (size_t i = 0; i < p->ops.size(); ++i) {
        if (done[i]) continue;
        // one launch covers every plane that runs on the same kernel instance
        FusedLaunchParams lp{};
        int tiles = 0;
        SmemLayout lay{0, 0, 0, 0, 0};
        const int vt = kInstances[p->ops[i].instance].vt;
        for (size_t j = i; j < p->ops.size() && lp.n_ops < kMaxFusedOps; ++j) {
            const FusedOp &fo = p->ops[j];
            if (done[j] || fo.instance != p->ops[i].instance || fo.out_mode != p->ops[i].out_mode) continue;
            done[j] = true;
            const PlaneOp &op = plan.ops[fo.op_index];
            const SampleView &s = src_views[fo.op_index], &d = dst_views[fo.op_index];
            FusedOpParams &q = lp.op[lp.n_ops++];
            q.src = s.base; q.src_pitch = s.pitch; q.src_fstride = s.frame_stride; q.src_xstep = s.xstep;
            q.dst = d.base; q.dst_pitch = d.pitch; q.dst_fstride = d.frame_stride; q.dst_xstep = d.xstep;
            q.sw = op.sw; q.sh = op.sh; q.dw = op.dw; q.dh = op.dh;
            q.th = fo.th; q.tiles_x = fo.tiles_x; q.tiles_y = fo.tiles_y; q.tile_begin = tiles;
            q.n_groups = fo.n_groups;
            q.hdense = reinterpret_cast<const float *>(p->blob + fo.hdense_off);
            q.hbase = reinterpret_cast<const int *>(p->blob + fo.hbase_off);
            q.vpos = reinterpret_cast<const int *>(p->blob + fo.vpos_off);
            q.vring = reinterpret_cast<const float *>(p->blob + fo.vring_off);
            q.vring_i = reinterpret_cast<const int *>(p->blob + fo.vring_i_off);
            q.vtaps = fo.vtaps;
            q.s_pitch = fo.s_pitch; q.s_chunks = fo.s_chunks; q.s_chunks_log2 = fo.s_chunks_log2; q.m_rows = fo.m_rows;
            q.src_lo = s.base - op.src_offset;
            q.src_extent = static_cast<size_t>(op.sh - 1) * static_cast<size_t>(s.pitch) + static_cast<size_t>(op.sw) * op.src_xstep;
            q.v_thr = fo.v_thr; q.v_init = fo.v_init; q.v_mono = fo.v_mono;
            q.src_mode = fo.src_mode; q.src_scale = fo.src_scale; q.h_addend = fo.h_addend;
            q.dst_shift = op.dst_shift; q.dither_xoff = op.dither_xoff;
            std::copy(fo.tile_c0.begin(), fo.tile_c0.end(), q.tile_c0);
            std::copy(fo.tile_nc.begin(), fo.tile_nc.end(), q.tile_nc);
            std::copy(fo.tile_r0.begin(), fo.tile_r0.end(), q.tile_r0);
            std::copy(fo.tile_r1.begin(), fo.tile_r1.end(), q.tile_r1);
            tiles += fo.tiles_x * fo.tiles_y;
            const SmemLayout l = smem_layout(fo.m_rows, fo.s_pitch, fo.s_chunks, fo.th, vt);
            lay.m_floats = std::max(lay.m_floats, l.m_floats);
            lay.s_floats = std::max(lay.s_floats, l.s_floats);
            lay.raw_words = std::max(lay.raw_words, l.raw_words);
            lay.cv_pairs = std::max(lay.cv_pairs, l.cv_pairs);
            lay.vp_ints = std::max(lay.vp_ints, l.vp_ints);
        }
        lp.tiles_per_frame = tiles;
        lp.total_tiles = tiles * n_frames;
        lp.m_floats = lay.m_floats; lp.s_floats = lay.s_floats; lp.raw_words = lay.raw_words;
        lp.cv_pairs = lay.cv_pairs; lp.vp_ints = lay.vp_ints;
#ifdef TVS_PROFILING   // phase-skipping switch for profiling builds only: a production library never reads it
        if (const char *dbg = getenv("TVS_DEBUG_SKIP")) lp.debug_skip = atoi(dbg);
#endif
        const size_t smem = layout_bytes(lay, kInstances[p->ops[i].instance].span);
        // persistent CTAs: two per SM, each walking a contiguous share of the tiles
        const int grid = std::min(lp.total_tiles, kCtasPerSM * p->sm_count);
        kInstances[p->ops[i].instance].fn[p->ops[i].out_mode]<<<grid, kThreads, smem, stream>>>(lp);
        if (cudaPeekAtLastError() != cudaSuccess) return -1;
        ++launched;
    }
    return launched;
}

}  // namespace tvs
