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
#include <cstring>

#include "tvs_internal.h"

namespace tvs {
namespace {

constexpr int kHalfW = 128;               // columns per half tile: lane * 4 + k
constexpr int kTileW = 2 * kHalfW;
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kGroup = 4;                 // adjacent output columns per lane
constexpr int kGroupsPerHalf = kHalfW / kGroup;
constexpr int kMaxFusedOps = 3;
constexpr int kMaxTilesX = 64, kMaxTilesY = 160;   // tile descriptors travel in the kernel parameters
constexpr size_t kSmemBudget = 112 * 1024;  // two CTAs per SM

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
    int s_words;           // 32-bit source words staged per row
    int s_words_log2;      // log2 of the power of two >= s_words (index split in the staging loop)
    int m_rows;            // rows of the shared-memory tiles
    float v_thr;           // |residual| at or above this needs the exact path
    // per tile column: first source column and column count of the footprint; per tile row: source row range
    short tile_c0[kMaxTilesX], tile_nc[kMaxTilesX], tile_r0[kMaxTilesY], tile_r1[kMaxTilesY];
};

struct FusedLaunchParams {
    FusedOpParams op[kMaxFusedOps];
    int n_ops;
    int tiles_per_frame, total_tiles;
    // shared-memory layout, sized for the largest plane of the launch (element counts)
    int m_floats, s_floats, raw_words, cv_pairs, vp_ints;
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

// Pull intermediate rows [next_row, need_end) into the register rings (left and right half).  Ring slot =
// row % VT, so the slot a row lands in is a compile-time constant inside each case; the switch enters at the
// current slot and falls through, which keeps the rings in registers (no dynamic indexing) at the cost of one
// uniform branch per call.  Rows past the tile's last real row are zero-filled in shared memory, so the loads
// need no guard.
template <int VT>
__device__ __forceinline__ void load_ring_rows(Quad (&wa)[VT], Quad (&wb)[VT], int &slot, int &next_row,
                                               int need_end, const float *&mp) {
#define TVS_RING_CASE(s)                                                                           \
    case s:                                                                                        \
        if constexpr (s < VT) {                                                                    \
            wa[s] = load_quad(mp);                                                                 \
            wb[s] = load_quad(mp + kHalfW);                                                        \
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

// One output row of 4 columns: float sum, rounded quotient, packed bytes, and the exact redo where needed.
// `gg` holds each ring coefficient twice, (g, g), so one FFMA2 serves two columns.
template <int VT>
__device__ __forceinline__ unsigned vertical_quad(const Quad (&win)[VT], const f32x2 (&gg)[VT], float v_thr,
                                                  const int *exact_coef) {
    f32x2 a01 = 0ull, a23 = 0ull;
#pragma unroll
    for (int s = 0; s < VT; ++s) {
        a01 = fma2(win[s].lo, gg[s], a01);
        a23 = fma2(win[s].hi, gg[s], a23);
    }
    const f32x2 scale = pack2(0x1p-126f, 0x1p-126f), back = pack2(-0x1p126f, -0x1p126f);
    const f32x2 q01 = fma2(a01, scale, 0ull), q23 = fma2(a23, scale, 0ull);   // round(acc / 2^19) * 2^-149
    const f32x2 r01 = fma2(q01, back, a01), r23 = fma2(q23, back, a23);       // exact residual * 2^-42
    float q0, q1, q2, q3, r0, r1, r2, r3;
    unpack2(q01, q0, q1); unpack2(q23, q2, q3);
    unpack2(r01, r0, r1); unpack2(r23, r2, r3);
    unsigned word = pack_sat_u8(__float_as_int(q3), __float_as_int(q2), 0u);
    word = pack_sat_u8(__float_as_int(q1), __float_as_int(q0), word);
    const float worst = fmaxf(fmaxf(fabsf(r0), fabsf(r1)), fmaxf(fabsf(r2), fabsf(r3)));
    if (worst >= v_thr) {
        // within the float error band of a rounding boundary: redo those samples in int32
        const float r[4] = {r0, r1, r2, r3};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (fabsf(r[k]) >= v_thr) {
                int a = 0;
#pragma unroll
                for (int s = 0; s < VT; ++s) a += subnormal_to_int(quad_elem(win[s], k)) * __ldg(exact_coef + s);
                const unsigned v = static_cast<unsigned>(min(max((a + (1 << 18)) >> 19, 0), 255));
                word = (word & ~(0xffu << (8 * k))) | (v << (8 * k));
            }
        }
    }
    return word;
}

__device__ __forceinline__ void store_quad_slow(uint8_t *row, int x, int dw, int xstep, unsigned word) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (x + k < dw) row[static_cast<ptrdiff_t>(x + k) * xstep] = static_cast<uint8_t>(word >> (8 * k));
}

__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(static_cast<unsigned>(__cvta_generic_to_shared(smem_dst))), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<unsigned>(__cvta_generic_to_shared(smem_dst))), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

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

// First byte of the tile's source footprint (sample c0 of row r0 of frame f).
__device__ __forceinline__ const uint8_t *tile_src_base(const FusedOpParams &op, const TileCoord &c) {
    return op.src + c.f * op.src_fstride + static_cast<ptrdiff_t>(op.tile_r0[c.ty]) * op.src_pitch +
           static_cast<ptrdiff_t>(op.tile_c0[c.tx]) * op.src_xstep;
}

// Start the asynchronous copy of everything tile `c` needs from global memory: the raw source footprint (as
// aligned 32-bit words), the ring-ordered row coefficients and the row start positions.
template <int VT>
__device__ __forceinline__ void prefetch_tile(const FusedLaunchParams &P, const TileCoord &c, unsigned *raw, f32x2 *cv,
                                              int *vp, int tid) {
    const FusedOpParams &op = P.op[c.oi];
    const int c0 = op.tile_c0[c.tx], ncols = op.tile_nc[c.tx];
    const int nrows = op.tile_r1[c.ty] - op.tile_r0[c.ty];
    const int xs = op.src_xstep;
    const int nv = min(ncols, op.sw - c0);
    const int wlog = op.s_words_log2, wmask = (1 << wlog) - 1, s_words = op.s_words;
    const uint8_t *src_base = tile_src_base(op, c);
    const ptrdiff_t src_pitch = op.src_pitch;
    const int total = nrows << wlog;
    for (int idx = tid; idx < total; idx += kThreads) {
        const int row = idx >> wlog, w = idx & wmask;
        const uint8_t *rp = src_base + row * src_pitch;
        const unsigned mis = static_cast<unsigned>(reinterpret_cast<uintptr_t>(rp)) & 3u;
        const int nwords = (static_cast<int>(mis) + (nv - 1) * xs) / 4 + 1;
        if (w < nwords) cp_async4(raw + row * s_words + w, rp - mis + 4 * w);
    }
    const int y0 = c.ty * op.th;
    const char *cvg = reinterpret_cast<const char *>(op.vring) + static_cast<size_t>(y0) * VT * sizeof(f32x2);
    for (int i = tid; i < op.th * VT / 2; i += kThreads) cp_async16(reinterpret_cast<char *>(cv) + 16 * i, cvg + 16 * i);
    const char *vpg = reinterpret_cast<const char *>(op.vpos + y0);
    for (int i = tid; i < op.th / 4; i += kThreads) cp_async16(reinterpret_cast<char *>(vp) + 16 * i, vpg + 16 * i);
    cp_async_commit();
}

// Persistent kernel: each CTA walks a contiguous range of tiles.  While tile i is filtered, the source bytes and
// row tables of tile i+1 stream into the other half of a double buffer (cp.async), so no phase waits on HBM.
template <int SPAN, int VT>
__global__ void __launch_bounds__(kThreads, 2)
fused_scale_kernel(const __grid_constant__ FusedLaunchParams P) {
    static_assert(SPAN % 2 == 0, "the packed horizontal pass pairs up source samples");
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int half = warp >> 2;

    float *M = smem;                                             // [m_rows + VT][256] intermediate t * 2^-149
    float *S = M + P.m_floats;                                   // [m_rows][s_pitch]  source samples as floats
    unsigned *RAW = reinterpret_cast<unsigned *>(S + P.s_floats);   // 2 x [m_rows][s_words] raw source words
    f32x2 *CVB = reinterpret_cast<f32x2 *>(RAW + 2 * P.raw_words);  // 2 x [th + 1][VT] ring-ordered (g, g) pairs
    int *VPB = reinterpret_cast<int *>(CVB + 2 * P.cv_pairs);       // 2 x [th]      first source row per output row

    // contiguous share of the tile list
    const int per = P.total_tiles / gridDim.x, extra = P.total_tiles % gridDim.x;
    const int t_begin = blockIdx.x * per + min(static_cast<int>(blockIdx.x), extra);
    const int t_end = t_begin + per + (static_cast<int>(blockIdx.x) < extra ? 1 : 0);
    if (t_begin >= t_end) return;

    f32x2 cd[kGroup][SPAN / 2];      // this lane's dense horizontal coefficients, (c[i], c[i+1]) pairs
    int hb = 0;                      // first source column of this lane's span
    bool h_active = false;
    int cd_oi = -1, cd_tx = -1;

    int buf = 0;
    TileCoord cur = decode_tile(P, t_begin);
    prefetch_tile<VT>(P, cur, RAW, CVB, VPB, tid);

    for (int tile = t_begin; tile < t_end; ++tile, buf ^= 1) {
        const FusedOpParams &op = P.op[cur.oi];
        cp_async_wait_all();
        __syncthreads();             // tile's inputs have landed; everyone is done with the previous tile's M / S
        TileCoord nxt = cur;
        if (tile + 1 < t_end) {
            nxt = decode_tile(P, tile + 1);
            prefetch_tile<VT>(P, nxt, RAW + (buf ^ 1) * P.raw_words, CVB + (buf ^ 1) * P.cv_pairs, VPB + (buf ^ 1) * P.vp_ints, tid);
        }

        const int th = op.th, dw = op.dw, n_groups = op.n_groups;
        const int x0 = cur.tx * kTileW;
        const int y0 = cur.ty * th, y1 = min(y0 + th, op.dh);
        const int c0 = op.tile_c0[cur.tx], ncols = op.tile_nc[cur.tx];
        const int r0 = op.tile_r0[cur.ty], r1 = op.tile_r1[cur.ty];
        const int nrows = r1 - r0;
        const int s_pitch = op.s_pitch, xs = op.src_xstep;
        const uint8_t *src_base = tile_src_base(op, cur);
        const ptrdiff_t src_pitch = op.src_pitch;
        const unsigned *raw = RAW + buf * P.raw_words;
        const f32x2 *CV = CVB + buf * P.cv_pairs;
        const int *VP = VPB + buf * P.vp_ints;

        // ---- a new tile column: fetch this lane's horizontal coefficients (kept across the column's tiles) ----
        if (cur.oi != cd_oi || cur.tx != cd_tx) {
            cd_oi = cur.oi; cd_tx = cur.tx;
            const int gb = x0 / kGroup + half * kGroupsPerHalf;
            h_active = gb < n_groups;
            const int g = min(gb + lane, n_groups - 1);
            const float *hd = op.hdense + static_cast<size_t>(g) * (kGroup * SPAN);
            f32x2 *flat = &cd[0][0];
            if constexpr ((kGroup * SPAN) % 4 == 0) {
#pragma unroll
                for (int i = 0; i < kGroup * SPAN / 4; ++i) {
                    const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2 *>(hd) + i);
                    flat[2 * i] = v.x; flat[2 * i + 1] = v.y;
                }
            } else {
#pragma unroll
                for (int i = 0; i < kGroup * SPAN / 2; ++i) flat[i] = __ldg(reinterpret_cast<const f32x2 *>(hd) + i);
            }
            hb = __ldg(op.hbase + g) - c0;
        }

        // ---- raw words -> float samples.  The sample that lands in S[row][j] is sample j - (mis / xs) of the footprint,
        // where mis is the misalignment of the row's first byte (rows are read as aligned words). ----
        {
            const int nv = min(ncols, op.sw - c0);    // samples that exist; the rest carry zero weight
            const int wlog = op.s_words_log2, wmask = (1 << wlog) - 1, s_words = op.s_words;
            const int total = nrows << wlog;
            for (int idx = tid; idx < total; idx += kThreads) {
                const int row = idx >> wlog, w = idx & wmask;
                if (w >= s_words) continue;
                const unsigned mis = static_cast<unsigned>(reinterpret_cast<uintptr_t>(src_base + row * src_pitch)) & 3u;
                const int nwords = (static_cast<int>(mis) + (nv - 1) * xs) / 4 + 1;
                const unsigned v = w < nwords ? raw[row * s_words + w] : 0u;
                if (xs == 1) {
                    float4 o;
                    o.x = byte_to_float(v, 0x7540u); o.y = byte_to_float(v, 0x7541u);
                    o.z = byte_to_float(v, 0x7542u); o.w = byte_to_float(v, 0x7543u);
                    *reinterpret_cast<float4 *>(S + row * s_pitch + 4 * w) = o;
                } else {
                    const unsigned p = mis & 1u;
                    float2 o;
                    o.x = byte_to_float(v, 0x7540u | p); o.y = byte_to_float(v, 0x7542u | p);
                    *reinterpret_cast<float2 *>(S + row * s_pitch + 2 * w) = o;
                }
            }
            // rows the vertical ring may touch beyond the last real row read as zero
            for (int i = tid; i < VT * (kTileW / 4); i += kThreads)
                reinterpret_cast<float4 *>(M + nrows * kTileW)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        __syncthreads();

        // ---- horizontal pass: S -> M ----------------------------------------------------------------
        if (h_active) {
            const float t_max = __int_as_float(32767);
            float *mcol = M + half * kHalfW + 4 * lane;
            for (int row = warp & 3; row < nrows; row += 4) {
                const unsigned mis = static_cast<unsigned>(reinterpret_cast<uintptr_t>(src_base + row * src_pitch)) & 3u;
                const float *sp = S + row * s_pitch + hb + (xs == 1 ? mis : (mis >> 1));
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
                    ov[k] = fminf(fma_rm(even + odd, 0x1p-126f, 0.0f), t_max);   // floor(sum / 128), clamped above
                }
                *reinterpret_cast<float4 *>(mcol + row * kTileW) = o;
            }
        }
        __syncthreads();

        // ---- vertical pass: M -> destination --------------------------------------------------------
        const int strip = th / kWarps;
        const int ya = y0 + warp * strip, yb = min(ya + strip, y1);
        if (ya < yb) {
            const bool right = x0 + kHalfW < dw;        // the right half of the last tile may be empty
            const int xa = x0 + 4 * lane, xb = xa + kHalfW;
            const int dst_xstep = op.dst_xstep;
            const ptrdiff_t dst_pitch = op.dst_pitch;
            const float v_thr = op.v_thr;
            uint8_t *drow = op.dst + cur.f * op.dst_fstride + static_cast<ptrdiff_t>(ya) * dst_pitch;
            // 32-bit stores need every row of this tile column to start on a 4-byte boundary
            const bool word_ok = dst_xstep == 1 && ((reinterpret_cast<uintptr_t>(drow) + x0) & 3u) == 0 && (dst_pitch & 3) == 0;
            const bool fast_a = word_ok && xa + 3 < dw, fast_b = word_ok && xb + 3 < dw;
            const int *exact = op.vring_i + static_cast<size_t>(ya) * VT;
            const f32x2 *cv = CV + (ya - y0) * VT;
            const int *vp = VP + (ya - y0);
            const float *m_col = M + 4 * lane;
            Quad wa[VT], wb[VT];
#pragma unroll
            for (int s = 0; s < VT; ++s) wa[s] = wb[s] = Quad{0ull, 0ull};
            int first = vp[0];
            int next_row = first;
            int slot = next_row % VT;
            const float *mp = m_col + (next_row - r0) * kTileW;
            f32x2 gg[VT];
#pragma unroll
            for (int s = 0; s < VT; ++s) gg[s] = cv[s];
            for (int y = ya; y < yb; ++y, drow += dst_pitch, exact += VT) {
                // next row's start position and coefficients are requested now and consumed next iteration
                // (both tables carry one padding row)
                ++vp; cv += VT;
                const int first_next = *vp;
                f32x2 gg_next[VT];
                if constexpr (VT % 2 == 0) {
#pragma unroll
                    for (int s = 0; s < VT; s += 2) {
                        const ulonglong2 c2 = *reinterpret_cast<const ulonglong2 *>(cv + s);
                        gg_next[s] = c2.x; gg_next[s + 1] = c2.y;
                    }
                } else {
#pragma unroll
                    for (int s = 0; s < VT; ++s) gg_next[s] = cv[s];
                }
                if (next_row < first) {   // steps larger than the ring: skip rows nobody needs
                    next_row = first; slot = first % VT; mp = m_col + (first - r0) * kTileW;
                }
                if (next_row < first + VT) load_ring_rows<VT>(wa, wb, slot, next_row, first + VT, mp);
                const unsigned word_a = vertical_quad<VT>(wa, gg, v_thr, exact);
                if (fast_a) *reinterpret_cast<unsigned *>(drow + xa) = word_a;
                else store_quad_slow(drow, xa, dw, dst_xstep, word_a);
                if (right) {
                    const unsigned word_b = vertical_quad<VT>(wb, gg, v_thr, exact);
                    if (fast_b) *reinterpret_cast<unsigned *>(drow + xb) = word_b;
                    else store_quad_slow(drow, xb, dw, dst_xstep, word_b);
                }
                first = first_next;
#pragma unroll
                for (int s = 0; s < VT; ++s) gg[s] = gg_next[s];
            }
        }
        cur = nxt;
    }
}

using KernelFn = void (*)(const FusedLaunchParams);
struct Instance { int span, vt; KernelFn fn; };
#define TVS_INST(S, V) {S, V, fused_scale_kernel<S, V>}
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
    int th = 0, tiles_x = 0, tiles_y = 0, n_groups = 0, s_pitch = 0, s_words = 0, s_words_log2 = 0, m_rows = 0;
    size_t smem_bytes = 0;
    float v_thr = 0;
    // offsets (in bytes) into the device blob
    size_t hdense_off = 0, hbase_off = 0, vpos_off = 0, vring_off = 0, vring_i_off = 0;
    int vtaps = 0;
    std::vector<short> tile_c0, tile_nc, tile_r0, tile_r1;
};

struct FusedPlan {
    std::vector<FusedOp> ops;
    uint8_t *blob = nullptr;
    size_t blob_bytes = 0;
    int device = 0;
    int sm_count = 148;
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
SmemLayout smem_layout(int m_rows, int s_pitch, int s_words, int th, int vt) {
    SmemLayout l;
    l.m_floats = (m_rows + vt) * kTileW;
    l.s_floats = m_rows * s_pitch;
    l.raw_words = (m_rows * s_words + 3) & ~3;
    l.cv_pairs = ((th + 1) * vt + 1) & ~1;
    l.vp_ints = (th + 1 + 3) & ~3;
    return l;
}
size_t layout_bytes(const SmemLayout &l) {
    return static_cast<size_t>(l.m_floats + l.s_floats + 2 * l.raw_words + 2 * l.vp_ints) * 4 + static_cast<size_t>(2 * l.cv_pairs) * 8;
}
size_t fused_smem_bytes(int m_rows, int s_pitch, int s_words, int th, int vt) {
    return layout_bytes(smem_layout(m_rows, s_pitch, s_words, th, vt));
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

FusedPlan *fused_plan_create(const Plan &plan, const AxisTable host_tables[4], int device, std::string &err) {
    std::vector<uint8_t> blob;
    auto fp = new FusedPlan;
    fp->device = device;
    cudaDeviceGetAttribute(&fp->sm_count, cudaDevAttrMultiProcessorCount, device);
    for (size_t oi = 0; oi < plan.ops.size(); ++oi) {
        const PlaneOp &op = plan.ops[oi];
        if (op.kind != kOpScale) continue;
        if (op.src_bits != 8 || op.dst_bits != 8 || op.rule != kOut8) { err = "only 8-bit to 8-bit so far"; delete fp; return nullptr; }
        const AxisTable &h = host_tables[op.htable], &v = host_tables[op.vtable];

        // float exactness of the horizontal sum: rows must sum to 2^14 and 255 * sum|c| must stay below 2^24
        const int n_groups = (op.dw + kGroup - 1) / kGroup;
        int span = 0;
        long long worst_abs = 0;
        for (int x = 0; x < op.dw; ++x) {
            long long s = 0, a = 0;
            for (int j = 0; j < h.taps; ++j) { const int c = h.coef[static_cast<size_t>(x) * h.taps + j]; s += c; a += std::abs(c); }
            if (s != (1 << 14)) { err = "horizontal row does not sum to one"; delete fp; return nullptr; }
            worst_abs = std::max(worst_abs, a);
        }
        if (255 * worst_abs >= (1LL << 24)) { err = "horizontal sum not exact in fp32"; delete fp; return nullptr; }
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
        const double e = vertical_error_bound(v, VT);
        if (e > 4096.0) { err = "vertical error bound too large"; delete fp; return nullptr; }

        FusedOp fo;
        fo.op_index = static_cast<int>(oi);
        fo.instance = inst;
        fo.n_groups = n_groups;
        fo.vtaps = v.taps;
        fo.tiles_x = (op.dw + kTileW - 1) / kTileW;
        fo.v_thr = static_cast<float>(std::ldexp(static_cast<double>(1 << 18) - e - 1.0, -42));

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
                    if (c) hdense[(static_cast<size_t>(g) * kGroup + k) * SPAN + (h.pos[x] - hbase[g] + j)] = std::ldexp(static_cast<float>(c), -30);
                }
            }
        }
        int ncols_max = 0;
        for (int tx = 0; tx < fo.tiles_x; ++tx) {
            const int ga = tx * 2 * kGroupsPerHalf, gb = std::min(ga + 2 * kGroupsPerHalf - 1, n_groups - 1);
            ncols_max = std::max(ncols_max, hbase[gb] + SPAN - hbase[ga]);
        }
        // words per staged row for the worst alignment, and the float pitch that holds them
        fo.s_words = (3 + (ncols_max - 1) * op.src_xstep) / 4 + 1;
        fo.s_words_log2 = 0;
        while ((1 << fo.s_words_log2) < fo.s_words) ++fo.s_words_log2;
        fo.s_pitch = (fo.s_words * (4 / op.src_xstep) + 3) & ~3;   // rows stay 16-byte aligned
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
            fo.smem_bytes = fused_smem_bytes(fo.m_rows, fo.s_pitch, fo.s_words, th, VT);
            if (fo.smem_bytes <= kSmemBudget || th == kWarps) break;
        }
        if (fo.smem_bytes > 220 * 1024) { err = "tile does not fit in shared memory"; delete fp; return nullptr; }
        fo.th = th;
        if (fo.tiles_x > kMaxTilesX || fo.tiles_y > kMaxTilesY) { err = "too many tiles for the descriptor table"; delete fp; return nullptr; }

        // vertical coefficients indexed by ring slot, and start rows; both padded to whole tiles so the kernel
        // can copy a tile's worth without bounds checks
        const int padded_rows = fo.tiles_y * th;
        std::vector<float> vring(static_cast<size_t>(padded_rows) * VT * 2, 0.0f);   // (g, g) pairs
        std::vector<int> vring_i(static_cast<size_t>(padded_rows) * VT, 0);
        std::vector<int> vpos_padded(padded_rows, v.pos[op.dh - 1]);
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
            const int ga = tx * 2 * kGroupsPerHalf, gb = std::min(ga + 2 * kGroupsPerHalf - 1, n_groups - 1);
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
        fp->ops.push_back(fo);
    }
    if (fp->ops.empty()) { err = "nothing to scale"; delete fp; return nullptr; }
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
        if (fo.smem_bytes > 48 * 1024)
            cudaFuncSetAttribute(reinterpret_cast<const void *>(kInstances[fo.instance].fn),
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    }
    return fp;
}

void fused_plan_destroy(FusedPlan *p) {
    if (!p) return;
    if (p->blob) cudaFree(p->blob);
    delete p;
}

size_t fused_plan_device_bytes(const FusedPlan *p) { return p ? p->blob_bytes : 0; }

int fused_launch(const FusedPlan *p, const Plan &plan, const SampleView src_views[], const SampleView dst_views[],
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
            if (done[j] || fo.instance != p->ops[i].instance) continue;
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
            q.s_pitch = fo.s_pitch; q.s_words = fo.s_words; q.s_words_log2 = fo.s_words_log2; q.m_rows = fo.m_rows;
            q.v_thr = fo.v_thr;
            std::copy(fo.tile_c0.begin(), fo.tile_c0.end(), q.tile_c0);
            std::copy(fo.tile_nc.begin(), fo.tile_nc.end(), q.tile_nc);
            std::copy(fo.tile_r0.begin(), fo.tile_r0.end(), q.tile_r0);
            std::copy(fo.tile_r1.begin(), fo.tile_r1.end(), q.tile_r1);
            tiles += fo.tiles_x * fo.tiles_y;
            const SmemLayout l = smem_layout(fo.m_rows, fo.s_pitch, fo.s_words, fo.th, vt);
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
        const size_t smem = layout_bytes(lay);
        // persistent CTAs: two per SM, each walking a contiguous share of the tiles
        const int grid = std::min(lp.total_tiles, 2 * p->sm_count);
        kInstances[p->ops[i].instance].fn<<<grid, kThreads, smem, stream>>>(lp);
        if (cudaPeekAtLastError() != cudaSuccess) return -1;
        ++launched;
    }
    return launched;
}

}  // namespace tvs
