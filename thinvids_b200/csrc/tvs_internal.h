// Internal types shared by the C-ABI layer and the CUDA kernels.  Not installed; the public surface
// is include/tvscale.h.
#pragma once
#include <cuda_runtime.h>

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <algorithm>
#include <cstdlib>
#include <string>
#include <vector>

#include "filter_tables.h"

namespace tvs {

// Where the samples of ONE logical plane (Y, U or V) of frame 0 live.  NV12/P010 chroma is one stored
// plane seen through two views that differ in `base` by one sample and step over the other component.
struct SampleView {
    uint8_t *base = nullptr;     // first sample of row 0, frame 0
    ptrdiff_t pitch = 0;         // bytes between rows
    ptrdiff_t frame_stride = 0;  // bytes between frames
    int xstep = 1;               // bytes between consecutive samples of this logical plane
};

// Device-resident copy of one AxisTable.
struct DeviceTable {
    const int16_t *coef = nullptr;
    const int32_t *pos = nullptr;
    int taps = 0;
    int length = 0;
};

// How the vertical pass turns its 32-bit sum into an output sample.
enum OutputRule : int {
    kOut8 = 0,        // 8-bit source -> 8-bit: (acc + (64<<12)) >> 19, clip to 0..255
    kOut8Dither = 1,  // >8-bit source -> 8-bit: ordered 8x8 dither instead of the constant 64
    kOut10 = 2        // -> 10-bit: (acc + (1<<16)) >> 17, clip to 0..1023 (P010 then shifts left by 6)
};

enum PlaneOpKind : int {
    kOpScale = 0,   // horizontal + vertical filter (also used with 1-tap tables for depth changes)
    kOpCopy = 1,    // same-size, same-depth sample copy (re-interleave / de-interleave)
    kOpDither2 = 2, // same-size 10 -> 8 bit copy with libswscale's 2-bit ordered dither
    // same-size 10 -> 8 bit through the scaler proper when both of its filters are the 1-tap identity: per sample
    // t = min(32 v, 32767), out = min((t + D128[y & 7][(x + dither_xoff) & 7]) >> 7, 255)  (SURVEY.md A.3 with one tap)
    kOpDither128 = 3
};

// One unit of work of a frame: produce one logical (or, for the P010->NV12 copy, stored) plane.
struct PlaneOp {
    PlaneOpKind kind = kOpScale;
    int src_plane = 0, dst_plane = 0;      // index into the caller's plane pointer arrays
    int src_offset = 0, dst_offset = 0;    // byte offset of this component inside an interleaved sample group
    int src_xstep = 1, dst_xstep = 1;      // bytes between samples
    int src_bits = 8, dst_bits = 8;        // significant bits
    int src_shift = 0, dst_shift = 0;      // P010 keeps its 10 bits in the top of the word
    int sw = 0, sh = 0, dw = 0, dh = 0;    // logical plane sizes
    int htable = 0, vtable = 0;            // which DeviceTable (0 hLum, 1 hChr, 2 vLum, 3 vChr)
    int hshift = 7;                        // right shift after the horizontal sum
    OutputRule rule = kOut8;
    int dither_xoff = 0;                   // 3 for the V component when dithering
    int rc_coeff = 0, rc_offset = 0;       // range conversion of the intermediate: t' = min((t * rc_coeff + rc_offset) >> 14, 32767); 0 = none
    int raw_tail_from = -1;                // kOpDither2 only: columns >= this use the un-shifted P010 word
    size_t mid_offset = 0;                 // element offset of this plane's intermediate inside a frame's scratch block
};

struct Plan {
    std::vector<PlaneOp> ops;
    size_t mid_elems_per_frame = 0;        // int16 elements of two-pass scratch per frame
    bool any_scale = false;
};

// ---- the float vertical sum (shared by the tile and the streaming kernels) ---------------------------------
// Both kernels accumulate acc * 2^-42 in FP32 and take q = RN(acc / 2^shift).  Two regimes:
//  * general tables: every FMA may round; the accumulated error is at most `e` units of acc, so a sample whose exact
//    residual is within e of a rounding boundary (|r| >= thr) is recomputed in int32;
//  * tables whose coefficients are all multiples of 2^k with the largest partial sum below 2^(23+k) (area 2:1, any
//    1-tap axis, ...): the float sum is exact, but exact ties (acc = odd * 2^(shift-1)) are then COMMON (1 sample in
//    256 for a 2:1 box filter) and RN-to-even is not libswscale's round-half-up.  Starting the accumulator at half a
//    unit, 2^(k-1), makes ties impossible and turns RN into round-half-up; no residual check is needed (thr = huge).
struct VerticalFloatRule {
    bool ok;      // false: error bound too large for this output rule, use the two-pass kernels
    float thr;    // residual threshold, scaled by 2^-42
    float init;   // accumulator start value, scaled by 2^-42
};
inline VerticalFloatRule vertical_float_rule(const AxisTable &v, int vt_inst, bool out10, bool dither) {
    long long worst = 0;
    int k = 31;
    for (int y = 0; y < v.length; ++y) {
        long long s = 0;
        for (int j = 0; j < v.taps; ++j) {
            const int c = std::abs(static_cast<int>(v.coef[static_cast<size_t>(y) * v.taps + j]));
            s += c;
            if (c) k = std::min(k, __builtin_ctz(static_cast<unsigned>(c)));
        }
        worst = std::max(worst, s);
    }
    const double bound = 32767.0 * static_cast<double>(worst) + 1.0;
    const int bits = static_cast<int>(std::floor(std::log2(bound))) + 1;   // bits of the largest partial sum
    const double half_step = out10 ? 65536.0 : 262144.0;                    // half of 2^17 / 2^19
    if (dither) k = std::min(k, 13);                                        // dither offsets are multiples of 2^13
    if (k >= 1 && k < 31 && bits - k + 1 <= 24) return {true, 3.0e38f, std::ldexp(1.0f, k - 1 - 42)};
    const double e = 0.5 * std::ldexp(1.0, bits - 1 - 23) * vt_inst;        // half an ulp of the bound per FMA
    if (e > half_step / 8) return {false, 0.f, 0.f};
    return {true, static_cast<float>(std::ldexp(half_step - e - 1.0, -42)), 0.0f};
}

// The same rule for the streaming kernel, whose vertical sum runs in tap order with even and odd taps in separate chains
// (joined by one addition) when there are at least 4 tap slots.  Each FMA's rounding error is at most half an ulp of ITS OWN
// partial sum, bounded by 32767 * (sum of |g| accumulated so far in that chain): following the actual order row by row gives
// a bound about a third of "taps * half an ulp of the largest possible sum" for lanczos / bicubic tables, and the exact path
// runs that much less often.  `init_abs` bounds the accumulator's start value (ordered dither), in units of the integer sum.
inline VerticalFloatRule vertical_float_rule_split(const AxisTable &v, int vt_inst, bool out10, bool dither) {
    const VerticalFloatRule coarse = vertical_float_rule(v, vt_inst + 1, out10, dither);
    if (!coarse.ok || coarse.thr > 1.0e38f) return coarse;   // exact tables keep their own treatment
    auto half_ulp = [](double bound) { return bound < 1.0 ? 0.0 : 0.5 * std::ldexp(1.0, static_cast<int>(std::floor(std::log2(bound))) + 1 - 24); };
    const double init_abs = dither ? 64.0 * 4096.0 : 0.0;
    const bool split = vt_inst >= 4;
    double e_max = 0.0;
    for (int y = 0; y < v.length; ++y) {
        double run[2] = {init_abs, 0.0}, e = 0.0;
        for (int j = 0; j < v.taps; ++j) {
            const int c = std::abs(static_cast<int>(v.coef[static_cast<size_t>(y) * v.taps + j]));
            if (!c) continue;   // a zero coefficient leaves the accumulator unchanged, exactly
            double &r = run[split ? (j & 1) : 0];
            r += 32767.0 * c;
            e += half_ulp(r);
        }
        if (split) e += half_ulp(run[0] + run[1]);
        e_max = std::max(e_max, e);
    }
    const double half_step = out10 ? 65536.0 : 262144.0;
    if (e_max > half_step / 8) return {false, 0.f, 0.f};
    return {true, static_cast<float>(std::ldexp(half_step - e_max - 1.0, -42)), coarse.init};
}

// ---- kernel launchers (kernels_generic.cu) -------------------------------------------------------
// Each returns the number of kernels it launched, or -1 after a launch error (cudaGetLastError kept).
int launch_hpass_generic(const PlaneOp &op, const SampleView &src, const DeviceTable &h, int16_t *mid,
                         size_t mid_frame_stride, int n_frames, cudaStream_t stream);
int launch_vpass_generic(const PlaneOp &op, const SampleView &dst, const DeviceTable &v, const int16_t *mid,
                         size_t mid_frame_stride, int n_frames, cudaStream_t stream);
int launch_repack(const PlaneOp &op, const SampleView &src, const SampleView &dst, int n_frames, cudaStream_t stream);
// memcpy-class special cases of the same-size converters
int launch_copy_rows(const SampleView &src, const SampleView &dst, int row_bytes, int rows, int n_frames, cudaStream_t stream);
int launch_p010_copy(const SampleView &src, const SampleView &dst, int words, int rows, int n_frames, cudaStream_t stream);
// vector forms of kOpDither2 (contiguous 16-bit plane -> byte plane; planar U + V -> interleaved bytes); they return 0
// without launching when the surfaces are not aligned for them
int launch_dither2_rows(const PlaneOp &op, const SampleView &src, const SampleView &dst, int n_frames, cudaStream_t stream);
int launch_dither2_interleave(const PlaneOp &op, const SampleView &u, const SampleView &v, const SampleView &dst, int n_frames,
                              cudaStream_t stream);
// kOpDither128 for planar U + V -> interleaved bytes (returns 0 when the surfaces are not aligned for it)
int launch_dither128_interleave(const PlaneOp &op, const SampleView &u, const SampleView &v, const SampleView &dst, int n_frames,
                                cudaStream_t stream);
// libswscale's input stage for packed rgb24 (rgb24ToY_c, rgb24ToUV_c / rgb24ToUV_half_c, BT.601 limited range): per frame
// `planes` receives Y [h][w], then U and V [h][cw] as uint16, cw = w / 2 (half) or w
int launch_rgb_input(const uint8_t *rgb, ptrdiff_t pitch, ptrdiff_t frame_stride, uint8_t *planes, size_t planes_frame_stride,
                     size_t u_offset_bytes, size_t c_plane_bytes, int w, int h, bool half, bool bgr, int wide, int n_frames, cudaStream_t stream);
// exact 2:1 area down-scale of an 8-bit plane (pair = interleaved UV bytes)
int launch_box2(const SampleView &src, const SampleView &dst, bool pair, int out_bytes, int rows, int n_frames, cudaStream_t stream);
int launch_interleave(const SampleView &u, const SampleView &v, const SampleView &dst, int sample_bytes, int w, int rows, int n_frames,
                      cudaStream_t stream);

// ---- fused tile kernel (kernels_fused_impl.cuh, built as two variants) ---------------------------------
struct FusedPlan;  // opaque, owned by the context
#define TVS_DECLARE_FUSED(SFX)                                                                                              \
    FusedPlan *fused_plan_create##SFX(const Plan &plan, const AxisTable host_tables[4], int device, std::string &err);      \
    void fused_plan_destroy##SFX(FusedPlan *p);                                                                             \
    size_t fused_plan_device_bytes##SFX(const FusedPlan *p);                                                                \
    bool fused_plan_prefers_two_pass##SFX(const FusedPlan *p);                                                              \
    /* Launches the fused kernels for every scale op of the plan.  Returns kernels launched or -1. */                       \
    int fused_launch##SFX(const FusedPlan *p, const Plan &plan, const SampleView src_views[], const SampleView dst_views[], \
                          int n_frames, cudaStream_t stream);
TVS_DECLARE_FUSED(_w256)   // 256-column tiles, 2 CTAs x 256 threads per SM
TVS_DECLARE_FUSED(_w128)   // 128-column tiles, 4 CTAs x 128 threads per SM
#undef TVS_DECLARE_FUSED

// ---- streaming down-scale kernel (kernels_stream.cu) ------------------------------------------------------
struct StreamPlan;  // opaque, owned by the context
StreamPlan *stream_plan_create(const Plan &plan, const AxisTable host_tables[4], int device, std::string &err);
void stream_plan_destroy(StreamPlan *p);
// Returns kernels launched, or a negative value after a launch error.
int stream_launch(const StreamPlan *p, const Plan &plan, const SampleView src_views[], const SampleView dst_views[], int n_frames,
                  cudaStream_t stream);

// ---- bwdif de-interlacer (kernels_bwdif.cu): one plane of a batch; returns kernels launched or -1 ----
int launch_bwdif_plane(const uint8_t *src, ptrdiff_t src_pitch, ptrdiff_t src_fstride, uint8_t *dst, ptrdiff_t dst_pitch,
                       ptrdiff_t dst_fstride, int w, int h, int depth, int n_frames, int tff, bool stream_start, bool stream_end,
                       cudaStream_t stream);

// ---- slab kernel (kernels_slab.cu): warp-specialised row streaming for planes that are not shrunk vertically ----
struct SlabPlan;  // opaque, owned by the context
SlabPlan *slab_plan_create(const Plan &plan, const AxisTable host_tables[4], int device, std::string &err);
void slab_plan_destroy(SlabPlan *p);
size_t slab_plan_device_bytes(const SlabPlan *p);
// Returns kernels launched, or a negative value after a launch error.
int slab_launch(const SlabPlan *p, const Plan &plan, const SampleView src_views[], const SampleView dst_views[], int n_frames,
                cudaStream_t stream);

}  // namespace tvs
