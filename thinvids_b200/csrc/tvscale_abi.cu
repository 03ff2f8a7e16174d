// C ABI of the scale / pixel-format stage (include/tvscale.h): context life-cycle, launch planning,
// the device-pointer batch call and the host-buffer drop-in for sws_scale().
#include <cerrno>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/tvscale.h"
#include "tvs_internal.h"

using namespace tvs;

namespace {

thread_local std::string g_last_error;

int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

int cuda_fail(cudaError_t e, const char *what) {
    return fail(-EIO, "%s: %s", what, cudaGetErrorString(e));
}

#define TVS_CUDA(call, what)                                    \
    do {                                                        \
        cudaError_t e_ = (call);                                \
        if (e_ != cudaSuccess) return cuda_fail(e_, what);      \
    } while (0)

struct FormatInfo {
    int bits;       // significant bits per sample
    int bytes;      // bytes per sample
    int shift;      // left shift of the significant bits inside the word (P010: 6)
    bool semi;      // chroma interleaved in one plane
    int planes;     // stored planes
    bool rgb;       // one plane of packed R, G, B bytes (source only)
    bool bgr = false;   // ... stored B, G, R
    bool be = false;    // 16-bit components stored big-endian (rgb48be)
};

bool format_info(int fmt, FormatInfo &fi) {
    switch (fmt) {
    case TVS_FMT_YUV420P: fi = {8, 1, 0, false, 3, false}; return true;
    case TVS_FMT_NV12: fi = {8, 1, 0, true, 2, false}; return true;
    case TVS_FMT_YUV420P10LE: fi = {10, 2, 0, false, 3, false}; return true;
    case TVS_FMT_P010LE: fi = {10, 2, 6, true, 2, false}; return true;
    case TVS_FMT_RGB24: fi = {8, 1, 0, false, 1, true}; return true;
    case TVS_FMT_BGR24: fi = {8, 1, 0, false, 1, true, true}; return true;
    case TVS_FMT_RGB48LE: fi = {16, 2, 0, false, 1, true, false, false}; return true;
    case TVS_FMT_RGB48BE: fi = {16, 2, 0, false, 1, true, false, true}; return true;
    case TVS_FMT_BGR48LE: fi = {16, 2, 0, false, 1, true, true, false}; return true;
    case TVS_FMT_BGR48BE: fi = {16, 2, 0, false, 1, true, true, true}; return true;
    default: return false;
    }
}

size_t tight_frame_bytes(const FormatInfo &fi, int w, int h) {
    if (fi.rgb) return static_cast<size_t>(3) * w * h * fi.bytes;
    const size_t cw = (w + 1) / 2, ch = (h + 1) / 2;
    return (static_cast<size_t>(w) * h + 2 * cw * ch) * fi.bytes;
}

}  // namespace

struct tvs_ctx {
    int device = 0;
    int src_w = 0, src_h = 0, dst_w = 0, dst_h = 0, src_fmt = 0, dst_fmt = 0, flags = 0;
    // Range conversion (libswscale's lumConvertRange / chrConvertRange on the 15-bit intermediate): 0 none, 1 full ->
    // limited, 2 limited -> full.  Constants {luma, chroma} as ff_sws_init_range_convert derives them (oracle/swscale_port.c).
    int range_convert = 0;
    int rc_coeff[2] = {0, 0}, rc_offset[2] = {0, 0};
    FormatInfo sfi{}, dfi{};
    AxisTable host_tables[4];
    DeviceTable dev_tables[4];
    void *dev_table_blob = nullptr;
    Plan plan;
    FusedPlan *fused = nullptr;
    bool fused_narrow = false;   // which build of the fused kernel owns `fused`: 128-column (true) or 256-column tiles
    StreamPlan *streaming = nullptr;   // vertical down-scales: the line-buffer kernel
    SlabPlan *slab = nullptr;          // planes that are not shrunk vertically: the warp-specialised slab kernel
    bool box2 = false;                 // exact 2:1 area on 8-bit planes with matching layouts: the box-filter kernel
    int forced_path = -1;
    // two-pass scratch
    int16_t *mid = nullptr;
    size_t mid_capacity_elems = 0;
    // RGB sources: the input stage's 16-bit Y, U, V planes (libswscale's formatConvBuffer, for whole frames)
    bool rgb_half = false;        // chroma made at half width (rgb24ToUV_half_c) or at full width
    int rgb_cw = 0;               // its width
    uint8_t *rgb_planes = nullptr;
    size_t rgb_capacity = 0;
    // host-buffer path
    static constexpr int kSlots = 3;                            // staging buffers in flight (copy in / scale / copy out)
    cudaStream_t copy_streams[3] = {nullptr, nullptr, nullptr};   // host-to-device, kernels, device-to-host
    cudaEvent_t stage_events[kSlots][3] = {};                     // per slot: source landed, kernels done, result copied out
    uint8_t *stage_src[kSlots] = {nullptr, nullptr, nullptr};
    uint8_t *stage_dst[kSlots] = {nullptr, nullptr, nullptr};
    size_t stage_src_bytes = 0, stage_dst_bytes = 0;
    long long launches = 0;
    // Cascade (libswscale's RETCODE_USE_CASCADE): the conversion runs as two contexts, src -> tmp -> dst, tmp being
    // yuv420p frames of int(sqrt(srcW * dstW)) x int(sqrt(srcH * dstH)) in a device buffer of `cascade_chunk` frames.
    tvs_ctx *cascade[2] = {nullptr, nullptr};
    uint8_t *cascade_tmp = nullptr;
    int cascade_w = 0, cascade_h = 0, cascade_chunk = 0;
};

namespace {

// The two builds of the fused kernel, selected per context.
bool fused_prefers_two_pass(const tvs_ctx *c) {
    return c->fused_narrow ? fused_plan_prefers_two_pass_w128(c->fused) : fused_plan_prefers_two_pass_w256(c->fused);
}
// forced_path: -1 automatic, 0 fused tile kernel, 1 two-pass kernels, 2 streaming kernel, 3 slab kernel
// Vertical down-scales: the streaming kernel's raw-byte pass wins on 8-bit sources, the slab kernel on 16-bit planar
// ones (measured: 4K yuv420p10le -> 1080p lanczos 37.8 K vs 29.5 K frames/s, profiles/r02_slab_history.md).
bool slab_beats_stream(const tvs_ctx *c) {
    if (!c->slab) return false;
    for (const PlaneOp &op : c->plan.ops)
        if (op.kind == kOpScale && !(op.src_bits > 8 && op.src_xstep == 2)) return false;
    return true;
}
bool use_stream_path(const tvs_ctx *c) {
    return c->streaming && (c->forced_path == 2 || (c->forced_path < 0 && !slab_beats_stream(c)));
}
bool use_slab_path(const tvs_ctx *c) {
    if (use_stream_path(c)) return false;
    return c->slab && (c->forced_path == 3 || c->forced_path < 0);
}
bool use_fused_path(const tvs_ctx *c) {
    if (use_stream_path(c) || use_slab_path(c)) return false;
    return c->fused && (c->forced_path == 0 || (c->forced_path < 0 && !fused_prefers_two_pass(c)));
}

// Logical plane `pl` (0 Y, 1 U, 2 V) of a format: which stored plane, byte offset inside the sample group,
// bytes between samples.
void logical_plane(const FormatInfo &fi, int pl, int &stored, int &offset, int &xstep) {
    if (fi.semi && pl > 0) {
        stored = 1;
        offset = (pl - 1) * fi.bytes;
        xstep = 2 * fi.bytes;
    } else {
        stored = pl;
        offset = 0;
        xstep = fi.bytes;
    }
}

// every output sample is one input sample times `one` (what initFilter yields for equal sizes and equal siting)
bool identity_table(const AxisTable &t, int one) {
    if (t.taps != 1) return false;
    for (int i = 0; i < t.length; ++i)
        if (t.pos[i] != i || t.coef[i] != one) return false;
    return true;
}

void build_plan(tvs_ctx *c) {
    Plan &p = c->plan;
    p.ops.clear();
    const FormatInfo &s = c->sfi, &d = c->dfi;
    const int sw = c->src_w, sh = c->src_h, dw = c->dst_w, dh = c->dst_h;
    const int csw = (sw + 1) / 2, csh = (sh + 1) / 2, cdw = (dw + 1) / 2, cdh = (dh + 1) / 2;
    const bool same = sw == dw && sh == dh && !c->range_convert;   // a range change always runs the scaler proper, as in libswscale

    auto logical_op = [&](int pl, PlaneOpKind kind) {
        PlaneOp op;
        op.kind = kind;
        logical_plane(s, pl, op.src_plane, op.src_offset, op.src_xstep);
        logical_plane(d, pl, op.dst_plane, op.dst_offset, op.dst_xstep);
        op.src_bits = s.bits; op.dst_bits = d.bits; op.src_shift = s.shift; op.dst_shift = d.shift;
        op.sw = pl ? csw : sw; op.sh = pl ? csh : sh; op.dw = pl ? cdw : dw; op.dh = pl ? cdh : dh;
        op.htable = pl ? 1 : 0; op.vtable = pl ? 3 : 2;
        op.hshift = s.bits == 8 ? 7 : s.bits - 1;
        op.rule = d.bits > 8 ? kOut10 : (s.bits > 8 ? kOut8Dither : kOut8);
        op.dither_xoff = pl == 2 ? 3 : 0;
        if (c->range_convert) { op.rc_coeff = c->rc_coeff[pl ? 1 : 0]; op.rc_offset = c->rc_offset[pl ? 1 : 0]; }
        // Note: libswscale's same-size planar -> P010 converter writes only floor(w/2) chroma pairs and
        // leaves the last pair of an odd-width frame untouched; we write all ceil(w/2) pairs (the value the
        // scaler proper would produce), so every destination byte is defined.
        return op;
    };

    if (s.rgb) {
        // The scaler sees the input stage's planar 16-bit lines (virtual planes 0..2 in c->rgb_planes): luma sw x sh,
        // chroma rgb_cw x sh (never subsampled vertically), 14/15-bit values, horizontal shift 13, no ordered dither; from
        // 16-bit components (rgb48) full 16-bit values, shift 15, and the ordered dither of every source deeper than 8 bits.
        for (int pl = 0; pl < 3; ++pl) {
            PlaneOp op;
            op.kind = kOpScale;
            op.src_plane = pl; op.src_offset = 0; op.src_xstep = 2;
            logical_plane(d, pl, op.dst_plane, op.dst_offset, op.dst_xstep);
            op.src_bits = 16; op.dst_bits = d.bits; op.src_shift = 0; op.dst_shift = d.shift;
            op.sw = pl ? c->rgb_cw : sw; op.sh = sh; op.dw = pl ? cdw : dw; op.dh = pl ? cdh : dh;
            op.htable = pl ? 1 : 0; op.vtable = pl ? 3 : 2;
            op.hshift = s.bits == 16 ? 15 : 13;
            op.rule = d.bits > 8 ? kOut10 : s.bits == 16 ? kOut8Dither : kOut8;
            op.dither_xoff = s.bits == 16 && pl == 2 ? 3 : 0;
            p.ops.push_back(op);
        }
    } else if (same && s.bits == 10 && d.bits == 8 && s.semi && d.semi) {
        // P010 -> NV12: each STORED plane is copied with the 2-bit dither, the UV plane as one plane of
        // 2*cw samples (SURVEY.md A.4 as extended in oracle/swscale_port.c).
        for (int st = 0; st < 2; ++st) {
            PlaneOp op;
            op.kind = kOpDither2;
            op.src_plane = op.dst_plane = st;
            op.src_xstep = 2; op.dst_xstep = 1;
            op.src_bits = 10; op.dst_bits = 8; op.src_shift = 6; op.dst_shift = 0;
            op.sw = op.dw = st ? 2 * cdw : dw;
            op.sh = op.dh = st ? cdh : dh;
            op.raw_tail_from = op.dw & ~7;
            p.ops.push_back(op);
        }
    } else {
        for (int pl = 0; pl < 3; ++pl) {
            PlaneOpKind kind = kOpScale;
            if (same) {
                if (s.bits == d.bits) kind = kOpCopy;
                else if (s.bits == 10 && !s.semi && !d.semi) kind = kOpDither2;        // yuv420p10le -> yuv420p
                else if (s.bits == 10 && !s.semi && d.semi && pl == 0) kind = kOpDither2;  // -> nv12 luma
                else if (s.bits == 10 && !s.semi && d.semi && d.bits == 8 && identity_table(c->host_tables[1], 1 << 14) &&
                         identity_table(c->host_tables[3], 1 << 12))
                    kind = kOpDither128;                                                   // -> nv12 chroma: a 1-tap scaler pass
            }
            p.ops.push_back(logical_op(pl, kind));
        }
    }
    size_t off = 0;
    p.any_scale = false;
    for (PlaneOp &op : p.ops) {
        if (op.kind != kOpScale) continue;
        p.any_scale = true;
        op.mid_offset = off;
        off += static_cast<size_t>(op.sh) * op.dw;
    }
    p.mid_elems_per_frame = off;
}

int upload_tables(tvs_ctx *c) {
    size_t bytes = 0;
    size_t coef_off[4], pos_off[4];
    for (int i = 0; i < 4; ++i) {
        coef_off[i] = bytes;
        bytes += (c->host_tables[i].coef.size() * sizeof(int16_t) + 255) & ~static_cast<size_t>(255);
        pos_off[i] = bytes;
        bytes += (c->host_tables[i].pos.size() * sizeof(int32_t) + 255) & ~static_cast<size_t>(255);
    }
    TVS_CUDA(cudaMalloc(&c->dev_table_blob, bytes), "cudaMalloc(tables)");
    uint8_t *base = static_cast<uint8_t *>(c->dev_table_blob);
    for (int i = 0; i < 4; ++i) {
        const AxisTable &t = c->host_tables[i];
        TVS_CUDA(cudaMemcpy(base + coef_off[i], t.coef.data(), t.coef.size() * sizeof(int16_t), cudaMemcpyHostToDevice),
                 "cudaMemcpy(coef)");
        TVS_CUDA(cudaMemcpy(base + pos_off[i], t.pos.data(), t.pos.size() * sizeof(int32_t), cudaMemcpyHostToDevice),
                 "cudaMemcpy(pos)");
        c->dev_tables[i].coef = reinterpret_cast<const int16_t *>(base + coef_off[i]);
        c->dev_tables[i].pos = reinterpret_cast<const int32_t *>(base + pos_off[i]);
        c->dev_tables[i].taps = t.taps;
        c->dev_tables[i].length = t.length;
    }
    return 0;
}

int check_planes(const FormatInfo &fi, int w, int h, const void *const planes[3], const ptrdiff_t pitch[3],
                 const char *side) {
    const int cw = (w + 1) / 2;
    for (int i = 0; i < fi.planes; ++i) {
        if (!planes[i]) return fail(-EINVAL, "%s plane %d is NULL", side, i);
        const ptrdiff_t row = fi.rgb ? 3 * static_cast<ptrdiff_t>(w) * fi.bytes : static_cast<ptrdiff_t>(i == 0 ? w : (fi.semi ? 2 * cw : cw)) * fi.bytes;
        if (pitch[i] < row) return fail(-EINVAL, "%s pitch[%d]=%td is smaller than a row (%td bytes)", side, i, pitch[i], row);
        if (fi.bytes == 2 && ((reinterpret_cast<uintptr_t>(planes[i]) | static_cast<uintptr_t>(pitch[i])) & 1))
            return fail(-EINVAL, "%s plane %d: 16-bit samples need 2-byte aligned pointer and pitch", side, i);
    }
    return 0;
}

void bind_views(const tvs_ctx *c, const void *const src[3], const ptrdiff_t sp[3], const ptrdiff_t sf[3],
                void *const dst[3], const ptrdiff_t dp[3], const ptrdiff_t df[3],
                std::vector<SampleView> &sv, std::vector<SampleView> &dv) {
    const size_t n = c->plan.ops.size();
    sv.resize(n);
    dv.resize(n);
    for (size_t i = 0; i < n; ++i) {
        const PlaneOp &op = c->plan.ops[i];
        sv[i].base = const_cast<uint8_t *>(static_cast<const uint8_t *>(src[op.src_plane])) + op.src_offset;
        sv[i].pitch = sp[op.src_plane];
        sv[i].frame_stride = sf[op.src_plane];
        sv[i].xstep = op.src_xstep;
        dv[i].base = static_cast<uint8_t *>(dst[op.dst_plane]) + op.dst_offset;
        dv[i].pitch = dp[op.dst_plane];
        dv[i].frame_stride = df[op.dst_plane];
        dv[i].xstep = op.dst_xstep;
    }
}

int ensure_mid(tvs_ctx *c, size_t elems) {
    if (c->mid_capacity_elems >= elems) return 0;
    if (c->mid) {
        TVS_CUDA(cudaFree(c->mid), "cudaFree(scratch)");
        c->mid = nullptr;
        c->mid_capacity_elems = 0;
    }
    TVS_CUDA(cudaMalloc(&c->mid, elems * sizeof(int16_t)), "cudaMalloc(scratch)");
    c->mid_capacity_elems = elems;
    return 0;
}

// Enqueue the whole batch on `stream`.  Pointers are device pointers.
int enqueue_planar(tvs_ctx *c, const void *const src[3], const ptrdiff_t sp[3], const ptrdiff_t sf[3],
                  void *const dst[3], const ptrdiff_t dp[3], const ptrdiff_t df[3], int n_frames, cudaStream_t stream) {
    std::vector<SampleView> sv, dv;
    bind_views(c, src, sp, sf, dst, dp, df, sv, dv);
    const Plan &plan = c->plan;

    for (size_t i = 0; i < plan.ops.size(); ++i) {
        const PlaneOp &op = plan.ops[i];
        if (op.kind == kOpScale) continue;
        int rc;
        const int bytes = op.src_bits > 8 ? 2 : 1;
        const bool plain = op.kind == kOpCopy && op.src_bits == op.dst_bits && op.src_shift == op.dst_shift;
        const bool p010_both = plain && op.src_shift == 6 && op.src_bits == 10;   // p010le -> p010le: the library's copy has a quirk
        if (plain && op.src_xstep == bytes && op.dst_xstep == bytes) {
            rc = p010_both ? launch_p010_copy(sv[i], dv[i], op.dw, op.dh, n_frames, stream)
                           : launch_copy_rows(sv[i], dv[i], op.dw * bytes, op.dh, n_frames, stream);     // plane -> plane
        } else if (plain && op.src_xstep == bytes && op.dst_xstep == 2 * bytes && op.dst_offset == 0 && i + 1 < plan.ops.size() &&
                   plan.ops[i + 1].kind == kOpCopy && plan.ops[i + 1].dst_plane == op.dst_plane && plan.ops[i + 1].src_xstep == bytes) {
            rc = launch_interleave(sv[i], sv[i + 1], dv[i], bytes, op.dw, op.dh, n_frames, stream);       // U, V -> UV
            ++i;                                                                                          // consumed the V op too
        } else if (plain && op.src_xstep == 2 * bytes && op.dst_xstep == 2 * bytes && op.src_offset == 0 && op.dst_offset == 0 &&
                   i + 1 < plan.ops.size() && plan.ops[i + 1].src_plane == op.src_plane && plan.ops[i + 1].dst_plane == op.dst_plane) {
            rc = p010_both ? launch_p010_copy(sv[i], dv[i], 2 * op.dw, op.dh, n_frames, stream)
                           : launch_copy_rows(sv[i], dv[i], 2 * op.dw * bytes, op.dh, n_frames, stream);  // UV -> UV as one byte plane
            ++i;
        } else {
            rc = 0;
            if (op.kind == kOpDither2 && op.src_xstep == 2 && op.dst_xstep == 1) {
                rc = launch_dither2_rows(op, sv[i], dv[i], n_frames, stream);                             // plane -> plane, 10 -> 8 bit
            } else if (op.kind == kOpDither2 && op.src_xstep == 2 && op.dst_xstep == 2 && op.dst_offset == 0 && op.raw_tail_from < 0 &&
                       i + 1 < plan.ops.size() && plan.ops[i + 1].kind == kOpDither2 && plan.ops[i + 1].dst_plane == op.dst_plane &&
                       plan.ops[i + 1].src_xstep == 2 && plan.ops[i + 1].dst_offset == 1 && plan.ops[i + 1].src_shift == op.src_shift) {
                rc = launch_dither2_interleave(op, sv[i], sv[i + 1], dv[i], n_frames, stream);            // U, V -> UV, 10 -> 8 bit
                if (rc > 0) ++i;
            }
            if (op.kind == kOpDither128 && op.src_xstep == 2 && op.dst_xstep == 2 && op.dst_offset == 0 && op.dither_xoff == 0 &&
                i + 1 < plan.ops.size() && plan.ops[i + 1].kind == kOpDither128 && plan.ops[i + 1].dst_plane == op.dst_plane &&
                plan.ops[i + 1].dst_offset == 1 && plan.ops[i + 1].dither_xoff == 3) {
                rc = launch_dither128_interleave(op, sv[i], sv[i + 1], dv[i], n_frames, stream);          // U, V -> UV through the 1-tap scaler rule
                if (rc > 0) ++i;
            }
            if (rc == 0) rc = launch_repack(op, sv[i], dv[i], n_frames, stream);
        }
        if (rc < 0) return cuda_fail(cudaGetLastError(), "repack kernel launch");
        ++c->launches;
    }
    if (!plan.any_scale) return 0;

    if (c->box2 && c->forced_path < 0) {
        // one launch per stored plane: Y, then U and V (planar) or the interleaved UV plane
        for (size_t i = 0; i < plan.ops.size(); ++i) {
            const PlaneOp &op = plan.ops[i];
            const bool pair = op.src_xstep == 2;
            if (pair && op.src_offset != 0) continue;   // the V view of an interleaved plane: done with its U view
            if (launch_box2(sv[i], dv[i], pair, pair ? 2 * op.dw : op.dw, op.dh, n_frames, stream) < 0)
                return cuda_fail(cudaGetLastError(), "box-filter kernel launch");
            ++c->launches;
        }
        return 0;
    }
    if (c->forced_path == 2 && !c->streaming) return fail(-ENOSYS, "streaming path requested but this shape is not supported by it");
    if (use_stream_path(c)) {
        const int n = stream_launch(c->streaming, plan, sv.data(), dv.data(), n_frames, stream);
        if (n < 0) return cuda_fail(cudaGetLastError(), "streaming kernel launch");
        c->launches += n;
        return 0;
    }
    if (c->forced_path == 3 && !c->slab) return fail(-ENOSYS, "slab path requested but this shape is not supported by it");
    if (use_slab_path(c)) {
        const int n = slab_launch(c->slab, plan, sv.data(), dv.data(), n_frames, stream);
        if (n < 0) return cuda_fail(cudaGetLastError(), "slab kernel launch");
        c->launches += n;
        return 0;
    }
    const bool use_fused = use_fused_path(c);
    if (c->forced_path == 0 && !c->fused) return fail(-ENOSYS, "fused path requested but this shape is not supported by it");
    if (use_fused) {
        const int n = c->fused_narrow ? fused_launch_w128(c->fused, plan, sv.data(), dv.data(), n_frames, stream)
                                      : fused_launch_w256(c->fused, plan, sv.data(), dv.data(), n_frames, stream);
        if (n < 0) return cuda_fail(cudaGetLastError(), "fused kernel launch");
        c->launches += n;
        return 0;
    }

    // Two-pass: bound the scratch to ~256 MiB and walk the batch in chunks of frames.
    const size_t per_frame = plan.mid_elems_per_frame;
    size_t chunk = (static_cast<size_t>(256) << 20) / (per_frame * sizeof(int16_t));
    if (chunk < 1) chunk = 1;
    if (chunk > static_cast<size_t>(n_frames)) chunk = n_frames;
    if (int rc = ensure_mid(c, per_frame * chunk)) return rc;
    for (int f0 = 0; f0 < n_frames; f0 += static_cast<int>(chunk)) {
        const int nf = std::min<int>(static_cast<int>(chunk), n_frames - f0);
        for (size_t i = 0; i < plan.ops.size(); ++i) {
            const PlaneOp &op = plan.ops[i];
            if (op.kind != kOpScale) continue;
            SampleView s = sv[i], d = dv[i];
            s.base += static_cast<ptrdiff_t>(f0) * s.frame_stride;
            d.base += static_cast<ptrdiff_t>(f0) * d.frame_stride;
            if (launch_hpass_generic(op, s, c->dev_tables[op.htable], c->mid, per_frame, nf, stream) < 0)
                return cuda_fail(cudaGetLastError(), "horizontal kernel launch");
            if (launch_vpass_generic(op, d, c->dev_tables[op.vtable], c->mid, per_frame, nf, stream) < 0)
                return cuda_fail(cudaGetLastError(), "vertical kernel launch");
            c->launches += 2;
        }
    }
    return 0;
}

// ---- fixed scratch for tvs_scale_batch_static ------------------------------------------------------------------
// The two-pass kernels park libswscale's int16 intermediate in global memory between the horizontal and the
// vertical launch.  tvs_scale_batch gets that buffer from cudaMalloc; the entry point below uses this array
// instead, so it never has to ask the driver for memory.  Every GPU that loads the library carries the array
// (TVS_STATIC_SCRATCH_BYTES) until the process exits.
static __device__ __align__(256) int16_t g_static_mid[TVS_STATIC_SCRATCH_BYTES / sizeof(int16_t)];

constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mid_mutex;
int16_t *g_static_mid_ptr[kMaxStaticDevices];   // device address of g_static_mid, per device, once resolved

// `g_static_mid` is only a host-side handle here; the kernels need the address the symbol has on `device`.
// Resolved once per device, with that device current.
int static_mid_for_device(int device, int16_t **out) {
    if (device < 0 || device >= kMaxStaticDevices) return fail(-EINVAL, "device %d is outside the static-scratch table", device);
    std::lock_guard<std::mutex> lock(g_static_mid_mutex);
    if (!g_static_mid_ptr[device]) {
        TVS_CUDA(cudaSetDevice(device), "cudaSetDevice");
        void *p = nullptr;
        TVS_CUDA(cudaGetSymbolAddress(&p, g_static_mid), "cudaGetSymbolAddress(static scratch)");
        g_static_mid_ptr[device] = static_cast<int16_t *>(p);
    }
    *out = g_static_mid_ptr[device];
    return 0;
}

// RGB sources first go through libswscale's input stage (rgb_input_kernel) into 16-bit planes owned by the context,
// a bounded number of frames at a time, and then through the same launches as planar sources.
int enqueue_batch(tvs_ctx *c, const void *const src[3], const ptrdiff_t sp[3], const ptrdiff_t sf[3],
                  void *const dst[3], const ptrdiff_t dp[3], const ptrdiff_t df[3], int n_frames, cudaStream_t stream) {
    if (c->cascade[0]) {
        // src -> tmp -> dst, a bounded number of frames at a time through the context's intermediate frames
        const int tw = c->cascade_w, th = c->cascade_h, cw = (tw + 1) / 2, ch = (th + 1) / 2;
        const size_t tframe = static_cast<size_t>(tw) * th + 2 * static_cast<size_t>(cw) * ch;
        void *tplanes[3] = {c->cascade_tmp, c->cascade_tmp + static_cast<size_t>(tw) * th,
                            c->cascade_tmp + static_cast<size_t>(tw) * th + static_cast<size_t>(cw) * ch};
        const ptrdiff_t tpitch[3] = {tw, cw, cw};
        const ptrdiff_t tfs[3] = {static_cast<ptrdiff_t>(tframe), static_cast<ptrdiff_t>(tframe), static_cast<ptrdiff_t>(tframe)};
        for (int f0 = 0; f0 < n_frames; f0 += c->cascade_chunk) {
            const int nf = std::min(c->cascade_chunk, n_frames - f0);
            const void *s3[3] = {nullptr, nullptr, nullptr};
            void *d3[3] = {nullptr, nullptr, nullptr};
            for (int i = 0; i < c->sfi.planes; ++i) s3[i] = static_cast<const uint8_t *>(src[i]) + static_cast<ptrdiff_t>(f0) * sf[i];
            for (int i = 0; i < c->dfi.planes; ++i) d3[i] = static_cast<uint8_t *>(dst[i]) + static_cast<ptrdiff_t>(f0) * df[i];
            const long long before = c->cascade[0]->launches + c->cascade[1]->launches;
            if (int rc = enqueue_batch(c->cascade[0], s3, sp, sf, tplanes, tpitch, tfs, nf, stream)) return rc;
            if (int rc = enqueue_batch(c->cascade[1], tplanes, tpitch, tfs, d3, dp, df, nf, stream)) return rc;
            c->launches += c->cascade[0]->launches + c->cascade[1]->launches - before;
        }
        return 0;
    }
    if (!c->sfi.rgb) return enqueue_planar(c, src, sp, sf, dst, dp, df, n_frames, stream);
    const size_t y_bytes = static_cast<size_t>(c->src_w) * c->src_h * 2, c_bytes = static_cast<size_t>(c->rgb_cw) * c->src_h * 2;
    const size_t per_frame = y_bytes + 2 * c_bytes;
    size_t chunk = std::max<size_t>(1, (static_cast<size_t>(128) << 20) / per_frame);
    chunk = std::min<size_t>(chunk, static_cast<size_t>(n_frames));
    if (c->rgb_capacity < chunk * per_frame) {
        if (c->rgb_planes) TVS_CUDA(cudaFree(c->rgb_planes), "cudaFree(rgb planes)");
        c->rgb_planes = nullptr; c->rgb_capacity = 0;
        TVS_CUDA(cudaMalloc(&c->rgb_planes, chunk * per_frame), "cudaMalloc(rgb planes)");
        c->rgb_capacity = chunk * per_frame;
    }
    const void *vsrc[3] = {c->rgb_planes, c->rgb_planes + y_bytes, c->rgb_planes + y_bytes + c_bytes};
    const ptrdiff_t vpitch[3] = {2 * static_cast<ptrdiff_t>(c->src_w), 2 * static_cast<ptrdiff_t>(c->rgb_cw), 2 * static_cast<ptrdiff_t>(c->rgb_cw)};
    const ptrdiff_t vfs[3] = {static_cast<ptrdiff_t>(per_frame), static_cast<ptrdiff_t>(per_frame), static_cast<ptrdiff_t>(per_frame)};
    for (int f0 = 0; f0 < n_frames; f0 += static_cast<int>(chunk)) {
        const int nf = std::min<int>(static_cast<int>(chunk), n_frames - f0);
        if (launch_rgb_input(static_cast<const uint8_t *>(src[0]) + static_cast<ptrdiff_t>(f0) * sf[0], sp[0], sf[0], c->rgb_planes, per_frame,
                             y_bytes, c_bytes, c->src_w, c->src_h, c->rgb_half, c->sfi.bgr, c->sfi.bits == 16 ? (c->sfi.be ? 2 : 1) : 0, nf, stream) < 0)
            return cuda_fail(cudaGetLastError(), "rgb input kernel launch");
        ++c->launches;
        void *vdst[3] = {nullptr, nullptr, nullptr};
        for (int i = 0; i < c->dfi.planes; ++i) vdst[i] = static_cast<uint8_t *>(dst[i]) + static_cast<ptrdiff_t>(f0) * df[i];
        if (int rc = enqueue_planar(c, vsrc, vpitch, vfs, vdst, dp, df, nf, stream)) return rc;
    }
    return 0;
}

}  // namespace

extern "C" {

const char *tvs_last_error(void) { return g_last_error.c_str(); }

const char *tvs_version(void) { return "tvscale 0.1 (sm_100a)"; }

int tvs_filter_table(int src_len, int dst_len, int alg, int one, int src_pos, int dst_pos,
                     int16_t *coef, size_t coef_cap, int32_t *pos, size_t pos_cap) {
    AxisTable t;
    if (!build_axis_table(t, src_len, dst_len, static_cast<Algorithm>(alg), one, src_pos, dst_pos))
        return fail(-EINVAL, "cannot build a filter table for %d -> %d with algorithm 0x%x", src_len, dst_len, alg);
    if (coef) {
        if (coef_cap < t.coef.size() || !pos || pos_cap < t.pos.size()) return fail(-ENOSPC, "table buffers too small");
        memcpy(coef, t.coef.data(), t.coef.size() * sizeof(int16_t));
        memcpy(pos, t.pos.data(), t.pos.size() * sizeof(int32_t));
    }
    return t.taps;
}

void *tvs_host_alloc(size_t bytes) {
    void *p = nullptr;
    const cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) {
        cuda_fail(e, "cudaHostAlloc");
        cudaGetLastError();
        return nullptr;
    }
    return p;
}

void tvs_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

tvs_ctx *tvs_create(int device, int src_w, int src_h, int src_fmt, int dst_w, int dst_h, int dst_fmt, int flags,
                    const int chr_pos[4]) {
    return tvs_create_ex(device, src_w, src_h, src_fmt, dst_w, dst_h, dst_fmt, flags, chr_pos, nullptr);
}

tvs_ctx *tvs_create_ex(int device, int src_w, int src_h, int src_fmt, int dst_w, int dst_h, int dst_fmt, int flags,
                       const int chr_pos[4], const double param[2]) {
    if (param && (!(param[0] == param[0]) || !(param[1] == param[1]) || (param[0] != TVS_PARAM_DEFAULT && std::fabs(param[0]) > 64) ||
                  (param[1] != TVS_PARAM_DEFAULT && std::fabs(param[1]) > 64))) {
        fail(-EINVAL, "scaler parameters out of range");
        return nullptr;
    }
    if (param && (flags & TVS_LANCZOS) && param[0] != TVS_PARAM_DEFAULT && param[0] < 1.0) {
        fail(-EINVAL, "lanczos needs at least one lobe (param0 = %g)", param[0]);
        return nullptr;
    }
    // The deprecated "J" format is yuv420p carrying full-range samples (libswscale's handle_jpeg)
    bool src_full = (flags & TVS_SRC_FULL_RANGE) != 0, dst_full = (flags & TVS_DST_FULL_RANGE) != 0;
    if (src_fmt == TVS_FMT_YUVJ420P) { src_fmt = TVS_FMT_YUV420P; src_full = true; }
    if (dst_fmt == TVS_FMT_YUVJ420P) { dst_fmt = TVS_FMT_YUV420P; dst_full = true; }
    flags &= ~(TVS_SRC_FULL_RANGE | TVS_DST_FULL_RANGE);
    FormatInfo sfi, dfi;
    if (!format_info(src_fmt, sfi) || !format_info(dst_fmt, dfi)) {
        fail(-EINVAL, "unsupported pixel format (src %d, dst %d)", src_fmt, dst_fmt);
        return nullptr;
    }
    if (src_w < 1 || src_h < 1 || dst_w < 1 || dst_h < 1 || src_w > 32768 || src_h > 32768 || dst_w > 32768 || dst_h > 32768) {
        fail(-EINVAL, "bad size %dx%d -> %dx%d", src_w, src_h, dst_w, dst_h);
        return nullptr;
    }
    if (dfi.rgb) {
        fail(-ENOSYS, "packed RGB formats are supported as source formats only");
        return nullptr;
    }
    if (sfi.rgb && src_full != dst_full) {
        // rgb24 is full range by nature and the input stage's table already produces limited-range YUV; what libswscale does
        // with an explicit destination range there has not been restated
        fail(-ENOSYS, "range flags are not implemented for packed RGB sources");
        return nullptr;
    }
    if (sfi.rgb && (src_w & 1)) {
        // libswscale's half-width chroma input (rgb24ToUV_half_c) reads a pixel past the row when the width is odd
        fail(-EINVAL, "packed RGB sources need an even width (got %d)", src_w);
        return nullptr;
    }
    const int alg_bits = flags & (TVS_BILINEAR | TVS_BICUBIC | TVS_AREA | TVS_LANCZOS);
    int alg = alg_bits ? alg_bits : TVS_BICUBIC;  // ffmpeg's default when flags= is absent
    if (alg & (alg - 1)) {
        fail(-EINVAL, "flags 0x%x select more than one scaling algorithm", flags);
        return nullptr;
    }
    if (flags & ~(TVS_BILINEAR | TVS_BICUBIC | TVS_AREA | TVS_LANCZOS | TVS_ACCURATE_RND | TVS_BITEXACT)) {
        fail(-ENOSYS, "flags 0x%x contain options this scaler does not implement", flags);
        return nullptr;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || device < 0 || device >= count) {
        fail(-ENODEV, "CUDA device %d not available (%s); this library has no CPU path", device,
             e != cudaSuccess ? cudaGetErrorString(e) : "index out of range");
        return nullptr;
    }
    if ((e = cudaSetDevice(device)) != cudaSuccess) {
        cuda_fail(e, "cudaSetDevice");
        return nullptr;
    }

    {
        // Resolve the fixed scratch array of tvs_scale_batch_static on this device now.  With lazy module loading this
        // is what brings the array's module (and its TVS_STATIC_SCRATCH_BYTES) onto the device, so it happens here,
        // where allocation is expected, and never inside a scale call.  A failure is not fatal for the other entry points.
        int16_t *unused = nullptr;
        if (static_mid_for_device(device, &unused) != 0) cudaGetLastError();
    }

    tvs_ctx *c = new tvs_ctx;
    c->device = device;
    c->src_w = src_w; c->src_h = src_h; c->dst_w = dst_w; c->dst_h = dst_h;
    c->src_fmt = src_fmt; c->dst_fmt = dst_fmt; c->flags = flags;
    c->sfi = sfi; c->dfi = dfi;
    if (src_full != dst_full) {
        // ff_sws_init_range_convert: the constants follow the DESTINATION depth; t' = min((t * coeff + offset) >> 14, 32767)
        const int depth = dfi.bits, src_shift = 15 - depth, total = 14 + src_shift;
        const long long mpeg_min = 16LL << (depth - 8), mpeg_lum = 235LL << (depth - 8), mpeg_chr = 240LL << (depth - 8), jpeg_max = (1LL << depth) - 1;
        auto solve = [&](long long smin, long long smax, long long dmin, long long dmax, int &coeff, int &offset) {
            const unsigned long long q = (static_cast<unsigned long long>(dmax - dmin) << total) / static_cast<unsigned long long>(smax - smin);
            coeff = static_cast<int>((q + ((1ULL << src_shift) - 1)) >> src_shift);
            offset = static_cast<int>((dmax << total) - (smax << src_shift) * coeff + (1 << 13));
        };
        c->range_convert = src_full ? 1 : 2;
        if (src_full) {
            solve(0, jpeg_max, mpeg_min, mpeg_lum, c->rc_coeff[0], c->rc_offset[0]);
            solve(0, jpeg_max, mpeg_min, mpeg_chr, c->rc_coeff[1], c->rc_offset[1]);
        } else {
            solve(mpeg_min, mpeg_lum, 0, jpeg_max, c->rc_coeff[0], c->rc_offset[0]);
            solve(mpeg_min, mpeg_chr, 0, jpeg_max, c->rc_coeff[1], c->rc_offset[1]);
        }
    }
    static const int centred[4] = {TVS_CHR_POS_DEFAULT, TVS_CHR_POS_DEFAULT, TVS_CHR_POS_DEFAULT, TVS_CHR_POS_DEFAULT};
    if (!chr_pos) chr_pos = centred;
    int csw = (src_w + 1) / 2, csh = (src_h + 1) / 2;
    const int cdw = (dst_w + 1) / 2, cdh = (dst_h + 1) / 2;
    int src_hsub = 1, src_vsub = 1;   // log2 chroma subsampling of the source as the scaler sees it
    if (sfi.rgb) {
        // libswscale drops every other pixel for chroma "unless the destination chroma is wider than half the source"
        // (sws_init_context: (dstW >> chrDstHSubSample) <= (srcW >> 1)); RGB chroma is never subsampled vertically.
        c->rgb_half = (dst_w >> 1) <= (src_w >> 1);
        c->rgb_cw = csw = c->rgb_half ? src_w / 2 : src_w;
        csh = src_h;
        src_hsub = c->rgb_half ? 1 : 0;
        src_vsub = 0;
    }
    const Algorithm a = static_cast<Algorithm>(alg);
    const int luma = axis_position(TVS_CHR_POS_DEFAULT, 0);
    const bool built[4] = {
        build_axis_table(c->host_tables[0], src_w, dst_w, a, 1 << 14, luma, luma, param),
        build_axis_table(c->host_tables[1], csw, cdw, a, 1 << 14, axis_position(chr_pos[0], src_hsub), axis_position(chr_pos[2], 1), param),
        build_axis_table(c->host_tables[2], src_h, dst_h, a, 1 << 12, luma, luma, param),
        build_axis_table(c->host_tables[3], csh, cdh, a, 1 << 12, axis_position(chr_pos[1], src_vsub), axis_position(chr_pos[3], 1), param)};
    for (int i = 0; i < 4; ++i) {
        if (!built[i] && c->host_tables[i].taps <= 256) {   // (a table wider than 256 taps is the cascade's business, below)
            fail(-EINVAL, "cannot build the filter tables for %dx%d -> %dx%d", src_w, src_h, dst_w, dst_h);
            delete c;
            return nullptr;
        }
    }
    // With accurate_rnd libswscale does not run a filter that, padded to the x86 build's table alignment (4 taps
    // horizontally, 2 vertically), reaches 256 * 16 / 24 = 170 taps (initFilter's RETCODE_USE_CASCADE): the context
    // becomes two, src -> tmp -> dst, with tmp a yuv420p frame of int(sqrt(srcW * dstW)) x int(sqrt(srcH * dstH)) and both
    // inner contexts at the library's default chroma positions (oracle/swscale_port.c, pinned against the genuine library).
    if (((c->host_tables[0].taps + 3) & ~3) >= 170 || ((c->host_tables[1].taps + 3) & ~3) >= 170 ||
        ((c->host_tables[2].taps + 1) & ~1) >= 170 || ((c->host_tables[3].taps + 1) & ~1) >= 170) {
        if (static_cast<long long>(src_w) * src_h <= 4LL * dst_w * dst_h) {
            fail(-EINVAL, "%dx%d -> %dx%d: libswscale refuses this conversion (it would cascade, and the source is not four times the destination)",
                 src_w, src_h, dst_w, dst_h);
            delete c;
            return nullptr;
        }
        // (the two inner contexts are created with the default ranges, so a cascade converts no range: libswscale's behaviour,
        // pinned in tests/test_oracle.py)
        c->range_convert = 0;
        c->cascade_w = static_cast<int>(std::sqrt(static_cast<double>(src_w) * dst_w));
        c->cascade_h = static_cast<int>(std::sqrt(static_cast<double>(src_h) * dst_h));
        c->cascade[0] = tvs_create_ex(device, src_w, src_h, src_fmt, c->cascade_w, c->cascade_h, TVS_FMT_YUV420P, flags, nullptr, param);
        c->cascade[1] = c->cascade[0] ? tvs_create_ex(device, c->cascade_w, c->cascade_h, TVS_FMT_YUV420P, dst_w, dst_h, dst_fmt, flags, nullptr, param) : nullptr;
        const size_t tframe = static_cast<size_t>(c->cascade_w) * c->cascade_h +
                              2 * static_cast<size_t>((c->cascade_w + 1) / 2) * ((c->cascade_h + 1) / 2);
        c->cascade_chunk = static_cast<int>(std::min<size_t>(65535, std::max<size_t>(1, (static_cast<size_t>(64) << 20) / tframe)));
        if (!c->cascade[0] || !c->cascade[1] ||
            cudaMalloc(&c->cascade_tmp, tframe * static_cast<size_t>(c->cascade_chunk)) != cudaSuccess) {
            const std::string why = g_last_error;
            cudaGetLastError();
            tvs_destroy(c);
            fail(-EIO, "cascaded context: %s", why.empty() ? "device allocation failed" : why.c_str());
            return nullptr;
        }
        return c;
    }
    build_plan(c);
    if (upload_tables(c) != 0) {
        tvs_destroy(c);
        return nullptr;
    }
    // Exact 2:1 area: every table row is the pair [one/2, one/2] on samples 2i, 2i+1 (see box2_kernel).
    {
        auto is_box = [](const AxisTable &t, int one) {
            if (t.taps != 2) return false;
            for (int i = 0; i < t.length; ++i)
                if (t.pos[i] != 2 * i || t.coef[2 * i] != one / 2 || t.coef[2 * i + 1] != one / 2) return false;
            return true;
        };
        bool ok = c->plan.any_scale && !c->range_convert && sfi.bits == 8 && dfi.bits == 8 && src_w == 2 * dst_w && src_h == 2 * dst_h &&
                  (src_w % 4) == 0 && (src_h % 4) == 0 && sfi.semi == dfi.semi;
        for (int i = 0; ok && i < 4; ++i) ok = is_box(c->host_tables[i], i < 2 ? 1 << 14 : 1 << 12);
        c->box2 = ok;
    }
    // Range changes run on the integer two-pass kernels (the conversion is an integer multiply-add on the intermediate, wider
    // than the float kernels keep exact); RGB sources on the slab kernel's split-sample pass, else on the two-pass kernels
    if (c->plan.any_scale && !sfi.rgb && !c->range_convert) {
        std::string why;
        // Down-scales stage a wide source footprint and do better with 128-column tiles; up-scales amortise the
        // vertical pass over 256 columns.  If the preferred build cannot take the shape, try the other one;
        // nullptr = the shape is handled by the two-pass kernels.
        if (src_h > dst_h) c->streaming = stream_plan_create(c->plan, c->host_tables, device, why);
        c->slab = slab_plan_create(c->plan, c->host_tables, device, why);   // nullptr unless no plane is shrunk vertically
        c->fused_narrow = src_w > dst_w;
        for (int attempt = 0; attempt < 2 && !c->fused; ++attempt, c->fused_narrow = !c->fused_narrow) {
            c->fused = c->fused_narrow ? fused_plan_create_w128(c->plan, c->host_tables, device, why)
                                       : fused_plan_create_w256(c->plan, c->host_tables, device, why);
            if (c->fused) break;
        }
    }
    if (c->plan.any_scale && sfi.rgb) {
        // the input stage's 15-bit planes go through the slab kernel's split-sample horizontal pass when it takes the shape
        std::string why;
        c->slab = slab_plan_create(c->plan, c->host_tables, device, why);
    }
    return c;
}

void tvs_destroy(tvs_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    for (int i = 0; i < 2; ++i) if (c->cascade[i]) tvs_destroy(c->cascade[i]);
    if (c->cascade_tmp) cudaFree(c->cascade_tmp);
    cudaSetDevice(c->device);
    if (c->fused) { if (c->fused_narrow) fused_plan_destroy_w128(c->fused); else fused_plan_destroy_w256(c->fused); }
    if (c->streaming) stream_plan_destroy(c->streaming);
    if (c->slab) slab_plan_destroy(c->slab);
    if (c->dev_table_blob) cudaFree(c->dev_table_blob);
    if (c->mid) cudaFree(c->mid);
    if (c->rgb_planes) cudaFree(c->rgb_planes);
    for (int i = 0; i < tvs_ctx::kSlots; ++i) {
        if (c->stage_src[i]) cudaFree(c->stage_src[i]);
        if (c->stage_dst[i]) cudaFree(c->stage_dst[i]);
        for (int k = 0; k < 3; ++k)
            if (c->stage_events[i][k]) cudaEventDestroy(c->stage_events[i][k]);
    }
    for (int i = 0; i < 3; ++i)
        if (c->copy_streams[i]) cudaStreamDestroy(c->copy_streams[i]);
    delete c;
}

int tvs_scale_batch(tvs_ctx *c, const void *const src[3], const ptrdiff_t src_pitch[3], const ptrdiff_t src_frame_stride[3],
                    void *const dst[3], const ptrdiff_t dst_pitch[3], const ptrdiff_t dst_frame_stride[3],
                    int n_frames, void *stream) {
    if (!c) return fail(-EINVAL, "NULL context");
    if (n_frames < 0) return fail(-EINVAL, "n_frames < 0");
    if (n_frames > 65535) return fail(-EINVAL, "at most 65535 frames per call");
    if (int rc = check_planes(c->sfi, c->src_w, c->src_h, src, src_pitch, "src")) return rc;
    if (int rc = check_planes(c->dfi, c->dst_w, c->dst_h, reinterpret_cast<const void *const *>(dst), dst_pitch, "dst")) return rc;
    if (n_frames == 0) return c->dst_h;
    TVS_CUDA(cudaSetDevice(c->device), "cudaSetDevice");
    if (int rc = enqueue_batch(c, src, src_pitch, src_frame_stride, dst, dst_pitch, dst_frame_stride, n_frames,
                               static_cast<cudaStream_t>(stream)))
        return rc;
    return c->dst_h;
}

int tvs_scale_host(tvs_ctx *c, const void *const src[3], const ptrdiff_t src_pitch[3], const ptrdiff_t src_frame_stride[3],
                   void *const dst[3], const ptrdiff_t dst_pitch[3], const ptrdiff_t dst_frame_stride[3], int n_frames) {
    if (!c) return fail(-EINVAL, "NULL context");
    if (n_frames < 0) return fail(-EINVAL, "n_frames < 0");
    if (int rc = check_planes(c->sfi, c->src_w, c->src_h, src, src_pitch, "src")) return rc;
    if (int rc = check_planes(c->dfi, c->dst_w, c->dst_h, reinterpret_cast<const void *const *>(dst), dst_pitch, "dst")) return rc;
    if (n_frames == 0) return c->dst_h;
    TVS_CUDA(cudaSetDevice(c->device), "cudaSetDevice");

    // Device staging keeps planes tightly packed, frame after frame: [frame][plane].
    const int cw_s = (c->src_w + 1) / 2, ch_s = (c->src_h + 1) / 2, cw_d = (c->dst_w + 1) / 2, ch_d = (c->dst_h + 1) / 2;
    auto plane_rows = [](int i, int h, int ch) { return i == 0 ? h : ch; };
    auto plane_row_bytes = [](const FormatInfo &fi, int i, int w, int cw) {
        if (fi.rgb) return static_cast<size_t>(3) * w * fi.bytes;
        return static_cast<size_t>(i == 0 ? w : (fi.semi ? 2 * cw : cw)) * fi.bytes;
    };
    const size_t sframe = tight_frame_bytes(c->sfi, c->src_w, c->src_h);
    const size_t dframe = tight_frame_bytes(c->dfi, c->dst_w, c->dst_h);
    // Three stages - host-to-device copy, kernels, device-to-host copy - each on its own stream, chunks of about
    // 48 MiB walking through kSlots staging buffers: while chunk i is scaled, chunk i+1 is coming in and chunk i-1
    // is going out, so both copy engines and the SMs are busy at once.
    size_t chunk = (static_cast<size_t>(48) << 20) / (sframe + dframe);
    if (chunk < 1) chunk = 1;
    if (chunk > static_cast<size_t>(n_frames)) chunk = n_frames;
    if (chunk > 65535) chunk = 65535;   // the kernels index frames through a 16-bit grid dimension
    if (c->stage_src_bytes < chunk * sframe || c->stage_dst_bytes < chunk * dframe) {
        for (int i = 0; i < tvs_ctx::kSlots; ++i) {
            if (c->stage_src[i]) cudaFree(c->stage_src[i]);
            if (c->stage_dst[i]) cudaFree(c->stage_dst[i]);
            c->stage_src[i] = c->stage_dst[i] = nullptr;
        }
        c->stage_src_bytes = c->stage_dst_bytes = 0;
        for (int i = 0; i < tvs_ctx::kSlots; ++i) {
            TVS_CUDA(cudaMalloc(&c->stage_src[i], chunk * sframe), "cudaMalloc(src staging)");
            TVS_CUDA(cudaMalloc(&c->stage_dst[i], chunk * dframe), "cudaMalloc(dst staging)");
        }
        c->stage_src_bytes = chunk * sframe;
        c->stage_dst_bytes = chunk * dframe;
    }
    for (int i = 0; i < 3; ++i)
        if (!c->copy_streams[i]) TVS_CUDA(cudaStreamCreateWithFlags(&c->copy_streams[i], cudaStreamNonBlocking), "cudaStreamCreate");
    for (int i = 0; i < tvs_ctx::kSlots; ++i)
        for (int k = 0; k < 3; ++k)
            if (!c->stage_events[i][k]) TVS_CUDA(cudaEventCreateWithFlags(&c->stage_events[i][k], cudaEventDisableTiming), "cudaEventCreate");
    cudaStream_t st_in = c->copy_streams[0], st_exec = c->copy_streams[1], st_out = c->copy_streams[2];

    // Device-side plane table of a staged chunk.
    ptrdiff_t s_pitch[3] = {0, 0, 0}, s_fs[3], d_pitch[3] = {0, 0, 0}, d_fs[3];
    size_t s_off[3] = {0, 0, 0}, d_off[3] = {0, 0, 0};
    {
        size_t o = 0;
        for (int i = 0; i < c->sfi.planes; ++i) {
            s_off[i] = o;
            s_pitch[i] = plane_row_bytes(c->sfi, i, c->src_w, cw_s);
            o += s_pitch[i] * plane_rows(i, c->src_h, ch_s);
        }
        o = 0;
        for (int i = 0; i < c->dfi.planes; ++i) {
            d_off[i] = o;
            d_pitch[i] = plane_row_bytes(c->dfi, i, c->dst_w, cw_d);
            o += d_pitch[i] * plane_rows(i, c->dst_h, ch_d);
        }
        for (int i = 0; i < 3; ++i) { s_fs[i] = sframe; d_fs[i] = dframe; }
    }
    // A host frame that is itself tightly packed [Y|U|V] with frames back to back moves as ONE copy per chunk.
    auto packed = [&](const void *const pl[3], const ptrdiff_t pitch[3], const ptrdiff_t fs[3], const FormatInfo &fi,
                      const size_t off[3], const ptrdiff_t tp[3], size_t frame) {
        for (int i = 0; i < fi.planes; ++i) {
            if (pitch[i] != tp[i] || fs[i] != static_cast<ptrdiff_t>(frame)) return false;
            if (static_cast<const uint8_t *>(pl[i]) != static_cast<const uint8_t *>(pl[0]) + off[i]) return false;
        }
        return true;
    };
    const bool src_packed = packed(src, src_pitch, src_frame_stride, c->sfi, s_off, s_pitch, sframe);
    const bool dst_packed = packed(reinterpret_cast<const void *const *>(dst), dst_pitch, dst_frame_stride, c->dfi, d_off, d_pitch, dframe);

    // One chunk: its copies and kernels, ordered against its neighbours through the slot's three events
    // (0: source landed, 1: kernels done, 2: result copied out).
    auto enqueue_chunk = [&](int index, int f0, int nf) -> int {
        const int slot = index % tvs_ctx::kSlots;
        cudaEvent_t ev_in = c->stage_events[slot][0], ev_exec = c->stage_events[slot][1], ev_out = c->stage_events[slot][2];
        const bool reused = index >= tvs_ctx::kSlots;
        if (reused) TVS_CUDA(cudaStreamWaitEvent(st_in, ev_exec, 0), "stream wait");   // the slot's source buffer has been consumed
        if (src_packed) {
            TVS_CUDA(cudaMemcpyAsync(c->stage_src[slot], static_cast<const uint8_t *>(src[0]) + static_cast<size_t>(f0) * sframe,
                                     static_cast<size_t>(nf) * sframe, cudaMemcpyHostToDevice, st_in), "H2D copy");
        } else {
            for (int f = 0; f < nf; ++f)
                for (int i = 0; i < c->sfi.planes; ++i)
                    TVS_CUDA(cudaMemcpy2DAsync(c->stage_src[slot] + f * sframe + s_off[i], s_pitch[i],
                                               static_cast<const uint8_t *>(src[i]) + static_cast<ptrdiff_t>(f0 + f) * src_frame_stride[i],
                                               src_pitch[i], s_pitch[i], plane_rows(i, c->src_h, ch_s), cudaMemcpyHostToDevice, st_in),
                             "H2D plane copy");
        }
        TVS_CUDA(cudaEventRecord(ev_in, st_in), "event record");
        TVS_CUDA(cudaStreamWaitEvent(st_exec, ev_in, 0), "stream wait");
        if (reused) TVS_CUDA(cudaStreamWaitEvent(st_exec, ev_out, 0), "stream wait");  // the slot's result buffer has been drained
        const void *sp[3] = {c->stage_src[slot] + s_off[0], c->stage_src[slot] + s_off[1], c->stage_src[slot] + s_off[2]};
        void *dp[3] = {c->stage_dst[slot] + d_off[0], c->stage_dst[slot] + d_off[1], c->stage_dst[slot] + d_off[2]};
        if (int rc = enqueue_batch(c, sp, s_pitch, s_fs, dp, d_pitch, d_fs, nf, st_exec)) return rc;
        TVS_CUDA(cudaEventRecord(ev_exec, st_exec), "event record");
        TVS_CUDA(cudaStreamWaitEvent(st_out, ev_exec, 0), "stream wait");
        if (dst_packed) {
            TVS_CUDA(cudaMemcpyAsync(static_cast<uint8_t *>(dst[0]) + static_cast<size_t>(f0) * dframe, c->stage_dst[slot],
                                     static_cast<size_t>(nf) * dframe, cudaMemcpyDeviceToHost, st_out), "D2H copy");
        } else {
            for (int f = 0; f < nf; ++f)
                for (int i = 0; i < c->dfi.planes; ++i)
                    TVS_CUDA(cudaMemcpy2DAsync(static_cast<uint8_t *>(dst[i]) + static_cast<ptrdiff_t>(f0 + f) * dst_frame_stride[i], dst_pitch[i],
                                               c->stage_dst[slot] + f * dframe + d_off[i], d_pitch[i], d_pitch[i],
                                               plane_rows(i, c->dst_h, ch_d), cudaMemcpyDeviceToHost, st_out),
                             "D2H plane copy");
        }
        TVS_CUDA(cudaEventRecord(ev_out, st_out), "event record");
        return 0;
    };
    int rc = 0, index = 0;
    for (int f0 = 0; f0 < n_frames && rc == 0; f0 += static_cast<int>(chunk), ++index)
        rc = enqueue_chunk(index, f0, std::min<int>(static_cast<int>(chunk), n_frames - f0));
    // Whatever happened, nothing may still be writing the caller's buffers when we return.
    const std::string first_error = rc ? g_last_error : std::string();
    for (int i = 0; i < 3; ++i) {
        const cudaError_t e = cudaStreamSynchronize(c->copy_streams[i]);
        if (e != cudaSuccess && rc == 0) rc = cuda_fail(e, "stream sync");
    }
    if (!first_error.empty()) g_last_error = first_error;
    if (rc) return rc;
    return c->dst_h;
}

int tvs_context_table(const tvs_ctx *c, int which, int16_t *coef, size_t coef_cap, int32_t *pos, size_t pos_cap) {
    if (!c || which < 0 || which > 3) return fail(-EINVAL, "bad table index");
    const AxisTable &t = c->host_tables[which];
    if (coef) {
        if (coef_cap < t.coef.size() || !pos || pos_cap < t.pos.size()) return fail(-ENOSPC, "table buffers too small");
        memcpy(coef, t.coef.data(), t.coef.size() * sizeof(int16_t));
        memcpy(pos, t.pos.data(), t.pos.size() * sizeof(int32_t));
    }
    return t.taps;
}

long long tvs_info(const tvs_ctx *c, int what) {
    if (!c) return fail(-EINVAL, "NULL context");
    switch (what) {
    case TVS_INFO_KERNEL_LAUNCHES: return c->launches;
    case TVS_INFO_PATH:
        if (c->cascade[0]) return 6;
        if (!c->plan.any_scale) return 2;
        if (c->box2 && c->forced_path < 0) return 4;
        return use_stream_path(c) ? 3 : use_slab_path(c) ? 5 : use_fused_path(c) ? 0 : 1;
    case TVS_INFO_SCRATCH_BYTES:
        return static_cast<long long>(c->mid_capacity_elems * sizeof(int16_t) + tvs_ctx::kSlots * (c->stage_src_bytes + c->stage_dst_bytes)) +
               (c->cascade[0] ? tvs_info(c->cascade[0], what) + tvs_info(c->cascade[1], what) : 0);
    case TVS_INFO_SRC_FRAME_BYTES: return static_cast<long long>(tight_frame_bytes(c->sfi, c->src_w, c->src_h));
    case TVS_INFO_DST_FRAME_BYTES: return static_cast<long long>(tight_frame_bytes(c->dfi, c->dst_w, c->dst_h));
    default: return fail(-EINVAL, "unknown info key %d", what);
    }
}

int tvs_set_path(tvs_ctx *c, int path) {
    if (!c || path < -1 || path > 3) return fail(-EINVAL, "bad path");
    if (c->cascade[0]) {
        // both stages take the request, or neither does
        if (tvs_set_path(c->cascade[0], path) == 0 && tvs_set_path(c->cascade[1], path) == 0) return 0;
        const std::string why = g_last_error;
        tvs_set_path(c->cascade[0], -1);
        tvs_set_path(c->cascade[1], -1);
        return fail(-ENOSYS, "cascaded context: %s", why.c_str());
    }
    if (path == 3 && c->plan.any_scale && !c->slab) return fail(-ENOSYS, "this shape is not supported by the slab kernel");
    if (path == 0 && c->plan.any_scale && !c->fused) return fail(-ENOSYS, "this shape is not supported by the fused kernel");
    if (path == 2 && c->plan.any_scale && !c->streaming) return fail(-ENOSYS, "this shape is not supported by the streaming kernel");
    c->forced_path = path;
    return 0;
}

int tvs_bwdif_batch(int device, int fmt, int width, int height, const void *const src[3], const ptrdiff_t src_pitch[3],
                    const ptrdiff_t src_frame_stride[3], void *const dst[3], const ptrdiff_t dst_pitch[3],
                    const ptrdiff_t dst_frame_stride[3], int n_frames, int tff, int flags, void *stream) {
    FormatInfo fi;
    if (!format_info(fmt, fi) || fi.semi || fi.rgb) return fail(-EINVAL, "bwdif takes planar 4:2:0 frames (yuv420p, yuv420p10le)");
    if (width < 1 || height < 1 || n_frames < 1) return fail(-EINVAL, "bad size or frame count");
    if (int rc = check_planes(fi, width, height, src, src_pitch, "src")) return rc;
    if (int rc = check_planes(fi, width, height, reinterpret_cast<const void *const *>(dst), dst_pitch, "dst")) return rc;
    TVS_CUDA(cudaSetDevice(device), "cudaSetDevice");
    const int cw = (width + 1) / 2, ch = (height + 1) / 2;
    int written = 0;
    for (int pl = 0; pl < 3; ++pl) {
        const int rc = launch_bwdif_plane(static_cast<const uint8_t *>(src[pl]), src_pitch[pl], src_frame_stride[pl],
                                          static_cast<uint8_t *>(dst[pl]), dst_pitch[pl], dst_frame_stride[pl], pl ? cw : width,
                                          pl ? ch : height, fi.bits, n_frames, tff, (flags & TVS_BWDIF_STREAM_START) != 0,
                                          (flags & TVS_BWDIF_STREAM_END) != 0, static_cast<cudaStream_t>(stream));
        if (rc < 0) return cuda_fail(cudaGetLastError(), "bwdif kernel launch");
    }
    written = n_frames - ((flags & TVS_BWDIF_STREAM_START) ? 0 : 1) - ((flags & TVS_BWDIF_STREAM_END) ? 0 : 1);
    return written < 0 ? 0 : written;
}

int tvs_bwdif_host(int device, int fmt, int width, int height, const void *src, void *dst, int n_frames, int tff, int flags) {
    FormatInfo fi;
    if (!format_info(fmt, fi) || fi.semi || fi.rgb) return fail(-EINVAL, "bwdif takes planar 4:2:0 frames (yuv420p, yuv420p10le)");
    if (width < 1 || height < 1 || n_frames < 1 || !src || !dst) return fail(-EINVAL, "bad arguments");
    TVS_CUDA(cudaSetDevice(device), "cudaSetDevice");
    const size_t frame = tight_frame_bytes(fi, width, height);
    const int cw = (width + 1) / 2, ch = (height + 1) / 2;
    uint8_t *d_src = nullptr, *d_dst = nullptr;
    if (cudaMalloc(&d_src, frame * n_frames) != cudaSuccess || cudaMalloc(&d_dst, frame * n_frames) != cudaSuccess) {
        const cudaError_t e = cudaGetLastError();
        if (d_src) cudaFree(d_src);
        return cuda_fail(e, "cudaMalloc(bwdif staging)");
    }
    int rc = 0;
    cudaError_t e = cudaMemcpy(d_src, src, frame * n_frames, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        const size_t y = static_cast<size_t>(width) * height * fi.bytes, c = static_cast<size_t>(cw) * ch * fi.bytes;
        const void *sp[3] = {d_src, d_src + y, d_src + y + c};
        void *dp[3] = {d_dst, d_dst + y, d_dst + y + c};
        const ptrdiff_t pitch[3] = {static_cast<ptrdiff_t>(width) * fi.bytes, static_cast<ptrdiff_t>(cw) * fi.bytes, static_cast<ptrdiff_t>(cw) * fi.bytes};
        const ptrdiff_t fs[3] = {static_cast<ptrdiff_t>(frame), static_cast<ptrdiff_t>(frame), static_cast<ptrdiff_t>(frame)};
        rc = tvs_bwdif_batch(device, fmt, width, height, sp, pitch, fs, dp, pitch, fs, n_frames, tff, flags, nullptr);
        if (rc >= 0) {
            // frames that are context only are not written on the device; hand back only the written range
            const int first = (flags & TVS_BWDIF_STREAM_START) ? 0 : 1;
            if (rc > 0) e = cudaMemcpy(static_cast<uint8_t *>(dst) + first * frame, d_dst + first * frame, frame * rc, cudaMemcpyDeviceToHost);
        }
    }
    cudaFree(d_src);
    cudaFree(d_dst);
    if (e != cudaSuccess) return cuda_fail(e, "bwdif copies");
    return rc;
}

int tvs_scale_batch_static(tvs_ctx *c, const void *const src[3], const ptrdiff_t src_pitch[3], const ptrdiff_t src_frame_stride[3],
                           void *const dst[3], const ptrdiff_t dst_pitch[3], const ptrdiff_t dst_frame_stride[3],
                           int n_frames, void *stream) {
    if (!c) return fail(-EINVAL, "NULL context");
    if (n_frames < 0 || n_frames > 65535) return fail(-EINVAL, "n_frames must be 0..65535");
    if (c->sfi.rgb) return fail(-ENOSYS, "rgb24 sources need the context's input-stage buffer; use tvs_scale_batch");
    if (c->cascade[0]) return fail(-ENOSYS, "a cascaded context needs its intermediate frames; use tvs_scale_batch");
    if (int rc = check_planes(c->sfi, c->src_w, c->src_h, src, src_pitch, "src")) return rc;
    if (int rc = check_planes(c->dfi, c->dst_w, c->dst_h, reinterpret_cast<const void *const *>(dst), dst_pitch, "dst")) return rc;
    const Plan &plan = c->plan;
    const size_t need = plan.mid_elems_per_frame * static_cast<size_t>(n_frames) * sizeof(int16_t);
    if (need > TVS_STATIC_SCRATCH_BYTES)
        return fail(-E2BIG, "%d frame(s) of this shape need %zu bytes of intermediate, the static scratch holds %u "
                            "(at most %zu frame(s) per call)", n_frames, need, static_cast<unsigned>(TVS_STATIC_SCRATCH_BYTES),
                    plan.mid_elems_per_frame ? TVS_STATIC_SCRATCH_BYTES / (plan.mid_elems_per_frame * sizeof(int16_t)) : size_t(0));
    if (n_frames == 0) return c->dst_h;
    if (plan.ops.size() > 3) return fail(-EINVAL, "unexpected plan");
    int16_t *mid = nullptr;
    if (int rc = static_mid_for_device(c->device, &mid)) return rc;   // the first call on a device resolves the address
    TVS_CUDA(cudaSetDevice(c->device), "cudaSetDevice");
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    // Same launches as the two-pass branch of tvs_scale_batch, with the array above as the intermediate.  Nothing
    // in it has to start at zero: the horizontal kernel writes every element the vertical kernel reads.
    for (size_t i = 0; i < plan.ops.size(); ++i) {
        const PlaneOp &op = plan.ops[i];
        SampleView s, d;
        s.base = const_cast<uint8_t *>(static_cast<const uint8_t *>(src[op.src_plane])) + op.src_offset;
        s.pitch = src_pitch[op.src_plane]; s.frame_stride = src_frame_stride[op.src_plane]; s.xstep = op.src_xstep;
        d.base = static_cast<uint8_t *>(dst[op.dst_plane]) + op.dst_offset;
        d.pitch = dst_pitch[op.dst_plane]; d.frame_stride = dst_frame_stride[op.dst_plane]; d.xstep = op.dst_xstep;
        if (op.kind != kOpScale) {
            if (launch_repack(op, s, d, n_frames, st) < 0) return cuda_fail(cudaGetLastError(), "repack kernel launch");
            ++c->launches;
            continue;
        }
        if (launch_hpass_generic(op, s, c->dev_tables[op.htable], mid, plan.mid_elems_per_frame, n_frames, st) < 0)
            return cuda_fail(cudaGetLastError(), "horizontal kernel launch");
        if (launch_vpass_generic(op, d, c->dev_tables[op.vtable], mid, plan.mid_elems_per_frame, n_frames, st) < 0)
            return cuda_fail(cudaGetLastError(), "vertical kernel launch");
        c->launches += 2;
    }
    return c->dst_h;
}

}  // extern "C"
