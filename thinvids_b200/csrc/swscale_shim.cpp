// libswscale's public trio - sws_getContext / sws_scale / sws_freeContext, what ffmpeg's vf_scale calls for the reference's
// `scale=...` (worker/tasks.py:62-65,444-449,1565-1580; tools/upscale_benchmark.py:163,182) - on top of libtvscale.
// Built as its own small library (libswscale_tvs.so) so that a program written against libswscale's API can be pointed at
// the GPU path without touching its call sites: same arguments, same return values (output rows, or a negative AVERROR),
// HOST plane pointers and int strides.  Output is what libswscale gives for flags | SWS_ACCURATE_RND | SWS_BITEXACT.
//
// Not covered (NULL / AVERROR(EINVAL) / AVERROR(ENOSYS), never a silent approximation): source or destination filters,
// slices (srcSliceY != 0 or srcSliceH != srcH), negative strides, pixel formats outside include/tvscale.h.
// AVPixelFormat values are those of libavutil 60 (FFmpeg 8.0); they move between major releases, so
// tvs_shim_map_pix_fmt() lets the host program (which has the real headers) override the table.
#include <cerrno>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "../../include/tvscale.h"

#define SHIM_API extern "C" __attribute__((visibility("default")))

namespace {

struct PixMap { int av, tvs; };
PixMap g_map[16] = {{0, TVS_FMT_YUV420P}, {23, TVS_FMT_NV12}, {62, TVS_FMT_YUV420P10LE}, {158, TVS_FMT_P010LE},
                    {2, TVS_FMT_RGB24},  {3, TVS_FMT_BGR24}, {12, TVS_FMT_YUVJ420P}, {35, TVS_FMT_RGB48LE}, {34, TVS_FMT_RGB48BE},
                    {58, TVS_FMT_BGR48LE}, {57, TVS_FMT_BGR48BE}};
int g_map_n = 11;

int to_tvs(int av) {
    for (int i = 0; i < g_map_n; ++i)
        if (g_map[i].av == av) return g_map[i].tvs;
    return -1;
}

int planes_of(int tvs_fmt) {
    switch (tvs_fmt) {
    case TVS_FMT_NV12: case TVS_FMT_P010LE: return 2;
    case TVS_FMT_RGB24: case TVS_FMT_BGR24: case TVS_FMT_RGB48LE: case TVS_FMT_RGB48BE: case TVS_FMT_BGR48LE: case TVS_FMT_BGR48BE: return 1;
    default: return 3;
    }
}

}  // namespace

struct SwsContext {
    tvs_ctx *ctx;
    int src_w, src_h, dst_w, dst_h, src_planes, dst_planes;
};
struct SwsFilter;

// Register (or replace) one AVPixelFormat value -> TVS_FMT_* pair.  Returns 0, or -1 when the table is full.
SHIM_API int tvs_shim_map_pix_fmt(int av_pix_fmt, int tvs_fmt) {
    for (int i = 0; i < g_map_n; ++i)
        if (g_map[i].av == av_pix_fmt) { g_map[i].tvs = tvs_fmt; return 0; }
    if (g_map_n == 16) return -1;
    g_map[g_map_n++] = {av_pix_fmt, tvs_fmt};
    return 0;
}

SHIM_API SwsContext *sws_getContext(int srcW, int srcH, int srcFormat, int dstW, int dstH, int dstFormat, int flags,
                                    SwsFilter *srcFilter, SwsFilter *dstFilter, const double *param) {
    if (srcFilter || dstFilter) return nullptr;
    const int sf = to_tvs(srcFormat), df = to_tvs(dstFormat);
    if (sf < 0 || df < 0) return nullptr;
    // the algorithm bits pass through; everything else libswscale accepts here (SWS_PRINT_INFO, SWS_FULL_CHR_*, SWS_DIRECT_BGR,
    // SWS_ERROR_DIFFUSION ...) either does not apply to these formats or asks for the rounding this library always uses
    int alg = flags & (TVS_BILINEAR | TVS_BICUBIC | TVS_AREA | TVS_LANCZOS);
    if (!alg) alg = TVS_BICUBIC;
    const char *dev = getenv("TVS_DEVICE");
    tvs_ctx *ctx = tvs_create_ex(dev ? atoi(dev) : 0, srcW, srcH, sf, dstW, dstH, df, alg, nullptr, param);
    if (!ctx) return nullptr;
    SwsContext *c = new SwsContext{ctx, srcW, srcH, dstW, dstH, planes_of(sf), planes_of(df)};
    return c;
}

SHIM_API int sws_scale(SwsContext *c, const uint8_t *const srcSlice[], const int srcStride[], int srcSliceY, int srcSliceH,
                       uint8_t *const dst[], const int dstStride[]) {
    if (!c || !srcSlice || !dst || !srcStride || !dstStride) return -EINVAL;
    if (srcSliceY != 0 || srcSliceH != c->src_h) return -ENOSYS;   // whole frames only
    const void *sp[3] = {nullptr, nullptr, nullptr};
    void *dp[3] = {nullptr, nullptr, nullptr};
    ptrdiff_t ss[3] = {0, 0, 0}, ds[3] = {0, 0, 0}, zero[3] = {0, 0, 0};
    for (int i = 0; i < c->src_planes; ++i) {
        if (srcStride[i] < 0) return -ENOSYS;
        sp[i] = srcSlice[i]; ss[i] = srcStride[i];
    }
    for (int i = 0; i < c->dst_planes; ++i) {
        if (dstStride[i] < 0) return -ENOSYS;
        dp[i] = dst[i]; ds[i] = dstStride[i];
    }
    return tvs_scale_host(c->ctx, sp, ss, zero, dp, ds, zero, 1);
}

SHIM_API void sws_freeContext(SwsContext *c) {
    if (!c) return;
    tvs_destroy(c->ctx);
    delete c;
}
