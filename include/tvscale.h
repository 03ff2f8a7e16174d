/*
 * tvscale.h — C ABI of the B200-native scale / pixel-format stage for thinvids.
 *
 * What it replaces.  The reference's `encode` task (worker/tasks.py:1354) spawns one ffmpeg child per
 * segment with `-vf setpts=PTS-STARTPTS,[bwdif,]scale=-2:H[,format=nv12,hwupload]`
 * (worker/tasks.py:62-65,436-449,1558-1586); tools/upscale_benchmark.py:163,182 does the same with
 * `scale=-2:H:flags=lanczos,format=yuv420p|nv12`.  Inside ffmpeg that filter is libswscale's public trio
 * sws_getContext / sws_scale / sws_freeContext (libswscale/swscale.h; third-party, not vendored under
 * /root/reference).  The entry points below are that trio re-cut for device memory and batches of
 * frames; a binding that used to call libswscale binds these instead.
 *
 * Semantics.  Output bytes are identical to libswscale 9.1.100 run with
 * flags = alg | SWS_ACCURATE_RND | SWS_BITEXACT, for 4:2:0 8- and 10-bit planar and
 * semi-planar formats, packed RGB sources, full <-> limited range conversion, scaler parameters and the cascade
 * libswscale runs for extreme down-scales.  There is no CPU fallback: every pixel is produced by sm_100a CUDA kernels, and
 * every call fails with a negative code when no usable device is present.
 *
 * Conventions.  All functions return 0 (or a positive count where stated) on success and a negative
 * errno-style code on failure; tvs_last_error() then holds a human-readable message for the calling
 * thread.  A context belongs to one device and, like a SwsContext, must not be used by two threads at
 * once.  Plain C types only: no CUDA or torch types appear in any signature (streams travel as void*).
 * Every entry point that takes a context makes that context's device the calling thread's current CUDA
 * device (cudaSetDevice) and leaves it current; a host thread that drives several GPUs must not assume
 * its previous device is still selected after a call.
 *
 * One deliberate difference from libswscale's bytes: for a same-size planar -> p010le conversion of an
 * ODD-width frame libswscale writes only floor(w/2) chroma pairs and leaves the last pair of every chroma
 * row untouched; this library writes all ceil(w/2) pairs (the value the scaler proper produces), so that
 * every destination byte is defined.  Everything else is byte-identical.
 */
#ifndef TVSCALE_H
#define TVSCALE_H

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define TVS_API __attribute__((visibility("default")))
#else
#define TVS_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* Pixel formats (the 4:2:0 subset of AVPixelFormat the reference's filter strings produce). */
enum {
    TVS_FMT_YUV420P     = 0, /* AV_PIX_FMT_YUV420P      planes Y, U, V            8 bit  */
    TVS_FMT_NV12        = 1, /* AV_PIX_FMT_NV12         planes Y, UV interleaved  8 bit  */
    TVS_FMT_YUV420P10LE = 2, /* AV_PIX_FMT_YUV420P10LE  planes Y, U, V            10 bit in the low bits of LE16  */
    TVS_FMT_P010LE      = 3, /* AV_PIX_FMT_P010LE       planes Y, UV interleaved  10 bit in the high bits of LE16; as a source the
                                six bits below may hold anything (they are shifted away, as in libswscale; the same-size
                                p010le -> p010le copy keeps them except in every fourth word and the row tail, as libswscale's
                                does).  Samples of
                                TVS_FMT_YUV420P10LE sources must not exceed 1023: libswscale feeds larger words to its filters
                                as they are, the float kernels here assume 10 bits */
    TVS_FMT_RGB24       = 4, /* AV_PIX_FMT_RGB24        one plane of R, G, B bytes; SOURCE only, even widths (what
                                tools/upscale_benchmark.py:159-182 feeds its lanczos stage); BT.601 limited-range output */
    TVS_FMT_YUVJ420P    = 5, /* AV_PIX_FMT_YUVJ420P     yuv420p holding full-range (JPEG) samples: what ffmpeg's decoders hand
                                the scale filter for MJPEG and most phone / camera H.264.  As source it implies
                                TVS_SRC_FULL_RANGE, as destination TVS_DST_FULL_RANGE (libswscale's handle_jpeg) */
    TVS_FMT_BGR24       = 6, /* AV_PIX_FMT_BGR24        as TVS_FMT_RGB24 with B, G, R byte order (OpenCV's frames); SOURCE only */
    TVS_FMT_RGB48LE     = 7, /* AV_PIX_FMT_RGB48LE      R, G, B of 16 bits each, little-endian; SOURCE only, even widths */
    TVS_FMT_RGB48BE     = 8, /* AV_PIX_FMT_RGB48BE      the same big-endian (what ffmpeg decodes a 16-bit PNG to); SOURCE only */
    TVS_FMT_BGR48LE     = 9, /* AV_PIX_FMT_BGR48LE      B, G, R of 16 bits each (OpenCV's 16-bit frames); SOURCE only */
    TVS_FMT_BGR48BE     = 10 /* AV_PIX_FMT_BGR48BE      SOURCE only */
};

/* Scaler flags: numerically equal to libswscale's SWS_* so ffmpeg `flags=` values pass through. */
enum {
    TVS_BILINEAR     = 2,
    TVS_BICUBIC      = 4,      /* ffmpeg's default when the filter string has no flags= (worker/tasks.py:448) */
    TVS_AREA         = 0x20,
    TVS_LANCZOS      = 0x200,  /* tools/upscale_benchmark.py:163,182 */
    TVS_ACCURATE_RND = 0x40000,/* accepted; results are always accurate_rnd+bitexact */
    TVS_BITEXACT     = 0x80000,
    /* Sample ranges: libswscale's src_range / dst_range options (what vf_scale's in_range / out_range, or the frame's own
     * colour range, set).  Not SWS_* bits.  When exactly one is set, the 15-bit intermediate is converted between full
     * (JPEG) and limited (MPEG) range after the horizontal pass, with libswscale 9.1's constants and rounding (restated in
     * oracle/swscale_port.c, checked against the genuine library); such contexts run on the integer two-pass kernels.
     * A context that cascades (TVS_INFO_PATH 6) converts no range - neither does libswscale.  -ENOSYS for rgb24 sources. */
    TVS_SRC_FULL_RANGE = 0x10000000,
    TVS_DST_FULL_RANGE = 0x20000000
};

/* Chroma siting sentinel: "use libswscale's default" (centre), as sws_getContext does. */
#define TVS_CHR_POS_DEFAULT (-513)

typedef struct tvs_ctx tvs_ctx;

/*
 * Replaces sws_getContext(srcW, srcH, srcFormat, dstW, dstH, dstFormat, flags, NULL, NULL, NULL)
 * (libswscale/swscale.h), i.e. what ffmpeg's vf_scale does for `scale=W:H:flags=...` at
 * worker/tasks.py:1565-1567,1580.  Builds the four fixed-point coefficient tables on the host
 * (SURVEY.md A.2), uploads them to `device`, and plans the kernel launches.
 * chr_pos = {src_h, src_v, dst_h, dst_v} in 1/256 luma-pixel units (ffmpeg's *_chr_pos options) or NULL.
 * Returns NULL on failure.
 */
TVS_API tvs_ctx *tvs_create(int device, int src_w, int src_h, int src_fmt, int dst_w, int dst_h, int dst_fmt,
                    int flags, const int chr_pos[4]);

/* tvs_create with libswscale's two scaler parameters (sws_getContext's `param`, ffmpeg's scale=...:param0=..:param1=..):
 * bicubic B = param[0], C = param[1] (defaults 0 and 0.6); lanczos lobes = param[0] (default 3, at least 1).  An entry equal to
 * TVS_PARAM_DEFAULT (= SWS_PARAM_DEFAULT), or param == NULL, selects the default.  The other algorithms ignore them. */
#define TVS_PARAM_DEFAULT 123456.0
TVS_API tvs_ctx *tvs_create_ex(int device, int src_w, int src_h, int src_fmt, int dst_w, int dst_h, int dst_fmt,
                               int flags, const int chr_pos[4], const double param[2]);

/* Replaces sws_freeContext(). */
TVS_API void tvs_destroy(tvs_ctx *ctx);

/*
 * Replaces sws_scale(ctx, src, srcStride, 0, srcH, dst, dstStride) for `n_frames` frames that are
 * already in DEVICE memory (decoder surfaces, or a previous filter's output).
 *   src[i] / dst[i]          device pointer to plane i of frame 0 (2 planes for NV12/P010, else 3)
 *   *_pitch[i]               bytes between rows of plane i (any value >= the row size; NVDEC pitches fit)
 *   *_frame_stride[i]        bytes between the same plane of consecutive frames
 *   stream                   a cudaStream_t passed as void* (NULL = default stream)
 * Asynchronous with respect to the host: returns once the work is enqueued on `stream`.
 * Returns dst_h (like sws_scale) or a negative error code.
 */
TVS_API int tvs_scale_batch(tvs_ctx *ctx,
                    const void *const src[3], const ptrdiff_t src_pitch[3], const ptrdiff_t src_frame_stride[3],
                    void *const dst[3], const ptrdiff_t dst_pitch[3], const ptrdiff_t dst_frame_stride[3],
                    int n_frames, void *stream);

/*
 * The literal drop-in for sws_scale(): HOST planes in, HOST planes out, same plane/stride meaning,
 * `n_frames` frames spaced by *_frame_stride bytes.  Copies the frames to the context's device,
 * scales them there and copies the result back, in chunks that overlap copies with kernels on two
 * streams; returns after the last byte has landed in dst.  Host buffers may be pageable; pinned
 * buffers (cudaHostAlloc / torch pin_memory) are copied without an intermediate staging step.
 * Returns dst_h or a negative error code.
 */
TVS_API int tvs_scale_host(tvs_ctx *ctx,
                   const void *const src[3], const ptrdiff_t src_pitch[3], const ptrdiff_t src_frame_stride[3],
                   void *const dst[3], const ptrdiff_t dst_pitch[3], const ptrdiff_t dst_frame_stride[3],
                   int n_frames);

/*
 * No counterpart in the reference or in libswscale.
 *
 * tvs_scale_batch() for callers that must not allocate device memory when they scale.  Two kinds of caller are in
 * that position: a worker whose GPU memory has been handed out in full before the first frame arrives (decoder
 * surface pools, the encoder's buffers, another framework's arena), for whom a late cudaMalloc can simply fail; and
 * a frame loop that cannot afford the device-wide synchronisation a cudaMalloc / cudaFree implies while its other
 * streams are busy.  This call runs the two-pass kernels (horizontal launch, vertical launch) and keeps the int16
 * intermediate between them in a fixed array that is part of the library image, TVS_STATIC_SCRATCH_BYTES long, so it
 * performs no cudaMalloc, cudaFree or other allocating call.  Same arguments, result bytes and return value as
 * tvs_scale_batch(); the caller still provides source and destination planes in device memory.
 *
 * Limits and rules:
 *   - n_frames x (intermediate of one frame: 2 bytes x dst_w x src_h for luma plus the two chroma planes) must fit in
 *     TVS_STATIC_SCRATCH_BYTES, e.g. 12 frames of 720x480 -> 1920x1080 or 2 frames of 1920x1080 -> 3840x2160.
 *     Larger requests fail with -E2BIG and a message that states the per-call maximum; split the batch.
 *   - The array's device address is looked up once per device (cudaGetSymbolAddress, with that device current).
 *     tvs_create() does this, so the array's TVS_STATIC_SCRATCH_BYTES are taken from the device when the first
 *     context is created there - the one moment at which this library allocates - and not by a scale call.  Should
 *     that lookup have failed at creation, the first call here repeats it; make that first call before starting
 *     any stream capture.
 *   - There is ONE array per device, shared by every context and every call in the process, and it keeps whatever
 *     the last call left in it.  Nothing in it needs clearing (each call overwrites what it later reads), but two
 *     calls whose kernels can run at the same time on one GPU - different streams, or different threads - race on
 *     it.  Issue these calls on one stream, or order them with events.
 *   - Every GPU that loads the library holds the array until the process exits.
 */
#define TVS_STATIC_SCRATCH_BYTES (32u << 20)
TVS_API int tvs_scale_batch_static(tvs_ctx *ctx,
                    const void *const src[3], const ptrdiff_t src_pitch[3], const ptrdiff_t src_frame_stride[3],
                    void *const dst[3], const ptrdiff_t dst_pitch[3], const ptrdiff_t dst_frame_stride[3],
                    int n_frames, void *stream);

/*
 * Replaces ffmpeg's `bwdif=mode=send_frame:parity=auto:deint=all` (libavfilter; the filter the reference puts in front of
 * `scale` on its DVD path: worker/tasks.py:62-63,446-447, tools/upscale_benchmark.py:128-132).  PARITY UNPINNED: there is no
 * libavfilter in the build image to compare with; the kernel follows the published algorithm and is tested against a
 * numpy restatement of it (oracle/bwdif_port.py) and against its invariants.  The `encode` task keeps ffmpeg's own
 * bwdif in its decode child unless asked otherwise.
 *
 * `src` / `dst`: planar 4:2:0 frames (TVS_FMT_YUV420P or TVS_FMT_YUV420P10LE), `n_frames` of them in stream order, same
 * plane / pitch / frame-stride meaning as tvs_scale_batch.  Frame i is interpolated from frames i - 1, i, i + 1.  `tff`:
 * the frames are top-field-first (the top field is kept), else bottom-field-first.  `flags`: TVS_BWDIF_STREAM_START =
 * frame 0 is the first frame of the stream (it has no predecessor and gets the intra filter), otherwise frame 0 is only
 * context for frame 1 and is not written; TVS_BWDIF_STREAM_END likewise for the last frame.  A stream is processed in
 * batches by repeating one frame of context on either side.  Returns the number of frames written, or a negative code.
 * tvs_bwdif_host is the same on tightly packed host frames (synchronous, allocates its device staging).
 */
enum { TVS_BWDIF_STREAM_START = 1, TVS_BWDIF_STREAM_END = 2 };
TVS_API int tvs_bwdif_batch(int device, int fmt, int width, int height,
                    const void *const src[3], const ptrdiff_t src_pitch[3], const ptrdiff_t src_frame_stride[3],
                    void *const dst[3], const ptrdiff_t dst_pitch[3], const ptrdiff_t dst_frame_stride[3],
                    int n_frames, int tff, int flags, void *stream);
TVS_API int tvs_bwdif_host(int device, int fmt, int width, int height, const void *src, void *dst, int n_frames, int tff, int flags);

/*
 * No counterpart in libswscale (which works on pageable memory): page-locked host buffers for tvs_scale_host().
 * Frames read into / written from such a buffer cross PCIe without the driver's intermediate staging copy and let the
 * host entry point overlap its copies with its kernels.  thinvids_b200/worker.py reads the decoder's raw video into a
 * small ring of these.  Needs a CUDA device (returns NULL and sets tvs_last_error() without one).
 */
TVS_API void *tvs_host_alloc(size_t bytes);
TVS_API void tvs_host_free(void *p);

/*
 * Host-only: one axis of libswscale's initFilter.  No device needed.
 *   inc = ((src_len << 16) + (dst_len >> 1)) / dst_len, one = 1<<14 (horizontal) or 1<<12 (vertical),
 *   src_pos / dst_pos = 128 for luma and centred chroma.
 * Writes taps*dst_len int16 coefficients and dst_len int32 start positions when the buffers are large
 * enough (coef_cap / pos_cap count elements); returns the tap count, or a negative error code.
 * Calling with coef == NULL just returns the tap count.
 */
TVS_API int tvs_filter_table(int src_len, int dst_len, int alg, int one, int src_pos, int dst_pos,
                     int16_t *coef, size_t coef_cap, int32_t *pos, size_t pos_cap);

/* Read back a context's table: which = 0 hLum, 1 hChr, 2 vLum, 3 vChr.  Same buffer rules as above. */
TVS_API int tvs_context_table(const tvs_ctx *ctx, int which, int16_t *coef, size_t coef_cap, int32_t *pos, size_t pos_cap);

/*
 * Plan introspection for benchmarks and tests.
 *   TVS_INFO_KERNEL_LAUNCHES  kernels launched by this context so far (monotonic)
 *   TVS_INFO_PATH             0 = fused tile kernel, 1 = two-pass kernels through device scratch,
 *                             2 = same-size repack only, 3 = streaming (line-buffer) kernel for vertical down-scales,
 *                             4 = box-filter kernel (exact 2:1 area on 8-bit planes),
 *                             5 = slab kernel (one autonomous warp per 128-column slab),
 *                             6 = cascade: two inner contexts, src -> tmp -> dst, as libswscale does for filters that reach
 *                                 170 taps (extreme down-scales); each stage picks its own kernel.
 *                             Shapes that fall back to 1 (two-pass): contexts that convert between full and limited
 *                             range, filters wider than the slab / fused / streaming instances cover (more than 24
 *                             horizontal or 12 vertical taps), tables that fail the FP32 exactness preconditions, and
 *                             footprints that do not fit in shared memory.
 *   TVS_INFO_SCRATCH_BYTES    device scratch currently owned by the context
 *   TVS_INFO_SRC_FRAME_BYTES / TVS_INFO_DST_FRAME_BYTES   tightly packed frame sizes
 */
enum { TVS_INFO_KERNEL_LAUNCHES = 0, TVS_INFO_PATH = 1, TVS_INFO_SCRATCH_BYTES = 2,
       TVS_INFO_SRC_FRAME_BYTES = 3, TVS_INFO_DST_FRAME_BYTES = 4 };
TVS_API long long tvs_info(const tvs_ctx *ctx, int what);

/* Force a kernel path for testing: -1 = automatic, 0 = fused tiles, 1 = two-pass, 2 = streaming, 3 = slab.
 * Returns 0, -EINVAL, or -ENOSYS when the requested kernel does not handle this context's shape. */
TVS_API int tvs_set_path(tvs_ctx *ctx, int path);

/* Message describing the last failure on the calling thread ("" if none). */
TVS_API const char *tvs_last_error(void);

/* Library version string, e.g. "tvscale 0.1 (sm_100a)". */
TVS_API const char *tvs_version(void);

#ifdef __cplusplus
}
#endif
#endif /* TVSCALE_H */
