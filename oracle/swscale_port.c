/*
 * oracle/swscale_port.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, single-threaded restatement of the arithmetic that the reference
 * hands to ffmpeg's `scale` filter (FFmpeg libswscale 9.1.100 / FFmpeg 8.0.1,
 * legacy scaler, flags = alg | SWS_ACCURATE_RND | SWS_BITEXACT).  The reference
 * reaches it at worker/tasks.py:62-65,444-449,1565-1567,1580 and
 * tools/upscale_benchmark.py:163,182.  libswscale's source is NOT vendored under
 * /root/reference (it is the distro `ffmpeg` package, ansible_workers.yml:41), so
 * this file restates the published algorithm as specified in SURVEY.md Appendix A
 * (upstream names: libswscale/utils.c initFilter, swscale.c hScale8To15_c /
 * hScale16To15_c, output.c yuv2planeX_8_c / yuv2planeX_10_c / yuv2nv12cX_c /
 * yuv2p010*, input.c nv12ToUV_c / p010LEToY_c, swscale_unscaled.c wrappers).
 *
 * Parity is PINNED: tests/test_oracle.py checks this port byte-for-byte against
 * the genuine libswscale binary bundled in this image (oracle/ref_swscale.py) and
 * against the golden SHA-256 vectors in tests/golden/ (SURVEY.md Appendix B).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it.
 * The product (thinvids_b200/) never links, imports or calls anything in oracle/.
 *
 * Build:  gcc -O2 -ffp-contract=off -fPIC -shared -o _build/libtvoracle.so swscale_port.c -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { FMT_YUV420P = 0, FMT_NV12 = 1, FMT_YUV420P10 = 2, FMT_P010 = 3, FMT_RGB24 = 4 /* source only */, FMT_BGR24 = 6 /* source only */,
       FMT_RGB48LE = 7, FMT_RGB48BE = 8, FMT_BGR48LE = 9, FMT_BGR48BE = 10 /* 16 bits per component, source only */ };
static int fmt_rgb48(int f) { return f >= FMT_RGB48LE && f <= FMT_BGR48BE; }
static int fmt_rgb(int f) { return f == FMT_RGB24 || f == FMT_BGR24 || fmt_rgb48(f); }
enum { ALG_BILINEAR = 2, ALG_BICUBIC = 4, ALG_AREA = 0x20, ALG_LANCZOS = 0x200 };

typedef struct {
    int n;        /* rows (= destination length) */
    int taps;     /* taps per row */
    int16_t *f;   /* n * taps */
    int32_t *pos; /* n */
} o_filter;

typedef struct {
    int sw, sh, dw, dh, sfmt, dfmt, alg;
    int csw, csh, cdw, cdh;
    int rgb_half;       /* FMT_RGB24 source: chroma is made at half width */
    uint16_t *rgb[3];   /* FMT_RGB24 source: the input stage's Y, U, V lines for the frame being scaled */
    o_filter hl, hc, vl, vc;
    /* Cascade (libswscale utils.c, sws_init_context's RETCODE_USE_CASCADE branch): with accurate_rnd a filter that, padded
     * to the table alignment of the x86 build (4 taps horizontally, 2 vertically), reaches 256 * 16 / 24 = 171 taps is not
     * run; the context becomes two contexts, src -> tmp -> dst, with tmp a yuv420p frame of int(sqrt(srcW * dstW)) x
     * int(sqrt(srcH * dstH)).  The two inner contexts take the library's DEFAULT chroma positions whatever the outer
     * context was given (pinned against the genuine library, tests/test_oracle.py). */
    void *cascade[2];
    uint8_t *tmp[3];
    int tmp_stride[3];
    /* Range conversion (swscale_internal.h lumConvertRange / chrConvertRange, FFmpeg >= 7.1 form): applied to the
     * 15-bit intermediate right after the horizontal pass when source and destination ranges differ,
     *   t' = min((t * coeff + offset) >> 14, 32767).
     * Constants as ff_sws_init_range_convert derives them from the DESTINATION bit depth; the form and the values were
     * checked against the genuine library by tools/fit_range_convert.py (unique coeff, offset inside a 60-wide window). */
    int range_convert;            /* 0 none, 1 full -> limited, 2 limited -> full */
    int rc_coeff[2];              /* luma, chroma */
    int64_t rc_offset[2];
} o_ctx;

static int ilog2(unsigned v) { int r = 0; while (v > 1) { v >>= 1; r++; } return r; }
static int64_t iabs64(int64_t v) { return v < 0 ? -v : v; }
static int fmt_bits(int f) { return (f == FMT_YUV420P || f == FMT_NV12 || f == FMT_RGB24 || f == FMT_BGR24) ? 8 : (f >= FMT_RGB48LE && f <= FMT_BGR48BE) ? 16 : 10; }
static int fmt_semi(int f) { return f == FMT_NV12 || f == FMT_P010; }

/* SURVEY.md A.2 — coefficient table for one axis. */
/* libswscale's SWS_PARAM_DEFAULT: "use the algorithm's own default" for a scaler parameter (vf_scale's param0 / param1):
 * bicubic B = param[0] (0), C = param[1] (0.6); lanczos lobes = param[0] (3), support ceil(2 * lobes). */
#define PARAM_DEFAULT 123456.0
static double g_param[2] = { PARAM_DEFAULT, PARAM_DEFAULT };   /* set around tvo_create_ex; the port is single-threaded test code */

static int make_filter(o_filter *out, int inc, int src_len, int dst_len, int one, int alg,
                       int src_pos, int dst_pos)
{
    const int ratio_log = ilog2((unsigned)(src_len / dst_len));
    const int64_t fone = (int64_t)1 << (54 - (ratio_log < 8 ? ratio_log : 8));
    int taps;
    int64_t *w = NULL;
    int32_t *pos = (int32_t *)calloc((size_t)dst_len, sizeof(int32_t));
    if (!pos) return -1;

    if (abs(inc - 0x10000) < 10 && src_pos == dst_pos) {
        taps = 1;
        w = (int64_t *)calloc((size_t)dst_len, sizeof(int64_t));
        for (int i = 0; i < dst_len; i++) { w[i] = fone; pos[i] = i; }
    } else if (alg == ALG_AREA && inc <= 0x10000) {
        /* area used for enlarging degenerates to a 2-tap linear kernel */
        taps = 2;
        w = (int64_t *)calloc((size_t)dst_len * taps, sizeof(int64_t));
        int64_t centre = (((int64_t)dst_pos * inc) >> 8) - (((int64_t)src_pos * 0x8000) >> 7);
        for (int i = 0; i < dst_len; i++) {
            int64_t first = (centre - ((int64_t)(taps - 1) << 15) + (1 << 15)) >> 16;
            pos[i] = (int32_t)first;
            for (int j = 0; j < taps; j++) {
                int64_t c = fone - iabs64(first * 65536 - centre) * (fone >> 16);
                w[(size_t)i * taps + j] = c < 0 ? 0 : c;
                first++;
            }
            centre += inc;
        }
    } else {
        int support = alg == ALG_BICUBIC ? 4 : alg == ALG_LANCZOS ? (g_param[0] != PARAM_DEFAULT ? (int)ceil(2 * g_param[0]) : 6) : alg == ALG_AREA ? 1 : 2;
        if (inc <= 0x10000) taps = 1 + support;
        else taps = 1 + (int)(((int64_t)support * src_len + dst_len - 1) / dst_len);
        if (taps > src_len - 2) taps = src_len - 2;
        if (taps < 1) taps = 1;
        w = (int64_t *)calloc((size_t)dst_len * taps, sizeof(int64_t));
        int64_t centre = (((int64_t)dst_pos * inc) >> 7) - (((int64_t)src_pos * 0x10000) >> 7);
        for (int i = 0; i < dst_len; i++) {
            int64_t first = (centre - (int64_t)(taps - 2) * 65536) / 131072;
            pos[i] = (int32_t)first;
            for (int j = 0; j < taps; j++) {
                int64_t d = iabs64(first * 131072 - centre) << 13;
                int64_t c;
                if (inc > 0x10000) d = d * dst_len / src_len;
                double fd = (double)d * (1.0 / (1 << 30));
                if (alg == ALG_BICUBIC) {
                    const int64_t B = (int64_t)((g_param[0] != PARAM_DEFAULT ? g_param[0] : 0) * (1 << 24));
                    const int64_t C = (int64_t)((g_param[1] != PARAM_DEFAULT ? g_param[1] : 0.6) * (1 << 24));
                    if (d >= ((int64_t)1 << 31)) {
                        c = 0;
                    } else {
                        int64_t dd = (d * d) >> 30;
                        int64_t ddd = (dd * d) >> 30;
                        if (d < ((int64_t)1 << 30))
                            c = (12 * (1 << 24) - 9 * B - 6 * C) * ddd +
                                (-18 * (1 << 24) + 12 * B + 6 * C) * dd +
                                (6 * (1 << 24) - 2 * B) * ((int64_t)1 << 30);
                        else
                            c = (-B - 6 * C) * ddd + (6 * B + 30 * C) * dd +
                                (-12 * B - 48 * C) * d + (8 * B + 24 * C) * ((int64_t)1 << 30);
                    }
                    c /= ((int64_t)1 << 54) / fone;
                } else if (alg == ALG_AREA) {
                    int64_t d2 = d - (1 << 29);
                    if (d2 * inc < -((int64_t)1 << 45)) c = (int64_t)1 << 46;
                    else if (d2 * inc < ((int64_t)1 << 45)) c = -d2 * inc + ((int64_t)1 << 45);
                    else c = 0;
                    c *= fone >> 46;
                } else if (alg == ALG_LANCZOS) {
                    const double p = g_param[0] != PARAM_DEFAULT ? g_param[0] : 3.0;
                    double v = d ? sin(fd * M_PI) * sin(fd * M_PI / p) / (fd * fd * M_PI * M_PI / p) : 1.0;
                    c = (int64_t)(v * (double)fone);
                    if (fd > p) c = 0;
                } else { /* bilinear */
                    c = ((int64_t)1 << 30) - d;
                    if (c < 0) c = 0;
                    c *= fone >> 30;
                }
                w[(size_t)i * taps + j] = c;
                first++;
            }
            centre += 2 * (int64_t)inc;
        }
    }

    /* step 1: drop negligible taps (walk from the last row to the first) */
    const double cut = 0.002 * (double)fone;
    int keep = 0;
    for (int i = dst_len - 1; i >= 0; i--) {
        int64_t *row = w + (size_t)i * taps;
        int64_t acc = 0;
        for (int rep = 0; rep < taps; rep++) {
            acc += iabs64(row[0]);
            if ((double)acc > cut) break;
            if (i < dst_len - 1 && pos[i] >= pos[i + 1]) break;
            memmove(row, row + 1, (size_t)(taps - 1) * sizeof(int64_t));
            row[taps - 1] = 0;
            pos[i]++;
        }
        acc = 0;
        int m = taps;
        for (int j = taps - 1; j > 0; j--) {
            acc += iabs64(row[j]);
            if ((double)acc > cut) break;
            m--;
        }
        if (m > keep) keep = m;
    }
    const int nt = keep; /* filterAlign = 1 under BITEXACT */
    int64_t *v = (int64_t *)calloc((size_t)dst_len * nt, sizeof(int64_t));
    for (int i = 0; i < dst_len; i++)
        memcpy(v + (size_t)i * nt, w + (size_t)i * taps, (size_t)nt * sizeof(int64_t));
    free(w);

    /* step 2: fold taps that fall outside the source onto the edge samples */
    for (int i = 0; i < dst_len; i++) {
        int64_t *row = v + (size_t)i * nt;
        if (pos[i] < 0) {
            for (int j = 1; j < nt; j++) {
                int left = j + pos[i];
                if (left < 0) left = 0;
                row[left] += row[j];
                row[j] = 0;
            }
            pos[i] = 0;
        }
        if (pos[i] + nt > src_len) {
            int over = nt - src_len;
            int shift = pos[i] + (over < 0 ? over : 0);
            int64_t acc = 0;
            for (int j = nt - 1; j >= 0; j--)
                if (pos[i] + j >= src_len) { acc += row[j]; row[j] = 0; }
            for (int j = nt - 1; j >= 0; j--)
                row[j] = j < shift ? 0 : row[j - shift];
            pos[i] -= shift;
            row[src_len - 1 - pos[i]] += acc;
        }
    }

    /* step 3: normalise to `one` with error feedback along the row */
    out->f = (int16_t *)calloc((size_t)dst_len * nt, sizeof(int16_t));
    for (int i = 0; i < dst_len; i++) {
        int64_t *row = v + (size_t)i * nt;
        int64_t sum = 0, err = 0;
        for (int j = 0; j < nt; j++) sum += row[j];
        sum = (sum + one / 2) / one;
        if (!sum) sum = 1;
        for (int j = 0; j < nt; j++) {
            int64_t val = row[j] + err;
            int q = (int)(val >= 0 ? (val + (sum >> 1)) / sum : (val - (sum >> 1)) / sum);
            out->f[(size_t)i * nt + j] = (int16_t)q;
            err = val - (int64_t)q * sum;
        }
    }
    free(v);
    out->n = dst_len;
    out->taps = nt;
    out->pos = pos;
    return 0;
}

static int local_pos(int p, int sub)
{
    if (p == -1 || p <= -513) p = (128 << sub) - 128;
    p += 128;
    return p >> sub;
}

static int step16(int s, int d) { return (int)((((int64_t)s << 16) + (d >> 1)) / d); }

void tvo_destroy(void *p);

/* RETCODE_USE_CASCADE of initFilter, for flags that include SWS_ACCURATE_RND on an x86 build: filterAlign is 4 for the
 * horizontal tables and 2 for the vertical ones, the limit MAX_FILTER_SIZE * 16 / APCK_SIZE = 256 * 16 / 24. */
int tvo_needs_cascade(int h_lum_taps, int h_chr_taps, int v_lum_taps, int v_chr_taps)
{
    const int limit = 256 * 16 / 24;   /* 170 in integer arithmetic: a padded filter size >= 170 cascades */
    return ((h_lum_taps + 3) & ~3) >= limit || ((h_chr_taps + 3) & ~3) >= limit ||
           ((v_lum_taps + 1) & ~1) >= limit || ((v_chr_taps + 1) & ~1) >= limit;
}

void *tvo_create(int sw, int sh, int sfmt, int dw, int dh, int dfmt, int alg, const int *chr_pos);
/* Turn `c` into a cascade when one of its filters is too long (see o_ctx); 0 = fine, -1 = libswscale refuses. */
static int tvo_setup_cascade(o_ctx *c)
{
    const int sw = c->sw, sh = c->sh, dw = c->dw, dh = c->dh, sfmt = c->sfmt, dfmt = c->dfmt, alg = c->alg;
    if (!tvo_needs_cascade(c->hl.taps, c->hc.taps, c->vl.taps, c->vc.taps)) return 0;
    const int tw = (int)sqrt((double)sw * dw), th = (int)sqrt((double)sh * dh);
    if ((int64_t)sw * sh <= 4LL * dw * dh) return -1;   /* the library gives up here (EINVAL) */
    c->cascade[0] = tvo_create(sw, sh, sfmt, tw, th, FMT_YUV420P, alg, NULL);
    c->cascade[1] = tvo_create(tw, th, FMT_YUV420P, dw, dh, dfmt, alg, NULL);
    if (!c->cascade[0] || !c->cascade[1]) return -1;
    const int cw = (tw + 1) >> 1, ch = (th + 1) >> 1;
    c->tmp_stride[0] = tw; c->tmp_stride[1] = c->tmp_stride[2] = cw;
    c->tmp[0] = (uint8_t *)malloc((size_t)tw * th);
    c->tmp[1] = (uint8_t *)malloc((size_t)cw * ch);
    c->tmp[2] = (uint8_t *)malloc((size_t)cw * ch);
    return 0;
}

static void solve_range(int src_min, int src_max, int dst_min, int dst_max, int src_shift, int mult_shift, int *coeff, int64_t *offset)
{
    const int total = mult_shift + src_shift;
    const uint64_t q = ((uint64_t)(dst_max - dst_min) << total) / (uint64_t)(src_max - src_min);
    *coeff = (int)((q + ((1u << src_shift) - 1)) >> src_shift);           /* AV_CEIL_RSHIFT */
    *offset = ((int64_t)dst_max << total) - ((int64_t)src_max << src_shift) * *coeff + (1 << (mult_shift - 1));
}

/* src_full / dst_full: 1 = full (JPEG) range, 0 = limited (MPEG) range: libswscale's src_range / dst_range options, which
 * ffmpeg's scale filter sets from in_range / out_range or the frame's colour range, and which the deprecated yuvj420p
 * pixel format implies.  Returns 0, or -1 where this restatement was not checked (RGB sources). */
int tvo_set_ranges(void *p, int src_full, int dst_full)
{
    o_ctx *c = (o_ctx *)p;
    c->range_convert = 0;
    if (!!src_full == !!dst_full) return 0;
    /* a cascade's two inner contexts are created with the default ranges, so the library converts nothing there
     * (pinned against it in tests/test_oracle.py) */
    if (c->cascade[0]) return 0;
    if (fmt_rgb(c->sfmt)) return -1;
    const int depth = fmt_bits(c->dfmt);
    const int src_shift = 15 - depth, mult_shift = 14;
    const int mpeg_min = 16 << (depth - 8), mpeg_max_lum = 235 << (depth - 8), mpeg_max_chr = 240 << (depth - 8);
    const int jpeg_max = (1 << depth) - 1;
    if (src_full) {
        c->range_convert = 1;
        solve_range(0, jpeg_max, mpeg_min, mpeg_max_lum, src_shift, mult_shift, &c->rc_coeff[0], &c->rc_offset[0]);
        solve_range(0, jpeg_max, mpeg_min, mpeg_max_chr, src_shift, mult_shift, &c->rc_coeff[1], &c->rc_offset[1]);
    } else {
        c->range_convert = 2;
        solve_range(mpeg_min, mpeg_max_lum, 0, jpeg_max, src_shift, mult_shift, &c->rc_coeff[0], &c->rc_offset[0]);
        solve_range(mpeg_min, mpeg_max_chr, 0, jpeg_max, src_shift, mult_shift, &c->rc_coeff[1], &c->rc_offset[1]);
    }
    return 0;
}

/* chr_pos: {src_h, src_v, dst_h, dst_v}, -513 = library default; may be NULL. */
void *tvo_create(int sw, int sh, int sfmt, int dw, int dh, int dfmt, int alg, const int *chr_pos)
{
    static const int dflt[4] = { -513, -513, -513, -513 };
    if (!chr_pos) chr_pos = dflt;
    o_ctx *c = (o_ctx *)calloc(1, sizeof(*c));
    c->sw = sw; c->sh = sh; c->dw = dw; c->dh = dh; c->sfmt = sfmt; c->dfmt = dfmt; c->alg = alg;
    c->csw = (sw + 1) >> 1; c->csh = (sh + 1) >> 1; c->cdw = (dw + 1) >> 1; c->cdh = (dh + 1) >> 1;
    if (fmt_rgb(sfmt)) {
        /* RGB has no chroma planes of its own: the input stage makes them, at full height and -- unless the
         * destination chroma is wider than half the source -- at half width (rgb24ToUV_half_c, the sum of two
         * pixels); both cases sit at libswscale's default (centred) positions. */
        c->rgb_half = (dw >> 1) <= (sw >> 1);
        c->csw = c->rgb_half ? (sw + 1) >> 1 : sw;
        c->csh = sh;
        make_filter(&c->hl, step16(sw, dw), sw, dw, 1 << 14, alg, local_pos(-513, 0), local_pos(-513, 0));
        make_filter(&c->hc, step16(c->csw, c->cdw), c->csw, c->cdw, 1 << 14, alg,
                    local_pos(chr_pos[0], c->rgb_half ? 1 : 0), local_pos(chr_pos[2], 1));
        make_filter(&c->vl, step16(sh, dh), sh, dh, 1 << 12, alg, local_pos(-513, 0), local_pos(-513, 0));
        make_filter(&c->vc, step16(c->csh, c->cdh), c->csh, c->cdh, 1 << 12, alg, local_pos(chr_pos[1], 0), local_pos(chr_pos[3], 1));
        if (tvo_setup_cascade(c)) { tvo_destroy(c); return NULL; }
        return c;
    }
    make_filter(&c->hl, step16(sw, dw), sw, dw, 1 << 14, alg, local_pos(-513, 0), local_pos(-513, 0));
    make_filter(&c->hc, step16(c->csw, c->cdw), c->csw, c->cdw, 1 << 14, alg,
                local_pos(chr_pos[0], 1), local_pos(chr_pos[2], 1));
    make_filter(&c->vl, step16(sh, dh), sh, dh, 1 << 12, alg, local_pos(-513, 0), local_pos(-513, 0));
    make_filter(&c->vc, step16(c->csh, c->cdh), c->csh, c->cdh, 1 << 12, alg,
                local_pos(chr_pos[1], 1), local_pos(chr_pos[3], 1));
    if (tvo_setup_cascade(c)) { tvo_destroy(c); return NULL; }
    return c;
}

/* tvo_create with scaler parameters (NULL = defaults); a cascade's inner contexts inherit them, as in libswscale. */
void *tvo_create_ex(int sw, int sh, int sfmt, int dw, int dh, int dfmt, int alg, const int *chr_pos, const double *param)
{
    if (param) { g_param[0] = param[0]; g_param[1] = param[1]; }
    void *c = tvo_create(sw, sh, sfmt, dw, dh, dfmt, alg, chr_pos);
    g_param[0] = g_param[1] = PARAM_DEFAULT;
    return c;
}

void tvo_destroy(void *p)
{
    o_ctx *c = (o_ctx *)p;
    if (!c) return;
    o_filter *fs[4] = { &c->hl, &c->hc, &c->vl, &c->vc };
    for (int i = 0; i < 4; i++) { free(fs[i]->f); free(fs[i]->pos); }
    for (int i = 0; i < 2; i++) if (c->cascade[i]) tvo_destroy(c->cascade[i]);
    for (int i = 0; i < 3; i++) free(c->tmp[i]);
    free(c);
}

/* which: 0 hLum, 1 hChr, 2 vLum, 3 vChr.  Returns taps; fills n.  Pointers stay owned by ctx. */
int tvo_table(void *p, int which, int *n, const int16_t **f, const int32_t **pos)
{
    o_ctx *c = (o_ctx *)p;
    o_filter *t = which == 0 ? &c->hl : which == 1 ? &c->hc : which == 2 ? &c->vl : &c->vc;
    *n = t->n; *f = t->f; *pos = t->pos;
    return t->taps;
}

static const uint8_t D128[8][8] = {
    {  36, 68,  60, 92,  34, 66,  58, 90 }, { 100,  4, 124, 28,  98,  2, 122, 26 },
    {  52, 84,  44, 76,  50, 82,  42, 74 }, { 116, 20, 108, 12, 114, 18, 106, 10 },
    {  32, 64,  56, 88,  38, 70,  62, 94 }, {  96,  0, 120, 24, 102,  6, 126, 30 },
    {  48, 80,  40, 72,  54, 86,  46, 78 }, { 112, 16, 104,  8, 118, 22, 110, 14 },
};

/* Read sample x of row r of logical plane `pl` (0 Y, 1 U, 2 V) from the source frame. */
static inline int src_sample(const o_ctx *c, const uint8_t *const src[], const int stride[], int pl, int r, int x)
{
    switch (c->sfmt) {
    case FMT_RGB24: case FMT_BGR24: case FMT_RGB48LE: case FMT_RGB48BE: case FMT_BGR48LE: case FMT_BGR48BE: return c->rgb[pl][(size_t)r * (pl ? c->csw : c->sw) + x];
    case FMT_YUV420P: return src[pl][(size_t)r * stride[pl] + x];
    case FMT_NV12:
        return pl == 0 ? src[0][(size_t)r * stride[0] + x] : src[1][(size_t)r * stride[1] + 2 * x + (pl - 1)];
    case FMT_YUV420P10: { uint16_t v; memcpy(&v, src[pl] + (size_t)r * stride[pl] + 2 * x, 2); return v; }
    default: { /* P010: 10 significant bits in the top of each 16-bit word */
        uint16_t v;
        if (pl == 0) memcpy(&v, src[0] + (size_t)r * stride[0] + 2 * x, 2);
        else memcpy(&v, src[1] + (size_t)r * stride[1] + 4 * x + 2 * (pl - 1), 2);
        return v >> 6; }
    }
}

static inline void dst_store(const o_ctx *c, uint8_t *const dst[], const int stride[], int pl, int r, int x, int v)
{
    switch (c->dfmt) {
    case FMT_YUV420P: dst[pl][(size_t)r * stride[pl] + x] = (uint8_t)v; break;
    case FMT_NV12:
        if (pl == 0) dst[0][(size_t)r * stride[0] + x] = (uint8_t)v;
        else dst[1][(size_t)r * stride[1] + 2 * x + (pl - 1)] = (uint8_t)v;
        break;
    case FMT_YUV420P10: { uint16_t w = (uint16_t)v; memcpy(dst[pl] + (size_t)r * stride[pl] + 2 * x, &w, 2); break; }
    default: {
        uint16_t w = (uint16_t)(v << 6);
        if (pl == 0) memcpy(dst[0] + (size_t)r * stride[0] + 2 * x, &w, 2);
        else memcpy(dst[1] + (size_t)r * stride[1] + 4 * x + 2 * (pl - 1), &w, 2);
        break; }
    }
}

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }

/* SURVEY.md A.3 for one logical plane. */
static int scale_plane(const o_ctx *c, const uint8_t *const src[], const int sstride[],
                       uint8_t *const dst[], const int dstride[], int pl)
{
    const o_filter *hf = pl ? &c->hc : &c->hl;
    const o_filter *vf = pl ? &c->vc : &c->vl;
    const int sw = pl ? c->csw : c->sw, sh = pl ? c->csh : c->sh;
    const int dw = pl ? c->cdw : c->dw, dh = pl ? c->cdh : c->dh;
    const int sbits = fmt_bits(c->sfmt), dbits = fmt_bits(c->dfmt);
    /* RGB input reaches the scaler as 14/15-bit lines (srcBpc = 16) and hScale16To15 then shifts by 13 */
    /* ... and 16-bit RGB as 16-bit lines, shifted by 15 (hScale16To15_c: depth - 1) */
    const int hshift = fmt_rgb48(c->sfmt) ? 15 : fmt_rgb(c->sfmt) ? 13 : sbits == 8 ? 7 : sbits - 1;
    int16_t *mid = (int16_t *)malloc((size_t)sh * dw * sizeof(int16_t));
    if (!mid) return -1;

    for (int r = 0; r < sh; r++) {
        for (int x = 0; x < dw; x++) {
            int32_t acc = 0;
            const int16_t *f = hf->f + (size_t)x * hf->taps;
            for (int j = 0; j < hf->taps; j++) {
                int sx = hf->pos[x] + j;
                if (sx >= sw) continue; /* only when taps > sw; coefficient is zero there */
                acc += src_sample(c, src, sstride, pl, r, sx) * f[j];
            }
            acc >>= hshift;
            if (acc > 32767) acc = 32767;
            if (c->range_convert) {
                const int64_t v = ((int64_t)acc * c->rc_coeff[pl ? 1 : 0] + c->rc_offset[pl ? 1 : 0]) >> 14;
                acc = (int32_t)(v > 32767 ? 32767 : v);
            }
            mid[(size_t)r * dw + x] = (int16_t)acc;
        }
    }
    for (int y = 0; y < dh; y++) {
        const int16_t *g = vf->f + (size_t)y * vf->taps;
        for (int x = 0; x < dw; x++) {
            int32_t acc = 0;
            for (int j = 0; j < vf->taps; j++) {
                int sr = vf->pos[y] + j;
                if (sr >= sh) continue;
                acc += mid[(size_t)sr * dw + x] * g[j];
            }
            int v;
            if (dbits == 8) {
                int d = sbits == 8 ? 64 : D128[y & 7][(x + (pl == 2 ? 3 : 0)) & 7];
                v = clampi((acc + (d << 12)) >> 19, 0, 255);
            } else {
                v = clampi((acc + (1 << 16)) >> 17, 0, 1023);
            }
            dst_store(c, dst, dstride, pl, y, x, v);
        }
    }
    free(mid);
    (void)sw;
    return 0;
}

/* the same-size 10 -> 8 bit planar copy uses its own 2-bit dither. */
static void copy_plane_10to8(const o_ctx *c, const uint8_t *const src[], const int sstride[],
                             uint8_t *const dst[], const int dstride[], int pl)
{
    static const uint8_t d2[2][8] = { { 1, 2, 1, 2, 1, 2, 1, 2 }, { 3, 0, 3, 0, 3, 0, 3, 0 } };
    const int w = pl ? c->cdw : c->dw, h = pl ? c->cdh : c->dh;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            int t = (src_sample(c, src, sstride, pl, y, x) + d2[y & 1][x & 7]) >> 2;
            dst_store(c, dst, dstride, pl, y, x, t - (t >> 8));
        }
}

static void copy_plane(const o_ctx *c, const uint8_t *const src[], const int sstride[],
                       uint8_t *const dst[], const int dstride[], int pl)
{
    const int w = pl ? c->cdw : c->dw, h = pl ? c->cdh : c->dh;
    if (c->sfmt == FMT_P010 && c->dfmt == FMT_P010) {
        /* Same format: the library copies each STORED plane word by word - but the first word of every group of four, and
         * the row's last (words % 4) words, lose their six padding bits (observed on libswscale 9.1.100, SIMD and plain C
         * alike, once the fuzzer filled P010's padding with noise; conforming frames, padding zero, are copied unchanged).
         * The interleaved UV plane is one plane of 2 * cw words; it is written when pl == 1, nothing is left for pl == 2. */
        if (pl == 2) return;
        const int words = pl ? 2 * w : w;
        for (int y = 0; y < h; y++) {
            const uint8_t *sr = src[pl] + (size_t)y * sstride[pl];
            uint8_t *dr = dst[pl] + (size_t)y * dstride[pl];
            for (int i = 0; i < words; i++) {
                uint16_t v;
                memcpy(&v, sr + 2 * i, 2);
                if ((i & 3) == 0 || i >= (words & ~3)) v &= 0xffc0;
                memcpy(dr + 2 * i, &v, 2);
            }
        }
        return;
    }
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++)
            dst_store(c, dst, dstride, pl, y, x, src_sample(c, src, sstride, pl, y, x));
}

/*
 * p010le -> nv12 at equal size: libswscale copies each STORED plane (the interleaved UV plane
 * as one plane of 2*cw samples) with the 2-bit dither above.  Its 8-samples-at-a-time main loop
 * drops the 6 padding bits of the P010 word first; its scalar tail loop (the last cols % 8
 * samples of each row) does not, so those samples come out as the low byte of the un-shifted
 * arithmetic.  Observed on libswscale 9.1.100, SIMD and plain C alike; reproduced for parity.
 */
static void copy_stored_p010_to_nv12(const o_ctx *c, const uint8_t *const src[], const int sstride[],
                                     uint8_t *const dst[], const int dstride[], int stored)
{
    static const uint8_t d2[2][8] = { { 1, 2, 1, 2, 1, 2, 1, 2 }, { 3, 0, 3, 0, 3, 0, 3, 0 } };
    const int cols = stored ? 2 * c->cdw : c->dw, rows = stored ? c->cdh : c->dh;
    const int main_cols = cols & ~7;
    for (int y = 0; y < rows; y++)
        for (int x = 0; x < cols; x++) {
            uint16_t raw;
            memcpy(&raw, src[stored] + (size_t)y * sstride[stored] + 2 * x, 2);
            unsigned v = x < main_cols ? (unsigned)(raw >> 6) : raw;
            unsigned t = (v + d2[y & 1][x & 7]) >> 2;
            dst[stored][(size_t)y * dstride[stored] + x] = (uint8_t)(t - (t >> 8));
        }
}

/*
 * How each logical plane is produced when source and destination sizes are equal
 * (libswscale's "unscaled" special converters; SURVEY.md A.4, extended by probing the genuine
 * library over all 16 format pairs).  0 = generic scaler (A.3 with 1-tap tables; this also
 * reproduces the 8->10 shift and every mixed-layout 10-bit copy exactly), 1 = plain copy,
 * 2 = 10->8 two-bit dither copy, 4 = handled per stored plane by copy_stored_p010_to_nv12.
 */
static int unscaled_rule(const o_ctx *c, int pl)
{
    const int sb = fmt_bits(c->sfmt), db = fmt_bits(c->dfmt);
    const int ssemi = fmt_semi(c->sfmt), dsemi = fmt_semi(c->dfmt);
    if (sb == db) return 1;
    if (sb == 8) return 0;
    if (!ssemi && !dsemi) return 2;            /* yuv420p10le -> yuv420p */
    if (!ssemi && dsemi) return pl == 0 ? 2 : 0; /* yuv420p10le -> nv12: chroma via the scaler */
    if (ssemi && dsemi) return 4;              /* p010le -> nv12 */
    return 0;                                  /* p010le -> yuv420p */
}

/* libswscale's input stage for packed RGB (input.c: rgb24ToY_c, rgb24ToUV_c, rgb24ToUV_half_c) with the default
 * ITU-R BT.601 limited-range table; RGB2YUV_SHIFT = 15.  Restated and checked against the genuine library on 300
 * random even-width cases (tests/golden/make_golden.py --rgb). */
enum { RY = 8414, GY = 16519, BY = 3208, RU = -4865, GU = -9528, BU = 14392, RV = 14392, GV = -12061, BV = -2332 };

static int rgb_input_stage(o_ctx *c, const uint8_t *rgb, int stride)
{
    for (int pl = 0; pl < 3; pl++) {
        c->rgb[pl] = (uint16_t *)malloc((size_t)c->sh * (pl ? c->csw : c->sw) * sizeof(uint16_t));
        if (!c->rgb[pl]) return -1;
    }
    if (fmt_rgb48(c->sfmt)) {
        /* rgb48ToY_c / rgb48ToUV_c / rgb48ToUV_half_c: 16-bit components (either byte order), RGB2YUV_SHIFT = 15, the result
         * a full 16-bit line; the half-width chroma averages two pixels with rounding first */
        const int be = c->sfmt == FMT_RGB48BE || c->sfmt == FMT_BGR48BE;
        const int ro48 = (c->sfmt == FMT_BGR48LE || c->sfmt == FMT_BGR48BE) ? 2 : 0, bo48 = 2 - ro48;   /* bgr48: R and B swapped in memory */
        for (int r = 0; r < c->sh; r++) {
            const uint8_t *row = rgb + (size_t)r * stride;
#define PX16(i) (be ? (unsigned)(row[2 * (i)] << 8 | row[2 * (i) + 1]) : (unsigned)(row[2 * (i) + 1] << 8 | row[2 * (i)]))
            for (int x = 0; x < c->sw; x++) {
                const unsigned R = PX16(3 * x + ro48), G = PX16(3 * x + 1), B = PX16(3 * x + bo48);
                c->rgb[0][(size_t)r * c->sw + x] = (uint16_t)(((int)(RY * R + GY * G + BY * B) + (0x2001 << 14)) >> 15);
                if (!c->rgb_half) {
                    c->rgb[1][(size_t)r * c->csw + x] = (uint16_t)(((int)(RU * (int)R + GU * (int)G + BU * (int)B) + (0x10001 << 14)) >> 15);
                    c->rgb[2][(size_t)r * c->csw + x] = (uint16_t)(((int)(RV * (int)R + GV * (int)G + BV * (int)B) + (0x10001 << 14)) >> 15);
                }
            }
            if (c->rgb_half)
                for (int x = 0; x < c->csw; x++) {
                    const int R = (int)((PX16(6 * x + ro48) + PX16(6 * x + 3 + ro48) + 1) >> 1), G = (int)((PX16(6 * x + 1) + PX16(6 * x + 4) + 1) >> 1),
                              B = (int)((PX16(6 * x + bo48) + PX16(6 * x + 3 + bo48) + 1) >> 1);
                    c->rgb[1][(size_t)r * c->csw + x] = (uint16_t)((RU * R + GU * G + BU * B + (0x10001 << 14)) >> 15);
                    c->rgb[2][(size_t)r * c->csw + x] = (uint16_t)((RV * R + GV * G + BV * B + (0x10001 << 14)) >> 15);
                }
#undef PX16
        }
        return 0;
    }
    const int ro = c->sfmt == FMT_BGR24 ? 2 : 0, bo = 2 - ro;   /* bgr24ToY_c / bgr24ToUV*_c: the same sums with R and B swapped in memory */
    for (int r = 0; r < c->sh; r++) {
        const uint8_t *row = rgb + (size_t)r * stride;
        for (int x = 0; x < c->sw; x++) {
            const int R = row[3 * x + ro], G = row[3 * x + 1], B = row[3 * x + bo];
            c->rgb[0][(size_t)r * c->sw + x] = (uint16_t)((RY * R + GY * G + BY * B + (32 << 14) + (1 << 8)) >> 9);
            if (!c->rgb_half) {
                c->rgb[1][(size_t)r * c->csw + x] = (uint16_t)((RU * R + GU * G + BU * B + (256 << 14) + (1 << 8)) >> 9);
                c->rgb[2][(size_t)r * c->csw + x] = (uint16_t)((RV * R + GV * G + BV * B + (256 << 14) + (1 << 8)) >> 9);
            }
        }
        if (c->rgb_half)
            for (int x = 0; x < c->csw; x++) {
                const int R = row[6 * x + ro] + row[6 * x + 3 + ro], G = row[6 * x + 1] + row[6 * x + 4], B = row[6 * x + bo] + row[6 * x + 3 + bo];
                c->rgb[1][(size_t)r * c->csw + x] = (uint16_t)((RU * R + GU * G + BU * B + (256 << 15) + (1 << 9)) >> 10);
                c->rgb[2][(size_t)r * c->csw + x] = (uint16_t)((RV * R + GV * G + BV * B + (256 << 15) + (1 << 9)) >> 10);
            }
    }
    return 0;
}

int tvo_scale(void *p, const uint8_t *const src[], const int sstride[], uint8_t *const dst[], const int dstride[])
{
    o_ctx *c = (o_ctx *)p;
    if (c->cascade[0]) {
        uint8_t *const tmp[3] = { c->tmp[0], c->tmp[1], c->tmp[2] };
        if (tvo_scale(c->cascade[0], src, sstride, tmp, c->tmp_stride) < 0) return -1;
        const uint8_t *const tsrc[3] = { c->tmp[0], c->tmp[1], c->tmp[2] };
        return tvo_scale(c->cascade[1], tsrc, c->tmp_stride, dst, dstride) < 0 ? -1 : c->dh;
    }
    if (fmt_rgb(c->sfmt)) {
        /* even widths only: with an odd width the library's half-width chroma reads one pixel past the row */
        if ((c->sw & 1) || fmt_bits(c->dfmt) < 8 || fmt_rgb(c->dfmt)) return -1;
        int rc = c->dh;
        if (rgb_input_stage(c, src[0], sstride[0])) rc = -1;
        for (int pl = 0; pl < 3 && rc >= 0; pl++)
            if (scale_plane(c, src, sstride, dst, dstride, pl)) rc = -1;
        for (int pl = 0; pl < 3; pl++) { free(c->rgb[pl]); c->rgb[pl] = NULL; }
        return rc;
    }
    const int same = c->sw == c->dw && c->sh == c->dh;
    /* planar -> p010le at equal size: libswscale's converter writes floor(w/2) chroma pairs, so
     * with an odd width the last pair of every chroma row is left untouched. */
    const int cdw_full = c->cdw;
    if (same && !c->range_convert && c->dfmt == FMT_P010 && !fmt_semi(c->sfmt)) c->cdw = c->dw >> 1;
    int rc = c->dh;
    if (same && !c->range_convert && unscaled_rule(c, 0) == 4) {
        copy_stored_p010_to_nv12(c, src, sstride, dst, dstride, 0);
        copy_stored_p010_to_nv12(c, src, sstride, dst, dstride, 1);
    } else {
        for (int pl = 0; pl < 3; pl++) {
            int rule = same && !c->range_convert ? unscaled_rule(c, pl) : 0;   /* a range change always runs the generic scaler */
            if (rule == 1) copy_plane(c, src, sstride, dst, dstride, pl);
            else if (rule == 2) copy_plane_10to8(c, src, sstride, dst, dstride, pl);
            else if (scale_plane(c, src, sstride, dst, dstride, pl)) rc = -1;
        }
    }
    c->cdw = cdw_full;
    return rc;
}
