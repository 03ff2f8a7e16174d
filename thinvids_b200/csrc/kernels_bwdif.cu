// bwdif de-interlacer (ffmpeg's `bwdif=mode=send_frame:parity=auto:deint=all`, the other per-pixel stage of the reference's
// DVD path: worker/tasks.py:62-63,446-447), as a CUDA stage ahead of the scaler.
//
// PARITY UNPINNED: libavfilter is neither under /root/reference nor in this image, so this kernel is checked against
// oracle/bwdif_port.py - a restatement of the published algorithm written alongside it - and against properties
// (kept field copied, static scenes unchanged), not against ffmpeg.  See that file for the algorithm.
//
// One thread per output sample; a launch covers one plane of a batch of frames.  The work is a few dozen integer
// operations and ~17 cached loads per interpolated sample on DVD-sized frames: memory-shaped, nothing to tile.
#include "tvs_internal.h"

namespace tvs {
namespace {

struct BwdifParams {
    const uint8_t *src; ptrdiff_t src_pitch, src_fstride;
    uint8_t *dst; ptrdiff_t dst_pitch, dst_fstride;
    int w, h;
    int n_frames;        // frames in `src`
    int first_out;       // first frame that is output (1 when frame 0 is context only)
    int n_out;
    int intra_first, intra_last;   // frame 0 / frame n_frames - 1 is the start / end of the stream: intra filter
    int parity;          // rows with ((y ^ parity) & 1) are interpolated
    int clip_max, df;
};

template <typename T>
__device__ __forceinline__ int px(const uint8_t *plane, ptrdiff_t pitch, int y, int x) {
    return static_cast<int>(reinterpret_cast<const T *>(plane + static_cast<ptrdiff_t>(y) * pitch)[x]);
}

template <typename T>
__global__ void __launch_bounds__(128) bwdif_kernel(const BwdifParams P) {
    const int x = blockIdx.x * 128 + threadIdx.x, y = blockIdx.y;
    if (x >= P.w) return;
    const int i = P.first_out + blockIdx.z;   // frame index inside the batch
    const uint8_t *cur = P.src + static_cast<ptrdiff_t>(i) * P.src_fstride;
    const uint8_t *prev = i > 0 ? cur - P.src_fstride : cur;
    const uint8_t *next = i + 1 < P.n_frames ? cur + P.src_fstride : cur;
    T *out = reinterpret_cast<T *>(P.dst + static_cast<ptrdiff_t>(i) * P.dst_fstride + static_cast<ptrdiff_t>(y) * P.dst_pitch) + x;
    const ptrdiff_t sp = P.src_pitch;
    const int h = P.h, df = P.df;
    if (!((y ^ P.parity) & 1)) { *out = static_cast<T>(px<T>(cur, sp, y, x)); return; }   // the kept field

    const bool intra = (i == 0 && P.intra_first) || (i == P.n_frames - 1 && P.intra_last);
    int p1 = (y + df) < h ? y + 1 : y - 1, m1 = y > (df - 1) ? y - 1 : y + 1;
    if (intra) {
        int p3 = (y + 3 * df) < h ? y + 3 : y - 3, m3 = y > (3 * df - 1) ? y - 3 : y + 3;
        p1 = min(max(p1, 0), h - 1); m1 = min(max(m1, 0), h - 1); p3 = min(max(p3, 0), h - 1); m3 = min(max(m3, 0), h - 1);
        const int v = (5077 * (px<T>(cur, sp, m1, x) + px<T>(cur, sp, p1, x)) - 981 * (px<T>(cur, sp, m3, x) + px<T>(cur, sp, p3, x))) >> 13;
        *out = static_cast<T>(min(max(v, 0), P.clip_max));
        return;
    }
    const bool edge = (y < 4) || ((y + 5) > h);
    if (!edge) { p1 = y + 1; m1 = y - 1; }
    p1 = min(max(p1, 0), h - 1); m1 = min(max(m1, 0), h - 1);
    const uint8_t *prev2 = prev, *next2 = cur;   // send_frame: the missing field lies between the previous frame and this one
    const int c = px<T>(cur, sp, m1, x), e = px<T>(cur, sp, p1, x);
    const int a2 = px<T>(prev2, sp, y, x), b2 = px<T>(next2, sp, y, x);
    const int d = (a2 + b2) >> 1;
    const int td0 = abs(a2 - b2);
    const int td1 = (abs(px<T>(prev, sp, m1, x) - c) + abs(px<T>(prev, sp, p1, x) - e)) >> 1;
    const int td2 = (abs(px<T>(next, sp, m1, x) - c) + abs(px<T>(next, sp, p1, x) - e)) >> 1;
    int diff = max(max(td0 >> 1, td1), td2);
    if (diff == 0) { *out = static_cast<T>(d); return; }
    const bool spat = !edge || !((y < 2) || ((y + 3) > h));
    int up2 = 0, dn2 = 0;
    if (spat) {
        up2 = px<T>(prev2, sp, y - 2, x) + px<T>(next2, sp, y - 2, x);
        dn2 = px<T>(prev2, sp, y + 2, x) + px<T>(next2, sp, y + 2, x);
        const int b = (up2 >> 1) - c, f = (dn2 >> 1) - e, dc = d - c, de = d - e;
        const int mx = max(max(de, dc), min(b, f)), mn = min(min(de, dc), max(b, f));
        diff = max(max(diff, mn), -mx);
    }
    int interpol;
    if (edge) {
        interpol = (c + e) >> 1;
    } else {
        const int c3e3 = px<T>(cur, sp, y - 3, x) + px<T>(cur, sp, y + 3, x);
        if (abs(c - e) > td0) {
            const int far4 = px<T>(prev2, sp, y - 4, x) + px<T>(next2, sp, y - 4, x) + px<T>(prev2, sp, y + 4, x) + px<T>(next2, sp, y + 4, x);
            interpol = (((5570 * (a2 + b2) - 3801 * (up2 + dn2) + 1016 * far4) >> 2) + 4309 * (c + e) - 213 * c3e3) >> 13;
        } else {
            interpol = (5077 * (c + e) - 981 * c3e3) >> 13;
        }
    }
    interpol = interpol > d + diff ? d + diff : interpol < d - diff ? d - diff : interpol;
    *out = static_cast<T>(min(max(interpol, 0), P.clip_max));
}

}  // namespace

// One plane of a batch.  Returns kernels launched, or -1 after a launch error.
int launch_bwdif_plane(const uint8_t *src, ptrdiff_t src_pitch, ptrdiff_t src_fstride, uint8_t *dst, ptrdiff_t dst_pitch,
                       ptrdiff_t dst_fstride, int w, int h, int depth, int n_frames, int tff, bool stream_start, bool stream_end,
                       cudaStream_t stream) {
    BwdifParams P{};
    P.src = src; P.src_pitch = src_pitch; P.src_fstride = src_fstride;
    P.dst = dst; P.dst_pitch = dst_pitch; P.dst_fstride = dst_fstride;
    P.w = w; P.h = h; P.n_frames = n_frames;
    P.first_out = stream_start ? 0 : 1;
    const int last_out = stream_end ? n_frames - 1 : n_frames - 2;
    P.n_out = last_out - P.first_out + 1;
    if (P.n_out <= 0) return 0;
    P.intra_first = stream_start; P.intra_last = stream_end;
    P.parity = (tff ? 1 : 0) ^ 1;
    P.clip_max = (1 << depth) - 1;
    P.df = (depth + 7) / 8;
    const dim3 grid((w + 127) / 128, h, P.n_out);
    if (P.n_out > 65535 || h > 65535) return -1;
    if (depth > 8) bwdif_kernel<uint16_t><<<grid, 128, 0, stream>>>(P);
    else bwdif_kernel<uint8_t><<<grid, 128, 0, stream>>>(P);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tvs
