// bwdif de-interlacer (ffmpeg's `bwdif=mode=send_frame:parity=auto:deint=all`, the other per-pixel stage of the reference's
// DVD path: worker/tasks.py:62-63,446-447), as a CUDA stage ahead of the scaler.
//
// PARITY UNPINNED: libavfilter is neither under /root/reference nor in this image, so this kernel is checked against
// oracle/bwdif_port.py - a restatement of the published algorithm written alongside it - and against properties
// (kept field copied, static scenes unchanged), not against ffmpeg.  See that file for the algorithm.
//
// A launch covers one plane of a batch of frames.  Per interpolated sample the work is a few dozen integer operations
// and up to 18 cached loads: `bwdif_quad_kernel` gives a thread four adjacent samples so that each of those rows is one
// vector load (2.5-3x the one-sample-per-thread kernel, which stays for unaligned planes and widths not divisible by 4).
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

// Four adjacent samples per thread (planes whose width is a multiple of 4 and whose rows and base are aligned for the
// vector loads; everything else takes the kernel above).  Same arithmetic sample by sample; the 18 rows a quad of
// interpolated samples may touch are each read with one 32- or 64-bit load instead of four scalar ones, and kept rows are
// copied a quad at a time.  Rows (y, quad) are flattened so that narrow planes (a DVD's 360-sample chroma) fill their CTAs.
template <typename T> struct Quad4 { int v[4]; };
template <typename T>
__device__ __forceinline__ Quad4<T> load4(const uint8_t *plane, ptrdiff_t pitch, int y, int x) {
    Quad4<T> q;
    const uint8_t *p = plane + static_cast<ptrdiff_t>(y) * pitch + static_cast<ptrdiff_t>(x) * sizeof(T);
    if (sizeof(T) == 1) {
        const unsigned w = __ldg(reinterpret_cast<const unsigned *>(p));
        q.v[0] = w & 0xff; q.v[1] = (w >> 8) & 0xff; q.v[2] = (w >> 16) & 0xff; q.v[3] = w >> 24;
    } else {
        const uint2 w = __ldg(reinterpret_cast<const uint2 *>(p));
        q.v[0] = w.x & 0xffff; q.v[1] = w.x >> 16; q.v[2] = w.y & 0xffff; q.v[3] = w.y >> 16;
    }
    return q;
}
template <typename T>
__device__ __forceinline__ void store4(uint8_t *p, const int (&v)[4]) {
    if (sizeof(T) == 1) *reinterpret_cast<unsigned *>(p) = static_cast<unsigned>(v[0]) | (static_cast<unsigned>(v[1]) << 8) | (static_cast<unsigned>(v[2]) << 16) | (static_cast<unsigned>(v[3]) << 24);
    else *reinterpret_cast<uint2 *>(p) = make_uint2(static_cast<unsigned>(v[0]) | (static_cast<unsigned>(v[1]) << 16), static_cast<unsigned>(v[2]) | (static_cast<unsigned>(v[3]) << 16));
}

template <typename T>
__global__ void __launch_bounds__(256) bwdif_quad_kernel(const BwdifParams P) {
    // a thread takes a quad of one row PAIR: it copies the kept row's quad and interpolates the other row's, so that no
    // warp mixes the two kinds of rows
    const int quads = P.w >> 2;
    const int idx = blockIdx.x * 256 + threadIdx.x;
    const int pairs = (P.h + 1) >> 1;
    if (idx >= quads * pairs) return;
    const int pr = idx / quads, x = (idx - pr * quads) << 2;
    const int i = P.first_out + blockIdx.y;   // frame index inside the batch
    const uint8_t *cur = P.src + static_cast<ptrdiff_t>(i) * P.src_fstride;
    const uint8_t *prev = i > 0 ? cur - P.src_fstride : cur;
    const uint8_t *next = i + 1 < P.n_frames ? cur + P.src_fstride : cur;
    const ptrdiff_t sp = P.src_pitch;
    const int h = P.h, df = P.df;
    uint8_t *const oplane = P.dst + static_cast<ptrdiff_t>(i) * P.dst_fstride + static_cast<ptrdiff_t>(x) * sizeof(T);
    const int y = 2 * pr + (P.parity ? 0 : 1), yk = y ^ 1;   // rows with ((y ^ parity) & 1) are interpolated
    if (yk < h) store4<T>(oplane + static_cast<ptrdiff_t>(yk) * P.dst_pitch, load4<T>(cur, sp, yk, x).v);   // the kept field
    if (y >= h) return;
    uint8_t *out = oplane + static_cast<ptrdiff_t>(y) * P.dst_pitch;

    const bool intra = (i == 0 && P.intra_first) || (i == P.n_frames - 1 && P.intra_last);
    int p1 = (y + df) < h ? y + 1 : y - 1, m1 = y > (df - 1) ? y - 1 : y + 1;
    int res[4];
    if (intra) {
        int p3 = (y + 3 * df) < h ? y + 3 : y - 3, m3 = y > (3 * df - 1) ? y - 3 : y + 3;
        p1 = min(max(p1, 0), h - 1); m1 = min(max(m1, 0), h - 1); p3 = min(max(p3, 0), h - 1); m3 = min(max(m3, 0), h - 1);
        const Quad4<T> a = load4<T>(cur, sp, m1, x), b = load4<T>(cur, sp, p1, x), c3 = load4<T>(cur, sp, m3, x), d3 = load4<T>(cur, sp, p3, x);
#pragma unroll
        for (int k = 0; k < 4; ++k) res[k] = min(max((5077 * (a.v[k] + b.v[k]) - 981 * (c3.v[k] + d3.v[k])) >> 13, 0), P.clip_max);
        store4<T>(out, res);
        return;
    }
    const bool edge = (y < 4) || ((y + 5) > h);
    if (!edge) { p1 = y + 1; m1 = y - 1; }
    p1 = min(max(p1, 0), h - 1); m1 = min(max(m1, 0), h - 1);
    const uint8_t *prev2 = prev, *next2 = cur;   // send_frame: the missing field lies between the previous frame and this one
    const Quad4<T> C = load4<T>(cur, sp, m1, x), E = load4<T>(cur, sp, p1, x);
    const Quad4<T> A2 = load4<T>(prev2, sp, y, x), B2 = load4<T>(next2, sp, y, x);
    const Quad4<T> PM = load4<T>(prev, sp, m1, x), PP = load4<T>(prev, sp, p1, x), NM = load4<T>(next, sp, m1, x), NP = load4<T>(next, sp, p1, x);
    const bool spat = !edge || !((y < 2) || ((y + 3) > h));
    Quad4<T> U2a{}, U2b{}, D2a{}, D2b{}, C3{}, E3{}, F4a{}, F4b{}, F4c{}, F4d{};
    if (spat) {
        U2a = load4<T>(prev2, sp, y - 2, x); U2b = load4<T>(next2, sp, y - 2, x);
        D2a = load4<T>(prev2, sp, y + 2, x); D2b = load4<T>(next2, sp, y + 2, x);
    }
    if (!edge) {
        C3 = load4<T>(cur, sp, y - 3, x); E3 = load4<T>(cur, sp, y + 3, x);
        F4a = load4<T>(prev2, sp, y - 4, x); F4b = load4<T>(next2, sp, y - 4, x);
        F4c = load4<T>(prev2, sp, y + 4, x); F4d = load4<T>(next2, sp, y + 4, x);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int c = C.v[k], e = E.v[k], a2 = A2.v[k], b2 = B2.v[k];
        const int d = (a2 + b2) >> 1;
        const int td0 = abs(a2 - b2);
        const int td1 = (abs(PM.v[k] - c) + abs(PP.v[k] - e)) >> 1;
        const int td2 = (abs(NM.v[k] - c) + abs(NP.v[k] - e)) >> 1;
        int diff = max(max(td0 >> 1, td1), td2);
        if (diff == 0) { res[k] = d; continue; }
        const int up2 = U2a.v[k] + U2b.v[k], dn2 = D2a.v[k] + D2b.v[k];
        if (spat) {
            const int b = (up2 >> 1) - c, f = (dn2 >> 1) - e, dc = d - c, de = d - e;
            const int mx = max(max(de, dc), min(b, f)), mn = min(min(de, dc), max(b, f));
            diff = max(max(diff, mn), -mx);
        }
        int interpol;
        if (edge) {
            interpol = (c + e) >> 1;
        } else {
            const int c3e3 = C3.v[k] + E3.v[k];
            if (abs(c - e) > td0) {
                const int far4 = F4a.v[k] + F4b.v[k] + F4c.v[k] + F4d.v[k];
                interpol = (((5570 * (a2 + b2) - 3801 * (up2 + dn2) + 1016 * far4) >> 2) + 4309 * (c + e) - 213 * c3e3) >> 13;
            } else {
                interpol = (5077 * (c + e) - 981 * c3e3) >> 13;
            }
        }
        interpol = interpol > d + diff ? d + diff : interpol < d - diff ? d - diff : interpol;
        res[k] = min(max(interpol, 0), P.clip_max);
    }
    store4<T>(out, res);
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
    if (P.n_out > 65535 || h > 65535) return -1;
    const int sample = depth > 8 ? 2 : 1, vec = 4 * sample;
    const bool quad_ok = (w & 3) == 0 && (src_pitch % vec) == 0 && (dst_pitch % vec) == 0 && (src_fstride % vec) == 0 && (dst_fstride % vec) == 0 &&
                         (reinterpret_cast<uintptr_t>(src) % vec) == 0 && (reinterpret_cast<uintptr_t>(dst) % vec) == 0 &&
                         static_cast<long long>(w >> 2) * h < (1LL << 31) - 256;
    if (quad_ok) {
        const dim3 qgrid(static_cast<unsigned>((static_cast<long long>(w >> 2) * ((h + 1) >> 1) + 255) / 256), P.n_out);
        if (depth > 8) bwdif_quad_kernel<uint16_t><<<qgrid, 256, 0, stream>>>(P);
        else bwdif_quad_kernel<uint8_t><<<qgrid, 256, 0, stream>>>(P);
        return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
    }
    const dim3 grid((w + 127) / 128, h, P.n_out);
    if (depth > 8) bwdif_kernel<uint16_t><<<grid, 128, 0, stream>>>(P);
    else bwdif_kernel<uint8_t><<<grid, 128, 0, stream>>>(P);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace tvs
