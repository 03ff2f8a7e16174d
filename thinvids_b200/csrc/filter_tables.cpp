// See filter_tables.h.  Behavioural specification: SURVEY.md Appendix A.1-A.2 (libswscale initFilter).
// Compile with -ffp-contract=off: the lanczos weights depend on plain IEEE double arithmetic.
#include "filter_tables.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>

namespace tvs {
namespace {

using i64 = long long;

constexpr i64 kOne16 = 1LL << 16;
constexpr double kPi = 3.14159265358979323846;

inline i64 magnitude(i64 v) { return v < 0 ? -v : v; }

inline int floor_log2(unsigned v) {
    int r = 0;
    while (v >>= 1) ++r;
    return r;
}

// Kernel weight at distance `d` (30 fractional bits) from the sample centre, scaled to `unity`.
struct KernelEval {
    Algorithm alg;
    i64 unity;
    int inc;
    double p0, p1;   // libswscale's scaler parameters (kParamDefault = the algorithm's own default)

    i64 operator()(i64 d) const {
        switch (alg) {
        case kBicubic: {
            const i64 b = static_cast<i64>((p0 != kParamDefault ? p0 : 0.0) * (1 << 24));
            const i64 c = static_cast<i64>((p1 != kParamDefault ? p1 : 0.6) * (1 << 24));
            const i64 q30 = 1LL << 30, q24 = 1LL << 24;
            i64 w = 0;
            if (d < (1LL << 31)) {
                const i64 d2 = (d * d) >> 30;
                const i64 d3 = (d2 * d) >> 30;
                if (d < q30)
                    w = (12 * q24 - 9 * b - 6 * c) * d3 + (-18 * q24 + 12 * b + 6 * c) * d2 + (6 * q24 - 2 * b) * q30;
                else
                    w = (-b - 6 * c) * d3 + (6 * b + 30 * c) * d2 + (-12 * b - 48 * c) * d + (8 * b + 24 * c) * q30;
            }
            return w / ((1LL << 54) / unity);
        }
        case kArea: {
            const i64 e = d - (1LL << 29);
            const i64 edge = 1LL << 45;
            i64 w;
            if (e * inc < -edge) w = 1LL << 46;
            else if (e * inc < edge) w = -e * inc + edge;
            else w = 0;
            return w * (unity >> 46);
        }
        case kLanczos: {
            const double lobes = p0 != kParamDefault ? p0 : 3.0;
            const double x = static_cast<double>(d) * (1.0 / (1 << 30));
            if (x > lobes) return 0;
            const double v = d ? std::sin(x * kPi) * std::sin(x * kPi / lobes) / (x * x * kPi * kPi / lobes) : 1.0;
            return static_cast<i64>(v * static_cast<double>(unity));
        }
        default: {  // bilinear
            const i64 w = (1LL << 30) - d;
            return (w < 0 ? 0 : w) * (unity >> 30);
        }
        }
    }
};

int kernel_support(Algorithm alg, double p0) {
    switch (alg) {
    case kBicubic: return 4;
    case kLanczos: return p0 != kParamDefault ? static_cast<int>(std::ceil(2 * p0)) : 6;
    case kArea: return 1;
    default: return 2;
    }
}

}  // namespace

int AxisTable::max_pos_step() const {
    int m = 0;
    for (int i = 1; i < length; ++i) m = std::max(m, pos[i] - pos[i - 1]);
    return m;
}

int axis_increment(int src_len, int dst_len) {
    return static_cast<int>(((static_cast<i64>(src_len) << 16) + (dst_len >> 1)) / dst_len);
}

int axis_position(int siting, int subsample_log2) {
    if (siting == -1 || siting <= -513) siting = (128 << subsample_log2) - 128;
    return (siting + 128) >> subsample_log2;
}

bool build_axis_table(AxisTable &out, int src_len, int dst_len, Algorithm alg, int one, int src_pos, int dst_pos, const double *param) {
    const double p0 = param ? param[0] : kParamDefault, p1 = param ? param[1] : kParamDefault;
    if (src_len <= 0 || dst_len <= 0) return false;
    if (alg != kBilinear && alg != kBicubic && alg != kArea && alg != kLanczos) return false;

    const int inc = axis_increment(src_len, dst_len);
    const i64 unity = 1LL << (54 - std::min(floor_log2(static_cast<unsigned>(src_len / dst_len)), 8));
    const bool enlarging = inc <= kOne16;

    std::vector<i64> raw;          // dst_len x width, before trimming
    std::vector<int32_t> first(dst_len);
    int width;

    if (std::abs(inc - static_cast<int>(kOne16)) < 10 && src_pos == dst_pos) {
        // identity axis
        width = 1;
        raw.assign(dst_len, unity);
        for (int i = 0; i < dst_len; ++i) first[i] = i;
    } else if (alg == kArea && enlarging) {
        // "area" has no meaning when enlarging: libswscale falls back to a 2-tap triangle
        width = 2;
        raw.assign(static_cast<size_t>(dst_len) * 2, 0);
        i64 centre = ((static_cast<i64>(dst_pos) * inc) >> 8) - ((static_cast<i64>(src_pos) * 0x8000) >> 7);
        for (int i = 0; i < dst_len; ++i, centre += inc) {
            const i64 left = (centre - (static_cast<i64>(width - 1) << 15) + (1 << 15)) >> 16;
            first[i] = static_cast<int32_t>(left);
            for (int j = 0; j < 2; ++j) {
                const i64 w = unity - magnitude((left + j) * kOne16 - centre) * (unity >> 16);
                raw[static_cast<size_t>(i) * 2 + j] = std::max<i64>(w, 0);
            }
        }
    } else {
        const int support = kernel_support(alg, p0);
        width = enlarging ? 1 + support
                          : 1 + static_cast<int>((static_cast<i64>(support) * src_len + dst_len - 1) / dst_len);
        width = std::max(std::min(width, src_len - 2), 1);
        raw.assign(static_cast<size_t>(dst_len) * width, 0);
        const KernelEval weight{alg, unity, inc, p0, p1};
        // centre carries 17 fractional bits here
        i64 centre = ((static_cast<i64>(dst_pos) * inc) >> 7) - ((static_cast<i64>(src_pos) * kOne16) >> 7);
        for (int i = 0; i < dst_len; ++i, centre += 2LL * inc) {
            const i64 left = (centre - static_cast<i64>(width - 2) * kOne16) / (2 * kOne16);
            first[i] = static_cast<int32_t>(left);
            for (int j = 0; j < width; ++j) {
                i64 d = magnitude((left + j) * 2 * kOne16 - centre) << 13;
                if (!enlarging) d = d * dst_len / src_len;
                raw[static_cast<size_t>(i) * width + j] = weight(d);
            }
        }
    }

    // Trim: slide each row right past negligible leading weights (keeping starts monotone), and find the
    // widest run that still matters once negligible trailing weights are ignored.  Last row first.
    const double negligible = 0.002 * static_cast<double>(unity);
    int kept = 0;
    for (int i = dst_len - 1; i >= 0; --i) {
        i64 *row = &raw[static_cast<size_t>(i) * width];
        i64 seen = 0;
        for (int n = 0; n < width; ++n) {
            seen += magnitude(row[0]);
            if (static_cast<double>(seen) > negligible) break;
            if (i + 1 < dst_len && first[i] >= first[i + 1]) break;
            std::rotate(row, row + 1, row + width);
            row[width - 1] = 0;
            ++first[i];
        }
        seen = 0;
        int live = width;
        for (int j = width - 1; j > 0; --j) {
            seen += magnitude(row[j]);
            if (static_cast<double>(seen) > negligible) break;
            --live;
        }
        kept = std::max(kept, live);
    }
    if (kept > 256) {   // far past the size at which libswscale cascades two contexts: report the size, build nothing
        out.length = dst_len;
        out.taps = kept;
        out.coef.clear();
        out.pos.clear();
        return false;
    }

    const int taps = kept;
    out.length = dst_len;
    out.taps = taps;
    out.coef.assign(static_cast<size_t>(dst_len) * taps, 0);
    out.pos = first;

    std::vector<i64> row(taps);
    for (int i = 0; i < dst_len; ++i) {
        std::copy_n(&raw[static_cast<size_t>(i) * width], taps, row.begin());
        int32_t &start = out.pos[i];

        // Edge handling lives in the table: weights that would read outside [0, src_len) are folded onto
        // the first / last source sample, so the kernels never clamp coordinates.
        if (start < 0) {
            for (int j = 1; j < taps; ++j) {
                row[std::max(j + start, 0)] += row[j];
                row[j] = 0;
            }
            start = 0;
        }
        if (start + taps > src_len) {
            const int slide = start + std::min(taps - src_len, 0);
            i64 spill = 0;
            for (int j = taps - 1; j >= 0; --j)
                if (start + j >= src_len) { spill += row[j]; row[j] = 0; }
            for (int j = taps - 1; j >= 0; --j) row[j] = j < slide ? 0 : row[j - slide];
            start -= slide;
            row[src_len - 1 - start] += spill;
        }

        // Quantise to `one` with the rounding error carried along the row, so every row sums to `one`.
        i64 total = 0;
        for (i64 w : row) total += w;
        i64 unit = (total + one / 2) / one;
        if (unit == 0) unit = 1;
        i64 carry = 0;
        for (int j = 0; j < taps; ++j) {
            const i64 v = row[j] + carry;
            const i64 q = v >= 0 ? (v + (unit >> 1)) / unit : (v - (unit >> 1)) / unit;
            out.coef[static_cast<size_t>(i) * taps + j] = static_cast<int16_t>(q);
            carry = v - static_cast<int>(q) * unit;
        }
    }
    return true;
}

}  // namespace tvs
