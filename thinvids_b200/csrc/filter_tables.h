// Host-side fixed-point resampling tables (one per axis and per luma/chroma).
//
// These reproduce, to the last LSB, the coefficient tables libswscale's initFilter() builds for
// `scale=W:H:flags=<alg>+accurate_rnd+bitexact` — the filter the reference asks ffmpeg for at
// worker/tasks.py:62-65,444-449 and tools/upscale_benchmark.py:163,182.  Specification: SURVEY.md A.1-A.2.
// Everything here runs on the host (lanczos needs glibc's double sin()); the kernels only read the result.
#pragma once
#include <cstdint>
#include <vector>

namespace tvs {

enum Algorithm : int { kBilinear = 2, kBicubic = 4, kArea = 0x20, kLanczos = 0x200 };

struct AxisTable {
    int length = 0;              // number of output samples along this axis
    int taps = 0;                // coefficients per output sample
    std::vector<int16_t> coef;   // length * taps, row-major
    std::vector<int32_t> pos;    // first source sample of each output sample

    int max_pos_step() const;    // largest pos[i+1]-pos[i]
};

// 16.16 source step per destination sample, as libswscale computes it.
int axis_increment(int src_len, int dst_len);

// Sub-sample position passed to the table builder: `siting` in 1/256 luma-pixel units,
// -513 (or -1) meaning "centred"; `subsample_log2` is 0 for luma and 1 for 4:2:0 chroma.
int axis_position(int siting, int subsample_log2);

// Build one table.  `one` is 1<<14 for horizontal and 1<<12 for vertical tables.
// Returns false for arguments the scaler cannot handle (non-positive sizes, unknown algorithm,
// or a kernel wider than 256 taps: `out.taps` then holds the size and no coefficients - libswscale cascades two
// contexts long before that, and so does tvs_create).
// `param`: libswscale's two scaler parameters (ffmpeg's param0 / param1: bicubic B and C, lanczos lobes), or nullptr;
// a value of kParamDefault (SWS_PARAM_DEFAULT) means the algorithm's own default (B 0, C 0.6; 3 lobes).
constexpr double kParamDefault = 123456.0;
bool build_axis_table(AxisTable &out, int src_len, int dst_len, Algorithm alg, int one, int src_pos, int dst_pos, const double *param = nullptr);

}  // namespace tvs
