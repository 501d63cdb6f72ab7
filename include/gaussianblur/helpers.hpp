// include/gaussianblur/helpers.hpp — public types of the B200-native LibGaussianBlur hot path.
//
// Same names, members and meaning as the reference's types
// (/root/reference/include/gaussianblur/helpers.hpp:97-132) so that code written against
// `ImgGeom`, `Image`, `TrailingZeros` and `KernelDFT` keeps compiling. The reference's CPU
// helpers in that header (hybrid_loop, flip_block, (de)interleave_channels, PFAlloc) have no
// counterpart here: their work is fused into the CUDA kernels behind include/gblur_capi.h.
#pragma once
#include <cstdint>
#include <memory>
#include <vector>

struct gblur_plan;  // opaque GPU plan (include/gblur_capi.h)

// Image geometry: rows x cols pixels, `channels` interleaved bytes per pixel.
typedef struct {
  int rows;
  int cols;
  int channels;
} ImgGeom;

// Row-major interleaved (HWC) 8-bit image; blurred in place, caller owns the buffer.
typedef struct {
  std::vector<uint8_t> data;
  ImgGeom geom;
} Image;

// Zeros appended after the reflect pad so each transform length is 32 * 2^a 3^b 5^c.
typedef struct {
  int rows;
  int cols;
} TrailingZeros;

// In the reference these two members own PFFFT_Setup objects; here both refer to one GPU plan
// (twiddles, permuted spectra, workspace). They may be null until the plan is first needed.
using GblurPlanHandle = std::shared_ptr<gblur_plan>;

// Precomputed kernel spectra + plan. kerf_1D_col is the spectrum used along a row (length
// cols + 2*pad + trailing_zeros.cols), kerf_1D_row the one used along a column — the reference's
// naming. Layout: pffft "ordered" real spectrum [K0, K_{N/2}, Re1, Im1, ...].
typedef struct {
  std::vector<float> kerf_1D_row;
  std::vector<float> kerf_1D_col;
  GblurPlanHandle rows_setup;
  GblurPlanHandle cols_setup;
  int pad;
  TrailingZeros trailing_zeros;
  ImgGeom geom;  // additive: geometry and sigma the plan was made for
  float sigma;
} KernelDFT;
