// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Thin extern "C" shim around the UNMODIFIED reference sources, which are compiled where
// they lie under /root/reference by oracle/Makefile into oracle/_ref/libgblur_ref.so.
// Nothing here restates the algorithm: every entry point calls straight into the
// reference's own functions:
//   gaussianblur::gaussianblur            /root/reference/src/gaussianblur.cpp:307
//   gaussianblur::prepare_kernel_DFT      /root/reference/src/gaussianblur.cpp:100
//   gaussianblur::process_channel_tiles   /root/reference/src/gaussianblur.cpp:202  (externally
//       linkable but un-headered; declared below exactly as defined there)
//   deinterleave_channels<C>              /root/reference/include/gaussianblur/helpers.hpp:218
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load the resulting library.
#include <gaussianblur/gaussianblur.h>
#include <gaussianblur/helpers.hpp>

#include <chrono>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

extern "C" {
#include <pffft_pommier/pffft.h>
}

namespace gaussianblur {
// src/gaussianblur.cpp:202-208 (definition has external linkage, no header declaration)
void process_channel_tiles(const int channel, const int tiles, const int tile_size, const int pad,
                           const int trailing_zeros, PFFFT_Setup *setup,
                           const AlignedVector<float> &kernel, AlignedVector<float> &tmp,
                           AlignedVector<float> &tile, AlignedVector<float> &work,
                           AlignedVector<float> &resf, DeinterleavedChs &deinterleaved_channels,
                           float scaler);
}  // namespace gaussianblur

extern "C" {

// Threads the reference fans out to (helpers.hpp:143).
int ref_hw_threads(void) { return (int)std::thread::hardware_concurrency(); }

// The reference's public entry point on a caller-owned HWC u8 buffer, in place.
// Returns wall seconds spent inside gaussianblur::gaussianblur only.
double ref_gaussianblur(uint8_t *hwc, int rows, int cols, int channels, float sigma,
                        int apply_to_alpha) {
  Image img;
  img.geom = ImgGeom{rows, cols, channels};
  const size_t n = (size_t)rows * cols * channels;
  img.data.assign(hwc, hwc + n);
  auto t0 = std::chrono::steady_clock::now();
  gaussianblur::gaussianblur(img, sigma, apply_to_alpha != 0);
  auto t1 = std::chrono::steady_clock::now();
  std::memcpy(hwc, img.data.data(), n);
  return std::chrono::duration<double>(t1 - t0).count();
}

// prepare_kernel_DFT facts. kerf_* may be NULL (size query). Naming follows the reference:
// kerf_col has n_rowpass floats (transform along the column index, i.e. of a row),
// kerf_row has n_colpass floats.
int ref_kernel_dft(int rows, int cols, float sigma, int *pad, int *n_rowpass, int *n_colpass,
                   int *tz_rows, int *tz_cols, float *kerf_col, float *kerf_row) {
  KernelDFT k = gaussianblur::prepare_kernel_DFT(ImgGeom{rows, cols, 3}, sigma);
  if (pad) *pad = k.pad;
  if (n_rowpass) *n_rowpass = (int)k.kerf_1D_col.size();
  if (n_colpass) *n_colpass = (int)k.kerf_1D_row.size();
  if (tz_rows) *tz_rows = k.trailing_zeros.rows;
  if (tz_cols) *tz_cols = k.trailing_zeros.cols;
  if (kerf_col) std::memcpy(kerf_col, k.kerf_1D_col.data(), k.kerf_1D_col.size() * sizeof(float));
  if (kerf_row) std::memcpy(kerf_row, k.kerf_1D_row.data(), k.kerf_1D_row.size() * sizeof(float));
  return 0;
}

// One float plane (rows x cols, row-major) through the reference's two process_channel_tiles
// calls (row pass, column pass) exactly as gaussianblur::pffft drives them
// (src/gaussianblur.cpp:261-280). If inter_T is non-NULL it receives the plane after pass 1,
// which the reference stores transposed (cols x rows). plane is blurred in place (fp32).
// This is the oracle route for 1-channel images and for >2^31-pixel bands.
double ref_blur_plane_f32(float *plane, int rows, int cols, float sigma, float *inter_T) {
  KernelDFT k = gaussianblur::prepare_kernel_DFT(ImgGeom{rows, cols, 1}, sigma);
  const size_t n = (size_t)rows * cols;
  DeinterleavedChs chs(1, std::vector<float>(plane, plane + n));
  AlignedVector<float> tmp, tile, work;
  const size_t maxsize = std::max(k.kerf_1D_row.size(), k.kerf_1D_col.size());
  tmp.reserve(maxsize);
  tile.reserve(maxsize);
  work.reserve(maxsize);
  AlignedVector<float> resf(n);
  auto t0 = std::chrono::steady_clock::now();
  gaussianblur::process_channel_tiles(0, rows, cols, k.pad, k.trailing_zeros.cols,
                                      k.cols_setup.get(), k.kerf_1D_col, tmp, tile, work, resf,
                                      chs, 1.0F / k.kerf_1D_col.size());
  if (inter_T) std::memcpy(inter_T, chs[0].data(), n * sizeof(float));
  gaussianblur::process_channel_tiles(0, cols, rows, k.pad, k.trailing_zeros.rows,
                                      k.rows_setup.get(), k.kerf_1D_row, tmp, tile, work, resf,
                                      chs, 1.0F / k.kerf_1D_row.size());
  auto t1 = std::chrono::steady_clock::now();
  std::memcpy(plane, chs[0].data(), n * sizeof(float));
  return std::chrono::duration<double>(t1 - t0).count();
}

// u8 single plane: u8 -> f32, ref_blur_plane_f32, uint8_t(v + 0.5f) as interleave_channels
// does (helpers.hpp:251,270).
double ref_blur_plane_u8(uint8_t *plane, int rows, int cols, float sigma) {
  const size_t n = (size_t)rows * cols;
  std::vector<float> f(n);
  for (size_t i = 0; i < n; ++i) f[i] = plane[i];
  double t = ref_blur_plane_f32(f.data(), rows, cols, sigma, nullptr);
  for (size_t i = 0; i < n; ++i) plane[i] = (uint8_t)(f[i] + 0.5F);
  return t;
}

// Row-pass intermediate of an interleaved image for channel `ch`, returned in image
// orientation (rows x cols, row-major), i.e. the reference's transposed pass-1 plane
// transposed back. Test hook for the fp32 intermediate comparison.
int ref_rowpass_plane(const uint8_t *hwc, int rows, int cols, int channels, int ch, float sigma,
                      float *out_rows_cols) {
  const size_t n = (size_t)rows * cols;
  std::vector<float> plane(n), interT(n);
  for (size_t i = 0; i < n; ++i) plane[i] = hwc[i * channels + ch];
  ref_blur_plane_f32(plane.data(), rows, cols, sigma, interT.data());
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < cols; ++c) out_rows_cols[(size_t)r * cols + c] = interT[(size_t)c * rows + r];
  return 0;
}

// A strip of ncols pixel columns [c0, c0 + ncols) of the blurred plane of a (rows x cols) single-channel image that
// may be larger than the reference's own envelope (int rows*cols, src/gaussianblur.cpp:177: BASELINE.json config 5 is
// 32768 x 65536). Every row and every column is an independent line in the reference (src/gaussianblur.cpp:202-244), so
// the strip of a hypothetical monolithic run is obtained exactly by
//   pass 1 in row bands of band_rows rows through process_channel_tiles as gaussianblur::pffft drives it (:268-272),
//          keeping only the strip's columns of each band's (transposed) result, then
//   pass 2 on those ncols complete columns (:275-279),
// with the kernel spectra of the FULL geometry (prepare_kernel_DFT never multiplies rows by cols).
// plane: rows x cols u8, row-major. out: rows x ncols u8. Returns wall seconds.
double ref_band_strip_u8(const uint8_t *plane, int rows, int cols, float sigma, int c0, int ncols, int band_rows,
                         uint8_t *out) {
  KernelDFT k = gaussianblur::prepare_kernel_DFT(ImgGeom{rows, cols, 1}, sigma);
  AlignedVector<float> tmp, tile, work;
  const size_t maxsize = std::max(k.kerf_1D_row.size(), k.kerf_1D_col.size());
  tmp.reserve(maxsize);
  tile.reserve(maxsize);
  work.reserve(maxsize);
  auto t0 = std::chrono::steady_clock::now();
  std::vector<float> strip_T((size_t)ncols * rows);  // [column of the strip][row]: the orientation pass 2 reads
  for (int r0 = 0; r0 < rows; r0 += band_rows) {
    const int br = std::min(band_rows, rows - r0);
    const size_t n = (size_t)br * cols;
    DeinterleavedChs chs(1, std::vector<float>(n));
    const uint8_t *src = plane + (size_t)r0 * cols;
    for (size_t i = 0; i < n; ++i) chs[0][i] = src[i];
    AlignedVector<float> resf(n);
    gaussianblur::process_channel_tiles(0, br, cols, k.pad, k.trailing_zeros.cols, k.cols_setup.get(), k.kerf_1D_col, tmp,
                                        tile, work, resf, chs, 1.0F / k.kerf_1D_col.size());
    // chs[0] is now cols x br (transposed)
    for (int c = 0; c < ncols; ++c)
      std::memcpy(&strip_T[(size_t)c * rows + r0], &chs[0][(size_t)(c0 + c) * br], (size_t)br * sizeof(float));
  }
  {
    const size_t n = (size_t)ncols * rows;
    DeinterleavedChs chs(1, std::move(strip_T));
    AlignedVector<float> resf(n);
    gaussianblur::process_channel_tiles(0, ncols, rows, k.pad, k.trailing_zeros.rows, k.rows_setup.get(), k.kerf_1D_row, tmp,
                                        tile, work, resf, chs, 1.0F / k.kerf_1D_row.size());
    // chs[0] is rows x ncols again
    for (size_t i = 0; i < n; ++i) out[i] = (uint8_t)(chs[0][i] + 0.5F);
  }
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
