// hostsem.hpp — host-side semantics the reference fixes and the GPU path must keep:
// kernel window, FFT-length selection, tap generation and the kernel spectrum.
// Written from the behaviour of (paths under /root/reference/):
//   gaussian_window          src/gaussianblur.cpp:11-26
//   is_valid_size            src/gaussianblur.cpp:82-89
//   nearest_transform_size   src/gaussianblur.cpp:91-98
//   get_gaussian             src/gaussianblur.cpp:28-60
//   prepare_kernel_DFT       src/gaussianblur.cpp:100-172 (sizes / trailing zeros)
// The spectrum itself is evaluated in double from the float taps (the reference runs pffft in
// fp32; the two agree to ~1e-7 absolute) and rounded to fp32.
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

namespace gblur {
namespace host {

constexpr int kNMin = 32;

// radius = sigma*sqrt(2 log 255) - 1 in double, narrowed to float; width = int(radius*2 + 0.5f)
inline int gaussian_window(float sigma, int max_width = 0) {
  const float radius = (float)std::max((double)sigma * std::sqrt(2.0 * std::log(255.0)) - 1.0, 0.0);
  int width = (int)(radius * 2.0f + 0.5f);
  if (max_width) width = std::min(width, max_width);
  if (width % 2 == 0) ++width;
  return width;
}

inline bool is_valid_size(int n) {
  int r = n;
  while (r >= 5 * kNMin && r % 5 == 0) r /= 5;
  while (r >= 3 * kNMin && r % 3 == 0) r /= 3;
  while (r >= 2 * kNMin && r % 2 == 0) r /= 2;
  return r == kNMin;
}

inline int nearest_transform_size(int n) {
  if (n < kNMin) n = kNMin;
  n = kNMin * ((n + kNMin - 1) / kNMin);
  while (!is_valid_size(n)) n += kNMin;
  return n;
}

struct Sizes {
  int ksize, pad;
  int n_rowpass;  // transform length along a row  (reference: sizes[1], kerf_1D_col, cols_setup)
  int n_colpass;  // transform length along a column (reference: sizes[0], kerf_1D_row, rows_setup)
  int tz_rows, tz_cols;
};

inline Sizes plan_sizes(int rows, int cols, float sigma) {
  Sizes s{};
  s.ksize = gaussian_window(sigma, std::max(rows, cols));
  s.pad = (s.ksize - 1) / 2;
  int need[2] = {rows + 2 * s.pad, cols + 2 * s.pad};
  int tz[2] = {0, 0};
  for (int i = 0; i < 2; ++i) {
    if (!is_valid_size(need[i])) {
      const int n = nearest_transform_size(need[i]);
      tz[i] = n - need[i];
      need[i] = n;
    }
  }
  s.n_colpass = need[0];
  s.n_rowpass = need[1];
  s.tz_rows = tz[0];
  s.tz_cols = tz[1];
  return s;
}

// Normalised taps exactly as the reference's float loop produces them: y, y*y, s are float,
// exp() runs in double on the float argument, the quotient by float(pi)*s is narrowed on store,
// the sum is a sequential float accumulate and every tap is scaled by float(1/sum).
inline std::vector<float> gaussian_taps(float sigma, int width) {
  std::vector<float> t((size_t)width, 0.0f);
  const float mid = (width - 1) / 2.0f;
  const float s = 2.0f * sigma * sigma;
  const float pis = 3.14159265358979323846f * s;
  int i = 0;
  for (float y = -mid; y <= mid && i < width; ++y, ++i)
    t[(size_t)i] = (float)(std::exp((double)(-(y * y) / s)) / (double)pis);
  float acc = 0.0f;
  for (float v : t) acc += v;
  const float inv = 1.0f / acc;
  for (float& v : t) v *= inv;
  return t;
}

// Real gains K[k], k = 0..N/2, of the centred (rotated) kernel: K[k] = sum_i t[i] cos(2 pi k (i-mid)/N).
inline std::vector<double> kernel_gains(const std::vector<float>& taps, int N) {
  const int w = (int)taps.size();
  const int half = w / 2;
  std::vector<double> g((size_t)N / 2 + 1);
  const double step = 2.0 * 3.14159265358979323846264338327950288 / (double)N;
  for (int k = 0; k <= N / 2; ++k) {
    double acc = taps[(size_t)half];
    for (int d = 1; d <= half; ++d) {
      // exact argument reduction: (k*d) mod N keeps cos() accurate for large products
      const long long kd = ((long long)k * d) % N;
      acc += ((double)taps[(size_t)(half + d)] + (double)taps[(size_t)(half - d)]) * std::cos(step * (double)kd);
    }
    g[(size_t)k] = acc;
  }
  return g;
}

// one gain K[k] of the same sum (the Nyquist coefficient needs only K[0] and K[N/2])
inline double kernel_gain_at(const std::vector<float>& taps, int N, int k) {
  const int half = (int)taps.size() / 2;
  const double step = 2.0 * 3.14159265358979323846264338327950288 / (double)N;
  double acc = taps[(size_t)half];
  for (int d = 1; d <= half; ++d) {
    const long long kd = ((long long)k * d) % N;
    acc += ((double)taps[(size_t)(half + d)] + (double)taps[(size_t)(half - d)]) * std::cos(step * (double)kd);
  }
  return acc;
}

// pffft "ordered" real-spectrum layout: [K0, K_{N/2}, Re1, Im1, Re2, Im2, ...], Im = 0
inline std::vector<float> kerf_ordered(const std::vector<double>& gains, int N) {
  std::vector<float> o((size_t)N, 0.0f);
  o[0] = (float)gains[0];
  o[1] = (float)gains[(size_t)N / 2];
  for (int k = 1; k < N / 2; ++k) o[(size_t)2 * k] = (float)gains[(size_t)k];
  return o;
}

}  // namespace host
}  // namespace gblur
