// ols_host.hpp — host-only tables of the overlap-save row pass (ols_core.cuh): the 32 x 32 twiddles, the
// block gains K_1024[k] / 1024 of the reference's centred taps (src/gaussianblur.cpp:28-60) and the
// coefficient of the reference's Nyquist term for its own transform length N (src/gaussianblur.cpp:62-79
// with pffft.h:143-146).
#pragma once
#include <cmath>
#include <vector>

#include "hostsem.hpp"

namespace gblur {
namespace ols {

struct HostTables {
  std::vector<float> tw;     // 32 rows of 34 complex values, tw[lane*34 + k] = exp(-2 pi i lane k / 1024), k < 32
  std::vector<float> gains;  // 32 rows of 36 floats, gains[k1*36 + k2] = K_1024[k1 + 32 k2] / 1024
  float q = 0.f;             // (K_N[0] - K_N[N/2]) / N for the reference's transform length along a row
  float q_col = 0.f;         // the same for its transform length along a column
};

// taps.size() must not exceed 1024 - 512 + 1 (pad <= 256)
// dil > 1: the block gains of the taps spread dil samples apart (rows of dil interleaved channels run as one row); the
// Nyquist coefficients stay those of the taps themselves at the reference's lengths
inline HostTables build_tables(const std::vector<float>& taps, int n_ref, int n_ref_col = 0, int dil = 1) {
  constexpr int NB = 1024, R = 32, GS = 36, TS = 34;
  constexpr double two_pi = 6.28318530717958647692528676655900577;
  HostTables t;
  t.tw.assign((size_t)2 * R * TS, 0.0f);
  for (int a = 0; a < R; ++a)
    for (int b = 0; b < R; ++b) {
      const double ang = two_pi * (double)((a * b) % NB) / (double)NB;
      t.tw[2 * (a * TS + b)] = (float)std::cos(ang);
      t.tw[2 * (a * TS + b) + 1] = (float)(-std::sin(ang));
    }
  std::vector<float> spread(((size_t)taps.size() - 1) * (size_t)dil + 1, 0.0f);
  for (size_t j = 0; j < taps.size(); ++j) spread[j * (size_t)dil] = taps[j];
  const std::vector<double> kb = host::kernel_gains(spread, NB);
  t.gains.assign((size_t)R * GS, 0.0f);
  for (int k1 = 0; k1 < R; ++k1)
    for (int k2 = 0; k2 < R; ++k2) {
      const int k = k1 + R * k2;
      const int kk = k <= NB / 2 ? k : NB - k;
      t.gains[(size_t)k1 * GS + k2] = (float)(kb[(size_t)kk] / (double)NB);
    }
  t.q = (float)((host::kernel_gain_at(taps, n_ref, 0) - host::kernel_gain_at(taps, n_ref, n_ref / 2)) / (double)n_ref);
  if (n_ref_col > 0)
    t.q_col = (float)((host::kernel_gain_at(taps, n_ref_col, 0) - host::kernel_gain_at(taps, n_ref_col, n_ref_col / 2)) /
                      (double)n_ref_col);
  return t;
}

}  // namespace ols
}  // namespace gblur
