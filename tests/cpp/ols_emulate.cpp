// ols_emulate.cpp — host emulation of the overlap-save row-pass kernel (csrc/ols_kernels.cu): runs the very
// lane functions of csrc/ols_core.cuh for the 32 lanes of a warp one after the other, phase by phase (the
// phases are separated by __syncwarp in the kernel), on a panel intermediate laid out like the real one.
// Test infrastructure only: built by tests/test_ols_emulation.py with g++, never part of the product library.
#include <math.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

#include "ols_core.cuh"
#include "ols_host.hpp"

using namespace gblur;

extern "C" {

// in: rows x T RGBA u8. out: rows x T x 4 floats, the row-pass result (reference layout: element (r, 4x + ch)).
// n_ref: the reference's transform length along a row; n_col_slots / padq: panel geometry to emulate.
// Returns 0, or a negative code when a mirrored border slot disagrees with its source row.
int ols_emulate_rowpass(const uint8_t* in, int rows, int T, float sigma, int ksize, int n_ref, int padq, float* out) {
  const std::vector<float> taps = host::gaussian_taps(sigma, ksize);
  const ols::HostTables tb = ols::build_tables(taps, n_ref);
  const int pad = (ksize - 1) / 2;
  ols::RowParams P{};
  P.in = in;
  P.in_pitch = (long long)T * 4;
  P.T = T;
  P.pad = pad;
  P.L = ols::kNB - 2 * pad;
  P.nblk = (T + P.L - 1) / P.L;
  P.nitems = (long long)rows * P.nblk * 2;
  P.rows_per_frame = rows;
  P.row_offset = 0;
  P.blk_padq = padq;
  P.blk_np = ols::phys_slot(rows + 2 * pad, padq) + 3;
  const long long panels = ((long long)T * 4 + 31) / 32;
  P.blk_frame_stride = panels * P.blk_np * 32;
  std::vector<float> inter((size_t)P.blk_frame_stride, -12345.0f);
  P.inter = inter.data();
  // alternating sums exactly as altsum_rows_kernel forms them (integer sums, one float product)
  std::vector<float4> qa((size_t)rows);
  for (int r = 0; r < rows; ++r) {
    int a[4] = {0, 0, 0, 0};
    for (int lane = 0; lane < 32; ++lane)
      ols::altsum_lane(reinterpret_cast<const uint32_t*>(in + (size_t)r * T * 4), T, pad, lane, a);
    int chk[4] = {0, 0, 0, 0};  // the definition, pixel by pixel
    for (int x = 0; x < T; ++x)
      for (int ch = 0; ch < 4; ++ch) chk[ch] += ols::altsum_weight(x, pad, T) * (int)in[((size_t)r * T + x) * 4 + ch];
    for (int ch = 0; ch < 4; ++ch)
      if (chk[ch] != a[ch]) return -3;
    qa[(size_t)r] = make_float4(tb.q * (float)a[0], tb.q * (float)a[1], tb.q * (float)a[2], tb.q * (float)a[3]);
  }
  P.qa = qa.data();
  P.tw = reinterpret_cast<const float2*>(tb.tw.data());
  P.gains = tb.gains.data();
  std::vector<float4> exbuf((size_t)ols::kExch / 2 + 8);
  float2* ex = reinterpret_cast<float2*>(exbuf.data());
  for (long long it = 0; it < P.nitems; ++it) {
    const int pair = (int)(it & 1);
    const long long bj = it >> 1;
    const long long r = bj / P.nblk;
    const int j = (int)(bj - r * P.nblk);
    const int s = j * P.L;
    const uint32_t* row = reinterpret_cast<const uint32_t*>(P.in + r * P.in_pitch);
    const bool interior = s - P.pad >= 0 && s - P.pad + ols::kNB <= P.T;
    static float2 regs[32][ols::kR];  // the registers of the 32 lanes, kept across the __syncwarp points
    for (int lane = 0; lane < 32; ++lane) {
      if (interior)
        ols::load_block<true>(regs[lane], P, row, s, pair, lane);
      else
        ols::load_block<false>(regs[lane], P, row, s, pair, lane);
    }
    for (int q = 0; q < 4; ++q) {
      for (int lane = 0; lane < 32; ++lane) dft<ols::kR, false>(regs[lane]);
      if ((q & 1) == 0) {
        for (int lane = 0; lane < 32; ++lane) ols::tw_write_col(regs[lane], lane, ex, P.tw);
        for (int lane = 0; lane < 32; ++lane) ols::read_row(regs[lane], lane, ex);
      } else if (q == 1) {
        for (int lane = 0; lane < 32; ++lane) ols::mid_gains(regs[lane], lane, P.gains);
      }
    }
    const ols::RowDst d = ols::make_row_dst(P, r);
    const float4 q4 = P.qa[r];
    for (int lane = 0; lane < 32; ++lane)
      ols::store_block(regs[lane], P, d, pair ? make_float2(q4.z, q4.w) : make_float2(q4.x, q4.y), s, pair, lane);
  }
  // read the panel layout back: slot(pad + r) is row r; mirrored border slots must hold the same values
  auto at = [&](int slot_logical, int w) -> float {
    const int slot = ols::phys_slot(slot_logical, padq);
    return inter[(size_t)(w >> 5) * P.blk_np * 32 + (size_t)slot * 32 + (w & 31)];
  };
  for (int r = 0; r < rows; ++r)
    for (int w = 0; w < T * 4; ++w) {
      const float v = at(pad + r, w);
      out[(size_t)r * T * 4 + w] = v;
      if (r >= 1 && r <= pad && at(pad - r, w) != v) return -1;
      if (r >= rows - 1 - pad && r <= rows - 2 && at(pad + 2 * rows - 2 - r, w) != v) return -2;
    }
  return 0;
}

}  // extern "C"
