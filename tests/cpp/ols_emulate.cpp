// ols_emulate.cpp — host emulation of the overlap-save row-pass kernel (csrc/ols_kernels.cu): runs the very
// lane functions of csrc/ols_core.cuh for the 32 lanes of a warp one after the other, phase by phase (the
// phases are separated by __syncwarp in the kernel), on a panel intermediate laid out like the real one.
// Test infrastructure only: built by tests/test_ols_emulation.py with g++, never part of the product library.
#include <math.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <algorithm>
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
    // the row-per-CTA kernel's split of the same sums over 2 * nblk warps
    int b[4] = {0, 0, 0, 0};
    const int nthreads = 64 * P.nblk;
    for (int tid = 0; tid < nthreads; ++tid)
      ols::altsum_part(reinterpret_cast<const uint32_t*>(in + (size_t)r * T * 4), T, pad, tid, nthreads, b);
    for (int ch = 0; ch < 4; ++ch)
      if (chk[ch] != b[ch]) return -4;
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

// Whole blur through the emulated kernels: row pass into the panel intermediate, column sums -> synthetic row u ->
// its row pass -> column terms, then the column pass (four warps per quarter panel, 128 emulated threads).
// in / out: rows x cols RGBA u8. Returns 0 or a negative code.
template <int CW>
static int emulate_blur(const uint8_t* in, int rows, int cols, float sigma, int ksize, int n_row, int n_col, int padq,
                        uint8_t* out) {
  const std::vector<float> taps = host::gaussian_taps(sigma, ksize);
  const ols::HostTables tb = ols::build_tables(taps, n_row, n_col);
  const int pad = (ksize - 1) / 2, T = cols;
  const float2* tw = reinterpret_cast<const float2*>(tb.tw.data());
  ols::RowParams P{};
  P.in = in;
  P.in_pitch = (long long)T * 4;
  P.T = T;
  P.pad = pad;
  P.L = ols::kNB - 2 * pad;
  P.nblk = (T + P.L - 1) / P.L;
  P.nitems = (long long)rows * P.nblk * 2;
  P.rows_per_frame = rows;
  P.blk_padq = padq;
  P.blk_np = ols::phys_slot(rows + 2 * pad, padq) + 5;
  const long long panels = ((long long)T * 4 + 31) / 32;
  P.blk_frame_stride = panels * P.blk_np * 32;
  std::vector<float> inter((size_t)P.blk_frame_stride, 0.0f);
  P.inter = inter.data();
  std::vector<float4> qa((size_t)rows);
  for (int r = 0; r < rows; ++r) {
    int a[4] = {0, 0, 0, 0};
    for (int lane = 0; lane < 32; ++lane)
      ols::altsum_lane(reinterpret_cast<const uint32_t*>(in + (size_t)r * T * 4), T, pad, lane, a);
    qa[(size_t)r] = make_float4(tb.q * (float)a[0], tb.q * (float)a[1], tb.q * (float)a[2], tb.q * (float)a[3]);
  }
  P.qa = qa.data();
  P.tw = tw;
  P.gains = tb.gains.data();
  std::vector<float4> exbuf((size_t)CW * ols::kExch / 2 + 8);
  float2* ex_all = reinterpret_cast<float2*>(exbuf.data());
  static float2 regs[CW][32][ols::kR];
  auto passes = [&](float2 (*rg)[ols::kR], float2* ex) {
    for (int q = 0; q < 4; ++q) {
      for (int lane = 0; lane < 32; ++lane) dft<ols::kR, false>(rg[lane]);
      if ((q & 1) == 0) {
        for (int lane = 0; lane < 32; ++lane) ols::tw_write_col(rg[lane], lane, ex, tw);
        for (int lane = 0; lane < 32; ++lane) ols::read_row(rg[lane], lane, ex);
      } else if (q == 1) {
        for (int lane = 0; lane < 32; ++lane) ols::mid_gains(rg[lane], lane, P.gains);
      }
    }
  };
  // ---- row pass ----
  for (long long it = 0; it < P.nitems; ++it) {
    const int pair = (int)(it & 1);
    const long long bj = it >> 1;
    const long long r = bj / P.nblk;
    const int s = (int)(bj - r * P.nblk) * P.L;
    const uint32_t* row = reinterpret_cast<const uint32_t*>(P.in + r * P.in_pitch);
    const bool interior = s - P.pad >= 0 && s - P.pad + ols::kNB <= P.T;
    for (int lane = 0; lane < 32; ++lane) {
      if (interior)
        ols::load_block<true>(regs[0][lane], P, row, s, pair, lane);
      else
        ols::load_block<false>(regs[0][lane], P, row, s, pair, lane);
    }
    passes(regs[0], ex_all);
    const ols::RowDst d = ols::make_row_dst(P, r);
    const float4 q4 = P.qa[r];
    for (int lane = 0; lane < 32; ++lane)
      ols::store_block(regs[0][lane], P, d, pair ? make_float2(q4.z, q4.w) : make_float2(q4.x, q4.y), s, pair, lane);
  }
  // ---- synthetic row u (colsum_kernel: 64 image rows per thread, integer atomics) ----
  std::vector<int> u((size_t)T * 4, 0);
  for (int r0 = 0; r0 < rows; r0 += 64)
    for (int x = 0; x < T; ++x) {
      int a[4] = {0, 0, 0, 0};
      ols::colsum_lane(in, P.in_pitch, x, r0, std::min(rows, r0 + 64), pad, rows, a);
      for (int c = 0; c < 4; ++c) u[(size_t)4 * x + c] += a[c];
    }
  // ---- its row pass (u_rowpass_kernel) ----
  long long au[4] = {0, 0, 0, 0};
  for (int x = 0; x < T; ++x)
    for (int c = 0; c < 4; ++c) au[c] += (long long)ols::altsum_weight(x, pad, T) * u[(size_t)4 * x + c];
  float qa_u[4];
  for (int c = 0; c < 4; ++c) qa_u[c] = tb.q * (float)au[c];
  std::vector<float2> qc((size_t)T * 2);
  for (int it = 0; it < P.nblk * 2; ++it) {
    const int pair = it & 1, s = (it >> 1) * P.L;
    for (int lane = 0; lane < 32; ++lane) ols::load_block_u(regs[0][lane], u.data(), T, pad, s, pair, lane);
    passes(regs[0], ex_all);
    for (int lane = 0; lane < 32; ++lane)
      ols::store_block_u(regs[0][lane], make_float2(qa_u[2 * pair], qa_u[2 * pair + 1]), tb.q_col, T, pad, P.L, s, pair,
                         lane, qc.data());
  }
  // ---- column pass ----
  if (T % (CW / 2)) return -10;
  ols::ColParams C{};
  C.inter = inter.data();
  C.panel_stride = (long long)P.blk_np * 32;
  C.padq = padq;
  C.T = rows;
  C.pad = pad;
  C.L = P.L;
  C.nblk = (rows + C.L - 1) / C.L;
  C.nunits = (T + CW / 2 - 1) / (CW / 2);
  C.nitems = (long long)C.nunits * C.nblk;
  C.cols = T;
  C.out = out;
  C.out_pitch = (long long)T * 4;
  C.qc = qc.data();
  C.tw = tw;
  C.gains = tb.gains.data();
  std::vector<uint64_t> tilebuf(1024 * CW / 4);
  uint16_t* tile = reinterpret_cast<uint16_t*>(tilebuf.data());
  for (long long it = 0; it < C.nitems; ++it) {
    const int j = (int)(it / C.nunits);
    const int unit = (int)(it - (long long)j * C.nunits);
    const int s = j * C.L;
    for (int t = 0; t < 32 * CW; ++t) ols::col_fetch<CW>(C, unit, s, t, ex_all);
    for (int w = 0; w < CW; ++w) {
      float2* ex = ex_all + w * ols::kExch;
      for (int lane = 0; lane < 32; ++lane) ols::read_col(regs[w][lane], lane, ex);
      passes(regs[w], ex);
      for (int lane = 0; lane < 32; ++lane) ols::col_to_tile<CW>(regs[w][lane], C.qc[unit * CW + w], s, w, lane, tile);
    }
    for (int t = 0; t < 32 * CW; ++t) ols::col_store<CW>(C, unit, s, t, tile);
  }
  return 0;
}

extern "C" int ols_emulate_blur(const uint8_t* in, int rows, int cols, float sigma, int ksize, int n_row, int n_col,
                                int padq, int cw, uint8_t* out) {
  return cw == 8 ? emulate_blur<8>(in, rows, cols, sigma, ksize, n_row, n_col, padq, out)
                 : emulate_blur<4>(in, rows, cols, sigma, ksize, n_row, n_col, padq, out);
}

