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
  P.T = P.Tpx = T;
  P.pad = P.padpx = pad;
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
        for (int lane = 0; lane < 32; ++lane) ols::tw_write_col(regs[lane], lane, ex, P.tw + lane * ols::kTwStride);
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

// One-channel rows (C == 1) and rows of three interleaved channels (C == 3, run as byte rows with the taps spread three
// bytes apart) through the lane functions of ols_row_c1_kernel<.., DIL = C>: a complex sequence carries image rows r and
// r + 1, both staged as the kernel's bulk copies stage them (bytes [x_lo, x_hi) of the row, 16-byte granules).
// in: rows x (T * C) u8. out: rows x (T * C) floats (element (r, C x + ch)). Returns 0 or a negative code.
template <int DIL>
static int emulate_rowpass_c(const uint8_t* in, int rows, int T, float sigma, int ksize, int n_ref, int padq, float* out) {
  const std::vector<float> taps = host::gaussian_taps(sigma, ksize);
  const ols::HostTables tb = ols::build_tables(taps, n_ref, 0, DIL);
  const int padpx = (ksize - 1) / 2;
  ols::RowParams P{};
  P.in = in;
  P.in_pitch = (long long)T * DIL;
  P.Tpx = T;
  P.padpx = padpx;
  P.T = T * DIL;
  P.pad = padpx * DIL;
  if (P.pad > ols::kMaxPad) return -20;
  P.L = ols::kNB - 2 * P.pad;
  P.nblk = (P.T + P.L - 1) / P.L;
  P.rows_per_frame = rows;
  P.blk_padq = padq;
  P.blk_np = ols::phys_slot(rows + 2 * padpx, padq) + 3;
  const long long panels = ((long long)P.T + 31) / 32;
  P.blk_frame_stride = panels * P.blk_np * 32;
  std::vector<float> inter((size_t)P.blk_frame_stride, -12345.0f);
  P.inter = inter.data();
  std::vector<float4> qa((size_t)rows);
  for (int r = 0; r < rows; ++r) {
    const uint8_t* row = in + (size_t)r * P.in_pitch;
    int chk[3] = {0, 0, 0};
    for (int x = 0; x < T; ++x)
      for (int ch = 0; ch < DIL; ++ch) chk[ch] += ols::altsum_weight(x, padpx, T) * (int)row[(size_t)x * DIL + ch];
    int a[3] = {0, 0, 0};
    if (DIL == 1) {
      for (int lane = 0; lane < 32; ++lane) a[0] += ols::altsum_lane_c1(row, T, padpx, lane);
    } else {
      for (int lane = 0; lane < 32; ++lane) ols::altsum_lane_c3(row, T, padpx, lane, a);
    }
    for (int ch = 0; ch < DIL; ++ch)
      if (chk[ch] != a[ch]) return -3;
    qa[(size_t)r] = make_float4(tb.q * (float)a[0], tb.q * (float)a[1], tb.q * (float)a[2], 0.f);
  }
  P.tw = reinterpret_cast<const float2*>(tb.tw.data());
  P.gains = tb.gains.data();
  std::vector<float4> exbuf((size_t)ols::kExch / 2 + 8);
  float2* ex = reinterpret_cast<float2*>(exbuf.data());
  std::vector<uint8_t> st0(ols::kStageC1), st1(ols::kStageC1);
  const long long npairs = (rows + 1) / 2;
  static float2 regs[32][ols::kR];
  for (long long rp = 0; rp < npairs; ++rp)
    for (int j = 0; j < P.nblk; ++j) {
      const int s = j * P.L;
      const long long r = 2 * rp;
      const bool second = r + 1 < rows;
      // the bulk copies of the kernel's issue(): bytes [x_lo, x_hi) of both rows, stale bytes elsewhere
      const int x_lo = ols::stage_x_lo_c1(s, P.pad);
      int x_hi = s - P.pad + ols::kNB;
      x_hi = ((x_hi < P.T ? x_hi : P.T) + 15) & ~15;
      if (x_hi - x_lo > ols::kStageC1 || x_hi > (int)P.in_pitch) return -21;
      std::fill(st0.begin(), st0.end(), (uint8_t)0xA5);
      std::fill(st1.begin(), st1.end(), (uint8_t)0x5A);
      std::memcpy(st0.data(), in + (size_t)r * P.in_pitch + x_lo, (size_t)(x_hi - x_lo));
      if (second) std::memcpy(st1.data(), in + (size_t)(r + 1) * P.in_pitch + x_lo, (size_t)(x_hi - x_lo));
      const bool interior = s - P.pad >= 0 && s - P.pad + ols::kNB <= P.T;
      for (int lane = 0; lane < 32; ++lane) {
        if (interior)
          ols::load_block_staged_c1<true, DIL>(regs[lane], P, st0.data(), st1.data(), s, lane);
        else
          ols::load_block_staged_c1<false, DIL>(regs[lane], P, st0.data(), st1.data(), s, lane);
      }
      for (int q = 0; q < 4; ++q) {
        for (int lane = 0; lane < 32; ++lane) dft<ols::kR, false>(regs[lane]);
        if ((q & 1) == 0) {
          for (int lane = 0; lane < 32; ++lane) ols::tw_write_col(regs[lane], lane, ex, P.tw + lane * ols::kTwStride);
          for (int lane = 0; lane < 32; ++lane) ols::read_row(regs[lane], lane, ex);
        } else if (q == 1) {
          for (int lane = 0; lane < 32; ++lane) ols::mid_gains(regs[lane], lane, P.gains);
        }
      }
      const ols::RowDst d0 = ols::make_row_dst(P, r), d1 = ols::make_row_dst(P, second ? r + 1 : r);
      const float4 q0 = qa[(size_t)r], q1 = second ? qa[(size_t)r + 1] : make_float4(0.f, 0.f, 0.f, 0.f);
      for (int lane = 0; lane < 32; ++lane) ols::store_block_c1<DIL>(regs[lane], P, d0, d1, q0, q1, s, lane, second);
    }
  auto at = [&](int slot_logical, int w) -> float {
    const int slot = ols::phys_slot(slot_logical, padq);
    return inter[(size_t)(w >> 5) * P.blk_np * 32 + (size_t)slot * 32 + (w & 31)];
  };
  for (int r = 0; r < rows; ++r)
    for (int w = 0; w < P.T; ++w) {
      const float v = at(padpx + r, w);
      out[(size_t)r * P.T + w] = v;
      if (r >= 1 && r <= padpx && at(padpx - r, w) != v) return -1;
      if (r >= rows - 1 - padpx && r <= rows - 2 && at(padpx + 2 * rows - 2 - r, w) != v) return -2;
    }
  return 0;
}

extern "C" int ols_emulate_rowpass_c(const uint8_t* in, int rows, int T, int C, float sigma, int ksize, int n_ref, int padq,
                                     float* out) {
  if (C == 1) return emulate_rowpass_c<1>(in, rows, T, sigma, ksize, n_ref, padq, out);
  if (C == 3) return emulate_rowpass_c<3>(in, rows, T, sigma, ksize, n_ref, padq, out);
  return -30;
}

// Whole blur through the emulated kernels: row pass into the panel intermediate (no bank-padding holes), column sums
// -> synthetic row u -> its row pass -> column terms, then the column pass: the four TMA tile loads of an item are
// emulated as a copy into the landing area under the 64-byte swizzle (col_landing_offset is the kernel's own map of it),
// the eight warps run the kernel's lane functions, and the TMA tile stores as a clipped copy of the u8 tile.
// in / out: rows x cols RGBA u8, cols a multiple of 4. Returns 0 or a negative code.
// alpha_keep: RGBA with alpha passed through (ols_col_kernel<true, true>): the item's strip of the INPUT image is put into
// the tile first (the kernel's TMA tile loads, rows past the image zero-filled) and the (B, A) warps write only their
// first byte.
static int emulate_blur(const uint8_t* in, int rows, int cols, float sigma, int ksize, int n_row, int n_col, uint8_t* out,
                        bool alpha_keep = false) {
  constexpr int CW = ols::kColWarps;
  const std::vector<float> taps = host::gaussian_taps(sigma, ksize);
  const ols::HostTables tb = ols::build_tables(taps, n_row, n_col);
  const int pad = (ksize - 1) / 2, T = cols;
  const float2* tw = reinterpret_cast<const float2*>(tb.tw.data());
  ols::RowParams P{};
  P.in = in;
  P.in_pitch = (long long)T * 4;
  P.T = P.Tpx = T;
  P.pad = P.padpx = pad;
  P.L = ols::kNB - 2 * pad;
  P.nblk = (T + P.L - 1) / P.L;
  P.nitems = (long long)rows * P.nblk * 2;
  P.rows_per_frame = rows;
  P.blk_padq = 0;
  P.pair_major = 1;
  P.blk_np = rows + 2 * pad + 5;
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
  std::vector<float4> regionbuf((size_t)ols::kColRegion / 16 + 8);
  unsigned char* region = reinterpret_cast<unsigned char*>(regionbuf.data());
  static float2 regs[CW][32][ols::kR];
  auto passes = [&](float2 (*rg)[ols::kR], float2* ex) {
    for (int q = 0; q < 4; ++q) {
      for (int lane = 0; lane < 32; ++lane) dft<ols::kR, false>(rg[lane]);
      if ((q & 1) == 0) {
        for (int lane = 0; lane < 32; ++lane) ols::tw_write_col(rg[lane], lane, ex, tw + lane * ols::kTwStride);
        for (int lane = 0; lane < 32; ++lane) ols::read_row(rg[lane], lane, ex);
      } else if (q == 1) {
        for (int lane = 0; lane < 32; ++lane) ols::mid_gains(rg[lane], lane, P.gains);
      }
    }
  };
  float2* ex0 = reinterpret_cast<float2*>(region);
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
    passes(regs[0], ex0);
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
    passes(regs[0], ex0);
    for (int lane = 0; lane < 32; ++lane)
      ols::store_block_u(regs[0][lane], make_float2(qa_u[2 * pair], qa_u[2 * pair + 1]), tb.q_col, T, pad, P.L, s, pair,
                         lane, qc.data());
  }
  // ---- column pass ----
  if ((T * 4) % ols::kColFloats) return -10;
  const int L = P.L, nblk = (rows + L - 1) / L, nunits = T * 4 / ols::kColFloats;
  std::vector<uint64_t> tilebuf(ols::kColTileBytes / 8);
  unsigned char* tile = reinterpret_cast<unsigned char*>(tilebuf.data());
  for (int j = 0; j < nblk; ++j)
    for (int unit = 0; unit < nunits; ++unit) {
      const int s = j * L;
      // the landing boxes: {8 floats, 256 slots} x 4 per pair, zero beyond the tensor's np slots. Box of pair g starts
      // at float 16 g + 8 (unit & 1) of the panel row; float 2 k + c of the box is (pixel k of the unit, component c)
      for (int n = 0; n < ols::kNB; ++n)
        for (int w = 0; w < CW; ++w)
          for (int c = 0; c < 2; ++c) {
            const int slot = s + n, g = w >> 2, k = w & 3;
            const float v = slot < P.blk_np ? inter[(size_t)(unit >> 1) * P.blk_np * 32 + (size_t)slot * 32 + 16 * g +
                                                    8 * (unit & 1) + 2 * k + c]
                                            : 0.0f;
            std::memcpy(region + ols::col_landing_offset(n, w) + c * 4, &v, 4);
          }
      for (int w = 0; w < CW; ++w)
        for (int lane = 0; lane < 32; ++lane) ols::read_col(regs[w][lane], lane, w, region);
      std::memset(tile, 0xEE, ols::kColTileBytes);
      if (alpha_keep)
        for (int k = 0; k < L; ++k) {
          if (s + k < rows)
            std::memcpy(tile + (size_t)k * ols::kColFloats, in + (size_t)(s + k) * T * 4 + (size_t)unit * ols::kColFloats,
                        ols::kColFloats);
          else
            std::memset(tile + (size_t)k * ols::kColFloats, 0, ols::kColFloats);
        }
      for (int w = 0; w < CW; ++w) {
        passes(regs[w], reinterpret_cast<float2*>(region) + w * ols::kExch);
        for (int lane = 0; lane < 32; ++lane)
          ols::col_to_tile(regs[w][lane], qc[((size_t)unit * 4 + (w & 3)) * 2 + (w >> 2)], s, 0, pad, L,
                           4 * (w & 3) + 2 * (w >> 2), lane, tile, alpha_keep && (w >> 2) == 1);
      }
      // the tile stores: rows [s, s + L) clipped to the image
      for (int k = 0; k < L && s + k < rows; ++k)
        std::memcpy(out + (size_t)(s + k) * T * 4 + (size_t)unit * ols::kColFloats, tile + (size_t)k * ols::kColFloats,
                    ols::kColFloats);
    }
  return 0;
}

extern "C" int ols_emulate_blur(const uint8_t* in, int rows, int cols, float sigma, int ksize, int n_row, int n_col,
                                uint8_t* out) {
  return emulate_blur(in, rows, cols, sigma, ksize, n_row, n_col, out);
}

extern "C" int ols_emulate_blur_keep_alpha(const uint8_t* in, int rows, int cols, float sigma, int ksize, int n_row, int n_col,
                                           uint8_t* out) {
  return emulate_blur(in, rows, cols, sigma, ksize, n_row, n_col, out, true);
}
