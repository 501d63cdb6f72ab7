// ols_core.cuh — overlap-save row pass: the lane-level code of the warp-per-block kernel.
//
// Same result as the whole-line transform of the reference
//   process_channel_tiles              /root/reference/src/gaussianblur.cpp:202-244
//   pffft_sorted_optimized_convolution /root/reference/src/gaussianblur.cpp:62-79
// obtained differently: the padded line is cut into blocks of kNB = 1024 samples that
// overlap by 2*pad; each block is convolved circularly with the centred taps through a 32 x 32 FFT held by
// ONE WARP (32 points per lane in registers, one transpose through shared memory per direction, no CTA
// barrier), and the pad samples at either end of a block are dropped. That is the exact linear convolution;
// the reference's result differs from it by one rank-1 term, its Nyquist behaviour (bin N/2 multiplied by
// K[0] instead of K[N/2]):  y_ref[n] = y_lin[n] + (K[0] - K[N/2]) / N * (-1)^n * sum_m (-1)^m b[m],
// b = the reflect_101-padded line. The alternating sum is an exact integer on u8 input and is produced
// per row by a small pre-kernel; the block kernel adds the term on its last store.
//
// Everything here is written so that it also compiles as plain host C++ (g++ with the CUDA include
// directory): tests/cpp/ols_emulate.cpp runs these very functions lane by lane and the CPU test suite checks
// them against the oracle, so the arithmetic and all index maps are verified without a GPU.
#pragma once
#include <math.h>

#include <algorithm>
#include <cstdint>

#include "fft_radix.cuh"

#if defined(__CUDA_ARCH__)
#define GBLUR_OLS_LDG(p) __ldg(p)
#else
#define GBLUR_OLS_LDG(p) (*(p))
#endif

namespace gblur {
namespace ols {

#if !defined(__CUDACC__)
using std::min;
#endif

constexpr int kNB = 1024;        // block transform length
constexpr int kR = 32;           // radix of both stages = lanes per warp = points per lane
constexpr int kRowStride = 34;   // float2 slots per exchange row: 272 bytes keeps STS.64 columns and LDS.128 rows conflict-free
constexpr int kExch = kR * kRowStride;  // exchange slots (float2) per warp
constexpr int kGainStride = 36;  // floats per row of the gain table (LDS.128 by a quarter warp hits 8 distinct bank groups)
constexpr int kMaxPad = 256;     // at least half of every block is output

struct RowParams {
  const uint8_t* in;       // HWC u8, C == 4, frames stacked as extra rows
  long long in_pitch;      // bytes per image row (multiple of 4)
  int T, pad;              // pixels per row, reflect pad
  int L;                   // outputs per block = kNB - 2 * pad
  int nblk;                // blocks per row = ceil(T / L)
  long long nitems;        // rows_total * nblk * 2 (two channel pairs: R + iG, B + iA)
  // panel intermediate of the column pass (problems.cuh): [frame][panel = w/32][slot][32 floats], w = 4x + ch
  float* inter;
  int rows_per_frame;
  int row_offset;
  int blk_np, blk_padq;
  long long blk_frame_stride;
  const float4* qa;        // per image row: (K[0] - K[N/2]) / N * alternating sum, one float per channel
  const float2* tw;        // tw[a * 32 + b] = exp(-2 pi i a b / 1024)
  const float* gains;      // gains[k1 * kGainStride + k2] = K_1024[k1 + 32 k2] / 1024
};

__device__ __forceinline__ int phys_slot(int p, int padq) { return padq ? p + p / padq : p; }

// reflect_101 source pixel of padded position n, -1 in the zero tail (src/gaussianblur.cpp:217-228)
__device__ __forceinline__ int reflect_src(int n, int pad, int T) {
  int t = n - pad;
  if (t < 0) t = -t;
  if (t >= T) t = 2 * T - 2 - t;
  return (n < T + 2 * pad) ? t : -1;
}

// weight of pixel x in the alternating sum over the padded line: (-1)^(pad + x) times the number of padded
// positions that read x (itself, its left mirror if 1 <= x <= pad, its right mirror if T-1-pad <= x <= T-2)
__device__ __forceinline__ int altsum_weight(int x, int pad, int T) {
  const int m = 1 + ((x >= 1 && x <= pad) ? 1 : 0) + ((x >= T - 1 - pad && x <= T - 2) ? 1 : 0);
  return ((pad + x) & 1) ? -m : m;
}

__device__ __forceinline__ void altsum_acc(uint32_t w, int x, int pad, int T, int (&a)[4]) {
  const int m = altsum_weight(x, pad, T);
  a[0] += m * (int)(w & 255u);
  a[1] += m * (int)((w >> 8) & 255u);
  a[2] += m * (int)((w >> 16) & 255u);
  a[3] += m * (int)(w >> 24);
}

// One lane's share of a row's alternating sums (the warp adds its 32 shares): 16-byte loads of four pixels when
// the row is 16-byte aligned, groups without a mirrored pixel take the +-1 shortcut.
__device__ __forceinline__ void altsum_lane(const uint32_t* __restrict__ row, int T, int pad, int lane, int (&a)[4]) {
  int x_scalar = 0;  // first pixel the scalar loop still has to take
  if ((reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    const uint4* row4 = reinterpret_cast<const uint4*>(row);
    const int n4 = T >> 2;
#pragma unroll 4
    for (int g = lane; g < n4; g += 32) {
      const uint4 v = GBLUR_OLS_LDG(row4 + g);
      const int x = 4 * g;
      if (x > pad && x + 3 < T - 1 - pad) {
        const int sg = ((pad + x) & 1) ? -1 : 1;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int sh = 8 * c;
          a[c] += sg * ((int)((v.x >> sh) & 255u) - (int)((v.y >> sh) & 255u) + (int)((v.z >> sh) & 255u) -
                        (int)((v.w >> sh) & 255u));
        }
      } else {
        altsum_acc(v.x, x, pad, T, a);
        altsum_acc(v.y, x + 1, pad, T, a);
        altsum_acc(v.z, x + 2, pad, T, a);
        altsum_acc(v.w, x + 3, pad, T, a);
      }
    }
    x_scalar = 4 * n4;
  }
  for (int x = x_scalar + lane; x < T; x += 32) altsum_acc(GBLUR_OLS_LDG(row + x), x, pad, T, a);
}

// bytes (2 * pair, 2 * pair + 1) of an RGBA word as a complex sample
__device__ __forceinline__ float2 u8_pair(uint32_t w, int pair) {
#if defined(__CUDA_ARCH__)
  // 0x4B000000 | b is the float 8388608 + b: no conversion-pipe instruction
  const uint32_t sel = 0x7540u + 2u * (uint32_t)pair;
  const uint32_t lo = __byte_perm(w, 0x4B000000u, sel);
  const uint32_t hi = __byte_perm(w, 0x4B000000u, sel + 1u);
  return make_float2(__uint_as_float(lo) - 8388608.0f, __uint_as_float(hi) - 8388608.0f);
#else
  return make_float2((float)((w >> (16 * pair)) & 255u), (float)((w >> (16 * pair + 8)) & 255u));
#endif
}

// Where image row r lands in the panel intermediate (same geometry as make_blocked_dst, problems.cuh)
struct RowDst {
  float* base;        // inter + frame offset + slot(pad + r) * 32
  int dl, dr;         // float offsets of the mirrored border slots, 0 = none
  long long gstride;  // floats between consecutive panels
};
__device__ __forceinline__ RowDst make_row_dst(const RowParams& P, long long r_global) {
  RowDst d;
  const long long f = (r_global + P.row_offset) / P.rows_per_frame;
  const int r = (int)(r_global + P.row_offset - f * P.rows_per_frame);
  const int Tc = P.rows_per_frame, pad = P.pad;
  const int p1 = phys_slot(pad + r, P.blk_padq);
  d.base = P.inter + f * P.blk_frame_stride + (long long)p1 * 32;
  d.gstride = (long long)P.blk_np * 32;
  d.dl = (r >= 1 && r <= pad) ? (phys_slot(pad - r, P.blk_padq) - p1) * 32 : 0;
  d.dr = (r >= Tc - 1 - pad && r <= Tc - 2) ? (phys_slot(pad + 2 * Tc - 2 - r, P.blk_padq) - p1) * 32 : 0;
  return d;
}

// One item = one 1024-sample block of one channel pair. Per lane, 32 complex values stay in registers:
//   load_block : c[i] = b[s + lane + 32 i]                                           (lane = l, register = i)
//   then four passes q = 0..3 of ONE rolled loop, each starting with the same 32-point DFT over the register index:
//     q = 0: twiddle w^(lane k), column `lane` of the exchange, __syncwarp, row `lane` back   (lane = k1, register = l)
//     q = 1: X[lane + 32 k2] is complete: gains, complex conjugate
//     q = 2: as q = 0 — the inverse transform is conj(DFT(conj(.))) and conj(a conj(w)) = conj(a) w, so the inverse
//            half is THE SAME CODE on conjugated data                                         (lane = l, register = k1)
//     q = 3: nothing more: c[i] = conj(1024 y_lin[s + lane + 32 i])
//   store_block: conjugate back, add the Nyquist term, crop, panel stores
// The loop is kept rolled on purpose: the warps of an SM run out of phase here, so the instructions of an item must
// fit the SM's 32 KB instruction cache. The fully unrolled first version (four DFT copies, 44 KB) ran no faster than
// the whole-line kernel and slowed down with more independent CTAs (profiles/r01/ols_rowpass_v1_check.jsonl).
// Exchange pattern: S[row][col], row stride 34 slots. tw_write_col writes S[k][lane] (STS.64, conflict-free),
// read_row reads S[lane][0..31] (LDS.128, conflict-free). __syncwarp between them and after the reads.
template <bool INTERIOR>
__device__ __forceinline__ void load_block(float2 (&c)[kR], const RowParams& P, const uint32_t* __restrict__ row, int s,
                                           int pair, int lane) {
  if constexpr (INTERIOR) {
    const uint32_t* p = row + (s - P.pad + lane);
#pragma unroll
    for (int i = 0; i < kR; ++i) c[i] = u8_pair(GBLUR_OLS_LDG(p + kR * i), pair);
  } else {
#pragma unroll
    for (int i = 0; i < kR; ++i) {
      const int x = reflect_src(s + lane + kR * i, P.pad, P.T);
      const float2 v = u8_pair(GBLUR_OLS_LDG(row + (x < 0 ? 0 : x)), pair);
      c[i] = (x < 0) ? make_float2(0.f, 0.f) : v;
    }
  }
}

// twiddle w^(lane * k) on register k, then column `lane` of the exchange
__device__ __forceinline__ void tw_write_col(const float2 (&c)[kR], int lane, float2* __restrict__ ex,
                                             const float2* __restrict__ tw) {
  ex[lane] = c[0];
#pragma unroll
  for (int k = 1; k < kR; ++k) ex[k * kRowStride + lane] = cmul(c[k], tw[k * kR + lane]);
}

// row `lane` of the exchange
__device__ __forceinline__ void read_row(float2 (&c)[kR], int lane, const float2* __restrict__ ex) {
  const float4* rowp = reinterpret_cast<const float4*>(ex + lane * kRowStride);
#pragma unroll
  for (int m = 0; m < kR / 2; ++m) {
    const float4 t = rowp[m];
    c[2 * m] = make_float2(t.x, t.y);
    c[2 * m + 1] = make_float2(t.z, t.w);
  }
}

// c[k2] = conj(X[lane + 32 k2] * gain)
__device__ __forceinline__ void mid_gains(float2 (&c)[kR], int lane, const float* __restrict__ gains) {
  const float4* g = reinterpret_cast<const float4*>(gains + lane * kGainStride);
#pragma unroll
  for (int m = 0; m < kR / 4; ++m) {
    const float4 t = g[m];
    c[4 * m].x *= t.x, c[4 * m].y *= -t.x;
    c[4 * m + 1].x *= t.y, c[4 * m + 1].y *= -t.y;
    c[4 * m + 2].x *= t.z, c[4 * m + 2].y *= -t.z;
    c[4 * m + 3].x *= t.w, c[4 * m + 3].y *= -t.w;
  }
}

// c[i] = conj(1024 * y_lin[s + lane + 32 i]) (the 1/1024 sits in the gains)
__device__ __forceinline__ void store_block(const float2 (&c)[kR], const RowParams& P, const RowDst& d, float2 qa, int s,
                                            int pair, int lane) {
  // (-1)^n for n = s + lane + 32 i does not depend on i
  const float sg = ((s + lane) & 1) ? -1.0f : 1.0f;
  const float2 add = make_float2(qa.x * sg, qa.y * sg);
  // register i holds block index lane + 32 i; outputs are the indices [pad, min(pad + L, T - s + pad))
  const int lo = P.pad - lane;
  const int hi = min(P.pad + P.L, P.T - s + P.pad) - lane;
  // pixel x = x0 + 32 i lands in panel (x >> 3) = (x0 >> 3) + 4 i at float offset (x0 & 7) * 4 (+ 2 for pair 1)
  const int x0 = s + lane - P.pad;
  float* dp = d.base + 2 * pair + (long long)(x0 >> 3) * d.gstride + ((x0 & 7) << 2);
  const long long step = 4 * d.gstride;
  if ((d.dl | d.dr) == 0) {
#pragma unroll
    for (int i = 0; i < kR; ++i) {
      if (kR * i >= lo && kR * i < hi) *reinterpret_cast<float2*>(dp) = make_float2(c[i].x + add.x, add.y - c[i].y);
      dp += step;
    }
  } else {  // rows inside the reflect_101 border also fill the column pass's mirrored slots
#pragma unroll
    for (int i = 0; i < kR; ++i) {
      if (kR * i >= lo && kR * i < hi) {
        const float2 v = make_float2(c[i].x + add.x, add.y - c[i].y);
        *reinterpret_cast<float2*>(dp) = v;
        if (d.dl) *reinterpret_cast<float2*>(dp + d.dl) = v;
        if (d.dr) *reinterpret_cast<float2*>(dp + d.dr) = v;
      }
      dp += step;
    }
  }
}

}  // namespace ols
}  // namespace gblur
