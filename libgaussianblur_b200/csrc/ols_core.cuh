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

#include <cstdint>

#include "fft_radix.cuh"

#if defined(__CUDA_ARCH__)
#define GBLUR_OLS_LDG(p) __ldg(p)
#else
#define GBLUR_OLS_LDG(p) (*(p))
#endif

namespace gblur {
namespace ols {

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

// bytes (2 * pair, 2 * pair + 1) of an RGBA word as a complex sample
__device__ __forceinline__ float2 u8_pair(uint32_t w, int pair) {
#if defined(__CUDA_ARCH__)
  // 0x4B000000 | b is the float 8388608 + b: no conversion-pipe instruction
  const uint32_t lo = pair ? __byte_perm(w, 0x4B000000u, 0x7542u) : __byte_perm(w, 0x4B000000u, 0x7540u);
  const uint32_t hi = pair ? __byte_perm(w, 0x4B000000u, 0x7543u) : __byte_perm(w, 0x4B000000u, 0x7541u);
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

// ---- phase A (lane = l): samples b[s + l + 32 i] -> DFT over i -> twiddle w^(l k1) -> column l of the exchange ----
template <bool INTERIOR>
__device__ __forceinline__ void phase_a(const RowParams& P, const uint32_t* __restrict__ row, int s, int pair, int lane,
                                        float2* __restrict__ ex, const float2* __restrict__ tw) {
  float2 c[kR];
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
  dft<kR, false>(c);
  ex[lane] = c[0];
#pragma unroll
  for (int k = 1; k < kR; ++k) ex[k * kRowStride + lane] = cmul(c[k], tw[k * kR + lane]);
}

// ---- phase B (lane = k1): row k1 -> DFT over l -> gains -> inverse DFT over k2 -> conj twiddle -> row k1 ----
__device__ __forceinline__ void phase_b(int lane, float2* __restrict__ ex, const float2* __restrict__ tw,
                                        const float* __restrict__ gains) {
  float2 c[kR];
  float4* rowp = reinterpret_cast<float4*>(ex + lane * kRowStride);
#pragma unroll
  for (int m = 0; m < kR / 2; ++m) {
    const float4 t = rowp[m];
    c[2 * m] = make_float2(t.x, t.y);
    c[2 * m + 1] = make_float2(t.z, t.w);
  }
  dft<kR, false>(c);  // c[k2] = X[k1 + 32 k2]
  const float4* g = reinterpret_cast<const float4*>(gains + lane * kGainStride);
#pragma unroll
  for (int m = 0; m < kR / 4; ++m) {
    const float4 t = g[m];
    c[4 * m].x *= t.x, c[4 * m].y *= t.x;
    c[4 * m + 1].x *= t.y, c[4 * m + 1].y *= t.y;
    c[4 * m + 2].x *= t.z, c[4 * m + 2].y *= t.z;
    c[4 * m + 3].x *= t.w, c[4 * m + 3].y *= t.w;
  }
  dft<kR, true>(c);  // c[l] = sum_k2 X[k1 + 32 k2] w32^(-l k2)
#pragma unroll
  for (int l = 1; l < kR; ++l) c[l] = cmulc(c[l], tw[l * kR + lane]);
#pragma unroll
  for (int m = 0; m < kR / 2; ++m) rowp[m] = make_float4(c[2 * m].x, c[2 * m].y, c[2 * m + 1].x, c[2 * m + 1].y);
}

// ---- phase C (lane = l): column l -> inverse DFT over k1 -> + Nyquist term -> crop -> panel stores ----
__device__ __forceinline__ void phase_c(const RowParams& P, const RowDst& d, float2 qa, int s, int pair, int lane,
                                        const float2* __restrict__ ex) {
  float2 c[kR];
#pragma unroll
  for (int k = 0; k < kR; ++k) c[k] = ex[k * kRowStride + lane];
  dft<kR, true>(c);  // c[i] = 1024 * y_lin[s + lane + 32 i] (the 1/1024 sits in the gains)
  // (-1)^n for n = s + lane + 32 i does not depend on i
  const float sg = ((s + lane) & 1) ? -1.0f : 1.0f;
  const float2 add = make_float2(qa.x * sg, qa.y * sg);
  const int lo = P.pad, hi = P.pad + P.L;
  float* colbase = d.base + 2 * pair;
#pragma unroll
  for (int i = 0; i < kR; ++i) {
    const int idx = lane + kR * i;
    const int x = s + idx - P.pad;
    if (idx >= lo && idx < hi && x < P.T) {
      const float2 v = make_float2(c[i].x + add.x, c[i].y + add.y);
      float* dp = colbase + (long long)(x >> 3) * d.gstride + ((x & 7) << 2);
      *reinterpret_cast<float2*>(dp) = v;
      if (d.dl) *reinterpret_cast<float2*>(dp + d.dl) = v;
      if (d.dr) *reinterpret_cast<float2*>(dp + d.dr) = v;
    }
  }
}

}  // namespace ols
}  // namespace gblur
