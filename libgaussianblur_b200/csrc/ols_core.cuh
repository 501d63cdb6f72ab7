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

// four consecutive pixels starting at x (a multiple of 4)
__device__ __forceinline__ void altsum_group(const uint4& v, int x, int pad, int T, int (&a)[4]) {
  if (x > pad && x + 3 < T - 1 - pad) {  // no mirrored pixel in the group: weights +-1, alternating
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

// One lane's share of a row's alternating sums (the warp adds its 32 shares): 16-byte loads of four pixels when
// the row is 16-byte aligned, four of them issued before the first is used (the kernel is a pure stream over the
// input: memory-level parallelism is all that matters), groups without a mirrored pixel take the +-1 shortcut.
__device__ __forceinline__ void altsum_lane(const uint32_t* __restrict__ row, int T, int pad, int lane, int (&a)[4]) {
  int x_scalar = 0;  // first pixel the scalar loop still has to take
  if ((reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    const uint4* row4 = reinterpret_cast<const uint4*>(row);
    const int n4 = T >> 2;  // whole groups of four pixels
    int g = lane;
    for (; g + 96 < n4; g += 128) {  // groups g, g + 32, g + 64, g + 96 all exist
      const uint4 v0 = GBLUR_OLS_LDG(row4 + g), v1 = GBLUR_OLS_LDG(row4 + g + 32);
      const uint4 v2 = GBLUR_OLS_LDG(row4 + g + 64), v3 = GBLUR_OLS_LDG(row4 + g + 96);
      altsum_group(v0, 4 * g, pad, T, a);
      altsum_group(v1, 4 * (g + 32), pad, T, a);
      altsum_group(v2, 4 * (g + 64), pad, T, a);
      altsum_group(v3, 4 * (g + 96), pad, T, a);
    }
    for (; g < n4; g += 32) altsum_group(GBLUR_OLS_LDG(row4 + g), 4 * g, pad, T, a);
    x_scalar = 4 * n4;
  }
  for (int x = x_scalar + lane; x < T; x += 32) altsum_acc(GBLUR_OLS_LDG(row + x), x, pad, T, a);
}

// The same sums split over `nthreads` cooperating threads (row-per-CTA kernel): thread `tid` takes the groups of four
// pixels tid, tid + nthreads, ...
__device__ __forceinline__ void altsum_part(const uint32_t* __restrict__ row, int T, int pad, int tid, int nthreads,
                                            int (&a)[4]) {
  int x_scalar = 0;
  if ((reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    const uint4* row4 = reinterpret_cast<const uint4*>(row);
    const int n4 = T >> 2;
    for (int g = tid; g < n4; g += nthreads) altsum_group(GBLUR_OLS_LDG(row4 + g), 4 * g, pad, T, a);
    x_scalar = 4 * n4;
  }
  for (int x = x_scalar + tid; x < T; x += nthreads) altsum_acc(GBLUR_OLS_LDG(row + x), x, pad, T, a);
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

// ------------------------------------------------------------------------------------------------------------
// Column pass: fp32 panel intermediate -> u8 image (replaces the second process_channel_tiles call and
// interleave_channels, /root/reference/src/gaussianblur.cpp:261-305, helpers.hpp:248-274).
// One CTA of CW = 4 (or 8) warps takes a QUARTER (HALF) PANEL — 8 (16) float columns = 2 (4) RGBA pixels = one (two)
// 32-byte sectors per slot — and one 1024-slot block of the padded column; warp w transforms column pair w (floats
// 2w, 2w+1 of the unit) with the same four-pass loop as the row pass. The slots are fetched cooperatively with 16-byte loads straight into the four
// exchange buffers (column layout, so every lane later reads exactly what it will overwrite), the u8 results are
// collected in a shared tile and leave as one 8-byte store per image row.
// The reference's Nyquist term along a column needs sum_n (-1)^n I[n][w] over the padded column of the row-pass
// result I. By linearity that is the row pass applied to ONE synthetic row u = sum_r weight(r) x[r] (exact integers
// on u8 input, colsum_lane below), so it costs one more line instead of a pass over the 531 MB intermediate.
// ------------------------------------------------------------------------------------------------------------
struct ColParams {
  const float* inter;       // panel intermediate [panel][slot][32 floats]
  long long panel_stride;   // floats per panel = np * 32
  int padq;                 // slot holes of the whole-line column kernel's layout (0 = none)
  int T, pad;               // image rows, reflect pad
  int L, nblk;              // outputs per block, blocks per column
  int nunits;               // units (CW column pairs = CW / 2 pixels) with at least one valid pixel
  long long nitems;         // nunits * nblk
  int cols;
  uint8_t* out;             // HWC u8, 4 channels
  long long out_pitch;      // bytes per image row
  const float2* qc;         // per column pair (2x + pair): (K[0] - K[N/2]) / N of the column length * row pass of u
  const float2* tw;
  const float* gains;
};

// CW = warps per CTA = column pairs per unit: 4 (quarter panel, 32 bytes per slot, 8-byte stores) or 8 (half panel,
// 64 bytes per slot, 16-byte stores: half as many L1 line-operations per byte on both sides)
// cooperative fetch of one item: CW/2 threads share a slot (16 bytes each); 64 slots per pass, 16 passes
template <int CW>
__device__ __forceinline__ void col_fetch(const ColParams& P, int unit, int s, int t, float2* __restrict__ ex_all) {
  constexpr int TPS = CW / 2;         // threads per slot
  constexpr int UPP = 16 / CW;        // units per panel
  const int part = t % TPS, sl = t / TPS;
  const float* src = P.inter + (long long)(unit / UPP) * P.panel_stride + (unit % UPP) * (2 * CW) + part * 4;
  float2* e0 = ex_all + (2 * part) * kExch;  // exchange buffers of warps 2 * part and 2 * part + 1
  float2* e1 = e0 + kExch;
  const int nvalid = P.T + 2 * P.pad;
  float4 v[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int n = s + k * 64 + sl;
    v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (n < nvalid) v[k] = GBLUR_OLS_LDG(reinterpret_cast<const float4*>(src + (long long)phys_slot(n, P.padq) * 32));
  }
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int nl = k * 64 + sl;
    const int slot = (nl >> 5) * kRowStride + (nl & 31);
    e0[slot] = make_float2(v[k].x, v[k].y);
    e1[slot] = make_float2(v[k].z, v[k].w);
  }
}

// column `lane` of the exchange: c[i] = sample s + lane + 32 i
__device__ __forceinline__ void read_col(float2 (&c)[kR], int lane, const float2* __restrict__ ex) {
#pragma unroll
  for (int i = 0; i < kR; ++i) c[i] = ex[i * kRowStride + lane];
}

// uint8_t(v + 0.5f): fp32 add, truncate (helpers.hpp:251-252,270); clamped to [0, 255]
__device__ __forceinline__ uint32_t to_u8(float v) {
  const float f = fminf(fmaxf(v + 0.5f, 0.0f), 255.0f);
#if defined(__CUDA_ARCH__)
  return __float2uint_rz(f);
#else
  return (uint32_t)f;
#endif
}

// c[i] = conj(1024 y_lin[s + lane + 32 i]) -> two bytes per block index into the shared tile [1024][CW pairs]
template <int CW>
__device__ __forceinline__ void col_to_tile(const float2 (&c)[kR], float2 qc, int s, int w, int lane,
                                            uint16_t* __restrict__ tile) {
  const float sg = ((s + lane) & 1) ? -1.0f : 1.0f;
  const float2 add = make_float2(qc.x * sg, qc.y * sg);
#pragma unroll
  for (int i = 0; i < kR; ++i) {
    const uint32_t b0 = to_u8(c[i].x + add.x), b1 = to_u8(add.y - c[i].y);
    tile[(lane + kR * i) * CW + w] = (uint16_t)(b0 | (b1 << 8));
  }
}

// thread t of 32 CW stores block indices t + 32 CW k: 2 CW bytes (CW / 2 RGBA pixels) per image row
template <int CW>
__device__ __forceinline__ void col_store(const ColParams& P, int unit, int s, int t, const uint16_t* __restrict__ tile) {
  constexpr int PX = CW / 2;  // pixels per unit
  const int hi = min(P.pad + P.L, P.T - s + P.pad);
  const int npx = min(PX, P.cols - unit * PX);  // valid pixels of this unit (the last one may be cut)
  uint8_t* base = P.out + (long long)unit * (4 * PX);
#pragma unroll
  for (int k = 0; k < 1024 / (32 * CW); ++k) {
    const int idx = t + 32 * CW * k;
    if (idx >= P.pad && idx < hi) {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(tile + idx * CW);
      uint8_t* dp = base + (long long)(s + idx - P.pad) * P.out_pitch;
      if (npx == PX) {
        if constexpr (CW == 4)
          *reinterpret_cast<uint2*>(dp) = *reinterpret_cast<const uint2*>(src);
        else
          *reinterpret_cast<uint4*>(dp) = *reinterpret_cast<const uint4*>(src);
      } else {
        for (int p = 0; p < npx; ++p) reinterpret_cast<uint32_t*>(dp)[p] = src[p];
      }
    }
  }
}

// ---- the synthetic row u and its row pass (one CTA) ----
// u[4x + ch] = sum over image rows r of altsum_weight(r) * x[r][x][ch]; one thread per pixel column and row chunk
__device__ __forceinline__ void colsum_lane(const uint8_t* __restrict__ in, long long in_pitch, int x, int r0, int r1,
                                            int pad, int rows, int (&a)[4]) {
  const uint32_t* p = reinterpret_cast<const uint32_t*>(in + (long long)r0 * in_pitch) + x;
  for (int r = r0; r < r1; ++r) {
    const uint32_t w = GBLUR_OLS_LDG(p);
    const int m = altsum_weight(r, pad, rows);
    a[0] += m * (int)(w & 255u);
    a[1] += m * (int)((w >> 8) & 255u);
    a[2] += m * (int)((w >> 16) & 255u);
    a[3] += m * (int)(w >> 24);
    p = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(p) + in_pitch);
  }
}

// block loader for the synthetic row (int32 sums, always through the reflect map)
__device__ __forceinline__ void load_block_u(float2 (&c)[kR], const int* __restrict__ u, int T, int pad, int s, int pair,
                                             int lane) {
#pragma unroll
  for (int i = 0; i < kR; ++i) {
    const int x = reflect_src(s + lane + kR * i, pad, T);
    const int xx = x < 0 ? 0 : x;
    const float2 v = make_float2((float)u[4 * xx + 2 * pair], (float)u[4 * xx + 2 * pair + 1]);
    c[i] = (x < 0) ? make_float2(0.f, 0.f) : v;
  }
}

// c[i] = conj(1024 * linear convolution of u at s + lane + 32 i): add u's own Nyquist term, scale by the column
// coefficient and store per column pair (2x + pair)
__device__ __forceinline__ void store_block_u(const float2 (&c)[kR], float2 qa_u, float q_col, int T, int pad, int L, int s,
                                              int pair, int lane, float2* __restrict__ qc) {
  const float sg = ((s + lane) & 1) ? -1.0f : 1.0f;
  const int hi = min(pad + L, T - s + pad);
#pragma unroll
  for (int i = 0; i < kR; ++i) {
    const int idx = lane + kR * i;
    if (idx >= pad && idx < hi) {
      const int x = s + idx - pad;
      qc[2 * x + pair] = make_float2(q_col * (c[i].x + qa_u.x * sg), q_col * (qa_u.y * sg - c[i].y));
    }
  }
}

}  // namespace ols
}  // namespace gblur
