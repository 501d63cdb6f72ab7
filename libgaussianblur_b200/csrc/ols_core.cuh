// ols_core.cuh — overlap-save passes: the lane-level code of the warp-per-block kernels (ols_kernels.cu).
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
constexpr int kTwStride = 34;    // float2 per lane row of the twiddle table (LDS.128 rows, conflict-free like the exchange)
constexpr int kTwSlots = kR * kTwStride;
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
  int pair_major;          // 1: the 32 floats of a panel row are ordered [pair][pixel][2] instead of [pixel][pair][2]
  int slot_origin;         // the panel buffer starts at this slot of the padded column (halo-sharded runs), else 0
  long long split_at, split_gap;  // a launch over two row ranges: logical rows >= split_at lie split_gap rows further on
  int Tpx, padpx;          // interleaved 3-channel rows run as one byte row (T, pad, L in BYTES): pixels per row, pad in pixels
  long long blk_frame_stride;
  const float4* qa;        // per image row: (K[0] - K[N/2]) / N * alternating sum, one float per channel
  const float2* tw;        // tw[lane * kTwStride + k] = exp(-2 pi i lane k / 1024)
  const float* gains;      // gains[k1 * kGainStride + k2] = K_1024[k1 + 32 k2] / 1024
};

__device__ __forceinline__ int phys_slot(int p, int padq) { return padq ? p + p / padq : p; }

// Row of the input buffer (and of the frame stack) behind logical row r of the launch. One launch can cover two row
// ranges — the halo rows above and below a rank's band (capi.cpp gblur_halo_rowpass2_device): two short launches cost
// 0.125 ms where the same rows inside one launch cost 0.05 (profiles/r02/scale).
__device__ __forceinline__ long long phys_row(const RowParams& P, long long r) {
  return r + ((P.split_gap != 0 && r >= P.split_at) ? P.split_gap : 0);
}

// reflect_101 source pixel of padded position n, -1 in the zero tail (src/gaussianblur.cpp:217-228)
__device__ __forceinline__ int reflect_src(int n, int pad, int T) {
  int t = n - pad;
  if (t < 0) t = -t;
  if (t >= T) t = 2 * T - 2 - t;
  return (n < T + 2 * pad) ? t : -1;
}

// The same for a row of DIL interleaved channels seen as one byte row: padded byte n is channel n % DIL of padded pixel
// n / DIL, so its source is byte DIL * reflect + channel. (An RGB row convolved per channel = the byte row convolved with
// the taps spread DIL bytes apart: one transform serves all channels and every store is a contiguous run of floats.)
template <int DIL>
__device__ __forceinline__ int reflect_src_dil(int n, int padpx, int Tpx) {
  const int pp = n / DIL, ch = n - DIL * pp;  // n >= 0
  const int t = reflect_src(pp, padpx, Tpx);
  return t < 0 ? -1 : DIL * t + ch;
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

// bytes (2 * pair, 2 * pair + 1) of an RGBA word as a complex sample
__device__ __forceinline__ float2 u8_pair(uint32_t w, int pair) {
#if defined(__CUDA_ARCH__)
  // 0x4B000000 | b is the float 8388608 + b: no conversion-pipe instruction
  const uint32_t sel = 0x7540u + 2u * (uint32_t)pair;
  const uint32_t lo = __byte_perm(w, 0x4B000000u, sel);
  const uint32_t hi = __byte_perm(w, 0x4B000000u, sel + 1u);
  return pk_add(make_float2(__uint_as_float(lo), __uint_as_float(hi)), pk_bc(-8388608.0f));
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
  const int Tc = P.rows_per_frame, pad = P.padpx;  // the pad along a column is in pixels also where the row runs in bytes
  const int p1 = phys_slot(pad + r, P.blk_padq) - P.slot_origin;
  d.base = P.inter + f * P.blk_frame_stride + (long long)p1 * 32;
  d.gstride = (long long)P.blk_np * 32;
  d.dl = (r >= 1 && r <= pad) ? (phys_slot(pad - r, P.blk_padq) - P.slot_origin - p1) * 32 : 0;
  d.dr = (r >= Tc - 1 - pad && r <= Tc - 2) ? (phys_slot(pad + 2 * Tc - 2 - r, P.blk_padq) - P.slot_origin - p1) * 32 : 0;
  if (P.slot_origin) {  // a halo-sharded buffer holds only its own slots: mirrors outside it belong to another rank
    const int ml = p1 + d.dl / 32, mr = p1 + d.dr / 32;
    if (ml < 0 || ml >= P.blk_np) d.dl = 0;
    if (mr < 0 || mr >= P.blk_np) d.dr = 0;
  }
  return d;
}

// One item = one 1024-sample block of one complex sequence (two real lines). Per lane, 32 complex values stay in
// registers:
//   load       : c[i] = a[lane + 32 i]                                                (lane = l, register = i)
//   then four passes q = 0..3 of ONE rolled loop, each starting with the same 32-point DFT over the register index:
//     q = 0: twiddle w^(lane k), column `lane` of the exchange, __syncwarp, row `lane` back   (lane = k1, register = l)
//     q = 1: A[lane + 32 k2] is complete: gains
//     q = 2: as q = 0. Passes 0-1 are a complete forward DFT with natural order in (index = lane + 32 register) and
//            out; running THE SAME CODE on the product B = A G gives C[m] = sum_k B[k] w^(km) = 1024 y[-m mod 1024],
//            the inverse transform with its index negated — no conjugation and no re/im exchange anywhere.
//     q = 3: nothing more: lane l, register j hold 1024 y[n], n = (32 - l) + 32 (31 - j)  (mod 1024)
//   store      : registers in descending order, lanes in descending order of n; add the Nyquist term, crop.
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

// The same block out of a staged copy of the row's pixels [x_lo, x_lo + kStagePx) in shared memory (the row kernel
// bulk-copies it there one item ahead, so no global-load latency sits between two items of a warp).
constexpr int kStagePx = kNB + 4;  // 1024 pixels from a 4-pixel (16-byte) aligned start
__device__ __forceinline__ int stage_x_lo(int s, int pad) { return (s - pad > 0 ? s - pad : 0) & ~3; }
template <bool INTERIOR>
__device__ __forceinline__ void load_block_staged(float2 (&c)[kR], const RowParams& P, const uint32_t* __restrict__ stage,
                                                  int s, int pair, int lane) {
  const int x_lo = stage_x_lo(s, P.pad);
  if constexpr (INTERIOR) {
    const uint32_t* p = stage + (s - P.pad - x_lo + lane);
#pragma unroll
    for (int i = 0; i < kR; ++i) c[i] = u8_pair(p[kR * i], pair);
  } else {
#pragma unroll
    for (int i = 0; i < kR; ++i) {
      const int x = reflect_src(s + lane + kR * i, P.pad, P.T);
      const float2 v = u8_pair(stage[x < 0 ? 0 : x - x_lo], pair);
      c[i] = (x < 0) ? make_float2(0.f, 0.f) : v;
    }
  }
}

// ---- one-channel images: a complex sequence carries image rows r (real part) and r + 1 (imaginary part) ----
// staged copies of both rows' bytes [x_lo, x_lo + kStageC1) (16-byte aligned start)
constexpr int kStageC1 = kNB + 32;
__device__ __forceinline__ int stage_x_lo_c1(int s, int pad) { return (s - pad > 0 ? s - pad : 0) & ~15; }
__device__ __forceinline__ float2 u8x2_to_f32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  return pk_add(make_float2(__uint_as_float(0x4B000000u | a), __uint_as_float(0x4B000000u | b)), pk_bc(-8388608.0f));
#else
  return make_float2((float)a, (float)b);
#endif
}
template <bool INTERIOR, int DIL = 1>
__device__ __forceinline__ void load_block_staged_c1(float2 (&c)[kR], const RowParams& P, const uint8_t* __restrict__ st0,
                                                     const uint8_t* __restrict__ st1, int s, int lane) {
  const int x_lo = stage_x_lo_c1(s, P.pad);
  if constexpr (INTERIOR) {
    const int o = s - P.pad - x_lo + lane;
#pragma unroll
    for (int i = 0; i < kR; ++i) c[i] = u8x2_to_f32(st0[o + kR * i], st1[o + kR * i]);
  } else {
#pragma unroll
    for (int i = 0; i < kR; ++i) {
      const int x = DIL == 1 ? reflect_src(s + lane + kR * i, P.pad, P.T) : reflect_src_dil<DIL>(s + lane + kR * i, P.padpx, P.Tpx);
      const int o = x < 0 ? 0 : x - x_lo;
      const float2 v = u8x2_to_f32(st0[o], st1[o]);
      c[i] = (x < 0) ? make_float2(0.f, 0.f) : v;
    }
  }
}

// lane l, register j hold 1024 (y_r[n] + i y_{r+1}[n]), n = (32 - l) + 32 (31 - j): two scalar stores per sample, one
// into each row's slot (32 lanes = 128 contiguous bytes of a panel row). second = false: row r + 1 does not exist.
// qa0 / qa1: the rows' Nyquist terms; DIL == 1 uses .x, DIL == 3 one component per channel — byte b = 3 x + ch gets
// (-1)^(padpx + x) qa[ch], and b mod 6 (which fixes both ch and the parity of x) advances by 32 = 2 mod 6 from one
// register to the next, so three add vectors serve the whole loop.
template <int DIL = 1>
__device__ __forceinline__ void store_block_c1(const float2 (&c)[kR], const RowParams& P, const RowDst& d0, const RowDst& d1,
                                               float4 qa0, float4 qa1, int s, int lane, bool second) {
  static_assert(DIL == 1 || DIL == 3, "one channel, or three interleaved");
  const int l1 = kR - lane;
  const int x0 = s + l1 - P.pad;  // float column x = x0 + 32 i: panel (x0 >> 5) + i, float x0 & 31
  float2 add[DIL];
  if constexpr (DIL == 1) {
    const float sg = ((s + l1) & 1) ? -1.0f : 1.0f;
    add[0] = make_float2(qa0.x * sg, qa1.x * sg);
  } else {
    int m0 = x0 % 6;
    if (m0 < 0) m0 += 6;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      int m = m0 + 2 * k;
      if (m >= 6) m -= 6;
      const int ch = m >= 3 ? m - 3 : m;
      const bool neg = ((m >= 3) != ((P.padpx & 1) != 0));
      const float a0 = ch == 0 ? qa0.x : (ch == 1 ? qa0.y : qa0.z);
      const float a1 = ch == 0 ? qa1.x : (ch == 1 ? qa1.y : qa1.z);
      add[k] = make_float2(neg ? -a0 : a0, neg ? -a1 : a1);
    }
  }
  const int lo = P.pad - l1;
  const int hi = min(P.pad + P.L, P.T - s + P.pad) - l1;
  const long long off = (long long)(x0 >> 5) * d0.gstride + (x0 & 31);
  float* p0 = d0.base + off;
  float* p1 = d1.base + off;
#pragma unroll
  for (int i = 0; i < kR; ++i) {
    if (kR * i >= lo && kR * i < hi) {
      const float2 v = pk_add(c[kR - 1 - i], add[DIL == 1 ? 0 : i % 3]);
      *p0 = v.x;
      if (d0.dl) p0[d0.dl] = v.x;
      if (d0.dr) p0[d0.dr] = v.x;
      if (second) {
        *p1 = v.y;
        if (d1.dl) p1[d1.dl] = v.y;
        if (d1.dr) p1[d1.dr] = v.y;
      }
    }
    p0 += d0.gstride;
    p1 += d0.gstride;
  }
  if (DIL == 1 && P.pad == 0 && lane == 0 && s < P.T) {  // a one-tap window: block index 0 sits at "index 1024" of lane 0
    const float sg0 = (s & 1) ? -1.0f : 1.0f;
    const long long o0 = (long long)(s >> 5) * d0.gstride + (s & 31);
    d0.base[o0] = c[0].x + qa0.x * sg0;
    if (second) d1.base[o0] = c[0].y + qa1.x * sg0;
  }
}

// one lane's share of the alternating sums of a row of three interleaved channels (16 bytes per load, rows 16-byte
// aligned and a multiple of 16 bytes long): a[ch] += weight(x) * byte(3 x + ch)
__device__ __forceinline__ void altsum_lane_c3(const uint8_t* __restrict__ row, int Tpx, int padpx, int lane, int (&a)[3]) {
  const uint4* row4 = reinterpret_cast<const uint4*>(row);
  const int n16 = (3 * Tpx) >> 4;
  for (int g = lane; g < n16; g += 32) {
    const uint4 v = GBLUR_OLS_LDG(row4 + g);
    const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
    int x = (16 * g) / 3, ch = 16 * g - 3 * x;
    int m = altsum_weight(x, padpx, Tpx);
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const int byte = (int)((w4[k >> 2] >> (8 * (k & 3))) & 255u);
      if (ch == 0) a[0] += m * byte;
      if (ch == 1) a[1] += m * byte;
      if (ch == 2) a[2] += m * byte;
      if (++ch == 3) {
        ch = 0;
        ++x;
        m = altsum_weight(x < Tpx ? x : Tpx - 1, padpx, Tpx);
      }
    }
  }
}

// one lane's share of a one-channel row's alternating sum: 16 pixels per load; groups without a mirrored pixel are
// four DP4A with the byte weights (+1, -1, +1, -1)
__device__ __forceinline__ int altsum_lane_c1(const uint8_t* __restrict__ row, int T, int pad, int lane) {
  int a = 0;
  int x_scalar = 0;
  if ((reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    const uint4* row4 = reinterpret_cast<const uint4*>(row);
    const int n16 = T >> 4;
    for (int g = lane; g < n16; g += 32) {
      const uint4 v = GBLUR_OLS_LDG(row4 + g);
      const int x = 16 * g;
      if (x > pad && x + 15 < T - 1 - pad) {
        const int sgn = ((pad + x) & 1) ? -1 : 1;
        int t = 0;
#if defined(__CUDA_ARCH__)
        // unsigned pixel bytes times signed weights
        asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(t) : "r"(v.x), "r"(0xFF01FF01));
        asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(t) : "r"(v.y), "r"(0xFF01FF01));
        asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(t) : "r"(v.z), "r"(0xFF01FF01));
        asm("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(t) : "r"(v.w), "r"(0xFF01FF01));
#else
        const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
        for (int k = 0; k < 16; ++k) t += ((k & 1) ? -1 : 1) * (int)((w4[k >> 2] >> (8 * (k & 3))) & 255u);
#endif
        a += sgn * t;
      } else {
        const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
        for (int k = 0; k < 16; ++k) a += altsum_weight(x + k, pad, T) * (int)((w4[k >> 2] >> (8 * (k & 3))) & 255u);
      }
    }
    x_scalar = 16 * n16;
  }
  for (int x = x_scalar + lane; x < T; x += 32) a += altsum_weight(x, pad, T) * (int)row[x];
  return a;
}

// 16 bytes of the twiddle table. On the device this is an ordered (volatile) shared-memory load: left to itself the
// compiler sinks every table load to its first use, two instructions before the dependent multiply, and the warp then
// eats the full shared-memory latency sixteen times per exchange (ncu: the hottest short-scoreboard stalls of the kernel)
__device__ __forceinline__ float4 tw_load(const float4* p) {
#if defined(__CUDA_ARCH__)
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"((uint32_t)__cvta_generic_to_shared(p))
               : "memory");
  return v;
#else
  return *p;
#endif
}

// twiddle w^(lane * k) on register k, then column `lane` of the exchange; twl = this lane's row of the table.
// The table loads run kTwAhead entries ahead of their use (registers freed by the stores pay for them).
constexpr int kTwAhead = 4;
__device__ __forceinline__ void tw_write_col(const float2 (&c)[kR], int lane, float2* __restrict__ ex,
                                             const float2* __restrict__ twl) {
  const float4* t4 = reinterpret_cast<const float4*>(twl);
  float4 t[kR / 2];
#pragma unroll
  for (int m = 0; m < kTwAhead; ++m) t[m] = tw_load(t4 + m);
#pragma unroll
  for (int m = 0; m < kR / 2; ++m) {
    if (m + kTwAhead < kR / 2) t[m + kTwAhead] = tw_load(t4 + m + kTwAhead);  // twiddles of registers 2m, 2m + 1
    ex[(2 * m) * kRowStride + lane] = m == 0 ? c[0] : cmul(c[2 * m], make_float2(t[m].x, t[m].y));
    ex[(2 * m + 1) * kRowStride + lane] = cmul(c[2 * m + 1], make_float2(t[m].z, t[m].w));
  }
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

// c[k2] = A[lane + 32 k2] * gain
__device__ __forceinline__ void mid_gains(float2 (&c)[kR], int lane, const float* __restrict__ gains) {
  const float4* g = reinterpret_cast<const float4*>(gains + lane * kGainStride);
#pragma unroll
  for (int m = 0; m < kR / 4; ++m) {
    const float4 t = g[m];
    c[4 * m] = cscale(c[4 * m], t.x);
    c[4 * m + 1] = cscale(c[4 * m + 1], t.y);
    c[4 * m + 2] = cscale(c[4 * m + 2], t.z);
    c[4 * m + 3] = cscale(c[4 * m + 3], t.w);
  }
}

// The four passes. In: c[i] = a[lane + 32 i]. Out: c[j] = 1024 y[(32 - lane) + 32 (31 - j) mod 1024] (the 1/1024 sits in
// the gains), y = the circular convolution of the block with the centred taps.
// after_exchanges(): called once, when this warp has made its last access to the exchange buffer (the column kernel
// hands the buffer back to the TMA loads of the next item there, one 32-point DFT before the end of the item).
struct NoHook {
  __device__ __forceinline__ void operator()() const {}
};
template <class Hook = NoHook>
__device__ __forceinline__ void block_passes(float2 (&c)[kR], int lane, float2* __restrict__ ex,
                                             const float2* __restrict__ twl, const float* __restrict__ gains,
                                             Hook after_exchanges = Hook()) {
#pragma unroll 1
  for (int q = 0; q < 4; ++q) {
    dft<kR, false>(c);
    if ((q & 1) == 0) {
      tw_write_col(c, lane, ex, twl);
#if defined(__CUDA_ARCH__)
      __syncwarp();
#endif
      read_row(c, lane, ex);
#if defined(__CUDA_ARCH__)
      __syncwarp();  // every row has been read before the next pass overwrites columns
#endif
      if (q == 2) after_exchanges();
    } else if (q == 1) {
      mid_gains(c, lane, gains);
    }
  }
}

// block index of (lane, descending register position i): n = (32 - lane) + 32 i, value in c[31 - i]
__device__ __forceinline__ void store_block(const float2 (&c)[kR], const RowParams& P, const RowDst& d, float2 qa, int s,
                                            int pair, int lane) {
  const int l1 = kR - lane;  // 1 .. 32
  // (-1)^n for n = s + l1 + 32 i does not depend on i
  const float sg = ((s + l1) & 1) ? -1.0f : 1.0f;
  const float2 add = make_float2(qa.x * sg, qa.y * sg);
  // outputs are the block indices [pad, min(pad + L, T - s + pad)); index 1024 (lane 0, i = 31) is index 0
  const int lo = P.pad - l1;
  const int hi = min(P.pad + P.L, P.T - s + P.pad) - l1;
  // pixel x = x0 + 32 i lands in panel (x >> 3) = (x0 >> 3) + 4 i at float offset (x0 & 7) * 4 (+ 2 for pair 1), or,
  // pair-major (the overlap-save column pass reads it): at pair * 16 + (x0 & 7) * 2 — eight lanes then fill two whole
  // 32-byte sectors, where the pixel-major order leaves every sector half written until the other pair's warp comes
  // by (measured: 3x the DRAM reads and 1.5x the writes once the two pairs are stored 10 us apart)
  const int x0 = s + l1 - P.pad;
  float* dp = d.base + (long long)(x0 >> 3) * d.gstride +
              (P.pair_major ? 16 * pair + ((x0 & 7) << 1) : 2 * pair + ((x0 & 7) << 2));
  const long long step = 4 * d.gstride;
  if ((d.dl | d.dr) == 0) {
#pragma unroll
    for (int i = 0; i < kR; ++i) {
      if (kR * i >= lo && kR * i < hi) *reinterpret_cast<float2*>(dp) = pk_add(c[kR - 1 - i], add);
      dp += step;
    }
  } else {  // rows inside the reflect_101 border also fill the column pass's mirrored slots
#pragma unroll
    for (int i = 0; i < kR; ++i) {
      if (kR * i >= lo && kR * i < hi) {
        const float2 v = pk_add(c[kR - 1 - i], add);
        *reinterpret_cast<float2*>(dp) = v;
        if (d.dl) *reinterpret_cast<float2*>(dp + d.dl) = v;
        if (d.dr) *reinterpret_cast<float2*>(dp + d.dr) = v;
      }
      dp += step;
    }
  }
  // pad == 0 (a one-tap window): block index 0 is an output and sits at "index 1024" of lane 0
  if (P.pad == 0 && lane == 0 && s < P.T) {
    float* d0 = d.base + (long long)(s >> 3) * d.gstride + (P.pair_major ? 16 * pair + ((s & 7) << 1) : 2 * pair + ((s & 7) << 2));
    const float sg0 = (s & 1) ? -1.0f : 1.0f;
    *reinterpret_cast<float2*>(d0) = make_float2(c[0].x + qa.x * sg0, c[0].y + qa.y * sg0);
  }
}

// ------------------------------------------------------------------------------------------------------------
// Column pass: fp32 panel intermediate -> u8 image (replaces the second process_channel_tiles call and
// interleave_channels, /root/reference/src/gaussianblur.cpp:261-305, helpers.hpp:248-274).
// One CTA of kColWarps = 8 warps takes FOUR PIXELS — 16 float columns, 2 x 32 bytes per slot — and one 1024-slot
// block of the padded column; warp w transforms one channel pair of one pixel with the same four passes as the row
// pass.
//   * in : eight TMA tile loads {8 floats, 256 slots} with the 32-byte swizzle land the block in the CTA's exchange
//          region (no LSU instruction, no register staging); a lane reads its 32 samples back with LDS.64 (the swizzle
//          leaves a 2-way bank conflict, 4 wavefronts per load instead of 2).
//   * out: the u8 results (F2IP.U8.F32.TRUNC: truncate, saturate and pack two channels in one instruction, the + 0.5
//          rides on the Nyquist add) are collected in a [row][16 bytes] tile and leave through TMA tile stores —
//          16 bytes per image row and CTA, no LSU instruction on the global side.
// The reference's Nyquist term along a column needs sum_n (-1)^n I[n][w] over the padded column of the row-pass
// result I. By linearity that is the row pass applied to ONE synthetic row u = sum_r weight(r) x[r] (exact integers
// on u8 input, colsum_lane below), so it costs one more line instead of a pass over the 531 MB intermediate.
// ------------------------------------------------------------------------------------------------------------
constexpr int kColWarps = 8;
constexpr int kColFloats = 2 * kColWarps;          // float columns per unit = bytes per output row of a unit (4 RGBA pixels)
constexpr int kColBoxSlots = 256;                  // slots per TMA load box
constexpr int kColBoxFloats = 8;                   // floats per slot and box: one pair of the unit's four pixels (32 bytes)
constexpr int kColRegion = kColWarps * kExch * 8;  // bytes: exchange slices of the 8 warps; the landing area (64 KB) aliases it
constexpr int kColTileBytes = kNB * kColFloats;    // u8 tile [1024 rows][16 bytes]

struct ColParams {
  int T, pad;               // image rows, reflect pad
  int L, nblk;              // outputs per block, blocks per column
  int unit0, nunits;        // first unit and unit count of this launch (a unit = 4 pixels = 16 float columns)
  int frames;
  long long nitems;         // frames * nunits * nblk
  int panels_per_frame;
  int st_na, st_rows_a, st_rows_b;  // store boxes: st_na boxes of st_rows_a rows, then one of st_rows_b rows (0 = none)
  int sign_origin;          // parity offset of the block positions (halo-sharded runs: the buffer's first slot), else 0
  long long qc_frame_stride;        // float2 per frame of qc
  const uint8_t* alpha_src; // RGBA with alpha passed through: the image the alpha bytes come from (else null)
  long long alpha_pitch;
  const float2* qc;         // per column pair 2x + pair: (K[0] - K[N/2]) / N of the column length * row pass of u
  const float2* tw;
  const float* gains;
};

// The panel rows are pair-major (store_block): floats [16 g + 2 k, + 2) belong to pair g of pixel k. Warp w takes
// pair g = w >> 2 of the unit's pixel k = w & 3. The landing area holds, per pair, 1024 slots of 32 bytes (the four
// pixels of the unit) under the 32-byte TMA swizzle: bit 4 of the offset (which 16-byte half) is XORed with bit 7
// (= bit 2 of the slot index).
__device__ __forceinline__ int col_landing_offset(int n, int w) {
  const int g = w >> 2, k = w & 3;
  return g * (kNB * 32) + n * 32 + ((((k >> 1) ^ (n >> 2)) & 1) << 4) + ((k & 1) << 3);
}

// c[i] = sample lane + 32 i of warp w's column pair
__device__ __forceinline__ void read_col(float2 (&c)[kR], int lane, int w, const unsigned char* __restrict__ landing) {
  const unsigned char* p = landing + col_landing_offset(lane, w);
#pragma unroll
  for (int i = 0; i < kR; ++i) c[i] = *reinterpret_cast<const float2*>(p + i * (kR * 32));  // 32 slots further: + 1024 bytes
}

// Natural column order (any channel count with C == Cp; used for one-channel images): the unit's 16 float columns are
// 64 contiguous bytes of the panel row, warp w takes floats 2w, 2w + 1, and the landing area holds 1024 slots of 64
// bytes under the 64-byte TMA swizzle: the 16-byte chunk index (w >> 1) is XORed with bits 7-8 of the offset.
__device__ __forceinline__ int col_landing_offset_nat(int n, int w) {
  return n * 64 + ((((w >> 1) ^ (n >> 1)) & 3) << 4) + ((w & 1) << 3);
}
__device__ __forceinline__ void read_col_nat(float2 (&c)[kR], int lane, int w, const unsigned char* __restrict__ landing) {
  const unsigned char* p = landing + col_landing_offset_nat(lane, w);
#pragma unroll
  for (int i = 0; i < kR; ++i) c[i] = *reinterpret_cast<const float2*>(p + i * (kR * 64));
}

// two channels -> two bytes, each uint8_t(v) (truncation: helpers.hpp:251-252,270; the caller has added 0.5), saturated
__device__ __forceinline__ uint32_t to_u8x2(float2 v) {
#if defined(__CUDA_ARCH__)
  const int a = __float2int_rz(v.x), b = __float2int_rz(v.y);
  uint32_t r;
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(b), "r"(a), "r"(0));  // fuses to F2IP.U8.F32.TRUNC
  return r;
#else
  const float x = std::min(std::max(v.x, 0.0f), 255.0f), y = std::min(std::max(v.y, 0.0f), 255.0f);
  return (uint32_t)x | ((uint32_t)y << 8);
#endif
}

// lane l, register j hold 1024 y[n], n = (32 - l) + 32 (31 - j): two bytes per output row into tile[n - pad][2w, 2w+1]
// wb: byte offset of the warp's two outputs in the 16-byte tile row (pair-major RGBA: 4 (w & 3) + 2 (w >> 2): pixel
// w & 3, channels 2g, 2g + 1; natural order: 2 w)
// keep_hi: write only the first of the two bytes — the second one (alpha of an RGBA image whose alpha is passed through)
// is already in the tile, put there by a TMA load of the input image
__device__ __forceinline__ void col_to_tile(const float2 (&c)[kR], float2 qc, int s, int sign_origin, int pad, int L, int wb,
                                            int lane, unsigned char* __restrict__ tile, bool keep_hi = false) {
  const int l1 = kR - lane;
  const float sg = ((s + sign_origin + l1) & 1) ? -1.0f : 1.0f;
  const float2 add = make_float2(fmaf(qc.x, sg, 0.5f), fmaf(qc.y, sg, 0.5f));
  const int lo = pad - l1, hi = pad + L - l1;
  unsigned char* tp = tile + (l1 - pad) * kColFloats + wb;
#pragma unroll
  for (int i = 0; i < kR; ++i) {
    if (kR * i >= lo && kR * i < hi) {
      const uint32_t v = to_u8x2(pk_add(c[kR - 1 - i], add));
      if (keep_hi)
        tp[i * (kR * kColFloats)] = (unsigned char)v;
      else
        *reinterpret_cast<uint16_t*>(tp + i * (kR * kColFloats)) = (uint16_t)v;
    }
  }
  if (pad == 0 && lane == 0) {
    const float sg0 = ((s + sign_origin) & 1) ? -1.0f : 1.0f;
    const uint32_t v = to_u8x2(make_float2(c[0].x + fmaf(qc.x, sg0, 0.5f), c[0].y + fmaf(qc.y, sg0, 0.5f)));
    if (keep_hi)
      tile[wb] = (unsigned char)v;
    else
      *reinterpret_cast<uint16_t*>(tile + wb) = (uint16_t)v;
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

// the same for 16 consecutive byte columns (one 16-byte load per image row, four rows in flight): the column-sum kernel
// is a pure stream over the u8 input, so bytes in flight per thread are what sets its speed
__device__ __forceinline__ void colsum_lane16(const uint8_t* __restrict__ in, long long in_pitch, int x16, int r0, int r1,
                                              int pad, int rows, int (&a)[16]) {
  const uint8_t* p = in + (long long)r0 * in_pitch + 16LL * x16;
  auto acc = [&](const uint4& v, int r) {
    const int m = altsum_weight(r, pad, rows);
    const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 16; ++k) a[k] += m * (int)((w4[k >> 2] >> (8 * (k & 3))) & 255u);
  };
  int r = r0;
  for (; r + 4 <= r1; r += 4) {
    const uint4 v0 = GBLUR_OLS_LDG(reinterpret_cast<const uint4*>(p));
    const uint4 v1 = GBLUR_OLS_LDG(reinterpret_cast<const uint4*>(p + in_pitch));
    const uint4 v2 = GBLUR_OLS_LDG(reinterpret_cast<const uint4*>(p + 2 * in_pitch));
    const uint4 v3 = GBLUR_OLS_LDG(reinterpret_cast<const uint4*>(p + 3 * in_pitch));
    acc(v0, r);
    acc(v1, r + 1);
    acc(v2, r + 2);
    acc(v3, r + 3);
    p += 4 * in_pitch;
  }
  for (; r < r1; ++r) {
    acc(GBLUR_OLS_LDG(reinterpret_cast<const uint4*>(p)), r);
    p += in_pitch;
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

// lane l, register j hold 1024 * (linear convolution of u) at block index (32 - l) + 32 (31 - j): add u's own Nyquist
// term, scale by the column coefficient and store per column pair (2x + pair)
__device__ __forceinline__ void store_block_u(const float2 (&c)[kR], float2 qa_u, float q_col, int T, int pad, int L, int s,
                                              int pair, int lane, float2* __restrict__ qc) {
  const int l1 = kR - lane;
  const float sg = ((s + l1) & 1) ? -1.0f : 1.0f;
  const int hi = min(pad + L, T - s + pad);
#pragma unroll
  for (int i = 0; i < kR; ++i) {
    const int idx = l1 + kR * i;
    if (idx >= pad && idx < hi) {
      const int x = s + idx - pad;
      qc[2 * x + pair] = make_float2(q_col * (c[kR - 1 - i].x + qa_u.x * sg), q_col * (c[kR - 1 - i].y + qa_u.y * sg));
    }
  }
  if (pad == 0 && lane == 0 && s < T) {
    const float sg0 = (s & 1) ? -1.0f : 1.0f;
    qc[2 * s + pair] = make_float2(q_col * (c[0].x + qa_u.x * sg0), q_col * (c[0].y + qa_u.y * sg0));
  }
}

}  // namespace ols
}  // namespace gblur
