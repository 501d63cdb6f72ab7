// ols_kernels.cu — overlap-save passes for sm_100a: one warp per 1024-sample block of one complex sequence.
// See ols_core.cuh for the algorithm and its relation to the reference (src/gaussianblur.cpp:62-79,202-244).
//
// Why: the whole-line kernels (line_kernels.cuh) are issue/latency bound with 13 warps per SM and a CTA
// barrier per stage (profiles/r01/ncu_full_8k_v3_summary.txt: 2.2 of 4 warp instructions per clock, 17 % of
// the stall samples on the barrier). Here a transform never leaves its warp: 32 points per lane, two
// radix-32 stages, one shared-memory transpose per direction ordered by __syncwarp only, 8.7 KB of shared
// memory per warp. The row pass has no CTA-wide synchronisation in its loop at all; the column pass
// synchronises its 8 warps three times per item around the TMA transfers they share.
#include <cstdlib>
#include <mutex>

#include "async_copy.cuh"
#include "ols.h"
#include "ols_core.cuh"
#include "ols_host.hpp"

namespace gblur {

void count_launch();  // kernels.cu

namespace ols {

struct RowPlan {
  float2* d_tw = nullptr;
  float* d_gains = nullptr;
  float q = 0.f;
  int ksize = 0;
  int dil = 1;              // 3: the block gains belong to taps spread three bytes apart (interleaved RGB rows)
  float4* d_qa = nullptr;
  long long qa_rows = 0;
  // column pass
  float q_col = 0.f;
  int* d_u = nullptr;       // synthetic row u: 4 * cols integer sums
  float2* d_qc = nullptr;   // 2 * cols column-pair terms
  int u_cols = 0;
  // tensor maps of the column pass, rebuilt when the buffers or the geometry change
  CUtensorMap in_map, out_a, out_b, alpha_a, alpha_b;
  const void* map_inter = nullptr;
  const void* map_out = nullptr;
  const void* map_alpha = nullptr;
  long long map_key[6] = {0, 0, 0, 0, 0, 0};
};

constexpr size_t kTableBytes = (size_t)kTwSlots * 8 + (size_t)kR * kGainStride * 4;

__device__ __forceinline__ void load_tables(float2* tw, float* gains, const float2* g_tw, const float* g_gains, int tid,
                                            int nthreads) {
  for (int i = tid; i < kTwSlots; i += nthreads) tw[i] = __ldg(g_tw + i);
  for (int i = tid; i < kR * kGainStride; i += nthreads) gains[i] = __ldg(g_gains + i);
}

// (K[0] - K[N/2]) / N * sum over the reflect_101-padded row of (-1)^n b[n], per channel; exact integer sums.
// One warp per row, 16-byte loads (four in flight per lane) when the row allows them.
constexpr int kAltsumWarps = 8;

__global__ void __launch_bounds__(kAltsumWarps * 32)
    altsum_rows_kernel(const uint8_t* __restrict__ in, long long in_pitch, int T, int pad, float q, long long rows_total,
                       float4* __restrict__ qa, long long split_at, long long split_gap) {
  const int lane = threadIdx.x & 31;
  const long long r = (long long)blockIdx.x * kAltsumWarps + (threadIdx.x >> 5);
  if (r >= rows_total) return;
  const uint32_t* row = reinterpret_cast<const uint32_t*>(in + (r + (r >= split_at ? split_gap : 0)) * in_pitch);
  int a[4] = {0, 0, 0, 0};
  altsum_lane(row, T, pad, lane, a);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int c = 0; c < 4; ++c) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
  }
  if (lane == 0) qa[r] = make_float4(q * (float)a[0], q * (float)a[1], q * (float)a[2], q * (float)a[3]);
}

template <int WARPS, int MAXREG>
__global__ void __maxnreg__(MAXREG) ols_row_rgba_kernel(const __grid_constant__ RowParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);
  float* gains = reinterpret_cast<float*>(tw + kTwSlots);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* ex = reinterpret_cast<float2*>(gains + kR * kGainStride) + warp * kExch;
  load_tables(tw, gains, P.tw, P.gains, threadIdx.x, WARPS * 32);
  __syncthreads();
  const float2* twl = tw + lane * kTwStride;
  const long long stride = (long long)gridDim.x * WARPS;
  for (long long it = (long long)blockIdx.x * WARPS + warp; it < P.nitems; it += stride) {
    const int pair = (int)(it & 1);
    const long long bj = it >> 1;
    const long long r = bj / P.nblk;
    const int j = (int)(bj - r * P.nblk);
    const int s = j * P.L;
    const long long rphys = phys_row(P, r);
    const uint32_t* row = reinterpret_cast<const uint32_t*>(P.in + rphys * P.in_pitch);
    float2 c[kR];
    if (s - P.pad >= 0 && s - P.pad + kNB <= P.T)
      load_block<true>(c, P, row, s, pair, lane);
    else
      load_block<false>(c, P, row, s, pair, lane);
    block_passes(c, lane, ex, twl, gains);
    const RowDst d = make_row_dst(P, rphys);
    const float4 q4 = __ldg(P.qa + r);
    store_block(c, P, d, pair ? make_float2(q4.z, q4.w) : make_float2(q4.x, q4.y), s, pair, lane);
  }
}

// The same pass with the input staged: one item = one block of one row, BOTH channel pairs, and the block's 4 KB of
// pixels arrive in the warp's staging buffer by a bulk asynchronous copy (UBLKCP) issued one item ahead — as soon as
// the second pair has been read out of the buffer — so a warp never waits for a global load between two items
// (ncu of the kernel above: 16 % of its stall samples sit on the first use of the loaded pixels).
// Needs 16-byte aligned rows of a multiple of four pixels. 16 warps per SM (12.8 KB of shared memory per warp).
template <int WARPS, int MAXREG>
__global__ void __maxnreg__(MAXREG) ols_row_rgba_staged_kernel(const __grid_constant__ RowParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);
  float* gains = reinterpret_cast<float*>(tw + kTwSlots);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* ex = reinterpret_cast<float2*>(gains + kR * kGainStride) + warp * kExch;
  uint32_t* stage = reinterpret_cast<uint32_t*>(reinterpret_cast<float2*>(gains + kR * kGainStride) + WARPS * kExch) +
                    warp * kStagePx;
  uint64_t* bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint32_t*>(reinterpret_cast<float2*>(gains + kR * kGainStride) +
                                                                            WARPS * kExch) +
                                              WARPS * kStagePx) +
                  warp;
  load_tables(tw, gains, P.tw, P.gains, threadIdx.x, WARPS * 32);
  if (lane == 0) {
    mbar_init(bar, 1);
    fence_proxy_async();
  }
  __syncthreads();
  const float2* twl = tw + lane * kTwStride;
  const int nblocks = (int)(P.nitems >> 1);
  const int stride = (int)gridDim.x * WARPS;
  auto issue = [&](int b) {  // one lane: the pixels the block reads, clipped to the row
    const int r = b / P.nblk, s = (b - r * P.nblk) * P.L;
    const int x_lo = stage_x_lo(s, P.pad);
    int x_hi = s - P.pad + kNB;
    x_hi = ((x_hi < P.T ? x_hi : P.T) + 3) & ~3;
    const uint32_t bytes = (uint32_t)(x_hi - x_lo) * 4u;
    mbar_expect_tx(bar, bytes);
    bulk_g2s(stage, P.in + phys_row(P, r) * P.in_pitch + (long long)x_lo * 4, bytes, bar);
  };
  int b = (int)blockIdx.x * WARPS + warp;
  if (lane == 0 && b < nblocks) issue(b);
  uint32_t parity = 0;
  for (; b < nblocks; b += stride) {
    const int r = b / P.nblk;
    const int s = (b - r * P.nblk) * P.L;
    const bool interior = s - P.pad >= 0 && s - P.pad + kNB <= P.T;
    const float4 q4 = __ldg(P.qa + r);
    const RowDst d = make_row_dst(P, phys_row(P, r));
    mbar_wait(bar, parity);
    parity ^= 1;
    // both channel pairs through ONE copy of the code (rolled: the item's instructions must stay inside the SM's
    // instruction cache, see ols_core.cuh; the first staged version inlined everything twice and ran 25 % slower)
#pragma unroll 1
    for (int pair = 0; pair < 2; ++pair) {
      float2 c[kR];
      if (interior)
        load_block_staged<true>(c, P, stage, s, pair, lane);
      else
        load_block_staged<false>(c, P, stage, s, pair, lane);
      if (pair == 1) {
        __syncwarp();  // every lane has read the staging buffer for the last time
        if (lane == 0 && b + stride < nblocks) issue(b + stride);
      }
      block_passes(c, lane, ex, twl, gains);
      store_block(c, P, d, pair ? make_float2(q4.z, q4.w) : make_float2(q4.x, q4.y), s, pair, lane);
    }
  }
}

// ---- one-channel images (BASELINE.json config 4) ----
__global__ void __launch_bounds__(kAltsumWarps * 32)
    altsum_rows_c1_kernel(const uint8_t* __restrict__ in, long long in_pitch, int T, int pad, float q, long long rows_total,
                          float* __restrict__ qa, long long split_at, long long split_gap) {
  const int lane = threadIdx.x & 31;
  const long long r = (long long)blockIdx.x * kAltsumWarps + (threadIdx.x >> 5);
  if (r >= rows_total) return;
  int a = altsum_lane_c1(in + (r + (r >= split_at ? split_gap : 0)) * in_pitch, T, pad, lane);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  if (lane == 0) qa[r] = q * (float)a;
}

// rows of three interleaved channels (RGB): per row (q * sum_ch) as a float4 (w unused)
__global__ void __launch_bounds__(kAltsumWarps * 32)
    altsum_rows_c3_kernel(const uint8_t* __restrict__ in, long long in_pitch, int Tpx, int padpx, float q, long long rows_total,
                          float4* __restrict__ qa, long long split_at, long long split_gap) {
  const int lane = threadIdx.x & 31;
  const long long r = (long long)blockIdx.x * kAltsumWarps + (threadIdx.x >> 5);
  if (r >= rows_total) return;
  int a[3] = {0, 0, 0};
  altsum_lane_c3(in + (r + (r >= split_at ? split_gap : 0)) * in_pitch, Tpx, padpx, lane, a);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a[0] += __shfl_xor_sync(0xffffffffu, a[0], o);
    a[1] += __shfl_xor_sync(0xffffffffu, a[1], o);
    a[2] += __shfl_xor_sync(0xffffffffu, a[2], o);
  }
  if (lane == 0) qa[r] = make_float4(q * (float)a[0], q * (float)a[1], q * (float)a[2], 0.f);
}

// One item = one block of a PAIR of image rows (row r is the real part of the complex sequence, row r + 1 the imaginary
// part); both rows' bytes are staged by bulk copies one item ahead, as in the RGBA kernel above.
// DIL == 3: the rows hold three interleaved channels and run as byte rows (P.T, P.pad, P.L in bytes; ols_core.cuh
// reflect_src_dil): the block gains are those of the taps spread three bytes apart, the Nyquist term is per channel.
template <int WARPS, int MAXREG, int DIL>
__global__ void __maxnreg__(MAXREG) ols_row_c1_kernel(const __grid_constant__ RowParams P, long long rows_total) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);
  float* gains = reinterpret_cast<float*>(tw + kTwSlots);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* ex0 = reinterpret_cast<float2*>(gains + kR * kGainStride);
  float2* ex = ex0 + warp * kExch;
  uint8_t* stage_all = reinterpret_cast<uint8_t*>(ex0 + WARPS * kExch);
  uint8_t* st0 = stage_all + warp * (2 * kStageC1);
  uint8_t* st1 = st0 + kStageC1;
  uint64_t* bar = reinterpret_cast<uint64_t*>(stage_all + WARPS * (2 * kStageC1)) + warp;
  load_tables(tw, gains, P.tw, P.gains, threadIdx.x, WARPS * 32);
  if (lane == 0) {
    mbar_init(bar, 1);
    fence_proxy_async();
  }
  __syncthreads();
  const float2* twl = tw + lane * kTwStride;
  const float* qa = reinterpret_cast<const float*>(P.qa);
  // 32-bit item arithmetic (launch_blocks_c1 refuses more than 2^31 items: a 64-bit division is ~70 instructions)
  const unsigned npairs = (unsigned)((rows_total + 1) >> 1);
  const unsigned nitems = npairs * (unsigned)P.nblk;
  const unsigned stride = gridDim.x * WARPS;
  auto issue = [&](unsigned b) {  // one lane: the bytes the block reads from both rows, clipped to the row
    const unsigned rp = b / (unsigned)P.nblk;
    const int s = (int)(b - rp * (unsigned)P.nblk) * P.L;
    const int x_lo = stage_x_lo_c1(s, P.pad);
    int x_hi = s - P.pad + kNB;
    x_hi = ((x_hi < P.T ? x_hi : P.T) + 15) & ~15;
    const uint32_t bytes = (uint32_t)(x_hi - x_lo);
    const bool second = 2LL * rp + 1 < rows_total;
    mbar_expect_tx(bar, second ? 2 * bytes : bytes);
    bulk_g2s(st0, P.in + phys_row(P, 2LL * rp) * P.in_pitch + x_lo, bytes, bar);
    if (second) bulk_g2s(st1, P.in + phys_row(P, 2LL * rp + 1) * P.in_pitch + x_lo, bytes, bar);
  };
  unsigned b = blockIdx.x * WARPS + warp;
  if (lane == 0 && b < nitems) issue(b);
  uint32_t parity = 0;
  for (; b < nitems; b += stride) {
    const unsigned rp = b / (unsigned)P.nblk;
    const int s = (int)(b - rp * (unsigned)P.nblk) * P.L;
    const long long r = 2LL * rp;
    const bool second = r + 1 < rows_total;
    const bool interior = s - P.pad >= 0 && s - P.pad + kNB <= P.T;
    float4 qa0 = make_float4(0.f, 0.f, 0.f, 0.f), qa1 = qa0;
    if constexpr (DIL == 1) {
      qa0.x = __ldg(qa + r);
      if (second) qa1.x = __ldg(qa + r + 1);
    } else {
      qa0 = __ldg(P.qa + r);
      if (second) qa1 = __ldg(P.qa + r + 1);
    }
    const RowDst d0 = make_row_dst(P, phys_row(P, r)), d1 = make_row_dst(P, phys_row(P, second ? r + 1 : r));
    mbar_wait(bar, parity);
    parity ^= 1;
    float2 c[kR];
    // without a second row its staging buffer holds stale bytes: the imaginary part is computed and dropped
    if (interior)
      load_block_staged_c1<true, DIL>(c, P, st0, st1, s, lane);
    else
      load_block_staged_c1<false, DIL>(c, P, st0, st1, s, lane);
    __syncwarp();  // every lane has read the staging buffers
    if (lane == 0 && b + stride < nitems) issue(b + stride);
    block_passes(c, lane, ex, twl, gains);
    store_block_c1<DIL>(c, P, d0, d1, qa0, qa1, s, lane, second);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// column pass
// ---------------------------------------------------------------------------------------------------------------
constexpr int kColsumRows = 64;  // image rows per thread of the column-sum kernel

// rows [row_begin, row_end) of an image of `rows` rows; `in` points at image row row_base
__global__ void __launch_bounds__(128) colsum_kernel(const uint8_t* __restrict__ in, long long in_pitch, int cols, int rows,
                                                     int pad, int* __restrict__ u, int row_begin, int row_end, int row_base) {
  const int x = blockIdx.x * 128 + threadIdx.x;
  if (x >= cols) return;
  const int r0 = row_begin + blockIdx.y * kColsumRows;
  const int r1 = min(row_end, r0 + kColsumRows);
  int a[4] = {0, 0, 0, 0};
  colsum_lane(in - (long long)row_base * in_pitch, in_pitch, x, r0, r1, pad, rows, a);
#pragma unroll
  for (int c = 0; c < 4; ++c) atomicAdd(u + 4 * x + c, a[c]);  // integer atomics: exact, order-independent
}

// 16 byte columns per thread (rows of whole 16-byte groups, 16-byte aligned). Measured in the fused 8K step, where this
// kernel runs beside the row pass's own pre-kernel over the same input: 0.5785 -> 0.5690 ms per step. Rows per thread
// swept afterwards (profiles/r02/colsum_variants.jsonl): 16 -> 0.610, 32 -> 0.581, 64 -> 0.5645, 128 -> 0.5636,
// 256 -> 0.575 ms (the global atomics are the cost: fewer rows per thread = more of them); four row groups per CTA
// that meet in shared memory before one atomic per column did not beat 64 rows per thread (0.5687).
__global__ void __launch_bounds__(128) colsum16_kernel(const uint8_t* __restrict__ in, long long in_pitch, int groups16, int rows,
                                                       int pad, int* __restrict__ u, int row_begin, int row_end, int row_base) {
  const int x = blockIdx.x * 128 + threadIdx.x;
  if (x >= groups16) return;
  const int r0 = row_begin + blockIdx.y * kColsumRows;
  const int r1 = min(row_end, r0 + kColsumRows);
  int a[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) a[k] = 0;
  colsum_lane16(in - (long long)row_base * in_pitch, in_pitch, x, r0, r1, pad, rows, a);
#pragma unroll
  for (int k = 0; k < 16; ++k) atomicAdd(u + 16 * x + k, a[k]);
}

// The row pass of the synthetic row: a few CTAs of kUWarps warps over the nblk * 2 items (every CTA forms u's own
// alternating sum itself: T integers). Writes the column terms qc.
constexpr int kUWarps = 16;

__global__ void __launch_bounds__(kUWarps * 32, 1)
    u_rowpass_kernel(const int* __restrict__ u, int T, int pad, float q_row, float q_col, const float2* __restrict__ g_tw,
                     const float* __restrict__ g_gains, float2* __restrict__ qc) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);
  float* gains = reinterpret_cast<float*>(tw + kTwSlots);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* ex = reinterpret_cast<float2*>(gains + kR * kGainStride) + warp * kExch;
  __shared__ long long part[kUWarps][4];
  __shared__ float qa_u[4];
  load_tables(tw, gains, g_tw, g_gains, threadIdx.x, kUWarps * 32);
  // u's own alternating sum over its reflect_101-padded line (64-bit integers: exact)
  long long a[4] = {0, 0, 0, 0};
  for (int x = threadIdx.x; x < T; x += kUWarps * 32) {
    const long long m = altsum_weight(x, pad, T);
#pragma unroll
    for (int c = 0; c < 4; ++c) a[c] += m * (long long)u[4 * x + c];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int c = 0; c < 4; ++c) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
  }
  if (lane == 0) {
#pragma unroll
    for (int c = 0; c < 4; ++c) part[warp][c] = a[c];
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    long long t = 0;
    for (int w = 0; w < kUWarps; ++w) t += part[w][threadIdx.x];
    qa_u[threadIdx.x] = q_row * (float)t;
  }
  __syncthreads();
  const int L = kNB - 2 * pad;
  const int nitems = ((T + L - 1) / L) * 2;
  const float2* twl = tw + lane * kTwStride;
  for (int it = blockIdx.x * kUWarps + warp; it < nitems; it += gridDim.x * kUWarps) {
    const int pair = it & 1, s = (it >> 1) * L;
    float2 c[kR];
    load_block_u(c, u, T, pad, s, pair, lane);
    block_passes(c, lane, ex, twl, gains);
    store_block_u(c, make_float2(qa_u[2 * pair], qa_u[2 * pair + 1]), q_col, T, pad, L, s, pair, lane, qc);
  }
}

// One-channel variant: u is ONE real line (u[x], x < T); it rides as the real part of a complex sequence with a zero
// imaginary part. Writes the column terms V[x] (the block kernel reads them as float2 per column pair).
__global__ void __launch_bounds__(kUWarps * 32, 1)
    u_rowpass_c1_kernel(const int* __restrict__ u, int T, int pad, float q_row, float q_col, const float2* __restrict__ g_tw,
                        const float* __restrict__ g_gains, float* __restrict__ V) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);
  float* gains = reinterpret_cast<float*>(tw + kTwSlots);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* ex = reinterpret_cast<float2*>(gains + kR * kGainStride) + warp * kExch;
  __shared__ long long part[kUWarps];
  __shared__ float qa_u;
  load_tables(tw, gains, g_tw, g_gains, threadIdx.x, kUWarps * 32);
  long long a = 0;
  for (int x = threadIdx.x; x < T; x += kUWarps * 32) a += (long long)altsum_weight(x, pad, T) * (long long)u[x];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  if (lane == 0) part[warp] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t = 0;
    for (int w = 0; w < kUWarps; ++w) t += part[w];
    qa_u = q_row * (float)t;
  }
  __syncthreads();
  const int L = kNB - 2 * pad;
  const int nblk = (T + L - 1) / L;
  const float2* twl = tw + lane * kTwStride;
  for (int j = blockIdx.x * kUWarps + warp; j < nblk; j += gridDim.x * kUWarps) {
    const int s = j * L;
    float2 c[kR];
#pragma unroll
    for (int i = 0; i < kR; ++i) {
      const int x = reflect_src(s + lane + kR * i, pad, T);
      c[i] = make_float2(x < 0 ? 0.f : (float)u[x], 0.f);
    }
    block_passes(c, lane, ex, twl, gains);
    const int l1 = kR - lane;
    const float sg = ((s + l1) & 1) ? -1.0f : 1.0f;
    const int hi = min(pad + L, T - s + pad);
#pragma unroll
    for (int i = 0; i < kR; ++i) {
      const int idx = l1 + kR * i;
      if (idx >= pad && idx < hi) V[s + idx - pad] = q_col * (c[kR - 1 - i].x + qa_u * sg);
    }
    if (pad == 0 && lane == 0 && s < T) V[s] = q_col * (c[0].x + qa_u * ((s & 1) ? -1.0f : 1.0f));
  }
}

struct ColKernelParams {
  ColParams P;
  CUtensorMap in_map;  // panel intermediate {32 floats, np slots, panels}, box {8, 256, 1}, 32-byte swizzle
  CUtensorMap out_a;   // u8 image {row bytes, rows, frames}, box {16, st_rows_a, 1}
  CUtensorMap out_b;   // same tensor, box {16, st_rows_b, 1}
  CUtensorMap alpha_a; // the INPUT image with the boxes of out_a / out_b: RGBA with alpha passed through (P.alpha_src)
  CUtensorMap alpha_b;
};

constexpr size_t kColSmem = (size_t)kColRegion + kColTileBytes + kTableBytes + 16;

// ALPHA_KEEP (RGBA with alpha passed through) is a template parameter: as a run-time flag its branches in the epilogue
// cost the plain kernel 1.6 % (0.5667 -> 0.5760 ms per 8K step)
template <bool PAIR_MAJOR, bool ALPHA_KEEP>
__global__ void __launch_bounds__(kColWarps * 32, 2) ols_col_kernel(const __grid_constant__ ColKernelParams K) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* region = smem_raw;                 // landing area (first 64 KB), then the warps' exchange slices
  unsigned char* tile = smem_raw + kColRegion;      // u8 results [row][16 bytes]
  float2* tw = reinterpret_cast<float2*>(tile + kColTileBytes);
  float* gains = reinterpret_cast<float*>(tw + kTwSlots);
  uint64_t* bar = reinterpret_cast<uint64_t*>(gains + kR * kGainStride);
  const ColParams& P = K.P;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  float2* ex = reinterpret_cast<float2*>(region) + w * kExch;

  // consecutive items = the blocks of one unit, then the neighbouring unit, dealt out round-robin: at any moment the
  // CTAs of the grid work on neighbouring units and blocks. Two units share every 32-byte sector of the OUTPUT rows and
  // four units every 128-byte line of the panel rows, so their partners must be in flight together. (Measured: giving
  // every CTA a contiguous run of items instead — to make the second fetch of the 2 * pad slots two blocks share an L2
  // hit — took the kernel from 0.29 to 0.48 ms, DRAM reads from 612 MB to 1.39 GB and writes from 129 to 260 MB:
  // half-written output sectors are read back, and every 32-byte box read pulls a whole line.)
  // (32-bit item arithmetic: launch_col refuses more than 2^31 items; a 64-bit division is ~70 instructions a thread)
  auto decode = [&](unsigned it, int& f, int& unit, int& s) {
    const unsigned fu = it / (unsigned)P.nblk;
    s = (int)(it - fu * (unsigned)P.nblk) * P.L;
    f = (int)(fu / (unsigned)P.nunits);
    unit = P.unit0 + (int)(fu - (unsigned)f * (unsigned)P.nunits);
  };
  auto request = [&](unsigned it) {  // one thread: the four landing boxes of item `it`
    int f, unit, s;
    decode(it, f, unit, s);
    mbar_expect_tx(bar, kNB * kColFloats * 4);
    if constexpr (PAIR_MAJOR) {
#pragma unroll
      for (int g = 0; g < 2; ++g)
#pragma unroll
        for (int q = 0; q < kNB / kColBoxSlots; ++q)
          tma_load_3d(region + g * (kNB * 32) + q * (kColBoxSlots * 32), &K.in_map, g * 16 + (unit & 1) * kColBoxFloats,
                      s + q * kColBoxSlots, f * P.panels_per_frame + (unit >> 1), bar);
    } else {
#pragma unroll
      for (int q = 0; q < kNB / kColBoxSlots; ++q)
        tma_load_3d(region + q * (kColBoxSlots * 64), &K.in_map, (unit & 1) * kColFloats, s + q * kColBoxSlots,
                    f * P.panels_per_frame + (unit >> 1), bar);
    }
  };

  uint64_t* bar2 = bar + 1;  // alpha pass-through: the item's strip of the input image has landed in the tile
  constexpr bool alpha_keep = PAIR_MAJOR && ALPHA_KEEP;
  if (tid == 0) {
    mbar_init(bar, 1);
    mbar_init(bar2, 1);
    fence_proxy_async();
  }
  load_tables(tw, gains, P.tw, P.gains, tid, kColWarps * 32);
  __syncthreads();
  const unsigned nitems = (unsigned)P.nitems;
  unsigned it = blockIdx.x;
  if (tid == 0 && it < nitems) request(it);
  const float2* twl = tw + lane * kTwStride;
  uint32_t parity = 0, parity2 = 0;
  for (; it < nitems; it += gridDim.x) {
    int f, unit, s;
    decode(it, f, unit, s);
    // the item's column term: requested now, used after the last DFT (it was the kernel's only exposed global load)
    const int pairidx = PAIR_MAJOR ? (unit * 4 + (w & 3)) * 2 + (w >> 2) : unit * kColWarps + w;
    const float2 qc = P.qc ? __ldg(P.qc + f * P.qc_frame_stride + pairidx) : make_float2(0.f, 0.f);
    mbar_wait(bar, parity);
    parity ^= 1;
    float2 c[kR];
    if constexpr (PAIR_MAJOR)
      read_col(c, lane, w, region);
    else
      read_col_nat(c, lane, w, region);
    __syncthreads();  // the landing area is consumed: it becomes the exchange slices
    const unsigned next = it + gridDim.x;
    block_passes(c, lane, ex, twl, gains, [&] {
      if (tid == 0) bulk_wait_read<0>();  // the previous item's stores have read the tile
      fence_proxy_async();                // order this thread's exchange traffic before the next TMA writes
      __syncthreads();                    // exchanges done: the region is free, and so is the tile
      if (tid == 0) {
        if constexpr (alpha_keep) {
          // RGBA, alpha not blurred: this item's [rows][16 bytes] strip of the INPUT image goes into the tile (the boxes
          // of the stores below, rows past the image zero-filled); the epilogue then writes the three blurred bytes of
          // every pixel around the alpha byte. In place this reads exactly the bytes the item's own stores overwrite.
          mbar_expect_tx(bar2, (uint32_t)P.L * kColFloats);
          const int c0 = unit * kColFloats;
#pragma unroll 1
          for (int b = 0; b < P.st_na; ++b)
            tma_load_3d(tile + b * P.st_rows_a * kColFloats, &K.alpha_a, c0, s + b * P.st_rows_a, f, bar2);
          if (P.st_rows_b)
            tma_load_3d(tile + P.st_na * P.st_rows_a * kColFloats, &K.alpha_b, c0, s + P.st_na * P.st_rows_a, f, bar2);
        }
        if (next < nitems) request(next);  // lands under the last DFT and the epilogue
      }
    });
    if constexpr (alpha_keep) {
      mbar_wait(bar2, parity2);
      parity2 ^= 1;
    }
    col_to_tile(c, qc, s, P.sign_origin, P.pad, P.L, PAIR_MAJOR ? 4 * (w & 3) + 2 * (w >> 2) : 2 * w, lane, tile,
                alpha_keep && (w >> 2) == 1);
    fence_proxy_async();  // the tile writes become visible to the async proxy
    __syncthreads();
    if (tid == 0) {
      const int c0 = unit * kColFloats;
#pragma unroll 1
      for (int b = 0; b < P.st_na; ++b) tma_store_3d(&K.out_a, c0, s + b * P.st_rows_a, f, tile + b * P.st_rows_a * kColFloats);
      if (P.st_rows_b) tma_store_3d(&K.out_b, c0, s + P.st_na * P.st_rows_a, f, tile + P.st_na * P.st_rows_a * kColFloats);
      bulk_commit();
    }
  }
  if (tid == 0) bulk_wait<0>();
}

cudaError_t row_plan_create(RowPlan** out, const std::vector<float>& taps, int n_ref, int n_ref_col, int dil) {
  *out = nullptr;
  if (dil != 1 && dil != 3) return cudaErrorInvalidValue;
  // window too wide for 1024-sample blocks (spread taps: dil * pad samples on either side): not eligible
  if (((int)taps.size() - 1) * dil > 2 * kMaxPad || (dil != 1 && taps.size() < 3)) return cudaSuccess;
  const HostTables t = build_tables(taps, n_ref, n_ref_col, dil);
  RowPlan* p = new RowPlan();
  p->dil = dil;
  p->q = t.q;
  p->q_col = t.q_col;
  p->ksize = (int)taps.size();
  cudaError_t e = cudaMalloc(&p->d_tw, t.tw.size() * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&p->d_gains, t.gains.size() * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpy(p->d_tw, t.tw.data(), t.tw.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess)
    e = cudaMemcpy(p->d_gains, t.gains.data(), t.gains.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    row_plan_destroy(p);
    return e;
  }
  *out = p;
  return cudaSuccess;
}

void row_plan_destroy(RowPlan* p) {
  if (!p) return;
  if (p->d_tw) cudaFree(p->d_tw);
  if (p->d_gains) cudaFree(p->d_gains);
  if (p->d_qa) cudaFree(p->d_qa);
  if (p->d_u) cudaFree(p->d_u);
  if (p->d_qc) cudaFree(p->d_qc);
  delete p;
}

bool row_eligible(const RowPlan* p, const RowLaunch& a) {
  if (!p || !a.blk.enabled || a.npeers != 0 || a.split_cols != 0 || 2 * a.pad + 1 != p->ksize || a.pad > kMaxPad || a.T < 1 ||
      a.rows_total < 1 || (reinterpret_cast<uintptr_t>(a.inter) % 16) != 0)
    return false;
  if (a.C == 4 && a.Cp == 4)
    return a.blk.L == 4 && (a.in_pitch % 4) == 0 && (reinterpret_cast<uintptr_t>(a.in) % 4) == 0;
  if (p->dil != (a.C == 3 ? 3 : 1)) return false;
  if (a.C == 1 && a.Cp == 1)  // one channel: staged rows only (16-byte aligned rows of whole 16-pixel groups), natural order
    return a.blk.pair_major == 0 && (a.in_pitch % 16) == 0 && (reinterpret_cast<uintptr_t>(a.in) % 16) == 0 && (a.T % 16) == 0;
  if (a.C == 3 && a.Cp == 3) {  // RGB as byte rows: the same, plus a last block that still holds a whole pixel
    const int TB = 3 * a.T, L = kNB - 6 * a.pad;
    return a.blk.pair_major == 0 && (a.in_pitch % 16) == 0 && (reinterpret_cast<uintptr_t>(a.in) % 16) == 0 && (TB % 16) == 0 &&
           3 * a.pad <= kMaxPad && a.pad >= 1 && (TB % L == 0 || TB % L >= 3);
  }
  return false;
}

bool row_profitable(const RowLaunch& a, int n_ref) {
  // RGB byte rows are never picked automatically: measured on 512 x 1080p RGB sigma=10 (B200, profiles/r02/
  // bench_video_rgb_ols_rowpass_{on,off}.json) 45.3 Gpixel/s against 58.8 with the whole-line kernels — the short
  // 2048-point lines run four CTAs per SM there, the byte rows transform 17 % more points and the one-channel block
  // kernel pays two byte loads and two scalar stores per sample. GBLUR_ROW_OLS=1 still forces them (parity-tested).
  if (a.C == 3) return false;
  const int L = kNB - 2 * a.pad;
  if (L < 1 || a.T < 4096) return false;
  const long long nblk = (a.T + L - 1) / L;
  // lines that need a thread-block cluster (n_ref > ~14000 points as two sequences) run at a tenth of the roofline as
  // whole lines: the blocks win there even when they transform 50 % more points
  if (n_ref > 14000) return nblk * kNB * 2 <= (long long)n_ref * 3;
  return nblk * kNB * 5 <= (long long)n_ref * 6;
}

namespace {

// The dynamic shared-memory limit of a kernel is one attribute per function and device: set once, to the size the
// kernel always uses, under a process-wide once-flag per device (not per thread: ADVICE r1).
struct PerDevice {
  std::once_flag once[64];
  cudaError_t err[64];
  int per_sm[64];
};

template <int WARPS, int MAXREG>
cudaError_t launch_blocks(const RowParams& P, int num_sms, cudaStream_t stream) {
  constexpr size_t smem = kTableBytes + (size_t)WARPS * kExch * 8;
  auto k = ols_row_rgba_kernel<WARPS, MAXREG>;
  static PerDevice pd;
  int device = 0;
  cudaGetDevice(&device);
  device &= 63;
  std::call_once(pd.once[device], [&] {
    pd.per_sm[device] = 0;
    pd.err[device] = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (pd.err[device] == cudaSuccess)
      pd.err[device] = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pd.per_sm[device], k, WARPS * 32, smem);
    if (pd.err[device] == cudaSuccess && pd.per_sm[device] < 1) pd.err[device] = cudaErrorLaunchOutOfResources;
  });
  if (pd.err[device] != cudaSuccess) return pd.err[device];
  long long grid = (long long)num_sms * pd.per_sm[device];
  const long long need = (P.nitems + WARPS - 1) / WARPS;
  if (grid > need) grid = need;
  k<<<(unsigned)grid, WARPS * 32, smem, stream>>>(P);
  count_launch();
  return cudaGetLastError();
}

template <int WARPS, int MAXREG>
cudaError_t launch_blocks_staged(const RowParams& P, int num_sms, cudaStream_t stream) {
  constexpr size_t smem = kTableBytes + (size_t)WARPS * (kExch * 8 + kStagePx * 4 + 8);
  auto k = ols_row_rgba_staged_kernel<WARPS, MAXREG>;
  static PerDevice pd;
  int device = 0;
  cudaGetDevice(&device);
  device &= 63;
  std::call_once(pd.once[device], [&] {
    pd.per_sm[device] = 0;
    pd.err[device] = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (pd.err[device] == cudaSuccess)
      pd.err[device] = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pd.per_sm[device], k, WARPS * 32, smem);
    if (pd.err[device] == cudaSuccess && pd.per_sm[device] < 1) pd.err[device] = cudaErrorLaunchOutOfResources;
  });
  if (pd.err[device] != cudaSuccess) return pd.err[device];
  long long grid = (long long)num_sms * pd.per_sm[device];
  const long long need = ((P.nitems >> 1) + WARPS - 1) / WARPS;
  if (grid > need) grid = need;
  k<<<(unsigned)grid, WARPS * 32, smem, stream>>>(P);
  count_launch();
  return cudaGetLastError();
}

}  // namespace

namespace {
template <int WARPS, int MAXREG, int DIL>
cudaError_t launch_blocks_c1(const RowParams& P, long long rows_total, int num_sms, cudaStream_t stream) {
  constexpr size_t smem = kTableBytes + (size_t)WARPS * (kExch * 8 + 2 * kStageC1 + 8);
  auto k = ols_row_c1_kernel<WARPS, MAXREG, DIL>;
  static PerDevice pd;
  int device = 0;
  cudaGetDevice(&device);
  device &= 63;
  std::call_once(pd.once[device], [&] {
    pd.per_sm[device] = 0;
    pd.err[device] = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (pd.err[device] == cudaSuccess)
      pd.err[device] = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pd.per_sm[device], k, WARPS * 32, smem);
    if (pd.err[device] == cudaSuccess && pd.per_sm[device] < 1) pd.err[device] = cudaErrorLaunchOutOfResources;
  });
  if (pd.err[device] != cudaSuccess) return pd.err[device];
  long long grid = (long long)num_sms * pd.per_sm[device];
  const long long need = (((rows_total + 1) >> 1) * P.nblk + WARPS - 1) / WARPS;
  if ((need + grid) * WARPS >= (1LL << 31)) return cudaErrorNotSupported;  // the kernel counts items in 32 bits
  if (grid > need) grid = need;
  k<<<(unsigned)grid, WARPS * 32, smem, stream>>>(P, rows_total);
  count_launch();
  return cudaGetLastError();
}
}  // namespace

cudaError_t launch_row(RowPlan* p, const RowLaunch& a, int num_sms, cudaStream_t stream) {
  if (p->qa_rows < a.rows_total) {
    if (p->d_qa) cudaFree(p->d_qa);
    p->d_qa = nullptr;
    p->qa_rows = 0;
    cudaError_t e = cudaMalloc(&p->d_qa, (size_t)a.rows_total * sizeof(float4));
    if (e != cudaSuccess) return e;
    p->qa_rows = a.rows_total;
  }
  RowParams P{};
  P.in = a.in;
  P.in_pitch = a.in_pitch;
  P.T = a.T;
  P.pad = a.pad;
  P.Tpx = a.T;
  P.padpx = a.pad;
  P.L = kNB - 2 * a.pad;
  P.nblk = (a.T + P.L - 1) / P.L;
  P.nitems = a.rows_total * P.nblk * 2;
  P.inter = a.inter;
  P.rows_per_frame = a.blk.rows_per_frame;
  P.row_offset = a.row_offset;
  P.blk_np = a.blk.np;
  P.blk_padq = a.blk.padq;
  P.blk_frame_stride = a.blk.frame_stride;
  P.pair_major = a.blk.pair_major;
  P.slot_origin = a.blk.slot_origin;
  P.split_at = a.split_gap ? a.split_at : (1LL << 62);
  P.split_gap = a.split_gap;
  P.qa = p->d_qa;
  P.tw = p->d_tw;
  P.gains = p->d_gains;

  const unsigned nb = (unsigned)((a.rows_total + kAltsumWarps - 1) / kAltsumWarps);
  if (a.C == 1) {
    altsum_rows_c1_kernel<<<nb, kAltsumWarps * 32, 0, stream>>>(a.in, a.in_pitch, a.T, a.pad, p->q, a.rows_total,
                                                                reinterpret_cast<float*>(p->d_qa), P.split_at, P.split_gap);
    count_launch();
    const cudaError_t e1 = cudaGetLastError();
    if (e1 != cudaSuccess) return e1;
    return launch_blocks_c1<16, 128, 1>(P, a.rows_total, num_sms, stream);
  }
  if (a.C == 3) {  // byte rows
    P.Tpx = a.T;
    P.padpx = a.pad;
    P.T = 3 * a.T;
    P.pad = 3 * a.pad;
    P.L = kNB - 2 * P.pad;
    P.nblk = (P.T + P.L - 1) / P.L;
    altsum_rows_c3_kernel<<<nb, kAltsumWarps * 32, 0, stream>>>(a.in, a.in_pitch, a.T, a.pad, p->q, a.rows_total, p->d_qa,
                                                                P.split_at, P.split_gap);
    count_launch();
    const cudaError_t e3 = cudaGetLastError();
    if (e3 != cudaSuccess) return e3;
    return launch_blocks_c1<16, 128, 3>(P, a.rows_total, num_sms, stream);
  }
  altsum_rows_kernel<<<nb, kAltsumWarps * 32, 0, stream>>>(a.in, a.in_pitch, a.T, a.pad, p->q, a.rows_total, p->d_qa,
                                                           P.split_at, P.split_gap);
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;

  // kernel variant (process-wide tuning knob, read once). Staged input (default where the rows allow it): 0 = one CTA of
  // 16 warps x 128 registers, 1 = 2 x 8 x 128, 2 = 1 x 12 x 168. Direct loads: 10 = 2 CTAs x 10 warps x 96 registers
  // (the fallback), 11 = 2 x 8 x 128, 12 = 1 x 16 x 128
  static const int variant = [] {
    const char* v = getenv("GBLUR_OLS_VARIANT");
    return v ? atoi(v) : 0;
  }();
  // the staged kernel wants 16-byte aligned rows of whole 4-pixel groups and block counts that fit an int
  const bool can_stage = (a.in_pitch % 16) == 0 && (reinterpret_cast<uintptr_t>(a.in) % 16) == 0 && (a.T % 4) == 0 &&
                         P.nitems < (1LL << 31);
  if (can_stage && variant >= 0 && variant < 10) {
    switch (variant) {
      case 1:
        return launch_blocks_staged<8, 128>(P, num_sms, stream);  // 2 CTAs x 8 warps
      case 2:
        return launch_blocks_staged<12, 168>(P, num_sms, stream);
      default:
        return launch_blocks_staged<16, 128>(P, num_sms, stream);
    }
  }
  switch (variant) {
    case 11:
      return launch_blocks<8, 128>(P, num_sms, stream);
    case 12:
      return launch_blocks<16, 128>(P, num_sms, stream);
    default:
      return launch_blocks<10, 96>(P, num_sms, stream);
  }
}

bool col_eligible(const RowPlan* p, const ColLaunch& a) {
  const int w_end = a.w_end ? a.w_end : a.W;
  const bool layout_ok = (a.C == 4 && a.blk.pair_major == 1) || (a.C == 1 && a.blk.pair_major == 0);
  return p && a.C == a.Cp && layout_ok && a.blk.enabled && a.blk.padq == 0 && (a.W % kColFloats) == 0 &&
         (a.w_begin % kColFloats) == 0 && (w_end % kColFloats) == 0 && w_end > a.w_begin &&
         (reinterpret_cast<uintptr_t>(a.out) % 16) == 0 && (a.out_pitch % 16) == 0 && (a.out_frame_stride % 16) == 0 &&
         (reinterpret_cast<uintptr_t>(a.inter) % 128) == 0 && 2 * a.pad + 1 == p->ksize && a.pad <= kMaxPad &&
         a.T >= 1 && a.frames == 1;  // the column Nyquist terms are built from one frame so far
}

// Both passes as blocks (the pair-major / hole-free panel layout ties them together). Measured on B200, RGBA, against the
// whole-line kernels (profiles/r02/gates.jsonl): 3840 x 2160 0.370 -> 0.181 ms (sigma 5 / 10), -> 0.189 ms at sigma 20
// (blocks: 1.28 x and 1.33 x the reference's transform points), 2560 x 1440 0.179 -> 0.095 ms, but 1920 x 1080 sigma 5
// 0.0615 -> 0.0712 ms: two blocks for 1080 rows against the reference's 1152 points, and too few items to fill the
// persistent grids (2160 row items for 2368 warps). Rule: at least 3 Mpixel, both sides >= 1024, and neither pass more
// than 1.6 x the reference's points (1.5 x where the whole line would need a thread-block cluster, as before).
bool both_profitable(int row_len, int col_len, int pad, int n_ref_row, int n_ref_col) {
  const int L = kNB - 2 * pad;
  if (L < 1 || row_len < 1024 || col_len < 1024 || n_ref_row < 1 || n_ref_col < 1) return false;
  if ((long long)row_len * col_len < 3000000LL) return false;
  const double rw = (double)((row_len + L - 1) / L) * kNB / n_ref_row, cw = (double)((col_len + L - 1) / L) * kNB / n_ref_col;
  return rw <= (n_ref_row > 14000 ? 1.5 : 1.6) && cw <= (n_ref_col > 14000 ? 1.5 : 1.6);
}

bool col_profitable(const ColLaunch& a, int n_ref_col) {
  const int L = kNB - 2 * a.pad;
  if (L < 1 || a.T < 2048) return false;
  const long long nblk = (a.T + L - 1) / L;
  if (n_ref_col > 14000) return nblk * kNB * 2 <= (long long)n_ref_col * 3;
  return nblk * kNB * 5 <= (long long)n_ref_col * 6;
}

cudaError_t launch_colsum(const ColLaunch& a, const uint8_t* d_in, long long in_pitch, int* d_u, int row_begin, int row_end,
                          int row_base, int rows, cudaStream_t stream) {
  const int cols = a.W / 4;
  if (row_end <= row_begin) return cudaSuccess;
  const unsigned gy = (unsigned)((row_end - row_begin + kColsumRows - 1) / kColsumRows);
  if ((a.W % 16) == 0 && (in_pitch % 16) == 0 && (reinterpret_cast<uintptr_t>(d_in) % 16) == 0) {
    const int g16 = a.W / 16;
    colsum16_kernel<<<dim3((unsigned)((g16 + 127) / 128), gy), 128, 0, stream>>>(d_in, in_pitch, g16, rows, a.pad, d_u, row_begin,
                                                                                  row_end, row_base);
    count_launch();
    return cudaGetLastError();
  }
  const dim3 cgrid((unsigned)((cols + 127) / 128), gy);
  colsum_kernel<<<cgrid, 128, 0, stream>>>(d_in, in_pitch, cols, rows, a.pad, d_u, row_begin, row_end, row_base);
  count_launch();
  return cudaGetLastError();
}

namespace {
cudaError_t ensure_terms_buffers(RowPlan* p, int cols) {
  if (p->u_cols >= cols) return cudaSuccess;
  if (p->d_u) cudaFree(p->d_u);
  if (p->d_qc) cudaFree(p->d_qc);
  p->d_u = nullptr;
  p->d_qc = nullptr;
  p->u_cols = 0;
  cudaError_t e = cudaMalloc(&p->d_u, (size_t)cols * 4 * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc(&p->d_qc, (size_t)cols * 2 * sizeof(float2));
  if (e != cudaSuccess) return e;
  p->u_cols = cols;
  return cudaSuccess;
}
}  // namespace

cudaError_t launch_col_terms_from_u(RowPlan* p, const ColLaunch& a, const int* d_u, cudaStream_t stream) {
  cudaError_t e = ensure_terms_buffers(p, a.W / 4);
  if (e != cudaSuccess) return e;
  constexpr size_t usmem = kTableBytes + (size_t)kUWarps * kExch * 8;
  static PerDevice upd;
  int device = 0;
  cudaGetDevice(&device);
  device &= 63;
  std::call_once(upd.once[device], [&] {
    upd.err[device] = cudaFuncSetAttribute(u_rowpass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)usmem);
    if (upd.err[device] == cudaSuccess)
      upd.err[device] = cudaFuncSetAttribute(u_rowpass_c1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)usmem);
  });
  if (upd.err[device] != cudaSuccess) return upd.err[device];
  const int L = kNB - 2 * a.pad;
  if (a.C == 1) {
    const int items = (a.W + L - 1) / L;
    u_rowpass_c1_kernel<<<(items + kUWarps - 1) / kUWarps, kUWarps * 32, usmem, stream>>>(
        d_u, a.W, a.pad, p->q, p->q_col, p->d_tw, p->d_gains, reinterpret_cast<float*>(p->d_qc));
  } else {
    const int items = 2 * ((a.W / 4 + L - 1) / L);
    u_rowpass_kernel<<<(items + kUWarps - 1) / kUWarps, kUWarps * 32, usmem, stream>>>(d_u, a.W / 4, a.pad, p->q, p->q_col, p->d_tw,
                                                                                       p->d_gains, p->d_qc);
  }
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_col_terms(RowPlan* p, const ColLaunch& a, const uint8_t* d_in, long long in_pitch, cudaStream_t stream) {
  // the column sums run over BYTE columns in groups of four (a.W = cols * C bytes per row, a multiple of 16): for RGBA
  // a group is a pixel, for one channel four pixels — the sums are the same either way
  const int cols = a.W / 4, rows = a.T;
  cudaError_t e = ensure_terms_buffers(p, cols);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(p->d_u, 0, (size_t)cols * 4 * sizeof(int), stream);
  if (e != cudaSuccess) return e;
  e = launch_colsum(a, d_in, in_pitch, p->d_u, 0, rows, 0, rows, stream);
  if (e != cudaSuccess) return e;
  return launch_col_terms_from_u(p, a, p->d_u, stream);
}

cudaError_t launch_col(RowPlan* p, const ColLaunch& a, int num_sms, cudaStream_t stream) {
  const int rows = a.T;
  static PerDevice cpd;
  int device = 0;
  cudaGetDevice(&device);
  device &= 63;
  ColKernelParams K{};
  ColParams& P = K.P;
  P.T = rows;
  P.pad = a.pad;
  P.L = kNB - 2 * a.pad;
  P.nblk = (rows + P.L - 1) / P.L;
  const int w_end = a.w_end ? a.w_end : a.W;
  P.unit0 = a.w_begin / kColFloats;
  P.nunits = (w_end - a.w_begin) / kColFloats;
  P.frames = a.frames;
  P.nitems = (long long)P.frames * P.nunits * P.nblk;
  if (P.nitems + 2 * (long long)num_sms >= (1LL << 31)) return cudaErrorNotSupported;  // the kernel counts items in 32 bits
  P.panels_per_frame = (int)(a.blk.frame_stride / ((long long)a.blk.np * 32));
  P.st_rows_a = 256;
  P.st_na = P.L / 256;
  P.st_rows_b = P.L % 256;
  P.qc_frame_stride = 0;
  P.sign_origin = a.blk.slot_origin;
  P.qc = p->d_qc;
  P.tw = p->d_tw;
  P.gains = p->d_gains;
  P.alpha_src = (a.C == 4 && a.blk.pair_major) ? a.alpha_src : nullptr;
  if (P.alpha_src && ((reinterpret_cast<uintptr_t>(P.alpha_src) % 16) != 0 || (a.alpha_pitch % 16) != 0))
    return cudaErrorNotSupported;  // TMA tiles of the input image
  P.alpha_pitch = a.alpha_pitch;

  const long long panels_total = (long long)P.panels_per_frame * a.frames;
  const long long key[6] = {a.blk.np, panels_total, rows, a.out_pitch, a.out_frame_stride,
                            ((long long)a.frames << 40) | ((long long)a.blk.pair_major << 36) | (unsigned)a.W};
  const uint8_t* alpha_src = (a.C == 4 && a.blk.pair_major) ? a.alpha_src : nullptr;
  bool same = p->map_inter == a.inter && p->map_out == a.out && p->map_alpha == alpha_src;
  for (int i = 0; i < 6; ++i) same = same && p->map_key[i] == key[i];
  if (!same) {
    const unsigned long long idims[3] = {32, (unsigned long long)a.blk.np, (unsigned long long)panels_total};
    const unsigned long long istr[2] = {128, (unsigned long long)a.blk.np * 128};
    const unsigned ibox[3] = {(unsigned)(a.blk.pair_major ? kColBoxFloats : kColFloats), (unsigned)kColBoxSlots, 1};
    const unsigned long long odims[3] = {(unsigned long long)a.W, (unsigned long long)rows, (unsigned long long)a.frames};
    const unsigned long long ostr[2] = {(unsigned long long)a.out_pitch, (unsigned long long)a.out_frame_stride};
    const unsigned obox_a[3] = {(unsigned)kColFloats, (unsigned)P.st_rows_a, 1};
    const unsigned obox_b[3] = {(unsigned)kColFloats, (unsigned)(P.st_rows_b ? P.st_rows_b : 1), 1};
    if (!encode_tensor_map(&p->in_map, 4, 3, a.inter, idims, istr, ibox, a.blk.pair_major ? 32 : 64) ||
        !encode_tensor_map(&p->out_a, 1, 3, a.out, odims, ostr, obox_a, 0) ||
        !encode_tensor_map(&p->out_b, 1, 3, a.out, odims, ostr, obox_b, 0)) {
      p->map_inter = nullptr;
      return cudaErrorNotSupported;
    }
    if (alpha_src) {
      // the input image seen through the boxes of the output tile (same geometry: a.alpha_pitch bytes per row)
      const unsigned long long astr[2] = {(unsigned long long)a.alpha_pitch, (unsigned long long)a.alpha_pitch * (unsigned long long)rows};
      if (!encode_tensor_map(&p->alpha_a, 1, 3, alpha_src, odims, astr, obox_a, 0) ||
          !encode_tensor_map(&p->alpha_b, 1, 3, alpha_src, odims, astr, obox_b, 0)) {
        p->map_inter = nullptr;
        return cudaErrorNotSupported;
      }
    }
    p->map_inter = a.inter;
    p->map_out = a.out;
    p->map_alpha = alpha_src;
    for (int i = 0; i < 6; ++i) p->map_key[i] = key[i];
  }
  K.in_map = p->in_map;
  K.out_a = p->out_a;
  K.out_b = p->out_b;
  if (alpha_src) {
    K.alpha_a = p->alpha_a;
    K.alpha_b = p->alpha_b;
  }

  std::call_once(cpd.once[device], [&] {
    cpd.per_sm[device] = 0;
    cpd.err[device] = cudaFuncSetAttribute(ols_col_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kColSmem);
    if (cpd.err[device] == cudaSuccess)
      cpd.err[device] = cudaFuncSetAttribute(ols_col_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kColSmem);
    if (cpd.err[device] == cudaSuccess)
      cpd.err[device] = cudaFuncSetAttribute(ols_col_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kColSmem);
    if (cpd.err[device] == cudaSuccess)
      cpd.err[device] =
          cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cpd.per_sm[device], ols_col_kernel<true, true>, kColWarps * 32, kColSmem);
    if (cpd.err[device] == cudaSuccess && cpd.per_sm[device] < 1) cpd.err[device] = cudaErrorLaunchOutOfResources;
  });
  if (cpd.err[device] != cudaSuccess) return cpd.err[device];
  long long grid = (long long)num_sms * cpd.per_sm[device];
  if (grid > P.nitems) grid = P.nitems;
  if (a.blk.pair_major && alpha_src)
    ols_col_kernel<true, true><<<(unsigned)grid, kColWarps * 32, kColSmem, stream>>>(K);
  else if (a.blk.pair_major)
    ols_col_kernel<true, false><<<(unsigned)grid, kColWarps * 32, kColSmem, stream>>>(K);
  else
    ols_col_kernel<false, false><<<(unsigned)grid, kColWarps * 32, kColSmem, stream>>>(K);
  count_launch();
  return cudaGetLastError();
}

}  // namespace ols
}  // namespace gblur
