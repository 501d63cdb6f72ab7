// ols_kernels.cu — overlap-save row pass for sm_100a: one warp per (row, 1024-sample block, channel pair).
// See ols_core.cuh for the algorithm and its relation to the reference (src/gaussianblur.cpp:62-79,202-244).
//
// Why: the whole-line kernels (line_kernels.cuh) are issue/latency bound with 13 warps per SM and a CTA
// barrier per stage (profiles/r01/ncu_full_8k_v3_summary.txt: 2.2 of 4 warp instructions per clock, 17 % of
// the stall samples on the barrier). Here a transform never leaves its warp: 32 points per lane, two
// radix-32 stages, one shared-memory transpose per direction ordered by __syncwarp only, 8.7 KB of shared
// memory per warp, so 16+ independent warps per SM with no CTA-wide synchronisation in the loop.
#include <cstdlib>

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
  float4* d_qa = nullptr;
  long long qa_rows = 0;
  bool qa_valid = false;
  // column pass
  float q_col = 0.f;
  int* d_u = nullptr;       // synthetic row u: 4 * cols integer sums
  float2* d_qc = nullptr;   // 2 * cols column-pair terms
  int u_cols = 0;
};

// (K[0] - K[N/2]) / N * sum over the reflect_101-padded row of (-1)^n b[n], per channel; exact integer sums.
// One warp per row, 16-byte loads (four in flight per lane) when the row allows them.
constexpr int kAltsumWarps = 8;

__global__ void __launch_bounds__(kAltsumWarps * 32)
    altsum_rows_kernel(const uint8_t* __restrict__ in, long long in_pitch, int T, int pad, float q, long long rows_total,
                       float4* __restrict__ qa) {
  const int lane = threadIdx.x & 31;
  const long long r = (long long)blockIdx.x * kAltsumWarps + (threadIdx.x >> 5);
  if (r >= rows_total) return;
  const uint32_t* row = reinterpret_cast<const uint32_t*>(in + r * in_pitch);
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
  float* gains = reinterpret_cast<float*>(tw + kNB);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* ex = reinterpret_cast<float2*>(gains + kR * kGainStride) + warp * kExch;
  for (int i = threadIdx.x; i < kNB; i += WARPS * 32) tw[i] = __ldg(P.tw + i);
  for (int i = threadIdx.x; i < kR * kGainStride; i += WARPS * 32) gains[i] = __ldg(P.gains + i);
  __syncthreads();
  const long long stride = (long long)gridDim.x * WARPS;
  for (long long it = (long long)blockIdx.x * WARPS + warp; it < P.nitems; it += stride) {
    const int pair = (int)(it & 1);
    const long long bj = it >> 1;
    const long long r = bj / P.nblk;
    const int j = (int)(bj - r * P.nblk);
    const int s = j * P.L;
    const uint32_t* row = reinterpret_cast<const uint32_t*>(P.in + r * P.in_pitch);
    float2 c[kR];
    if (s - P.pad >= 0 && s - P.pad + kNB <= P.T)
      load_block<true>(c, P, row, s, pair, lane);
    else
      load_block<false>(c, P, row, s, pair, lane);
#pragma unroll 1
    for (int q = 0; q < 4; ++q) {
      dft<kR, false>(c);
      if ((q & 1) == 0) {
        tw_write_col(c, lane, ex, tw);
        __syncwarp();
        read_row(c, lane, ex);
        __syncwarp();  // every row has been read before the next pass overwrites columns
      } else if (q == 1) {
        mid_gains(c, lane, gains);
      }
    }
    const RowDst d = make_row_dst(P, r);
    const float4 q4 = __ldg(P.qa + r);
    store_block(c, P, d, pair ? make_float2(q4.z, q4.w) : make_float2(q4.x, q4.y), s, pair, lane);
  }
}

// Variant 5 (GBLUR_OLS_VARIANT=5, not measured yet): one CTA per row with one warp per item of the row (2 * nblk <= 20
// warps), no pre-kernel. The warps first add up the row's alternating sums together (each reads 1 / nwarps of the row,
// which the block loads touch again right afterwards), one CTA barrier per row, then run their blocks as above.
__global__ void __maxnreg__(96) ols_row_cta_kernel(const __grid_constant__ RowParams P, float q_row) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);
  float* gains = reinterpret_cast<float*>(tw + kNB);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  float2* ex = reinterpret_cast<float2*>(gains + kR * kGainStride) + warp * kExch;
  __shared__ int part[2][20][4];
  for (int i = threadIdx.x; i < kNB; i += blockDim.x) tw[i] = __ldg(P.tw + i);
  for (int i = threadIdx.x; i < kR * kGainStride; i += blockDim.x) gains[i] = __ldg(P.gains + i);
  const long long nrows = P.nitems / (2 * P.nblk);
  const int pair = warp & 1, s = (warp >> 1) * P.L;
  int par = 0;
  for (long long r = blockIdx.x; r < nrows; r += gridDim.x, par ^= 1) {
    const uint32_t* row = reinterpret_cast<const uint32_t*>(P.in + r * P.in_pitch);
    int a[4] = {0, 0, 0, 0};
    altsum_part(row, P.T, P.pad, (int)threadIdx.x, (int)blockDim.x, a);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int c = 0; c < 4; ++c) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
    }
    if (lane == 0) {
#pragma unroll
      for (int c = 0; c < 4; ++c) part[par][warp][c] = a[c];
    }
    float2 c[kR];
    if (s - P.pad >= 0 && s - P.pad + kNB <= P.T)
      load_block<true>(c, P, row, s, pair, lane);
    else
      load_block<false>(c, P, row, s, pair, lane);
    // the barrier also publishes the tables on the first row; part[par] is rewritten two rows later, after the
    // barrier of the next row, which every warp reaches only when it is done with this row's sums
    __syncthreads();
    int s0 = 0, s1 = 0;
    for (int w = 0; w < nwarps; ++w) {
      s0 += part[par][w][2 * pair];
      s1 += part[par][w][2 * pair + 1];
    }
#pragma unroll 1
    for (int q = 0; q < 4; ++q) {
      dft<kR, false>(c);
      if ((q & 1) == 0) {
        tw_write_col(c, lane, ex, tw);
        __syncwarp();
        read_row(c, lane, ex);
        __syncwarp();
      } else if (q == 1) {
        mid_gains(c, lane, gains);
      }
    }
    const RowDst d = make_row_dst(P, r);
    store_block(c, P, d, make_float2(q_row * (float)s0, q_row * (float)s1), s, pair, lane);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// column pass
// ---------------------------------------------------------------------------------------------------------------
constexpr int kColsumRows = 64;  // image rows per thread of the column-sum kernel

__global__ void __launch_bounds__(128) colsum_kernel(const uint8_t* __restrict__ in, long long in_pitch, int cols, int rows,
                                                     int pad, int* __restrict__ u) {
  const int x = blockIdx.x * 128 + threadIdx.x;
  if (x >= cols) return;
  const int r0 = blockIdx.y * kColsumRows;
  const int r1 = min(rows, r0 + kColsumRows);
  int a[4] = {0, 0, 0, 0};
  colsum_lane(in, in_pitch, x, r0, r1, pad, rows, a);
#pragma unroll
  for (int c = 0; c < 4; ++c) atomicAdd(u + 4 * x + c, a[c]);  // integer atomics: exact, order-independent
}

// The row pass of the synthetic row: one CTA, kUWarps warps over the nblk * 2 items. Writes the column terms qc.
constexpr int kUWarps = 16;

__global__ void __launch_bounds__(kUWarps * 32, 1)
    u_rowpass_kernel(const int* __restrict__ u, int T, int pad, float q_row, float q_col, const float2* __restrict__ g_tw,
                     const float* __restrict__ g_gains, float2* __restrict__ qc) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);
  float* gains = reinterpret_cast<float*>(tw + kNB);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* ex = reinterpret_cast<float2*>(gains + kR * kGainStride) + warp * kExch;
  __shared__ long long part[kUWarps][4];
  __shared__ float qa_u[4];
  for (int i = threadIdx.x; i < kNB; i += kUWarps * 32) tw[i] = __ldg(g_tw + i);
  for (int i = threadIdx.x; i < kR * kGainStride; i += kUWarps * 32) gains[i] = __ldg(g_gains + i);
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
  for (int it = warp; it < nitems; it += kUWarps) {
    const int pair = it & 1, s = (it >> 1) * L;
    float2 c[kR];
    load_block_u(c, u, T, pad, s, pair, lane);
#pragma unroll 1
    for (int q = 0; q < 4; ++q) {
      dft<kR, false>(c);
      if ((q & 1) == 0) {
        tw_write_col(c, lane, ex, tw);
        __syncwarp();
        read_row(c, lane, ex);
        __syncwarp();
      } else if (q == 1) {
        mid_gains(c, lane, gains);
      }
    }
    store_block_u(c, make_float2(qa_u[2 * pair], qa_u[2 * pair + 1]), q_col, T, pad, L, s, pair, lane, qc);
  }
}

template <int CW, int MAXREG>
__global__ void __maxnreg__(MAXREG) ols_col_rgba_kernel(const __grid_constant__ ColParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);
  float* gains = reinterpret_cast<float*>(tw + kNB);
  float2* ex_all = reinterpret_cast<float2*>(gains + kR * kGainStride);
  uint16_t* tile = reinterpret_cast<uint16_t*>(ex_all + CW * kExch);
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  float2* ex = ex_all + warp * kExch;
  for (int i = t; i < kNB; i += CW * 32) tw[i] = __ldg(P.tw + i);
  for (int i = t; i < kR * kGainStride; i += CW * 32) gains[i] = __ldg(P.gains + i);
  // no barrier needed here: the first one inside the loop also publishes the tables
  for (long long it = blockIdx.x; it < P.nitems; it += gridDim.x) {
    // consecutive items = neighbouring units of the same block: their sectors share 128-byte lines
    const int j = (int)(it / P.nunits);
    const int unit = (int)(it - (long long)j * P.nunits);
    const int s = j * P.L;
    col_fetch<CW>(P, unit, s, t, ex_all);
    __syncthreads();
    float2 c[kR];
    read_col(c, lane, ex);
#pragma unroll 1
    for (int q = 0; q < 4; ++q) {
      dft<kR, false>(c);
      if ((q & 1) == 0) {
        tw_write_col(c, lane, ex, tw);
        __syncwarp();
        read_row(c, lane, ex);
        __syncwarp();
      } else if (q == 1) {
        mid_gains(c, lane, gains);
      }
    }
    col_to_tile<CW>(c, __ldg(P.qc + unit * CW + warp), s, warp, lane, tile);
    __syncthreads();
    col_store<CW>(P, unit, s, t, tile);
    // the next fetch overwrites exchange buffers every warp has finished with (barrier above); the next
    // col_to_tile comes after the next barrier, when every thread has left this col_store
  }
}

cudaError_t row_plan_create(RowPlan** out, const std::vector<float>& taps, int n_ref, int n_ref_col) {
  *out = nullptr;
  if ((int)taps.size() > 2 * kMaxPad + 1) return cudaSuccess;  // window too wide for 1024-sample blocks: not eligible
  const HostTables t = build_tables(taps, n_ref, n_ref_col);
  RowPlan* p = new RowPlan();
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
  return p && a.C == 4 && a.Cp == 4 && a.blk.enabled && a.blk.L == 4 && a.npeers == 0 && a.split_cols == 0 &&
         (a.in_pitch % 4) == 0 && (reinterpret_cast<uintptr_t>(a.in) % 4) == 0 &&
         (reinterpret_cast<uintptr_t>(a.inter) % 16) == 0 && 2 * a.pad + 1 == p->ksize && a.pad <= kMaxPad &&
         a.T >= 1 && a.rows_total >= 1;
}

bool row_profitable(const RowLaunch& a, int n_ref) {
  const int L = kNB - 2 * a.pad;
  if (L < 1 || a.T < 4096) return false;
  const long long nblk = (a.T + L - 1) / L;
  return nblk * kNB * 5 <= (long long)n_ref * 6;
}

namespace {

template <int WARPS, int MAXREG>
cudaError_t launch_blocks(const RowParams& P, int num_sms, cudaStream_t stream) {
  constexpr size_t smem = (size_t)kNB * 8 + (size_t)kR * kGainStride * 4 + (size_t)WARPS * kExch * 8;
  auto k = ols_row_rgba_kernel<WARPS, MAXREG>;
  static thread_local int cached_device = -1, cached_per_sm = 0;
  int device = 0;
  cudaGetDevice(&device);
  if (cached_device != device) {
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, WARPS * 32, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    cached_device = device;
    cached_per_sm = per_sm;
  }
  long long grid = (long long)num_sms * cached_per_sm;
  const long long need = (P.nitems + WARPS - 1) / WARPS;
  if (grid > need) grid = need;
  k<<<(unsigned)grid, WARPS * 32, smem, stream>>>(P);
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
    p->qa_valid = false;
  }
  RowParams P{};
  P.in = a.in;
  P.in_pitch = a.in_pitch;
  P.T = a.T;
  P.pad = a.pad;
  P.L = kNB - 2 * a.pad;
  P.nblk = (a.T + P.L - 1) / P.L;
  P.nitems = a.rows_total * P.nblk * 2;
  P.inter = a.inter;
  P.rows_per_frame = a.blk.rows_per_frame;
  P.row_offset = a.row_offset;
  P.blk_np = a.blk.np;
  P.blk_padq = a.blk.padq;
  P.blk_frame_stride = a.blk.frame_stride;
  P.qa = p->d_qa;
  P.tw = p->d_tw;
  P.gains = p->d_gains;

  // warps per CTA / register cap; resident warps per SM: 20 in two CTAs (default: 0.362 ms on 8K RGBA against 0.375
  // for 16 in one CTA and 0.38 for 16 in two or four, profiles/r01/ols_rowpass_v2_check.jsonl), 16, 16, 24, 16;
  // 5 = one CTA per row without the pre-kernel (rows of at most 10 blocks)
  static const int variant = [] {
    const char* v = getenv("GBLUR_OLS_VARIANT");
    return v ? atoi(v) : 1;
  }();
  if (variant == 5 && 2 * P.nblk <= 20) {
    const int nwarps = 2 * P.nblk;
    const size_t smem = (size_t)kNB * 8 + (size_t)kR * kGainStride * 4 + (size_t)nwarps * kExch * 8;
    static thread_local int cta_device = -1;
    int device = 0;
    cudaGetDevice(&device);
    if (cta_device != device) {
      cudaError_t ea = cudaFuncSetAttribute(ols_row_cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)((size_t)kNB * 8 + (size_t)kR * kGainStride * 4 + (size_t)20 * kExch * 8));
      if (ea != cudaSuccess) return ea;
      cta_device = device;
    }
    long long grid = num_sms;
    if (grid > a.rows_total) grid = a.rows_total;
    ols_row_cta_kernel<<<(unsigned)grid, nwarps * 32, smem, stream>>>(P, p->q);
    count_launch();
    return cudaGetLastError();
  }

  // timing aid only: GBLUR_OLS_NO_ALTSUM=1 reuses the sums of the previous call on this plan (same input assumed)
  static const bool reuse_sums = getenv("GBLUR_OLS_NO_ALTSUM") != nullptr;
  cudaError_t e = cudaSuccess;
  if (!(reuse_sums && p->qa_valid)) {
    const unsigned nb = (unsigned)((a.rows_total + kAltsumWarps - 1) / kAltsumWarps);
    altsum_rows_kernel<<<nb, kAltsumWarps * 32, 0, stream>>>(a.in, a.in_pitch, a.T, a.pad, p->q, a.rows_total, p->d_qa);
    count_launch();
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    p->qa_valid = true;
  }

  switch (variant) {
    case 0:
      return launch_blocks<8, 128>(P, num_sms, stream);
    case 2:
      return launch_blocks<4, 128>(P, num_sms, stream);
    case 3:
      return launch_blocks<24, 80>(P, num_sms, stream);
    case 4:
      return launch_blocks<16, 128>(P, num_sms, stream);
    default:
      return launch_blocks<10, 96>(P, num_sms, stream);
  }
}

namespace {
template <int CW>
cudaError_t launch_col_blocks(ColParams P, int num_sms, cudaStream_t stream) {
  constexpr size_t tables = (size_t)kNB * 8 + (size_t)kR * kGainStride * 4;
  constexpr size_t csmem = tables + (size_t)CW * kExch * 8 + (size_t)kNB * CW * sizeof(uint16_t);
  P.nunits = (P.cols + CW / 2 - 1) / (CW / 2);
  P.nitems = (long long)P.nunits * P.nblk;
  auto k = ols_col_rgba_kernel<CW, 128>;
  static thread_local int c_device = -1, c_per_sm = 0;
  int device = 0;
  cudaGetDevice(&device);
  if (c_device != device) {
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csmem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, CW * 32, csmem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    c_device = device;
    c_per_sm = per_sm;
  }
  long long grid = (long long)num_sms * c_per_sm;
  if (grid > P.nitems) grid = P.nitems;
  k<<<(unsigned)grid, CW * 32, csmem, stream>>>(P);
  count_launch();
  return cudaGetLastError();
}
}  // namespace

bool col_eligible(const RowPlan* p, const ColLaunch& a) {
  return p && p->q_col != 0.f && a.C == 4 && a.Cp == 4 && a.frames == 1 && a.blk.enabled && a.panel_in == 0 &&
         a.w_begin == 0 && (a.w_end == 0 || a.w_end == a.W) && (a.W % 16) == 0 &&
         (reinterpret_cast<uintptr_t>(a.out) % 16) == 0 && (a.out_pitch % 16) == 0 &&
         (reinterpret_cast<uintptr_t>(a.inter) % 16) == 0 && 2 * a.pad + 1 == p->ksize && a.pad <= kMaxPad && a.T >= 1;
}

bool col_profitable(const ColLaunch& a, int n_ref_col) {
  const int L = kNB - 2 * a.pad;
  if (L < 1 || a.T < 4096) return false;
  const long long nblk = (a.T + L - 1) / L;
  return nblk * kNB * 5 <= (long long)n_ref_col * 6;
}

cudaError_t launch_col(RowPlan* p, const ColLaunch& a, const uint8_t* d_in, long long in_pitch, int num_sms,
                       cudaStream_t stream) {
  const int cols = a.W / 4, rows = a.T;
  cudaError_t e;
  if (p->u_cols < cols) {
    if (p->d_u) cudaFree(p->d_u);
    if (p->d_qc) cudaFree(p->d_qc);
    p->d_u = nullptr;
    p->d_qc = nullptr;
    p->u_cols = 0;
    e = cudaMalloc(&p->d_u, (size_t)cols * 4 * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc(&p->d_qc, (size_t)cols * 2 * sizeof(float2));
    if (e != cudaSuccess) return e;
    p->u_cols = cols;
  }
  e = cudaMemsetAsync(p->d_u, 0, (size_t)cols * 4 * sizeof(int), stream);
  if (e != cudaSuccess) return e;
  const dim3 cgrid((unsigned)((cols + 127) / 128), (unsigned)((rows + kColsumRows - 1) / kColsumRows));
  colsum_kernel<<<cgrid, 128, 0, stream>>>(d_in, in_pitch, cols, rows, a.pad, p->d_u);
  count_launch();
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;

  constexpr size_t tables = (size_t)kNB * 8 + (size_t)kR * kGainStride * 4;
  constexpr size_t usmem = tables + (size_t)kUWarps * kExch * 8;
  static thread_local int u_device = -1;
  int device = 0;
  cudaGetDevice(&device);
  if (u_device != device) {
    e = cudaFuncSetAttribute(u_rowpass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)usmem);
    if (e != cudaSuccess) return e;
    u_device = device;
  }
  u_rowpass_kernel<<<1, kUWarps * 32, usmem, stream>>>(p->d_u, cols, a.pad, p->q, p->q_col, p->d_tw, p->d_gains, p->d_qc);
  count_launch();
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;

  ColParams P{};
  P.inter = a.inter;
  P.panel_stride = (long long)a.blk.np * 32;
  P.padq = a.blk.padq;
  P.T = rows;
  P.pad = a.pad;
  P.L = kNB - 2 * a.pad;
  P.nblk = (rows + P.L - 1) / P.L;
  P.cols = cols;
  P.out = a.out;
  P.out_pitch = a.out_pitch;
  P.qc = p->d_qc;
  P.tw = p->d_tw;
  P.gains = p->d_gains;
  // GBLUR_OLS_COL_VARIANT: 0 = four warps per quarter panel (4 CTAs per SM), 1 = eight warps per half panel (2 CTAs)
  static const int variant = [] {
    const char* v = getenv("GBLUR_OLS_COL_VARIANT");
    return v ? atoi(v) : 0;
  }();
  return variant == 1 ? launch_col_blocks<8>(P, num_sms, stream) : launch_col_blocks<4>(P, num_sms, stream);
}

}  // namespace ols
}  // namespace gblur
