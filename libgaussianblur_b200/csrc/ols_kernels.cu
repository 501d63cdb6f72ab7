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

cudaError_t row_plan_create(RowPlan** out, const std::vector<float>& taps, int n_ref) {
  *out = nullptr;
  if ((int)taps.size() > 2 * kMaxPad + 1) return cudaSuccess;  // window too wide for 1024-sample blocks: not eligible
  const HostTables t = build_tables(taps, n_ref);
  RowPlan* p = new RowPlan();
  p->q = t.q;
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

  // warps per CTA / register cap; resident warps per SM: 20 in two CTAs (default: 0.362 ms on 8K RGBA against 0.375
  // for 16 in one CTA and 0.38 for 16 in two or four, profiles/r01/ols_rowpass_v2_check.jsonl), 16, 16, 24, 16
  static const int variant = [] {
    const char* v = getenv("GBLUR_OLS_VARIANT");
    return v ? atoi(v) : 1;
  }();
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

}  // namespace ols
}  // namespace gblur
