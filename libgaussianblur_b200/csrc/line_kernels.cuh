// line_kernels.cuh — the two kernel templates (runtime-radix and compile-time shape) and the
// persistent-grid launcher shared by every translation unit that instantiates them.
#pragma once
#include <mutex>
#include <cuda_runtime.h>

#include "async_copy.cuh"
#include "fftconv.cuh"
#include "kernels.h"
#include "problems.cuh"

namespace gblur {

void count_launch();  // kernels.cu

// which specialised row problem the launch arguments allow
constexpr int kRowFastNone = 0, kRowFastRGBA = 1, kRowFastRGB3 = 2;

// Stage one unit into shared memory (BULK problems): np pieces of one exchange-buffer element each,
// 128 bytes apart in the panel intermediate, gathered with cp.async (no registers, all in flight
// at once). Every thread waits for its own copies; the caller's barrier publishes them.
template <int NSEQ, class Problem, class Params>
__device__ __forceinline__ void gather_stage_unit(const Params& P, int unit, typename SmemElem<NSEQ>::type* sm,
                                                  int tid, int nthreads) {
  const float* src = Problem::bulk_src(P, unit);
  for (int s = tid; s < P.np; s += nthreads) cp_async<8 * NSEQ>(sm + s, src + (long long)s * kPanelFloats);
  cp_async_commit();
}
// The copies of this thread (issued by gather_stage_unit or by the previous unit's last stage) have
// landed; the barrier makes everybody's visible.
__device__ __forceinline__ void gather_wait() {
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
}

template <int NSEQ, class Problem, class Params>
__global__ void __launch_bounds__(512, 1)
    line_kernel_generic(const __grid_constant__ LineShape sh, const __grid_constant__ Params P) {
  extern __shared__ float4 smem_raw[];
  auto* sm = reinterpret_cast<typename SmemElem<NSEQ>::type*>(smem_raw);
  if constexpr (Problem::kBulkIn) {
    if ((int)blockIdx.x < P.nunits)
      gather_stage_unit<NSEQ, Problem>(P, (int)blockIdx.x, sm, (int)threadIdx.x, (int)blockDim.x);
  }
  for (int unit = blockIdx.x; unit < P.nunits; unit += gridDim.x) {
    typename Problem::Loader ld;
    typename Problem::Storer st;
    Problem::make(P, unit, ld, st);
    if constexpr (Problem::kBulkIn) gather_wait();
    run_line_generic<NSEQ, Problem::kBulkIn>(sh, ld, st, sm, (int)threadIdx.x, (int)blockDim.x);
  }
}

// MAXREG is given to ptxas directly: with __launch_bounds__ it rounds the CTA size up (400 -> 512
// threads) and caps at 128 registers with a worse allocation.
// Shared memory: [exchange buffer][factored twiddle tables][permuted spectrum]; the tables are
// copied once per (persistent) CTA and then read with LDS — they never compete for L1.
template <int NSEQ, class SS, int NT, int MAXREG, class Problem, class Params>
__global__ void __maxnreg__(MAXREG)
    line_kernel_static(const __grid_constant__ LineShape sh, const __grid_constant__ Params P) {
  extern __shared__ __align__(128) float4 smem_raw[];  // TMA destinations need 128-byte alignment
  using Elem = typename SmemElem<NSEQ>::type;
  auto* sm = reinterpret_cast<Elem*>(smem_raw);
  float2* twsm = reinterpret_cast<float2*>(sm + SS::EXCH);
  float* kpsm = reinterpret_cast<float*>(twsm + SS::TWTOT);
  for (int i = threadIdx.x; i < SS::TWTOT; i += NT) twsm[i] = __ldg(sh.tw[0] + i);
  for (int i = threadIdx.x; i < SS::N; i += NT) kpsm[i] = __ldg(sh.kperm + i);
  // the mbarrier lives behind the tables in dynamic shared memory (a static __shared__ would shift the base)
  uint64_t& tma_bar = *reinterpret_cast<uint64_t*>(
      (reinterpret_cast<uintptr_t>(kpsm + SS::N) + 15) & ~static_cast<uintptr_t>(15));
  uint32_t tma_parity = 0;
  // staged row input (RowProblemRGBA<.., true>): [mbarrier][row bytes] behind the TMA barrier slot
  uint64_t* stage_bar = &tma_bar + 1;
  uint32_t* stage = reinterpret_cast<uint32_t*>(&tma_bar + 2);
  if constexpr (Problem::kStageIn) {
    if (threadIdx.x == 0) {
      mbar_init(stage_bar, 1);
      fence_proxy_async();
      if ((int)blockIdx.x < P.nunits) {
        const uint32_t bytes = Problem::stage_bytes(P, (long long)blockIdx.x);
        mbar_expect_tx(stage_bar, bytes);
        bulk_g2s(stage, Problem::stage_src(P, (long long)blockIdx.x), bytes, stage_bar);
      }
    }
  }
  if constexpr (Problem::kTmaIn) {
    static_assert(SS::M0 % NT == 0, "TMA prefetch needs whole passes of last-stage butterflies");
    if (threadIdx.x == 0) {
      mbar_init(&tma_bar, 1);
      fence_proxy_async();
      if ((int)blockIdx.x < P.nunits) Problem::tma_request(P, (int)blockIdx.x, sm, &tma_bar);
    }
  } else if constexpr (Problem::kBulkIn) {
    if ((int)blockIdx.x < P.nunits) gather_stage_unit<NSEQ, Problem>(P, (int)blockIdx.x, sm, (int)threadIdx.x, NT);
  }
  __syncthreads();
  for (int unit = blockIdx.x; unit < P.nunits; unit += gridDim.x) {
    typename Problem::Loader ld;
    typename Problem::Storer st;
    Problem::make(P, unit, ld, st);
    if constexpr (Problem::kTmaIn) {
      st.smbase = sm;
      st.bar = &tma_bar;
      mbar_wait(&tma_bar, tma_parity);  // this unit's tiles have landed
      tma_parity ^= 1u;
    } else if constexpr (Problem::kBulkIn) {
      gather_wait();
    }
    if constexpr (Problem::kStageIn) {
      ld.set_stage(stage, stage_bar);
      mbar_wait(stage_bar, tma_parity);  // this row's bytes have landed in the staging buffer
      tma_parity ^= 1u;
    }
    run_line_static<NSEQ, SS, Problem::kBulkIn>(twsm, kpsm, ld, st, sm, (int)threadIdx.x, NT);
  }
}

inline LineShape to_device_shape(const LinePlan& lp) {
  LineShape sh{};
  sh.N = lp.N;
  sh.nstages = lp.nstages;
  for (int i = 0; i < kMaxStages; ++i) {
    sh.radix[i] = lp.radix[i];
    sh.M[i] = lp.M[i];
    sh.tw[i] = lp.d_tw[i];
  }
  sh.padq = lp.padq;
  sh.kperm = lp.d_kperm;
  return sh;
}

// Persistent grid: one CTA per resident slot (SMs x occupancy), each looping over line units.
// The occupancy query and the smem opt-in are done once per (kernel, smem size).
template <auto k, class Params>
cudaError_t launch_line(const LinePlan& lp_in, const Params& P, int nthreads, int num_sms, cudaStream_t stream,
                        size_t extra_smem = 0) {
  if (P.nunits <= 0) return cudaSuccess;
  LinePlan lp = lp_in;
  lp.smem_bytes += extra_smem;
  // The dynamic shared-memory limit is ONE attribute per function and device, shared by every host thread: it is set
  // once, to the device's opt-in maximum, so a thread launching this instantiation with a larger FFT length can never
  // find it lowered by another thread. Only the occupancy result is cached per thread.
  static std::once_flag attr_once[64];
  static cudaError_t attr_err[64];
  static thread_local size_t cached_smem = (size_t)-1;
  static thread_local int cached_threads = -1;
  static thread_local int cached_device = -1;
  static thread_local int cached_per_sm = 0;
  int device = 0;
  cudaGetDevice(&device);
  std::call_once(attr_once[device & 63], [&] {
    int optin = 0;
    attr_err[device & 63] = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (attr_err[device & 63] == cudaSuccess)
      attr_err[device & 63] = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
  });
  if (attr_err[device & 63] != cudaSuccess) return attr_err[device & 63];
  if (cached_smem != lp.smem_bytes || cached_threads != nthreads || cached_device != device) {
    cudaError_t e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, nthreads, lp.smem_bytes);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    cached_smem = lp.smem_bytes;
    cached_threads = nthreads;
    cached_device = device;
    cached_per_sm = per_sm;
  }
  long long grid = (long long)num_sms * cached_per_sm;
  if (grid > P.nunits) grid = P.nunits;
  LineShape sh = to_device_shape(lp);
  k<<<(unsigned)grid, nthreads, lp.smem_bytes, stream>>>(sh, P);
  count_launch();
  return cudaGetLastError();
}

// entry points implemented one per translation unit
// rgba: uchar4/float4 fast path; blocked: blocked intermediate; packed: aligned u16/u32 output stores
template <int ID>
cudaError_t launch_row_static(const LinePlan&, const RowParams&, int fast, bool blocked, int num_sms, cudaStream_t);
// input: 0 = row-major loader, 1 = panel layout gathered with cp.async, 2 = panel layout through TMA tiles
template <int ID>
cudaError_t launch_col_static(const LinePlan&, const ColParams&, bool packed, int input, int num_sms, cudaStream_t);
cudaError_t launch_row_generic(const LinePlan&, const RowParams&, int fast, bool blocked, int num_sms, cudaStream_t);
cudaError_t launch_col_generic(const LinePlan&, const ColParams&, bool packed, bool blocked, int num_sms,
                               cudaStream_t);
// cluster_kernels.cu: lines split over the shared memories of a thread-block cluster
cudaError_t launch_row_cluster(const LinePlan&, const RowParams&, bool blocked, cudaStream_t);
cudaError_t launch_col_cluster(const LinePlan&, const ColParams&, bool packed, cudaStream_t);

}  // namespace gblur
