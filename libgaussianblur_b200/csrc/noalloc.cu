// noalloc.cu — gblur_colsum_noalloc: the weighted column sums of an image with NO allocation on the call path.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The planned path computes these sums
// (the synthetic row u behind the reference's Nyquist term along a column, src/gaussianblur.cpp:62-79; ols_core.cuh
// colsum_lane) into a buffer the plan cudaMalloc's lazily (ols_kernels.cu ensure_terms_buffers). This variant of the
// same kernel accumulates into a file-scope __device__ array instead: every GPU that loads the library holds
// kNoAllocScratchBytes (1 MiB + 2 KiB) until exit.
//
//   * limit: cols * channels <= GBLUR_NOALLOC_MAX_ROW_BYTES (262144 bytes per image row); larger rows are rejected.
//   * the call issues one cudaMemsetAsync on the scratch and one kernel; no cudaMalloc/cudaFree/allocating routine.
//     The set-up (cudaGetSymbolAddress, once per device) allocates nothing either.
//   * every call on a device shares the one array: calls that overlap on the GPU (two streams, two host threads) RACE.
//     Callers serialise them on one stream or with events. The scratch keeps the last call's contents, so each call
//     clears the part it uses first, in stream order.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <mutex>

#include "../../include/gblur_capi.h"
#include "hostsem.hpp"
#include "ols_core.cuh"

namespace gblur {
void count_launch();                      // kernels.cu
void set_last_error(const char* message); // capi.cpp: the text gblur_last_error() returns on this thread
}  // namespace gblur

namespace {

using namespace gblur;
using namespace gblur::ols;

constexpr int kMaxRowBytes = GBLUR_NOALLOC_MAX_ROW_BYTES;  // one int32 accumulator per byte column
constexpr int kGroupCols = 128;                             // 4-byte column groups per CTA (= threads per CTA)
constexpr int kRowsPerCta = 64;                             // image rows per CTA
constexpr int kMaxTickets = kMaxRowBytes / 4 / kGroupCols;  // one ticket per CTA column
constexpr size_t kNoAllocScratchBytes = (size_t)(kMaxRowBytes + kMaxTickets) * sizeof(int);

// [0, kMaxRowBytes): accumulators; [kMaxRowBytes, +kMaxTickets): arrival counters of the CTA columns.
// In HOST code this name is only a stand-in for the symbol, never a device address: the kernel gets the pointer that
// cudaGetSymbolAddress returned for the current device (scratch_for_device below).
static __device__ __align__(256) int g_noalloc_scratch[kMaxRowBytes + kMaxTickets];
static_assert(sizeof(g_noalloc_scratch) == kNoAllocScratchBytes && kNoAllocScratchBytes == (1u << 20) + 2048,
              "the static scratch is 1 MiB + 2 KiB per device (include/gblur_capi.h states it)");

// Same arithmetic as colsum_kernel (ols_kernels.cu): thread x owns byte columns [4x, 4x + 4), CTA row blockIdx.y the
// image rows [64 y, 64 y + 64). Integer atomics into the scratch (exact, order-independent); the last CTA of a CTA
// column to arrive copies the finished sums to the caller's buffer.
__global__ void __launch_bounds__(kGroupCols)
    colsum_noalloc_kernel(const uint8_t* __restrict__ in, long long in_pitch, int groups, int rows, int pad, int* scratch,
                          int* __restrict__ out) {
  __shared__ int s_last;
  const int x = blockIdx.x * kGroupCols + threadIdx.x;
  int* acc = scratch;
  int* ticket = scratch + kMaxRowBytes + blockIdx.x;
  if (x < groups) {
    const int r0 = blockIdx.y * kRowsPerCta;
    const int r1 = min(rows, r0 + kRowsPerCta);
    int a[4] = {0, 0, 0, 0};
    colsum_lane(in, in_pitch, x, r0, r1, pad, rows, a);
#pragma unroll
    for (int c = 0; c < 4; ++c) atomicAdd(acc + 4 * x + c, a[c]);
  }
  __threadfence();  // this thread's atomics are visible device-wide before the CTA takes its ticket
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1) == (int)gridDim.y - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (x < groups) {
    int4 v;
    v.x = __ldcg(acc + 4 * x);
    v.y = __ldcg(acc + 4 * x + 1);
    v.z = __ldcg(acc + 4 * x + 2);
    v.w = __ldcg(acc + 4 * x + 3);
    *reinterpret_cast<int4*>(out + 4 * x) = v;
  }
}

struct PerDeviceScratch {
  std::once_flag once[64];
  cudaError_t err[64];
  int* ptr[64];
};
PerDeviceScratch g_pd;

// device address of g_noalloc_scratch on the CURRENT device (each device has its own instance of the array)
cudaError_t scratch_for_device(int device, int** out) {
  std::call_once(g_pd.once[device], [&] {
    void* p = nullptr;
    g_pd.err[device] = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    g_pd.ptr[device] = static_cast<int*>(p);
    if (g_pd.err[device] == cudaSuccess && p == nullptr) g_pd.err[device] = cudaErrorInvalidSymbol;
  });
  *out = g_pd.ptr[device];
  return g_pd.err[device];
}

int na_fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  gblur::set_last_error(buf);
  return code;
}

struct CurrentDevice {  // make `device` current for the call, restore the caller's on the way out
  int prev = -1;
  cudaError_t err;
  explicit CurrentDevice(int device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != device) err = cudaSetDevice(device);
  }
  ~CurrentDevice() {
    int now = -1;
    if (prev >= 0 && cudaGetDevice(&now) == cudaSuccess && now != prev) cudaSetDevice(prev);
  }
};

}  // namespace

extern "C" int gblur_colsum_noalloc(int device, const uint8_t* d_in, int rows, int cols, int channels, float sigma,
                                    int32_t* d_colsum, void* cuda_stream) {
  if (!d_in || !d_colsum) return na_fail(GBLUR_E_INVALID, "gblur_colsum_noalloc: null device pointer");
  if (rows < 1 || cols < 1) return na_fail(GBLUR_E_INVALID, "gblur_colsum_noalloc: empty image");
  if (!(sigma > 0.0f)) return na_fail(GBLUR_E_INVALID, "Invalid smoothing factor");
  if (channels != 1 && channels != 3 && channels != 4) return na_fail(GBLUR_E_INVALID, "Unsupported number of channels");
  if (device < 0 || device >= 64) return na_fail(GBLUR_E_INVALID, "gblur_colsum_noalloc: device %d outside [0, 64)", device);
  const long long row_bytes = (long long)cols * channels;
  if (row_bytes > kMaxRowBytes)
    return na_fail(GBLUR_E_UNSUPPORTED,
                   "gblur_colsum_noalloc: %lld bytes per row exceed the static scratch (GBLUR_NOALLOC_MAX_ROW_BYTES = %d)",
                   row_bytes, kMaxRowBytes);
  if (row_bytes % 4 != 0 || (reinterpret_cast<uintptr_t>(d_in) & 3) || (reinterpret_cast<uintptr_t>(d_colsum) & 15))
    return na_fail(GBLUR_E_UNSUPPORTED,
                   "gblur_colsum_noalloc: cols * channels must be a multiple of 4, d_in 4-byte and d_colsum 16-byte aligned");
  const host::Sizes sz = host::plan_sizes(rows, cols, sigma);
  if (sz.pad > std::min(rows, cols) - 1)
    return na_fail(GBLUR_E_PRECOND, "reflect_101 needs pad (%d) <= min(rows, cols) - 1", sz.pad);

  CurrentDevice cur(device);
  if (cur.err != cudaSuccess) return na_fail(GBLUR_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(cur.err));
  int* scratch = nullptr;
  cudaError_t e = scratch_for_device(device, &scratch);
  if (e != cudaSuccess) return na_fail(GBLUR_E_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch): %s", cudaGetErrorString(e));

  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  const int groups = (int)(row_bytes / 4);
  const dim3 grid((unsigned)((groups + kGroupCols - 1) / kGroupCols), (unsigned)((rows + kRowsPerCta - 1) / kRowsPerCta));
  // the array keeps the last call's sums and tickets: clear what this call uses, in stream order
  e = cudaMemsetAsync(scratch, 0, (size_t)row_bytes * sizeof(int), stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(scratch + kMaxRowBytes, 0, (size_t)grid.x * sizeof(int), stream);
  if (e != cudaSuccess) return na_fail(GBLUR_E_CUDA, "cudaMemsetAsync(scratch): %s", cudaGetErrorString(e));
  colsum_noalloc_kernel<<<grid, kGroupCols, 0, stream>>>(d_in, row_bytes, groups, rows, sz.pad, scratch, d_colsum);
  count_launch();
  e = cudaGetLastError();
  if (e != cudaSuccess) return na_fail(GBLUR_E_CUDA, "colsum_noalloc_kernel: %s", cudaGetErrorString(e));
  return GBLUR_OK;
}
