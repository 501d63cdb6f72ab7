// ols_gpu_check.cu — one-shot GPU check of the overlap-save row pass against the whole-line kernels, through the
// C ABI only (no Python): parity of the fp32 intermediate and of the final u8 image on the headline shape
// (7680x4320 RGBA sigma=20) and CUDA-event timings of both row passes. Build: tools/build_ols_check.sh.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "gblur_capi.h"

static uint32_t fmix32(uint32_t h) {
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
#define CK(x) do { int rc_ = (x); if (rc_ != 0) { printf("FAIL %s -> %d: %s\n", #x, rc_, gblur_last_error()); return 1; } } while (0)
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA FAIL %s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

static float time_stage(gblur_plan* p, const uint8_t* d_in, uint8_t* d_out, int stage, int iters) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) gblur_run_stage_device(p, d_in, d_out, 1, 1, nullptr, stage);
  cudaDeviceSynchronize();
  cudaEventRecord(a, 0);
  for (int i = 0; i < iters; ++i) gblur_run_stage_device(p, d_in, d_out, 1, 1, nullptr, stage);
  cudaEventRecord(b, 0);
  cudaEventSynchronize(b);
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  return ms / iters;
}

int main(int argc, char** argv) {
  const int rows = argc > 1 ? atoi(argv[1]) : 4320, cols = argc > 2 ? atoi(argv[2]) : 7680;
  const float sigma = argc > 3 ? (float)atof(argv[3]) : 20.0f;
  const size_t n = (size_t)rows * cols * 4;
  std::vector<uint8_t> img(n), out_a(n), out_b(n);
  for (size_t i = 0; i < n; ++i) img[i] = (uint8_t)(fmix32((uint32_t)i ^ 0x1236u) >> 24);
  uint8_t *d_in, *d_a, *d_b;
  CU(cudaMalloc(&d_in, n)); CU(cudaMalloc(&d_a, n)); CU(cudaMalloc(&d_b, n));
  CU(cudaMemcpy(d_in, img.data(), n, cudaMemcpyHostToDevice));

  setenv("GBLUR_ROW_OLS", "0", 1);
  setenv("GBLUR_COL_OLS", "0", 1);
  gblur_plan* pa = nullptr;
  CK(gblur_plan_create(&pa, 0, rows, cols, 4, sigma));
  CK(gblur_run_device(pa, d_in, d_a, 1, nullptr));
  CU(cudaDeviceSynchronize());
  std::vector<float> ia((size_t)rows * cols * 4), ib((size_t)rows * cols * 4);
  CK(gblur_dump_intermediate(pa, ia.data(), 1));
  CU(cudaMemcpy(out_a.data(), d_a, n, cudaMemcpyDeviceToHost));
  const float row_a = time_stage(pa, d_in, d_a, 1, 20), col_a = time_stage(pa, d_in, d_a, 2, 20);
  printf("{\"path\": \"whole_line\", \"shape\": [%d, %d, 4], \"sigma\": %g, \"row_ms\": %.4f, \"col_ms\": %.4f}\n", rows, cols,
         sigma, row_a, col_a);
  fflush(stdout);

  // the kernel variant (GBLUR_OLS_VARIANT) is latched per process on first use: one variant per run
  // argv[4] = "auto": the library's own choice (GBLUR_ROW_OLS unset, default variant)
  const char* variant = argc > 4 ? argv[4] : "auto";
  // argv[5] = 1: overlap-save column pass as well
  if (argc > 5 && atoi(argv[5]) > 0) setenv("GBLUR_COL_OLS", "1", 1); else unsetenv("GBLUR_COL_OLS");
  if (variant[0] == 'a') {
    unsetenv("GBLUR_ROW_OLS");
  } else {
    setenv("GBLUR_ROW_OLS", "1", 1);
    setenv("GBLUR_OLS_VARIANT", variant, 1);
  }
  gblur_plan* pb = nullptr;
  CK(gblur_plan_create(&pb, 0, rows, cols, 4, sigma));
  CK(gblur_run_device(pb, d_in, d_b, 1, nullptr));
  CU(cudaDeviceSynchronize());
  CK(gblur_dump_intermediate(pb, ib.data(), 1));
  CU(cudaMemcpy(out_b.data(), d_b, n, cudaMemcpyDeviceToHost));
  double max_i = 0;
  for (size_t i = 0; i < ia.size(); ++i) max_i = std::fmax(max_i, std::fabs((double)ia[i] - (double)ib[i]));
  int max_u = 0;
  size_t ndiff = 0;
  for (size_t i = 0; i < n; ++i) {
    const int d = std::abs((int)out_a[i] - (int)out_b[i]);
    if (d) ++ndiff;
    if (d > max_u) max_u = d;
  }
  const float row_b = time_stage(pb, d_in, d_b, 1, 20), col_b = time_stage(pb, d_in, d_b, 2, 20);
  printf("{\"path\": \"overlap_save\", \"variant\": \"%s\", \"row_ms\": %.4f, \"col_ms\": %.4f, \"inter_max_abs_diff\": %.3g, "
         "\"u8_max_diff\": %d, \"u8_ndiff\": %zu, \"launches\": %llu}\n",
         variant, row_b, col_b, max_i, max_u, ndiff, gblur_launch_count());
  fflush(stdout);
  gblur_plan_destroy(pb);
  gblur_plan_destroy(pa);
  return 0;
}
