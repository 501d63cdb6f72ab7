// microbenchmark: scalar vs packed fp32 issue throughput on sm_100a
#include <cstdio>
#include <cuda_runtime.h>
#include "pk.cuh"
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("err %s line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

template <int MODE>
__global__ void __launch_bounds__(256) k_chain(float* out, int iters, float s, float t) {
  // 16 independent accumulators (pairs)
  float2 a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  float2 m = make_float2(s, s), n = make_float2(t, t);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0) { a[i].x = fmaf(a[i].x, s, a[(i + 5) & 15].y); a[i].y = fmaf(a[i].y, t, a[(i + 3) & 15].x); }   // 2 FFMA (3 distinct regs)
      if (MODE == 1) { a[i].x = a[i].x + a[(i + 5) & 15].y; a[i].y = a[i].y + a[(i + 3) & 15].x; }                     // 2 FADD
      if (MODE == 2) { a[i] = __ffma2_rn(a[i], m, a[(i + 5) & 15]); }                                                  // 1 FFMA2
      if (MODE == 3) { a[i] = __fadd2_rn(a[i], a[(i + 5) & 15]); }                                                     // 1 FADD2
      if (MODE == 4) { a[i] = __ffma2_rn(a[i], n, make_float2(a[(i + 5) & 15].y, -a[(i + 5) & 15].x)); }               // FFMA2 with swap
      if (MODE == 5) { a[i].x = fmaf(a[i].x, 1.0001f, a[(i + 5) & 15].y); a[i].y = fmaf(a[i].y, 0.9999f, a[(i + 3) & 15].x); } // FFMA imm
    }
  }
  float r = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) r += a[i].x + a[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <bool PACKED, int MAXREG>
__global__ void __maxnreg__(MAXREG) k_dft(float2* out, int iters) {
  float2 c[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) c[i] = make_float2(threadIdx.x * 0.01f + i, i * 0.25f - threadIdx.x);
  for (int it = 0; it < iters; ++it) {
    if (PACKED) pk::Dft<32, false>::run(c); else gblur::dft<32, false>(c);
#pragma unroll
    for (int i = 0; i < 32; ++i) { c[i].x *= 0.17f; c[i].y *= 0.17f; }
  }
  float2 r = make_float2(0, 0);
#pragma unroll
  for (int i = 0; i < 32; ++i) { r.x += c[i].x; r.y += c[i].y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// DFT + smem transpose like the OLS kernel (twiddle + STS.64 col / LDS.128 row)
template <bool PACKED, int MAXREG>
__global__ void __maxnreg__(MAXREG) k_dft_ex(float2* out, int iters, const float2* gtw) {
  extern __shared__ __align__(16) float2 sm[];
  float2* tw = sm;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* ex = sm + 1024 + warp * (32 * 34);
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tw[i] = gtw[i];
  __syncthreads();
  float2 c[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) c[i] = make_float2(threadIdx.x * 0.01f + i, i * 0.25f - threadIdx.x);
  for (int it = 0; it < iters; ++it) {
    if (PACKED) pk::Dft<32, false>::run(c); else gblur::dft<32, false>(c);
    ex[lane] = c[0];
#pragma unroll
    for (int k = 1; k < 32; ++k) ex[k * 34 + lane] = PACKED ? pk::cmul(c[k], tw[k * 32 + lane]) : gblur::cmul(c[k], tw[k * 32 + lane]);
    __syncwarp();
    const float4* rowp = reinterpret_cast<const float4*>(ex + lane * 34);
#pragma unroll
    for (int m = 0; m < 16; ++m) { float4 t = rowp[m]; c[2 * m] = make_float2(t.x * 0.17f, t.y * 0.17f); c[2 * m + 1] = make_float2(t.z * 0.17f, t.w * 0.17f); }
    __syncwarp();
  }
  float2 r = make_float2(0, 0);
#pragma unroll
  for (int i = 0; i < 32; ++i) { r.x += c[i].x; r.y += c[i].y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <typename F>
float time_ms(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

int main() {
  cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
  int sms = pr.multiProcessorCount; int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", pr.name, sms, clk);
  float* out; CK(cudaMalloc(&out, 1 << 26));
  float2* tw; CK(cudaMalloc(&tw, 8192)); CK(cudaMemset(tw, 0, 8192));
  const int iters = 4000;
  const char* names[] = {"FFMA x2 (3 src regs)", "FADD x2", "FFMA2", "FADD2", "FFMA2 swap", "FFMA imm x2"};
  for (int wps : {8, 16, 32}) {  // warps per SM
    int threads = 256, blocks = sms * wps * 32 / threads;
    float ms[6];
    ms[0] = time_ms([&] { k_chain<0><<<blocks, threads>>>(out, iters, 1.0001f, 0.9999f); });
    ms[1] = time_ms([&] { k_chain<1><<<blocks, threads>>>(out, iters, 1.0001f, 0.9999f); });
    ms[2] = time_ms([&] { k_chain<2><<<blocks, threads>>>(out, iters, 1.0001f, 0.9999f); });
    ms[3] = time_ms([&] { k_chain<3><<<blocks, threads>>>(out, iters, 1.0001f, 0.9999f); });
    ms[4] = time_ms([&] { k_chain<4><<<blocks, threads>>>(out, iters, 1.0001f, 0.9999f); });
    ms[5] = time_ms([&] { k_chain<5><<<blocks, threads>>>(out, iters, 1.0001f, 0.9999f); });
    for (int m = 0; m < 6; ++m) {
      double lane_ops = (double)blocks * threads * iters * 16 * 2;  // scalar-equivalent fp ops
      printf("{\"bench\": \"chain\", \"mode\": \"%s\", \"warps_per_sm\": %d, \"ms\": %.4f, \"scalar_ops_per_clk_per_sm\": %.1f}\n", names[m], wps, ms[m],
             lane_ops / (ms[m] * 1e-3) / sms / (clk * 1e3));
    }
  }
  CK(cudaGetLastError());
  // DFT-32 throughput
  const int it2 = 2000;
  auto rep = [&](const char* name, int wps, float ms) {
    double dfts = (double)sms * wps * 32 * it2;
    printf("{\"bench\": \"%s\", \"warps_per_sm\": %d, \"ms\": %.4f, \"clk_per_dft32_per_sm_x32lanes\": %.1f, \"ns_per_1024pt_stage\": %.3f}\n", name, wps, ms,
           (ms * 1e-3) * (clk * 1e3) / (wps * it2), ms * 1e6 / ((double)sms * wps * it2));
  };
  for (int wps : {8, 12, 16, 20, 24}) {
    int blocks = sms * wps / 4;
    if (wps <= 16) rep("dft32 scalar r128", wps, time_ms([&] { k_dft<false, 128><<<blocks, 128>>>((float2*)out, it2); }));
    rep("dft32 scalar r96", wps, time_ms([&] { k_dft<false, 96><<<blocks, 128>>>((float2*)out, it2); }));
    if (wps <= 16) rep("dft32 packed r128", wps, time_ms([&] { k_dft<true, 128><<<blocks, 128>>>((float2*)out, it2); }));
    rep("dft32 packed r96", wps, time_ms([&] { k_dft<true, 96><<<blocks, 128>>>((float2*)out, it2); }));
    size_t sm = 8192 + 4 * 32 * 34 * 8;
    cudaFuncSetAttribute(k_dft_ex<false, 96>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    cudaFuncSetAttribute(k_dft_ex<true, 96>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    rep("dft32+exch scalar r96", wps, time_ms([&] { k_dft_ex<false, 96><<<blocks, 128, sm>>>((float2*)out, it2, tw); }));
    rep("dft32+exch packed r96", wps, time_ms([&] { k_dft_ex<true, 96><<<blocks, 128, sm>>>((float2*)out, it2, tw); }));
  }
  CK(cudaDeviceSynchronize());
  return 0;
}
