// microbenchmark 2: does a packed f32x2 instruction free an issue slot? FP work mixed with independent integer ALU
// work (IADD3/LOP3 on the alu pipe) and with shared-memory loads; plus the operand-modifier variants of the packed ops.
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("err %s line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

// MODE: 0 scalar FADD pair, 1 FADD2; KI = integer ops per complex add
template <int MODE, int KI>
__global__ void __launch_bounds__(256) k_mix(float* out, int iters, int seed) {
  float2 a[16];
  int b[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) { a[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f); b[i] = seed * i + threadIdx.x; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0) { a[i].x = a[i].x + a[(i + 5) & 15].x; a[i].y = a[i].y + a[(i + 5) & 15].y; }
      if (MODE == 1) { a[i] = __fadd2_rn(a[i], a[(i + 5) & 15]); }
#pragma unroll
      for (int k = 0; k < KI; ++k) b[i] = (b[i] ^ b[(i + 3 + k) & 15]) + it;
    }
  }
  float r = 0; int q = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) { r += a[i].x + a[i].y; q += b[i]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = r + q;
}

// packed-operand forms: 0 plain fadd2, 1 fadd2 swapped b, 2 fadd2 swapped+neg-half b, 3 fmul2 by scalar reg, 4 ffma2 swapped multiplicand,
// 5 ffma2 plain w/ scalar, 6 fadd2 negated b
template <int FORM>
__global__ void __launch_bounds__(256) k_form(float* out, int iters, float s, float t) {
  float2 a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f + 1.0f);
  const float sv = s + threadIdx.x * 1e-9f, tv = t + threadIdx.x * 1e-9f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float2 o = a[(i + 5) & 15];
      if (FORM == 0) a[i] = __fadd2_rn(a[i], o);
      if (FORM == 1) a[i] = __fadd2_rn(a[i], make_float2(o.y, o.x));
      if (FORM == 2) a[i] = __fadd2_rn(a[i], make_float2(o.y, -o.x));
      if (FORM == 3) a[i] = __fmul2_rn(o, make_float2(sv, sv));
      if (FORM == 4) a[i] = __ffma2_rn(make_float2(-o.y, o.x), make_float2(tv, tv), a[i]);
      if (FORM == 5) a[i] = __ffma2_rn(o, make_float2(tv, tv), a[i]);
      if (FORM == 6) a[i] = __fadd2_rn(a[i], make_float2(-o.x, -o.y));
    }
  }
  float r = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) r += a[i].x + a[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// FP + shared-memory loads: KL LDS.64 per 4 complex adds
template <int MODE, int KL>
__global__ void __launch_bounds__(256) k_lds(float* out, int iters) {
  __shared__ float2 sm[2048];
  for (int i = threadIdx.x; i < 2048; i += 256) sm[i] = make_float2(i, -i);
  __syncthreads();
  float2 a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  float2 acc = make_float2(0, 0);
  int idx = threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0) { a[i].x = a[i].x + a[(i + 5) & 15].x; a[i].y = a[i].y + a[(i + 5) & 15].y; }
      if (MODE == 1) { a[i] = __fadd2_rn(a[i], a[(i + 5) & 15]); }
      if ((i & 3) == 0) {
#pragma unroll
        for (int k = 0; k < KL; ++k) { float2 v = sm[(idx + 32 * (i + k)) & 2047]; acc.x += v.x; acc.y += v.y; }  // 2 scalar FADD per LDS too
      }
    }
    idx += 7;
  }
  float r = acc.x + acc.y;
#pragma unroll
  for (int i = 0; i < 16; ++i) r += a[i].x + a[i].y;
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
  float* out; CK(cudaMalloc(&out, 1 << 26));
  const int iters = 4000, wps = 16, threads = 256, blocks = sms * wps * 32 / threads;
  auto rep = [&](const char* name, int k, float ms) {
    // cycles per (complex add + k extras) per SMSP-warp: total warp-iterations per SMSP = wps/4 * iters * 16
    double cyc = (ms * 1e-3) * (clk * 1e3) / ((double)(wps / 4) * iters * 16);
    printf("{\"bench\": \"%s\", \"extras\": %d, \"ms\": %.4f, \"clk_per_complex_op_per_smsp\": %.3f}\n", name, k, ms, cyc);
  };
#define MIX(K) rep("scalar FADDx2 + K int", K, time_ms([&] { k_mix<0, K><<<blocks, threads>>>(out, iters, 3); })); \
               rep("FADD2 + K int", K, time_ms([&] { k_mix<1, K><<<blocks, threads>>>(out, iters, 3); }));
  MIX(0) MIX(1) MIX(2) MIX(3) MIX(4)
  const char* fn[] = {"fadd2 plain", "fadd2 swap", "fadd2 swap+NP", "fmul2 scalar-bc", "ffma2 swapNP x bc + c", "ffma2 plain x bc + c", "fadd2 neg"};
  rep(fn[0], 0, time_ms([&] { k_form<0><<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); }));
  rep(fn[1], 0, time_ms([&] { k_form<1><<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); }));
  rep(fn[2], 0, time_ms([&] { k_form<2><<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); }));
  rep(fn[3], 0, time_ms([&] { k_form<3><<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); }));
  rep(fn[4], 0, time_ms([&] { k_form<4><<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); }));
  rep(fn[5], 0, time_ms([&] { k_form<5><<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); }));
  rep(fn[6], 0, time_ms([&] { k_form<6><<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); }));
#define LDSX(K) rep("scalar FADDx2 + K/4 lds", K, time_ms([&] { k_lds<0, K><<<blocks, threads>>>(out, iters); })); \
                rep("FADD2 + K/4 lds", K, time_ms([&] { k_lds<1, K><<<blocks, threads>>>(out, iters); }));
  LDSX(0) LDSX(1) LDSX(2)
  CK(cudaDeviceSynchronize());
  return 0;
}
