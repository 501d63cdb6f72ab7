#pragma once
#include <cuda_runtime.h>
#include "../../libgaussianblur_b200/csrc/fft_radix.cuh"
namespace pk {
using gblur::ct_cos2pi; using gblur::ct_sin2pi; using gblur::ct_gcd; using gblur::ct_modinv;
__device__ __forceinline__ float2 neg(float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return __fadd2_rn(a, neg(b)); }
__device__ __forceinline__ float2 bc(float s) { return make_float2(s, s); }
// (v.y, -v.x) = -i v
__device__ __forceinline__ float2 mi(float2 v) { return make_float2(v.y, -v.x); }
__device__ __forceinline__ float2 pi_(float2 v) { return make_float2(-v.y, v.x); }
__device__ __forceinline__ float2 cmul(float2 v, float2 w) {  // v*w = v*w.x + (i v) * w.y
  return __ffma2_rn(pi_(v), bc(w.y), __fmul2_rn(v, bc(w.x)));
}
template <bool INV> __device__ __forceinline__ float2 rot90(float2 v) { return INV ? pi_(v) : mi(v); }
template <int K, int R, bool INV>
__device__ __forceinline__ float2 mul_w(float2 v) {
  constexpr int k = ((K % R) + R) % R;
  if constexpr (k == 0) return v;
  else if constexpr (4 * k == R) return rot90<INV>(v);
  else if constexpr (2 * k == R) return neg(v);
  else if constexpr (4 * k == 3 * R) return rot90<!INV>(v);
  else {
    constexpr float c = float(ct_cos2pi(k, R));
    constexpr float s = float(ct_sin2pi(k, R));
    // forward W = c - i s: v*c + (-i v) * s
    return __ffma2_rn(rot90<INV>(v), bc(s), __fmul2_rn(v, bc(c)));
  }
}
template <int R, bool INV> struct Dft;
template <bool INV> struct Dft<1, INV> { static __device__ __forceinline__ void run(float2 (&)[1]) {} };
template <bool INV> struct Dft<2, INV> {
  static __device__ __forceinline__ void run(float2 (&c)[2]) { float2 a = c[0], b = c[1]; c[0] = cadd(a, b); c[1] = csub(a, b); }
};
template <bool INV> struct Dft<4, INV> {
  static __device__ __forceinline__ void run(float2 (&c)[4]) {
    float2 t0 = cadd(c[0], c[2]), t1 = csub(c[0], c[2]);
    float2 t2 = cadd(c[1], c[3]), t3 = csub(c[1], c[3]);
    float2 r = rot90<INV>(t3);
    c[0] = cadd(t0, t2); c[2] = csub(t0, t2); c[1] = cadd(t1, r); c[3] = csub(t1, r);
  }
};
template <int R> struct Split;
template <> struct Split<8> { static constexpr int A = 2, B = 4; };
template <> struct Split<16> { static constexpr int A = 4, B = 4; };
template <> struct Split<32> { static constexpr int A = 4, B = 8; };
template <int R, bool INV>
struct Dft {
  static constexpr int A = Split<R>::A, B = Split<R>::B;
  template <int a>
  static __device__ __forceinline__ void inner(const float2 (&c)[R], float2 (&t)[A][B]) {
    if constexpr (a < A) {
      float2 v[B];
#pragma unroll
      for (int b = 0; b < B; ++b) v[b] = c[a + A * b];
      Dft<B, INV>::run(v);
      twiddle_row<a, 0>(v);
#pragma unroll
      for (int b = 0; b < B; ++b) t[a][b] = v[b];
      inner<a + 1>(c, t);
    }
  }
  template <int a, int kb>
  static __device__ __forceinline__ void twiddle_row(float2 (&v)[B]) {
    if constexpr (kb < B) { v[kb] = mul_w<a * kb, R, INV>(v[kb]); twiddle_row<a, kb + 1>(v); }
  }
  static __device__ __forceinline__ void run(float2 (&c)[R]) {
    float2 t[A][B];
    inner<0>(c, t);
#pragma unroll
    for (int kb = 0; kb < B; ++kb) {
      float2 v[A];
#pragma unroll
      for (int a = 0; a < A; ++a) v[a] = t[a][kb];
      Dft<A, INV>::run(v);
#pragma unroll
      for (int ka = 0; ka < A; ++ka) c[kb + B * ka] = v[ka];
    }
  }
};
}
