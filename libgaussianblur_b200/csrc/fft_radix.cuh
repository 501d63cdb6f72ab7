// fft_radix.cuh — register-resident small DFTs for sm_100a.
//
// Replaces the radix passes of the reference's FFT primitive (pffft radf2/3/4/5_ps and
// radb2/3/4/5_ps, /root/reference/include/pffft_pommier/pffft.c:492-976) with fully unrolled
// complex butterflies of radix 2..32 that live in registers. Composite radices are built at
// compile time from the 2/3/4/5 base cases: Good–Thomas (prime-factor) index maps when the two
// factors are coprime (no internal twiddles), Cooley–Tukey with compile-time twiddle constants
// otherwise. The forward transform uses exp(-2*pi*i*nk/R); INV selects the conjugate kernel
// (unnormalised, like pffft: BACKWARD(FORWARD(x)) = N*x, pffft.h:117).
#pragma once
#include <cuda_runtime.h>

namespace gblur {

// ------------------------------------------------------------------------------------------
// compile-time cos/sin of 2*pi*k/n (double precision Taylor series after reduction to [-pi,pi])
// ------------------------------------------------------------------------------------------
constexpr double kPi = 3.14159265358979323846264338327950288;

__host__ __device__ constexpr double ct_sin_series(double x) {
  double term = x, sum = x;
  for (int i = 1; i < 40; ++i) {
    term *= -x * x / double((2 * i) * (2 * i + 1));
    sum += term;
  }
  return sum;
}
__host__ __device__ constexpr double ct_cos_series(double x) {
  double term = 1.0, sum = 1.0;
  for (int i = 1; i < 40; ++i) {
    term *= -x * x / double((2 * i - 1) * (2 * i));
    sum += term;
  }
  return sum;
}
// exact octant reduction on the integer ratio k/n keeps the series argument in [0, pi/4]
__host__ __device__ constexpr double ct_cos2pi(long long k, long long n) {
  k %= n;
  if (k < 0) k += n;
  if (2 * k > n) k = n - k;                  // cos(2pi - t) = cos t      -> t in [0, pi]
  if (4 * k > n) return -ct_cos2pi(n - 2 * k, 2 * n);  // cos t = -cos(pi - t), pi - t = 2pi (n-2k)/(2n)
  if (8 * k > n) return ct_sin_series(2.0 * kPi * double(n - 4 * k) / double(4 * n));  // cos t = sin(pi/2 - t)
  return ct_cos_series(2.0 * kPi * double(k) / double(n));
}
__host__ __device__ constexpr double ct_sin2pi(long long k, long long n) {
  // sin t = cos(t - pi/2) = cos(2pi (4k - n) / (4n))
  return ct_cos2pi(4 * k - n, 4 * n);
}

__host__ __device__ constexpr int ct_gcd(int a, int b) { return b == 0 ? a : ct_gcd(b, a % b); }
__host__ __device__ constexpr int ct_modinv(int a, int m) {  // a^{-1} mod m, gcd(a,m)=1, m small
  for (int x = 1; x < m; ++x)
    if ((a * x) % m == 1) return x;
  return 1;
}

// ------------------------------------------------------------------------------------------
// complex helpers
//
// sm_100a issues fp32 either as scalar FADD/FMUL/FFMA or as the packed pair FADD2/FMUL2/FFMA2 (PTX
// add/mul/fma.rn.f32x2) on a 64-bit register pair. Measured on B200 (profiles/r02/ubench_fp32_v1.jsonl,
// ubench_issue_v1.jsonl): both forms deliver the same 128 results per clock and SM — a packed instruction holds
// the fp32 pipe for two cycles — but it takes ONE issue slot for two results, and these kernels are issue-bound
// (ncu: "not selected" is the top stall reason). A complex value (re, im) is such a pair, so a complex add is one
// instruction and a real scale one. The catch: ptxas folds a component swap of a packed operand into the
// instruction (R.F32x2.LO_HI) and that form costs 3.5-4 pipe cycles instead of 2, which made the first all-packed
// version of these butterflies SLOWER than the scalar one (row blocks 0.317 -> 0.350 ms on 8K RGBA). So every
// operation that needs the cross term of a complex product or a quarter turn (+-i) is written in scalar form,
// where the swap is a register renaming, and only the swap-free operations are packed:
//   cadd / csub / cscale : FADD2 / FMUL2                                  (1 slot,  2 pipe cycles)
//   a +- (+-i) b         : two scalar FADD                                (2 slots, 2 pipe cycles)
//   v * (wr + i wi)      : FMUL2 by wr, two scalar FFMA for the wi terms  (3 slots, 4 pipe cycles; scalar: 4 slots)
// Host builds (tests/cpp/ols_emulate.cpp) take the scalar definitions below.
// ------------------------------------------------------------------------------------------
#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ >= 1000) && !defined(GBLUR_NO_PACKED_F32)
#define GBLUR_PACKED_F32 1
__device__ __forceinline__ float2 pk_add(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 pk_mul(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 pk_fma(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
#else
__device__ __forceinline__ float2 pk_add(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 pk_mul(float2 a, float2 b) { return make_float2(a.x * b.x, a.y * b.y); }
__device__ __forceinline__ float2 pk_fma(float2 a, float2 b, float2 c) {
  return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y));
}
#endif
__device__ __forceinline__ float2 pk_bc(float s) { return make_float2(s, s); }  // scalar on both halves
__device__ __forceinline__ float2 pk_neg(float2 a) { return make_float2(-a.x, -a.y); }

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return pk_add(a, b); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return pk_add(a, pk_neg(b)); }
// v * s, s real
__device__ __forceinline__ float2 cscale(float2 v, float s) { return pk_mul(v, pk_bc(s)); }
// a + i b, a - i b (scalar on purpose, see above)
__device__ __forceinline__ float2 cadd_pi(float2 a, float2 b) { return make_float2(a.x - b.y, a.y + b.x); }
__device__ __forceinline__ float2 cadd_mi(float2 a, float2 b) { return make_float2(a.x + b.y, a.y - b.x); }
// a + t * (+-i b) with a real factor t: two scalar FFMA
__device__ __forceinline__ float2 cfma_pi(float2 a, float t, float2 b) {
  return make_float2(fmaf(-b.y, t, a.x), fmaf(b.x, t, a.y));
}
__device__ __forceinline__ float2 cfma_mi(float2 a, float t, float2 b) {
  return make_float2(fmaf(b.y, t, a.x), fmaf(-b.x, t, a.y));
}
// v * (wr + i wi) = v wr + (i v) wi
__device__ __forceinline__ float2 cmul(float2 v, float2 w) { return cfma_pi(cscale(v, w.x), w.y, v); }
// v * conj(w) = v wr - (i v) wi
__device__ __forceinline__ float2 cmulc(float2 v, float2 w) { return cfma_mi(cscale(v, w.x), w.y, v); }
// multiply by -i (forward quarter turn) or +i (inverse): a value whose halves are exchanged; a packed consumer
// pays for the swap, so the butterflies below avoid it wherever they can (cadd_rot / csub_rot)
template <bool INV>
__device__ __forceinline__ float2 rot90(float2 v) {
  return INV ? make_float2(-v.y, v.x) : make_float2(v.y, -v.x);
}
// a + rot90<INV>(b), a - rot90<INV>(b)
template <bool INV>
__device__ __forceinline__ float2 cadd_rot(float2 a, float2 b) {
  return INV ? cadd_pi(a, b) : cadd_mi(a, b);
}
template <bool INV>
__device__ __forceinline__ float2 csub_rot(float2 a, float2 b) {
  return INV ? cadd_mi(a, b) : cadd_pi(a, b);
}
// a + t * rot90<INV>(b)
template <bool INV>
__device__ __forceinline__ float2 cfma_rot(float2 a, float t, float2 b) {
  return INV ? cfma_pi(a, t, b) : cfma_mi(a, t, b);
}

// v * W_R^K with a compile-time twiddle, W_R = exp(-+ 2 pi i / R); trivial turns cost nothing
template <int K, int R, bool INV>
__device__ __forceinline__ float2 mul_w(float2 v) {
  constexpr int k = ((K % R) + R) % R;
  if constexpr (k == 0) {
    return v;
  } else if constexpr (4 * k == R) {
    return rot90<INV>(v);
  } else if constexpr (2 * k == R) {
    return pk_neg(v);
  } else if constexpr (4 * k == 3 * R) {
    return rot90<!INV>(v);
  } else {
    constexpr float c = float(ct_cos2pi(k, R));
    constexpr float s = float(ct_sin2pi(k, R));  // W = c - i s (forward), c + i s (inverse)
    return cfma_rot<INV>(cscale(v, c), s, v);
  }
}

// ------------------------------------------------------------------------------------------
// base butterflies
// ------------------------------------------------------------------------------------------
template <int R, bool INV>
struct Dft;

template <bool INV>
struct Dft<1, INV> {
  static __device__ __forceinline__ void run(float2 (&)[1]) {}
};

template <bool INV>
struct Dft<2, INV> {
  static __device__ __forceinline__ void run(float2 (&c)[2]) {
    float2 a = c[0], b = c[1];
    c[0] = cadd(a, b);
    c[1] = csub(a, b);
  }
};

template <bool INV>
struct Dft<3, INV> {
  static __device__ __forceinline__ void run(float2 (&c)[3]) {
    constexpr float s = float(ct_sin2pi(1, 3));
    float2 t = cadd(c[1], c[2]);
    float2 d = csub(c[1], c[2]);
    float2 m = pk_fma(t, pk_bc(-0.5f), c[0]);
    c[0] = cadd(c[0], t);
    // forward: X1 = m - i s d, X2 = m + i s d ; inverse swaps them
    c[1] = cfma_rot<INV>(m, s, d);
    c[2] = cfma_rot<!INV>(m, s, d);
  }
};

template <bool INV>
struct Dft<4, INV> {
  static __device__ __forceinline__ void run(float2 (&c)[4]) {
    float2 t0 = cadd(c[0], c[2]), t1 = csub(c[0], c[2]);
    float2 t2 = cadd(c[1], c[3]), t3 = csub(c[1], c[3]);
    c[0] = cadd(t0, t2);
    c[2] = csub(t0, t2);
    c[1] = cadd_rot<INV>(t1, t3);  // forward: t1 - i t3
    c[3] = csub_rot<INV>(t1, t3);
  }
};

template <bool INV>
struct Dft<5, INV> {
  static __device__ __forceinline__ void run(float2 (&c)[5]) {
    constexpr float c1 = float(ct_cos2pi(1, 5)), c2 = float(ct_cos2pi(2, 5));
    constexpr float s1 = float(ct_sin2pi(1, 5)), s2 = float(ct_sin2pi(2, 5));
    float2 t1 = cadd(c[1], c[4]), t2 = cadd(c[2], c[3]);
    float2 t3 = csub(c[1], c[4]), t4 = csub(c[2], c[3]);
    float2 a0 = c[0];
    c[0] = cadd(cadd(a0, t1), t2);
    float2 m1 = pk_fma(t1, pk_bc(c1), pk_fma(t2, pk_bc(c2), a0));
    float2 m2 = pk_fma(t1, pk_bc(c2), pk_fma(t2, pk_bc(c1), a0));
    float2 u1 = pk_fma(t3, pk_bc(s1), cscale(t4, s2));
    float2 u2 = pk_fma(t3, pk_bc(s2), cscale(t4, -s1));
    // forward: X1 = m1 - i u1, X4 = m1 + i u1, X2 = m2 - i u2, X3 = m2 + i u2 ; inverse swaps the signs
    c[1] = cadd_rot<INV>(m1, u1);
    c[4] = csub_rot<INV>(m1, u1);
    c[2] = cadd_rot<INV>(m2, u2);
    c[3] = csub_rot<INV>(m2, u2);
  }
};

// ------------------------------------------------------------------------------------------
// composite radices R = A * B
// ------------------------------------------------------------------------------------------
template <int R>
struct Split;  // A = inner stride / outer DFT, B = first (inner) DFT length
#define GBLUR_SPLIT(R_, A_, B_) \
  template <>                   \
  struct Split<R_> {            \
    static constexpr int A = A_, B = B_; \
  };
GBLUR_SPLIT(6, 2, 3)
GBLUR_SPLIT(8, 2, 4)
GBLUR_SPLIT(9, 3, 3)
GBLUR_SPLIT(10, 2, 5)
GBLUR_SPLIT(12, 3, 4)
GBLUR_SPLIT(15, 3, 5)
GBLUR_SPLIT(16, 4, 4)
GBLUR_SPLIT(18, 2, 9)
GBLUR_SPLIT(20, 4, 5)
GBLUR_SPLIT(24, 3, 8)
GBLUR_SPLIT(25, 5, 5)
GBLUR_SPLIT(27, 3, 9)
GBLUR_SPLIT(30, 5, 6)
GBLUR_SPLIT(32, 4, 8)
#undef GBLUR_SPLIT

template <int R, bool INV>
struct Dft {
  static constexpr int A = Split<R>::A, B = Split<R>::B;
  static constexpr bool kPfa = (ct_gcd(A, B) == 1);

  // Good–Thomas output map coefficients (CRT), evaluated once at compile time
  static constexpr int kP = kPfa ? B * ct_modinv(B % A, A) : B;
  static constexpr int kQ = kPfa ? A * ct_modinv(A % B, B) : 1;

  // input index of element (a, b) and output index of element (ka, kb)
  static __device__ __forceinline__ constexpr int in_idx(int a, int b) {
    return kPfa ? (B * a + A * b) % R : (a + A * b);
  }
  static __device__ __forceinline__ constexpr int out_idx(int ka, int kb) { return (kP * ka + kQ * kb) % R; }

  template <int a>
  static __device__ __forceinline__ void inner(const float2 (&c)[R], float2 (&t)[A][B]) {
    if constexpr (a < A) {
      float2 v[B];
#pragma unroll
      for (int b = 0; b < B; ++b) v[b] = c[in_idx(a, b)];
      Dft<B, INV>::run(v);
      twiddle_row<a, 0>(v);
#pragma unroll
      for (int b = 0; b < B; ++b) t[a][b] = v[b];
      inner<a + 1>(c, t);
    }
  }
  // Cooley–Tukey only: t[a][kb] *= W_R^{a*kb}
  template <int a, int kb>
  static __device__ __forceinline__ void twiddle_row(float2 (&v)[B]) {
    if constexpr (!kPfa && kb < B) {
      v[kb] = mul_w<a * kb, R, INV>(v[kb]);
      twiddle_row<a, kb + 1>(v);
    }
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
      for (int ka = 0; ka < A; ++ka) c[out_idx(ka, kb)] = v[ka];
    }
  }
};

template <int R, bool INV>
__device__ __forceinline__ void dft(float2 (&c)[R]) {
  Dft<R, INV>::run(c);
}

}  // namespace gblur
