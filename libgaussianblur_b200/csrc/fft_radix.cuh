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
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
// v * (wr + i wi)
__device__ __forceinline__ float2 cmul(float2 v, float2 w) {
  return make_float2(fmaf(v.x, w.x, -v.y * w.y), fmaf(v.x, w.y, v.y * w.x));
}
// v * conj(w)
__device__ __forceinline__ float2 cmulc(float2 v, float2 w) {
  return make_float2(fmaf(v.x, w.x, v.y * w.y), fmaf(v.y, w.x, -v.x * w.y));
}
// multiply by -i (forward quarter turn) or +i (inverse)
template <bool INV>
__device__ __forceinline__ float2 rot90(float2 v) {
  return INV ? make_float2(-v.y, v.x) : make_float2(v.y, -v.x);
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
    return make_float2(-v.x, -v.y);
  } else if constexpr (4 * k == 3 * R) {
    return rot90<!INV>(v);
  } else {
    constexpr float c = float(ct_cos2pi(k, R));
    constexpr float s = float(ct_sin2pi(k, R));  // W = c - i s (forward), c + i s (inverse)
    if constexpr (!INV)
      return make_float2(fmaf(v.x, c, v.y * s), fmaf(v.y, c, -v.x * s));
    else
      return make_float2(fmaf(v.x, c, -v.y * s), fmaf(v.y, c, v.x * s));
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
    float2 m = make_float2(fmaf(-0.5f, t.x, c[0].x), fmaf(-0.5f, t.y, c[0].y));
    c[0] = cadd(c[0], t);
    // forward: X1 = m - i s d, X2 = m + i s d ; inverse swaps them
    float2 e = make_float2(s * d.y, -s * d.x);  // -i s d
    if constexpr (!INV) {
      c[1] = cadd(m, e);
      c[2] = csub(m, e);
    } else {
      c[1] = csub(m, e);
      c[2] = cadd(m, e);
    }
  }
};

template <bool INV>
struct Dft<4, INV> {
  static __device__ __forceinline__ void run(float2 (&c)[4]) {
    float2 t0 = cadd(c[0], c[2]), t1 = csub(c[0], c[2]);
    float2 t2 = cadd(c[1], c[3]), t3 = csub(c[1], c[3]);
    float2 r = rot90<INV>(t3);  // forward: -i t3
    c[0] = cadd(t0, t2);
    c[2] = csub(t0, t2);
    c[1] = cadd(t1, r);
    c[3] = csub(t1, r);
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
    c[0] = make_float2(a0.x + t1.x + t2.x, a0.y + t1.y + t2.y);
    float2 m1 = make_float2(fmaf(c1, t1.x, fmaf(c2, t2.x, a0.x)), fmaf(c1, t1.y, fmaf(c2, t2.y, a0.y)));
    float2 m2 = make_float2(fmaf(c2, t1.x, fmaf(c1, t2.x, a0.x)), fmaf(c2, t1.y, fmaf(c1, t2.y, a0.y)));
    float2 u1 = make_float2(fmaf(s1, t3.x, s2 * t4.x), fmaf(s1, t3.y, s2 * t4.y));
    float2 u2 = make_float2(fmaf(s2, t3.x, -s1 * t4.x), fmaf(s2, t3.y, -s1 * t4.y));
    // forward: X1 = m1 - i u1, X4 = m1 + i u1, X2 = m2 - i u2, X3 = m2 + i u2
    float2 e1 = make_float2(u1.y, -u1.x), e2 = make_float2(u2.y, -u2.x);
    if constexpr (!INV) {
      c[1] = cadd(m1, e1);
      c[4] = csub(m1, e1);
      c[2] = cadd(m2, e2);
      c[3] = csub(m2, e2);
    } else {
      c[1] = csub(m1, e1);
      c[4] = cadd(m1, e1);
      c[2] = csub(m2, e2);
      c[3] = cadd(m2, e2);
    }
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
