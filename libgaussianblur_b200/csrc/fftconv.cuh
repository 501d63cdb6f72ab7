// fftconv.cuh — one CTA = one "line unit": NSEQ complex sequences (2*NSEQ real image lines packed
// two-for-one) that are padded, forward-transformed, multiplied by the real kernel spectrum,
// inverse-transformed and cropped without leaving the SM.
//
// This is the B200 replacement for the reference's hot loop
//   process_channel_tiles                     /root/reference/src/gaussianblur.cpp:202-244
//   pffft_transform_ordered (FORWARD/BACKWARD) /root/reference/include/pffft_pommier/pffft.c:1631-1698,1904
//   pffft_sorted_optimized_convolution         /root/reference/src/gaussianblur.cpp:62-79
// Structure (not a translation of pffft's SIMD-4 real FFT):
//   * two real lines a,b ride one complex FFT as a+ib; because the kernel spectrum is real and even
//     the product needs no real-FFT split/merge pass: IFFT(FFT(a+ib)*K) = (a conv g) + i (b conv g).
//   * forward = decimation in frequency (digit-reversed spectrum), inverse = decimation in time
//     starting from that same digit-reversed order, so no reordering pass (pffft_zreorder) exists;
//     the kernel spectrum is pre-permuted on the host instead.
//   * every stage is a radix-R butterfly held in registers; the stages exchange through shared
//     memory in place (a thread rewrites exactly the slots it read), the last forward stage, the
//     spectrum multiply and the first inverse stage are fused on the same registers.
#pragma once
#include <cstdint>

#include "fft_radix.cuh"

namespace gblur {

constexpr int kMaxStages = 5;

template <int NSEQ>
struct Pt {
  float2 v[NSEQ];
};

// shared-memory element: the NSEQ sequences of one position sit side by side (8 or 16 bytes)
template <int NSEQ>
struct SmemElem;
template <>
struct SmemElem<1> {
  using type = float2;
  static __device__ __forceinline__ Pt<1> ld(const float2* p) {
    Pt<1> r;
    r.v[0] = *p;
    return r;
  }
  static __device__ __forceinline__ void st(float2* p, const Pt<1>& v) { *p = v.v[0]; }
};
template <>
struct SmemElem<2> {
  using type = float4;
  static __device__ __forceinline__ Pt<2> ld(const float4* p) {
    float4 t = *p;
    Pt<2> r;
    r.v[0] = make_float2(t.x, t.y);
    r.v[1] = make_float2(t.z, t.w);
    return r;
  }
  static __device__ __forceinline__ void st(float4* p, const Pt<2>& v) {
    *p = make_float4(v.v[0].x, v.v[0].y, v.v[1].x, v.v[1].y);
  }
};

// One slot of padding after every `padq` positions keeps the contiguous-per-thread accesses of
// the last stage (stride R_last between threads) off the same banks. padq == 0: no padding.
__device__ __forceinline__ int phys(int p, int padq) { return padq ? p + p / padq : p; }

// Runtime description of one FFT length (also used, constant-folded, by the static kernels).
struct LineShape {
  int N;
  int nstages;
  int radix[kMaxStages];
  int M[kMaxStages];  // M[s] = N / (radix[0] * ... * radix[s]); M[nstages-1] == 1
  int padq;           // radix[nstages-1] if that radix is even, else 0
  const float2* tw[kMaxStages];  // tw[s][(j-1)*M[s] + m] = exp(-2 pi i m j / (radix[s]*M[s])), s < nstages-1
  const float* kperm;            // spectrum gains * 1/N in digit-reversed position order
};

// --------------------------------------------------------------------------------------------
// stages
// --------------------------------------------------------------------------------------------
// forward stage s < last: (loader | smem) -> radix-R DFT -> twiddle -> smem
template <int R, int NSEQ, bool FIRST, class Loader>
__device__ __forceinline__ void fwd_stage(const Loader& ld, typename SmemElem<NSEQ>::type* sm,
                                          const float2* __restrict__ tw, int N, int M, int padq,
                                          int tid, int nthreads) {
  using SE = SmemElem<NSEQ>;
  const int nb = N / R;
  const int pstep = phys(M, padq);  // M is a multiple of padq for every non-last stage
  for (int u = tid; u < nb; u += nthreads) {
    const int b = u / M, m = u - b * M;
    const int base = b * (R * M) + m;
    const int pb = phys(base, padq);
    float2 c[NSEQ][R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      Pt<NSEQ> p;
      if constexpr (FIRST)
        p = ld.load(base + j * M);
      else
        p = SE::ld(sm + pb + j * pstep);
#pragma unroll
      for (int q = 0; q < NSEQ; ++q) c[q][j] = p.v[q];
    }
#pragma unroll
    for (int q = 0; q < NSEQ; ++q) dft<R, false>(c[q]);
#pragma unroll
    for (int j = 0; j < R; ++j) {
      Pt<NSEQ> p;
      if (j == 0) {
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) p.v[q] = c[q][0];
      } else {
        const float2 w = __ldg(tw + (j - 1) * M + m);
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) p.v[q] = cmul(c[q][j], w);
      }
      SE::st(sm + pb + j * pstep, p);
    }
  }
}

// last forward stage + spectrum multiply + first inverse stage, all on one register set
template <int R, int NSEQ>
__device__ __forceinline__ void fused_last_stage(typename SmemElem<NSEQ>::type* sm,
                                                 const float* __restrict__ kperm, int N, int padq,
                                                 int tid, int nthreads) {
  using SE = SmemElem<NSEQ>;
  const int nb = N / R;
  for (int u = tid; u < nb; u += nthreads) {
    const int pb = phys(u * R, padq);
    float2 c[NSEQ][R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      Pt<NSEQ> p = SE::ld(sm + pb + j);
#pragma unroll
      for (int q = 0; q < NSEQ; ++q) c[q][j] = p.v[q];
    }
#pragma unroll
    for (int q = 0; q < NSEQ; ++q) dft<R, false>(c[q]);
    const float* kp = kperm + u * R;
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const float g = __ldg(kp + j);
#pragma unroll
      for (int q = 0; q < NSEQ; ++q) {
        c[q][j].x *= g;
        c[q][j].y *= g;
      }
    }
#pragma unroll
    for (int q = 0; q < NSEQ; ++q) dft<R, true>(c[q]);
#pragma unroll
    for (int j = 0; j < R; ++j) {
      Pt<NSEQ> p;
#pragma unroll
      for (int q = 0; q < NSEQ; ++q) p.v[q] = c[q][j];
      SE::st(sm + pb + j, p);
    }
  }
}

// inverse stage s < last: smem -> conj twiddle -> inverse radix-R DFT -> (smem | storer)
template <int R, int NSEQ, bool FINAL, class Storer>
__device__ __forceinline__ void inv_stage(const Storer& st, typename SmemElem<NSEQ>::type* sm,
                                          const float2* __restrict__ tw, int N, int M, int padq,
                                          int tid, int nthreads) {
  using SE = SmemElem<NSEQ>;
  const int nb = N / R;
  const int pstep = phys(M, padq);
  for (int u = tid; u < nb; u += nthreads) {
    const int b = u / M, m = u - b * M;
    const int base = b * (R * M) + m;
    const int pb = phys(base, padq);
    float2 c[NSEQ][R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      Pt<NSEQ> p = SE::ld(sm + pb + j * pstep);
      if (j == 0) {
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) c[q][0] = p.v[q];
      } else {
        const float2 w = __ldg(tw + (j - 1) * M + m);
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) c[q][j] = cmulc(p.v[q], w);
      }
    }
#pragma unroll
    for (int q = 0; q < NSEQ; ++q) dft<R, true>(c[q]);
#pragma unroll
    for (int j = 0; j < R; ++j) {
      Pt<NSEQ> p;
#pragma unroll
      for (int q = 0; q < NSEQ; ++q) p.v[q] = c[q][j];
      if constexpr (FINAL)
        st.store(base + j * M, p);
      else
        SE::st(sm + pb + j * pstep, p);
    }
  }
}

// --------------------------------------------------------------------------------------------
// drivers: one padded line unit through all stages
// --------------------------------------------------------------------------------------------
// radices the runtime-dispatched (any valid N) kernel may use
#define GBLUR_FOR_EACH_GENERIC_RADIX(X) X(2) X(3) X(4) X(5) X(6) X(8) X(9) X(10) X(12) X(15) X(16)

// SMEM_IN: the padded line already sits in shared memory (bulk-copied there), so stage 0 reads it
// in place instead of calling the loader.
template <int NSEQ, bool SMEM_IN, class Loader, class Storer>
__device__ __forceinline__ void run_line_generic(const LineShape& sh, const Loader& ld, const Storer& st,
                                                 typename SmemElem<NSEQ>::type* sm, int tid, int nthreads) {
  const int S = sh.nstages;
  // forward stages 0 .. S-2
  for (int s = 0; s < S - 1; ++s) {
    switch (sh.radix[s]) {
#define X(R_)                                                                                        \
  case R_:                                                                                           \
    if (s == 0 && !SMEM_IN)                                                                          \
      fwd_stage<R_, NSEQ, true>(ld, sm, sh.tw[s], sh.N, sh.M[s], sh.padq, tid, nthreads);            \
    else                                                                                             \
      fwd_stage<R_, NSEQ, false>(ld, sm, sh.tw[s], sh.N, sh.M[s], sh.padq, tid, nthreads);           \
    break;
      GBLUR_FOR_EACH_GENERIC_RADIX(X)
#undef X
    }
    __syncthreads();
  }
  switch (sh.radix[S - 1]) {
#define X(R_)                                                             \
  case R_:                                                                \
    fused_last_stage<R_, NSEQ>(sm, sh.kperm, sh.N, sh.padq, tid, nthreads); \
    break;
    GBLUR_FOR_EACH_GENERIC_RADIX(X)
#undef X
  }
  __syncthreads();
  for (int s = S - 2; s >= 0; --s) {
    switch (sh.radix[s]) {
#define X(R_)                                                                                        \
  case R_:                                                                                           \
    if (s == 0)                                                                                      \
      inv_stage<R_, NSEQ, true>(st, sm, sh.tw[s], sh.N, sh.M[s], sh.padq, tid, nthreads);            \
    else                                                                                             \
      inv_stage<R_, NSEQ, false>(st, sm, sh.tw[s], sh.N, sh.M[s], sh.padq, tid, nthreads);           \
    break;
      GBLUR_FOR_EACH_GENERIC_RADIX(X)
#undef X
    }
    __syncthreads();
  }
}

// compile-time shapes: R3 == 1 -> three stages, R2 == R3 == 1 -> two stages
template <int R0, int R1, int R2, int R3>
struct StaticShape {
  static constexpr int N = R0 * R1 * R2 * R3;
  static constexpr int S = (R3 > 1) ? 4 : (R2 > 1) ? 3 : 2;
  static constexpr int RL = (S == 4) ? R3 : (S == 3) ? R2 : R1;
  static constexpr int PADQ = (RL % 2 == 0) ? RL : 0;
  static constexpr int M0 = N / R0, M1 = M0 / R1, M2 = M1 / R2;
};

template <int NSEQ, class SS, bool SMEM_IN, class Loader, class Storer>
__device__ __forceinline__ void run_line_static(const LineShape& sh, const Loader& ld, const Storer& st,
                                                typename SmemElem<NSEQ>::type* sm, int tid, int nthreads) {
  constexpr int N = SS::N, PQ = SS::PADQ;
  fwd_stage<N / SS::M0, NSEQ, !SMEM_IN>(ld, sm, sh.tw[0], N, SS::M0, PQ, tid, nthreads);
  __syncthreads();
  if constexpr (SS::S >= 3) {
    fwd_stage<SS::M0 / SS::M1, NSEQ, false>(ld, sm, sh.tw[1], N, SS::M1, PQ, tid, nthreads);
    __syncthreads();
  }
  if constexpr (SS::S >= 4) {
    fwd_stage<SS::M1 / SS::M2, NSEQ, false>(ld, sm, sh.tw[2], N, SS::M2, PQ, tid, nthreads);
    __syncthreads();
  }
  fused_last_stage<SS::RL, NSEQ>(sm, sh.kperm, N, PQ, tid, nthreads);
  __syncthreads();
  if constexpr (SS::S >= 4) {
    inv_stage<SS::M1 / SS::M2, NSEQ, false>(st, sm, sh.tw[2], N, SS::M2, PQ, tid, nthreads);
    __syncthreads();
  }
  if constexpr (SS::S >= 3) {
    inv_stage<SS::M0 / SS::M1, NSEQ, false>(st, sm, sh.tw[1], N, SS::M1, PQ, tid, nthreads);
    __syncthreads();
  }
  inv_stage<N / SS::M0, NSEQ, true>(st, sm, sh.tw[0], N, SS::M0, PQ, tid, nthreads);
  __syncthreads();
}

}  // namespace gblur
