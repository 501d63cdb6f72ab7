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
  const float2* tw[kMaxStages];  // factored stage tables (TwSplit), s < nstages-1; contiguous in stage order
  const float* kperm;            // spectrum gains * 1/N in digit-reversed position order
};

// Stage twiddles are stored factored: for radix R = A*B only the B-1 "fine" powers w^1..w^(B-1)
// and the A-1 "coarse" powers w^B, w^2B, .. of w = exp(-2 pi i m / (R*M)) are tabulated
// (A+B-2 entries per butterfly column m instead of R-1); the rest are one complex product each.
// The tables are then small enough to sit in shared memory next to the exchange buffer, which
// keeps them out of the L1 that the streaming loads/stores thrash.
//   layout: tw[e*M + m], e < B-1 -> j = e+1 ; e >= B-1 -> j = (e-(B-1)+1)*B
template <int R>
struct TwSplit {
  static constexpr int B = (R % 4 == 0) ? 4 : (R % 5 == 0) ? 5 : (R % 3 == 0) ? 3 : (R % 2 == 0) ? 2 : R;
  static constexpr int A = R / B;
  static constexpr int NBASE = (B - 1) + (A - 1);
};

template <int R, bool TSM>
struct StageTwiddles {
  using S = TwSplit<R>;
  float2 fine[S::B > 1 ? S::B - 1 : 1];
  float2 coarse[S::A > 1 ? S::A - 1 : 1];
  __device__ __forceinline__ void load(const float2* __restrict__ tw, int M, int m) {
#pragma unroll
    for (int b = 1; b < S::B; ++b) fine[b - 1] = TSM ? tw[(b - 1) * M + m] : __ldg(tw + (b - 1) * M + m);
#pragma unroll
    for (int a = 1; a < S::A; ++a)
      coarse[a - 1] = TSM ? tw[(S::B - 2 + a) * M + m] : __ldg(tw + (S::B - 2 + a) * M + m);
  }
  // w^j, j compile-time after unrolling
  __device__ __forceinline__ float2 get(int j) const {
    const int a = j / S::B, b = j % S::B;
    if (a == 0) return fine[b - 1];
    if (b == 0) return coarse[a - 1];
    return cmul(coarse[a - 1], fine[b - 1]);
  }
};

// --------------------------------------------------------------------------------------------
// stages
// --------------------------------------------------------------------------------------------
// forward stage s < last: (loader | smem) -> radix-R DFT -> twiddle -> smem
template <int R, int NSEQ, bool FIRST, bool TSM, class Loader>
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
    StageTwiddles<R, TSM> tws;
    tws.load(tw, M, m);
#pragma unroll
    for (int q = 0; q < NSEQ; ++q) dft<R, false>(c[q]);
#pragma unroll
    for (int j = 0; j < R; ++j) {
      Pt<NSEQ> p;
      if (j == 0) {
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) p.v[q] = c[q][0];
      } else {
        const float2 w = tws.get(j);
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) p.v[q] = cmul(c[q][j], w);
      }
      SE::st(sm + pb + j * pstep, p);
    }
  }
}

// last forward stage + spectrum multiply + first inverse stage, all on one register set
template <int R, int NSEQ, bool TSM>
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
      const float g = TSM ? kp[j] : __ldg(kp + j);
#pragma unroll
      for (int q = 0; q < NSEQ; ++q) c[q][j] = cscale(c[q][j], g);
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
template <int R, int NSEQ, bool FINAL, bool TSM, class Storer>
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
    StageTwiddles<R, TSM> tws;
    tws.load(tw, M, m);
#pragma unroll
    for (int j = 0; j < R; ++j) {
      Pt<NSEQ> p = SE::ld(sm + pb + j * pstep);
      if (j == 0) {
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) c[q][0] = p.v[q];
      } else {
        const float2 w = tws.get(j);
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) c[q][j] = cmulc(p.v[q], w);
      }
    }
    // the buffer is dead once the LAST pass has read its slots (nb is a multiple of the CTA size for these
    // problems, so every thread reaches the barrier inside after_reads)
    if constexpr (FINAL && Storer::kTmaNext) {
      if (u + nthreads >= nb) st.after_reads();
    }
    if constexpr (FINAL && Storer::kPrefetch) {
      // the slots this thread just read are dead for the current unit: refill them, asynchronously,
      // with the next unit's samples while the last butterflies and the global stores run
#pragma unroll
      for (int j = 0; j < R; ++j) st.prefetch(sm + pb + j * pstep, pb + j * pstep);
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
      fwd_stage<R_, NSEQ, true, false>(ld, sm, sh.tw[s], sh.N, sh.M[s], sh.padq, tid, nthreads);            \
    else                                                                                             \
      fwd_stage<R_, NSEQ, false, false>(ld, sm, sh.tw[s], sh.N, sh.M[s], sh.padq, tid, nthreads);           \
    break;
      GBLUR_FOR_EACH_GENERIC_RADIX(X)
#undef X
    }
    __syncthreads();
  }
  switch (sh.radix[S - 1]) {
#define X(R_)                                                             \
  case R_:                                                                \
    fused_last_stage<R_, NSEQ, false>(sm, sh.kperm, sh.N, sh.padq, tid, nthreads); \
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
      inv_stage<R_, NSEQ, true, false>(st, sm, sh.tw[s], sh.N, sh.M[s], sh.padq, tid, nthreads);            \
    else                                                                                             \
      inv_stage<R_, NSEQ, false, false>(st, sm, sh.tw[s], sh.N, sh.M[s], sh.padq, tid, nthreads);           \
    break;
      GBLUR_FOR_EACH_GENERIC_RADIX(X)
#undef X
    }
    __syncthreads();
  }
}

// Stages 1 .. S-1 forward, the fused last stage and the inverse stages S-2 .. 1 on a buffer that
// holds `nloc` consecutive positions of the line (cluster kernels: stage 0 crosses CTAs and is
// done by the caller; after it the positions [c*nloc, (c+1)*nloc) are self-contained).
// `kperm` already points at this range's gains.
template <int NSEQ, class Loader, class Storer>
__device__ __forceinline__ void run_local_stages(const LineShape& sh, int nloc, const float* kperm, const Loader& ld,
                                                 const Storer& st, typename SmemElem<NSEQ>::type* sm, int tid,
                                                 int nthreads) {
  const int S = sh.nstages;
  for (int s = 1; s < S - 1; ++s) {
    switch (sh.radix[s]) {
#define X(R_)                                                                                  \
  case R_:                                                                                     \
    fwd_stage<R_, NSEQ, false, false>(ld, sm, sh.tw[s], nloc, sh.M[s], sh.padq, tid, nthreads); \
    break;
      GBLUR_FOR_EACH_GENERIC_RADIX(X)
#undef X
    }
    __syncthreads();
  }
  switch (sh.radix[S - 1]) {
#define X(R_)                                                                \
  case R_:                                                                   \
    fused_last_stage<R_, NSEQ, false>(sm, kperm, nloc, sh.padq, tid, nthreads); \
    break;
    GBLUR_FOR_EACH_GENERIC_RADIX(X)
#undef X
  }
  __syncthreads();
  for (int s = S - 2; s >= 1; --s) {
    switch (sh.radix[s]) {
#define X(R_)                                                                                  \
  case R_:                                                                                     \
    inv_stage<R_, NSEQ, false, false>(st, sm, sh.tw[s], nloc, sh.M[s], sh.padq, tid, nthreads); \
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
  static constexpr int EXCH = N + (PADQ ? N / PADQ : 0);  // exchange-buffer slots
  // factored twiddle tables of the non-last stages (float2 entries), kept in shared memory
  static constexpr int TW0 = TwSplit<R0>::NBASE * M0;
  static constexpr int TW1 = (S >= 3) ? TwSplit<R1>::NBASE * M1 : 0;
  static constexpr int TW2 = (S >= 4) ? TwSplit<R2>::NBASE * M2 : 0;
  static constexpr int TWTOT = TW0 + TW1 + TW2;
};

// twsm / kpsm: the stage tables and the permuted spectrum, already copied to shared memory
template <int NSEQ, class SS, bool SMEM_IN, class Loader, class Storer>
__device__ __forceinline__ void run_line_static(const float2* twsm, const float* kpsm, const Loader& ld,
                                                const Storer& st, typename SmemElem<NSEQ>::type* sm, int tid,
                                                int nthreads) {
  constexpr int N = SS::N, PQ = SS::PADQ;
  const float2* tw0 = twsm;
  const float2* tw1 = twsm + SS::TW0;
  const float2* tw2 = twsm + SS::TW0 + SS::TW1;
  fwd_stage<N / SS::M0, NSEQ, !SMEM_IN, true>(ld, sm, tw0, N, SS::M0, PQ, tid, nthreads);
  __syncthreads();
  if constexpr (Loader::kStageNext) ld.after_first_stage();  // staged input consumed: fetch the next line
  if constexpr (SS::S >= 3) {
    fwd_stage<SS::M0 / SS::M1, NSEQ, false, true>(ld, sm, tw1, N, SS::M1, PQ, tid, nthreads);
    __syncthreads();
  }
  if constexpr (SS::S >= 4) {
    fwd_stage<SS::M1 / SS::M2, NSEQ, false, true>(ld, sm, tw2, N, SS::M2, PQ, tid, nthreads);
    __syncthreads();
  }
  fused_last_stage<SS::RL, NSEQ, true>(sm, kpsm, N, PQ, tid, nthreads);
  __syncthreads();
  if constexpr (SS::S >= 4) {
    inv_stage<SS::M1 / SS::M2, NSEQ, false, true>(st, sm, tw2, N, SS::M2, PQ, tid, nthreads);
    __syncthreads();
  }
  if constexpr (SS::S >= 3) {
    inv_stage<SS::M0 / SS::M1, NSEQ, false, true>(st, sm, tw1, N, SS::M1, PQ, tid, nthreads);
    __syncthreads();
  }
  inv_stage<N / SS::M0, NSEQ, true, true>(st, sm, tw0, N, SS::M0, PQ, tid, nthreads);
  // SMEM_IN: the buffer is refilled asynchronously and the caller waits for that (plus a barrier or an
  // mbarrier) before the next unit; otherwise the next unit's first stage overwrites slots that slower
  // threads may still be reading
  if constexpr (!SMEM_IN) __syncthreads();
}

}  // namespace gblur
