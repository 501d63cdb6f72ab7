// cluster_kernels.cu — lines too long for one CTA's shared memory (FFT lengths 34560, 69120, ...).
//
// One line unit is handled by a thread-block cluster of G CTAs whose shared memories together hold
// the padded line (distributed shared memory). The first forward stage (radix 16) reads the line
// from global memory and scatters its outputs to the CTA that owns them; after that stage the line
// is 16 independent sub-transforms of length N/16, 16/G per CTA, so every middle stage, the
// spectrum multiply and the matching inverse stages run on local shared memory exactly as in the
// single-CTA kernel. The last inverse stage gathers its inputs from all CTAs of the cluster and
// writes the cropped result to global memory. Two cluster barriers per unit order the exchanges.
// Replaces, for these lengths, the same reference code as fftconv.cuh
// (/root/reference/src/gaussianblur.cpp:202-244, include/pffft_pommier/pffft.c:1631-1698).
#include <cooperative_groups.h>

#include <type_traits>

#include "line_kernels.cuh"

namespace cg = cooperative_groups;

namespace gblur {

constexpr int kClusterR0 = 16;

template <int NSEQ, int G, class Loader>
__device__ __forceinline__ void cluster_fwd_stage0(const Loader& ld, typename SmemElem<NSEQ>::type* const (&peer)[G],
                                                   const float2* __restrict__ tw, int M0, int padq, int rank, int tid,
                                                   int nthreads) {
  using SE = SmemElem<NSEQ>;
  constexpr int R = kClusterR0, JPER = R / G;
  const int pstep = phys(M0, padq);
  for (int u = rank * nthreads + tid; u < M0; u += G * nthreads) {
    float2 c[NSEQ][R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const Pt<NSEQ> p = ld.load(u + j * M0);
#pragma unroll
      for (int q = 0; q < NSEQ; ++q) c[q][j] = p.v[q];
    }
    StageTwiddles<R, false> tws;
    tws.load(tw, M0, u);
#pragma unroll
    for (int q = 0; q < NSEQ; ++q) dft<R, false>(c[q]);
    const int pu = phys(u, padq);
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
      SE::st(peer[j / JPER] + (j % JPER) * pstep + pu, p);  // owner of sub-transform j
    }
  }
}

template <int NSEQ, int G, class Storer>
__device__ __forceinline__ void cluster_inv_stage0(const Storer& st, typename SmemElem<NSEQ>::type* const (&peer)[G],
                                                   const float2* __restrict__ tw, int M0, int padq, int rank, int tid,
                                                   int nthreads) {
  using SE = SmemElem<NSEQ>;
  constexpr int R = kClusterR0, JPER = R / G;
  const int pstep = phys(M0, padq);
  for (int u = rank * nthreads + tid; u < M0; u += G * nthreads) {
    float2 c[NSEQ][R];
    StageTwiddles<R, false> tws;
    tws.load(tw, M0, u);
    const int pu = phys(u, padq);
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const Pt<NSEQ> p = SE::ld(peer[j / JPER] + (j % JPER) * pstep + pu);
      if (j == 0) {
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) c[q][0] = p.v[q];
      } else {
        const float2 w = tws.get(j);
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
      st.store(u + j * M0, p);
    }
  }
}

// Compile-time stage list after the radix-16 cluster stage (R = 1: stage absent). The runtime-radix
// path (SS = void) handles every other length.
template <int R1_, int R2_, int R3_, int R4_>
struct ClusterShape {
  static constexpr int R1 = R1_, R2 = R2_, R3 = R3_, R4 = R4_;
  static constexpr int N = kClusterR0 * R1 * R2 * R3 * R4;
  static constexpr int M0 = N / kClusterR0, M1 = M0 / R1, M2 = M1 / R2, M3 = M2 / R3;
  static constexpr int RL = R4 > 1 ? R4 : (R3 > 1 ? R3 : (R2 > 1 ? R2 : R1));
  static constexpr int PADQ = (RL % 2 == 0) ? RL : 0;
};

// Local stages with every stride, radix and the bank-padding period constant-folded.
template <int NSEQ, int G, class SS, class Loader, class Storer>
__device__ __forceinline__ void run_local_static(const LineShape& sh, const float* kperm, const Loader& ld,
                                                 const Storer& st, typename SmemElem<NSEQ>::type* sm, int tid) {
  constexpr int NT = 512, NLOC = SS::N / G, PQ = SS::PADQ;
  constexpr bool S3 = SS::R2 > 1, S4 = SS::R3 > 1, S5 = SS::R4 > 1;  // stage 2 / 3 / 4 present
  // forward: every present stage except the last one
  if constexpr (S3) {
    fwd_stage<SS::R1, NSEQ, false, false>(ld, sm, sh.tw[1], NLOC, SS::M1, PQ, tid, NT);
    __syncthreads();
  }
  if constexpr (S4) {
    fwd_stage<SS::R2, NSEQ, false, false>(ld, sm, sh.tw[2], NLOC, SS::M2, PQ, tid, NT);
    __syncthreads();
  }
  if constexpr (S5) {
    fwd_stage<SS::R3, NSEQ, false, false>(ld, sm, sh.tw[3], NLOC, SS::M3, PQ, tid, NT);
    __syncthreads();
  }
  fused_last_stage<SS::RL, NSEQ, false>(sm, kperm, NLOC, PQ, tid, NT);
  __syncthreads();
  if constexpr (S5) {
    inv_stage<SS::R3, NSEQ, false, false>(st, sm, sh.tw[3], NLOC, SS::M3, PQ, tid, NT);
    __syncthreads();
  }
  if constexpr (S4) {
    inv_stage<SS::R2, NSEQ, false, false>(st, sm, sh.tw[2], NLOC, SS::M2, PQ, tid, NT);
    __syncthreads();
  }
  if constexpr (S3) {
    inv_stage<SS::R1, NSEQ, false, false>(st, sm, sh.tw[1], NLOC, SS::M1, PQ, tid, NT);
    __syncthreads();
  }
}

template <int NSEQ, int G, class SS, class Problem, class Params>
__global__ void __launch_bounds__(512, 1)
    line_kernel_cluster(const __grid_constant__ LineShape sh, const __grid_constant__ Params P) {
  extern __shared__ float4 smem_raw[];
  using Elem = typename SmemElem<NSEQ>::type;
  Elem* sm = reinterpret_cast<Elem*>(smem_raw);
  cg::cluster_group cl = cg::this_cluster();
  const int rank = (int)cl.block_rank();
  Elem* peer[G];
#pragma unroll
  for (int g = 0; g < G; ++g) peer[g] = cl.map_shared_rank(sm, g);
  const int nclusters = (int)gridDim.x / G, cid = (int)blockIdx.x / G;
  constexpr bool kStatic = !std::is_void<SS>::value;
  int nloc, M0, padq, nt;
  if constexpr (kStatic) {
    nloc = SS::N / G;
    M0 = SS::M0;
    padq = SS::PADQ;
    nt = 512;
  } else {
    nloc = sh.N / G;
    M0 = sh.M[0];
    padq = sh.padq;
    nt = (int)blockDim.x;
  }
  const float* kperm = sh.kperm + (long long)rank * nloc;
  const int tid = (int)threadIdx.x;
  cl.sync();  // every CTA of the cluster is resident before anybody writes into a peer
  for (int unit = cid; unit < P.nunits; unit += nclusters) {
    typename Problem::Loader ld;
    typename Problem::Storer st;
    Problem::make(P, unit, ld, st);
    cluster_fwd_stage0<NSEQ, G>(ld, peer, sh.tw[0], M0, padq, rank, tid, nt);
    cl.sync();
    if constexpr (kStatic)
      run_local_static<NSEQ, G, SS>(sh, kperm, ld, st, sm, tid);
    else
      run_local_stages<NSEQ>(sh, nloc, kperm, ld, st, sm, tid, nt);
    cl.sync();
    cluster_inv_stage0<NSEQ, G>(st, peer, sh.tw[0], M0, padq, rank, tid, nt);
    cl.sync();  // peers have finished reading this CTA's buffer
  }
}

// FFT lengths of BASELINE.json configs 4 and 5 with the planner's radices (planner.hpp: plan_line)
using Shape34560 = ClusterShape<15, 12, 12, 1>;  // 16 * 15 * 12 * 12, G = 2
using Shape69120 = ClusterShape<10, 9, 6, 8>;    // 16 * 10 * 9 * 6 * 8, G = 4

template <class SS>
static bool plan_matches(const LinePlan& lp, int G) {
  const int r[4] = {SS::R1, SS::R2, SS::R3, SS::R4};
  if (lp.N != SS::N || lp.cluster != G || lp.nthreads != 512 || lp.radix[0] != kClusterR0) return false;
  int s = 1;
  for (int i = 0; i < 4; ++i)
    if (r[i] > 1) {
      if (s >= lp.nstages || lp.radix[s] != r[i]) return false;
      ++s;
    }
  return s == lp.nstages;
}

template <auto k, int G, class Params>
static cudaError_t launch_cluster(const LinePlan& lp, const Params& P, cudaStream_t stream) {
  if (P.nunits <= 0) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lp.smem_bytes);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3((unsigned)lp.nthreads);
  cfg.gridDim = dim3((unsigned)G);
  cfg.dynamicSmemBytes = lp.smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = G;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int nclusters = 0;
  e = cudaOccupancyMaxActiveClusters(&nclusters, k, &cfg);
  if (e != cudaSuccess) return e;
  if (nclusters < 1) return cudaErrorLaunchOutOfResources;
  if (nclusters > P.nunits) nclusters = P.nunits;
  cfg.gridDim = dim3((unsigned)(nclusters * G));
  LineShape sh = to_device_shape(lp);
  e = cudaLaunchKernelEx(&cfg, k, sh, P);
  count_launch();
  return e != cudaSuccess ? e : cudaGetLastError();
}

template <int Q, int G, class SS>
static cudaError_t row_gs(const LinePlan& lp, const RowParams& P, bool blocked, cudaStream_t s) {
  if (blocked) return launch_cluster<line_kernel_cluster<Q, G, SS, RowProblemGeneric<Q, true>, RowParams>, G>(lp, P, s);
  return launch_cluster<line_kernel_cluster<Q, G, SS, RowProblemGeneric<Q, false>, RowParams>, G>(lp, P, s);
}
template <int Q, int G, class SS>
static cudaError_t col_gs(const LinePlan& lp, const ColParams& P, bool packed, cudaStream_t s) {
  if (packed) return launch_cluster<line_kernel_cluster<Q, G, SS, ColProblem<Q, true, false>, ColParams>, G>(lp, P, s);
  return launch_cluster<line_kernel_cluster<Q, G, SS, ColProblem<Q, false, false>, ColParams>, G>(lp, P, s);
}
// compile-time stage lists exist for (34560: G = 2 one sequence, G = 4 two sequences) and (69120: G = 4 one sequence)
template <int G>
static cudaError_t row_g(const LinePlan& lp, const RowParams& P, bool blocked, cudaStream_t s) {
  if constexpr (G == 2)
    if (lp.nseq == 1 && plan_matches<Shape34560>(lp, 2)) return row_gs<1, 2, Shape34560>(lp, P, blocked, s);
  if constexpr (G == 4) {
    if (lp.nseq == 2 && plan_matches<Shape34560>(lp, 4)) return row_gs<2, 4, Shape34560>(lp, P, blocked, s);
    if (lp.nseq == 1 && plan_matches<Shape69120>(lp, 4)) return row_gs<1, 4, Shape69120>(lp, P, blocked, s);
  }
  return lp.nseq == 2 ? row_gs<2, G, void>(lp, P, blocked, s) : row_gs<1, G, void>(lp, P, blocked, s);
}
template <int G>
static cudaError_t col_g(const LinePlan& lp, const ColParams& P, bool packed, cudaStream_t s) {
  if constexpr (G == 2)
    if (lp.nseq == 1 && plan_matches<Shape34560>(lp, 2)) return col_gs<1, 2, Shape34560>(lp, P, packed, s);
  if constexpr (G == 4) {
    if (lp.nseq == 2 && plan_matches<Shape34560>(lp, 4)) return col_gs<2, 4, Shape34560>(lp, P, packed, s);
    if (lp.nseq == 1 && plan_matches<Shape69120>(lp, 4)) return col_gs<1, 4, Shape69120>(lp, P, packed, s);
  }
  return lp.nseq == 2 ? col_gs<2, G, void>(lp, P, packed, s) : col_gs<1, G, void>(lp, P, packed, s);
}

cudaError_t launch_row_cluster(const LinePlan& lp, const RowParams& P, bool blocked, cudaStream_t s) {
  switch (lp.cluster) {
    case 2: return row_g<2>(lp, P, blocked, s);
    case 4: return row_g<4>(lp, P, blocked, s);
    case 8: return row_g<8>(lp, P, blocked, s);
  }
  return cudaErrorInvalidValue;
}

// The column pass of a cluster plan reads the row-major intermediate (no panel gather).
cudaError_t launch_col_cluster(const LinePlan& lp, const ColParams& P, bool packed, cudaStream_t s) {
  switch (lp.cluster) {
    case 2: return col_g<2>(lp, P, packed, s);
    case 4: return col_g<4>(lp, P, packed, s);
    case 8: return col_g<8>(lp, P, packed, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace gblur
