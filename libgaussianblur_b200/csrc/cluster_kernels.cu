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

template <int NSEQ, int G, class Problem, class Params>
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
  const int nloc = sh.N / G, M0 = sh.M[0];
  const float* kperm = sh.kperm + (long long)rank * nloc;
  const int tid = (int)threadIdx.x, nt = (int)blockDim.x;
  cl.sync();  // every CTA of the cluster is resident before anybody writes into a peer
  for (int unit = cid; unit < P.nunits; unit += nclusters) {
    typename Problem::Loader ld;
    typename Problem::Storer st;
    Problem::make(P, unit, ld, st);
    cluster_fwd_stage0<NSEQ, G>(ld, peer, sh.tw[0], M0, sh.padq, rank, tid, nt);
    cl.sync();
    run_local_stages<NSEQ>(sh, nloc, kperm, ld, st, sm, tid, nt);
    cl.sync();
    cluster_inv_stage0<NSEQ, G>(st, peer, sh.tw[0], M0, sh.padq, rank, tid, nt);
    cl.sync();  // peers have finished reading this CTA's buffer
  }
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

template <int G>
static cudaError_t row_g(const LinePlan& lp, const RowParams& P, bool blocked, cudaStream_t s) {
  if (blocked) return launch_cluster<line_kernel_cluster<1, G, RowProblemGeneric<1, true>, RowParams>, G>(lp, P, s);
  return launch_cluster<line_kernel_cluster<1, G, RowProblemGeneric<1, false>, RowParams>, G>(lp, P, s);
}
template <int G>
static cudaError_t col_g(const LinePlan& lp, const ColParams& P, bool packed, cudaStream_t s) {
  if (packed) return launch_cluster<line_kernel_cluster<1, G, ColProblem<1, true, false>, ColParams>, G>(lp, P, s);
  return launch_cluster<line_kernel_cluster<1, G, ColProblem<1, false, false>, ColParams>, G>(lp, P, s);
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
