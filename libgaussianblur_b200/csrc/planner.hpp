// planner.hpp — host-only: choose stage radices for an FFT length and build the twiddle and
// permuted-spectrum tables the kernels consume. Replaces pffft_new_setup's decompose()/rffti1_ps
// (/root/reference/include/pffft_pommier/pffft.c:1073-1152,1231-1280) for the GPU path.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdlib>
#include <vector>

#include "hostsem.hpp"
#include "kernels.h"

namespace gblur {
namespace host {

constexpr size_t kMaxSmemBytes = 227 * 1024 - 64;  // 64 bytes left for static shared (mbarrier)
constexpr double kTwoPi = 6.28318530717958647692528676655900577;

struct Factorization {
  int nstages = 0;
  int radix[kHostMaxStages] = {0, 0, 0, 0, 0};
};

// radices the runtime-radix kernel instantiates (keep in sync with GBLUR_FOR_EACH_GENERIC_RADIX)
inline const std::vector<int>& generic_radices() {
  static const std::vector<int> r = {16, 15, 12, 10, 9, 8, 6, 5, 4, 3, 2};
  return r;
}

namespace detail {
inline void search(int rem, int max_r, int depth, int* cur, Factorization& best, long long& best_cost) {
  if (rem == 1) {
    if (depth < 2) return;
    long long sum = 0;
    for (int i = 0; i < depth; ++i) sum += cur[i];
    const long long cost = (long long)depth * 1000 + sum;  // fewest stages, then most balanced
    if (cost < best_cost) {
      best_cost = cost;
      best.nstages = depth;
      for (int i = 0; i < kHostMaxStages; ++i) best.radix[i] = i < depth ? cur[i] : 1;
    }
    return;
  }
  if (depth == kHostMaxStages) return;
  for (int r : generic_radices()) {
    if (r > max_r || rem % r) continue;
    cur[depth] = r;
    search(rem / r, r, depth + 1, cur, best, best_cost);
  }
}
}  // namespace detail

// Fewest stages, most balanced radices; a radix divisible by 4 goes last when there is one
// (its per-thread spectrum gains are then 16-byte vectorisable).
inline bool factorize_generic(int N, Factorization& f) {
  int cur[kHostMaxStages];
  long long best_cost = 1LL << 60;
  detail::search(N, 16, 0, cur, f, best_cost);
  if (f.nstages == 0) return false;
  for (int i = f.nstages - 1; i >= 0; --i) {
    if (f.radix[i] % 4 == 0) {
      std::swap(f.radix[i], f.radix[f.nstages - 1]);
      break;
    }
  }
  return true;
}

inline int exch_elems_for(int N, int rlast) {
  const int padq = (rlast % 2 == 0) ? rlast : 0;
  return N + (padq ? N / padq : 0);
}

// factored twiddle table of a radix-R stage: (B-1) fine + (A-1) coarse powers (fftconv.cuh TwSplit)
inline void tw_split(int R, int& A, int& B) {
  B = (R % 4 == 0) ? 4 : (R % 5 == 0) ? 5 : (R % 3 == 0) ? 3 : (R % 2 == 0) ? 2 : R;
  A = R / B;
}
inline int tw_nbase(int R) {
  int A, B;
  tw_split(R, A, B);
  return (B - 1) + (A - 1);
}

// Fill everything of a LinePlan that does not live on the device. Returns false when the length
// does not fit one CTA's shared memory even with a single sequence.
// cluster_nseq: preferred sequences per unit when the line needs a cluster (the row pass runs best with one —
// byte loads from 2 rows, G = 2 — the column pass with two — 16-byte gathers, 4-byte stores); 0 = no preference
inline bool plan_line(int N, LinePlan& lp, int cluster_nseq = 0) {
  lp = LinePlan{};
  lp.N = N;
  for (int nseq = 2; nseq >= 1; --nseq) {
    Factorization f;
    int st_r[4], st_threads = 0;
    const int sid = find_static_shape(N, nseq, st_r, &st_threads);
    if (sid == -2) continue;  // a tuning variant with the other sequence count was forced
    if (sid >= 0) {
      f.nstages = 0;
      for (int i = 0; i < 4; ++i)
        if (st_r[i] > 1) f.radix[f.nstages++] = st_r[i];
    } else if (!factorize_generic(N, f)) {
      return false;
    }
    const int rlast = f.radix[f.nstages - 1];
    const int exch = exch_elems_for(N, rlast);
    size_t smem = (size_t)exch * 8 * (size_t)nseq;
    if (sid >= 0) {  // compile-time kernels keep the stage tables and the spectrum in shared memory
      int m = N;
      for (int s = 0; s < f.nstages - 1; ++s) {
        m /= f.radix[s];
        smem += (size_t)tw_nbase(f.radix[s]) * (size_t)m * 8;
      }
      smem += (size_t)N * 4 + 48;  // + alignment slack and the two mbarriers (TMA gather, staged row input)
    }
    if (smem > kMaxSmemBytes) continue;
    lp.exch_elems = exch;
    lp.nseq = nseq;
    lp.nstages = f.nstages;
    int m = N, max_nb = 1;
    for (int s = 0; s < f.nstages; ++s) {
      lp.radix[s] = f.radix[s];
      m /= f.radix[s];
      lp.M[s] = m;
      max_nb = std::max(max_nb, N / f.radix[s]);
    }
    lp.padq = (rlast % 2 == 0) ? rlast : 0;
    lp.smem_bytes = smem;
    lp.static_id = sid;
    lp.nthreads = sid >= 0 ? st_threads : std::min(512, std::max(32, ((max_nb + 31) / 32) * 32));
    return true;
  }
  // Too long for one CTA: split the line over a thread-block cluster of G CTAs (distributed
  // shared memory). Stage 0 is radix 16 and crosses CTAs; after it the line falls apart into 16
  // independent sub-transforms of length N/16, 16/G of them per CTA, which run locally.
  // Two sequences per cluster (16-byte pieces for the column gather, shared twiddles) when that fits a
  // cluster of at most 4 CTAs, else one sequence. GBLUR_CLUSTER_NSEQ=1|2 forces the choice (tuning).
  static const int forced_nseq = [] {
    const char* e = getenv("GBLUR_CLUSTER_NSEQ");
    return e ? atoi(e) : 0;
  }();
  struct Cand {
    int G, nseq;
  };
  const Cand order2[] = {{2, 2}, {4, 2}, {2, 1}, {4, 1}, {8, 2}, {8, 1}};
  const Cand order1[] = {{2, 1}, {4, 1}, {2, 2}, {4, 2}, {8, 1}, {8, 2}};
  const Cand* order = cluster_nseq == 1 ? order1 : order2;
  for (int ci = 0; ci < 6; ++ci) {
    const Cand cand = order[ci];
    const int G = cand.G;
    if (forced_nseq && cand.nseq != forced_nseq) continue;
    if (N % (16 * G)) continue;
    Factorization rest;
    int cur[kHostMaxStages];
    long long best_cost = 1LL << 60;
    // at most kHostMaxStages - 1 further stages
    struct Local {
      static void go(int rem, int max_r, int depth, int* cur, Factorization& best, long long& best_cost) {
        if (rem == 1) {
          if (depth < 1) return;
          long long sum = 0;
          for (int i = 0; i < depth; ++i) sum += cur[i];
          const long long cost = (long long)depth * 1000 + sum;
          if (cost < best_cost) {
            best_cost = cost;
            best.nstages = depth;
            for (int i = 0; i < kHostMaxStages; ++i) best.radix[i] = i < depth ? cur[i] : 1;
          }
          return;
        }
        if (depth == kHostMaxStages - 1) return;
        for (int r : generic_radices()) {
          if (r > max_r || rem % r) continue;
          cur[depth] = r;
          go(rem / r, r, depth + 1, cur, best, best_cost);
        }
      }
    };
    Local::go(N / 16, 16, 0, cur, rest, best_cost);
    if (rest.nstages == 0) continue;
    for (int i = rest.nstages - 1; i >= 0; --i) {
      if (rest.radix[i] % 4 == 0) {
        std::swap(rest.radix[i], rest.radix[rest.nstages - 1]);
        break;
      }
    }
    const int rlast = rest.radix[rest.nstages - 1];
    const int nloc = N / G;
    const int exch = exch_elems_for(nloc, rlast);
    const size_t smem = (size_t)exch * 8 * (size_t)cand.nseq;
    if (smem > kMaxSmemBytes) continue;
    lp.cluster = G;
    lp.nseq = cand.nseq;
    lp.exch_elems = exch;
    lp.nstages = rest.nstages + 1;
    lp.radix[0] = 16;
    lp.M[0] = N / 16;
    int m = N / 16;
    for (int s = 0; s < rest.nstages; ++s) {
      lp.radix[s + 1] = rest.radix[s];
      m /= rest.radix[s];
      lp.M[s + 1] = m;
    }
    lp.padq = (rlast % 2 == 0) ? rlast : 0;
    lp.smem_bytes = smem;
    lp.static_id = -1;
    lp.nthreads = 512;
    return true;
  }
  return false;
}

// Factored stage tables, concatenated over s < S-1: tw[s][e*M + m] = exp(-2 pi i m j_e / (R*M))
// with j_e = e+1 for e < B-1 (fine powers) and j_e = (e-(B-1)+1)*B after that (coarse powers).
inline void build_twiddles(const LinePlan& lp, std::vector<float>& flat, size_t offsets[kHostMaxStages]) {
  flat.clear();
  for (int s = 0; s < lp.nstages - 1; ++s) {
    const int R = lp.radix[s], M = lp.M[s];
    const long long BM = (long long)R * M;
    int A, B;
    tw_split(R, A, B);
    offsets[s] = flat.size() / 2;
    for (int e = 0; e < (B - 1) + (A - 1); ++e) {
      const int j = e < B - 1 ? e + 1 : (e - (B - 1) + 1) * B;
      for (int m = 0; m < M; ++m) {
        const long long k = ((long long)m * j) % BM;
        const double a = kTwoPi * (double)k / (double)BM;
        flat.push_back((float)std::cos(a));
        flat.push_back((float)(-std::sin(a)));
      }
    }
  }
}

// Gains in the digit-reversed position order the forward (decimation-in-frequency) stages leave
// the spectrum in, pre-scaled by 1/N, with the reference's Nyquist behaviour: the bin N/2 is
// multiplied by K[0] (src/gaussianblur.cpp:62-79 with pffft.h:143-146).
inline std::vector<float> build_kperm(const LinePlan& lp, const std::vector<double>& gains) {
  const int N = lp.N;
  std::vector<float> kp((size_t)N);
  for (int p = 0; p < N; ++p) {
    int rem = p;
    long long k = 0, weight = 1;
    for (int s = 0; s < lp.nstages; ++s) {
      const int digit = rem / lp.M[s];
      rem -= digit * lp.M[s];
      k += digit * weight;
      weight *= lp.radix[s];
    }
    const int kk = (int)std::min<long long>(k, N - k);
    const double g = (kk == N / 2) ? gains[0] : gains[(size_t)kk];
    kp[(size_t)p] = (float)(g / (double)N);
  }
  return kp;
}

}  // namespace host
}  // namespace gblur
