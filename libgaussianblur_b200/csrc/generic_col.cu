// generic_col.cu — runtime-radix column-pass kernels (any valid FFT length).
#include "line_kernels.cuh"

namespace gblur {
template <int Q>
static cudaError_t launch_q(const LinePlan& lp, const ColParams& P, bool packed, bool blocked, int num_sms,
                            cudaStream_t s) {
  const int nt = lp.nthreads;
  if (blocked) {
    if (packed) return launch_line<line_kernel_generic<Q, ColProblem<Q, true, true>, ColParams>>(lp, P, nt, num_sms, s);
    return launch_line<line_kernel_generic<Q, ColProblem<Q, false, true>, ColParams>>(lp, P, nt, num_sms, s);
  }
  if (packed) return launch_line<line_kernel_generic<Q, ColProblem<Q, true, false>, ColParams>>(lp, P, nt, num_sms, s);
  return launch_line<line_kernel_generic<Q, ColProblem<Q, false, false>, ColParams>>(lp, P, nt, num_sms, s);
}

cudaError_t launch_col_generic(const LinePlan& lp, const ColParams& P, bool packed, bool blocked, int num_sms,
                               cudaStream_t s) {
  return lp.nseq == 2 ? launch_q<2>(lp, P, packed, blocked, num_sms, s) : launch_q<1>(lp, P, packed, blocked, num_sms, s);
}
}  // namespace gblur
