// generic_col.cu — runtime-radix column-pass kernels (any valid FFT length).
#include "line_kernels.cuh"

namespace gblur {
cudaError_t launch_col_generic(const LinePlan& lp, const ColParams& P, bool packed, int num_sms, cudaStream_t s) {
  if (lp.nseq == 2) {
    if (packed) return launch_line<line_kernel_generic<2, ColProblem<2, true>, ColParams>>(lp, P, lp.nthreads, num_sms, s);
    return launch_line<line_kernel_generic<2, ColProblem<2, false>, ColParams>>(lp, P, lp.nthreads, num_sms, s);
  }
  if (packed) return launch_line<line_kernel_generic<1, ColProblem<1, true>, ColParams>>(lp, P, lp.nthreads, num_sms, s);
  return launch_line<line_kernel_generic<1, ColProblem<1, false>, ColParams>>(lp, P, lp.nthreads, num_sms, s);
}
}  // namespace gblur
