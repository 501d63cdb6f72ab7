// generic_row.cu — runtime-radix row-pass kernels (any valid FFT length).
#include "line_kernels.cuh"

namespace gblur {
cudaError_t launch_row_generic(const LinePlan& lp, const RowParams& P, bool rgba, int num_sms, cudaStream_t s) {
  if (lp.nseq == 2) {
    if (rgba) return launch_line<line_kernel_generic<2, RowProblemRGBA, RowParams>>(lp, P, lp.nthreads, num_sms, s);
    return launch_line<line_kernel_generic<2, RowProblemGeneric<2>, RowParams>>(lp, P, lp.nthreads, num_sms, s);
  }
  return launch_line<line_kernel_generic<1, RowProblemGeneric<1>, RowParams>>(lp, P, lp.nthreads, num_sms, s);
}
}  // namespace gblur
