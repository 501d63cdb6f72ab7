// generic_row.cu — runtime-radix row-pass kernels (any valid FFT length).
#include "line_kernels.cuh"

namespace gblur {
cudaError_t launch_row_generic(const LinePlan& lp, const RowParams& P, int fast, bool blocked, int num_sms,
                               cudaStream_t s) {
  const int nt = lp.nthreads;
  const bool rgba = fast == kRowFastRGBA;
  if (lp.nseq == 2) {
    if (fast == kRowFastRGB3 && blocked)
      return launch_line<line_kernel_generic<2, RowProblemRGB3<false>, RowParams>>(lp, P, nt, num_sms, s);
    if (rgba && blocked) return launch_line<line_kernel_generic<2, RowProblemRGBA<true>, RowParams>>(lp, P, nt, num_sms, s);
    if (rgba) return launch_line<line_kernel_generic<2, RowProblemRGBA<false>, RowParams>>(lp, P, nt, num_sms, s);
    if (blocked) return launch_line<line_kernel_generic<2, RowProblemGeneric<2, true>, RowParams>>(lp, P, nt, num_sms, s);
    return launch_line<line_kernel_generic<2, RowProblemGeneric<2, false>, RowParams>>(lp, P, nt, num_sms, s);
  }
  if (blocked) return launch_line<line_kernel_generic<1, RowProblemGeneric<1, true>, RowParams>>(lp, P, nt, num_sms, s);
  return launch_line<line_kernel_generic<1, RowProblemGeneric<1, false>, RowParams>>(lp, P, nt, num_sms, s);
}
}  // namespace gblur
