// static_impl.cu — compiled once per entry of static_shapes.h with -DGBLUR_TU_SHAPE=<id>.
#include "line_kernels.cuh"
#include "static_shapes.h"

#ifndef GBLUR_TU_SHAPE
#error "compile with -DGBLUR_TU_SHAPE=<id>"
#endif

namespace gblur {

template <int ID>
struct ShapeOf;
#define X(ID, NSEQ_, R0, R1, R2, R3, NT_, MINB_)     \
  template <>                                        \
  struct ShapeOf<ID> {                               \
    using SS = StaticShape<R0, R1, R2, R3>;          \
    static constexpr int NSEQ = NSEQ_, NT = NT_, MINB = MINB_; \
  };
GBLUR_STATIC_SHAPES(X)
#undef X

template <int ID>
cudaError_t launch_row_static(const LinePlan& lp, const RowParams& P, bool rgba, int num_sms, cudaStream_t s) {
  using S = ShapeOf<ID>;
  if constexpr (S::NSEQ == 2) {
    if (rgba)
      return launch_line<line_kernel_static<2, typename S::SS, S::NT, S::MINB, RowProblemRGBA, RowParams>>(lp, P,
                         S::NT, num_sms, s);
  }
  return launch_line<line_kernel_static<S::NSEQ, typename S::SS, S::NT, S::MINB, RowProblemGeneric<S::NSEQ>, RowParams>>(lp, P,
      S::NT, num_sms, s);
}

template <int ID>
cudaError_t launch_col_static(const LinePlan& lp, const ColParams& P, bool packed, int num_sms, cudaStream_t s) {
  using S = ShapeOf<ID>;
  if (packed)
    return launch_line<line_kernel_static<S::NSEQ, typename S::SS, S::NT, S::MINB, ColProblem<S::NSEQ, true>, ColParams>>(lp, P,
        S::NT, num_sms, s);
  return launch_line<line_kernel_static<S::NSEQ, typename S::SS, S::NT, S::MINB, ColProblem<S::NSEQ, false>, ColParams>>(lp, P,
      S::NT, num_sms, s);
}

template cudaError_t launch_row_static<GBLUR_TU_SHAPE>(const LinePlan&, const RowParams&, bool, int, cudaStream_t);
template cudaError_t launch_col_static<GBLUR_TU_SHAPE>(const LinePlan&, const ColParams&, bool, int, cudaStream_t);

}  // namespace gblur
