// static_impl.cu — compiled once per entry of static_shapes.h with -DGBLUR_TU_SHAPE=<id>.
#include <cstdlib>

#include "line_kernels.cuh"
#include "static_shapes.h"

#ifndef GBLUR_TU_SHAPE
#error "compile with -DGBLUR_TU_SHAPE=<id>"
#endif

namespace gblur {

template <int ID>
struct ShapeOf;
#define X(ID, NSEQ_, R0, R1, R2, R3, NT_, MAXREG_)     \
  template <>                                        \
  struct ShapeOf<ID> {                               \
    using SS = StaticShape<R0, R1, R2, R3>;          \
    static constexpr int NSEQ = NSEQ_, NT = NT_, MAXREG = MAXREG_; \
  };
GBLUR_STATIC_SHAPES(X)
#undef X

template <int ID>
cudaError_t launch_row_static(const LinePlan& lp, const RowParams& P, int fast, bool blocked, int num_sms,
                              cudaStream_t s) {
  using S = ShapeOf<ID>;
  using SS = typename S::SS;
  const bool rgba = fast == kRowFastRGBA;
  if constexpr (S::NSEQ == 2) {
    if (fast == kRowFastRGB3 && blocked) {
      // staged input: rows r0 and r0 + 1 of the next unit, contiguous in memory
      const size_t stage_bytes = (size_t)P.in_pitch * 2;
      const bool staged = (P.in_pitch % 16) == 0 && P.in_pitch == (long long)P.T * P.C &&
                          (reinterpret_cast<uintptr_t>(P.in) % 16) == 0 &&
                          lp.smem_bytes + stage_bytes <= (size_t)227 * 1024 - 64 && !getenv("GBLUR_NO_STAGE");
      if (staged)
        return launch_line<line_kernel_static<2, SS, S::NT, S::MAXREG, RowProblemRGB3<true>, RowParams>>(
            lp, P, S::NT, num_sms, s, stage_bytes);
      return launch_line<line_kernel_static<2, SS, S::NT, S::MAXREG, RowProblemRGB3<false>, RowParams>>(lp, P, S::NT,
                                                                                                        num_sms, s);
    }
    // staged input: the next row is bulk-copied into shared memory behind the tables while this one is
    // transformed; needs 16-byte aligned rows and room for T*4 bytes
    const size_t stage_bytes = (size_t)P.T * 4;
    const bool staged = rgba && blocked && (stage_bytes % 16) == 0 && (P.in_pitch % 16) == 0 &&
                        (reinterpret_cast<uintptr_t>(P.in) % 16) == 0 &&
                        lp.smem_bytes + stage_bytes <= (size_t)227 * 1024 - 64 && !getenv("GBLUR_NO_STAGE");
    if (staged)
      return launch_line<line_kernel_static<2, SS, S::NT, S::MAXREG, RowProblemRGBA<true, true>, RowParams>>(
          lp, P, S::NT, num_sms, s, stage_bytes);
    if (rgba && blocked)
      return launch_line<line_kernel_static<2, SS, S::NT, S::MAXREG, RowProblemRGBA<true>, RowParams>>(lp, P, S::NT,
                                                                                                       num_sms, s);
    if (rgba)
      return launch_line<line_kernel_static<2, SS, S::NT, S::MAXREG, RowProblemRGBA<false>, RowParams>>(lp, P, S::NT,
                                                                                                        num_sms, s);
  }
  if constexpr (S::NSEQ == 1) {
    if (rgba && blocked)
      return launch_line<line_kernel_static<1, SS, S::NT, S::MAXREG, RowProblemRGBA1<true>, RowParams>>(lp, P, S::NT,
                                                                                                        num_sms, s);
    if (rgba)
      return launch_line<line_kernel_static<1, SS, S::NT, S::MAXREG, RowProblemRGBA1<false>, RowParams>>(lp, P, S::NT,
                                                                                                         num_sms, s);
  }
  if (blocked)
    return launch_line<line_kernel_static<S::NSEQ, SS, S::NT, S::MAXREG, RowProblemGeneric<S::NSEQ, true>, RowParams>>(
        lp, P, S::NT, num_sms, s);
  return launch_line<line_kernel_static<S::NSEQ, SS, S::NT, S::MAXREG, RowProblemGeneric<S::NSEQ, false>, RowParams>>(
      lp, P, S::NT, num_sms, s);
}

template <int ID>
cudaError_t launch_col_static(const LinePlan& lp, const ColParams& P, bool packed, int input, int num_sms,
                              cudaStream_t s) {
  using S = ShapeOf<ID>;
  using SS = typename S::SS;
  constexpr int Q = S::NSEQ;
  if constexpr (SS::M0 % S::NT == 0) {  // whole passes of last-stage butterflies: the TMA prefetch protocol applies
    if (input == 2) {
      if (packed)
        return launch_line<line_kernel_static<Q, SS, S::NT, S::MAXREG, ColProblem<Q, true, true, true>, ColParams>>(
            lp, P, S::NT, num_sms, s);
      return launch_line<line_kernel_static<Q, SS, S::NT, S::MAXREG, ColProblem<Q, false, true, true>, ColParams>>(
          lp, P, S::NT, num_sms, s);
    }
  }
  if (input >= 1) {
    if (packed)
      return launch_line<line_kernel_static<Q, SS, S::NT, S::MAXREG, ColProblem<Q, true, true>, ColParams>>(
          lp, P, S::NT, num_sms, s);
    return launch_line<line_kernel_static<Q, SS, S::NT, S::MAXREG, ColProblem<Q, false, true>, ColParams>>(
        lp, P, S::NT, num_sms, s);
  }
  if (packed)
    return launch_line<line_kernel_static<Q, SS, S::NT, S::MAXREG, ColProblem<Q, true, false>, ColParams>>(
        lp, P, S::NT, num_sms, s);
  return launch_line<line_kernel_static<Q, SS, S::NT, S::MAXREG, ColProblem<Q, false, false>, ColParams>>(
      lp, P, S::NT, num_sms, s);
}

template cudaError_t launch_row_static<GBLUR_TU_SHAPE>(const LinePlan&, const RowParams&, int, bool, int,
                                                       cudaStream_t);
template cudaError_t launch_col_static<GBLUR_TU_SHAPE>(const LinePlan&, const ColParams&, bool, int, int,
                                                       cudaStream_t);

}  // namespace gblur
