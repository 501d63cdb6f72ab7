// ols.h — host interface of the overlap-save row pass (ols_kernels.cu, ols_core.cuh).
// Experimental in round 1: selected with GBLUR_ROW_OLS=1 (capi.cpp), default off.
#pragma once
#include <cuda_runtime.h>

#include <vector>

#include "kernels.h"

namespace gblur {
namespace ols {

struct RowPlan;  // device tables: twiddles, block gains, Nyquist coefficient, per-row alternating sums

// taps: the reference's normalised taps; n_ref: the reference's transform length along a row
cudaError_t row_plan_create(RowPlan** out, const std::vector<float>& taps, int n_ref);
void row_plan_destroy(RowPlan* p);

// RGBA with all four channels processed, panel intermediate for a two-sequence column plan, no peer/split stores
bool row_eligible(const RowPlan* p, const RowLaunch& a);

// two launches: alternating sums per row, then one warp per (row, block, channel pair)
cudaError_t launch_row(RowPlan* p, const RowLaunch& a, int num_sms, cudaStream_t stream);

}  // namespace ols
}  // namespace gblur
