// ols.h — host interface of the overlap-save row pass (ols_kernels.cu, ols_core.cuh).
// capi.cpp picks it automatically for single-frame fused runs where row_profitable() holds; GBLUR_ROW_OLS=1 forces
// it wherever row_eligible() holds, GBLUR_ROW_OLS=0 turns it off.
#pragma once
#include <cuda_runtime.h>

#include <vector>

#include "kernels.h"

namespace gblur {
namespace ols {

struct RowPlan;  // device tables: twiddles, block gains, Nyquist coefficient, per-row alternating sums

// taps: the reference's normalised taps; n_ref: the reference's transform length along a row
// n_ref_col: its transform length along a column (0: the column pass is not wanted)
// dil: 1, or 3 for rows of three interleaved channels (RGB) run as byte rows — the block gains then belong to the taps
// spread three samples apart (row pass only)
cudaError_t row_plan_create(RowPlan** out, const std::vector<float>& taps, int n_ref, int n_ref_col = 0, int dil = 1);
void row_plan_destroy(RowPlan* p);

// RGBA with all four channels processed (panel intermediate for a two-sequence column plan), one channel, or RGB with
// a plan made with dil = 3; no peer/split stores
bool row_eligible(const RowPlan* p, const RowLaunch& a);

// The 1024-sample blocks cost nblk * 1024 transform points per row against the whole line's n_ref: worth it for wide
// rows (>= 4096 pixels, where the whole-line kernel is down to 13 warps per SM) when the blocks waste <= 20 %.
// Measured on 7680x4320 RGBA sigma=20: 0.36 ms against 0.417 ms (profiles/r01/ols_rowpass_v2_check.jsonl).
bool row_profitable(const RowLaunch& a, int n_ref);

// two launches: alternating sums per row, then one warp per (row, block, channel pair)
cudaError_t launch_row(RowPlan* p, const RowLaunch& a, int num_sms, cudaStream_t stream);

// ---- column pass (same plan object) ----
// C == Cp == 4, one frame, panel intermediate without bank-padding holes, the float-column range on 16-column
// boundaries, 16-byte aligned output rows
bool col_eligible(const RowPlan* p, const ColLaunch& a);
bool col_profitable(const ColLaunch& a, int n_ref_col);
// both passes as overlap-save blocks (what the automatic rule of fused single-frame runs asks): row_len pixels per row,
// col_len rows, the reference's transform lengths along a row / a column
bool both_profitable(int row_len, int col_len, int pad, int n_ref_row, int n_ref_col);
// The column terms of the reference's Nyquist behaviour, once per image (two launches: column sums of the u8 input =
// the synthetic row u, then the row pass of u on one CTA). d_in / in_pitch: the image the intermediate comes from.
cudaError_t launch_col_terms(RowPlan* p, const ColLaunch& a, const uint8_t* d_in, long long in_pitch, cudaStream_t stream);
// The two halves of launch_col_terms for halo-sharded runs: the column sums of image rows [row_begin, row_end) (d_in
// points at image row row_base; rows = the image's row count) ADDED into d_u, and the terms from a complete d_u.
cudaError_t launch_colsum(const ColLaunch& a, const uint8_t* d_in, long long in_pitch, int* d_u, int row_begin, int row_end,
                          int row_base, int rows, cudaStream_t stream);
cudaError_t launch_col_terms_from_u(RowPlan* p, const ColLaunch& a, const int* d_u, cudaStream_t stream);
// The block kernel over the float columns [a.w_begin, a.w_end) (TMA tile loads and stores, ols_core.cuh).
cudaError_t launch_col(RowPlan* p, const ColLaunch& a, int num_sms, cudaStream_t stream);

}  // namespace ols
}  // namespace gblur
