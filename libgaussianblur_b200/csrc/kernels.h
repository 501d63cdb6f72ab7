// kernels.h — host-visible launch interface of the sm_100a line-convolution kernels.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

namespace gblur {

constexpr int kHostMaxStages = 5;

// Host mirror of the device LineShape plus launch facts, filled by the planner (plan.cpp).
struct LinePlan {
  int N = 0;
  int nstages = 0;
  int radix[kHostMaxStages] = {0, 0, 0, 0, 0};
  int M[kHostMaxStages] = {0, 0, 0, 0, 0};
  int padq = 0;
  const float2* d_tw[kHostMaxStages] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  const float* d_kperm = nullptr;
  int nseq = 2;          // complex sequences per line unit (1 or 2)
  int nthreads = 0;      // CTA size for the runtime-radix kernel
  int static_id = -1;    // index into the compile-time shape table, -1 = runtime-radix kernel
  size_t smem_bytes = 0; // dynamic shared memory per CTA (exchange buffer + in-smem tables)
  int exch_elems = 0;    // exchange-buffer slots (N + bank-padding slots); per CTA for cluster plans
  int cluster = 1;       // CTAs per line unit: > 1 = thread-block cluster, the line is split over their shared memories
};

// Panel intermediate [frame][panel = w/32][np slots][32 floats] (see problems.cuh): the row pass
// writes whole 128-byte lines, the column pass gathers its L floats per slot with cp.async.
// enabled == 0: row-major [rows][pitch].
struct BlockedLayout {
  int enabled = 0;
  int L = 4;                   // floats per column group = 2 * column nseq
  int np = 0;                  // slots per unit (column FFT length + bank padding slots)
  int padq = 0;                // column plan's padq
  int pair_major = 0;          // 1: panel rows ordered [pair][pixel][2] (overlap-save row AND column pass, ols_core.cuh)
  int ols_pair = 0;            // the layout was chosen for the overlap-save column pass: the row pass runs as blocks as well
  int slot_origin = 0;         // halo-sharded runs: the buffer holds the padded-column slots [slot_origin, slot_origin + np)
  int rows_per_frame = 0;
  long long frame_stride = 0;  // floats per frame = panels * np * 32
};

struct RowLaunch {
  const uint8_t* in;
  long long in_pitch;
  int C, Cp;
  int T, pad;
  long long rows_total;
  float* inter;
  long long inter_pitch;
  BlockedLayout blk;
  int split_cols = 0;          // row-major only: see RowParams::split_cols
  long long split_stride = 0;
  float* peer[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  int npeers = 0;              // panel layout only: see RowParams::peer
  int peer_panels = 0;
  int row_offset = 0;
  // overlap-save row pass only: the launch covers two row ranges; logical rows >= split_at lie split_gap rows further on
  // in the input (and in the frame's row numbering). split_gap == 0: one range.
  long long split_at = 0, split_gap = 0;
};

struct ColLaunch {
  const float* inter;
  long long pitch;
  long long frame_stride;
  int frames;
  int T, pad;
  int W;            // valid float columns of the whole image (cols * Cp)
  int w_begin = 0;  // column-band launch: float columns [w_begin, w_end) only, w_begin a multiple of 32
  int w_end = 0;    // 0 = W
  uint8_t* out;
  long long out_pitch;
  long long out_frame_stride;
  int C, Cp;
  BlockedLayout blk;
  int panel_in = 0;  // the column pass of a cluster plan reads the panel layout through its loader
  int row_origin = 0;  // halo-sharded runs: image row of the first row of `out` (and of slot blk.slot_origin + pad)
  // overlap-save column pass of an RGBA image whose alpha is NOT blurred: all four channels run through both passes and
  // the column kernel puts the ORIGINAL alpha byte (row r, pixel x: alpha_src[r * alpha_pitch + 4 x + 3]) into its output
  // tile; alpha_src may be the output buffer itself (in place: an item reads exactly the bytes it is about to overwrite)
  const uint8_t* alpha_src = nullptr;
  long long alpha_pitch = 0;
  const CUtensorMap* tmap = nullptr;  // host copy of the tensor map over the panel intermediate (TMA gather), or null
  int tma_rows = 0;
};

// Returns the static shape id for (N, nseq) or -1; fills radices when found.
int find_static_shape(int N, int nseq, int* radices /*[4]*/, int* nthreads);

// Tensor map over a panel intermediate for column units of L floats: dims {32, np, panels}, box {L, rows, 1}.
// Returns false when the driver entry point is missing or np has no usable divisor (then cp.async is used).
bool make_panel_tensor_map(CUtensorMap* out, const float* base, int np, long long panels, int L, int* box_rows);

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link dependency on libcuda). rank <= 3;
// dims / box fastest dimension first, strides_bytes for dimensions 1 .. rank-1. swizzle_bytes: 0, 32, 64 or 128.
// elem_bytes: 4 = FLOAT32, 1 = UINT8. Returns false when the entry point is missing or the encode fails.
bool encode_tensor_map(CUtensorMap* out, int elem_bytes, int rank, const void* base, const unsigned long long* dims,
                       const unsigned long long* strides_bytes, const unsigned* box, int swizzle_bytes);

// One-time per-process kernel attribute setup (max dynamic smem). Returns cudaSuccess or error.
cudaError_t kernels_init(int device);

cudaError_t launch_rowpass(const LinePlan& lp, const RowLaunch& a, int num_sms, cudaStream_t stream);
cudaError_t launch_colpass(const LinePlan& lp, const ColLaunch& a, int num_sms, cudaStream_t stream);

// Copies channel `ch` of every pixel from src to dst (alpha pass-through when out != in).
cudaError_t launch_copy_channel(const uint8_t* src, uint8_t* dst, long long npixels, int C, int ch,
                                cudaStream_t stream);

// counts of kernel launches issued through this interface (for bench.py's gpu_launches)
unsigned long long launch_counter();

}  // namespace gblur
