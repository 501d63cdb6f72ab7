/* gblur_capi.h — C ABI of the B200-native LibGaussianBlur hot path.
 *
 * This is the drop-in boundary: the only way host code (the C++ gaussianblur:: API in
 * include/gaussianblur/, the pybind11 module, the Python/ctypes layer, a foreign-language FFI)
 * reaches the sm_100a kernels. Plain pointers and sizes only; opaque plan handle; int status
 * (0 = ok, otherwise a GBLUR_E_* code with text from gblur_last_error()). There is NO CPU
 * fallback behind any of these calls: without a CUDA device they fail with GBLUR_E_CUDA.
 *
 * What each entry point replaces in the reference (paths under /root/reference/):
 *   gblur_plan_create        prepare_kernel_DFT                src/gaussianblur.cpp:100-172
 *                            (+ pffft_new_setup twiddles       include/pffft_pommier/pffft.c:1231-1280)
 *   gblur_plan_info          KernelDFT.pad / .trailing_zeros   include/gaussianblur/helpers.hpp:125-132
 *   gblur_plan_spectrum      KernelDFT.kerf_1D_col / _row      include/gaussianblur/helpers.hpp:126-127
 *   gblur_run_host           gaussianblur::gaussianblur        src/gaussianblur.cpp:307-325
 *   gblur_run_device         the same, device-resident buffers (deinterleave :174-200, pffft driver
 *                            :246-287, process_channel_tiles :202-244, interleave :289-305)
 *   gblur_run_batch_device   a loop of the above over frames with ONE kernel spectrum (additive)
 *   gblur_rowpass_device     process_channel_tiles, first call  src/gaussianblur.cpp:268-272
 *   gblur_rowpass_split_device  the same first call, output laid out per destination rank (additive)
 *   gblur_colpass_device     process_channel_tiles, second call src/gaussianblur.cpp:275-279
 *   gblur_dump_intermediate  (test hook) the fp32 plane after pass 1
 *   gblur_mg_*               hybrid_loop, the reference's in-library fan-out  include/gaussianblur/helpers.hpp:134-172
 */
#ifndef GBLUR_CAPI_H_
#define GBLUR_CAPI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gblur_plan gblur_plan;

enum {
  GBLUR_OK = 0,
  GBLUR_E_INVALID = 1,     /* bad argument (null pointer, sigma <= 0, channels not in {1,3,4}, ...) */
  GBLUR_E_PRECOND = 2,     /* pad > min(rows, cols) - 1: the reference reads out of bounds here      */
  GBLUR_E_UNSUPPORTED = 3, /* FFT length does not fit one CTA's shared memory (yet)                  */
  GBLUR_E_CUDA = 4         /* CUDA runtime failure, including "no device"                            */
};

/* Geometry + sigma -> immutable plan on `device`: reference window/pad/FFT-length rules, kernel
 * spectra, twiddles, workspace for `max_frames` frames. channels in {1,3,4} (1 is additive). */
int gblur_plan_create(gblur_plan** out, int device, int rows, int cols, int channels, float sigma);
int gblur_plan_create_batch(gblur_plan** out, int device, int rows, int cols, int channels, float sigma,
                            int max_frames);
/* Band plan for row-band sharding of one (full_rows x full_cols) image: window, pad and FFT
 * lengths come from the FULL geometry; the row pass runs on `band_rows` rows, the column pass on
 * `band_cols` pixel columns. */
int gblur_plan_create_band(gblur_plan** out, int device, int full_rows, int full_cols, int channels, float sigma,
                           int band_rows, int band_cols);
void gblur_plan_destroy(gblur_plan* plan);

/* Any out pointer may be NULL. n_rowpass = FFT length along a row (the reference's sizes[1]),
 * n_colpass = along a column (sizes[0]); tz_* are KernelDFT.trailing_zeros.{rows,cols}. */
int gblur_plan_info(const gblur_plan* plan, int* ksize, int* pad, int* n_rowpass, int* n_colpass, int* tz_rows,
                    int* tz_cols);
/* axis 0: kerf_1D_col (n_rowpass floats), axis 1: kerf_1D_row (n_colpass floats); pffft "ordered"
 * layout [K0, K_{N/2}, Re1, Im1, ...] (include/pffft_pommier/pffft.h:143-146). */
int gblur_plan_spectrum(const gblur_plan* plan, int axis, float* host_out);
/* Stage radices chosen for an axis (for logs/tests): writes up to 5 ints, returns count via *n. */
int gblur_plan_radices(const gblur_plan* plan, int axis, int* radices, int* n, int* static_kernel);
/* Which kernels the last row pass / column pass launched on this plan used (for logs/tests/bench.py):
 * 0 = whole-line transform of the reference's length (line_kernels.cuh), 1 = overlap-save over 1024-sample blocks
 * (ols_kernels.cu; picked automatically for single-frame runs of wide RGBA images, GBLUR_ROW_OLS / GBLUR_COL_OLS = 0|1
 * at plan creation override). Both produce the reference's result (src/gaussianblur.cpp:62-79,202-244). */
int gblur_plan_last_kernels(const gblur_plan* plan, int* row_kind, int* col_kind);

/* Host buffer in place: H2D, blur, D2H, synchronous on return. Large single images are pipelined against
 * PCIe (row chunks run under the H2D copy, column bands leave under the column pass); a batch longer than the
 * plan's max_frames is streamed in chunks of max_frames (H2D || blur || D2H). Pinned host memory is needed for
 * the overlap (pageable memory works but serialises). */
int gblur_run_host(gblur_plan* plan, uint8_t* hwc_inout, int apply_to_alpha);
int gblur_run_batch_host(gblur_plan* plan, uint8_t* hwc_inout, int n_frames, int apply_to_alpha);
/* Device buffers (d_in may equal d_out); asynchronous on `cuda_stream` (a cudaStream_t, may be 0). */
int gblur_run_device(gblur_plan* plan, const uint8_t* d_in, uint8_t* d_out, int apply_to_alpha, void* cuda_stream);
int gblur_run_batch_device(gblur_plan* plan, const uint8_t* d_in, uint8_t* d_out, int n_frames, int apply_to_alpha,
                           void* cuda_stream);

/* Profiling hook: the same fused run restricted to stages = 1 (row pass into the plan's workspace),
 * 2 (column pass out of it), 3 (both) or 4 (the column kernel alone: an overlap-save column pass reuses the column
 * terms of its previous run on this plan, so only the block kernel is launched; otherwise the same as 2); lets bench.py
 * time each kernel with CUDA events. */
int gblur_run_stage_device(gblur_plan* plan, const uint8_t* d_in, uint8_t* d_out, int n_frames, int apply_to_alpha,
                           void* cuda_stream, int stages);

/* The two passes separately, for band sharding. Row pass: u8 [band_rows][cols*channels] ->
 * fp32 [band_rows][inter_pitch] with element (r, x*Cp + ch), Cp = processed channels.
 * Column pass: fp32 [full_rows][inter_pitch] holding band_cols*Cp valid columns -> u8
 * [full_rows][out_pitch bytes] with band_cols*channels valid bytes per row. */
int gblur_rowpass_device(gblur_plan* plan, const uint8_t* d_in, float* d_inter, long long inter_pitch_floats,
                         int apply_to_alpha, void* cuda_stream);
/* Row pass that writes the send buffer of the band-sharded all-to-all directly: pixel columns are cut in
 * bands of split_cols; element (r, x, ch) lands at
 *   d_inter[(x / split_cols) * split_stride_floats + r * split_cols * Cp + (x % split_cols) * Cp + ch],
 * i.e. one contiguous [band_rows][split_cols * Cp] block per destination rank. */
int gblur_rowpass_split_device(gblur_plan* plan, const uint8_t* d_in, float* d_inter, int split_cols,
                               long long split_stride_floats, int apply_to_alpha, void* cuda_stream);
int gblur_colpass_device(gblur_plan* plan, const float* d_inter, long long inter_pitch_floats, uint8_t* d_out,
                         long long out_pitch_bytes, int apply_to_alpha, void* cuda_stream);

/* ---- band sharding with the exchange fused into the row pass (NVLink peer stores, no collective) ----
 * Every rank owns a band of pixel columns and a PANEL buffer for it ([panel of 32 float columns][np slots]
 * [32 floats]), allocated by gblur_band_panel_alloc on its column-band plan and exported to
 * the other processes with gblur_ipc_export / gblur_ipc_open (CUDA IPC). gblur_rowpass_peer_device runs a row
 * band and stores each 128-byte line (and the mirrored reflect_101 border rows) straight into the panel buffer
 * of the rank that owns those columns — peer_panels[h] is rank h's buffer as mapped in THIS process. After a
 * cross-rank barrier, gblur_colpass_panel_device blurs the columns out of the local panel buffer.
 * cols*Cp must be a multiple of 32*n_peers. */
int gblur_plan_panel_geometry(const gblur_plan* col_plan, int* np, int* padq);
int gblur_band_panel_alloc(gblur_plan* col_plan, int apply_to_alpha, float** d_panel, long long* n_floats);
int gblur_ipc_export(const void* d_ptr, unsigned char handle[64]);
int gblur_ipc_open(const unsigned char handle[64], void** d_ptr);
int gblur_ipc_close(void* d_ptr);
int gblur_rowpass_peer_device(gblur_plan* row_plan, const uint8_t* d_in, float* const* peer_panels, int n_peers,
                              int row_offset, int apply_to_alpha, void* cuda_stream);
int gblur_colpass_panel_device(gblur_plan* col_plan, const float* d_panel, uint8_t* d_out, long long out_pitch_bytes,
                               int apply_to_alpha, void* cuda_stream);
/* Staged variant of the same exchange (copy engines instead of peer stores): the row pass stores into LOCAL
 * buffers with the panel geometry of each destination rank (pass them to gblur_rowpass_peer_device as the
 * "peers"), and this call then moves the slots [slot_begin, slot_end) of the padded column lines (slot = pad +
 * image row; reflect_101 mirror slots lie in [0, pad) and [pad + rows, pad + rows + pad)) of every panel from
 * src_panel to dst_panel — one strided device-to-device copy, asynchronous on cuda_stream, dst_panel usually the
 * IPC mapping of the owner's panel buffer (NVLink). Row chunks can travel while the next chunk is computed. */
int gblur_panel_push_device(const gblur_plan* col_plan, int apply_to_alpha, const float* src_panel, float* dst_panel,
                            int slot_begin, int slot_end, void* cuda_stream);

/* ---- halo sharding: row bands for BOTH passes, no exchange of the fp32 intermediate (third row) ----
 * The column pass runs as overlap-save blocks (1024 slots, L = 1024 - 2 pad outputs each), so a rank that owns whole
 * blocks of the padded column needs from the others only `pad` input rows on either side of its band and one integer
 * per byte column: the weighted column sums behind the reference's Nyquist term (src/gaussianblur.cpp:62-79), which are
 * summed over the ranks (e.g. ncclAllReduce of cols*channels int32). Every row and column stays an independent line as
 * in the reference (src/gaussianblur.cpp:202-244). 1 or 4 channels (all blurred), pad <= 256, cols*channels % 16 == 0.
 *   gblur_halo_partition      rows this rank writes [out_row0, out_row1) and reads [in_row0, in_row1) (host only)
 *   gblur_plan_create_halo    plan + local panel buffer for that rank
 *   gblur_halo_colsum_device  adds the rank's own rows into d_u (cols*channels ints, zeroed by the caller)
 *   gblur_halo_run_device     row pass over rows + halo, column terms from the SUMMED d_u, column blocks -> d_out */
int gblur_halo_partition(int rows, int cols, float sigma, int n_ranks, int rank, int* out_row0, int* out_row1, int* in_row0,
                         int* in_row1);
int gblur_plan_create_halo(gblur_plan** out, int device, int full_rows, int full_cols, int channels, float sigma, int n_ranks,
                           int rank);
int gblur_halo_info(const gblur_plan* plan, int* out_row0, int* out_row1, int* in_row0, int* in_row1);
int gblur_halo_colsum_device(gblur_plan* plan, const uint8_t* d_in_rows, int* d_u, void* cuda_stream);
int gblur_halo_run_device(gblur_plan* plan, const uint8_t* d_in_rows, const int* d_u_total, uint8_t* d_out, void* cuda_stream);
/* the two halves of gblur_halo_run_device: the row pass of a sub-range of the rank's rows (so it can start on the rows
 * it holds while the halo is on the link), and column terms + column blocks */
int gblur_halo_rowpass_device(gblur_plan* plan, const uint8_t* d_in_rows, int row_begin, int row_end, void* cuda_stream);
/* the same over two row ranges (end_a <= begin_b; either may be empty) in ONE launch: the halo rows above and below the
 * rank's band, once they have arrived (two short launches cost 2.5 x what the same rows cost inside one) */
int gblur_halo_rowpass2_device(gblur_plan* plan, const uint8_t* d_in_rows, int begin_a, int end_a, int begin_b, int end_b,
                               void* cuda_stream);
int gblur_halo_colpass_device(gblur_plan* plan, const int* d_u_total, uint8_t* d_out, void* cuda_stream);

/* ---- several GPUs of one box behind ONE handle, single process (csrc/multi_gpu.cpp) ----
 * The reference fans its passes out over std::threads inside the library (include/gaussianblur/helpers.hpp:134-172,
 * called from src/gaussianblur.cpp:212); these entry points fan out over GPUs the same way, so a C or C++ caller needs
 * no torch, no NCCL and no second process to use more than one device.
 *   gblur_mg_create      n_gpus devices (devices == NULL: 0 .. n_gpus-1), peer access enabled between every pair where
 *                        the hardware allows; geometry and sigma as in gblur_plan_create. Plans are built lazily.
 *   gblur_mg_run_frames  n_frames contiguous HWC frames in host memory, blurred in place; frames are cut into
 *                        contiguous blocks, one per GPU, each GPU streams its block (H2D || blur || D2H): no exchange.
 *   gblur_mg_run_bands   ONE image in host memory, blurred in place, cut into row bands.
 *                        Halo mode (1 channel, or 4 with alpha blurred; pad <= 256; cols * channels % 16 == 0): row
 *                        bands for BOTH passes — every GPU takes its rows plus `pad` rows on either side from the host
 *                        image, the host adds the GPUs' integer column sums (cols * channels values each) and hands the
 *                        total back, every GPU returns its output rows. Nothing crosses between the GPUs and no peer
 *                        access is needed (gblur_halo_*; GBLUR_MG_BANDS=exchange turns it off).
 *                        Otherwise the exchange: every GPU runs the row pass of its band and stores the fp32
 *                        intermediate straight into the panel buffer of the GPU owning those pixel columns (NVLink peer
 *                        stores), then every GPU runs the column pass of its column band. Needs peer access,
 *                        cols % n_gpus == 0 and (cols / n_gpus) * Cp % 32 == 0; otherwise the whole image runs on the
 *                        first device. gblur_mg_info tells which, per alpha mode, after the first run: 2 = halo,
 *                        1 = exchange, 0 = one device.
 * One call at a time per handle (calls on one handle serialise); errors: gblur_mg_last_error(). */
typedef struct gblur_mg gblur_mg;
int gblur_mg_create(gblur_mg** out, int n_gpus, const int* devices, int rows, int cols, int channels, float sigma);
void gblur_mg_destroy(gblur_mg* mg);
int gblur_mg_info(const gblur_mg* mg, int* n_gpus, int* peer_access, int* bands_sharded_alpha0, int* bands_sharded_alpha1);
int gblur_mg_run_frames(gblur_mg* mg, uint8_t* hwc_frames_inout, int n_frames, int apply_to_alpha);
int gblur_mg_run_bands(gblur_mg* mg, uint8_t* hwc_inout, int apply_to_alpha);
const char* gblur_mg_last_error(void);

/* ---- an entry point that allocates nothing ----
 * NO counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * gblur_colsum_noalloc computes what gblur_halo_colsum_device computes for a whole image — per byte column w of the
 * rows x (cols * channels) u8 image, d_colsum[w] = sum over rows r of weight(r) * in[r][w], weight = (-1)^(pad + r) times
 * the number of reflect_101-padded positions that read row r: the integer sums behind the Nyquist term of the column
 * pass — but it accumulates in a file-scope static __device__ array (1 MiB + 2 KiB per GPU that loads the library, held
 * until exit) instead of a buffer the plan allocates. The call makes `device` current for its duration, issues two
 * cudaMemsetAsync on that array and one kernel on `cuda_stream`, and calls no cudaMalloc / cudaFree / allocating routine.
 *   limits   cols * channels <= GBLUR_NOALLOC_MAX_ROW_BYTES (GBLUR_E_UNSUPPORTED beyond it, with a message), a multiple
 *            of 4; d_in 4-byte aligned, d_colsum (cols * channels int32, fully overwritten) 16-byte aligned.
 *   sharing  every call on a device uses the SAME array, which keeps the last call's contents (each call clears what it
 *            uses first, in stream order): calls that overlap on the GPU — different streams, different host threads —
 *            race. Serialise them on one stream or with events. */
#define GBLUR_NOALLOC_MAX_ROW_BYTES 262144
int gblur_colsum_noalloc(int device, const uint8_t* d_in, int rows, int cols, int channels, float sigma, int32_t* d_colsum,
                         void* cuda_stream);

/* Test hook: fp32 row-pass output of the last gblur_run_* call on this plan (frame 0), copied to
 * host as [rows][cols*Cp]. */
int gblur_dump_intermediate(gblur_plan* plan, float* host_out, int apply_to_alpha);

/* Kernels launched through this library since load (bench.py reports the delta). */
unsigned long long gblur_launch_count(void);
/* Thread-local message for the last non-zero status. */
const char* gblur_last_error(void);
/* Number of CUDA devices visible, or -1 with an error message. */
int gblur_device_count(void);

/* ---- host-only helpers: no CUDA device needed (reference semantics, used by CPU-side tests) ----
 * gblur_host_plan_sizes  gaussian_window + sizes/trailing zeros   src/gaussianblur.cpp:11-26,104-123
 * gblur_host_taps        get_gaussian's normalised float taps     src/gaussianblur.cpp:28-48
 * gblur_host_spectrum    kerf_1D_* for (sigma, ksize, n), ordered layout, computed in double
 * gblur_host_line_plan   stage radices / CTA shape chosen for FFT length n
 * gblur_host_kperm       1/n-scaled gains in the digit-reversed order the kernels use, with the
 *                        reference's Nyquist-bin behaviour (bin n/2 gets K[0])  src/gaussianblur.cpp:62-79 */
int gblur_host_plan_sizes(int rows, int cols, float sigma, int* ksize, int* pad, int* n_rowpass, int* n_colpass,
                          int* tz_rows, int* tz_cols);
int gblur_host_taps(float sigma, int ksize, float* out);
int gblur_host_spectrum(float sigma, int ksize, int n, float* out_ordered);
int gblur_host_line_plan(int n, int* radices, int* nstages, int* nseq, int* static_kernel, long long* smem_bytes,
                         int* nthreads);
int gblur_host_kperm(float sigma, int ksize, int n, float* out_n);
/* Which kernel family a single-frame fused run takes when nothing is forced: *blocks = 2 both passes as overlap-save
 * 1024-sample blocks, 1 the row pass only, 0 whole-line transforms of the reference's length (DESIGN.md section 2,
 * "Selection"; buffers that are not 16-byte aligned keep the whole-line kernels regardless). */
int gblur_host_kernel_choice(int rows, int cols, int channels, float sigma, int apply_to_alpha, int* blocks);

#ifdef __cplusplus
}
#endif
#endif /* GBLUR_CAPI_H_ */
