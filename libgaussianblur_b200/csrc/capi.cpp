// capi.cpp — implementation of include/gblur_capi.h: plans, workspaces and the pass drivers.
// Host semantics follow the reference (see hostsem.hpp / planner.hpp for file:line citations);
// all pixel work happens in the sm_100a kernels (kernels.h). No CPU fallback exists.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <algorithm>
#include <cstdlib>
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gblur_capi.h"
#include "hostsem.hpp"
#include "kernels.h"
#include "ols.h"
#include "planner.hpp"

using namespace gblur;

namespace {

// default of the overlap-save column pass when GBLUR_COL_OLS is unset: 0 = off, 2 = automatic (ols::col_profitable)
constexpr int kColOlsAuto = 2;

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define CU(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) return fail(GBLUR_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) == cudaSuccess) ok = (cudaSetDevice(dev) == cudaSuccess);
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

struct LineTables {
  LinePlan lp;
  float* d_tw = nullptr;
  float* d_kperm = nullptr;
  std::vector<float> kerf;  // host copy, pffft ordered layout
};

}  // namespace

struct gblur_plan {
  std::mutex host_mu;  // the host-buffer entry points use the plan's own stream, staging buffers and workspace: one at a time
  int device = 0;
  int num_sms = 148;
  int channels = 0;
  float sigma = 0.f;
  int full_rows = 0, full_cols = 0;  // geometry the window / FFT lengths come from
  int rows = 0, cols = 0;            // rows the row pass covers, pixel columns the column pass covers
  int max_frames = 1;
  bool band = false;
  host::Sizes sz{};
  LineTables row, col;  // row pass (N = n_rowpass), column pass (N = n_colpass)
  float* d_inter = nullptr;
  size_t inter_floats = 0;
  std::vector<void*> owned;    // device buffers handed out by gblur_band_panel_alloc
  CUtensorMap tmap;            // TMA view of the panel intermediate for the column pass
  const float* tmap_base = nullptr;
  int tmap_np = 0, tmap_L = 0, tmap_rows = 0;
  long long tmap_panels = 0;
  uint8_t* d_stage = nullptr;  // device image for gblur_run_host
  size_t stage_bytes = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t s_in = nullptr, s_out = nullptr;  // copy streams of the pipelined host path
  uint8_t* h_ring[3] = {nullptr, nullptr, nullptr};  // pinned staging slots for pageable host buffers
  size_t h_ring_bytes = 0;
  cudaStream_t s_side = nullptr;                 // the column Nyquist terms run here, beside the row pass
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  std::vector<cudaEvent_t> events;               // reused by the pipelined host path
  uint8_t* d_stage2 = nullptr;                   // second and third staging buffer (batches)
  size_t stage2_bytes = 0;
  uint8_t* d_stage3 = nullptr;
  size_t stage3_bytes = 0;
  int last_cp = 0;       // processed channels of the last run (layout of d_inter)
  int inter_cp = 0;      // processed-channel count the blocked workspace is currently laid out for
  int last_frames = 0;
  // overlap-save row pass (ols.h). GBLUR_ROW_OLS at plan creation: unset = automatic (single-frame fused runs of
  // wide RGBA images whose blocks waste little, see ols::row_profitable), 1 = wherever eligible, 0 = never
  ols::RowPlan* ols_row = nullptr;
  int ols_mode = 2;
  // overlap-save column pass, same plan object. GBLUR_COL_OLS: 1 = wherever eligible, 0 = never, unset = kColOlsAuto
  int ols_col_mode = 0;
  bool last_alpha_restore = false;  // the last fused run took that route (gblur_dump_intermediate)
  bool ols_alpha_restore = true;  // RGBA, alpha not blurred: four channels through the overlap-save passes + alpha restore
  int last_row_kind = 0, last_col_kind = 0;  // kernels of the last pass launched: 0 = whole line, 1 = overlap-save
  // halo-sharded plan (gblur_plan_create_halo): this rank owns the column-pass blocks [blk0, blk1) of the padded column
  bool halo = false;
  int h_L = 0, h_blk0 = 0, h_blk1 = 0, h_ranks = 1;
  int h_out0 = 0, h_out1 = 0;  // image rows this rank writes
  int h_in0 = 0, h_in1 = 0;    // image rows this rank reads (its own plus the halo of pad rows on either side)
  int h_np = 0;                // slots of the local panel buffer
  float* d_halo_panel = nullptr;
};

namespace {

int build_tables(LineTables& t, int N, const std::vector<float>& taps, int cluster_nseq) {
  if (!host::plan_line(N, t.lp, cluster_nseq))
    return fail(GBLUR_E_UNSUPPORTED, "FFT length %d does not fit one CTA's shared memory", N);
  const std::vector<double> gains = host::kernel_gains(taps, N);
  t.kerf = host::kerf_ordered(gains, N);
  std::vector<float> tw;
  size_t off[kHostMaxStages] = {0, 0, 0, 0, 0};
  host::build_twiddles(t.lp, tw, off);
  const std::vector<float> kp = host::build_kperm(t.lp, gains);
  if (!tw.empty()) {
    CU(cudaMalloc(&t.d_tw, tw.size() * sizeof(float)));
    CU(cudaMemcpy(t.d_tw, tw.data(), tw.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  CU(cudaMalloc(&t.d_kperm, kp.size() * sizeof(float)));
  CU(cudaMemcpy(t.d_kperm, kp.data(), kp.size() * sizeof(float), cudaMemcpyHostToDevice));
  for (int s = 0; s < t.lp.nstages - 1; ++s) t.lp.d_tw[s] = reinterpret_cast<const float2*>(t.d_tw) + off[s];
  t.lp.d_kperm = t.d_kperm;
  return GBLUR_OK;
}

void free_tables(LineTables& t) {
  if (t.d_tw) cudaFree(t.d_tw);
  if (t.d_kperm) cudaFree(t.d_kperm);
  t.d_tw = t.d_kperm = nullptr;
}

int processed_channels(const gblur_plan* p, int apply_to_alpha) {
  return (p->channels == 4 && !apply_to_alpha) ? 3 : p->channels;
}

long long default_pitch(long long w) { return (w + 3) / 4 * 4; }

// Panel intermediate of the fused run: [panel of 32 float columns][np slots][32 floats].
// ols_col: the overlap-save column pass will read it (TMA boxes of consecutive slots): no bank-padding holes.
BlockedLayout blocked_layout(const gblur_plan* p, int cp, bool ols_col = false) {
  BlockedLayout b;
  const LinePlan& c = p->col.lp;
  b.enabled = 1;
  b.L = 2 * c.nseq;
  // cluster column plans read the panel through their loader: natural slot order, no padding holes
  b.np = (c.cluster > 1 || ols_col) ? c.N : c.exch_elems;
  b.padq = (c.cluster > 1 || ols_col) ? 0 : c.padq;
  // RGBA: written by the overlap-save row pass, read by the overlap-save column pass (one channel keeps natural order)
  b.pair_major = (ols_col && p->channels == 4) ? 1 : 0;
  b.ols_pair = ols_col ? 1 : 0;
  b.rows_per_frame = p->full_rows;
  const long long panels = ((long long)p->cols * cp + 31) / 32;
  b.frame_stride = panels * b.np * 32;
  return b;
}

int ensure_inter(gblur_plan* p, size_t floats) {
  if (p->inter_floats >= floats) return GBLUR_OK;
  if (p->d_inter) cudaFree(p->d_inter);
  p->d_inter = nullptr;
  p->inter_floats = 0;
  CU(cudaMalloc(&p->d_inter, floats * sizeof(float)));
  // pad columns of the intermediate are read (never stored) by the column pass: keep them finite
  CU(cudaMemset(p->d_inter, 0, floats * sizeof(float)));
  CU(cudaDeviceSynchronize());
  p->inter_floats = floats;
  return GBLUR_OK;
}

int create_common(gblur_plan** out, int device, int full_rows, int full_cols, int channels, float sigma, int rows,
                  int cols, int max_frames, bool band) {
  if (!out) return fail(GBLUR_E_INVALID, "null plan out pointer");
  *out = nullptr;
  if (!(sigma > 0.0f)) return fail(GBLUR_E_INVALID, "Invalid smoothing factor");
  if (channels != 1 && channels != 3 && channels != 4) return fail(GBLUR_E_INVALID, "Unsupported number of channels");
  if (full_rows < 1 || full_cols < 1 || rows < 1 || cols < 1 || max_frames < 1)
    return fail(GBLUR_E_INVALID, "non-positive geometry");
  const host::Sizes sz = host::plan_sizes(full_rows, full_cols, sigma);
  if (sz.pad > std::min(full_rows, full_cols) - 1)
    return fail(GBLUR_E_PRECOND, "pad %d exceeds min(rows, cols) - 1 = %d (reference precondition)", sz.pad,
                std::min(full_rows, full_cols) - 1);
  int ndev = 0;
  CU(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return fail(GBLUR_E_CUDA, "device %d not available (%d visible)", device, ndev);
  DeviceGuard guard(device);
  if (!guard.ok) return fail(GBLUR_E_CUDA, "cudaSetDevice(%d) failed", device);

  gblur_plan* p = new gblur_plan();
  p->device = device;
  p->channels = channels;
  p->sigma = sigma;
  p->full_rows = full_rows;
  p->full_cols = full_cols;
  p->rows = rows;
  p->cols = cols;
  p->max_frames = max_frames;
  p->band = band;
  p->sz = sz;
  cudaDeviceGetAttribute(&p->num_sms, cudaDevAttrMultiProcessorCount, device);
  const std::vector<float> taps = host::gaussian_taps(sigma, sz.ksize);
  int rc = GBLUR_OK;
  if (cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking) != cudaSuccess)
    rc = fail(GBLUR_E_CUDA, "cudaStreamCreate failed");
  if (rc == GBLUR_OK) rc = build_tables(p->row, sz.n_rowpass, taps, 1);
  if (rc == GBLUR_OK) rc = build_tables(p->col, sz.n_colpass, taps, 2);
  if (rc == GBLUR_OK) {
    const char* e = getenv("GBLUR_ROW_OLS");
    p->ols_mode = e ? (atoi(e) > 0 ? 1 : 0) : 2;
    if (p->ols_mode != 0 && !band) {
      const char* ec = getenv("GBLUR_COL_OLS");
      p->ols_col_mode = ec ? (atoi(ec) > 0 ? 1 : 0) : kColOlsAuto;
      if (const char* ea = getenv("GBLUR_OLS_ALPHA_RESTORE")) p->ols_alpha_restore = atoi(ea) > 0;
      // three channels: the row pass only, on byte rows (block gains of the taps spread three bytes apart)
      const cudaError_t ce = channels == 3 ? ols::row_plan_create(&p->ols_row, taps, sz.n_rowpass, 0, 3)
                                           : ols::row_plan_create(&p->ols_row, taps, sz.n_rowpass, sz.n_colpass);
      if (ce != cudaSuccess) rc = fail(GBLUR_E_CUDA, "overlap-save row plan: %s", cudaGetErrorString(ce));
    }
  }
  if (rc != GBLUR_OK) {
    std::string keep = g_err;
    gblur_plan_destroy(p);
    g_err = keep;
    return rc;
  }
  *out = p;
  return GBLUR_OK;
}

// Will the overlap-save column pass (ols.h) take a fused run with `cp` processed channels and `frames` frames? Decided
// before the row pass, because the panel layout it reads has no bank-padding holes.
bool ols_col_wanted(const gblur_plan* p, int cp, int frames) {
  if (!p->ols_row || p->ols_col_mode == 0 || p->band || (p->channels != 4 && p->channels != 1) || cp != p->channels ||
      frames != 1)
    return false;
  if (((long long)p->cols * cp) % 16 || ((long long)p->full_cols * p->channels) % 16) return false;
  if (p->ols_col_mode == 1) return true;
  // automatic: the two passes together must gain (the pair-major panel rows tie them together)
  return p->ols_mode != 0 && ols::both_profitable(p->full_cols, p->full_rows, p->sz.pad, p->sz.n_rowpass, p->sz.n_colpass);
}

// rows_sub > 0: only rows [row0, row0 + rows_sub) of a single frame (d_in / d_inter still point at row 0)
int run_rowpass(gblur_plan* p, const uint8_t* d_in, float* d_inter, long long pitch, int frames, int alpha,
                cudaStream_t s, const BlockedLayout* blk = nullptr, int split_cols = 0, long long split_stride = 0,
                int row0 = 0, int rows_sub = 0, bool fused_single = false) {
  const int cp = processed_channels(p, alpha);
  RowLaunch a{};
  a.in = d_in;
  a.in_pitch = (long long)p->full_cols * p->channels;
  a.C = p->channels;
  a.Cp = cp;
  a.T = p->full_cols;
  a.pad = p->sz.pad;
  a.rows_total = (long long)p->rows * frames;
  a.inter = d_inter;
  a.inter_pitch = pitch;
  if (blk) a.blk = *blk;
  a.split_cols = split_cols;
  a.split_stride = split_stride;
  if (rows_sub > 0) {
    a.in = d_in + (long long)row0 * a.in_pitch;
    a.rows_total = rows_sub;
    if (blk)
      a.row_offset = row0;  // panel layout: the slot index carries the row
    else
      a.inter = d_inter + (long long)row0 * pitch;
  }
  if (blk && blk->pair_major && !(p->ols_row && ols::row_eligible(p->ols_row, a)))
    return fail(GBLUR_E_UNSUPPORTED, "pair-major panel layout, but the overlap-save row pass is not eligible");
  if (p->ols_row && ols::row_eligible(p->ols_row, a) &&
      ((blk && (blk->pair_major || blk->ols_pair)) || p->ols_mode == 1 ||
       (fused_single && ols::row_profitable(a, p->sz.n_rowpass)))) {
    // (Leaving a few SMs out of this persistent grid for the column-term kernels on the side stream was measured and
    // dropped: 0.5785 ms per 8K step with 0 SMs reserved, 0.5790 with 2, 0.5886 with 4 — profiles/r02/row_reserve.jsonl.)
    CU(ols::launch_row(p->ols_row, a, p->num_sms, s));
    p->last_row_kind = 1;
    return GBLUR_OK;
  }
  p->last_row_kind = 0;
  CU(launch_rowpass(p->row.lp, a, p->num_sms, s));
  return GBLUR_OK;
}

// (Re)build the cached tensor map when the buffer, the layout or the extent changed.
bool panel_tensor_map(gblur_plan* p, const float* base, const BlockedLayout& b, long long panels) {
  if (p->tmap_rows && p->tmap_base == base && p->tmap_np == b.np && p->tmap_L == b.L && p->tmap_panels == panels)
    return true;
  int rows = 0;
  if (!make_panel_tensor_map(&p->tmap, base, b.np, panels, b.L, &rows)) {
    p->tmap_rows = 0;
    p->tmap_base = nullptr;
    return false;
  }
  p->tmap_base = base;
  p->tmap_np = b.np;
  p->tmap_L = b.L;
  p->tmap_panels = panels;
  p->tmap_rows = rows;
  return true;
}

// w_end > 0: only the float columns [w_begin, w_end) (w_begin a multiple of 32) of a single frame
// ols_in: the u8 image the intermediate was computed from (single-frame fused runs only): the overlap-save column pass
// forms the column term of the reference's Nyquist behaviour from it (ols_core.cuh); ols_terms = false reuses the terms
// of the previous launch on this plan (column bands of one image)
int run_colpass(gblur_plan* p, const float* d_inter, long long pitch, uint8_t* d_out, long long out_pitch, int frames,
                int alpha, cudaStream_t s, const BlockedLayout* blk = nullptr, int w_begin = 0, int w_end = 0,
                const uint8_t* ols_in = nullptr, bool ols_terms = true, const uint8_t* alpha_src = nullptr) {
  const int cp = processed_channels(p, alpha);
  ColLaunch a{};
  a.inter = d_inter;
  a.pitch = pitch;
  a.frame_stride = (long long)p->full_rows * pitch;
  a.frames = frames;
  a.T = p->full_rows;
  a.pad = p->sz.pad;
  a.W = p->cols * cp;
  a.out = d_out;
  a.out_pitch = out_pitch;
  a.out_frame_stride = (long long)p->full_rows * out_pitch;
  a.C = p->channels;
  a.Cp = cp;
  a.w_begin = w_begin;
  a.w_end = w_end;
  if (blk) {
    a.blk = *blk;
    a.panel_in = p->col.lp.cluster > 1 ? 1 : 0;
    const long long panels = ((long long)p->cols * cp + 31) / 32 * frames;
    if (p->col.lp.cluster == 1 && blk->padq == p->col.lp.padq && panel_tensor_map(p, d_inter, *blk, panels)) {
      a.tmap = &p->tmap;
      a.tma_rows = p->tmap_rows;
    }
  }
  a.alpha_src = alpha_src;
  a.alpha_pitch = (long long)p->full_cols * p->channels;
  if (alpha_src && !(ols_in && blk && blk->padq == 0 && ols_col_wanted(p, cp, frames) && ols::col_eligible(p->ols_row, a)))
    return fail(GBLUR_E_UNSUPPORTED, "alpha pass-through on four channels needs the overlap-save column pass");
  if (ols_in && blk && blk->padq == 0 && ols_col_wanted(p, cp, frames) && ols::col_eligible(p->ols_row, a)) {
    if (ols_terms) CU(ols::launch_col_terms(p->ols_row, a, ols_in, (long long)p->full_cols * p->channels, s));
    CU(ols::launch_col(p->ols_row, a, p->num_sms, s));
    p->last_col_kind = 1;
    return GBLUR_OK;
  }
  if (blk && blk->padq != p->col.lp.padq && p->col.lp.cluster == 1)
    return fail(GBLUR_E_UNSUPPORTED, "panel layout without padding holes, but the overlap-save column pass is not eligible");
  p->last_col_kind = 0;
  CU(launch_colpass(p->col.lp, a, p->num_sms, s));
  return GBLUR_OK;
}

}  // namespace

namespace gblur {
// for entry points that live in other translation units (noalloc.cu): the text gblur_last_error() returns on this thread
void set_last_error(const char* message) { g_err = message ? message : ""; }
}  // namespace gblur

extern "C" {

const char* gblur_last_error(void) { return g_err.c_str(); }

unsigned long long gblur_launch_count(void) { return launch_counter(); }

int gblur_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    fail(GBLUR_E_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
    return -1;
  }
  return n;
}

int gblur_plan_create(gblur_plan** out, int device, int rows, int cols, int channels, float sigma) {
  return create_common(out, device, rows, cols, channels, sigma, rows, cols, 1, false);
}

int gblur_plan_create_batch(gblur_plan** out, int device, int rows, int cols, int channels, float sigma,
                            int max_frames) {
  return create_common(out, device, rows, cols, channels, sigma, rows, cols, max_frames, false);
}

int gblur_plan_create_band(gblur_plan** out, int device, int full_rows, int full_cols, int channels, float sigma,
                           int band_rows, int band_cols) {
  if (band_rows > full_rows || band_cols > full_cols) return fail(GBLUR_E_INVALID, "band larger than the image");
  return create_common(out, device, full_rows, full_cols, channels, sigma, band_rows, band_cols, 1, true);
}

void gblur_plan_destroy(gblur_plan* p) {
  if (!p) return;
  DeviceGuard guard(p->device);
  if (p->stream) cudaStreamSynchronize(p->stream);
  free_tables(p->row);
  free_tables(p->col);
  ols::row_plan_destroy(p->ols_row);
  if (p->d_inter) cudaFree(p->d_inter);
  if (p->d_halo_panel) cudaFree(p->d_halo_panel);
  if (p->d_stage) cudaFree(p->d_stage);
  if (p->d_stage2) cudaFree(p->d_stage2);
  if (p->d_stage3) cudaFree(p->d_stage3);
  for (cudaEvent_t e : p->events) cudaEventDestroy(e);
  for (uint8_t* h : p->h_ring)
    if (h) cudaFreeHost(h);
  if (p->ev_fork) cudaEventDestroy(p->ev_fork);
  if (p->ev_join) cudaEventDestroy(p->ev_join);
  if (p->s_side) cudaStreamDestroy(p->s_side);
  if (p->s_in) cudaStreamDestroy(p->s_in);
  if (p->s_out) cudaStreamDestroy(p->s_out);
  for (void* q : p->owned) cudaFree(q);
  if (p->stream) cudaStreamDestroy(p->stream);
  delete p;
}

int gblur_plan_info(const gblur_plan* p, int* ksize, int* pad, int* n_rowpass, int* n_colpass, int* tz_rows,
                    int* tz_cols) {
  if (!p) return fail(GBLUR_E_INVALID, "null plan");
  if (ksize) *ksize = p->sz.ksize;
  if (pad) *pad = p->sz.pad;
  if (n_rowpass) *n_rowpass = p->sz.n_rowpass;
  if (n_colpass) *n_colpass = p->sz.n_colpass;
  if (tz_rows) *tz_rows = p->sz.tz_rows;
  if (tz_cols) *tz_cols = p->sz.tz_cols;
  return GBLUR_OK;
}

int gblur_plan_spectrum(const gblur_plan* p, int axis, float* host_out) {
  if (!p || !host_out || (axis != 0 && axis != 1)) return fail(GBLUR_E_INVALID, "bad spectrum query");
  const std::vector<float>& k = axis == 0 ? p->row.kerf : p->col.kerf;
  std::memcpy(host_out, k.data(), k.size() * sizeof(float));
  return GBLUR_OK;
}

int gblur_plan_radices(const gblur_plan* p, int axis, int* radices, int* n, int* static_kernel) {
  if (!p || (axis != 0 && axis != 1)) return fail(GBLUR_E_INVALID, "bad radices query");
  const LinePlan& lp = axis == 0 ? p->row.lp : p->col.lp;
  if (radices)
    for (int s = 0; s < lp.nstages; ++s) radices[s] = lp.radix[s];
  if (n) *n = lp.nstages;
  if (static_kernel) *static_kernel = lp.static_id >= 0 ? 1 : 0;
  return GBLUR_OK;
}

int gblur_plan_last_kernels(const gblur_plan* p, int* row_kind, int* col_kind) {
  if (!p) return fail(GBLUR_E_INVALID, "null plan");
  if (row_kind) *row_kind = p->last_row_kind;
  if (col_kind) *col_kind = p->last_col_kind;
  return GBLUR_OK;
}

namespace {
struct FusedCtx {
  BlockedLayout blk;
  const BlockedLayout* bp = nullptr;  // null: row-major intermediate
  long long pitch = 0;
  int cp = 0;
  // RGBA with alpha passed through, on the overlap-save passes: all four channels are transformed (cp = 4, the passes
  // are called with alpha = 1) and the column kernel restores the input's alpha byte in its output tile
  bool alpha_restore = false;
  int alpha = 0;  // what the passes are called with
  bool ols_col = false;  // the column pass runs as overlap-save blocks (pair-major / hole-free panel layout)
};

// Workspace, layout choice and (when the layout changes) the zero fill, ordered on stream s.
// io_aligned: the caller's image pointers are 16-byte aligned (the overlap-save column pass moves its tiles with TMA);
// unaligned buffers keep the whole-line kernels instead of failing later
int fused_prepare(gblur_plan* p, int n_frames, int apply_to_alpha, cudaStream_t s, FusedCtx& c, bool io_aligned = true) {
  if (p->band) return fail(GBLUR_E_INVALID, "band plans run through gblur_rowpass_device / gblur_colpass_device");
  if (n_frames < 1 || n_frames > p->max_frames)
    return fail(GBLUR_E_INVALID, "n_frames %d outside [1, %d]", n_frames, p->max_frames);
  c.cp = processed_channels(p, apply_to_alpha);
  c.alpha = apply_to_alpha ? 1 : 0;
  if (p->channels == 4 && !apply_to_alpha && n_frames == 1 && p->ols_alpha_restore && io_aligned &&
      ols_col_wanted(p, 4, n_frames)) {
    // a fourth channel costs the block kernels a third more work and they are still well ahead of the whole-line
    // kernels on three channels (GBLUR_OLS_ALPHA_RESTORE=0 at plan creation keeps the whole-line path)
    c.alpha_restore = true;
    c.alpha = 1;
    c.cp = 4;
  }
  // workspace sized for the widest layout (all channels processed)
  const BlockedLayout widest = blocked_layout(p, p->channels);
  const size_t rowmajor_floats = (size_t)p->rows * (size_t)default_pitch((long long)p->cols * p->channels);
  int rc = ensure_inter(p, (size_t)p->max_frames * std::max((size_t)widest.frame_stride, rowmajor_floats));
  if (rc != GBLUR_OK) return rc;
  // tuning switch: GBLUR_LAYOUT=rowmajor keeps the fp32 intermediate row-major (column pass gathers
  // 16-byte pieces); default is the panel layout
  static const bool rowmajor = [] {
    const char* e = getenv("GBLUR_LAYOUT");
    return e && std::strcmp(e, "rowmajor") == 0;
  }();
  const bool ols_col = io_aligned && ols_col_wanted(p, c.cp, n_frames);
  c.ols_col = ols_col;
  c.blk = blocked_layout(p, c.cp, ols_col);
  // single-GPU run with a cluster column pass: the row-major intermediate is a little faster (cheaper stores)
  static const bool panel_forced = [] {
    const char* e = getenv("GBLUR_LAYOUT");
    return e && std::strcmp(e, "panel") == 0;
  }();
  const bool use_rowmajor = !ols_col && (rowmajor || (p->col.lp.cluster > 1 && !panel_forced));
  c.bp = use_rowmajor ? nullptr : &c.blk;
  c.pitch = default_pitch((long long)p->cols * c.cp);
  const int layout_tag = use_rowmajor ? -c.cp : (ols_col ? c.cp + 16 : c.cp);
  if (p->inter_cp != 0 && p->inter_cp != layout_tag) {
    // the zero tail / padding slots of the new layout may hold samples of the old one
    CU(cudaMemsetAsync(p->d_inter, 0, p->inter_floats * sizeof(float), s));
  }
  p->inter_cp = layout_tag;
  p->last_cp = c.cp;
  p->last_alpha_restore = c.alpha_restore;
  p->last_frames = n_frames;
  return GBLUR_OK;
}
}  // namespace

static int run_fused(gblur_plan* p, const uint8_t* d_in, uint8_t* d_out, int n_frames, int apply_to_alpha,
                     void* cuda_stream, int stages) {
  if (!p || !d_in || !d_out) return fail(GBLUR_E_INVALID, "null argument");
  DeviceGuard guard(p->device);
  cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
  FusedCtx c;
  const bool io_aligned = ((reinterpret_cast<uintptr_t>(d_in) | reinterpret_cast<uintptr_t>(d_out)) & 15) == 0;
  int rc = fused_prepare(p, n_frames, apply_to_alpha, s, c, io_aligned);
  if (rc != GBLUR_OK) return rc;
  // The column terms of the overlap-save column pass depend on the input image only: in a two-stage run they are
  // computed on a side stream while the row pass runs (one pass over the u8 input, hidden under a pass that is not
  // memory-bound), and joined before the column blocks.
  bool terms_forked = false;
  if (stages == 3 && n_frames == 1 && c.bp && c.blk.padq == 0 && c.ols_col) {
    if (!p->s_side) CU(cudaStreamCreateWithFlags(&p->s_side, cudaStreamNonBlocking));
    if (!p->ev_fork) CU(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
    if (!p->ev_join) CU(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
    ColLaunch a{};
    a.T = p->full_rows;
    a.pad = p->sz.pad;
    a.W = p->cols * c.cp;
    a.C = p->channels;
    a.Cp = c.cp;
    CU(cudaEventRecord(p->ev_fork, s));
    CU(cudaStreamWaitEvent(p->s_side, p->ev_fork, 0));
    CU(ols::launch_col_terms(p->ols_row, a, d_in, (long long)p->full_cols * p->channels, p->s_side));
    CU(cudaEventRecord(p->ev_join, p->s_side));
    terms_forked = true;
  }
  if (stages & 1) {
    rc = run_rowpass(p, d_in, p->d_inter, c.pitch, n_frames, c.alpha, s, c.bp, 0, 0, 0, 0, n_frames == 1);
    if (rc != GBLUR_OK) return rc;
  }
  // stages == 4: the column kernel alone — an overlap-save column pass reuses the column terms its last run on this
  // plan left behind (same image), so the launch that is timed is the block kernel and nothing else
  const bool col_kernel_only = stages == 4;
  if (!(stages & 2) && !col_kernel_only) return GBLUR_OK;
  if (terms_forked) CU(cudaStreamWaitEvent(s, p->ev_join, 0));
  if (d_in != d_out && c.cp < p->channels)
    CU(launch_copy_channel(d_in, d_out, (long long)p->rows * p->cols * n_frames, p->channels, 3, s));
  // alpha restore: the column kernel takes the alpha bytes from the input image (d_out itself when the call is in place:
  // an item reads exactly the bytes it then overwrites); a stage-2-only call finds them in d_in as well
  return run_colpass(p, p->d_inter, c.pitch, d_out, (long long)p->cols * p->channels, n_frames, c.alpha, s, c.bp,
                     0, 0, n_frames == 1 ? d_in : nullptr, !terms_forked && !col_kernel_only, c.alpha_restore ? d_in : nullptr);
}

int gblur_run_batch_device(gblur_plan* p, const uint8_t* d_in, uint8_t* d_out, int n_frames, int apply_to_alpha,
                           void* cuda_stream) {
  return run_fused(p, d_in, d_out, n_frames, apply_to_alpha, cuda_stream, 3);
}

int gblur_run_stage_device(gblur_plan* p, const uint8_t* d_in, uint8_t* d_out, int n_frames, int apply_to_alpha,
                           void* cuda_stream, int stages) {
  if (stages < 1 || stages > 4)
    return fail(GBLUR_E_INVALID, "stages must be 1 (row pass), 2 (column pass), 3 (both) or 4 (column kernel alone)");
  return run_fused(p, d_in, d_out, n_frames, apply_to_alpha, cuda_stream, stages);
}

int gblur_run_device(gblur_plan* p, const uint8_t* d_in, uint8_t* d_out, int apply_to_alpha, void* cuda_stream) {
  return gblur_run_batch_device(p, d_in, d_out, 1, apply_to_alpha, cuda_stream);
}

namespace {
int ensure_stage(uint8_t** buf, size_t* have, size_t bytes) {
  if (*have >= bytes) return GBLUR_OK;
  if (*buf) cudaFree(*buf);
  *buf = nullptr;
  *have = 0;
  CU(cudaMalloc(buf, bytes));
  *have = bytes;
  return GBLUR_OK;
}

int ensure_pipe(gblur_plan* p, size_t nevents) {
  if (!p->s_in) CU(cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking));
  if (!p->s_out) CU(cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking));
  while (p->events.size() < nevents) {
    cudaEvent_t e;
    CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    p->events.push_back(e);
  }
  return GBLUR_OK;
}

// A few persistent host threads that copy between a caller's PAGEABLE buffer and pinned staging slots. cudaMemcpyAsync
// on pageable memory goes through the driver's own small staging buffer, synchronously and at 6-12 GB/s (measured
// through the drop-in call: 23 ms per 8K RGBA image against 5.4 ms from pinned memory); several threads copying into
// our own pinned ring keep both PCIe directions busy instead.
class HostPool {
 public:
  static HostPool& get() {
    static HostPool p;
    return p;
  }
  // fn(part) for part in [0, parts), spread over the pool and the calling thread; returns when all are done
  void run(int parts, const std::function<void(int)>& fn) {
    if (parts <= 1 || workers_.empty()) {
      for (int i = 0; i < parts; ++i) fn(i);
      return;
    }
    std::unique_lock<std::mutex> call_lock(call_mu_);  // one job at a time
    {
      std::lock_guard<std::mutex> lk(mu_);
      fn_ = &fn;
      parts_ = parts;
      next_.store(0);
      pending_ = parts;
      ++generation_;
    }
    cv_.notify_all();
    work();
    std::unique_lock<std::mutex> lk(mu_);
    done_cv_.wait(lk, [&] { return pending_ == 0; });
    fn_ = nullptr;
  }
  int size() const { return (int)workers_.size() + 1; }

 private:
  HostPool() {
    unsigned hw = std::thread::hardware_concurrency();
    int n = (int)std::min(8u, hw > 2 ? hw / 2 : 1u) - 1;
    if (const char* e = getenv("GBLUR_HOST_THREADS")) n = std::max(0, atoi(e) - 1);
    for (int i = 0; i < n; ++i) workers_.emplace_back([this] { loop(); });
  }
  ~HostPool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  void work() {
    const std::function<void(int)>* fn;
    int parts;
    {
      std::lock_guard<std::mutex> lk(mu_);
      fn = fn_;
      parts = parts_;
    }
    if (!fn) return;
    int done = 0;
    for (int i = next_.fetch_add(1); i < parts; i = next_.fetch_add(1)) {
      (*fn)(i);
      ++done;
    }
    if (done) {
      std::lock_guard<std::mutex> lk(mu_);
      pending_ -= done;
      if (pending_ == 0) done_cv_.notify_all();
    }
  }
  void loop() {
    unsigned long long seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return stop_ || generation_ != seen; });
        if (stop_) return;
        seen = generation_;
      }
      work();
    }
  }
  std::vector<std::thread> workers_;
  std::mutex mu_, call_mu_;
  std::condition_variable cv_, done_cv_;
  const std::function<void(int)>* fn_ = nullptr;
  int parts_ = 0, pending_ = 0;
  std::atomic<int> next_{0};
  unsigned long long generation_ = 0;
  bool stop_ = false;
};

// rows x width bytes between two pitched host buffers, split over the pool
void host_copy_2d(uint8_t* dst, size_t dst_pitch, const uint8_t* src, size_t src_pitch, size_t width, size_t rows) {
  HostPool& pool = HostPool::get();
  const int parts = (int)std::min<size_t>((size_t)pool.size() * 2, std::max<size_t>(1, rows * width / (1u << 20)));
  pool.run(parts, [&](int i) {
    const size_t r0 = rows * (size_t)i / parts, r1 = rows * (size_t)(i + 1) / parts;
    if (dst_pitch == width && src_pitch == width) {
      std::memcpy(dst + r0 * width, src + r0 * width, (r1 - r0) * width);
    } else {
      for (size_t r = r0; r < r1; ++r) std::memcpy(dst + r * dst_pitch, src + r * src_pitch, width);
    }
  });
}

bool host_pointer_is_pinned(const void* ptr) {
  cudaPointerAttributes at;
  const cudaError_t e = cudaPointerGetAttributes(&at, ptr);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

// One image, pipelined against PCIe: the rows arrive in K chunks and each chunk's row pass runs while the
// next one is still on the bus; the column pass runs in K bands of pixel columns and each band leaves for
// the host (strided 2-D copy) while the next one is computed.
int run_host_single_pipelined(gblur_plan* p, uint8_t* hwc, int alpha, int K) {
  const size_t row_bytes = (size_t)p->cols * p->channels;
  int rc = ensure_stage(&p->d_stage, &p->stage_bytes, row_bytes * p->rows);
  if (rc != GBLUR_OK) return rc;
  rc = ensure_pipe(p, (size_t)3 * K + 2);
  if (rc != GBLUR_OK) return rc;
  cudaStream_t sc = p->stream;
  FusedCtx c;
  rc = fused_prepare(p, 1, alpha, sc, c);
  if (rc != GBLUR_OK) return rc;
  // a pageable buffer travels through three pinned staging slots (host threads copy, the copy engines stream)
  const int px_per_band = std::max(32, ((p->cols + K - 1) / K + 31) / 32 * 32);
  const bool staged_host = !host_pointer_is_pinned(hwc) && getenv("GBLUR_NO_HOST_RING") == nullptr;
  if (staged_host) {
    const size_t slot = std::max(((size_t)(p->rows + K - 1) / K + 1) * row_bytes, (size_t)px_per_band * p->channels * p->rows);
    if (p->h_ring_bytes < slot) {
      for (uint8_t*& h : p->h_ring) {
        if (h) cudaFreeHost(h);
        h = nullptr;
      }
      p->h_ring_bytes = 0;
      for (uint8_t*& h : p->h_ring) CU(cudaHostAlloc(&h, slot, cudaHostAllocDefault));
      p->h_ring_bytes = slot;
    }
  }
  // the copy stream must not run ahead of a previous call's D2H (same staging buffer)
  for (int k = 0; k < K; ++k) {
    const int r0 = (int)((long long)p->rows * k / K), r1 = (int)((long long)p->rows * (k + 1) / K);
    if (r1 <= r0) continue;
    const uint8_t* src = hwc + r0 * row_bytes;
    if (staged_host) {
      uint8_t* slot = p->h_ring[k % 3];
      if (k >= 3) CU(cudaEventSynchronize(p->events[k - 3]));  // the slot's previous H2D has drained
      host_copy_2d(slot, row_bytes, src, row_bytes, row_bytes, (size_t)(r1 - r0));
      src = slot;
    }
    CU(cudaMemcpyAsync(p->d_stage + r0 * row_bytes, src, (size_t)(r1 - r0) * row_bytes, cudaMemcpyHostToDevice, p->s_in));
    CU(cudaEventRecord(p->events[k], p->s_in));
    CU(cudaStreamWaitEvent(sc, p->events[k], 0));
    rc = run_rowpass(p, p->d_stage, p->d_inter, c.pitch, 1, c.alpha, sc, c.bp, 0, 0, r0, r1 - r0, true);
    if (rc != GBLUR_OK) return rc;
  }
  // column bands on 32-pixel boundaries (=> float-column boundaries on multiples of 32 for any Cp)
  int e = K;
  int nb = 0;  // bands issued
  struct Band {
    int x0, x1;
  } bands[64];
  auto drain = [&](int b) -> int {  // staged host: band b has landed in its slot, copy it out to the caller's image
    CU(cudaEventSynchronize(p->events[2 * K + 1 + b]));
    const size_t w = (size_t)(bands[b].x1 - bands[b].x0) * p->channels;
    host_copy_2d(hwc + (size_t)bands[b].x0 * p->channels, row_bytes, p->h_ring[b % 3], w, w, (size_t)p->rows);
    return GBLUR_OK;
  };
  for (int x0 = 0; x0 < p->cols && nb < 64; x0 += px_per_band, ++e, ++nb) {
    const int x1 = std::min(p->cols, x0 + px_per_band);
    bands[nb] = {x0, x1};
    rc = run_colpass(p, p->d_inter, c.pitch, p->d_stage, (long long)row_bytes, 1, c.alpha, sc, c.bp, x0 * c.cp, x1 * c.cp,
                     p->d_stage, x0 == 0, c.alpha_restore ? p->d_stage : nullptr);
    if (rc != GBLUR_OK) return rc;
    CU(cudaEventRecord(p->events[e], sc));
    CU(cudaStreamWaitEvent(p->s_out, p->events[e], 0));
    const size_t w = (size_t)(x1 - x0) * p->channels;
    if (staged_host) {
      if (nb >= 3) {  // the slot is reused: its previous band must have been copied out
        rc = drain(nb - 3);
        if (rc != GBLUR_OK) return rc;
      }
      CU(cudaMemcpy2DAsync(p->h_ring[nb % 3], w, p->d_stage + (size_t)x0 * p->channels, row_bytes, w, (size_t)p->rows,
                           cudaMemcpyDeviceToHost, p->s_out));
      CU(cudaEventRecord(p->events[2 * K + 1 + nb], p->s_out));
    } else {
      CU(cudaMemcpy2DAsync(hwc + (size_t)x0 * p->channels, row_bytes, p->d_stage + (size_t)x0 * p->channels, row_bytes, w,
                           (size_t)p->rows, cudaMemcpyDeviceToHost, p->s_out));
    }
  }
  if (staged_host)
    for (int b = std::max(0, nb - 3); b < nb; ++b) {
      rc = drain(b);
      if (rc != GBLUR_OK) return rc;
    }
  CU(cudaStreamSynchronize(p->s_out));
  CU(cudaStreamSynchronize(sc));
  return GBLUR_OK;
}

// A batch, pipelined by chunks of max_frames frames through a ring of three staging buffers:
// H2D(chunk k+1) || blur(chunk k) || D2H(chunk k-1). Three, not two: with two, the H2D of chunk k+2 waits for the D2H
// of chunk k, which started one blur after the H2D of chunk k+1 did — the inbound link idles for one blur per chunk
// (8K RGBA: 3.3 ms per frame instead of the 2.7 ms one PCIe direction takes).
int run_host_batch_pipelined(gblur_plan* p, uint8_t* hwc, int n_frames, int alpha) {
  const size_t frame_bytes = (size_t)p->rows * p->cols * p->channels;
  const int chunk = p->max_frames;
  constexpr int kRing = 3;
  int rc = ensure_stage(&p->d_stage, &p->stage_bytes, frame_bytes * chunk);
  if (rc == GBLUR_OK) rc = ensure_stage(&p->d_stage2, &p->stage2_bytes, frame_bytes * chunk);
  if (rc == GBLUR_OK && n_frames > 2 * chunk) rc = ensure_stage(&p->d_stage3, &p->stage3_bytes, frame_bytes * chunk);
  if (rc == GBLUR_OK) rc = ensure_pipe(p, 3 * kRing);
  if (rc != GBLUR_OK) return rc;
  uint8_t* buf[kRing] = {p->d_stage, p->d_stage2, p->d_stage3};
  const int ring = n_frames > 2 * chunk ? kRing : 2;
  cudaEvent_t *ev_in = &p->events[0], *ev_cmp = &p->events[kRing], *ev_out = &p->events[2 * kRing];
  int k = 0;
  for (int f = 0; f < n_frames; f += chunk, ++k) {
    const int b = k % ring, n = std::min(chunk, n_frames - f);
    if (k >= ring) CU(cudaStreamWaitEvent(p->s_in, ev_out[b], 0));  // the buffer's previous D2H has drained
    CU(cudaMemcpyAsync(buf[b], hwc + (size_t)f * frame_bytes, frame_bytes * n, cudaMemcpyHostToDevice, p->s_in));
    CU(cudaEventRecord(ev_in[b], p->s_in));
    CU(cudaStreamWaitEvent(p->stream, ev_in[b], 0));
    rc = gblur_run_batch_device(p, buf[b], buf[b], n, alpha, p->stream);
    if (rc != GBLUR_OK) return rc;
    CU(cudaEventRecord(ev_cmp[b], p->stream));
    CU(cudaStreamWaitEvent(p->s_out, ev_cmp[b], 0));
    CU(cudaMemcpyAsync(hwc + (size_t)f * frame_bytes, buf[b], frame_bytes * n, cudaMemcpyDeviceToHost, p->s_out));
    CU(cudaEventRecord(ev_out[b], p->s_out));
  }
  CU(cudaStreamSynchronize(p->s_out));
  CU(cudaStreamSynchronize(p->stream));
  return GBLUR_OK;
}
}  // namespace

int gblur_run_batch_host(gblur_plan* p, uint8_t* hwc, int n_frames, int apply_to_alpha) {
  if (!p || !hwc) return fail(GBLUR_E_INVALID, "null argument");
  if (p->band) return fail(GBLUR_E_INVALID, "band plans run through gblur_rowpass_device / gblur_colpass_device");
  if (n_frames < 1) return fail(GBLUR_E_INVALID, "n_frames %d < 1", n_frames);
  std::lock_guard<std::mutex> host_lock(p->host_mu);
  DeviceGuard guard(p->device);
  const size_t frame_bytes = (size_t)p->rows * p->cols * p->channels;
  static const int pipe_chunks = [] {
    const char* e = getenv("GBLUR_PIPE_CHUNKS");
    return e ? atoi(e) : 8;
  }();
  if (n_frames > p->max_frames) return run_host_batch_pipelined(p, hwc, n_frames, apply_to_alpha);
  // a single large image on a panel/row-major single-CTA plan: overlap PCIe with the passes
  if (n_frames == 1 && pipe_chunks > 1 && frame_bytes >= (8u << 20) && p->cols >= 64 * pipe_chunks &&
      p->rows >= 8 * pipe_chunks && p->col.lp.cluster == 1)
    return run_host_single_pipelined(p, hwc, apply_to_alpha, std::min(pipe_chunks, 16));
  const size_t bytes = frame_bytes * (size_t)n_frames;
  int rc = ensure_stage(&p->d_stage, &p->stage_bytes, bytes);
  if (rc != GBLUR_OK) return rc;
  CU(cudaMemcpyAsync(p->d_stage, hwc, bytes, cudaMemcpyHostToDevice, p->stream));
  rc = gblur_run_batch_device(p, p->d_stage, p->d_stage, n_frames, apply_to_alpha, p->stream);
  if (rc != GBLUR_OK) return rc;
  CU(cudaMemcpyAsync(hwc, p->d_stage, bytes, cudaMemcpyDeviceToHost, p->stream));
  CU(cudaStreamSynchronize(p->stream));
  return GBLUR_OK;
}

int gblur_run_host(gblur_plan* p, uint8_t* hwc, int apply_to_alpha) {
  return gblur_run_batch_host(p, hwc, 1, apply_to_alpha);
}

int gblur_rowpass_device(gblur_plan* p, const uint8_t* d_in, float* d_inter, long long inter_pitch_floats,
                         int apply_to_alpha, void* cuda_stream) {
  if (!p || !d_in || !d_inter) return fail(GBLUR_E_INVALID, "null argument");
  const int cp = processed_channels(p, apply_to_alpha);
  if (inter_pitch_floats < (long long)p->full_cols * cp) return fail(GBLUR_E_INVALID, "intermediate pitch too small");
  DeviceGuard guard(p->device);
  return run_rowpass(p, d_in, d_inter, inter_pitch_floats, 1, apply_to_alpha, static_cast<cudaStream_t>(cuda_stream));
}

int gblur_rowpass_split_device(gblur_plan* p, const uint8_t* d_in, float* d_inter, int split_cols,
                               long long split_stride_floats, int apply_to_alpha, void* cuda_stream) {
  if (!p || !d_in || !d_inter) return fail(GBLUR_E_INVALID, "null argument");
  if (split_cols < 1 || split_cols > p->full_cols) return fail(GBLUR_E_INVALID, "split_cols outside [1, cols]");
  const int cp = processed_channels(p, apply_to_alpha);
  if (split_stride_floats < (long long)p->rows * split_cols * cp)
    return fail(GBLUR_E_INVALID, "split stride smaller than one band block");
  DeviceGuard guard(p->device);
  return run_rowpass(p, d_in, d_inter, (long long)split_cols * cp, 1, apply_to_alpha,
                     static_cast<cudaStream_t>(cuda_stream), nullptr, split_cols, split_stride_floats);
}

int gblur_colpass_device(gblur_plan* p, const float* d_inter, long long inter_pitch_floats, uint8_t* d_out,
                         long long out_pitch_bytes, int apply_to_alpha, void* cuda_stream) {
  if (!p || !d_inter || !d_out) return fail(GBLUR_E_INVALID, "null argument");
  const int cp = processed_channels(p, apply_to_alpha);
  if (inter_pitch_floats % 4 || inter_pitch_floats < default_pitch((long long)p->cols * cp))
    return fail(GBLUR_E_INVALID, "intermediate pitch must be a multiple of 4 floats covering band_cols*Cp rounded up to 4");
  if (out_pitch_bytes < (long long)p->cols * p->channels) return fail(GBLUR_E_INVALID, "output pitch too small");
  DeviceGuard guard(p->device);
  return run_colpass(p, d_inter, inter_pitch_floats, d_out, out_pitch_bytes, 1, apply_to_alpha,
                     static_cast<cudaStream_t>(cuda_stream));
}

/* ---- band-sharded fused exchange: the row pass stores straight into the column owners' panel buffers ---- */

namespace {
// Panel geometry every rank derives from the (identical) column plan: slots per panel and the
// bank-padding period. Cluster column plans read the panel through their loader: no padding holes.
void panel_geometry(const gblur_plan* p, int* np, int* padq) {
  const LinePlan& c = p->col.lp;
  *np = c.cluster > 1 ? c.N : c.exch_elems;
  *padq = c.cluster > 1 ? 0 : c.padq;
}
}  // namespace

int gblur_plan_panel_geometry(const gblur_plan* p, int* np, int* padq) {
  if (!p) return fail(GBLUR_E_INVALID, "null plan");
  int a, b;
  panel_geometry(p, &a, &b);
  if (np) *np = a;
  if (padq) *padq = b;
  return GBLUR_OK;
}

int gblur_band_panel_alloc(gblur_plan* p, int apply_to_alpha, float** d_panel, long long* n_floats) {
  if (!p || !d_panel) return fail(GBLUR_E_INVALID, "null argument");
  DeviceGuard guard(p->device);
  const int cp = processed_channels(p, apply_to_alpha);
  int np, padq;
  panel_geometry(p, &np, &padq);
  const long long panels = ((long long)p->cols * cp + 31) / 32;
  const long long floats = panels * np * 32;
  void* q = nullptr;
  CU(cudaMalloc(&q, (size_t)floats * sizeof(float)));
  CU(cudaMemset(q, 0, (size_t)floats * sizeof(float)));  // padding holes and the zero tail stay zero for ever
  CU(cudaDeviceSynchronize());
  p->owned.push_back(q);
  *d_panel = static_cast<float*>(q);
  if (n_floats) *n_floats = floats;
  return GBLUR_OK;
}

int gblur_ipc_export(const void* d_ptr, unsigned char handle[64]) {
  if (!d_ptr || !handle) return fail(GBLUR_E_INVALID, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr)));
  std::memcpy(handle, &h, 64);
  return GBLUR_OK;
}

int gblur_ipc_open(const unsigned char handle[64], void** d_ptr) {
  if (!handle || !d_ptr) return fail(GBLUR_E_INVALID, "null argument");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, 64);
  CU(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return GBLUR_OK;
}

int gblur_ipc_close(void* d_ptr) {
  if (!d_ptr) return GBLUR_OK;
  CU(cudaIpcCloseMemHandle(d_ptr));
  return GBLUR_OK;
}

int gblur_rowpass_peer_device(gblur_plan* p, const uint8_t* d_in, float* const* peer_panels, int n_peers,
                              int row_offset, int apply_to_alpha, void* cuda_stream) {
  if (!p || !d_in || !peer_panels) return fail(GBLUR_E_INVALID, "null argument");
  if (n_peers < 1 || n_peers > 8) return fail(GBLUR_E_INVALID, "n_peers outside [1, 8]");
  const int cp = processed_channels(p, apply_to_alpha);
  const long long w = (long long)p->full_cols * cp;
  if (w % (32LL * n_peers)) return fail(GBLUR_E_INVALID, "cols*Cp must be a multiple of 32*n_peers for the fused exchange");
  if (row_offset < 0 || row_offset + p->rows > p->full_rows) return fail(GBLUR_E_INVALID, "row band outside the image");
  DeviceGuard guard(p->device);
  int np, padq;
  panel_geometry(p, &np, &padq);
  BlockedLayout b;
  b.enabled = 1;
  b.L = 2 * p->col.lp.nseq;
  b.np = np;
  b.padq = padq;
  b.rows_per_frame = p->full_rows;
  b.frame_stride = 0;
  RowLaunch a{};
  a.in = d_in;
  a.in_pitch = (long long)p->full_cols * p->channels;
  a.C = p->channels;
  a.Cp = cp;
  a.T = p->full_cols;
  a.pad = p->sz.pad;
  a.rows_total = p->rows;
  a.inter = nullptr;
  a.inter_pitch = 0;
  a.blk = b;
  a.npeers = n_peers;
  a.peer_panels = (int)(w / 32 / n_peers);
  a.row_offset = row_offset;
  for (int i = 0; i < n_peers; ++i) {
    if (!peer_panels[i]) return fail(GBLUR_E_INVALID, "null peer panel pointer");
    a.peer[i] = peer_panels[i];
  }
  CU(launch_rowpass(p->row.lp, a, p->num_sms, static_cast<cudaStream_t>(cuda_stream)));
  return GBLUR_OK;
}

int gblur_panel_push_device(const gblur_plan* p, int apply_to_alpha, const float* src_panel, float* dst_panel,
                            int slot_begin, int slot_end, void* cuda_stream) {
  if (!p || !src_panel || !dst_panel) return fail(GBLUR_E_INVALID, "null argument");
  int np, padq;
  panel_geometry(p, &np, &padq);
  const int n_slots = p->sz.n_colpass;
  if (slot_begin < 0 || slot_end > n_slots || slot_begin > slot_end) return fail(GBLUR_E_INVALID, "slot range outside the padded line");
  if (slot_begin == slot_end) return GBLUR_OK;
  DeviceGuard guard(p->device);
  const int cp = processed_channels(p, apply_to_alpha);
  const long long panels = ((long long)p->cols * cp + 31) / 32;
  auto physical = [padq](long long n) { return padq ? n + n / padq : n; };  // fftconv.cuh phys()
  const long long first = physical(slot_begin), last = physical(slot_end - 1);  // padding holes inside travel too (zeros)
  const size_t pitch = (size_t)np * 128, width = (size_t)(last - first + 1) * 128;
  CU(cudaMemcpy2DAsync(dst_panel + first * 32, pitch, src_panel + first * 32, pitch, width, (size_t)panels,
                       cudaMemcpyDefault, static_cast<cudaStream_t>(cuda_stream)));
  return GBLUR_OK;
}

int gblur_colpass_panel_device(gblur_plan* p, const float* d_panel, uint8_t* d_out, long long out_pitch_bytes,
                               int apply_to_alpha, void* cuda_stream) {
  if (!p || !d_panel || !d_out) return fail(GBLUR_E_INVALID, "null argument");
  if (out_pitch_bytes < (long long)p->cols * p->channels) return fail(GBLUR_E_INVALID, "output pitch too small");
  DeviceGuard guard(p->device);
  const int cp = processed_channels(p, apply_to_alpha);
  int np, padq;
  panel_geometry(p, &np, &padq);
  BlockedLayout b;
  b.enabled = 1;
  b.L = 2 * p->col.lp.nseq;
  b.np = np;
  b.padq = padq;
  b.rows_per_frame = p->full_rows;
  b.frame_stride = (((long long)p->cols * cp + 31) / 32) * np * 32;
  ColLaunch a{};
  a.inter = d_panel;
  a.pitch = 0;
  a.frame_stride = b.frame_stride;
  a.frames = 1;
  a.T = p->full_rows;
  a.pad = p->sz.pad;
  a.W = p->cols * cp;
  a.out = d_out;
  a.out_pitch = out_pitch_bytes;
  a.out_frame_stride = (long long)p->full_rows * out_pitch_bytes;
  a.C = p->channels;
  a.Cp = cp;
  a.blk = b;
  a.panel_in = p->col.lp.cluster > 1 ? 1 : 0;
  if (!a.panel_in && panel_tensor_map(p, d_panel, b, b.frame_stride / ((long long)np * 32))) {
    a.tmap = &p->tmap;
    a.tma_rows = p->tmap_rows;
  }
  CU(launch_colpass(p->col.lp, a, p->num_sms, static_cast<cudaStream_t>(cuda_stream)));
  return GBLUR_OK;
}

/* ---- halo sharding: row bands for BOTH passes (third row) ----
 * With overlap-save column blocks a rank needs nothing from the others but pad input rows on either side of its band
 * and one integer per byte column (the column sums behind the reference's Nyquist term, src/gaussianblur.cpp:62-79):
 * it runs the row pass on its rows plus the halo, then the column blocks whose outputs it owns. No fp32 intermediate
 * crosses a link. Rank boundaries sit on block boundaries (multiples of L = 1024 - 2 pad rows). */
int gblur_halo_partition(int rows, int cols, float sigma, int n_ranks, int rank, int* out_row0, int* out_row1, int* in_row0,
                         int* in_row1) {
  if (rows < 1 || cols < 1 || !(sigma > 0.f) || n_ranks < 1 || rank < 0 || rank >= n_ranks)
    return fail(GBLUR_E_INVALID, "bad halo partition query");
  const host::Sizes sz = host::plan_sizes(rows, cols, sigma);
  const int L = 1024 - 2 * sz.pad;
  if (L < 512) return fail(GBLUR_E_UNSUPPORTED, "pad %d too wide for 1024-slot blocks", sz.pad);
  const int nblk = (rows + L - 1) / L;
  const int q = nblk / n_ranks, r = nblk % n_ranks;
  const int b0 = rank * q + std::min(rank, r), b1 = b0 + q + (rank < r ? 1 : 0);
  const int o0 = std::min(rows, b0 * L), o1 = std::min(rows, b1 * L);
  const int np = (b1 - b0) * L + 2 * sz.pad;
  if (out_row0) *out_row0 = o0;
  if (out_row1) *out_row1 = o1;
  if (in_row0) *in_row0 = b1 > b0 ? std::max(0, b0 * L - sz.pad) : 0;
  if (in_row1) *in_row1 = b1 > b0 ? std::min(rows, b0 * L + np - sz.pad) : 0;
  return GBLUR_OK;
}

int gblur_plan_create_halo(gblur_plan** out, int device, int full_rows, int full_cols, int channels, float sigma, int n_ranks,
                           int rank) {
  int o0, o1, i0, i1;
  int rc = gblur_halo_partition(full_rows, full_cols, sigma, n_ranks, rank, &o0, &o1, &i0, &i1);
  if (rc != GBLUR_OK) return rc;
  if (channels != 1 && channels != 4) return fail(GBLUR_E_UNSUPPORTED, "halo sharding: 1 or 4 channels");
  if (((long long)full_cols * channels) % 16) return fail(GBLUR_E_UNSUPPORTED, "halo sharding: cols * channels must be a multiple of 16");
  rc = create_common(out, device, full_rows, full_cols, channels, sigma, std::max(1, i1 - i0), full_cols, 1, true);
  if (rc != GBLUR_OK) return rc;
  gblur_plan* p = *out;
  if (!p->ols_row) {
    const std::vector<float> taps = host::gaussian_taps(sigma, p->sz.ksize);
    const cudaError_t ce = ols::row_plan_create(&p->ols_row, taps, p->sz.n_rowpass, p->sz.n_colpass);
    if (ce != cudaSuccess || !p->ols_row) {
      gblur_plan_destroy(p);
      *out = nullptr;
      return fail(GBLUR_E_UNSUPPORTED, "halo sharding needs the overlap-save kernels (pad <= 256)");
    }
  }
  p->halo = true;
  p->h_ranks = n_ranks;
  p->h_L = 1024 - 2 * p->sz.pad;
  p->h_blk0 = o0 / p->h_L;
  p->h_blk1 = o1 > o0 ? (o1 + p->h_L - 1) / p->h_L : p->h_blk0;
  p->h_out0 = o0;
  p->h_out1 = o1;
  p->h_in0 = i0;
  p->h_in1 = i1;
  p->h_np = (p->h_blk1 - p->h_blk0) * p->h_L + 2 * p->sz.pad;
  if (o1 > o0) {
    DeviceGuard guard(device);
    const long long panels = ((long long)full_cols * channels + 31) / 32;
    const size_t bytes = (size_t)panels * p->h_np * 32 * sizeof(float);
    CU(cudaMalloc(&p->d_halo_panel, bytes));
    CU(cudaMemset(p->d_halo_panel, 0, bytes));  // the zero tail beyond the padded column stays zero for ever
    CU(cudaDeviceSynchronize());
  }
  return GBLUR_OK;
}

int gblur_halo_info(const gblur_plan* p, int* out_row0, int* out_row1, int* in_row0, int* in_row1) {
  if (!p || !p->halo) return fail(GBLUR_E_INVALID, "not a halo plan");
  if (out_row0) *out_row0 = p->h_out0;
  if (out_row1) *out_row1 = p->h_out1;
  if (in_row0) *in_row0 = p->h_in0;
  if (in_row1) *in_row1 = p->h_in1;
  return GBLUR_OK;
}

namespace {
ColLaunch halo_col_launch(const gblur_plan* p, uint8_t* d_out) {
  ColLaunch a{};
  const int C = p->channels;
  const long long panels = ((long long)p->full_cols * C + 31) / 32;
  a.inter = p->d_halo_panel;
  a.frames = 1;
  a.T = p->h_out1 - p->h_out0;
  a.pad = p->sz.pad;
  a.W = p->full_cols * C;
  a.out = d_out;
  a.out_pitch = (long long)p->full_cols * C;
  a.out_frame_stride = (long long)std::max(1, a.T) * a.out_pitch;
  a.C = a.Cp = C;
  a.blk.enabled = 1;
  a.blk.L = 4;
  a.blk.np = p->h_np;
  a.blk.padq = 0;
  a.blk.pair_major = C == 4 ? 1 : 0;
  a.blk.slot_origin = p->h_blk0 * p->h_L;
  a.blk.rows_per_frame = p->full_rows;
  a.blk.frame_stride = panels * p->h_np * 32;
  a.row_origin = p->h_out0;
  return a;
}
}  // namespace

/* d_in_rows: the image rows [in_row0, in_row1) of this rank, HWC u8, pitch cols * channels. d_u: cols * channels ints the
 * caller zeroed; this rank's OWNED rows are added (integer atomics). Sum d_u over the ranks (all-reduce) before
 * gblur_halo_run_device. */
int gblur_halo_colsum_device(gblur_plan* p, const uint8_t* d_in_rows, int* d_u, void* cuda_stream) {
  if (!p || !p->halo || !d_u) return fail(GBLUR_E_INVALID, "bad halo colsum call");
  if (p->h_out1 <= p->h_out0) return GBLUR_OK;
  if (!d_in_rows) return fail(GBLUR_E_INVALID, "null input");
  DeviceGuard guard(p->device);
  const ColLaunch a = halo_col_launch(p, nullptr);
  CU(ols::launch_colsum(a, d_in_rows, (long long)p->full_cols * p->channels, d_u, p->h_out0, p->h_out1, p->h_in0,
                        p->full_rows, static_cast<cudaStream_t>(cuda_stream)));
  return GBLUR_OK;
}

/* Row pass of the image rows [row_begin, row_end) (a sub-range of [in_row0, in_row1); d_in_rows still points at
 * in_row0) into the rank's panel buffer. A rank can start on the rows it holds while the halo rows are still on the
 * link and run them in a second call. */
namespace {
int halo_rowpass(gblur_plan* p, const uint8_t* d_in_rows, int a0, int a1, int b0, int b1, void* cuda_stream) {
  DeviceGuard guard(p->device);
  const int C = p->channels;
  const ColLaunch c = halo_col_launch(p, nullptr);
  RowLaunch a{};
  a.in_pitch = (long long)p->full_cols * C;
  a.in = d_in_rows + (long long)(a0 - p->h_in0) * a.in_pitch;
  a.C = a.Cp = C;
  a.T = p->full_cols;
  a.pad = p->sz.pad;
  a.rows_total = (a1 - a0) + (b1 - b0);
  a.inter = p->d_halo_panel;
  a.blk = c.blk;
  a.row_offset = a0;
  if (b1 > b0) {
    a.split_at = a1 - a0;
    a.split_gap = b0 - a1;
  }
  if (!ols::row_eligible(p->ols_row, a))
    return fail(GBLUR_E_UNSUPPORTED, "halo sharding: geometry not eligible for the overlap-save kernels (alignment)");
  // The block kernel is persistent and takes every register of the SMs it runs on: communication kernels enqueued
  // beside it (the binding's halo send/recv and the all-reduce of the column sums) would wait for it to END. A few
  // SMs left free (GBLUR_HALO_RESERVE_SMS, read once; default 4) let them run under it: measured on 8 B200, 32768^2
  // sigma=50: 1.068 ms per step with 0, 1.036 with 4, 1.044 with 8 (profiles/r02/scale/n8_reserve*.json).
  static const int reserve = [] {
    const char* e = getenv("GBLUR_HALO_RESERVE_SMS");
    return e ? std::max(0, atoi(e)) : 4;
  }();
  const int sms = std::max(1, p->num_sms - (p->h_ranks > 1 ? reserve : 0));
  CU(ols::launch_row(p->ols_row, a, sms, static_cast<cudaStream_t>(cuda_stream)));
  p->last_row_kind = 1;
  return GBLUR_OK;
}
}  // namespace

int gblur_halo_rowpass_device(gblur_plan* p, const uint8_t* d_in_rows, int row_begin, int row_end, void* cuda_stream) {
  if (!p || !p->halo) return fail(GBLUR_E_INVALID, "not a halo plan");
  if (p->h_out1 <= p->h_out0 || row_end <= row_begin) return GBLUR_OK;
  if (!d_in_rows) return fail(GBLUR_E_INVALID, "null argument");
  if (row_begin < p->h_in0 || row_end > p->h_in1) return fail(GBLUR_E_INVALID, "rows outside [in_row0, in_row1)");
  return halo_rowpass(p, d_in_rows, row_begin, row_end, row_end, row_end, cuda_stream);
}

/* The same over TWO row ranges [begin_a, end_a) and [begin_b, end_b) (end_a <= begin_b) in ONE launch: the halo rows
 * above and below a rank's band once they have arrived. Two launches of 165 rows each cost 0.125 ms on a B200 where the
 * same rows inside one launch cost 0.05. Either range may be empty. */
int gblur_halo_rowpass2_device(gblur_plan* p, const uint8_t* d_in_rows, int begin_a, int end_a, int begin_b, int end_b,
                               void* cuda_stream) {
  if (!p || !p->halo) return fail(GBLUR_E_INVALID, "not a halo plan");
  if (p->h_out1 <= p->h_out0) return GBLUR_OK;
  if (!d_in_rows) return fail(GBLUR_E_INVALID, "null argument");
  if (end_a < begin_a) end_a = begin_a;
  if (end_b < begin_b) end_b = begin_b;
  if (end_a == begin_a && end_b == begin_b) return GBLUR_OK;
  if (end_a == begin_a) return gblur_halo_rowpass_device(p, d_in_rows, begin_b, end_b, cuda_stream);
  if (end_b == begin_b) return gblur_halo_rowpass_device(p, d_in_rows, begin_a, end_a, cuda_stream);
  if (begin_a < p->h_in0 || end_b > p->h_in1 || begin_b < end_a)
    return fail(GBLUR_E_INVALID, "row ranges must be ordered, disjoint and inside [in_row0, in_row1)");
  return halo_rowpass(p, d_in_rows, begin_a, end_a, begin_b, end_b, cuda_stream);
}

/* Column terms from the summed d_u_total, then the column blocks of the owned rows into d_out
 * ([out_row1 - out_row0][cols * channels] u8). */
int gblur_halo_colpass_device(gblur_plan* p, const int* d_u_total, uint8_t* d_out, void* cuda_stream) {
  if (!p || !p->halo) return fail(GBLUR_E_INVALID, "not a halo plan");
  if (p->h_out1 <= p->h_out0) return GBLUR_OK;
  if (!d_u_total || !d_out) return fail(GBLUR_E_INVALID, "null argument");
  DeviceGuard guard(p->device);
  cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
  const ColLaunch c = halo_col_launch(p, d_out);
  if (!ols::col_eligible(p->ols_row, c))
    return fail(GBLUR_E_UNSUPPORTED, "halo sharding: geometry not eligible for the overlap-save kernels (alignment)");
  CU(ols::launch_col_terms_from_u(p->ols_row, c, d_u_total, s));
  CU(ols::launch_col(p->ols_row, c, p->num_sms, s));
  p->last_col_kind = 1;
  return GBLUR_OK;
}

/* Both of the above over all of the rank's rows. All channels are blurred (apply_to_alpha semantics: true). */
int gblur_halo_run_device(gblur_plan* p, const uint8_t* d_in_rows, const int* d_u_total, uint8_t* d_out, void* cuda_stream) {
  if (!p || !p->halo) return fail(GBLUR_E_INVALID, "not a halo plan");
  const int rc = gblur_halo_rowpass_device(p, d_in_rows, p->h_in0, p->h_in1, cuda_stream);
  if (rc != GBLUR_OK) return rc;
  return gblur_halo_colpass_device(p, d_u_total, d_out, cuda_stream);
}

int gblur_dump_intermediate(gblur_plan* p, float* host_out, int apply_to_alpha) {
  if (!p || !host_out) return fail(GBLUR_E_INVALID, "null argument");
  if (!p->d_inter || p->inter_cp == 0) return fail(GBLUR_E_INVALID, "no run has happened on this plan");
  DeviceGuard guard(p->device);
  int cp = processed_channels(p, apply_to_alpha);
  // RGBA with alpha passed through on the overlap-save passes: the intermediate holds all four channels
  const bool drop_alpha = (cp == 3 && p->channels == 4 && p->inter_cp == 4 + 16 && p->last_alpha_restore);
  if (drop_alpha) cp = 4;
  if (cp != p->inter_cp && -cp != p->inter_cp && cp + 16 != p->inter_cp)
    return fail(GBLUR_E_INVALID, "last run used a different apply_to_alpha");
  if (p->inter_cp < 0) {  // row-major layout
    const long long w = (long long)p->cols * cp, pitch = default_pitch(w);
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy2D(host_out, (size_t)w * sizeof(float), p->d_inter, (size_t)pitch * sizeof(float),
                    (size_t)w * sizeof(float), (size_t)p->rows, cudaMemcpyDeviceToHost));
    return GBLUR_OK;
  }
  const BlockedLayout b = blocked_layout(p, cp, p->inter_cp > 16);
  std::vector<float> blk((size_t)b.frame_stride);
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(blk.data(), p->d_inter, blk.size() * sizeof(float), cudaMemcpyDeviceToHost));
  const long long w = (long long)p->cols * cp;
  for (int r = 0; r < p->rows; ++r) {
    const int n = p->sz.pad + r;
    const long long slot = b.padq ? n + n / b.padq : n;
    for (long long x = 0; x < w; ++x) {
      const int k = (int)(x % 32);  // pixel-major position 4 px + 2 pair + c  ->  pair-major 16 pair + 2 px + c
      const int kk = b.pair_major ? ((k >> 1) & 1) * 16 + (k >> 2) * 2 + (k & 1) : k;
      const float v = blk[(size_t)((x / 32) * b.np + slot) * 32 + (size_t)kk];
      if (!drop_alpha)
        host_out[(size_t)r * w + x] = v;
      else if ((x & 3) != 3)  // [rows][cols * 3]: the caller asked for the three blurred channels
        host_out[(size_t)r * (w / 4 * 3) + (x >> 2) * 3 + (x & 3)] = v;
    }
  }
  return GBLUR_OK;
}

/* ---- host-only helpers (no device needed): the semantics tests call these on CPU boxes ---- */

int gblur_host_plan_sizes(int rows, int cols, float sigma, int* ksize, int* pad, int* n_rowpass, int* n_colpass,
                          int* tz_rows, int* tz_cols) {
  const host::Sizes s = host::plan_sizes(rows, cols, sigma);
  if (ksize) *ksize = s.ksize;
  if (pad) *pad = s.pad;
  if (n_rowpass) *n_rowpass = s.n_rowpass;
  if (n_colpass) *n_colpass = s.n_colpass;
  if (tz_rows) *tz_rows = s.tz_rows;
  if (tz_cols) *tz_cols = s.tz_cols;
  return GBLUR_OK;
}

int gblur_host_taps(float sigma, int ksize, float* out) {
  if (!out || ksize < 1) return fail(GBLUR_E_INVALID, "bad taps query");
  const std::vector<float> t = host::gaussian_taps(sigma, ksize);
  std::memcpy(out, t.data(), t.size() * sizeof(float));
  return GBLUR_OK;
}

int gblur_host_spectrum(float sigma, int ksize, int n, float* out_ordered) {
  if (!out_ordered || n < 32 || ksize < 1) return fail(GBLUR_E_INVALID, "bad spectrum query");
  const std::vector<float> k = host::kerf_ordered(host::kernel_gains(host::gaussian_taps(sigma, ksize), n), n);
  std::memcpy(out_ordered, k.data(), k.size() * sizeof(float));
  return GBLUR_OK;
}

/* Stage plan for FFT length n: radices[<=5], *nstages, *nseq, *static_kernel, *smem_bytes, *nthreads.
 * Returns GBLUR_E_UNSUPPORTED when n does not fit one CTA. */
int gblur_host_line_plan(int n, int* radices, int* nstages, int* nseq, int* static_kernel, long long* smem_bytes,
                         int* nthreads) {
  LinePlan lp;
  if (!host::is_valid_size(n)) return fail(GBLUR_E_INVALID, "%d is not a valid transform size", n);
  if (!host::plan_line(n, lp)) return fail(GBLUR_E_UNSUPPORTED, "FFT length %d does not fit one CTA's shared memory", n);
  if (radices)
    for (int s = 0; s < lp.nstages; ++s) radices[s] = lp.radix[s];
  if (nstages) *nstages = lp.nstages;
  if (nseq) *nseq = lp.nseq;
  if (static_kernel) *static_kernel = lp.static_id >= 0;
  if (smem_bytes) *smem_bytes = (long long)lp.smem_bytes;
  if (nthreads) *nthreads = lp.nthreads;
  return GBLUR_OK;
}

/* Digit-reversed, 1/N-scaled gains for FFT length n exactly as uploaded to the device, plus the
 * twiddle table; lets CPU tests check the permutation against a numpy FFT. */
int gblur_host_kperm(float sigma, int ksize, int n, float* out_n) {
  LinePlan lp;
  if (!out_n) return fail(GBLUR_E_INVALID, "null out");
  if (!host::plan_line(n, lp)) return fail(GBLUR_E_UNSUPPORTED, "FFT length %d does not fit one CTA's shared memory", n);
  const std::vector<float> kp = host::build_kperm(lp, host::kernel_gains(host::gaussian_taps(sigma, ksize), n));
  std::memcpy(out_n, kp.data(), kp.size() * sizeof(float));
  return GBLUR_OK;
}

/* Which kernel family a single-frame fused run of this shape takes when nothing is forced (no device needed; the same
 * rule gblur_run_* applies, apart from the 16-byte alignment of the caller's buffers and the GBLUR_*_OLS overrides):
 * *blocks = 2 both passes as overlap-save blocks, 1 the row pass only, 0 whole-line transforms of the reference's length. */
int gblur_host_kernel_choice(int rows, int cols, int channels, float sigma, int apply_to_alpha, int* blocks) {
  if (!blocks) return fail(GBLUR_E_INVALID, "null out");
  if (rows < 1 || cols < 1 || !(sigma > 0.f)) return fail(GBLUR_E_INVALID, "bad shape or sigma");
  if (channels != 1 && channels != 3 && channels != 4) return fail(GBLUR_E_INVALID, "Unsupported number of channels");
  const host::Sizes sz = host::plan_sizes(rows, cols, sigma);
  *blocks = 0;
  (void)apply_to_alpha;  // RGBA runs all four channels through the blocks either way (alpha restored in the epilogue)
  if (sz.pad > 256 /* ols::kMaxPad */ || channels == 3) return GBLUR_OK;
  const bool groups16 = ((long long)cols * channels) % 16 == 0;  // TMA tiles / staged rows want whole 16-byte groups
  if (groups16 && ols::both_profitable(cols, rows, sz.pad, sz.n_rowpass, sz.n_colpass)) {
    *blocks = 2;
    return GBLUR_OK;
  }
  if (channels == 1 && !groups16) return GBLUR_OK;
  RowLaunch r{};
  r.T = cols;
  r.pad = sz.pad;
  r.C = channels;
  if ((channels == 1 || (channels == 4 && apply_to_alpha)) && ols::row_profitable(r, sz.n_rowpass)) *blocks = 1;
  return GBLUR_OK;
}

}  // extern "C"
