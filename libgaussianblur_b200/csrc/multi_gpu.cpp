// multi_gpu.cpp — the multi-GPU entry points of the C ABI (include/gblur_capi.h: gblur_mg_*), single process.
//
// The reference's own fan-out lives inside its C++ library: hybrid_loop splits the rows of every pass over
// std::threads (/root/reference/include/gaussianblur/helpers.hpp:134-172, called from src/gaussianblur.cpp:212), and
// its setups are shareable between them (include/pffft_pommier/pffft.h:100-103). Here the same two axes of
// parallelism go to the GPUs of one box, behind plain C entry points a C++ caller of gaussianblur() can reach:
//   * frames  — a batch is cut into contiguous blocks of frames, one block per GPU, no exchange at all
//               (BASELINE.json config 3); one host thread per GPU drives that GPU's H2D || blur || D2H pipeline.
//   * bands   — ONE image is cut into row bands (config 4/5). Where the overlap-save kernels take the shape the bands
//               serve BOTH passes (halo mode, run_bands_halo below: pad rows from the host image + the host's sum of
//               the integer column sums; nothing crosses between the GPUs). Otherwise: every GPU runs the row pass of its band and stores each
//               128-byte line of the fp32 intermediate straight into the panel buffer of the GPU that owns those pixel
//               columns (peer stores over NVLink, cudaDeviceEnablePeerAccess; the exchange IS the row pass's store, no
//               collective), events order the passes across devices, and every GPU then runs the column pass of its
//               column band (src/gaussianblur.cpp:261-280: all rows first, then all columns).
// No torch, no NCCL: plans, streams and events only. libgaussianblur_b200/multi.py is the multi-process binding of the
// same kernels (one process per GPU, CUDA IPC instead of peer pointers) that bench.py drives under torchrun.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gblur_capi.h"

namespace {

struct Range {
  int lo, hi;
  int size() const { return hi - lo; }
};
Range shard(int n, int world, int rank) {  // contiguous blocks, the first n % world ranks get one more
  const int q = n / world, r = n % world;
  const int lo = rank * q + std::min(rank, r);
  return {lo, lo + q + (rank < r ? 1 : 0)};
}

struct DevState {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev_row = nullptr;   // this device's row pass (all its peer stores) is complete
  gblur_plan* frames_plan = nullptr;
  int frames_cap = 0;
  // band mode, per alpha mode (index = apply_to_alpha)
  gblur_plan* row_plan[2] = {nullptr, nullptr};
  gblur_plan* col_plan[2] = {nullptr, nullptr};
  float* panel[2] = {nullptr, nullptr};
  uint8_t* d_in = nullptr;   // the row band
  uint8_t* d_out = nullptr;  // the column band [rows][band_cols * C]
  size_t in_bytes = 0, out_bytes = 0;
  // halo mode (row bands for both passes)
  gblur_plan* halo_plan = nullptr;
  int out0 = 0, out1 = 0, in0 = 0, in1 = 0;
  uint8_t* d_hin = nullptr;   // image rows [in0, in1)
  uint8_t* d_hout = nullptr;  // image rows [out0, out1)
  int* d_u = nullptr;         // column sums: this device's rows, then the total
  int* h_u = nullptr;         // pinned: this device's partial sums
  cudaEvent_t ev_in = nullptr, ev_u = nullptr;
};

}  // namespace

struct gblur_mg {
  int n = 0;
  int rows = 0, cols = 0, channels = 0;
  float sigma = 0.f;
  std::vector<DevState> dev;
  bool peer_ok = false;  // every pair of devices can address each other's memory
  int halo_state = 0;    // 0 = not tried, 1 = ready, -1 = not eligible
  int* h_usum = nullptr; // pinned: column sums over all devices
  std::mutex mu;         // one call at a time per handle
  std::string err;
};

namespace {

thread_local std::string g_mg_err;

int mg_fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_mg_err = buf;
  return code;
}

#define MG_CU(call)                                                                              \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess) return mg_fail(GBLUR_E_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
  } while (0)
#define MG_RC(call)                                           \
  do {                                                        \
    int rc_ = (call);                                         \
    if (rc_ != GBLUR_OK) {                                    \
      g_mg_err = gblur_last_error();                          \
      return rc_;                                             \
    }                                                         \
  } while (0)

int processed(int channels, int alpha) { return (channels == 4 && !alpha) ? 3 : channels; }

// the fused peer-store exchange wants whole panels per owner and equal column bands
bool bands_divisible(const gblur_mg* m, int alpha) {
  const long long w = (long long)m->cols * processed(m->channels, alpha);
  return m->cols % m->n == 0 && w % (32LL * m->n) == 0 && (((long long)(m->cols / m->n) * processed(m->channels, alpha)) % 32) == 0;
}

int ensure_band_state(gblur_mg* m, int alpha) {
  for (int g = 0; g < m->n; ++g) {
    DevState& d = m->dev[g];
    if (d.row_plan[alpha]) continue;
    const Range rr = shard(m->rows, m->n, g), cr = shard(m->cols, m->n, g);
    MG_CU(cudaSetDevice(d.device));
    MG_RC(gblur_plan_create_band(&d.row_plan[alpha], d.device, m->rows, m->cols, m->channels, m->sigma, rr.size(), m->cols));
    MG_RC(gblur_plan_create_band(&d.col_plan[alpha], d.device, m->rows, m->cols, m->channels, m->sigma, m->rows, cr.size()));
    MG_RC(gblur_band_panel_alloc(d.col_plan[alpha], alpha, &d.panel[alpha], nullptr));
    const size_t inb = (size_t)rr.size() * m->cols * m->channels, outb = (size_t)m->rows * cr.size() * m->channels;
    if (d.in_bytes < inb) {
      if (d.d_in) cudaFree(d.d_in);
      d.d_in = nullptr;
      MG_CU(cudaMalloc(&d.d_in, inb));
      d.in_bytes = inb;
    }
    if (d.out_bytes < outb) {
      if (d.d_out) cudaFree(d.d_out);
      d.d_out = nullptr;
      MG_CU(cudaMalloc(&d.d_out, outb));
      d.out_bytes = outb;
    }
  }
  return GBLUR_OK;
}


// ---- halo mode: row bands for BOTH passes (include/gblur_capi.h gblur_halo_*). Every device takes the rows of its
// blocks plus `pad` rows on either side straight from the host image (the halo travels over PCIe twice: + 2 pad rows
// per band), forms the column sums of its own rows, the host adds the n partial sums (cols * channels integers each)
// and hands the total back, and every device runs its row pass and its column blocks. Nothing crosses between the
// GPUs: no peer access needed, any width that is a whole number of 16-byte groups, contiguous copies both ways.
// 1 channel, or 4 with alpha blurred; pad <= 256.
bool halo_shape_ok(const gblur_mg* m, int alpha) {
  if (m->n < 2) return false;
  if (const char* e = getenv("GBLUR_MG_BANDS"))
    if (std::strcmp(e, "exchange") == 0) return false;
  if (!(m->channels == 1 || (m->channels == 4 && alpha))) return false;
  if (((long long)m->cols * m->channels) % 16) return false;
  int o0, o1, i0, i1;
  return gblur_halo_partition(m->rows, m->cols, m->sigma, m->n, 0, &o0, &o1, &i0, &i1) == GBLUR_OK;
}

int ensure_halo_state(gblur_mg* m) {
  if (m->halo_state != 0) return GBLUR_OK;
  const size_t W = (size_t)m->cols * m->channels;
  m->halo_state = -1;
  for (int g = 0; g < m->n; ++g) {
    DevState& d = m->dev[(size_t)g];
    MG_CU(cudaSetDevice(d.device));
    const int rc = gblur_plan_create_halo(&d.halo_plan, d.device, m->rows, m->cols, m->channels, m->sigma, m->n, g);
    if (rc == GBLUR_E_UNSUPPORTED) return GBLUR_OK;  // stays -1: the caller falls back to the exchange
    if (rc != GBLUR_OK) {
      g_mg_err = gblur_last_error();
      return rc;
    }
    MG_RC(gblur_halo_info(d.halo_plan, &d.out0, &d.out1, &d.in0, &d.in1));
    if (d.out1 > d.out0) {
      MG_CU(cudaMalloc(&d.d_hin, (size_t)(d.in1 - d.in0) * W));
      MG_CU(cudaMalloc(&d.d_hout, (size_t)(d.out1 - d.out0) * W));
    }
    MG_CU(cudaMalloc(&d.d_u, W * sizeof(int)));
    MG_CU(cudaHostAlloc(&d.h_u, W * sizeof(int), cudaHostAllocPortable));
    MG_CU(cudaEventCreateWithFlags(&d.ev_in, cudaEventDisableTiming));
    MG_CU(cudaEventCreateWithFlags(&d.ev_u, cudaEventDisableTiming));
  }
  MG_CU(cudaHostAlloc(&m->h_usum, W * sizeof(int), cudaHostAllocPortable));
  m->halo_state = 1;
  return GBLUR_OK;
}

int run_bands_halo(gblur_mg* m, uint8_t* hwc) {
  const size_t W = (size_t)m->cols * m->channels;
  // 1. rows + halo host -> device, column sums of the own rows back to the host, row pass
  for (int g = 0; g < m->n; ++g) {
    DevState& d = m->dev[(size_t)g];
    MG_CU(cudaSetDevice(d.device));
    if (d.out1 > d.out0) {
      MG_CU(cudaMemcpyAsync(d.d_hin, hwc + (size_t)d.in0 * W, (size_t)(d.in1 - d.in0) * W, cudaMemcpyHostToDevice, d.stream));
      MG_CU(cudaEventRecord(d.ev_in, d.stream));
      MG_CU(cudaMemsetAsync(d.d_u, 0, W * sizeof(int), d.stream));
      MG_RC(gblur_halo_colsum_device(d.halo_plan, d.d_hin, d.d_u, d.stream));
      MG_CU(cudaMemcpyAsync(d.h_u, d.d_u, W * sizeof(int), cudaMemcpyDeviceToHost, d.stream));
      MG_CU(cudaEventRecord(d.ev_u, d.stream));
      MG_RC(gblur_halo_rowpass_device(d.halo_plan, d.d_hin, d.in0, d.in1, d.stream));
    }
  }
  // 2. the integer column sums of the whole image (exact: the order of the additions does not matter)
  std::memset(m->h_usum, 0, W * sizeof(int));
  for (int g = 0; g < m->n; ++g) {
    DevState& d = m->dev[(size_t)g];
    if (d.out1 <= d.out0) continue;
    MG_CU(cudaSetDevice(d.device));
    MG_CU(cudaEventSynchronize(d.ev_u));
    for (size_t i = 0; i < W; ++i) m->h_usum[i] += d.h_u[i];
  }
  // 3. column terms + column blocks; the output rows overwrite the input image, so no device may still be reading its
  //    halo from the host buffer: every copy back waits for every copy in
  for (int g = 0; g < m->n; ++g) {
    DevState& d = m->dev[(size_t)g];
    if (d.out1 <= d.out0) continue;
    MG_CU(cudaSetDevice(d.device));
    MG_CU(cudaMemcpyAsync(d.d_u, m->h_usum, W * sizeof(int), cudaMemcpyHostToDevice, d.stream));
    MG_RC(gblur_halo_colpass_device(d.halo_plan, d.d_u, d.d_hout, d.stream));
    for (int h = 0; h < m->n; ++h)
      if (h != g && m->dev[(size_t)h].out1 > m->dev[(size_t)h].out0)
        MG_CU(cudaStreamWaitEvent(d.stream, m->dev[(size_t)h].ev_in, 0));
    MG_CU(cudaMemcpyAsync(hwc + (size_t)d.out0 * W, d.d_hout, (size_t)(d.out1 - d.out0) * W, cudaMemcpyDeviceToHost, d.stream));
  }
  for (int g = 0; g < m->n; ++g) {
    MG_CU(cudaSetDevice(m->dev[(size_t)g].device));
    MG_CU(cudaStreamSynchronize(m->dev[(size_t)g].stream));
  }
  return GBLUR_OK;
}

}  // namespace

extern "C" {

const char* gblur_mg_last_error(void) { return g_mg_err.c_str(); }

int gblur_mg_create(gblur_mg** out, int n_gpus, const int* devices, int rows, int cols, int channels, float sigma) {
  if (!out) return mg_fail(GBLUR_E_INVALID, "null out pointer");
  *out = nullptr;
  if (n_gpus < 1 || n_gpus > 8) return mg_fail(GBLUR_E_INVALID, "n_gpus %d outside [1, 8]", n_gpus);
  if (!(sigma > 0.0f)) return mg_fail(GBLUR_E_INVALID, "Invalid smoothing factor");
  if (channels != 1 && channels != 3 && channels != 4) return mg_fail(GBLUR_E_INVALID, "Unsupported number of channels");
  if (rows < n_gpus || cols < n_gpus) return mg_fail(GBLUR_E_INVALID, "image smaller than the number of GPUs");
  int ndev = 0;
  MG_CU(cudaGetDeviceCount(&ndev));
  int prev = 0;
  cudaGetDevice(&prev);
  gblur_mg* m = new gblur_mg();
  m->n = n_gpus;
  m->rows = rows;
  m->cols = cols;
  m->channels = channels;
  m->sigma = sigma;
  m->dev.resize((size_t)n_gpus);
  int rc = GBLUR_OK;
  for (int g = 0; g < n_gpus && rc == GBLUR_OK; ++g) {
    const int dv = devices ? devices[g] : g;
    if (dv < 0 || dv >= ndev) {
      rc = mg_fail(GBLUR_E_CUDA, "device %d not available (%d visible)", dv, ndev);
      break;
    }
    for (int h = 0; h < g; ++h)
      if (m->dev[(size_t)h].device == dv) rc = mg_fail(GBLUR_E_INVALID, "device %d listed twice", dv);
    m->dev[(size_t)g].device = dv;
  }
  // peer access between every pair (the band exchange stores into peer memory)
  m->peer_ok = true;
  for (int g = 0; g < n_gpus && rc == GBLUR_OK; ++g) {
    DevState& d = m->dev[(size_t)g];
    if (cudaSetDevice(d.device) != cudaSuccess) {
      rc = mg_fail(GBLUR_E_CUDA, "cudaSetDevice(%d) failed", d.device);
      break;
    }
    for (int h = 0; h < n_gpus; ++h) {
      if (h == g) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, d.device, m->dev[(size_t)h].device);
      if (!can) {
        m->peer_ok = false;
        continue;
      }
      const cudaError_t e = cudaDeviceEnablePeerAccess(m->dev[(size_t)h].device, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled)
        cudaGetLastError();
      else if (e != cudaSuccess)
        m->peer_ok = false;
    }
    if (cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&d.ev_row, cudaEventDisableTiming) != cudaSuccess)
      rc = mg_fail(GBLUR_E_CUDA, "stream / event creation failed on device %d", d.device);
  }
  cudaSetDevice(prev);
  if (rc != GBLUR_OK) {
    const std::string keep = g_mg_err;
    gblur_mg_destroy(m);
    g_mg_err = keep;
    return rc;
  }
  *out = m;
  return GBLUR_OK;
}

void gblur_mg_destroy(gblur_mg* m) {
  if (!m) return;
  int prev = 0;
  cudaGetDevice(&prev);
  for (DevState& d : m->dev) {
    cudaSetDevice(d.device);
    if (d.stream) cudaStreamSynchronize(d.stream);
    if (d.frames_plan) gblur_plan_destroy(d.frames_plan);
    for (int a = 0; a < 2; ++a) {
      if (d.row_plan[a]) gblur_plan_destroy(d.row_plan[a]);
      if (d.col_plan[a]) gblur_plan_destroy(d.col_plan[a]);  // frees the panel buffer it handed out
    }
    if (d.d_in) cudaFree(d.d_in);
    if (d.d_out) cudaFree(d.d_out);
    if (d.halo_plan) gblur_plan_destroy(d.halo_plan);
    if (d.d_hin) cudaFree(d.d_hin);
    if (d.d_hout) cudaFree(d.d_hout);
    if (d.d_u) cudaFree(d.d_u);
    if (d.h_u) cudaFreeHost(d.h_u);
    if (d.ev_in) cudaEventDestroy(d.ev_in);
    if (d.ev_u) cudaEventDestroy(d.ev_u);
    if (d.ev_row) cudaEventDestroy(d.ev_row);
    if (d.stream) cudaStreamDestroy(d.stream);
  }
  if (m->h_usum) cudaFreeHost(m->h_usum);
  cudaSetDevice(prev);
  delete m;
}

int gblur_mg_info(const gblur_mg* m, int* n_gpus, int* peer_access, int* bands_fused_alpha0, int* bands_fused_alpha1) {
  if (!m) return mg_fail(GBLUR_E_INVALID, "null handle");
  if (n_gpus) *n_gpus = m->n;
  if (peer_access) *peer_access = m->peer_ok ? 1 : 0;
  // 2 = halo mode (row bands for both passes), 1 = row bands + peer-store exchange, 0 = whole image on one device
  for (int a = 0; a < 2; ++a) {
    int* out = a ? bands_fused_alpha1 : bands_fused_alpha0;
    if (!out) continue;
    if (halo_shape_ok(m, a) && m->halo_state >= 0)
      *out = 2;
    else
      *out = (m->n > 1 && m->peer_ok && bands_divisible(m, a)) ? 1 : 0;
  }
  return GBLUR_OK;
}

int gblur_mg_run_frames(gblur_mg* m, uint8_t* hwc_frames, int n_frames, int apply_to_alpha) {
  if (!m || !hwc_frames) return mg_fail(GBLUR_E_INVALID, "null argument");
  if (n_frames < 1) return mg_fail(GBLUR_E_INVALID, "n_frames %d < 1", n_frames);
  std::lock_guard<std::mutex> lock(m->mu);
  const size_t frame_bytes = (size_t)m->rows * m->cols * m->channels;
  const int used = std::min(m->n, n_frames);
  std::vector<int> rc((size_t)used, GBLUR_OK);
  std::vector<std::string> msg((size_t)used);
  std::vector<std::thread> th;
  for (int g = 0; g < used; ++g) {
    th.emplace_back([&, g] {
      DevState& d = m->dev[(size_t)g];
      const Range fr = shard(n_frames, used, g);
      const int cap = std::min(16, fr.size());
      if (!d.frames_plan || d.frames_cap < cap) {
        if (d.frames_plan) gblur_plan_destroy(d.frames_plan);
        d.frames_plan = nullptr;
        rc[(size_t)g] = gblur_plan_create_batch(&d.frames_plan, d.device, m->rows, m->cols, m->channels, m->sigma, cap);
        d.frames_cap = cap;
      }
      if (rc[(size_t)g] == GBLUR_OK)
        rc[(size_t)g] = gblur_run_batch_host(d.frames_plan, hwc_frames + (size_t)fr.lo * frame_bytes, fr.size(), apply_to_alpha);
      if (rc[(size_t)g] != GBLUR_OK) msg[(size_t)g] = gblur_last_error();  // thread-local in the worker: copy it out
    });
  }
  for (auto& t : th) t.join();
  for (int g = 0; g < used; ++g)
    if (rc[(size_t)g] != GBLUR_OK) {
      g_mg_err = msg[(size_t)g];
      return rc[(size_t)g];
    }
  return GBLUR_OK;
}

int gblur_mg_run_bands(gblur_mg* m, uint8_t* hwc, int apply_to_alpha) {
  if (!m || !hwc) return mg_fail(GBLUR_E_INVALID, "null argument");
  std::lock_guard<std::mutex> lock(m->mu);
  const int alpha = apply_to_alpha ? 1 : 0;
  int prev = 0;
  cudaGetDevice(&prev);
  struct Restore {
    int d;
    ~Restore() { cudaSetDevice(d); }
  } restore{prev};
  if (halo_shape_ok(m, alpha)) {
    MG_RC(ensure_halo_state(m));
    if (m->halo_state == 1) return run_bands_halo(m, hwc);
  }
  if (m->n == 1 || !m->peer_ok || !bands_divisible(m, alpha)) {
    // one GPU, no peer access, or a width the panel exchange cannot split evenly: the whole image on the first device
    DevState& d = m->dev[0];
    if (!d.frames_plan) {
      MG_RC(gblur_plan_create_batch(&d.frames_plan, d.device, m->rows, m->cols, m->channels, m->sigma, 1));
      d.frames_cap = 1;
    }
    MG_RC(gblur_run_host(d.frames_plan, hwc, apply_to_alpha));
    return GBLUR_OK;
  }
  MG_RC(ensure_band_state(m, alpha));
  const int C = m->channels;
  const size_t row_bytes = (size_t)m->cols * C;
  std::vector<float*> peers((size_t)m->n);
  for (int g = 0; g < m->n; ++g) peers[(size_t)g] = m->dev[(size_t)g].panel[alpha];
  // 1. every device: its row band host -> device, (alpha pass-through: its column band of the INPUT host -> device as
  //    the output's initial content), row pass with peer stores
  for (int g = 0; g < m->n; ++g) {
    DevState& d = m->dev[(size_t)g];
    const Range rr = shard(m->rows, m->n, g), cr = shard(m->cols, m->n, g);
    MG_CU(cudaSetDevice(d.device));
    MG_CU(cudaMemcpyAsync(d.d_in, hwc + (size_t)rr.lo * row_bytes, (size_t)rr.size() * row_bytes, cudaMemcpyHostToDevice, d.stream));
    if (processed(C, alpha) < C)
      MG_CU(cudaMemcpy2DAsync(d.d_out, (size_t)cr.size() * C, hwc + (size_t)cr.lo * C, row_bytes, (size_t)cr.size() * C,
                              (size_t)m->rows, cudaMemcpyHostToDevice, d.stream));
    MG_RC(gblur_rowpass_peer_device(d.row_plan[alpha], d.d_in, peers.data(), m->n, rr.lo, apply_to_alpha, d.stream));
    MG_CU(cudaEventRecord(d.ev_row, d.stream));
  }
  // 2. every device waits for ALL row passes (its panel is complete), then blurs its columns and returns them
  for (int g = 0; g < m->n; ++g) {
    DevState& d = m->dev[(size_t)g];
    const Range cr = shard(m->cols, m->n, g);
    MG_CU(cudaSetDevice(d.device));
    for (int h = 0; h < m->n; ++h)
      if (h != g) MG_CU(cudaStreamWaitEvent(d.stream, m->dev[(size_t)h].ev_row, 0));
    MG_RC(gblur_colpass_panel_device(d.col_plan[alpha], d.panel[alpha], d.d_out, (long long)cr.size() * C, apply_to_alpha, d.stream));
  }
  // the output overwrites the input image: no device may still be reading its row band from the host buffer — every
  // H2D copy precedes a row pass, and every column pass waits for every row pass, so the D2H copies below are safe
  for (int g = 0; g < m->n; ++g) {
    DevState& d = m->dev[(size_t)g];
    const Range cr = shard(m->cols, m->n, g);
    MG_CU(cudaSetDevice(d.device));
    MG_CU(cudaMemcpy2DAsync(hwc + (size_t)cr.lo * C, row_bytes, d.d_out, (size_t)cr.size() * C, (size_t)cr.size() * C,
                            (size_t)m->rows, cudaMemcpyDeviceToHost, d.stream));
  }
  for (int g = 0; g < m->n; ++g) {
    MG_CU(cudaSetDevice(m->dev[(size_t)g].device));
    MG_CU(cudaStreamSynchronize(m->dev[(size_t)g].stream));
  }
  return GBLUR_OK;
}

}  // extern "C"
