// gaussianblur_api.cpp — the reference's public C++ API (include/gaussianblur/gaussianblur.h)
// re-implemented as a thin host layer over the C ABI (include/gblur_capi.h).
// Behaviour mirrored from /root/reference/src/gaussianblur.cpp:
//   :307-325 gaussianblur()      sigma<=0 -> "Invalid smoothing factor", return, image untouched
//   :194-197 channel check       not 3/4  -> "Unsupported number of channels" on stderr, untouched
//   :100-172 prepare_kernel_DFT  pad, trailing zeros, spectra (host-only here)
// CUDA failures (impossible in the reference) surface as std::runtime_error; nothing is ever
// computed on the CPU.
#include <gaussianblur/gaussianblur.h>
#include <gblur_capi.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <list>
#include <mutex>
#include <stdexcept>
#include <string>

namespace gaussianblur {
namespace {

thread_local int t_device = 0;

struct PlanKey {
  int device, rows, cols, channels, frames;
  uint32_t sigma_bits;
  bool operator==(const PlanKey &o) const {
    return device == o.device && rows == o.rows && cols == o.cols && channels == o.channels && frames == o.frames &&
           sigma_bits == o.sigma_bits;
  }
};

// Small pool of GPU plans so repeated gaussianblur() calls on one geometry (video) do not rebuild spectra, twiddles
// and workspaces; the reference recomputes its kernel on every call (:319). A plan owns one workspace, one staging
// buffer and one stream, so it serves ONE call at a time: a caller checks a plan out exclusively (Lease) and a second
// thread blurring the same geometry at the same moment gets a second plan — the reference call is re-entrant
// (no globals, PFFFT_Setup read-only: include/pffft_pommier/pffft.h:100-103) and so is this one.
class PlanCache {
 public:
  struct Entry {
    PlanKey key;
    GblurPlanHandle plan;
    bool busy;
  };
  class Lease {
   public:
    Lease(PlanCache *c, GblurPlanHandle p) : cache_(c), plan_(std::move(p)) {}
    Lease(const Lease &) = delete;
    Lease &operator=(const Lease &) = delete;
    ~Lease() { cache_->release(plan_.get()); }
    gblur_plan *get() const { return plan_.get(); }
    const GblurPlanHandle &handle() const { return plan_; }

   private:
    PlanCache *cache_;
    GblurPlanHandle plan_;
  };

  GblurPlanHandle checkout(const PlanKey &key) {
    {
      std::lock_guard<std::mutex> lock(mu_);
      for (auto it = items_.begin(); it != items_.end(); ++it) {
        if (it->key == key && !it->busy) {
          it->busy = true;
          items_.splice(items_.begin(), items_, it);
          return items_.front().plan;
        }
      }
    }
    // plan creation (device allocations, table uploads) happens outside the lock
    float sigma;
    std::memcpy(&sigma, &key.sigma_bits, sizeof sigma);
    gblur_plan *raw = nullptr;
    const int rc = gblur_plan_create_batch(&raw, key.device, key.rows, key.cols, key.channels, sigma, key.frames);
    if (rc != GBLUR_OK) throw std::runtime_error(std::string("gaussianblur: ") + gblur_last_error());
    GblurPlanHandle h(raw, [](gblur_plan *p) { gblur_plan_destroy(p); });
    std::lock_guard<std::mutex> lock(mu_);
    items_.push_front(Entry{key, h, true});
    trim();
    return h;
  }

  void release(const gblur_plan *p) {
    std::lock_guard<std::mutex> lock(mu_);
    for (auto &e : items_)
      if (e.plan.get() == p) e.busy = false;
    trim();
  }

 private:
  void trim() {  // drop idle plans from the cold end; busy ones stay until they are released
    size_t idle = 0;
    for (auto &e : items_) idle += e.busy ? 0 : 1;
    for (auto it = items_.end(); idle > kCapacity && it != items_.begin();) {
      --it;
      if (!it->busy) {
        it = items_.erase(it);
        --idle;
      }
    }
  }
  static constexpr size_t kCapacity = 4;
  std::mutex mu_;
  std::list<Entry> items_;
};

PlanCache &cache() {
  static PlanCache c;
  return c;
}

PlanKey make_key(const ImgGeom &g, float sigma, int frames) {
  PlanKey k{t_device, g.rows, g.cols, g.channels, frames, 0};
  std::memcpy(&k.sigma_bits, &sigma, sizeof sigma);
  return k;
}

void check(int rc) {
  if (rc != GBLUR_OK) throw std::runtime_error(std::string("gaussianblur: ") + gblur_last_error());
}

bool valid_request(const ImgGeom &g, float sigma, bool allow_single_channel) {
  if (sigma <= 0) {
    printf("Invalid smoothing factor\n");
    return false;
  }
  const bool ok = g.channels == 3 || g.channels == 4 || (allow_single_channel && g.channels == 1);
  if (!ok) {
    std::cerr << "Unsupported number of channels" << std::endl;
    return false;
  }
  return true;
}

// ---- several GPUs (gblur_mg_*, csrc/multi_gpu.cpp) ----
// device count the batch / big-image entry points fan out over: GBLUR_DEVICES, else every visible device
int mg_device_count() {
  static const int n = [] {
    if (const char *e = getenv("GBLUR_DEVICES")) return std::max(1, atoi(e));
    const int v = gblur_device_count();
    return v > 8 ? 8 : (v < 1 ? 1 : v);
  }();
  return n;
}
int g_mg_override = 0;  // set_device_count()

struct MgSlot {
  std::mutex mu;
  gblur_mg *mg = nullptr;
  int n = 0;
  ImgGeom geom{0, 0, 0};
  uint32_t sigma_bits = 0;
  ~MgSlot() {
    if (mg) gblur_mg_destroy(mg);
  }
};
MgSlot &mg_slot() {
  static MgSlot s;
  return s;
}
// the process-wide multi-GPU handle for this geometry (rebuilt when geometry, sigma or device count change);
// returns with slot.mu held by `lock`
gblur_mg *mg_acquire(const ImgGeom &g, float sigma, int n, std::unique_lock<std::mutex> &lock) {
  MgSlot &s = mg_slot();
  lock = std::unique_lock<std::mutex>(s.mu);
  uint32_t bits;
  std::memcpy(&bits, &sigma, sizeof bits);
  if (s.mg && (s.n != n || s.geom.rows != g.rows || s.geom.cols != g.cols || s.geom.channels != g.channels || s.sigma_bits != bits)) {
    gblur_mg_destroy(s.mg);
    s.mg = nullptr;
  }
  if (!s.mg) {
    if (gblur_mg_create(&s.mg, n, nullptr, g.rows, g.cols, g.channels, sigma) != GBLUR_OK)
      throw std::runtime_error(std::string("gaussianblur: ") + gblur_mg_last_error());
    s.n = n;
    s.geom = g;
    s.sigma_bits = bits;
  }
  return s.mg;
}

}  // namespace

void set_device(int device) { t_device = device; }
void set_device_count(int n) { g_mg_override = n < 1 ? 0 : (n > 8 ? 8 : n); }

KernelDFT prepare_kernel_DFT(const ImgGeom image_geometry, const float sigma) {
  KernelDFT k{};
  k.geom = image_geometry;
  k.sigma = sigma;
  int ksize = 0, n_rp = 0, n_cp = 0;
  gblur_host_plan_sizes(image_geometry.rows, image_geometry.cols, sigma, &ksize, &k.pad, &n_rp, &n_cp,
                        &k.trailing_zeros.rows, &k.trailing_zeros.cols);
  k.kerf_1D_col.resize((size_t)n_rp);
  k.kerf_1D_row.resize((size_t)n_cp);
  check(gblur_host_spectrum(sigma, ksize, n_rp, k.kerf_1D_col.data()));
  if (n_cp == n_rp)
    k.kerf_1D_row = k.kerf_1D_col;
  else
    check(gblur_host_spectrum(sigma, ksize, n_cp, k.kerf_1D_row.data()));
  return k;
}

void gaussianblur_buffer(uint8_t *hwc, const ImgGeom geometry, const float sigma, const bool apply_to_alpha) {
  if (!valid_request(geometry, sigma, true)) return;
  // an image of a gigapixel-channel or more goes over every device as row bands (gblur_mg_run_bands falls back to one
  // device by itself when the width does not split)
  const int n = g_mg_override ? g_mg_override : mg_device_count();
  if (n > 1 && (size_t)geometry.rows * geometry.cols * geometry.channels >= ((size_t)1 << 30)) {
    std::unique_lock<std::mutex> lock;
    gblur_mg *mg = mg_acquire(geometry, sigma, n, lock);
    if (gblur_mg_run_bands(mg, hwc, apply_to_alpha ? 1 : 0) != GBLUR_OK)
      throw std::runtime_error(std::string("gaussianblur: ") + gblur_mg_last_error());
    return;
  }
  PlanCache::Lease plan(&cache(), cache().checkout(make_key(geometry, sigma, 1)));
  check(gblur_run_host(plan.get(), hwc, apply_to_alpha ? 1 : 0));
}

void gaussianblur(Image &image, const float sigma, const bool apply_to_alpha) {
  if (!valid_request(image.geom, sigma, false)) return;
  const size_t need = (size_t)image.geom.rows * image.geom.cols * image.geom.channels;
  if (image.data.size() < need) throw std::runtime_error("gaussianblur: image.data smaller than rows*cols*channels");
  PlanCache::Lease plan(&cache(), cache().checkout(make_key(image.geom, sigma, 1)));
  check(gblur_run_host(plan.get(), image.data.data(), apply_to_alpha ? 1 : 0));
}

void gaussianblur(Image &image, KernelDFT &kernel, const bool apply_to_alpha) {
  if (!valid_request(image.geom, kernel.sigma, false)) return;
  if (image.geom.rows != kernel.geom.rows || image.geom.cols != kernel.geom.cols)
    throw std::runtime_error("gaussianblur: KernelDFT was prepared for another geometry");
  const size_t need = (size_t)image.geom.rows * image.geom.cols * image.geom.channels;
  if (image.data.size() < need) throw std::runtime_error("gaussianblur: image.data smaller than rows*cols*channels");
  // The reference's KernelDFT does not depend on the channel count (src/gaussianblur.cpp:100-172), so one handle may
  // legitimately serve 3- and 4-channel images, from any thread and device: the spectra live in the handle, the GPU
  // plan behind it is checked out per call by (device, geometry, channels, sigma). The handle only remembers the first
  // plan (rows_setup / cols_setup become non-null like the reference's setups).
  PlanCache::Lease plan(&cache(), cache().checkout(make_key(image.geom, kernel.sigma, 1)));
  if (!kernel.rows_setup) {
    kernel.rows_setup = plan.handle();
    kernel.cols_setup = kernel.rows_setup;
  }
  check(gblur_run_host(plan.get(), image.data.data(), apply_to_alpha ? 1 : 0));
}

void gaussianblur_batch(uint8_t *frames, const int n_frames, const ImgGeom frame_geometry, const float sigma,
                        const bool apply_to_alpha) {
  if (n_frames <= 0) return;
  if (!valid_request(frame_geometry, sigma, true)) return;
  const int n = g_mg_override ? g_mg_override : mg_device_count();
  if (n > 1 && n_frames >= 2 * n) {  // frames shard over the devices, no exchange (BASELINE.json config 3)
    std::unique_lock<std::mutex> lock;
    gblur_mg *mg = mg_acquire(frame_geometry, sigma, n, lock);
    if (gblur_mg_run_frames(mg, frames, n_frames, apply_to_alpha ? 1 : 0) != GBLUR_OK)
      throw std::runtime_error(std::string("gaussianblur: ") + gblur_mg_last_error());
    return;
  }
  // the plan holds workspace for 16 frames; longer batches are streamed through it (H2D || blur || D2H)
  PlanCache::Lease plan(&cache(), cache().checkout(make_key(frame_geometry, sigma, n_frames < 16 ? n_frames : 16)));
  check(gblur_run_batch_host(plan.get(), frames, n_frames, apply_to_alpha ? 1 : 0));
}

}  // namespace gaussianblur
