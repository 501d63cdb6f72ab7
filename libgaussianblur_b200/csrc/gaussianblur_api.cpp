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

#include <cstdio>
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

// Small LRU of GPU plans so repeated gaussianblur() calls on one geometry (video) do not rebuild
// spectra, twiddles and workspaces; the reference recomputes its kernel on every call (:319).
class PlanCache {
 public:
  GblurPlanHandle get(const PlanKey &key) {
    std::lock_guard<std::mutex> lock(mu_);
    for (auto it = items_.begin(); it != items_.end(); ++it) {
      if (it->first == key) {
        items_.splice(items_.begin(), items_, it);
        return items_.front().second;
      }
    }
    float sigma;
    std::memcpy(&sigma, &key.sigma_bits, sizeof sigma);
    gblur_plan *raw = nullptr;
    const int rc = gblur_plan_create_batch(&raw, key.device, key.rows, key.cols, key.channels, sigma, key.frames);
    if (rc != GBLUR_OK) throw std::runtime_error(std::string("gaussianblur: ") + gblur_last_error());
    GblurPlanHandle h(raw, [](gblur_plan *p) { gblur_plan_destroy(p); });
    items_.emplace_front(key, h);
    while (items_.size() > kCapacity) items_.pop_back();
    return h;
  }

 private:
  static constexpr size_t kCapacity = 4;
  std::mutex mu_;
  std::list<std::pair<PlanKey, GblurPlanHandle>> items_;
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

}  // namespace

void set_device(int device) { t_device = device; }

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
  GblurPlanHandle plan = cache().get(make_key(geometry, sigma, 1));
  check(gblur_run_host(plan.get(), hwc, apply_to_alpha ? 1 : 0));
}

void gaussianblur(Image &image, const float sigma, const bool apply_to_alpha) {
  if (!valid_request(image.geom, sigma, false)) return;
  const size_t need = (size_t)image.geom.rows * image.geom.cols * image.geom.channels;
  if (image.data.size() < need) throw std::runtime_error("gaussianblur: image.data smaller than rows*cols*channels");
  GblurPlanHandle plan = cache().get(make_key(image.geom, sigma, 1));
  check(gblur_run_host(plan.get(), image.data.data(), apply_to_alpha ? 1 : 0));
}

void gaussianblur(Image &image, KernelDFT &kernel, const bool apply_to_alpha) {
  if (!valid_request(image.geom, kernel.sigma, false)) return;
  if (image.geom.rows != kernel.geom.rows || image.geom.cols != kernel.geom.cols)
    throw std::runtime_error("gaussianblur: KernelDFT was prepared for another geometry");
  const size_t need = (size_t)image.geom.rows * image.geom.cols * image.geom.channels;
  if (image.data.size() < need) throw std::runtime_error("gaussianblur: image.data smaller than rows*cols*channels");
  if (!kernel.rows_setup) {
    kernel.rows_setup = cache().get(make_key(image.geom, kernel.sigma, 1));
    kernel.cols_setup = kernel.rows_setup;
  }
  check(gblur_run_host(kernel.rows_setup.get(), image.data.data(), apply_to_alpha ? 1 : 0));
}

void gaussianblur_batch(uint8_t *frames, const int n_frames, const ImgGeom frame_geometry, const float sigma,
                        const bool apply_to_alpha) {
  if (n_frames <= 0) return;
  if (!valid_request(frame_geometry, sigma, true)) return;
  // the plan holds workspace for 16 frames; longer batches are streamed through it (H2D || blur || D2H)
  GblurPlanHandle plan = cache().get(make_key(frame_geometry, sigma, n_frames < 16 ? n_frames : 16));
  check(gblur_run_batch_host(plan.get(), frames, n_frames, apply_to_alpha ? 1 : 0));
}

}  // namespace gaussianblur
