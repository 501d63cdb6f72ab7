// mg_check.cpp — C++-only check of the multi-GPU C ABI (include/gblur_capi.h: gblur_mg_*), no Python, no torch, no
// NCCL: what a C++ user of the reference's library would link against. Test infrastructure: the checker is the
// compiled reference (oracle/_ref/libgblur_ref.so, dlopen'ed when argv[2] names it).
//   mg_check <n_gpus> [path/to/libgblur_ref.so] [plane_side]
//   1. frames: 2 n + 3 frames of 540 x 960 RGB, sigma 4, sharded over the GPUs == the same frames blurred one by one on
//      one plan (bit-identical), and frame 0 within 1 LSB of the reference;
//   2. bands : one plane_side^2 (default 16384) single-channel image, sigma 50, row bands with NVLink peer stores, within
//      1 LSB of the reference's two process_channel_tiles calls; RGBA 2048 x 4096 sigma 9 with alpha passed through.
#include <dlfcn.h>

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "gblur_capi.h"

static uint32_t fmix32(uint32_t h) {
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
static void fill(std::vector<uint8_t>& v, uint32_t seed) {
  for (size_t i = 0; i < v.size(); ++i) v[i] = (uint8_t)(fmix32((uint32_t)i ^ ((uint32_t)(i >> 32) * 0x9E3779B9u) ^ seed) >> 24);
}
static int max_diff(const uint8_t* a, const uint8_t* b, size_t n, size_t* ndiff) {
  int mx = 0;
  size_t nd = 0;
  for (size_t i = 0; i < n; ++i) {
    const int d = std::abs((int)a[i] - (int)b[i]);
    if (d) ++nd;
    if (d > mx) mx = d;
  }
  if (ndiff) *ndiff = nd;
  return mx;
}
#define MG(x) do { int rc_ = (x); if (rc_ != 0) { std::printf("FAIL %s -> %d: %s | %s\n", #x, rc_, gblur_mg_last_error(), gblur_last_error()); return 1; } } while (0)

typedef double (*ref_blur_fn)(uint8_t*, int, int, int, float, int);
typedef double (*ref_plane_fn)(uint8_t*, int, int, float);

int main(int argc, char** argv) {
  std::setvbuf(stdout, nullptr, _IOLBF, 0);
  const int n = argc > 1 ? std::atoi(argv[1]) : 2;
  ref_blur_fn ref_blur = nullptr;
  ref_plane_fn ref_plane = nullptr;
  if (argc > 2 && argv[2][0]) {
    // RTLD_DEEPBIND: the reference library and this program's libgblur_b200.so both define gaussianblur::gaussianblur
    // (the drop-in keeps the reference's symbol names); the checker must bind to its own copy
    void* h = dlopen(argv[2], RTLD_NOW | RTLD_LOCAL | RTLD_DEEPBIND);
    if (h) {
      ref_blur = (ref_blur_fn)dlsym(h, "ref_gaussianblur");
      ref_plane = (ref_plane_fn)dlsym(h, "ref_blur_plane_u8");
    }
  }
  const int side = argc > 3 ? std::atoi(argv[3]) : 16384;
  std::printf("devices visible: %d, using %d, reference checker: %s\n", gblur_device_count(), n, ref_blur ? "yes" : "no");

  {  // ---- frames ----
    const int rows = 540, cols = 960, C = 3, nf = 2 * n + 3;
    const size_t fb = (size_t)rows * cols * C;
    std::vector<uint8_t> in(fb * nf), a, b;
    fill(in, 0x1234);
    a = in;
    b = in;
    gblur_mg* mg = nullptr;
    MG(gblur_mg_create(&mg, n, nullptr, rows, cols, C, 4.0f));
    MG(gblur_mg_run_frames(mg, a.data(), nf, 0));
    gblur_mg_destroy(mg);
    gblur_plan* p = nullptr;
    MG(gblur_plan_create(&p, 0, rows, cols, C, 4.0f));
    for (int f = 0; f < nf; ++f) MG(gblur_run_host(p, b.data() + f * fb, 0));
    gblur_plan_destroy(p);
    if (std::memcmp(a.data(), b.data(), a.size()) != 0) {
      std::printf("FAIL frames: sharded result differs from the frame-by-frame result\n");
      return 1;
    }
    if (ref_blur) {
      std::vector<uint8_t> r(in.begin(), in.begin() + fb);
      ref_blur(r.data(), rows, cols, C, 4.0f, 0);
      size_t nd;
      const int mx = max_diff(r.data(), a.data(), fb, &nd);
      std::printf("frames: %d frames over %d GPUs, frame 0 vs reference max |delta| = %d (%zu bytes differ)\n", nf, n, mx, nd);
      if (mx > 1) return 1;
    } else {
      std::printf("frames: %d frames over %d GPUs identical to one-plan results\n", nf, n);
    }
  }
  for (int mode = 0; mode < 2; ++mode) {  // ---- bands, one channel: halo mode, then the peer-store exchange ----
    if (mode == 1) setenv("GBLUR_MG_BANDS", "exchange", 1);
    const int rows = side, cols = side;
    std::vector<uint8_t> img((size_t)rows * cols), want;
    fill(img, 0x77);
    want = img;
    gblur_mg* mg = nullptr;
    MG(gblur_mg_create(&mg, n, nullptr, rows, cols, 1, 50.0f));
    int ng = 0, peer = 0, sh0 = 0, sh1 = 0;
    MG(gblur_mg_run_bands(mg, img.data(), 0));  // first call builds the plans
    MG(gblur_mg_info(mg, &ng, &peer, &sh0, &sh1));
    std::vector<uint8_t> again = want;
    const auto t0 = std::chrono::steady_clock::now();
    MG(gblur_mg_run_bands(mg, again.data(), 0));
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    gblur_mg_destroy(mg);
    unsetenv("GBLUR_MG_BANDS");
    if (std::memcmp(img.data(), again.data(), img.size()) != 0) {
      std::printf("FAIL bands: two runs differ\n");
      return 1;
    }
    std::printf("bands: %d x %d x 1 sigma 50 on %d GPUs (peer access %d, sharding %d: 2 = halo, 1 = exchange, 0 = one GPU): %.1f ms host to host\n",
                rows, cols, ng, peer, sh0, secs * 1e3);
    if (ng > 1 && sh0 != (mode == 0 ? 2 : (peer ? 1 : 0))) {
      std::printf("FAIL bands: expected the %s path\n", mode == 0 ? "halo" : "exchange");
      return 1;
    }
    if (ref_plane) {
      ref_plane(want.data(), rows, cols, 50.0f);
      size_t nd;
      const int mx = max_diff(want.data(), img.data(), img.size(), &nd);
      std::printf("bands: vs reference max |delta| = %d (%zu of %zu bytes differ)\n", mx, nd, img.size());
      if (mx > 1) return 1;
    }
  }
  {  // ---- bands, RGBA with all four channels blurred: halo mode with pair-major panels ----
    const int rows = 6000, cols = 2064, C = 4;
    std::vector<uint8_t> img((size_t)rows * cols * C), want;
    fill(img, 0x55);
    want = img;
    gblur_mg* mg = nullptr;
    MG(gblur_mg_create(&mg, n, nullptr, rows, cols, C, 20.0f));
    MG(gblur_mg_run_bands(mg, img.data(), 1));
    int ng = 0, peer = 0, sh0 = 0, sh1 = 0;
    MG(gblur_mg_info(mg, &ng, &peer, &sh0, &sh1));
    gblur_mg_destroy(mg);
    if (ref_blur) {
      ref_blur(want.data(), rows, cols, C, 20.0f, 1);
      size_t nd;
      const int mx = max_diff(want.data(), img.data(), img.size(), &nd);
      std::printf("bands RGBA, alpha blurred (sharding %d): vs reference max |delta| = %d (%zu bytes differ)\n", sh1, mx, nd);
      if (mx > 1) return 1;
    }
  }
  {  // ---- bands, RGBA with alpha passed through ----
    const int rows = 2048, cols = 4096, C = 4;
    std::vector<uint8_t> img((size_t)rows * cols * C), want;
    fill(img, 0x99);
    want = img;
    gblur_mg* mg = nullptr;
    MG(gblur_mg_create(&mg, n, nullptr, rows, cols, C, 9.0f));
    MG(gblur_mg_run_bands(mg, img.data(), 0));
    gblur_mg_destroy(mg);
    for (size_t i = 3; i < img.size(); i += 4)
      if (img[i] != want[i]) {
        std::printf("FAIL bands RGBA: alpha byte changed\n");
        return 1;
      }
    if (ref_blur) {
      ref_blur(want.data(), rows, cols, C, 9.0f, 0);
      size_t nd;
      const int mx = max_diff(want.data(), img.data(), img.size(), &nd);
      std::printf("bands RGBA: vs reference max |delta| = %d (%zu bytes differ)\n", mx, nd);
      if (mx > 1) return 1;
    }
  }
  std::printf("mg check ok\n");
  return 0;
}
