// api_intents.cpp — the intents of the reference's own gtests (/root/reference/tests/test_gaussianblur.cpp),
// re-asserted against THIS repository's drop-in C++ API (include/gaussianblur/gaussianblur.h -> libgblur_b200.so):
//   PrepareKernelDFT (:12-28), BasicTestRGB (:31-59), InvalidChannelCount (:62-82),
//   BasicTestRGBAWithoutAlpha (:85-119), BasicTestRGBAWithAlpha (:167-205), StressTest (:123-164, < 10 s).
// Plain asserts, no test framework; exit code 0 = all passed. Needs a CUDA device (no CPU fallback exists).
#include <gaussianblur/gaussianblur.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <functional>
#include <thread>
#include <vector>

#define CHECK(cond)                                                          \
  do {                                                                       \
    if (!(cond)) {                                                           \
      std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
      std::exit(1);                                                          \
    }                                                                        \
  } while (0)

static double spread(const std::vector<uint8_t> &v, int channels, int first_n_channels) {
  // variance of the bytes of the first `first_n_channels` channels of every pixel
  double sum = 0, sq = 0;
  size_t n = 0;
  for (size_t i = 0; i < v.size(); ++i)
    if ((int)(i % channels) < first_n_channels) {
      sum += v[i];
      sq += double(v[i]) * v[i];
      ++n;
    }
  const double mean = sum / n;
  return sq / n - mean * mean;
}

static Image ramp_image(int rows, int cols, int channels, uint8_t alpha) {
  Image img;
  img.geom = ImgGeom{rows, cols, channels};
  img.data.resize((size_t)rows * cols * channels);
  uint32_t state = 12345u;
  for (size_t i = 0; i < img.data.size(); ++i) {
    state = state * 1664525u + 1013904223u;
    img.data[i] = (channels == 4 && i % 4 == 3) ? alpha : (uint8_t)(state >> 24);
  }
  return img;
}

int main() {
  {  // kernel spectra of a square image: both axes non-empty, same length, same values
    KernelDFT k = gaussianblur::prepare_kernel_DFT(ImgGeom{4, 4, 3}, 2.0f);
    CHECK(!k.kerf_1D_row.empty() && !k.kerf_1D_col.empty());
    CHECK(k.kerf_1D_row.size() == k.kerf_1D_col.size());
    for (size_t i = 0; i < k.kerf_1D_row.size(); ++i) CHECK(k.kerf_1D_row[i] == k.kerf_1D_col[i]);
    CHECK(k.kerf_1D_row.size() == 32 && k.pad == 2);  // N = 32, pad = 2 [probed]
  }
  {  // RGB: blurring lowers the variance
    Image img = ramp_image(3, 3, 3, 0);
    const double before = spread(img.data, 3, 3);
    gaussianblur::gaussianblur(img, 3.0f, false);
    CHECK(spread(img.data, 3, 3) < before);
  }
  {  // two channels: message on stderr, data untouched
    Image img;
    img.geom = ImgGeom{3, 3, 2};
    img.data.assign(18, 0);
    std::iota(img.data.begin(), img.data.end(), (uint8_t)1);
    const std::vector<uint8_t> copy = img.data;
    gaussianblur::gaussianblur(img, 3.0f, false);
    CHECK(img.data == copy);
    gaussianblur::gaussianblur(img, -1.0f, false);  // "Invalid smoothing factor", untouched
    CHECK(img.data == copy);
  }
  {  // RGBA, alpha not blurred: variance of RGB drops, every alpha byte still 128
    Image img = ramp_image(32, 32, 4, 128);
    for (size_t i = 3; i < img.data.size(); i += 8) img.data[i] = 128;
    const double before = spread(img.data, 4, 3);
    gaussianblur::gaussianblur(img, 3.0f, false);
    CHECK(spread(img.data, 4, 3) < before);
    for (size_t i = 3; i < img.data.size(); i += 4) CHECK(img.data[i] == 128);
  }
  {  // RGBA, alpha blurred too: some alpha byte changes
    Image img = ramp_image(32, 32, 4, 128);
    for (size_t i = 3; i < img.data.size(); i += 8) img.data[i] = 7;  // alpha alternates 128 / 7
    const double before = spread(img.data, 4, 4);
    gaussianblur::gaussianblur(img, 3.0f, true);
    CHECK(spread(img.data, 4, 4) < before);
    bool changed = false;
    for (size_t i = 3; i < img.data.size(); i += 4) changed = changed || (img.data[i] != 128 && img.data[i] != 7);
    CHECK(changed);
  }
  {  // the reference's stress shape: 10240 x 10240 RGBA, sigma = 5, alpha blurred, under 10 s
    Image img = ramp_image(10240, 10240, 4, 200);
    gaussianblur::gaussianblur(img, 5.0f, true);  // first call builds the plan and the CUDA context
    const auto t0 = std::chrono::steady_clock::now();
    gaussianblur::gaussianblur(img, 5.0f, true);
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("stress 10240x10240 RGBA sigma=5: %.3f s\n", secs);
    CHECK(secs < 10.0);
    // reuse through a prepared kernel (additive API)
    KernelDFT k = gaussianblur::prepare_kernel_DFT(img.geom, 5.0f);
    gaussianblur::gaussianblur(img, k, true);
    CHECK(k.rows_setup != nullptr);
  }
  {  // re-entrancy: two threads blur DIFFERENT images of the same geometry at the same time, several times over; every
     // result must equal the single-threaded result of the same input (the reference call is re-entrant)
    auto make = [](int seed) {
      Image img = ramp_image(1080, 1920, 4, 99);
      uint32_t h = (uint32_t)seed * 2654435761u;
      for (auto &b : img.data) {
        h = h * 1664525u + 1013904223u;
        b = (uint8_t)(h >> 24);
      }
      return img;
    };
    const Image a0 = make(1), b0 = make(2);
    Image wa = a0, wb = b0;
    gaussianblur::gaussianblur(wa, 7.0f, true);
    gaussianblur::gaussianblur(wb, 7.0f, true);
    bool same = true;
    auto worker = [&](const Image &src, const Image &want) {
      for (int it = 0; it < 6; ++it) {
        Image img = src;
        gaussianblur::gaussianblur(img, 7.0f, true);
        if (img.data != want.data) same = false;
      }
    };
    std::thread ta(worker, std::cref(a0), std::cref(wa)), tb(worker, std::cref(b0), std::cref(wb));
    ta.join();
    tb.join();
    CHECK(same);
    // one KernelDFT serving a 4-channel and then a 3-channel image (the reference's KernelDFT ignores channels)
    KernelDFT k = gaussianblur::prepare_kernel_DFT(a0.geom, 7.0f);
    Image c4 = a0;
    gaussianblur::gaussianblur(c4, k, true);
    CHECK(c4.data == wa.data);
    Image c3 = ramp_image(1080, 1920, 3, 0), c3b = c3;
    gaussianblur::gaussianblur(c3, k, false);
    gaussianblur::gaussianblur(c3b, 7.0f, false);
    CHECK(c3.data == c3b.data);
  }
  std::printf("api intents ok\n");
  return 0;
}
