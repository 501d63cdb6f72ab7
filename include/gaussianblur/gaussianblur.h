// include/gaussianblur/gaussianblur.h — public C++ entry points, B200-native.
//
// Drop-in for the reference header (/root/reference/include/gaussianblur/gaussianblur.h:20,30):
// same namespace, names, argument meaning and error behaviour. The work runs on the GPU through
// the C ABI in include/gblur_capi.h; there is no CPU fallback (a missing device or any CUDA
// failure throws std::runtime_error).
#pragma once

#include <gaussianblur/helpers.hpp>

namespace gaussianblur {

/**
 * Kernel width, reflect pad, FFT lengths (nearest 32*2^a*3^b*5^c size) and the real spectra of the
 * centred Gaussian for this geometry and sigma. Host-only: does not touch the GPU.
 */
KernelDFT prepare_kernel_DFT(const ImgGeom image_geometry, const float sigma);

/**
 * Blur `image` in place (reflect_101 borders). channels must be 3 or 4; with 4 channels the
 * alpha channel is blurred only if apply_to_alpha. sigma <= 0 prints "Invalid smoothing factor"
 * and returns; an unsupported channel count prints "Unsupported number of channels" to stderr
 * and returns; the image is left untouched in both cases (reference: src/gaussianblur.cpp:307-317).
 */
void gaussianblur(Image &image, const float sigma, const bool apply_to_alpha);

// ---- additive entry points (not in the reference) ----

/** Same blur, reusing a prepared kernel/plan across calls (video frames of one geometry). */
void gaussianblur(Image &image, KernelDFT &kernel, const bool apply_to_alpha);

/** n_frames contiguous HWC frames of one geometry blurred in place with ONE kernel spectrum. */
void gaussianblur_batch(uint8_t *frames, const int n_frames, const ImgGeom frame_geometry, const float sigma,
                        const bool apply_to_alpha);

/** Single-plane (1-channel) or HWC buffer owned by the caller, in place; channels in {1,3,4}. */
void gaussianblur_buffer(uint8_t *hwc, const ImgGeom geometry, const float sigma, const bool apply_to_alpha);

/** CUDA device used by the calls above for the calling thread (default 0). */
void set_device(int device);

// Additive: how many GPUs gaussianblur_batch (frames sharded, no exchange) and gaussianblur_buffer on gigapixel images
// (row bands, NVLink peer stores) fan out over — the reference's hybrid_loop fans out over host threads the same way
// (helpers.hpp:134-172). 0 = default: GBLUR_DEVICES from the environment, else every visible device; 1 = one device.
void set_device_count(int n);

}  // namespace gaussianblur
