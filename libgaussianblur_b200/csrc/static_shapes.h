// static_shapes.h — FFT lengths that get a fully compile-time kernel (all radices, strides and
// index arithmetic constant-folded). Every other valid length runs the runtime-radix kernel.
// Columns: id, NSEQ, R0, R1, R2, R3 (1 = stage absent), CTA threads, register cap per thread.
// libgaussianblur_b200/build.py compiles static_impl.cu once per id (GBLUR_TU_SHAPE=id).
//
// Register caps follow the per-scheduler register file (16384 registers per SM sub-partition):
// cap = floor(16384 / ceil(resident_warps / 4) / 32) rounded down to a multiple of 8.
// Several entries may share one FFT length: the first one listed is the default, the others are
// tuning variants selectable with the environment variable GBLUR_VARIANT_<N>=<id>.
#pragma once

#define GBLUR_STATIC_SHAPES(X)                                                                     \
  X(0, 2, 20, 20, 20, 1, 400, 128)  /* N = 8000 : 8K RGBA sigma=20 row pass, 1 CTA/SM (13 warps) */ \
  X(1, 2, 16, 16, 18, 1, 288, 96)   /* N = 4608 : 8K RGBA sigma=20 column pass, 2 CTAs/SM        */ \
  X(12, 2, 8, 16, 16, 1, 128, 128)  /* N = 2048 : 1080p row pass, one radix-16 butterfly per     */ \
                                    /*            thread, 4 CTAs/SM (44.5 vs 35.1 Gpx/s with 2+3) */ \
  X(11, 2, 12, 12, 8, 1, 96, 128)   /* N = 1152 : 1080p column pass, one radix-12 per thread     */ \
  X(2, 2, 16, 16, 8, 1, 256, 128)   /* N = 2048 : variant, half of the threads idle in 4 passes  */ \
  X(3, 2, 12, 12, 8, 1, 160, 128)   /* N = 1152 : variant                                        */ \
  X(7, 2, 16, 16, 18, 1, 288, 168)  /* N = 4608 : variant, 1 CTA/SM, no register pressure        */
