// static_shapes.h — FFT lengths that get a fully compile-time kernel (all radices, strides and
// index arithmetic constant-folded). Every other valid length runs the runtime-radix kernel.
// Columns: id, NSEQ, R0, R1, R2, R3 (1 = stage absent), CTA threads, min CTAs per SM.
// libgaussianblur_b200/build.py compiles static_impl.cu once per id (GBLUR_TU_SHAPE=id).
#pragma once

#define GBLUR_STATIC_SHAPES(X)                                                  \
  X(0, 2, 20, 20, 20, 1, 400, 1) /* N = 8000 : 8K RGBA sigma=20 row pass    */ \
  X(1, 2, 16, 16, 18, 1, 288, 2) /* N = 4608 : 8K RGBA sigma=20 column pass */ \
  X(2, 2, 16, 16, 8, 1, 256, 2)  /* N = 2048 : 1080p row pass               */ \
  X(3, 2, 12, 12, 8, 1, 160, 2)  /* N = 1152 : 1080p column pass            */

#define GBLUR_NUM_STATIC_SHAPES 4
