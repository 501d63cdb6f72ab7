// problems.cuh — where a line unit's samples come from and go to.
//
// Row pass   : u8 interleaved image -> fp32 intermediate. Fuses the reference's
//              deinterleave_channels (helpers.hpp:218-246), the reflect_101 + zero-tail copy
//              (src/gaussianblur.cpp:217-228) and the crop (:237-238) into the FFT kernel's
//              first load / last store.
// Column pass: fp32 intermediate -> u8 interleaved image. The intermediate is read down its
//              columns directly (no flip_block transpose, helpers.hpp:187-216) and the epilogue is
//              interleave_channels' uint8_t(v + 0.5f) (helpers.hpp:251-252,270) with a clamp.
// All citations are under /root/reference/.
#pragma once
#include <cstdint>

#include "fftconv.cuh"

namespace gblur {

// reflect_101 source index for padded coordinate n; returns -1 for the zero tail
__device__ __forceinline__ int reflect101(int n, int pad, int T) {
  int t = n - pad;
  if (t < 0) t = -t;
  if (t >= T) t = 2 * T - 2 - t;
  return (n < T + 2 * pad) ? t : -1;
}

// u8 -> f32 without the conversion pipe: 0x4B000000 | b is the float 8388608 + b
__device__ __forceinline__ float u8_to_f32(uint32_t word, int byte) {
  return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7540u | byte)) - 8388608.0f;
}

// uint8_t(v + 0.5f): fp32 add, truncate; clamped to [0,255] (NaN -> 0)
__device__ __forceinline__ uint32_t f32_to_u8(float v) {
  float f = fminf(fmaxf(v + 0.5f, 0.0f), 255.0f);
  return __float2uint_rz(f);
}

// --------------------------------------------------------------------------------------------
// Row pass
// --------------------------------------------------------------------------------------------
struct RowParams {
  const uint8_t* in;      // HWC u8, frames stacked as extra rows
  long long in_pitch;     // bytes per image row
  int C, Cp;              // channels in memory / channels processed (the first Cp of each pixel)
  int T, pad;             // line length (cols) and reflect pad
  long long nreal;        // total real lines = rows_total * Cp
  float* inter;           // [rows_total][inter_pitch] fp32, element (r, x*Cp + ch)
  long long inter_pitch;  // floats per row
  int nunits;
};

// Any C, Cp: real line sigma = r*Cp + ch; a unit takes 2*NSEQ consecutive lines.
template <int NSEQ>
struct RowProblemGeneric {
  static constexpr int L = 2 * NSEQ;
  struct Loader {
    const uint8_t* p[L];
    int C, pad, T;
    __device__ __forceinline__ Pt<NSEQ> load(int n) const {
      const int x = reflect101(n, pad, T);
      Pt<NSEQ> r;
      const long long off = (long long)(x < 0 ? 0 : x) * C;
#pragma unroll
      for (int q = 0; q < NSEQ; ++q) {
        float a = (float)__ldg(p[2 * q] + off);
        float b = (float)__ldg(p[2 * q + 1] + off);
        r.v[q] = (x < 0) ? make_float2(0.f, 0.f) : make_float2(a, b);
      }
      return r;
    }
  };
  struct Storer {
    float* q[L];  // nullptr: line does not exist (tail unit)
    int Cp, pad, T;
    __device__ __forceinline__ void store(int n, const Pt<NSEQ>& v) const {
      const int x = n - pad;
      if (x < 0 || x >= T) return;
      const long long off = (long long)x * Cp;
#pragma unroll
      for (int s = 0; s < NSEQ; ++s) {
        if (q[2 * s]) q[2 * s][off] = v.v[s].x;
        if (q[2 * s + 1]) q[2 * s + 1][off] = v.v[s].y;
      }
    }
  };
  static __device__ __forceinline__ void make(const RowParams& P, int unit, Loader& ld, Storer& st) {
    ld.C = P.C;
    ld.pad = st.pad = P.pad;
    ld.T = st.T = P.T;
    st.Cp = P.Cp;
#pragma unroll
    for (int k = 0; k < L; ++k) {
      long long sig = (long long)unit * L + k;
      const bool ok = sig < P.nreal;
      if (!ok) sig = (long long)unit * L;  // read a real line, never store it
      const long long r = sig / P.Cp;
      const int ch = (int)(sig - r * P.Cp);
      ld.p[k] = P.in + r * P.in_pitch + ch;
      st.q[k] = ok ? P.inter + r * P.inter_pitch + ch : nullptr;
    }
  }
};

// C == Cp == 4, two sequences: one uchar4 load feeds (R + iG, B + iA), one float4 store per pixel
struct RowProblemRGBA {
  struct Loader {
    const uint32_t* row;
    int pad, T;
    __device__ __forceinline__ Pt<2> load(int n) const {
      const int x = reflect101(n, pad, T);
      const uint32_t w = __ldg(row + (x < 0 ? 0 : x));
      Pt<2> r;
      r.v[0] = make_float2(u8_to_f32(w, 0), u8_to_f32(w, 1));
      r.v[1] = make_float2(u8_to_f32(w, 2), u8_to_f32(w, 3));
      if (x < 0) r.v[0] = r.v[1] = make_float2(0.f, 0.f);
      return r;
    }
  };
  struct Storer {
    float4* row;
    int pad, T;
    __device__ __forceinline__ void store(int n, const Pt<2>& v) const {
      const int x = n - pad;
      if (x < 0 || x >= T) return;
      row[x] = make_float4(v.v[0].x, v.v[0].y, v.v[1].x, v.v[1].y);
    }
  };
  static __device__ __forceinline__ void make(const RowParams& P, int unit, Loader& ld, Storer& st) {
    ld.pad = st.pad = P.pad;
    ld.T = st.T = P.T;
    ld.row = reinterpret_cast<const uint32_t*>(P.in + (long long)unit * P.in_pitch);
    st.row = reinterpret_cast<float4*>(P.inter + (long long)unit * P.inter_pitch);
  }
};

// --------------------------------------------------------------------------------------------
// Column pass
// --------------------------------------------------------------------------------------------
struct ColParams {
  const float* inter;      // [frames][T rows][pitch] fp32
  long long pitch;         // floats per row (multiple of 4)
  long long frame_stride;  // floats per frame
  int T, pad;              // line length (rows) and reflect pad
  int W;                   // valid float columns per row (cols * Cp)
  int groups_per_frame;    // ceil(W / (2*NSEQ))
  int nunits;              // frames * groups_per_frame
  uint8_t* out;            // HWC u8
  long long out_pitch;     // bytes per image row
  long long out_frame_stride;
  int C, Cp;
};

// PACKED: Cp == C and every row/pitch is aligned so that the 2*NSEQ bytes of a unit are one
// aligned 16/32-bit store; otherwise bytes are scattered individually.
template <int NSEQ, bool PACKED>
struct ColProblem {
  static constexpr int L = 2 * NSEQ;
  struct Loader {
    const float* col;  // inter + frame*frame_stride + w0
    long long pitch;
    int pad, T;
    __device__ __forceinline__ Pt<NSEQ> load(int n) const {
      const int r = reflect101(n, pad, T);
      const float* p = col + (long long)(r < 0 ? 0 : r) * pitch;
      Pt<NSEQ> o;
      if constexpr (NSEQ == 2) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p));
        o.v[0] = make_float2(t.x, t.y);
        o.v[1] = make_float2(t.z, t.w);
      } else {
        o.v[0] = __ldg(reinterpret_cast<const float2*>(p));
      }
      if (r < 0) {
#pragma unroll
        for (int q = 0; q < NSEQ; ++q) o.v[q] = make_float2(0.f, 0.f);
      }
      return o;
    }
  };
  struct Storer {
    uint8_t* base;  // out + frame*out_frame_stride (+ w0 when PACKED)
    long long pitch;
    int off[L];     // byte offset of each line inside a row, -1 = not stored (generic path)
    int pad, T;
    __device__ __forceinline__ void store(int n, const Pt<NSEQ>& v) const {
      const int r = n - pad;
      if (r < 0 || r >= T) return;
      uint8_t* row = base + (long long)r * pitch;
      if constexpr (PACKED) {
        if constexpr (NSEQ == 2) {
          uint32_t w = f32_to_u8(v.v[0].x) | (f32_to_u8(v.v[0].y) << 8) | (f32_to_u8(v.v[1].x) << 16) |
                       (f32_to_u8(v.v[1].y) << 24);
          *reinterpret_cast<uint32_t*>(row) = w;
        } else {
          uint16_t w = (uint16_t)(f32_to_u8(v.v[0].x) | (f32_to_u8(v.v[0].y) << 8));
          *reinterpret_cast<uint16_t*>(row) = w;
        }
      } else {
#pragma unroll
        for (int s = 0; s < NSEQ; ++s) {
          if (off[2 * s] >= 0) row[off[2 * s]] = (uint8_t)f32_to_u8(v.v[s].x);
          if (off[2 * s + 1] >= 0) row[off[2 * s + 1]] = (uint8_t)f32_to_u8(v.v[s].y);
        }
      }
    }
  };
  static __device__ __forceinline__ void make(const ColParams& P, int unit, Loader& ld, Storer& st) {
    const int f = unit / P.groups_per_frame;
    const int g = unit - f * P.groups_per_frame;
    const int w0 = g * L;
    ld.pad = st.pad = P.pad;
    ld.T = st.T = P.T;
    ld.pitch = P.pitch;
    ld.col = P.inter + (long long)f * P.frame_stride + w0;
    st.pitch = P.out_pitch;
    st.base = P.out + (long long)f * P.out_frame_stride + (PACKED ? w0 : 0);
#pragma unroll
    for (int k = 0; k < L; ++k) {
      const int w = w0 + k;
      st.off[k] = (w < P.W) ? (w / P.Cp) * P.C + (w % P.Cp) : -1;
    }
  }
};

}  // namespace gblur
