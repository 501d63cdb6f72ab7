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
#include <cuda.h>

#include <cstdint>

#include "async_copy.cuh"
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
  // row-major only: pixel columns split in bands of split_cols (0 = off); band h of every row lives at
  // inter + h * split_stride with row pitch split_cols * Cp (send buffer of the band-sharded all-to-all)
  int split_cols;
  long long split_stride;
  // band-sharded fused exchange (BLOCKED problems): panel h*peer_panels .. of every row goes to the panel
  // buffer of rank h, peer[h] (NVLink peer memory; peer[own rank] is local). npeers == 0: off, `inter` is used.
  float* peer[8];
  int npeers;
  int peer_panels;
  unsigned peer_magic;  // floor(2^32 / peer_panels) + 1: panel / peer_panels == umulhi(panel, magic) for panel < 2^16
  int row_offset;  // first image row of this band (slot = pad + row_offset + r)
  // panel intermediate (BLOCKED problems): [frame][panel = w/32][slot physc(n)][32 floats] with
  // w = x*Cp + ch the float column: every row writes whole 128-byte lines, and a column unit
  // (L adjacent float columns) finds its padded line at a fixed 128-byte stride; see BlockedDst
  int rows_per_frame;          // column-pass line length (rows of one frame)
  int blk_L;                   // floats per column group = 2 * (column plan's NSEQ)
  int blk_np;                  // slots per panel (N_col + padding slots)
  int blk_padq;                // column plan's padq
  long long blk_frame_stride;  // floats per frame = panels * np * 32
};

// Where one image row lands in the blocked intermediate: slot n = pad + r plus, for rows inside
// the reflect_101 border, the mirrored slots (the row pass writes the column pass's padding, so
// the column pass needs no index arithmetic at all; the zero tail is never written).
constexpr int kPanelFloats = 32;  // one 128-byte line per (panel, slot)
// peer mode: RowParams::inter is this fake base, so Storer::q[] carries a non-null offset into every rank's panel buffer
constexpr unsigned long long kPeerFakeBase = 4096;

struct BlockedDst {
  float* base;      // inter + frame offset + physc(pad + r) * 32
  int dl, dr;       // float offsets of the left / right mirror slot relative to base, 0 = none
  long long gstride;  // floats between consecutive panels = np * 32
};

__device__ __forceinline__ BlockedDst make_blocked_dst(const RowParams& P, long long r_global) {
  BlockedDst d;
  const long long f = (r_global + P.row_offset) / P.rows_per_frame;
  const int r = (int)(r_global + P.row_offset - f * P.rows_per_frame);
  const int Tc = P.rows_per_frame, pad = P.pad, L = kPanelFloats;
  const int p1 = phys(pad + r, P.blk_padq);
  d.base = P.inter + f * P.blk_frame_stride + (long long)p1 * L;
  d.gstride = (long long)P.blk_np * L;
  d.dl = (r >= 1 && r <= pad) ? (phys(pad - r, P.blk_padq) - p1) * L : 0;
  d.dr = (r >= Tc - 1 - pad && r <= Tc - 2) ? (phys(pad + 2 * Tc - 2 - r, P.blk_padq) - p1) * L : 0;
  return d;
}

// Any C, Cp: real line sigma = r*Cp + ch; a unit takes 2*NSEQ consecutive lines.
// BLOCKED selects the blocked intermediate (scalar scattered stores) instead of row-major.
template <int NSEQ, bool BLOCKED = false>
struct RowProblemGeneric {
  static constexpr int L = 2 * NSEQ;
  static constexpr bool kBulkIn = false;
  static constexpr bool kTmaIn = false;
  static constexpr bool kStageIn = false;
  struct Loader {
    static constexpr bool kStageNext = false;
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
    static constexpr bool kPrefetch = false;
    static constexpr bool kTmaNext = false;
    float* q[L];  // nullptr: line does not exist (tail unit); BLOCKED: BlockedDst::base
    int dl[L], dr[L];
    long long gstride;
    int ch[L];
    int Cp, pad, T;
    const RowParams* prm;  // rarely used fields (split store, peer panels) are read from the parameter bank
    __device__ __forceinline__ void put(int k, int x, int w0, float v) const {
      if (!q[k]) return;
      if constexpr (!BLOCKED) {
        if (prm->split_cols) {
          const int h = x / prm->split_cols;
          q[k][(long long)h * prm->split_stride + (long long)(x - h * prm->split_cols) * Cp] = v;
        } else {
          q[k][(long long)x * Cp] = v;
        }
      } else {
        const int w = w0 + ch[k];
        int panel = w >> 5;
        float* d = q[k];
        if (prm->npeers) {  // q[k] holds the offset inside a panel buffer; pick the owner of this panel
          const int pp = prm->peer_panels;
          const int h = pp == 1 ? panel : (int)__umulhi((unsigned)panel, prm->peer_magic);
          panel -= h * pp;
          d = prm->peer[h] + ((reinterpret_cast<uintptr_t>(q[k]) - kPeerFakeBase) >> 2);
        }
        d += (long long)panel * (int)gstride + (w & 31);
        *d = v;
        if (dl[k]) d[dl[k]] = v;
        if (dr[k]) d[dr[k]] = v;
      }
    }
    // one plane (Cp == 1), panel layout: every line of the unit lands in the same panel column, so the panel
    // arithmetic (and the owner lookup of the fused exchange) is done once per position
    __device__ __forceinline__ void store_planar(int x, const Pt<NSEQ>& v) const {
      int panel = x >> 5;
      long long off = x & 31;
      float* peer_base = nullptr;
      if (prm->npeers) {
        const int pp = prm->peer_panels;
        const int h = pp == 1 ? panel : (int)__umulhi((unsigned)panel, prm->peer_magic);
        panel -= h * pp;
        peer_base = prm->peer[h];
      }
      off += (long long)panel * (int)gstride;
#pragma unroll
      for (int k = 0; k < L; ++k) {
        if (!q[k]) continue;
        float* d = (peer_base ? peer_base + ((reinterpret_cast<uintptr_t>(q[k]) - kPeerFakeBase) >> 2) : q[k]) + off;
        const float val = (k & 1) ? v.v[k >> 1].y : v.v[k >> 1].x;
        *d = val;
        if (dl[k]) d[dl[k]] = val;
        if (dr[k]) d[dr[k]] = val;
      }
    }
    __device__ __forceinline__ void store(int n, const Pt<NSEQ>& v) const {
      const int x = n - pad;
      if (x < 0 || x >= T) return;
      if constexpr (BLOCKED) {
        if (Cp == 1) {
          store_planar(x, v);
          return;
        }
      }
      const int w0 = x * Cp;
#pragma unroll
      for (int s = 0; s < NSEQ; ++s) {
        put(2 * s, x, w0, v.v[s].x);
        put(2 * s + 1, x, w0, v.v[s].y);
      }
    }
  };
  static __device__ __forceinline__ void make(const RowParams& P, int unit, Loader& ld, Storer& st) {
    ld.C = P.C;
    ld.pad = st.pad = P.pad;
    ld.T = st.T = P.T;
    st.Cp = P.Cp;
    st.gstride = 0;
    st.prm = &P;
#pragma unroll
    for (int k = 0; k < L; ++k) {
      long long sig = (long long)unit * L + k;
      const bool ok = sig < P.nreal;
      if (!ok) sig = (long long)unit * L;  // read a real line, never store it
      const long long r = sig / P.Cp;
      const int ch = (int)(sig - r * P.Cp);
      ld.p[k] = P.in + r * P.in_pitch + ch;
      st.ch[k] = ch;
      st.dl[k] = st.dr[k] = 0;
      if constexpr (!BLOCKED) {
        st.q[k] = ok ? P.inter + r * P.inter_pitch + ch : nullptr;
      } else {
        const BlockedDst d = make_blocked_dst(P, r);
        st.q[k] = ok ? d.base : nullptr;
        st.dl[k] = d.dl;
        st.dr[k] = d.dr;
        st.gstride = d.gstride;
      }
    }
  }
};

// C == Cp == 4, two sequences: one uchar4 load feeds (R + iG, B + iA), one float4 store per pixel.
// BLOCKED (needs a column plan with two sequences, L = 4): pixel x is column group x.
// STAGED: the row's bytes are not loaded from global memory by the first stage; a bulk asynchronous copy
// (TMA 1-D) has put them into a shared-memory staging buffer while the previous row was being transformed,
// and the copy of the NEXT row is started as soon as the first stage has consumed this one.
template <bool BLOCKED = false, bool STAGED = false>
struct RowProblemRGBA {
  static constexpr bool kBulkIn = false;
  static constexpr bool kTmaIn = false;
  static constexpr bool kStageIn = STAGED;
  struct Loader {
    static constexpr bool kStageNext = STAGED;
    const uint32_t* row;  // global row, or the staging buffer when STAGED
    int pad, T;
    uint64_t* bar;         // STAGED only
    const void* next_row;  // global address of the row this CTA transforms next, or nullptr
    uint32_t row_bytes;
    __device__ __forceinline__ void set_stage(void* buf, uint64_t* b) {
      row = static_cast<const uint32_t*>(buf);
      bar = b;
    }
    __device__ __forceinline__ Pt<2> load(int n) const {
      const int x = reflect101(n, pad, T);
      const uint32_t w = STAGED ? row[x < 0 ? 0 : x] : __ldg(row + (x < 0 ? 0 : x));
      Pt<2> r;
      r.v[0] = make_float2(u8_to_f32(w, 0), u8_to_f32(w, 1));
      r.v[1] = make_float2(u8_to_f32(w, 2), u8_to_f32(w, 3));
      if (x < 0) r.v[0] = r.v[1] = make_float2(0.f, 0.f);
      return r;
    }
    // every thread calls this after the barrier that ends the first stage
    __device__ __forceinline__ void after_first_stage() const {
      if (threadIdx.x == 0 && next_row) {
        fence_proxy_async();
        mbar_expect_tx(bar, row_bytes);
        bulk_g2s(const_cast<uint32_t*>(row), next_row, row_bytes, bar);
      }
    }
  };
  struct Storer {
    static constexpr bool kPrefetch = false;
    static constexpr bool kTmaNext = false;
    float* row;  // row-major: start of the row; blocked: BlockedDst::base
    int dl, dr;
    long long gstride;
    int pad, T;
    __device__ __forceinline__ void store(int n, const Pt<2>& v) const {
      const int x = n - pad;
      if (x < 0 || x >= T) return;
      const float4 o = make_float4(v.v[0].x, v.v[0].y, v.v[1].x, v.v[1].y);
      if constexpr (!BLOCKED) {
        reinterpret_cast<float4*>(row)[x] = o;
      } else {
        float* d = row + (long long)(x >> 3) * gstride + ((x & 7) << 2);
        *reinterpret_cast<float4*>(d) = o;
        if (dl) *reinterpret_cast<float4*>(d + dl) = o;
        if (dr) *reinterpret_cast<float4*>(d + dr) = o;
      }
    }
  };
  // STAGED: what the bulk copy for `unit` moves (one image row)
  static __device__ __forceinline__ const void* stage_src(const RowParams& P, long long unit) {
    return P.in + unit * P.in_pitch;
  }
  static __device__ __forceinline__ uint32_t stage_bytes(const RowParams& P, long long) { return (uint32_t)P.T * 4u; }
  static __device__ __forceinline__ void make(const RowParams& P, int unit, Loader& ld, Storer& st) {
    ld.pad = st.pad = P.pad;
    ld.T = st.T = P.T;
    ld.row = reinterpret_cast<const uint32_t*>(P.in + (long long)unit * P.in_pitch);
    ld.bar = nullptr;
    ld.next_row = nullptr;
    ld.row_bytes = (uint32_t)P.T * 4u;
    if constexpr (STAGED) {
      const long long next = (long long)unit + gridDim.x;
      if (next < P.nunits) ld.next_row = P.in + next * P.in_pitch;
    }
    st.dl = st.dr = st.gstride = 0;
    if constexpr (!BLOCKED) {
      st.row = P.inter + (long long)unit * P.inter_pitch;
    } else {
      const BlockedDst d = make_blocked_dst(P, unit);
      st.row = d.base;
      st.dl = d.dl;
      st.dr = d.dr;
      st.gstride = d.gstride;
    }
  }
};

// C == Cp == 4, ONE sequence per unit: unit = (row, channel pair p), (R + iG) for p = 0 and
// (B + iA) for p = 1. 16-bit loads, float2 stores (row-major: at 16-byte stride; blocked with
// L = 4: the half of group x that belongs to pair p).
template <bool BLOCKED = false>
struct RowProblemRGBA1 {
  static constexpr bool kBulkIn = false;
  static constexpr bool kTmaIn = false;
  static constexpr bool kStageIn = false;
  struct Loader {
    static constexpr bool kStageNext = false;
    const uint16_t* row;  // already offset by the pair
    int pad, T;
    __device__ __forceinline__ Pt<1> load(int n) const {
      const int x = reflect101(n, pad, T);
      const uint32_t w = __ldg(row + 2 * (x < 0 ? 0 : x));
      Pt<1> r;
      r.v[0] = (x < 0) ? make_float2(0.f, 0.f) : make_float2(u8_to_f32(w, 0), u8_to_f32(w, 1));
      return r;
    }
  };
  struct Storer {
    static constexpr bool kPrefetch = false;
    static constexpr bool kTmaNext = false;
    float* row;
    int dl, dr;
    long long gstride;
    int pad, T;
    __device__ __forceinline__ void store(int n, const Pt<1>& v) const {
      const int x = n - pad;
      if (x < 0 || x >= T) return;
      if constexpr (!BLOCKED) {
        *reinterpret_cast<float2*>(row + (long long)x * 4) = v.v[0];
      } else {
        float* d = row + (long long)(x >> 3) * gstride + ((x & 7) << 2);
        *reinterpret_cast<float2*>(d) = v.v[0];
        if (dl) *reinterpret_cast<float2*>(d + dl) = v.v[0];
        if (dr) *reinterpret_cast<float2*>(d + dr) = v.v[0];
      }
    }
  };
  static __device__ __forceinline__ void make(const RowParams& P, int unit, Loader& ld, Storer& st) {
    const int r = unit >> 1, p = unit & 1;
    ld.pad = st.pad = P.pad;
    ld.T = st.T = P.T;
    ld.row = reinterpret_cast<const uint16_t*>(P.in + (long long)r * P.in_pitch) + p;
    st.dl = st.dr = 0;
    if constexpr (!BLOCKED) {
      st.row = P.inter + (long long)r * P.inter_pitch + 2 * p;
      st.gstride = 0;
    } else {
      const BlockedDst d = make_blocked_dst(P, r);
      st.row = d.base + 2 * p;
      st.dl = d.dl;
      st.dr = d.dr;
      st.gstride = d.gstride;
    }
  }
};

// Cp == 3 (RGB, or RGBA with the alpha plane left alone), two sequences, panel intermediate only.
// A unit is the same four consecutive lines sigma = 4u .. 4u+3 that RowProblemGeneric<2> takes (so
// the pairing (a + ib) and therefore every rounding is identical), but they always lie in the image
// rows r0 = 4u/3 and r0 + 1: the loader is four byte loads at fixed offsets from one base, the storer
// four scalar stores with one 32-bit panel computation each. Rows inside the reflect border and
// the tail unit take a checked slow path.
// STAGED: rows r0 and r0+1 are bulk-copied into shared memory while the previous unit runs.
template <bool STAGED = false>
struct RowProblemRGB3 {
  static constexpr bool kBulkIn = false;
  static constexpr bool kTmaIn = false;
  static constexpr bool kStageIn = STAGED;
  struct Loader {
    static constexpr bool kStageNext = STAGED;
    const uint8_t* row;  // global row r0, or the staging buffer when STAGED
    int off[4];          // byte offset of line k's channel: (row select) * pitch + channel
    int C, pad, T;
    uint64_t* bar;         // STAGED only
    const void* next_row;  // global address of the rows this CTA stages next, or nullptr
    uint32_t next_bytes;
    __device__ __forceinline__ void set_stage(void* buf, uint64_t* b) {
      row = static_cast<const uint8_t*>(buf);
      bar = b;
    }
    static __device__ __forceinline__ float byte_to_f32(uint32_t b) {
      return __uint_as_float(b | 0x4B000000u) - 8388608.0f;
    }
    __device__ __forceinline__ Pt<2> load(int n) const {
      const int x = reflect101(n, pad, T);
      const int xo = (x < 0 ? 0 : x) * C;
      float f[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint32_t b;
        if constexpr (STAGED) {
          asm volatile("ld.shared.u8 %0, [%1];" : "=r"(b) : "r"(smem_addr(row) + (uint32_t)(off[k] + xo)));
        } else {
          b = __ldg(row + (off[k] + xo));
        }
        f[k] = byte_to_f32(b);
      }
      Pt<2> r;
      r.v[0] = make_float2(f[0], f[1]);
      r.v[1] = make_float2(f[2], f[3]);
      if (x < 0) r.v[0] = r.v[1] = make_float2(0.f, 0.f);
      return r;
    }
    __device__ __forceinline__ void after_first_stage() const {
      if (threadIdx.x == 0 && next_row) {
        fence_proxy_async();
        mbar_expect_tx(bar, next_bytes);
        bulk_g2s(const_cast<uint8_t*>(row), next_row, next_bytes, bar);
      }
    }
  };
  struct Storer {
    static constexpr bool kPrefetch = false;
    static constexpr bool kTmaNext = false;
    float* q[4];    // BlockedDst::base of line k's image row
    int ch[4];      // channel of line k
    int dl0, dr0, dl1, dr1;  // mirror slots of rows r0, r0 + 1 (float offsets, 0 = none)
    int m0;         // lines k < m0 are in row r0, the others in row r0 + 1
    int nvalid;     // lines k >= nvalid do not exist (tail unit)
    int gs32;       // floats between consecutive panels minus 32
    int slow;       // any mirror slot or missing line: take the checked path
    int pad, T;
    __device__ __forceinline__ void store(int n, const Pt<2>& v) const {
      const int x = n - pad;
      if (x < 0 || x >= T) return;
      const int w0 = 3 * x;
      const float val[4] = {v.v[0].x, v.v[0].y, v.v[1].x, v.v[1].y};
      if (!slow) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int w = w0 + ch[k];
          q[k][w + (w >> 5) * gs32] = val[k];
        }
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (k >= nvalid) continue;
          const int w = w0 + ch[k];
          float* d = q[k] + (w + (w >> 5) * gs32);
          const int ml = k >= m0 ? dl1 : dl0, mr = k >= m0 ? dr1 : dr0;  // no dynamic indexing: keeps the struct in registers
          *d = val[k];
          if (ml) d[ml] = val[k];
          if (mr) d[mr] = val[k];
        }
      }
    }
  };
  static __device__ __forceinline__ long long first_row(long long unit) { return (unit * 4) / 3; }
  // STAGED: the bulk copy for `unit` moves rows r0 and (when it exists) r0 + 1; needs in_pitch == T * C
  static __device__ __forceinline__ const void* stage_src(const RowParams& P, long long unit) {
    return P.in + first_row(unit) * P.in_pitch;
  }
  static __device__ __forceinline__ uint32_t stage_bytes(const RowParams& P, long long unit) {
    const long long rows_total = P.nreal / 3;
    return (uint32_t)P.in_pitch * (first_row(unit) + 1 < rows_total ? 2u : 1u);
  }
  static __device__ __forceinline__ void make(const RowParams& P, int unit, Loader& ld, Storer& st) {
    const long long sig0 = (long long)unit * 4;
    const long long r0 = sig0 / 3;
    const int c0 = (int)(sig0 - r0 * 3);
    ld.C = P.C;
    ld.pad = st.pad = P.pad;
    ld.T = st.T = P.T;
    ld.row = P.in + r0 * P.in_pitch;
    ld.bar = nullptr;
    ld.next_row = nullptr;
    ld.next_bytes = 0;
    if constexpr (STAGED) {
      const long long next = (long long)unit + gridDim.x;
      if (next < P.nunits) {
        ld.next_row = stage_src(P, next);
        ld.next_bytes = stage_bytes(P, next);
      }
    }
    const long long left = P.nreal - sig0;
    st.nvalid = left < 4 ? (int)left : 4;
    st.m0 = 3 - c0;
    const BlockedDst d0 = make_blocked_dst(P, r0);
    BlockedDst d1 = d0;
    if (st.nvalid > st.m0) d1 = make_blocked_dst(P, r0 + 1);
    st.dl0 = d0.dl;
    st.dr0 = d0.dr;
    st.dl1 = d1.dl;
    st.dr1 = d1.dr;
    st.gs32 = (int)d0.gstride - kPanelFloats;
    st.slow = (st.nvalid < 4) | d0.dl | d0.dr | d1.dl | d1.dr;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bool ok = k < st.nvalid;
      const int rs = (ok && k >= st.m0) ? 1 : 0;
      const int ch = ok ? (rs ? k - st.m0 : c0 + k) : c0;  // a missing line re-reads line 0 and is never stored
      ld.off[k] = rs * (int)P.in_pitch + ch;
      st.ch[k] = ch;
      st.q[k] = rs ? d1.base : d0.base;
    }
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
  int W;                   // end of the valid float columns (cols * Cp): columns >= W are not stored
  int w_begin;             // first float column of this launch (a multiple of 32; column-band launches)
  int groups_per_frame;    // ceil(W / (2*NSEQ))
  int nunits;              // frames * groups_per_frame
  uint8_t* out;            // HWC u8
  long long out_pitch;     // bytes per image row
  long long out_frame_stride;
  int C, Cp;
  // BULK problems: inter is the panel layout [frame][panel][np slots][32 floats]; a unit's padded
  // line is np pieces of L floats at a 128-byte stride, gathered with cp.async into the exchange buffer
  long long panel_stride;  // floats between consecutive panels = np * 32
  int np;                  // slots per unit
  int panel_in;            // non-BULK problems: inter is the panel layout too (cluster column pass)
  // TMA problems: tensor map over the panel intermediate, dims {32 floats, np slots, panels*frames}, box
  // {L floats, tma_rows slots, 1}: one unit = np / tma_rows strided tiles of 4*L-byte rows
  CUtensorMap tmap;
  int tma_rows;
  int panels_per_frame;
};

// PACKED: Cp == C and every row/pitch is aligned so that the 2*NSEQ bytes of a unit are one
// aligned 16/32-bit store; otherwise bytes are scattered individually.
// BULK: the unit is gathered with cp.async from the panel intermediate (padding and mirrored
// border rows were already written by the row pass), so the Loader is never called.
// TMA (with BULK): the gather is done by the TMA unit instead — strided tiles through a tensor map —
// and the next unit's tiles are requested as soon as the last inverse stage has read the buffer.
template <int NSEQ, bool PACKED, bool BULK = false, bool TMA = false>
struct ColProblem {
  static constexpr int L = 2 * NSEQ;
  static constexpr bool kBulkIn = BULK;
  static constexpr bool kTmaIn = BULK && TMA;
  static constexpr bool kStageIn = false;
  // request every tile of `unit` (one elected thread); completion is counted in bytes on `bar`
  static __device__ __forceinline__ void tma_request(const ColParams& P, int unit, void* smbase, uint64_t* bar) {
    const int f = unit / P.groups_per_frame;
    const int g = unit - f * P.groups_per_frame;
    const int w0 = P.w_begin + g * L;
    const int panel = f * P.panels_per_frame + (w0 >> 5);
    fence_proxy_async();
    mbar_expect_tx(bar, (uint32_t)P.np * (uint32_t)(4 * L));
    char* dst = static_cast<char*>(smbase);
    for (int r = 0; r < P.np; r += P.tma_rows) tma_load_3d(dst + (size_t)r * (4 * L), &P.tmap, w0 & 31, r, panel, bar);
  }
  static __device__ __forceinline__ const float* bulk_src(const ColParams& P, int unit) {
    const int f = unit / P.groups_per_frame;
    const int g = unit - f * P.groups_per_frame;
    const int w0 = P.w_begin + g * L;
    return P.inter + (long long)f * P.frame_stride + (long long)(w0 >> 5) * P.panel_stride + (w0 & 31);
  }
  struct Loader {
    static constexpr bool kStageNext = false;
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
    static constexpr bool kPrefetch = BULK && !TMA;
    static constexpr bool kTmaNext = BULK && TMA;
    const ColParams* params;  // TMA: request the next unit's tiles from here
    void* smbase;
    uint64_t* bar;
    int next_unit;            // -1: none
    // called by every thread once the last inverse stage has read its slots
    __device__ __forceinline__ void after_reads() const {
      __syncthreads();
      if (threadIdx.x == 0 && next_unit >= 0) tma_request(*params, next_unit, smbase, bar);
    }
    uint8_t* base;  // out + frame*out_frame_stride (+ w0 when PACKED)
    long long pitch;
    int off[L];     // byte offset of each line inside a row, -1 = not stored (generic path)
    int pad, T;
    const float* next_src;  // BULK: padded line of the unit this CTA processes next, or nullptr
    __device__ __forceinline__ void prefetch(typename SmemElem<NSEQ>::type* dst, int phys_slot) const {
      if (next_src) cp_async<8 * NSEQ>(dst, next_src + (long long)phys_slot * kPanelFloats);
    }
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
    const int w0 = P.w_begin + g * L;
    ld.pad = st.pad = P.pad;
    ld.T = st.T = P.T;
    ld.pitch = P.pitch;
    ld.col = P.inter + (long long)f * P.frame_stride + w0;
    if (!BULK && P.panel_in) {  // slot n of the padded line sits at n * 128 bytes; borders are already mirrored
      ld.pad = 0;
      ld.T = P.np;
      ld.pitch = kPanelFloats;
      ld.col = bulk_src(P, unit);
    }
    st.pitch = P.out_pitch;
    st.base = P.out + (long long)f * P.out_frame_stride + (PACKED ? w0 : 0);
    st.next_src = nullptr;
    st.params = &P;
    st.smbase = nullptr;
    st.bar = nullptr;
    st.next_unit = -1;
    if constexpr (BULK) {
      const int next = unit + (int)gridDim.x;
      if (next < P.nunits) {
        st.next_src = bulk_src(P, next);
        st.next_unit = next;
      }
    }
#pragma unroll
    for (int k = 0; k < L; ++k) {
      const int w = w0 + k;
      st.off[k] = (w < P.W) ? (w / P.Cp) * P.C + (w % P.Cp) : -1;
    }
  }
};

}  // namespace gblur
