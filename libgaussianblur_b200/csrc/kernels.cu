// kernels.cu — dispatch from the host plan to the kernel instantiations (static_impl.cu per
// compile-time shape, generic_row.cu / generic_col.cu for every other valid FFT length).
#include "kernels.h"

#include <atomic>
#include <cstdio>
#include <cstdlib>

#include "line_kernels.cuh"
#include "static_shapes.h"

namespace gblur {

static std::atomic<unsigned long long> g_launches{0};
unsigned long long launch_counter() { return g_launches.load(); }
void count_launch() { g_launches.fetch_add(1); }

__global__ void copy_channel_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                    long long npixels, int C, int ch) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < npixels;
       i += (long long)gridDim.x * blockDim.x)
    dst[i * C + ch] = src[i * C + ch];
}

struct StaticEntry {
  int id, nseq, r[4], nthreads;
};
static const StaticEntry kStaticTable[] = {
#define X(ID, NSEQ, R0, R1, R2, R3, NT, MINB) {ID, NSEQ, {R0, R1, R2, R3}, NT},
    GBLUR_STATIC_SHAPES(X)
#undef X
};

int find_static_shape(int N, int nseq, int* radices, int* nthreads) {
  // tuning override: GBLUR_VARIANT_<N>=<id> forces one table entry for that FFT length
  char name[64];
  snprintf(name, sizeof name, "GBLUR_VARIANT_%d", N);
  const char* forced = getenv(name);
  const int forced_id = forced ? atoi(forced) : -1;
  for (const auto& e : kStaticTable) {
    if (e.r[0] * e.r[1] * e.r[2] * e.r[3] != N) continue;
    if (forced_id >= 0 ? (e.id != forced_id || e.nseq != nseq) : (e.nseq != nseq)) continue;
    for (int i = 0; i < 4; ++i) radices[i] = e.r[i];
    *nthreads = e.nthreads;
    return e.id;
  }
  if (forced_id >= 0) {
    // a forced variant with the other sequence count must win over the default of this count
    for (const auto& e : kStaticTable)
      if (e.id == forced_id && e.r[0] * e.r[1] * e.r[2] * e.r[3] == N) return -2;
  }
  return -1;
}

cudaError_t kernels_init(int) { return cudaSuccess; }

namespace {
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeFn tensor_map_encoder() {
  static EncodeFn encode = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      fn = nullptr;
    return reinterpret_cast<EncodeFn>(fn);
  }();
  return encode;
}
}  // namespace

bool encode_tensor_map(CUtensorMap* out, int elem_bytes, int rank, const void* base, const unsigned long long* dims,
                       const unsigned long long* strides_bytes, const unsigned* box, int swizzle_bytes) {
  EncodeFn encode = tensor_map_encoder();
  if (!encode || rank < 1 || rank > 3) return false;
  cuuint64_t d[3], st[2];
  cuuint32_t b[3], es[3] = {1, 1, 1};
  for (int i = 0; i < rank; ++i) d[i] = dims[i], b[i] = box[i];
  for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
  const CUtensorMapSwizzle sw = swizzle_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                                : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                      : CU_TENSOR_MAP_SWIZZLE_NONE;
  const CUtensorMapDataType dt = elem_bytes == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  return encode(out, dt, (cuuint32_t)rank, const_cast<void*>(base), d, st, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool make_panel_tensor_map(CUtensorMap* out, const float* base, int np, long long panels, int L, int* box_rows) {
  EncodeFn encode = tensor_map_encoder();
  if (!encode || getenv("GBLUR_NO_TMA")) return false;
  int rows = 0;
  for (int r = 256; r >= 32; --r)
    if (np % r == 0) {
      rows = r;
      break;
    }
  if (!rows || panels < 1) return false;
  const cuuint64_t dims[3] = {32, (cuuint64_t)np, (cuuint64_t)panels};
  const cuuint64_t strides[2] = {128, (cuuint64_t)np * 128};  // bytes, dims 1 and 2
  const cuuint32_t box[3] = {(cuuint32_t)L, (cuuint32_t)rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  // L2 promotion (64/128/256 B) was measured on the 8K and 1080p column passes: no effect, left off
  const CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return false;
  *box_rows = rows;
  return true;
}

cudaError_t launch_rowpass(const LinePlan& lp, const RowLaunch& a, int num_sms, cudaStream_t stream) {
  RowParams P{};
  P.in = a.in;
  P.in_pitch = a.in_pitch;
  P.C = a.C;
  P.Cp = a.Cp;
  P.T = a.T;
  P.pad = a.pad;
  P.nreal = a.rows_total * a.Cp;
  P.inter = a.inter;
  P.inter_pitch = a.inter_pitch;
  const long long L = 2 * lp.nseq;
  P.nunits = (int)((P.nreal + L - 1) / L);
  const bool blocked = a.blk.enabled != 0;
  P.split_cols = blocked ? 0 : a.split_cols;
  P.split_stride = a.split_stride;
  P.npeers = blocked ? a.npeers : 0;
  P.peer_panels = a.peer_panels;
  P.peer_magic = a.peer_panels > 0 ? (unsigned)(0x100000000ULL / (unsigned)a.peer_panels) + 1u : 0u;
  P.row_offset = a.row_offset;
  for (int i = 0; i < 8; ++i) P.peer[i] = a.peer[i];
  if (P.npeers) P.inter = reinterpret_cast<float*>(kPeerFakeBase);
  P.rows_per_frame = blocked ? a.blk.rows_per_frame : 1;
  P.blk_L = a.blk.L;
  P.blk_np = a.blk.np;
  P.blk_padq = a.blk.padq;
  P.blk_frame_stride = a.blk.frame_stride;
  // uchar4/float4 (two sequences) or u16/float2 (one sequence, static shapes only) fast path
  const bool rgba = P.split_cols == 0 && P.npeers == 0 && (lp.nseq == 2 || lp.static_id >= 0) && a.C == 4 && a.Cp == 4 && (a.in_pitch % 4) == 0 &&
                    (reinterpret_cast<uintptr_t>(a.in) % 4) == 0 && (reinterpret_cast<uintptr_t>(a.inter) % 16) == 0 &&
                    (blocked || (a.inter_pitch % 4) == 0);
  // three processed channels (RGB, or RGBA without alpha), panel layout, two sequences: byte loads at fixed
  // offsets, 32-bit panel arithmetic (so every float offset inside a frame's panels must fit an int)
  const bool rgb3 = blocked && P.npeers == 0 && lp.nseq == 2 && a.Cp == 3 && (a.C == 3 || a.C == 4) &&
                    a.in_pitch < (1LL << 30) && !getenv("GBLUR_NO_RGB3") &&
                    (long long)((3LL * a.T + 31) / 32 + 1) * a.blk.np * kPanelFloats < (1LL << 31);
  const int fast = rgba ? kRowFastRGBA : rgb3 ? kRowFastRGB3 : kRowFastNone;
  if (lp.cluster > 1) return launch_row_cluster(lp, P, blocked, stream);
  switch (lp.static_id) {
#define X(ID, NSEQ, R0, R1, R2, R3, NT, MINB) \
  case ID:                                    \
    return launch_row_static<ID>(lp, P, fast, blocked, num_sms, stream);
    GBLUR_STATIC_SHAPES(X)
#undef X
    default:
      return launch_row_generic(lp, P, fast, blocked, num_sms, stream);
  }
}

cudaError_t launch_colpass(const LinePlan& lp, const ColLaunch& a, int num_sms, cudaStream_t stream) {
  ColParams P{};
  P.inter = a.inter;
  P.pitch = a.pitch;
  P.frame_stride = a.frame_stride;
  P.T = a.T;
  P.pad = a.pad;
  const int w_end = a.w_end > 0 ? a.w_end : a.W;
  P.W = w_end;
  P.w_begin = a.w_begin;
  const int L = 2 * lp.nseq;
  P.groups_per_frame = (w_end - a.w_begin + L - 1) / L;
  P.nunits = a.frames * P.groups_per_frame;
  P.out = a.out;
  P.out_pitch = a.out_pitch;
  P.out_frame_stride = a.out_frame_stride;
  P.C = a.C;
  P.Cp = a.Cp;
  const bool blocked = a.blk.enabled != 0;
  P.panel_stride = (long long)a.blk.np * 32;
  P.np = a.blk.np;
  P.panel_in = a.panel_in;
  P.tma_rows = 0;
  P.panels_per_frame = (a.W + 31) / 32;
  const bool tma = blocked && a.tmap && a.tma_rows > 0;
  if (tma) {
    P.tmap = *a.tmap;
    P.tma_rows = a.tma_rows;
  }
  if (blocked) P.frame_stride = a.blk.frame_stride;
  const bool packed = a.C == a.Cp && (a.W % L) == 0 && (a.out_pitch % L) == 0 && (a.out_frame_stride % L) == 0 &&
                      (reinterpret_cast<uintptr_t>(a.out) % L) == 0;
  if (lp.cluster > 1) return (blocked && !a.panel_in) ? cudaErrorInvalidValue : launch_col_cluster(lp, P, packed, stream);
  switch (lp.static_id) {
#define X(ID, NSEQ, R0, R1, R2, R3, NT, MINB) \
  case ID:                                    \
    return launch_col_static<ID>(lp, P, packed, blocked ? (tma ? 2 : 1) : 0, num_sms, stream);
    GBLUR_STATIC_SHAPES(X)
#undef X
    default:
      return launch_col_generic(lp, P, packed, blocked, num_sms, stream);
  }
}

cudaError_t launch_copy_channel(const uint8_t* src, uint8_t* dst, long long npixels, int C, int ch,
                                cudaStream_t stream) {
  if (npixels <= 0) return cudaSuccess;
  long long blocks = (npixels + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  copy_channel_kernel<<<(unsigned)blocks, 256, 0, stream>>>(src, dst, npixels, C, ch);
  count_launch();
  return cudaGetLastError();
}

}  // namespace gblur
