// Library plumbing + layout packing kernels of libsgnn_b200.
#include <stdarg.h>
#include <stdio.h>

#include "common.cuh"

static thread_local char g_last_error[512] = "";

void sgnn_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
}

extern "C" const char* sgnn_last_error(void) { return g_last_error; }
extern "C" int sgnn_version(void) { return 100; }
extern "C" int sgnn_sm_arch(void) { return 100; }

// Number of SMs / compute capability of the current device (used by the host side to size
// persistent grids and to refuse non-sm_100 devices).
extern "C" int sgnn_device_info(int* sm_count, int* cc_major, int* cc_minor) {
  int dev = 0;
  SGNN_CUDA(cudaGetDevice(&dev));
  SGNN_CUDA(cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev));
  SGNN_CUDA(cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev));
  SGNN_CUDA(cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
  return SGNN_OK;
}

namespace sgnn {

// W[row0 + n, col0 + k] (row-major, leading dimension ld) -> K-major SWIZZLE_128B image with
// n rows and kpad columns (zero padded), optionally scaled and rounded to TF32.
__global__ void pack_weight_kernel(const float* __restrict__ w, int ld, int row0, int n, int col0,
                                   int k, int kpad, float scale, int round_tf32,
                                   float* __restrict__ dst) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * kpad) return;
  const int r = idx / kpad, c = idx - r * kpad;
  float v = c < k ? w[(long)(row0 + r) * ld + col0 + c] * scale : 0.f;
  if (round_tf32) v = to_tf32(v);
  *reinterpret_cast<float*>(reinterpret_cast<char*>(dst) + sw128_offset(r, c, n)) = v;
}

// rows[perm[e], 0:width] (row-major) -> 128-row tiles in the A-operand layout, one tile after
// another (tile t = rows 128t .. 128t+127 of the permuted order); tail rows zero filled.
__global__ void pack_rows_to_tiles_kernel(const float* __restrict__ rows, long ld,
                                          const int* __restrict__ perm, int n_rows, int width,
                                          int round_tf32, float* __restrict__ tiles) {
  const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int chunks = width >> 2;
  const long total = (long)sgnn_ceil_div(n_rows, kTileRows) * kTileRows * chunks;
  if (idx >= total) return;
  const long row = idx / chunks;
  const int c = (int)(idx - row * chunks);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (row < n_rows) {
    const long g = perm ? perm[row] : row;
    v = *(reinterpret_cast<const float4*>(rows + g * ld) + c);
  }
  if (round_tf32) v = make_float4(to_tf32(v.x), to_tf32(v.y), to_tf32(v.z), to_tf32(v.w));
  const long tile = row / kTileRows;
  const int r = (int)(row - tile * kTileRows);
  char* base = reinterpret_cast<char*>(tiles) + tile * (long)kTileRows * width * 4;
  *reinterpret_cast<float4*>(base + sw128_chunk_offset(r, c, kTileRows)) = v;
}

// inverse of the above (tile layout -> row-major rows[perm[e]]), times `scale`
__global__ void unpack_tiles_to_rows_kernel(const float* __restrict__ tiles, const int* __restrict__ perm,
                                            int n_rows, int width, float scale,
                                            float* __restrict__ rows, long ld) {
  const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int chunks = width >> 2;
  if (idx >= (long)n_rows * chunks) return;
  const long row = idx / chunks;
  const int c = (int)(idx - row * chunks);
  const long tile = row / kTileRows;
  const int r = (int)(row - tile * kTileRows);
  const char* base = reinterpret_cast<const char*>(tiles) + tile * (long)kTileRows * width * 4;
  float4 v = *reinterpret_cast<const float4*>(base + sw128_chunk_offset(r, c, kTileRows));
  v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
  const long g = perm ? perm[row] : row;
  *(reinterpret_cast<float4*>(rows + g * ld) + c) = v;
}

}  // namespace sgnn

using namespace sgnn;

extern "C" int sgnn_pack_weight(const float* w, int ld, int row0, int n, int col0, int k, int kpad,
                                float scale, int round_tf32, float* dst, void* stream) {
  SGNN_CHECK_ARG(n > 0 && n % 8 == 0, "pack_weight: n=%d must be a positive multiple of 8", n);
  SGNN_CHECK_ARG(kpad % 32 == 0 && k <= kpad && k > 0, "pack_weight: bad k=%d kpad=%d", k, kpad);
  pack_weight_kernel<<<sgnn_ceil_div((long)n * kpad, 256), 256, 0, (cudaStream_t)stream>>>(
      w, ld, row0, n, col0, k, kpad, scale, round_tf32, dst);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

extern "C" int sgnn_pack_rows_to_tiles(const float* rows, long ld, const int* perm, int n_rows,
                                       int width, int round_tf32, float* tiles, void* stream) {
  SGNN_CHECK_ARG(width % 32 == 0 && width > 0, "pack_rows_to_tiles: width=%d must be a multiple of 32", width);
  if (n_rows == 0) return SGNN_OK;
  const long total = (long)sgnn_ceil_div(n_rows, kTileRows) * kTileRows * (width / 4);
  pack_rows_to_tiles_kernel<<<sgnn_ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(
      rows, ld, perm, n_rows, width, round_tf32, tiles);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

extern "C" int sgnn_unpack_tiles_to_rows(const float* tiles, const int* perm, int n_rows, int width,
                                         float scale, float* rows, long ld, void* stream) {
  SGNN_CHECK_ARG(width % 32 == 0 && width > 0, "unpack_tiles_to_rows: width=%d must be a multiple of 32", width);
  if (n_rows == 0) return SGNN_OK;
  unpack_tiles_to_rows_kernel<<<sgnn_ceil_div((long)n_rows * (width / 4), 256), 256, 0,
                                (cudaStream_t)stream>>>(tiles, perm, n_rows, width, scale, rows, ld);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
