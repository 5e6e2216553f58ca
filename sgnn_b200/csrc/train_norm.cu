// LayerNorm forward / backward of the TRAINING path (the nn.LayerNorm that closes every encoder, edge
// and node MLP: sgnn/single_scale/graph_network.py:86-96,139-147; multi_scale_gnn.py:236-259).
//
// ATen's LayerNorm kernels are built for wide rows; on the [E, 64] edge latents of a training step
// (E = 190k) torch.profiler shows its forward at 153 us (98 MB -> 0.64 TB/s) and its backward at 260 us
// per call, 32 % of the whole step (profiles/README.md, r01n).  These rows are 128 .. 512 bytes: LPR =
// width / 4 lanes own one row (one float4 each), so a warp covers 32 / LPR consecutive rows with fully
// coalesced 16-byte accesses and the row statistics are 3 .. 5 shuffles.  Pure HBM streaming:
//   forward   reads x, writes y (+ mean, rstd)                  8 * width bytes per row
//   backward  reads dy, x (+ mean, rstd), writes dx             12 * width bytes per row
// The column sums of the backward (dgamma, dbeta) stay in registers across a thread's rows, are
// combined per CTA in a fixed order through shared memory and finished by a second tiny kernel over the
// per-CTA partials: no atomics, bit-reproducible gradients.
#include "common.cuh"

namespace sgnn {

constexpr int kLnThreads = 256;
constexpr int kLnMaxBlocks = 148 * 4;

template <int LPR>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int LPR>
__global__ void __launch_bounds__(kLnThreads)
ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
              float eps, long n_rows, float* __restrict__ y, float* __restrict__ mean,
              float* __restrict__ rstd) {
  constexpr int W = 4 * LPR, RPW = 32 / LPR;
  const int lane = threadIdx.x & 31, sub = lane % LPR, rw = lane / LPR;
  const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + sub);
  const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + sub);
  const long warps = (long)gridDim.x * (kLnThreads / 32);
  for (long r0 = ((long)blockIdx.x * (kLnThreads / 32) + (threadIdx.x >> 5)) * RPW; r0 < n_rows; r0 += warps * RPW) {
    const long r = r0 + rw;
    const bool ok = r < n_rows;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ok) v = __ldcs(reinterpret_cast<const float4*>(x) + r * LPR + sub);
    const float m = group_sum<LPR>((v.x + v.y) + (v.z + v.w)) * (1.0f / W);
    const float4 d = make_float4(v.x - m, v.y - m, v.z - m, v.w - m);
    const float var = group_sum<LPR>((d.x * d.x + d.y * d.y) + (d.z * d.z + d.w * d.w)) * (1.0f / W);
    const float rs = 1.0f / sqrtf(var + eps);
    if (ok) {
      __stcs(reinterpret_cast<float4*>(y) + r * LPR + sub,
             make_float4(fmaf(d.x * rs, g.x, b.x), fmaf(d.y * rs, g.y, b.y), fmaf(d.z * rs, g.z, b.z),
                         fmaf(d.w * rs, g.w, b.w)));
      if (sub == 0) { mean[r] = m; rstd[r] = rs; }
    }
  }
}

// part: [gridDim.x][2][W]  (dgamma partial, dbeta partial)
template <int LPR>
__global__ void __launch_bounds__(kLnThreads)
ln_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean,
              const float* __restrict__ rstd, const float* __restrict__ gamma, long n_rows,
              float* __restrict__ dx, float* __restrict__ part, const int* __restrict__ dy_index = nullptr,
              const int* __restrict__ n_rows_dev = nullptr) {
  constexpr int W = 4 * LPR, RPW = 32 / LPR;
  if (n_rows_dev != nullptr) {                // training step under a CUDA graph: the row count lives on the device
    const long nd = *n_rows_dev;
    n_rows = nd < 0 ? 0 : (nd < n_rows ? nd : n_rows);
  }
  __shared__ float4 s_dg[kLnThreads], s_db[kLnThreads];
  const int lane = threadIdx.x & 31, sub = lane % LPR, rw = lane / LPR;
  const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + sub);
  float4 dg = make_float4(0.f, 0.f, 0.f, 0.f), db = make_float4(0.f, 0.f, 0.f, 0.f);
  const long warps = (long)gridDim.x * (kLnThreads / 32);
  for (long r0 = ((long)blockIdx.x * (kLnThreads / 32) + (threadIdx.x >> 5)) * RPW; r0 < n_rows; r0 += warps * RPW) {
    const long r = r0 + rw;
    const bool ok = r < n_rows;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f), d = v;
    float m = 0.f, rs = 0.f;
    if (ok) {
      v = __ldcs(reinterpret_cast<const float4*>(x) + r * LPR + sub);
      const long rd = dy_index != nullptr ? (long)__ldg(dy_index + r) : r;      // dy rows gathered (edges: dagg[tgt])
      d = dy_index != nullptr ? __ldg(reinterpret_cast<const float4*>(dy) + rd * LPR + sub)
                              : __ldcs(reinterpret_cast<const float4*>(dy) + rd * LPR + sub);
      m = __ldg(mean + r);
      rs = __ldg(rstd + r);
    }
    const float4 xh = make_float4((v.x - m) * rs, (v.y - m) * rs, (v.z - m) * rs, (v.w - m) * rs);
    dg.x = fmaf(d.x, xh.x, dg.x); dg.y = fmaf(d.y, xh.y, dg.y); dg.z = fmaf(d.z, xh.z, dg.z); dg.w = fmaf(d.w, xh.w, dg.w);
    db.x += d.x; db.y += d.y; db.z += d.z; db.w += d.w;
    const float4 gd = make_float4(d.x * g.x, d.y * g.y, d.z * g.z, d.w * g.w);
    const float s1 = group_sum<LPR>((gd.x + gd.y) + (gd.z + gd.w)) * (1.0f / W);
    const float s2 = group_sum<LPR>((gd.x * xh.x + gd.y * xh.y) + (gd.z * xh.z + gd.w * xh.w)) * (1.0f / W);
    if (ok)
      __stcs(reinterpret_cast<float4*>(dx) + r * LPR + sub,
             make_float4(rs * (gd.x - s1 - xh.x * s2), rs * (gd.y - s1 - xh.y * s2), rs * (gd.z - s1 - xh.z * s2),
                         rs * (gd.w - s1 - xh.w * s2)));
  }
  s_dg[threadIdx.x] = dg;
  s_db[threadIdx.x] = db;
  __syncthreads();
  if (threadIdx.x < LPR) {                  // thread t owns columns 4t .. 4t+3: same `sub` every LPR threads
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), c = a;
    for (int k = threadIdx.x; k < kLnThreads; k += LPR) {
      const float4 p = s_dg[k], q = s_db[k];
      a.x += p.x; a.y += p.y; a.z += p.z; a.w += p.w;
      c.x += q.x; c.y += q.y; c.z += q.z; c.w += q.w;
    }
    float4* out = reinterpret_cast<float4*>(part + (long)blockIdx.x * 2 * W);
    out[threadIdx.x] = a;
    out[LPR + threadIdx.x] = c;
  }
}

// dgamma[c] = sum_b part[b][0][c], dbeta[c] = sum_b part[b][1][c]: 8 interleaved partial sums per column
// (block b goes to partial b % 8), combined in a fixed order
__global__ void __launch_bounds__(256)
ln_bwd_finish_kernel(const float* __restrict__ part, int n_blocks, int width, float* __restrict__ dgamma,
                     float* __restrict__ dbeta) {
  __shared__ float s[8][33];
  const int t = blockIdx.x * 32 + threadIdx.x;
  float a = 0.f;
  if (t < 2 * width)
    for (int b = threadIdx.y; b < n_blocks; b += 8) a += part[(long)b * 2 * width + t];
  s[threadIdx.y][threadIdx.x] = a;
  __syncthreads();
  if (threadIdx.y == 0 && t < 2 * width) {
    float r = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) r += s[k][threadIdx.x];
    if (t < width) dgamma[t] = r; else dbeta[t - width] = r;
  }
}

static int ln_blocks(long n_rows, int width) {
  const int rows_per_block = (kLnThreads / 32) * (32 / (width / 4));
  long b = (n_rows + rows_per_block - 1) / rows_per_block;
  if (b > kLnMaxBlocks) b = kLnMaxBlocks;
  return b < 1 ? 1 : (int)b;
}

// Scratch of sgnn_layer_norm_bwd_static: the per-CTA (dgamma, dbeta) partials of the widest supported row
// (kLnMaxBlocks * 2 * kLnStaticMaxWidth floats = 592 KiB), resident on every GPU that loads the library.
// Host code must not pass this name to a kernel (it is the host-side shadow of the symbol): the entry point
// resolves the device address with cudaGetSymbolAddress, once per device.
constexpr int kLnStaticMaxWidth = 128;
constexpr int kLnStaticMaxDevices = 64;
static __device__ __align__(256) float g_ln_bwd_scratch[kLnMaxBlocks * 2 * kLnStaticMaxWidth];
static float* g_ln_bwd_scratch_addr[kLnStaticMaxDevices];       // device address per device ordinal, NULL = not resolved

}  // namespace sgnn

using namespace sgnn;

extern "C" int sgnn_layer_norm_supported(int width) { return width == 32 || width == 64 || width == 128; }

extern "C" size_t sgnn_layer_norm_bwd_workspace_bytes(int width) {
  return sizeof(float) * (size_t)kLnMaxBlocks * 2 * (size_t)(width > 0 ? width : 1);
}

extern "C" int sgnn_layer_norm_fwd(const float* x, long n_rows, int width, const float* gamma, const float* beta,
                                   float eps, float* y, float* mean, float* rstd, void* stream) {
  SGNN_CHECK_ARG(sgnn_layer_norm_supported(width), "layer_norm_fwd: width must be 32, 64 or 128, got %d", width);
  if (n_rows <= 0) return SGNN_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = ln_blocks(n_rows, width);
  if (width == 32) ln_fwd_kernel<8><<<grid, kLnThreads, 0, st>>>(x, gamma, beta, eps, n_rows, y, mean, rstd);
  else if (width == 64) ln_fwd_kernel<16><<<grid, kLnThreads, 0, st>>>(x, gamma, beta, eps, n_rows, y, mean, rstd);
  else ln_fwd_kernel<32><<<grid, kLnThreads, 0, st>>>(x, gamma, beta, eps, n_rows, y, mean, rstd);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

extern "C" int sgnn_layer_norm_bwd(const float* dy, const float* x, const float* mean, const float* rstd,
                                   const float* gamma, long n_rows, int width, float* dx, float* dgamma,
                                   float* dbeta, void* workspace, size_t workspace_bytes, void* stream) {
  SGNN_CHECK_ARG(sgnn_layer_norm_supported(width), "layer_norm_bwd: width must be 32, 64 or 128, got %d", width);
  SGNN_CHECK_ARG(workspace_bytes >= sgnn_layer_norm_bwd_workspace_bytes(width), "layer_norm_bwd: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_rows <= 0) {
    SGNN_CUDA(cudaMemsetAsync(dgamma, 0, sizeof(float) * width, st));
    SGNN_CUDA(cudaMemsetAsync(dbeta, 0, sizeof(float) * width, st));
    return SGNN_OK;
  }
  const int grid = ln_blocks(n_rows, width);
  float* part = (float*)workspace;
  if (width == 32) ln_bwd_kernel<8><<<grid, kLnThreads, 0, st>>>(dy, x, mean, rstd, gamma, n_rows, dx, part);
  else if (width == 64) ln_bwd_kernel<16><<<grid, kLnThreads, 0, st>>>(dy, x, mean, rstd, gamma, n_rows, dx, part);
  else ln_bwd_kernel<32><<<grid, kLnThreads, 0, st>>>(dy, x, mean, rstd, gamma, n_rows, dx, part);
  SGNN_LAUNCH_CHECK();
  ln_bwd_finish_kernel<<<sgnn_ceil_div(2 * width, 32), dim3(32, 8), 0, st>>>(part, grid, width, dgamma, dbeta);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

// Variant of sgnn_layer_norm_bwd for callers that cannot allocate (CUDA graph capture, serving loops): the
// partials live in g_ln_bwd_scratch instead of a caller workspace.  Calls no cudaMalloc / cudaFree.  Every
// call on a device shares the array: calls that overlap on the GPU (two streams) race.  Nothing in it has to
// start at zero -- each CTA of a launch overwrites its own slot and the finish kernel reads those slots only.
extern "C" int sgnn_layer_norm_bwd_static(const float* dy, const float* x, const float* mean, const float* rstd,
                                          const float* gamma, long n_rows, int width, float* dx, float* dgamma,
                                          float* dbeta, void* stream) {
  SGNN_CHECK_ARG(width <= kLnStaticMaxWidth,
                 "layer_norm_bwd_static: width %d exceeds the static scratch (max width %d, %zu bytes)", width,
                 kLnStaticMaxWidth, sizeof(float) * (size_t)kLnMaxBlocks * 2 * kLnStaticMaxWidth);
  SGNN_CHECK_ARG(sgnn_layer_norm_supported(width), "layer_norm_bwd_static: width must be 32, 64 or 128, got %d", width);
  int dev = -1;
  SGNN_CUDA(cudaGetDevice(&dev));
  SGNN_CHECK_ARG(dev >= 0 && dev < kLnStaticMaxDevices, "layer_norm_bwd_static: device ordinal %d out of range", dev);
  float* part = __atomic_load_n(&g_ln_bwd_scratch_addr[dev], __ATOMIC_ACQUIRE);
  if (part == nullptr) {                       // set-up, once per device, with that device current
    SGNN_CUDA(cudaGetSymbolAddress((void**)&part, g_ln_bwd_scratch));
    SGNN_CHECK_ARG(part != nullptr, "layer_norm_bwd_static: cudaGetSymbolAddress returned NULL");
    __atomic_store_n(&g_ln_bwd_scratch_addr[dev], part, __ATOMIC_RELEASE);
  }
  return sgnn_layer_norm_bwd(dy, x, mean, rstd, gamma, n_rows, width, dx, dgamma, dbeta, part,
                             sizeof(float) * (size_t)kLnMaxBlocks * 2 * kLnStaticMaxWidth, stream);
}

// sgnn_layer_norm_bwd for the graph-captured training step: row r of the upstream gradient is dy[dy_index[r]]
// (NULL: dy[r]) -- the backward of the sum-aggregation's gather folded into this pass -- and the row count is
// min(max_rows, *n_rows_dev) read on the device (NULL: max_rows).
extern "C" int sgnn_layer_norm_bwd_rows(const float* dy, const int* dy_index, const float* x, const float* mean,
                                        const float* rstd, const float* gamma, long max_rows, const int* n_rows_dev,
                                        int width, float* dx, float* dgamma, float* dbeta, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  SGNN_CHECK_ARG(sgnn_layer_norm_supported(width), "layer_norm_bwd_rows: width must be 32, 64 or 128, got %d", width);
  SGNN_CHECK_ARG(workspace_bytes >= sgnn_layer_norm_bwd_workspace_bytes(width), "layer_norm_bwd_rows: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (max_rows <= 0) {
    SGNN_CUDA(cudaMemsetAsync(dgamma, 0, sizeof(float) * width, st));
    SGNN_CUDA(cudaMemsetAsync(dbeta, 0, sizeof(float) * width, st));
    return SGNN_OK;
  }
  const int grid = ln_blocks(max_rows, width);
  float* part = (float*)workspace;
  if (width == 32) ln_bwd_kernel<8><<<grid, kLnThreads, 0, st>>>(dy, x, mean, rstd, gamma, max_rows, dx, part, dy_index, n_rows_dev);
  else if (width == 64) ln_bwd_kernel<16><<<grid, kLnThreads, 0, st>>>(dy, x, mean, rstd, gamma, max_rows, dx, part, dy_index, n_rows_dev);
  else ln_bwd_kernel<32><<<grid, kLnThreads, 0, st>>>(dy, x, mean, rstd, gamma, max_rows, dx, part, dy_index, n_rows_dev);
  SGNN_LAUNCH_CHECK();
  ln_bwd_finish_kernel<<<sgnn_ceil_div(2 * width, 32), dim3(32, 8), 0, st>>>(part, grid, width, dgamma, dbeta);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
