// Weight / bias gradients of the Linear layers of the TRAINING path (loss.backward() through every
// build_mlp Linear: sgnn/single_scale/graph_network.py:21-58, train.py:300-330 of the reference).
//
//   dW[m][k] = sum_r dy[r][m] * x[r][k],     db[m] = sum_r dy[r][m]        (r over nodes or edges)
//
// These products are 64 x 64 outputs over 38 400 .. 200 000 rows.  cuBLAS picks ONE 64 x 64 (or two
// 128 x 32) output tiles for them and walks all rows inside one or two CTAs: torch.profiler shows 107 us
// and 224 us per call, 41 calls per step, 52 % of the GPU time of a training step (profiles/README.md,
// r01n).  Here the ROWS are split over the grid: a CTA owns a slab of rows and one 64 x 64 output tile,
// streams 32-row chunks of dy and x through a 3-stage cp.async ring and keeps an 8 x 4 (128 threads) or
// 8 x 8 (64 threads) block of the tile per thread in registers.  Per-slab partial tiles go to a workspace
// (L2-resident: <= 592 x 16 KB) and a second kernel adds them up in a fixed order: no atomics,
// bit-reproducible gradients.  The bias gradient rides along (column sums of the dy chunk in shared
// memory), which also removes ATen's separate reduction kernel per Linear.
//
// fp32 FFMA on purpose: the gradient parity tests pin the fp32 path against the reference in fp64.
// Bytes: reads 4 (M + K) per row once (M x K beyond 64 x 64: each tile re-reads its operand slabs);
// flops: 2 * 64 * 64 per row and tile -- 16 flops per byte, above the fp32 SIMT ridge of B200
// (~72 TFLOP/s / 6.5 TB/s = 11).  Measured (profiles/README.md r01p): 45 TFLOP/s at 1M rows, 33 at 190 k
// rows (47 us against 293 us for torch.mm + sum).  ncu at 190 k rows (profiles/r01p_linear_wgrad_190k_ncu_raw.csv):
// FMA pipe 56 % active, shared-memory wavefronts 38 % of peak, issue slots 42 % busy with 8 warps per SM
// (48 KB of shared memory per CTA -> 4 CTAs of 2 warps); top stalls are fixed-latency waits and instruction
// fetch.  Ring shapes (32 x 2, 16 x 4, 16 x 3) and 6 or 8 resident CTAs per SM all measured the same (r01p).  The product is issued
// as packed FFMA2 (same results as two FFMA, half the issue slots).
#include "common.cuh"

namespace sgnn {

constexpr int kWgTile = 64;                 // output tile is kWgTile x kWgTile
#ifndef WG_ROWS
#define WG_ROWS 32
#endif
#ifndef WG_STAGES
#define WG_STAGES 3
#endif
constexpr int kWgRows = WG_ROWS;            // rows per chunk
constexpr int kWgStages = WG_STAGES;
#ifndef WG_CTAS
#define WG_CTAS 4
#endif
constexpr int kWgMaxSplits = 148 * WG_CTAS;   // 48 KB of shared memory per CTA -> 4 CTAs per SM
constexpr int kWgChunksPerSplit = 2;

__device__ __forceinline__ void wg_cp16(float* dst, const float* src, bool ok) {
  const int n = ok ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void wg_cp4(float* dst, const float* src, bool ok) {
  const int n = ok ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void wg_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void wg_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// One operand chunk: rows [r0, r0 + 32) x columns [c0, c0 + wt) of a row-major [n_rows, ld] matrix into
// dst[32][64].  Rows past the end are zero-filled; columns >= wt are never written (zeroed once).
template <int NT>
__device__ __forceinline__ void wg_load_chunk(float* dst, const float* __restrict__ a, long ld, long n_rows,
                                              long r0, int c0, int wt, bool vec) {
  const int tid = threadIdx.x;
  if (vec) {
    const int q = wt >> 2;                            // float4 per row
    if (q == 16) {
#pragma unroll
      for (int i = 0; i < kWgRows * 16 / NT; ++i) {
        const int idx = tid + i * NT, row = idx >> 4, c4 = idx & 15;
        const bool ok = r0 + row < n_rows;
        wg_cp16(dst + row * kWgTile + c4 * 4, a + (ok ? (r0 + row) * ld + c0 + c4 * 4 : 0), ok);
      }
    } else {
      for (int idx = tid; idx < kWgRows * q; idx += NT) {
        const int row = idx / q, c4 = idx - row * q;
        const bool ok = r0 + row < n_rows;
        wg_cp16(dst + row * kWgTile + c4 * 4, a + (ok ? (r0 + row) * ld + c0 + c4 * 4 : 0), ok);
      }
    }
  } else {
    for (int idx = tid; idx < kWgRows * wt; idx += NT) {
      const int row = idx / wt, c = idx - row * wt;
      const bool ok = r0 + row < n_rows;
      wg_cp4(dst + row * kWgTile + c, a + (ok ? (r0 + row) * ld + c0 + c : 0), ok);
    }
  }
}

// NT = 64: thread (ty, tx) of an 8 x 8 thread grid owns output rows {4 ty + i, 32 + 4 ty + i} and columns
// {4 tx + j, 32 + 4 tx + j}, i, j < 4 (64 FFMA per 4 shared-memory loads).  NT = 128: an 8 x 16 grid, the
// same rows and columns {4 tx + j} (32 FFMA per 3 loads, but twice the warps to hide latencies).  Every
// shared-memory read is one conflict-free 16-byte access (x reads of a warp cover contiguous bytes, dy
// reads are broadcasts).  The bias gradient is a column sum of the dy chunk: thread t < 64 adds column t.
template <int NT, bool BIAS>
__global__ void __launch_bounds__(NT)
linear_wgrad_kernel(const float* __restrict__ dy, const float* __restrict__ x, long n_rows, int M, int K,
                    int chunks_per_split, int vec_dy, int vec_x, float* __restrict__ part, long part_stride,
                    long ld_dy, long ld_x, const int* __restrict__ n_rows_dev) {
  constexpr int JH = NT == 64 ? 2 : 1;              // groups of 4 columns per thread
  constexpr int TXB = NT == 64 ? 3 : 4;             // log2 of threads along the columns
  __shared__ __align__(16) float sm[kWgStages][2][kWgRows * kWgTile];
  const int tid = threadIdx.x, ty = tid >> TXB, tx = tid & ((1 << TXB) - 1);
  const int k_tiles = (K + kWgTile - 1) / kWgTile;
  const int mt = blockIdx.y / k_tiles, kt = blockIdx.y - mt * k_tiles;
  const int m0 = mt * kWgTile, k0 = kt * kWgTile;
  const int wm = M - m0 < kWgTile ? M - m0 : kWgTile, wk = K - k0 < kWgTile ? K - k0 : kWgTile;
  if (n_rows_dev != nullptr) {               // row count on the device: the slabs are cut here, over the whole grid
    const long nd = *n_rows_dev;
    n_rows = nd < 0 ? 0 : (nd < n_rows ? nd : n_rows);
    const long ch = (n_rows + kWgRows - 1) / kWgRows;
    chunks_per_split = (int)((ch + gridDim.x - 1) / gridDim.x);
  }
  const long n_chunks = (n_rows + kWgRows - 1) / kWgRows;
  const long c_beg = (long)blockIdx.x * chunks_per_split;
  long c_end = c_beg + chunks_per_split;
  if (c_end > n_chunks) c_end = n_chunks;
  const int n_my = c_end > c_beg ? (int)(c_end - c_beg) : 0;

  if (wm < kWgTile || wk < kWgTile) {   // narrow tile: columns past the operand widths must read as zero
    float4* z = reinterpret_cast<float4*>(&sm[0][0][0]);
    for (int i = tid; i < kWgStages * 2 * kWgRows * kWgTile / 4; i += NT) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
  }

  // accumulators as fp32 pairs: one packed FFMA2 (two IEEE fp32 FMAs, same results as two FFMA) per pair
  // halves the issue slots of the product, which leaves the shared-memory pipe as the only loaded unit
  float2 acc[8][2 * JH];
  float bsum = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 2 * JH; ++j) acc[i][j] = make_float2(0.f, 0.f);

#pragma unroll
  for (int s = 0; s < kWgStages - 1; ++s) {
    if (s < n_my) {
      wg_load_chunk<NT>(sm[s][0], dy, ld_dy, n_rows, (c_beg + s) * kWgRows, m0, wm, vec_dy);
      wg_load_chunk<NT>(sm[s][1], x, ld_x, n_rows, (c_beg + s) * kWgRows, k0, wk, vec_x);
    }
    wg_commit();
  }
  for (int c = 0; c < n_my; ++c) {
    wg_wait<kWgStages - 2>();
    __syncthreads();                       // chunk c has landed for everyone; chunk c - 1 is no longer read
    {
      const int cn = c + kWgStages - 1, st = cn % kWgStages;
      if (cn < n_my) {
        wg_load_chunk<NT>(sm[st][0], dy, ld_dy, n_rows, (c_beg + cn) * kWgRows, m0, wm, vec_dy);
        wg_load_chunk<NT>(sm[st][1], x, ld_x, n_rows, (c_beg + cn) * kWgRows, k0, wk, vec_x);
      }
      wg_commit();
    }
    const float* ds = sm[c % kWgStages][0] + ty * 4;
    const float* xs = sm[c % kWgStages][1] + tx * 4;
#pragma unroll 4
    for (int r = 0; r < kWgRows; ++r) {
      const float4 a0 = *reinterpret_cast<const float4*>(ds + r * kWgTile);
      const float4 a1 = *reinterpret_cast<const float4*>(ds + r * kWgTile + 32);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float2 b[2 * JH];
#pragma unroll
      for (int h = 0; h < JH; ++h) {
        const float4 bv = *reinterpret_cast<const float4*>(xs + r * kWgTile + 32 * h);
        b[2 * h] = make_float2(bv.x, bv.y);
        b[2 * h + 1] = make_float2(bv.z, bv.w);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float2 aa = make_float2(a[i], a[i]);
#pragma unroll
        for (int j = 0; j < 2 * JH; ++j) acc[i][j] = __ffma2_rn(aa, b[j], acc[i][j]);
      }
    }
    if (BIAS && tid < kWgTile) {
      const float* dc = sm[c % kWgStages][0] + tid;
#pragma unroll 8
      for (int r = 0; r < kWgRows; ++r) bsum += dc[r * kWgTile];
    }
  }
  wg_wait<0>();

  // partial tile of this row slab: part[split][M * K (+ M)]
  float* out = part + (long)blockIdx.x * part_stride;
  const bool vec_out = (K & 3) == 0 && (part_stride & 3) == 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + (i < 4 ? 4 * ty + i : 32 + 4 * ty + (i - 4));
    if (m >= M) continue;
#pragma unroll
    for (int h = 0; h < JH; ++h) {
      const int k = k0 + h * 32 + 4 * tx;
      if (vec_out && k + 3 < K) {
        *reinterpret_cast<float4*>(out + (long)m * K + k) =
            make_float4(acc[i][2 * h].x, acc[i][2 * h].y, acc[i][2 * h + 1].x, acc[i][2 * h + 1].y);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (k + j < K) out[(long)m * K + k + j] = (j & 1) ? acc[i][2 * h + (j >> 1)].y : acc[i][2 * h + (j >> 1)].x;
      }
    }
  }
  if (BIAS && kt == 0 && tid < kWgTile && m0 + tid < M) out[(long)M * K + m0 + tid] = bsum;
}

// out[e] = sum over splits of part[split][e], in a fixed order: 32 lanes of splits per element, each
// adding its splits in increasing order, then a fixed tree over the lanes.
__global__ void __launch_bounds__(1024)
linear_wgrad_finish_kernel(const float* __restrict__ part, long part_stride, int n_splits, int n_w, int n_b,
                           float* __restrict__ dw, float* __restrict__ db, int K = 0, long ld_dw = 0, float scale = 1.0f) {
  __shared__ float red[32][33];
  const int e = blockIdx.x * 32 + threadIdx.x, sl = threadIdx.y;
  const int total = n_w + n_b;
  float v = 0.f;
  if (e < total)
    for (int s = sl; s < n_splits; s += 32) v += __ldg(part + (long)s * part_stride + e);
  red[sl][threadIdx.x] = v;
  __syncthreads();
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    if (sl < o) red[sl][threadIdx.x] += red[sl + o][threadIdx.x];
    __syncthreads();
  }
  if (sl == 0 && e < total) {
    if (e >= n_w) db[e - n_w] = red[0][threadIdx.x];
    else if (K == 0) dw[e] = red[0][threadIdx.x];
    else dw[(long)(e / K) * ld_dw + (e % K)] = red[0][threadIdx.x] * scale;      // a column block of a wider gradient
  }
}

struct WgPlan {
  int splits, chunks_per_split, tiles;
  long part_stride;
};
static WgPlan wg_plan(long n_rows, int M, int K) {
  WgPlan p;
  p.tiles = ((M + kWgTile - 1) / kWgTile) * ((K + kWgTile - 1) / kWgTile);
  const long chunks = (n_rows + kWgRows - 1) / kWgRows;
  int max_splits = kWgMaxSplits / p.tiles;
  if (max_splits < 1) max_splits = 1;
  long splits = (chunks + kWgChunksPerSplit - 1) / kWgChunksPerSplit;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  p.chunks_per_split = (int)((chunks + splits - 1) / splits);
  if (p.chunks_per_split < 1) p.chunks_per_split = 1;
  p.splits = (int)((chunks + p.chunks_per_split - 1) / p.chunks_per_split);
  if (p.splits < 1) p.splits = 1;
  p.part_stride = (((long)M * K + M) + 3) & ~3L;
  return p;
}

}  // namespace sgnn

using namespace sgnn;

extern "C" size_t sgnn_linear_wgrad_workspace_bytes(long n_rows, int out_features, int in_features) {
  if (n_rows < 0 || out_features <= 0 || in_features <= 0) return 0;
  const WgPlan p = wg_plan(n_rows, out_features, in_features);
  return sizeof(float) * (size_t)p.splits * (size_t)p.part_stride;
}

extern "C" int sgnn_linear_wgrad(const float* dy, const float* x, long n_rows, int out_features, int in_features,
                                 float* dw, float* db, void* workspace, size_t workspace_bytes, void* stream) {
  const int M = out_features, K = in_features;
  SGNN_CHECK_ARG(M > 0 && K > 0 && n_rows >= 0, "linear_wgrad: bad shape rows=%ld out=%d in=%d", n_rows, M, K);
  SGNN_CHECK_ARG(workspace_bytes >= sgnn_linear_wgrad_workspace_bytes(n_rows, M, K), "linear_wgrad: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_rows == 0) {
    SGNN_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)M * K, st));
    if (db) SGNN_CUDA(cudaMemsetAsync(db, 0, sizeof(float) * M, st));
    return SGNN_OK;
  }
  const WgPlan p = wg_plan(n_rows, M, K);
  const int vec_dy = (M & 3) == 0 && ((uintptr_t)dy & 15) == 0;
  const int vec_x = (K & 3) == 0 && ((uintptr_t)x & 15) == 0;
  float* part = (float*)workspace;
  const dim3 grid(p.splits, p.tiles);
  // Thread layout (r01p, wgrad_micro): 8 x 8 outputs per thread is ~10 % faster once every CTA walks many
  // chunks (190 k rows: 47 vs 52 us; 1M rows: 183 vs 205 us); 8 x 4 with twice the warps wins on short
  // slabs, where latency rather than shared-memory bandwidth decides (38 400 rows: 18.3 vs 20.1 us).
  const bool wide = p.chunks_per_split >= 8;
#define WG_LAUNCH(NT, B)                                                                                          \
  linear_wgrad_kernel<NT, B><<<grid, NT, 0, st>>>(dy, x, n_rows, M, K, p.chunks_per_split, vec_dy, vec_x, part,   \
                                                  p.part_stride, M, K, nullptr)
  if (wide) { if (db) WG_LAUNCH(64, true); else WG_LAUNCH(64, false); }
  else      { if (db) WG_LAUNCH(128, true); else WG_LAUNCH(128, false); }
#undef WG_LAUNCH
  SGNN_LAUNCH_CHECK();
  const int n_w = M * K, n_b = db ? M : 0;
  linear_wgrad_finish_kernel<<<(n_w + n_b + 31) / 32, dim3(32, 32), 0, st>>>(part, p.part_stride, p.splits, n_w, n_b,
                                                                            dw, db);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

// sgnn_linear_wgrad for the graph-captured training step: operands with leading dimensions (column blocks of
// wider matrices), the row count min(max_rows, *n_rows_dev) read on the device (NULL: max_rows; the slabs are
// then cut on the device over the grid planned for max_rows), and the result written as scale * dW into a
// column block of a wider gradient (row stride ld_dw).  Workspace: sgnn_linear_wgrad_workspace_bytes(max_rows, ..).
extern "C" int sgnn_linear_wgrad_ex(const float* dy, long ld_dy, const float* x, long ld_x, long max_rows,
                                    const int* n_rows_dev, int out_features, int in_features, float* dw, long ld_dw,
                                    float scale, float* db, void* workspace, size_t workspace_bytes, void* stream) {
  const int M = out_features, K = in_features;
  SGNN_CHECK_ARG(M > 0 && K > 0 && max_rows > 0, "linear_wgrad_ex: bad shape rows=%ld out=%d in=%d", max_rows, M, K);
  SGNN_CHECK_ARG(ld_dy >= M && ld_x >= K && ld_dw >= K, "linear_wgrad_ex: leading dimensions too small");
  SGNN_CHECK_ARG(workspace_bytes >= sgnn_linear_wgrad_workspace_bytes(max_rows, M, K), "linear_wgrad_ex: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const WgPlan p = wg_plan(max_rows, M, K);
  const int vec_dy = (M & 3) == 0 && (ld_dy & 3) == 0 && ((uintptr_t)dy & 15) == 0;
  const int vec_x = (K & 3) == 0 && (ld_x & 3) == 0 && ((uintptr_t)x & 15) == 0;
  float* part = (float*)workspace;
  const dim3 grid(p.splits, p.tiles);
  const bool wide = p.chunks_per_split >= 32;      // planned for the capacity: the live row count is usually a fraction of it
#define WG_LAUNCH(NT, B)                                                                                          \
  linear_wgrad_kernel<NT, B><<<grid, NT, 0, st>>>(dy, x, max_rows, M, K, p.chunks_per_split, vec_dy, vec_x, part, \
                                                  p.part_stride, ld_dy, ld_x, n_rows_dev)
  if (wide) { if (db) WG_LAUNCH(64, true); else WG_LAUNCH(64, false); }
  else      { if (db) WG_LAUNCH(128, true); else WG_LAUNCH(128, false); }
#undef WG_LAUNCH
  SGNN_LAUNCH_CHECK();
  const int n_w = M * K, n_b = db ? M : 0;
  linear_wgrad_finish_kernel<<<(n_w + n_b + 31) / 32, dim3(32, 32), 0, st>>>(part, p.part_stride, p.splits, n_w, n_b,
                                                                            dw, db, K, ld_dw, scale);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
