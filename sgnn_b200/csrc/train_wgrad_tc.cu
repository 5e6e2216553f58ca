// Weight / bias gradients of the 64 x 64 Linear layers of the training step on the tensor cores
// (loss.backward() through build_mlp: sgnn/single_scale/graph_network.py:21-58, train.py:300-330 of the
// reference):   dW[m][k] = sum_r dy[r][m] x[r][k],   db[m] = sum_r dy[r][m].
//
// The contraction runs over the ROWS (edges or nodes), so both operands are "MN-major" for tcgen05: a row of
// dy / x is 64 consecutive floats = two 128-byte lines, one per 32-column sub-tile.  The only shared-memory layout
// tcgen05 takes for MN-major 32-bit operands is the 128-byte swizzle with a 32-byte base (descriptor layout type 1):
// 32-byte unit u of the 128-byte line of row r sits at unit u ^ (r & 3), atoms are 4 rows = 512 B (with the plain
// 128-byte swizzle the MMAs complete and write zeros).  A stage of the ring holds 32 rows as four sub-tiles
// [dy 0-31 | dy 32-63 | x 0-31 | x 32-63]; one tcgen05.mma (K = 8 rows) reads A = the four sub-tiles from dy on
// (M = 128: rows 64..127 of the accumulator are x^T x, never read -- M = 64 would cost the same tensor time) and
// B = the two x sub-tiles (N = 64).  The bias gradient is a second MMA of the same A against a tile of ones (N = 16).
// Operands go global -> shared with cp.async (zero-filled past the device-side row count) and never pass through
// registers in TF32 mode (the tensor core ignores the 13 low mantissa bits: truncation).  In split mode (fp32-grade
// products) every thread rewrites the chunks it loaded as hi = tf32(v) (rounded, in place) and lo = tf32(v - hi)
// (second image of the stage): hi * hi + lo * hi + hi * lo, errors unbiased -- but the tensor core adds into its fp32
// accumulator with truncation, so the error of a long row slab grows with its length (3e-7 of max|dW| at 1 k rows,
// 4.6e-6 at 173 k, FFMA: 3e-7): the training engine uses this kernel in TF32 mode only (train_engine._wgrad_flush).
// A CTA walks a contiguous slab of rows, accumulates in tensor memory, and writes one partial tile; a second
// kernel adds the partials in a fixed order (deterministic, no atomics).  Up to six products that share the row
// set go in one launch (grid.y).  Bytes: every operand row is read once: 512 B per row and product.
#include "common.cuh"
#include "params.h"

namespace sgnn {

constexpr int kWtRows = 32;                      // rows per ring stage
constexpr int kWtSub = kWtRows * 128;            // bytes of one 32-column sub-tile
constexpr int kWtRaw = 4 * kWtSub;               // dy and x sub-tiles of a stage
constexpr int kWtThreads = 128;
constexpr int kWtTmemCols = 128;                 // D: columns 0..63, bias: 64..79
constexpr int kWtPart = 64 * 64 + 64;            // floats of one partial (dW then db)
constexpr int kWtCtas = 2 * 148;                 // two CTAs per SM, one wave

template <bool SPLIT> struct WtCfg {
  static constexpr int stages = SPLIT ? 3 : 5;
  static constexpr int stage_bytes = SPLIT ? 2 * kWtRaw : kWtRaw;         // split: raw tiles, then their lo tiles
  static constexpr int smem_bytes = stages * stage_bytes + 1024 + 1024 + 128;   // + alignment slack, ones, barriers
};

// MN-major operand, SWIZZLE_128B with 32-byte base: 32-element groups along M / N are `lbo` bytes apart,
// 4-row K groups 512 B
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)1 << 61;
  return d;
}
__device__ __forceinline__ uint32_t umma_idesc_tf32_mn(int m, int n) {
  return umma_idesc_tf32(m, n) | (1u << 15) | (1u << 16);
}
__device__ __forceinline__ void wt_cp16(uint32_t dst, const float* src, bool ok) {
  const int n = ok ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}

template <bool SPLIT>
__global__ void __launch_bounds__(kWtThreads, 2)
wgrad_tc_kernel(const __grid_constant__ SgnnWgradBatch b, float* __restrict__ part) {
  constexpr int S = WtCfg<SPLIT>::stages;
  constexpr int SB = WtCfg<SPLIT>::stage_bytes;
  extern __shared__ char wt_smem_raw[];
  char* sm = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(wt_smem_raw) + 1023) & ~uintptr_t(1023));
  char* ones = sm + S * SB;
  uint64_t* bar = reinterpret_cast<uint64_t*>(ones + 1024);          // bar[s]: the MMAs that read stage s are done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + S);
  const int tid = threadIdx.x, warp = tid >> 5;
  const SgnnWgradItem& it = b.item[blockIdx.y];

  long n_rows = b.max_rows;
  if (b.n_rows_dev != nullptr) {
    const long nd = *b.n_rows_dev;
    n_rows = nd < 0 ? 0 : (nd < n_rows ? nd : n_rows);
  }
  const long tiles = (n_rows + kWtRows - 1) / kWtRows;
  const long per = (tiles + gridDim.x - 1) / gridDim.x;
  const long t_beg = (long)blockIdx.x * per;
  const long t_end = t_beg + per < tiles ? t_beg + per : tiles;
  const int n_my = t_end > t_beg ? (int)(t_end - t_beg) : 0;
  float* out = part + ((long)blockIdx.y * gridDim.x + blockIdx.x) * kWtPart;

  if (n_my == 0) {                                  // nothing to add: the finish kernel still reads this slot
    for (int i = tid; i < kWtPart; i += kWtThreads) out[i] = 0.f;
    return;
  }
  if (warp == 0) tmem_alloc(tmem_slot, kWtTmemCols);
  if (tid < S) mbar_init(&bar[tid], 1);
  reinterpret_cast<float2*>(ones)[tid] = make_float2(1.f, 1.f);      // 128 threads x 8 B = 1 KB of 1.0f
  mbar_fence_init();
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  // this thread's chunks of a stage: 16-byte chunk c16 of rows (tid >> 4) + 8 j, j < 4, of both operands
  const int c16 = tid & 15, rb = tid >> 4;
  const uint32_t off = (uint32_t)((c16 >> 3) * kWtSub + rb * 128 + ((((c16 & 7) >> 1) ^ (rb & 3)) << 5) + ((c16 & 1) << 4));
  const float* gdy = it.dy + c16 * 4;
  const float* gx = it.x + c16 * 4;
  const uint32_t sm_u = smem_u32(sm);

  auto load = [&](int t) {
    const uint32_t st = sm_u + (uint32_t)((t % S) * SB) + off;
    const long r0 = (t_beg + t) * kWtRows + rb;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long r = r0 + 8 * j;
      const bool ok = r < n_rows;
      wt_cp16(st + j * 1024, ok ? gdy + r * it.ld_dy : it.dy, ok);
      wt_cp16(st + 2 * kWtSub + j * 1024, ok ? gx + r * it.ld_x : it.x, ok);
    }
  };

#pragma unroll
  for (int s = 0; s < S - 1; ++s) {
    if (s < n_my) load(s);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  const uint32_t idesc = umma_idesc_tf32_mn(128, 64), idesc_b = umma_idesc_tf32_mn(128, 16);
  const uint64_t d_ones = umma_desc_mn_sw128(smem_u32(ones), 1024);

  for (int c = 0; c < n_my; ++c) {
    asm volatile("cp.async.wait_group %0;" ::"n"(S - 2) : "memory");
    char* stg = sm + (c % S) * SB;
    if (SPLIT) {                                   // hi = tf32(v) in place, lo = tf32(v - hi), for the chunks this thread loaded
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int o = 0; o < 2; ++o) {
          const uint32_t a = off + o * 2 * kWtSub + j * 1024;
          const float4 v = *reinterpret_cast<const float4*>(stg + a);
          float4 h, l;
          h.x = to_tf32(v.x); h.y = to_tf32(v.y); h.z = to_tf32(v.z); h.w = to_tf32(v.w);
          l.x = to_tf32(v.x - h.x); l.y = to_tf32(v.y - h.y); l.z = to_tf32(v.z - h.z); l.w = to_tf32(v.w - h.w);
          *reinterpret_cast<float4*>(stg + a) = h;
          *reinterpret_cast<float4*>(stg + kWtRaw + a) = l;
        }
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0) {
      tc_fence_after_sync();
      if (elect_one()) {
        const uint32_t s0 = sm_u + (uint32_t)((c % S) * SB);
#pragma unroll
        for (int ks = 0; ks < kWtRows / 8; ++ks) {
          const uint32_t acc = (c > 0 || ks > 0) ? 1u : 0u;
          const uint64_t a_hi = umma_desc_mn_sw128(s0 + ks * 1024, kWtSub);
          const uint64_t b_hi = umma_desc_mn_sw128(s0 + 2 * kWtSub + ks * 1024, kWtSub);
          umma_tf32(tmem, a_hi, b_hi, idesc, acc);
          umma_tf32(tmem + 64, a_hi, d_ones, idesc_b, acc);
          if (SPLIT) {
            const uint64_t a_lo = umma_desc_mn_sw128(s0 + kWtRaw + ks * 1024, kWtSub);
            const uint64_t b_lo = umma_desc_mn_sw128(s0 + kWtRaw + 2 * kWtSub + ks * 1024, kWtSub);
            umma_tf32(tmem, a_lo, b_hi, idesc, 1u);
            umma_tf32(tmem, a_hi, b_lo, idesc, 1u);
            umma_tf32(tmem + 64, a_lo, d_ones, idesc_b, 1u);
          }
        }
        tc_commit(&bar[c % S]);
      }
      __syncwarp();
    }
    const int cn = c + S - 1;
    if (cn < n_my) {
      if (c >= 1) mbar_wait(&bar[(c - 1) % S], (uint32_t)(((c - 1) / S) & 1));    // tile c - 1 has left that stage
      load(cn);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  mbar_wait(&bar[(n_my - 1) % S], (uint32_t)(((n_my - 1) / S) & 1));     // commits complete in order: all MMAs are done
  tc_fence_after_sync();

  if (warp < 2) {                                   // accumulator rows 0..63 = lanes 0..63
    float v[64], bv[32];
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
    tmem_ld64(lane_base, v);
    tmem_ld32(lane_base + 64, bv);
    float4* o = reinterpret_cast<float4*>(out + tid * 64);
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    out[64 * 64 + tid] = bv[0];
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, kWtTmemCols);
}

// dw / db of every item = sum over its splits, in a fixed order (32 lanes of splits per element, each adding its
// splits in increasing order, then a fixed tree over the lanes)
__global__ void __launch_bounds__(1024)
wgrad_tc_finish_kernel(const __grid_constant__ SgnnWgradBatch b, const float* __restrict__ part, int n_splits) {
  __shared__ float red[32][33];
  const SgnnWgradItem& it = b.item[blockIdx.y];
  const int e = blockIdx.x * 32 + threadIdx.x, sl = threadIdx.y;
  const float* p = part + (long)blockIdx.y * n_splits * kWtPart;
  float v = 0.f;
  if (e < kWtPart)
    for (int s = sl; s < n_splits; s += 32) v += __ldg(p + (long)s * kWtPart + e);
  red[sl][threadIdx.x] = v;
  __syncthreads();
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    if (sl < o) red[sl][threadIdx.x] += red[sl + o][threadIdx.x];
    __syncthreads();
  }
  if (sl == 0 && e < kWtPart) {
    if (e >= 64 * 64) { if (it.db != nullptr) it.db[e - 64 * 64] = red[0][threadIdx.x]; }
    else it.dw[(long)(e >> 6) * it.ld_dw + (e & 63)] = red[0][threadIdx.x] * it.scale;
  }
}

static int wt_splits(int n_items, long max_rows) {
  int s = kWtCtas / n_items;
  const long tiles = (max_rows + kWtRows - 1) / kWtRows;
  if (s > tiles) s = (int)tiles;
  return s < 1 ? 1 : s;
}

}  // namespace sgnn

using namespace sgnn;

extern "C" size_t sgnn_linear_wgrad_tc_workspace_bytes(int n_items, long max_rows) {
  if (n_items < 1 || n_items > SGNN_WGRAD_MAX_ITEMS || max_rows < 1) return 0;
  return sizeof(float) * (size_t)n_items * (size_t)wt_splits(n_items, max_rows) * kWtPart;
}

extern "C" int sgnn_linear_wgrad_tc(const SgnnWgradBatch* batch, void* stream) {
  SGNN_CHECK_ARG(batch != nullptr, "linear_wgrad_tc: null batch");
  const SgnnWgradBatch& b = *batch;
  SGNN_CHECK_ARG(b.n_items >= 1 && b.n_items <= SGNN_WGRAD_MAX_ITEMS && b.max_rows >= 1,
                 "linear_wgrad_tc: bad batch items=%d rows=%ld", b.n_items, b.max_rows);
  for (int i = 0; i < b.n_items; ++i) {
    const SgnnWgradItem& it = b.item[i];
    SGNN_CHECK_ARG(it.dy && it.x && it.dw, "linear_wgrad_tc: item %d has a null operand", i);
    SGNN_CHECK_ARG(it.ld_dy >= 64 && it.ld_x >= 64 && it.ld_dw >= 64 && (it.ld_dy & 3) == 0 && (it.ld_x & 3) == 0,
                   "linear_wgrad_tc: item %d: leading dimensions must be >= 64 and multiples of 4", i);
    SGNN_CHECK_ARG((((uintptr_t)it.dy | (uintptr_t)it.x) & 15) == 0, "linear_wgrad_tc: item %d: operands must be 16-byte aligned", i);
  }
  SGNN_CHECK_ARG(b.workspace != nullptr && b.workspace_bytes >= sgnn_linear_wgrad_tc_workspace_bytes(b.n_items, b.max_rows),
                 "linear_wgrad_tc: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  const int splits = wt_splits(b.n_items, b.max_rows);
  float* part = (float*)b.workspace;
  if (b.split) {
    SGNN_ENSURE_SMEM(wgrad_tc_kernel<true>, WtCfg<true>::smem_bytes);
    wgrad_tc_kernel<true><<<dim3(splits, b.n_items), kWtThreads, WtCfg<true>::smem_bytes, st>>>(b, part);
  } else {
    SGNN_ENSURE_SMEM(wgrad_tc_kernel<false>, WtCfg<false>::smem_bytes);
    wgrad_tc_kernel<false><<<dim3(splits, b.n_items), kWtThreads, WtCfg<false>::smem_bytes, st>>>(b, part);
  }
  SGNN_LAUNCH_CHECK();
  wgrad_tc_finish_kernel<<<dim3((kWtPart + 31) / 32, b.n_items), dim3(32, 32), 0, st>>>(b, part, splits);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
