// Training-path MLP blocks on the tensor cores: forward WITH the activations the backward needs, and the
// backward's chain of input-gradient products.  Reference: every build_mlp Sequential + LayerNorm of the
// single-scale network under loss.backward() (sgnn/single_scale/graph_network.py:21-58,86-96,139-147,
// 178-222; sgnn/single_scale/train.py:230-280).
//
//   forward   h1 = relu(A W1^T + b1 [+ P[tgt] + Q[src]]),  y2 = h1 W2^T + b2,  y = LN(y2) [+ res]
//             (n_linear 1: y = A W1^T + b1 -- the per-node P | Q product of an edge block)
//   backward  dpre = (dy W2) * (h1 > 0)  [n_linear 1: dpre = dy],  dx = dpre W1 (routed to two destinations)
//
// A CTA of 128 threads owns 128-row tiles (persistent loop, row count read on the device); thread t owns
// row t in every epilogue: TMEM lane t <-> row t, bias / ReLU / LayerNorm / the ReLU mask are thread-local.
// The products are tcgen05.mma kind::tf32 with the A operand (activations) in TENSOR MEMORY, the weight images in
// shared memory (K-major SWIZZLE_128B) and fp32 accumulators in TMEM.  Keeping the activations out of shared
// memory is what lets two CTAs share an SM with hi / lo operands (64 KB of images + one 32 KB staging tile each).
//
// Precision.  TF32 keeps 10 mantissa bits; the gradient parity tests pin the training step against the
// reference in fp64 with a bar of 20 x the reference's own fp32-vs-fp64 gap.  With split = 1 every operand is
// cut into hi = tf32(v) and lo = v - hi (exact in fp32) and every product runs as lo*hi + hi*lo + hi*hi on the
// tensor pipe: the dropped lo*lo term is 2^-22 relative, i.e. fp32-grade results at a third of the TF32 rate --
// still far above what the rows' HBM traffic allows (a 128 x 64 x 64 product is 0.14 us of tensor time per SM,
// its rows take ~1.5 us of the SM's HBM share).  split = 0 is plain TF32.
//
// Bytes per row (H = 64, fp32): forward of an edge block reads 256 (e0) + 512 (gathered P, Q rows, L2) and
// writes 256 x 3 (h1, y2, y) + 8; its backward reads 256 x 2 (dy, h1) and writes 256 (dpre) + 256 (de0 +=).
#include "mlp_tile.cuh"
#include "params.h"

namespace sgnn {

constexpr int kTrKChunk = 64;       // the A operand is staged in chunks of at most 64 columns
// 256 threads per CTA: threads t and t + 128 own the SAME row (TMEM lane t & 127 -- a warp reaches the lane
// quadrant 32 * (warp % 4)) and alternate over its 32-column chunks.  With 128 threads and one CTA per SM there
// was ONE warp per scheduler: ncu showed 15 % issue utilisation with every dependent-issue and shared-memory
// latency exposed (profiles/README.md, r02b).
constexpr int kTrThreads = 256;
// TMEM columns of a CTA: [0, 64) A hi, [64, 128) A lo, [128, 256) accumulators (two of 64, or one of 128)
constexpr int kTrTmemCols = 256;
constexpr int kTrAHi = 0, kTrALo = 64, kTrAcc = 128;
__device__ __forceinline__ int tr_row() { return (int)threadIdx.x & (kTileRows - 1); }
__device__ __forceinline__ int tr_half() { return (int)threadIdx.x >> 7; }
__device__ __forceinline__ int tr_log2(int x) { return 31 - __clz(x); }

struct TrMisc {
  uint64_t bar_w, bar_mma;
  uint32_t tmem_slot;
  int pad;
  int tgt[kTileRows], src[kTileRows];
  float2 red[2][kTileRows];           // LayerNorm: (mean, sum of squared deviations) of each half of a row
};

// ---- exact fp32 staging tile (K-major SWIZZLE_128B addressing: conflict-free 16-byte row-owner accesses) ----
// Everything a thread reads or writes per ROW goes through this tile: the global side is a cooperative loop in
// which a warp covers whole contiguous rows (4 L1 wavefronts per 512 bytes instead of 32 for row-owner accesses).
__device__ __forceinline__ void tr_row_store32(char* t, int col, const float* v) {
#pragma unroll
  for (int j = 0; j < 8; ++j)
    *reinterpret_cast<float4*>(t + sw128_chunk_offset(tr_row(), col / 4 + j, kTileRows)) =
        make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}
__device__ __forceinline__ void tr_row_load32(const char* t, int col, float* v) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 a = *reinterpret_cast<const float4*>(t + sw128_chunk_offset(tr_row(), col / 4 + j, kTileRows));
    v[4 * j] = a.x; v[4 * j + 1] = a.y; v[4 * j + 2] = a.z; v[4 * j + 3] = a.w;
  }
}
// tile[r][0:width) = a[ia[r]][c0 : c0 + width) (+ b[ib[r]][..]) for r < rows_valid; width is 32, 64 or 128 and
// c0, ld multiples of 4.  ia / ib: shared index arrays (NULL: row0 + r).  Columns [width, wpad) are zeroed.
__device__ __forceinline__ void tr_tile_gather(char* t, const float* __restrict__ a, const float* __restrict__ b, long ld,
                                               const int* ia, const int* ib, long row0, int c0, int width, int wpad,
                                               int rows_valid) {
  const int chunks = wpad >> 2, wch = width >> 2, sh = tr_log2(chunks);
  const int total = rows_valid * chunks;
  constexpr int U = 4;
  for (int idx0 = threadIdx.x; idx0 < total; idx0 += U * kTrThreads) {
    float4 va[U], vb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = idx0 + u * kTrThreads;
      va[u] = vb[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      const int r = idx >> sh, c = idx & (chunks - 1);
      if (idx < total && c < wch) {
        va[u] = __ldg(reinterpret_cast<const float4*>(a + (ia ? (long)ia[r] : row0 + r) * ld + c0) + c);
        if (b != nullptr) vb[u] = __ldg(reinterpret_cast<const float4*>(b + (ib ? (long)ib[r] : row0 + r) * ld + c0) + c);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = idx0 + u * kTrThreads;
      if (idx < total) {
        const int r = idx >> sh, c = idx & (chunks - 1);
        *reinterpret_cast<float4*>(t + sw128_chunk_offset(r, c, kTileRows)) =
            make_float4(va[u].x + vb[u].x, va[u].y + vb[u].y, va[u].z + vb[u].z, va[u].w + vb[u].w);
      }
    }
  }
}
// dst[row0 + r][0:width) (=|+=) tile[r] (+ add[row0 + r]); coalesced; width is 32, 64 or 128
__device__ __forceinline__ void tr_tile_store(float* __restrict__ dst, long ld, long row0, const char* t, int width,
                                              int rows_valid, bool accumulate, const float* __restrict__ add, long ld_add) {
  const int chunks = width >> 2, sh = tr_log2(chunks);
  const int total = rows_valid * chunks;
  for (int idx = threadIdx.x; idx < total; idx += kTrThreads) {
    const int r = idx >> sh, c = idx & (chunks - 1);
    float4 v = *reinterpret_cast<const float4*>(t + sw128_chunk_offset(r, c, kTileRows));
    if (add != nullptr) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(add + (row0 + r) * ld_add) + c);
      v = make_float4(v.x + a.x, v.y + a.y, v.z + a.z, v.w + a.w);
    }
    float4* d = reinterpret_cast<float4*>(dst + (row0 + r) * ld) + c;
    if (accumulate) {
      const float4 z = *d;
      v = make_float4(v.x + z.x, v.y + z.y, v.z + z.z, v.w + z.w);
    }
    *d = v;
  }
}

// ---- tensor memory ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t tr_lane_base() { return (uint32_t)(((threadIdx.x >> 5) & 3) * 32) << 16; }
// 32 accumulator columns of this thread's row (lane quadrant of the warp: warp % 4)
__device__ __forceinline__ void tr_acc_load32(const TileAcc<true>& acc, int col, float* v) {
  tmem_ld32(acc.tmem_base + tr_lane_base() + (uint32_t)col, v);
}
// 32 columns [col, col + 32) of this thread's row of the A operand: hi = tf32(v) (and lo = v - hi, exact)
template <bool SPLIT>
__device__ __forceinline__ void tr_a_store32(const TileAcc<true>& acc, int col, const float* v) {
  float h[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) h[j] = to_tf32(v[j]);
  tmem_st32(acc.tmem_base + tr_lane_base() + (uint32_t)(kTrAHi + col), h);
  if constexpr (SPLIT) {
#pragma unroll
    for (int j = 0; j < 32; ++j) h[j] = v[j] - h[j];
    tmem_st32(acc.tmem_base + tr_lane_base() + (uint32_t)(kTrALo + col), h);
  }
}
// staging tile columns [0, kw) of every row -> A operand (rows >= rows_valid and columns >= kw as zeros), times scale
template <bool SPLIT>
__device__ __forceinline__ void tr_a_from_tile(const TileAcc<true>& acc, const char* t, int kw, bool valid, float scale) {
  for (int c = tr_half() * 32; c < kw; c += 64) {
    float v[32];
    tr_row_load32(t, c, v);
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = valid ? v[j] * scale : 0.f;
    tr_a_store32<SPLIT>(acc, c, v);
  }
}
// the A operand is complete: tensor-memory stores -> visible to the MMAs of the elected thread
__device__ __forceinline__ void tr_a_publish() {
  tmem_st_wait();
  tc_fence_before_sync();
  __syncthreads();
}

// D[128 x n] (+)= A[128 x k] W[n x k]^T, A in tensor memory; with SPLIT the two cross terms go first.
// All threads call; completion through gemm_wait<true>.
__device__ __forceinline__ void tr_umma_ts(uint32_t d, uint32_t a, uint32_t w_base, int k, int n, bool accumulate) {
  const uint32_t idesc = umma_idesc_tf32(kTileRows, n);
  uint32_t acc = accumulate ? 1u : 0u;
  for (int kb = 0; kb < k / kKBlock; ++kb) {
    const uint32_t w_kb = w_base + kb * n * kKBlockBytes;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      umma_tf32_ts(d, a + kb * kKBlock + j * 8, umma_desc_sw128(w_kb + j * 32), idesc, acc);
      acc = 1u;
    }
  }
}
template <bool SPLIT>
__device__ __forceinline__ void tr_gemm(TileAcc<true>& acc, const char* w_hi, const char* w_lo, int k, int n,
                                        int acc_col, bool accumulate) {
  if (threadIdx.x < 32) {
    tc_fence_after_sync();
    if (elect_one()) {
      const uint32_t d = acc.tmem_base + (uint32_t)acc_col;
      if constexpr (SPLIT) {
        tr_umma_ts(d, acc.tmem_base + kTrALo, smem_u32(w_hi), k, n, accumulate);
        tr_umma_ts(d, acc.tmem_base + kTrAHi, smem_u32(w_lo), k, n, true);
        tr_umma_ts(d, acc.tmem_base + kTrAHi, smem_u32(w_hi), k, n, true);
      } else {
        tr_umma_ts(d, acc.tmem_base + kTrAHi, smem_u32(w_hi), k, n, accumulate);
      }
    }
    __syncwarp();
  }
}

__device__ __forceinline__ int tr_rows(int max_rows, const int* n_rows_dev) {
  if (n_rows_dev == nullptr) return max_rows;
  const int n = *n_rows_dev;
  return n < 0 ? 0 : (n > max_rows ? max_rows : n);
}

// stage `bytes` (multiple of 16) with bulk copies on a barrier whose expect_tx the caller has set
__device__ __forceinline__ void tr_stage(char* dst, const float* src, int bytes, uint64_t* bar) {
  for (int off = 0; off < bytes; off += 32768)
    tma_bulk_g2s(dst + off, reinterpret_cast<const char*>(src) + off, (uint32_t)(bytes - off < 32768 ? bytes - off : 32768),
                 bar);
}

struct TrFwdSmem {
  int w1, w2, g, g_bytes, misc, total;
};
__host__ __device__ inline TrFwdSmem tr_fwd_smem(int kpad, int hidden, int out, int n_linear, int split) {
  TrFwdSmem s;
  const int out_pad = (out + 15) & ~15;
  s.w1 = hidden * kpad * 4;
  s.w2 = n_linear == 2 ? out_pad * hidden * 4 : 0;
  s.g = (s.w1 + s.w2) * (split ? 2 : 1);
  int gw = hidden > kTrKChunk ? hidden : kTrKChunk;          // input chunk, gathered terms, h1, y2 / y, [P | Q]
  s.g_bytes = kTileRows * gw * 4;
  s.misc = s.g + s.g_bytes;
  s.total = s.misc + (int)sizeof(TrMisc) + 1024;
  return s;
}

template <bool SPLIT>
__global__ void __launch_bounds__(kTrThreads, 2)
train_fwd_kernel(const SgnnTrainFwd p) {
  char* smem = smem_base_1024();
  const TrFwdSmem L = tr_fwd_smem(p.kpad, p.hidden, p.out, p.n_linear, SPLIT);
  const int out_pad = (p.out + 15) & ~15;
  char* w1h = smem;
  char* w1l = SPLIT ? w1h + L.w1 : w1h;
  char* w2h = w1h + L.w1 * (SPLIT ? 2 : 1);
  char* w2l = SPLIT ? w2h + L.w2 : w2h;
  char* g_t = smem + L.g;
  TrMisc* misc = reinterpret_cast<TrMisc*>(smem + L.misc);
  const int tid = threadIdx.x, row = tr_row(), half = tr_half();

  if (tid == 0) {
    mbar_init(&misc->bar_w, 1);
    mbar_fence_init();
    mbar_arrive_expect_tx(&misc->bar_w, (uint32_t)((L.w1 + L.w2) * (SPLIT ? 2 : 1)));
    tr_stage(w1h, p.w1_hi, L.w1, &misc->bar_w);
    if (SPLIT) tr_stage(w1l, p.w1_lo, L.w1, &misc->bar_w);
    if (L.w2) {
      tr_stage(w2h, p.w2_hi, L.w2, &misc->bar_w);
      if (SPLIT) tr_stage(w2l, p.w2_lo, L.w2, &misc->bar_w);
    }
  }
  TileAcc<true> acc;
  const int acc0 = kTrAcc, acc1 = kTrAcc + 64;                   // n_linear 1: one accumulator of up to 128 columns
  acc_setup<true>(acc, kTrTmemCols, &misc->bar_mma, &misc->tmem_slot, nullptr);
  mbar_wait(&misc->bar_w, 0);

  const int rows = tr_rows(p.max_rows, p.n_rows_dev);
  const int n_tiles = (rows + kTileRows - 1) / kTileRows;
  const bool gathered = p.add_p != nullptr;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long row0 = (long)tile * kTileRows;
    const int rows_valid = rows - row0 < kTileRows ? (int)(rows - row0) : kTileRows;
    const long r = row0 + row;
    const bool valid = row < rows_valid;
    int tg = 0, sr = 0;
    if (valid && p.tgt != nullptr) { tg = __ldg(p.tgt + r); sr = __ldg(p.src + r); }
    if (half == 0) { misc->tgt[row] = tg; misc->src[row] = sr; }

    // ---- first Linear, K staged in chunks of 64 columns: rows -> staging tile -> A operand in tensor memory ----
    for (int kc = 0; kc < p.kpad; kc += kTrKChunk) {
      const int kw = p.kpad - kc < kTrKChunk ? p.kpad - kc : kTrKChunk;
      if (p.prologue == SGNN_TPRO_EDGE_FEATURES) {
        if (half == 0) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
          if (valid) {
            float sq = 0.f;
            for (int d = 0; d < p.dim; ++d) {
              const float f = __fdiv_rn(__fsub_rn(__ldg(p.pos + (long)sr * p.pos_stride + d),
                                                  __ldg(p.pos + (long)tg * p.pos_stride + d)), p.radius);
              v[d] = f;
              sq = __fadd_rn(sq, __fmul_rn(f, f));
            }
            v[p.dim] = sqrtf(sq);
            if (p.feat_out != nullptr)
              for (int d = 0; d <= p.dim; ++d) p.feat_out[r * (p.dim + 1) + d] = v[d];
          }
          tr_a_store32<SPLIT>(acc, 0, v);
        }
      } else {
        if (p.prologue == SGNN_TPRO_ROWS2 && kc >= p.k0) {
          const int w = p.k1 - (kc - p.k0) < kw ? p.k1 - (kc - p.k0) : kw;
          tr_tile_gather(g_t, p.a1, nullptr, p.ld1, nullptr, nullptr, row0, kc - p.k0, w, kw, rows_valid);
        } else {
          const int w = p.k0 - kc < kw ? p.k0 - kc : kw;
          tr_tile_gather(g_t, p.a0, nullptr, p.ld0, nullptr, nullptr, row0, kc, w, kw, rows_valid);
        }
        __syncthreads();
        tr_a_from_tile<SPLIT>(acc, g_t, kw, valid, (p.prologue == SGNN_TPRO_ROWS2 && kc >= p.k0) ? 1.0f : p.scale0);
      }
      tr_a_publish();                       // (also: everybody has read the staging tile)
      const int woff = (kc / kKBlock) * p.hidden * kKBlockBytes;
      tr_gemm<SPLIT>(acc, w1h + woff, w1l + woff, kw, p.hidden, acc0, kc > 0);
      // while the last chunk's product runs: receiver + sender terms of the rows, gathered with whole-row requests
      if (gathered && kc + kTrKChunk >= p.kpad)
        tr_tile_gather(g_t, p.add_p, p.add_q, p.ld_add, misc->tgt, misc->src, row0, 0, p.hidden, p.hidden, rows_valid);
      gemm_wait<true>(acc);
    }
    if (gathered) __syncthreads();

    // ---- bias (+ gathered terms), ReLU; h1 exact into the staging tile, hi / lo into the A operand ----
    for (int c = half * 32; c < p.hidden; c += 64) {
      float v[32];
      tr_acc_load32(acc, acc0 + c, v);
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (p.b1 != nullptr && c + j < p.b1_cols) v[j] += __ldg(p.b1 + c + j);
      if (gathered && valid) {
        float gsum[32];
        tr_row_load32(g_t, c, gsum);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] += gsum[j];
      }
      if (p.n_linear == 2) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = valid ? fmaxf(v[j], 0.f) : 0.f;
        tr_a_store32<SPLIT>(acc, c, v);
      }
      tr_row_store32(g_t, c, v);
    }
    if (p.n_linear != 2) {
      tc_fence_before_sync();
      __syncthreads();
      tr_tile_store(p.y, p.ld_y, row0, g_t, p.hidden, rows_valid, false, nullptr, 0);
      __syncthreads();                      // staging tile and accumulator free for the next tile
      continue;
    }
    tr_a_publish();
    tr_gemm<SPLIT>(acc, w2h, w2l, p.hidden, out_pad, acc1, false);
    tr_tile_store(p.h1, p.hidden, row0, g_t, p.hidden, rows_valid, false, nullptr, 0);     // while the second product runs
    gemm_wait<true>(acc);
    __syncthreads();                        // the staging tile is free again

    // ---- second Linear's bias, LayerNorm (statistics kept), output ----
    if (p.has_ln) {                         // out is 64 or 128: each half of the row owns out / 2 columns
      const float n_loc = 0.5f * (float)p.out;
      float s4[4] = {0.f, 0.f, 0.f, 0.f};
      for (int c = half * 32; c < p.out; c += 64) {
        float v[32];
        tr_acc_load32(acc, acc1 + c, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) s4[j & 3] += v[j] + __ldg(p.b2 + c + j);
      }
      const float m_loc = ((s4[0] + s4[1]) + (s4[2] + s4[3])) / n_loc;
      float q4[4] = {0.f, 0.f, 0.f, 0.f};
      for (int c = half * 32; c < p.out; c += 64) {
        float v[32];
        tr_acc_load32(acc, acc1 + c, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float d = v[j] + __ldg(p.b2 + c + j) - m_loc;
          q4[j & 3] = fmaf(d, d, q4[j & 3]);
        }
        tr_row_store32(g_t, c, v);          // (y2 without its bias, for now)
      }
      misc->red[half][row] = make_float2(m_loc, (q4[0] + q4[1]) + (q4[2] + q4[3]));
      __syncthreads();
      const float2 oth = misc->red[half ^ 1][row];          // pairwise combination of the two halves' moments
      const float dm = oth.x - m_loc;
      const float mean = 0.5f * (m_loc + oth.x);
      const float m2 = ((q4[0] + q4[1]) + (q4[2] + q4[3])) + oth.y + dm * dm * (0.5f * n_loc);
      const float rstd = 1.0f / sqrtf(m2 / (float)p.out + 1e-5f);
      if (valid && half == 0) { p.mean[r] = mean; p.rstd[r] = rstd; }
      float y[2][32];                       // this thread's chunks of the output (out <= 128: at most two)
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int c = half * 32 + 64 * k;
        if (c < p.out) {
          float v[32];
          tr_row_load32(g_t, c, v);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            v[j] += __ldg(p.b2 + c + j);
            y[k][j] = (v[j] - mean) * rstd * __ldg(p.gamma + c + j) + __ldg(p.beta + c + j);
          }
          tr_row_store32(g_t, c, v);        // y2, the LayerNorm's input
        }
      }
      tc_fence_before_sync();
      __syncthreads();
      tr_tile_store(p.y2, p.out, row0, g_t, p.out, rows_valid, false, nullptr, 0);
      __syncthreads();
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int c = half * 32 + 64 * k;
        if (c < p.out) tr_row_store32(g_t, c, y[k]);
      }
      __syncthreads();
      tr_tile_store(p.y, p.ld_y, row0, g_t, p.out, rows_valid, false, p.res, p.ld_res);     // (+ residual on the way out)
    } else {
      for (int c = half * 32; c < out_pad; c += 64) {
        float v[32];
        tr_acc_load32(acc, acc1 + c, v);
        if (valid)
          for (int j = 0; j < 32 && c + j < p.out; ++j) p.y[r * p.ld_y + c + j] = v[j] + __ldg(p.b2 + c + j);
      }
      tc_fence_before_sync();
    }
    __syncthreads();
  }
  acc_teardown<true>(acc);
}

// ------------------------------------------------------------------------------------------ backward
struct TrBwdSmem {
  int w2t, w1t, g, g_bytes, misc, total;
};
__host__ __device__ inline TrBwdSmem tr_bwd_smem(int k_dy, int n_linear, int hidden, int want_dx, int kx, int split) {
  TrBwdSmem s;
  const int kdp = (k_dy + 31) & ~31;
  const int k2 = n_linear == 2 ? hidden : kdp;             // K of the dx product
  s.w2t = n_linear == 2 ? hidden * kdp * 4 : 0;
  s.w1t = want_dx ? kx * k2 * 4 : 0;
  s.g = (s.w2t + s.w1t) * (split ? 2 : 1);
  s.g_bytes = kTileRows * kTrKChunk * 4;                   // dy chunk, h1 tile, dpre, one 64-column half of dx
  s.misc = s.g + s.g_bytes;
  s.total = s.misc + (int)sizeof(TrMisc) + 1024;
  return s;
}

template <bool SPLIT>
__global__ void __launch_bounds__(kTrThreads, 2)
train_bwd_kernel(const SgnnTrainBwd p) {
  char* smem = smem_base_1024();
  const TrBwdSmem L = tr_bwd_smem(p.k_dy, p.n_linear, p.hidden, p.want_dx, p.kx, SPLIT);
  const int kdp = (p.k_dy + 31) & ~31;
  char* w2h = smem;
  char* w2l = SPLIT ? w2h + L.w2t : w2h;
  char* w1h = w2h + L.w2t * (SPLIT ? 2 : 1);
  char* w1l = SPLIT ? w1h + L.w1t : w1h;
  char* g_t = smem + L.g;
  TrMisc* misc = reinterpret_cast<TrMisc*>(smem + L.misc);
  const int tid = threadIdx.x, row = tr_row(), half = tr_half();

  if (tid == 0) {
    mbar_init(&misc->bar_w, 1);
    mbar_fence_init();
    mbar_arrive_expect_tx(&misc->bar_w, (uint32_t)((L.w2t + L.w1t) * (SPLIT ? 2 : 1)));
    if (L.w2t) {
      tr_stage(w2h, p.w2t_hi, L.w2t, &misc->bar_w);
      if (SPLIT) tr_stage(w2l, p.w2t_lo, L.w2t, &misc->bar_w);
    }
    if (L.w1t) {
      tr_stage(w1h, p.w1t_hi, L.w1t, &misc->bar_w);
      if (SPLIT) tr_stage(w1l, p.w1t_lo, L.w1t, &misc->bar_w);
    }
  }
  TileAcc<true> acc;
  const int acch = kTrAcc, accx = kTrAcc;                        // dh is consumed before the dx product starts: same columns
  acc_setup<true>(acc, kTrTmemCols, &misc->bar_mma, &misc->tmem_slot, nullptr);
  mbar_wait(&misc->bar_w, 0);

  const int rows = tr_rows(p.max_rows, p.n_rows_dev);
  const int n_tiles = (rows + kTileRows - 1) / kTileRows;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long row0 = (long)tile * kTileRows;
    const int rows_valid = rows - row0 < kTileRows ? (int)(rows - row0) : kTileRows;
    const bool valid = row < rows_valid;

    // ---- first product: dh = dy W2 (n_linear 2) or dx = dy W1 (n_linear 1); K = width of dy, in chunks of 64 ----
    const int n1 = p.n_linear == 2 ? p.hidden : p.kx;
    for (int kc = 0; kc < kdp; kc += kTrKChunk) {
      const int kw = kdp - kc < kTrKChunk ? kdp - kc : kTrKChunk;
      const int w = p.k_dy - kc < kw ? p.k_dy - kc : kw;
      tr_tile_gather(g_t, p.dy, nullptr, p.ld_dy, nullptr, nullptr, row0, kc, w, kw, rows_valid);
      __syncthreads();
      tr_a_from_tile<SPLIT>(acc, g_t, kw, valid, 1.0f);
      tr_a_publish();
      const int woff = (kc / kKBlock) * n1 * kKBlockBytes;
      if (p.n_linear == 2) {
        tr_gemm<SPLIT>(acc, w2h + woff, w2l + woff, kw, n1, acch, kc > 0);
        if (kc + kTrKChunk >= kdp)          // while the product runs: this tile of h1, whole rows per request
          tr_tile_gather(g_t, p.h1, nullptr, p.hidden, nullptr, nullptr, row0, 0, p.hidden, p.hidden, rows_valid);
      } else {
        tr_gemm<SPLIT>(acc, w1h + woff, w1l + woff, kw, n1, accx, kc > 0);
      }
      gemm_wait<true>(acc);
    }
    if (p.n_linear == 2) {
      __syncthreads();
      // ---- ReLU mask from the h1 tile; dpre exact into the staging tile, hi / lo into the A operand ----
      for (int c = half * 32; c < p.hidden; c += 64) {
        float v[32], h[32];
        tr_acc_load32(acc, acch + c, v);
        tr_row_load32(g_t, c, h);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = (valid && h[j] > 0.f) ? v[j] : 0.f;
        if (p.want_dx) tr_a_store32<SPLIT>(acc, c, v);
        tr_row_store32(g_t, c, v);
      }
      if (p.want_dx) tmem_st_wait();
      tc_fence_before_sync();
      __syncthreads();
      if (p.want_dx) tr_gemm<SPLIT>(acc, w1h, w1l, p.hidden, p.kx, accx, false);
      tr_tile_store(p.dpre, p.hidden, row0, g_t, p.hidden, rows_valid, false, nullptr, 0);
      if (!p.want_dx) {
        __syncthreads();
        continue;
      }
      gemm_wait<true>(acc);
    }
    // ---- dx by 64-column halves through the staging tile: [0, k0) -> dx0, [k0, k0 + k1) -> dx1 ----
    for (int c0 = 0; c0 < p.kx; c0 += kTrKChunk) {
      __syncthreads();                      // the staging tile is free
      const int cw = p.kx - c0 < kTrKChunk ? p.kx - c0 : kTrKChunk;
      for (int c = half * 32; c < cw; c += 64) {
        float v[32];
        tr_acc_load32(acc, accx + c0 + c, v);
        tr_row_store32(g_t, c, v);
      }
      tc_fence_before_sync();
      __syncthreads();
      if (c0 < p.k0) tr_tile_store(p.dx0 + c0, p.ld_dx0, row0, g_t, cw, rows_valid, p.acc0 != 0, nullptr, 0);
      else tr_tile_store(p.dx1 + (c0 - p.k0), p.ld_dx1, row0, g_t, cw, rows_valid, p.acc1 != 0,
                         p.add1 ? p.add1 + (c0 - p.k0) : nullptr, p.ld_add1);
    }
    __syncthreads();
  }
  acc_teardown<true>(acc);
}

// ------------------------------------------------------------------------------------------ weight images
// All images of a training step in ONE launch (the weights change every step): blockIdx.y = descriptor.
__global__ void __launch_bounds__(256)
train_pack_kernel(const SgnnPackDesc* __restrict__ descs, float* __restrict__ img_hi, float* __restrict__ img_lo) {
  const SgnnPackDesc d = descs[blockIdx.y];
  const int total = d.n * d.k;
  for (int idx = blockIdx.x * 256 + threadIdx.x; idx < total; idx += gridDim.x * 256) {
    const int r = idx / d.k, c = idx - r * d.k;
    const float v = __ldg(d.src + r * d.stride_n + c * d.stride_k) * d.scale;
    const float h = to_tf32(v);
    const uint32_t off = sw128_offset(d.row0 + r, d.col0 + c, d.n_pad) >> 2;
    img_hi[d.dst_off + off] = h;
    img_lo[d.dst_off + off] = v - h;
  }
}

static int tr_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      sms = 148;
  }
  return sms;
}

}  // namespace sgnn

using namespace sgnn;

static bool al16(const void* p) { return ((uintptr_t)p & 15) == 0; }

extern "C" int sgnn_train_pack(const SgnnPackDesc* descs_dev, int n_descs, float* img_hi, float* img_lo, void* stream) {
  if (n_descs <= 0) return SGNN_OK;
  train_pack_kernel<<<dim3(16, n_descs), 256, 0, (cudaStream_t)stream>>>(descs_dev, img_hi, img_lo);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

extern "C" int sgnn_train_mlp_fwd(const SgnnTrainFwd* p, void* stream) {
  SGNN_CHECK_ARG(p->n_linear == 1 || p->n_linear == 2, "train_mlp_fwd: n_linear must be 1 or 2");
  SGNN_CHECK_ARG(p->kpad > 0 && p->kpad % 32 == 0 && p->kpad <= 256, "train_mlp_fwd: bad kpad %d", p->kpad);
  SGNN_CHECK_ARG(p->hidden >= 16 && p->hidden % 32 == 0 && p->hidden <= 128, "train_mlp_fwd: hidden %d must be 32, 64, 96 or 128",
                 p->hidden);
  SGNN_CHECK_ARG(p->n_linear == 1 || (p->out >= 1 && p->out <= 128), "train_mlp_fwd: bad out %d", p->out);
  SGNN_CHECK_ARG(!p->has_ln || (p->n_linear == 2 && p->out == 64), "train_mlp_fwd: LayerNorm needs out 64");
  SGNN_CHECK_ARG(p->hidden == 64 || (p->n_linear == 1 && p->hidden == 128), "train_mlp_fwd: hidden must be 64 (a single Linear: 64 or 128)");
  SGNN_CHECK_ARG(p->prologue == SGNN_TPRO_EDGE_FEATURES || (p->a0 && al16(p->a0) && p->ld0 % 4 == 0 && p->k0 % 4 == 0),
                 "train_mlp_fwd: input rows must be 16-byte aligned, leading dimension and width multiples of 4");
  SGNN_CHECK_ARG(p->prologue != SGNN_TPRO_ROWS2 || (p->a1 && al16(p->a1) && p->ld1 % 4 == 0 && p->k1 % 4 == 0),
                 "train_mlp_fwd: second input must be 16-byte aligned, leading dimension and width multiples of 4");
  SGNN_CHECK_ARG(p->n_linear == 1 || p->out <= 64, "train_mlp_fwd: out must be <= 64");
  SGNN_CHECK_ARG(p->n_linear == 2 || p->hidden % 32 == 0, "train_mlp_fwd: single Linear needs width %% 32 == 0");
  SGNN_CHECK_ARG(p->prologue != SGNN_TPRO_ROWS2 || (p->k0 % kTrKChunk == 0 && p->k0 + p->k1 <= p->kpad),
                 "train_mlp_fwd: ROWS2 needs k0 %% 64 == 0");
  SGNN_CHECK_ARG(p->prologue != SGNN_TPRO_EDGE_FEATURES || (p->kpad == 32 && p->dim >= 1 && p->dim <= 3 && p->src && p->tgt),
                 "train_mlp_fwd: EDGE_FEATURES needs kpad 32, dim 1..3 and the edge lists");
  SGNN_CHECK_ARG(p->add_p == nullptr || (p->add_q && p->src && p->tgt && p->ld_add % 4 == 0 && al16(p->add_p) && al16(p->add_q)),
                 "train_mlp_fwd: gathered terms need both matrices, the edge lists and 16-byte alignment");
  SGNN_CHECK_ARG(al16(p->y) && p->ld_y % 4 == 0 && (p->h1 == nullptr || al16(p->h1)) && (p->y2 == nullptr || al16(p->y2)) &&
                     (p->res == nullptr || (al16(p->res) && p->ld_res % 4 == 0)),
                 "train_mlp_fwd: outputs must be 16-byte aligned with leading dimensions %% 4 == 0");
  SGNN_CHECK_ARG(p->n_linear == 1 || p->h1 != nullptr, "train_mlp_fwd: h1 output missing");
  SGNN_CHECK_ARG(!p->has_ln || (p->y2 && p->mean && p->rstd && p->gamma && p->beta), "train_mlp_fwd: LayerNorm outputs missing");
  if (p->max_rows <= 0) return SGNN_OK;
  const TrFwdSmem s = tr_fwd_smem(p->kpad, p->hidden, p->out, p->n_linear, p->split);
  SGNN_CHECK_ARG(s.total <= 227 * 1024, "train_mlp_fwd: needs %d bytes of shared memory", s.total);
  const int tiles = (p->max_rows + kTileRows - 1) / kTileRows;
  int per_sm = (227 * 1024) / s.total;
  if (per_sm > 512 / kTrTmemCols) per_sm = 512 / kTrTmemCols;  // resident CTAs share the SM's 512 TMEM columns
  int grid = tr_sm_count() * per_sm;
  if (grid > tiles) grid = tiles;
  if (p->split) {
    SGNN_ENSURE_SMEM(train_fwd_kernel<true>, s.total);
    train_fwd_kernel<true><<<grid, kTrThreads, s.total, (cudaStream_t)stream>>>(*p);
  } else {
    SGNN_ENSURE_SMEM(train_fwd_kernel<false>, s.total);
    train_fwd_kernel<false><<<grid, kTrThreads, s.total, (cudaStream_t)stream>>>(*p);
  }
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

extern "C" int sgnn_train_mlp_bwd(const SgnnTrainBwd* p, void* stream) {
  SGNN_CHECK_ARG(p->n_linear == 1 || p->n_linear == 2, "train_mlp_bwd: n_linear must be 1 or 2");
  SGNN_CHECK_ARG(p->k_dy >= 1 && p->k_dy <= 256, "train_mlp_bwd: bad k_dy %d", p->k_dy);
  SGNN_CHECK_ARG(p->n_linear == 1 || (p->hidden % 32 == 0 && p->hidden >= 32 && p->hidden <= 128 && p->h1 && p->dpre &&
                                      al16(p->h1) && al16(p->dpre)),
                 "train_mlp_bwd: hidden %d must be 32..128 (multiple of 32) with h1 / dpre given", p->hidden);
  SGNN_CHECK_ARG(p->n_linear == 2 || p->want_dx, "train_mlp_bwd: a single Linear has only the dx product");
  SGNN_CHECK_ARG(p->n_linear == 1 || p->hidden == 64, "train_mlp_bwd: hidden must be 64");
  SGNN_CHECK_ARG(!p->want_dx || (p->k0 == 64 && (p->k1 == 0 || p->k1 == 64)), "train_mlp_bwd: dx widths must be 64 (+ 64)");
  SGNN_CHECK_ARG(p->dy && al16(p->dy) && p->ld_dy % 4 == 0 && p->k_dy % 4 == 0,
                 "train_mlp_bwd: dy must be 16-byte aligned, leading dimension and width multiples of 4");
  if (p->want_dx) {
    SGNN_CHECK_ARG(p->kx == p->k0 + p->k1 && p->kx % 16 == 0 && p->kx <= 128 && p->k0 % 32 == 0 && p->k0 > 0,
                   "train_mlp_bwd: dx widths k0 %d k1 %d", p->k0, p->k1);
    SGNN_CHECK_ARG(p->dx0 && al16(p->dx0) && p->ld_dx0 % 4 == 0 && (p->k1 == 0 || (p->dx1 && al16(p->dx1) && p->ld_dx1 % 4 == 0 && p->k1 % 16 == 0)),
                   "train_mlp_bwd: dx destinations must be 16-byte aligned");
    SGNN_CHECK_ARG(p->add1 == nullptr || (al16(p->add1) && p->ld_add1 % 4 == 0), "train_mlp_bwd: add1 alignment");
  }
  if (p->max_rows <= 0) return SGNN_OK;
  const TrBwdSmem s = tr_bwd_smem(p->k_dy, p->n_linear, p->hidden, p->want_dx, p->kx, p->split);
  SGNN_CHECK_ARG(s.total <= 227 * 1024, "train_mlp_bwd: needs %d bytes of shared memory", s.total);
  const int tiles = (p->max_rows + kTileRows - 1) / kTileRows;
  int per_sm = (227 * 1024) / s.total;
  if (per_sm > 512 / kTrTmemCols) per_sm = 512 / kTrTmemCols;
  int grid = tr_sm_count() * per_sm;
  if (grid > tiles) grid = tiles;
  if (p->split) {
    SGNN_ENSURE_SMEM(train_bwd_kernel<true>, s.total);
    train_bwd_kernel<true><<<grid, kTrThreads, s.total, (cudaStream_t)stream>>>(*p);
  } else {
    SGNN_ENSURE_SMEM(train_bwd_kernel<false>, s.total);
    train_bwd_kernel<false><<<grid, kTrThreads, s.total, (cudaStream_t)stream>>>(*p);
  }
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
