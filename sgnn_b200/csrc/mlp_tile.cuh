// Row-tile MLP machinery shared by the encoder / message-passing / decoder kernels.
//
// A CTA of 128 * P threads (P = 1, 2 or 4 "parts") owns a tile of 128 rows (edges or nodes); thread t works on
// row t & 127 in every epilogue (TMEM lane <-> accumulator row) and, of a row's 32-column chunks, on chunks
// t >> 7, (t >> 7) + P, ...: bias / ReLU / LayerNorm affine are thread-local, the LayerNorm moments are combined
// over the parts through shared memory.  One thread per row (P = 1) made every epilogue of a 128-wide layer ONE
// dependent instruction stream per scheduler (11.9 k instructions per warp and tile in the multi-scale node
// kernel, r02d; 46 us per tile with every phase latency-bound, r02f): the parts split it four ways.
// A GEMM step is  D[128 x n] (+)= A[128 x k] * W[n x k]^T  with A and W in shared memory in
// the K-major SWIZZLE_128B layout of common.cuh.
//
// Two arithmetic modes, selected at compile time, share every line of prologue/epilogue code:
//   TC = true  : tcgen05.mma kind::tf32, accumulators in TMEM (the product path)
//   TC = false : fp32 FFMA by the row-owning thread, accumulators in shared memory
//                (precision="fp32": tight-tolerance parity mode, ~30x slower)
#pragma once
#include <stdlib.h>

#include "common.cuh"
#include "params.h"

namespace sgnn {

constexpr int kThreads = 128;                // threads per part
constexpr int kMaxParts = 4;
__device__ __forceinline__ int t_row() { return threadIdx.x & (kTileRows - 1); }
__device__ __forceinline__ int t_part() { return threadIdx.x >> 7; }
__device__ __forceinline__ int t_parts() { return blockDim.x >> 7; }
// Parts for a layer whose widest row is `width` columns: four at 128 columns (the multi-scale network).  Narrower
// tensor-core layers keep one thread per row: that mapping adds the LayerNorm moments in the order of the 64-wide
// throughput kernels, so which of the two runs a layer does not change a bit of the result
// (tests/test_edge_fast_gpu.py); the FFMA parity mode has no TF32 rounding to amplify the reordering and takes two.
// SGNN_GENERIC_PARTS=1|2|4 overrides (capped by the width).
__host__ inline int sgnn_tile_parts(int width, bool tc) {
  int parts = width >= 128 ? 4 : (width >= 64 && !tc ? 2 : 1);
  const int cap = width >= 128 ? 4 : (width >= 64 ? 2 : 1);
  if (const char* e = getenv("SGNN_GENERIC_PARTS")) {
    const int v = atoi(e);
    if (v == 1 || v == 2 || v == 4) parts = v < cap ? v : cap;
  }
  return parts;
}
// this thread's 32-column chunks of a `width`-wide row
#define SGNN_FOR_MY_CHUNKS(c, width) \
  for (int c = 32 * t_part(), c##_step_ = 32 * t_parts(); c < (width); c += c##_step_)

template <bool TC>
struct TileAcc {
  uint32_t tmem_base;   // TC: TMEM base address (lane 0, column 0 of this CTA's allocation)
  float* simt_acc;      // !TC: [128][acc_cols] fp32, XOR-swizzled in 16-byte chunks
  int acc_cols;         // total accumulator columns (power of two >= 32 for TMEM)
  uint64_t* bar;        // TC: mbarrier signalled by tcgen05.commit
  uint32_t phase;
};

// ---- fp32 SIMT accumulator addressing (thread == row; conflict-free 16-byte accesses) ----
// Columns wrap modulo acc_cols (a power of two): the FFMA mode keeps ONE live accumulator region
// at a time, so the ping-pong regions of the tensor-core mode may alias.
__device__ __forceinline__ float4* simt_acc_chunk(float* base, int row, int chunk, int acc_cols) {
  chunk &= (acc_cols >> 2) - 1;
  return reinterpret_cast<float4*>(base + (size_t)row * acc_cols) + (chunk ^ (row & 7));
}

// One GEMM step.  Called by ALL threads of the CTA (uniform control flow).
//   a_tile / w_img : shared-memory operands (1024-byte aligned), k multiple of 32, n multiple of 16
//   acc_col        : first accumulator column of D
// TC: thread 0 issues the MMAs; completion is observed with gemm_wait().
template <bool TC>
__device__ __forceinline__ void gemm_step(TileAcc<TC>& acc, const void* a_tile, const void* w_img,
                                          int k, int n, int acc_col, bool accumulate) {
  if constexpr (TC) {
    if (threadIdx.x < 32) {            // warp 0, converged: the elected lane issues from uniform registers
      tc_fence_after_sync();
      if (elect_one()) umma_gemm_tf32(acc.tmem_base + acc_col, smem_u32(a_tile), smem_u32(w_img), k, n, accumulate);
      __syncwarp();
    }
  } else {
    const int row = t_row();
    const char* a = reinterpret_cast<const char*>(a_tile);
    const char* w = reinterpret_cast<const char*>(w_img);
    // a part computes the 32-column chunks its epilogue reads (n is a multiple of 32 in every caller with P > 1)
    for (int n0 = 0; n0 < n; n0 += 16) {
      if (((n0 >> 5) % t_parts()) != t_part()) continue;
      float d[16];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float4 v = accumulate ? *simt_acc_chunk(acc.simt_acc, row, (acc_col + n0) / 4 + j, acc.acc_cols)
                              : make_float4(0.f, 0.f, 0.f, 0.f);
        d[4 * j] = v.x; d[4 * j + 1] = v.y; d[4 * j + 2] = v.z; d[4 * j + 3] = v.w;
      }
      for (int kc = 0; kc < k / 4; ++kc) {
        const float4 av = *reinterpret_cast<const float4*>(a + sw128_chunk_offset(row, kc, kTileRows));
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float4 wv = *reinterpret_cast<const float4*>(w + sw128_chunk_offset(n0 + j, kc, n));
          d[j] = fmaf(av.x, wv.x, d[j]);
          d[j] = fmaf(av.y, wv.y, d[j]);
          d[j] = fmaf(av.z, wv.z, d[j]);
          d[j] = fmaf(av.w, wv.w, d[j]);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *simt_acc_chunk(acc.simt_acc, row, (acc_col + n0) / 4 + j, acc.acc_cols) =
            make_float4(d[4 * j], d[4 * j + 1], d[4 * j + 2], d[4 * j + 3]);
    }
  }
}

// Make all GEMM steps issued so far visible to the epilogue (all threads call).
template <bool TC>
__device__ __forceinline__ void gemm_wait(TileAcc<TC>& acc) {
  if constexpr (TC) {
    if (threadIdx.x < 32) {            // same warp (and, elect.sync picking the lowest lane, same thread) as the MMAs
      if (elect_one()) tc_commit(acc.bar);
      __syncwarp();
    }
    mbar_wait(acc.bar, acc.phase);
    acc.phase ^= 1u;
    tc_fence_after_sync();
  } else {
    if (t_parts() > 1) __syncthreads();      // FFMA mode: the other parts of a row have finished reading the A tile
  }
}

// ---- weight images: resident in shared memory, or streamed per GEMM ----------------------
// Blob layout (params.h): [0, kVecRegionBytes) fp32 vectors, then the GEMM images.  When the
// whole blob fits next to the kernel's tiles it is staged once and stays resident.  Otherwise
// (H = 128 with two hidden layers: 192-384 KB of images) only the vector region is resident and
// every GEMM pulls its image from global memory (L2-resident) into one 64 KB stage buffer by TMA,
// in chunks of whole K-blocks.
constexpr int kVecRegionBytes = SGNN_VEC_REGION_BYTES;
constexpr int kWbufBytesTc = 65536;     // tensor-core mode: up to 4 K-blocks of a 128-row image
constexpr int kWbufBytesSimt = 16384;   // FFMA parity mode: shared memory also holds the accumulators
__host__ __device__ inline int wbuf_bytes(bool tc) { return tc ? kWbufBytesTc : kWbufBytesSimt; }

// K-blocks of an n-row image that fit the stage buffer: wbuf / (128 n), without the ~40-instruction integer division
// on the path of every GEMM (n <= 256 and a multiple of 16: the quotient's fraction is a multiple of 1 / 256)
__device__ __forceinline__ int stage_kblocks(bool tc, int n) {
  return (int)(__fdividef((float)(wbuf_bytes(tc) / kKBlockBytes), (float)n) + 0.001f);
}

struct WeightStage {
  const char* gblob;   // the blob in global memory
  char* smem_blob;     // shared copy: whole blob (resident) or the vector region (streamed)
  char* wbuf;          // streamed: stage buffer
  uint64_t* bar;       // streamed: TMA completion barrier
  uint32_t phase;
  int streamed;
  int busy;            // streamed: MMAs reading wbuf issued and not yet waited for
  int pf_off;          // streamed: blob offset of the image whose first chunk weights_prefetch put in flight (-1: none)
};

// gemm_wait for GEMMs issued through gemm_weights
template <bool TC>
__device__ __forceinline__ void gemm_weights_wait(TileAcc<TC>& acc, WeightStage& ws) {
  gemm_wait<TC>(acc);
  ws.busy = 0;
}

// First stage-buffer chunk of the image at w_off, put in flight ahead of its gemm_weights call (the weights do not
// depend on the epilogue that runs in between, nor on the previous kernel).  All threads call, with no MMAs reading
// the stage buffer (right after gemm_weights_wait, or before the first GEMM).  k, n: the arguments of that call.
template <bool TC>
__device__ __forceinline__ void weights_prefetch(WeightStage& ws, int w_off, int k, int n) {
  if (!ws.streamed || ws.busy || ws.pf_off >= 0) return;
  if constexpr (!TC) __syncthreads();       // FFMA mode: every thread is done with wbuf
  const int kb_total = k / kKBlock;
  int kb_chunk = stage_kblocks(TC, n);
  if (kb_chunk > kb_total) kb_chunk = kb_total;
  if (threadIdx.x == 0) {
    const int bytes = kb_chunk * n * kKBlockBytes;
    mbar_arrive_expect_tx(ws.bar, (uint32_t)bytes);
    const char* src = ws.gblob + w_off;
    for (int off = 0; off < bytes; off += 32768)
      tma_bulk_g2s(ws.wbuf + off, src + off, (uint32_t)(bytes - off < 32768 ? bytes - off : 32768), ws.bar);
  }
  ws.pf_off = w_off;
}
// a prefetch nobody consumed has to land before the CTA exits or reuses the stage buffer
__device__ __forceinline__ void weights_prefetch_drain(WeightStage& ws) {
  if (ws.pf_off >= 0) {
    mbar_wait(ws.bar, ws.phase);
    ws.phase ^= 1u;
    ws.pf_off = -1;
  }
}

// D[128 x n] (+)= A[128 x k] * W^T with W the image at byte offset w_off of the blob.
// All threads call (uniform control flow).  Completion is observed with gemm_wait().
// (Inlined on purpose: as one shared __noinline__ copy the kernels shrank from 278 / 121 KB to 142 / 73 KB of SASS but
// the multi-scale step got 10 % slower -- acc / ws then live in local memory; r02f.)
template <bool TC>
__device__ __forceinline__ void gemm_weights(TileAcc<TC>& acc, WeightStage& ws, const void* a_tile,
                                             int w_off, int k, int n, int acc_col, bool accumulate) {
  if (!ws.streamed) {
    gemm_step<TC>(acc, a_tile, ws.smem_blob + w_off, k, n, acc_col, accumulate);
    return;
  }
  const int kb_total = k / kKBlock;
  int kb_chunk = stage_kblocks(TC, n);
  if (kb_chunk > kb_total) kb_chunk = kb_total;
  for (int kb0 = 0; kb0 < kb_total; kb0 += kb_chunk) {
    const int nkb = kb_total - kb0 < kb_chunk ? kb_total - kb0 : kb_chunk;
    const bool prefetched = kb0 == 0 && ws.pf_off == w_off;
    if (!prefetched) {
      weights_prefetch_drain(ws);                    // a prefetch of another image: let it land, then overwrite
      if (ws.busy) gemm_weights_wait<TC>(acc, ws);   // earlier MMAs have finished reading wbuf
      __syncthreads();                      // FFMA mode: every thread is done with wbuf
      if (threadIdx.x == 0) {
        const int bytes = nkb * n * kKBlockBytes;
        mbar_arrive_expect_tx(ws.bar, (uint32_t)bytes);
        const char* src = ws.gblob + w_off + (long)kb0 * n * kKBlockBytes;
        for (int off = 0; off < bytes; off += 32768)
          tma_bulk_g2s(ws.wbuf + off, src + off, (uint32_t)(bytes - off < 32768 ? bytes - off : 32768), ws.bar);
      }
    }
    ws.pf_off = -1;
    mbar_wait(ws.bar, ws.phase);
    ws.phase ^= 1u;
    gemm_step<TC>(acc, reinterpret_cast<const char*>(a_tile) + kb0 * kTileRows * kKBlockBytes, ws.wbuf,
                  nkb * kKBlock, n, acc_col, accumulate || kb0 > 0);
    ws.busy = 1;
  }
}

// Load 32 consecutive accumulator columns of this thread's row.
template <bool TC>
__device__ __forceinline__ void acc_load32(const TileAcc<TC>& acc, int col, float* v) {
  if constexpr (TC) {
    const uint32_t lane_base = (uint32_t)(((threadIdx.x >> 5) & 3) * 32) << 16;
    tmem_ld32(acc.tmem_base + lane_base + (uint32_t)col, v);
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 t = *simt_acc_chunk(acc.simt_acc, t_row(), col / 4 + j, acc.acc_cols);
      v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
    }
  }
}

// Epilogue -> next GEMM's A operand.  After the writers finish: generic->async proxy fence,
// tcgen05 fence, CTA barrier.  (All threads call.)
template <bool TC>
__device__ __forceinline__ void a_tile_publish() {
  if constexpr (TC) {
    fence_proxy_async_smem();
    tc_fence_before_sync();
  }
  __syncthreads();
}

// Store 32 columns [col, col+32) of this thread's row into an A tile (rounded to TF32 for TC).
template <bool TC>
__device__ __forceinline__ void a_tile_store32(void* a_tile, int col, const float* v) {
  char* a = reinterpret_cast<char*>(a_tile);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float4 t;
    if constexpr (TC) {
      t = make_float4(to_tf32(v[4 * j]), to_tf32(v[4 * j + 1]), to_tf32(v[4 * j + 2]), to_tf32(v[4 * j + 3]));
    } else {
      t = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
    *reinterpret_cast<float4*>(a + sw128_chunk_offset(t_row(), col / 4 + j, kTileRows)) = t;
  }
}

// Cooperative, coalesced load of `rows_valid` fp32 rows of width `width` (multiple of 4) from
// global memory into columns [col0, col0 + width) of an A tile.  Row r of the tile comes from
// src + row_index(r) * ld.  Rows >= rows_valid are zero-filled.  A warp covers a row with
// 16-byte lanes, so each global request is one contiguous row segment.
template <bool TC, class RowIndex>
__device__ __forceinline__ void a_tile_load_rows(void* a_tile, int col0, const float* __restrict__ src,
                                                 long ld, int width, int rows_valid, RowIndex row_index,
                                                 float scale = 1.0f) {
  char* a = reinterpret_cast<char*>(a_tile);
  const int chunks = width >> 2;                   // 16-byte chunks per row
  const int total = kTileRows * chunks;
  // independent loads in flight per thread: one load per iteration made this loop a chain of global round trips
  // (32 iterations x ~0.45 us for a 128-wide tile, the multi-scale kernels); four per iteration left sixteen round
  // trips for the [agg | x] tile of a 128-wide node kernel (~16 of its 61 us at 4800 nodes, r02d): eight now
  constexpr int U = 8;
  const int nthr = blockDim.x;
  for (int idx0 = threadIdx.x; idx0 < total; idx0 += U * nthr) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = idx0 + u * nthr;
      const int r = idx / chunks, c = idx - r * chunks;
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (idx < total && r < rows_valid) {
        const long g = row_index(r);
        if (g >= 0) v[u] = __ldg(reinterpret_cast<const float4*>(src + g * ld) + c);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int idx = idx0 + u * nthr;
      if (idx < total) {
        const int r = idx / chunks, c = idx - r * chunks;
        float4 t = make_float4(v[u].x * scale, v[u].y * scale, v[u].z * scale, v[u].w * scale);
        if constexpr (TC) t = make_float4(to_tf32(t.x), to_tf32(t.y), to_tf32(t.z), to_tf32(t.w));
        *reinterpret_cast<float4*>(a + sw128_chunk_offset(r, (col0 >> 2) + c, kTileRows)) = t;
      }
    }
  }
}

// The same rows through cp.async: every 16-byte chunk of the tile is in flight at once (ONE global round trip instead
// of total / (8 * 128) of them), nothing passes through registers.  Chunk idx = threadIdx.x + blockDim.x * i belongs to
// this thread in a_tile_fix_own() too, so the fix-up needs no barrier after cp_async_wait_all().
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool ok) {
  const int n = ok ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <class RowIndex>
__device__ __forceinline__ void a_tile_load_rows_async(void* a_tile, int col0, const float* __restrict__ src,
                                                       long ld, int width, int rows_valid, RowIndex row_index) {
  const uint32_t a = smem_u32(a_tile);
  const int chunks = width >> 2, total = kTileRows * chunks, nthr = blockDim.x;
  const int dr = nthr / chunks, dc = nthr - dr * chunks;
  int r = threadIdx.x / chunks, c = threadIdx.x - r * chunks;
  for (int idx = threadIdx.x; idx < total; idx += nthr) {
    const long g = r < rows_valid ? (long)row_index(r) : -1L;
    cp_async16(a + sw128_chunk_offset(r, (col0 >> 2) + c, kTileRows), g >= 0 ? src + g * ld + 4 * c : src, g >= 0);
    r += dr; c += dc;
    if (c >= chunks) { c -= chunks; ++r; }
  }
}
// In-place pass over the chunks a_tile_load_rows_async() gave this thread: fix(r, c, t) may patch the chunk
// (returns nothing), then TF32 rounding in tensor-core mode.
template <bool TC, class Fix>
__device__ __forceinline__ void a_tile_fix_own(void* a_tile, int col0, int width, Fix fix) {
  char* a = reinterpret_cast<char*>(a_tile);
  const int chunks = width >> 2, total = kTileRows * chunks, nthr = blockDim.x;
  const int dr = nthr / chunks, dc = nthr - dr * chunks;
  int r = threadIdx.x / chunks, c = threadIdx.x - r * chunks;
  for (int idx = threadIdx.x; idx < total; idx += nthr) {
    float4* q = reinterpret_cast<float4*>(a + sw128_chunk_offset(r, (col0 >> 2) + c, kTileRows));
    float4 t = *q;
    fix(r, c, t);
    if constexpr (TC) t = make_float4(to_tf32(t.x), to_tf32(t.y), to_tf32(t.z), to_tf32(t.w));
    *q = t;
    r += dr; c += dc;
    if (c >= chunks) { c -= chunks; ++r; }
  }
}

// ---- row-major staging tile for coalesced global I/O of thread-per-row epilogues -----------------------
// [128 rows][width] fp32, 16-byte chunk c of row r at chunk c ^ (r & 7) (width a multiple of 32): a lane reading
// or writing ITS row and a warp copying row segments are both conflict-free.  A warp owns the 32 x 32 blocks
// (its 32 rows, the 32-column chunks of its part) in both views, so __syncwarp() is the only barrier between them.
__device__ __forceinline__ float4* stage_chunk(char* stage, int row, int chunk, int width) {
  return reinterpret_cast<float4*>(stage) + row * (width >> 2) + (chunk ^ (row & 7));
}
// this warp's rows, columns [c0, c0 + 32) of src (leading dimension ld) -> stage, by cp.async; rows >= nvalid: skipped
__device__ __forceinline__ void stage_load_block_async(char* stage, const float* __restrict__ src, long ld, int width,
                                                       long row0, int nvalid, int c0) {
  const int lane = threadIdx.x & 31, wr0 = ((threadIdx.x >> 5) & 3) * 32;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = wr0 + 4 * i + (lane >> 3), ch = (c0 >> 2) + (lane & 7);
    if (r < nvalid) cp_async16(smem_u32(stage_chunk(stage, r, ch, width)), src + (row0 + r) * ld + 4 * ch, true);
  }
}
// this warp's rows, columns [c0, c0 + 32) of the stage -> dst (leading dimension ld): 128-byte row segments
__device__ __forceinline__ void stage_store_block(char* stage, float* __restrict__ dst, long ld, int width, long row0,
                                                  int nvalid, int c0) {
  const int lane = threadIdx.x & 31, wr0 = ((threadIdx.x >> 5) & 3) * 32;
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = wr0 + 4 * i + (lane >> 3), ch = (c0 >> 2) + (lane & 7);
    if (r < nvalid) *reinterpret_cast<float4*>(dst + (row0 + r) * ld + 4 * ch) = *stage_chunk(stage, r, ch, width);
  }
  __syncwarp();
}

// LayerNorm statistics of this thread's row over accumulator columns [col, col+n), after
// adding bias (torch semantics: biased variance, eps inside the sqrt).
// xch: [2][kMaxParts][128] floats of shared memory for the exchange between the parts of a row (all threads call).
template <bool TC>
__device__ __forceinline__ void row_ln_stats(const TileAcc<TC>& acc, int col, int n,
                                             const float* __restrict__ bias, float& mean, float& rstd, float* xch) {
  // 4 interleaved partial sums (column j goes to accumulator j & 3), as in edge_layer_fast.cu
  const int parts = t_parts(), row = t_row();
  float s4[4] = {0.f, 0.f, 0.f, 0.f};
  SGNN_FOR_MY_CHUNKS(c, n) {
    float v[32];
    acc_load32<TC>(acc, col + c, v);
#pragma unroll
    for (int j = 0; j < 32; ++j) s4[j & 3] += v[j] + bias[c + j];
  }
  float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
  if (parts > 1) {
    xch[t_part() * kTileRows + row] = sum;
    __syncthreads();
    sum = 0.f;
    for (int q = 0; q < parts; ++q) sum += xch[q * kTileRows + row];       // the same order in every part
  }
  mean = sum / (float)n;
  float q4[4] = {0.f, 0.f, 0.f, 0.f};
  SGNN_FOR_MY_CHUNKS(c, n) {
    float v[32];
    acc_load32<TC>(acc, col + c, v);
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float d = v[j] + bias[c + j] - mean;
      q4[j & 3] = fmaf(d, d, q4[j & 3]);
    }
  }
  float q = (q4[0] + q4[1]) + (q4[2] + q4[3]);
  if (parts > 1) {
    float* x2 = xch + kMaxParts * kTileRows;
    x2[t_part() * kTileRows + row] = q;
    __syncthreads();
    q = 0.f;
    for (int k = 0; k < parts; ++k) q += x2[k * kTileRows + row];
  }
  rstd = 1.0f / sqrtf(q / (float)n + 1e-5f);   // IEEE sqrt and divide, as torch
}

// Layers 1 .. n_hidden of an MLP whose layer-0 activations (after ReLU) already sit in `hid`
// (A-operand layout, width m.hidden).  Hidden layers: bias + ReLU written back into `hid`.
// Returns the accumulator column where the last Linear's output (bias not yet added) starts.
// `w` = accumulator region width, `region` = region used by layer 0.  All threads call.
// next_off / next_k / next_n: the streamed image the caller multiplies by next (-1: none), prefetched behind the last GEMM.
template <bool TC>
__device__ __forceinline__ int mlp_run_rest(TileAcc<TC>& acc, WeightStage& ws, char* hid,
                                            const SgnnMlp& m, int w, int region, int next_off = -1,
                                            int next_k = 0, int next_n = 0) {
  const float* blob_f = reinterpret_cast<const float*>(ws.smem_blob);
  const int n_out = (m.out + 15) & ~15;      // UMMA N granularity at M = 128
  for (int l = 1; l < m.n_hidden; ++l) {
    a_tile_publish<TC>();
    region ^= 1;
    gemm_weights<TC>(acc, ws, hid, m.w_off[l], m.hidden, m.hidden, region * w, false);
    gemm_weights_wait<TC>(acc, ws);
    weights_prefetch<TC>(ws, m.w_off[l + 1], m.hidden, l + 1 < m.n_hidden ? m.hidden : n_out);
    const float* b = blob_f + m.b_off[l];
    SGNN_FOR_MY_CHUNKS(c, m.hidden) {
      float v[32];
      acc_load32<TC>(acc, region * w + c, v);
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + b[c + j], 0.f);
      a_tile_store32<TC>(hid, c, v);
    }
  }
  a_tile_publish<TC>();
  region ^= 1;
  gemm_weights<TC>(acc, ws, hid, m.w_off[m.n_hidden], m.hidden, n_out, region * w, false);
  gemm_weights_wait<TC>(acc, ws);
  if (next_off >= 0) weights_prefetch<TC>(ws, next_off, next_k, next_n);
  return region * w;
}

// 32 output columns [c, c+32) of this thread's row: bias, then LayerNorm affine if has_ln
template <bool TC>
__device__ __forceinline__ void mlp_out_chunk(const TileAcc<TC>& acc, int acc_col, int c,
                                              const float* blob_f, const SgnnMlp& m, float mean,
                                              float rstd, float* v) {
  acc_load32<TC>(acc, acc_col + c, v);
  const float* b = blob_f + m.b_off[m.n_hidden];
  if (m.has_ln) {
    const float* gam = blob_f + m.gamma_off;
    const float* bet = blob_f + m.beta_off;
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = (v[j] + b[c + j] - mean) * rstd * gam[c + j] + bet[c + j];
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] += b[c + j];
  }
}

// ---- TMA bulk store shared -> global ----------------------------------------------------
__device__ __forceinline__ void tma_bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// stage `bytes` (multiple of 16) from global to shared with TMA bulk copies on one mbarrier
__device__ __forceinline__ void tma_load_span(char* dst, const char* src, int bytes, uint64_t* bar) {
  mbar_arrive_expect_tx(bar, (uint32_t)bytes);
  for (int off = 0; off < bytes; off += 32768) {
    const int n = bytes - off < 32768 ? bytes - off : 32768;
    tma_bulk_g2s(dst + off, src + off, (uint32_t)n, bar);
  }
}

__host__ __device__ inline int align_up(int x, int a) { return (x + a - 1) / a * a; }
__host__ __device__ inline int pow2_cols(int c) {
  int p = 32;
  while (p < c) p <<= 1;
  return p;
}

// ---- per-CTA setup / teardown of the accumulator --------------------------------------
template <bool TC>
__device__ __forceinline__ void acc_setup(TileAcc<TC>& acc, int acc_cols, uint64_t* bar,
                                          uint32_t* tmem_slot, float* simt_acc) {
  acc.acc_cols = acc_cols;
  acc.bar = bar;
  acc.phase = 0;
  acc.simt_acc = simt_acc;
  acc.tmem_base = 0;
  if constexpr (TC) {
    if (threadIdx.x == 0) {
      mbar_init(bar, 1);
      mbar_fence_init();
    }
    if (threadIdx.x < 32) tmem_alloc(tmem_slot, (uint32_t)acc_cols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    acc.tmem_base = *tmem_slot;
  } else {
    __syncthreads();
  }
}
template <bool TC>
__device__ __forceinline__ void acc_teardown(TileAcc<TC>& acc) {
  if constexpr (TC) {
    tc_fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(acc.tmem_base, (uint32_t)acc.acc_cols);
  }
}

// Stage the blob (resident: all of it; streamed: the vector region) and describe it in `ws`.
// Called by all threads before acc_setup's barrier; wait with mbar_wait(bar_w, 0) afterwards.
__device__ __forceinline__ void weights_setup(WeightStage& ws, const void* gblob, int blob_bytes,
                                              char* smem_blob, char* wbuf, uint64_t* bar_w,
                                              uint64_t* bar_stream, int streamed) {
  ws.gblob = reinterpret_cast<const char*>(gblob);
  ws.smem_blob = smem_blob;
  ws.wbuf = wbuf;
  ws.bar = bar_stream;
  ws.phase = 0;
  ws.streamed = streamed;
  ws.busy = 0;
  ws.pf_off = -1;
  if (threadIdx.x == 0) {
    mbar_init(bar_w, 1);
    mbar_init(bar_stream, 1);
    mbar_fence_init();
    int bytes = blob_bytes;
    if (streamed && bytes > kVecRegionBytes) bytes = kVecRegionBytes;
    tma_load_span(smem_blob, ws.gblob, bytes, bar_w);
  }
}
// shared-memory bytes taken by the blob / the stage buffer
__host__ __device__ inline int blob_smem_bytes(int blob_bytes, bool streamed) {
  return streamed ? kVecRegionBytes : blob_bytes;
}

// dynamic shared memory base, aligned to 1024 bytes (SWIZZLE_128B atoms).  Declared with the
// alignment instead of rounding a generic pointer up, so the compiler keeps the shared address
// space (LDS/STS instead of generic LD/ST).  None of the tile kernels has static shared memory.
__device__ __forceinline__ char* smem_base_1024() {
  extern __shared__ __align__(1024) char sgnn_dyn_smem[];
  return sgnn_dyn_smem;
}

}  // namespace sgnn
