// Edge side of one message-passing layer, throughput version (same maths and the same results as
// edge_layer.cu; reference: InteractionNetwork.message + PyG sum-aggregation,
// sgnn/single_scale/graph_network.py:173,178-199).
//
// What changed against the one-tile-per-CTA kernel (profiles/README.md r01a: instruction- and
// latency-bound at 14 % of the HBM roofline, 8 warps per SM):
//   * ONE persistent CTA per SM with 4 tile workers of 128 threads that share a single resident
//     weight image; each worker runs its own TMA -> MMA -> epilogue chain, so 16 warps hide each
//     other's TMA / tensor-core / L2-gather latencies.
//   * hidden activations go from the epilogue to the next GEMM through TENSOR MEMORY
//     (tcgen05.st, then tcgen05.mma with the A operand in TMEM): no shared-memory round trip, no
//     swizzle arithmetic, and the worker's only big buffer (the e0 tile) is refilled by TMA as
//     soon as the first GEMM has consumed it -- a whole tile of prefetch distance.
//   * all widths are template parameters: every loop unrolls, the 64 outputs of a row stay in
//     registers through LayerNorm.
//   * segmented sums: each warp transposes its 32 rows through a 2.3 KB padded scratch in four
//     16-row x 32-column passes (warp-local __syncwarp only); lane = column, rows scanned in
//     order with a compile-time unrolled, warp-uniform segment mask.  Deterministic.
#include "mlp_tile.cuh"
#include "params.h"

namespace sgnn {

constexpr int kWorkers = 4;
constexpr int kFastThreads = kWorkers * kThreads;   // 512
constexpr int kScrRows = 16;
constexpr int kScrStride = 36;                       // floats per scratch row: 32 + 4 pad (16-byte aligned)
constexpr int kScrFloats = kScrRows * kScrStride;    // per warp

template <int OUT>
struct FastWorker {
  uint64_t bar_e, bar_m;
  int tgt_s[kTileRows];         // receiver of each row of the current tile
  int piece_tgt[8];
  int piece_flag[8];            // [2g] head piece of warp g (1 = exists, 3 = runs through), [2g+1] tail
  float pieces[8 * OUT];
};
template <int OUT>
struct FastMisc {
  uint64_t bar_w;
  uint32_t tmem_slot;
  int pad;
  FastWorker<OUT> w[kWorkers];
};

struct FastSmem {
  int e_tile, scratch, misc, total;
};
template <int HE, int OUT>
__host__ __device__ inline FastSmem edge_fast_smem(int blob_bytes) {
  FastSmem s;
  s.e_tile = align_up(blob_bytes, 1024);
  s.scratch = s.e_tile + kWorkers * kTileRows * HE * 4;
  s.misc = align_up(s.scratch + kWorkers * 4 * kScrFloats * 4, 16);
  s.total = s.misc + (int)sizeof(FastMisc<OUT>) + 1024;
  return s;
}

// round-to-nearest TF32 for a TMEM A operand: add half an ulp; the tensor core drops the 13 low
// mantissa bits itself, so the mask of to_tf32() is not needed here (same products, bit for bit)
__device__ __forceinline__ float tf32_bias(float x) { return __uint_as_float(__float_as_uint(x) + 0x1000u); }

__device__ __forceinline__ void worker_bar(int w) {
  asm volatile("bar.sync %0, %1;" ::"r"(w + 1), "r"(kThreads) : "memory");
}

template <int HE, int HID, int OUT>
__global__ void __launch_bounds__(kFastThreads, 1) edge_layer_fast_kernel(const SgnnEdgeLayerParams p) {
  constexpr int REG = HID > OUT ? HID : OUT;          // TMEM columns per accumulator region
  constexpr int TMEM_COLS = 512;
  static_assert(kWorkers * 2 * REG <= TMEM_COLS, "TMEM budget");
  static_assert(HE % 32 == 0 && HID % 32 == 0 && OUT % 32 == 0, "widths");
  constexpr int CPL = OUT / 32;
  constexpr unsigned FULL = 0xffffffffu;

  char* smem = smem_base_1024();
  const FastSmem L = edge_fast_smem<HE, OUT>(p.blob_bytes);
  const float* blob_f = reinterpret_cast<const float*>(smem);
  FastMisc<OUT>* misc = reinterpret_cast<FastMisc<OUT>*>(smem + L.misc);
  const int tid = threadIdx.x, w = tid >> 7, wt = tid & 127, lane = tid & 31, wq = wt >> 5;
  FastWorker<OUT>* wk = &misc->w[w];
  char* e_buf = smem + L.e_tile + w * (kTileRows * HE * 4);
  float* scr = reinterpret_cast<float*>(smem + L.scratch) + (w * 4 + wq) * kScrFloats;
  constexpr int tile_bytes = kTileRows * HE * 4;
  const int nh = p.mlp.n_hidden;

  pdl_launch_dependents();
  if (tid == 0) {
    mbar_init(&misc->bar_w, 1);
    for (int i = 0; i < kWorkers; ++i) {
      mbar_init(&misc->w[i].bar_e, 1);
      mbar_init(&misc->w[i].bar_m, 1);
    }
    mbar_fence_init();
    tma_load_span(smem, reinterpret_cast<const char*>(p.blob), p.blob_bytes, &misc->bar_w);
  }
  if (tid < 32) tmem_alloc(&misc->tmem_slot, TMEM_COLS);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_wait();          // everything above touched only shared/tensor memory and the constant weights
  const int n_edges = p.rowptr[p.n_nodes];
  const int n_tiles = (n_edges + kTileRows - 1) / kTileRows;
  const uint32_t tmem_w = misc->tmem_slot + (uint32_t)(w * 2 * REG);     // this worker's columns
  const uint32_t lane_base = (uint32_t)(wq * 32) << 16;

  // tiles go round-robin over CTAs first, then over a CTA's workers: small problems spread over
  // all SMs (one tile per SM) instead of packing four tiles onto a quarter of them
  const int tile0 = blockIdx.x + w * gridDim.x, stride = gridDim.x * kWorkers;
  if (wt == 0 && tile0 < n_tiles)
    tma_load_span(e_buf, reinterpret_cast<const char*>(p.e_tiles) + (long)tile0 * tile_bytes, tile_bytes,
                  &wk->bar_e);
  mbar_wait(&misc->bar_w, 0);
  uint32_t phase_e = 0, phase_m = 0;
  // sender / receiver of this thread's row, loaded one tile ahead so that the P/Q rows of the NEXT
  // tile can be pulled into L2 while this tile computes (PQ is 512 B per node: at 1M particles it
  // does not fit in L2 and a cold gather costs a full HBM round trip with the whole worker stalled)
  int s_nx = 0, t_nx = -3;
  if (tile0 < n_tiles && tile0 * kTileRows + wt < n_edges) {
    s_nx = __ldg(p.src + tile0 * kTileRows + wt);
    t_nx = __ldg(p.tgt + tile0 * kTileRows + wt);
  }

  for (int tile = tile0; tile < n_tiles; tile += stride) {
    const int e0 = tile * kTileRows;
    const int nvalid = n_edges - e0 < kTileRows ? n_edges - e0 : kTileRows;
    const bool valid = wt < nvalid;
    const int s = s_nx, t = t_nx;
    {
      const long en = (long)(tile + stride) * kTileRows + wt;
      s_nx = 0; t_nx = -3;
      if (en < n_edges) {
        s_nx = __ldg(p.src + en);
        t_nx = __ldg(p.tgt + en);
      }
    }
    if (wt < 8) wk->piece_flag[wt] = 0;
    wk->tgt_s[wt] = t;             // read back by this row's own warp only (after __syncwarp)

    // ---- segment structure of this warp's 32 rows (warp-uniform) ----
    const int nrows = nvalid - wq * 32 < 0 ? 0 : (nvalid - wq * 32 > 32 ? 32 : nvalid - wq * 32);
    int t_prev = __shfl_up_sync(FULL, t, 1);
    if (lane == 0) {
      const int idx = e0 + wq * 32 - 1;
      t_prev = (idx >= 0 && idx < n_edges) ? __ldg(p.tgt + idx) : -2;
    }
    const unsigned startmask = __ballot_sync(FULL, t != t_prev);   // bit r: row r starts a segment
    int t_after = -3;
    {
      const int idx = e0 + wq * 32 + nrows;
      if (nrows > 0 && idx < n_edges) t_after = __ldg(p.tgt + idx);
    }
    const int t_last = __shfl_sync(FULL, t, nrows > 0 ? nrows - 1 : 0);
    const bool first_cont = nrows > 0 && !(startmask & 1u);
    const bool cont_next = nrows > 0 && t_after == t_last;

    // ---- P[tgt] + Q[src]: 2 x HID fp32 per row, L2-resident gathers, issued before any wait ----
    float pq[HID];
    {
      const float4* prow = reinterpret_cast<const float4*>(p.pq + (long)(valid ? t : 0) * 2 * HID);
      const float4* qrow = reinterpret_cast<const float4*>(p.pq + (long)s * 2 * HID + HID);
#pragma unroll
      for (int j = 0; j < HID / 4; ++j) {
        const float4 a = __ldg(prow + j), b = __ldg(qrow + j);
        pq[4 * j] = a.x + b.x; pq[4 * j + 1] = a.y + b.y;
        pq[4 * j + 2] = a.z + b.z; pq[4 * j + 3] = a.w + b.w;
      }
    }

    if (t_nx >= 0) {
      const char* pn = reinterpret_cast<const char*>(p.pq + (long)t_nx * 2 * HID);
      const char* qn = reinterpret_cast<const char*>(p.pq + (long)s_nx * 2 * HID + HID);
#pragma unroll
      for (int b = 0; b < HID * 4; b += 128) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(pn + b));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(qn + b));
      }
    }

    // ---- stage 0: e0 (2^l W1c)^T on the tensor core, accumulator in region 0.  One thread waits
    //      on the mbarriers; the other 127 sleep in the worker's hardware barrier. ----
    if (wt < 32) {             // the worker's first warp, converged (see elect_one in common.cuh)
      mbar_wait(&wk->bar_e, phase_e);
      phase_e ^= 1u;
      tc_fence_after_sync();
      if (elect_one()) {
        umma_gemm_tf32(tmem_w, smem_u32(e_buf), smem_u32(smem + p.mlp.w_off[0]), HE, HID);
        tc_commit(&wk->bar_m);
      }
      __syncwarp();
      mbar_wait(&wk->bar_m, phase_m);
      phase_m ^= 1u;
      if (tile + stride < n_tiles && elect_one())  // the e0 buffer is free: prefetch the next tile
        tma_load_span(e_buf, reinterpret_cast<const char*>(p.e_tiles) + (long)(tile + stride) * tile_bytes,
                      tile_bytes, &wk->bar_e);
      __syncwarp();
    }
    worker_bar(w);
    tc_fence_after_sync();
#pragma unroll
    for (int c = 0; c < HID; c += 32) {
      float v[32];
      tmem_ld32(tmem_w + lane_base + c, v);
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = tf32_bias(fmaxf(v[j] + pq[c + j], 0.f));
      tmem_st32(tmem_w + lane_base + c, v);        // in place: A operand of the next GEMM
    }
    tmem_st_wait();
    tc_fence_before_sync();
    worker_bar(w);

    // ---- hidden -> hidden stages (nmlp_layers > 1): A in region ra, D in region 1 - ra ----
    int ra = 0;
    for (int l = 1; l < nh; ++l) {
      if (wt < 32) {
        tc_fence_after_sync();
        if (elect_one()) {
          umma_gemm_tf32_ts(tmem_w + (1 - ra) * REG, tmem_w + ra * REG, smem_u32(smem + p.mlp.w_off[l]), HID, HID);
          tc_commit(&wk->bar_m);
        }
        __syncwarp();
        mbar_wait(&wk->bar_m, phase_m);
        phase_m ^= 1u;
      }
      worker_bar(w);
      tc_fence_after_sync();
      ra = 1 - ra;
      const float* b = blob_f + p.mlp.b_off[l];
#pragma unroll
      for (int c = 0; c < HID; c += 32) {
        float v[32];
        tmem_ld32(tmem_w + lane_base + ra * REG + c, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = tf32_bias(fmaxf(v[j] + b[c + j], 0.f));
        tmem_st32(tmem_w + lane_base + ra * REG + c, v);
      }
      tmem_st_wait();
      tc_fence_before_sync();
      worker_bar(w);
    }

    // ---- output stage + LayerNorm: the row's OUT values stay in registers ----
    if (wt < 32) {
      tc_fence_after_sync();
      if (elect_one()) {
        umma_gemm_tf32_ts(tmem_w + (1 - ra) * REG, tmem_w + ra * REG, smem_u32(smem + p.mlp.w_off[nh]), HID, OUT);
        tc_commit(&wk->bar_m);
      }
      __syncwarp();
      mbar_wait(&wk->bar_m, phase_m);
      phase_m ^= 1u;
    }
    worker_bar(w);
    tc_fence_after_sync();
    float y[OUT];
    {
      const uint32_t ro = (uint32_t)((1 - ra) * REG);
#pragma unroll
      for (int c = 0; c < OUT; c += 32) tmem_ld32(tmem_w + lane_base + ro + c, y + c);
      const float* b = blob_f + p.mlp.b_off[nh];
      const float* gam = blob_f + p.mlp.gamma_off;
      const float* bet = blob_f + p.mlp.beta_off;
#pragma unroll
      for (int j = 0; j < OUT; ++j) y[j] += b[j];
      // 4 partial accumulators: a single dependent chain of 64 adds would idle the schedulers
      float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < OUT; ++j) s4[j & 3] += y[j];
      const float mean = ((s4[0] + s4[1]) + (s4[2] + s4[3])) / (float)OUT;
      float q4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < OUT; ++j) {
        y[j] -= mean;
        q4[j & 3] = fmaf(y[j], y[j], q4[j & 3]);
      }
      const float q = (q4[0] + q4[1]) + (q4[2] + q4[3]);
      const float rstd = 1.0f / sqrtf(q / (float)OUT + 1e-5f);
#pragma unroll
      for (int j = 0; j < OUT; ++j) y[j] = y[j] * rstd * gam[j] + bet[j];
    }

    // ---- segmented sums over this warp's rows: lane = column, rows in order ----
    if (nrows > 0) {
      __syncwarp();
      const int* tg = wk->tgt_s + wq * 32;
      const bool full = nrows == 32;
#pragma unroll
      for (int c0 = 0; c0 < OUT; c0 += 32) {
        float sum = 0.f;
        bool first_seg = true;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          if ((lane >> 4) == half) {
            float4* dst = reinterpret_cast<float4*>(scr + (lane & 15) * kScrStride);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              dst[j] = make_float4(y[c0 + 4 * j], y[c0 + 4 * j + 1], y[c0 + 4 * j + 2], y[c0 + 4 * j + 3]);
          }
          __syncwarp();
          float v[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = scr[i * kScrStride + lane];     // 16 independent LDS
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int r = half * 16 + i;
            if (full || r < nrows) {
              if (r > 0 && ((startmask >> r) & 1u)) {       // the segment ending at row r-1 is complete
                if (first_seg && first_cont) {
                  wk->pieces[(2 * wq) * OUT + c0 + lane] = sum;
                  if (lane == 0 && c0 == 0) { wk->piece_tgt[2 * wq] = tg[r - 1]; wk->piece_flag[2 * wq] = 1; }
                } else {
                  p.agg[(long)tg[r - 1] * OUT + c0 + lane] = sum;
                }
                first_seg = false;
                sum = 0.f;
              }
              sum += v[i];
            }
          }
          __syncwarp();
        }
        if (first_seg && first_cont) {
          wk->pieces[(2 * wq) * OUT + c0 + lane] = sum;
          if (lane == 0 && c0 == 0) { wk->piece_tgt[2 * wq] = t_last; wk->piece_flag[2 * wq] = cont_next ? 3 : 1; }
        } else if (cont_next) {
          wk->pieces[(2 * wq + 1) * OUT + c0 + lane] = sum;
          if (lane == 0 && c0 == 0) { wk->piece_tgt[2 * wq + 1] = t_last; wk->piece_flag[2 * wq + 1] = 1; }
        } else {
          p.agg[(long)t_last * OUT + c0 + lane] = sum;
        }
      }
    }
    tc_fence_before_sync();      // this tile's TMEM reads are retired before the next tile's MMAs
    worker_bar(w);

    // ---- stitch the pieces of CSR rows that straddle warps / tiles (warp 0 of the worker) ----
    if (wq == 0) {
      float carry[CPL];
      bool has = false, from_prev = false;
      int ctgt = -1;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const int f0 = wk->piece_flag[2 * g];
        if (f0) {
          if (!has) {
            has = true; from_prev = true; ctgt = wk->piece_tgt[2 * g];
#pragma unroll
            for (int j = 0; j < CPL; ++j) carry[j] = 0.f;
          }
#pragma unroll
          for (int j = 0; j < CPL; ++j) carry[j] += wk->pieces[(2 * g) * OUT + 32 * j + lane];
          if (!(f0 & 2)) {
            float* dst = from_prev ? p.head + (long)tile * OUT : p.agg + (long)ctgt * OUT;
#pragma unroll
            for (int j = 0; j < CPL; ++j) dst[32 * j + lane] = carry[j];
            has = false;
          }
        }
        if (wk->piece_flag[2 * g + 1]) {
          has = true; from_prev = false; ctgt = wk->piece_tgt[2 * g + 1];
#pragma unroll
          for (int j = 0; j < CPL; ++j) carry[j] = wk->pieces[(2 * g + 1) * OUT + 32 * j + lane];
        }
      }
      if (has) {   // the row continues into the next tile: its remainder arrives as that tile's head
        float* dst = from_prev ? p.head + (long)tile * OUT : p.agg + (long)ctgt * OUT;
#pragma unroll
        for (int j = 0; j < CPL; ++j) dst[32 * j + lane] = carry[j];
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (tid < 32) tmem_dealloc(misc->tmem_slot, TMEM_COLS);
}

}  // namespace sgnn

using namespace sgnn;

// Launches the throughput kernel when the layer shape has an instantiation and its weights fit;
// returns 1 if launched, 0 if the caller should use the generic kernel, < 0 on error.
int sgnn_edge_layer_fast_try(const SgnnEdgeLayerParams* p, void* stream) {
  if (!(p->he == 64 && p->mlp.hidden == 64 && p->mlp.out == 64)) return 0;
  const int smem = edge_fast_smem<64, 64>(p->blob_bytes).total;
  if (smem > 227 * 1024) return 0;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    SGNN_CUDA(cudaGetDevice(&dev));
    SGNN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  int grid = p->max_tiles;
  if (grid > sms) grid = sms;
  if (grid < 1) grid = 1;
  SGNN_ENSURE_SMEM((edge_layer_fast_kernel<64, 64, 64>), smem);
  SGNN_CUDA(sgnn_launch((edge_layer_fast_kernel<64, 64, 64>), dim3(grid), dim3(kFastThreads), smem, (cudaStream_t)stream, *p));
  return 1;
}
