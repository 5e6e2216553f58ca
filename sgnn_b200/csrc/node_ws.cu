// Node side of one message-passing layer, warp-specialised version for 64-wide latents with one hidden
// layer (reference: InteractionNetwork.update + residual, sgnn/single_scale/graph_network.py:176,
// 201-222; Decoder :321-333 and the integrator learned_simulator.py:381-411 for the last layer).
//
//   x'_i = x_i + LN(V2 relu(V1 [agg_i | x_i] + c1) + c2);  tail: P = x' W1a^T + b1 and the TF32-rounded
//   x' for the next edge layer (SGNN_TAIL_PX), or decoder MLP + Euler integrator (SGNN_TAIL_DECODE).
//
// node_fast.cu runs this as four independent serial chains per SM (row loads -> TMEM -> MMA -> ... ,
// five MMA round trips per tile): bound by global-load and round-trip latency at 41 % of the HBM
// roofline.  Here, as in edge_layer_ws.cu:
//   warp 16      producer: the tile's [agg | x] operand (K = 128) by FOUR TMA tile loads (2D tensor maps
//                over the row-major agg and rounded-x matrices, box 128 rows x one K-block, SWIZZLE_128B:
//                the copy engine does the row-major -> K-major swizzled transposition) into one of two
//                64 KB stages; coordinates are warp-uniform, so no per-lane issue loops
//   warps 17-19  one issuer warp per GEMM (first Linear from shared memory, second Linear and the
//                tail GEMM with A in TMEM), tcgen05 under elect.sync in converged code
//   warps 0-15   four epilogue groups of 128 threads (thread = row = TMEM lane); a group owns TMEM
//                columns [128g, 128g + 128) and every 4th tile
// The agg rows arrive unpatched: a receiver whose CSR row straddles edge tiles still lacks the `head`
// partials and a receiver without edges has no agg row at all.  The owning thread patches its row in
// the stage (rare) before the group releases the tile to the first issuer.
#include "mlp_tile.cuh"
#include "params.h"
#include "tensormap.cuh"

namespace sgnn {

constexpr int kNwGroups = 4;
constexpr int kNwThreads = kNwGroups * kThreads + 128;     // + one warpgroup: producer and three issuer warps
constexpr int kNwEpiRegs = 104, kNwAuxRegs = 48;           // setmaxnreg split (see edge_layer_ws.cu)
constexpr int kNwW = 64;
constexpr int kNwSlotBytes = kTileRows * kKBlockBytes;     // 16 KB
constexpr int kNwStageBytes = 4 * kNwSlotBytes;            // agg K-blocks 0,1 then x K-blocks 0,1
constexpr int kNwStages = 2;

struct NwGroup {
  // `full` is per GROUP, not per stage: a group sees only every 4th tile, and a parity wait is only
  // unambiguous on a barrier whose every phase the waiter consumes
  uint64_t full, patched, acc0_ready, hid_ready, acc1_ready, xo_ready, acc2_ready, region_free;
};
struct NwMisc {
  uint64_t bar_w;
  uint64_t empty[kNwStages];
  uint32_t tmem_slot;
  int pad;
  NwGroup g[kNwGroups];
};
struct NwSmem {
  int ring, misc, total;
};
__host__ __device__ inline NwSmem node_ws_smem(int blob_bytes) {
  NwSmem s;
  s.ring = align_up(blob_bytes, 1024);
  s.misc = s.ring + kNwStages * kNwStageBytes;
  s.total = s.misc + (int)sizeof(NwMisc) + 1024;
  return s;
}

__device__ __forceinline__ float nw_tf32r(float x) { return __uint_as_float(__float_as_uint(x) + 0x1000u); }
__device__ __forceinline__ float2 nf2(float a, float b) { return make_float2(a, b); }

// ---------------------------------------------------------------------------------- producer
__device__ __forceinline__ void nw_producer(const CUtensorMap* agg_map, const CUtensorMap* x_map, char* ring,
                                            NwMisc* misc, int n_my) {
  for (int i = 0; i < n_my; ++i) {
    const int tile = blockIdx.x + i * gridDim.x;
    const int st = i & 1, use = i >> 1;
    if (use > 0) mbar_wait_spin(&misc->empty[st], (uint32_t)(use - 1) & 1u);
    char* stage = ring + st * kNwStageBytes;
    uint64_t* full = &misc->g[i & 3].full;
    if (elect_one()) {
      mbar_arrive_expect_tx(full, kNwStageBytes);
      tma_tile_2d(stage, agg_map, 0, tile * kTileRows, full);
      tma_tile_2d(stage + kNwSlotBytes, agg_map, kKBlock, tile * kTileRows, full);
      tma_tile_2d(stage + 2 * kNwSlotBytes, x_map, 0, tile * kTileRows, full);
      tma_tile_2d(stage + 3 * kNwSlotBytes, x_map, kKBlock, tile * kTileRows, full);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------- MMA issuers
// acc0 = [agg | x] W0^T, K = 128 from the stage (W0 image: K-blocks 0,1 = agg columns, 2,3 = x columns)
__device__ __forceinline__ void nw_mma0(uint32_t ring_u32, uint32_t w0, NwMisc* misc, int n_my, uint32_t tmem) {
  const uint32_t idesc = umma_idesc_tf32(kTileRows, kNwW);
  for (int a = 0; a < n_my; ++a) {
    const int g = a & 3, st = a & 1;
    if (a >= kNwGroups) mbar_wait_spin(&misc->g[g].region_free, (uint32_t)((a >> 2) - 1) & 1u);
    mbar_wait_spin(&misc->g[g].patched, (uint32_t)(a >> 2) & 1u);     // implies the group's `full`
    tc_fence_after_sync();
    const uint32_t d = tmem + (uint32_t)(g * 2 * kNwW);
    const uint32_t a_st = ring_u32 + st * kNwStageBytes;
    if (elect_one()) {
#pragma unroll
      for (int kb = 0; kb < 4; ++kb)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          umma_tf32(d, umma_desc_sw128(a_st + kb * kNwSlotBytes + j * 32),
                    umma_desc_sw128(w0 + kb * kNwW * kKBlockBytes + j * 32), idesc, (kb > 0 || j > 0) ? 1u : 0u);
      tc_commit(&misc->empty[st]);
      tc_commit(&misc->g[g].acc0_ready);
    }
    __syncwarp();
  }
}
// D (region 1) = A (region 0, TMEM) W^T for the second Linear (which == 0) and the tail GEMM (which == 1)
__device__ __forceinline__ void nw_mma_ts(uint32_t w_img, NwMisc* misc, int n_my, uint32_t tmem, int which) {
  for (int b = 0; b < n_my; ++b) {
    NwGroup* gr = &misc->g[b & 3];
    mbar_wait_spin(which == 0 ? &gr->hid_ready : &gr->xo_ready, (uint32_t)(b >> 2) & 1u);
    tc_fence_after_sync();
    const uint32_t base = tmem + (uint32_t)((b & 3) * 2 * kNwW);
    if (elect_one()) {
      umma_gemm_tf32_ts(base + kNwW, base, w_img, kNwW, kNwW);
      tc_commit(which == 0 ? &gr->acc1_ready : &gr->acc2_ready);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------- epilogue
__device__ __forceinline__ void nw_epilogue(const SgnnNodeParams& p, const float* blob_f, char* ring, NwMisc* misc,
                                            int g, int wt, int n_my, uint32_t tmem) {
  constexpr int W = kNwW;
  const int wq = wt >> 5;
  NwGroup* gr = &misc->g[g];
  const uint32_t lane_base = (uint32_t)(wq * 32) << 16;
  const uint32_t t0 = tmem + (uint32_t)(g * 2 * W) + lane_base, t1 = t0 + W;
  const float4* b0 = reinterpret_cast<const float4*>(blob_f + p.mlp.b_off[0]);
  const float4* b1 = reinterpret_cast<const float4*>(blob_f + p.mlp.b_off[1]);
  const float4* gam4 = reinterpret_cast<const float4*>(blob_f + p.mlp.gamma_off);
  const float4* bet4 = reinterpret_cast<const float4*>(blob_f + p.mlp.beta_off);

  // CSR bounds of this thread's row, loaded one tile ahead
  int beg_nx = 0, end_nx = 0;
  if (g < n_my) {
    const long r = (long)(blockIdx.x + g * gridDim.x) * kTileRows + wt;
    if (r < p.n_rows) { beg_nx = __ldg(p.rowptr + r); end_nx = __ldg(p.rowptr + r + 1); }
  }

#pragma unroll 1
  for (int i = g; i < n_my; i += kNwGroups) {
    const uint32_t par = (uint32_t)(i >> 2) & 1u;
    const int tile = blockIdx.x + i * gridDim.x;
    const long r = (long)tile * kTileRows + wt;
    const bool valid = r < p.n_rows;
    const int beg = beg_nx, end = end_nx;
    {
      const long rn = (long)(tile + kNwGroups * gridDim.x) * kTileRows + wt;
      beg_nx = end_nx = 0;
      if (i + kNwGroups < n_my && rn < p.n_rows) { beg_nx = __ldg(p.rowptr + rn); end_nx = __ldg(p.rowptr + rn + 1); }
    }

    // ---- this thread's agg row in the stage: complete it (rare) and round it to TF32 (the copy engine
    //      delivered raw fp32; the x half was rounded by its producer), then release the tile ----
    mbar_wait_sleep(&gr->full, par);
    {
      char* agg_rows = ring + (i & 1) * kNwStageBytes;
      const bool has = valid && end > beg;           // no edges: the edge kernel never wrote this agg row
      const int t_lo = (beg >> 7) + 1, t_hi = has ? ((end - 1) >> 7) : 0;   // later edge tiles of a straddling CSR row
#pragma unroll 4
      for (int c = 0; c < W / 4; ++c) {
        float4* q = reinterpret_cast<float4*>(agg_rows + sw128_chunk_offset(wt, c, kTileRows));
        float4 a = has ? *q : make_float4(0.f, 0.f, 0.f, 0.f);
        for (int t = t_lo; t <= t_hi; ++t) {
          const float4 h = __ldg(reinterpret_cast<const float4*>(p.head + (long)t * W) + c);
          a.x += h.x; a.y += h.y; a.z += h.z; a.w += h.w;
        }
        *q = make_float4(nw_tf32r(a.x), nw_tf32r(a.y), nw_tf32r(a.z), nw_tf32r(a.w));
      }
    }
    fence_proxy_async_smem();                    // generic-proxy patch -> visible to tcgen05.mma
    mbar_arrive(&gr->patched);

    // ---- hidden = relu(acc0 + c1), rounded to TF32, back into the same TMEM columns ----
    mbar_wait_sleep(&gr->acc0_ready, par);
    tc_fence_after_sync();
#pragma unroll
    for (int c = 0; c < W; c += 32) {
      float v[32];
      tmem_ld32(t0 + c, v);
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 b = b0[(c + j) >> 2];
        const float2 s0 = __fadd2_rn(nf2(v[j], v[j + 1]), nf2(b.x, b.y));
        const float2 s1 = __fadd2_rn(nf2(v[j + 2], v[j + 3]), nf2(b.z, b.w));
        v[j] = nw_tf32r(fmaxf(s0.x, 0.f)); v[j + 1] = nw_tf32r(fmaxf(s0.y, 0.f));
        v[j + 2] = nw_tf32r(fmaxf(s1.x, 0.f)); v[j + 3] = nw_tf32r(fmaxf(s1.y, 0.f));
      }
      tmem_st32(t0 + c, v);
    }
    tmem_st_wait();
    tc_fence_before_sync();
    mbar_arrive(&gr->hid_ready);

    // ---- x' = x + LN(acc1 + c2) ----
    mbar_wait_sleep(&gr->acc1_ready, par);
    tc_fence_after_sync();
    float y[W];
    tmem_ld64(t1, y);
    {
      float2 s01 = nf2(0.f, 0.f), s23 = nf2(0.f, 0.f);
#pragma unroll
      for (int j = 0; j < W / 4; ++j) {
        const float4 b = b1[j];
        const float2 a = __fadd2_rn(nf2(y[4 * j], y[4 * j + 1]), nf2(b.x, b.y));
        const float2 c = __fadd2_rn(nf2(y[4 * j + 2], y[4 * j + 3]), nf2(b.z, b.w));
        y[4 * j] = a.x; y[4 * j + 1] = a.y; y[4 * j + 2] = c.x; y[4 * j + 3] = c.y;
        s01 = __fadd2_rn(s01, a);
        s23 = __fadd2_rn(s23, c);
      }
      const float nmean = -(((s01.x + s01.y) + (s23.x + s23.y)) / (float)W);
      const float2 nm = nf2(nmean, nmean);
      float2 q01 = nf2(0.f, 0.f), q23 = nf2(0.f, 0.f);
#pragma unroll
      for (int j = 0; j < W / 4; ++j) {
        const float2 a = __fadd2_rn(nf2(y[4 * j], y[4 * j + 1]), nm);
        const float2 c = __fadd2_rn(nf2(y[4 * j + 2], y[4 * j + 3]), nm);
        y[4 * j] = a.x; y[4 * j + 1] = a.y; y[4 * j + 2] = c.x; y[4 * j + 3] = c.y;
        q01 = __ffma2_rn(a, a, q01);
        q23 = __ffma2_rn(c, c, q23);
      }
      const float rstd = 1.0f / sqrtf(((q01.x + q01.y) + (q23.x + q23.y)) / (float)W + 1e-5f);
      const float2 r2 = nf2(rstd, rstd);
#pragma unroll
      for (int j = 0; j < W / 4; ++j) {
        const float4 ga = gam4[j], be = bet4[j];
        const float2 a = __ffma2_rn(__fmul2_rn(nf2(y[4 * j], y[4 * j + 1]), r2), nf2(ga.x, ga.y), nf2(be.x, be.y));
        const float2 c = __ffma2_rn(__fmul2_rn(nf2(y[4 * j + 2], y[4 * j + 3]), r2), nf2(ga.z, ga.w), nf2(be.z, be.w));
        y[4 * j] = a.x; y[4 * j + 1] = a.y; y[4 * j + 2] = c.x; y[4 * j + 3] = c.y;
      }
    }
    if (valid) {
      const float4* xr = reinterpret_cast<const float4*>(p.x_in + r * W);
#pragma unroll
      for (int j = 0; j < W / 4; ++j) {
        const float4 x = __ldg(xr + j);
        y[4 * j] += x.x; y[4 * j + 1] += x.y; y[4 * j + 2] += x.z; y[4 * j + 3] += x.w;
      }
      if (p.x_out) {
        float4* xo = reinterpret_cast<float4*>(p.x_out + r * W);
#pragma unroll
        for (int j = 0; j < W / 4; ++j) xo[j] = make_float4(y[4 * j], y[4 * j + 1], y[4 * j + 2], y[4 * j + 3]);
      }
    }
    // x' rounded to TF32: A operand of the tail GEMM (TMEM region 0) and, for the next edge layer's
    // sender gather, the Q half of the pq row
#pragma unroll
    for (int j = 0; j < W; ++j) y[j] = nw_tf32r(y[j]);
    tmem_st32(t0, y);
    tmem_st32(t0 + 32, y + 32);
    if (p.tail == SGNN_TAIL_PX && valid) {
      float4* dst = reinterpret_cast<float4*>(p.pq_out + r * 2 * W + W);
#pragma unroll
      for (int j = 0; j < W / 4; ++j) dst[j] = make_float4(y[4 * j], y[4 * j + 1], y[4 * j + 2], y[4 * j + 3]);
    }
    tmem_st_wait();
    tc_fence_before_sync();
    mbar_arrive(&gr->xo_ready);

    // ---- tail ----
    mbar_wait_sleep(&gr->acc2_ready, par);
    tc_fence_after_sync();
    if (p.tail == SGNN_TAIL_PX) {
      float o[W];
      tmem_ld64(t1, o);
      tc_fence_before_sync();
      mbar_arrive(&gr->region_free);
      if (valid) {
        const float4* pb = reinterpret_cast<const float4*>(blob_f + p.pq_bias_off);
        float4* dst = reinterpret_cast<float4*>(p.pq_out + r * 2 * W);
#pragma unroll
        for (int j = 0; j < W / 4; ++j) {
          const float4 b = pb[j];
          dst[j] = make_float4(o[4 * j] + b.x, o[4 * j + 1] + b.y, o[4 * j + 2] + b.z, o[4 * j + 3] + b.w);
        }
      }
    } else {      // SGNN_TAIL_DECODE with one hidden layer: relu(acc2 + d1) -> Linear(64 -> out <= 8) in FFMA -> integrator
      float o[8];
      const int nout = p.dec.out;
      const float* bl = blob_f + p.dec.b_off[1];
      const float* wl = blob_f + p.dec_wlast_off;
      const float* bh = blob_f + p.dec.b_off[0];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = 0.f;
#pragma unroll
      for (int c = 0; c < W; c += 32) {
        float h[32];
        tmem_ld32(t1 + c, h);
#pragma unroll
        for (int j = 0; j < 32; ++j) h[j] = fmaxf(h[j] + bh[c + j], 0.f);
        for (int k = 0; k < nout; ++k) {
          float acc = o[k];
#pragma unroll
          for (int j = 0; j < 32; ++j) acc = fmaf(h[j], wl[k * W + c + j], acc);
          o[k] = acc;
        }
      }
      tc_fence_before_sync();
      mbar_arrive(&gr->region_free);
      if (valid) {
        for (int k = 0; k < nout; ++k) o[k] += bl[k];
        if (p.dec_out)
          for (int k = 0; k < nout; ++k) p.dec_out[r * nout + k] = o[k];
        if (p.strain_out) p.strain_out[r] = o[nout - 1];
        if (p.next_pos) {
          const float* ps = p.pos_seq + r * p.pos_seq_stride;
          for (int d = 0; d < p.dim; ++d) {
            const float a = __fadd_rn(__fmul_rn(o[d], p.acc_std[d]), p.acc_mean[d]);
            const float last = ps[(p.seq_len - 1) * p.dim + d];
            const float vel = __fsub_rn(last, ps[(p.seq_len - 2) * p.dim + d]);
            p.next_pos[r * p.next_pos_stride + d] = __fadd_rn(last, __fadd_rn(vel, a));
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------- kernel
__global__ void __launch_bounds__(kNwThreads, 1)
node_ws_kernel(const SgnnNodeParams p, const __grid_constant__ CUtensorMap agg_map,
               const __grid_constant__ CUtensorMap x_map) {
  constexpr int TMEM_COLS = 512;
  char* smem = smem_base_1024();
  const NwSmem L = node_ws_smem(p.blob_bytes);
  NwMisc* misc = reinterpret_cast<NwMisc*>(smem + L.misc);
  const int tid = threadIdx.x, warp = tid >> 5;

  pdl_launch_dependents();
  if (tid == 0) {
    mbar_init(&misc->bar_w, 1);
    for (int s = 0; s < kNwStages; ++s) mbar_init(&misc->empty[s], 1);
    for (int g = 0; g < kNwGroups; ++g) {
      mbar_init(&misc->g[g].full, 1);
      mbar_init(&misc->g[g].patched, kThreads);
      mbar_init(&misc->g[g].acc0_ready, 1);
      mbar_init(&misc->g[g].hid_ready, kThreads);
      mbar_init(&misc->g[g].acc1_ready, 1);
      mbar_init(&misc->g[g].xo_ready, kThreads);
      mbar_init(&misc->g[g].acc2_ready, 1);
      mbar_init(&misc->g[g].region_free, kThreads);
    }
    mbar_fence_init();
    tma_load_span(smem, reinterpret_cast<const char*>(p.blob), p.blob_bytes, &misc->bar_w);
    tensormap_prefetch(&agg_map);
    tensormap_prefetch(&x_map);
  }
  if (tid < 32) tmem_alloc(&misc->tmem_slot, TMEM_COLS);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_wait();          // everything above touched only shared/tensor memory and the constant weights
  const int n_tiles = (p.n_rows + kTileRows - 1) / kTileRows;
  const int n_my = (int)blockIdx.x < n_tiles ? (n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const uint32_t tmem = misc->tmem_slot;
  const uint32_t sW = smem_u32(smem);

  if (warp < 16) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kNwEpiRegs));
    mbar_wait(&misc->bar_w, 0);
    nw_epilogue(p, reinterpret_cast<const float*>(smem), smem + L.ring, misc, warp >> 2, tid & 127, n_my, tmem);
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kNwAuxRegs));
    if (warp == 16) {
      nw_producer(&agg_map, &x_map, smem + L.ring, misc, n_my);
    } else {
      mbar_wait(&misc->bar_w, 0);
      if (warp == 17) nw_mma0(smem_u32(smem + L.ring), sW + p.mlp.w_off[0], misc, n_my, tmem);
      else if (warp == 18) nw_mma_ts(sW + p.mlp.w_off[1], misc, n_my, tmem, 0);
      else nw_mma_ts(sW + (p.tail == SGNN_TAIL_PX ? p.wp_off : p.dec.w_off[0]), misc, n_my, tmem, 1);
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (tid < 32) tmem_dealloc(tmem, TMEM_COLS);
}

}  // namespace sgnn

using namespace sgnn;

// 1 = launched, 0 = shape / mode not covered (caller falls back to node_fast_kernel), < 0 = error
int sgnn_node_ws_try(const SgnnNodeParams* p, void* stream) {
  if (getenv("SGNN_NODE_FAST")) return 0;                       // A/B switch: keep the 4-worker kernel
  if (p->prologue != SGNN_PRO_AGG_X || !p->xr_in || !p->has_mlp || !p->residual || !p->mlp.has_ln) return 0;
  if (p->agg_w != kNwW || p->x_w != kNwW || p->mlp.hidden != kNwW || p->mlp.out != kNwW || p->mlp.n_hidden != 1) return 0;
  if (p->tail == SGNN_TAIL_PX) {
    if (p->pq_hidden != kNwW) return 0;
  } else if (p->tail == SGNN_TAIL_DECODE) {
    if (p->dec.hidden != kNwW || p->dec.n_hidden != 1 || p->dec.out > 8) return 0;
  } else {
    return 0;
  }
  const NwSmem s = node_ws_smem(p->blob_bytes);
  if (s.total > 227 * 1024 || tensormap_encoder() == nullptr) return 0;
  CUtensorMap agg_map, x_map;
  int rc = make_row_tile_map(&agg_map, p->agg, p->n_rows, kNwW, kNwW * 4, kTileRows);
  if (rc != SGNN_OK) return rc;
  rc = make_row_tile_map(&x_map, p->xr_in, p->n_rows, kNwW, p->xr_ld * 4, kTileRows);
  if (rc != SGNN_OK) return rc;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    SGNN_CUDA(cudaGetDevice(&dev));
    SGNN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  const int n_tiles = sgnn_ceil_div(p->n_rows, kTileRows);
  int grid = n_tiles < sms ? n_tiles : sms;
  if (grid < 1) grid = 1;
  SGNN_ENSURE_SMEM(node_ws_kernel, s.total);
  SGNN_CUDA(sgnn_launch(node_ws_kernel, dim3(grid), dim3(kNwThreads), s.total, (cudaStream_t)stream, *p, agg_map, x_map));
  return 1;
}
