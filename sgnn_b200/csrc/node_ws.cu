// Node side of one message-passing layer, warp-specialised version for 64-wide latents with one hidden
// layer (reference: InteractionNetwork.update + residual, sgnn/single_scale/graph_network.py:176,
// 201-222; Decoder :321-333 and the integrator learned_simulator.py:381-411 for the last layer).
//
//   x'_i = x_i + LN(V2 relu(V1 [agg_i | x_i] + c1) + c2);  tail: P = x' W1a^T + b1 and the TF32-rounded
//   x' for the next edge layer (SGNN_TAIL_PX), or decoder MLP + Euler integrator (SGNN_TAIL_DECODE).
//
// node_fast.cu runs this as four independent serial chains per SM with every row read and written by
// its own thread (16-byte accesses 256 or 512 bytes apart): ncu shows the L1 data pipe at 67 % with the
// issue slots 13 % busy -- the kernel is bound by uncoalesced LSU wavefronts, not by HBM.  Here NO tile
// data goes through the LSU to global memory at all:
//   * the latents travel between layers only as the "rounded" copy in the Q half of pq: bits(x) + 0x1000,
//     low mantissa bits kept (the tensor core ignores them).  That is exactly invertible, so the same
//     copy serves as the TF32 operand of both GEMMs that consume x AND as the fp32 residual
//     (x = bits - 0x1000): 1024 algorithmic bytes per node and layer instead of 1792
//   * warp 16 (producer) brings the tile's [agg | x] operand (K = 128) in by four TMA tile loads (2D
//     tensor maps, box 128 rows x one 128-byte K-block, SWIZZLE_128B: the copy engine does the row-major
//     -> K-major transposition); agg into one of two short-lived slots (free again after the first GEMM),
//     x into the slot of the epilogue group that owns the tile, where it stays for the whole chain
//   * warps 17-19: one issuer warp per GEMM (first Linear from shared memory, second Linear and the tail
//     GEMM with A in TMEM), tcgen05 under elect.sync in converged code
//   * warps 0-11: three epilogue groups of 128 threads (thread = row = TMEM lane), each owning TMEM
//     columns [128g, 128g + 128) and every 3rd tile.  The residual is read from the group's x slot, x'
//     is written back over it and leaves by a TMA tensor store; P follows through the same slot
// The agg rows arrive unpatched: a receiver whose CSR row straddles edge tiles still lacks the `head`
// partials and a receiver without edges has no agg row at all.  The owning thread completes and rounds
// its row in the slot before the group releases the tile to the first issuer.
#include "mlp_tile.cuh"
#include "params.h"
#include "tensormap.cuh"

namespace sgnn {

// optional in-kernel timeline (SGNN_TIMELINE builds; profiles/node_ws_timeline.py): CTA 0 / thread 0 stamps
// %globaltimer into slots [0,16) and clock64 into [16,32) of a user buffer, PX layers only
__device__ unsigned long long* g_nw_timeline = nullptr;
__device__ __forceinline__ void nw_stamp(int slot, bool on) {
#ifdef SGNN_TIMELINE
  if (on && g_nw_timeline != nullptr && blockIdx.x == 0 && (threadIdx.x == 0 || slot >= 11)) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    g_nw_timeline[slot] = t;
    g_nw_timeline[16 + slot] = (unsigned long long)clock64();
  }
#else
  (void)slot; (void)on;
#endif
}

constexpr int kNwGroups = 3;
constexpr int kNwThreads = kNwGroups * kThreads + 128;     // + one warpgroup: producer and three issuer warps
// setmaxnreg: 512 threads launch at 128; (128 - 56) * 128 released = (152 - 128) * 384 acquired
constexpr int kNwEpiRegs = 152, kNwAuxRegs = 56;
constexpr int kNwW = 64;
constexpr int kNwKbBytes = kTileRows * kKBlockBytes;       // 16 KB: one K-block of a 128-row operand tile
constexpr int kNwSlotBytes = 2 * kNwKbBytes;               // 32 KB: 64 columns
constexpr int kNwAggSlots = 2;
constexpr int kNwVecBytes = 2048;                          // resident part of the blob's vector region
constexpr int kNwImgBytes = 4 * kNwW * kNwW * 4;           // W0 (K = 128) + W1 + tail image

struct NwGroup {
  // `full` is per GROUP, not per slot: a parity wait is only unambiguous on a barrier whose every phase
  // the waiter consumes, and a group sees only every 3rd tile
  uint64_t full, patched, acc0_ready, hid_ready, acc1_ready, xo_ready, acc2_ready, region_free, slot_free;
};
struct NwMisc {
  uint64_t bar_w;
  uint64_t agg_empty[kNwAggSlots];
  uint32_t tmem_slot;
  int pad;
  NwGroup g[kNwGroups];
};
constexpr int kNwImgOff = kNwVecBytes;
constexpr int kNwAggOff = kNwImgOff + kNwImgBytes;
constexpr int kNwXOff = kNwAggOff + kNwAggSlots * kNwSlotBytes;
constexpr int kNwMiscOff = kNwXOff + kNwGroups * kNwSlotBytes;
constexpr int kNwSmemBytes = kNwMiscOff + 1024;
static_assert(sizeof(NwMisc) <= 1024, "misc block");
static_assert(kNwSmemBytes <= 227 * 1024, "shared memory budget");

__device__ __forceinline__ float nw_tf32r(float x) { return __uint_as_float(__float_as_uint(x) + 0x1000u); }
__device__ __forceinline__ float nw_unround(float x) { return __uint_as_float(__float_as_uint(x) - 0x1000u); }
__device__ __forceinline__ float2 nf2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ void nw_group_bar(int g) {
  asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(kThreads) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, int col, int row, const void* smem_src) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(smem_u32(smem_src)), "r"(col), "r"(row)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }

// ---------------------------------------------------------------------------------- producer
__device__ __forceinline__ void nw_producer(const CUtensorMap* agg_map, const CUtensorMap* x_map, char* smem,
                                            NwMisc* misc, int n_my) {
  for (int i = 0; i < n_my; ++i) {
    const int tile = blockIdx.x + i * gridDim.x;
    const int g = i % kNwGroups, use_g = i / kNwGroups, sa = i & 1;
    uint64_t* full = &misc->g[g].full;
    char* agg_slot = smem + kNwAggOff + sa * kNwSlotBytes;
    char* x_slot = smem + kNwXOff + g * kNwSlotBytes;
    if (i >= kNwAggSlots) mbar_wait_spin(&misc->agg_empty[sa], (uint32_t)((i >> 1) - 1) & 1u);
    if (elect_one()) {
      mbar_arrive_expect_tx(full, 2 * kNwSlotBytes);
      tma_tile_2d(agg_slot, agg_map, 0, tile * kTileRows, full);
      tma_tile_2d(agg_slot + kNwKbBytes, agg_map, kKBlock, tile * kTileRows, full);
    }
    __syncwarp();
    if (use_g > 0) mbar_wait_spin(&misc->g[g].slot_free, (uint32_t)(use_g - 1) & 1u);
    if (elect_one()) {
      tma_tile_2d(x_slot, x_map, 0, tile * kTileRows, full);
      tma_tile_2d(x_slot + kNwKbBytes, x_map, kKBlock, tile * kTileRows, full);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------- MMA issuers
// acc0 = [agg | x] W0^T, K = 128 (W0 image: K-blocks 0,1 = agg columns, 2,3 = x columns)
__device__ __forceinline__ void nw_mma0(uint32_t smem_u, uint32_t w0, NwMisc* misc, int n_my, uint32_t tmem, bool px) {
  const uint32_t idesc = umma_idesc_tf32(kTileRows, kNwW);
  for (int a = 0; a < n_my; ++a) {
    const int g = a % kNwGroups, use_g = a / kNwGroups, sa = a & 1;
    if (use_g > 0) mbar_wait_spin(&misc->g[g].region_free, (uint32_t)(use_g - 1) & 1u);
    mbar_wait_spin(&misc->g[g].patched, (uint32_t)use_g & 1u);     // implies the group's `full`
    nw_stamp(11, px && (threadIdx.x & 31) == 0);
    tc_fence_after_sync();
    const uint32_t d = tmem + (uint32_t)(g * 2 * kNwW);
    const uint32_t a_agg = smem_u + kNwAggOff + sa * kNwSlotBytes;
    const uint32_t a_x = smem_u + kNwXOff + g * kNwSlotBytes;
    if (elect_one()) {
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) {
        const uint32_t a_kb = (kb < 2 ? a_agg : a_x) + (kb & 1) * kNwKbBytes;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          umma_tf32(d, umma_desc_sw128(a_kb + j * 32), umma_desc_sw128(w0 + kb * kNwW * kKBlockBytes + j * 32), idesc,
                    (kb > 0 || j > 0) ? 1u : 0u);
        if (kb == 1) tc_commit(&misc->agg_empty[sa]);
      }
      tc_commit(&misc->g[g].acc0_ready);
    }
    __syncwarp();
    nw_stamp(12, px && (threadIdx.x & 31) == 0);
  }
}
// D (region 1) = A (region 0, TMEM) W^T for the second Linear (which == 0) and the tail GEMM (which == 1)
__device__ __forceinline__ void nw_mma_ts(uint32_t w_img, NwMisc* misc, int n_my, uint32_t tmem, int which) {
  for (int b = 0; b < n_my; ++b) {
    const int g = b % kNwGroups;
    NwGroup* gr = &misc->g[g];
    mbar_wait_spin(which == 0 ? &gr->hid_ready : &gr->xo_ready, (uint32_t)(b / kNwGroups) & 1u);
    tc_fence_after_sync();
    const uint32_t base = tmem + (uint32_t)(g * 2 * kNwW);
    if (elect_one()) {
      umma_gemm_tf32_ts(base + kNwW, base, w_img, kNwW, kNwW);
      tc_commit(which == 0 ? &gr->acc1_ready : &gr->acc2_ready);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------- epilogue
__device__ __forceinline__ void nw_epilogue(const SgnnNodeParams& p, const CUtensorMap* pq_map, char* smem,
                                            NwMisc* misc, int g, int wt, int n_my, uint32_t tmem) {
  constexpr int W = kNwW;
  const int wq = wt >> 5;
  NwGroup* gr = &misc->g[g];
  const uint32_t lane_base = (uint32_t)(wq * 32) << 16;
  const uint32_t t0 = tmem + (uint32_t)(g * 2 * W) + lane_base, t1 = t0 + W;
  const float* vec = reinterpret_cast<const float*>(smem);
  const float4* b0 = reinterpret_cast<const float4*>(vec + p.mlp.b_off[0]);
  const float4* b1 = reinterpret_cast<const float4*>(vec + p.mlp.b_off[1]);
  const float4* gam4 = reinterpret_cast<const float4*>(vec + p.mlp.gamma_off);
  const float4* bet4 = reinterpret_cast<const float4*>(vec + p.mlp.beta_off);
  const float4* ag4 = reinterpret_cast<const float4*>(vec + (p.agg_affine ? p.agg_gamma_off : 0));
  const float4* ab4 = reinterpret_cast<const float4*>(vec + (p.agg_affine ? p.agg_beta_off : 0));
  char* x_slot = smem + kNwXOff + g * kNwSlotBytes;
  const bool px = p.tail == SGNN_TAIL_PX;
  const bool poll = false;      // polling epilogue warps measured slower even with one tile per group (r01n)

  // CSR bounds of this thread's row, loaded one tile ahead
  int beg_nx = 0, end_nx = 0;
  if (g < n_my) {
    const long r = (long)(blockIdx.x + g * gridDim.x) * kTileRows + wt;
    if (r < p.n_rows) { beg_nx = __ldg(p.rowptr + r); end_nx = __ldg(p.rowptr + r + 1); }
  }

#pragma unroll 1
  for (int i = g; i < n_my; i += kNwGroups) {
    const uint32_t par = (uint32_t)(i / kNwGroups) & 1u;
    const int tile = blockIdx.x + i * gridDim.x;
    const long r = (long)tile * kTileRows + wt;
    const bool valid = r < p.n_rows;
    const int beg = beg_nx, end = end_nx;
    {
      const long rn = (long)(tile + kNwGroups * gridDim.x) * kTileRows + wt;
      beg_nx = end_nx = 0;
      if (i + kNwGroups < n_my && rn < p.n_rows) { beg_nx = __ldg(p.rowptr + rn); end_nx = __ldg(p.rowptr + rn + 1); }
    }

    // ---- this thread's agg row in its slot: complete it and round it to TF32 (the copy engine delivered
    //      raw fp32; the x half was rounded by its producer), then release the tile.  A receiver whose CSR
    //      row straddles two edge tiles still lacks the later tile's `head` partial: rare per ROW but nearly
    //      certain per TILE (one tile in five ends inside a row at degree 5), and the loads sat on the
    //      tile's serial chain, 1.8 us after the tile had landed (in-kernel stamps, r01n) -- so the partial
    //      is fetched BEFORE the wait for the tile, next to the TMA loads it does not depend on ----
    const bool has = valid && end > beg;             // no edges: the edge kernel never wrote this agg row
    const float cnt = has ? (float)(end - beg) : 0.f;
    const int t_lo = (beg >> 7) + 1, t_hi = has ? ((end - 1) >> 7) : 0;     // later edge tiles of a straddling CSR row
    float4 hp[W / 4];
    if (t_hi >= t_lo) {
      const float4* h4 = reinterpret_cast<const float4*>(p.head + (long)t_lo * W);
#pragma unroll
      for (int c = 0; c < W / 4; ++c) hp[c] = __ldg(h4 + c);
    } else {
#pragma unroll
      for (int c = 0; c < W / 4; ++c) hp[c] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    mbar_wait_mode(&gr->full, par, poll);
    nw_stamp(3, px);
    {
      char* agg_rows = smem + kNwAggOff + (i & 1) * kNwSlotBytes;
      if (t_hi <= t_lo) {
#pragma unroll
        for (int c = 0; c < W / 4; ++c) {
          float4* q = reinterpret_cast<float4*>(agg_rows + sw128_chunk_offset(wt, c, kTileRows));
          float4 a = has ? *q : make_float4(0.f, 0.f, 0.f, 0.f);
          a.x += hp[c].x; a.y += hp[c].y; a.z += hp[c].z; a.w += hp[c].w;
          if (p.agg_affine) {                        // the edge MLP's LayerNorm affine on the complete sum
            const float4 ga = ag4[c], be = ab4[c];
            a.x = fmaf(a.x, ga.x, cnt * be.x); a.y = fmaf(a.y, ga.y, cnt * be.y);
            a.z = fmaf(a.z, ga.z, cnt * be.z); a.w = fmaf(a.w, ga.w, cnt * be.w);
          }
          *q = make_float4(nw_tf32r(a.x), nw_tf32r(a.y), nw_tf32r(a.z), nw_tf32r(a.w));
        }
      } else {                                       // rows longer than a tile (arbitrary edge lists only)
#pragma unroll 1
        for (int c = 0; c < W / 4; ++c) {
          float4* q = reinterpret_cast<float4*>(agg_rows + sw128_chunk_offset(wt, c, kTileRows));
          float4 a = *q;
          for (int t = t_lo; t <= t_hi; ++t) {
            const float4 h = __ldg(reinterpret_cast<const float4*>(p.head + (long)t * W) + c);
            a.x += h.x; a.y += h.y; a.z += h.z; a.w += h.w;
          }
          if (p.agg_affine) {
            const float4 ga = ag4[c], be = ab4[c];
            a.x = fmaf(a.x, ga.x, cnt * be.x); a.y = fmaf(a.y, ga.y, cnt * be.y);
            a.z = fmaf(a.z, ga.z, cnt * be.z); a.w = fmaf(a.w, ga.w, cnt * be.w);
          }
          *q = make_float4(nw_tf32r(a.x), nw_tf32r(a.y), nw_tf32r(a.z), nw_tf32r(a.w));
        }
      }
    }
    nw_stamp(10, px);
    fence_proxy_async_smem();                    // generic-proxy writes -> visible to tcgen05.mma
    mbar_arrive(&gr->patched);

    // ---- hidden = relu(acc0 + c1), rounded to TF32, back into the same TMEM columns ----
    mbar_wait_mode(&gr->acc0_ready, par, poll);
    nw_stamp(4, px);
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
    mbar_wait_mode(&gr->acc1_ready, par, poll);
    nw_stamp(5, px);
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
    // residual from the group's x slot (rounded copy -> exact fp32), x' (rounded) back in its place:
    // A operand of the tail GEMM through TMEM, sender/next-layer copy through the slot and a TMA store
    {
      float4* xo = (p.x_out != nullptr && valid) ? reinterpret_cast<float4*>(p.x_out + r * W) : nullptr;
#pragma unroll
      for (int c = 0; c < W / 4; ++c) {
        float4* q = reinterpret_cast<float4*>(x_slot + sw128_chunk_offset(wt, c, kTileRows));
        const float4 xr = *q;
        const float a0 = y[4 * c] + nw_unround(xr.x), a1 = y[4 * c + 1] + nw_unround(xr.y);
        const float a2 = y[4 * c + 2] + nw_unround(xr.z), a3 = y[4 * c + 3] + nw_unround(xr.w);
        if (xo) xo[c] = make_float4(a0, a1, a2, a3);          // only on request (tests, last layer)
        y[4 * c] = nw_tf32r(a0); y[4 * c + 1] = nw_tf32r(a1); y[4 * c + 2] = nw_tf32r(a2); y[4 * c + 3] = nw_tf32r(a3);
        if (px) *q = make_float4(y[4 * c], y[4 * c + 1], y[4 * c + 2], y[4 * c + 3]);
      }
    }
    tmem_st32(t0, y);
    tmem_st32(t0 + 32, y + 32);
    tmem_st_wait();
    tc_fence_before_sync();
    mbar_arrive(&gr->xo_ready);
    nw_stamp(6, px);
    fence_proxy_async_smem();
    nw_group_bar(g);                             // every row of x' is in the slot / has been read from it

    if (px) {
      if (wq == 0) {
        if (elect_one()) {
          tma_store_2d(pq_map, W, tile * kTileRows, x_slot);
          tma_store_2d(pq_map, W + kKBlock, tile * kTileRows, x_slot + kNwKbBytes);
          tma_store_commit();
        }
        __syncwarp();
      }
      mbar_wait_mode(&gr->acc2_ready, par, poll);
      nw_stamp(7, px);
      tc_fence_after_sync();
      float o[W];
      tmem_ld64(t1, o);
      tc_fence_before_sync();
      mbar_arrive(&gr->region_free);
      if (wq == 0) {
        if (elect_one()) tma_store_wait_read();  // the copy engine has read x' out of the slot
        __syncwarp();
      }
      nw_group_bar(g);
      const float4* pb = reinterpret_cast<const float4*>(vec + p.pq_bias_off);
#pragma unroll
      for (int c = 0; c < W / 4; ++c) {
        const float4 b = pb[c];
        *reinterpret_cast<float4*>(x_slot + sw128_chunk_offset(wt, c, kTileRows)) =
            make_float4(o[4 * c] + b.x, o[4 * c + 1] + b.y, o[4 * c + 2] + b.z, o[4 * c + 3] + b.w);
      }
      fence_proxy_async_smem();
      nw_group_bar(g);
      if (wq == 0) {
        if (elect_one()) {
          tma_store_2d(pq_map, 0, tile * kTileRows, x_slot);
          tma_store_2d(pq_map, kKBlock, tile * kTileRows, x_slot + kNwKbBytes);
          tma_store_commit();
          tma_store_wait_read();
          mbar_arrive(&gr->slot_free);
        }
        __syncwarp();
      }
    } else {      // SGNN_TAIL_DECODE with one hidden layer: relu(acc2 + d1) -> Linear(64 -> out <= 8) in FFMA -> integrator
      if (wq == 0) {
        if (elect_one()) mbar_arrive(&gr->slot_free);
        __syncwarp();
      }
      mbar_wait_mode(&gr->acc2_ready, par, poll);
      tc_fence_after_sync();
      float o[8];
      const int nout = p.dec.out;
      const float* bl = p.blob + p.dec.b_off[1];        // decoder vectors: read-only, L1-resident
      const float* wl = p.blob + p.dec_wlast_off;
      const float4* bh = reinterpret_cast<const float4*>(p.blob + p.dec.b_off[0]);
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = 0.f;
#pragma unroll
      for (int c = 0; c < W; c += 32) {
        float h[32];
        tmem_ld32(t1 + c, h);
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 b = __ldg(bh + ((c + j) >> 2));
          h[j] = fmaxf(h[j] + b.x, 0.f); h[j + 1] = fmaxf(h[j + 1] + b.y, 0.f);
          h[j + 2] = fmaxf(h[j + 2] + b.z, 0.f); h[j + 3] = fmaxf(h[j + 3] + b.w, 0.f);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (k < nout) {
            const float4* w4 = reinterpret_cast<const float4*>(wl + k * W + c);
            float acc = o[k];
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 w = __ldg(w4 + (j >> 2));
              acc = fmaf(h[j], w.x, acc); acc = fmaf(h[j + 1], w.y, acc);
              acc = fmaf(h[j + 2], w.z, acc); acc = fmaf(h[j + 3], w.w, acc);
            }
            o[k] = acc;
          }
        }
      }
      tc_fence_before_sync();
      mbar_arrive(&gr->region_free);
      if (valid) {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (k < nout) {
            o[k] += __ldg(bl + k);
            if (p.dec_out) p.dec_out[r * nout + k] = o[k];
            if (p.strain_out && k == nout - 1) p.strain_out[r] = o[k];
          }
        if (p.next_pos) {
          const float* ps = p.pos_seq + r * p.pos_seq_stride;
#pragma unroll
          for (int d = 0; d < 3; ++d)
            if (d < p.dim) {
              const float a = __fadd_rn(__fmul_rn(o[d], p.acc_std[d]), p.acc_mean[d]);
              const float last = ps[(p.seq_len - 1) * p.dim + d];
              const float vel = __fsub_rn(last, ps[(p.seq_len - 2) * p.dim + d]);
              p.next_pos[r * p.next_pos_stride + d] = __fadd_rn(last, __fadd_rn(vel, a));
            }
        }
      }
    }
  }
  if (px && wq == 0) {
    // shared memory must outlive the copy engine's READS only; the writes are ordered before the grid's
    // completion (and a dependent's griddepcontrol.wait) by the hardware
    if (elect_one()) tma_store_wait_read();
    __syncwarp();
  }
  nw_stamp(8, px);
}

// ---------------------------------------------------------------------------------- kernel
__global__ void __launch_bounds__(kNwThreads, 1)
node_ws_kernel(const SgnnNodeParams p, const __grid_constant__ CUtensorMap agg_map,
               const __grid_constant__ CUtensorMap x_map, const __grid_constant__ CUtensorMap pq_map) {
  constexpr int TMEM_COLS = 512;
  char* smem = smem_base_1024();
  NwMisc* misc = reinterpret_cast<NwMisc*>(smem + kNwMiscOff);
  const int tid = threadIdx.x, warp = tid >> 5;

  pdl_launch_dependents();
  nw_stamp(0, p.tail == SGNN_TAIL_PX);
  if (tid == 0) {
    mbar_init(&misc->bar_w, 1);
    for (int s = 0; s < kNwAggSlots; ++s) mbar_init(&misc->agg_empty[s], 1);
    for (int g = 0; g < kNwGroups; ++g) {
      mbar_init(&misc->g[g].full, 1);
      mbar_init(&misc->g[g].patched, kThreads);
      mbar_init(&misc->g[g].acc0_ready, 1);
      mbar_init(&misc->g[g].hid_ready, kThreads);
      mbar_init(&misc->g[g].acc1_ready, 1);
      mbar_init(&misc->g[g].xo_ready, kThreads);
      mbar_init(&misc->g[g].acc2_ready, 1);
      mbar_init(&misc->g[g].region_free, kThreads);
      mbar_init(&misc->g[g].slot_free, 1);
    }
    mbar_fence_init();
    // resident weights: the head of the vector region and the three GEMM images that follow it
    const char* blob = reinterpret_cast<const char*>(p.blob);
    mbar_arrive_expect_tx(&misc->bar_w, kNwVecBytes + kNwImgBytes);
    tma_bulk_g2s(smem, blob, kNwVecBytes, &misc->bar_w);
    tma_bulk_g2s(smem + kNwImgOff, blob + SGNN_VEC_REGION_BYTES, 32768, &misc->bar_w);
    tma_bulk_g2s(smem + kNwImgOff + 32768, blob + SGNN_VEC_REGION_BYTES + 32768, 32768, &misc->bar_w);
    tensormap_prefetch(&agg_map);
    tensormap_prefetch(&x_map);
    tensormap_prefetch(&pq_map);
  }
  if (tid < 32) tmem_alloc(&misc->tmem_slot, TMEM_COLS);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  nw_stamp(1, p.tail == SGNN_TAIL_PX);
  pdl_wait();          // everything above touched only shared/tensor memory and the constant weights
  nw_stamp(2, p.tail == SGNN_TAIL_PX);
  const int n_tiles = (p.n_rows + kTileRows - 1) / kTileRows;
  const int n_my = (int)blockIdx.x < n_tiles ? (n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const uint32_t tmem = misc->tmem_slot;
  const uint32_t sU = smem_u32(smem);
  const uint32_t img = sU + kNwImgOff - SGNN_VEC_REGION_BYTES;      // image at blob byte offset `o` lives at img + o

  if (warp < 4 * kNwGroups) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kNwEpiRegs));
    mbar_wait(&misc->bar_w, 0);
    nw_epilogue(p, &pq_map, smem, misc, warp >> 2, tid & 127, n_my, tmem);
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kNwAuxRegs));
    const int role = warp - 4 * kNwGroups;
    if (role == 0) {
      nw_producer(&agg_map, &x_map, smem, misc, n_my);
    } else {
      mbar_wait(&misc->bar_w, 0);
      nw_stamp(13, p.tail == SGNN_TAIL_PX && role == 1 && (threadIdx.x & 31) == 0);
      if (role == 1) nw_mma0(sU, img + p.mlp.w_off[0], misc, n_my, tmem, p.tail == SGNN_TAIL_PX);
      else if (role == 2) nw_mma_ts(img + p.mlp.w_off[1], misc, n_my, tmem, 0);
      else nw_mma_ts(img + (p.tail == SGNN_TAIL_PX ? p.wp_off : p.dec.w_off[0]), misc, n_my, tmem, 1);
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  nw_stamp(9, p.tail == SGNN_TAIL_PX);
  if (tid < 32) tmem_dealloc(tmem, TMEM_COLS);
}

}  // namespace sgnn

using namespace sgnn;

// debug: device buffer of >= 32 uint64 receiving the stamps of node_ws_kernel (NULL = off)
extern "C" int sgnn_debug_node_ws_timeline(unsigned long long* buf) {
  SGNN_CUDA(cudaMemcpyToSymbol(g_nw_timeline, &buf, sizeof(buf)));
  return SGNN_OK;
}

static bool nw_vec_resident(int off_floats, int n) { return off_floats >= 0 && (off_floats + n) * 4 <= kNwVecBytes; }
static bool nw_img_resident(int off_bytes, int bytes) {
  return off_bytes >= SGNN_VEC_REGION_BYTES && off_bytes % 1024 == 0 &&
         off_bytes + bytes <= SGNN_VEC_REGION_BYTES + kNwImgBytes;
}

// 1 = launched, 0 = shape / mode not covered (caller falls back to node_fast_kernel), < 0 = error
int sgnn_node_ws_try(const SgnnNodeParams* p, void* stream) {
  if (getenv("SGNN_NODE_FAST")) return 0;                       // A/B switch: keep the 4-worker kernel
  if (p->prologue != SGNN_PRO_AGG_X || !p->xr_in || !p->has_mlp || !p->residual || !p->mlp.has_ln) return 0;
  if (p->agg_w != kNwW || p->x_w != kNwW || p->mlp.hidden != kNwW || p->mlp.out != kNwW || p->mlp.n_hidden != 1) return 0;
  const int wb = kNwW * kNwW * 4;
  if (!nw_vec_resident(p->mlp.b_off[0], kNwW) || !nw_vec_resident(p->mlp.b_off[1], kNwW) ||
      !nw_vec_resident(p->mlp.gamma_off, kNwW) || !nw_vec_resident(p->mlp.beta_off, kNwW))
    return 0;
  if (!nw_img_resident(p->mlp.w_off[0], 2 * wb) || !nw_img_resident(p->mlp.w_off[1], wb)) return 0;
  if (p->agg_affine && (!nw_vec_resident(p->agg_gamma_off, kNwW) || !nw_vec_resident(p->agg_beta_off, kNwW))) return 0;
  if (p->blob_bytes < SGNN_VEC_REGION_BYTES + kNwImgBytes) return 0;
  if (p->tail == SGNN_TAIL_PX) {
    if (p->pq_hidden != kNwW || !p->pq_out || !nw_vec_resident(p->pq_bias_off, kNwW) || !nw_img_resident(p->wp_off, wb))
      return 0;
    // x' replaces x in place: the rounded input must BE the Q half of the output rows
    if (p->xr_in != p->pq_out + kNwW || p->xr_ld != 2 * kNwW) return 0;
  } else if (p->tail == SGNN_TAIL_DECODE) {
    if (p->dec.hidden != kNwW || p->dec.n_hidden != 1 || p->dec.out > 8 || p->dim > 3 || !nw_img_resident(p->dec.w_off[0], wb))
      return 0;
  } else {
    return 0;
  }
  if (tensormap_encoder() == nullptr) return 0;
  CUtensorMap agg_map, x_map, pq_map;
  int rc = make_row_tile_map(&agg_map, p->agg, p->n_rows, kNwW, kNwW * 4, kTileRows);
  if (rc != SGNN_OK) return rc;
  rc = make_row_tile_map(&x_map, p->xr_in, p->n_rows, kNwW, p->xr_ld * 4, kTileRows);
  if (rc != SGNN_OK) return rc;
  pq_map = x_map;
  if (p->tail == SGNN_TAIL_PX) {
    rc = make_row_tile_map(&pq_map, p->pq_out, p->n_rows, 2 * kNwW, 2 * kNwW * 4, kTileRows);
    if (rc != SGNN_OK) return rc;
  }
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    SGNN_CUDA(cudaGetDevice(&dev));
    SGNN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  const int n_tiles = sgnn_ceil_div(p->n_rows, kTileRows);
  int grid = n_tiles < sms ? n_tiles : sms;
  if (grid < 1) grid = 1;
  SGNN_ENSURE_SMEM(node_ws_kernel, kNwSmemBytes);
  SGNN_CUDA(sgnn_launch(node_ws_kernel, dim3(grid), dim3(kNwThreads), kNwSmemBytes, (cudaStream_t)stream, *p, agg_map,
                        x_map, pq_map));
  return 1;
}
