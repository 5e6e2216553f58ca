// Edge side of one message-passing layer, warp-specialised version for 64-wide latents with one
// hidden layer (reference: InteractionNetwork.message + PyG sum-aggregation,
// sgnn/single_scale/graph_network.py:173,178-199).
//
//   m_e = LN(W2 relu(W1 [x_tgt | x_src | e] + b1) + b2),   agg_i = sum_{e: tgt = i} m_e
//
// Split of the first Linear: the receiver term is per NODE (P = x W1a^T + b1, written by the node
// kernel, read here with near-broadcast loads because edges are receiver-sorted); the sender term
// and the edge term go through the tensor core as ONE K = 128 product [e0 | x_src] [W1c | W1b]^T.
// The x_src rows are gathered with cp.async (16 bytes per lane, per-lane addresses) straight into the
// swizzled operand layout by two producer warps, so the epilogue warps issue no gather load; TMA
// tile::gather4 does the same job and is kept behind a debug flag, but it was slower end to end
// (profiles/README.md r01h).  The per-receiver sums are an in-register segmented shuffle scan
// over the 32 rows a warp holds (no shared-memory scratch).
//
// Roles (640 threads, one persistent CTA per SM; registers re-split with setmaxnreg):
//   warps 0-15  four epilogue groups of 128 threads; group g owns TMEM columns [128g, 128g + 128)
//               (accumulator 0 / hidden activations in place, accumulator 1) and every 4th tile
//   warps 16,19 producers: a tile's e0 image by TMA bulk copy and its x_src K-blocks by cp.async
//               into one of two 64 KB tile stages
//   warps 17,18 one thread each issues the tcgen05.mma of the first / the second GEMM, so that the
//               second GEMM of tile t never waits behind the loads of tile t + 1
// Hand-offs are mbarriers only (full/empty per stage; acc0_ready, hid_ready, acc1_ready,
// region_free per group); the only CTA-wide barriers are at start-up and tear-down.
#include "mlp_tile.cuh"
#include "params.h"
#include "tensormap.cuh"

namespace sgnn {

constexpr int kWsGroups = 4;
constexpr int kWsEpiThreads = kWsGroups * kThreads;        // 512
constexpr int kWsThreads = kWsEpiThreads + 128;            // + one warpgroup: producer warp, MMA warp, 2 idle
constexpr int kWsEpiRegs = 104, kWsAuxRegs = 48;           // setmaxnreg moves registers INSIDE the CTA only: launch at 96; (96-48)*128 released >= (104-96)*512 taken
constexpr int kWsW = 64;                                   // he = hidden = out
constexpr int kWsSlotBytes = kTileRows * kKBlockBytes;     // 16 KB: 128 rows x one K-block
constexpr int kWsStageBytes = 4 * kWsSlotBytes;            // one tile: e0 K-blocks 0,1 then x_src K-blocks 0,1
constexpr int kWsStages = 2;
constexpr int kWsVecBytes = 1024;                          // resident part of the blob's vector region

struct WsGroup {
  uint64_t acc0_ready, hid_ready, acc1_ready, region_free;
  int tgt_s[kTileRows];
  int piece_tgt[8];
  int piece_flag[8];
  float pieces[8 * kWsW];
};
struct WsMisc {
  uint64_t bar_w;
  uint64_t full[kWsStages], empty[kWsStages][4];     // full per tile stage, empty per K-block slot of a stage
  uint32_t tmem_slot;
  int pad;
  WsGroup g[kWsGroups];
};
// Shared memory: [vectors 1 KB | weight images | 2 tile stages | barriers + pieces].
// Only the first kWsVecBytes of the blob's 8 KB vector region are used by this layer (b2, gamma,
// beta) and only they are staged; the images follow at their blob offset minus kWsImgShift.
constexpr int kWsImgShift = SGNN_VEC_REGION_BYTES - kWsVecBytes;
struct WsSmem {
  int ring, misc, total;
};
__host__ __device__ inline WsSmem edge_ws_smem(int blob_bytes) {
  WsSmem s;
  s.ring = align_up(blob_bytes - kWsImgShift, 1024);
  s.misc = align_up(s.ring + kWsStages * kWsStageBytes, 16);
  s.total = s.misc + (int)sizeof(WsMisc) + 1024;
  return s;
}

__device__ __forceinline__ float ws_tf32r(float x) { return __uint_as_float(__float_as_uint(x) + 0x1000u); }
__device__ __forceinline__ void ws_group_bar(int g) {
  asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(kThreads) : "memory");
}

// measurement switches (sgnn_debug_ws_flags; results are WRONG when set; bits 3, 4: see the epilogue): bit 0 = sender K-blocks by
// plain bulk copy instead of gather4 (same bytes), bit 2 = epilogue stops after reading acc1
#ifdef SGNN_TIMELINE
__device__ int g_ws_flags = 0;
#define WS_FLAGS g_ws_flags
#else
#define WS_FLAGS 0          // release builds: the wrong-answer switches do not exist
#endif
// timeline of the MMA issuer of CTA 0 (SGNN_TIMELINE builds only): per tile 16 SM-clock stamps
// [region free, slot0 full, slot1 full, slot2 full, slot3 full / MMA0 committed, hid_ready seen, -, -]
__device__ unsigned long long* g_ws_timeline = nullptr;
__device__ __forceinline__ void ws_stamp(int tile_i, int k) {
#ifdef SGNN_TIMELINE
  if (g_ws_timeline && blockIdx.x == 0 && tile_i < 256 && (threadIdx.x & 31) == 0) {
    g_ws_timeline[tile_i * 16 + k] = (unsigned long long)clock64();     // (globaltimer reads cost ~0.25 us each)
  }
#endif
}

// ---------------------------------------------------------------------------------- producers
// Two warps fill a stage: `part` 0 = sender K-block 0, `part` 1 = sender K-block 1; the tile's e0 image
// (one 32 KB TMA bulk copy) is issued by the two in turn.  Sender rows are gathered with cp.async (16 bytes per lane, 8 lanes
// per 128-byte row segment, 4 rows per instruction, L2 -> shared memory without registers), written
// straight into the SWIZZLE_128B operand layout; every lane's copies arrive on the stage's mbarrier.
// TMA tile::gather4 does the same job in one instruction per 4 rows but takes its operands from
// UNIFORM registers: issued per lane it compiles to a 70-cycle ELECT / R2UR loop, issued by one lane
// from shuffled indices to ~25 dependent instructions -- 2.0 us per tile either way, the bottleneck
// of the kernel (profiles/README.md r01h); the LSU takes per-lane addresses natively.  The gather4
// path is kept behind measurement switch bit 1.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void ws_producer(const SgnnEdgeLayerParams& p, const CUtensorMap* xmap, char* ring,
                                            WsMisc* misc, int n_my, int n_edges, int lane, int part) {
  constexpr unsigned FULL = 0xffffffffu;
  const int flags = WS_FLAGS;
  const int rg = lane >> 3, ch = lane & 7;            // row within a group of 4, 16-byte chunk of the K-block
  const char* xbase = reinterpret_cast<const char*>(p.pq) + (kWsW + part * kKBlock) * 4 + ch * 16;
  // destination of (row 4j + rg, chunk ch) inside the K-block image: row * 128 + ((ch ^ (row & 7)) << 4)
  const uint32_t d_even = (uint32_t)(rg * 128 + ((ch ^ rg) << 4)), d_odd = (uint32_t)(rg * 128 + ((ch ^ (4 + rg)) << 4));
  int rows_nx[4];                                      // sender ids of rows lane, lane + 32, lane + 64, lane + 96,
  {                                                    // loaded one tile ahead (a cold read is ~1 us on this serial chain)
    // the FIRST tile's ids are fetched before the edge count is looked at (the count is itself a load issued
    // just before: one round trip instead of two at the head of the kernel's serial chain).  Slots past the
    // end of the edge list hold ids of earlier steps or zeros -- valid rows either way -- and are masked
    // once the count is known; the last tile of the capacity may be partial, so it is not read ahead.
    const long e0 = (long)blockIdx.x * kTileRows;
    const bool ahead = (long)blockIdx.x + 1 < p.max_tiles;
#pragma unroll
    for (int c = 0; c < 4; ++c) rows_nx[c] = ahead ? __ldg(p.src + e0 + c * 32 + lane) : 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const bool in = n_my > 0 && e0 + c * 32 + lane < n_edges;
      if (!ahead && in) rows_nx[c] = __ldg(p.src + e0 + c * 32 + lane);
      if (!in) rows_nx[c] = 0;
    }
  }
  for (int i = 0; i < n_my; ++i) {
    const long tile = blockIdx.x + (long)i * gridDim.x;
    int rows[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) rows[c] = rows_nx[c];
    {
      const long e0 = (tile + gridDim.x) * kTileRows;
#pragma unroll
      for (int c = 0; c < 4; ++c) rows_nx[c] = (i + 1 < n_my && e0 + c * 32 + lane < n_edges) ? __ldg(p.src + e0 + c * 32 + lane) : 0;
    }
    const int st = i & 1, use = i >> 1;
    // The tile's e0 image (two waits, expect_tx, one bulk copy: ~0.65 us of a producer's 2.2 us per tile in the
    // 1M timeline, r01p) is issued by the two producers in turn, so neither carries it on every tile.
    const bool do_e0 = part == st;
    if (use > 0) {      // a K-block slot is free as soon as ITS four MMAs are done: refills start mid-tile
      if (do_e0) {
        mbar_wait_spin(&misc->empty[st][0], (uint32_t)(use - 1) & 1u);
        mbar_wait_spin(&misc->empty[st][1], (uint32_t)(use - 1) & 1u);
      }
    }
    ws_stamp(i, 8 + 4 * part);
    char* stage = ring + st * kWsStageBytes;
    uint64_t* full = &misc->full[st];
    char* dst = stage + (2 + part) * kWsSlotBytes;
    if (use > 0 && (flags & 3)) mbar_wait_spin(&misc->empty[st][2 + part], (uint32_t)(use - 1) & 1u);
    if (do_e0 || (flags & 3)) {
      if (elect_one()) {
        uint32_t tx = do_e0 ? 2 * kWsSlotBytes : 0;
        if (flags & 3) tx += kWsSlotBytes;
        mbar_arrive_expect_tx(full, tx);
        if (do_e0)
          tma_bulk_g2s(stage, reinterpret_cast<const char*>(p.e_tiles) + tile * (2 * kWsSlotBytes), 2 * kWsSlotBytes, full);
        if (flags & 1)
          tma_bulk_g2s(dst, reinterpret_cast<const char*>(p.e_tiles) + tile * (2 * kWsSlotBytes) + part * kWsSlotBytes,
                       kWsSlotBytes, full);
      }
      __syncwarp();
    }
    if (use > 0) mbar_wait_spin(&misc->empty[st][2 + part], (uint32_t)(use - 1) & 1u);
    ws_stamp(i, 9 + 4 * part);
    if (flags & 1) {
      // sender K-block replaced by a bulk copy of the same size (measurement only)
    } else if (flags & 2) {
#pragma unroll 4
      for (int j = 0; j < 32; ++j) {
        const int c = j >> 3, l0 = (4 * j) & 31;
        const int r0 = __shfl_sync(FULL, rows[c], l0), r1 = __shfl_sync(FULL, rows[c], l0 + 1);
        const int r2 = __shfl_sync(FULL, rows[c], l0 + 2), r3 = __shfl_sync(FULL, rows[c], l0 + 3);
        if (elect_one()) tma_gather4(dst + j * 512, xmap, part * kKBlock, r0, r1, r2, r3, full);
      }
      __syncwarp();
    } else {
      const uint32_t d0 = smem_u32(dst);
      // Row offsets in 16-byte units, computed by the lane that LOADED the id, so that the per-copy chain is
      // shuffle -> one IMAD.WIDE onto the 64-bit base -> LDGSTS.  Shuffling the ids and scaling afterwards
      // compiled to a 6-instruction dependent chain per copy through one register pair (48-register budget).
      // (Measured equal at 1M, r01p: the copies are paced by the LSU queue, ~60 cycles per LDGSTS.128.)
      uint32_t off16[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) off16[c] = (uint32_t)rows[c] * (uint32_t)(2 * kWsW * 4 / 16);
      const char* xb = xbase;
      asm volatile("" : "+l"(xb));                   // keep the 64-bit base in a register pair (not re-derived from the constant bank per copy)
#pragma unroll
      for (int jb = 0; jb < 32; jb += 4) {           // shuffles first, then the copies: the volatile copies pin the order
        uint32_t o[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) o[u] = __shfl_sync(FULL, off16[(jb + u) >> 3], (4 * (jb + u) + rg) & 31);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          cp_async16(d0 + (jb + u) * 512 + (((jb + u) & 1) ? d_odd : d_even), xb + (size_t)o[u] * 16);
      }
    }
    ws_stamp(i, 10 + 4 * part);
    if (!(flags & 3)) {
      cp_async_arrive_noinc(full);       // counts once this lane's copies have landed
    } else if (do_e0 || lane != 0) {
      mbar_arrive(full);                 // same arrival count as the cp.async path (the warp without e0: lane 0 arrived with expect_tx)
    }
  }
}

// ---------------------------------------------------------------------------------- MMA issuers
// Two warps, each BLOCKING on its own mbarriers; the tcgen05 instructions sit under elect.sync in
// converged code so that their operands stay in uniform registers.  One wait and one commit per
// tile: every wait / fence / commit on this serial chain costs 100-200 cycles.
// First GEMM: acc0 = [e0 | x_src] [W1c | W1b]^T, K = 128 = the four K-blocks of a stage.
__device__ __forceinline__ void ws_mma0(const SgnnEdgeLayerParams& p, uint32_t ring_u32, uint32_t smem_base_u32,
                                        WsMisc* misc, int n_my, uint32_t tmem, int wxs_off) {
  const uint32_t idesc = umma_idesc_tf32(kTileRows, kWsW);
  const uint32_t w_e = smem_base_u32 + p.mlp.w_off[0] - kWsImgShift, w_x = smem_base_u32 + wxs_off - kWsImgShift;
  for (int a = 0; a < n_my; ++a) {
    if (a >= kWsGroups) mbar_wait_spin(&misc->g[a & 3].region_free, (uint32_t)((a >> 2) - 1) & 1u);
    ws_stamp(a, 0);
    const uint32_t d = tmem + (uint32_t)((a & 3) * 2 * kWsW);
    const int st = a & 1;
    mbar_wait_spin(&misc->full[st], (uint32_t)(a >> 1) & 1u);
    fence_proxy_async_smem();          // cp.async wrote through the generic proxy, tcgen05.mma reads through the async one
    tc_fence_after_sync();
    ws_stamp(a, 1);
    const uint32_t a_st = ring_u32 + st * kWsStageBytes;
    if (elect_one()) {
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) {
        const uint32_t a_kb = a_st + kb * kWsSlotBytes;
        const uint32_t w_kb = (kb < 2 ? w_e : w_x) + (uint32_t)((kb & 1) * kWsW * kKBlockBytes);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          umma_tf32(d, umma_desc_sw128(a_kb + j * 32), umma_desc_sw128(w_kb + j * 32), idesc,
                    (kb > 0 || j > 0) ? 1u : 0u);
        tc_commit(&misc->empty[st][kb]);
      }
      tc_commit(&misc->g[a & 3].acc0_ready);
    }
    __syncwarp();
    ws_stamp(a, 2);
  }
}
// Second GEMM: acc1 = hidden W2^T with the hidden activations read from TMEM.
__device__ __forceinline__ void ws_mma1(const SgnnEdgeLayerParams& p, uint32_t smem_base_u32, WsMisc* misc, int n_my,
                                        uint32_t tmem) {
  const uint32_t w_2 = smem_base_u32 + p.mlp.w_off[1] - kWsImgShift;
  for (int b = 0; b < n_my; ++b) {
    mbar_wait_spin(&misc->g[b & 3].hid_ready, (uint32_t)(b >> 2) & 1u);
    tc_fence_after_sync();
    ws_stamp(b, 5);
    const uint32_t base = tmem + (uint32_t)((b & 3) * 2 * kWsW);
    if (elect_one()) {
      umma_gemm_tf32_ts(base + kWsW, base, w_2, kWsW, kWsW);
      tc_commit(&misc->g[b & 3].acc1_ready);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------- epilogue
// Thread wt of group g owns row wt of the group's tiles (TMEM lane wt).
//
// Code size matters here: the whole per-tile path must stay inside the 32 KB SM instruction cache.
// The first version of this epilogue was 54 KB of unrolled code (36 KB of it the segmented sums);
// with 16 warps in different places the instruction-cache hit rate was 52 % and the GPC-level
// instruction cache ran at 94 % of its request peak -- "no instruction" was 46 % of all stall
// samples (profiles/README.md r01h).  Hence: branch-free predicated segment scan, packed fp32x2
// arithmetic for the LayerNorm, no duplicated special cases.
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }

template <bool DEFER>
__device__ __forceinline__ void ws_epilogue(const SgnnEdgeLayerParams& p, const float* blob_f, WsMisc* misc,
                                            int g, int wt, int n_my, int n_edges, uint32_t tmem) {
  constexpr int W = kWsW;
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = wt & 31, wq = wt >> 5;
  WsGroup* wk = &misc->g[g];
  const uint32_t lane_base = (uint32_t)(wq * 32) << 16;
  const uint32_t t0 = tmem + (uint32_t)(g * 2 * W) + lane_base, t1 = t0 + W;
  const int stride_tiles = gridDim.x * kWsGroups;
  const bool poll = false;      // polling epilogue warps measured slower even with one tile per group (r01n)
  const float4* b4 = reinterpret_cast<const float4*>(blob_f + p.mlp.b_off[1]);
  // The LayerNorm affine is applied AFTER the sums: sum_e (z_e * gamma + beta) = gamma * sum_e z_e +
  // count * beta, per receiver (and per piece of a receiver, the pieces add up).  In the scan a lane
  // owns columns 2*lane, 2*lane + 1.

  // receiver of this thread's row, loaded one tile ahead (a cold read costs ~1 us on the group's serial chain)
  // (with the receivers of the rows before and after it: the segment structure then needs no load)
  int t_nx = -3, tb_nx = -2, ta_nx = -3;
  {
    // as in the producers: the first tile's receiver ids are fetched before the edge count is looked at
    const long tile0 = (long)blockIdx.x + (long)g * gridDim.x;
    const long e = tile0 * kTileRows + wt;
    const bool ahead = tile0 + 1 < p.max_tiles;
    int a = 0, b = 0, c = 0;
    if (ahead) {
      a = __ldg(p.tgt + e);
      b = e >= 1 ? __ldg(p.tgt + e - 1) : 0;
      c = __ldg(p.tgt + e + 1);
    }
    if (g < n_my) {
      if (e < n_edges) t_nx = ahead ? a : __ldg(p.tgt + e);
      if (e - 1 >= 0 && e - 1 < n_edges) tb_nx = ahead ? b : __ldg(p.tgt + e - 1);
      if (e + 1 < n_edges) ta_nx = ahead ? c : __ldg(p.tgt + e + 1);
    }
  }

#pragma unroll 1
  for (int i = g; i < n_my; i += kWsGroups) {
    const uint32_t par = (uint32_t)(i >> 2) & 1u;
    const int tile = blockIdx.x + i * gridDim.x;
    const int e0 = tile * kTileRows;
    const int nvalid = n_edges - e0 < kTileRows ? n_edges - e0 : kTileRows;
    const bool valid = wt < nvalid;
    const int t = t_nx, t_prev = tb_nx, t_next = ta_nx;
    {
      const long en = (long)(tile + stride_tiles) * kTileRows + wt;
      const bool more = i + kWsGroups < n_my;
      t_nx = (more && en < n_edges) ? __ldg(p.tgt + en) : -3;
      tb_nx = (more && en - 1 < n_edges) ? __ldg(p.tgt + en - 1) : -2;
      ta_nx = (more && en + 1 < n_edges) ? __ldg(p.tgt + en + 1) : -3;
    }

    // ---- segment structure of this warp's 32 rows (warp-uniform) ----
    const int nrows = nvalid - wq * 32 < 0 ? 0 : (nvalid - wq * 32 > 32 ? 32 : nvalid - wq * 32);
    // bit r: row r starts a new receiver; rows past the end of the edge list never do
    const unsigned starts = __ballot_sync(FULL, t != t_prev) & (nrows == 32 ? FULL : ((1u << nrows) - 1u));
    const int t_first = __shfl_sync(FULL, t, 0);
    const int t_last = __shfl_sync(FULL, t, nrows > 0 ? nrows - 1 : 0);
    const int t_after = __shfl_sync(FULL, t_next, nrows > 0 ? nrows - 1 : 0);
    const bool first_cont = nrows > 0 && !(starts & 1u);      // row 0 continues the previous warp's receiver
    const bool cont_next = nrows > 0 && t_after == t_last;    // the last receiver continues in the next warp
    const bool one_piece = first_cont && starts == 0u;        // the whole warp is the middle of one receiver

    // ---- P[tgt]: W fp32 per row; consecutive rows share receivers, so a warp touches few lines ----
    float pr[W];
    {
      // measurement switch bit 3: every row reads receiver 0 (one L1 line per load instruction instead of one per
      // distinct receiver of the warp); wrong results
      // 32-byte loads: half the load instructions, each touching one L1 line per distinct receiver of the warp
      const float* prow = p.pq + (long)((valid && !(WS_FLAGS & 8)) ? t : 0) * 2 * W;
#pragma unroll
      for (int j = 0; j < W / 8; ++j) ldg256_nc(prow + 8 * j, pr + 8 * j);
    }

    // ---- hidden = relu(acc0 + P), rounded to TF32, back into the same TMEM columns ----
    mbar_wait_mode(&wk->acc0_ready, par, poll);
    tc_fence_after_sync();
    if (wt == 0) ws_stamp(i, 6);
#pragma unroll
    for (int c = 0; c < W; c += 32) {
      float v[32];
      tmem_ld32(t0 + c, v);
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        const float2 s = __fadd2_rn(f2(v[j], v[j + 1]), f2(pr[c + j], pr[c + j + 1]));
        v[j] = ws_tf32r(fmaxf(s.x, 0.f));
        v[j + 1] = ws_tf32r(fmaxf(s.y, 0.f));
      }
      tmem_st32(t0 + c, v);
    }
    tmem_st_wait();
    tc_fence_before_sync();
    mbar_arrive(&wk->hid_ready);

    {   // while the second GEMM runs: pull the P rows of this group's next tile into L2 (one lane per
        // receiver: a prefetch instruction issues once per active lane)
      const int tp = __shfl_up_sync(FULL, t_nx, 1);
      if (t_nx >= 0 && (lane == 0 || t_nx != tp)) {
        const char* pn = reinterpret_cast<const char*>(p.pq + (long)t_nx * 2 * W);
        asm volatile("prefetch.global.L2 [%0];" ::"l"(pn));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(pn + 128));
      }
    }

    // ---- output Linear's accumulator -> registers; the TMEM region is free again at once ----
    mbar_wait_mode(&wk->acc1_ready, par, poll);
    if (wt == 0) ws_stamp(i, 3);
    tc_fence_after_sync();
    float y[W];
    tmem_ld64(t1, y);
    tc_fence_before_sync();
    mbar_arrive(&wk->region_free);
    if (WS_FLAGS & 4) {      // measurement switch: skip LayerNorm, sums and stitch (wrong results)
      if (y[0] == 123.456f) p.agg[wt] = y[1];
      continue;
    }
    {   // bias + LayerNorm without its affine, in packed fp32x2
      float2 s01 = f2(0.f, 0.f), s23 = f2(0.f, 0.f);
#pragma unroll
      for (int j = 0; j < W / 4; ++j) {
        const float4 b = b4[j];
        const float2 a = __fadd2_rn(f2(y[4 * j], y[4 * j + 1]), f2(b.x, b.y));
        const float2 c = __fadd2_rn(f2(y[4 * j + 2], y[4 * j + 3]), f2(b.z, b.w));
        y[4 * j] = a.x; y[4 * j + 1] = a.y; y[4 * j + 2] = c.x; y[4 * j + 3] = c.y;
        s01 = __fadd2_rn(s01, a);
        s23 = __fadd2_rn(s23, c);
      }
      const float nmean = -(((s01.x + s01.y) + (s23.x + s23.y)) / (float)W);
      const float2 nm = f2(nmean, nmean);
      float2 q01 = f2(0.f, 0.f), q23 = f2(0.f, 0.f);
#pragma unroll
      for (int j = 0; j < W / 4; ++j) {
        const float2 a = __fadd2_rn(f2(y[4 * j], y[4 * j + 1]), nm);
        const float2 c = __fadd2_rn(f2(y[4 * j + 2], y[4 * j + 3]), nm);
        y[4 * j] = a.x; y[4 * j + 1] = a.y; y[4 * j + 2] = c.x; y[4 * j + 3] = c.y;
        q01 = __ffma2_rn(a, a, q01);
        q23 = __ffma2_rn(c, c, q23);
      }
      // rows past the end of the edge list contribute zeros to the sums below
      const float rstd = valid ? 1.0f / sqrtf(((q01.x + q01.y) + (q23.x + q23.y)) / (float)W + 1e-5f) : 0.f;
      const float2 r2 = f2(rstd, rstd);
#pragma unroll
      for (int j = 0; j < W / 2; ++j) {
        const float2 a = __fmul2_rn(f2(y[2 * j], y[2 * j + 1]), r2);
        y[2 * j] = a.x; y[2 * j + 1] = a.y;
      }
    }

    if (wt == 0) ws_stamp(i, 4);
    // ---- segmented sums over this warp's 32 rows, in place: thread = row keeps its 64 columns, a
    //      Hillis-Steele scan over the lanes (shfl.up by 1, 2, 4, ... up to the longest receiver of the
    //      warp) leaves every receiver's sum in its LAST row, whose thread applies the LayerNorm affine
    //      and writes the receiver's 256-byte agg row (or the warp's head / tail piece).  64 independent
    //      chains per level instead of one 32-step serial walk per column pair behind a shared-memory
    //      transposition: in-kernel stamps put that walk at 2.8 us per tile, 40 % of the kernel at 4800
    //      particles and the bound of the epilogue at 1M (r01n).  Deterministic (fixed tree order).
    //      (The shared piece slots and flags are written only here, after acc1_ready: warp 0 arrived
    //      on hid_ready after it finished stitching the previous tile, so nothing it reads is touched.)
    if (lane == 0) {
      wk->piece_tgt[2 * wq] = t_first;
      wk->piece_flag[2 * wq] = first_cont ? ((one_piece && cont_next) ? 3 : 1) : 0;
      wk->piece_tgt[2 * wq + 1] = t_last;
      wk->piece_flag[2 * wq + 1] = (cont_next && !one_piece) ? 1 : 0;
    }
    if (nrows > 0) {
      // first row of this lane's receiver within the warp (row 0 opens a segment for the scan in any case)
      const unsigned below = (starts | 1u) & (FULL >> (31 - lane));
      const int seg0 = 31 - __clz((int)below);
      const bool in_rows = lane < nrows;
      const bool is_end = in_rows && (lane == nrows - 1 || ((starts >> (lane + 1)) & 1u));
      const int cnt_i = lane - seg0 + 1;
      const int maxlen = __reduce_max_sync(FULL, is_end ? cnt_i : 0);
#pragma unroll 1
      for (int d = 1; d < maxlen; d <<= 1) {
        const float tk = lane - d >= seg0 ? 1.0f : 0.0f;
        const float2 tk2 = f2(tk, tk);
#pragma unroll
        for (int j = 0; j < W; j += 2) {           // two shuffles + ONE packed FMA: y += up * {1, 0} (exact; a predicated
          const float2 up = f2(__shfl_up_sync(FULL, y[j], d), __shfl_up_sync(FULL, y[j + 1], d));   // add costs two)
          const float2 r = __ffma2_rn(up, tk2, f2(y[j], y[j + 1]));
          y[j] = r.x; y[j + 1] = r.y;
        }
      }
      if (is_end && !(WS_FLAGS & 16)) {          // measurement switch bit 4: no agg rows written (wrong results)
        const bool head_piece = first_cont && seg0 == 0;                        // continues the previous warp's receiver
        const bool tail_piece = cont_next && !one_piece && lane == nrows - 1;   // continues in the next warp
        if constexpr (DEFER) {
          // The node kernel applies gamma / count * beta to the receiver's complete sum.  The sums pass through one
          // packed multiply by an opaque 1.0f: with NO arithmetic between the scan and the stores ptxas spilled 388
          // bytes around the scan in this instantiation (64 with the affine, 68 with this), and the kernel ran 30 %
          // SLOWER than with the affine (0.77 vs 0.59 ms at 1M, r02c).  x * 1.0f is exact.
          float one = 1.0f;
          asm volatile("" : "+f"(one));
          const float2 o2 = f2(one, one);
#pragma unroll
          for (int j = 0; j < W / 2; ++j) {
            const float2 a = __fmul2_rn(f2(y[2 * j], y[2 * j + 1]), o2);
            y[2 * j] = a.x; y[2 * j + 1] = a.y;
          }
        }
        if constexpr (!DEFER) {                    // defer_affine: the node kernel applies gamma / count * beta to the
          const float cnt = (float)cnt_i;          // receiver's complete sum (one pass per NODE row there instead of this
          const float4* g4 = reinterpret_cast<const float4*>(blob_f + p.mlp.gamma_off);   // divergent block per edge tile:
          const float4* be4 = reinterpret_cast<const float4*>(blob_f + p.mlp.beta_off);   // -19 % at 1M, r01p switch bit 4)
          const float2 c2 = f2(cnt, cnt);
#pragma unroll
          for (int j = 0; j < W / 4; ++j) {        // packed fp32x2 (same roundings as fmaf(y, gamma, cnt * beta)): every
            const float4 ga = g4[j], be = be4[j];  // instruction of this block costs ~7 cycles of a contended scheduler
            const float2 a = __ffma2_rn(f2(y[4 * j], y[4 * j + 1]), f2(ga.x, ga.y), __fmul2_rn(c2, f2(be.x, be.y)));
            const float2 b = __ffma2_rn(f2(y[4 * j + 2], y[4 * j + 3]), f2(ga.z, ga.w), __fmul2_rn(c2, f2(be.z, be.w)));
            y[4 * j] = a.x; y[4 * j + 1] = a.y; y[4 * j + 2] = b.x; y[4 * j + 3] = b.y;
          }
        }
        if (!tail_piece && !head_piece && !(WS_FLAGS & 32)) {   // (measurement switch bit 5: rows go to shared memory)
          // a receiver's agg row leaves as eight 32-byte stores: every store fills a whole sector (sixteen
          // 16-byte stores per row cost 19 % of the kernel at 1M: r01p, switch bit 4)
          float* dst = p.agg + (long)t * W;
#pragma unroll
          for (int j = 0; j < W / 8; ++j)
            stg256(dst + 8 * j, y[8 * j], y[8 * j + 1], y[8 * j + 2], y[8 * j + 3], y[8 * j + 4], y[8 * j + 5],
                   y[8 * j + 6], y[8 * j + 7]);
        } else {
          float4* dst = reinterpret_cast<float4*>(&wk->pieces[(2 * wq + (tail_piece ? 1 : 0)) * W]);
#pragma unroll
          for (int j = 0; j < W / 4; ++j) dst[j] = make_float4(y[4 * j], y[4 * j + 1], y[4 * j + 2], y[4 * j + 3]);
        }
      }
    }
    if (wt == 0) ws_stamp(i, 11);
    ws_group_bar(g);
    if (wt == 0) ws_stamp(i, 7);

    // ---- stitch the pieces of CSR rows that straddle warps / tiles (warp 0 of the group).  The
    //      next tile's pieces are written only after all 128 threads arrived on hid_ready, i.e.
    //      after this warp left the stitch. ----
    if (wq == 0) {
      float carry[W / 32];
      bool has = false, from_prev = false;
      int ctgt = -1;
#pragma unroll 1
      for (int q = 0; q < 4; ++q) {
        const int f0 = wk->piece_flag[2 * q];
        if (f0) {
          if (!has) {
            has = true; from_prev = true; ctgt = wk->piece_tgt[2 * q];
#pragma unroll
            for (int j = 0; j < W / 32; ++j) carry[j] = 0.f;
          }
#pragma unroll
          for (int j = 0; j < W / 32; ++j) carry[j] += wk->pieces[(2 * q) * W + 32 * j + lane];
          if (!(f0 & 2)) {
            float* dst = from_prev ? p.head + (long)tile * W : p.agg + (long)ctgt * W;
#pragma unroll
            for (int j = 0; j < W / 32; ++j) dst[32 * j + lane] = carry[j];
            has = false;
          }
        }
        if (wk->piece_flag[2 * q + 1]) {
          has = true; from_prev = false; ctgt = wk->piece_tgt[2 * q + 1];
#pragma unroll
          for (int j = 0; j < W / 32; ++j) carry[j] = wk->pieces[(2 * q + 1) * W + 32 * j + lane];
        }
      }
      if (has) {   // the row continues into the next tile: its remainder arrives as that tile's head
        float* dst = from_prev ? p.head + (long)tile * W : p.agg + (long)ctgt * W;
#pragma unroll
        for (int j = 0; j < W / 32; ++j) dst[32 * j + lane] = carry[j];
      }
      __syncwarp();
    }
  }
}

// ---------------------------------------------------------------------------------- kernel
template <bool DEFER>
__global__ void __launch_bounds__(kWsThreads, 1)
edge_layer_ws_kernel(const SgnnEdgeLayerParams p, const __grid_constant__ CUtensorMap xmap) {
  constexpr int TMEM_COLS = 512;
  char* smem = smem_base_1024();
  const WsSmem L = edge_ws_smem(p.blob_bytes);
  WsMisc* misc = reinterpret_cast<WsMisc*>(smem + L.misc);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  pdl_launch_dependents();
  if (tid == 0) ws_stamp(255, 0);              // kernel-level stamps live in row 255
  if (tid == 0) {
    mbar_init(&misc->bar_w, 1);
    for (int s = 0; s < kWsStages; ++s) {
      mbar_init(&misc->full[s], 65);           // 2 x 32 producer lanes + the expect_tx arrival of the e0 copy
      for (int k = 0; k < 4; ++k) mbar_init(&misc->empty[s][k], 1);
    }
    for (int g = 0; g < kWsGroups; ++g) {
      mbar_init(&misc->g[g].acc0_ready, 1);
      mbar_init(&misc->g[g].hid_ready, kThreads);
      mbar_init(&misc->g[g].acc1_ready, 1);
      mbar_init(&misc->g[g].region_free, kThreads);
    }
    mbar_fence_init();
    mbar_arrive_expect_tx(&misc->bar_w, (uint32_t)(p.blob_bytes - kWsImgShift));
    tma_bulk_g2s(smem, p.blob, kWsVecBytes, &misc->bar_w);
    for (int off = SGNN_VEC_REGION_BYTES; off < p.blob_bytes; off += 16384)
      tma_bulk_g2s(smem + off - kWsImgShift, reinterpret_cast<const char*>(p.blob) + off,
                   (uint32_t)(p.blob_bytes - off < 16384 ? p.blob_bytes - off : 16384), &misc->bar_w);
    tensormap_prefetch(&xmap);
  }
  if (tid < 32) tmem_alloc(&misc->tmem_slot, TMEM_COLS);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  if (tid == 0) ws_stamp(255, 1);
  pdl_wait();          // everything above touched only shared/tensor memory and the constant weights
  if (tid == 0) ws_stamp(255, 2);
  const int n_edges = p.rowptr[p.n_nodes];
  const int n_tiles = (n_edges + kTileRows - 1) / kTileRows;
  const int n_my = (int)blockIdx.x < n_tiles ? (n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const uint32_t tmem = misc->tmem_slot;
  if (tid == 0 && n_edges >= 0) ws_stamp(255, 3);

  if (warp < 16) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kWsEpiRegs));
    mbar_wait(&misc->bar_w, 0);
    ws_epilogue<DEFER>(p, reinterpret_cast<const float*>(smem), misc, warp >> 2, tid & 127, n_my, n_edges, tmem);
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kWsAuxRegs));
  }
  if (warp == 16) {
    ws_producer(p, &xmap, smem + L.ring, misc, n_my, n_edges, lane, 0);
  } else if (warp == 19) {
    ws_producer(p, &xmap, smem + L.ring, misc, n_my, n_edges, lane, 1);
  } else if (warp == 17) {
    mbar_wait(&misc->bar_w, 0);
    ws_mma0(p, smem_u32(smem + L.ring), smem_u32(smem), misc, n_my, tmem, p.wxs_off);
  } else if (warp == 18) {
    mbar_wait(&misc->bar_w, 0);
    ws_mma1(p, smem_u32(smem), misc, n_my, tmem);
  }
  tc_fence_before_sync();
  __syncthreads();
  if (tid == 0) ws_stamp(255, 4);
  if (tid < 32) tmem_dealloc(tmem, TMEM_COLS);
}

}  // namespace sgnn

using namespace sgnn;

// 1 if the layer can run in gather mode (the node kernels must then put the TF32-rounded node
// latents in the Q half of the pq buffer: tail SGNN_TAIL_PX)
extern "C" int sgnn_edge_layer_gather_supported(int he, int hidden, int out, int n_hidden, int blob_bytes) {
  if (!(he == kWsW && hidden == kWsW && out == kWsW && n_hidden == 1)) return 0;
  const WsSmem s = edge_ws_smem(blob_bytes);
  return s.total <= 227 * 1024 && tensormap_encoder() != nullptr;
}

extern "C" int sgnn_debug_ws_timeline(void* buf) {
  SGNN_CUDA(cudaMemcpyToSymbol(g_ws_timeline, &buf, sizeof(buf)));
  return SGNN_OK;
}
extern "C" int sgnn_debug_ws_flags(int flags) {
#ifdef SGNN_TIMELINE
  SGNN_CUDA(cudaMemcpyToSymbol(g_ws_flags, &flags, sizeof(flags)));
  return SGNN_OK;
#else
  SGNN_CHECK_ARG(flags == 0, "debug_ws_flags: the measurement switches exist in SGNN_TIMELINE builds only");
  return SGNN_OK;
#endif
}

int sgnn_edge_layer_gather_launch(const SgnnEdgeLayerParams* p, void* stream) {
  SGNN_CHECK_ARG(p->he == kWsW && p->mlp.hidden == kWsW && p->mlp.out == kWsW && p->mlp.n_hidden == 1,
                 "edge_layer (gather mode): widths must be 64 with one hidden layer");
  const WsSmem s = edge_ws_smem(p->blob_bytes);
  SGNN_CHECK_ARG(s.total <= 227 * 1024, "edge_layer (gather mode): needs %d bytes of shared memory", s.total);
  SGNN_CHECK_ARG((p->mlp.b_off[1] + kWsW) * 4 <= kWsVecBytes && (p->mlp.gamma_off + kWsW) * 4 <= kWsVecBytes &&
                     (p->mlp.beta_off + kWsW) * 4 <= kWsVecBytes,
                 "edge_layer (gather mode): the layer's vectors must lie in the first %d bytes of the blob", kWsVecBytes);
  SGNN_CHECK_ARG(p->mlp.w_off[0] >= SGNN_VEC_REGION_BYTES && p->mlp.w_off[1] >= SGNN_VEC_REGION_BYTES &&
                     p->wxs_off >= SGNN_VEC_REGION_BYTES, "edge_layer (gather mode): image offsets inside the vector region");
  CUtensorMap map;
  const int rc = make_row_gather_map(&map, p->pq + kWsW, p->n_nodes, kWsW, 2 * kWsW * 4);
  if (rc != SGNN_OK) return rc;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    SGNN_CUDA(cudaGetDevice(&dev));
    SGNN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  int grid = p->max_tiles < sms ? p->max_tiles : sms;
  if (grid < 1) grid = 1;
  if (p->defer_affine) {
    SGNN_ENSURE_SMEM(edge_layer_ws_kernel<true>, s.total);
    SGNN_CUDA(sgnn_launch(edge_layer_ws_kernel<true>, dim3(grid), dim3(kWsThreads), s.total, (cudaStream_t)stream, *p, map));
  } else {
    SGNN_ENSURE_SMEM(edge_layer_ws_kernel<false>, s.total);
    SGNN_CUDA(sgnn_launch(edge_layer_ws_kernel<false>, dim3(grid), dim3(kWsThreads), s.total, (cudaStream_t)stream, *p, map));
  }
  return SGNN_OK;
}
