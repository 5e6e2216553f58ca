// Edge side of one message-passing layer (InteractionNetwork.message + PyG sum-aggregation,
// sgnn/single_scale/graph_network.py:178-199; G2M/M2M/M2G blocks, multi_scale_gnn.py:84-205).
//
//   m_e   = LN( MLP( [x_tgt | x_src | 2^l e0] ) )          reference: concat then Linear
//         = LN( MLP'( e0 (2^l W1c)^T + P[tgt] + Q[src] ) )  here: P = x W1a^T + b1, Q = x W1b^T
//   agg_i = sum_{e : tgt(e) = i} m_e                        (deterministic, CSR row order)
//
// One persistent CTA of 128 threads per SM slot walks 128-edge tiles of the receiver-sorted
// edge array.  Per tile: TMA bulk load of the e0 tile (already in A-operand layout) ->
// tcgen05 GEMMs with the layer's weights resident in shared memory -> thread-per-row epilogues
// (gather P/Q, ReLU, LayerNorm) -> segmented column sums -> agg rows.  Neither the [E,3H]
// concat, nor the hidden activations, nor the per-edge messages ever touch HBM, and because of
// the reference's e' = 2e quirk (graph_network.py:173-176) no edge latent is written at all.
//
// Algorithmic HBM bytes per edge: 4*he (e0) + 8 (src,tgt) ; per node 4*out (agg) ; P/Q rows
// are gathers that mostly hit L2 (counted once per node: 8*hidden).
#include <stdlib.h>

#include "mlp_tile.cuh"
#include "params.h"

namespace sgnn {

struct EdgeMisc {
  uint64_t bar_w, bar_e, bar_mma, bar_stream;
  uint32_t tmem_slot;
  int n_edges;
  int tgt_s[kTileRows + 4];     // [0] = receiver of the edge before the tile, [1+r] = row r
  int piece_tgt[8];
  int piece_flag[8];            // [2g] head piece of warp g (1 = exists, 3 = runs through), [2g+1] tail
  float pieces[8 * 128];
  float xch[2 * kMaxParts * kTileRows];     // LayerNorm moments between the parts of a row
};

struct EdgeLayerSmem {
  int blob, wbuf, e_tile, hid, misc, simt, total;
};
__host__ __device__ inline EdgeLayerSmem edge_layer_smem(int blob_bytes, int he, int hidden, int out,
                                                         bool tc, bool streamed) {
  EdgeLayerSmem s;
  const int w = hidden > out ? hidden : out;
  s.blob = 0;
  s.wbuf = align_up(blob_smem_bytes(blob_bytes, streamed), 1024);
  s.e_tile = s.wbuf + (streamed ? wbuf_bytes(tc) : 0);
  s.hid = s.e_tile + kTileRows * he * 4;
  s.misc = s.hid + kTileRows * w * 4;
  s.simt = align_up(s.misc + (int)sizeof(EdgeMisc), 1024);
  s.total = s.simt + (tc ? 0 : kTileRows * pow2_cols(w) * 4) + 1024;  // +1024: base alignment
  return s;
}
// resident weights when everything fits in 227 KB, else streamed
__host__ inline bool edge_layer_streamed(int blob_bytes, int he, int hidden, int out, bool tc) {
  return edge_layer_smem(blob_bytes, he, hidden, out, tc, false).total > 227 * 1024;
}

__device__ __forceinline__ float m_tile_read(const float* m, int row, int col, int width) {
  return m[row * width + ((((col >> 2) ^ (row & 7))) << 2) + (col & 3)];
}

// optional in-kernel timeline (SGNN_TIMELINE builds; profiles/node_generic_timeline.py): %globaltimer stamps of CTA 0
__device__ unsigned long long* g_edge_generic_timeline = nullptr;
__device__ __forceinline__ void estamp(int slot) {
#ifdef SGNN_TIMELINE
  if (g_edge_generic_timeline != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    g_edge_generic_timeline[slot] = t;
  }
#else
  (void)slot;
#endif
}

template <bool TC>
__global__ void __launch_bounds__(kThreads * kMaxParts) edge_layer_kernel(const SgnnEdgeLayerParams p,
                                                              const int streamed) {
  pdl_launch_dependents();
  estamp(0);
  char* smem = smem_base_1024();
  const int hidden = p.mlp.hidden, out = p.mlp.out, he = p.he, nh = p.mlp.n_hidden;
  const int w = hidden > out ? hidden : out;
  const EdgeLayerSmem L = edge_layer_smem(p.blob_bytes, he, hidden, out, TC, streamed != 0);
  const float* blob_f = reinterpret_cast<const float*>(smem + L.blob);
  char* e_tile = smem + L.e_tile;
  char* hid = smem + L.hid;
  EdgeMisc* misc = reinterpret_cast<EdgeMisc*>(smem + L.misc);
  const int tid = t_row(), lane = threadIdx.x & 31, warp = (threadIdx.x >> 5) & 3;   // row; lane; warp within its part
  const int part = t_part();
  const bool t0 = threadIdx.x == 0;
  const int tile_bytes = kTileRows * he * 4;

  // the blob is constant: its vector region and (streamed weights) the first image are in flight before
  // griddepcontrol.wait; the edge count and the e0 tiles come from the kernels before this one
  WeightStage ws;
  if (t0) mbar_init(&misc->bar_e, 1);
  weights_setup(ws, p.blob, p.blob_bytes, smem + L.blob, smem + L.wbuf, &misc->bar_w,
                &misc->bar_stream, streamed);
  TileAcc<TC> acc;
  acc_setup<TC>(acc, TC ? pow2_cols(2 * w) : pow2_cols(w), &misc->bar_mma, &misc->tmem_slot,
                reinterpret_cast<float*>(smem + L.simt));
  weights_prefetch<TC>(ws, p.mlp.w_off[0], he, hidden);
  estamp(1);
  pdl_wait();
  const int n_edges = p.rowptr[p.n_nodes];
  const int n_tiles = (n_edges + kTileRows - 1) / kTileRows;
  mbar_wait(&misc->bar_w, 0);
  estamp(2);
  if ((int)blockIdx.x >= n_tiles) {         // uniform per CTA
    weights_prefetch_drain(ws);
    acc_teardown<TC>(acc);
    return;
  }
  if (t0)
    tma_load_span(e_tile, reinterpret_cast<const char*>(p.e_tiles) + (long)blockIdx.x * tile_bytes,
                  tile_bytes, &misc->bar_e);
  uint32_t phase_e = 0;

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int e0 = tile * kTileRows;
    const int nvalid = n_edges - e0 < kTileRows ? n_edges - e0 : kTileRows;
    const int eidx = e0 + tid;
    const int s = tid < nvalid ? p.src[eidx] : -1;
    const int t = tid < nvalid ? p.tgt[eidx] : -3;
    if (part == 0) misc->tgt_s[1 + tid] = t;
    if (t0) {
      misc->tgt_s[0] = tile > 0 ? p.tgt[e0 - 1] : -2;
      misc->tgt_s[1 + kTileRows] = e0 + kTileRows < n_edges ? p.tgt[e0 + kTileRows] : -3;
    }
    if (threadIdx.x < 8) misc->piece_flag[threadIdx.x] = 0;

    // ---- stage 0: e0 (2^l W1c)^T, then + P[tgt] + Q[src], ReLU ----
    mbar_wait(&misc->bar_e, phase_e);
    phase_e ^= 1u;
    gemm_weights<TC>(acc, ws, e_tile, p.mlp.w_off[0], he, hidden, 0, false);
    gemm_weights_wait<TC>(acc, ws);
    weights_prefetch<TC>(ws, p.mlp.w_off[1], hidden, nh > 1 ? hidden : out);
    estamp(3);
    if constexpr (!TC) __syncthreads();      // every thread has finished reading the e0 tile
    if (t0 && tile + (int)gridDim.x < n_tiles)
      tma_load_span(e_tile, reinterpret_cast<const char*>(p.e_tiles) +
                                (long)(tile + gridDim.x) * tile_bytes, tile_bytes, &misc->bar_e);
    {
      const float* prow = p.pq + (long)(t < 0 ? 0 : t) * 2 * hidden;
      const float* qrow = p.pq + (long)(s < 0 ? 0 : s) * 2 * hidden + hidden;
      SGNN_FOR_MY_CHUNKS(c, hidden) {
        float v[32];
        acc_load32<TC>(acc, c, v);
        if (tid < nvalid) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(prow + c) + j);
            const float4 b = __ldg(reinterpret_cast<const float4*>(qrow + c) + j);
            v[4 * j] += a.x + b.x; v[4 * j + 1] += a.y + b.y;
            v[4 * j + 2] += a.z + b.z; v[4 * j + 3] += a.w + b.w;
          }
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        a_tile_store32<TC>(hid, c, v);
      }
    }
    estamp(4);
    // ---- hidden -> hidden stages (nmlp_layers > 1) ----
    int region = 0;
    for (int l = 1; l < nh; ++l) {
      a_tile_publish<TC>();
      region ^= 1;
      gemm_weights<TC>(acc, ws, hid, p.mlp.w_off[l], hidden, hidden, region * w, false);
      gemm_weights_wait<TC>(acc, ws);
      weights_prefetch<TC>(ws, p.mlp.w_off[l + 1], hidden, l + 1 < nh ? hidden : out);
      const float* b = blob_f + p.mlp.b_off[l];
      SGNN_FOR_MY_CHUNKS(c, hidden) {
        float v[32];
        acc_load32<TC>(acc, region * w + c, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + b[c + j], 0.f);
        a_tile_store32<TC>(hid, c, v);
      }
    }
    estamp(5);
    // ---- output stage + LayerNorm -> message tile M (aliases hid) ----
    a_tile_publish<TC>();
    region ^= 1;
    gemm_weights<TC>(acc, ws, hid, p.mlp.w_off[nh], hidden, out, region * w, false);
    gemm_weights_wait<TC>(acc, ws);
    if (tile + (int)gridDim.x < n_tiles) weights_prefetch<TC>(ws, p.mlp.w_off[0], he, hidden);
    estamp(6);
    float* m_tile = reinterpret_cast<float*>(hid);
    {
      const float* b = blob_f + p.mlp.b_off[nh];
      const float* gam = blob_f + p.mlp.gamma_off;
      const float* bet = blob_f + p.mlp.beta_off;
      float mean, rstd;
      row_ln_stats<TC>(acc, region * w, out, b, mean, rstd, misc->xch);
      SGNN_FOR_MY_CHUNKS(c, out) {
        float v[32];
        acc_load32<TC>(acc, region * w + c, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float4 y;
          y.x = (v[4 * j] + b[c + 4 * j] - mean) * rstd * gam[c + 4 * j] + bet[c + 4 * j];
          y.y = (v[4 * j + 1] + b[c + 4 * j + 1] - mean) * rstd * gam[c + 4 * j + 1] + bet[c + 4 * j + 1];
          y.z = (v[4 * j + 2] + b[c + 4 * j + 2] - mean) * rstd * gam[c + 4 * j + 2] + bet[c + 4 * j + 2];
          y.w = (v[4 * j + 3] + b[c + 4 * j + 3] - mean) * rstd * gam[c + 4 * j + 3] + bet[c + 4 * j + 3];
          const int chunk = (c >> 2) + j;
          *reinterpret_cast<float4*>(m_tile + tid * out + ((chunk ^ (tid & 7)) << 2)) = y;
        }
      }
    }
    if constexpr (TC) tc_fence_before_sync();
    __syncthreads();   // tgt_s / piece_flag visible; all TMEM reads of this tile retired
    estamp(7);

    // ---- segmented sums over the receiver-sorted rows: warp g of part q owns rows 32g..32g+31 and the columns
    //      [q out/nq, (q+1) out/nq), a lane cpl consecutive ones of them; control flow is warp-uniform ----
    const int cpl_w = out >> 5;                // columns per lane over the full width: 1, 2, 3 or 4
    int nq = t_parts();
    while (cpl_w % nq) nq >>= 1;               // parts that take a share of the columns
    const int cpl = cpl_w / nq;
    const int colb = part * (out / nq) + lane * cpl;
    const int r0 = warp * 32;
    if (r0 < nvalid && part < nq) {
      // The receivers of the warp's 32 rows sit in its lanes (segment starts as one ballot mask) and a column's 32
      // messages are fetched up front, so the walk over the rows is register arithmetic under warp-uniform branches
      // (the row-by-row version -- shared-memory loads of receiver and message, a dynamically indexed sum[] in local
      // memory -- took 7.5 of the kernel's 21 us at 128-wide messages, r02f).  Additions in row order, as before.
      const int nr = r0 + 32 < nvalid ? 32 : nvalid - r0;
      const int my_t = misc->tgt_s[1 + r0 + lane];
      const unsigned start = __ballot_sync(0xffffffffu, my_t != misc->tgt_s[r0 + lane]);   // bit r: row r0 + r opens a receiver
      const bool first_cont = !(start & 1u);
      const int last_t = __shfl_sync(0xffffffffu, my_t, nr - 1);
      const bool cont_next = misc->tgt_s[1 + r0 + nr] == last_t;
      const bool flagger = lane == 0 && part == 0;
      for (int j = 0; j < cpl; ++j) {
        const int col = colb + j;
        float sum = 0.f, nxt = m_tile_read(m_tile, r0, col, out);
        bool first_seg = true;
        estamp(10);
#pragma unroll 1
        for (int r = 0; r < nr; ++r) {
          const float mv = nxt;
          if (r + 1 < nr) nxt = m_tile_read(m_tile, r0 + r + 1, col, out);
          if (r > 0 && ((start >> r) & 1u)) {          // the receiver of rows .. r - 1 is complete within this warp
            const int cur = __shfl_sync(0xffffffffu, my_t, r - 1);
            if (first_seg && first_cont) {
              misc->pieces[(2 * warp) * 128 + col] = sum;
              if (flagger && j == 0) { misc->piece_tgt[2 * warp] = cur; misc->piece_flag[2 * warp] = 1; }
            } else {
              p.agg[(long)cur * out + col] = sum;
            }
            first_seg = false;
            sum = 0.f;
          }
          sum += mv;
        }
        if (first_seg && first_cont) {
          misc->pieces[(2 * warp) * 128 + col] = sum;
          if (flagger && j == 0) { misc->piece_tgt[2 * warp] = last_t; misc->piece_flag[2 * warp] = cont_next ? 3 : 1; }
        } else if (cont_next) {
          misc->pieces[(2 * warp + 1) * 128 + col] = sum;
          if (flagger && j == 0) { misc->piece_tgt[2 * warp + 1] = last_t; misc->piece_flag[2 * warp + 1] = 1; }
        } else {
          p.agg[(long)last_t * out + col] = sum;
        }
      }
      estamp(11);
    }
    __syncthreads();
    estamp(8);
    // ---- stitch the pieces of rows that straddle warp groups / tiles (warp 0) ----
    if (threadIdx.x < 32) {
      float carry[4] = {0.f, 0.f, 0.f, 0.f};
      bool has = false, from_prev = false;
      int ctgt = -1;
      for (int g = 0; g < 4; ++g) {
        const int f0 = misc->piece_flag[2 * g];
        if (f0) {
          if (!has) { has = true; from_prev = true; ctgt = misc->piece_tgt[2 * g];
                      carry[0] = carry[1] = carry[2] = carry[3] = 0.f; }
          for (int j = 0; j < cpl_w; ++j) carry[j] += misc->pieces[(2 * g) * 128 + lane * cpl_w + j];
          if (!(f0 & 2)) {
            float* dst = from_prev ? p.head + (long)tile * out : p.agg + (long)ctgt * out;
            for (int j = 0; j < cpl_w; ++j) dst[lane * cpl_w + j] = carry[j];
            has = false;
          }
        }
        if (misc->piece_flag[2 * g + 1]) {
          has = true; from_prev = false; ctgt = misc->piece_tgt[2 * g + 1];
          for (int j = 0; j < cpl_w; ++j) carry[j] = misc->pieces[(2 * g + 1) * 128 + lane * cpl_w + j];
        }
      }
      if (has) {   // the row continues into the next tile: its remainder arrives as that tile's head
        float* dst = from_prev ? p.head + (long)tile * out : p.agg + (long)ctgt * out;
        for (int j = 0; j < cpl_w; ++j) dst[lane * cpl_w + j] = carry[j];
      }
    }
    // next iteration's first __syncthreads (a_tile_publish) orders the piece buffers
  }
  weights_prefetch_drain(ws);
  acc_teardown<TC>(acc);
  estamp(9);
}

}  // namespace sgnn

// debug: device buffer of >= 16 uint64 receiving the stamps of edge_layer_kernel (NULL = off); SGNN_TIMELINE builds
extern "C" int sgnn_debug_edge_generic_timeline(unsigned long long* buf) {
  SGNN_CUDA(cudaMemcpyToSymbol(sgnn::g_edge_generic_timeline, &buf, sizeof(buf)));
  return SGNN_OK;
}

using namespace sgnn;

static int check_mlp(const SgnnMlp& m, const char* who) {
  SGNN_CHECK_ARG(m.n_hidden >= 1 && m.n_hidden <= SGNN_MAX_HIDDEN, "%s: n_hidden=%d unsupported", who, m.n_hidden);
  SGNN_CHECK_ARG(m.hidden % 32 == 0 && m.hidden >= 32 && m.hidden <= 128,
                 "%s: hidden width %d must be a multiple of 32 in [32,128]", who, m.hidden);
  return SGNN_OK;
}

extern "C" size_t sgnn_edge_layer_smem_bytes(int blob_bytes, int he, int hidden, int out, int tc) {
  const bool st = edge_layer_streamed(blob_bytes, he, hidden, out, tc != 0);
  return (size_t)edge_layer_smem(blob_bytes, he, hidden, out, tc != 0, st).total;
}

int sgnn_edge_layer_fast_try(const SgnnEdgeLayerParams* p, void* stream);   // edge_layer_fast.cu
int sgnn_edge_layer_gather_launch(const SgnnEdgeLayerParams* p, void* stream);   // edge_layer_ws.cu

extern "C" int sgnn_edge_layer_fwd(const SgnnEdgeLayerParams* p, int use_tensor_cores, int grid,
                                   void* stream) {
  int rc = check_mlp(p->mlp, "edge_layer");
  if (rc) return rc;
  SGNN_CHECK_ARG(p->mlp.out % 32 == 0 && p->mlp.out >= 32 && p->mlp.out <= 128,
                 "edge_layer: message width %d must be a multiple of 32 in [32,128]", p->mlp.out);
  SGNN_CHECK_ARG(p->he % 32 == 0 && p->he >= 32 && p->he <= 128, "edge_layer: he=%d unsupported", p->he);
  SGNN_CHECK_ARG(p->blob_bytes % 16 == 0, "edge_layer: blob size must be a multiple of 16");
  SGNN_CHECK_ARG(p->mlp.has_ln, "edge_layer: the edge MLP ends in a LayerNorm in the reference");
  if (p->max_tiles <= 0 || p->n_nodes <= 0) return SGNN_OK;
  if (p->wxs_off > 0) {      // gather mode: the Q half of pq holds TF32-rounded node latents, not x W1b^T
    SGNN_CHECK_ARG(use_tensor_cores, "edge_layer: gather mode needs the tensor-core path");
    return sgnn_edge_layer_gather_launch(p, stream);
  }
  if (use_tensor_cores && !getenv("SGNN_EDGE_GENERIC")) {     // throughput kernel where instantiated
    const int r = sgnn_edge_layer_fast_try(p, stream);
    if (r != 0) return r < 0 ? r : SGNN_OK;
  }
  const int streamed = edge_layer_streamed(p->blob_bytes, p->he, p->mlp.hidden, p->mlp.out, use_tensor_cores);
  const int smem = edge_layer_smem(p->blob_bytes, p->he, p->mlp.hidden, p->mlp.out, use_tensor_cores, streamed).total;
  SGNN_CHECK_ARG(smem <= 227 * 1024, "edge_layer: needs %d bytes of shared memory (> 227 KB) even with "
                 "streamed weights; this layer shape is not supported", smem);
  SGNN_CHECK_ARG(p->blob_bytes >= SGNN_VEC_REGION_BYTES, "edge_layer: blob lacks the vector region");
  if (grid > p->max_tiles) grid = p->max_tiles;
  if (grid < 1) grid = 1;
  const int threads = kThreads * sgnn_tile_parts(p->mlp.hidden > p->mlp.out ? p->mlp.hidden : p->mlp.out, use_tensor_cores != 0);
  if (use_tensor_cores) {
    SGNN_ENSURE_SMEM(edge_layer_kernel<true>, smem);
    SGNN_CUDA(sgnn_launch((edge_layer_kernel<true>), dim3(grid), dim3(threads), smem, (cudaStream_t)stream, *p, streamed));
  } else {
    SGNN_ENSURE_SMEM(edge_layer_kernel<false>, smem);
    SGNN_CUDA(sgnn_launch((edge_layer_kernel<false>), dim3(grid), dim3(threads), smem, (cudaStream_t)stream, *p, streamed));
  }
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
