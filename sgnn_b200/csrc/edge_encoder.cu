// Edge encoder: per-edge features -> MLP -> LayerNorm -> e0, written as 128-row tiles that are
// already in the tensor-core A-operand layout (so every message-passing layer can fetch a
// tile with one TMA bulk copy).
//   features: (p[src] - p[tgt]) / R and its L2 norm   (learned_simulator.py:299-312,
//             multi_scale_simulator.py:222-232)
//   network : Encoder.edge_fn (graph_network.py:92-96) / *_edge_encoder (multi_scale_gnn.py:248-259)
// The first Linear has k_in = dim + 1 = 3 inputs: it is evaluated with FFMA by the row's
// thread (a K = 3 GEMM would waste a whole 32-wide K-block); the remaining layers are tcgen05 GEMMs.
// Algorithmic HBM bytes per edge: 8 (src, tgt) + 4 * out written; positions are L2-resident gathers.
#include <stdlib.h>

#include "mlp_tile.cuh"

namespace sgnn {

struct EncMisc {
  uint64_t bar_w, bar_mma, bar_stream;
  uint32_t tmem_slot;
  int pad;
  float xch[2 * kMaxParts * kTileRows];     // LayerNorm moments between the parts of a row
};

struct EdgeEncSmem {
  int wbuf, hid, misc, simt, total;
};
__host__ __device__ inline EdgeEncSmem edge_enc_smem(int blob_bytes, int hidden, int out, bool tc,
                                                     bool streamed) {
  EdgeEncSmem s;
  const int w = hidden > out ? hidden : out;
  s.wbuf = align_up(blob_smem_bytes(blob_bytes, streamed), 1024);
  s.hid = s.wbuf + (streamed ? wbuf_bytes(tc) : 0);
  s.misc = s.hid + kTileRows * w * 4;
  s.simt = align_up(s.misc + (int)sizeof(EncMisc), 1024);
  s.total = s.simt + (tc ? 0 : kTileRows * pow2_cols(w) * 4) + 1024;
  return s;
}

__host__ inline bool edge_enc_streamed(int blob_bytes, int hidden, int out, bool tc) {
  return edge_enc_smem(blob_bytes, hidden, out, tc, false).total > 227 * 1024;
}

constexpr int kMaxEdgeFeat = 8;

template <bool TC>
__global__ void __launch_bounds__(kThreads * kMaxParts) edge_encoder_kernel(const SgnnEdgeEncoderParams p,
                                                                const int streamed) {
  pdl_launch_dependents();
  char* smem = smem_base_1024();
  const int hidden = p.mlp.hidden, out = p.mlp.out;
  const int w = hidden > out ? hidden : out;
  const EdgeEncSmem L = edge_enc_smem(p.blob_bytes, hidden, out, TC, streamed != 0);
  const float* blob_f = reinterpret_cast<const float*>(smem);
  char* hid = smem + L.hid;
  EncMisc* misc = reinterpret_cast<EncMisc*>(smem + L.misc);
  const int tid = t_row();
  const bool t0 = threadIdx.x == 0, part0 = t_part() == 0;

  // the blob is constant: staged (and the first streamed image put in flight) ahead of griddepcontrol.wait
  WeightStage ws;
  weights_setup(ws, p.blob, p.blob_bytes, smem, smem + L.wbuf, &misc->bar_w, &misc->bar_stream, streamed);
  TileAcc<TC> acc;
  acc_setup<TC>(acc, TC ? pow2_cols(2 * w) : pow2_cols(w), &misc->bar_mma, &misc->tmem_slot,
                reinterpret_cast<float*>(smem + L.simt));
  const int n_out16 = (out + 15) & ~15;
  weights_prefetch<TC>(ws, p.mlp.w_off[1], hidden, p.mlp.n_hidden > 1 ? hidden : n_out16);
  pdl_wait();
  const int n_edges = p.rowptr ? p.rowptr[p.n_nodes] : p.n_nodes;   // rowptr NULL: n_nodes = E
  const int n_tiles = (n_edges + kTileRows - 1) / kTileRows;
  mbar_wait(&misc->bar_w, 0);
  if ((int)blockIdx.x >= n_tiles) {
    weights_prefetch_drain(ws);
    acc_teardown<TC>(acc);
    return;
  }
  const int tile_bytes = kTileRows * out * 4;

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int eidx = tile * kTileRows + tid;
    const bool valid = eidx < n_edges;
    // ---- features ----
    float f[kMaxEdgeFeat];
#pragma unroll
    for (int k = 0; k < kMaxEdgeFeat; ++k) f[k] = 0.f;
    if (valid) {
      if (p.feat) {
        for (int k = 0; k < p.k_in; ++k) f[k] = p.feat[(long)eidx * p.k_in + k];
      } else {
        const int s = p.src[eidx], t = p.tgt[eidx];
        float sq = 0.f;
        for (int d = 0; d < p.dim; ++d) {
          const float diff = __fsub_rn(p.pos[(long)s * p.pos_stride + d], p.pos[(long)t * p.pos_stride + d]);
          f[d] = __fdiv_rn(diff, p.radius);
          sq = __fadd_rn(sq, __fmul_rn(f[d], f[d]));
        }
        f[p.dim] = sqrtf(sq);
      }
      if (p.feat_out && part0)
        for (int k = 0; k < p.k_in; ++k) p.feat_out[(long)eidx * p.k_in + k] = f[k];
    }
    // ---- layer 0 (k_in -> hidden) in fp32 FFMA, ReLU ----
    {
      const float* w0 = blob_f + p.w0_off;
      const float* b0 = blob_f + p.mlp.b_off[0];
      SGNN_FOR_MY_CHUNKS(c, hidden) {
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float a = 0.f;
          for (int k = 0; k < p.k_in; ++k) a = fmaf(f[k], w0[(c + j) * p.k_in + k], a);
          v[j] = fmaxf(a + b0[c + j], 0.f);
        }
        a_tile_store32<TC>(hid, c, v);
      }
    }
    const int acc_col = tile + (int)gridDim.x < n_tiles
                            ? mlp_run_rest<TC>(acc, ws, hid, p.mlp, w, 1, p.mlp.w_off[1], hidden, p.mlp.n_hidden > 1 ? hidden : n_out16)
                            : mlp_run_rest<TC>(acc, ws, hid, p.mlp, w, 1);
    // ---- LayerNorm -> tile (A-operand layout) staged in `hid`, then one TMA bulk store ----
    float mean = 0.f, rstd = 1.f;
    if (p.mlp.has_ln) row_ln_stats<TC>(acc, acc_col, out, blob_f + p.mlp.b_off[p.mlp.n_hidden], mean, rstd, misc->xch);
    SGNN_FOR_MY_CHUNKS(c, out) {
      float v[32];
      mlp_out_chunk<TC>(acc, acc_col, c, blob_f, p.mlp, mean, rstd, v);
      if (!valid) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.f;
      }
      a_tile_store32<TC>(hid, c, v);
    }
    fence_proxy_async_smem();
    if constexpr (TC) tc_fence_before_sync();
    __syncthreads();
    if (t0) {
      tma_bulk_s2g(reinterpret_cast<char*>(p.e_tiles) + (long)tile * tile_bytes, hid, (uint32_t)tile_bytes);
      tma_store_wait_read();       // `hid` is rewritten by the next tile's layer 0
    }
    __syncthreads();
  }
  if (t0) tma_store_wait_all();
  weights_prefetch_drain(ws);
  acc_teardown<TC>(acc);
}

}  // namespace sgnn

using namespace sgnn;

extern "C" size_t sgnn_edge_encoder_smem_bytes(int blob_bytes, int hidden, int out, int tc) {
  const bool st = edge_enc_streamed(blob_bytes, hidden, out, tc != 0);
  return (size_t)edge_enc_smem(blob_bytes, hidden, out, tc != 0, st).total;
}

int sgnn_edge_encoder_fast_try(const SgnnEdgeEncoderParams* p, void* stream);   // edge_encoder_fast.cu

extern "C" int sgnn_edge_encoder_fwd(const SgnnEdgeEncoderParams* p, int use_tensor_cores, int grid,
                                     void* stream) {
  SGNN_CHECK_ARG(p->mlp.n_hidden >= 1 && p->mlp.n_hidden <= SGNN_MAX_HIDDEN, "edge_encoder: n_hidden=%d", p->mlp.n_hidden);
  SGNN_CHECK_ARG(p->mlp.hidden % 32 == 0 && p->mlp.hidden >= 32 && p->mlp.hidden <= 128,
                 "edge_encoder: hidden width %d must be a multiple of 32 in [32,128]", p->mlp.hidden);
  SGNN_CHECK_ARG(p->mlp.out % 32 == 0 && p->mlp.out >= 32 && p->mlp.out <= 128,
                 "edge_encoder: output width %d must be a multiple of 32 in [32,128]", p->mlp.out);
  SGNN_CHECK_ARG(p->k_in >= 1 && p->k_in <= kMaxEdgeFeat, "edge_encoder: k_in=%d must be in [1,%d]", p->k_in, kMaxEdgeFeat);
  SGNN_CHECK_ARG(p->feat != nullptr || p->k_in == p->dim + 1, "edge_encoder: k_in must be dim+1");
  SGNN_CHECK_ARG(p->blob_bytes % 16 == 0, "edge_encoder: blob size must be a multiple of 16");
  if (p->max_tiles <= 0) return SGNN_OK;
  if (use_tensor_cores && !getenv("SGNN_ENC_GENERIC")) {
    const int r = sgnn_edge_encoder_fast_try(p, stream);
    if (r != 0) return r < 0 ? r : SGNN_OK;
  }
  const int streamed = edge_enc_streamed(p->blob_bytes, p->mlp.hidden, p->mlp.out, use_tensor_cores);
  const int smem = edge_enc_smem(p->blob_bytes, p->mlp.hidden, p->mlp.out, use_tensor_cores, streamed).total;
  SGNN_CHECK_ARG(p->blob_bytes >= SGNN_VEC_REGION_BYTES, "edge_encoder: blob lacks the vector region");
  SGNN_CHECK_ARG(smem <= 227 * 1024, "edge_encoder: needs %d bytes of shared memory (> 227 KB)", smem);
  if (grid > p->max_tiles) grid = p->max_tiles;
  if (grid < 1) grid = 1;
  const int threads = kThreads * sgnn_tile_parts(p->mlp.hidden > p->mlp.out ? p->mlp.hidden : p->mlp.out, use_tensor_cores != 0);
  if (use_tensor_cores) {
    SGNN_ENSURE_SMEM(edge_encoder_kernel<true>, smem);
    SGNN_CUDA(sgnn_launch((edge_encoder_kernel<true>), dim3(grid), dim3(threads), smem, (cudaStream_t)stream, *p, streamed));
  } else {
    SGNN_ENSURE_SMEM(edge_encoder_kernel<false>, smem);
    SGNN_CUDA(sgnn_launch((edge_encoder_kernel<false>), dim3(grid), dim3(threads), smem, (cudaStream_t)stream, *p, streamed));
  }
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
