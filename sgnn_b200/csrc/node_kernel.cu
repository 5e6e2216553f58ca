// Node side of the Encode-Process-Decode network: one kernel, three prologues, three tails.
//
//   prologue FEATURES : node features from the position window (time_diff, velocity
//                       normalisation, wall distance, type embedding) -- learned_simulator.py:
//                       257-290,540-550 / multi_scale_simulator.py:173-201   -> Encoder.node_fn
//            AGG_X    : [aggregate | x] of InteractionNetwork.update (graph_network.py:201-222);
//                       the aggregate is agg[i] (+ the head partials of CSR rows that straddle
//                       edge tiles), or 0 for receivers without edges
//            ROWS     : explicit input rows (module-level API: Encoder / Decoder called directly)
//   body              : Linear-ReLU-...-Linear (tcgen05 GEMMs, weights resident in shared memory)
//                       -> LayerNorm -> (+ x residual) -> x_out
//   tail     PQ       : P = x_out W1a^T + b1, Q = x_out W1b^T for the NEXT layer's edge kernel
//            DECODE   : Decoder MLP (graph_network.py:321-333) with the last dim+1-wide Linear in
//                       FFMA, then the Euler integrator (_decoder_postprocessor, :381-411)
// A persistent CTA of 128 threads walks 128-node tiles; thread t owns row t in all epilogues.
// Algorithmic HBM bytes per node and layer: 4*(agg_w + x_w) read + 4*out + 8*pq_hidden written.
#include <stdlib.h>

#include "mlp_tile.cuh"

namespace sgnn {

// optional in-kernel timeline (SGNN_TIMELINE builds; profiles/node_generic_timeline.py): %globaltimer stamps of
// CTA 0 / thread 0, slots 0..15 for PQ / plain tails, 16..31 for the DECODE tail
__device__ unsigned long long* g_node_generic_timeline = nullptr;
__device__ __forceinline__ void gstamp(const SgnnNodeParams& p, int slot) {
#ifdef SGNN_TIMELINE
  if (g_node_generic_timeline != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    g_node_generic_timeline[slot + (p.tail == SGNN_TAIL_DECODE ? 16 : 0)] = t;
  }
#else
  (void)p; (void)slot;
#endif
}

struct NodeMisc {
  uint64_t bar_w, bar_mma, bar_stream;
  uint32_t tmem_slot;
  int pad;
  int rowptr[kTileRows + 1];     // CSR bounds of the tile's rows (AGG_X prologue)
  float xch[2 * kMaxParts * kTileRows];     // LayerNorm moments / decoder partial sums between the parts of a row
};

struct NodeSmem {
  int wbuf, a, misc, simt, acc_cols, total;
};
__host__ __device__ inline int imax(int a, int b) { return a > b ? a : b; }

__host__ __device__ inline NodeSmem node_smem(const SgnnNodeParams& p, bool tc, bool streamed) {
  NodeSmem s;
  int ka = p.k_pad, wacc = 32;
  if (p.has_mlp) {
    ka = imax(ka, imax(p.mlp.hidden, p.mlp.out));
    wacc = imax(p.mlp.hidden, (p.mlp.out + 15) & ~15);
  }
  int cols = 2 * wacc;
  if (p.tail == SGNN_TAIL_PQ) cols = imax(cols, 2 * p.pq_hidden);
  if (p.tail == SGNN_TAIL_DECODE) {
    ka = imax(ka, p.dec.hidden);
    cols = imax(cols, 2 * p.dec.hidden);
  }
  if (!tc) cols = (cols + 1) / 2;      // FFMA mode: one live region (columns wrap), P and Q in turn
  s.acc_cols = pow2_cols(cols);
  s.wbuf = align_up(blob_smem_bytes(p.blob_bytes, streamed), 1024);
  s.a = s.wbuf + (streamed ? wbuf_bytes(tc) : 0);
  s.misc = s.a + kTileRows * ka * 4;
  s.simt = align_up(s.misc + (int)sizeof(NodeMisc), 1024);
  s.total = s.simt + (tc ? 0 : kTileRows * s.acc_cols * 4) + 1024;
  return s;
}

__host__ inline bool node_streamed(const SgnnNodeParams& p, bool tc) {
  return node_smem(p, tc, false).total > 227 * 1024;
}

template <bool TC>
__device__ __forceinline__ void a_store_scalar(char* a_tile, int row, int col, float v) {
  if constexpr (TC) v = to_tf32(v);
  *reinterpret_cast<float*>(a_tile + sw128_offset(row, col, kTileRows)) = v;
}

// FEATURES prologue: one thread per row (part 0)
template <bool TC>
__device__ __forceinline__ void prologue_features(const SgnnNodeParams& p, char* a_tile, long i, bool valid) {
  if (t_part() != 0) return;
  const int row = t_row();
  int k = 0;
  if (valid) {
    const float* ps = p.pos_seq + i * p.pos_seq_stride;
    for (int t = 0; t + 1 < p.seq_len; ++t)
      for (int d = 0; d < p.dim; ++d) {
        const float vel = __fsub_rn(ps[(t + 1) * p.dim + d], ps[t * p.dim + d]);
        const float f = __fdiv_rn(__fsub_rn(vel, p.vel_mean[d]), p.vel_std[d]);
        if (p.feat_out) p.feat_out[i * p.k_in + k] = f;
        a_store_scalar<TC>(a_tile, row, k++, f);
      }
    float wd = __fadd_rn(ps[(p.seq_len - 1) * p.dim], p.wall_offset);
    wd = fminf(fmaxf(wd, 0.f), p.wall_max);
    wd = __fdiv_rn(wd, p.wall_div);
    if (p.feat_out) p.feat_out[i * p.k_in + k] = wd;
    a_store_scalar<TC>(a_tile, row, k++, wd);
    if (p.emb) {
      long ty = p.types[i];      // nn.Embedding raises on ids outside the table; here they are clamped
      if (p.n_types > 0) ty = ty < 0 ? 0 : (ty >= p.n_types ? p.n_types - 1 : ty);
      const float* e = p.emb + ty * p.emb_dim;
      for (int j = 0; j < p.emb_dim; ++j) {
        if (p.feat_out) p.feat_out[i * p.k_in + k] = e[j];
        a_store_scalar<TC>(a_tile, row, k++, e[j]);
      }
    }
  }
  for (; k < p.k_pad; ++k) a_store_scalar<TC>(a_tile, row, k, 0.f);
}

// AGG_X prologue: cooperative, coalesced row segments.  The tile's CSR bounds are fetched ONCE into shared memory
// (one round trip) and the agg chunks then go eight at a time: with the bounds looked up per chunk every batch of
// four loads was two dependent round trips, 8 batches for a 128-wide agg tile (~10 of the kernel's 61 us at 4800
// nodes, multi-scale configs[3], r02d).
template <bool TC>
__device__ __forceinline__ void prologue_agg_x(const SgnnNodeParams& p, char* a_tile, long row0, int nvalid, int* s_rowptr) {
  for (int r = threadIdx.x; r <= kTileRows; r += blockDim.x) s_rowptr[r] = r <= nvalid ? __ldg(p.rowptr + row0 + r) : 0;
  __syncthreads();
  // every chunk of the [agg | x] tile in flight at once (cp.async): the register version walked the tile in 16
  // dependent round trips of 8 loads per thread, 17.6 of the kernel's 46 us at 4800 nodes and 128-wide latents
  // (profiles/node_generic_timeline.py, r02f).  Receivers without edges: the edge kernel never wrote the row.
  a_tile_load_rows_async(a_tile, 0, p.agg, p.agg_w, p.agg_w, nvalid,
                         [=](int r) { return s_rowptr[r + 1] > s_rowptr[r] ? row0 + r : -1L; });
  a_tile_load_rows_async(a_tile, p.agg_w, p.x_in, p.x_w, p.x_w, nvalid, [=](int r) { return row0 + r; });
  cp_async_commit();
  cp_async_wait_all();
  const float* head = p.head;
  const int agg_w = p.agg_w;
  a_tile_fix_own<TC>(a_tile, 0, p.agg_w, [=](int r, int c, float4& t) {
    if (r < nvalid && s_rowptr[r + 1] > s_rowptr[r]) {
      const int t0 = s_rowptr[r] >> 7, t1 = (s_rowptr[r + 1] - 1) >> 7;
      for (int tt = t0 + 1; tt <= t1; ++tt) {          // CSR row straddling edge tiles
        const float4 h = __ldg(reinterpret_cast<const float4*>(head + (long)tt * agg_w) + c);
        t.x += h.x; t.y += h.y; t.z += h.z; t.w += h.w;
      }
    }
  });
  if constexpr (TC) a_tile_fix_own<TC>(a_tile, p.agg_w, p.x_w, [](int, int, float4&) {});
}

// ROWS prologue: arbitrary width / leading dimension (scalar, coalesced along the row)
template <bool TC>
__device__ __forceinline__ void prologue_rows(const SgnnNodeParams& p, char* a_tile, long row0, int nvalid) {
  for (int idx = threadIdx.x; idx < kTileRows * p.k_pad; idx += blockDim.x) {
    const int r = idx / p.k_pad, c = idx - r * p.k_pad;
    float v = 0.f;
    if (r < nvalid && c < p.k_in) v = __ldg(p.rows_in + (row0 + r) * p.rows_ld + c);
    a_store_scalar<TC>(a_tile, r, c, v);
  }
}

template <bool TC>
__global__ void __launch_bounds__(kThreads * kMaxParts) node_kernel(const SgnnNodeParams p, const int streamed) {
  pdl_launch_dependents();
  gstamp(p, 0);
  char* smem = smem_base_1024();
  const NodeSmem L = node_smem(p, TC, streamed != 0);
  const float* blob_f = reinterpret_cast<const float*>(smem);
  char* a_tile = smem + L.a;
  NodeMisc* misc = reinterpret_cast<NodeMisc*>(smem + L.misc);
  const int tid = t_row();       // the row this thread works on
  const int n_tiles = (p.n_rows + kTileRows - 1) / kTileRows;
  if ((int)blockIdx.x >= n_tiles) return;

  WeightStage ws;
  weights_setup(ws, p.blob, p.blob_bytes, smem, smem + L.wbuf, &misc->bar_w, &misc->bar_stream, streamed);
  TileAcc<TC> acc;
  acc_setup<TC>(acc, L.acc_cols, &misc->bar_mma, &misc->tmem_slot, reinterpret_cast<float*>(smem + L.simt));
  const int w = p.has_mlp ? imax(p.mlp.hidden, (p.mlp.out + 15) & ~15) : 32;
  // streamed weights: the first image of a tile is in flight before the tile's rows are (the blob is constant: it
  // may be read ahead of griddepcontrol.wait)
  const int first_off = p.has_mlp ? p.mlp.w_off[0] : (p.tail == SGNN_TAIL_PQ ? p.wp_off : p.tail == SGNN_TAIL_DECODE ? p.dec.w_off[0] : -1);
  const int first_n = p.has_mlp ? p.mlp.hidden : (p.tail == SGNN_TAIL_PQ ? p.pq_hidden : p.dec.hidden);
  if (first_off >= 0) weights_prefetch<TC>(ws, first_off, p.k_pad, first_n);
  // Staging tile for coalesced row I/O: the A-tile K-blocks behind the columns the later GEMMs read, free once
  // GEMM 0 has consumed the input tile (thread-per-row 16-byte stores ran the 64 KB x' tile out in 7.6 us and the
  // 128 KB P | Q tile in 5.7 us at 4800 nodes, r02f).
  const int used_cols = p.has_mlp ? imax(imax(p.mlp.hidden, p.mlp.out), p.tail == SGNN_TAIL_DECODE ? p.dec.hidden : 0) : 0;
  const int stage_w = p.has_mlp ? imax(p.mlp.out, p.tail == SGNN_TAIL_PQ ? p.pq_hidden : 0) : 0;
  char* stage = a_tile + kTileRows * used_cols * 4;
  const bool staged = p.has_mlp && (L.misc - L.a) - kTileRows * used_cols * 4 >= kTileRows * stage_w * 4;
  gstamp(p, 1);
  pdl_wait();
  mbar_wait(&misc->bar_w, 0);
  gstamp(p, 2);

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long row0 = (long)tile * kTileRows;
    const int nvalid = p.n_rows - row0 < kTileRows ? (int)(p.n_rows - row0) : kTileRows;
    const long i = row0 + tid;
    const bool valid = tid < nvalid;
    // previous tile's MMAs (async proxy) have completed (gemm_wait); its TMEM reads retired below
    if (p.prologue == SGNN_PRO_FEATURES) prologue_features<TC>(p, a_tile, i, valid);
    else if (p.prologue == SGNN_PRO_AGG_X) prologue_agg_x<TC>(p, a_tile, row0, nvalid, misc->rowptr);
    else prologue_rows<TC>(p, a_tile, row0, nvalid);

    gstamp(p, 3);
    int x_width = p.k_pad;       // width of the tail's input held in the A tile
    if (p.has_mlp) {
      const int hidden = p.mlp.hidden, out = p.mlp.out;
      a_tile_publish<TC>();
      gemm_weights<TC>(acc, ws, a_tile, p.mlp.w_off[0], p.k_pad, hidden, 0, false);
      gemm_weights_wait<TC>(acc, ws);
      weights_prefetch<TC>(ws, p.mlp.w_off[1], hidden, p.mlp.n_hidden > 1 ? hidden : (out + 15) & ~15);
      if (staged && p.residual) {            // the K-blocks behind `stage` are free now: raw residual rows, coalesced
        if constexpr (!TC) __syncthreads();  // FFMA mode: every thread is done with its row of the input tile
        SGNN_FOR_MY_CHUNKS(c, out) stage_load_block_async(stage, p.x_in, p.x_w, out, row0, nvalid, c);
        cp_async_commit();
      }
      gstamp(p, 4);
      {
        const float* b = blob_f + p.mlp.b_off[0];
        SGNN_FOR_MY_CHUNKS(c, hidden) {
          float v[32];
          acc_load32<TC>(acc, c, v);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + b[c + j], 0.f);
          a_tile_store32<TC>(a_tile, c, v);
        }
      }
      gstamp(p, 5);
      const bool more = tile + (int)gridDim.x < n_tiles;
      const int acc_col = p.tail == SGNN_TAIL_PQ       ? mlp_run_rest<TC>(acc, ws, a_tile, p.mlp, w, 0, p.wp_off, out, p.pq_hidden)
                          : p.tail == SGNN_TAIL_DECODE ? mlp_run_rest<TC>(acc, ws, a_tile, p.mlp, w, 0, p.dec.w_off[0], out, p.dec.hidden)
                          : more                       ? mlp_run_rest<TC>(acc, ws, a_tile, p.mlp, w, 0, p.mlp.w_off[0], p.k_pad, hidden)
                                                       : mlp_run_rest<TC>(acc, ws, a_tile, p.mlp, w, 0);
      gstamp(p, 6);
      float mean = 0.f, rstd = 1.f;
      if (p.mlp.has_ln) row_ln_stats<TC>(acc, acc_col, out, blob_f + p.mlp.b_off[p.mlp.n_hidden], mean, rstd, misc->xch);
      gstamp(p, 7);
      if (staged) {
        cp_async_wait_all();
        __syncwarp();
      }
      SGNN_FOR_MY_CHUNKS(c, out) {
        float v[32];
        mlp_out_chunk<TC>(acc, acc_col, c, blob_f, p.mlp, mean, rstd, v);
        if (staged) {                 // residual from / result into the staging tile; the block leaves as row segments
          if (valid && p.residual) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 x = *stage_chunk(stage, tid, (c >> 2) + j, out);
              v[4 * j] += x.x; v[4 * j + 1] += x.y; v[4 * j + 2] += x.z; v[4 * j + 3] += x.w;
            }
          }
          if (p.x_out) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *stage_chunk(stage, tid, (c >> 2) + j, out) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            stage_store_block(stage, p.x_out, out, out, row0, nvalid, c);
          }
        } else if (valid) {
          if (p.residual) {
            const float4* xr = reinterpret_cast<const float4*>(p.x_in + i * p.x_w + c);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 x = __ldg(xr + j);
              v[4 * j] += x.x; v[4 * j + 1] += x.y; v[4 * j + 2] += x.z; v[4 * j + 3] += x.w;
            }
          }
          if (p.x_out) {
            float4* xo = reinterpret_cast<float4*>(p.x_out + i * out + c);
#pragma unroll
            for (int j = 0; j < 8; ++j) xo[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          }
        }
        if (p.tail != SGNN_TAIL_NONE) a_tile_store32<TC>(a_tile, c, v);
      }
      x_width = out;
    }
    gstamp(p, 8);

    if (p.tail == SGNN_TAIL_PQ) {
      const int hq = p.pq_hidden;
      a_tile_publish<TC>();
      if constexpr (TC) {
        gemm_weights<TC>(acc, ws, a_tile, p.wp_off, x_width, hq, 0, false);
        gemm_weights<TC>(acc, ws, a_tile, p.wq_off, x_width, hq, hq, false);
        gemm_weights_wait<TC>(acc, ws);
        if (first_off >= 0 && tile + (int)gridDim.x < n_tiles) weights_prefetch<TC>(ws, first_off, p.k_pad, first_n);
      }
      gstamp(p, 9);
      const float* b1 = blob_f + p.pq_bias_off;
      for (int half = 0; half < 2; ++half) {
        if constexpr (!TC) {      // one accumulator region: P, then Q
          gemm_weights<TC>(acc, ws, a_tile, half ? p.wq_off : p.wp_off, x_width, hq, 0, false);
          gemm_weights_wait<TC>(acc, ws);
        }
        const int base = TC ? half * hq : 0;
        SGNN_FOR_MY_CHUNKS(c, hq) {
          float v[32];
          acc_load32<TC>(acc, base + c, v);
          if (half == 0) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += b1[c + j];
          }
          if (staged) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *stage_chunk(stage, tid, (c >> 2) + j, hq) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            stage_store_block(stage, p.pq_out + half * hq, 2 * hq, hq, row0, nvalid, c);
          } else if (valid) {
            float4* dst = reinterpret_cast<float4*>(p.pq_out + i * 2 * hq + half * hq + c);
#pragma unroll
            for (int j = 0; j < 8; ++j) dst[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          }
        }
      }
    } else if (p.tail == SGNN_TAIL_DECODE) {
      const int hd = p.dec.hidden, nout = p.dec.out;
      int region = 1, kin = x_width;
      for (int l = 0; l < p.dec.n_hidden; ++l) {
        a_tile_publish<TC>();
        region ^= 1;
        gemm_weights<TC>(acc, ws, a_tile, p.dec.w_off[l], kin, hd, region * hd, false);
        gemm_weights_wait<TC>(acc, ws);
        if (l + 1 < p.dec.n_hidden) weights_prefetch<TC>(ws, p.dec.w_off[l + 1], hd, hd);
        else if (first_off >= 0 && tile + (int)gridDim.x < n_tiles) weights_prefetch<TC>(ws, first_off, p.k_pad, first_n);
        kin = hd;
        if (l + 1 < p.dec.n_hidden) {
          const float* b = blob_f + p.dec.b_off[l];
          SGNN_FOR_MY_CHUNKS(c, hd) {
            float v[32];
            acc_load32<TC>(acc, region * hd + c, v);
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + b[c + j], 0.f);
            a_tile_store32<TC>(a_tile, c, v);
          }
        }
      }
      // last hidden activation stays in registers; final Linear (hd -> nout <= 8) in FFMA
      float o[8];
      const float* bl = blob_f + p.dec.b_off[p.dec.n_hidden];
      const float* wl = blob_f + p.dec_wlast_off;
      const float* bh = blob_f + p.dec.b_off[p.dec.n_hidden - 1];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = 0.f;
      SGNN_FOR_MY_CHUNKS(c, hd) {
        float v[32];
        acc_load32<TC>(acc, region * hd + c, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + bh[c + j], 0.f);
        for (int k = 0; k < nout; ++k) {
          float a = o[k];
#pragma unroll
          for (int j = 0; j < 32; ++j) a = fmaf(v[j], wl[k * hd + c + j], a);
          o[k] = a;
        }
      }
      if (t_parts() > 1) {        // partial dot products of the parts, added in part order (the A tile is free: the
        float* xo = reinterpret_cast<float*>(a_tile);       // MMAs that read it have completed)
        for (int k = 0; k < nout; ++k) xo[(t_part() * 8 + k) * kTileRows + tid] = o[k];
        __syncthreads();
        if (t_part() == 0)
          for (int k = 0; k < nout; ++k) {
            float a = 0.f;
            for (int q = 0; q < t_parts(); ++q) a += xo[(q * 8 + k) * kTileRows + tid];
            o[k] = a;
          }
      }
      if (valid && t_part() == 0) {
        for (int k = 0; k < nout; ++k) o[k] += bl[k];
        if (p.dec_out)
          for (int k = 0; k < nout; ++k) p.dec_out[i * nout + k] = o[k];
        if (p.strain_out) p.strain_out[i] = o[nout - 1];
        if (p.next_pos) {
          const float* ps = p.pos_seq + i * p.pos_seq_stride;
          for (int d = 0; d < p.dim; ++d) {
            const float a = __fadd_rn(__fmul_rn(o[d], p.acc_std[d]), p.acc_mean[d]);
            const float last = ps[(p.seq_len - 1) * p.dim + d];
            const float vel = __fsub_rn(last, ps[(p.seq_len - 2) * p.dim + d]);
            p.next_pos[i * p.next_pos_stride + d] = __fadd_rn(last, __fadd_rn(vel, a));
          }
        }
      }
    }
    gstamp(p, 10);
    if constexpr (TC) tc_fence_before_sync();
    __syncthreads();     // TMEM reads and A-tile reads of this tile retired before the next prologue
  }
  weights_prefetch_drain(ws);
  acc_teardown<TC>(acc);
  gstamp(p, 11);
}

}  // namespace sgnn

using namespace sgnn;

// debug: device buffer of >= 32 uint64 receiving the stamps of node_kernel (NULL = off); SGNN_TIMELINE builds
extern "C" int sgnn_debug_node_generic_timeline(unsigned long long* buf) {
  SGNN_CUDA(cudaMemcpyToSymbol(g_node_generic_timeline, &buf, sizeof(buf)));
  return SGNN_OK;
}

extern "C" size_t sgnn_node_smem_bytes(const SgnnNodeParams* p, int tc) {
  return (size_t)node_smem(*p, tc != 0, node_streamed(*p, tc != 0)).total;
}

int sgnn_node_fast_try(const SgnnNodeParams* p, void* stream);     // node_fast.cu
int sgnn_node_ws_try(const SgnnNodeParams* p, void* stream);       // node_ws.cu

extern "C" int sgnn_node_fwd(const SgnnNodeParams* p, int use_tensor_cores, int grid, void* stream) {
  SGNN_CHECK_ARG(p->k_pad % 32 == 0 && p->k_pad >= 32 && p->k_pad <= 256, "node: k_pad=%d must be a multiple of 32 in [32,256]", p->k_pad);
  SGNN_CHECK_ARG(p->k_in <= p->k_pad, "node: k_in > k_pad");
  if (p->has_mlp) {
    SGNN_CHECK_ARG(p->mlp.n_hidden >= 1 && p->mlp.n_hidden <= SGNN_MAX_HIDDEN, "node: n_hidden=%d", p->mlp.n_hidden);
    SGNN_CHECK_ARG(p->mlp.hidden % 32 == 0 && p->mlp.hidden >= 32 && p->mlp.hidden <= 128, "node: hidden width %d unsupported", p->mlp.hidden);
    SGNN_CHECK_ARG(p->mlp.out % 32 == 0 && p->mlp.out >= 32 && p->mlp.out <= 128, "node: output width %d unsupported (use the DECODE tail for narrow outputs)", p->mlp.out);
    SGNN_CHECK_ARG(!p->residual || p->x_w == p->mlp.out, "node: residual needs x_w == out");
  }
  if (p->prologue == SGNN_PRO_AGG_X)
    SGNN_CHECK_ARG(p->agg_w % 32 == 0 && p->x_w % 32 == 0 && p->k_pad == p->agg_w + p->x_w, "node: AGG_X widths must be multiples of 32");
  if (p->tail == SGNN_TAIL_PQ)
    SGNN_CHECK_ARG(p->pq_hidden % 32 == 0 && p->pq_hidden >= 32 && p->pq_hidden <= 128, "node: pq_hidden=%d unsupported", p->pq_hidden);
  if (p->tail == SGNN_TAIL_DECODE) {
    SGNN_CHECK_ARG(p->dec.out >= 1 && p->dec.out <= 8, "node: decoder output width %d must be in [1,8]", p->dec.out);
    SGNN_CHECK_ARG(p->dec.hidden % 32 == 0 && p->dec.hidden <= 128 && p->dec.n_hidden >= 1, "node: decoder hidden width %d unsupported", p->dec.hidden);
    SGNN_CHECK_ARG(p->next_pos == nullptr || p->dec.out >= p->dim, "node: decoder output narrower than dim");
  }
  SGNN_CHECK_ARG(p->blob_bytes % 16 == 0, "node: blob size must be a multiple of 16");
  if (p->n_rows <= 0) return SGNN_OK;
  if (use_tensor_cores && !getenv("SGNN_NODE_GENERIC")) {      // throughput kernel for 64-wide latents
    int r = sgnn_node_ws_try(p, stream);
    if (r != 0) return r < 0 ? r : SGNN_OK;
    r = sgnn_node_fast_try(p, stream);
    if (r != 0) return r < 0 ? r : SGNN_OK;
  }
  SGNN_CHECK_ARG(!p->agg_affine, "node: agg_affine (edge kernel with defer_affine) is implemented by the 64-wide "
                 "warp-specialised node kernel only; this layer shape / switch must not defer the affine");
  SGNN_CHECK_ARG(p->tail != SGNN_TAIL_PX, "node: tail PX (gather-mode edge kernel) exists only in the 64-wide "
                 "tensor-core kernel; this layer shape / precision must use tail PQ");
  const int streamed = node_streamed(*p, use_tensor_cores);
  const NodeSmem L = node_smem(*p, use_tensor_cores, streamed);
  SGNN_CHECK_ARG(p->blob_bytes >= SGNN_VEC_REGION_BYTES, "node: blob lacks the vector region");
  SGNN_CHECK_ARG(L.acc_cols <= 512, "node: needs %d TMEM columns (> 512)", L.acc_cols);
  SGNN_CHECK_ARG(L.total <= 227 * 1024, "node: needs %d bytes of shared memory (> 227 KB); layer shape unsupported", L.total);
  const int n_tiles = sgnn_ceil_div(p->n_rows, kTileRows);
  if (grid > n_tiles) grid = n_tiles;
  if (grid < 1) grid = 1;
  const int width = p->has_mlp ? imax(p->mlp.hidden, p->mlp.out) : (p->tail == SGNN_TAIL_PQ ? p->pq_hidden : p->dec.hidden);
  const int threads = kThreads * sgnn_tile_parts(width, use_tensor_cores != 0);
  if (use_tensor_cores) {
    SGNN_ENSURE_SMEM(node_kernel<true>, L.total);
    SGNN_CUDA(sgnn_launch((node_kernel<true>), dim3(grid), dim3(threads), L.total, (cudaStream_t)stream, *p, streamed));
  } else {
    SGNN_ENSURE_SMEM(node_kernel<false>, L.total);
    SGNN_CUDA(sgnn_launch((node_kernel<false>), dim3(grid), dim3(threads), L.total, (cudaStream_t)stream, *p, streamed));
  }
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
