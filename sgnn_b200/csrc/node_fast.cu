// Node side of the Encode-Process-Decode network, throughput version for 64-wide latents (same
// maths as node_kernel.cu; reference: Encoder.node_fn graph_network.py:86-90, InteractionNetwork
// .update + residual :176,201-222, Decoder :321-333, Euler integrator learned_simulator.py:381-411).
//
// ONE persistent CTA per SM, 4 tile workers of 128 threads sharing one resident weight image.
// Thread t of a worker owns node row t of the worker's 128-row tile from the first load to the
// last store: it reads its own agg / x rows (full 128-byte lines), rounds them to TF32 and parks
// them in TENSOR MEMORY as the A operand (tcgen05.st); every GEMM of the chain takes A from TMEM
// and B (weights) from shared memory, so there is NO activation tile in shared memory at all and
// no swizzle arithmetic.  TMEM per worker: region A (64 columns) + region D (64 columns);
//   [agg] -> A, MMA (K=64) -> D ; [x] -> A, MMA accumulate -> D ; bias+ReLU in place (D becomes A)
//   hidden..out GEMMs ping-pong between the two regions ; LayerNorm + residual in registers
//   x' (TF32) -> A ; P = x' W1a^T -> D, then Q = x' W1b^T -> D   (or the decoder MLP + integrator)
#include "mlp_tile.cuh"
#include "params.h"

namespace sgnn {

// optional in-kernel timeline (compile with -DSGNN_TIMELINE; profiles/node_timeline.py):
// %globaltimer stamps of CTA 0 / thread 0 into a user buffer.  Compiled out of the product build.
__device__ unsigned long long* g_node_timeline = nullptr;
__device__ __forceinline__ void stamp(int slot) {
#ifdef SGNN_TIMELINE
  if (g_node_timeline != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    g_node_timeline[slot] = t;
  }
#else
  (void)slot;
#endif
}

constexpr int kNodeWorkers = 4;
constexpr int kNodeFastThreads = kNodeWorkers * kThreads;
constexpr int NW = 64;                               // every width handled by this kernel

struct NodeFastMisc {
  uint64_t bar_w;
  uint64_t bar_m[kNodeWorkers];
  uint32_t tmem_slot;
  int pad;
};

__host__ __device__ inline int node_fast_smem(int blob_bytes) {
  return align_up(blob_bytes, 16) + (int)sizeof(NodeFastMisc) + 1024;
}

__device__ __forceinline__ float tf32r(float x) { return __uint_as_float(__float_as_uint(x) + 0x1000u); }

__device__ __forceinline__ void nworker_bar(int w) {
  asm volatile("bar.sync %0, %1;" ::"r"(w + 1), "r"(kThreads) : "memory");
}

// The worker's threads have written an A operand into TMEM: publish it, let one lane of warp 0 issue the
// MMAs described by `issue`, wait for them, and make the accumulator readable by everyone.
template <class Issue>
__device__ __forceinline__ void run_mma(int w, int wt, uint64_t* bar, uint32_t& phase, Issue issue) {
  tmem_st_wait();
  tc_fence_before_sync();
  nworker_bar(w);
  if (wt < 32) {             // the worker's first warp, converged (see elect_one in common.cuh)
    tc_fence_after_sync();
    if (elect_one()) {
      issue();
      tc_commit(bar);
    }
    __syncwarp();
    mbar_wait(bar, phase);
  }
  phase ^= 1u;
  nworker_bar(w);
  tc_fence_after_sync();
}

// 64 fp32 of a row -> TF32 -> TMEM columns [col, col + 64)
__device__ __forceinline__ void row_to_tmem(uint32_t taddr, float* v) {
#pragma unroll
  for (int j = 0; j < NW; ++j) v[j] = tf32r(v[j]);
  tmem_st32(taddr, v);
  tmem_st32(taddr + 32, v + 32);
}

__device__ __forceinline__ void load_row64(const float* __restrict__ src, float* v) {
  const float4* s4 = reinterpret_cast<const float4*>(src);
#pragma unroll
  for (int j = 0; j < NW / 4; ++j) {
    const float4 t = __ldg(s4 + j);
    v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
  }
}

__global__ void __launch_bounds__(kNodeFastThreads, 1) node_fast_kernel(const SgnnNodeParams p) {
  constexpr int TMEM_COLS = 512;
  char* smem = smem_base_1024();
  const float* blob_f = reinterpret_cast<const float*>(smem);
  NodeFastMisc* misc = reinterpret_cast<NodeFastMisc*>(smem + align_up(p.blob_bytes, 16));
  const int tid = threadIdx.x, w = tid >> 7, wt = tid & 127, wq = wt >> 5;
  const int n_tiles = (p.n_rows + kTileRows - 1) / kTileRows;

  pdl_launch_dependents();
  stamp(0);
  if (tid == 0) {
    mbar_init(&misc->bar_w, 1);
    for (int i = 0; i < kNodeWorkers; ++i) mbar_init(&misc->bar_m[i], 1);
    mbar_fence_init();
    tma_load_span(smem, reinterpret_cast<const char*>(p.blob), p.blob_bytes, &misc->bar_w);
  }
  if (tid < 32) tmem_alloc(&misc->tmem_slot, TMEM_COLS);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t lane_base = (uint32_t)(wq * 32) << 16;
  const uint32_t RA = misc->tmem_slot + (uint32_t)(w * 2 * NW);       // column addresses (MMA operands)
  const uint32_t RD = RA + NW;
  const uint32_t tA = RA + lane_base, tD = RD + lane_base;            // this thread's lanes
  uint64_t* bar = &misc->bar_m[w];
  uint32_t phase = 0;
  stamp(1);
  mbar_wait(&misc->bar_w, 0);
  stamp(2);
  pdl_wait();
  stamp(3);          // everything above touched only shared/tensor memory and the constant weights
  const uint32_t sW = smem_u32(smem);
  const int nh = p.mlp.n_hidden;

  for (int tile = blockIdx.x + w * gridDim.x; tile < n_tiles; tile += gridDim.x * kNodeWorkers) {
    const long i = (long)tile * kTileRows + wt;
    const bool valid = i < p.n_rows;
    float v[NW];

    // ---- first Linear: A from the prologue, D = region D ----
    if (p.prologue == SGNN_PRO_AGG_X) {
#pragma unroll
      for (int j = 0; j < NW; ++j) v[j] = 0.f;
      if (valid) {
        const int beg = p.rowptr[i], end = p.rowptr[i + 1];
        if (end > beg) {
          load_row64(p.agg + i * NW, v);
          const int t0 = beg >> 7, t1 = (end - 1) >> 7;
          for (int t = t0 + 1; t <= t1; ++t) {           // CSR row straddling edge tiles (rare)
            const float4* h4 = reinterpret_cast<const float4*>(p.head + (long)t * NW);
#pragma unroll
            for (int j = 0; j < NW / 4; ++j) {
              const float4 h = __ldg(h4 + j);
              v[4 * j] += h.x; v[4 * j + 1] += h.y; v[4 * j + 2] += h.z; v[4 * j + 3] += h.w;
            }
          }
        }
      }
      stamp(4);
      row_to_tmem(tA, v);
      run_mma(w, wt, bar, phase, [&] { umma_gemm_tf32_ts(RD, RA, sW + p.mlp.w_off[0], NW, NW); });
      stamp(5);
#pragma unroll
      for (int j = 0; j < NW; ++j) v[j] = 0.f;
      if (valid) {
        // xr_in: the rounded copy bits(x) + 0x1000 a PX tail left in the Q half of pq -- exactly invertible
        if (p.xr_in) {
          load_row64(p.xr_in + i * p.xr_ld, v);
#pragma unroll
          for (int j = 0; j < NW; ++j) v[j] = __uint_as_float(__float_as_uint(v[j]) - 0x1000u);
        } else {
          load_row64(p.x_in + i * NW, v);
        }
      }
      row_to_tmem(tA, v);
      stamp(6);
      run_mma(w, wt, bar, phase, [&] {           // x columns of W0: K-blocks 2,3 ; accumulate
        const uint32_t idesc = umma_idesc_tf32(kTileRows, NW);
        const uint32_t wb = sW + p.mlp.w_off[0] + 2 * NW * kKBlockBytes;
        for (int kb = 0; kb < 2; ++kb)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            umma_tf32_ts(RD, RA + kb * kKBlock + j * 8, umma_desc_sw128(wb + kb * NW * kKBlockBytes + j * 32),
                         idesc, 1u);
      });
    } else {   // SGNN_PRO_FEATURES: (S-1)*dim velocities + wall distance (+ embedding), K padded to 32
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = 0.f;
      const float* ps = p.pos_seq + (valid ? i : 0) * p.pos_seq_stride;
      // dim == 2, windows of up to 16 frames (the reference's S = 11): the whole window in one batch of
      // independent 8-byte loads and static register indices.  The generic loop below walks the window
      // with dependent scalar loads and a 32-way select per feature: 6 us of the encoder kernel's 11 us
      // at 4800 particles (in-kernel stamps, r01n).  Same arithmetic, same order: bit-identical features.
      const bool win2 = p.dim == 2 && p.seq_len >= 2 && p.seq_len <= 16 && p.emb == nullptr &&
                        (reinterpret_cast<uintptr_t>(p.pos_seq) & 7) == 0 && (p.pos_seq_stride & 1) == 0;
      if (valid && win2) {
        const float2* ps2 = reinterpret_cast<const float2*>(ps);
        float2 wv[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) wv[t] = t < p.seq_len ? __ldg(ps2 + t) : make_float2(0.f, 0.f);
        const float m0 = p.vel_mean[0], m1 = p.vel_mean[1], s0 = p.vel_std[0], s1 = p.vel_std[1];
#pragma unroll
        for (int t = 0; t < 15; ++t)
          if (t + 1 < p.seq_len) {
            f[2 * t] = __fdiv_rn(__fsub_rn(__fsub_rn(wv[t + 1].x, wv[t].x), m0), s0);
            f[2 * t + 1] = __fdiv_rn(__fsub_rn(__fsub_rn(wv[t + 1].y, wv[t].y), m1), s1);
          }
        float last_x = 0.f;
#pragma unroll
        for (int t = 0; t < 16; ++t) if (t == p.seq_len - 1) last_x = wv[t].x;
        float wd = __fadd_rn(last_x, p.wall_offset);
        wd = __fdiv_rn(fminf(fmaxf(wd, 0.f), p.wall_max), p.wall_div);
        const int k = 2 * (p.seq_len - 1);
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j == k) f[j] = wd;
        if (p.feat_out)
          for (int j = 0; j < p.k_in; ++j) {
            float val = 0.f;
#pragma unroll
            for (int q = 0; q < 32; ++q) if (q == j) val = f[q];
            p.feat_out[i * p.k_in + j] = val;
          }
      } else if (valid) {
        int k = 0;
        for (int t = 0; t + 1 < p.seq_len; ++t)
          for (int d = 0; d < p.dim; ++d) {
            const float vel = __fsub_rn(ps[(t + 1) * p.dim + d], ps[t * p.dim + d]);
            const float val = __fdiv_rn(__fsub_rn(vel, p.vel_mean[d]), p.vel_std[d]);
#pragma unroll
            for (int j = 0; j < 32; ++j) if (j == k) f[j] = val;      // registers: static indices only
            ++k;
          }
        float wd = __fadd_rn(ps[(p.seq_len - 1) * p.dim], p.wall_offset);
        wd = __fdiv_rn(fminf(fmaxf(wd, 0.f), p.wall_max), p.wall_div);
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j == k) f[j] = wd;
        if (p.feat_out)
          for (int j = 0; j < p.k_in; ++j) {
            float val = 0.f;
#pragma unroll
            for (int q = 0; q < 32; ++q) if (q == j) val = f[q];
            p.feat_out[i * p.k_in + j] = val;
          }
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = tf32r(f[j]);
      tmem_st32(tA, f);
      run_mma(w, wt, bar, phase, [&] { umma_gemm_tf32_ts(RD, RA, sW + p.mlp.w_off[0], 32, NW); });
    }

    stamp(7);
    // ---- bias + ReLU in place, hidden layers ping-pong A <-> D, then the output Linear ----
    uint32_t cur = RD, oth = RA;                         // cur: region holding the fresh accumulator
    for (int l = 0; l < nh; ++l) {
      const float* b = blob_f + p.mlp.b_off[l];
      const uint32_t tc = cur + lane_base;
#pragma unroll
      for (int c = 0; c < NW; c += 32) {
        float h[32];
        tmem_ld32(tc + c, h);
#pragma unroll
        for (int j = 0; j < 32; ++j) h[j] = tf32r(fmaxf(h[j] + b[c + j], 0.f));
        tmem_st32(tc + c, h);
      }
      const uint32_t a = cur, d = oth;
      const int woff = p.mlp.w_off[l + 1];
      run_mma(w, wt, bar, phase, [&] { umma_gemm_tf32_ts(d, a, sW + woff, NW, NW); });
      cur = d; oth = a;
    }

    stamp(8);
    // ---- LayerNorm, residual, x' ----
    {
      const uint32_t tc = cur + lane_base;
      tmem_ld32(tc, v);
      tmem_ld32(tc + 32, v + 32);
      // bias + LayerNorm in packed fp32x2 (the same operations in the same order as the scalar form:
      // identical bits, half the issue slots -- one warp per scheduler runs this chain at small N)
      const float4* b4 = reinterpret_cast<const float4*>(blob_f + p.mlp.b_off[nh]);
      if (p.mlp.has_ln) {
        const float4* gam4 = reinterpret_cast<const float4*>(blob_f + p.mlp.gamma_off);
        const float4* bet4 = reinterpret_cast<const float4*>(blob_f + p.mlp.beta_off);
        float2 s01 = make_float2(0.f, 0.f), s23 = make_float2(0.f, 0.f);
#pragma unroll
        for (int j = 0; j < NW / 4; ++j) {
          const float4 b = b4[j];
          const float2 a = __fadd2_rn(make_float2(v[4 * j], v[4 * j + 1]), make_float2(b.x, b.y));
          const float2 c = __fadd2_rn(make_float2(v[4 * j + 2], v[4 * j + 3]), make_float2(b.z, b.w));
          v[4 * j] = a.x; v[4 * j + 1] = a.y; v[4 * j + 2] = c.x; v[4 * j + 3] = c.y;
          s01 = __fadd2_rn(s01, a);
          s23 = __fadd2_rn(s23, c);
        }
        const float nmean = -(((s01.x + s01.y) + (s23.x + s23.y)) / (float)NW);
        const float2 nm = make_float2(nmean, nmean);
        float2 q01 = make_float2(0.f, 0.f), q23 = make_float2(0.f, 0.f);
#pragma unroll
        for (int j = 0; j < NW / 4; ++j) {
          const float2 a = __fadd2_rn(make_float2(v[4 * j], v[4 * j + 1]), nm);
          const float2 c = __fadd2_rn(make_float2(v[4 * j + 2], v[4 * j + 3]), nm);
          v[4 * j] = a.x; v[4 * j + 1] = a.y; v[4 * j + 2] = c.x; v[4 * j + 3] = c.y;
          q01 = __ffma2_rn(a, a, q01);
          q23 = __ffma2_rn(c, c, q23);
        }
        const float rstd = 1.0f / sqrtf(((q01.x + q01.y) + (q23.x + q23.y)) / (float)NW + 1e-5f);
        const float2 r2 = make_float2(rstd, rstd);
#pragma unroll
        for (int j = 0; j < NW / 4; ++j) {
          const float4 ga = gam4[j], be = bet4[j];
          const float2 a = __ffma2_rn(__fmul2_rn(make_float2(v[4 * j], v[4 * j + 1]), r2), make_float2(ga.x, ga.y),
                                      make_float2(be.x, be.y));
          const float2 c = __ffma2_rn(__fmul2_rn(make_float2(v[4 * j + 2], v[4 * j + 3]), r2), make_float2(ga.z, ga.w),
                                      make_float2(be.z, be.w));
          v[4 * j] = a.x; v[4 * j + 1] = a.y; v[4 * j + 2] = c.x; v[4 * j + 3] = c.y;
        }
      } else {
#pragma unroll
        for (int j = 0; j < NW / 4; ++j) {
          const float4 b = b4[j];
          v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
        }
      }
      if (valid) {
        if (p.residual) {
          const bool rounded = p.prologue == SGNN_PRO_AGG_X && p.xr_in != nullptr;
          const float4* xr = reinterpret_cast<const float4*>(rounded ? p.xr_in + i * p.xr_ld : p.x_in + i * NW);
          const uint32_t unr = rounded ? 0x1000u : 0u;
#pragma unroll
          for (int j = 0; j < NW / 4; ++j) {
            const float4 x = __ldg(xr + j);
            v[4 * j] += __uint_as_float(__float_as_uint(x.x) - unr);
            v[4 * j + 1] += __uint_as_float(__float_as_uint(x.y) - unr);
            v[4 * j + 2] += __uint_as_float(__float_as_uint(x.z) - unr);
            v[4 * j + 3] += __uint_as_float(__float_as_uint(x.w) - unr);
          }
        }
        if (p.x_out) {
          float4* xo = reinterpret_cast<float4*>(p.x_out + i * NW);
#pragma unroll
          for (int j = 0; j < NW / 4; ++j) xo[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
      }
    }

    stamp(9);
    // ---- tail: x' (TF32) becomes the A operand in the region that is NOT `cur` ... both are free
    //      now (the accumulator has been read), use region A for x' and region D for the products ----
    if (p.tail == SGNN_TAIL_PQ || p.tail == SGNN_TAIL_PX) {
      row_to_tmem(tA, v);              // v is now rounded to TF32 (low mantissa bits are ignored by the MMA)
      if (p.tail == SGNN_TAIL_PX && valid) {     // gather mode: the edge kernel multiplies x_src by W1b itself
        float4* dst = reinterpret_cast<float4*>(p.pq_out + i * 2 * NW + NW);
#pragma unroll
        for (int j = 0; j < NW / 4; ++j) dst[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
      const float* b1 = blob_f + p.pq_bias_off;
      const int n_half = p.tail == SGNN_TAIL_PQ ? 2 : 1;
      for (int half = 0; half < n_half; ++half) {
        const int woff = half ? p.wq_off : p.wp_off;
        run_mma(w, wt, bar, phase, [&] { umma_gemm_tf32_ts(RD, RA, sW + woff, NW, NW); });
        float o[NW];
        tmem_ld32(tD, o);
        tmem_ld32(tD + 32, o + 32);
        if (half == 0) {
#pragma unroll
          for (int j = 0; j < NW; ++j) o[j] += b1[j];
        }
        if (valid) {
          float4* dst = reinterpret_cast<float4*>(p.pq_out + i * 2 * NW + half * NW);
#pragma unroll
          for (int j = 0; j < NW / 4; ++j) dst[j] = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
        }
        tc_fence_before_sync();        // region D is read before the next product overwrites it
      }
    } else if (p.tail == SGNN_TAIL_DECODE) {
      row_to_tmem(tA, v);
      uint32_t a = RA, d = RD;
      for (int l = 0; l < p.dec.n_hidden; ++l) {
        const int woff = p.dec.w_off[l];
        const uint32_t aa = a, dd = d;
        run_mma(w, wt, bar, phase, [&] { umma_gemm_tf32_ts(dd, aa, sW + woff, NW, NW); });
        if (l + 1 < p.dec.n_hidden) {
          const float* b = blob_f + p.dec.b_off[l];
          const uint32_t td = d + lane_base;
#pragma unroll
          for (int c = 0; c < NW; c += 32) {
            float h[32];
            tmem_ld32(td + c, h);
#pragma unroll
            for (int j = 0; j < 32; ++j) h[j] = tf32r(fmaxf(h[j] + b[c + j], 0.f));
            tmem_st32(td + c, h);
          }
        }
        const uint32_t t = a; a = d; d = t;
      }
      // `a` now names the region with the last hidden pre-activation; final Linear in FFMA
      float o[8];
      const int nout = p.dec.out;
      const float* bl = blob_f + p.dec.b_off[p.dec.n_hidden];
      const float* wl = blob_f + p.dec_wlast_off;
      const float* bh = blob_f + p.dec.b_off[p.dec.n_hidden - 1];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = 0.f;
      const uint32_t ta = a + lane_base;
#pragma unroll
      for (int c = 0; c < NW; c += 32) {
        float h[32];
        tmem_ld32(ta + c, h);
#pragma unroll
        for (int j = 0; j < 32; ++j) h[j] = fmaxf(h[j] + bh[c + j], 0.f);
        for (int k = 0; k < nout; ++k) {
          float acc = o[k];
#pragma unroll
          for (int j = 0; j < 32; ++j) acc = fmaf(h[j], wl[k * NW + c + j], acc);
          o[k] = acc;
        }
      }
      if (valid) {
        for (int k = 0; k < nout; ++k) o[k] += bl[k];
        if (p.dec_out)
          for (int k = 0; k < nout; ++k) p.dec_out[i * nout + k] = o[k];
        if (p.strain_out) p.strain_out[i] = o[nout - 1];
        if (p.next_pos) {
          const float* ps = p.pos_seq + i * p.pos_seq_stride;
          for (int d2 = 0; d2 < p.dim; ++d2) {
            const float acc = __fadd_rn(__fmul_rn(o[d2], p.acc_std[d2]), p.acc_mean[d2]);
            const float last = ps[(p.seq_len - 1) * p.dim + d2];
            const float vel = __fsub_rn(last, ps[(p.seq_len - 2) * p.dim + d2]);
            p.next_pos[i * p.next_pos_stride + d2] = __fadd_rn(last, __fadd_rn(vel, acc));
          }
        }
      }
    }
    stamp(10);
    tc_fence_before_sync();
    nworker_bar(w);                  // all TMEM reads of this tile retired before the next tile's stores
  }
  tc_fence_before_sync();
  __syncthreads();
  if (tid < 32) tmem_dealloc(misc->tmem_slot, TMEM_COLS);
  stamp(11);
}

}  // namespace sgnn

using namespace sgnn;

// debug: device buffer of >= 16 uint64 receiving %globaltimer stamps of the node kernel (NULL = off)
extern "C" int sgnn_debug_node_timeline(unsigned long long* buf) {
  SGNN_CUDA(cudaMemcpyToSymbol(g_node_timeline, &buf, sizeof(buf)));
  return SGNN_OK;
}

// 1 = launched, 0 = shape not covered (caller uses the generic kernel), < 0 = error
int sgnn_node_fast_try(const SgnnNodeParams* p, void* stream) {
  if (!p->has_mlp || p->mlp.hidden != NW || p->mlp.out != NW) return 0;
  if (p->prologue == SGNN_PRO_AGG_X) {
    if (p->agg_w != NW || p->x_w != NW) return 0;
  } else if (p->prologue == SGNN_PRO_FEATURES) {
    if (p->k_pad != 32 || p->emb != nullptr || p->dim > 3) return 0;
  } else {
    return 0;
  }
  if (p->residual && p->x_w != NW) return 0;
  if ((p->tail == SGNN_TAIL_PQ || p->tail == SGNN_TAIL_PX) && p->pq_hidden != NW) return 0;
  if (p->tail == SGNN_TAIL_DECODE && (p->dec.hidden != NW || p->dec.out > 8)) return 0;
  const int smem = node_fast_smem(p->blob_bytes);
  if (smem > 227 * 1024) return 0;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    SGNN_CUDA(cudaGetDevice(&dev));
    SGNN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  const int n_tiles = sgnn_ceil_div(p->n_rows, kTileRows);
  int grid = n_tiles < sms ? n_tiles : sms;
  if (grid < 1) grid = 1;
  SGNN_ENSURE_SMEM(node_fast_kernel, smem);
  SGNN_CUDA(sgnn_launch(node_fast_kernel, dim3(grid), dim3(kNodeFastThreads), smem, (cudaStream_t)stream, *p));
  return 1;
}
