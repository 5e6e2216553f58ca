// Edge encoder, throughput version for 64-wide latents (same maths as edge_encoder.cu; reference:
// edge features learned_simulator.py:299-312 / multi_scale_simulator.py:203-238 and Encoder.edge_fn
// graph_network.py:92-96).  4 tile workers per CTA share the resident weights; thread t owns edge
// row t: features and the K=3 first Linear in FFMA, hidden activations to the tensor core through
// TMEM, LayerNorm in registers, and the row is written straight into its place in the
// SWIZZLE_128B tile image (the 8 chunks of a K-block row are one full 128-byte line, permuted),
// so there is no shared-memory staging and no TMA store.
#include "mlp_tile.cuh"
#include "params.h"

namespace sgnn {

constexpr int kEncWorkers = 4;
constexpr int kEncFastThreads = kEncWorkers * kThreads;
constexpr int EW = 64;

struct EncFastMisc {
  uint64_t bar_w;
  uint64_t bar_m[kEncWorkers];
  uint32_t tmem_slot;
  int pad;
};

__device__ __forceinline__ float enc_tf32r(float x) { return __uint_as_float(__float_as_uint(x) + 0x1000u); }
__device__ __forceinline__ void enc_worker_bar(int w) {
  asm volatile("bar.sync %0, %1;" ::"r"(w + 1), "r"(kThreads) : "memory");
}

template <int KIN>
__global__ void __launch_bounds__(kEncFastThreads, 1) edge_encoder_fast_kernel(const SgnnEdgeEncoderParams p) {
  constexpr int TMEM_COLS = 512;
  char* smem = smem_base_1024();
  const float* blob_f = reinterpret_cast<const float*>(smem);
  const int stage_off = align_up(p.blob_bytes, 1024);
  EncFastMisc* misc = reinterpret_cast<EncFastMisc*>(smem + stage_off + kEncWorkers * kTileRows * EW * 4);
  const int tid = threadIdx.x, w = tid >> 7, wt = tid & 127, wq = wt >> 5;
  char* stage = smem + stage_off + w * (kTileRows * EW * 4);     // this worker's output tile image
  pdl_launch_dependents();
  if (tid == 0) {
    mbar_init(&misc->bar_w, 1);
    for (int i = 0; i < kEncWorkers; ++i) mbar_init(&misc->bar_m[i], 1);
    mbar_fence_init();
    tma_load_span(smem, reinterpret_cast<const char*>(p.blob), p.blob_bytes, &misc->bar_w);
  }
  if (tid < 32) tmem_alloc(&misc->tmem_slot, TMEM_COLS);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t lane_base = (uint32_t)(wq * 32) << 16;
  const uint32_t RA = misc->tmem_slot + (uint32_t)(w * 2 * EW), RD = RA + EW;
  uint64_t* bar = &misc->bar_m[w];
  uint32_t phase = 0;
  mbar_wait(&misc->bar_w, 0);
  pdl_wait();          // everything above touched only shared/tensor memory and the constant weights
  const int n_edges = p.rowptr ? p.rowptr[p.n_nodes] : p.n_nodes;
  const int n_tiles = (n_edges + kTileRows - 1) / kTileRows;
  const uint32_t sW = smem_u32(smem);
  const int nh = p.mlp.n_hidden;
  const int tile_floats = kTileRows * EW;

  // Sender / receiver ids and their positions are fetched ONE TILE AHEAD (ids at the start of the previous
  // tile, positions at its end): a worker is one serial chain per tile, and two dependent global round
  // trips at its head (~1 us) are a quarter of it.
  const int tile_step = gridDim.x * kEncWorkers;
  const bool gather = p.feat == nullptr;
  int s_nx = 0, t_nx = 0;
  float ps_nx[KIN - 1], pt_nx[KIN - 1];
#pragma unroll
  for (int d = 0; d < KIN - 1; ++d) ps_nx[d] = pt_nx[d] = 0.f;
  {
    const long e = (long)(blockIdx.x + w * gridDim.x) * kTileRows + wt;
    if (gather && e < n_edges) {
      s_nx = __ldg(p.src + e);
      t_nx = __ldg(p.tgt + e);
#pragma unroll
      for (int d = 0; d < KIN - 1; ++d) {
        ps_nx[d] = __ldg(p.pos + (long)s_nx * p.pos_stride + d);
        pt_nx[d] = __ldg(p.pos + (long)t_nx * p.pos_stride + d);
      }
    }
  }
  bool stage_busy = false;            // a TMA store issued from the stage buffer has not been waited for

  for (int tile = blockIdx.x + w * gridDim.x; tile < n_tiles; tile += tile_step) {
    const long eidx = (long)tile * kTileRows + wt;
    const bool valid = eidx < n_edges;
    // ---- features ----
    float f[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = 0.f;
    float ps[KIN - 1], pt[KIN - 1];
#pragma unroll
    for (int d = 0; d < KIN - 1; ++d) { ps[d] = ps_nx[d]; pt[d] = pt_nx[d]; }
    const long e_next = eidx + (long)tile_step * kTileRows;
    const bool next_valid = gather && tile + tile_step < n_tiles && e_next < n_edges;
    if (next_valid) {
      s_nx = __ldg(p.src + e_next);
      t_nx = __ldg(p.tgt + e_next);
    }
    if (valid) {
      if (p.feat) {
#pragma unroll
        for (int k = 0; k < KIN; ++k) f[k] = p.feat[eidx * KIN + k];
      } else {
        float sq = 0.f;
#pragma unroll
        for (int d = 0; d < KIN - 1; ++d) {
          const float diff = __fsub_rn(ps[d], pt[d]);
          f[d] = __fdiv_rn(diff, p.radius);
          sq = __fadd_rn(sq, __fmul_rn(f[d], f[d]));
        }
        f[KIN - 1] = sqrtf(sq);
      }
      if (p.feat_out) {
#pragma unroll
        for (int k = 0; k < KIN; ++k) p.feat_out[eidx * KIN + k] = f[k];
      }
    }
    // ---- layer 0 (k_in -> 64) in fp32 FFMA, ReLU -> TMEM region A ----
    {
      // parameters as 16-byte broadcast loads: the scalar version issued ~450 LDS.32 per thread and tile
      // (first-layer weights, four bias / affine vectors), an LSU load of its own at 16 warps per SM
      const float4* w04 = reinterpret_cast<const float4*>(blob_f + p.w0_off);      // [64][KIN] row-major
      const float4* b04 = reinterpret_cast<const float4*>(blob_f + p.mlp.b_off[0]);
#pragma unroll
      for (int c = 0; c < EW; c += 32) {
        float v[32];
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {         // 4 output columns: KIN 16-byte loads of weights, one of bias
          float wv[4 * KIN];
#pragma unroll
          for (int q = 0; q < KIN; ++q) {
            const float4 t4 = w04[((c + 4 * j4) * KIN) / 4 + q];
            wv[4 * q] = t4.x; wv[4 * q + 1] = t4.y; wv[4 * q + 2] = t4.z; wv[4 * q + 3] = t4.w;
          }
          const float4 bb = b04[c / 4 + j4];
          const float bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            float a = 0.f;
#pragma unroll
            for (int k = 0; k < KIN; ++k) a = fmaf(f[k], wv[u * KIN + k], a);
            v[4 * j4 + u] = enc_tf32r(fmaxf(a + bv[u], 0.f));
          }
        }
        tmem_st32(RA + lane_base + c, v);
      }
    }
    uint32_t a = RA, d = RD;
    for (int l = 1; l <= nh; ++l) {
      tmem_st_wait();
      tc_fence_before_sync();
      enc_worker_bar(w);
      if (wt < 32) {           // the worker's first warp, converged (see elect_one in common.cuh)
        tc_fence_after_sync();
        if (elect_one()) {
          umma_gemm_tf32_ts(d, a, sW + p.mlp.w_off[l], EW, EW);
          tc_commit(bar);
        }
        __syncwarp();
        mbar_wait(bar, phase);
      }
      phase ^= 1u;
      enc_worker_bar(w);
      tc_fence_after_sync();
      if (l < nh) {
        const float4* b4 = reinterpret_cast<const float4*>(blob_f + p.mlp.b_off[l]);
#pragma unroll
        for (int c = 0; c < EW; c += 32) {
          float h[32];
          tmem_ld32(d + lane_base + c, h);
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 bb = b4[(c + j) >> 2];
            h[j] = enc_tf32r(fmaxf(h[j] + bb.x, 0.f)); h[j + 1] = enc_tf32r(fmaxf(h[j + 1] + bb.y, 0.f));
            h[j + 2] = enc_tf32r(fmaxf(h[j + 2] + bb.z, 0.f)); h[j + 3] = enc_tf32r(fmaxf(h[j + 3] + bb.w, 0.f));
          }
          tmem_st32(d + lane_base + c, h);
        }
      }
      const uint32_t t = a; a = d; d = t;
    }
    // ---- `a` holds the output Linear's accumulator: bias, LayerNorm, TF32, tile image ----
    float y[EW];
    tmem_ld32(a + lane_base, y);
    tmem_ld32(a + lane_base + 32, y + 32);
    // positions of the next tile's edges: the ids asked for at the top of this tile have arrived by now
    if (next_valid) {
#pragma unroll
      for (int d = 0; d < KIN - 1; ++d) {
        ps_nx[d] = __ldg(p.pos + (long)s_nx * p.pos_stride + d);
        pt_nx[d] = __ldg(p.pos + (long)t_nx * p.pos_stride + d);
      }
    }
    {
      const float4* b4 = reinterpret_cast<const float4*>(blob_f + p.mlp.b_off[nh]);
#pragma unroll
      for (int j = 0; j < EW; j += 4) {
        const float4 bb = b4[j >> 2];
        y[j] += bb.x; y[j + 1] += bb.y; y[j + 2] += bb.z; y[j + 3] += bb.w;
      }
      if (p.mlp.has_ln) {
        const float4* gam4 = reinterpret_cast<const float4*>(blob_f + p.mlp.gamma_off);
        const float4* bet4 = reinterpret_cast<const float4*>(blob_f + p.mlp.beta_off);
        float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < EW; ++j) s4[j & 3] += y[j];
        const float mean = ((s4[0] + s4[1]) + (s4[2] + s4[3])) / (float)EW;
        float q4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < EW; ++j) {
          y[j] -= mean;
          q4[j & 3] = fmaf(y[j], y[j], q4[j & 3]);
        }
        const float rstd = 1.0f / sqrtf(((q4[0] + q4[1]) + (q4[2] + q4[3])) / (float)EW + 1e-5f);
#pragma unroll
        for (int j = 0; j < EW; j += 4) {
          const float4 ga = gam4[j >> 2], be = bet4[j >> 2];
          y[j] = y[j] * rstd * ga.x + be.x; y[j + 1] = y[j + 1] * rstd * ga.y + be.y;
          y[j + 2] = y[j + 2] * rstd * ga.z + be.z; y[j + 3] = y[j + 3] * rstd * ga.w + be.w;
        }
      }
    }
    // tile image in shared memory (conflict-free 16-byte stores), then ONE coalesced TMA bulk store.  The
    // copy engine's read of the PREVIOUS tile's image is waited for only here, a whole tile later.
    if (stage_busy) {
      if (wt == 0) tma_store_wait_read();
      enc_worker_bar(w);
    }
#pragma unroll
    for (int ch = 0; ch < EW / 4; ++ch) {
      float4 o = valid ? make_float4(to_tf32(y[4 * ch]), to_tf32(y[4 * ch + 1]), to_tf32(y[4 * ch + 2]),
                                     to_tf32(y[4 * ch + 3]))
                       : make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<float4*>(stage + sw128_chunk_offset(wt, ch, kTileRows)) = o;
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    enc_worker_bar(w);           // also: every TMEM read of this tile retired before the next tile's tmem_st
    if (wt == 0) tma_bulk_s2g(p.e_tiles + (long)tile * tile_floats, stage, (uint32_t)(tile_floats * 4));
    stage_busy = true;
  }
  if (wt == 0) tma_store_wait_all();
  tc_fence_before_sync();
  __syncthreads();
  if (tid < 32) tmem_dealloc(misc->tmem_slot, TMEM_COLS);
}

}  // namespace sgnn

using namespace sgnn;

int sgnn_edge_encoder_fast_try(const SgnnEdgeEncoderParams* p, void* stream) {
  if (p->mlp.hidden != EW || p->mlp.out != EW) return 0;
  if (!((p->k_in == 3 && (p->feat || p->dim == 2)) || (p->k_in == 4 && (p->feat || p->dim == 3)))) return 0;
  const int smem = align_up(p->blob_bytes, 1024) + kEncWorkers * kTileRows * EW * 4 + (int)sizeof(EncFastMisc) + 1024;
  if (smem > 227 * 1024) return 0;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    SGNN_CUDA(cudaGetDevice(&dev));
    SGNN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  }
  int grid = p->max_tiles < sms ? p->max_tiles : sms;
  if (grid < 1) grid = 1;
  if (p->k_in == 3) {
    SGNN_ENSURE_SMEM(edge_encoder_fast_kernel<3>, smem);
    SGNN_CUDA(sgnn_launch((edge_encoder_fast_kernel<3>), dim3(grid), dim3(kEncFastThreads), smem, (cudaStream_t)stream, *p));
  } else {
    SGNN_ENSURE_SMEM(edge_encoder_fast_kernel<4>, smem);
    SGNN_CUDA(sgnn_launch((edge_encoder_fast_kernel<4>), dim3(grid), dim3(kEncFastThreads), smem, (cudaStream_t)stream, *p));
  }
  return 1;
}
