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

  for (int tile = blockIdx.x + w * gridDim.x; tile < n_tiles; tile += gridDim.x * kEncWorkers) {
    const long eidx = (long)tile * kTileRows + wt;
    const bool valid = eidx < n_edges;
    // ---- features ----
    float f[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = 0.f;
    if (valid) {
      if (p.feat) {
#pragma unroll
        for (int k = 0; k < KIN; ++k) f[k] = p.feat[eidx * KIN + k];
      } else {
        const long s = p.src[eidx], t = p.tgt[eidx];
        float sq = 0.f;
#pragma unroll
        for (int d = 0; d < KIN - 1; ++d) {
          const float diff = __fsub_rn(p.pos[s * p.pos_stride + d], p.pos[t * p.pos_stride + d]);
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
      const float* w0 = blob_f + p.w0_off;
      const float* b0 = blob_f + p.mlp.b_off[0];
#pragma unroll
      for (int c = 0; c < EW; c += 32) {
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float a = 0.f;
#pragma unroll
          for (int k = 0; k < KIN; ++k) a = fmaf(f[k], w0[(c + j) * KIN + k], a);
          v[j] = enc_tf32r(fmaxf(a + b0[c + j], 0.f));
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
        const float* b = blob_f + p.mlp.b_off[l];
#pragma unroll
        for (int c = 0; c < EW; c += 32) {
          float h[32];
          tmem_ld32(d + lane_base + c, h);
#pragma unroll
          for (int j = 0; j < 32; ++j) h[j] = enc_tf32r(fmaxf(h[j] + b[c + j], 0.f));
          tmem_st32(d + lane_base + c, h);
        }
      }
      const uint32_t t = a; a = d; d = t;
    }
    // ---- `a` holds the output Linear's accumulator: bias, LayerNorm, TF32, tile image ----
    float y[EW];
    tmem_ld32(a + lane_base, y);
    tmem_ld32(a + lane_base + 32, y + 32);
    {
      const float* b = blob_f + p.mlp.b_off[nh];
#pragma unroll
      for (int j = 0; j < EW; ++j) y[j] += b[j];
      if (p.mlp.has_ln) {
        const float* gam = blob_f + p.mlp.gamma_off;
        const float* bet = blob_f + p.mlp.beta_off;
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
        for (int j = 0; j < EW; ++j) y[j] = y[j] * rstd * gam[j] + bet[j];
      }
    }
    // tile image in shared memory (conflict-free 16-byte stores), then ONE coalesced TMA bulk store
#pragma unroll
    for (int ch = 0; ch < EW / 4; ++ch) {
      float4 o = valid ? make_float4(to_tf32(y[4 * ch]), to_tf32(y[4 * ch + 1]), to_tf32(y[4 * ch + 2]),
                                     to_tf32(y[4 * ch + 3]))
                       : make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<float4*>(stage + sw128_chunk_offset(wt, ch, kTileRows)) = o;
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    enc_worker_bar(w);
    if (wt == 0) {
      tma_bulk_s2g(p.e_tiles + (long)tile * tile_floats, stage, (uint32_t)(tile_floats * 4));
      tma_store_wait_read();          // the stage buffer is rewritten by the next tile
    }
    tc_fence_before_sync();
    enc_worker_bar(w);
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
