// Self-test entry: D[128 x n] = A[128 x k] * W[n x k]^T through exactly the operand layout,
// descriptors and accumulator path the network kernels use (tcgen05 when tc != 0, FFMA else).
// Used by tests/test_gemm_tile.py to validate the UMMA descriptors in isolation.
#include "mlp_tile.cuh"

namespace sgnn {

struct DbgMisc { uint64_t bar_mma; uint32_t tmem_slot; int pad; };

template <bool TC>
__global__ void __launch_bounds__(kThreads) debug_gemm_kernel(const float* __restrict__ a,
                                                              const float* __restrict__ wimg, int k,
                                                              int n, float* __restrict__ d) {
  char* smem = smem_base_1024();
  char* a_tile = smem;
  char* w_s = smem + kTileRows * k * 4;
  DbgMisc* misc = reinterpret_cast<DbgMisc*>(w_s + n * k * 4);
  float* simt = reinterpret_cast<float*>(smem + align_up(kTileRows * k * 4 + n * k * 4 + (int)sizeof(DbgMisc), 1024));
  TileAcc<TC> acc;
  acc_setup<TC>(acc, pow2_cols(n), &misc->bar_mma, &misc->tmem_slot, simt);
  a_tile_load_rows<TC>(a_tile, 0, a, k, k, kTileRows, [](int r) { return (long)r; });
  for (int i = threadIdx.x; i < n * k; i += kThreads) reinterpret_cast<float*>(w_s)[i] = wimg[i];
  a_tile_publish<TC>();
  gemm_step<TC>(acc, a_tile, w_s, k, n, 0, false);
  gemm_wait<TC>(acc);
  for (int c = 0; c < n; c += 32) {
    float v[32];
    acc_load32<TC>(acc, c, v);
    for (int j = 0; j < 32 && c + j < n; ++j) d[threadIdx.x * n + c + j] = v[j];
  }
  acc_teardown<TC>(acc);
}

// Same product with the A operand staged in TENSOR MEMORY (tcgen05.st by the row-owning thread,
// tcgen05.mma with A from TMEM): the hand-off the fused kernels use between two GEMMs of an MLP.
__global__ void __launch_bounds__(kThreads) debug_gemm_tmem_a_kernel(const float* __restrict__ a,
                                                                     const float* __restrict__ wimg,
                                                                     int k, int n, float* __restrict__ d) {
  char* smem = smem_base_1024();
  char* w_s = smem;
  DbgMisc* misc = reinterpret_cast<DbgMisc*>(w_s + n * k * 4);
  TileAcc<true> acc;
  acc_setup<true>(acc, 512, &misc->bar_mma, &misc->tmem_slot, nullptr);
  for (int i = threadIdx.x; i < n * k; i += kThreads) reinterpret_cast<float*>(w_s)[i] = wimg[i];
  const uint32_t lane_base = (uint32_t)((threadIdx.x >> 5) * 32) << 16;
  const uint32_t a_col = 256;                       // A lives in columns [256, 256 + k)
  for (int c = 0; c < k; c += 32) {
    float v[32];
    for (int j = 0; j < 32; ++j) v[j] = to_tf32(a[threadIdx.x * k + c + j]);
    tmem_st32(acc.tmem_base + lane_base + a_col + c, v);
  }
  tmem_st_wait();
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  if (threadIdx.x == 0) {
    tc_fence_after_sync();
    umma_gemm_tf32_ts(acc.tmem_base, acc.tmem_base + a_col, smem_u32(w_s), k, n);
  }
  gemm_wait<true>(acc);
  for (int c = 0; c < n; c += 32) {
    float v[32];
    acc_load32<true>(acc, c, v);
    for (int j = 0; j < 32 && c + j < n; ++j) d[threadIdx.x * n + c + j] = v[j];
  }
  acc_teardown<true>(acc);
}

}  // namespace sgnn

using namespace sgnn;

// a: [128, k] row-major; wimg: image produced by sgnn_pack_weight (n rows, k cols); d: [128, n]
extern "C" int sgnn_debug_gemm(const float* a, const float* wimg, int k, int n, float* d, int tc,
                               void* stream) {
  SGNN_CHECK_ARG(k % 32 == 0 && k >= 32 && k <= 256, "debug_gemm: bad k");
  SGNN_CHECK_ARG(n % 32 == 0 && n >= 32 && n <= 256, "debug_gemm: bad n");
  const int smem = align_up(kTileRows * k * 4 + n * k * 4 + 64, 1024) + (tc ? 0 : kTileRows * pow2_cols(n) * 4) + 1024;
  SGNN_CHECK_ARG(smem <= 227 * 1024, "debug_gemm: too much shared memory");
  if (tc == 2) {
    SGNN_ENSURE_SMEM(debug_gemm_tmem_a_kernel, smem);
    debug_gemm_tmem_a_kernel<<<1, kThreads, smem, (cudaStream_t)stream>>>(a, wimg, k, n, d);
  } else if (tc) {
    SGNN_ENSURE_SMEM(debug_gemm_kernel<true>, smem);
    debug_gemm_kernel<true><<<1, kThreads, smem, (cudaStream_t)stream>>>(a, wimg, k, n, d);
  } else {
    SGNN_ENSURE_SMEM(debug_gemm_kernel<false>, smem);
    debug_gemm_kernel<false><<<1, kThreads, smem, (cudaStream_t)stream>>>(a, wimg, k, n, d);
  }
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
