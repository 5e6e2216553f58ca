// Exclusive prefix sum over int32 (device-wide), used by the neighbour search twice:
// cell counts -> cell starts, per-receiver neighbour counts -> CSR row pointers.
// out has n + 1 entries (out[n] = total).  In-place (out == in) is allowed.
// HBM-bound integer work: n <= 8192 is one CTA; larger inputs are reduce / scan-of-sums / add.
#pragma once
#include "common.cuh"

namespace sgnn {

constexpr int kScanThreads = 1024;
constexpr int kScanItems = 16;
constexpr int kScanChunk = kScanThreads * kScanItems;  // 16384

// block-wide exclusive scan of one value per thread; returns the exclusive prefix, total via ref
__device__ __forceinline__ int block_exclusive_scan(int v, int& total) {
  __shared__ int warp_sums[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = lane < (blockDim.x >> 5) ? warp_sums[lane] : 0;
    int wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    warp_sums[lane] = wi - w;  // exclusive prefix of warp sums
  }
  __syncthreads();
  const int prefix = warp_sums[warp] + incl - v;
  // total = last warp's exclusive prefix + its sum: have the last thread publish it
  __shared__ int s_total;
  if (threadIdx.x == blockDim.x - 1) s_total = prefix + v;
  __syncthreads();
  total = s_total;
  return prefix;
}

// One chunk of kScanChunk items per CTA.  mode 0: single-CTA full scan (writes out[n]).
// mode 1: chunk-local scan, chunk total -> sums[blockIdx.x].
template <int MODE>
__global__ void __launch_bounds__(kScanThreads) scan_chunk_kernel(const int* __restrict__ in,
                                                                  int* __restrict__ out, int n,
                                                                  int* __restrict__ sums) {
  pdl_launch_dependents();
  pdl_wait();
  const int base = blockIdx.x * kScanChunk + threadIdx.x * kScanItems;
  int v[kScanItems];
  int s = 0;
  // a thread's 16 items are one 64-byte line: four 16-byte loads / stores when it is whole and aligned
  const bool vec = base + kScanItems <= n && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  if (vec) {
#pragma unroll
    for (int i = 0; i < kScanItems / 4; ++i) {
      const int4 q = reinterpret_cast<const int4*>(in + base)[i];
      v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) v[i] = (base + i < n) ? in[base + i] : 0;
  }
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) s += v[i];
  int total;
  int prefix = block_exclusive_scan(s, total);
  if (vec) {
#pragma unroll
    for (int i = 0; i < kScanItems / 4; ++i) {
      int4 q;
      q.x = prefix; prefix += v[4 * i];
      q.y = prefix; prefix += v[4 * i + 1];
      q.z = prefix; prefix += v[4 * i + 2];
      q.w = prefix; prefix += v[4 * i + 3];
      reinterpret_cast<int4*>(out + base)[i] = q;
    }
  } else {
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
      if (base + i < n) out[base + i] = prefix;
      prefix += v[i];
    }
  }
  if (threadIdx.x == 0) {
    if (MODE == 0) out[n] = total;
    else sums[blockIdx.x] = total;
  }
}

static __global__ void scan_add_kernel(int* __restrict__ out, int n, const int* __restrict__ sums,
                                int nblocks) {
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.x;
  const int off = sums[b];
  for (int i = threadIdx.x; i < kScanChunk; i += blockDim.x) {
    int idx = b * kScanChunk + i;
    if (idx < n) out[idx] += off;
  }
  if (b == nblocks - 1 && threadIdx.x == 0) out[n] = sums[nblocks];
}

inline size_t scan_tmp_ints(int n) { return (size_t)sgnn_ceil_div(n, kScanChunk) + 2; }

// tmp: scan_tmp_ints(n) ints.  Launches 1 kernel (n <= 8192) or 3.
inline int exclusive_scan_i32(const int* in, int* out, int n, int* tmp, cudaStream_t st) {
  if (n <= kScanChunk) {
    SGNN_CUDA(sgnn_launch((scan_chunk_kernel<0>), dim3(1), dim3(kScanThreads), 0, st, in, out, n, nullptr));
    SGNN_LAUNCH_CHECK();
    return SGNN_OK;
  }
  const int nb = sgnn_ceil_div(n, kScanChunk);
  SGNN_CHECK_ARG(nb <= kScanChunk, "exclusive_scan_i32: n=%d too large", n);
  SGNN_CUDA(sgnn_launch((scan_chunk_kernel<1>), dim3(nb), dim3(kScanThreads), 0, st, in, out, n, tmp));
  SGNN_LAUNCH_CHECK();
  SGNN_CUDA(sgnn_launch((scan_chunk_kernel<0>), dim3(1), dim3(kScanThreads), 0, st, tmp, tmp, nb, nullptr));
  SGNN_LAUNCH_CHECK();
  SGNN_CUDA(sgnn_launch(scan_add_kernel, dim3(nb), dim3(1024), 0, st, out, n, tmp, nb));
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

}  // namespace sgnn
