// Common device/host helpers for libsgnn_b200 (sm_100a only).
//
// Tensor-core GEMMs in this library are tcgen05.mma kind::tf32 with both operands in shared
// memory in the canonical K-major SWIZZLE_128B layout and fp32 accumulators in TMEM:
//   * a K-block is 32 fp32 (128 bytes) wide; a tile of R rows x 32 floats is R*128 bytes,
//     rows grouped by 8 (1024-byte swizzle atoms); inside an atom the 16-byte chunk index is
//     XORed with (row & 7)  [Swizzle<3,4,3> on byte addresses]
//   * a wider operand is a sequence of K-blocks, each its own [R x 128 B] tile
//   * one tcgen05.mma consumes K = 8 fp32 (32 bytes); stepping K inside a K-block advances
//     the descriptor start address by 32 bytes
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define SGNN_OK 0
#define SGNN_ERR_ARG -1
#define SGNN_ERR_CUDA -2
#define SGNN_ERR_UNSUPPORTED -3

void sgnn_set_error(const char* fmt, ...);

#define SGNN_CHECK_ARG(cond, ...)        \
  do {                                   \
    if (!(cond)) {                       \
      sgnn_set_error(__VA_ARGS__);       \
      return SGNN_ERR_ARG;               \
    }                                    \
  } while (0)

#define SGNN_CUDA(expr)                                                         \
  do {                                                                          \
    cudaError_t _e = (expr);                                                    \
    if (_e != cudaSuccess) {                                                    \
      sgnn_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),    \
                     __FILE__, __LINE__);                                       \
      return SGNN_ERR_CUDA;                                                     \
    }                                                                           \
  } while (0)

#define SGNN_LAUNCH_CHECK()  SGNN_CUDA(cudaGetLastError())

// Raise a kernel's dynamic shared-memory limit once (never inside a CUDA-graph capture after
// the first, un-captured call).
#define SGNN_ENSURE_SMEM(kernel, bytes)                                                     \
  do {                                                                                      \
    static int _cur = 0;                                                                    \
    if ((bytes) > _cur) {                                                                   \
      SGNN_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                     (bytes)));                                             \
      _cur = (bytes);                                                                       \
    }                                                                                       \
  } while (0)

#ifdef __CUDACC__
#define SGNN_HD __host__ __device__
#else
#define SGNN_HD
#endif
SGNN_HD static inline int sgnn_ceil_div(long a, long b) { return (int)((a + b - 1) / b); }

#ifdef __CUDACC__
#include <stdlib.h>
// Programmatic dependent launch: every hot-path kernel is launched with the programmatic stream
// serialization attribute, calls pdl_launch_dependents() first (so the NEXT kernel's CTAs may be
// scheduled and run their prologue -- barrier init, TMEM allocation, weight staging -- while
// this one still computes) and pdl_wait() before touching anything a predecessor produced or
// still reads.  At 4.8k particles a step is 21 launches of ~10 us: the launch gaps matter.
// SGNN_NO_PDL=1 falls back to plain launches.
template <class... KArgs, class... Args>
static inline cudaError_t sgnn_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                      cudaStream_t st, Args... args) {
  static const bool pdl = getenv("SGNN_NO_PDL") == nullptr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

namespace sgnn {

__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

constexpr int kTileRows = 128;        // rows (edges or nodes) per tensor-core tile = UMMA M
constexpr int kKBlock = 32;           // fp32 per 128-byte swizzled row
constexpr int kKBlockBytes = 128;

// ---- TF32 ---------------------------------------------------------------------------
// round-to-nearest (ties away) to TF32, kept in an fp32 container. The tensor core ignores
// the 13 low mantissa bits; rounding first halves the representation error and unbiases it.
__device__ __forceinline__ float to_tf32(float x) {
  // cvt.rna.tf32.f32 expands to ~5 instructions on sm_100; for every non-NaN input it equals
  // "add half an ulp of the 10-bit mantissa to the magnitude, clear the 13 low bits" (carries
  // into the exponent round up to the next binade, the largest finite values round to inf).
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}

// ---- swizzled K-major tile addressing -------------------------------------------------
// byte offset of element (row, col) of an operand with `rows` rows; col over the full K.
__device__ __forceinline__ uint32_t sw128_offset(int row, int col, int rows) {
  const int kb = col >> 5, c = col & 31;
  return (uint32_t)(kb * rows * kKBlockBytes + (row >> 3) * 1024 + (row & 7) * 128 +
                    ((((c >> 2) ^ (row & 7)) & 7) << 4) + ((c & 3) << 2));
}
// byte offset of the 16-byte chunk `chunk` (4 floats; chunk index over the full K) of `row`
__device__ __forceinline__ uint32_t sw128_chunk_offset(int row, int chunk, int rows) {
  const int kb = chunk >> 3, c = chunk & 7;
  return (uint32_t)(kb * rows * kKBlockBytes + (row >> 3) * 1024 + (row & 7) * 128 +
                    (((c ^ (row & 7)) & 7) << 4));
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier -----------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// wait with a long hardware suspend hint: 16 epilogue warps that SPIN on try_wait compete with the
// working warps for instruction fetch (ncu r01h: "no instruction" was 46 % of all stall samples)
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WS_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
      "@p bra WS_DONE;\n\t"
      "bra WS_WAIT;\n\t"
      "WS_DONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity), "r"(0x989680u)
      : "memory");
}
// The single-thread issuers and the producers poll instead: waking from the hardware suspend costs
// ~0.4 us (in-kernel timeline), which the serial issue chains cannot afford five times per tile.
__device__ __forceinline__ void mbar_wait_spin(uint64_t* bar, uint32_t parity) {
  while (!mbar_test(bar, parity)) {
  }
}

// Epilogue waits: suspend in the throughput regime (many tiles per CTA in flight, the suspended warps
// leave the issue slots and the instruction fetch to the working ones), poll in the latency regime (a
// CTA owns one or two tiles -- 4800-particle steps -- and its kernel is ONE serial chain of five such
// waits: in-kernel stamps put ~0.4 us of wake-up latency on every suspended wait, r01n).
// 256-bit global accesses (sm_100: LDG / STG .ENL2.256): one full 32-byte sector per lane and instruction.
// The address must be 32-byte aligned.
__device__ __forceinline__ void ldg256_nc(const float* src, float* v) {
  asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "l"(src));
}
__device__ __forceinline__ void stg256(float* dst, float a, float b, float c, float d, float e, float f, float g,
                                       float h) {
  asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "f"(a), "f"(b), "f"(c), "f"(d),
               "f"(e), "f"(f), "f"(g), "f"(h)
               : "memory");
}

__device__ __forceinline__ void mbar_wait_mode(uint64_t* bar, uint32_t parity, bool poll) {
  if (poll) mbar_wait_spin(bar, parity);
  else mbar_wait_sleep(bar, parity);
}

// ---- TMA bulk copy (global -> shared, completes on an mbarrier) ------------------------
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// generic-proxy writes to shared memory -> visible to the async proxy (tcgen05.mma, TMA)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- tcgen05 / TMEM -------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_result)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before_sync() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after_sync() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// all previously issued tcgen05.mma of this thread arrive on `bar` when complete
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}

// One elected lane of a CONVERGED warp (elect.sync).  tcgen05.mma / commit / TMA take their operands
// from uniform registers; under `if (threadIdx.x == 0)` the warp is divergent, the compiler may not
// use the uniform datapath there and wraps EVERY such instruction in a ~20-instruction ELECT /
// R2UR.BROADCAST / BRA.U.ANY loop (~110 cycles per MMA, measured with the in-kernel timeline of
// edge_layer_ws.cu).  Issue sites therefore run in whole, converged warps and put only the
// tcgen05 / TMA instructions under this predicate.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\t"
      "elect.sync rx|px, %1;\n\t"
      "@px mov.s32 %0, 1;\n\t}"
      : "+r"(pred)
      : "r"(0xffffffffu));
  return pred != 0;
}

// shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);   // start address  [0,14)
  d |= (uint64_t)1 << 16;                        // leading byte offset (unused for SW128 K-major)
  d |= (uint64_t)(1024 >> 4) << 32;              // stride byte offset [32,46)
  d |= (uint64_t)1 << 46;                        // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                        // layout type: SWIZZLE_128B
  return d;
}
// instruction descriptor: D fp32, A/B tf32, both K-major, shape M x N
__device__ __forceinline__ uint32_t umma_idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(m >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]^T, K = 8 (tf32)
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Issue the MMAs of one GEMM: D[128 x n] = A[128 x k] * W[n x k]^T (k multiple of 32).
// a_base / w_base: shared addresses of the operands (1024-byte aligned K-block tiles).
__device__ __forceinline__ void umma_gemm_tf32(uint32_t tmem_d, uint32_t a_base, uint32_t w_base,
                                               int k, int n, bool accumulate_first = false) {
  const uint32_t idesc = umma_idesc_tf32(kTileRows, n);
  uint32_t acc = accumulate_first ? 1u : 0u;
  for (int kb = 0; kb < k / kKBlock; ++kb) {
    const uint32_t a_kb = a_base + kb * kTileRows * kKBlockBytes;
    const uint32_t w_kb = w_base + kb * n * kKBlockBytes;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      umma_tf32(tmem_d, umma_desc_sw128(a_kb + j * 32), umma_desc_sw128(w_kb + j * 32), idesc, acc);
      acc = 1u;
    }
  }
}

// D[tmem] (+)= A[tmem] * B[smem]^T, K = 8 (tf32).  A: 128 lanes x 8 consecutive 32-bit columns.
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b,
                                             uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[128 x n] = A[128 x k] * W[n x k]^T with A in TMEM (columns tmem_a .. tmem_a + k)
__device__ __forceinline__ void umma_gemm_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t w_base,
                                                  int k, int n) {
  const uint32_t idesc = umma_idesc_tf32(kTileRows, n);
  uint32_t acc = 0u;
  for (int kb = 0; kb < k / kKBlock; ++kb) {
    const uint32_t w_kb = w_base + kb * n * kKBlockBytes;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      umma_tf32_ts(tmem_d, tmem_a + kb * kKBlock + j * 8, umma_desc_sw128(w_kb + j * 32), idesc, acc);
      acc = 1u;
    }
  }
}
// registers -> TMEM: this warp's 32 lanes x 32 consecutive 32-bit columns starting at `taddr`
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
      "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
      "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])),
      "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
      "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])),
      "r"(__float_as_uint(v[15])), "r"(__float_as_uint(v[16])), "r"(__float_as_uint(v[17])),
      "r"(__float_as_uint(v[18])), "r"(__float_as_uint(v[19])), "r"(__float_as_uint(v[20])),
      "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])), "r"(__float_as_uint(v[23])),
      "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])), "r"(__float_as_uint(v[26])),
      "r"(__float_as_uint(v[27])), "r"(__float_as_uint(v[28])), "r"(__float_as_uint(v[29])),
      "r"(__float_as_uint(v[30])), "r"(__float_as_uint(v[31]))
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// TMEM -> registers: this warp's 32 lanes x 32 consecutive fp32 columns starting at `taddr`
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// two 32-column loads in flight, one wait: a row's 64 consecutive fp32 columns starting at `taddr`
__device__ __forceinline__ void tmem_ld64(uint32_t taddr, float* v) {
  uint32_t r[64];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
      "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
      "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]),
        "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]),
        "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]),
        "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]),
        "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 64; ++i) v[i] = __uint_as_float(r[i]);
}

}  // namespace sgnn
#endif  // __CUDACC__
