// Neighbour search: cell list -> receiver-sorted CSR edge set, bit-exact to the reference's
// `radius_graph(pos, r, batch, loop, max_num_neighbors)` as it runs on CUDA
// (call sites: sgnn/single_scale/learned_simulator.py:116-117, K=20;
//  sgnn/multi_scale/multi_scale_graph.py:222-227,269-274, K=24).
//
// Rule reproduced (torch_cluster radius_kernel, restated in oracle/radius_oracle.c): for every
// target i, the first K sources j in ascending index order within the same example with
// fp32  d2 = fma(dy,dy, dx*dx) ... < (float)(r*r);  output sorted by target, sources ascending.
// The reference does O(N^2) work; this does O(N * candidates of 3^dim cells) and selects the K
// smallest indices with a bounded insertion list, which yields the same set.
//
// All of it is HBM/latency-bound integer work: coalesced loads, one 16-byte load per candidate.
#include "common.cuh"
#include "scan.cuh"

namespace sgnn {

constexpr int kMaxK = 32;  // max_num_neighbors supported (reference uses 20 and 24)

struct CellGrid {
  float lo[3];
  double inv_h;
  int nc[3];
};

// order-preserving float <-> int encoding for atomicMin/atomicMax
__device__ __forceinline__ int f2ord(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// Every thread derives the same grid from the bounding box (a few flops; no extra launch).
__device__ __forceinline__ CellGrid make_grid(const int* __restrict__ bbox, int dim, float h0,
                                              int n_examples, int cell_cap) {
  CellGrid g;
  double h = (double)h0;
  float ext[3] = {0.f, 0.f, 0.f};
  for (int d = 0; d < 3; ++d) {
    g.lo[d] = d < dim ? ord2f(bbox[d]) : 0.f;
    ext[d] = d < dim ? ord2f(bbox[4 + d]) - g.lo[d] : 0.f;
  }
  for (int it = 0; it < 64; ++it) {
    double total = (double)n_examples;
    for (int d = 0; d < 3; ++d) {
      g.nc[d] = d < dim ? (int)floor((double)ext[d] / h) + 1 : 1;
      total *= (double)g.nc[d];
    }
    if (total <= (double)cell_cap) break;
    h *= 1.25;
  }
  g.inv_h = 1.0 / h;
  return g;
}

__device__ __forceinline__ int cell_coord(float x, float lo, double inv_h, int nc) {
  int c = (int)floor(((double)x - (double)lo) * inv_h);
  return min(max(c, 0), nc - 1);
}

// also zeroes the cell counters (replaces a memset node between two kernels)
__global__ void bbox_init_kernel(int* bbox, int* cell_count, int n_cells) {
  pdl_launch_dependents();
  pdl_wait();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_cells; i += gridDim.x * blockDim.x) cell_count[i] = 0;
  if (blockIdx.x == 0 && threadIdx.x < 4) {
    bbox[threadIdx.x] = 0x7fffffff;
    bbox[4 + threadIdx.x] = (int)0x80000000;
  }
}

__global__ void __launch_bounds__(256) bbox_kernel(const float* __restrict__ pos, long stride, int n,
                                                   int dim, int* __restrict__ bbox) {
  pdl_launch_dependents();
  pdl_wait();
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    for (int d = 0; d < dim; ++d) {
      float v = pos[(long)i * stride + d];
      mn[d] = fminf(mn[d], v);
      mx[d] = fmaxf(mx[d], v);
    }
  }
  for (int d = 0; d < dim; ++d) {
    for (int o = 16; o > 0; o >>= 1) {
      mn[d] = fminf(mn[d], __shfl_xor_sync(0xffffffffu, mn[d], o));
      mx[d] = fmaxf(mx[d], __shfl_xor_sync(0xffffffffu, mx[d], o));
    }
    if ((threadIdx.x & 31) == 0 && mn[d] <= mx[d]) {
      atomicMin(&bbox[d], f2ord(mn[d]));
      atomicMax(&bbox[4 + d], f2ord(mx[d]));
    }
  }
}

__device__ __forceinline__ int example_of(const int* __restrict__ ptr, int n_examples, int i) {
  int lo = 0, hi = n_examples;  // largest b with ptr[b] <= i
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (ptr[mid] <= i) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ void cell_count_one(int i, const float* __restrict__ pos, long stride, int dim,
                                               const int* __restrict__ ex_ptr, int n_examples,
                                               const CellGrid& g, int* __restrict__ cell_of,
                                               int* __restrict__ rank, int* __restrict__ cell_count) {
  int c = example_of(ex_ptr, n_examples, i);
  for (int d = dim - 1; d >= 0; --d)
    c = c * g.nc[d] + cell_coord(pos[(long)i * stride + d], g.lo[d], g.inv_h, g.nc[d]);
  cell_of[i] = c;
  rank[i] = atomicAdd(&cell_count[c], 1);
}

__global__ void __launch_bounds__(256)
cell_count_kernel(const float* __restrict__ pos, long stride, int n, int dim,
                  const int* __restrict__ ex_ptr, int n_examples, const int* __restrict__ bbox,
                  float h0, int cell_cap, int* __restrict__ cell_of, int* __restrict__ rank,
                  int* __restrict__ cell_count) {
  pdl_launch_dependents();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const CellGrid g = make_grid(bbox, dim, h0, n_examples, cell_cap);
  cell_count_one(i, pos, stride, dim, ex_ptr, n_examples, g, cell_of, rank, cell_count);
}

__device__ __forceinline__ void cell_fill_one(int i, const float* __restrict__ pos, long stride, int dim,
                                              const int* __restrict__ cell_of, const int* __restrict__ rank,
                                              const int* __restrict__ cell_start, float4* __restrict__ sorted) {
  float4 e;
  e.x = pos[(long)i * stride];
  e.y = dim > 1 ? pos[(long)i * stride + 1] : 0.f;
  e.z = dim > 2 ? pos[(long)i * stride + 2] : 0.f;
  e.w = __int_as_float(i);
  sorted[cell_start[cell_of[i]] + rank[i]] = e;
}

__global__ void __launch_bounds__(256)
cell_fill_kernel(const float* __restrict__ pos, long stride, int n, int dim,
                 const int* __restrict__ cell_of, const int* __restrict__ rank,
                 const int* __restrict__ cell_start, float4* __restrict__ sorted) {
  pdl_launch_dependents();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  cell_fill_one(i, pos, stride, dim, cell_of, rank, cell_start, sorted);
}

// Neighbours of the t-th particle in cell-sorted order (neighbouring threads share candidates).
__device__ __forceinline__ void neighbor_one(int t, int dim, const CellGrid& g, float r2, int kq, int loop,
                                             int slab, const int* __restrict__ cell_of,
                                             const int* __restrict__ cell_start,
                                             const float4* __restrict__ sorted, int* __restrict__ cnt,
                                             int* __restrict__ nbr) {
  const float4 q = sorted[t];
  const int i = __float_as_int(q.w);
  int c = cell_of[i];
  const int cx = c % g.nc[0]; c /= g.nc[0];
  const int cy = c % g.nc[1]; c /= g.nc[1];
  const int cz = c % g.nc[2];
  const int ex = c / g.nc[2];
  int best[kMaxK + 1];
  int nb = 0;
  const int zlo = dim > 2 ? max(cz - 1, 0) : 0, zhi = dim > 2 ? min(cz + 1, g.nc[2] - 1) : 0;
  const int ylo = dim > 1 ? max(cy - 1, 0) : 0, yhi = dim > 1 ? min(cy + 1, g.nc[1] - 1) : 0;
  const int xlo = max(cx - 1, 0), xhi = min(cx + 1, g.nc[0] - 1);
  for (int z = zlo; z <= zhi; ++z)
    for (int y = ylo; y <= yhi; ++y) {
      // cells xlo..xhi of one (example, z, y) row are contiguous in the sorted array
      const int row = ((ex * g.nc[2] + z) * g.nc[1] + y) * g.nc[0];
      const int s0 = cell_start[row + xlo], s1 = cell_start[row + xhi + 1];
      for (int s = s0; s < s1; ++s) {
        const float4 e = __ldg(&sorted[s]);
        const int j = __float_as_int(e.w);
        if (nb == kq && j > best[nb - 1]) continue;
        float diff = e.x - q.x;
        float d2 = __fmaf_rn(diff, diff, 0.0f);
        if (dim > 1) { diff = e.y - q.y; d2 = __fmaf_rn(diff, diff, d2); }
        if (dim > 2) { diff = e.z - q.z; d2 = __fmaf_rn(diff, diff, d2); }
        if (!(d2 < r2)) continue;
        int p = nb < kq ? nb++ : nb - 1;
        while (p > 0 && best[p - 1] > j) { best[p] = best[p - 1]; --p; }
        best[p] = j;
      }
    }
  int kept = 0;
  for (int u = 0; u < nb; ++u)
    if (loop || best[u] != i) nbr[(long)i * slab + kept++] = best[u];
  cnt[i] = kept;
}

__global__ void __launch_bounds__(128)
neighbor_kernel(int n, int dim, const int* __restrict__ ex_ptr, int n_examples,
                const int* __restrict__ bbox, float h0, int cell_cap, float r2, int kq, int loop,
                int slab, const int* __restrict__ cell_of, const int* __restrict__ cell_start,
                const float4* __restrict__ sorted, int* __restrict__ cnt, int* __restrict__ nbr) {
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const CellGrid g = make_grid(bbox, dim, h0, n_examples, cell_cap);
  neighbor_one(t, dim, g, r2, kq, loop, slab, cell_of, cell_start, sorted, cnt, nbr);
}

__global__ void __launch_bounds__(256)
compact_kernel(int n, int slab, const int* __restrict__ rowptr, const int* __restrict__ nbr,
               int* __restrict__ src, int* __restrict__ tgt) {
  pdl_launch_dependents();
  pdl_wait();
  // 8 lanes per receiver: only the filled part of a receiver's slab is read (one thread per slab SLOT
  // cost 101 us at 1M particles, a third of the whole neighbour search: 22 M threads for 5.2 M edges)
  const long gid = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int i = (int)(gid >> 3);
  if (i >= n) return;
  const int beg = rowptr[i], deg = rowptr[i + 1] - beg;
  for (int k = (int)(gid & 7); k < deg; k += 8) {
    src[beg + k] = nbr[(long)i * slab + k];
    if (tgt) tgt[beg + k] = i;
  }
}

struct RadiusWorkspace {
  int* bbox;        // 8
  int* cell_of;     // n
  int* rank;        // n
  int* cell_count;  // cell_cap + 1 (scanned in place -> cell_start)
  int* cnt;         // n + 1
  int* nbr;         // n * slab
  int* scan_tmp;
  float4* sorted;   // n
  int cell_cap;
  size_t bytes;
};

static size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

static RadiusWorkspace carve(void* base, int n, int slab) {
  RadiusWorkspace w;
  w.cell_cap = 2 * n + 1024;
  size_t off = 0;
  char* b = (char*)base;
  auto take = [&](size_t bytes) { char* p = b + off; off += align16(bytes); return p; };
  w.sorted = (float4*)take(sizeof(float4) * (size_t)n);
  w.bbox = (int*)take(sizeof(int) * 8);
  w.cell_of = (int*)take(sizeof(int) * (size_t)n);
  w.rank = (int*)take(sizeof(int) * (size_t)n);
  w.cell_count = (int*)take(sizeof(int) * ((size_t)w.cell_cap + 2));
  w.cnt = (int*)take(sizeof(int) * ((size_t)n + 2));
  w.nbr = (int*)take(sizeof(int) * (size_t)n * slab);
  w.scan_tmp = (int*)take(sizeof(int) * scan_tmp_ints(w.cell_cap + n + 2));
  w.bytes = off;
  return w;
}

}  // namespace sgnn

using namespace sgnn;

extern "C" size_t sgnn_radius_graph_workspace_bytes(int n, int max_neighbors) {
  if (n <= 0) return 256;
  return carve(nullptr, n, max_neighbors + 1).bytes + 256;
}

// number of kernels one sgnn_radius_graph_csr call launches (host-side launch accounting)
extern "C" int sgnn_radius_graph_launches(int n, int max_neighbors) {
  if (n <= 0) return 0;
  const int cell_items = 2 * n + 1024 + 1;
  return 6 + (cell_items <= kScanChunk ? 1 : 3) + (n <= kScanChunk ? 1 : 3);
}

extern "C" int sgnn_radius_graph_csr(const float* pos, long pos_stride, int n, int dim,
                                     const int* example_ptr, int n_examples, double r,
                                     int max_neighbors, int loop, void* workspace,
                                     size_t workspace_bytes, int* rowptr, int* src, int* tgt,
                                     void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  SGNN_CHECK_ARG(n >= 0 && dim >= 1 && dim <= 3, "radius_graph: need 1 <= dim <= 3, got %d", dim);
  SGNN_CHECK_ARG(max_neighbors >= 1 && max_neighbors <= kMaxK,
                 "radius_graph: max_num_neighbors must be in [1,%d], got %d", kMaxK, max_neighbors);
  SGNN_CHECK_ARG(n_examples >= 1 && example_ptr != nullptr, "radius_graph: example_ptr required");
  SGNN_CHECK_ARG(r > 0.0, "radius_graph: radius must be positive");
  if (n == 0) {
    SGNN_CUDA(cudaMemsetAsync(rowptr, 0, sizeof(int), st));
    return SGNN_OK;
  }
  const int slab = max_neighbors + 1;
  SGNN_CHECK_ARG(workspace_bytes >= sgnn_radius_graph_workspace_bytes(n, max_neighbors),
                 "radius_graph: workspace too small");
  SGNN_CHECK_ARG(((uintptr_t)workspace & 15) == 0, "radius_graph: workspace must be 16-byte aligned");
  RadiusWorkspace w = carve(workspace, n, slab);
  const float r2 = (float)(r * r);                 // as torch_cluster: double product, cast once
  const float h0 = (float)(r * (1.0 + 1e-5)) ;     // cell edge just above r
  const int kq = loop ? max_neighbors : max_neighbors + 1;

  SGNN_CUDA(sgnn_launch(bbox_init_kernel, dim3(sgnn_ceil_div(w.cell_cap + 2, 1024)), dim3(256), 0, st, w.bbox,
                        w.cell_count, w.cell_cap + 2));
  const int nb256 = sgnn_ceil_div(n, 256);
  SGNN_CUDA(sgnn_launch(bbox_kernel, dim3(nb256 < 592 ? nb256 : 592), dim3(256), 0, st, pos, pos_stride, n, dim, w.bbox));
  SGNN_CUDA(sgnn_launch(cell_count_kernel, dim3(nb256), dim3(256), 0, st, pos, pos_stride, n, dim, example_ptr, n_examples, w.bbox,
                                           h0, w.cell_cap, w.cell_of, w.rank, w.cell_count));
  SGNN_LAUNCH_CHECK();
  int rc = exclusive_scan_i32(w.cell_count, w.cell_count, w.cell_cap + 1, w.scan_tmp, st);
  if (rc) return rc;
  SGNN_CUDA(sgnn_launch(cell_fill_kernel, dim3(nb256), dim3(256), 0, st, pos, pos_stride, n, dim, w.cell_of, w.rank, w.cell_count,
                                          w.sorted));
  SGNN_CUDA(sgnn_launch(neighbor_kernel, dim3(sgnn_ceil_div(n, 128)), dim3(128), 0, st, n, dim, example_ptr, n_examples, w.bbox, h0,
                                                         w.cell_cap, r2, kq, loop, slab, w.cell_of,
                                                         w.cell_count, w.sorted, w.cnt, w.nbr));
  SGNN_LAUNCH_CHECK();
  rc = exclusive_scan_i32(w.cnt, rowptr, n, w.scan_tmp, st);
  if (rc) return rc;
  SGNN_CUDA(sgnn_launch(compact_kernel, dim3(sgnn_ceil_div((long)n * 8, 256)), dim3(256), 0, st, n, slab, rowptr, w.nbr, src, tgt));
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
