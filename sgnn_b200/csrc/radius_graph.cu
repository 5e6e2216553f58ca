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
#include <cooperative_groups.h>

#include "common.cuh"
#include "scan.cuh"

namespace sgnn {

constexpr int kMaxK = 32;  // max_num_neighbors supported (reference uses 20 and 24)

struct CellGrid {
  float lo[3];
  double inv_h;
  int nc[3];
  int bad;      // the bounding box is not finite (diverged rollout, Inf / NaN input): no edges are emitted
};

// order-preserving float <-> int encoding for atomicMin/atomicMax
__device__ __forceinline__ int f2ord(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// Every thread derives the same grid from the bounding box (a few flops; no extra launch).  The grid is
// valid for ANY input: cells per axis are formed in double and clamped to [1, cell_cap] before the int
// cast (an extent of 1e30 or Inf would overflow it), the cell edge grows until the grid fits the cap, and
// a non-finite box (NaN positions never enter it, Inf ones do) marks the grid `bad`: one cell per example,
// in-range cell indices, and the search emits no edges instead of walking every particle of the example.
__device__ __forceinline__ CellGrid make_grid_bounds(const float* lo, const float* hi, int dim, float h0,
                                                     int n_examples, int cell_cap) {
  CellGrid g;
  double h = (double)h0;
  double ext[3] = {0.0, 0.0, 0.0};
  g.bad = 0;
  for (int d = 0; d < 3; ++d) {
    g.lo[d] = 0.f;
    if (d < dim) {
      const bool fin = isfinite(lo[d]) && isfinite(hi[d]);
      if (!fin && !(lo[d] > hi[d])) g.bad = 1;      // (lo > hi: no particle at all -- an empty box, not a bad one)
      if (fin && hi[d] >= lo[d]) {
        g.lo[d] = lo[d];
        ext[d] = (double)hi[d] - (double)lo[d];
      }
    }
  }
  if (!(h > 0.0) || !isfinite(h)) g.bad = 1;
  const double cap = (double)(cell_cap > 0 ? cell_cap : 1);
  bool fits = false;
  for (int it = 0; it < 400 && !g.bad; ++it) {      // 1.25^64 covers sane inputs; doubling afterwards covers FLT_MAX
    double total = (double)n_examples;
    for (int d = 0; d < 3; ++d) {
      double q = d < dim ? floor(ext[d] / h) + 1.0 : 1.0;
      q = q < 1.0 ? 1.0 : (q > cap ? cap + 1.0 : q);
      g.nc[d] = (int)q;
      total *= q;
    }
    if (total <= cap) { fits = true; break; }
    h *= it < 64 ? 1.25 : 2.0;
  }
  if (!fits) {
    g.nc[0] = g.nc[1] = g.nc[2] = 1;
    if (!g.bad && (double)n_examples > cap) g.bad = 1;
  }
  g.inv_h = g.bad ? 0.0 : 1.0 / h;
  return g;
}
__device__ __forceinline__ CellGrid make_grid(const int* __restrict__ bbox, int dim, float h0,
                                              int n_examples, int cell_cap) {
  float lo[3] = {0.f, 0.f, 0.f}, hi[3] = {0.f, 0.f, 0.f};
  for (int d = 0; d < dim; ++d) { lo[d] = ord2f(bbox[d]); hi[d] = ord2f(bbox[4 + d]); }
  return make_grid_bounds(lo, hi, dim, h0, n_examples, cell_cap);
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
  if (g.bad) { cnt[i] = 0; return; }
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


// ---------------------------------------------------------------------------------------------
// Small systems (n <= 8192): the whole search as ONE kernel on ONE thread-block cluster of 16 CTAs x
// 512 threads, one particle per thread.  At 4800 particles the nine launches of the general path are
// nine ~2.6 us latency chains (23.5 us per rollout step under graph replay, 18 % of the step); a
// fused kernel with software grid barriers measured slower (30 us: every barrier is two global round
// trips).  Here the phases are separated by the hardware cluster barrier (~0.2 us) and everything the
// phases hand to each other -- bounding-box partials, cell counters / starts, the cell-sorted particle
// array, scan totals -- lives in distributed shared memory (a remote access is ~215 cycles against
// ~800 for L2): cells and sorted slots are dealt to the CTAs in contiguous blocks.
namespace cg = cooperative_groups;
// optional phase stamps of CTA 0 / thread 0 (SGNN_TIMELINE builds; profiles/radius_timeline.py)
__device__ unsigned long long* g_rg_timeline = nullptr;
__device__ __forceinline__ void rg_stamp(int slot) {
#ifdef SGNN_TIMELINE
  if (g_rg_timeline != nullptr && blockIdx.x == 0 && threadIdx.x == 0) g_rg_timeline[slot] = (unsigned long long)clock64();
#else
  (void)slot;
#endif
}
constexpr int kClMaxN = 8192;
constexpr int kClThreads = 512;
constexpr int kClCtas = 16;
constexpr int kClCellsPerCta = 2048 + 2;       // power-of-two block of cells per CTA (>= (2 * 8192 + 1024) / 16)

struct ClusterSmem {
  float bb_lo[kClCtas][4], bb_hi[kClCtas][4];   // every CTA's partial box, written by its owner into all copies
  int tot_cells[kClCtas], tot_cnt[kClCtas];     // block totals of the two scans, same scheme
  float w_lo[kClThreads / 32][4], w_hi[kClThreads / 32][4];
  int cell_start[kClCellsPerCta];               // this CTA's block of cell counters, scanned in place
  float4 sorted[kClThreads];                    // this CTA's block of the cell-sorted array (x, y, z, id)
  int sorted_cell[kClThreads];
  int cnt_s[kClThreads];                        // neighbour count / list of THIS CTA's particles (by id), written
  int nbr_s[kClThreads][kMaxK + 1];             // remotely by whichever thread searched for them
};

__global__ void __launch_bounds__(kClThreads)
radius_cluster_kernel(const float* __restrict__ pos, long stride, int n, int dim, const int* __restrict__ ex_ptr,
                      int n_examples, float h0, int cell_cap, float r2, int kq, int loop,
                      int* __restrict__ rowptr, int* __restrict__ src, int* __restrict__ tgt) {
  extern __shared__ __align__(16) unsigned char cl_smem_raw[];
  ClusterSmem& sm = *reinterpret_cast<ClusterSmem*>(cl_smem_raw);
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int i = rank * kClThreads + tid;
  const bool live = i < n;
  pdl_launch_dependents();
  rg_stamp(0);
  for (int k = tid; k < kClCellsPerCta; k += kClThreads) sm.cell_start[k] = 0;
  pdl_wait();
  rg_stamp(1);

  // ---- A: bounding box.  CTA partial -> slot `rank` of every CTA's copy ----
  float4 e = make_float4(0.f, 0.f, 0.f, __int_as_float(i));
  if (live) {
    e.x = pos[(long)i * stride];
    if (dim > 1) e.y = pos[(long)i * stride + 1];
    if (dim > 2) e.z = pos[(long)i * stride + 2];
  }
  {
    float mn[3] = {live ? e.x : INFINITY, live ? e.y : INFINITY, live ? e.z : INFINITY};
    float mx[3] = {live ? e.x : -INFINITY, live ? e.y : -INFINITY, live ? e.z : -INFINITY};
#pragma unroll
    for (int d = 0; d < 3; ++d)
      for (int o = 16; o > 0; o >>= 1) {
        mn[d] = fminf(mn[d], __shfl_xor_sync(0xffffffffu, mn[d], o));
        mx[d] = fmaxf(mx[d], __shfl_xor_sync(0xffffffffu, mx[d], o));
      }
    if (lane == 0)
      for (int d = 0; d < 3; ++d) { sm.w_lo[warp][d] = mn[d]; sm.w_hi[warp][d] = mx[d]; }
    __syncthreads();
    if (tid < kClCtas) {                              // thread t delivers this CTA's partial to CTA t
      float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
      for (int w = 0; w < kClThreads / 32; ++w)
        for (int d = 0; d < 3; ++d) { lo[d] = fminf(lo[d], sm.w_lo[w][d]); hi[d] = fmaxf(hi[d], sm.w_hi[w][d]); }
      ClusterSmem* dst = cluster.map_shared_rank(&sm, tid);
      for (int d = 0; d < 3; ++d) { dst->bb_lo[rank][d] = lo[d]; dst->bb_hi[rank][d] = hi[d]; }
    }
  }
  rg_stamp(2);
  cluster.sync();
  rg_stamp(3);

  // ---- B: grid, cell of every particle, rank within its cell (counter blocks dealt to the CTAs) ----
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int r = 0; r < kClCtas; ++r)
    for (int d = 0; d < 3; ++d) { lo[d] = fminf(lo[d], sm.bb_lo[r][d]); hi[d] = fmaxf(hi[d], sm.bb_hi[r][d]); }
  const CellGrid g = make_grid_bounds(lo, hi, dim, h0, n_examples, cell_cap);
  const int n_cells = n_examples * g.nc[0] * g.nc[1] * g.nc[2];
  // cells per CTA, rounded up to a power of two: owner and slot of a cell are a shift and a mask (the search
  // looks up 6 .. 18 cell starts per particle)
  int csh = 0;
  while ((kClCtas << csh) < n_cells) ++csh;
  const int cpc = 1 << csh, cmask = cpc - 1;
  int c = 0, crank = 0;
  if (live) {
    const float v[3] = {e.x, e.y, e.z};
    c = example_of(ex_ptr, n_examples, i);
    for (int d = dim - 1; d >= 0; --d) c = c * g.nc[d] + cell_coord(v[d], g.lo[d], g.inv_h, g.nc[d]);
    crank = atomicAdd(cluster.map_shared_rank(&sm.cell_start[c & cmask], c >> csh), 1);
  }
  rg_stamp(4);
  cluster.sync();
  rg_stamp(5);

  // ---- C: cell starts: block-local exclusive scan, block totals to everybody, offsets ----
  {
    const int chunk = (cpc + kClThreads - 1) / kClThreads;
    const int k0 = min(tid * chunk, cpc), k1 = min(k0 + chunk, cpc);
    int sum = 0;
    for (int k = k0; k < k1; ++k) sum += sm.cell_start[k];
    int total;
    int run = block_exclusive_scan(sum, total);
    for (int k = k0; k < k1; ++k) { const int t = sm.cell_start[k]; sm.cell_start[k] = run; run += t; }
    if (tid < kClCtas) cluster.map_shared_rank(&sm, tid)->tot_cells[rank] = total;
  }
  cluster.sync();
  {
    int off = 0, mine = 0;
    for (int r = 0; r < kClCtas; ++r) { const int t = sm.tot_cells[r]; if (r < rank) off += t; if (r == rank) mine = t; }
    for (int k = tid; k < cpc; k += kClThreads) sm.cell_start[k] += off;
    if (tid == 0) sm.cell_start[cpc] = off + mine;             // start of the next CTA's first cell
  }
  rg_stamp(6);
  cluster.sync();
  rg_stamp(7);

  // ---- D: particles in cell order, slot s lives in CTA s / 512 ----
  if (live) {
    const int s = *cluster.map_shared_rank(&sm.cell_start[c & cmask], c >> csh) + crank;
    ClusterSmem* dst = cluster.map_shared_rank(&sm, s / kClThreads);
    dst->sorted[s % kClThreads] = e;
    dst->sorted_cell[s % kClThreads] = c;
  }
  rg_stamp(8);
  cluster.sync();
  rg_stamp(9);

  // ---- E: neighbours of the i-th particle in cell order (same rule and order as neighbor_one) ----
  if (live) {
    const float4 q = sm.sorted[tid];
    const int pid = __float_as_int(q.w);
    int cc = sm.sorted_cell[tid];
    const int cx = cc % g.nc[0]; cc /= g.nc[0];
    const int cy = cc % g.nc[1]; cc /= g.nc[1];
    const int cz = cc % g.nc[2];
    const int ex = cc / g.nc[2];
    auto start_of = [&](int cell) -> int {          // cell in [0, n_cells]
      if (cell >= n_cells) return n;
      return *cluster.map_shared_rank(&sm.cell_start[cell & cmask], cell >> csh);
    };
    int best[kMaxK + 1];
    int nb = 0;
    const int zlo = dim > 2 ? max(cz - 1, 0) : 0, zhi = dim > 2 ? min(cz + 1, g.nc[2] - 1) : 0;
    const int ylo = dim > 1 ? max(cy - 1, 0) : 0, yhi = dim > 1 ? min(cy + 1, g.nc[1] - 1) : 0;
    const int xlo = max(cx - 1, 0), xhi = min(cx + 1, g.nc[0] - 1);
    // bounds of all (up to nine) cell rows first: independent remote loads, one latency instead of nine
    int rs0[9], rs1[9];
#pragma unroll
    for (int u = 0; u < 9; ++u) {
      const int z = zlo + u / 3, y = ylo + u % 3;
      rs0[u] = rs1[u] = 0;
      if (z <= zhi && y <= yhi) {
        const int row = ((ex * g.nc[2] + z) * g.nc[1] + y) * g.nc[0];
        rs0[u] = start_of(row + xlo);
        rs1[u] = start_of(row + xhi + 1);
      }
    }
#pragma unroll
    for (int u = 0; u < 9; ++u) {
      if (g.bad) break;                                // non-finite bounding box: no edges (see make_grid_bounds)
      for (int s = rs0[u]; s < rs1[u]; s += 4) {       // four candidates in flight
        float4 o4[4];
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const int sv = min(s + v, rs1[u] - 1);
          o4[v] = cluster.map_shared_rank(&sm, sv / kClThreads)->sorted[sv % kClThreads];
        }
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          if (s + v >= rs1[u]) continue;
          const float4 o = o4[v];
          const int j = __float_as_int(o.w);
          if (nb == kq && j > best[nb - 1]) continue;
          float diff = o.x - q.x;
          float d2 = __fmaf_rn(diff, diff, 0.0f);
          if (dim > 1) { diff = o.y - q.y; d2 = __fmaf_rn(diff, diff, d2); }
          if (dim > 2) { diff = o.z - q.z; d2 = __fmaf_rn(diff, diff, d2); }
          if (!(d2 < r2)) continue;
          int p = nb < kq ? nb++ : nb - 1;
          while (p > 0 && best[p - 1] > j) { best[p] = best[p - 1]; --p; }
          best[p] = j;
        }
      }
    }
    // the list goes to the CTA that owns particle `pid` (its row pointer and edges are written from there)
    ClusterSmem* own = cluster.map_shared_rank(&sm, pid / kClThreads);
    int kept = 0;
    for (int u = 0; u < nb; ++u)
      if (loop || best[u] != pid) own->nbr_s[pid % kClThreads][kept++] = best[u];
    own->cnt_s[pid % kClThreads] = kept;
  }
  rg_stamp(10);
  cluster.sync();
  rg_stamp(11);

  // ---- F: row pointers (scan in particle order) and the compacted edge list ----
  {
    const int v = live ? sm.cnt_s[tid] : 0;
    int total;
    const int prefix = block_exclusive_scan(v, total);
    if (tid < kClCtas) cluster.map_shared_rank(&sm, tid)->tot_cnt[rank] = total;
    cluster.sync();
    int base = prefix;
    for (int r = 0; r < rank; ++r) base += sm.tot_cnt[r];
    if (live) {
      rowptr[i] = base;
      if (i == n - 1) rowptr[n] = base + v;
      for (int k = 0; k < v; ++k) {
        src[base + k] = sm.nbr_s[tid][k];
        if (tgt) tgt[base + k] = i;
      }
    }
  }
  rg_stamp(12);
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
// debug: device buffer of >= 16 uint64 receiving the phase stamps of radius_cluster_kernel (NULL = off)
extern "C" int sgnn_debug_radius_timeline(unsigned long long* buf) {
  SGNN_CUDA(cudaMemcpyToSymbol(g_rg_timeline, &buf, sizeof(buf)));
  return SGNN_OK;
}

// Which path small systems take.  Measured at 4800 particles under graph replay (profiles/README.md, r01o): the
// single-cluster kernel takes 18.6 us against 23.5 us for the nine launches on a regular, static geometry
// (teacher-forced steps), but in an autoregressive rollout -- irregular cells, 30 k instead of 24 k edges -- the
// whole step is 7 us SLOWER with it: a cluster is at most 16 SMs, and 4800 searches on 10 SMs lose against 38
// CTAs on 38 SMs once the search phase is the long one.  So the general path is the default and the cluster
// kernel is opt-in: SGNN_RADIUS_CLUSTER=1 or sgnn_radius_graph_mode(1).
static int g_radius_mode = -1;
static bool use_cluster_path(int n) {
  if (g_radius_mode < 0) g_radius_mode = getenv("SGNN_RADIUS_CLUSTER") != nullptr ? 1 : 0;
  return g_radius_mode == 1 && n <= kClMaxN;
}

// 0 = general nine-launch path for every size (default), 1 = single-cluster kernel for n <= 8192; returns the old mode
extern "C" int sgnn_radius_graph_mode(int mode) {
  if (g_radius_mode < 0) g_radius_mode = getenv("SGNN_RADIUS_CLUSTER") != nullptr ? 1 : 0;
  const int old = g_radius_mode;
  if (mode == 0 || mode == 1) g_radius_mode = mode;
  return old;
}

extern "C" int sgnn_radius_graph_launches(int n, int max_neighbors) {
  if (n <= 0) return 0;
  if (use_cluster_path(n)) return 1;
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

  if (use_cluster_path(n)) {
    static bool attr_set = false;
    if (!attr_set) {
      SGNN_CUDA(cudaFuncSetAttribute(radius_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      SGNN_CUDA(cudaFuncSetAttribute(radius_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(ClusterSmem)));
      attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kClCtas);
    cfg.blockDim = dim3(kClThreads);
    cfg.dynamicSmemBytes = sizeof(ClusterSmem);
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kClCtas;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = getenv("SGNN_NO_PDL") ? 1 : 2;
    SGNN_CUDA(cudaLaunchKernelEx(&cfg, radius_cluster_kernel, pos, pos_stride, n, dim, example_ptr, n_examples, h0,
                                 w.cell_cap, r2, kq, loop, rowptr, src, tgt));
    SGNN_LAUNCH_CHECK();
    return SGNN_OK;
  }
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
