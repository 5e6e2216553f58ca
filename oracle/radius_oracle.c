/*
 * TEST INFRASTRUCTURE ONLY -- CPU oracle for the neighbour search of the sgnn hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this. The product (sgnn_b200/) never does.
 *
 * What it restates
 * ----------------
 * The reference calls `torch_geometric.nn.radius_graph(pos, r, batch, loop=True,
 * max_num_neighbors=K)` at
 *   sgnn/single_scale/learned_simulator.py:116-117      (K = 20)
 *   sgnn/multi_scale/multi_scale_graph.py:222-227,269-274 (K = 24)
 * which forwards to torch_cluster.radius.  torch_cluster is a third-party dependency that
 * is NOT vendored under /root/reference and NOT pinned (requirements.txt lists neither it
 * nor torch_geometric; README.md:19-25 suggests the wheels current for torch 2.5.0+cu118).
 * This file restates the published algorithm of torch_cluster's CUDA `radius_kernel`
 * (csrc/cuda/radius_cuda.cu), which is the only well-defined neighbour-truncation rule
 * (the CPU nanoflann path returns traversal-order dependent subsets):
 *
 *   for every query y (ascending), in the x-rows of the same example (ascending):
 *       dist = 0; for d in dim: dist += (x[d]-y[d])*(x[d]-y[d])     // fp32, nvcc contracts to FMA
 *       if (dist < r2) { emit (y, x); if (++count >= K') break; }
 *   r2 = (float)(r*r) with r*r formed in double; K' = K if loop else K+1;
 *   radius_graph then returns stack([x_index (source), y_index (target)]) and, when
 *   loop == False, drops pairs with source == target.
 *
 * "parity unpinned" for this third-party piece: the reference's own tests hold no edge-value
 * vectors; the rule above is recalled from torch_cluster's source.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>


/* fp32 squared distance with the FMA contraction nvcc applies to `dist += a*a`. */
static inline float dist2_fma(const float *a, const float *b, int dim) {
  float dist = 0.0f;
  for (int d = 0; d < dim; ++d) {
    float diff = a[d] - b[d];
    dist = fmaf(diff, diff, dist);
  }
  return dist;
}

/* fp32 squared distance without contraction (separate multiply and add roundings). */
static inline float dist2_nofma(const float *a, const float *b, int dim) {
  volatile float dist = 0.0f;
  for (int d = 0; d < dim; ++d) {
    float diff = a[d] - b[d];
    volatile float sq = diff * diff;
    dist = dist + sq;
  }
  return dist;
}

/*
 * Brute-force radius graph, one example after another.
 *   pos      [n, dim] fp32 row-major
 *   ptr      [n_examples + 1] example boundaries (ptr[0] = 0, ptr[n_examples] = n)
 *   r        connectivity radius (double, as Python passes it)
 *   k        max_num_neighbors
 *   loop     1 keeps self pairs (subject to truncation), 0 applies the K+1-then-drop rule
 *   use_fma  1 = contracted distance (default build of torch_cluster), 0 = uncontracted
 *   nbr      [n, k + 1] int32 scratch/output slab: sources of target i in ascending order
 *   cnt      [n] int32 number of sources kept for target i
 *   q_lo,q_hi  half-open range of queries handled by this call (the Python caller splits
 *              the queries over host threads; there is no OpenMP runtime in the image)
 * Returns the number of edges of that range.  The caller compacts nbr/cnt into edge_index.
 */
long sgnn_oracle_radius_graph(const float *pos, int n, int dim, const int64_t *ptr,
                              int n_examples, double r, int k, int loop, int use_fma,
                              int q_lo, int q_hi, int32_t *nbr, int32_t *cnt) {
  const float r2 = (float)(r * r);
  const int kq = loop ? k : k + 1;
  const int slab = k + 1;
  long total = 0;
  for (int b = 0; b < n_examples; ++b) {
    const int lo = (int)ptr[b], hi = (int)ptr[b + 1];
    for (int i = (lo > q_lo ? lo : q_lo); i < (hi < q_hi ? hi : q_hi); ++i) {
      int c = 0, kept = 0;
      for (int j = lo; j < hi; ++j) {
        float d2 = use_fma ? dist2_fma(pos + (size_t)j * dim, pos + (size_t)i * dim, dim)
                           : dist2_nofma(pos + (size_t)j * dim, pos + (size_t)i * dim, dim);
        if (d2 < r2) {
          if (loop || j != i) nbr[(size_t)i * slab + kept++] = j;
          if (++c >= kq) break;
        }
      }
      cnt[i] = kept;
      total += kept;
    }
  }
  return total;
}

/*
 * Same rule evaluated through a uniform cell list (cell edge >= r), for inputs too large
 * for the O(n^2) loop (the 1M-particle configuration).  Candidates of the 3^dim
 * neighbouring cells are merged by ascending index with a bounded insertion list, which
 * yields exactly the "first K by index" subset.  dim must be 2 or 3.  order[] is the
 * particle list sorted by (example, cell); cell_start[] has n_examples * ncell + 1 entries.
 * These are built by the Python caller with numpy (argsort), keeping this file small.
 */
long sgnn_oracle_radius_graph_cells(const float *pos, int n, int dim, const int32_t *example_of,
                                    const int32_t *cell_xyz, /* [n,3] integer cell coords */
                                    const int32_t *order, const int64_t *cell_start,
                                    const int32_t *ncell_xyz, /* [3] */
                                    double r, int k, int loop, int use_fma, int q_lo, int q_hi,
                                    int32_t *nbr, int32_t *cnt) {
  const float r2 = (float)(r * r);
  const int kq = loop ? k : k + 1;
  const int slab = k + 1;
  const long ncell = (long)ncell_xyz[0] * ncell_xyz[1] * ncell_xyz[2];
  long total = 0;
  for (int i = q_lo; i < (q_hi < n ? q_hi : n); ++i) {
    int32_t best[64];
    int nb = 0; /* sorted ascending, at most kq entries */
    const int cx = cell_xyz[3 * i], cy = cell_xyz[3 * i + 1], cz = cell_xyz[3 * i + 2];
    const int zlo = dim == 3 ? -1 : 0, zhi = dim == 3 ? 1 : 0;
    for (int dz = zlo; dz <= zhi; ++dz)
      for (int dy = -1; dy <= 1; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
          int x = cx + dx, y = cy + dy, z = cz + dz;
          if (x < 0 || y < 0 || z < 0 || x >= ncell_xyz[0] || y >= ncell_xyz[1] ||
              z >= ncell_xyz[2])
            continue;
          long c = (long)example_of[i] * ncell + ((long)z * ncell_xyz[1] + y) * ncell_xyz[0] + x;
          for (int64_t s = cell_start[c]; s < cell_start[c + 1]; ++s) {
            int j = order[s];
            if (nb == kq && j > best[nb - 1]) continue;
            float d2 = use_fma ? dist2_fma(pos + (size_t)j * dim, pos + (size_t)i * dim, dim)
                               : dist2_nofma(pos + (size_t)j * dim, pos + (size_t)i * dim, dim);
            if (!(d2 < r2)) continue;
            int p = nb < kq ? nb++ : nb - 1;
            while (p > 0 && best[p - 1] > j) {
              best[p] = best[p - 1];
              --p;
            }
            best[p] = j;
          }
        }
    int kept = 0;
    for (int t = 0; t < nb; ++t)
      if (loop || best[t] != i) nbr[(size_t)i * slab + kept++] = best[t];
    cnt[i] = kept;
    total += kept;
  }
  return total;
}
