// Graph primitives of the TRAINING path (forward and backward of the message-passing glue) and
// the stand-alone feature builders.  All of it is HBM-bound row traffic: one warp per row (or per
// CSR segment), 16-byte coalesced accesses, no atomics -> deterministic gradients.
//
//   gather_add   out[e] = P[tgt[e]] + Q[src[e]]            PyG propagate's index_select x2 + the
//                                                          x_i / x_j columns of the edge MLP's
//                                                          first Linear (graph_network.py:173,197)
//   segment_sum  out[i] = sum_{k in row i} m[perm[k]]      scatter_add of propagate (aggregate),
//                                                          and the backward of both gathers
//   gather_rows  out[e] = rows[idx[e]]                     backward of segment_sum
//   node / edge features                                   learned_simulator.py:257-312,
//                                                          multi_scale_simulator.py:173-238
#include "common.cuh"
#include "scan.cuh"

namespace sgnn {

// one warp per edge row; width multiple of 4; ACC: out[e] += ... (the edge term of the first Linear is there already)
template <bool ACC>
__global__ void __launch_bounds__(256)
gather_add_kernel(const float* __restrict__ P, long ldp, const float* __restrict__ Q, long ldq,
                  const int* __restrict__ tgt, const int* __restrict__ src, int n_edges, int width,
                  float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int chunks = width >> 2;
  for (long e = (long)blockIdx.x * 8 + (threadIdx.x >> 5); e < n_edges; e += (long)gridDim.x * 8) {
    const float4* p = reinterpret_cast<const float4*>(P + (long)tgt[e] * ldp);
    const float4* q = reinterpret_cast<const float4*>(Q + (long)src[e] * ldq);
    float4* o = reinterpret_cast<float4*>(out + e * width);
    for (int c = lane; c < chunks; c += 32) {
      const float4 a = __ldg(p + c), b = __ldg(q + c);
      float4 r = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
      if (ACC) {
        const float4 z = o[c];
        r = make_float4(z.x + r.x, z.y + r.y, z.z + r.z, z.w + r.w);
      }
      o[c] = r;
    }
  }
}

// one warp per segment, rows summed in order (deterministic); empty segments give zeros
__global__ void __launch_bounds__(256)
segment_sum_kernel(const float* __restrict__ m, const int* __restrict__ rowptr,
                   const int* __restrict__ perm, int n_rows, int width, float* __restrict__ out, long ld_out) {
  const int lane = threadIdx.x & 31;
  const int chunks = width >> 2;
  for (long i = (long)blockIdx.x * 8 + (threadIdx.x >> 5); i < n_rows; i += (long)gridDim.x * 8) {
    const int beg = rowptr[i], end = rowptr[i + 1];
    for (int c = lane; c < chunks; c += 32) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = beg; k < end; ++k) {
        const long r = perm ? perm[k] : k;
        const float4 v = __ldg(reinterpret_cast<const float4*>(m + r * width) + c);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      reinterpret_cast<float4*>(out + i * ld_out)[c] = s;
    }
  }
}

__global__ void __launch_bounds__(256)
gather_rows_kernel(const float* __restrict__ rows, long ld, const int* __restrict__ idx, int n_out,
                   int width, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int chunks = width >> 2;
  for (long e = (long)blockIdx.x * 8 + (threadIdx.x >> 5); e < n_out; e += (long)gridDim.x * 8) {
    const float4* r = reinterpret_cast<const float4*>(rows + (long)idx[e] * ld);
    float4* o = reinterpret_cast<float4*>(out + e * width);
    for (int c = lane; c < chunks; c += 32) o[c] = __ldg(r + c);
  }
}

// node features: normalised velocities of the window + wall distance (thread per particle)
__global__ void __launch_bounds__(256)
node_features_kernel(const float* __restrict__ pos_seq, long stride, int n, int seq_len, int dim,
                     const float* __restrict__ vel_mean, const float* __restrict__ vel_std,
                     float wall_offset, float wall_max, float wall_div, float* __restrict__ out,
                     long ld_out) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* ps = pos_seq + i * stride;
  float* o = out + i * ld_out;
  int k = 0;
  for (int t = 0; t + 1 < seq_len; ++t)
    for (int d = 0; d < dim; ++d) {
      const float vel = __fsub_rn(ps[(t + 1) * dim + d], ps[t * dim + d]);
      o[k++] = __fdiv_rn(__fsub_rn(vel, vel_mean[d]), vel_std[d]);
    }
  float wd = __fadd_rn(ps[(seq_len - 1) * dim], wall_offset);
  wd = fminf(fmaxf(wd, 0.f), wall_max);
  o[k] = __fdiv_rn(wd, wall_div);
}

// edge features: (p[src] - p[tgt]) / R and its norm (thread per edge)
__global__ void __launch_bounds__(256)
edge_features_kernel(const float* __restrict__ pos, long stride, int dim, float radius,
                     const int* __restrict__ src, const int* __restrict__ tgt, int n_edges,
                     float* __restrict__ out) {
  const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_edges) return;
  const long s = src[e], t = tgt[e];
  float sq = 0.f;
  for (int d = 0; d < dim; ++d) {
    const float f = __fdiv_rn(__fsub_rn(pos[s * stride + d], pos[t * stride + d]), radius);
    out[e * (dim + 1) + d] = f;
    sq = __fadd_rn(sq, __fmul_rn(f, f));
  }
  out[e * (dim + 1) + dim] = sqrtf(sq);
}

// ---- edges grouped by SENDER (the transposed CSR), built without a sort of the whole edge list ----
// count: senders' out-degrees (integer atomics: order-independent); scan -> rp; fill: every edge takes a slot of
// its sender's segment in arrival order; order: each sender's few slots (<= ~40) are sorted by edge id, so the
// result -- and every sum taken over it -- is the same on every run.
__global__ void __launch_bounds__(256)
by_source_count_kernel(const int* __restrict__ src, const int* __restrict__ n_edges_dev, int max_edges, int n_nodes,
                       int* __restrict__ count) {
  const int e_total = min(max(*n_edges_dev, 0), max_edges);
  for (int e = blockIdx.x * 256 + threadIdx.x; e < e_total; e += gridDim.x * 256) {
    const int j = src[e];
    if (j >= 0 && j < n_nodes) atomicAdd(count + j, 1);
  }
}
__global__ void __launch_bounds__(256)
by_source_fill_kernel(const int* __restrict__ src, const int* __restrict__ n_edges_dev, int max_edges, int n_nodes,
                      const int* __restrict__ rp, int* __restrict__ cursor, int* __restrict__ perm) {
  const int e_total = min(max(*n_edges_dev, 0), max_edges);
  for (int e = blockIdx.x * 256 + threadIdx.x; e < e_total; e += gridDim.x * 256) {
    const int j = src[e];
    if (j >= 0 && j < n_nodes) perm[rp[j] + atomicAdd(cursor + j, 1)] = e;
  }
}
__global__ void __launch_bounds__(256)
by_source_order_kernel(const int* __restrict__ rp, int n_nodes, int* __restrict__ perm) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= n_nodes) return;
  const int beg = rp[j], end = rp[j + 1];
  for (int a = beg + 1; a < end; ++a) {              // insertion sort of a short segment
    const int v = perm[a];
    int b = a - 1;
    while (b >= beg && perm[b] > v) { perm[b + 1] = perm[b]; --b; }
    perm[b + 1] = v;
  }
}

static int warp_grid(long rows) {
  long g = (rows + 7) / 8;
  if (g > 148L * 64) g = 148L * 64;
  return g < 1 ? 1 : (int)g;
}

}  // namespace sgnn

using namespace sgnn;

static int gather_add_launch(const float* P, long ldp, const float* Q, long ldq, const int* tgt, const int* src,
                             int n_edges, int width, float* out, bool acc, void* stream) {
  SGNN_CHECK_ARG(width > 0 && width % 4 == 0 && ldp % 4 == 0 && ldq % 4 == 0,
                 "gather_add_rows: widths must be multiples of 4");
  if (n_edges <= 0) return SGNN_OK;
  if (acc)
    gather_add_kernel<true><<<warp_grid(n_edges), 256, 0, (cudaStream_t)stream>>>(P, ldp, Q, ldq, tgt, src, n_edges,
                                                                                 width, out);
  else
    gather_add_kernel<false><<<warp_grid(n_edges), 256, 0, (cudaStream_t)stream>>>(P, ldp, Q, ldq, tgt, src, n_edges,
                                                                                  width, out);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

extern "C" int sgnn_gather_add_rows(const float* P, long ldp, const float* Q, long ldq, const int* tgt,
                                    const int* src, int n_edges, int width, float* out, void* stream) {
  return gather_add_launch(P, ldp, Q, ldq, tgt, src, n_edges, width, out, false, stream);
}

extern "C" int sgnn_gather_add_rows_acc(const float* P, long ldp, const float* Q, long ldq, const int* tgt,
                                        const int* src, int n_edges, int width, float* out, void* stream) {
  return gather_add_launch(P, ldp, Q, ldq, tgt, src, n_edges, width, out, true, stream);
}

extern "C" int sgnn_segment_sum_rows_ld(const float* m, const int* rowptr, const int* perm, int n_rows, int width,
                                        float* out, long ld_out, void* stream) {
  SGNN_CHECK_ARG(width > 0 && width % 4 == 0 && ld_out % 4 == 0 && ld_out >= width,
                 "segment_sum_rows: width and ld_out must be multiples of 4");
  if (n_rows <= 0) return SGNN_OK;
  segment_sum_kernel<<<warp_grid(n_rows), 256, 0, (cudaStream_t)stream>>>(m, rowptr, perm, n_rows, width, out, ld_out);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

extern "C" int sgnn_segment_sum_rows(const float* m, const int* rowptr, const int* perm, int n_rows,
                                     int width, float* out, void* stream) {
  return sgnn_segment_sum_rows_ld(m, rowptr, perm, n_rows, width, out, width, stream);
}

extern "C" int sgnn_gather_rows(const float* rows, long ld, const int* idx, int n_out, int width,
                                float* out, void* stream) {
  SGNN_CHECK_ARG(width > 0 && width % 4 == 0 && ld % 4 == 0, "gather_rows: width must be a multiple of 4");
  if (n_out <= 0) return SGNN_OK;
  gather_rows_kernel<<<warp_grid(n_out), 256, 0, (cudaStream_t)stream>>>(rows, ld, idx, n_out, width, out);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

extern "C" int sgnn_node_features(const float* pos_seq, long stride, int n, int seq_len, int dim,
                                  const float* vel_mean, const float* vel_std, float wall_offset,
                                  float wall_max, float wall_div, float* out, long ld_out, void* stream) {
  SGNN_CHECK_ARG(seq_len >= 2 && dim >= 1 && dim <= 3, "node_features: bad seq_len/dim");
  if (n <= 0) return SGNN_OK;
  node_features_kernel<<<sgnn_ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(
      pos_seq, stride, n, seq_len, dim, vel_mean, vel_std, wall_offset, wall_max, wall_div, out, ld_out);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

extern "C" int sgnn_edge_features(const float* pos, long stride, int dim, float radius, const int* src,
                                  const int* tgt, int n_edges, float* out, void* stream) {
  SGNN_CHECK_ARG(dim >= 1 && dim <= 3, "edge_features: bad dim");
  if (n_edges <= 0) return SGNN_OK;
  edge_features_kernel<<<sgnn_ceil_div(n_edges, 256), 256, 0, (cudaStream_t)stream>>>(pos, stride, dim, radius,
                                                                                     src, tgt, n_edges, out);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}

// Edges of a receiver-sorted CSR grouped by sender: rp_src [n_nodes + 1], perm_src [max_edges] (edge ids, ascending
// inside a sender's segment).  Backward of the sender gather of PyG's propagate (graph_network.py:173): the
// scatter-add over senders becomes sgnn_segment_sum_rows over (rp_src, perm_src).  The edge count is read on the
// device (n_edges_dev, e.g. rowptr + n_nodes).  Workspace: sgnn_csr_by_source_workspace_bytes(n_nodes).
extern "C" size_t sgnn_csr_by_source_workspace_bytes(int n_nodes) {
  return sizeof(int) * ((size_t)2 * (n_nodes + 2) + scan_tmp_ints(n_nodes + 1) + 8);
}
extern "C" int sgnn_csr_by_source(const int* src, const int* n_edges_dev, int max_edges, int n_nodes, int* rp_src,
                                  int* perm_src, void* workspace, size_t workspace_bytes, void* stream) {
  SGNN_CHECK_ARG(n_nodes > 0 && max_edges > 0, "csr_by_source: empty graph");
  SGNN_CHECK_ARG(workspace_bytes >= sgnn_csr_by_source_workspace_bytes(n_nodes), "csr_by_source: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  int* count = (int*)workspace;
  int* cursor = count + (n_nodes + 2);
  int* tmp = cursor + (n_nodes + 2);
  SGNN_CUDA(cudaMemsetAsync(count, 0, sizeof(int) * 2 * (size_t)(n_nodes + 2), st));
  int grid = sgnn_ceil_div(max_edges, 256);
  if (grid > 148 * 8) grid = 148 * 8;
  by_source_count_kernel<<<grid, 256, 0, st>>>(src, n_edges_dev, max_edges, n_nodes, count);
  SGNN_LAUNCH_CHECK();
  const int rc = exclusive_scan_i32(count, rp_src, n_nodes, tmp, st);
  if (rc != SGNN_OK) return rc;
  by_source_fill_kernel<<<grid, 256, 0, st>>>(src, n_edges_dev, max_edges, n_nodes, rp_src, cursor, perm_src);
  SGNN_LAUNCH_CHECK();
  by_source_order_kernel<<<sgnn_ceil_div(n_nodes, 256), 256, 0, st>>>(rp_src, n_nodes, perm_src);
  SGNN_LAUNCH_CHECK();
  return SGNN_OK;
}
