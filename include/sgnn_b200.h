/*
 * libsgnn_b200 -- C ABI of the B200 (sm_100a) kernels behind the sgnn learned-simulator path.
 *
 * The reference (xrkong/sgnn) is pure Python: it has no FFI.  Its hot path bottoms out in
 * third-party Python/C++ entry points; each function below names the call it replaces
 * (paths relative to the reference tree).  Host code (sgnn_b200/*.py) binds these with ctypes.
 *
 * Conventions: every entry returns 0 on success or a negative code (message through
 * sgnn_last_error()); pointers are DEVICE pointers unless stated; the caller owns all
 * buffers incl. workspaces; every call is asynchronous on `stream` (a cudaStream_t passed
 * as void*), allocates nothing and never synchronises with the host.
 */
#ifndef SGNN_B200_H
#define SGNN_B200_H
#include <stddef.h>
#include "../sgnn_b200/csrc/params.h"

#ifdef __cplusplus
extern "C" {
#endif

const char* sgnn_last_error(void);
int sgnn_version(void);
int sgnn_sm_arch(void);                               /* 100: built for sm_100a only */
int sgnn_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* Neighbour search -> receiver-sorted CSR, bit-exact to
 *   torch_geometric.nn.radius_graph(pos, r, batch, loop, max_num_neighbors)
 * as called at sgnn/single_scale/learned_simulator.py:116-117 (K=20) and
 * sgnn/multi_scale/multi_scale_graph.py:222-227,269-274 (K=24).
 *   pos/pos_stride : particle i at pos + i*pos_stride (floats); dim in [1,3]
 *   example_ptr    : [n_examples+1] int32 boundaries (replaces the `batch` vector built at
 *                    learned_simulator.py:104-106)
 *   rowptr [n+1], src/tgt [n*(K + !loop)] : outputs; E = rowptr[n] stays on the device     */
size_t sgnn_radius_graph_workspace_bytes(int n, int max_neighbors);
int sgnn_radius_graph_launches(int n, int max_neighbors);   /* kernels launched per call */
int sgnn_radius_graph_mode(int mode);   /* 0: general path (default); 1: one 16-CTA cluster kernel for n <= 8192; returns the old mode */
int sgnn_radius_graph_csr(const float* pos, long pos_stride, int n, int dim, const int* example_ptr,
                          int n_examples, double r, int max_neighbors, int loop, void* workspace,
                          size_t workspace_bytes, int* rowptr, int* src, int* tgt, void* stream);

/* nn.Linear weight block -> tensor-core operand image (K-major SWIZZLE_128B); replaces the
 * implicit cuBLAS operand handling of every nn.Linear at graph_network.py:41-42.           */
int sgnn_pack_weight(const float* w, int ld, int row0, int n, int col0, int k, int kpad, float scale,
                     int round_tf32, float* dst, void* stream);
/* row-major latents <-> 128-row operand tiles (module-level API of InteractionNetwork)      */
int sgnn_pack_rows_to_tiles(const float* rows, long ld, const int* perm, int n_rows, int width,
                            int round_tf32, float* tiles, void* stream);
int sgnn_unpack_tiles_to_rows(const float* tiles, const int* perm, int n_rows, int width, float scale,
                              float* rows, long ld, void* stream);

/* Edge features + Encoder.edge_fn: learned_simulator.py:299-312, graph_network.py:92-96,
 * multi_scale_simulator.py:203-238, multi_scale_gnn.py:248-259.                             */
size_t sgnn_edge_encoder_smem_bytes(int blob_bytes, int hidden, int out, int tc);
int sgnn_edge_encoder_fwd(const SgnnEdgeEncoderParams* p, int use_tensor_cores, int grid, void* stream);

/* InteractionNetwork.message + PyG aggregate (index_select x2, cat, edge MLP, LayerNorm,
 * scatter_add): graph_network.py:173,178-199; multi_scale_gnn.py:84-110,132-157,179-205.   */
size_t sgnn_edge_layer_smem_bytes(int blob_bytes, int he, int hidden, int out, int tc);
int sgnn_edge_layer_fwd(const SgnnEdgeLayerParams* p, int use_tensor_cores, int grid, void* stream);
/* 1 if a layer of this shape can run in GATHER mode (SgnnEdgeLayerParams.wxs_off > 0: the sender
 * rows are gathered by TMA and multiplied by W1b inside the first GEMM; the producing node kernels
 * must then use tail SGNN_TAIL_PX).  Same reference lines as sgnn_edge_layer_fwd.               */
int sgnn_edge_layer_gather_supported(int he, int hidden, int out, int n_hidden, int blob_bytes);

/* Node features + Encoder.node_fn (learned_simulator.py:257-290, graph_network.py:86-90);
 * InteractionNetwork.update + residual (graph_network.py:176,201-222); Decoder + Euler
 * integrator (graph_network.py:321-333, learned_simulator.py:381-411).                      */
size_t sgnn_node_smem_bytes(const SgnnNodeParams* p, int tc);
int sgnn_node_fwd(const SgnnNodeParams* p, int use_tensor_cores, int grid, void* stream);

/* Training-path graph primitives (deterministic, no atomics) and stand-alone feature builders.
 *   gather_add_rows : out[e] = P[tgt[e]] + Q[src[e]]   -- PyG propagate's index_select x2 feeding
 *                     the x_i | x_j columns of edge_fn's first Linear (graph_network.py:173,197)
 *   segment_sum_rows: out[i] = sum over CSR row i of m[perm ? perm[k] : k] -- propagate's
 *                     scatter_add aggregate (graph_network.py:173); also the backward of the gathers
 *   gather_rows     : out[e] = rows[idx[e]]             -- backward of segment_sum
 *   node_features / edge_features: learned_simulator.py:257-312, multi_scale_simulator.py:173-238 */
int sgnn_gather_add_rows(const float* P, long ldp, const float* Q, long ldq, const int* tgt,
                         const int* src, int n_edges, int width, float* out, void* stream);
int sgnn_segment_sum_rows(const float* m, const int* rowptr, const int* perm, int n_rows, int width,
                          float* out, void* stream);
/* variants for the fused first Linear of the edge MLP (train_ops._EdgePre): out[e] += P[tgt[e]] + Q[src[e]],
 * and segment sums written into a column block of a wider matrix (row stride ld_out floats).        */
int sgnn_gather_add_rows_acc(const float* P, long ldp, const float* Q, long ldq, const int* tgt,
                             const int* src, int n_edges, int width, float* out, void* stream);
int sgnn_segment_sum_rows_ld(const float* m, const int* rowptr, const int* perm, int n_rows, int width,
                             float* out, long ld_out, void* stream);
int sgnn_gather_rows(const float* rows, long ld, const int* idx, int n_out, int width, float* out,
                     void* stream);

/* nn.LayerNorm over rows of 32 / 64 / 128 floats with autograd support (the LayerNorm closing every
 * encoder / edge / node MLP, graph_network.py:86-96,139-147): forward keeps mean and rstd per row,
 * backward returns dx and the deterministic column sums dgamma, dbeta.                            */
int sgnn_layer_norm_supported(int width);
size_t sgnn_layer_norm_bwd_workspace_bytes(int width);
int sgnn_layer_norm_fwd(const float* x, long n_rows, int width, const float* gamma, const float* beta,
                        float eps, float* y, float* mean, float* rstd, void* stream);
int sgnn_layer_norm_bwd(const float* dy, const float* x, const float* mean, const float* rstd,
                        const float* gamma, long n_rows, int width, float* dx, float* dgamma, float* dbeta,
                        void* workspace, size_t workspace_bytes, void* stream);
/* No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (CUDA graph capture, tight serving loops).  Same result as sgnn_layer_norm_bwd, but the
 * per-CTA dgamma / dbeta partials live in a static __device__ array inside the library (592 KiB per GPU,
 * sized for the maximum width 128; wider rows are rejected) instead of a caller workspace.  Allocates
 * nothing.  All calls on one device share that array: calls that overlap on the GPU race.             */
int sgnn_layer_norm_bwd_static(const float* dy, const float* x, const float* mean, const float* rstd,
                               const float* gamma, long n_rows, int width, float* dx, float* dgamma,
                               float* dbeta, void* stream);
/* Weight and bias gradients of a Linear layer of the training path (autograd of every build_mlp Linear,
 * graph_network.py:21-58): dw[out][in] = sum_r dy[r][out] x[r][in], db[out] = sum_r dy[r][out] (db may be
 * NULL).  dy [n_rows, out_features] and x [n_rows, in_features] are contiguous row-major fp32.  Rows are
 * split over the grid, partial tiles are added in a fixed order: deterministic.                      */
size_t sgnn_linear_wgrad_workspace_bytes(long n_rows, int out_features, int in_features);
int sgnn_linear_wgrad(const float* dy, const float* x, long n_rows, int out_features, int in_features,
                      float* dw, float* db, void* workspace, size_t workspace_bytes, void* stream);
/* Training step as hand-written kernels (sgnn_b200/train_engine.py; csrc/train_mlp.cu): the autograd of every
 * build_mlp Sequential + LayerNorm under loss.backward() (graph_network.py:21-58,86-96,139-222; train.py:230-280).
 *   sgnn_train_pack      all forward / transposed weight images of a step (hi = TF32 part, lo = remainder), one launch
 *   sgnn_train_mlp_fwd   Linear or Linear-ReLU-Linear[-LayerNorm] over row tiles on tcgen05, keeping h1, the LayerNorm
 *                        input and statistics; gathered P[tgt] + Q[src] terms of the edge MLP's first Linear
 *   sgnn_train_mlp_bwd   dpre = (dy W2) * (h1 > 0), dx = dpre W1 routed to two destinations (store / accumulate)
 *                        -- the interaction layer's input-gradient chain ("sgnn_interaction_layer_bwd" of SURVEY 8(b))
 *   sgnn_layer_norm_bwd_rows  LayerNorm backward with the upstream rows gathered by an index (backward of the
 *                        sum-aggregation) and the row count read on the device ("sgnn_mlp_ln_bwd")
 *   sgnn_linear_wgrad_ex weight / bias gradients on operand column blocks, device-side row count, scaled result
 *                        written into a column block of a wider gradient                                      */
/* Edges of a receiver-sorted CSR grouped by sender (deterministic, no full sort): backward of propagate's sender
 * gather (graph_network.py:173); edge count read on the device.                                              */
size_t sgnn_csr_by_source_workspace_bytes(int n_nodes);
int sgnn_csr_by_source(const int* src, const int* n_edges_dev, int max_edges, int n_nodes, int* rp_src, int* perm_src,
                       void* workspace, size_t workspace_bytes, void* stream);
int sgnn_train_pack(const SgnnPackDesc* descs_dev, int n_descs, float* img_hi, float* img_lo, void* stream);
int sgnn_train_mlp_fwd(const SgnnTrainFwd* p, void* stream);
int sgnn_train_mlp_bwd(const SgnnTrainBwd* p, void* stream);
int sgnn_layer_norm_bwd_rows(const float* dy, const int* dy_index, const float* x, const float* mean,
                             const float* rstd, const float* gamma, long max_rows, const int* n_rows_dev, int width,
                             float* dx, float* dgamma, float* dbeta, void* workspace, size_t workspace_bytes,
                             void* stream);
int sgnn_linear_wgrad_ex(const float* dy, long ld_dy, const float* x, long ld_x, long max_rows, const int* n_rows_dev,
                         int out_features, int in_features, float* dw, long ld_dw, float scale, float* db,
                         void* workspace, size_t workspace_bytes, void* stream);
 /* The 64 x 64 weight / bias gradients of up to SGNN_WGRAD_MAX_ITEMS Linears over one row set as ONE tcgen05 launch
 * (csrc/train_wgrad_tc.cu; same autograd as sgnn_linear_wgrad: graph_network.py:21-58 under loss.backward()):
 * operands staged MN-major in shared memory, TF32 products (split = 1: hi / lo-split, fp32-grade), fp32 accumulation
 * in tensor memory, per-slab partials added in a fixed order.                                                 */
size_t sgnn_linear_wgrad_tc_workspace_bytes(int n_items, long max_rows);
int sgnn_linear_wgrad_tc(const SgnnWgradBatch* batch, void* stream);
int sgnn_node_features(const float* pos_seq, long stride, int n, int seq_len, int dim,
                       const float* vel_mean, const float* vel_std, float wall_offset, float wall_max,
                       float wall_div, float* out, long ld_out, void* stream);
int sgnn_edge_features(const float* pos, long stride, int dim, float radius, const int* src,
                       const int* tgt, int n_edges, float* out, void* stream);

/* debug: %globaltimer stamps of the throughput node kernel's phases (CTA 0); NULL switches it off */
int sgnn_debug_node_timeline(unsigned long long* buf);
int sgnn_debug_node_ws_timeline(unsigned long long* buf);
int sgnn_debug_radius_timeline(unsigned long long* buf);    /* radius_cluster_kernel, >= 16 slots; SGNN_TIMELINE builds */   /* node_ws_kernel, >= 32 slots; SGNN_TIMELINE builds */

/* self-test of the GEMM tile primitive (tests/test_gemm_tile_gpu.py)                        */
int sgnn_debug_gemm(const float* a, const float* wimg, int k, int n, float* d, int tc, void* stream);

#ifdef __cplusplus
}
#endif
#endif
