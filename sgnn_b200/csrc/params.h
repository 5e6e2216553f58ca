// Plain-C parameter blocks of the network kernels.  They are part of the C ABI
// (include/sgnn_b200.h re-exports them): the Python host side fills them with ctypes.
//
// "blob": one contiguous device buffer per kernel holding every GEMM weight as a K-major
// SWIZZLE_128B image (sgnn_pack_weight; 1024-byte aligned offsets, in BYTES) followed by the
// fp32 vectors (biases, LayerNorm gamma/beta, small dense layers; offsets in FLOATS).  It is
// staged into shared memory once per CTA with TMA bulk copies and stays resident.
#pragma once
#include <stdint.h>

#define SGNN_MAX_HIDDEN 3   /* hidden layers per MLP (reference uses 1 and 2) */
/* Blob layout: bytes [0, SGNN_VEC_REGION_BYTES) hold the fp32 vectors, the GEMM images follow at
 * 1024-byte aligned offsets.  Kernels whose images do not fit in shared memory keep only the
 * vector region resident and stream the images per GEMM. */
#define SGNN_VEC_REGION_BYTES 8192

#ifdef __cplusplus
extern "C" {
#endif

/* Linear(k_in -> hidden) ReLU [Linear(hidden->hidden) ReLU]* Linear(hidden -> out) [LayerNorm] */
typedef struct {
  int n_hidden;                      /* number of hidden layers (nmlp_layers) >= 1              */
  int hidden;                        /* hidden width, multiple of 32                            */
  int out;                           /* output width                                            */
  int has_ln;                        /* LayerNorm(out) after the last Linear                    */
  int w_off[SGNN_MAX_HIDDEN + 1];    /* byte offsets of the GEMM images, layer 0 .. n_hidden    */
  int b_off[SGNN_MAX_HIDDEN + 1];    /* float offsets of the biases (-1: no bias / folded)      */
  int gamma_off, beta_off;           /* float offsets of the LayerNorm affine parameters        */
} SgnnMlp;

/* Per message-passing layer, edge side:  m_e = LN(MLP(e W1c^T + P[tgt] + Q[src])),
 * agg[i] = sum of m_e over the CSR row of receiver i.  Reference: InteractionNetwork.message +
 * PyG aggregate (sgnn/single_scale/graph_network.py:178-199). */
typedef struct {
  const float* e_tiles;   /* edge latents e0, 128-row tiles in A-operand layout, width he      */
  const float* pq;        /* [n_nodes, 2*hidden]: P = x W1a^T + b1 | Q = x W1b^T               */
  const int* src;         /* [E] sender of each edge (CSR order)                               */
  const int* tgt;         /* [E] receiver of each edge, ascending                              */
  const int* rowptr;      /* [n_nodes + 1]; E = rowptr[n_nodes] is read on the device          */
  int n_nodes;
  int he;                 /* edge latent width (K of the first GEMM), multiple of 32           */
  const float* blob;
  int blob_bytes;
  SgnnMlp mlp;            /* layer 0 image = W1c * 2^layer (the e' = 2e quirk), no bias        */
  float* agg;             /* [n_nodes, out]; rows of receivers without edges are NOT written   */
  float* head;            /* [n_tiles, out]: partial sum of a CSR row continuing from the      */
                          /* previous tile (added back by the node kernel)                     */
  int max_tiles;
  int wxs_off;            /* 0: PQ mode.  > 0: GATHER mode -- byte offset of the W1b image; the Q half */
                          /* of `pq` then holds the TF32-rounded node latents x (node tail            */
                          /* SGNN_TAIL_PX) and the sender term x_src W1b^T is part of the first GEMM  */
  int defer_affine;       /* GATHER mode, 1: agg / head hold the plain sums of the normalised messages; the    */
                          /* LayerNorm affine (gamma * sum + count * beta) is applied by the node kernel      */
                          /* (SgnnNodeParams.agg_affine)                                                       */
} SgnnEdgeLayerParams;

/* Edge encoder: features from positions -> MLP -> LN -> e0 tiles.  Reference:
 * LearnedSimulator._encoder_preprocessor edge part (learned_simulator.py:299-312) and
 * Encoder.edge_fn (graph_network.py:92-96). */
typedef struct {
  const float* pos;       /* most recent positions: particle i at pos + i * pos_stride         */
  long pos_stride;
  int dim;
  float inv_radius_unused;
  float radius;           /* displacements are DIVIDED by this (fp32 division, as torch)       */
  const int* src;
  const int* tgt;
  const int* rowptr;
  int n_nodes;
  const float* feat;      /* optional explicit edge features [E, k_in] (module API); else NULL */
  int k_in;               /* dim + 1 when computed from positions                              */
  const float* blob;
  int blob_bytes;
  SgnnMlp mlp;            /* layer 0 is evaluated with FFMA from w0_off (plain [hidden,k_in])  */
  int w0_off;             /* float offset of the plain row-major first-layer weight            */
  float* e_tiles;         /* out: tiles, width mlp.out                                         */
  float* feat_out;        /* optional [E, k_in] copy of the features (tests / module API)      */
  int max_tiles;
} SgnnEdgeEncoderParams;

enum { SGNN_PRO_FEATURES = 0, SGNN_PRO_AGG_X = 1, SGNN_PRO_ROWS = 2 };
enum { SGNN_TAIL_NONE = 0, SGNN_TAIL_PQ = 1, SGNN_TAIL_DECODE = 2, SGNN_TAIL_PX = 3 };

/* Node-side kernel: prologue -> MLP (-> LN) (-> + residual) -> x_out, then an optional tail
 * GEMM on x_out: PQ for the next layer's edge kernel (PX: P and the TF32-rounded x_out in Q's
 * place, for an edge kernel in gather mode), or the decoder + Euler integrator. */
typedef struct {
  int n_rows;
  int prologue;
  /* FEATURES: node features from the position window (learned_simulator.py:257-290,
   * multi_scale_simulator.py:173-201) */
  const float* pos_seq;   /* window of particle i: pos_seq + i * pos_seq_stride, [S, dim] */
  long pos_seq_stride;    /* floats between particles (S*dim when contiguous)          */
  int seq_len, dim;
  const float* vel_mean;  /* [dim] */
  const float* vel_std;   /* [dim] */
  float wall_offset;      /* 2.0: distance to the wall at x = -2 */
  float wall_max;         /* clamp upper bound (connectivity radius / grid radius) */
  float wall_div;         /* 1.0 single-scale; grid radius multi-scale (fp32 division) */
  const int64_t* types;   /* [n] particle types, used when emb != NULL */
  const float* emb;       /* [n_types, emb_dim] or NULL */
  int emb_dim;
  float* feat_out;        /* optional [n, k_in] copy of the features */
  /* AGG_X: [aggregate | x] of InteractionNetwork.update (graph_network.py:201-222) */
  const float* agg;       /* [n, agg_w] */
  const float* head;      /* [n_tiles_e, agg_w] */
  const int* rowptr;      /* [n + 1] */
  int agg_w;
  const float* x_in;      /* [n, x_w] (also the residual) */
  int x_w;
  /* ROWS: explicit input rows */
  const float* rows_in;   /* [n, k_in] */
  long rows_ld;
  int k_in;               /* logical input width of the MLP (features / rows) */
  int k_pad;              /* k_in rounded up to 32 (AGG_X: agg_w + x_w, both multiples of 32) */
  const float* blob;
  int blob_bytes;
  int has_mlp;            /* 0: skip the MLP (tail applied directly to the ROWS input)         */
  SgnnMlp mlp;
  int residual;           /* x_out = x_in + LN(MLP(..)) */
  float* x_out;           /* [n, mlp.out] or NULL */
  /* tail */
  int tail;
  int pq_hidden;          /* PQ: width of P and of Q (next layer's hidden width)               */
  int wp_off, wq_off;     /* byte offsets of the W1a / W1b images (N = pq_hidden, K = out)     */
  int pq_bias_off;        /* float offset of b1 (added to P)                                   */
  float* pq_out;          /* [n, 2*pq_hidden]                                                  */
  SgnnMlp dec;            /* DECODE: hidden layers as GEMMs; the last Linear (hidden -> out,   */
  int dec_wlast_off;      /* out <= 8) is a plain row-major [out, hidden] fp32 matrix (floats) */
  float* dec_out;         /* optional raw network output [n, dec.out]                          */
  const float* acc_mean;  /* integrator (learned_simulator.py:381-411); NULL = skip            */
  const float* acc_std;
  float* next_pos;        /* [n, dim] */
  long next_pos_stride;
  float* strain_out;      /* [n] */
  const float* xr_in;     /* AGG_X, optional: x_in rounded to TF32 (row i at xr_in + i * xr_ld), as written */
  long xr_ld;             /* by a node kernel with tail PX into the Q half of pq; enables node_ws_kernel    */
  int n_types;            /* rows of `emb`; type ids are clamped to [0, n_types) (0: not clamped)           */
  int agg_affine;         /* AGG_X, 1: agg / head are plain sums (edge kernel with defer_affine); the row becomes    */
  int agg_gamma_off;      /* gamma * (agg + head) + (rowptr[i+1] - rowptr[i]) * beta with the EDGE MLP's LayerNorm  */
  int agg_beta_off;       /* parameters at these float offsets of the blob (node_ws_kernel only)                    */
} SgnnNodeParams;


/* ---- training path: row-tile kernels with activations kept for the backward pass (csrc/train_mlp.cu) ----
 * One launch = one Linear (n_linear 1) or Linear-ReLU-Linear[-LayerNorm] (n_linear 2) of build_mlp
 * (graph_network.py:21-58) over `rows` rows, row-major fp32 in and out.  The row count may live on the
 * device (n_rows_dev: the edge count rowptr[n] of this step), so that a whole training step has no host
 * read and can be captured in a CUDA graph; buffers are sized by max_rows.
 * Weights are K-major SWIZZLE_128B images made by sgnn_train_pack (hi = TF32-rounded value, lo = the
 * remainder).  split = 1 runs every product as hi*hi + lo*hi + hi*lo (fp32-grade accuracy on the
 * tensor cores), split = 0 as hi*hi only (plain TF32). */
enum { SGNN_TPRO_ROWS = 0, SGNN_TPRO_ROWS2 = 1, SGNN_TPRO_EDGE_FEATURES = 2 };

typedef struct {
  int max_rows;
  const int* n_rows_dev;   /* NULL: max_rows rows; else min(*n_rows_dev, max_rows), read on the device   */
  int prologue;
  const float* a0;         /* ROWS / ROWS2: input columns [0, k0): row r at a0 + r * ld0, times scale0    */
  long ld0;
  int k0;
  float scale0;
  const float* a1;         /* ROWS2: input columns [k0, k0 + k1) (k0 a multiple of 64)                   */
  long ld1;
  int k1;
  const float* pos;        /* EDGE_FEATURES: (pos[src] - pos[tgt]) / radius and its norm, dim + 1 columns */
  long pos_stride;
  int dim;
  float radius;
  const int* src;          /* [rows] sender / receiver of each row (EDGE_FEATURES, add_p / add_q)         */
  const int* tgt;
  float* feat_out;         /* EDGE_FEATURES: [max_rows, dim + 1] copy of the features (weight gradient)   */
  const float* add_p;      /* optional: first Linear's output += add_p[tgt[r]] + add_q[src[r]]            */
  const float* add_q;      /* (the x_i / x_j column blocks of the edge MLP, graph_network.py:173,197)     */
  long ld_add;
  int kpad;                /* input width rounded up to 32                                                */
  int hidden;              /* output width of the first Linear (multiple of 16, <= 128)                  */
  int out;                 /* output width of the second Linear (<= 128)                                  */
  int n_linear;
  int has_ln;
  int split;
  const float* w1_hi;      /* image N = hidden, K = kpad                                                  */
  const float* w1_lo;
  const float* w2_hi;      /* image N = out rounded up to 16, K = hidden                                  */
  const float* w2_lo;
  const float* b1;         /* bias of the first Linear, added to columns [0, b1_cols)                     */
  int b1_cols;
  const float* b2;
  const float* gamma;
  const float* beta;
  const float* res;        /* optional residual added to the output: row r at res + r * ld_res            */
  long ld_res;
  float* h1;               /* out [rows, hidden]: ReLU output of the first Linear (n_linear 2)            */
  float* y2;               /* out [rows, out]: input of the LayerNorm (has_ln)                            */
  float* mean;             /* out [rows]                                                                  */
  float* rstd;             /* out [rows]                                                                  */
  float* y;                /* out: row r at y + r * ld_y; width out (n_linear 2) or hidden (n_linear 1)   */
  long ld_y;
} SgnnTrainFwd;

/* Backward of the same block from dy = gradient of the LAST Linear's output (after the LayerNorm backward,
 * sgnn_layer_norm_bwd_rows):  n_linear 2: dpre = (dy W2) * (h1 > 0), written for the weight gradients;
 * n_linear 1: dpre = dy.  Then optionally dx = dpre W1, routed by columns to two destinations. */
typedef struct {
  int max_rows;
  const int* n_rows_dev;
  const float* dy;
  long ld_dy;
  int k_dy;                /* width of dy; its images use K = k_dy rounded up to 32                       */
  int n_linear;
  int hidden;              /* n_linear 2: width of h1 / dpre                                              */
  const float* h1;
  float* dpre;             /* out [rows, hidden] (n_linear 2)                                             */
  const float* w2t_hi;     /* image of W2^T: N = hidden, K = pad32(k_dy)                                  */
  const float* w2t_lo;
  int want_dx;
  int kx;                  /* width of dx = k0 + k1, multiple of 16                                       */
  const float* w1t_hi;     /* image of W1^T: N = kx, K = hidden (n_linear 2) or pad32(k_dy) (n_linear 1)  */
  const float* w1t_lo;
  float* dx0;              /* columns [0, k0) (k0 a multiple of 32)                                       */
  long ld_dx0;
  int k0;
  int acc0;                /* 1: dx0 += ...; 0: dx0 = ...                                                 */
  float* dx1;              /* columns [k0, k0 + k1)                                                       */
  long ld_dx1;
  int k1;
  int acc1;
  const float* add1;       /* optional [rows, k1] added into dx1 (gradient arriving over the residual)    */
  long ld_add1;
  int split;
} SgnnTrainBwd;

/* One block of a weight image of the training step: image element (row0 + r, col0 + c), r < n, c < k, of an
 * image with n_pad rows gets scale * src[r * stride_n + c * stride_k] (stride swap = transposed image; several
 * blocks may fill one image, e.g. [Wa; Wb]).  Images are zeroed once at set-up: padding is never written. */
typedef struct {
  const float* src;
  long stride_n, stride_k;
  int n, k, row0, col0, n_pad, pad_;
  float scale;
  int pad2_;
  long dst_off;            /* float offset of the image in the hi and the lo image buffers                */
} SgnnPackDesc;

/* One 64 x 64 weight gradient of a batch (csrc/train_wgrad_tc.cu): dw[0:64, col0 .. col0+64) of a row-major
 * gradient with row stride ld_dw = scale * dy[:, 0:64]^T x[:, 0:64], db[0:64] = column sums of dy (NULL: none).
 * dy / x are column blocks of row-major fp32 matrices (16-byte aligned, leading dimensions multiples of 4).   */
typedef struct {
  const float* dy;
  const float* x;
  float* dw;
  float* db;
  long ld_dy, ld_x, ld_dw;
  float scale;
  int pad_;
} SgnnWgradItem;
#define SGNN_WGRAD_MAX_ITEMS 6
typedef struct {
  SgnnWgradItem item[SGNN_WGRAD_MAX_ITEMS];
  int n_items;
  int split;               /* 1: hi / lo-split TF32 operands (fp32-grade products), 0: plain TF32             */
  long max_rows;           /* rows of every item: min(max_rows, *n_rows_dev) (n_rows_dev NULL: max_rows)      */
  const int* n_rows_dev;
  void* workspace;
  size_t workspace_bytes;
} SgnnWgradBatch;

#ifdef __cplusplus
}
#endif
