/*
 * wdmpnn.h -- C ABI of the B200 (sm_100a) wD-MPNN encoder hot path.
 *
 * Drop-in boundary for chemprop/models/mpn.py (MPNEncoder.forward :66-173 and its autograd
 * backward).  The reference is pure Python/PyTorch and has no FFI of its own; these entry
 * points are what a binding for that path binds (ctypes stub: INTEGRATION.md, shipped binding:
 * polymer_chemprop_b200/_capi.py).  Plain pointers and sizes only -- no torch / C++ types.
 *
 * Ownership: the caller owns every buffer (inputs, outputs, gradients, workspace).  The library
 * allocates nothing persistent and keeps no state except a thread-local error string and a
 * launch counter.  All pointers are DEVICE pointers unless stated otherwise.  All kernels are
 * enqueued on `stream` (a cudaStream_t passed as void*); no call synchronises except
 * wd_csr_build_check().  Return value: 0 on success, negative WD_ERR_* otherwise, message via
 * wd_last_error().
 */
#ifndef WDMPNN_H_
#define WDMPNN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WD_OK 0
#define WD_ERR_INVALID_ARG (-1)
#define WD_ERR_CUDA (-2)
#define WD_ERR_UNSUPPORTED (-3)
#define WD_ERR_GRAPH (-4) /* BatchMolGraph invariants violated (see wd_csr_build_check) */

/* activation: chemprop/nn_utils.py:70-99 (PReLU has a learned slope and is not supported) */
enum { WD_ACT_RELU = 0, WD_ACT_LEAKY_RELU = 1, WD_ACT_TANH = 2, WD_ACT_SELU = 3, WD_ACT_ELU = 4 };
/* aggregation: chemprop/models/mpn.py:158-163 */
enum { WD_AGG_MEAN = 0, WD_AGG_SUM = 1, WD_AGG_NORM = 2 };
/* math_mode: arithmetic of the h x h / F x h contractions.
 *   FP32   : CUDA-core FFMA, fp32 accumulate                 (encoder output <= 1e-5 relative; the slow parity path)
 *   TF32   : tcgen05.mma kind::tf32, one pass, fp32 accumulate in TMEM
 *                                                            (encoder output <= 1e-3, predictions ~5e-4)
 *   TF32X3 : tcgen05.mma kind::tf32 with error compensation in the FORWARD contractions: both operands are split
 *            into tf32 hi + lo parts and a_hi w_hi + a_lo w_hi + a_hi w_lo accumulate in the same TMEM tile
 *            (fp32-grade forward: encoder output <= 1e-5, predictions <= 1e-4, the north-star bound); the backward
 *            contractions stay single-pass TF32.  The default of the Python drop-in.
 *   BF16X3 : the same scheme with the split point at bfloat16 and every product a tcgen05.mma kind::f16: a_hi w_hi +
 *            a_lo w_hi + a_hi w_lo over ONE bf16 image per operand (no tf32 image): 6 instead of 8 MMAs per K chunk,
 *            half the weight bytes on the SM; forward error ~4x that of TF32X3 (encoder output <= 2e-5, predictions
 *            <= 1e-4); backward contractions single-pass TF32 as in TF32X3. */
enum { WD_MATH_FP32 = 0, WD_MATH_TF32 = 1, WD_MATH_TF32X3 = 2, WD_MATH_BF16X3 = 3 };
/* WdConfig.flags */
#define WD_FLAG_NO_SIGN_BITS 1u /* tensor-core modes, ReLU: the backward reads the stored activations instead of the  */
                                /* 1-bit-per-element sign images the forward epilogues leave (A/B and test switch).   */
                                /* Part of the config, so the forward and the backward call of a step always agree.   */
#define WD_FLAG_NO_GRAPH 2u     /* never replay the call's launches through a CUDA graph (see wd_encoder_forward)     */
#define WD_FLAG_NO_SIDE_BRANCH 4u /* captured backward pass: keep the weight-gradient kernels in the main chain instead */
                                /* of on a parallel branch of the graph (A/B and test switch; results are identical)  */
#define WD_FLAG_NO_DEFER_SUM 8u /* backward, depth 3 without a W_h bias: materialise the running sum dI in the middle step */
                                /* as every other configuration does (A/B and test switch; results are identical)     */

/* Encoder configuration == the TrainArgs fields MPNEncoder.__init__ reads (mpn.py:24-35). */
typedef struct WdConfig {
  int32_t hidden_size;     /* args.hidden_size                                   */
  int32_t atom_fdim;       /* F_a (133 by default)                               */
  int32_t bond_fdim;       /* F_b = F_a + 14 (147 by default)                    */
  int32_t depth;           /* args.depth  (>= 1; depth-1 message-passing steps)  */
  int32_t activation;      /* WD_ACT_*                                           */
  int32_t aggregation;     /* WD_AGG_*                                           */
  int32_t undirected;      /* args.undirected                                    */
  int32_t bias;            /* args.bias: W_i / W_h carry a bias                  */
  int32_t math_mode;       /* WD_MATH_*                                          */
  float aggregation_norm;  /* args.aggregation_norm                              */
  float dropout;           /* args.dropout p in [0,1) for THIS call: 0 in eval mode (nn.Dropout, mpn.py:38).  */
  uint32_t dropout_seed_lo; /* Philox key of this call's masks; pass the SAME cfg to the backward call.        */
  uint32_t dropout_seed_hi; /* The reference's RNG stream cannot be matched: masks are Philox4x32-10 words of  */
                            /* (element index, site) -- restated in oracle/mpn_oracle.py for the parity tests.  */
  uint32_t flags;           /* WD_FLAG_* bits; 0 = defaults                                                     */
} WdConfig;

/* One BatchMolGraph on the device: get_components() tensors (featurization.py:815-846) plus the
 * int32 CSR arrays produced by wd_csr_build.  Row 0 of every per-atom / per-bond array is the
 * reference's padding row. */
typedef struct WdGraph {
  int64_t n_atoms;        /* N_a incl. padding row                  */
  int64_t n_bonds;        /* N_b incl. padding row                  */
  int64_t n_mols;         /* G = len(a_scope)                       */
  const float* f_atoms;   /* [N_a, F_a] row-major                   */
  const float* f_bonds;   /* [N_b, F_b] row-major                   */
  const float* w_atoms;   /* [N_a]                                  */
  const float* w_bonds;   /* [N_b]                                  */
  const int32_t* in_ptr;  /* [N_a+1] CSR of incoming bonds per atom (== out_ptr, see below) */
  const int32_t* in_idx;  /* [N_b-1] non-zero prefix of each a2b row, reference order        */
  const int32_t* out_idx; /* [N_b-1] bonds grouped by source atom b2a[b], ascending bond id  */
  const int32_t* src;     /* [N_b] int32 copy of b2a                */
  const int32_t* rev;     /* [N_b] int32 copy of b2revb             */
  const int32_t* dst;     /* [N_b] dst[b] = atom whose a2b row holds b (== b2a[b2revb[b]])   */
  const int32_t* mol_of;  /* [N_a] molecule index of each atom, -1 if in no a_scope          */
  const int32_t* a_start; /* [G] a_scope starts                     */
  const int32_t* a_size;  /* [G] a_scope sizes                      */
  const float* dop;       /* [G] degree_of_polym                    */
  int32_t ld_atoms;       /* row stride of f_atoms in floats; 0 = F_a (dense, the reference layout) */
  int32_t ld_bonds;       /* row stride of f_bonds in floats; 0 = F_b.  Strides that are multiples   */
                          /* of 4 put the feature rows on the 16-byte asynchronous-copy path         */
  int32_t rev_adjacent;   /* 1 iff wd_csr_build reported flags[2] == 0: b2revb[b] == ((b-1)^1)+1 for every real bond,  */
                          /* i.e. reverse bonds are adjacent pairs (how featurization.py:470-480 / :616-629 number     */
                          /* them).  The gathered depth step then fetches message[b2revb] as plain TMA boxes.  0 is    */
                          /* always safe (gathered copies).                                                           */
} WdGraph;

/* nn.Linear layout [out, in] row-major, exactly the reference's parameters (mpn.py:47-57).
 * b_i / b_h may be NULL when cfg.bias == 0. */
typedef struct WdWeights {
  const float* W_i; /* [h, F_b]      */
  const float* b_i; /* [h] or NULL   */
  const float* W_h; /* [h, h]        */
  const float* b_h; /* [h] or NULL   */
  const float* W_o; /* [h, F_a + h]  */
  const float* b_o; /* [h]           */
} WdWeights;

/* Gradients, same shapes as WdWeights; each buffer is OVERWRITTEN (not accumulated). */
typedef struct WdGrads {
  float* dW_i;
  float* db_i;
  float* dW_h;
  float* db_h;
  float* dW_o;
  float* db_o;
} WdGrads;

/* Rebuilds BatchMolGraph's padded a2b [N_a, K] int64 (featurization.py:805) into CSR on the
 * device and narrows b2a / b2revb to int32.  Because every out-edge of an atom is the reverse
 * of one of its in-edges, out-degree == in-degree and `in_ptr` also delimits `out_idx`.
 * a_start / a_size are DEVICE int32 [G]; mol_of [N_a] is filled from them.
 * `flags` is a DEVICE int32[4] written by the kernels: [0] number of invariant violations
 * (bond id out of range, bond listed in 0 or >1 rows, dst != b2a[b2revb], b2revb not an
 * involution, scope out of range), [1] max in-degree seen, [2] number of real bonds whose reverse is NOT the
 * adjacent one (b2revb[b] != ((b-1)^1)+1; 0 for every graph the reference builds), [3] reserved. */
int wd_csr_build(const int64_t* a2b, int64_t n_atoms, int64_t max_num_bonds, const int64_t* b2a,
                 const int64_t* b2revb, int64_t n_bonds, const int32_t* a_start,
                 const int32_t* a_size, int64_t n_mols, int32_t* in_ptr, int32_t* in_idx,
                 int32_t* out_idx, int32_t* src, int32_t* rev, int32_t* dst, int32_t* mol_of,
                 int32_t* flags, void* stream);

/* Synchronises `stream`, reads `flags` back and returns WD_ERR_GRAPH if flags[0] != 0. */
int wd_csr_build_check(const int32_t* flags, void* stream);

/* Rebuilds the get_components() feature matrices on the device from their compact form.  MolGraph
 * builds every bond row as f_atoms[a1] + f_bond (featurization.py:467-468; polymer edges :616-629), so
 * f_bonds[b, :F_a] == f_atoms[b2a[b]] and only the bond-only tail carries information:
 *   f_bonds_out[b, :] = [ f_atoms[b2a[b], 0:atom_fdim], bond_tail[b, 0:tail_fdim], 0-pad to ld_bonds ]
 *   f_atoms_out[a, :] = [ f_atoms[a, 0:atom_fdim], 0-pad to ld_atoms ]        (skipped if f_atoms_out == NULL)
 * f_atoms [N_a, atom_fdim] and bond_tail [N_b, tail_fdim] are DENSE row-major device arrays, b2a the
 * int64 device copy of BatchMolGraph.b2a; ld_* are multiples of 4 (WdGraph.ld_atoms / ld_bonds).
 * Byte copies only: bit-identical to uploading f_bonds itself.  The caller is responsible for the
 * prefix property (polymer_chemprop_b200.BatchMolGraph verifies it exactly per molecule). */
int wd_expand_features(const float* f_atoms, const float* bond_tail, const int64_t* b2a, int64_t n_atoms,
                       int64_t n_bonds, int32_t atom_fdim, int32_t tail_fdim, float* f_atoms_out,
                       int32_t ld_atoms, float* f_bonds_out, int32_t ld_bonds, void* stream);

/* A device-resident store of featurised graphs (the reference keeps them as host MolGraph objects cached in
 * SMILES_TO_GRAPH, chemprop/data/data.py:30-40, :277-314) in compact, per-graph LOCAL numbering (atoms and bonds of
 * a graph start at 0, as in MolGraph): all arrays are DEVICE pointers, graphs concatenated in store order. */
typedef struct WdStore {
  int64_t n_graphs;
  int32_t atom_fdim;           /* F_a                                                          */
  int32_t tail_fdim;           /* F_b - F_a: width of the bond-only part of a bond row          */
  const float* f_atoms;        /* [sum n_atoms, F_a] dense                                      */
  const float* bond_tail;      /* [sum n_bonds, F_b - F_a]: f_bonds[b] = [f_atoms[b2a[b]], bond_tail[b]] (featurization.py:467-468) */
  const float* w_atoms;        /* [sum n_atoms]                                                 */
  const float* w_bonds;        /* [sum n_bonds]                                                 */
  const int32_t* b2a;          /* [sum n_bonds] local source atom                               */
  const int32_t* b2revb;       /* [sum n_bonds] local reverse bond                              */
  const int32_t* in_idx;       /* [sum n_bonds] local ids of incoming bonds, a2b rows concatenated in atom order */
  const int32_t* out_idx;      /* [sum n_bonds] local ids of outgoing bonds (ascending), concatenated in atom order */
  const int32_t* in_ptr_local; /* [sum n_atoms] exclusive prefix of the in-degrees within the graph */
  const int64_t* atom_off;     /* [n_graphs + 1] first atom row of each graph in the store       */
  const int64_t* bond_off;     /* [n_graphs + 1] first bond row of each graph in the store       */
  const float* dop;            /* [n_graphs] degree_of_polym                                    */
} WdStore;

/* Assembles the batch made of graphs sel[0..n_mols) (DEVICE int64 store indices, repeats allowed) directly into the
 * arrays of `out` -- what BatchMolGraph.__init__ (featurization.py:757-813) + wd_expand_features + wd_csr_build
 * produce for the same graphs in the same order, bit for bit: row 0 padding, per-molecule offsets, padded feature
 * rows (out->ld_atoms / ld_bonds), int32 CSR / reverse-CSR, src / rev / dst, mol_of, dop.
 * a_start / a_size / b_start / b_size: DEVICE int32 [n_mols], the batch's scopes (a_scope / b_scope: starts are
 * 1 + exclusive prefix sums of the sizes; the caller computes them on the host, where it also needs the totals to
 * size the buffers); a_start / a_size are also what out->a_start / a_size point to.  bond_mol: DEVICE int32
 * [out->n_bonds] scratch.  Every array of `out` is caller-allocated and OVERWRITTEN (the const in WdGraph's
 * declarations is cast away). */
int wd_store_assemble(const WdStore* store, const int64_t* sel, const int32_t* a_start, const int32_t* a_size,
                      const int32_t* b_start, const int32_t* b_size, int64_t n_mols, const WdGraph* out,
                      int32_t* bond_mol, void* stream);

/* Bytes of caller-provided workspace.  The forward pass stores there what the backward pass
 * re-reads (I, M_t, A_t, H) plus scratch; pass the same buffer, untouched, to the backward. */
size_t wd_workspace_bytes(const WdConfig* cfg, int64_t n_atoms, int64_t n_bonds, int64_t n_mols);

/* out[G, h] = MPNEncoder.forward(graph)  (mpn.py:92-171, bond messages; dropout per cfg->dropout). */
int wd_encoder_forward(const WdGraph* graph, const WdWeights* weights, const WdConfig* cfg,
                       void* workspace, size_t workspace_bytes, float* out, void* stream);

/* Forward only (chemprop/train/predict.py:10-68, molecule_fingerprint.py:151-183 run the encoder under
 * torch.no_grad()): same kernels and results as wd_encoder_forward, but nothing is kept for a backward pass --
 * two alternating message buffers, one aggregate buffer, no S_t copy, no gradient scratch (~3.3 instead of ~10
 * [N_b, h] buffers at depth 3).  cfg->dropout is honoured as given (0 for model.eval()). */
size_t wd_infer_workspace_bytes(const WdConfig* cfg, int64_t n_atoms, int64_t n_bonds, int64_t n_mols);
int wd_encoder_infer(const WdGraph* graph, const WdWeights* weights, const WdConfig* cfg, void* workspace,
                     size_t workspace_bytes, float* out, void* stream);

/* Gradients of sum(out * d_out) w.r.t. the weights; deterministic (no atomics). */
int wd_encoder_backward(const WdGraph* graph, const WdWeights* weights, const WdConfig* cfg,
                        void* workspace, size_t workspace_bytes, const float* d_out,
                        const WdGrads* grads, void* stream);

/* HOST-buffer entry point (the reference's own calling convention: get_components() hands CPU
 * tensors to MPNEncoder.forward, mpn.py:81-87).  All graph pointers here are HOST pointers;
 * a_scope is [G,2] int64 (start,size) pairs, dop double[G].  Uploads, builds the CSR, runs
 * forward (+ backward when d_out != NULL) and copies out[G,h] (and gradients) back to the host.
 * Device memory is allocated per call with cudaMallocAsync on `stream`. */
int wd_encoder_host(const float* f_atoms, const float* f_bonds, const float* w_atoms,
                    const float* w_bonds, const int64_t* a2b, const int64_t* b2a,
                    const int64_t* b2revb, const int64_t* a_scope, const double* dop,
                    int64_t n_atoms, int64_t n_bonds, int64_t max_num_bonds, int64_t n_mols,
                    const WdWeights* host_weights, const WdConfig* cfg, float* out,
                    const float* d_out, const WdGrads* host_grads, void* stream);

/* Per-category device timing of this library's own launches, measured with CUDA events recorded
 * on the launching stream around each kernel (for bench.py's roofline line).  Off by default. */
enum {
  WD_PROF_INPUT_PROJ = 0,  /* I = W_i(f_bonds)                                   mpn.py:96      */
  WD_PROF_AGGREGATE = 1,   /* CSR gather a_message                               mpn.py:112-118 */
  WD_PROF_DEPTH_STEP = 2,  /* act(input + W_h(a_message[b2a] - message[b2revb]))  mpn.py:119-123 */
  WD_PROF_ATOM_UPDATE = 3, /* act(W_o([f_atoms, a_message]))                     mpn.py:132-133 */
  WD_PROF_READOUT = 4,     /* weighted per-molecule readout fwd+bwd              mpn.py:146-171 */
  WD_PROF_MSG_GRAD = 5,    /* backward message gradient + activation mask                        */
  WD_PROF_DW = 6,          /* weight-gradient contractions dW_h / dW_i / dW_o / biases           */
  WD_PROF_DX = 7,          /* dS = dZ W_h and dA_d = dZo W_o[:, F_a:]                            */
  WD_PROF_SCATTER = 8,     /* transpose-CSR scatter of dS onto source atoms                      */
  WD_PROF_MISC = 9,        /* CSR build, symmetrisation, weight packing, everything else         */
  WD_PROF_REDUCE = 10,     /* fixed-order reduction of split-K partials, bias-gradient column sums */
  WD_PROF_NUM = 11
};
/* ---- fused optimiser step ---------------------------------------------------------------
 * The reference updates the model with torch.optim.Adam / AdamW over model.parameters() (chemprop/utils.py:295-310,
 * build_optimizer: one parameter group, lr = init_lr, weight_decay), after an optional nn.utils.clip_grad_norm_
 * (chemprop/train/train.py:80-82); the learning rate is set per step by NoamLR (chemprop/nn_utils.py:115-194).
 * Here the parameters live in ONE flat fp32 buffer (the gradients already do: the data-parallel all-reduce packs them)
 * and one call does clip + Adam for all of them.  Same arithmetic as torch's single-tensor Adam (not amsgrad, not
 * maximize):
 *   scale = max_grad_norm > 0 ? min(1, max_grad_norm / (||grads||_2 + 1e-6)) : 1        (clip_grad_norm_)
 *   g = scale * grad;  decoupled == 0: g += weight_decay * p   |   decoupled != 0: p *= 1 - lr * weight_decay
 *   m += (1 - beta1) * (g - m);  v = beta2 * v + (1 - beta2) * g * g
 *   p -= (lr / (1 - beta1^step)) * m / (sqrt(v) / sqrt(1 - beta2^step) + eps)
 * The squared norm is reduced in a fixed order (deterministic).  grad_norm_out (DEVICE float, may be NULL) receives
 * ||grads||_2 when max_grad_norm > 0. */
typedef struct WdAdam {
  double lr, beta1, beta2, eps, weight_decay;   /* doubles, as torch holds them: 1 - beta, lr / (1 - beta1^step) and
                                                   1 / sqrt(1 - beta2^step) are formed in double, then rounded to fp32 */
  double max_grad_norm;        /* <= 0: no clipping                                                            */
  int64_t step;                /* 1-based count of this update                                                 */
  int32_t decoupled;           /* 0: Adam (L2 added to the gradient), 1: AdamW (decoupled decay)              */
} WdAdam;
/* scratch: DEVICE floats, >= wd_adam_scratch_floats() (norm partials); may be NULL when max_grad_norm <= 0. */
int64_t wd_adam_scratch_floats(void);
int wd_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, const WdAdam* cfg,
                 float* scratch, float* grad_norm_out, void* stream);

/* wd_adam_step for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * Same arithmetic and arguments as wd_adam_step minus `scratch`: the clipping scratch is a file-scope static device
 * array inside the library (2 KB per GPU that loads the library, held until the process exits).  The call allocates
 * nothing (no cudaMalloc / cudaFree / allocating library routine); the caller still owns params, grads and the moments.
 * Limit: n <= WD_ADAM_STATIC_MAX_PARAMS (wd_adam_static_max_params()), larger n is rejected with WD_ERR_INVALID_ARG.
 * All calls on one device SHARE the array, which keeps the last call's contents (every slot is rewritten before it is
 * read, so nothing needs clearing): calls that overlap on the GPU race -- issue them on one stream or order them.
 * The first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device current);
 * make that first call outside stream capture. */
#define WD_ADAM_STATIC_MAX_PARAMS (1LL << 28) /* 268 M fp32 parameters (1 GiB per flat buffer) */
int64_t wd_adam_static_max_params(void);
int wd_adam_step_static(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, const WdAdam* cfg,
                        float* grad_norm_out, void* stream);

/* ---- FFN head + masked MSE loss on the encoder output ------------------------------------------
 * MoleculeModel.ffn (chemprop/models/model.py:57-116: Sequential[Dropout, Linear, (act, Dropout, Linear)*], forward
 * :174-194) and the regression loss of chemprop/train/train.py:46-74,
 *   loss = sum((preds - targets)^2 * weights * mask) / sum(mask),     weights = target_weights * data_weights,
 * as two calls over caller-owned buffers (fp32 CUDA-core contractions, deterministic).  Dropout must be off.
 * nn.Linear layout: W[l] is [out_l, in_l] row-major, b[l] is [out_l]; layer 0 reads in_dim columns, the last layer writes
 * out_dim (= num_tasks), every layer in between is hidden_dim wide. */
#define WD_HEAD_MAX_LAYERS 8
typedef struct WdHead {
  int32_t n_layers;    /* args.ffn_num_layers                                   */
  int32_t activation;  /* WD_ACT_* (args.activation)                            */
  int32_t in_dim;      /* first_linear_dim (model.py:68-76)                     */
  int32_t hidden_dim;  /* args.ffn_hidden_size                                  */
  int32_t out_dim;     /* output_size = num_tasks                               */
  const float* W[WD_HEAD_MAX_LAYERS];
  const float* b[WD_HEAD_MAX_LAYERS];
} WdHead;
typedef struct WdHeadGrads { /* same shapes as the weights; OVERWRITTEN */
  float* dW[WD_HEAD_MAX_LAYERS];
  float* db[WD_HEAD_MAX_LAYERS];
} WdHeadGrads;
size_t wd_head_workspace_bytes(const WdHead* head, int64_t n_rows);
/* preds[n_rows, out_dim] = ffn(X[n_rows, ldx]).  With targets != NULL (training): also loss[0] (DEVICE float) and, kept
 * in the workspace for wd_head_backward, the layer outputs and d loss / d preds; weights may be NULL (all ones);
 * targets / weights / mask are dense [n_rows, out_dim] floats (mask: 1 where a target is present). */
int wd_head_forward(const WdHead* head, const float* X, int32_t ldx, int64_t n_rows, const float* targets,
                    const float* weights, const float* mask, void* workspace, size_t workspace_bytes, float* preds,
                    float* loss, void* stream);
/* Gradients of (dloss[0] * loss + sum(dpreds * preds)) w.r.t. every W[l], b[l] and X (dX[n_rows, lddx], may be NULL;
 * lddx >= in_dim, a multiple of 4, base 16-byte aligned: rows are written as float4, pad columns as zeros);
 * dloss (DEVICE float) NULL means 1, dpreds NULL means 0.  Same workspace as the forward call, untouched in between. */
int wd_head_backward(const WdHead* head, const float* X, int32_t ldx, int64_t n_rows, void* workspace, size_t workspace_bytes,
                     const float* dloss, const float* dpreds, float* dX, int32_t lddx, const WdHeadGrads* grads, void* stream);

void wd_profile_enable(int on); /* on != 0: record events around every launch from now on      */
void wd_profile_reset(void);    /* forget recorded launches (keeps the event pool)              */
/* Synchronises the recorded events and returns total milliseconds and launch count of category. */
int wd_profile_read(int category, double* total_ms, int64_t* launches);

/* Test hooks: the raw contractions behind the encoder on plain row-major operands, so that the
 * tensor-core kernels can be checked in isolation (tests/test_gpu_tf32.py).
 * wd_debug_gemm: C[M, ldc] = epi(A[M,K] * B^T + bias (+ R)); B(n,k) = b_kn ? W[k*ldb+n] : W[n*ldb+k];
 *   epi 0 = store, 1 = ReLU, 2 = ReLU(R + .).  Columns [N, ldc) of C are zero-filled.
 *   TF32 mode needs `scratch` >= 4 * round_up(N,304-ish tiles) * round_up(K,32) bytes (packed weights).
 * wd_debug_dw: out[N1, N2] = X[R, :N1]^T * Y[R, :N2]; `scratch` >= 160 * N1 * N2 * 4 bytes (partials). */
int wd_debug_gemm(const float* A, int lda, const float* W, int ldb, int b_kn, const float* bias,
                  const float* R, int ldr, float* C, int ldc, int64_t M, int N, int K, int epi,
                  int math_mode, void* scratch, size_t scratch_bytes, void* stream);
int wd_debug_dw(const float* X, int ldx, const float* Y, int ldy, float* out, int64_t R, int N1, int N2,
                int math_mode, void* scratch, size_t scratch_bytes, void* stream);
/* Wait-cycle counters of CTA 0 of the tcgen05 GEMM kernel (zero unless built with -DWD_T2_TIMING). */
int wd_debug_timing(uint64_t* out16, int reset);

/* Thread-local message of the last failing call on this thread. */
const char* wd_last_error(void);
/* Kernel launches issued by this library since load (for bench.py's gpu_launches). */
int64_t wd_launch_count(void);
/* Small batches (n_bonds <= 65536, env WD_GRAPH_MAX_BONDS) run each forward / infer / backward call as ONE CUDA graph
 * launch: the call's kernels are stream-captured and an executable graph cached per (device, pass, configuration) is
 * re-parameterised by cudaGraphExecUpdate, see csrc/encoder.cu GraphRun.  launches: graph launches so far; builds: how
 * many of them had to instantiate a new executable graph (the rest were parameter updates). */
void wd_graph_stats(int64_t* launches, int64_t* builds);
/* ABI version of this header. */
int wd_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* WDMPNN_H_ */
