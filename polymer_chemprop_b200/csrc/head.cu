// FFN head + masked MSE loss on the encoder output.
//
// The reference's MoleculeModel runs nn.Sequential[Dropout, Linear, (act, Dropout, Linear)*] on the [G, h] encodings
// (chemprop/models/model.py:57-116, forward :174-194) and the training loop forms
//   loss = (MSELoss(reduction='none')(preds, targets) * target_weights * data_weights * mask).sum() / mask.sum()
// (chemprop/train/train.py:46-74), i.e. ~25 small torch kernels forward + backward for G x 300 x 300 flops.  Here the
// head is two C calls over caller-owned buffers -- forward: one contraction per layer + one loss launch; backward: per
// layer a weight-gradient contraction, a bias column sum and the input-gradient contraction, one batched fixed-order
// reduction at the end -- on the fp32 CUDA-core kernels of simt_gemm.cuh (<= 1e-6 of the reference: predictions are
// what BASELINE.json's 1e-4 bound is stated on).  Deterministic, no atomics.  Dropout must be off (p = 0 / eval()).
#include <algorithm>

#include "common.cuh"
#include "operands.cuh"
#include "simt_gemm.cuh"

namespace wd {
namespace {

constexpr int kHeadColsumChunks = 64;

inline int64_t up64(int64_t n) { return round_up(std::max<int64_t>(n, 1), 64); }

struct HeadLayout {      // workspace, in floats
  int64_t Z[WD_HEAD_MAX_LAYERS];     // Z[l] = output of layer l (post-activation), l = 0 .. L-2
  int64_t dY[2];                     // ping-pong gradient buffers [G, max_dim]
  int64_t dP, msum;                  // d loss / d preds [G, T] (dense); sum of the mask
  int64_t Pp;                        // predictions with a padded row stride [G, ld4(T)] (the last contraction writes float4)
  int64_t PW[WD_HEAD_MAX_LAYERS], Pb[WD_HEAD_MAX_LAYERS];
  int zw[WD_HEAD_MAX_LAYERS], zb;
  int64_t total;
};

inline int ld4(int n) { return (n + 3) & ~3; }     // k_simt_gemm stores float4: every row stride it writes is a multiple of 4
inline int head_in(const WdHead& h, int l) { return l == 0 ? h.in_dim : h.hidden_dim; }
inline int head_out(const WdHead& h, int l) { return l == h.n_layers - 1 ? h.out_dim : h.hidden_dim; }

int head_dw_chunks(int64_t rows, int n1, int n2) {
  const int64_t tiles = ceil_div(n1, DW_BM) * ceil_div(n2, DW_BN);
  int64_t z = ceil_div(2 * kNumSMs, tiles);
  z = std::min<int64_t>(z, std::max<int64_t>(1, rows / 64));
  return (int)std::max<int64_t>(1, std::min<int64_t>(z, 64));
}

HeadLayout head_layout(const WdHead& h, int64_t G) {
  HeadLayout L{};
  int64_t off = 0;
  auto take = [&](int64_t n) { int64_t o = off; off += up64(n); return o; };
  const int maxd = std::max(std::max(h.in_dim, h.hidden_dim), h.out_dim);
  for (int l = 0; l + 1 < h.n_layers; ++l) L.Z[l] = take(G * ld4(h.hidden_dim));
  L.dY[0] = take(G * ld4(maxd));
  L.dY[1] = take(G * ld4(maxd));
  L.dP = take(G * h.out_dim);
  L.Pp = take(G * ld4(h.out_dim));
  L.msum = take(4);
  L.zb = (int)std::max<int64_t>(1, std::min<int64_t>(kHeadColsumChunks, G / 64));
  for (int l = 0; l < h.n_layers; ++l) {
    L.zw[l] = head_dw_chunks(G, head_out(h, l), head_in(h, l));
    L.PW[l] = take((int64_t)L.zw[l] * head_out(h, l) * head_in(h, l));
    L.Pb[l] = take((int64_t)L.zb * head_out(h, l));
  }
  L.total = off;
  return L;
}

int head_check(const WdHead* h) {
  if (!h) return fail(WD_ERR_INVALID_ARG, "head is NULL");
  if (h->n_layers < 1 || h->n_layers > WD_HEAD_MAX_LAYERS)
    return fail(WD_ERR_UNSUPPORTED, "ffn_num_layers %d not in [1, %d]", h->n_layers, WD_HEAD_MAX_LAYERS);
  if (h->in_dim < 1 || h->out_dim < 1 || (h->n_layers > 1 && h->hidden_dim < 1))
    return fail(WD_ERR_INVALID_ARG, "head dimensions must be >= 1");
  if (h->activation < WD_ACT_RELU || h->activation > WD_ACT_ELU)
    return fail(WD_ERR_UNSUPPORTED, "activation %d not supported (ReLU, LeakyReLU, tanh, SELU, ELU)", h->activation);
  for (int l = 0; l < h->n_layers; ++l)
    if (!h->W[l] || !h->b[l]) return fail(WD_ERR_INVALID_ARG, "head layer %d has NULL weights", l);
  return WD_OK;
}

// One block: msum = sum(mask); loss = sum((p - t)^2 w m) / msum; dP = 2 (p - t) w m / msum   (train.py:70-74).
// Fixed-order tree reductions: bitwise repeatable.  w may be NULL (all ones).
// Pp: predictions with row stride ldp (what the last contraction wrote); preds: the caller's dense [G, T] copy.
// T == nullptr (inference): only the copy.
__global__ void __launch_bounds__(1024) k_head_loss(const float* __restrict__ Pp, int ldp, int Tn, float* __restrict__ preds,
                                                   const float* __restrict__ T, const float* __restrict__ Wt,
                                                   const float* __restrict__ Mk, int64_t n, float* __restrict__ loss,
                                                   float* __restrict__ dP, float* __restrict__ msum_out) {
  pdl_prologue();
  __shared__ float sm[1024];
  __shared__ float s_msum;
  for (int64_t i = threadIdx.x; i < n; i += 1024) preds[i] = Pp[(i / Tn) * ldp + (i % Tn)];
  if (T == nullptr) return;
  float acc = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += 1024) acc += Mk[i];
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) { if ((int)threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) { s_msum = sm[0]; msum_out[0] = sm[0]; }
  __syncthreads();
  const float inv = 1.f / s_msum;
  acc = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += 1024) {
    const float d = Pp[(i / Tn) * ldp + (i % Tn)] - T[i], wm = (Wt ? Wt[i] : 1.f) * Mk[i];
    acc = fmaf(d * d, wm, acc);
    dP[i] = 2.f * d * wm * inv;
  }
  __syncthreads();
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) { if ((int)threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) loss[0] = sm[0] * inv;
}

// dY[i] = dY[i] * act'(Z[i]) (derivative through the stored activation output); optional scale by a device scalar
template <int ACT>
__global__ void k_head_act_grad(float* __restrict__ dY, const float* __restrict__ Z, int64_t n) {
  pdl_prologue();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dY[i] *= act_bwd_from_out<ACT>(Z[i]);
}

// dP[i] = dP[i] * dloss[0] + (dpreds ? dpreds[i] : 0): the incoming gradients of (loss, preds)
__global__ void k_head_seed(float* __restrict__ dP, const float* __restrict__ dloss, const float* __restrict__ dpreds, int64_t n) {
  pdl_prologue();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dP[i] = dP[i] * (dloss ? dloss[0] : 1.f) + (dpreds ? dpreds[i] : 0.f);
}

#define HEAD_TRY(expr) do { int _rc = (expr); if (_rc != WD_OK) return _rc; } while (0)

template <int ACT>
int head_forward_impl(const WdHead& h, const HeadLayout& L, const float* X, int ldx, int64_t G, const float* targets,
                      const float* weights, const float* mask, float* ws, float* preds, float* loss, cudaStream_t s) {
  const float* in = X;
  int ld_in = ldx;
  for (int l = 0; l < h.n_layers; ++l) {
    const bool last = l == h.n_layers - 1;
    const int K = head_in(h, l), N = head_out(h, l);
    float* out = last ? ws + L.Pp : ws + L.Z[l];
    const int ldo = ld4(N);
    AOperand a{in, nullptr, nullptr, nullptr, ld_in, 0, K};
    dim3 grid((unsigned)ceil_div(G, SG_BM), (unsigned)ceil_div(ldo, SG_BN));
    if (last) WD_LAUNCH((k_simt_gemm<A_PLAIN, B_NK, E_STORE, ACT, false>), grid, SG_THREADS, 0, s, a, h.W[l], K, h.b[l], nullptr, 0, out, ldo, G, N, K, Dropout{});
    else      WD_LAUNCH((k_simt_gemm<A_PLAIN, B_NK, E_ACT, ACT, false>), grid, SG_THREADS, 0, s, a, h.W[l], K, h.b[l], nullptr, 0, out, ldo, G, N, K, Dropout{});
    in = out;
    ld_in = ldo;
  }
  WD_LAUNCH(k_head_loss, 1, 1024, 0, s, (const float*)(ws + L.Pp), ld4(h.out_dim), h.out_dim, preds, targets, weights, mask,
            G * h.out_dim, loss, ws + L.dP, ws + L.msum);
  return WD_OK;
}

template <int ACT>
int head_backward_impl(const WdHead& h, const HeadLayout& L, const float* X, int ldx, int64_t G, float* ws, const float* dloss,
                       const float* dpreds, float* dX, int lddx, const WdHeadGrads& gr, cudaStream_t s) {
  const int64_t nP = G * h.out_dim;
  if (dloss != nullptr || dpreds != nullptr)
    WD_LAUNCH(k_head_seed, (unsigned)ceil_div(nP, 256), 256, 0, s, ws + L.dP, dloss, dpreds, nP);
  const float* dY = ws + L.dP;
  int ld_dy = h.out_dim;             // dP is dense; the internal gradient buffers have padded row strides
  ReduceJobs jobs{};
  unsigned blocks = 0;
  auto flush = [&]() -> int {
    if (jobs.count == 0) return WD_OK;
    WD_LAUNCH(k_reduce_partials<1>, blocks, RP_COLS * RP_GROUPS, 0, s, jobs);
    jobs.count = 0;
    blocks = 0;
    return WD_OK;
  };
  auto queue = [&](const float* P, float* out, int64_t n, int Z) -> int {
    if (jobs.count >= RP_MAX_JOBS) HEAD_TRY(flush());
    jobs.j[jobs.count++] = ReduceJob{P, out, n, Z, blocks};
    blocks += (unsigned)ceil_div(n, RP_COLS);
    return WD_OK;
  };
  for (int l = h.n_layers - 1; l >= 0; --l) {
    const int K = head_in(h, l), N = head_out(h, l);
    const float* Zin = l == 0 ? X : ws + L.Z[l - 1];
    const int ld_in = l == 0 ? ldx : ld4(K);
    {  // dW_l = dY^T Z_in
      const int Z = L.zw[l];
      const int64_t rpc = round_up(ceil_div(G, Z), DW_BK);
      AOperand y{Zin, nullptr, nullptr, nullptr, ld_in, 0, 0};
      dim3 grid((unsigned)ceil_div(K, DW_BN), (unsigned)ceil_div(N, DW_BM), (unsigned)Z);
      WD_LAUNCH((k_simt_dw<A_PLAIN, ACT, false>), grid, DW_THREADS, 0, s, dY, ld_dy, y, ws + L.PW[l], G, N, K, rpc);
      HEAD_TRY(queue(ws + L.PW[l], gr.dW[l], (int64_t)N * K, (int)ceil_div(G, rpc)));
    }
    {  // db_l = colsum(dY)
      const int Z = L.zb;
      const int64_t rpc = ceil_div(G, Z);
      WD_LAUNCH(k_colsum_partial, dim3((unsigned)ceil_div(N, 128), (unsigned)Z), 128, 0, s, dY, ld_dy, ws + L.Pb[l], G, N, rpc);
      HEAD_TRY(queue(ws + L.Pb[l], gr.db[l], N, (int)ceil_div(G, rpc)));
    }
    // d Z_in = dY W_l  (for l == 0: the gradient of the encodings)
    float* dst = l == 0 ? dX : ws + L.dY[l & 1];
    const int ld_dst = l == 0 ? lddx : ld4(K);
    if (dst != nullptr) {
      AOperand a{dY, nullptr, nullptr, nullptr, ld_dy, 0, N};
      dim3 grid((unsigned)ceil_div(G, SG_BM), (unsigned)ceil_div(ld_dst, SG_BN));
      WD_LAUNCH((k_simt_gemm<A_PLAIN, B_KN, E_STORE, ACT, false>), grid, SG_THREADS, 0, s, a, h.W[l], K, nullptr, nullptr, 0, dst, ld_dst, G, K, N, Dropout{});
      // same padded layout in both buffers; pad columns hold zeros (gradient) times act'(0): harmless, never read as data
      if (l > 0) WD_LAUNCH((k_head_act_grad<ACT>), (unsigned)ceil_div(G * ld_dst, 256), 256, 0, s, dst, (const float*)(ws + L.Z[l - 1]), G * (int64_t)ld_dst);
    }
    dY = dst;
    ld_dy = ld_dst;
  }
  return flush();
}

}  // namespace
}  // namespace wd

// ---- C ABI ----------------------------------------------------------------------------------------------------------
extern "C" size_t wd_head_workspace_bytes(const WdHead* head, int64_t n_rows) {
  if (wd::head_check(head) != WD_OK || n_rows < 1) return 0;
  return (size_t)wd::head_layout(*head, n_rows).total * sizeof(float);
}

extern "C" int wd_head_forward(const WdHead* head, const float* X, int32_t ldx, int64_t n_rows, const float* targets,
                               const float* weights, const float* mask, void* workspace, size_t workspace_bytes,
                               float* preds, float* loss, void* stream) {
  using namespace wd;
  HEAD_TRY(head_check(head));
  if (!X || !preds || n_rows < 1 || ldx < head->in_dim) return fail(WD_ERR_INVALID_ARG, "wd_head_forward: NULL pointer, no rows or ldx < in_dim");
  if (targets && (!mask || !loss)) return fail(WD_ERR_INVALID_ARG, "wd_head_forward: targets need mask and loss");
  if (!workspace || workspace_bytes < wd_head_workspace_bytes(head, n_rows)) return fail(WD_ERR_INVALID_ARG, "wd_head_forward: workspace too small");
  const HeadLayout L = head_layout(*head, n_rows);
  int rc = WD_OK;
  WD_DISPATCH_ACT(head->activation, rc = head_forward_impl<ACT>(*head, L, X, ldx, n_rows, targets, weights, mask,
                                                                reinterpret_cast<float*>(workspace), preds, loss, (cudaStream_t)stream));
  return rc;
}

extern "C" int wd_head_backward(const WdHead* head, const float* X, int32_t ldx, int64_t n_rows, void* workspace,
                                size_t workspace_bytes, const float* dloss, const float* dpreds, float* dX, int32_t lddx,
                                const WdHeadGrads* grads, void* stream) {
  using namespace wd;
  HEAD_TRY(head_check(head));
  if (!X || !grads || n_rows < 1) return fail(WD_ERR_INVALID_ARG, "wd_head_backward: NULL pointer or no rows");
  for (int l = 0; l < head->n_layers; ++l)
    if (!grads->dW[l] || !grads->db[l]) return fail(WD_ERR_INVALID_ARG, "wd_head_backward: gradient buffer of layer %d is NULL", l);
  if (dX && (lddx < head->in_dim || (lddx & 3) != 0 || (reinterpret_cast<uintptr_t>(dX) & 15) != 0))
    return fail(WD_ERR_INVALID_ARG, "wd_head_backward: dX needs lddx >= in_dim, lddx % 4 == 0 and a 16-byte aligned base");
  if (!workspace || workspace_bytes < wd_head_workspace_bytes(head, n_rows)) return fail(WD_ERR_INVALID_ARG, "wd_head_backward: workspace too small");
  const HeadLayout L = head_layout(*head, n_rows);
  int rc = WD_OK;
  WD_DISPATCH_ACT(head->activation, rc = head_backward_impl<ACT>(*head, L, X, ldx, n_rows, reinterpret_cast<float*>(workspace), dloss,
                                                                 dpreds, dX, lddx, *grads, (cudaStream_t)stream));
  return rc;
}
