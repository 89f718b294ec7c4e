// wd_encoder_host: the reference's own calling convention -- get_components() yields HOST
// tensors and MPNEncoder.forward uploads them on every call (chemprop/models/mpn.py:81-87).
// Upload -> device CSR build -> forward (-> backward) -> download, all stream-ordered.
#include <vector>

#include "common.cuh"

namespace {

struct DeviceArena {  // stream-ordered scratch, released in the destructor
  cudaStream_t s;
  std::vector<void*> ptrs;
  explicit DeviceArena(cudaStream_t st) : s(st) {}
  ~DeviceArena() { for (void* p : ptrs) cudaFreeAsync(p, s); }
  template <typename T>
  cudaError_t alloc(T** out, size_t n) {
    void* p = nullptr;
    cudaError_t e = cudaMallocAsync(&p, std::max<size_t>(n, 1) * sizeof(T), s);
    if (e == cudaSuccess) ptrs.push_back(p);
    *out = reinterpret_cast<T*>(p);
    return e;
  }
  template <typename T>
  cudaError_t upload(T** out, const T* host, size_t n) {
    cudaError_t e = alloc(out, n);
    if (e != cudaSuccess || n == 0) return e;
    return cudaMemcpyAsync(*out, host, n * sizeof(T), cudaMemcpyHostToDevice, s);
  }
  // dense host rows [n, f] -> device rows of stride ld >= f (pad columns zeroed): feature rows land on
  // 16-byte boundaries, which is what the asynchronous-copy loaders of the tensor-core kernels want
  cudaError_t upload_rows(float** out, const float* host, size_t n, int f, int ld) {
    if (ld == f) return upload(out, host, n * f);
    cudaError_t e = alloc(out, n * ld);
    if (e != cudaSuccess || n == 0) return e;
    e = cudaMemsetAsync(*out, 0, n * ld * sizeof(float), s);
    if (e != cudaSuccess) return e;
    return cudaMemcpy2DAsync(*out, (size_t)ld * sizeof(float), host, (size_t)f * sizeof(float), (size_t)f * sizeof(float), n,
                             cudaMemcpyHostToDevice, s);
  }
};

}  // namespace

extern "C" int wd_encoder_host(const float* f_atoms, const float* f_bonds, const float* w_atoms,
                               const float* w_bonds, const int64_t* a2b, const int64_t* b2a,
                               const int64_t* b2revb, const int64_t* a_scope, const double* dop,
                               int64_t Na, int64_t Nb, int64_t K, int64_t G, const WdWeights* hw,
                               const WdConfig* cfg, float* out, const float* d_out,
                               const WdGrads* hg, void* stream_) {
  using namespace wd;
  cudaStream_t s = (cudaStream_t)stream_;
  if (!cfg || !hw || !f_atoms || !f_bonds || !w_atoms || !w_bonds || !a2b || !b2a || !b2revb ||
      (G > 0 && (!a_scope || !dop || !out)))
    return fail(WD_ERR_INVALID_ARG, "wd_encoder_host: NULL host pointer");
  if (d_out && !hg) return fail(WD_ERR_INVALID_ARG, "wd_encoder_host: d_out given without grads");
  const size_t ws_bytes = wd_workspace_bytes(cfg, Na, Nb, G);
  if (ws_bytes == 0) return fail(WD_ERR_INVALID_ARG, "wd_encoder_host: invalid config or sizes");
  const int h = cfg->hidden_size, Fa = cfg->atom_fdim, Fb = cfg->bond_fdim, Fo = Fa + h;

  std::vector<int32_t> h_start(G), h_size(G);
  std::vector<float> h_dop(G);
  for (int64_t g = 0; g < G; ++g) {
    h_start[g] = (int32_t)a_scope[2 * g];
    h_size[g] = (int32_t)a_scope[2 * g + 1];
    h_dop[g] = (float)dop[g];
  }

  DeviceArena A(s);
  WdGraph dg{};
  WdWeights dw{};
  float *d_fa, *d_fb, *d_wa, *d_wb, *d_dop, *d_o = nullptr, *d_do = nullptr;
  int64_t *d_a2b, *d_b2a, *d_rev;
  int32_t *d_start, *d_size, *in_ptr, *in_idx, *out_idx, *src, *rev, *dst, *mol_of, *flags;
  void* ws;
  const int lda = (Fa + 3) & ~3, ldb = (Fb + 3) & ~3;
  WD_CUDA_OK(A.upload_rows(&d_fa, f_atoms, (size_t)Na, Fa, lda));
  WD_CUDA_OK(A.upload_rows(&d_fb, f_bonds, (size_t)Nb, Fb, ldb));
  WD_CUDA_OK(A.upload(&d_wa, w_atoms, (size_t)Na));
  WD_CUDA_OK(A.upload(&d_wb, w_bonds, (size_t)Nb));
  WD_CUDA_OK(A.upload(&d_a2b, a2b, (size_t)Na * K));
  WD_CUDA_OK(A.upload(&d_b2a, b2a, (size_t)Nb));
  WD_CUDA_OK(A.upload(&d_rev, b2revb, (size_t)Nb));
  WD_CUDA_OK(A.upload(&d_start, h_start.data(), (size_t)G));
  WD_CUDA_OK(A.upload(&d_size, h_size.data(), (size_t)G));
  WD_CUDA_OK(A.upload(&d_dop, h_dop.data(), (size_t)G));
  WD_CUDA_OK(A.alloc(&in_ptr, (size_t)Na + 1));
  WD_CUDA_OK(A.alloc(&in_idx, (size_t)Nb));
  WD_CUDA_OK(A.alloc(&out_idx, (size_t)Nb));
  WD_CUDA_OK(A.alloc(&src, (size_t)Nb));
  WD_CUDA_OK(A.alloc(&rev, (size_t)Nb));
  WD_CUDA_OK(A.alloc(&dst, (size_t)Nb));
  WD_CUDA_OK(A.alloc(&mol_of, (size_t)Na));
  WD_CUDA_OK(A.alloc(&flags, 4));
  {
    char* p;
    WD_CUDA_OK(A.alloc(&p, ws_bytes));
    ws = p;
  }
  float *W_i, *W_h, *W_o, *b_o, *b_i = nullptr, *b_h = nullptr;
  WD_CUDA_OK(A.upload(&W_i, hw->W_i, (size_t)h * Fb));
  WD_CUDA_OK(A.upload(&W_h, hw->W_h, (size_t)h * h));
  WD_CUDA_OK(A.upload(&W_o, hw->W_o, (size_t)h * Fo));
  WD_CUDA_OK(A.upload(&b_o, hw->b_o, (size_t)h));
  if (cfg->bias) {
    if (!hw->b_i || !hw->b_h) return fail(WD_ERR_INVALID_ARG, "cfg.bias set but b_i/b_h NULL");
    WD_CUDA_OK(A.upload(&b_i, hw->b_i, (size_t)h));
    WD_CUDA_OK(A.upload(&b_h, hw->b_h, (size_t)h));
  }
  dw = WdWeights{W_i, b_i, W_h, b_h, W_o, b_o};

  int rc = wd_csr_build(d_a2b, Na, K, d_b2a, d_rev, Nb, d_start, d_size, G, in_ptr, in_idx, out_idx, src,
                        rev, dst, mol_of, flags, s);
  if (rc != WD_OK) return rc;
  rc = wd_csr_build_check(flags, s);
  if (rc != WD_OK) return rc;
  int32_t h_flags[4] = {0, 0, 1, 0};
  WD_CUDA_OK(cudaMemcpyAsync(h_flags, flags, sizeof(h_flags), cudaMemcpyDeviceToHost, s));
  WD_CUDA_OK(cudaStreamSynchronize(s));
  dg.rev_adjacent = h_flags[2] == 0 ? 1 : 0;
  dg.n_atoms = Na; dg.n_bonds = Nb; dg.n_mols = G;
  dg.f_atoms = d_fa; dg.f_bonds = d_fb; dg.w_atoms = d_wa; dg.w_bonds = d_wb;
  dg.ld_atoms = lda; dg.ld_bonds = ldb;
  dg.in_ptr = in_ptr; dg.in_idx = in_idx; dg.out_idx = out_idx; dg.src = src; dg.rev = rev; dg.dst = dst;
  dg.mol_of = mol_of; dg.a_start = d_start; dg.a_size = d_size; dg.dop = d_dop;

  WD_CUDA_OK(A.alloc(&d_o, (size_t)G * h));
  rc = wd_encoder_forward(&dg, &dw, cfg, ws, ws_bytes, d_o, s);
  if (rc != WD_OK) return rc;
  if (G > 0) WD_CUDA_OK(cudaMemcpyAsync(out, d_o, sizeof(float) * G * h, cudaMemcpyDeviceToHost, s));
  if (d_out) {
    WdGrads g{};
    WD_CUDA_OK(A.upload(&d_do, d_out, (size_t)G * h));
    WD_CUDA_OK(A.alloc(&g.dW_i, (size_t)h * Fb));
    WD_CUDA_OK(A.alloc(&g.dW_h, (size_t)h * h));
    WD_CUDA_OK(A.alloc(&g.dW_o, (size_t)h * Fo));
    WD_CUDA_OK(A.alloc(&g.db_o, (size_t)h));
    if (cfg->bias) {
      WD_CUDA_OK(A.alloc(&g.db_i, (size_t)h));
      WD_CUDA_OK(A.alloc(&g.db_h, (size_t)h));
    }
    rc = wd_encoder_backward(&dg, &dw, cfg, ws, ws_bytes, d_do, &g, s);
    if (rc != WD_OK) return rc;
    auto down = [&](float* hp, const float* dp, size_t n) {
      return hp ? cudaMemcpyAsync(hp, dp, n * sizeof(float), cudaMemcpyDeviceToHost, s) : cudaSuccess;
    };
    WD_CUDA_OK(down(hg->dW_i, g.dW_i, (size_t)h * Fb));
    WD_CUDA_OK(down(hg->dW_h, g.dW_h, (size_t)h * h));
    WD_CUDA_OK(down(hg->dW_o, g.dW_o, (size_t)h * Fo));
    WD_CUDA_OK(down(hg->db_o, g.db_o, (size_t)h));
    if (cfg->bias) {
      WD_CUDA_OK(down(hg->db_i, g.db_i, (size_t)h));
      WD_CUDA_OK(down(hg->db_h, g.db_h, (size_t)h));
    }
  }
  WD_CUDA_OK(cudaStreamSynchronize(s));
  return WD_OK;
}
