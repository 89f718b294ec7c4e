// Device-side assembly of the padded feature matrices from the compact form of a BatchMolGraph.
// MolGraph builds every bond row as f_atoms[a1] + f_bond (chemprop/features/featurization.py:467-468,
// and :616-629 for the polymer edges), so f_bonds[b, :F_a] == f_atoms[b2a[b]] and only the bond-only
// tail [N_b, F_b - F_a] carries information: 15 MB instead of 159 MB at batch 4096.  The host keeps the
// tail (BatchMolGraph.f_bonds_tail, verified exactly per molecule when the graph is built), uploads it
// with the dense f_atoms, and these kernels rebuild the get_components() matrices in HBM, already in
// the 16-byte padded row layout the tensor-core kernels read (WdGraph.ld_atoms / ld_bonds).
// Byte copies only: the result is bit-identical to uploading f_bonds itself.
#include "common.cuh"

namespace wd {
namespace {

// out[r, :] = [ X[g(r), 0:Fx], T[r, 0:Ft], 0 ... ]  with g(r) = gather ? gather[r] : r; out rows are ld floats
// (ld % 4 == 0).  One thread per (row, float4 column): the writes are full 16-byte coalesced stores, the
// reads are scalar (dense 133-float rows are only 4-byte aligned) and served by L1/L2 (X is re-read ~3x).
__global__ void k_expand_rows(const float* __restrict__ X, int Fx, const int64_t* __restrict__ gather,
                              int64_t n_src, const float* __restrict__ T, int Ft, float* __restrict__ out,
                              int ld, int64_t n_rows) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  const int nc4 = ld / 4;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows * nc4) return;
  int64_t r = i / nc4;
  int c = (int)(i - r * nc4) * 4;
  int64_t g = gather ? gather[r] : r;
  if (g < 0 || g >= n_src) g = 0;   // an inconsistent graph is rejected by wd_csr_build's flags; stay in bounds here
  const float* x = X + g * Fx;
  const float* t = T ? T + r * Ft : nullptr;
  float v[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    int j = c + u;
    v[u] = j < Fx ? __ldg(x + j) : (j < Fx + Ft ? __ldg(t + (j - Fx)) : 0.f);
  }
  st4(out + r * ld + c, make_float4(v[0], v[1], v[2], v[3]));
}

}  // namespace
}  // namespace wd

extern "C" int wd_expand_features(const float* f_atoms, const float* bond_tail, const int64_t* b2a,
                                  int64_t n_atoms, int64_t n_bonds, int32_t atom_fdim, int32_t tail_fdim,
                                  float* f_atoms_out, int32_t ld_atoms, float* f_bonds_out, int32_t ld_bonds,
                                  void* stream_) {
  using namespace wd;
  cudaStream_t s = (cudaStream_t)stream_;
  if (!f_atoms || !b2a || !f_bonds_out || (tail_fdim > 0 && !bond_tail) || n_atoms < 1 || n_bonds < 1 ||
      atom_fdim < 1 || tail_fdim < 0)
    return fail(WD_ERR_INVALID_ARG, "wd_expand_features: NULL pointer or bad size");
  if (ld_bonds < atom_fdim + tail_fdim || (ld_bonds & 3) || (f_atoms_out && (ld_atoms < atom_fdim || (ld_atoms & 3))))
    return fail(WD_ERR_INVALID_ARG, "wd_expand_features: ld_atoms / ld_bonds must be multiples of 4 and >= the row width");
  if (f_atoms_out)
    WD_LAUNCH(k_expand_rows, (unsigned)ceil_div(n_atoms * (ld_atoms / 4), 256), 256, 0, s, f_atoms, (int)atom_fdim,
              (const int64_t*)nullptr, n_atoms, (const float*)nullptr, 0, f_atoms_out, (int)ld_atoms, n_atoms);
  WD_LAUNCH(k_expand_rows, (unsigned)ceil_div(n_bonds * (ld_bonds / 4), 256), 256, 0, s, f_atoms, (int)atom_fdim, b2a,
            n_atoms, bond_tail, (int)tail_fdim, f_bonds_out, (int)ld_bonds, n_bonds);
  return WD_OK;
}
