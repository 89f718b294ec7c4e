// Device-side batch assembly from a device-resident per-graph store.
//
// The reference assembles every batch on the host: MoleculeDataset.batch_graph() (chemprop/data/data.py:277-314)
// concatenates the cached MolGraphs with Python list appends in BatchMolGraph.__init__
// (chemprop/features/featurization.py:757-813: per-molecule index offsets, row 0 padding) and
// MPNEncoder.forward uploads the result on every call (chemprop/models/mpn.py:84-87).  Here the featurised
// graphs live in HBM once (compact: atom rows, bond-only tails, local indices), and one step's batch is
// assembled by these kernels straight into the arrays the encoder kernels read (WdGraph): the same
// concatenation, offsets and padding row, bit for bit (tests/test_gpu_store.py), with only the selected
// graph ids and their start offsets (a few KB) crossing PCIe.
// Pure byte / integer work, HBM-bound: one thread per (output row, 16-byte column), coalesced stores.
#include "common.cuh"

namespace wd {
namespace {

// one warp per selected molecule: which molecule every output atom / bond row belongs to
__global__ void k_store_owner(const int32_t* __restrict__ a_start, const int32_t* __restrict__ a_size,
                              const int32_t* __restrict__ b_start, const int32_t* __restrict__ b_size, int64_t n_mols,
                              int32_t* __restrict__ mol_of, int32_t* __restrict__ bond_mol) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (g >= n_mols) return;
  const int as = a_start[g], an = a_size[g], bs = b_start[g], bn = b_size[g];
  for (int i = lane; i < an; i += 32) mol_of[as + i] = (int32_t)g;
  for (int i = lane; i < bn; i += 32) bond_mol[bs + i] = (int32_t)g;
}

// atom rows: f_atoms (padded to ld floats), w_atoms, CSR row pointers
__global__ void k_store_atoms(const WdStore st, const int64_t* __restrict__ sel, const int32_t* __restrict__ a_start,
                              const int32_t* __restrict__ b_start, const int32_t* __restrict__ mol_of,
                              float* __restrict__ f_atoms, int ld, float* __restrict__ w_atoms,
                              int32_t* __restrict__ in_ptr, int64_t n_atoms, int64_t n_bonds) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  const int nc4 = ld / 4;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_atoms * nc4) return;
  int64_t r = i / nc4;
  int c = (int)(i - r * nc4) * 4;
  float v[4] = {0.f, 0.f, 0.f, 0.f};
  if (r >= 1) {                                            // row 0 is the padding row (featurization.py:775-781)
    const int g = mol_of[r];
    const int64_t sa = st.atom_off[sel[g]] + (r - a_start[g]);     // row of this atom in the store
    const float* x = st.f_atoms + sa * st.atom_fdim;
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = c + u < st.atom_fdim ? __ldg(x + c + u) : 0.f;
    if (c == 0) {
      w_atoms[r] = st.w_atoms[sa];
      in_ptr[r] = (b_start[g] - 1) + st.in_ptr_local[sa];  // bonds of molecule g occupy CSR slots [b_start-1, ...)
    }
  } else if (c == 0) {
    w_atoms[0] = 0.f;
    in_ptr[0] = 0;
    in_ptr[n_atoms] = (int32_t)(n_bonds - 1);
  }
  st4(f_atoms + r * ld + c, make_float4(v[0], v[1], v[2], v[3]));
}

// bond rows: f_bonds = [f_atoms[source atom], bond-only tail] (padded to ld floats), w_bonds, src / rev / dst and the
// in-edge / out-edge lists with the batch offsets applied (featurization.py:787-793)
__global__ void k_store_bonds(const WdStore st, const int64_t* __restrict__ sel, const int32_t* __restrict__ a_start,
                              const int32_t* __restrict__ b_start, const int32_t* __restrict__ bond_mol,
                              float* __restrict__ f_bonds, int ld, float* __restrict__ w_bonds,
                              int32_t* __restrict__ src, int32_t* __restrict__ rev, int32_t* __restrict__ dst,
                              int32_t* __restrict__ in_idx, int32_t* __restrict__ out_idx, int64_t n_bonds) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  const int nc4 = ld / 4;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_bonds * nc4) return;
  int64_t b = i / nc4;
  int c = (int)(i - b * nc4) * 4;
  float v[4] = {0.f, 0.f, 0.f, 0.f};
  if (b >= 1) {
    const int g = bond_mol[b];
    const int64_t s = sel[g];
    const int64_t a0 = st.atom_off[s], b0 = st.bond_off[s];
    const int64_t lb = b - b_start[g], sb = b0 + lb;       // local bond id, row of this bond in the store
    const int la = st.b2a[sb];
    const float* x = st.f_atoms + (a0 + la) * st.atom_fdim;
    const float* t = st.bond_tail + sb * st.tail_fdim;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = c + u;
      v[u] = j < st.atom_fdim ? __ldg(x + j) : (j < st.atom_fdim + st.tail_fdim ? __ldg(t + (j - st.atom_fdim)) : 0.f);
    }
    if (c == 0) {
      const int lr = st.b2revb[sb];
      w_bonds[b] = st.w_bonds[sb];
      src[b] = a_start[g] + la;
      rev[b] = b_start[g] + lr;
      dst[b] = a_start[g] + st.b2a[b0 + lr];
      in_idx[b - 1] = b_start[g] + st.in_idx[sb];          // slot (b_start - 1) + lb of the concatenated lists
      out_idx[b - 1] = b_start[g] + st.out_idx[sb];
    }
  } else if (c == 0) {
    w_bonds[0] = 0.f;
    src[0] = 0; rev[0] = 0; dst[0] = 0;
  }
  st4(f_bonds + b * ld + c, make_float4(v[0], v[1], v[2], v[3]));
}

__global__ void k_store_mols(const WdStore st, const int64_t* __restrict__ sel, float* __restrict__ dop, int64_t n_mols) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g < n_mols) dop[g] = st.dop[sel[g]];
}

}  // namespace
}  // namespace wd

extern "C" int wd_store_assemble(const WdStore* store, const int64_t* sel, const int32_t* a_start,
                                 const int32_t* a_size, const int32_t* b_start, const int32_t* b_size, int64_t n_mols,
                                 const WdGraph* out, int32_t* bond_mol, void* stream_) {
  using namespace wd;
  cudaStream_t s = (cudaStream_t)stream_;
  if (!store || !out || !sel || !a_start || !a_size || !b_start || !b_size || !bond_mol || n_mols < 1)
    return fail(WD_ERR_INVALID_ARG, "wd_store_assemble: NULL pointer or empty selection");
  const WdStore& st = *store;
  const WdGraph& g = *out;
  if (!st.f_atoms || !st.bond_tail || !st.w_atoms || !st.w_bonds || !st.b2a || !st.b2revb || !st.in_idx || !st.out_idx ||
      !st.in_ptr_local || !st.atom_off || !st.bond_off || !st.dop || st.atom_fdim < 1 || st.tail_fdim < 0)
    return fail(WD_ERR_INVALID_ARG, "wd_store_assemble: store has NULL device pointers");
  if (g.n_atoms < 1 || g.n_bonds < 1 || g.n_mols != n_mols || !g.f_atoms || !g.f_bonds || !g.w_atoms || !g.w_bonds ||
      !g.in_ptr || !g.in_idx || !g.out_idx || !g.src || !g.rev || !g.dst || !g.mol_of || !g.dop)
    return fail(WD_ERR_INVALID_ARG, "wd_store_assemble: output graph has NULL device pointers or wrong sizes");
  if (g.ld_atoms < st.atom_fdim || (g.ld_atoms & 3) || g.ld_bonds < st.atom_fdim + st.tail_fdim || (g.ld_bonds & 3))
    return fail(WD_ERR_INVALID_ARG, "wd_store_assemble: ld_atoms / ld_bonds must be multiples of 4 and >= the row width");
  int32_t* mol_of = const_cast<int32_t*>(g.mol_of);
  WD_CUDA_OK(cudaMemsetAsync(mol_of, 0xff, sizeof(int32_t) * 1, s));           // the padding atom is in no scope
  WD_LAUNCH(k_store_owner, (unsigned)ceil_div(n_mols * 32, 256), 256, 0, s, a_start, a_size, b_start, b_size, n_mols, mol_of,
            bond_mol);
  WD_LAUNCH(k_store_atoms, (unsigned)ceil_div(g.n_atoms * (g.ld_atoms / 4), 256), 256, 0, s, st, sel, a_start, b_start,
            (const int32_t*)mol_of, const_cast<float*>(g.f_atoms), (int)g.ld_atoms, const_cast<float*>(g.w_atoms),
            const_cast<int32_t*>(g.in_ptr), g.n_atoms, g.n_bonds);
  WD_LAUNCH(k_store_bonds, (unsigned)ceil_div(g.n_bonds * (g.ld_bonds / 4), 256), 256, 0, s, st, sel, a_start, b_start,
            (const int32_t*)bond_mol, const_cast<float*>(g.f_bonds), (int)g.ld_bonds, const_cast<float*>(g.w_bonds),
            const_cast<int32_t*>(g.src), const_cast<int32_t*>(g.rev), const_cast<int32_t*>(g.dst),
            const_cast<int32_t*>(g.in_idx), const_cast<int32_t*>(g.out_idx), g.n_bonds);
  WD_LAUNCH(k_store_mols, (unsigned)ceil_div(n_mols, 256), 256, 0, s, st, sel, const_cast<float*>(g.dop), n_mols);
  return WD_OK;
}
