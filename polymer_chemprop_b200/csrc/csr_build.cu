// Device rebuild of BatchMolGraph's padded a2b [N_a,K] int64 into int32 CSR / reverse-CSR
// (chemprop/features/featurization.py:757-813 builds the padded form on the host; the
// reference consumes it through index_select_ND, chemprop/nn_utils.py:50-67).
// Pure integer work: one thread per a2b row (K int64 ids, lanes K * 8 bytes apart -- rows are short and the whole
// matrix is read once, 6 MB at B=4096), a three-phase exclusive scan, no atomics on the
// data path (atomics only for the violation counter / max-degree statistic).
#include "common.cuh"

namespace wd {
namespace {

constexpr int kScanBlock = 256;

__device__ __forceinline__ int warp_incl_scan(int v) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int n = __shfl_up_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) >= o) v += n;
  }
  return v;
}

// block-wide exclusive scan of one int per thread (blockDim.x == kScanBlock); returns the
// exclusive prefix and the block total.
__device__ __forceinline__ int block_excl_scan(int v, int* total) {
  __shared__ int warp_sums[kScanBlock / 32];
  int incl = warp_incl_scan(v);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 31) warp_sums[w] = incl;
  __syncthreads();
  if (w == 0) {
    int s = lane < kScanBlock / 32 ? warp_sums[lane] : 0;
    s = warp_incl_scan(s);
    if (lane < kScanBlock / 32) warp_sums[lane] = s;
  }
  __syncthreads();
  int base = w ? warp_sums[w - 1] : 0;
  *total = warp_sums[kScanBlock / 32 - 1];
  __syncthreads();
  return base + incl - v;
}

// phase 1: in-degree per atom = non-zero slots of its a2b row; block-local exclusive prefix.
__global__ void k_csr_degree(const int64_t* __restrict__ a2b, int64_t n_atoms, int K,
                             int32_t* __restrict__ in_ptr, int32_t* __restrict__ block_sums,
                             int32_t* __restrict__ flags) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int64_t a = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
  int deg = 0;
  if (a < n_atoms) {
    const int64_t* row = a2b + a * K;
    for (int k = 0; k < K; ++k) deg += (row[k] != 0);
  }
  int total;
  int excl = block_excl_scan(deg, &total);
  if (a < n_atoms) in_ptr[a] = excl;
  if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
  int m = deg;
#pragma unroll
  for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(&flags[1], m);
}

// phase 2: one block scans the per-block totals in place (exclusive), carrying across chunks.
__global__ void k_csr_scan_blocks(int32_t* __restrict__ block_sums, int n_blocks) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n_blocks; base += kScanBlock) {
    int i = base + threadIdx.x;
    int v = i < n_blocks ? block_sums[i] : 0;
    int total;
    int excl = block_excl_scan(v, &total);
    if (i < n_blocks) block_sums[i] = carry + excl;
    __syncthreads();
    if (threadIdx.x == 0) carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) block_sums[n_blocks] = carry;
}

// phase 3: final row pointers, in-edge lists in reference order, dst scatter, and the
// out-edge list of each atom = reverses of its in-edges, sorted ascending (deterministic).
__global__ void k_csr_fill(const int64_t* __restrict__ a2b, const int64_t* __restrict__ b2revb,
                           int64_t n_atoms, int64_t n_bonds, int K,
                           const int32_t* __restrict__ block_sums, int32_t* __restrict__ in_ptr,
                           int32_t* __restrict__ in_idx, int32_t* __restrict__ out_idx,
                           int32_t* __restrict__ dst, int32_t* __restrict__ flags) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int64_t a = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
  if (a >= n_atoms) return;
  int base = in_ptr[a] + block_sums[blockIdx.x];
  in_ptr[a] = base;
  if (a == n_atoms - 1) in_ptr[n_atoms] = block_sums[gridDim.x];
  const int64_t* row = a2b + a * K;
  int n = 0, bad = 0;
  for (int k = 0; k < K; ++k) {
    int64_t b = row[k];
    if (b == 0) continue;
    if (b < 0 || b >= n_bonds) { ++bad; b = 0; }
    in_idx[base + n] = (int32_t)b;
    if (b) dst[b] = (int32_t)a;
    int64_t r = b ? b2revb[b] : 0;
    if (r < 0 || r >= n_bonds) { ++bad; r = 0; }
    // insertion sort of the reverse ids into out_idx[base .. base+n]
    int j = n;
    while (j > 0 && out_idx[base + j - 1] > (int32_t)r) {
      out_idx[base + j] = out_idx[base + j - 1];
      --j;
    }
    out_idx[base + j] = (int32_t)r;
    ++n;
  }
  if (bad) atomicAdd(&flags[0], bad);
}

// phase 4: int32 narrowing of b2a / b2revb and the invariant checks of SURVEY.md section 3.2.
__global__ void k_csr_bonds(const int64_t* __restrict__ b2a, const int64_t* __restrict__ b2revb,
                            int64_t n_atoms, int64_t n_bonds, const int32_t* __restrict__ in_ptr,
                            int32_t* __restrict__ src, int32_t* __restrict__ rev,
                            int32_t* __restrict__ dst, int32_t* __restrict__ flags) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_bonds) return;
  int64_t s = b2a[b], r = b2revb[b];
  int bad = 0;
  if (s < 0 || s >= n_atoms) { ++bad; s = 0; }
  if (r < 0 || r >= n_bonds) { ++bad; r = 0; }
  src[b] = (int32_t)s;
  rev[b] = (int32_t)r;
  if (b == 0) {
    dst[0] = 0;
    if (in_ptr[n_atoms] != (int32_t)(n_bonds - 1)) ++bad;  // every real bond listed exactly once
  } else {
    if (r != ((b - 1) ^ 1) + 1) atomicAdd(&flags[2], 1);    // not an adjacent reverse pair (never for reference-built graphs)
    if (b2revb[r] != b) ++bad;                              // reverse map is an involution
    int64_t sr = b2a[r];
    if (dst[b] < 0 || dst[b] != (int32_t)sr) ++bad;         // a2b row of b == source of rev(b)
  }
  if (bad) atomicAdd(&flags[0], bad);
}

__global__ void k_csr_mols(const int32_t* __restrict__ a_start, const int32_t* __restrict__ a_size,
                           int64_t n_mols, int64_t n_atoms, int32_t* __restrict__ mol_of,
                           int32_t* __restrict__ flags) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  // one warp per molecule, lanes stride over its atoms
  int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (g >= n_mols) return;
  int s = a_start[g], n = a_size[g];
  if (n < 0 || s < 0 || (int64_t)s + n > n_atoms) {
    if (lane == 0) atomicAdd(&flags[0], 1);
    return;
  }
  for (int i = lane; i < n; i += 32) mol_of[s + i] = (int32_t)g;
}

}  // namespace
}  // namespace wd

extern "C" int wd_csr_build(const int64_t* a2b, int64_t n_atoms, int64_t max_num_bonds,
                            const int64_t* b2a, const int64_t* b2revb, int64_t n_bonds,
                            const int32_t* a_start, const int32_t* a_size, int64_t n_mols,
                            int32_t* in_ptr, int32_t* in_idx, int32_t* out_idx, int32_t* src,
                            int32_t* rev, int32_t* dst, int32_t* mol_of, int32_t* flags,
                            void* stream_) {
  using namespace wd;
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n_atoms < 1 || n_bonds < 1 || max_num_bonds < 1 || n_mols < 0)
    return fail(WD_ERR_INVALID_ARG, "wd_csr_build: need n_atoms>=1, n_bonds>=1, K>=1 (got %lld, %lld, %lld)",
                (long long)n_atoms, (long long)n_bonds, (long long)max_num_bonds);
  if (n_bonds >= (1ll << 31) || n_atoms >= (1ll << 31) || n_atoms * max_num_bonds >= (1ll << 40))
    return fail(WD_ERR_UNSUPPORTED, "wd_csr_build: graph too large for int32 indices");
  int n_blocks = (int)ceil_div(n_atoms, kScanBlock);
  int32_t* block_sums = nullptr;
  WD_CUDA_OK(cudaMallocAsync((void**)&block_sums, sizeof(int32_t) * (n_blocks + 1), stream));
  WD_CUDA_OK(cudaMemsetAsync(flags, 0, 4 * sizeof(int32_t), stream));
  WD_CUDA_OK(cudaMemsetAsync(dst, 0xff, sizeof(int32_t) * n_bonds, stream));
  WD_CUDA_OK(cudaMemsetAsync(mol_of, 0xff, sizeof(int32_t) * n_atoms, stream));
  WD_LAUNCH(k_csr_degree, n_blocks, kScanBlock, 0, stream, a2b, n_atoms, (int)max_num_bonds, in_ptr,
            block_sums, flags);
  WD_LAUNCH(k_csr_scan_blocks, 1, kScanBlock, 0, stream, block_sums, n_blocks);
  WD_LAUNCH(k_csr_fill, n_blocks, kScanBlock, 0, stream, a2b, b2revb, n_atoms, n_bonds,
            (int)max_num_bonds, block_sums, in_ptr, in_idx, out_idx, dst, flags);
  WD_LAUNCH(k_csr_bonds, (int)ceil_div(n_bonds, 256), 256, 0, stream, b2a, b2revb, n_atoms, n_bonds,
            in_ptr, src, rev, dst, flags);
  if (n_mols > 0)
    WD_LAUNCH(k_csr_mols, (int)ceil_div(n_mols * 32, 256), 256, 0, stream, a_start, a_size, n_mols,
              n_atoms, mol_of, flags);
  WD_CUDA_OK(cudaFreeAsync(block_sums, stream));
  return WD_OK;
}

extern "C" int wd_csr_build_check(const int32_t* flags, void* stream_) {
  using namespace wd;
  cudaStream_t stream = (cudaStream_t)stream_;
  int32_t h[4] = {0, 0, 0, 0};
  WD_CUDA_OK(cudaMemcpyAsync(h, flags, sizeof(h), cudaMemcpyDeviceToHost, stream));
  WD_CUDA_OK(cudaStreamSynchronize(stream));
  if (h[0] != 0)
    return fail(WD_ERR_GRAPH,
                "BatchMolGraph invariants violated (%d): a2b/b2a/b2revb/a_scope are inconsistent "
                "(featurization.py:757-813 guarantees them)", h[0]);
  return WD_OK;
}
