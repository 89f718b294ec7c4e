// Decodes how tcgen05.mma kind::tf32 reads an MN-major SWIZZLE_128B shared-memory operand: a single 1.0 is
// walked through the operand region while the other operand (K-major, layout known-good from the GEMM
// kernels) holds B[n][k] = 8 n + k + 1, so D names the (m, k) the hardware associates with that float.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o gpurun_out/mn_probe scripts/mn_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../polymer_chemprop_b200/csrc/tc_ptx.cuh"

using namespace wd;

__device__ __forceinline__ uint64_t desc_mn(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

// mode 0: A MN-major probed, B K-major known.  mode 1: B MN-major probed, A K-major known.
__global__ void __launch_bounds__(128, 1) k_probe(int mode, uint32_t lbo, uint32_t sbo, uint32_t layout, int nfloats, int* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* reg = smem_raw + (((base + 1023u) & ~1023u) - base);
  float* probed = reinterpret_cast<float*>(reg);                 // 16 KB
  uint8_t* known = reg + 16384;                                  // K-major SW128: 128 rows x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(reg + 16384 + 16384);
  uint32_t* tbase = reinterpret_cast<uint32_t*>(bar + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
  // known operand: value(row, k) = 8 * row + k + 1 for k < 8 (exact in tf32 up to row 127: 1024)
  for (int i = threadIdx.x; i < 128 * 32; i += 128) {
    const int row = i >> 5, k = i & 31;
    *reinterpret_cast<float*>(known + ptx::sw128_off(row, k >> 2) + (k & 3) * 4) = k < 8 ? (float)(8 * row + k + 1) : 0.f;
  }
  if (warp == 0) ptx::tmem_alloc(tbase, 128);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tbase;
  const int M = 128, N = mode == 1 ? 128 : 16;
  uint32_t idesc = ptx::idesc_tf32(M, N) | (mode == 0 ? (1u << 15) : (mode == 1 ? (1u << 16) : 0u));
  uint32_t phase = 0;
  for (int f = 0; f < nfloats; ++f) {
    for (int i = threadIdx.x; i < 4096; i += 128) probed[i] = 0.f;
    __syncthreads();
    if (threadIdx.x == 0) probed[f] = 1.0f;
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    if (threadIdx.x == 0) {
      const uint64_t dp = desc_mn(ptx::smem_u32(probed), lbo, sbo, layout);
      const uint64_t dk = ptx::smem_desc_sw128(ptx::smem_u32(known));
      if (mode != 1) ptx::mma_tf32(tmem, dp, dk, idesc, 0u);
      else ptx::mma_tf32(tmem, dk, dp, idesc, 0u);
      ptx::tc_commit(bar);
    }
    ptx::mbar_wait(bar, phase);
    phase ^= 1;
    ptx::tc_fence_after();
    // D[m][n]: thread = row m.  mode 0: D[m][n] = sum_k A[m][k] (8 n + k + 1): row m* nonzero, D[m*][0] = k* + 1
    //                           mode 1: D[m][n] = sum_k (8 m + k + 1) B[n][k]: column n* nonzero, D[0][n*] = k* + 1
    float v[32];
    int found = -1;
    if (mode != 1) {
      ptx::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16), v);
      ptx::tmem_ld_wait();
      if (v[0] != 0.f) found = (threadIdx.x << 8) | (int)(v[0] - 1.f);
    } else {
      for (int c0 = 0; c0 < 128; c0 += 32) {
        ptx::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
        ptx::tmem_ld_wait();
        if (threadIdx.x == 0)
          for (int j = 0; j < 32; ++j) if (v[j] != 0.f) found = ((c0 + j) << 8) | (int)(v[j] - 1.f);
      }
    }
    if (found >= 0) out[f] = found;
    ptx::tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 128); }
}

int main() {
  int* d_out;
  const int NF = 2560;   // 10 KB of the probed region
  cudaMalloc(&d_out, NF * sizeof(int));
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000);
  struct V { int mode; uint32_t lbo, sbo, layout; const char* name; };
  const V vs[] = {{0, 1024, 512, 1, "A MN-major SW128_BASE32B LBO=1024 SBO=512"}, {1, 1024, 512, 1, "B MN-major SW128_BASE32B LBO=1024 SBO=512"}, {0, 512, 1024, 1, "A MN-major SW128_BASE32B LBO=512 SBO=1024"}, {2, 16, 1024, 2, "CONTROL: A K-major SW128 (LBO=16 SBO=1024)"}, {0, 1024, 4096, 2, "A MN-major SW128 LBO=1024 SBO=4096"}, {0, 4096, 1024, 2, "A MN-major SW128 LBO=4096 SBO=1024"},
                  {1, 1024, 4096, 2, "B MN-major SW128 LBO=1024 SBO=4096"}, {0, 1024, 4096, 0, "A MN-major NO-SWIZZLE LBO=1024 SBO=4096"},
                  {0, 128, 1024, 0, "A MN-major NO-SWIZZLE LBO=128 SBO=1024"}};
  for (const V& v : vs) {
    cudaMemset(d_out, 0xFF, NF * sizeof(int));
    k_probe<<<1, 128, 40000>>>(v.mode, v.lbo, v.sbo, v.layout, NF, d_out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("== %s: %s\n", v.name, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<int> h(NF);
    cudaMemcpy(h.data(), d_out, NF * sizeof(int), cudaMemcpyDeviceToHost);
    int hits = 0, match = 0;
    for (int f = 0; f < NF; ++f) {
      if (h[f] < 0) continue;
      ++hits;
      const int mn = h[f] >> 8, k = h[f] & 255;
      // expectation of tc_dw4.cuh: block j = f / 256 (LBO = 1024 apart), row k = (f % 256) / 32, 32-byte chunk ^ (k & 3)
      const int j = f / 256, kr = (f % 256) / 32, c = (f % 32) / 4, e4 = f % 4;
      if (mn == j * 32 + ((c ^ ((kr & 3) << 1)) << 2) + e4 && k == kr) ++match;
    }
    printf("   floats that reach the MMA: %d of %d; matching the tc_dw4 layout: %d\n", hits, NF, match);
    int shown = 0;
    for (int f = 0; f < NF && shown < 40; ++f)
      if (h[f] >= 0 && (f < 12 || f % 37 == 0)) { printf("   f=%4d -> mn=%3d k=%d\n", f, h[f] >> 8, h[f] & 255); ++shown; }
  }
  return 0;
}
