// Probe of the TMA row-gather (cp.async.bulk.tensor.2d.tile::gather4) for the gathered GEMM operand:
//  (a) what lands where in shared memory (SWIZZLE_128B, TFLOAT32 map, box 32 columns x 1 row),
//  (b) bytes per cycle per SM against 16-byte cp.async (LDGSTS) for 128 random rows x 128 bytes per stage.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o build/gather4_probe scripts/gather4_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../polymer_chemprop_b200/csrc/tc_tma.cuh"

thread_local char wd::g_err[512];
std::atomic<int64_t> wd::g_launches;

using namespace wd;

__device__ __forceinline__ void tma_gather4(void* dst, const CUtensorMap* map, int c0, int r0, int r1, int r2, int r3, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(
          ptx::smem_u32(dst)),
      "l"(map), "r"(ptx::smem_u32(bar)), "r"(c0), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
      : "memory");
}

constexpr int STAGES = 6, STAGE_BYTES = 16384;

// mode 0: gather4 (lane 0 of warp 0 issues 32 ops per stage); mode 1: LDGSTS by 128 threads (8 x 16 B each per stage)
__global__ void __launch_bounds__(160, 1) k_gather(const __grid_constant__ CUtensorMap map, const float* G, int ld, const int* idx,
                                                    int iters, int mode, float* dump, unsigned long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* ring = smem_raw + (((base + 1023u) & ~1023u) - base);
  uint64_t* full = reinterpret_cast<uint64_t*>(ring + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(&full[s], mode == 0 ? 1 : 128); ptx::mbar_init(&empty[s], 1); }
    ptx::fence_barrier_init();
  }
  __syncthreads();
  const long long t0 = clock64();
  const int* my = idx + (size_t)blockIdx.x * 128 * 64;     // 64 different row sets per CTA
  if (warp == 4) {            // consumer: just releases stages (and dumps the first one)
    for (int it = 0; it < iters; ++it) {
      const int s = it % STAGES;
      ptx::mbar_wait(&full[s], (it / STAGES) & 1);
      if (it == 0 && blockIdx.x == 0)
        for (int i = lane; i < STAGE_BYTES / 4; i += 32) dump[i] = reinterpret_cast<float*>(ring)[i];
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&empty[s]);
    }
  } else if (mode == 0) {
    if (warp == 0) {          // every lane issues one gather4 (4 rows) per stage
      for (int it = 0; it < iters; ++it) {
        const int s = it % STAGES;
        const int4 r4 = *reinterpret_cast<const int4*>(my + (it % 64) * 128 + 4 * lane);
        ptx::mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
        if (lane == 0) ptx::mbar_arrive_expect_tx(&full[s], STAGE_BYTES);
        __syncwarp();
        tma_gather4(ring + s * STAGE_BYTES + lane * 512, &map, 32 * (it % 8), r4.x, r4.y, r4.z, r4.w, &full[s]);
      }
    }
  } else {
    const int t = threadIdx.x, c = t & 7, rb = t >> 3;
    for (int it = 0; it < iters; ++it) {
      const int s = it % STAGES;
      ptx::mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
      const int* rows = my + (it % 64) * 128;
      const uint32_t dst = ptx::smem_u32(ring + s * STAGE_BYTES);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int rl = rb + 16 * j;
        ptx::cp_async16(dst + ptx::sw128_off(rl, c), G + (size_t)rows[rl] * ld + 32 * (it % 8) + c * 4, 16u);
      }
      ptx::cp_async_arrive(&full[s]);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = (unsigned long long)(clock64() - t0);
}

int main() {
  const int NR = 270000, LD = 304, K = 300;
  std::vector<float> h((size_t)NR * LD);
  for (int r = 0; r < NR; ++r)
    for (int c = 0; c < LD; ++c) h[(size_t)r * LD + c] = (float)((r % 32) * 32 + (c % 32)) + (c == 5 ? 0.37f : 0.f);
  float* G; cudaMalloc(&G, h.size() * 4); cudaMemcpy(G, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  std::vector<int> hidx((size_t)148 * 64 * 128);
  srand(1);
  for (auto& v : hidx) v = rand() % NR;
  int* idx; cudaMalloc(&idx, hidx.size() * 4); cudaMemcpy(idx, hidx.data(), hidx.size() * 4, cudaMemcpyHostToDevice);
  float* dump; cudaMalloc(&dump, STAGE_BYTES);
  unsigned long long* cyc; cudaMalloc(&cyc, 8);
  const int smem = 1024 + STAGES * STAGE_BYTES + 256;
  cudaFuncSetAttribute(k_gather, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int box_rows : {1}) {
    CUtensorMap map;
    if (tma_make_2d(&map, G, NR, K, LD, box_rows, TMA_SW128) != WD_OK) { printf("encode failed: %s\n", g_err); continue; }
    for (int mode = 0; mode < 2; ++mode) {
      if (mode == 1 && box_rows == 4) continue;
      const int iters = 2000;
      cudaMemset(dump, 0, STAGE_BYTES);
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaEventRecord(e0);
      k_gather<<<148, 160, smem>>>(map, G, LD, idx, iters, mode, dump, cyc);
      cudaEventRecord(e1);
      cudaError_t e = cudaDeviceSynchronize();
      float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
      unsigned long long c = 0; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
      printf("== %s box_rows=%d: %s; %.3f ms, %.1f B/cycle/SM, %.2f TB/s aggregate\n", mode == 0 ? "TMA gather4" : "LDGSTS 16B", box_rows,
             cudaGetErrorString(e), ms, (double)iters * STAGE_BYTES / (double)c, 148.0 * iters * STAGE_BYTES / (ms * 1e-3) / 1e12);
      if (e != cudaSuccess) return 1;
      std::vector<float> d(STAGE_BYTES / 4);
      cudaMemcpy(d.data(), dump, STAGE_BYTES, cudaMemcpyDeviceToHost);
      // expectation: stage row q*4+i = G[rows[q*4+i]][0..31], SWIZZLE_128B K-major rows of 128 bytes
      int good = 0;
      for (int rl = 0; rl < 128; ++rl)
        for (int k = 0; k < 32; ++k) {
          const float got = d[(ptx::sw128_off(rl, k >> 2) + (k & 3) * 4) / 4];
          const float want = h[(size_t)hidx[rl] * LD + k];
          if (got == want) ++good;
          else if (k == 5 && rl < 2) printf("   row %d col 5: got %.6f want %.6f (tf32 rounding)\n", rl, got, want);
        }
      printf("   stage 0 matches the SW128 row layout in %d of 4096 floats; d[0..3] = %.1f %.1f %.1f %.1f (row %d starts at %.1f)\n", good,
             d[0], d[1], d[2], d[3], hidx[0], h[(size_t)hidx[0] * LD]);
    }
  }
  return 0;
}
