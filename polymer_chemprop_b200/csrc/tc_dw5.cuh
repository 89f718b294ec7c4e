// tcgen05 weight-gradient contraction: TMA-fed MN-major operands.
//
//   P[z][n1, n2] = sum_{r in chunk z} X[r, n1] * Y[r, n2]
//
// What the wait-cycle counters and the ablations of generation 4 said (profiles/r1_g_*): the kernel was bound by
// work done ON the SM, not by HBM or L2 -- LDGSTS (cp.async) moves 16 bytes per cycle per SM at best, and the
// in-place rounding pass pushed every operand byte through shared memory two more times.  Here
//   * X and Y arrive by TMA tensor copies (cp.async.bulk.tensor, boxes of BK rows x 32 columns) written straight
//     into the MN-major SWIZZLE_128B_BASE32B operand layout, the tensor map's TFLOAT32 data type making the TMA
//     engine round fp32 -> tf32 in flight: no loader instructions, no transform pass, no proxy fence --
//     TMA -> mbarrier -> tcgen05.mma;
//   * the gathered operand of dW_h (S_t = a_message[b2a] - message[b2revb]) is a plain matrix too: the forward
//     depth step leaves it in the workspace (tc_gemm4.cuh).
// Tiling as tc_dw4.cuh: one CTA = up to 384 X columns x one Y tile of <= 160 columns x one row chunk z.
#pragma once
#include "tc_dw4.cuh"
#include "tc_tma.cuh"

namespace wd {

constexpr int D5_MAX_STAGES = 6;
constexpr int D5_MMA_WARP = 0, D5_TMA_WARP = 1, D5_EPI_W0 = 2;
constexpr int D5_BK = 32;                       // rows per stage
constexpr int D5_THREADS = 32 * (D5_EPI_W0 + 4);

struct D5Barriers {
  uint64_t full[D5_MAX_STAGES], empty[D5_MAX_STAGES];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

struct D5Params {
  CUtensorMap mx;       // X[R, N1], box BK x 32
  CUtensorMap my[2];    // Y segments (CONCAT: two), box BK x 32
  int N1;
  float* P;             // [Z][N1, ldp]
  int ldp;
  int64_t R, rows_per_z;
  int n1_groups, n2_tiles, stages;
  int XB, YB;           // 32-column blocks of the X / Y regions of a stage
  TcDwTile t[TC_DW_MAX_TILES];
};

inline int d5_stage_bytes(int XB, int YB) { return (XB + YB) * D5_BK * 128; }
inline int d5_smem_bytes(int XB, int YB, int stages) {
  return 1024 + stages * d5_stage_bytes(XB, YB) + TC_EPI_STAGE_BYTES + (int)sizeof(D5Barriers) + 64;
}

__global__ void __launch_bounds__(D5_THREADS, 1) k_tc_dw5(const __grid_constant__ D5Params p) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  constexpr int BK = D5_BK, BLK = BK * 128;                                   // bytes of one 32-column block
  extern __shared__ uint8_t smem_raw[];
  const int warp = ptx::uniform_warp_idx(), lane = threadIdx.x & 31;   // provably warp-uniform: see tc_ptx.cuh
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* ring = smem_raw + (((base + 1023u) & ~1023u) - base);
  const int XB = p.XB, YB = p.YB, SR = p.stages;
  const int x_bytes = XB * BLK, y_bytes = YB * BLK, stage_bytes = x_bytes + y_bytes;
  float* epi_st = reinterpret_cast<float*>(ring + SR * stage_bytes);
  D5Barriers* bar = reinterpret_cast<D5Barriers*>(ring + SR * stage_bytes + TC_EPI_STAGE_BYTES);
  long long tacc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};   // WD_T2_TIMING: 3 tma<-empty 5 mma<-full 7 epi<-tmem_full
  const long long t_start = clock64();

  if (threadIdx.x == 0) {
    for (int s = 0; s < SR; ++s) { ptx::mbar_init(&bar->full[s], 1); ptx::mbar_init(&bar->empty[s], 1); }
    ptx::mbar_init(&bar->tmem_full, 1);
    ptx::fence_barrier_init();
    ptx::tma_prefetch_desc(&p.mx);
  }
  if (warp == D5_MMA_WARP) ptx::tmem_alloc(&bar->tmem_base, TC_TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = bar->tmem_base;

  const int n1_0 = (int)(blockIdx.x % p.n1_groups) * D4_MROWS;
  const TcDwTile tl = p.t[blockIdx.x / p.n1_groups];
  const int n1_w = min(p.N1 - n1_0, D4_MROWS);
  const int n_mt = (int)ceil_div(n1_w, TC_BM);
  const int64_t r_begin = (int64_t)blockIdx.y * p.rows_per_z;                 // rows_per_z is a multiple of 32
  const int64_t r_end = min(p.R, r_begin + p.rows_per_z);
  const int KC = (int)ceil_div(r_end - r_begin, BK);
  const int xblocks = (int)ceil_div(n1_w, 32), yblocks = (int)ceil_div(tl.width, 32);

  if (warp == D5_MMA_WARP) {
    // ================= MMA issue =================
    const uint32_t idesc = idesc_tf32_mn(128, tl.BN);
    const uint32_t ring_u32 = ptx::uniform_u32(ptx::smem_u32(ring)), tmem_u = ptx::uniform_u32(tmem_base);
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % SR;
      T2_WAIT(5, &bar->full[s], (kc / SR) & 1);
      ptx::tc_fence_after();
      // convergent issue (the election is inside every statement); K group g = 8 rows = 1024 bytes = +64 in the low word
      const uint32_t xa = desc_lo_mn32(ring_u32 + s * stage_bytes, (uint32_t)BLK);
      const uint32_t ya = desc_lo_mn32(ring_u32 + s * stage_bytes + x_bytes, (uint32_t)BLK);
#pragma unroll
      for (int g = 0; g < BK / 8; ++g) {
        const uint32_t acc = (kc == 0 && g == 0) ? 0u : 1u;
        for (int mt = 0; mt < n_mt; ++mt)
          ptx::mma_elect<false>(tmem_u + mt * D4_ACC_COLS, xa + ((mt * 4 * BLK) >> 4) + g * 64, ya + g * 64, kDescHiMn32, idesc, acc);
      }
      ptx::tc_commit_elect(&bar->empty[s]);
      if (kc == KC - 1) ptx::tc_commit_elect(&bar->tmem_full);
      __syncwarp();
    }
  } else if (warp == D5_TMA_WARP) {
    // ================= TMA producer: X and Y boxes of every stage =================
    if (lane == 0) {
      const CUtensorMap* my = &p.my[tl.seg];
      const uint32_t bytes = (uint32_t)((xblocks + yblocks) * BLK);
      for (int kc = 0; kc < KC; ++kc) {
        const int s = kc % SR;
        const int r0 = (int)(r_begin + (int64_t)kc * BK);
        T2_WAIT(3, &bar->empty[s], ((kc / SR) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(&bar->full[s], bytes);
        uint8_t* xs = ring + s * stage_bytes;
        for (int b = 0; b < xblocks; ++b) ptx::tma_load_2d(xs + b * BLK, &p.mx, n1_0 + b * 32, r0, &bar->full[s]);
        for (int b = 0; b < yblocks; ++b) ptx::tma_load_2d(xs + x_bytes + b * BLK, my, tl.col0 + b * 32, r0, &bar->full[s]);
      }
    }
  } else {
    // ================= epilogue: warps 2-5, TMEM lane quarter = warp % 4 =================
    T2_WAIT(7, &bar->tmem_full, 0);
    ptx::tc_fence_after();
    const int we = warp & 3;
    float* Pz = p.P + (int64_t)blockIdx.y * p.N1 * p.ldp;
    for (int mt = 0; mt < n_mt; ++mt)
      d4_epilogue(tmem_base + mt * D4_ACC_COLS, we, lane, epi_st + we * 32 * 33, tl.BN, n1_0 + mt * TC_BM, p.N1,
                  tl.out_col0, tl.out_col0 + tl.width, Pz, p.ldp);
  }
#ifdef WD_T2_TIMING
  if (WD_TIMING_ON(15) && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0) {
    if (warp == D5_MMA_WARP || warp == D5_TMA_WARP || warp == D5_EPI_W0)
      for (int i = 0; i < 10; ++i) if (tacc[i]) atomicAdd(&g_t2_timing[i], (unsigned long long)tacc[i]);
    if (warp == D5_EPI_W0) atomicAdd(&g_t2_timing[15], (unsigned long long)(clock64() - t_start));
  }
#else
  (void)tacc; (void)t_start;
#endif
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == D5_MMA_WARP) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

}  // namespace wd
