// tcgen05 weight-gradient contraction over the row (bond / atom) dimension, deterministic split:
//   P[z][n1, n2] = sum_{r in chunk z} X[r, n1] * Y(r, n2)       (Y virtual, operands.cuh)
// followed by k_reduce_partials (fixed z order, no atomics).
//
// Here the reduction axis K is the ROW axis, so both operands are transposed on their way into
// the K-major SWIZZLE_128B stages: a warp reads 4 consecutive rows x 32 consecutive columns
// (coalesced 128-byte row segments) and each lane writes its column's 4 row values as one
// 16-byte chunk.  One CTA owns one (n1 tile of 128) x (n2 tile <= 304) accumulator in TMEM and
// walks its row chunk in steps of 32 rows; same warp roles and pipeline as tc_gemm.cuh.
#pragma once
#include "tc_gemm.cuh"

namespace wd {

constexpr int TC_DW_MAX_TILES = 24;

struct TcDwTile {
  int seg;       // 0: Y.X (or the gathered operand), 1: Y.X2 (second CONCAT segment)
  int col0;      // first column inside the segment
  int width;     // valid columns
  int BN;        // round_up(width, 16)
  int out_col0;  // first column in P rows
};

struct TcDwParams {
  const float* X;   // [R, ldx]; columns [0, N1) are used
  int ldx, N1;
  AOperand Y;
  float* P;         // [Z][N1, ldp]
  int ldp;
  int64_t R, rows_per_z;
  int n1_tiles, n2_tiles, stages, maxBN;
  TcDwTile t[TC_DW_MAX_TILES];
};

template <int YMODE, int ACT, bool PREACT>
__device__ __forceinline__ float tc_dw_y(const TcDwParams& p, const TcDwTile& t, int64_t r, int col, int s, int rv) {
  if (YMODE == A_GATHER_DIFF) {
    float u = __ldg(p.Y.X2 + (int64_t)rv * p.Y.ldx2 + col);
    if (PREACT) u = act_fwd<ACT>(u);
    return __ldg(p.Y.X + (int64_t)s * p.Y.ldx + col) - u;
  }
  if (YMODE == A_CONCAT && t.seg == 1) return __ldg(p.Y.X2 + r * p.Y.ldx2 + col);
  return __ldg(p.Y.X + r * p.Y.ldx + col);
}

template <int YMODE, int ACT, bool PREACT>
__global__ void __launch_bounds__(TC_THREADS, 1) k_tc_dw(const TcDwParams p) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* tiles = smem_raw + (((base + 1023u) & ~1023u) - base);
  const int stage_bytes = TC_A_STAGE + p.maxBN * 128;
  const int S = p.stages;
  float* epi_st = reinterpret_cast<float*>(tiles + S * stage_bytes);
  TcBarriers* bar = reinterpret_cast<TcBarriers*>(tiles + S * stage_bytes + TC_EPI_STAGE_BYTES);

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { ptx::mbar_init(&bar->full[s], TC_PROD_WARPS); ptx::mbar_init(&bar->empty[s], 1); }
    ptx::mbar_init(&bar->tmem_full[0], 1);
    ptx::fence_barrier_init();
  }
  if (warp == TC_MMA_WARP) ptx::tmem_alloc(&bar->tmem_base, TC_TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = bar->tmem_base;

  const int n1_0 = (int)(blockIdx.x % p.n1_tiles) * TC_BM;
  const TcDwTile t = p.t[blockIdx.x / p.n1_tiles];
  const int64_t r_begin = (int64_t)blockIdx.y * p.rows_per_z;
  const int64_t r_end = min(p.R, r_begin + p.rows_per_z);
  const int KC = (int)ceil_div(r_end - r_begin, TC_BK);   // host guarantees >= 1

  if (warp > TC_MMA_WARP) {
    // ================= producers: warp pw owns rows 4*pw .. 4*pw+3 of every 32-row stage =================
    const int pw = warp - TC_MMA_WARP - 1;
    const int groups = 4 + (int)ceil_div(t.BN, 32);   // 4 column groups of X, then the Y tile
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % S;
      const int64_t r0 = r_begin + (int64_t)kc * TC_BK + 4 * pw;
      int si[4], ri[4];
      bool ok[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        ok[i] = r0 + i < r_end;
        si[i] = ri[i] = 0;
        if (YMODE == A_GATHER_DIFF && ok[i]) { si[i] = __ldg(p.Y.src + r0 + i); ri[i] = __ldg(p.Y.rev + r0 + i); }
      }
      uint8_t* sa = tiles + s * stage_bytes;
      uint8_t* sb = sa + TC_A_STAGE;
      bool waited = false;
      for (int g0 = 0; g0 < groups; g0 += 4) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int g = g0 + u;
          float x[4] = {0.f, 0.f, 0.f, 0.f};
          if (g < 4) {
            const int col = n1_0 + g * 32 + lane;
            if (col < p.N1) {
#pragma unroll
              for (int i = 0; i < 4; ++i) if (ok[i]) x[i] = __ldg(p.X + (r0 + i) * p.ldx + col);
            }
          } else if (g < groups) {
            const int n = (g - 4) * 32 + lane;
            if (n < t.width) {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                if (ok[i]) x[i] = tc_dw_y<YMODE, ACT, PREACT>(p, t, r0 + i, t.col0 + n, si[i], ri[i]);
            }
          }
          v[u] = make_float4(ptx::to_tf32(x[0]), ptx::to_tf32(x[1]), ptx::to_tf32(x[2]), ptx::to_tf32(x[3]));
        }
        if (!waited) { ptx::mbar_wait(&bar->empty[s], ((kc / S) & 1) ^ 1); waited = true; }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int g = g0 + u;
          if (g < 4) {
            *reinterpret_cast<float4*>(sa + ptx::sw128_off(g * 32 + lane, pw)) = v[u];
          } else if (g < groups) {
            const int n = (g - 4) * 32 + lane;
            if (n < t.BN) *reinterpret_cast<float4*>(sb + ptx::sw128_off(n, pw)) = v[u];
          }
        }
      }
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&bar->full[s]);
    }
  } else if (warp == TC_MMA_WARP) {
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % S;
      ptx::mbar_wait(&bar->full[s], (kc / S) & 1);
      ptx::tc_fence_after();
      if (lane == 0) {
        const uint32_t a_addr = ptx::smem_u32(tiles + s * stage_bytes);
        tc_issue_chunk(a_addr, a_addr + TC_A_STAGE, tmem_base, t.BN, kc == 0);
        ptx::tc_commit(&bar->empty[s]);
        if (kc == KC - 1) ptx::tc_commit(&bar->tmem_full[0]);
      }
      __syncwarp();
    }
  } else {
    ptx::mbar_wait(&bar->tmem_full[0], 0);
    ptx::tc_fence_after();
    float* Pz = p.P + (int64_t)blockIdx.y * p.N1 * p.ldp;
    tc_epilogue<E_STORE, ACT>(tmem_base, warp, lane, epi_st + warp * 32 * 33, t.BN, n1_0, p.N1, t.out_col0,
                              t.out_col0 + t.width, t.out_col0 + t.width, nullptr, nullptr, 0, Pz, p.ldp);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == TC_MMA_WARP) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

}  // namespace wd
