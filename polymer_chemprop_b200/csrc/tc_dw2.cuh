// tcgen05 weight-gradient contraction, second generation (see tc_gemm2.cuh for the motivation):
//   P[z][n1, n2] = sum_{r in chunk z} X[r, n1] * Y(r, n2),   Y(r, :) = Y0[i0(r), :] - act?(Y1[i1(r), :])
// A loaders stream raw ROWS (32 rows per stage: 128 columns of X, <= 160 columns of Y per source) into
// a shared-memory ring with zero-filling cp.async; transform warps transpose them (the reduction axis
// is the row axis) into the K-major SWIZZLE_128B operand stages, subtracting / rounding to TF32 on
// the way; one elected lane issues the MMAs; the epilogue writes this CTA's partial tile.
#pragma once
#include "tc_dw.cuh"
#include "tc_gemm2.cuh"

namespace wd {

constexpr int D2_MAX_BN = 160;                       // Y columns per CTA
constexpr int D2_X_RAW = TC_BK * TC_BM * 4;          // 16 KB: 32 rows x 128 columns
constexpr int D2_OP_STAGES = 2, D2_MAX_RAW_STAGES = 3;
// warp roles: 0-3 epilogue, 4 MMA, 5 idle, 6.. loaders, then transform warps (8: the transposing
// transform is the instruction-heaviest role; measured 0.248 -> 0.200 ms per launch going from 4 to 8)
#ifndef WD_D2_XFORM_WARPS
#define WD_D2_XFORM_WARPS 8
#endif
constexpr int D2_LOAD_WARPS = 4, D2_XFORM_WARPS = WD_D2_XFORM_WARPS;
constexpr int D2_LOAD_T0 = 32 * 6, D2_LOAD_THREADS = 32 * D2_LOAD_WARPS;
constexpr int D2_XFORM_W0 = 6 + D2_LOAD_WARPS;
constexpr int D2_THREADS = 32 * (D2_XFORM_W0 + D2_XFORM_WARPS);

struct D2Barriers {
  uint64_t raw_full[D2_MAX_RAW_STAGES], raw_empty[D2_MAX_RAW_STAGES], op_full[D2_OP_STAGES], op_empty[D2_OP_STAGES];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

struct D2Params {
  const float* X;       // [R, ldx], 16-byte aligned rows
  int ldx, N1;
  const float* Y0[2];   // per segment (CONCAT: 2 segments; else segment 0 only)
  const float* Y1;      // subtrahend of segment 0 (NSRC == 2) or nullptr
  const int32_t* i0;    // row indirection of Y0[0] / Y1 (segment 0 only)
  const int32_t* i1;
  int ldy[2], ldy1;
  int align16[2];       // segment rows 16-byte aligned
  float* P;
  int ldp;
  int64_t R, rows_per_z;
  int n1_tiles, n2_tiles, maxBN, raw_stages;
  TcDwTile t[TC_DW_MAX_TILES];
};

inline int d2_raw_stage_bytes(int maxBN, int nsrc) { return D2_X_RAW + nsrc * TC_BK * maxBN * 4; }
inline int d2_smem_bytes(int maxBN, int nsrc, int raw_stages) {
  return 1024 + D2_OP_STAGES * (TC_A_STAGE + maxBN * 128) + raw_stages * d2_raw_stage_bytes(maxBN, nsrc) +
         TC_EPI_STAGE_BYTES + (int)sizeof(D2Barriers) + 64;
}

template <int NSRC, int ACT, bool PREACT>
__global__ void __launch_bounds__(D2_THREADS, 1) k_tc_dw2(const D2Params p) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* opA = smem_raw + (((base + 1023u) & ~1023u) - base);
  const int b_stage = p.maxBN * 128, op_stage = TC_A_STAGE + b_stage;
  const int y_raw = TC_BK * p.maxBN * 4;                       // bytes of one source's raw Y rows
  const int raw_stage = D2_X_RAW + NSRC * y_raw;
  const int SR = p.raw_stages;
  uint8_t* raw = opA + D2_OP_STAGES * op_stage;
  float* epi_st = reinterpret_cast<float*>(raw + SR * raw_stage);
  D2Barriers* bar = reinterpret_cast<D2Barriers*>(raw + SR * raw_stage + TC_EPI_STAGE_BYTES);

  if (threadIdx.x == 0) {
    for (int s = 0; s < SR; ++s) { ptx::mbar_init(&bar->raw_full[s], D2_LOAD_THREADS); ptx::mbar_init(&bar->raw_empty[s], D2_XFORM_WARPS); }
    for (int s = 0; s < D2_OP_STAGES; ++s) { ptx::mbar_init(&bar->op_full[s], D2_XFORM_WARPS); ptx::mbar_init(&bar->op_empty[s], 1); }
    ptx::mbar_init(&bar->tmem_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == T2_MMA_WARP) ptx::tmem_alloc(&bar->tmem_base, TC_TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = bar->tmem_base;

  const int n1_0 = (int)(blockIdx.x % p.n1_tiles) * TC_BM;
  const TcDwTile tl = p.t[blockIdx.x / p.n1_tiles];
  const int64_t r_begin = (int64_t)blockIdx.y * p.rows_per_z;
  const int64_t r_end = min(p.R, r_begin + p.rows_per_z);
  const int KC = (int)ceil_div(r_end - r_begin, TC_BK);        // host guarantees >= 1
  const int ystride = p.maxBN * 4;                             // raw Y row stride in bytes

  if (warp >= D2_XFORM_W0) {
    // ================= transform: transpose raw rows into K-major swizzled operand stages =================
    const int tw = warp - D2_XFORM_W0;
    const int ygroups = (int)ceil_div(tl.BN, 32);
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % SR, os = kc % D2_OP_STAGES;
      ptx::mbar_wait(&bar->raw_full[s], (kc / SR) & 1);
      ptx::mbar_wait(&bar->op_empty[os], ((kc / D2_OP_STAGES) & 1) ^ 1);
      const uint8_t* rx = raw + s * raw_stage;
      const uint8_t* ry = rx + D2_X_RAW;
      uint8_t* oa = opA + os * op_stage;
      uint8_t* ob = oa + TC_A_STAGE;
      // X: 4 column groups x 8 row quads
      for (int q = tw; q < 32; q += D2_XFORM_WARPS) {
        const int m = (q >> 3) * 32 + lane, kq = q & 7;
        float x[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) x[i] = *reinterpret_cast<const float*>(rx + (4 * kq + i) * (TC_BM * 4) + m * 4);
        tc_store_piece(oa, m, kq, make_float4(x[0], x[1], x[2], x[3]));
      }
      // Y: ygroups column groups x 8 row quads
      for (int q = tw; q < ygroups * 8; q += D2_XFORM_WARPS) {
        const int n = (q >> 3) * 32 + lane, kq = q & 7;
        if (n < tl.BN) {
          float y[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint8_t* a = ry + (4 * kq + i) * ystride + n * 4;
            float v = *reinterpret_cast<const float*>(a);
            if (NSRC == 2) {
              float u = *reinterpret_cast<const float*>(a + y_raw);
              if (PREACT) u = act_fwd<ACT>(u);
              v -= u;
            }
            y[i] = v;
          }
          tc_store_piece(ob, n, kq, make_float4(y[0], y[1], y[2], y[3]));
        }
      }
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) { ptx::mbar_arrive(&bar->op_full[os]); ptx::mbar_arrive(&bar->raw_empty[s]); }
    }
  } else if (warp >= 6) {
    // ================= loaders: thread owns one of the stage's 32 rows (LPR threads per row) =================
    constexpr int LPR = D2_LOAD_THREADS / TC_BK;               // loader threads per row
    const int t = threadIdx.x - D2_LOAD_T0, rl = t / LPR, q4 = t % LPR;
    const uint32_t raw_u32 = ptx::smem_u32(raw);
    const int seg = tl.seg;
    const float* Y0 = p.Y0[seg];
    const int ldy = p.ldy[seg];
    const bool al = p.align16[seg] != 0;
    const bool indirect = (seg == 0) && p.i0 != nullptr;
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % SR;
      const int64_t r = r_begin + (int64_t)kc * TC_BK + rl;
      const bool ok = r < r_end;
      int64_t ra = ok ? r : 0, rb = ra;
      if (ok && indirect) { ra = __ldg(p.i0 + r); if (NSRC == 2) rb = __ldg(p.i1 + r); }
      ptx::mbar_wait(&bar->raw_empty[s], ((kc / SR) & 1) ^ 1);
      const uint32_t dx = raw_u32 + s * raw_stage + rl * (TC_BM * 4);
      const uint32_t dy = raw_u32 + s * raw_stage + D2_X_RAW + rl * ystride;
      // X: 32 pieces of 16 bytes per row
      const float* xrow = p.X + (ok ? r : 0) * p.ldx + n1_0;
#pragma unroll
      for (int i = 0; i < 32 / LPR; ++i) {
        const int c = q4 + LPR * i;
        const int rem = (p.N1 - (n1_0 + c * 4)) * 4;
        const uint32_t v = !ok ? 0u : (rem >= 16 ? 16u : (rem > 0 ? (uint32_t)rem : 0u));
        ptx::cp_async16(dx + c * 16, v ? xrow + c * 4 : p.X, v);
      }
      // Y: BN/4 pieces per row and source
      const float* y0row = Y0 + ra * ldy + tl.col0;
      if (al) {
        for (int c = q4; c < tl.BN / 4; c += LPR) {
          const int rem = (tl.width - c * 4) * 4;
          const uint32_t v = !ok ? 0u : (rem >= 16 ? 16u : (rem > 0 ? (uint32_t)rem : 0u));
          ptx::cp_async16(dy + c * 16, v ? y0row + c * 4 : Y0, v);
          if (NSRC == 2) ptx::cp_async16(dy + y_raw + c * 16, v ? p.Y1 + rb * p.ldy1 + tl.col0 + c * 4 : p.Y1, v);
        }
      } else {
        for (int c = q4; c < tl.BN; c += LPR) {
          const uint32_t v = (ok && c < tl.width) ? 4u : 0u;
          ptx::cp_async4(dy + c * 4, v ? y0row + c : Y0, v);
        }
      }
      ptx::cp_async_arrive(&bar->raw_full[s]);
    }
  } else if (warp == T2_MMA_WARP) {
    for (int kc = 0; kc < KC; ++kc) {
      const int os = kc % D2_OP_STAGES;
      ptx::mbar_wait(&bar->op_full[os], (kc / D2_OP_STAGES) & 1);
      ptx::tc_fence_after();
      if (lane == 0) {
        const uint32_t a_addr = ptx::smem_u32(opA + os * op_stage);
        tc_issue_chunk(a_addr, a_addr + TC_A_STAGE, tmem_base, tl.BN, kc == 0);
        ptx::tc_commit(&bar->op_empty[os]);
        if (kc == KC - 1) ptx::tc_commit(&bar->tmem_full);
      }
      __syncwarp();
    }
  } else if (warp < T2_EPI_WARPS) {
    ptx::mbar_wait(&bar->tmem_full, 0);
    ptx::tc_fence_after();
    float* Pz = p.P + (int64_t)blockIdx.y * p.N1 * p.ldp;
    tc_epilogue<E_STORE, ACT>(tmem_base, warp, lane, epi_st + warp * 32 * 33, tl.BN, n1_0, p.N1, tl.out_col0,
                              tl.out_col0 + tl.width, tl.out_col0 + tl.width, nullptr, nullptr, 0, Pz, p.ldp);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == T2_MMA_WARP) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

}  // namespace wd
