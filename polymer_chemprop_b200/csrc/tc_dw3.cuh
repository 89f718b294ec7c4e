// tcgen05 weight-gradient contraction, third generation: L2-traffic-minimal tiling.
//
//   P[z][n1, n2] = sum_{r in chunk z} X[r, n1] * Y(r, n2),   Y(r, :) = Y0[i0(r), :] - act?(Y1[i1(r), :])
//
// ncu showed the second generation (tc_dw2.cuh) bound by L2 -> SM bandwidth (7.8 TB/s of a ~12 TB/s
// cap): with one 128-column X tile per CTA every Y row was fetched by 3 CTAs and every X row by 2.
// Here one CTA owns UP TO 384 X columns (three M=128 accumulators side by side in TMEM, 3 x 160 = 480
// columns) x one Y tile of <= 160 columns, so X rows are fetched once per Y tile and Y rows once:
// 156 KB instead of 336 KB of L2 reads per 32 rows at h = 300.
// To make that fit shared memory the reduction axis is walked in chunks of 16 rows and the operand
// stages use the 64-byte swizzle (rows of 16 tf32 = 64 bytes, SWIZZLE_64B K-major).
// Warp roles / pipeline as tc_dw2.cuh: loaders (cp.async raw rows) -> transform (transpose, subtract,
// round to TF32) -> MMA issue -> epilogue (one partial tile per accumulator).
#pragma once
#include "tc_dw2.cuh"

namespace wd {

constexpr int D3_BK = 16;                          // rows per chunk
constexpr int D3_MT = 3;                           // accumulators (M = 128 each) per CTA
constexpr int D3_MROWS = D3_MT * TC_BM;            // 384 operand rows (X columns), zero beyond N1
constexpr int D3_ACC_COLS = 160;                   // TMEM columns per accumulator
constexpr int D3_MAX_BN = 160;
constexpr int D3_A_STAGE = D3_MROWS * 64;          // 24 KB
constexpr int D3_OP_STAGES = 2, D3_MAX_RAW_STAGES = 4;
constexpr int D3_LOAD_WARPS = 4, D3_XFORM_WARPS = 8;
constexpr int D3_LOAD_T0 = 32 * 6, D3_LOAD_THREADS = 32 * D3_LOAD_WARPS;
constexpr int D3_XFORM_W0 = 6 + D3_LOAD_WARPS;
constexpr int D3_THREADS = 32 * (D3_XFORM_W0 + D3_XFORM_WARPS);

struct D3Barriers {
  uint64_t raw_full[D3_MAX_RAW_STAGES], raw_empty[D3_MAX_RAW_STAGES], op_full[D3_OP_STAGES], op_empty[D3_OP_STAGES];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

struct D3Params {
  const float* X;       // [R, ldx], 16-byte aligned rows
  int ldx, N1;
  const float* Y0[2];
  const float* Y1;
  const int32_t* i0;
  const int32_t* i1;
  int ldy[2], ldy1;
  int align16[2];
  float* P;             // [Z][N1, ldp]
  int ldp;
  int64_t R, rows_per_z;
  int n1_groups;        // groups of <= 384 X columns
  int n2_tiles, maxBN, raw_stages;
  int xw;               // raw X row width in floats: round_up(min(N1, 384), 32)
  TcDwTile t[TC_DW_MAX_TILES];
};

inline int d3_raw_stage_bytes(int xw, int maxBN, int nsrc) { return D3_BK * xw * 4 + nsrc * D3_BK * maxBN * 4; }
inline int d3_smem_bytes(int xw, int maxBN, int nsrc, int raw_stages) {
  return 1024 + D3_OP_STAGES * (D3_A_STAGE + maxBN * 64) + raw_stages * d3_raw_stage_bytes(xw, maxBN, nsrc) +
         TC_EPI_STAGE_BYTES + (int)sizeof(D3Barriers) + 64;
}

// byte offset of 16-byte chunk c (0..3) of row r in a SWIZZLE_64B K-major tile (64-byte rows):
// Swizzle<2,4,3>: address bits [4,6) ^= bits [7,9)  ==  chunk ^= (r >> 1) & 3
__device__ __forceinline__ uint32_t sw64_off(int r, int c) { return (uint32_t)r * 64u + (uint32_t)((c ^ ((r >> 1) & 3)) << 4); }

// smem descriptor, K-major SWIZZLE_64B: 8-row groups 512 bytes apart
__device__ __forceinline__ uint64_t smem_desc_sw64(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)4 << 61;              // SWIZZLE_64B
  return d;
}

__device__ __forceinline__ void d3_store_piece(uint8_t* stage, int r, int c, float4 v) {
  v = make_float4(ptx::to_tf32(v.x), ptx::to_tf32(v.y), ptx::to_tf32(v.z), ptx::to_tf32(v.w));
  *reinterpret_cast<float4*>(stage + sw64_off(r, c)) = v;
}

template <int NSRC, int ACT, bool PREACT>
__global__ void __launch_bounds__(D3_THREADS, 1) k_tc_dw3(const D3Params p) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* opA = smem_raw + (((base + 1023u) & ~1023u) - base);
  const int b_stage = p.maxBN * 64, op_stage = D3_A_STAGE + b_stage;
  const int x_raw = D3_BK * p.xw * 4;                          // bytes of the raw X rows of one stage
  const int y_raw = D3_BK * p.maxBN * 4;                       // bytes of one source's raw Y rows
  const int raw_stage = x_raw + NSRC * y_raw;
  const int SR = p.raw_stages;
  uint8_t* raw = opA + D3_OP_STAGES * op_stage;
  float* epi_st = reinterpret_cast<float*>(raw + SR * raw_stage);
  D3Barriers* bar = reinterpret_cast<D3Barriers*>(raw + SR * raw_stage + TC_EPI_STAGE_BYTES);
  long long tacc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};   // WD_T2_TIMING: 0 xform<-raw_full 1 xform<-op_empty 2 loader<-raw_empty
  const long long t_start = clock64();                   //               5 mma<-op_full 7 epi<-tmem_full

  if (threadIdx.x == 0) {
    for (int s = 0; s < SR; ++s) { ptx::mbar_init(&bar->raw_full[s], D3_LOAD_THREADS); ptx::mbar_init(&bar->raw_empty[s], D3_XFORM_WARPS); }
    for (int s = 0; s < D3_OP_STAGES; ++s) { ptx::mbar_init(&bar->op_full[s], D3_XFORM_WARPS); ptx::mbar_init(&bar->op_empty[s], 1); }
    ptx::mbar_init(&bar->tmem_full, 1);
    ptx::fence_barrier_init();
  }
  // operand rows that no transform task ever writes (X columns in [xw, 384)) must read as zero
  for (int i = threadIdx.x; i < D3_OP_STAGES * op_stage / 16; i += D3_THREADS)
    reinterpret_cast<float4*>(opA)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  ptx::fence_proxy_async_smem();
  if (warp == T2_MMA_WARP) ptx::tmem_alloc(&bar->tmem_base, TC_TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = bar->tmem_base;

  const int n1_0 = (int)(blockIdx.x % p.n1_groups) * D3_MROWS;
  const TcDwTile tl = p.t[blockIdx.x / p.n1_groups];
  const int n1_end = min(p.N1, n1_0 + D3_MROWS);               // X columns of this CTA: [n1_0, n1_end)
  const int n_mt = (int)ceil_div(n1_end - n1_0, TC_BM);         // accumulators in use
  const int64_t r_begin = (int64_t)blockIdx.y * p.rows_per_z;
  const int64_t r_end = min(p.R, r_begin + p.rows_per_z);
  const int KC = (int)ceil_div(r_end - r_begin, D3_BK);        // host guarantees >= 1
  const int xstride = p.xw * 4, ystride = p.maxBN * 4;         // raw row strides in bytes

  if (warp >= D3_XFORM_W0) {
    // ================= transform: transpose raw rows into K-major SWIZZLE_64B operand stages =================
    const int tw = warp - D3_XFORM_W0;
    const int xtasks = (p.xw / 32) * 4, ytasks = (int)ceil_div(tl.BN, 32) * 4;
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % SR, os = kc % D3_OP_STAGES;
      T2_WAIT(0, &bar->raw_full[s], (kc / SR) & 1);
      T2_WAIT(1, &bar->op_empty[os], ((kc / D3_OP_STAGES) & 1) ^ 1);
      const uint8_t* rx = raw + s * raw_stage;
      const uint8_t* ry = rx + x_raw;
      uint8_t* oa = opA + os * op_stage;
      uint8_t* ob = oa + D3_A_STAGE;
      for (int q = tw; q < xtasks + ytasks; q += D3_XFORM_WARPS) {
        if (q < xtasks) {
          const int m = (q >> 2) * 32 + lane, kq = q & 3;
          float x[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) x[i] = *reinterpret_cast<const float*>(rx + (4 * kq + i) * xstride + m * 4);
          d3_store_piece(oa, m, kq, make_float4(x[0], x[1], x[2], x[3]));
        } else {
          const int qq = q - xtasks;
          const int n = (qq >> 2) * 32 + lane, kq = qq & 3;
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
            d3_store_piece(ob, n, kq, make_float4(y[0], y[1], y[2], y[3]));
          }
        }
      }
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) { ptx::mbar_arrive(&bar->op_full[os]); ptx::mbar_arrive(&bar->raw_empty[s]); }
    }
  } else if (warp >= 6) {
    // ================= loaders: 8 threads per row of the 16-row stage =================
    constexpr int LPR = D3_LOAD_THREADS / D3_BK;
    const int t = threadIdx.x - D3_LOAD_T0, rl = t / LPR, q8 = t % LPR;
    const uint32_t raw_u32 = ptx::smem_u32(raw);
    const int seg = tl.seg;
    const float* Y0 = p.Y0[seg];
    const int ldy = p.ldy[seg];
    const bool al = p.align16[seg] != 0;
    const bool indirect = (seg == 0) && p.i0 != nullptr;
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % SR;
      const int64_t r = r_begin + (int64_t)kc * D3_BK + rl;
      const bool ok = r < r_end;
      int64_t ra = ok ? r : 0, rb = ra;
      if (ok && indirect) { ra = __ldg(p.i0 + r); if (NSRC == 2) rb = __ldg(p.i1 + r); }
      T2_WAIT(2, &bar->raw_empty[s], ((kc / SR) & 1) ^ 1);
      const uint32_t dx = raw_u32 + s * raw_stage + rl * xstride;
      const uint32_t dy = raw_u32 + s * raw_stage + x_raw + rl * ystride;
      const float* xrow = p.X + (ok ? r : 0) * p.ldx + n1_0;
      for (int c = q8; c < p.xw / 4; c += LPR) {
        const int rem = (n1_end - (n1_0 + c * 4)) * 4;
        const uint32_t v = !ok ? 0u : (rem >= 16 ? 16u : (rem > 0 ? (uint32_t)rem : 0u));
        ptx::cp_async16(dx + c * 16, v ? xrow + c * 4 : p.X, v);
      }
      const float* y0row = Y0 + ra * ldy + tl.col0;
      if (al) {
        for (int c = q8; c < tl.BN / 4; c += LPR) {
          const int rem = (tl.width - c * 4) * 4;
          const uint32_t v = !ok ? 0u : (rem >= 16 ? 16u : (rem > 0 ? (uint32_t)rem : 0u));
          ptx::cp_async16(dy + c * 16, v ? y0row + c * 4 : Y0, v);
          if (NSRC == 2) ptx::cp_async16(dy + y_raw + c * 16, v ? p.Y1 + rb * p.ldy1 + tl.col0 + c * 4 : p.Y1, v);
        }
      } else {
        for (int c = q8; c < tl.BN; c += LPR) {
          const uint32_t v = (ok && c < tl.width) ? 4u : 0u;
          ptx::cp_async4(dy + c * 4, v ? y0row + c : Y0, v);
        }
      }
      ptx::cp_async_arrive(&bar->raw_full[s]);
    }
  } else if (warp == T2_MMA_WARP) {
    const uint32_t idesc = ptx::idesc_tf32(128, tl.BN);
    for (int kc = 0; kc < KC; ++kc) {
      const int os = kc % D3_OP_STAGES;
      T2_WAIT(5, &bar->op_full[os], (kc / D3_OP_STAGES) & 1);
      ptx::tc_fence_after();
      if (lane == 0) {
        const uint32_t a_addr = ptx::smem_u32(opA + os * op_stage), b_addr = a_addr + D3_A_STAGE;
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
          const uint64_t bd = smem_desc_sw64(b_addr + ks * 32);
          const uint32_t acc = (kc == 0 && ks == 0) ? 0u : 1u;
          for (int mt = 0; mt < n_mt; ++mt)
            ptx::mma_tf32(tmem_base + mt * D3_ACC_COLS, smem_desc_sw64(a_addr + mt * TC_BM * 64 + ks * 32), bd, idesc, acc);
        }
        ptx::tc_commit(&bar->op_empty[os]);
        if (kc == KC - 1) ptx::tc_commit(&bar->tmem_full);
      }
      __syncwarp();
    }
  } else if (warp < T2_EPI_WARPS) {
    T2_WAIT(7, &bar->tmem_full, 0);
    ptx::tc_fence_after();
    float* Pz = p.P + (int64_t)blockIdx.y * p.N1 * p.ldp;
    for (int mt = 0; mt < n_mt; ++mt)
      tc_epilogue<E_STORE, ACT>(tmem_base + mt * D3_ACC_COLS, warp, lane, epi_st + warp * 32 * 33, tl.BN,
                                n1_0 + mt * TC_BM, p.N1, tl.out_col0, tl.out_col0 + tl.width, tl.out_col0 + tl.width,
                                nullptr, nullptr, 0, Pz, p.ldp);
  }
#ifdef WD_T2_TIMING
  if (WD_TIMING_ON(13) && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0) {
    if (warp == 0 || warp == T2_MMA_WARP || warp == 6 || warp == D3_XFORM_W0)
      for (int i = 0; i < 10; ++i) if (tacc[i]) atomicAdd(&g_t2_timing[i], (unsigned long long)tacc[i]);
    if (warp == 0) atomicAdd(&g_t2_timing[15], (unsigned long long)(clock64() - t_start));
  }
#else
  (void)tacc; (void)t_start;
#endif
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == T2_MMA_WARP) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

}  // namespace wd
