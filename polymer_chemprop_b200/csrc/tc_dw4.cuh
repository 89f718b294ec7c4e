// tcgen05 weight-gradient contraction with cp.async loaders: MN-major operands, no transpose.  The fallback of
// tc_dw5.cuh for Y rows that are only 4-byte aligned (dense 133 / 147-float feature rows).
//
//   P[z][n1, n2] = sum_{r in chunk z} X[r, n1] * Y[r, n2]
//
// The reduction runs over ROWS, so in the row-major activations the M / N index of both operands is the
// contiguous one: exactly the tensor core's MN-major operand form.  Generations 2 and 3 transposed every
// row chunk into K-major stages with scalar shared-memory reads; the wait-cycle counters (WD_T2_TIMING)
// showed those transform warps 95 % busy and everything else waiting on them.  Here
//   loaders    cp.async 16-byte pieces of raw rows straight into the MN-major operand stage
//              (SWIZZLE_128B_BASE32B: 8 rows x 128 bytes per 32-column block, blocks 1 KB apart),
//   transform  rounds to TF32 IN PLACE: pure 16-byte elementwise traffic, no address arithmetic, no second ring,
//   MMA        tcgen05.mma kind::tf32 with a_major = b_major = MN, up to three M=128 accumulators,
//   epilogue   the first four transform warps, once, after the last chunk.
// One CTA owns up to 384 X columns x one Y tile of <= 160 columns x one row chunk z (as generation 3).
#pragma once
#include "tc_core.cuh"

namespace wd {

constexpr int D4_BK = 16;                          // rows per stage: two K groups of 8
constexpr int D4_MT = 3, D4_MROWS = D4_MT * TC_BM, D4_ACC_COLS = 160, D4_MAX_BN = 160;
constexpr int D4_MAX_STAGES = 6;
constexpr int D4_MMA_WARP = 0, D4_LOAD_W0 = 1, D4_LOAD_WARPS = 4, D4_XFORM_W0 = D4_LOAD_W0 + D4_LOAD_WARPS, D4_XFORM_WARPS = 8;
constexpr int D4_LOAD_THREADS = 32 * D4_LOAD_WARPS, D4_XFORM_THREADS = 32 * D4_XFORM_WARPS;
constexpr int D4_THREADS = 32 * (D4_XFORM_W0 + D4_XFORM_WARPS);
static_assert(D4_LOAD_THREADS == 8 * D4_BK, "8 loader threads per stage row");

struct D4Barriers {
  uint64_t raw_full[D4_MAX_STAGES], op_full[D4_MAX_STAGES], empty[D4_MAX_STAGES];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

struct D4Params {
  const float* X;       // [R, ldx], 16-byte aligned rows
  int ldx, N1;
  const float* Y0[2];   // Y segments (CONCAT: two)
  int ldy[2];
  int align16[2];
  float* P;             // [Z][N1, ldp]
  int ldp;
  int64_t R, rows_per_z;
  int n1_groups;        // groups of <= 384 X columns
  int n2_tiles, stages;
  int XB, YB;           // 32-column blocks per 8-row group of the X / Y regions of a stage
  TcDwTile t[TC_DW_MAX_TILES];
};

inline int d4_stage_bytes(int XB, int YB) { return (2 * XB + 2 * YB) * 1024; }
inline int d4_smem_bytes(int XB, int YB, int stages) {
  return 1024 + stages * d4_stage_bytes(XB, YB) + TC_EPI_STAGE_BYTES + (int)sizeof(D4Barriers) + 64;
}

// MN-major descriptor for 32-bit operands.  The only MN-major layout the tensor core accepts for tf32 is
// SWIZZLE_128B_BASE32B (layout type 1; plain SWIZZLE_128B / no swizzle read as zeros -- scripts/mn_probe.cu
// decodes the mapping on the device): atoms of 4 K-rows x 128 bytes (32 MN elements), the 32-byte chunk
// index XORed with the row (address bits [5,7) ^= [7,9)).  LBO = distance of 32-element MN blocks,
// SBO = distance of the 4-row K atoms (512: the 8 rows of one MMA are contiguous here).
__device__ __forceinline__ uint64_t smem_desc_mn32(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)1 << 61;              // SWIZZLE_128B_BASE32B
  return d;
}
// the same descriptor as (low word, constant high word) for the convergent issue path (tc_ptx.cuh: mma_elect)
__device__ __forceinline__ uint32_t desc_lo_mn32(uint32_t smem_addr, uint32_t lbo_bytes) {
  return ((smem_addr & 0x3FFFFu) >> 4) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
constexpr uint32_t kDescHiMn32 = (512u >> 4) | (1u << 14) | (1u << 29);
__host__ __device__ inline uint32_t idesc_tf32_mn(int M, int N) { return ptx::idesc_tf32(M, N) | (1u << 15) | (1u << 16); }

// Store-only epilogue of one 128 x BN accumulator (partial tile): as tc_epilogue, but the write-out loop is
// batched (8 independent shared-memory reads, then 8 predicated stores) instead of one dependent chain per row.
__device__ __forceinline__ void d4_epilogue(uint32_t tmem_acc, int we, int lane, float* st /*[32][33]*/, int BN,
                                            int row0, int M, int col0, int cmax, float* C, int ldc) {
  for (int c0 = 0; c0 < BN; c0 += 32) {
    float v[32];
    const uint32_t ta = tmem_acc + ((uint32_t)(we * 32) << 16) + (uint32_t)c0;
    if (BN - c0 >= 32) {
      ptx::tmem_ld32(ta, v);
    } else {
      ptx::tmem_ld16(ta, v);
#pragma unroll
      for (int j = 16; j < 32; ++j) v[j] = 0.f;
    }
    ptx::tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) st[lane * 33 + j] = v[j];
    __syncwarp();
    const int col = col0 + c0 + lane;
    const bool in_c = col < cmax && c0 + lane < BN;
    float* cp = C + (int64_t)(row0 + we * 32) * ldc + col;
#pragma unroll
    for (int i0 = 0; i0 < 32; i0 += 8) {
      float x[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = st[(i0 + i) * 33 + lane];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (in_c && row0 + we * 32 + i0 + i < M) cp[(int64_t)(i0 + i) * ldc] = x[i];
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(D4_THREADS, 1) k_tc_dw4(const D4Params p) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* ring = smem_raw + (((base + 1023u) & ~1023u) - base);        // 1024-aligned: swizzle atoms
  const int XB = p.XB, YB = p.YB, SR = p.stages;
  const int x_bytes = 2 * XB * 1024, y_bytes = 2 * YB * 1024, stage_bytes = x_bytes + y_bytes;
  float* epi_st = reinterpret_cast<float*>(ring + SR * stage_bytes);
  D4Barriers* bar = reinterpret_cast<D4Barriers*>(ring + SR * stage_bytes + TC_EPI_STAGE_BYTES);
  long long tacc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};   // WD_T2_TIMING: 0 xform<-raw_full 2 loader<-empty 5 mma<-op_full
  const long long t_start = clock64();

  if (threadIdx.x == 0) {
    for (int s = 0; s < SR; ++s) {
      ptx::mbar_init(&bar->raw_full[s], D4_LOAD_THREADS);
      ptx::mbar_init(&bar->op_full[s], D4_XFORM_WARPS);
      ptx::mbar_init(&bar->empty[s], 1);
    }
    ptx::mbar_init(&bar->tmem_full, 1);
    ptx::fence_barrier_init();
  }
  // pieces no copy ever writes (column tails of the last block) must read as zero
  for (int i = threadIdx.x; i < SR * stage_bytes / 16; i += D4_THREADS)
    reinterpret_cast<float4*>(ring)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  ptx::fence_proxy_async_smem();
  if (warp == D4_MMA_WARP) ptx::tmem_alloc(&bar->tmem_base, TC_TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = bar->tmem_base;

  const int n1_0 = (int)(blockIdx.x % p.n1_groups) * D4_MROWS;
  const TcDwTile tl = p.t[blockIdx.x / p.n1_groups];
  const int n1_w = min(p.N1 - n1_0, D4_MROWS);                 // X columns of this CTA
  const int n_mt = (int)ceil_div(n1_w, TC_BM);                 // accumulators in use
  const int64_t r_begin = (int64_t)blockIdx.y * p.rows_per_z;
  const int64_t r_end = min(p.R, r_begin + p.rows_per_z);
  const int KC = (int)ceil_div(r_end - r_begin, D4_BK);        // host guarantees >= 1

  if (warp >= D4_XFORM_W0) {
    // ================= transform: round to TF32 in place (and fold the subtrahend source) =================
    const int t = threadIdx.x - 32 * D4_XFORM_W0;
    const int xp = 2 * XB * 64, yp = 2 * YB * 64;              // 16-byte pieces of the X / Y regions
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % SR;
      T2_WAIT(0, &bar->raw_full[s], (kc / SR) & 1);
      float4* sx = reinterpret_cast<float4*>(ring + s * stage_bytes);
      float4* sy = reinterpret_cast<float4*>(ring + s * stage_bytes + x_bytes);
      for (int i0 = t; i0 < xp; i0 += 4 * D4_XFORM_THREADS) {
        float4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) if (i0 + j * D4_XFORM_THREADS < xp) v[j] = sx[i0 + j * D4_XFORM_THREADS];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (i0 + j * D4_XFORM_THREADS < xp)
            sx[i0 + j * D4_XFORM_THREADS] = make_float4(ptx::to_tf32(v[j].x), ptx::to_tf32(v[j].y), ptx::to_tf32(v[j].z), ptx::to_tf32(v[j].w));
      }
      for (int i0 = t; i0 < yp; i0 += 2 * D4_XFORM_THREADS) {
        float4 v[2];
#pragma unroll
        for (int j = 0; j < 2; ++j)
          if (i0 + j * D4_XFORM_THREADS < yp) v[j] = sy[i0 + j * D4_XFORM_THREADS];
#pragma unroll
        for (int j = 0; j < 2; ++j)
          if (i0 + j * D4_XFORM_THREADS < yp) {
            sy[i0 + j * D4_XFORM_THREADS] = make_float4(ptx::to_tf32(v[j].x), ptx::to_tf32(v[j].y), ptx::to_tf32(v[j].z), ptx::to_tf32(v[j].w));
          }
      }
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&bar->op_full[s]);
    }
    // ================= epilogue: the first four transform warps, TMEM lane quarter = warp % 4 =================
    if (warp < D4_XFORM_W0 + 4) {
      T2_WAIT(7, &bar->tmem_full, 0);
      ptx::tc_fence_after();
      const int we = warp & 3;
      float* Pz = p.P + (int64_t)blockIdx.y * p.N1 * p.ldp;
      for (int mt = 0; mt < n_mt; ++mt)
        d4_epilogue(tmem_base + mt * D4_ACC_COLS, we, lane, epi_st + we * 32 * 33, tl.BN, n1_0 + mt * TC_BM, p.N1,
                    tl.out_col0, tl.out_col0 + tl.width, Pz, p.ldp);
    }
  } else if (warp >= D4_LOAD_W0) {
    // ================= loaders: 8 threads per row; thread q8 owns 16-byte chunk q8 of every 32-column block =================
    const int t = threadIdx.x - 32 * D4_LOAD_W0, rl = t >> 3, q8 = t & 7;
    const int kg = rl >> 3, k = rl & 7;
    const uint32_t ring_u32 = ptx::smem_u32(ring);
    const uint32_t piece = (uint32_t)(k * 128 + ((q8 ^ ((k & 3) << 1)) << 4));  // swizzled position inside the 8-row group
    const int seg = tl.seg;
    const float* Y0 = p.Y0[seg];
    const int ldy = p.ldy[seg];
    const bool al = p.align16[seg] != 0;
    const int xblocks = (int)ceil_div(n1_w, 32), yblocks = (int)ceil_div(tl.width, 32);
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % SR;
      const int64_t r = r_begin + (int64_t)kc * D4_BK + rl;
      const bool ok = r < r_end;
      int64_t ra = ok ? r : 0;
      const int64_t rx = (WD_ABL & 2) ? (r & 2047) : r;         // timing ablation: rows from an L2-resident window
      if (WD_ABL & 2) ra &= 2047;
      T2_WAIT(2, &bar->empty[s], ((kc / SR) & 1) ^ 1);
      if (WD_ABL & 32) { ptx::cp_async_arrive(&bar->raw_full[s]); continue; }
      const uint32_t dx = ring_u32 + s * stage_bytes + kg * XB * 1024 + piece;
      const uint32_t dy = ring_u32 + s * stage_bytes + x_bytes + kg * YB * 1024 + piece;
      const float* xrow = p.X + (ok ? rx : 0) * p.ldx + n1_0 + q8 * 4;
      for (int b = 0; b < xblocks; ++b) {
        const int rem = (n1_w - (b * 32 + q8 * 4)) * 4;
        if (rem <= 0) break;
        const uint32_t v = !ok ? 0u : (rem >= 16 ? 16u : (uint32_t)rem);
        ptx::cp_async16(dx + b * 1024, v ? xrow + b * 32 : p.X, v);
      }
      if (al) {
        const float* y0row = Y0 + ra * ldy + tl.col0 + q8 * 4;
        for (int b = 0; b < yblocks; ++b) {
          const int rem = (tl.width - (b * 32 + q8 * 4)) * 4;
          if (rem <= 0) break;
          const uint32_t v = !ok ? 0u : (rem >= 16 ? 16u : (uint32_t)rem);
          ptx::cp_async16(dy + b * 1024, v ? y0row + b * 32 : Y0, v);
        }
      } else {  // rows only 4-byte aligned (dense 133 / 147-float feature rows): 4-byte copies
        const float* y0row = Y0 + ra * ldy + tl.col0;
        const uint32_t dyr = ring_u32 + s * stage_bytes + x_bytes + kg * YB * 1024 + k * 128;
        for (int n = q8; n < tl.width; n += 8) {
          const uint32_t d = dyr + (n >> 5) * 1024 + ((((n & 31) >> 2) ^ ((k & 3) << 1)) << 4) + (n & 3) * 4;
          ptx::cp_async4(d, ok ? y0row + n : Y0, ok ? 4u : 0u);
        }
      }
      ptx::cp_async_arrive(&bar->raw_full[s]);
    }
  } else {
    // ================= MMA issue (warp 0) =================
    const uint32_t idesc = idesc_tf32_mn(128, tl.BN);
    const uint32_t ring_u32 = ptx::smem_u32(ring);
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % SR;
      T2_WAIT(5, &bar->op_full[s], (kc / SR) & 1);
      ptx::tc_fence_after();
      if (lane == 0) {
        const uint32_t xa = ring_u32 + s * stage_bytes, ya = xa + x_bytes;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const uint64_t bd = smem_desc_mn32(ya + g * YB * 1024, 1024u);
          const uint32_t acc = (kc == 0 && g == 0) ? 0u : 1u;
          for (int mt = 0; mt < n_mt; ++mt)
            ptx::mma_tf32(tmem_base + mt * D4_ACC_COLS,
                          smem_desc_mn32(xa + (g * XB + mt * 4) * 1024, 1024u), bd, idesc, acc);
        }
        ptx::tc_commit(&bar->empty[s]);
        if (kc == KC - 1) ptx::tc_commit(&bar->tmem_full);
      }
      __syncwarp();
    }
  }
#ifdef WD_T2_TIMING
  if (WD_TIMING_ON(14) && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0) {
    if (warp == D4_MMA_WARP || warp == D4_LOAD_W0 || warp == D4_XFORM_W0)
      for (int i = 0; i < 10; ++i) if (tacc[i]) atomicAdd(&g_t2_timing[i], (unsigned long long)tacc[i]);
    if (warp == D4_XFORM_W0) atomicAdd(&g_t2_timing[15], (unsigned long long)(clock64() - t_start));
  }
#else
  (void)tacc; (void)t_start;
#endif
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == D4_MMA_WARP) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

}  // namespace wd
