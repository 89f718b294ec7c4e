// tcgen05 (kind::tf32, fp32 accumulate in TMEM) contraction kernels of the encoder: math_mode TF32.
//
// Same virtual-operand contract as the SIMT kernels (operands.cuh): the gathered / concatenated A
// operand is never materialised in HBM.  Warp-specialised, persistent, one CTA per SM:
//   warps 0-3   epilogue   TMEM -> registers -> smem transpose -> coalesced global stores
//   warp  4     MMA issue  one elected lane issues tcgen05.mma, commits to mbarriers
//   warps 5-12  producers  gather A rows (float4, 8 lanes per 128-byte row chunk) into a
//                          SWIZZLE_128B K-major smem stage; one lane starts the bulk async copy
//                          (TMA engine) of the pre-swizzled weight chunk for the same stage
// K is walked in chunks of 32 tf32 (one 128-byte swizzle row); each chunk is 4 MMAs of K=8.
#pragma once
#include <cstdlib>

#include "operands.cuh"
#include "tc_ptx.cuh"

namespace wd {

constexpr int TC_BM = 128, TC_BK = 32;
constexpr int TC_EPI_WARPS = 4, TC_PROD_WARPS = 8, TC_MMA_WARP = 4;
constexpr int TC_THREADS = 32 * (TC_EPI_WARPS + 1 + TC_PROD_WARPS);
constexpr int TC_PROD_THREADS = 32 * TC_PROD_WARPS;
constexpr int TC_PROD_T0 = 32 * (TC_EPI_WARPS + 1);
constexpr int TC_MAX_BN = 304;                 // accumulator columns of one tile (<= 512 TMEM columns)
constexpr int TC_A_STAGE = TC_BM * 128;        // bytes of one A stage (128 rows x 32 tf32)
constexpr int TC_MAX_STAGES = 4;
constexpr int TC_TMEM_COLS = 512;
constexpr int TC_EPI_STAGE_BYTES = TC_EPI_WARPS * 32 * 33 * 4;
constexpr int TC_SMEM_LIMIT = 227 * 1024;

struct TcBarriers {
  uint64_t full[TC_MAX_STAGES], empty[TC_MAX_STAGES], tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

// N-tiling of an output dimension so that one tile's accumulator fits TMEM and its B stage smem.
struct TcTiling { int BN, n_tiles; };
inline int tc_max_bn() {   // WD_TC_MAX_BN: tuning knob (multiple of 16 in [16, 304])
  static int v = [] {
    const char* e = getenv("WD_TC_MAX_BN");
    int x = e ? atoi(e) : TC_MAX_BN;
    return (x >= 16 && x <= TC_MAX_BN && x % 16 == 0) ? x : TC_MAX_BN;
  }();
  return v;
}
inline TcTiling tc_tiling(int N) {
  int np = (int)round_up(N, 16);
  int nt = (int)ceil_div(np, tc_max_bn());
  int bn = (int)round_up(ceil_div(np, nt), 16);
  return {bn, nt};
}
inline int tc_stage_bytes(int BN) { return TC_A_STAGE + BN * 128; }
inline int tc_num_stages(int BN) {
  int s = (TC_SMEM_LIMIT - 2048 - TC_EPI_STAGE_BYTES - (int)sizeof(TcBarriers)) / tc_stage_bytes(BN);
  return s > TC_MAX_STAGES ? TC_MAX_STAGES : s;
}
inline int tc_smem_bytes(int BN, int stages) {
  return 1024 + stages * tc_stage_bytes(BN) + TC_EPI_STAGE_BYTES + (int)sizeof(TcBarriers) + 64;
}

// ---------------------------------------------------------------------------------------------
// Weight packing: W (nn.Linear [out,in] or its transpose) -> per (n_tile, k_chunk) smem images,
// [BN rows x 128 bytes] in SWIZZLE_128B order, tf32-rounded, zero padded.  The K axis is the
// concatenation of two segments (K1, K2), each padded to a multiple of 32.
// ---------------------------------------------------------------------------------------------
struct TcPackJob {
  const float* W;
  int ldb, bmode, N, K1, K2, BN, n_tiles, KC1, KC;
  float* Bp;
  int64_t first;     // index of this job's first element in the launch (jobs are laid end to end)
};
struct TcPackJobs {  // up to three weight matrices packed by ONE launch (forward: W_i, W_h, W_o; backward: W_h^T, W_o^T)
  TcPackJob job[3];
  int n;
  int64_t total;
};

__global__ void k_tc_pack_b(const TcPackJobs jobs) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= jobs.total) return;
  int ji = 0;
  if (jobs.n > 1 && e >= jobs.job[1].first) ji = 1;
  if (jobs.n > 2 && e >= jobs.job[2].first) ji = 2;
  const TcPackJob& j = jobs.job[ji];
  e -= j.first;
  const float* __restrict__ W = j.W;
  float* __restrict__ Bp = j.Bp;
  const int ldb = j.ldb, bmode = j.bmode, N = j.N, K1 = j.K1, K2 = j.K2, BN = j.BN, KC1 = j.KC1, KC = j.KC;
  int kk = (int)(e & 31);
  int nl = (int)((e >> 5) % BN);
  int kc = (int)((e / (32 * (int64_t)BN)) % KC);
  int nt = (int)(e / (32 * (int64_t)BN * KC));
  int n = nt * BN + nl;
  int kglob = -1;
  if (kc < KC1) { int k = kc * 32 + kk; if (k < K1) kglob = k; }
  else          { int k = (kc - KC1) * 32 + kk; if (k < K2) kglob = K1 + k; }
  float v = 0.f;
  if (n < N && kglob >= 0) v = bmode == B_NK ? __ldg(W + (int64_t)n * ldb + kglob) : __ldg(W + (int64_t)kglob * ldb + n);
  int64_t img = ((int64_t)nt * KC + kc) * BN * 32;
  Bp[img + (ptx::sw128_off(nl, kk >> 2) >> 2) + (kk & 3)] = ptx::to_tf32(v);
}

struct TcGemmParams {
  AOperand A;          // K1 = width of the first segment (== K for PLAIN / GATHER_DIFF)
  int K2;              // width of the second segment (CONCAT), else 0
  int vec0, vec1;      // segment rows are 16-byte aligned (float4 loads allowed)
  const float* Bp;     // packed weights (k_tc_pack_b)
  const float* bias;
  const float* R;      // residual [M, ldr] for E_RES_ACT
  int ldr;
  float* C;
  int ldc;
  int64_t M;
  int N, BN, n_tiles, KC1, KC, stages;
};

// one 16-byte piece (4 consecutive k) of a virtual A row
template <int AMODE, int ACT, bool PREACT>
__device__ __forceinline__ float4 tc_a_piece(const TcGemmParams& p, int kc, int c, int64_t r, int sr, int rr, bool ok) {
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (!ok) return v;
  if (AMODE == A_GATHER_DIFF) {
    int k = kc * 32 + c * 4;
    if (k < p.A.K1) {
      float4 a = ldg4(p.A.X + (int64_t)sr * p.A.ldx + k);
      float4 u = ldg4(p.A.X2 + (int64_t)rr * p.A.ldx2 + k);
      if (PREACT) u = f4_act<ACT>(u);
      v = make_float4(a.x - u.x, a.y - u.y, a.z - u.z, a.w - u.w);
    }
    return v;
  }
  const bool second = (AMODE == A_CONCAT) && kc >= p.KC1;
  const float* X = second ? p.A.X2 : p.A.X;
  const int ld = second ? p.A.ldx2 : p.A.ldx;
  const int Ks = second ? p.K2 : p.A.K1;
  const int k = (second ? kc - p.KC1 : kc) * 32 + c * 4;
  const float* row = X + r * ld;
  if (second ? p.vec1 : p.vec0) {
    if (k < Ks) v = ldg4(row + k);   // pad columns [Ks, ld) hold finite values; B is zero there
  } else {
    if (k + 0 < Ks) v.x = __ldg(row + k + 0);
    if (k + 1 < Ks) v.y = __ldg(row + k + 1);
    if (k + 2 < Ks) v.z = __ldg(row + k + 2);
    if (k + 3 < Ks) v.w = __ldg(row + k + 3);
  }
  return v;
}

__device__ __forceinline__ void tc_store_piece(uint8_t* stageA, int r_local, int c, float4 v) {
  v = make_float4(ptx::to_tf32(v.x), ptx::to_tf32(v.y), ptx::to_tf32(v.z), ptx::to_tf32(v.w));
  *reinterpret_cast<float4*>(stageA + ptx::sw128_off(r_local, c)) = v;
}

// issue the 4 K=8 MMAs of one 32-wide K chunk (N split in two instructions when BN > 256)
__device__ __forceinline__ void tc_issue_chunk(uint32_t a_addr, uint32_t b_addr, uint32_t d_tmem, int BN,
                                               bool first_chunk) {
  const int N1 = BN > 256 ? 256 : BN, N2 = BN - N1;
  const uint32_t id1 = ptx::idesc_tf32(128, N1), id2 = ptx::idesc_tf32(128, N2);
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    const uint32_t acc = (first_chunk && ks == 0) ? 0u : 1u;
    const uint64_t ad = ptx::smem_desc_sw128(a_addr + ks * 32);
    ptx::mma_tf32(d_tmem, ad, ptx::smem_desc_sw128(b_addr + ks * 32), id1, acc);
    if (N2 > 0) ptx::mma_tf32(d_tmem + N1, ad, ptx::smem_desc_sw128(b_addr + N1 * 128 + ks * 32), id2, acc);
  }
}

// Epilogue of one 128 x BN accumulator: warp `we` owns TMEM lanes [32*we, 32*we+32) == tile rows.
// 32-column chunks are transposed through a padded smem tile so that each global row segment is
// written (and the residual read) as one coalesced 128-byte access.
template <int EPI, int ACT>
__device__ __forceinline__ void tc_epilogue(uint32_t tmem_acc, int we, int lane, float* st /*[32][33]*/, int BN,
                                            int64_t row0, int64_t M, int col0, int N, int cmax,
                                            const float* bias, const float* R, int ldr, float* C, int ldc) {
  for (int c0 = 0; c0 < BN; c0 += 32) {
    const int col = col0 + c0 + lane;
    const bool in_n = col < N && c0 + lane < BN, in_c = col < cmax && c0 + lane < BN;
    // residual rows first: 32 independent coalesced loads in flight while TMEM is read
    float rv[32];
    if (EPI == E_RES_ACT) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int64_t r = row0 + we * 32 + i;
        rv[i] = (in_n && r < M) ? __ldg(R + r * ldr + col) : 0.f;
      }
    }
    const float bv = (bias != nullptr && in_n) ? __ldg(bias + col) : 0.f;
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
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const int64_t r = row0 + we * 32 + i;
      float x = st[i * 33 + lane] + bv;
      if (EPI == E_RES_ACT) x += rv[i];
      if (EPI != E_STORE) x = act_fwd<ACT>(x);
      if (r < M && in_c) C[r * ldc + col] = in_n ? x : 0.f;
    }
    __syncwarp();
  }
}

template <int AMODE, int EPI, int ACT, bool PREACT>
__global__ void __launch_bounds__(TC_THREADS, 1) k_tc_gemm(const TcGemmParams p) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* tiles = smem_raw + (((base + 1023u) & ~1023u) - base);
  const int stage_bytes = TC_A_STAGE + p.BN * 128;
  float* epi_st = reinterpret_cast<float*>(tiles + p.stages * stage_bytes);
  TcBarriers* bar = reinterpret_cast<TcBarriers*>(tiles + p.stages * stage_bytes + TC_EPI_STAGE_BYTES);
  const int S = p.stages;
  const int nacc = p.BN <= 256 ? 2 : 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { ptx::mbar_init(&bar->full[s], TC_PROD_WARPS + 1); ptx::mbar_init(&bar->empty[s], 1); }
    for (int a = 0; a < 2; ++a) { ptx::mbar_init(&bar->tmem_full[a], 1); ptx::mbar_init(&bar->tmem_empty[a], TC_EPI_WARPS); }
    ptx::fence_barrier_init();
  }
  if (warp == TC_MMA_WARP) ptx::tmem_alloc(&bar->tmem_base, TC_TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = bar->tmem_base;

  const int64_t m_tiles = ceil_div(p.M, TC_BM);
  const int64_t total = m_tiles * p.n_tiles;

  if (warp > TC_MMA_WARP) {
    // ================= producers =================
    const int pt = threadIdx.x - TC_PROD_T0, c = pt & 7, rb = pt >> 3;
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x) {
      const int64_t row0 = (tile / p.n_tiles) * TC_BM;
      const int nt = (int)(tile % p.n_tiles);
      int sr[4], rr[4];
      bool ok[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t r = row0 + rb + 32 * j;
        ok[j] = r < p.M;
        sr[j] = rr[j] = 0;
        if (AMODE == A_GATHER_DIFF && ok[j]) { sr[j] = __ldg(p.A.src + r); rr[j] = __ldg(p.A.rev + r); }
      }
      float4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = tc_a_piece<AMODE, ACT, PREACT>(p, 0, c, row0 + rb + 32 * j, sr[j], rr[j], ok[j]);
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int s = it % S;
        const uint32_t ph = (it / S) & 1;
        ptx::mbar_wait(&bar->empty[s], ph ^ 1);
        uint8_t* sa = tiles + s * stage_bytes;
        if (pt == 0) {
          const uint32_t bytes = (uint32_t)p.BN * 128u;
          ptx::mbar_arrive_expect_tx(&bar->full[s], bytes);
          ptx::bulk_g2s(sa + TC_A_STAGE, p.Bp + ((int64_t)nt * p.KC + kc) * p.BN * 32, bytes, &bar->full[s]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) tc_store_piece(sa, rb + 32 * j, c, v[j]);
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bar->full[s]);
        if (kc + 1 < p.KC) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            v[j] = tc_a_piece<AMODE, ACT, PREACT>(p, kc + 1, c, row0 + rb + 32 * j, sr[j], rr[j], ok[j]);
        }
      }
    }
  } else if (warp == TC_MMA_WARP) {
    // ================= MMA issue =================
    uint32_t it = 0, tl = 0;
    for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x, ++tl) {
      const int acc = tl % nacc;
      const uint32_t aph = (tl / nacc) & 1;
      ptx::mbar_wait(&bar->tmem_empty[acc], aph ^ 1);
      ptx::tc_fence_after();
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int s = it % S;
        ptx::mbar_wait(&bar->full[s], (it / S) & 1);
        ptx::tc_fence_after();
        if (lane == 0) {
          const uint32_t a_addr = ptx::smem_u32(tiles + s * stage_bytes);
          tc_issue_chunk(a_addr, a_addr + TC_A_STAGE, tmem_base + acc * 256, p.BN, kc == 0);
          ptx::tc_commit(&bar->empty[s]);
          if (kc == p.KC - 1) ptx::tc_commit(&bar->tmem_full[acc]);
        }
        __syncwarp();
      }
    }
  } else {
    // ================= epilogue =================
    uint32_t tl = 0;
    float* st = epi_st + warp * 32 * 33;
    for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x, ++tl) {
      const int acc = tl % nacc;
      const int64_t row0 = (tile / p.n_tiles) * TC_BM;
      const int col0 = (int)(tile % p.n_tiles) * p.BN;
      ptx::mbar_wait(&bar->tmem_full[acc], (tl / nacc) & 1);
      ptx::tc_fence_after();
      tc_epilogue<EPI, ACT>(tmem_base + acc * 256, warp, lane, st, p.BN, row0, p.M, col0, p.N, p.ldc, p.bias, p.R,
                            p.ldr, p.C, p.ldc);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&bar->tmem_empty[acc]);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == TC_MMA_WARP) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

}  // namespace wd
