// Shared pieces of the tcgen05 (kind::tf32, fp32 accumulate in TMEM) contraction kernels: tile constants, the packed
// weight images, the K-chunk MMA issue helper, the dW tile descriptor and the per-role wait-cycle counters.
//
// Same virtual-operand contract as the SIMT kernels (operands.cuh): the gathered / concatenated A operand is never
// materialised in HBM.  K is walked in chunks of 32 tf32 (one 128-byte swizzle row); each chunk is 4 MMAs of K=8.
#pragma once
#include <cstdlib>

#include "operands.cuh"
#include "tc_ptx.cuh"

namespace wd {

constexpr int TC_BM = 128, TC_BK = 32;
constexpr int TC_MAX_BN = 304;                 // accumulator columns of one tile (<= 512 TMEM columns)
constexpr int TC_A_STAGE = TC_BM * 128;        // bytes of one A stage (128 rows x 32 tf32)
constexpr int TC_TMEM_COLS = 512;
constexpr int TC_EPI_WARPS = 4;
constexpr int TC_EPI_STAGE_BYTES = TC_EPI_WARPS * 32 * 33 * 4;   // padded transpose tiles of the dW epilogues
constexpr int TC_SMEM_LIMIT = 227 * 1024;

#ifndef WD_ABL
#define WD_ABL 0   // timing ablations (results become wrong): 2 rows from an L2-resident window, 4 no stores, 32 no A copies;
                   // forward kernel (profiles/r2_fwd_abl.log): 64 no epilogue work, 128 no MMAs, 512 no correction MMAs,
                   // 1024 no transform work, 2048 no proxy fence in the epilogue, 4096 no weight copies;
                   // 8192 (results stay RIGHT): 3xBF16 without the early refill of the raw-row part of a stage (A/B switch)
#endif

// Per-role wait-time counters of CTA 0 (build with -DWD_T2_TIMING; read through wd_debug_timing):
//  0 xform<-raw_full 1 xform<-a_empty 2 loader<-raw_empty 3 bload<-b_empty 4 mma<-tmem_empty 5 mma<-a_full
//  6 mma<-b_full 7 epi<-tmem_full 8 epi<-r_full 9 rload<-r_empty 15 kernel cycles
__device__ unsigned long long g_t2_timing[16];
#ifndef WD_T2_TIMING_ONLY
#define WD_T2_TIMING_ONLY 0          // 0: every instrumented kernel records; n: only kernel n (3, 4 gemm; 14, 15 dW)
#endif
#define WD_TIMING_ON(n) (WD_T2_TIMING_ONLY == 0 || WD_T2_TIMING_ONLY == (n))
#ifdef WD_T2_TIMING
#define T2_WAIT(slot, b, par) do { const long long t0_ = clock64(); ptx::mbar_wait(b, par); tacc[slot] += clock64() - t0_; } while (0)
#else
#define T2_WAIT(slot, b, par) ptx::mbar_wait(b, par)
#endif

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
inline TcTiling tc_tiling(int N, int max_bn) {
  const int np = (int)round_up(N, 16);
  const int nt = (int)ceil_div(np, std::min(tc_max_bn(), max_bn));
  return {(int)round_up(ceil_div(np, nt), 16), nt};
}

// ---------------------------------------------------------------------------------------------
// Weight packing: W (nn.Linear [out,in] or its transpose) -> per (n_tile, k_chunk) smem images,
// [BN rows x 128 bytes] in SWIZZLE_128B order, tf32-rounded (round to nearest), zero padded.  The K axis is the
// concatenation of two segments (K1, K2), each padded to a multiple of 32.
// parts == 2 (3xTF32, math mode TF32X3): every (n_tile, k_chunk) holds TWO images back to back: hi = tf32(w), and the
//   bf16 CORRECTION image whose 128-byte rows are [bf16(hi) (k = 0..31) | bf16(w - hi) (k = 0..31)] (tc_gemm4.cuh).
// bf == 1 (3xBF16, math mode BF16X3; parts == 1): ONE image per (n_tile, k_chunk) in the correction-image format with the
//   split point at bfloat16: rows [bf16(w) | bf16(w - bf16(w))].
// ---------------------------------------------------------------------------------------------
struct TcPackJob {
  const float* W;
  int ldb, bmode, N, K1, K2, BN, n_tiles, KC1, KC, parts;
  float* Bp;
  int64_t first;     // index of this job's first element in the launch (jobs are laid end to end)
  int bf;            // 1: the single image is the bf16 hi | lo image of the 3xBF16 mode
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
  const int ldb = j.ldb, bmode = j.bmode, N = j.N, K1 = j.K1, K2 = j.K2, BN = j.BN, KC1 = j.KC1, KC = j.KC, parts = j.parts;
  const int kk = (int)(e & 31);
  const int nl = (int)((e >> 5) % BN);
  int64_t q = e / (32 * (int64_t)BN);          // (nt * KC + kc) * parts + part
  const int part = (int)(q % parts);
  q /= parts;
  const int kc = (int)(q % KC);
  const int nt = (int)(q / KC);
  const int n = nt * BN + nl;
  int kglob = -1;
  if (kc < KC1) { int k = kc * 32 + kk; if (k < K1) kglob = k; }
  else          { int k = (kc - KC1) * 32 + kk; if (k < K2) kglob = K1 + k; }
  float v = 0.f;
  if (n < N && kglob >= 0) v = bmode == B_NK ? __ldg(W + (int64_t)n * ldb + kglob) : __ldg(W + (int64_t)kglob * ldb + n);
  const float hi = j.bf ? __uint_as_float(ptx::pack_bf16(v, 0.f) << 16) : ptx::to_tf32(v);
  const int64_t img = (((int64_t)nt * KC + kc) * parts + part) * BN * 32;
  if (part == 0 && !j.bf) {
    Bp[img + (ptx::sw128_off(nl, kk >> 2) >> 2) + (kk & 3)] = hi;
  } else {   // bf16 element j of the row sits in 16-byte chunk j / 8: j = kk (hi) and j = 32 + kk (lo)
    uint16_t* row = reinterpret_cast<uint16_t*>(Bp + img);
    row[(ptx::sw128_off(nl, kk >> 3) >> 1) + (kk & 7)] = (uint16_t)(ptx::pack_bf16(hi, 0.f) & 0xffffu);
    row[(ptx::sw128_off(nl, 4 + (kk >> 3)) >> 1) + (kk & 7)] = (uint16_t)(ptx::pack_bf16(v - hi, 0.f) & 0xffffu);
  }
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

// the same K chunk issued from convergent code (all lanes call it; the election is inside every statement: tc_ptx.cuh)
__device__ __forceinline__ void tc_issue_chunk_elect(uint32_t a_addr, uint32_t b_addr, uint32_t d_tmem, int BN, bool first_chunk) {
  const int N1 = BN > 256 ? 256 : BN, N2 = BN - N1;
  const uint32_t id1 = ptx::idesc_tf32(128, N1), id2 = ptx::idesc_tf32(128, N2);
  const uint32_t a_lo = ptx::desc_lo_sw128(a_addr), b_lo = ptx::desc_lo_sw128(b_addr), b2_lo = ptx::desc_lo_sw128(b_addr + N1 * 128);
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    const uint32_t acc = (first_chunk && ks == 0) ? 0u : 1u;
    ptx::mma_elect<false>(d_tmem, a_lo + 2 * ks, b_lo + 2 * ks, ptx::kDescHiSw128, id1, acc);
    if (N2 > 0) ptx::mma_elect<false>(d_tmem + N1, a_lo + 2 * ks, b2_lo + 2 * ks, ptx::kDescHiSw128, id2, acc);
  }
}

// One tile of the Y operand of a weight-gradient contraction (tc_dw4.cuh / tc_dw5.cuh).
constexpr int TC_DW_MAX_TILES = 24;
struct TcDwTile {
  int seg;       // 0: Y.X, 1: Y.X2 (second CONCAT segment)
  int col0;      // first column inside the segment
  int width;     // valid columns
  int BN;        // round_up(width, 16)
  int out_col0;  // first column in P rows
};

}  // namespace wd
