// tcgen05 contraction kernel for PLAIN / CONCAT A operands (no gather), single-pass TF32:
//
//   C[M, ldc] = epi( A * B^T ),  A = [X0 | X1] row-major matrices with 16-byte aligned rows
//
// An earlier generation moved every A byte with cp.async (LDGSTS: measured ~12 bytes / cycle / SM, the binding
// limit once the epilogue was fixed) and then through a transform pass (round to TF32, swizzle).  A plain
// matrix needs neither: a TMA tensor copy (box 128 rows x 32 columns, SWIZZLE_128B) writes the K-major operand
// stage directly and its TFLOAT32 data type rounds fp32 -> tf32 in flight.  One producer lane issues the A box
// and the bulk copy of the packed weight chunk on one mbarrier; tcgen05.mma consumes it; the epilogue
// overlaps the next tile through the second TMEM accumulator.
//   warps 0-3 epilogue, warp 4 MMA issue, warp 5 producer.
// Used by the backward pass (dS = dZ W_h, dA = dZo W_o[:, F_a:]) in every tensor-core mode and by the forward input
// projection / atom update in the single-pass TF32 mode (the 3xTF32 mode needs the raw fp32 operand: tc_gemm4.cuh).
#pragma once
#include "tc_core.cuh"
#include "tc_tma.cuh"

namespace wd {

constexpr int T2_EPI_WARPS = 4, T2_MMA_WARP = 4;
constexpr int T2_MAX_BN = 160;                    // two accumulators in TMEM: epilogue overlaps the next tile
constexpr int T2_R_SLOTS = 3, T2_R_ROW = 144;     // transpose staging: 128 rows x 32 floats, rows padded to 144 B
constexpr int T2_R_SLOT_BYTES = TC_BM * T2_R_ROW;
constexpr int T2_R_BYTES = T2_R_SLOTS * T2_R_SLOT_BYTES;

constexpr int T3_STAGES = 4;
constexpr int T3_PROD_WARP = 5;
// optional second epilogue group (warps 6-9, TMEM lane quarter = warp % 4; see tc_gemm4.cuh): measured no gain for this
// kernel (dX 117.9 us either way), so the default stays one group
#ifndef WD_T3_EPI_GROUPS
#define WD_T3_EPI_GROUPS 1
#endif
constexpr int T3_EPI_GROUPS = WD_T3_EPI_GROUPS, T3_EPI2_W0 = 6;
constexpr int T3_THREADS = 32 * (6 + (T3_EPI_GROUPS - 1) * 4);

struct T3Barriers {
  uint64_t full[T3_STAGES], empty[T3_STAGES];
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

struct T3Params {
  CUtensorMap ma[2];    // K segments of A (box 128 x 32, SWIZZLE_128B, TFLOAT32)
  int KC0, KC;          // chunks of segment 0 / total
  const float* Bp;
  const float* bias;
  float* C;
  int ldc;
  int64_t M;
  int N, BN, n_tiles;
  Dropout drop;         // applied after the activation (EPI == E_ACT)
  SignBits sign;        // (stored value > 0) bits, written next to C when sign.p != nullptr
};

inline int t3_smem_bytes(int BN) {
  return 1024 + T3_STAGES * (TC_A_STAGE + BN * 128) + T2_R_BYTES + (int)sizeof(T3Barriers) + 64;
}

// Epilogue of one 128 x BN accumulator, 32 columns at a time.  The row-owning thread parks its accumulator row in
// a staging slot (conflict-free 16-byte accesses: rows are 144 bytes apart), then the warp writes the 32 rows out
// with 8 lanes per row (16 bytes each, 4 rows per instruction -> full 128-byte lines), applying bias + activation.
template <int EPI, int ACT>
__device__ __forceinline__ void t3_epilogue(uint32_t tmem_acc, int we, int lane, uint8_t* rring, uint32_t& rit, int BN,
                                            int64_t row0, int64_t M, int col0, int N, int cmax, const float* bias,
                                            float* C, int ldc, const Dropout& drop, const SignBits& sign, int grp,
                                            uint64_t* tmem_empty) {
  uint32_t sw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};   // sign words of this lane's chunk (common.cuh: sign_collect)
  const int rows_ok = (int)min((int64_t)32, M - (row0 + we * 32));         // rows of this warp inside the matrix (may be <= 0)
  const uint32_t rit0 = rit;
  // the accumulator goes back to the MMA warp right after this warp's last tcgen05.ld of the tile (tc_gemm4.cuh: t4_epilogue)
  auto release = [&]() {
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(tmem_empty);
  };
  const uint32_t nch = (uint32_t)((BN + 31) >> 5);
  uint32_t last = rit0 + nch - 1u;
  bool has_chunk = true;
  if (T3_EPI_GROUPS > 1 && (int)(last & 1u) != grp) { has_chunk = nch >= 2u; last -= 1u; }
  if (!has_chunk) release();
  for (int c0 = 0; c0 < BN; c0 += 32, ++rit) {
    if (T3_EPI_GROUPS > 1 && (int)(rit & 1u) != grp) continue;             // the other group's chunk
    const int slot = T3_EPI_GROUPS > 1 ? grp : (int)(rit % T2_R_SLOTS);     // staging only: one slot per group
    uint8_t* sl = rring + slot * T2_R_SLOT_BYTES + (we * 32) * T2_R_ROW;     // this warp's 32 rows
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
    if (has_chunk && rit == last) release();
    float4* myrow = reinterpret_cast<float4*>(sl + lane * T2_R_ROW);
#pragma unroll
    for (int j = 0; j < 8; ++j) myrow[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    __syncwarp();
    const int pc = lane & 7, rq = lane >> 3;
    const int colb = col0 + c0 + pc * 4;
    const int nval = (c0 + pc * 4 < BN) ? max(0, min(4, N - colb)) : 0;      // live columns of this piece
    const bool in_c = colb < cmax && c0 + pc * 4 < BN;
    const bool edge = c0 + 32 > BN || col0 + c0 + 32 > N;                   // warp-uniform: some lane has masked columns
    float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (bias != nullptr) {
      if (nval > 0) bv.x = __ldg(bias + colb);
      if (nval > 1) bv.y = __ldg(bias + colb + 1);
      if (nval > 2) bv.z = __ldg(bias + colb + 2);
      if (nval > 3) bv.w = __ldg(bias + colb + 3);
    }
    float4 xv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) xv[i] = *reinterpret_cast<const float4*>(sl + (rq + 4 * i) * T2_R_ROW + pc * 16);
    float* cp = C + (row0 + we * 32 + rq) * ldc + colb;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float4 x = xv[i];
      if (bias != nullptr) x = make_float4(x.x + bv.x, x.y + bv.y, x.z + bv.z, x.w + bv.w);       // warp-uniform
      if (EPI != E_STORE) x = f4_act<ACT>(x);
      if (EPI != E_STORE && drop.thresh != 0u) x = dropout4(x, drop, (row0 + we * 32 + rq + 4 * i) * ldc + colb);
      if (edge && nval < 4) { x.w = 0.f; if (nval < 3) x.z = 0.f; if (nval < 2) x.y = 0.f; if (nval < 1) x.x = 0.f; }
      const bool ok = in_c && rq + 4 * i < rows_ok && !(WD_ABL & 4);
      if (ok) *reinterpret_cast<float4*>(cp + (int64_t)(4 * i) * ldc) = x;
      if (sign.p != nullptr) sign_collect(sw[i], x, rq, pc, c0 >> 5);   // warp-uniform branch
    }
    __syncwarp();
  }
  // lane pc holds the word of chunk pc: with two groups only if this group wrote it
  if (sign.p != nullptr && (T3_EPI_GROUPS == 1 || (int)((rit0 + (uint32_t)(lane & 7)) & 1u) == grp))
    sign_flush(sign, sw, col0 / BN, row0 + we * 32, M, lane >> 3, lane & 7, (BN + 31) >> 5);
}

template <int EPI, int ACT>
__global__ void __launch_bounds__(T3_THREADS, 1) k_tc_gemm3(const __grid_constant__ T3Params p) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  static_assert(EPI != E_RES_ACT, "the residual epilogue belongs to the gathered kernel");
  extern __shared__ uint8_t smem_raw[];
  const int warp = ptx::uniform_warp_idx(), lane = threadIdx.x & 31;   // provably warp-uniform: see tc_ptx.cuh
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* ops = smem_raw + (((base + 1023u) & ~1023u) - base);        // 1024-aligned
  const int b_stage = p.BN * 128, stage = TC_A_STAGE + b_stage;          // multiples of 1024 (BN % 16 == 0 -> 2048)
  uint8_t* rring = ops + T3_STAGES * stage;
  T3Barriers* bar = reinterpret_cast<T3Barriers*>(rring + T2_R_BYTES);
  long long tacc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};   // WD_T2_TIMING: 3 producer<-empty 4 mma<-tmem_empty 5 mma<-full 7 epi<-tmem_full
  const long long t_start = clock64();

  if (threadIdx.x == 0) {
    for (int s = 0; s < T3_STAGES; ++s) { ptx::mbar_init(&bar->full[s], 1); ptx::mbar_init(&bar->empty[s], 1); }
    for (int a = 0; a < 2; ++a) { ptx::mbar_init(&bar->tmem_full[a], 1); ptx::mbar_init(&bar->tmem_empty[a], T2_EPI_WARPS * T3_EPI_GROUPS); }
    ptx::fence_barrier_init();
    ptx::tma_prefetch_desc(&p.ma[0]);
  }
  if (warp == T2_MMA_WARP) ptx::tmem_alloc(&bar->tmem_base, TC_TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = bar->tmem_base;
  const int64_t total = ceil_div(p.M, TC_BM) * p.n_tiles;

  if (warp == T3_PROD_WARP) {
    // ================= producer: A box by TMA + packed weight chunk by bulk copy, one barrier per stage =================
    if (lane == 0) {
      uint32_t it = 0;
      const uint32_t bytes = (uint32_t)stage;
      for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int row0 = (int)((tile / p.n_tiles) * TC_BM);
        const int nt = (int)(tile % p.n_tiles);
        for (int kc = 0; kc < p.KC; ++kc, ++it) {
          const int s = it % T3_STAGES;
          T2_WAIT(3, &bar->empty[s], ((it / T3_STAGES) & 1) ^ 1);
          ptx::mbar_arrive_expect_tx(&bar->full[s], bytes);
          const bool second = kc >= p.KC0;
          ptx::tma_load_2d(ops + s * stage, &p.ma[second ? 1 : 0], (second ? kc - p.KC0 : kc) * TC_BK, row0, &bar->full[s]);
          ptx::bulk_g2s(ops + s * stage + TC_A_STAGE, p.Bp + ((int64_t)nt * p.KC + kc) * p.BN * 32, (uint32_t)b_stage, &bar->full[s]);
        }
      }
    }
  } else if (warp == T2_MMA_WARP) {
    // ================= MMA issue =================
    uint32_t it = 0, tl = 0;
    const uint32_t ops_u = ptx::uniform_u32(ptx::smem_u32(ops)), tmem_u = ptx::uniform_u32(tmem_base);
    for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x, ++tl) {
      const int acc = tl & 1;
      T2_WAIT(4, &bar->tmem_empty[acc], ((tl >> 1) & 1) ^ 1);
      ptx::tc_fence_after();
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int s = it % T3_STAGES;
        T2_WAIT(5, &bar->full[s], (it / T3_STAGES) & 1);
        ptx::tc_fence_after();
        const uint32_t a_addr = ops_u + s * stage;         // convergent issue: the election is inside every statement
        tc_issue_chunk_elect(a_addr, a_addr + TC_A_STAGE, tmem_u + acc * 256, p.BN, kc == 0);
        ptx::tc_commit_elect(&bar->empty[s]);
        if (kc == p.KC - 1) ptx::tc_commit_elect(&bar->tmem_full[acc]);
        __syncwarp();
      }
    }
  } else if (warp < T2_EPI_WARPS || warp >= T3_EPI2_W0) {
    // ================= epilogue (warps 0-3, and 6-9 as the second group) =================
    uint32_t tl = 0, rit = 0;
    const int grp = warp >= T3_EPI2_W0 ? 1 : 0, wq = warp & 3;
    for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x, ++tl) {
      const int acc = tl & 1;
      const int64_t row0 = (tile / p.n_tiles) * TC_BM;
      const int col0 = (int)(tile % p.n_tiles) * p.BN;
      T2_WAIT(7, &bar->tmem_full[acc], (tl >> 1) & 1);
      ptx::tc_fence_after();
      t3_epilogue<EPI, ACT>(tmem_base + acc * 256, wq, lane, rring, rit, p.BN, row0, p.M, col0, p.N, p.ldc, p.bias,
                            p.C, p.ldc, p.drop, p.sign, grp, &bar->tmem_empty[acc]);
    }
  }
#ifdef WD_T2_TIMING
  if (WD_TIMING_ON(3) && blockIdx.x == 0 && lane == 0) {
    if (warp == 0 || warp == T2_MMA_WARP || warp == T3_PROD_WARP)
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
