// tcgen05 contraction kernel, second generation: latency hiding by a deep asynchronous copy ring.
//
//   C[M, ldc] = epi( A * B^T ),  A(r, :) = X0[i0(r), :] - act?(X1[i1(r), :])   (virtual, never in HBM)
//
// The first-generation kernel (tc_gemm.cuh) staged gathered rows through registers and was bound by
// load latency (ncu: long-scoreboard stalls, 0.3 TB/s).  Here the gather is shared-memory staged:
//   warps 0-3    epilogue   TMEM -> registers -> smem transpose -> coalesced stores (+ residual)
//   warp  4      MMA issue  tcgen05.mma kind::tf32, accumulators in TMEM
//   warp  5      B loader   bulk async copies (TMA engine) of pre-swizzled weight chunks, 2-stage ring
//   warps 6-9    A loaders  cp.async (16 B, zero-filling) of raw gathered row chunks into a 96 KB ring:
//                           nothing waits on a register, ~64-96 KB per SM are in flight
//   warps 10-13  transform  raw ring -> (subtract, activation, round to TF32) -> SWIZZLE_128B operand ring
// K is walked in chunks of 32 floats; the A operand may consist of two K segments (CONCAT).
#pragma once
#include "tc_gemm.cuh"

namespace wd {

#ifndef WD_T2_LOAD_WARPS
#define WD_T2_LOAD_WARPS 4
#endif
#ifndef WD_T2_XFORM_WARPS
#define WD_T2_XFORM_WARPS 4
#endif
constexpr int T2_EPI_WARPS = 4, T2_MMA_WARP = 4, T2_BLOAD_WARP = 5, T2_LOAD_WARPS = WD_T2_LOAD_WARPS,
              T2_XFORM_WARPS = WD_T2_XFORM_WARPS;
constexpr int T2_LOAD_T0 = 32 * 6, T2_XFORM_T0 = 32 * (6 + T2_LOAD_WARPS);
constexpr int T2_RLOAD_WARP = 6 + T2_LOAD_WARPS + T2_XFORM_WARPS;   // residual loader (one warp)
constexpr int T2_THREADS = 32 * (T2_RLOAD_WARP + 1);
constexpr int T2_MAX_BN = 160;                    // two accumulators in TMEM: epilogue overlaps the next tile
constexpr int T2_R_SLOTS = 3, T2_R_ROW = 144;     // residual / staging ring: 128 rows x 32 floats, rows padded to 144 B
constexpr int T2_R_SLOT_BYTES = TC_BM * T2_R_ROW;
constexpr int T2_R_BYTES = T2_R_SLOTS * T2_R_SLOT_BYTES;
constexpr int T2_LOAD_THREADS = 32 * T2_LOAD_WARPS, T2_XFORM_THREADS = 32 * T2_XFORM_WARPS;
constexpr int T2_LR = T2_LOAD_THREADS / 8, T2_LJ = TC_BM / T2_LR;      // loader: rows per pass / passes per 128-row chunk
constexpr int T2_XR = T2_XFORM_THREADS / 8, T2_XJ = TC_BM / T2_XR;     // transform: same
constexpr int T2_RAW_SLOTS = 6;                   // raw ring: 6 x 16 KB (3 stages of 2 sources or 6 of 1)
constexpr int T2_RAW_BYTES = T2_RAW_SLOTS * TC_A_STAGE;
constexpr int T2_OP_STAGES = 2;                   // operand rings (A swizzled, B weights)

struct T2Barriers {
  uint64_t raw_full[T2_RAW_SLOTS], raw_empty[T2_RAW_SLOTS];
  uint64_t a_full[T2_OP_STAGES], a_empty[T2_OP_STAGES], b_full[T2_OP_STAGES], b_empty[T2_OP_STAGES];
  uint64_t r_full[T2_R_SLOTS], r_empty[T2_R_SLOTS];
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

inline int t2_smem_bytes(int BN) {
  return 1024 + T2_OP_STAGES * (TC_A_STAGE + BN * 128) + T2_RAW_BYTES + T2_R_BYTES + (int)sizeof(T2Barriers) + 64;
}

// Epilogue of one 128 x BN accumulator, 32 columns at a time.  The residual chunk R[128, 32] has been
// prefetched into a ring slot by the residual-loader warp (cp.async); the row-owning thread adds its
// accumulator row in place (conflict-free 16-byte accesses: rows are 144 bytes apart), then the warp
// writes the 32 rows out with lane == column (coalesced 128-byte stores), applying bias + activation.
template <int EPI, int ACT>
__device__ __forceinline__ void t2_epilogue(uint32_t tmem_acc, int we, int lane, uint8_t* rring, T2Barriers* bar,
                                            uint32_t& rit, int BN, int64_t row0, int64_t M, int col0, int N, int cmax,
                                            const float* bias, float* C, int ldc) {
  for (int c0 = 0; c0 < BN; c0 += 32, ++rit) {
    const int slot = rit % T2_R_SLOTS;
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
    if (EPI == E_RES_ACT) ptx::mbar_wait(&bar->r_full[slot], (rit / T2_R_SLOTS) & 1);
    ptx::tmem_ld_wait();
    float4* myrow = reinterpret_cast<float4*>(sl + lane * T2_R_ROW);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 x = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      if (EPI == E_RES_ACT) {
        const float4 r = myrow[j];
        x = make_float4(x.x + r.x, x.y + r.y, x.z + r.z, x.w + r.w);
      }
      myrow[j] = x;
    }
    __syncwarp();
    const int col = col0 + c0 + lane;
    const bool in_n = col < N && c0 + lane < BN, in_c = col < cmax && c0 + lane < BN;
    const float bv = (bias != nullptr && in_n) ? __ldg(bias + col) : 0.f;
#pragma unroll 8
    for (int i = 0; i < 32; ++i) {
      const int64_t r = row0 + we * 32 + i;
      float x = *reinterpret_cast<const float*>(sl + i * T2_R_ROW + lane * 4) + bv;
      if (EPI != E_STORE) x = act_fwd<ACT>(x);
      if (r < M && in_c) C[r * ldc + col] = in_n ? x : 0.f;
    }
    __syncwarp();
    if (EPI == E_RES_ACT && lane == 0) ptx::mbar_arrive(&bar->r_empty[slot]);
  }
}

// One K segment of the virtual A operand.
struct T2Seg {
  const float* X0;      // minuend rows
  const float* X1;      // subtrahend rows (NSRC == 2) or nullptr
  const int32_t* i0;    // row indirection of X0 (nullptr: identity)
  const int32_t* i1;    // row indirection of X1
  int ld0, ld1, K;
  int align16;          // rows of every source are 16-byte aligned: 16-byte copies, else 4-byte copies
};

struct T2Params {
  T2Seg seg[2];         // seg[1].K == 0 when unused; indirection is honoured for seg[0] only
  int KC0, KC;          // chunks of segment 0 / total
  const float* Bp;
  const float* bias;
  const float* R;
  int ldr;
  float* C;
  int ldc;
  int64_t M;
  int N, BN, n_tiles;
};

template <int NSRC, int EPI, int ACT, bool PREACT>
__global__ void __launch_bounds__(T2_THREADS, 1) k_tc_gemm2(const T2Params p) {
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* opA = smem_raw + (((base + 1023u) & ~1023u) - base);        // 1024-aligned
  const int b_stage = p.BN * 128;
  uint8_t* opB = opA + T2_OP_STAGES * TC_A_STAGE;
  uint8_t* raw = opB + T2_OP_STAGES * b_stage;
  uint8_t* rring = raw + T2_RAW_BYTES;
  T2Barriers* bar = reinterpret_cast<T2Barriers*>(raw + T2_RAW_BYTES + T2_R_BYTES);
  constexpr int SR = T2_RAW_SLOTS / NSRC;                  // raw stages
  constexpr int RAW_STAGE = NSRC * TC_A_STAGE;
  const int nacc = p.BN <= 256 ? 2 : 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < SR; ++s) { ptx::mbar_init(&bar->raw_full[s], T2_LOAD_THREADS); ptx::mbar_init(&bar->raw_empty[s], T2_XFORM_WARPS); }
    for (int s = 0; s < T2_OP_STAGES; ++s) {
      ptx::mbar_init(&bar->a_full[s], T2_XFORM_WARPS); ptx::mbar_init(&bar->a_empty[s], 1);
      ptx::mbar_init(&bar->b_full[s], 1); ptx::mbar_init(&bar->b_empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) { ptx::mbar_init(&bar->tmem_full[a], 1); ptx::mbar_init(&bar->tmem_empty[a], T2_EPI_WARPS); }
    for (int s = 0; s < T2_R_SLOTS; ++s) { ptx::mbar_init(&bar->r_full[s], 32); ptx::mbar_init(&bar->r_empty[s], T2_EPI_WARPS); }
    ptx::fence_barrier_init();
  }
  if (warp == T2_MMA_WARP) ptx::tmem_alloc(&bar->tmem_base, TC_TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = bar->tmem_base;
  const int64_t total = ceil_div(p.M, TC_BM) * p.n_tiles;

  if (warp == T2_RLOAD_WARP) {
    // ================= residual loader: R[128 rows, 32 columns] chunks, ahead of the epilogue =================
    if (EPI == E_RES_ACT) {
      const uint32_t ring_u32 = ptx::smem_u32(rring);
      uint32_t rit = 0;
      for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int64_t row0 = (tile / p.n_tiles) * TC_BM;
        const int col0 = (int)(tile % p.n_tiles) * p.BN;
        for (int c0 = 0; c0 < p.BN; c0 += 32, ++rit) {
          const int slot = rit % T2_R_SLOTS;
          ptx::mbar_wait(&bar->r_empty[slot], ((rit / T2_R_SLOTS) & 1) ^ 1);
          const uint32_t dst = ring_u32 + slot * T2_R_SLOT_BYTES;
#pragma unroll 8
          for (int i = 0; i < 32; ++i) {
            const int q = lane + 32 * i, rl = q >> 3, pc = q & 7;
            const int64_t r = row0 + rl;
            const int col = col0 + c0 + pc * 4;
            const int rem = (p.N - col) * 4;
            const uint32_t vb = r >= p.M ? 0u : (rem >= 16 ? 16u : (rem > 0 ? (uint32_t)rem : 0u));
            ptx::cp_async16(dst + rl * T2_R_ROW + pc * 16, vb ? p.R + r * p.ldr + col : p.R, vb);
          }
          ptx::cp_async_arrive(&bar->r_full[slot]);
        }
      }
    }
  } else if (warp >= 6 + T2_LOAD_WARPS) {
    // ================= transform: raw ring -> swizzled TF32 operand ring =================
    const int t = threadIdx.x - T2_XFORM_T0, c = t & 7, rb = t >> 3;     // rows rb + T2_XR * j
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x) {
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int s = it % SR, as = it % T2_OP_STAGES;
        ptx::mbar_wait(&bar->raw_full[s], (it / SR) & 1);
        ptx::mbar_wait(&bar->a_empty[as], ((it / T2_OP_STAGES) & 1) ^ 1);
        const uint8_t* rs = raw + s * RAW_STAGE;
        uint8_t* oa = opA + as * TC_A_STAGE;
#pragma unroll
        for (int j = 0; j < T2_XJ; ++j) {
          const int rl = rb + T2_XR * j;
          float4 v = *reinterpret_cast<const float4*>(rs + rl * 128 + c * 16);
          if (NSRC == 2) {
            float4 u = *reinterpret_cast<const float4*>(rs + TC_A_STAGE + rl * 128 + c * 16);
            if (PREACT) u = f4_act<ACT>(u);
            v = make_float4(v.x - u.x, v.y - u.y, v.z - u.z, v.w - u.w);
          }
          tc_store_piece(oa, rl, c, v);
        }
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) { ptx::mbar_arrive(&bar->a_full[as]); ptx::mbar_arrive(&bar->raw_empty[s]); }
      }
    }
  } else if (warp >= 6) {
    // ================= A loaders: asynchronous gathered row copies =================
    const int t = threadIdx.x - T2_LOAD_T0;
    const uint32_t raw_u32 = ptx::smem_u32(raw);
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x) {
      const int64_t row0 = (tile / p.n_tiles) * TC_BM;
      // 16-byte path: thread owns chunk column c of rows rb + T2_LR * j
      const int c = t & 7, rb = t >> 3;
      int64_t o0[T2_LJ], o1[T2_LJ];
      uint32_t okmask = 0;
#pragma unroll
      for (int j = 0; j < T2_LJ; ++j) {
        const int64_t r = row0 + rb + T2_LR * j;
        const bool ok = r < p.M;
        okmask |= (ok ? 1u : 0u) << j;
        int64_t a = ok ? r : 0, b = a;
        if (ok && p.seg[0].i0) a = __ldg(p.seg[0].i0 + r);
        if (NSRC == 2 && ok && p.seg[0].i1) b = __ldg(p.seg[0].i1 + r);
        o0[j] = a;
        o1[j] = b;
      }
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int s = it % SR;
        ptx::mbar_wait(&bar->raw_empty[s], ((it / SR) & 1) ^ 1);
        const bool second = kc >= p.KC0;
        const T2Seg& sg = p.seg[second ? 1 : 0];
        const int kbase = (second ? kc - p.KC0 : kc) * TC_BK;
        const uint32_t dst = raw_u32 + s * RAW_STAGE;
        if (sg.align16) {
          const int k = kbase + c * 4;
          const int rem = (sg.K - k) * 4;
          const uint32_t valid = rem >= 16 ? 16u : (rem > 0 ? (uint32_t)rem : 0u);
#pragma unroll
          for (int j = 0; j < T2_LJ; ++j) {
            const uint32_t v = ((okmask >> j) & 1u) ? valid : 0u;
            const int64_t ra = second ? row0 + rb + T2_LR * j : o0[j];
            const uint32_t d = dst + (rb + T2_LR * j) * 128 + c * 16;
            ptx::cp_async16(d, v ? sg.X0 + ra * sg.ld0 + k : sg.X0, v);
            if (NSRC == 2) ptx::cp_async16(d + TC_A_STAGE, v ? sg.X1 + o1[j] * sg.ld1 + k : sg.X1, v);
          }
        } else {  // unaligned rows (identity row map only): 4-byte copies, lane = column
          const int col = t & 31, r4 = t >> 5, k = kbase + col;
          const bool vk = k < sg.K;
#pragma unroll 8
          for (int j = 0; j < TC_BM / T2_LOAD_WARPS; ++j) {
            const int rl = r4 + T2_LOAD_WARPS * j;
            const int64_t r = row0 + rl;
            const uint32_t v = (vk && r < p.M) ? 4u : 0u;
            ptx::cp_async4(dst + rl * 128 + col * 4, v ? sg.X0 + r * sg.ld0 + k : sg.X0, v);
          }
        }
        ptx::cp_async_arrive(&bar->raw_full[s]);
      }
    }
  } else if (warp == T2_BLOAD_WARP) {
    // ================= B loader: bulk async copies of packed weight chunks =================
    if (lane == 0) {
      uint32_t it = 0;
      const uint32_t bytes = (uint32_t)b_stage;
      for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int nt = (int)(tile % p.n_tiles);
        for (int kc = 0; kc < p.KC; ++kc, ++it) {
          const int bs = it % T2_OP_STAGES;
          ptx::mbar_wait(&bar->b_empty[bs], ((it / T2_OP_STAGES) & 1) ^ 1);
          ptx::mbar_arrive_expect_tx(&bar->b_full[bs], bytes);
          ptx::bulk_g2s(opB + bs * b_stage, p.Bp + ((int64_t)nt * p.KC + kc) * p.BN * 32, bytes, &bar->b_full[bs]);
        }
      }
    }
  } else if (warp == T2_MMA_WARP) {
    // ================= MMA issue =================
    uint32_t it = 0, tl = 0;
    for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x, ++tl) {
      const int acc = tl % nacc;
      ptx::mbar_wait(&bar->tmem_empty[acc], ((tl / nacc) & 1) ^ 1);
      ptx::tc_fence_after();
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int os = it % T2_OP_STAGES;
        const uint32_t ph = (it / T2_OP_STAGES) & 1;
        ptx::mbar_wait(&bar->a_full[os], ph);
        ptx::mbar_wait(&bar->b_full[os], ph);
        ptx::tc_fence_after();
        if (lane == 0) {
          tc_issue_chunk(ptx::smem_u32(opA + os * TC_A_STAGE), ptx::smem_u32(opB + os * b_stage),
                         tmem_base + acc * 256, p.BN, kc == 0);
          ptx::tc_commit(&bar->a_empty[os]);
          ptx::tc_commit(&bar->b_empty[os]);
          if (kc == p.KC - 1) ptx::tc_commit(&bar->tmem_full[acc]);
        }
        __syncwarp();
      }
    }
  } else {
    // ================= epilogue =================
    uint32_t tl = 0, rit = 0;
    for (int64_t tile = blockIdx.x; tile < total; tile += gridDim.x, ++tl) {
      const int acc = tl % nacc;
      const int64_t row0 = (tile / p.n_tiles) * TC_BM;
      const int col0 = (int)(tile % p.n_tiles) * p.BN;
      ptx::mbar_wait(&bar->tmem_full[acc], (tl / nacc) & 1);
      ptx::tc_fence_after();
      t2_epilogue<EPI, ACT>(tmem_base + acc * 256, warp, lane, rring, bar, rit, p.BN, row0, p.M, col0, p.N, p.ldc,
                            p.bias, p.C, p.ldc);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&bar->tmem_empty[acc]);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == T2_MMA_WARP) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

}  // namespace wd
