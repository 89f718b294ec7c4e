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
#ifndef WD_T2_BULK_A
#define WD_T2_BULK_A 0   // 1: aligned A rows by one 128-byte bulk copy per row and chunk (measured 2x SLOWER than cp.async)
#endif
#ifndef WD_ABL
#define WD_ABL 0   // timing ablations (results become wrong): 1 no B copies, 2 A rows from an L2-resident window,
#endif             // 4 no stores, 8 R rows from a window, 16 no MMA, 32 no A copies
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

// Per-role wait-time counters of CTA 0 (build with -DWD_T2_TIMING; read through wd_debug_timing):
//  0 xform<-raw_full 1 xform<-a_empty 2 loader<-raw_empty 3 bload<-b_empty 4 mma<-tmem_empty 5 mma<-a_full
//  6 mma<-b_full 7 epi<-tmem_full 8 epi<-r_full 9 rload<-r_empty 15 kernel cycles
__device__ unsigned long long g_t2_timing[16];
#ifndef WD_T2_TIMING_ONLY
#define WD_T2_TIMING_ONLY 0          // 0: every instrumented kernel records; n: only kernel generation n (2, 3, 4 gemm; 14, 15 dW)
#endif
#define WD_TIMING_ON(n) (WD_T2_TIMING_ONLY == 0 || WD_T2_TIMING_ONLY == (n))
#ifdef WD_T2_TIMING
#define T2_WAIT(slot, b, par) do { const long long t0_ = clock64(); ptx::mbar_wait(b, par); tacc[slot] += clock64() - t0_; } while (0)
#else
#define T2_WAIT(slot, b, par) ptx::mbar_wait(b, par)
#endif

static_assert(!WD_T2_BULK_A || T2_LOAD_THREADS == TC_BM, "bulk A path: one loader thread per tile row");

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
                                            const float* bias, float* C, int ldc, long long* tacc,
                                            const Dropout& drop, const SignBits& sign) {
  uint32_t sw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};   // sign words of this lane's chunk (common.cuh: sign_collect)
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
    if (EPI == E_RES_ACT) T2_WAIT(8, &bar->r_full[slot], (rit / T2_R_SLOTS) & 1);
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
    // write-out: 8 lanes per row (16 bytes each), 4 rows per instruction -> full 128-byte lines
    const int pc = lane & 7, rq = lane >> 3;
    const int colb = col0 + c0 + pc * 4;
    const int nval = (c0 + pc * 4 < BN) ? max(0, min(4, N - colb)) : 0;      // live columns of this piece
    const bool in_c = colb < cmax && c0 + pc * 4 < BN;
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
      float4 x = make_float4(xv[i].x + bv.x, xv[i].y + bv.y, xv[i].z + bv.z, xv[i].w + bv.w);
      if (EPI != E_STORE) x = f4_act<ACT>(x);
      if (EPI != E_STORE && drop.thresh != 0u) x = dropout4(x, drop, (row0 + we * 32 + rq + 4 * i) * ldc + colb);
      if (nval < 4) { x.w = 0.f; if (nval < 3) x.z = 0.f; if (nval < 2) x.y = 0.f; if (nval < 1) x.x = 0.f; }
      const bool ok = in_c && row0 + we * 32 + rq + 4 * i < M && !(WD_ABL & 4);
      if (ok) *reinterpret_cast<float4*>(cp + (int64_t)(4 * i) * ldc) = x;
      if (sign.p != nullptr) sign_collect(sw[i], x, rq, pc, c0 >> 5);   // warp-uniform branch
    }
    __syncwarp();
    if (EPI == E_RES_ACT && lane == 0) ptx::mbar_arrive(&bar->r_empty[slot]);
  }
  if (sign.p != nullptr) sign_flush(sign, sw, col0 / BN, row0 + we * 32, M, lane >> 3, lane & 7, (BN + 31) >> 5);
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
  float* S;             // optional copy of the virtual A operand (segment 0, before rounding), [M, lds]
  int lds;
  Dropout drop;         // applied after the activation (EPI != E_STORE)
  SignBits sign;        // (stored value > 0) bits, written next to C when sign.p != nullptr
};

template <int NSRC, int EPI, int ACT, bool PREACT>
__global__ void __launch_bounds__(T2_THREADS, 1) k_tc_gemm2(const T2Params p) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
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
  long long tacc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  const long long t_start = clock64();

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
          T2_WAIT(9, &bar->r_empty[slot], ((rit / T2_R_SLOTS) & 1) ^ 1);
          // 8 lanes per row, 4 rows per instruction, 32 instructions per chunk (per-row bulk copies were
          // measured 2x slower: the TMA engine does not like 128-byte copies)
          const int pc = lane & 7, rq = lane >> 3;
          const int col = col0 + c0 + pc * 4;
          const int rem = (p.N - col) * 4;
          const uint32_t vbc = rem >= 16 ? 16u : (rem > 0 ? (uint32_t)rem : 0u);
          const int nrows = p.M - row0 < TC_BM ? (int)(p.M - row0) : TC_BM;
          const uint32_t dst = ring_u32 + slot * T2_R_SLOT_BYTES + rq * T2_R_ROW + pc * 16;
          const int64_t rbase = (WD_ABL & 8) ? ((row0 + rq) & 1023) : row0 + rq;
          const float* src = p.R + rbase * p.ldr + col;
          const int64_t step = 4 * (int64_t)p.ldr;
          if (nrows == TC_BM && vbc == 16u) {
#pragma unroll 8
            for (int i = 0; i < 32; ++i) ptx::cp_async16(dst + i * (4 * T2_R_ROW), src + i * step, 16u);
          } else {
#pragma unroll 4
            for (int i = 0; i < 32; ++i) {
              const uint32_t vb = rq + 4 * i < nrows ? vbc : 0u;
              ptx::cp_async16(dst + i * (4 * T2_R_ROW), vb ? src + i * step : p.R, vb);
            }
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
      // the first column tile of every row tile also writes the operand rows out (read back by the dW pass)
      const int64_t srow0 = (tile / p.n_tiles) * TC_BM;
      const bool keep = p.S != nullptr && (tile % p.n_tiles) == 0;
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int s = it % SR, as = it % T2_OP_STAGES;
        T2_WAIT(0, &bar->raw_full[s], (it / SR) & 1);
        T2_WAIT(1, &bar->a_empty[as], ((it / T2_OP_STAGES) & 1) ^ 1);
        const uint8_t* rs = raw + s * RAW_STAGE;
        uint8_t* oa = opA + as * TC_A_STAGE;
        // bulk-copied rows carry stale bytes past the end of the K segment: this thread's piece may be dead
        const bool sec = kc >= p.KC0;
        const int kleft = p.seg[sec ? 1 : 0].K - (sec ? kc - p.KC0 : kc) * TC_BK;
        const bool dead = WD_T2_BULK_A && c * 4 >= kleft;
        // loads are batched ahead of the stores (the compiler cannot reorder them: both rings are smem)
        constexpr int XB = T2_XJ < 4 ? T2_XJ : 4;
#pragma unroll
        for (int j0 = 0; j0 < T2_XJ; j0 += XB) {
          float4 v[XB], u[XB];
#pragma unroll
          for (int j = 0; j < XB; ++j) {
            const int rl = rb + T2_XR * (j0 + j);
            v[j] = *reinterpret_cast<const float4*>(rs + rl * 128 + c * 16);
            if (NSRC == 2) u[j] = *reinterpret_cast<const float4*>(rs + TC_A_STAGE + rl * 128 + c * 16);
          }
#pragma unroll
          for (int j = 0; j < XB; ++j) {
            if (NSRC == 2) {
              if (PREACT) u[j] = f4_act<ACT>(u[j]);
              v[j] = make_float4(v[j].x - u[j].x, v[j].y - u[j].y, v[j].z - u[j].z, v[j].w - u[j].w);
            }
            if (dead) v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (keep && kc * TC_BK + c * 4 < p.lds && srow0 + rb + T2_XR * (j0 + j) < p.M)
              *reinterpret_cast<float4*>(p.S + (srow0 + rb + T2_XR * (j0 + j)) * p.lds + kc * TC_BK + c * 4) = v[j];
            tc_store_piece(oa, rb + T2_XR * (j0 + j), c, v[j]);
          }
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
        if (WD_ABL & 2) { a &= 2047; b &= 2047; }
        o0[j] = a;
        o1[j] = b;
      }
      int64_t g0 = 0, g1 = 0;                       // bulk path: source rows of chunk row t
      if (WD_T2_BULK_A && row0 + t < p.M) {
        g0 = g1 = row0 + t;
        if (p.seg[0].i0) g0 = __ldg(p.seg[0].i0 + row0 + t);
        if (NSRC == 2 && p.seg[0].i1) g1 = __ldg(p.seg[0].i1 + row0 + t);
        if (WD_ABL & 2) { g0 &= 2047; g1 &= 2047; }
      }
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int s = it % SR;
        T2_WAIT(2, &bar->raw_empty[s], ((it / SR) & 1) ^ 1);
        const bool second = kc >= p.KC0;
        const T2Seg& sg = p.seg[second ? 1 : 0];
        const int kbase = (second ? kc - p.KC0 : kc) * TC_BK;
        const uint32_t dst = raw_u32 + s * RAW_STAGE;
        if (WD_ABL & 32) {
          ptx::cp_async_arrive(&bar->raw_full[s]);
        } else if (sg.align16 && WD_T2_BULK_A) {
          // thread t owns row t of the chunk; columns past the segment end are masked by the transform warps
          const int kcols = min(TC_BK, ((sg.K + 3) & ~3) - kbase);
          const int64_t r = row0 + t;
          if (r < p.M) {
            const uint32_t bytes = (uint32_t)kcols * 4u;
            ptx::mbar_arrive_expect_tx(&bar->raw_full[s], bytes * NSRC);
            uint8_t* d = raw + s * RAW_STAGE + t * 128;
            ptx::bulk_g2s(d, sg.X0 + (second ? r : g0) * sg.ld0 + kbase, bytes, &bar->raw_full[s]);
            if (NSRC == 2) ptx::bulk_g2s(d + TC_A_STAGE, sg.X1 + g1 * sg.ld1 + kbase, bytes, &bar->raw_full[s]);
          } else {
            ptx::mbar_arrive(&bar->raw_full[s]);
          }
        } else if (sg.align16) {
          const int k = kbase + c * 4;
          const int rem = (sg.K - k) * 4;
          const uint32_t valid = rem >= 16 ? 16u : (rem > 0 ? (uint32_t)rem : 0u);
          const uint32_t d0 = dst + rb * 128 + c * 16;
          if (okmask == (1u << T2_LJ) - 1u && valid == 16u) {        // interior: no predication, no selects
#pragma unroll
            for (int j = 0; j < T2_LJ; ++j) {
              const int64_t ra = second ? row0 + rb + T2_LR * j : o0[j];
              ptx::cp_async16(d0 + j * (T2_LR * 128), sg.X0 + ra * sg.ld0 + k, 16u);
              if (NSRC == 2) ptx::cp_async16(d0 + j * (T2_LR * 128) + TC_A_STAGE, sg.X1 + o1[j] * sg.ld1 + k, 16u);
            }
          } else {
#pragma unroll
            for (int j = 0; j < T2_LJ; ++j) {
              const uint32_t v = ((okmask >> j) & 1u) ? valid : 0u;
              const int64_t ra = second ? row0 + rb + T2_LR * j : o0[j];
              ptx::cp_async16(d0 + j * (T2_LR * 128), v ? sg.X0 + ra * sg.ld0 + k : sg.X0, v);
              if (NSRC == 2) ptx::cp_async16(d0 + j * (T2_LR * 128) + TC_A_STAGE, v ? sg.X1 + o1[j] * sg.ld1 + k : sg.X1, v);
            }
          }
          ptx::cp_async_arrive(&bar->raw_full[s]);
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
          ptx::cp_async_arrive(&bar->raw_full[s]);
        }
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
          T2_WAIT(3, &bar->b_empty[bs], ((it / T2_OP_STAGES) & 1) ^ 1);
          if (WD_ABL & 1) { ptx::mbar_arrive(&bar->b_full[bs]); continue; }
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
      T2_WAIT(4, &bar->tmem_empty[acc], ((tl / nacc) & 1) ^ 1);
      ptx::tc_fence_after();
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int os = it % T2_OP_STAGES;
        const uint32_t ph = (it / T2_OP_STAGES) & 1;
        T2_WAIT(5, &bar->a_full[os], ph);
        T2_WAIT(6, &bar->b_full[os], ph);
        ptx::tc_fence_after();
        if (lane == 0) {
          if (!(WD_ABL & 16))
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
      T2_WAIT(7, &bar->tmem_full[acc], (tl / nacc) & 1);
      ptx::tc_fence_after();
      t2_epilogue<EPI, ACT>(tmem_base + acc * 256, warp, lane, rring, bar, rit, p.BN, row0, p.M, col0, p.N, p.ldc,
                            p.bias, p.C, p.ldc, tacc, p.drop, p.sign);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&bar->tmem_empty[acc]);
    }
  }
#ifdef WD_T2_TIMING
  if (WD_TIMING_ON(2) && blockIdx.x == 0 && lane == 0) {
    if (warp == 0 || warp == T2_MMA_WARP || warp == T2_BLOAD_WARP || warp == 6 || warp == 6 + T2_LOAD_WARPS || warp == T2_RLOAD_WARP)
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
