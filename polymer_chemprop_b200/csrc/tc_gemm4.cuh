// tcgen05 FORWARD contraction kernel (fourth generation): every forward contraction of the encoder that needs the
// raw fp32 operand on the SM -- the gathered depth step in every tensor-core mode, and all three forward
// contractions in the 3xTF32 mode --
//
//   GATHER : C[M, ldc] = act( R + (X0[i0(r), :] - act?(X1[i1(r), :])) * B^T + bias )     (mpn.py:119-123 as one kernel)
//   PLAIN  : C[M, ldc] = epi( [Xa | Xb] * B^T + bias )                                   (mpn.py:96, :132-133)
//
// Measured on the second generation (profiles/r1_g_*): cp.async (LDGSTS) moves ~12 bytes / cycle / SM and is the
// binding limit; with two column tiles of 160 every gathered row was fetched twice, and the residual tile
// went through LDGSTS as well.  Here
//   * one CTA tile spans up to 304 output columns (two accumulator halves of <= 160), so each gathered row is
//     fetched ONCE; three 160-column TMEM buffers rotate (tile i: buffers 2i, 2i+1 mod 3), the epilogue of
//     tile i overlaps the main loop of tile i+1;
//   * loaders cp.async the two gathered sources straight to their SWIZZLE_128B positions (PLAIN: one TMA box,
//     FLOAT32 map, no rounding in flight); the transform warps subtract / activate / round to TF32 IN PLACE (no
//     second operand ring) and, for the first column tile, write the operand rows S out for the weight-gradient
//     pass (after the hand-off to the MMA);
//   * the residual tile arrives by TMA (FLOAT32 map, SWIZZLE_128B) in a ring that doubles as the epilogue's
//     transpose staging; the packed weight chunk by bulk copies, one per accumulator half;
//   * reverse bonds are ADJACENT PAIRS in every graph the reference builds (b1 = n_bonds, b2 = b1 + 1,
//     featurization.py:470-480 and :616-629; with the padding bond 0 that is rev(b) = ((b-1)^1)+1).  When
//     wd_csr_build found that to hold for the whole batch (flags[2] == 0 -> WdGraph.rev_adjacent), the rows
//     X1[i1(r)] of a 128-row tile all lie in rows [row0-1, row0+129) of X1: ONE TMA box of 132 rows replaces
//     128 gathered cp.async rows, i.e. half of the LDGSTS traffic that bounds this kernel, and the transform
//     warps read the partner row of the box instead of their own.  Any other graph takes the cp.async path.
//
// 3xTF32 (T4Params.split, math mode WD_MATH_TF32X3).  Single-pass TF32 rounds both operands to 11 significant bits
// (2^-11 relative per element): 3.6e-4 on the encoder output and 5e-4 on predictions, above the 1e-4 bound.  With
//   a = a_hi + a_lo,  a_hi = tf32(a), a_lo = a - a_hi           (the transform warps already touch every element)
//   w = w_hi + w_lo                                              (k_tc_pack_b leaves both images, tc_core.cuh)
//   a w ~= a_hi w_hi + (a_lo w_hi + a_hi w_lo)                   (all into the SAME TMEM accumulator)
// the dropped term is ~2^-22 and the result is fp32-grade (2e-6 against fp64 at K = 300; what remains is the
// tensor core's own accumulation rounding, which grows with K).  The correction terms are 2^-11 of the product, so
// THEY only need 8 significant bits: they run as ONE kind::f16 (bf16 inputs) MMA group over "correction images"
// [bf16 a_lo | bf16 a_hi] and [bf16 w_hi | bf16 w_lo] -- K = 64 bf16 in the same 128-byte rows, 4 instructions per
// chunk instead of 8: the ablations (profiles/r2_fwd_ab.log) showed the kernel paced by the number of MMAs issued,
// not by the weight traffic.  No extra HBM pass: the correction image of A is written over the subtrahend part of
// the SAME smem stage (after a named barrier among the transform warps when that part holds the shared TMA box);
// the weight ring streams two images per chunk from L2.
// 3xBF16 (T4Params.split == 2, math mode WD_MATH_BF16X3): the same correction images with the split point at bfloat16,
//   a = a_hi + a_lo, a_hi = bf16(a), a_lo = bf16(a - a_hi);  w likewise;  a w ~= a_hi w_hi + a_lo w_hi + a_hi w_lo
// (dropped: ~2^-17 of the product; the exact-accumulation model gives 3e-6 on the encoder output at h = 300 against 4e-7 for
// 3xTF32 and 2e-4 for single-pass TF32, tests/test_oracle_golden.py).  There is NO tf32 image: the operand stage holds
// the raw rows and ONE 128-byte-row image [a_lo | a_hi], the weight ring one image [w_hi | w_lo] per chunk, and a K chunk
// is six kind::f16 MMAs per accumulator half -- the four of the correction group (a_lo w_hi + a_hi w_lo over the 64
// bf16 of a row) plus two that pair the a_hi half of the operand row with the w_hi half of the weight row (descriptor
// start addresses 64 bytes apart).  Against 3xTF32: 6 instead of 8 MMAs, half the weight-ring bytes, no hi-image write
// -- ~28 % fewer bytes through the shared-memory port, which is what bounds this kernel (profiles/r2_fwd_abl.log, L).
//   warps 0-3 epilogue, 4 MMA issue, 5 weight loader (lane 0; lane 1 issues the subtrahend boxes of the 3xBF16 early-refill
//   scheme, see early_x0 in t4_body), 6 residual loader, 7-10 operand loaders, 11-14 transform, 15-18 second epilogue group.
#pragma once
#include "tc_gemm3.cuh"

namespace wd {

#ifndef WD_T4_XFORM_WARPS
#define WD_T4_XFORM_WARPS 4
#endif
constexpr int T4_MMA_WARP = 4, T4_B_WARP = 5, T4_R_WARP = 6, T4_LOAD_W0 = 7, T4_LOAD_WARPS = 4, T4_XFORM_W0 = 11,
              T4_XFORM_WARPS = WD_T4_XFORM_WARPS;
constexpr int T4_XR = 4 * T4_XFORM_WARPS, T4_XJ = TC_BM / T4_XR;   // transform: rows per pass / passes per chunk
// Epilogue groups: 2 (the default) = warps 0-3 plus a second group of four warps after the transform warps (TMEM lane
// quarter = warp % 4 in both), the groups taking alternate 32-column chunks; 1 = warps 0-3 only.  Measured
// (profiles/r2_fwd_abl.log section G): depth-step contraction 328 -> 309 us, forward pass 1258 -> 1206 us, step 2.916 ->
// 2.864 ms.  The first version shared the residual ring's "box landed" barriers between the groups and failed the
// bitwise stress test once in ~60 steps (a parity wait that skips every second phase, see t4_rfull); with a barrier set
// per group it passes (6 x 240 steps).
#ifndef WD_T4_EPI_GROUPS
#define WD_T4_EPI_GROUPS 2
#endif
#ifndef WD_T4_MERGE
#define WD_T4_MERGE 1          // 1: one bulk copy per packed image (the rows of both accumulator halves are contiguous) instead of
                               // one per half.  Same bytes, half the copy operations: depth-step contraction 309 -> 288 us, step
                               // 2.87 -> 2.80 ms (profiles/r2_fwd_abl.log section J) -- the per-operation cost of the copy engine,
                               // not the bandwidth, is what the "no weight copies" ablation had measured
#endif
constexpr bool T4_MERGE = WD_T4_MERGE != 0;
constexpr int T4_EPI_GROUPS = WD_T4_EPI_GROUPS, T4_EPI2_W0 = T4_XFORM_W0 + T4_XFORM_WARPS;
constexpr int T4_THREADS = 32 * (T4_EPI2_W0 + (T4_EPI_GROUPS - 1) * 4);
#ifndef WD_T4_A_STAGES
#define WD_T4_A_STAGES 3
#endif
#ifndef WD_T4_MAX_BN
#define WD_T4_MAX_BN 304
#endif
constexpr int T4_A_STAGES = WD_T4_A_STAGES, T4_B_STAGES = 2, T4_R_SLOTS = 3;
constexpr int T4_B_SLOTS = 4 * T4_B_STAGES;      // weight-ring units: 2 per stage; a CTA pair keeps half of each unit, so twice as many fit
constexpr int T4_X1_ROWS = 132, T4_X1_LEAD = 2;   // TMA box of the subtrahend rows: [row0 - 2, row0 + 130)
constexpr int T4_X1_STAGE = 17 * 1024;            // >= T4_X1_ROWS * 128 bytes, keeps the stages 1024-byte aligned
constexpr int T4_A_STAGE = TC_A_STAGE + T4_X1_STAGE;   // minuend rows / a_hi (128 x 128 bytes) + subtrahend rows / a_lo
constexpr int T4_R_SLOT = TC_BM * 128;            // residual / staging slot: 128 rows x 32 floats, swizzled
constexpr int T4_HALF = 160, T4_MAX_BN = WD_T4_MAX_BN;     // accumulator half / widest column tile
static_assert(T4_LOAD_WARPS * 32 == TC_BM && (T4_XFORM_WARPS == 4 || T4_XFORM_WARPS == 8), "8 threads per row");

struct T4Barriers {
  uint64_t raw_full[T4_A_STAGES], op_full[T4_A_STAGES], a_empty[T4_A_STAGES];
  uint64_t x0_free[T4_A_STAGES];    // 3xBF16: the raw-row part of a stage is free again (transform done), see t4_body
  uint64_t b_full[T4_B_SLOTS], b_empty[T4_B_SLOTS];
  uint64_t r_full[T4_R_SLOTS * 2], r_empty[T4_R_SLOTS];    // r_full: one set per epilogue group (t4_rfull below)
  uint64_t tmem_full[3], tmem_empty[3];
  uint32_t tmem_base;
};

struct T4Params {
  CUtensorMap mr;       // residual R[M, N] (FLOAT32, box 128 x 32, SWIZZLE_128B); EPI == E_RES_ACT
  CUtensorMap mx1;      // GATHER: X1[M, K] (FLOAT32, box 132 rows x 32 columns, SWIZZLE_128B), used when x1_tma != 0
  CUtensorMap ma[2];    // PLAIN: the K segments of A (FLOAT32, box 128 x 32, SWIZZLE_128B), used when a_tma != 0
  const float* X0;      // GATHER: minuend rows (row map i0).  PLAIN: first K segment
  const float* X1;      // GATHER: subtrahend rows (row map i1).  PLAIN: second K segment (or nullptr)
  const int32_t* i0;
  const int32_t* i1;
  int ld0, ld1;
  int K, K2;            // widths of the K segments (K2 == 0: one segment)
  int KC0, KC;          // chunks of segment 0 / total
  const float* Bp;      // packed weights, `parts` images of [BN x 128 bytes] per (column tile, K chunk)
  const float* bias;
  float* C;
  int ldc;
  int64_t M;
  int N, BN, n_tiles;
  float* S;             // GATHER: optional copy of the operand rows, [M, lds]
  int lds;
  Dropout drop;         // applied after the activation
  SignBits sign;        // (stored value > 0) bits, written next to C when sign.p != nullptr
  int x1_tma;           // GATHER: 1: i1(r) == ((r-1)^1)+1 for every r >= 1 (adjacent reverse pairs): X1 rows arrive by TMA
  int a_tma;            // PLAIN: 1: rows are 16-byte aligned and arrive by TMA; 0: 4-byte cp.async copies
  int split;            // 0: single-pass TF32; 1: 3xTF32 (tf32 hi image + bf16 correction image, two MMA groups per K chunk);
                        // 2: 3xBF16 (ONE bf16 [lo | hi] image per operand, three bf16 products per K chunk, see below)
  CUtensorMap mb[2];    // CTA pair: the packed weight images as one [rows, 32] matrix, box = (W0 / 2, W1 / 2) rows, no swizzle
};

inline int t4_smem_bytes(int BN) {
  return 1024 + T4_A_STAGES * T4_A_STAGE + T4_B_STAGES * BN * 128 + T4_R_SLOTS * T4_R_SLOT + (int)sizeof(T4Barriers) + 64;
}

// Which "residual box has landed" barrier belongs to chunk `rit` (ring slot rit % 3), and the parity of its phase.  With
// two epilogue groups the uses of a slot alternate between the groups (chunks c, c + 3, c + 6 ...); a parity wait is only
// meaningful for a waiter that consumes EVERY phase, so each group has its own barrier per slot (the loader arms the
// one of the group that will read the box): a group's chunks in one slot are 6 apart.  (With a shared barrier a box of
// the other group that lands late lets the wait for c + 6 succeed on the phase of c: one corrupted step in ~60.)
__device__ __forceinline__ int t4_rfull(uint32_t rit) {
  return (int)(rit % T4_R_SLOTS) + (T4_EPI_GROUPS > 1 ? (int)(rit & 1u) * T4_R_SLOTS : 0);
}
__device__ __forceinline__ uint32_t t4_rfull_parity(uint32_t rit) {
  return (rit / (T4_R_SLOTS * T4_EPI_GROUPS)) & 1u;
}

// Epilogue of one accumulator half (128 rows x W columns), 32 columns at a time: the row-owning thread adds its
// accumulator row to the residual chunk in the (swizzled) ring slot (EPI == E_RES_ACT; otherwise the slot is only the
// transpose staging of this warp's 32 rows), then the warp writes the 32 rows out with 16-byte stores (8 lanes per
// row), applying bias + activation.
// The accumulator buffer is handed back to the MMA warp (`tmem_empty`) as soon as this warp's LAST tcgen05.ld of the half
// has completed -- the rest of that chunk (staging, activation, stores) works from registers.  (The MMA warp waited 29-33 %
// of the kernel for a free accumulator with the hand-back after the stores: profiles/r2_n_fwd_timing.log.)
template <int EPI, int ACT, bool PAIR>
__device__ __forceinline__ void t4_epilogue(uint32_t tmem_acc, int we, int lane, uint8_t* rring, T4Barriers* bar,
                                            uint32_t& rit, int W, int64_t row0, int64_t M, int col0, int N, int cmax,
                                            const float* bias, float* C, int ldc, long long* tacc,
                                            const Dropout& drop, const SignBits& sign, int sign_tile, int grp,
                                            uint64_t* tmem_empty) {
  uint32_t sw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};   // sign words of this lane's chunk (common.cuh: sign_collect)
  const int rows_ok = (int)min((int64_t)32, M - (row0 + we * 32));        // rows of this warp inside the matrix (may be <= 0)
  const uint32_t rit0 = rit;
  auto release = [&]() {
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) { if (PAIR) ptx::mbar_arrive_leader(tmem_empty); else ptx::mbar_arrive(tmem_empty); }
  };
  // this group's last chunk of the half (chunk numbers rit0 .. rit0 + nch - 1, a group takes those of its parity)
  const uint32_t nch = (uint32_t)((W + 31) >> 5);
  uint32_t last = rit0 + nch - 1u;
  bool has_chunk = true;
  if (T4_EPI_GROUPS > 1 && (int)(last & 1u) != grp) { has_chunk = nch >= 2u; last -= 1u; }
  if (!has_chunk) release();                       // nothing to read for this group: hand back at once
  for (int c0 = 0; c0 < W; c0 += 32, ++rit) {
    if (T4_EPI_GROUPS > 1 && (int)(rit & 1u) != grp) continue;            // the other group's chunk
    // the residual's ring slot (handed over chunk by chunk through r_full / r_empty); without a residual the slot is
    // only staging and nothing orders the two groups, so each keeps to a slot of its own
    const int slot = (EPI == E_RES_ACT || T4_EPI_GROUPS == 1) ? (int)(rit % T4_R_SLOTS) : grp;
    uint8_t* sl = rring + slot * T4_R_SLOT + (we * 32) * 128;               // this warp's 32 rows (row & 7 == lane & 7)
    if (WD_ABL & 64) {        // timing ablation: no epilogue work, only the slot protocol
      if (EPI == E_RES_ACT) {
        T2_WAIT(8, &bar->r_full[t4_rfull(rit)], t4_rfull_parity(rit));
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bar->r_empty[slot]);
      }
      if (has_chunk && rit == last) release();
      continue;
    }
    float v[32];
    const uint32_t ta = tmem_acc + ((uint32_t)(we * 32) << 16) + (uint32_t)c0;
    if (W - c0 >= 32) {
      ptx::tmem_ld32(ta, v);
    } else {
      ptx::tmem_ld16(ta, v);
#pragma unroll
      for (int j = 16; j < 32; ++j) v[j] = 0.f;
    }
    if (EPI == E_RES_ACT) T2_WAIT(8, &bar->r_full[t4_rfull(rit)], t4_rfull_parity(rit));
    ptx::tmem_ld_wait();
    if (has_chunk && rit == last) release();
    uint8_t* myrow = sl + lane * 128;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4* q = reinterpret_cast<float4*>(myrow + ((j ^ (lane & 7)) << 4));
      float4 x = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      if (EPI == E_RES_ACT) {
        const float4 r = *q;
        x = make_float4(x.x + r.x, x.y + r.y, x.z + r.z, x.w + r.w);
      }
      *q = x;
    }
    __syncwarp();
    const int pc = lane & 7, rq = lane >> 3;
    const int colb = col0 + c0 + pc * 4;
    const int nval = (c0 + pc * 4 < W) ? max(0, min(4, N - colb)) : 0;
    const bool in_c = colb < cmax && c0 + pc * 4 < W;
    const bool edge = c0 + 32 > W || col0 + c0 + 32 > N;                  // warp-uniform: some lane has masked columns
    float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (bias != nullptr) {
      if (nval > 0) bv.x = __ldg(bias + colb);
      if (nval > 1) bv.y = __ldg(bias + colb + 1);
      if (nval > 2) bv.z = __ldg(bias + colb + 2);
      if (nval > 3) bv.w = __ldg(bias + colb + 3);
    }
    float4 xv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int rr = rq + 4 * i;
      xv[i] = *reinterpret_cast<const float4*>(sl + rr * 128 + ((pc ^ (rr & 7)) << 4));
    }
    // The slot is handed back BEFORE the global stores are issued: everything this warp needs is in registers now,
    // and fence.proxy.async behaves like a MEMBAR that waits for the thread's outstanding global accesses -- placed
    // after the stores (as it was) it serialised every 32-column chunk on the store latency (role wait-cycle
    // counters, profiles/r1_k_t4_role_wait_breakdown.log: the epilogue was 77 % busy and the MMA waited 20 % of the
    // kernel for a free accumulator).
    if (EPI == E_RES_ACT) {
      if (!(WD_ABL & 2048)) ptx::fence_proxy_async_smem();   // generic accesses of the slot are ordered before the next TMA write
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&bar->r_empty[slot]);
    } else {
      __syncwarp();                           // the next chunk's row owners overwrite this warp's staging rows
    }
    float* cp = C + (row0 + we * 32 + rq) * ldc + colb;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float4 x = xv[i];
      if (bias != nullptr) x = make_float4(x.x + bv.x, x.y + bv.y, x.z + bv.z, x.w + bv.w);       // warp-uniform
      if (EPI != E_STORE) x = f4_act<ACT>(x);
      if (EPI != E_STORE && drop.thresh != 0u) x = dropout4(x, drop, (row0 + we * 32 + rq + 4 * i) * ldc + colb);
      if (edge && nval < 4) { x.w = 0.f; if (nval < 3) x.z = 0.f; if (nval < 2) x.y = 0.f; if (nval < 1) x.x = 0.f; }
      if (in_c && rq + 4 * i < rows_ok && !(WD_ABL & 4)) *reinterpret_cast<float4*>(cp + (int64_t)(4 * i) * ldc) = x;
      if (sign.p != nullptr) sign_collect(sw[i], x, rq, pc, c0 >> 5);   // warp-uniform branch
    }
  }
  // lane pc holds the word of chunk pc: with two groups only if this group wrote it
  if (sign.p != nullptr && (T4_EPI_GROUPS == 1 || (int)((rit0 + (uint32_t)(lane & 7)) & 1u) == grp))
    sign_flush(sign, sw, sign_tile, row0 + we * 32, M, lane >> 3, lane & 7, (W + 31) >> 5);
}

// PAIR: the kernel runs as clusters of two CTAs on the two SMs of a TPC (cta_group::2).  A pair owns 256 rows: the even
// CTA (the leader) rows [0, 128), the odd one rows [128, 256); each loads, transforms and writes out its own rows exactly
// as a single CTA does.  What changes is the weight operand and the MMA: each CTA fetches only HALF of the rows of every
// weight unit (N / 2) -- the L2 -> SM weight stream, the shared-memory writes and the ring footprint per SM halve, so the
// ring holds four K chunks' worth instead of two -- and the leader issues ONE M = 256 tcgen05.mma.cta_group::2 per K
// step for both SMs, each tensor core reading both halves.  Signals that feed the issuing warp (operand ready,
// accumulator free) go to the LEADER's barriers from both CTAs; the weight boxes arrive by cta_group::2 TMA loads that
// count their bytes on the leader's barrier; completion (stage free, accumulator full) is committed to both CTAs at once.
// Why: profiles/r2_fwd_abl.log -- in the 3xTF32 mode the weight stream (two packed images per K chunk, 778 KB from L2
// per 128-row tile) costs 72 us of the 324 us depth-step contraction and the kernel moves ~3.4 MB through shared memory
// per tile.
template <int AMODE, int EPI, int ACT, bool PREACT, bool PAIR>
__device__ __forceinline__ void t4_body(const T4Params& p) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  constexpr bool GATHER = AMODE == A_GATHER_DIFF;
  constexpr int TILE_ROWS = PAIR ? 2 * TC_BM : TC_BM;                  // rows of one scheduled tile (a pair's or a CTA's)
  const uint32_t rank = PAIR ? ptx::cluster_ctarank() : 0u;            // 0: leader (issues the MMAs)
  const int64_t t_first = PAIR ? (blockIdx.x >> 1) : blockIdx.x, t_step = PAIR ? (gridDim.x >> 1) : gridDim.x;
  const int64_t rank_rows = (int64_t)rank * TC_BM;
  static_assert(GATHER || !PREACT, "PREACT belongs to the gathered operand");
  extern __shared__ uint8_t smem_raw[];
  const int warp = ptx::uniform_warp_idx(), lane = threadIdx.x & 31;   // provably warp-uniform: see tc_ptx.cuh
  const uint32_t base = ptx::smem_u32(smem_raw);
  uint8_t* ringA = smem_raw + (((base + 1023u) & ~1023u) - base);      // 1024-aligned
  const int b_stage = PAIR ? p.BN * 64 : p.BN * 128;                   // one chunk image (PAIR: this CTA's half of its rows)
  constexpr int B_STAGES = PAIR ? 2 * T4_B_STAGES : T4_B_STAGES;       // same ring bytes either way
  uint8_t* ringB = ringA + T4_A_STAGES * T4_A_STAGE;
  uint8_t* rring = ringB + T4_B_STAGES * (p.BN * 128);
  T4Barriers* bar = reinterpret_cast<T4Barriers*>(rring + T4_R_SLOTS * T4_R_SLOT);
  long long tacc[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  const long long t_start = clock64();

  if (threadIdx.x == 0) {
    // arrivals on raw_full: the 128 loader threads' cp.async groups (+ the expect_tx arrival of the X1 box), or the
    // single expect_tx arrival of the PLAIN operand's TMA box
    const uint32_t raw_count = GATHER ? 32 * T4_LOAD_WARPS + (p.x1_tma ? 1 : 0) : (p.a_tma ? 1 : 32 * T4_LOAD_WARPS);
    for (int s = 0; s < T4_A_STAGES; ++s) {
      ptx::mbar_init(&bar->raw_full[s], raw_count);
      ptx::mbar_init(&bar->op_full[s], T4_XFORM_WARPS * (PAIR ? 2 : 1));     // PAIR: both CTAs report to the leader
      ptx::mbar_init(&bar->a_empty[s], 1);
      ptx::mbar_init(&bar->x0_free[s], T4_XFORM_WARPS);
    }
    for (int s = 0; s < T4_B_SLOTS; ++s) { ptx::mbar_init(&bar->b_full[s], 1); ptx::mbar_init(&bar->b_empty[s], 1); }
    for (int s = 0; s < T4_R_SLOTS; ++s) {
      ptx::mbar_init(&bar->r_full[s], 1);
      ptx::mbar_init(&bar->r_full[s + T4_R_SLOTS], 1);
      ptx::mbar_init(&bar->r_empty[s], T2_EPI_WARPS);
    }
    for (int a = 0; a < 3; ++a) {
      ptx::mbar_init(&bar->tmem_full[a], 1);
      ptx::mbar_init(&bar->tmem_empty[a], T2_EPI_WARPS * T4_EPI_GROUPS * (PAIR ? 2 : 1));
    }
    ptx::fence_barrier_init();
    if (EPI == E_RES_ACT) ptx::tma_prefetch_desc(&p.mr);
    if (GATHER && p.x1_tma) ptx::tma_prefetch_desc(&p.mx1);
    if (!GATHER && p.a_tma) ptx::tma_prefetch_desc(&p.ma[0]);
  }
  if (warp == T4_MMA_WARP) {
    if (PAIR) ptx::tmem_alloc_pair(&bar->tmem_base, TC_TMEM_COLS);
    else ptx::tmem_alloc(&bar->tmem_base, TC_TMEM_COLS);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (PAIR) ptx::cluster_sync();       // the partner's barriers are initialised before anything arrives on them
  ptx::tc_fence_after();
  const uint32_t tmem_base = bar->tmem_base;
  const int64_t total = ceil_div(p.M, TILE_ROWS) * p.n_tiles;
  const int W0 = min(p.BN, T4_HALF), W1 = p.BN - W0;                  // accumulator halves of a column tile
  // The weight ring holds T4_B_STAGES chunk images [BN x 128 bytes]; it is filled and drained in UNITS of one
  // accumulator half of one image (hi, then lo when split), each unit with its own full / empty barrier: unit u
  // lives in slot u % (2 nh) = (stage, half).
  const int nh = W1 > 0 ? 2 : 1, parts = p.split == 1 ? 2 : 1, nslots = B_STAGES * nh;
  const int h1_off = PAIR ? W0 * 64 : W0 * 128;                        // where the second half's unit starts inside a stage
  // 3xBF16: the MMAs read only the [lo | hi] image (the subtrahend part of the stage); the raw-row part is dead once the
  // transform warps have read it, so its loads for the stage's NEXT use are issued as soon as the transform is done
  // (barrier x0_free) instead of after the MMAs of this use (a_empty) -- the gathered rows are the slowest load of the
  // operand path.  The subtrahend box (GATHER) still waits for a_empty; it is issued by lane 1 of the weight-loader warp.
  const bool early_x0 = p.split == 2 && (GATHER ? p.x1_tma != 0 : p.a_tma != 0) && !(WD_ABL & 8192);

  if (warp >= T4_XFORM_W0 && warp < T4_EPI2_W0) {
    // ================= transform, in place: X0 <- tf32(a), X1 part <- tf32(a - tf32(a)),  a = X0 - act?(X1) =================
    // (Fetching the subtrahend rows with plain loads into registers instead was measured SLOWER: the proxy fence
    //  below is a MEMBAR that also waits for the thread's outstanding global loads, i.e. for the prefetch.)
    const int t = threadIdx.x - 32 * T4_XFORM_W0, c = t & 7, rb = t >> 3;
    const uint32_t po = (uint32_t)(rb * 128 + ((c ^ (rb & 7)) << 4));   // rows rb + T4_XR j share rb & 7
    uint32_t it = 0;
    for (int64_t tile = t_first; tile < total; tile += t_step) {
      const int64_t row0 = (tile / p.n_tiles) * TILE_ROWS + rank_rows;
      const bool keep = GATHER && p.S != nullptr && (tile % p.n_tiles) == 0;
      uint32_t uoff[T4_XJ];                       // where this thread's subtrahend pieces sit in the X1 part of a stage
#pragma unroll
      for (int j = 0; j < T4_XJ; ++j) {
        if (GATHER && p.x1_tma) {                 // box row of the partner bond: rev(r) - (row0 - T4_X1_LEAD)
          const int64_t r = row0 + rb + T4_XR * j;
          const int64_t rv = r >= 1 ? ((r - 1) ^ 1) + 1 : 0;
          const int pr = (int)(rv - row0) + T4_X1_LEAD;            // in [1, 130] by construction
          uoff[j] = (uint32_t)(pr * 128 + ((c ^ (pr & 7)) << 4));
        } else {
          uoff[j] = po + (uint32_t)(j * (T4_XR * 128));
        }
      }
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int s = it % T4_A_STAGES;
        T2_WAIT(0, &bar->raw_full[s], (it / T4_A_STAGES) & 1);
        // early refill, plain operand: nothing but the raw box gates raw_full, so the image part of the stage (still
        // read by the MMAs of its previous use) is waited for here
        if (!GATHER && early_x0) T2_WAIT(1, &bar->a_empty[s], ((it / T4_A_STAGES) & 1) ^ 1);
        uint8_t* st = ringA + s * T4_A_STAGE + po;
        uint8_t* x1 = ringA + s * T4_A_STAGE + TC_A_STAGE;
        if (WD_ABL & 1024) {      // timing ablation: no transform work
          __syncwarp();
          if (lane == 0) { if (PAIR) ptx::mbar_arrive_leader(&bar->op_full[s]); else ptx::mbar_arrive(&bar->op_full[s]); }
          if (lane == 0 && early_x0) ptx::mbar_arrive(&bar->x0_free[s]);
          continue;
        }
        float4 v[T4_XJ], u[T4_XJ];
#pragma unroll
        for (int j = 0; j < T4_XJ; ++j) {
          v[j] = *reinterpret_cast<const float4*>(st + j * (T4_XR * 128));
          if (GATHER) u[j] = *reinterpret_cast<const float4*>(x1 + uoff[j]);
        }
        // a_lo goes where the subtrahend rows are: with the shared TMA box another thread's partner row lies under
        // this thread's a_lo position, so every transform thread must have read its pieces first
        if (GATHER && p.split && p.x1_tma) ptx::bar_sync(1, 32 * T4_XFORM_WARPS);
#pragma unroll
        for (int j = 0; j < T4_XJ; ++j) {
          if (GATHER) {
            if (PREACT) u[j] = f4_act<ACT>(u[j]);
            v[j] = make_float4(v[j].x - u[j].x, v[j].y - u[j].y, v[j].z - u[j].z, v[j].w - u[j].w);
          }
          if (p.split == 2) {   // 3xBF16: one image, [a_lo | a_hi] with a_hi = bf16(a); the raw rows stay as they are
            const uint32_t h01 = ptx::pack_bf16(v[j].x, v[j].y), h23 = ptx::pack_bf16(v[j].z, v[j].w);
            const float4 hi = make_float4(__uint_as_float(h01 << 16), __uint_as_float(h01 & 0xffff0000u),
                                          __uint_as_float(h23 << 16), __uint_as_float(h23 & 0xffff0000u));
            uint8_t* row = x1 + (rb + T4_XR * j) * 128 + (c & 1) * 8;
            *reinterpret_cast<uint2*>(row + (((c >> 1) ^ (rb & 7)) << 4)) =
                make_uint2(ptx::pack_bf16(v[j].x - hi.x, v[j].y - hi.y), ptx::pack_bf16(v[j].z - hi.z, v[j].w - hi.w));
            *reinterpret_cast<uint2*>(row + (((4 + (c >> 1)) ^ (rb & 7)) << 4)) = make_uint2(h01, h23);
            continue;
          }
          const float4 hi = make_float4(ptx::to_tf32(v[j].x), ptx::to_tf32(v[j].y), ptx::to_tf32(v[j].z), ptx::to_tf32(v[j].w));
          *reinterpret_cast<float4*>(st + j * (T4_XR * 128)) = hi;
          if (p.split) {     // correction image row: 64 bf16 = [a_lo(k = 0..31) | a_hi(k = 0..31)], SWIZZLE_128B
            uint8_t* row = x1 + (rb + T4_XR * j) * 128 + (c & 1) * 8;
            *reinterpret_cast<uint2*>(row + (((c >> 1) ^ (rb & 7)) << 4)) =
                make_uint2(ptx::pack_bf16(v[j].x - hi.x, v[j].y - hi.y), ptx::pack_bf16(v[j].z - hi.z, v[j].w - hi.w));
            *reinterpret_cast<uint2*>(row + (((4 + (c >> 1)) ^ (rb & 7)) << 4)) =
                make_uint2(ptx::pack_bf16(hi.x, hi.y), ptx::pack_bf16(hi.z, hi.w));
          }
        }
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) { if (PAIR) ptx::mbar_arrive_leader(&bar->op_full[s]); else ptx::mbar_arrive(&bar->op_full[s]); }
        if (lane == 0 && early_x0) ptx::mbar_arrive(&bar->x0_free[s]);   // the raw rows may be refilled while the MMAs run
        const int kcol = kc * TC_BK + c * 4;                             // S goes out after the hand-off to the MMA
        if (keep && kcol < p.lds) {
#pragma unroll
          for (int j = 0; j < T4_XJ; ++j) {
            const int64_t r = row0 + rb + T4_XR * j;
            if (r < p.M) *reinterpret_cast<float4*>(p.S + r * p.lds + kcol) = v[j];
          }
        }
      }
    }
  } else if (warp >= T4_LOAD_W0 && warp < T4_XFORM_W0) {
    const int t = threadIdx.x - 32 * T4_LOAD_W0;
    const uint32_t ring_u32 = ptx::smem_u32(ringA);
    if (GATHER) {
      // ================= gather loaders: cp.async, 8 threads per row, rows rb + 16 j =================
      const int c = t & 7, rb = t >> 3;
      const uint32_t po = (uint32_t)(rb * 128 + ((c ^ (rb & 7)) << 4));
      uint32_t it = 0;
      for (int64_t tile = t_first; tile < total; tile += t_step) {
        const int64_t row0 = (tile / p.n_tiles) * TILE_ROWS + rank_rows;
        const float* s0[8];
        const float* s1[8];
        uint32_t okmask = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int64_t r = row0 + rb + 16 * j;
          const bool ok = r < p.M;
          okmask |= (ok ? 1u : 0u) << j;
          int64_t a = ok ? r : 0, b = a;
          if (ok && p.i0) a = __ldg(p.i0 + r);
          if (ok && p.i1 && !p.x1_tma) b = __ldg(p.i1 + r);
          s0[j] = p.X0 + a * p.ld0 + c * 4;
          s1[j] = p.X1 + b * p.ld1 + c * 4;
        }
        for (int kc = 0; kc < p.KC; ++kc, ++it) {
          const int s = it % T4_A_STAGES;
          T2_WAIT(2, early_x0 ? &bar->x0_free[s] : &bar->a_empty[s], ((it / T4_A_STAGES) & 1) ^ 1);
          const int k = kc * TC_BK + c * 4;
          const int rem = (p.K - k) * 4;
          const uint32_t valid = rem >= 16 ? 16u : (rem > 0 ? (uint32_t)rem : 0u);
          const uint32_t d0 = ring_u32 + s * T4_A_STAGE + po;
          if (WD_ABL & 32) {                        // timing ablation: no operand copies
            if (p.x1_tma && t == 0) ptx::mbar_arrive(&bar->raw_full[s]);
          } else if (p.x1_tma) {                    // subtrahend rows: one TMA box; minuend rows: gathered cp.async
            if (t == 0 && !early_x0) {
              ptx::mbar_arrive_expect_tx(&bar->raw_full[s], (uint32_t)(T4_X1_ROWS * 128));
              ptx::tma_load_2d(ringA + s * T4_A_STAGE + TC_A_STAGE, &p.mx1, kc * TC_BK, (int)row0 - T4_X1_LEAD,
                               &bar->raw_full[s]);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t v = ((okmask >> j) & 1u) ? valid : 0u;
              ptx::cp_async16(d0 + j * 2048, v ? s0[j] + kc * TC_BK : p.X0, v);
            }
          } else if (okmask == 0xFFu && valid == 16u) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              ptx::cp_async16(d0 + j * 2048, s0[j] + kc * TC_BK, 16u);
              ptx::cp_async16(d0 + j * 2048 + TC_A_STAGE, s1[j] + kc * TC_BK, 16u);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t v = ((okmask >> j) & 1u) ? valid : 0u;
              ptx::cp_async16(d0 + j * 2048, v ? s0[j] + kc * TC_BK : p.X0, v);
              ptx::cp_async16(d0 + j * 2048 + TC_A_STAGE, v ? s1[j] + kc * TC_BK : p.X1, v);
            }
          }
          ptx::cp_async_arrive(&bar->raw_full[s]);
        }
      }
    } else if (p.a_tma) {
      // ================= plain operand: one TMA box (128 rows x 32 columns, raw fp32) per K chunk =================
      if (t == 0) {
        uint32_t it = 0;
        for (int64_t tile = t_first; tile < total; tile += t_step) {
          const int row0 = (int)((tile / p.n_tiles) * TILE_ROWS + rank_rows);
          for (int kc = 0; kc < p.KC; ++kc, ++it) {
            const int s = it % T4_A_STAGES;
            T2_WAIT(2, early_x0 ? &bar->x0_free[s] : &bar->a_empty[s], ((it / T4_A_STAGES) & 1) ^ 1);
            const bool second = kc >= p.KC0;
            ptx::mbar_arrive_expect_tx(&bar->raw_full[s], (uint32_t)TC_A_STAGE);
            ptx::tma_load_2d(ringA + s * T4_A_STAGE, &p.ma[second ? 1 : 0], (second ? kc - p.KC0 : kc) * TC_BK, row0,
                             &bar->raw_full[s]);
          }
        }
      }
    } else {
      // ================= plain operand, rows only 4-byte aligned (dense 133 / 147-float feature rows) =================
      const int col = t & 31, r4 = t >> 5;
      uint32_t it = 0;
      for (int64_t tile = t_first; tile < total; tile += t_step) {
        const int64_t row0 = (tile / p.n_tiles) * TILE_ROWS + rank_rows;
        for (int kc = 0; kc < p.KC; ++kc, ++it) {
          const int s = it % T4_A_STAGES;
          T2_WAIT(2, &bar->a_empty[s], ((it / T4_A_STAGES) & 1) ^ 1);
          const bool second = kc >= p.KC0;
          const float* X = second ? p.X1 : p.X0;
          const int ld = second ? p.ld1 : p.ld0, Ks = second ? p.K2 : p.K;
          const int k = (second ? kc - p.KC0 : kc) * TC_BK + col;
          const bool vk = k < Ks;
          const uint32_t d0 = ring_u32 + s * T4_A_STAGE + (uint32_t)((col & 3) * 4);
#pragma unroll 8
          for (int j = 0; j < TC_BM / T4_LOAD_WARPS; ++j) {
            const int rl = r4 + T4_LOAD_WARPS * j;
            const int64_t r = row0 + rl;
            const uint32_t v = (vk && r < p.M) ? 4u : 0u;
            ptx::cp_async4(d0 + ptx::sw128_off(rl, col >> 2), v ? X + r * ld + k : X, v);
          }
          ptx::cp_async_arrive(&bar->raw_full[s]);
        }
      }
    }
  } else if (warp == T4_R_WARP) {
    // ================= residual loader: TMA boxes of 128 rows x 32 columns, ahead of the epilogue =================
    if (EPI == E_RES_ACT && lane == 0) {
      uint32_t rit = 0;
      for (int64_t tile = t_first; tile < total; tile += t_step) {
        const int row0 = (int)((tile / p.n_tiles) * TILE_ROWS + rank_rows);
        const int col0 = (int)(tile % p.n_tiles) * p.BN;
        for (int c0 = 0; c0 < p.BN; c0 += 32, ++rit) {
          // the halves are walked in order: columns [0, W0) then [W0, BN); W0 is a multiple of 32 whenever W1 > 0
          const int slot = rit % T4_R_SLOTS;
          T2_WAIT(9, &bar->r_empty[slot], ((rit / T4_R_SLOTS) & 1) ^ 1);
          ptx::mbar_arrive_expect_tx(&bar->r_full[t4_rfull(rit)], (uint32_t)T4_R_SLOT);
          ptx::tma_load_2d(rring + slot * T4_R_SLOT, &p.mr, col0 + c0, row0, &bar->r_full[t4_rfull(rit)]);
        }
      }
    }
  } else if (warp == T4_B_WARP) {
    // ================= weight loader: one bulk copy per unit (K chunk, hi | lo image, accumulator half) =================
    if (lane == 0) {
      uint32_t u = 0;
      for (int64_t tile = t_first; tile < total; tile += t_step) {
        const int nt = (int)(tile % p.n_tiles);
        for (int kc = 0; kc < p.KC; ++kc) {
          for (int part = 0; part < parts; ++part) {
            const float* img = p.Bp + (((int64_t)nt * p.KC + kc) * parts + part) * p.BN * 32;
            if (T4_MERGE && !PAIR) {          // one copy per image: both accumulator halves' rows are contiguous
              const int slot = u % B_STAGES;
              T2_WAIT(3, &bar->b_empty[slot], ((u / B_STAGES) & 1) ^ 1);
              if (WD_ABL & 4096) { ptx::mbar_arrive(&bar->b_full[slot]); ++u; continue; }     // timing ablation: no weight copies
              ptx::mbar_arrive_expect_tx(&bar->b_full[slot], (uint32_t)(p.BN * 128));
              ptx::bulk_g2s(ringB + slot * b_stage, img, (uint32_t)(p.BN * 128), &bar->b_full[slot]);
              ++u;
              continue;
            }
            for (int hb = 0; hb < nh; ++hb, ++u) {
              const int slot = u % nslots;
              T2_WAIT(3, &bar->b_empty[slot], ((u / nslots) & 1) ^ 1);
              const uint32_t bytes = (uint32_t)((hb ? W1 : W0) * 128);
              if (PAIR) {
                // this CTA's half of the unit's rows; both boxes count on the leader's barrier, which the leader arms
                const int wh = (hb ? W1 : W0) / 2;
                const int row = (int)((((int64_t)nt * p.KC + kc) * parts + part) * p.BN) + (hb ? W0 : 0) + (int)rank * wh;
                if (rank == 0) ptx::mbar_arrive_expect_tx(&bar->b_full[slot], bytes);
                ptx::tma_load_2d_pair(ringB + (slot / nh) * b_stage + (hb ? h1_off : 0), &p.mb[hb], 0, row, &bar->b_full[slot]);
                continue;
              }
              if (WD_ABL & 4096) { ptx::mbar_arrive(&bar->b_full[slot]); continue; }     // timing ablation: no weight copies
              ptx::mbar_arrive_expect_tx(&bar->b_full[slot], bytes);
              ptx::bulk_g2s(ringB + (slot / nh) * b_stage + (hb ? h1_off : 0), img + (hb ? W0 * 32 : 0), bytes,
                            &bar->b_full[slot]);
            }
          }
        }
      }
    } else if (lane == 1 && GATHER && early_x0) {
      // ================= subtrahend boxes of the early-refill scheme: these still wait for the MMAs (a_empty) =================
      uint32_t it = 0;
      for (int64_t tile = t_first; tile < total; tile += t_step) {
        const int64_t row0 = (tile / p.n_tiles) * TILE_ROWS + rank_rows;
        for (int kc = 0; kc < p.KC; ++kc, ++it) {
          const int s = it % T4_A_STAGES;
          ptx::mbar_wait(&bar->a_empty[s], ((it / T4_A_STAGES) & 1) ^ 1);
          ptx::mbar_arrive_expect_tx(&bar->raw_full[s], (uint32_t)(T4_X1_ROWS * 128));
          ptx::tma_load_2d(ringA + s * T4_A_STAGE + TC_A_STAGE, &p.mx1, kc * TC_BK, (int)row0 - T4_X1_LEAD, &bar->raw_full[s]);
        }
      }
    }
  } else if (warp == T4_MMA_WARP) {
    // ================= MMA issue: two accumulator halves per tile, three rotating TMEM buffers =================
    uint32_t it = 0, g = 0, u = 0;                                      // g: accumulator halves issued so far; u: weight units
    const uint32_t ringA_u = ptx::uniform_u32(ptx::smem_u32(ringA)), ringB_u = ptx::uniform_u32(ptx::smem_u32(ringB));
    const uint32_t tmem_u = ptx::uniform_u32(tmem_base);
    constexpr int MM = PAIR ? 256 : 128;                                // rows of one MMA (PAIR: both CTAs' tiles)
    const uint32_t idh[2] = {ptx::idesc_tf32(MM, W0), ptx::idesc_tf32(MM, W1 > 0 ? W1 : 16)};
    const uint32_t idc[2] = {ptx::idesc_bf16(MM, W0), ptx::idesc_bf16(MM, W1 > 0 ? W1 : 16)};
    for (int64_t tile = (PAIR && rank != 0) ? total : t_first; tile < total; tile += t_step) {     // PAIR: the leader issues
      const uint32_t g0 = g;
      g += nh;
      for (int kc = 0; kc < p.KC; ++kc, ++it) {
        const int s = it % T4_A_STAGES;
        T2_WAIT(5, &bar->op_full[s], (it / T4_A_STAGES) & 1);
        const uint32_t a_hi_lo = ptx::desc_lo_sw128(ringA_u + s * T4_A_STAGE);           // low descriptor words of the
        const uint32_t a_lo_lo = ptx::desc_lo_sw128(ringA_u + s * T4_A_STAGE + TC_A_STAGE);   // a_hi / correction images
        // Issue order: consecutive MMAs alternate between the two accumulator halves (independent chains; measured
        // 3 % faster than half by half).  part 0: a_hi w_hi in kind::tf32.  part 1 (3xTF32): ONE kind::f16 (bf16)
        // group over the correction images, K = 64 per chunk: [bf16 a_lo | bf16 a_hi] . [bf16 w_hi | bf16 w_lo].
        for (int part = 0; part < parts; ++part) {
          uint32_t b_addr[2] = {0u, 0u}, d[2] = {0u, 0u};
          int slots[2] = {0, 0};
          if (T4_MERGE && !PAIR) {
            const int slot = u % B_STAGES;
            T2_WAIT(6, &bar->b_full[slot], (u / B_STAGES) & 1);
            ++u;
            for (int hb = 0; hb < nh; ++hb) {
              slots[hb] = slot;
              const uint32_t gb = g0 + hb;
              if (kc == 0 && part == 0) T2_WAIT(4, &bar->tmem_empty[gb % 3], ((gb / 3) & 1) ^ 1);
              b_addr[hb] = ptx::desc_lo_sw128(ringB_u + slot * b_stage + (hb ? h1_off : 0));
              d[hb] = tmem_u + (gb % 3) * T4_HALF;
            }
          } else
          for (int hb = 0; hb < nh; ++hb, ++u) {
            const int slot = u % nslots;
            slots[hb] = slot;
            T2_WAIT(6, &bar->b_full[slot], (u / nslots) & 1);
            const uint32_t gb = g0 + hb;
            if (kc == 0 && part == 0) T2_WAIT(4, &bar->tmem_empty[gb % 3], ((gb / 3) & 1) ^ 1);
            b_addr[hb] = ptx::desc_lo_sw128(ringB_u + (slot / nh) * b_stage + (hb ? h1_off : 0));
            d[hb] = tmem_u + (gb % 3) * T4_HALF;
          }
          ptx::tc_fence_after();
          // issued from convergent code, the election is inside each statement (tc_ptx.cuh: mma_elect)
          if (p.split == 2) {       // 3xBF16: a_lo w_hi + a_hi w_lo (the row's 64 bf16), then a_hi w_hi (row halves paired crosswise)
#pragma unroll
            for (int ks = 0; ks < 6; ++ks)
              for (int hb = 0; hb < nh; ++hb) {
                const uint32_t ao = ks < 4 ? 2 * ks : 4 + 2 * (ks - 4), bo = ks < 4 ? 2 * ks : 2 * (ks - 4);
                const uint32_t acc = (kc == 0 && ks == 0) ? 0u : 1u;
                if (WD_ABL & 128) {
                } else if (PAIR) {
                  ptx::mma_elect_pair<true>(d[hb], a_lo_lo + ao, b_addr[hb] + bo, ptx::kDescHiSw128, idc[hb], acc);
                } else {
                  ptx::mma_elect<true>(d[hb], a_lo_lo + ao, b_addr[hb] + bo, ptx::kDescHiSw128, idc[hb], acc);
                }
              }
          } else if (part == 0) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
              for (int hb = 0; hb < nh; ++hb)
                if (WD_ABL & 128) {
                } else if (PAIR) {
                  ptx::mma_elect_pair<false>(d[hb], a_hi_lo + 2 * ks, b_addr[hb] + 2 * ks, ptx::kDescHiSw128, idh[hb],
                                             (kc == 0 && ks == 0) ? 0u : 1u);
                } else {
                  ptx::mma_elect<false>(d[hb], a_hi_lo + 2 * ks, b_addr[hb] + 2 * ks, ptx::kDescHiSw128, idh[hb],
                                        (kc == 0 && ks == 0) ? 0u : 1u);
                }
          } else {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
              for (int hb = 0; hb < nh; ++hb)
                if (WD_ABL & (128 | 512)) {
                } else if (PAIR) {
                  ptx::mma_elect_pair<true>(d[hb], a_lo_lo + 2 * ks, b_addr[hb] + 2 * ks, ptx::kDescHiSw128, idc[hb], 1u);
                } else {
                  ptx::mma_elect<true>(d[hb], a_lo_lo + 2 * ks, b_addr[hb] + 2 * ks, ptx::kDescHiSw128, idc[hb], 1u);
                }
          }
          // (PAIR: every completion below is committed to the barrier at the same offset in BOTH CTAs)
          if (T4_MERGE && !PAIR) ptx::tc_commit_elect(&bar->b_empty[slots[0]]);
          else
          for (int hb = 0; hb < nh; ++hb) { if (PAIR) ptx::tc_commit_elect_pair(&bar->b_empty[slots[hb]]); else ptx::tc_commit_elect(&bar->b_empty[slots[hb]]); }
          __syncwarp();
        }
        if (PAIR) ptx::tc_commit_elect_pair(&bar->a_empty[s]); else ptx::tc_commit_elect(&bar->a_empty[s]);
        if (kc == p.KC - 1) {
          if (PAIR) {
            ptx::tc_commit_elect_pair(&bar->tmem_full[g0 % 3]);
            if (nh > 1) ptx::tc_commit_elect_pair(&bar->tmem_full[(g0 + 1) % 3]);
          } else {
            ptx::tc_commit_elect(&bar->tmem_full[g0 % 3]);
            if (nh > 1) ptx::tc_commit_elect(&bar->tmem_full[(g0 + 1) % 3]);
          }
        }
        __syncwarp();
      }
    }
  } else {
    // ================= epilogue (warps 0-3; with two groups also the four warps after the transform warps) =================
    uint32_t g = 0, rit = 0;
    const int grp = warp >= T4_EPI2_W0 ? 1 : 0, wq = warp & 3;             // TMEM lane quarter / rows of the tile: warp % 4
    for (int64_t tile = t_first; tile < total; tile += t_step) {
      const int64_t row0 = (tile / p.n_tiles) * TILE_ROWS + rank_rows;
      const int col0 = (int)(tile % p.n_tiles) * p.BN;
      for (int hb = 0; hb < nh; ++hb, ++g) {
        const int buf = g % 3;
        T2_WAIT(7, &bar->tmem_full[buf], (g / 3) & 1);
        ptx::tc_fence_after();
        // (the accumulator is handed back inside, after the warp's last tcgen05.ld of the half)
        t4_epilogue<EPI, ACT, PAIR>(tmem_base + buf * T4_HALF, wq, lane, rring, bar, rit, hb == 0 ? W0 : W1, row0, p.M,
                                    col0 + hb * W0, p.N, p.ldc, p.bias, p.C, p.ldc, tacc, p.drop, p.sign,
                                    (int)(tile % p.n_tiles) * p.sign.tpn + hb, grp, &bar->tmem_empty[buf]);
      }
    }
  }
#ifdef WD_T2_TIMING
  if (WD_TIMING_ON(4) && blockIdx.x == 0 && lane == 0) {
    if (warp == 0 || warp == T4_MMA_WARP || warp == T4_B_WARP || warp == T4_R_WARP || warp == T4_LOAD_W0 || warp == T4_XFORM_W0)
      for (int i = 0; i < 10; ++i) if (tacc[i]) atomicAdd(&g_t2_timing[i], (unsigned long long)tacc[i]);
    if (warp == 0) atomicAdd(&g_t2_timing[15], (unsigned long long)(clock64() - t_start));
  }
#else
  (void)tacc; (void)t_start;
#endif
  ptx::tc_fence_before();
  __syncthreads();
  if (PAIR) ptx::cluster_sync();       // the partner's shared memory and barriers stay alive until both CTAs are done
  if (warp == T4_MMA_WARP) {
    ptx::tc_fence_after();
    if (PAIR) ptx::tmem_dealloc_pair(tmem_base, TC_TMEM_COLS);
    else ptx::tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

template <int AMODE, int EPI, int ACT, bool PREACT>
__global__ void __launch_bounds__(T4_THREADS, 1) k_tc_gemm4(const __grid_constant__ T4Params p) {
  t4_body<AMODE, EPI, ACT, PREACT, false>(p);
}
// the CTA-pair form: clusters of two CTAs, grid size even
template <int AMODE, int EPI, int ACT, bool PREACT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(T4_THREADS, 1) k_tc_gemm4_pair(const __grid_constant__ T4Params p) {
  t4_body<AMODE, EPI, ACT, PREACT, true>(p);
}

}  // namespace wd
