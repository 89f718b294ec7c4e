// Host-side launch helpers of the tcgen05 contraction kernels (tc_gemm3.cuh / tc_gemm4.cuh / tc_dw4.cuh / tc_dw5.cuh).
#pragma once
#include <map>
#include <mutex>
#include <utility>
#include "tc_gemm3.cuh"
#include "tc_gemm4.cuh"
#include "tc_dw4.cuh"
#include "tc_dw5.cuh"

namespace wd {

// A packed weight operand: geometry + device image (k_tc_pack_b layout).
struct TcPacked {
  int N, K1, K2, BN, n_tiles, KC1, KC;
  int parts;   // 1: tf32(w); 2: hi and lo images per (column tile, K chunk) for the 3xTF32 kernels
  float* img;
  int bf = 0;  // 1 (with parts == 1): the image is the bf16 [hi | lo] image of the 3xBF16 kernels (tc_core.cuh)
  int64_t floats() const { return (int64_t)n_tiles * KC * BN * 32 * parts; }
};

// max_bn: widest column tile the consuming kernel takes (160: k_tc_gemm3's two TMEM accumulators; 304: k_tc_gemm4)
inline TcPacked tc_packed_geometry(int N, int K1, int K2, int max_bn = T2_MAX_BN, int parts = 1, int bf = 0) {
  const TcTiling t = tc_tiling(N, max_bn);
  TcPacked g{N, K1, K2, t.BN, t.n_tiles, (int)ceil_div(K1, TC_BK), 0, parts, nullptr, bf};
  g.KC = g.KC1 + (int)ceil_div(K2, TC_BK);
  return g;
}

// Collects up to three (weight matrix -> packed image) jobs and packs them with one launch.
struct TcPacker {
  TcPackJobs jobs{};
  void add(const float* W, int ldb, int bmode, const TcPacked& g) {
    TcPackJob& j = jobs.job[jobs.n++];
    j = TcPackJob{W, ldb, bmode, g.N, g.K1, g.K2, g.BN, g.n_tiles, g.KC1, g.KC, g.parts, g.img, jobs.total, g.bf};
    jobs.total += g.floats();
  }
  int run(cudaStream_t s) {
    if (jobs.n == 0) return WD_OK;
    WD_LAUNCH(k_tc_pack_b, (unsigned)ceil_div(jobs.total, 256), 256, 0, s, jobs);
    return WD_OK;
  }
};

inline int tc_pack(cudaStream_t s, const float* W, int ldb, int bmode, const TcPacked& g) {
  TcPacker pk;
  pk.add(W, ldb, bmode, g);
  return pk.run(s);
}

inline bool tc_aligned16(const void* p, int ld) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && (ld & 3) == 0; }

// The dynamic shared memory limit of a kernel is a sticky per-device attribute: set it once per (kernel, device, size)
// instead of on every launch (one driver call per tensor-core launch on the host's critical path at small batches).
template <typename Kernel>
inline int tc_set_smem(Kernel k, int bytes) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, int> done;    // (kernel, device) -> bytes already granted
  int dev = 0;
  WD_CUDA_OK(cudaGetDevice(&dev));
  const std::pair<const void*, int> key(reinterpret_cast<const void*>(k), dev);
  std::lock_guard<std::mutex> lock(mu);
  auto it = done.find(key);
  if (it != done.end() && it->second == bytes) return WD_OK;
  WD_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  done[key] = bytes;
  return WD_OK;
}

// The forward kernel of tc_gemm4.cuh: gathered depth step (single-pass or 3xTF32) and, with `split`, the plain /
// concatenated forward contractions; also the single-pass fallback for feature rows that are only 4-byte aligned.
template <int AMODE, int EPI, int ACT, bool PREACT>
int tc_gemm4(cudaStream_t s, const AOperand& A, const TcPacked& B, const float* bias, const float* R, int ldr,
             float* C, int ldc, int64_t M, const Dropout& drop, const SignBits& sign) {
  constexpr bool GATHER = AMODE == A_GATHER_DIFF;
  constexpr int ACT2 = EPI == E_STORE ? WD_ACT_RELU : ACT;     // the store epilogue applies no activation: one instance
  T4Params p{};
  p.drop = drop;
  p.sign = sign;
  p.split = B.bf ? 2 : (B.parts == 2 ? 1 : 0);
  if (B.BN > T4_MAX_BN) return fail(WD_ERR_UNSUPPORTED, "tcgen05 forward kernel: unsupported tile geometry");
  if (!tc_aligned16(C, ldc)) return fail(WD_ERR_INVALID_ARG, "tcgen05 output rows must be 16-byte aligned");
  int rc = WD_OK;
  if (EPI == E_RES_ACT) {
    if (!tc_aligned16(R, ldr)) return fail(WD_ERR_INVALID_ARG, "tcgen05 residual rows must be 16-byte aligned");
    rc = tma_make_2d(&p.mr, R, M, B.N, ldr, TC_BM, TMA_SW128, /*round_tf32=*/false);
    if (rc != WD_OK) return rc;
  }
  p.X0 = A.X; p.X1 = A.X2; p.ld0 = A.ldx; p.ld1 = A.ldx2; p.K = B.K1; p.K2 = B.K2; p.KC0 = B.KC1; p.KC = B.KC;
  if (GATHER) {
    if (B.K2 != 0) return fail(WD_ERR_UNSUPPORTED, "tcgen05 gathered step: one K segment only");
    if (!tc_aligned16(A.X, A.ldx) || !tc_aligned16(A.X2, A.ldx2))
      return fail(WD_ERR_INVALID_ARG, "tcgen05 gathered step needs 16-byte aligned rows");
    p.i0 = A.src; p.i1 = A.rev;
    p.S = A.S_out; p.lds = A.ldx2;
    static const bool x1_tma_off = [] { const char* e = getenv("WD_T4_X1_TMA"); return e && atoi(e) == 0; }();   // A/B runs
    p.x1_tma = (A.rev_adjacent && A.rev != nullptr && !x1_tma_off) ? 1 : 0;
    if (p.x1_tma) {
      rc = tma_make_2d(&p.mx1, A.X2, M, B.K1, A.ldx2, T4_X1_ROWS, TMA_SW128, /*round_tf32=*/false);
      if (rc != WD_OK) return rc;
    }
  } else {
    p.a_tma = tc_aligned16(A.X, A.ldx) && (B.K2 == 0 || tc_aligned16(A.X2, A.ldx2)) ? 1 : 0;
    if (p.a_tma) {
      rc = tma_make_2d(&p.ma[0], A.X, M, B.K1, A.ldx, TC_BM, TMA_SW128, /*round_tf32=*/false);
      if (rc != WD_OK) return rc;
      if (B.K2 > 0) {
        rc = tma_make_2d(&p.ma[1], A.X2, M, B.K2, A.ldx2, TC_BM, TMA_SW128, /*round_tf32=*/false);
        if (rc != WD_OK) return rc;
      }
    }
  }
  p.Bp = B.img; p.bias = bias; p.C = C; p.ldc = ldc; p.M = M;
  p.N = B.N; p.BN = B.BN; p.n_tiles = B.n_tiles;
  const int smem = t4_smem_bytes(B.BN);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 tile does not fit shared memory");
  // CTA pairs (cta_group::2, tc_gemm4.cuh: t4_body<..., PAIR>): OPT-IN (WD_T4_PAIR=1).  Parity-green and bitwise equal to
  // the single-CTA kernel, but measured 1-12 % SLOWER on B200 at both h = 300 and h = 2400 (profiles/r2_fwd_abl.log,
  // section F): halving the weight stream per SM buys less than the lock-step of two SMs costs.
  const char* pair_env = getenv("WD_T4_PAIR");           // read per call: the A/B test flips it
  const bool pair_on = pair_env && atoi(pair_env) == 1;
  const int W0 = std::min(B.BN, T4_HALF), W1 = B.BN - W0;
  if (pair_on && ceil_div(M, TC_BM) >= 2 * kNumSMs && W0 % 16 == 0 && W1 % 16 == 0) {
    const int64_t img_rows = (int64_t)B.n_tiles * B.KC * B.parts * B.BN;
    rc = tma_make_2d(&p.mb[0], B.img, img_rows, 32, 32, W0 / 2, TMA_NO_SWIZZLE, /*round_tf32=*/false);
    if (rc == WD_OK && W1 > 0) rc = tma_make_2d(&p.mb[1], B.img, img_rows, 32, 32, W1 / 2, TMA_NO_SWIZZLE, /*round_tf32=*/false);
    if (rc != WD_OK) return rc;
    auto kern = k_tc_gemm4_pair<GATHER ? A_GATHER_DIFF : A_PLAIN, EPI, ACT2, GATHER && PREACT>;
    rc = tc_set_smem(kern, smem);
    if (rc != WD_OK) return rc;
    const int64_t tiles = ceil_div(M, 2 * TC_BM) * B.n_tiles;
    const unsigned grid = 2u * (unsigned)std::min<int64_t>(tiles, kNumSMs / 2);
    WD_LAUNCH(kern, grid, T4_THREADS, smem, s, p);
    return WD_OK;
  }
  auto kern = k_tc_gemm4<GATHER ? A_GATHER_DIFF : A_PLAIN, EPI, ACT2, GATHER && PREACT>;
  rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  const int64_t tiles = ceil_div(M, TC_BM) * B.n_tiles;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, kNumSMs);
  WD_LAUNCH(kern, grid, T4_THREADS, smem, s, p);
  return WD_OK;
}

// TMA-fed kernel for plain / concatenated A operands, single-pass TF32 (rounding in the TMA engine), no transform pass
template <int AMODE, int EPI, int ACT>
int tc_gemm3(cudaStream_t s, const AOperand& A, const TcPacked& B, const float* bias, float* C, int ldc, int64_t M,
             const Dropout& drop, const SignBits& sign) {
  constexpr int ACT2 = EPI == E_STORE ? WD_ACT_RELU : ACT;
  T3Params p{};
  p.drop = drop;
  p.sign = sign;
  int rc = tma_make_2d(&p.ma[0], A.X, M, B.K1, A.ldx, TC_BM, TMA_SW128);
  if (rc != WD_OK) return rc;
  if (AMODE == A_CONCAT) {
    rc = tma_make_2d(&p.ma[1], A.X2, M, B.K2, A.ldx2, TC_BM, TMA_SW128);
    if (rc != WD_OK) return rc;
  }
  p.KC0 = B.KC1; p.KC = B.KC;
  p.Bp = B.img; p.bias = bias; p.C = C; p.ldc = ldc; p.M = M;
  p.N = B.N; p.BN = B.BN; p.n_tiles = B.n_tiles;
  const int smem = t3_smem_bytes(B.BN);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 tile does not fit shared memory");
  auto kern = k_tc_gemm3<EPI, ACT2>;
  rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  const int64_t tiles = ceil_div(M, TC_BM) * B.n_tiles;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, kNumSMs);
  WD_LAUNCH(kern, grid, T3_THREADS, smem, s, p);
  return WD_OK;
}

// C[M, ldc] = epi( A[M, K1+K2] * B^T ) with B packed.  B.parts == 2 selects the 3xTF32 kernel, B.bf the 3xBF16 form of it.
template <int AMODE, int EPI, int ACT, bool PREACT>
int tc_gemm(cudaStream_t s, const AOperand& A, const TcPacked& B, const float* bias, const float* R, int ldr,
            float* C, int ldc, int64_t M, const Dropout& drop = Dropout{}, const SignBits& sign = SignBits{}) {
  if constexpr (AMODE == A_GATHER_DIFF || EPI == E_RES_ACT) {
    return tc_gemm4<AMODE, EPI, ACT, PREACT>(s, A, B, bias, R, ldr, C, ldc, M, drop, sign);
  } else {
    const bool aligned = tc_aligned16(A.X, A.ldx) && tc_aligned16(C, ldc) && (AMODE != A_CONCAT || tc_aligned16(A.X2, A.ldx2));
    if (B.parts == 1 && !B.bf && B.BN <= T2_MAX_BN && aligned) return tc_gemm3<AMODE, EPI, ACT>(s, A, B, bias, C, ldc, M, drop, sign);
    return tc_gemm4<AMODE, EPI, ACT, false>(s, A, B, bias, R, ldr, C, ldc, M, drop, sign);
  }
}

// n2 tiles of a dW contraction: each segment split into tiles of <= D4_MAX_BN columns
inline int tc_dw_tiles(int w0, int w1, TcDwTile* out) {
  const int max_bn = std::min(tc_max_bn(), D4_MAX_BN);
  int n = 0;
  const int widths[2] = {w0, w1};
  int out0 = 0;
  for (int seg = 0; seg < 2; ++seg) {
    const int w = widths[seg];
    if (w <= 0) continue;
    TcTiling t;
    t.n_tiles = (int)ceil_div(round_up(w, 16), max_bn);
    t.BN = (int)round_up(ceil_div(round_up(w, 16), t.n_tiles), 16);
    for (int j = 0; j < t.n_tiles; ++j) {
      const int col0 = j * t.BN;
      const int width = std::min(t.BN, w - col0);
      if (width <= 0) break;
      if (n >= TC_DW_MAX_TILES) return -1;
      out[n++] = TcDwTile{seg, col0, width, (int)round_up(width, 16), out0 + col0};
    }
    out0 += w;
  }
  return n;
}

inline int tc_dw_count_tiles(int w0, int w1) {
  TcDwTile tmp[TC_DW_MAX_TILES];
  return tc_dw_tiles(w0, w1, tmp);
}

// number of row chunks so that the grid roughly fills the GPU once
inline int tc_dw_z(int64_t rows, int n1, int n2_tiles) {
  const int64_t ctas = ceil_div(n1, D4_MROWS) * std::max(1, n2_tiles);
  int64_t z = std::max<int64_t>(1, kNumSMs / ctas);
  z = std::min<int64_t>(z, std::max<int64_t>(1, rows / 64));    // small batches: chunks of >= 64 rows keep every SM busy
  z = std::min<int64_t>(z, kNumSMs);
  const int64_t rpz = round_up(ceil_div(rows, z), TC_BK);
  return (int)ceil_div(rows, rpz);
}

// dW kernel with cp.async loaders and an in-place rounding pass: only for Y rows that are 4-byte aligned
// (dense 133 / 147-float feature rows handed in through the C ABI with ld_atoms / ld_bonds == 0)
template <int YMODE>
int tc_dw4(cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1, float* P, int64_t R, int N1,
           int Z) {
  D4Params p{};
  if (!tc_aligned16(X, ldx)) return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: X rows must be 16-byte aligned");
  p.X = X; p.ldx = ldx; p.N1 = N1; p.P = P; p.ldp = w0 + w1; p.R = R;
  p.Y0[0] = Y.X; p.ldy[0] = Y.ldx; p.align16[0] = tc_aligned16(Y.X, Y.ldx) ? 1 : 0;
  if (YMODE == A_CONCAT) { p.Y0[1] = Y.X2; p.ldy[1] = Y.ldx2; p.align16[1] = tc_aligned16(Y.X2, Y.ldx2) ? 1 : 0; }
  p.n2_tiles = tc_dw_tiles(w0, w1, p.t);
  if (p.n2_tiles <= 0) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW: too many column tiles");
  p.n1_groups = (int)ceil_div(N1, D4_MROWS);
  p.rows_per_z = round_up(ceil_div(R, Z), TC_BK);
  if (ceil_div(R, p.rows_per_z) != Z) return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: inconsistent row chunking");
  int maxBN = 16;
  for (int i = 0; i < p.n2_tiles; ++i) maxBN = std::max(maxBN, p.t[i].BN);
  p.XB = (int)ceil_div(std::min(N1, D4_MROWS), 32);
  p.YB = (int)ceil_div(maxBN, 32);
  p.stages = D4_MAX_STAGES;
  while (p.stages > 2 && d4_smem_bytes(p.XB, p.YB, p.stages) > TC_SMEM_LIMIT) --p.stages;
  const int smem = d4_smem_bytes(p.XB, p.YB, p.stages);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW tile does not fit shared memory");
  auto kern = k_tc_dw4;
  int rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  dim3 grid((unsigned)(p.n1_groups * p.n2_tiles), (unsigned)Z);
  WD_LAUNCH(kern, grid, D4_THREADS, smem, s, p);
  return WD_OK;
}

// dW kernel fed by TMA tensor copies (rounding to tf32 in flight) for X and Y; needs 16-byte aligned rows everywhere
template <int YMODE>
int tc_dw5(cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1, float* P, int64_t R, int N1,
           int Z) {
  D5Params p{};
  p.N1 = N1; p.P = P; p.ldp = w0 + w1; p.R = R;
  int rc = tma_make_2d(&p.mx, X, R, N1, ldx, D5_BK, TMA_SW128_ATOM32);
  if (rc != WD_OK) return rc;
  rc = tma_make_2d(&p.my[0], Y.X, R, w0, Y.ldx, D5_BK, TMA_SW128_ATOM32);
  if (rc != WD_OK) return rc;
  if (YMODE == A_CONCAT) {
    rc = tma_make_2d(&p.my[1], Y.X2, R, w1, Y.ldx2, D5_BK, TMA_SW128_ATOM32);
    if (rc != WD_OK) return rc;
  }
  p.n2_tiles = tc_dw_tiles(w0, w1, p.t);
  if (p.n2_tiles <= 0) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW: too many column tiles");
  p.n1_groups = (int)ceil_div(N1, D4_MROWS);
  p.rows_per_z = round_up(ceil_div(R, Z), TC_BK);
  if (ceil_div(R, p.rows_per_z) != Z) return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: inconsistent row chunking");
  int maxBN = 16;
  for (int i = 0; i < p.n2_tiles; ++i) maxBN = std::max(maxBN, p.t[i].BN);
  p.XB = (int)ceil_div(std::min(N1, D4_MROWS), 32);
  p.YB = (int)ceil_div(maxBN, 32);
  p.stages = D5_MAX_STAGES;
  while (p.stages > 2 && d5_smem_bytes(p.XB, p.YB, p.stages) > TC_SMEM_LIMIT) --p.stages;
  const int smem = d5_smem_bytes(p.XB, p.YB, p.stages);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW tile does not fit shared memory");
  auto kern = k_tc_dw5;
  rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  dim3 grid((unsigned)(p.n1_groups * p.n2_tiles), (unsigned)Z);
  WD_LAUNCH(kern, grid, D5_THREADS, smem, s, p);
  return WD_OK;
}

// P[0..Z) <- partial X^T Y over Z row chunks (Z from tc_dw_z); Y plain or two concatenated plain matrices
// (the gathered operand of dW_h is the S_t the forward depth step left in the workspace)
template <int YMODE>
int tc_dw(cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1, float* P, int64_t R, int N1,
          int Z) {
  static_assert(YMODE != A_GATHER_DIFF, "the tensor-core dW contraction reads S_t, not the gathered operand");
  const bool aligned = tc_aligned16(X, ldx) && tc_aligned16(Y.X, Y.ldx) && (YMODE == A_PLAIN || tc_aligned16(Y.X2, Y.ldx2));
  if (aligned) return tc_dw5<YMODE>(s, X, ldx, Y, w0, w1, P, R, N1, Z);
  return tc_dw4<YMODE>(s, X, ldx, Y, w0, w1, P, R, N1, Z);
}

}  // namespace wd
