// Host-side launch helpers of the tcgen05 contraction kernels (tc_gemm.cuh / tc_dw.cuh).
#pragma once
#include <map>
#include <mutex>
#include <utility>
#include "tc_gemm3.cuh"
#include "tc_gemm4.cuh"
#include "tc_dw.cuh"
#include "tc_dw2.cuh"
#include "tc_dw3.cuh"
#include "tc_dw4.cuh"
#include "tc_dw5.cuh"
#include "tc_gemm2.cuh"

namespace wd {

// A packed weight operand: geometry + device image (k_tc_pack_b layout).
struct TcPacked {
  int N, K1, K2, BN, n_tiles, KC1, KC;
  float* img;
  int64_t floats() const { return (int64_t)n_tiles * KC * BN * 32; }
};

inline bool tc_use_v1() {   // WD_TC_V1=1: first-generation (register-staged) kernels, for A/B comparisons
  static bool v = [] { const char* e = getenv("WD_TC_V1"); return e && atoi(e) != 0; }();
  return v;
}

inline int tc_dw_version() {   // WD_TC_DW=1..5: generation of the dW kernel (default 5), for A/B comparisons
  static int v = [] { const char* e = getenv("WD_TC_DW"); int x = e ? atoi(e) : 5; return (x >= 1 && x <= 5) ? x : 5; }();
  return tc_use_v1() ? 1 : v;
}

inline bool tc_use_gemm4() {   // WD_TC_GEMM4=0 keeps the gathered depth step on the second-generation kernel
  static bool v = [] { const char* e = getenv("WD_TC_GEMM4"); return !(e && atoi(e) == 0); }();
  return v && !tc_use_v1();
}

// max_bn: widest column tile the consuming kernel takes (160: two TMEM accumulators; 304: the gathered kernel)
inline TcPacked tc_packed_geometry(int N, int K1, int K2, int max_bn = T2_MAX_BN) {
  TcTiling t = tc_tiling(N);
  if (!tc_use_v1() && t.BN > max_bn) {
    const int np = (int)round_up(N, 16);
    t.n_tiles = (int)ceil_div(np, std::min(tc_max_bn(), max_bn));
    t.BN = (int)round_up(ceil_div(np, t.n_tiles), 16);
  }
  TcPacked g{N, K1, K2, t.BN, t.n_tiles, (int)ceil_div(K1, TC_BK), 0, nullptr};
  g.KC = g.KC1 + (int)ceil_div(K2, TC_BK);
  return g;
}

// Collects up to three (weight matrix -> packed image) jobs and packs them with one launch.
struct TcPacker {
  TcPackJobs jobs{};
  void add(const float* W, int ldb, int bmode, const TcPacked& g) {
    TcPackJob& j = jobs.job[jobs.n++];
    j = TcPackJob{W, ldb, bmode, g.N, g.K1, g.K2, g.BN, g.n_tiles, g.KC1, g.KC, g.img, jobs.total};
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

// second-generation kernel (cp.async raw ring + transform warps)
template <int AMODE, int EPI, int ACT, bool PREACT>
int tc_gemm2(cudaStream_t s, const AOperand& A, const TcPacked& B, const float* bias, const float* R, int ldr,
             float* C, int ldc, int64_t M, const Dropout& drop, const SignBits& sign) {
  T2Params p{};
  p.drop = drop;
  p.sign = sign;
  constexpr int NSRC = AMODE == A_GATHER_DIFF ? 2 : 1;
  if (AMODE == A_GATHER_DIFF) {
    p.seg[0] = T2Seg{A.X, A.X2, A.src, A.rev, A.ldx, A.ldx2, B.K1, 1};
    p.S = A.S_out; p.lds = A.ldx2;
    if (!tc_aligned16(A.X, A.ldx) || !tc_aligned16(A.X2, A.ldx2))
      return fail(WD_ERR_INVALID_ARG, "tcgen05 gather operand needs 16-byte aligned rows");
  } else {
    p.seg[0] = T2Seg{A.X, nullptr, nullptr, nullptr, A.ldx, 0, B.K1, tc_aligned16(A.X, A.ldx) ? 1 : 0};
    if (AMODE == A_CONCAT) p.seg[1] = T2Seg{A.X2, nullptr, nullptr, nullptr, A.ldx2, 0, B.K2, tc_aligned16(A.X2, A.ldx2) ? 1 : 0};
  }
  p.KC0 = B.KC1; p.KC = B.KC;
  p.Bp = B.img; p.bias = bias; p.R = R; p.ldr = ldr; p.C = C; p.ldc = ldc; p.M = M;
  p.N = B.N; p.BN = B.BN; p.n_tiles = B.n_tiles;
  const int smem = t2_smem_bytes(B.BN);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 tile does not fit shared memory");
  if (EPI == E_RES_ACT && (!tc_aligned16(R, ldr))) return fail(WD_ERR_INVALID_ARG, "tcgen05 residual rows must be 16-byte aligned");
  if (!tc_aligned16(C, ldc)) return fail(WD_ERR_INVALID_ARG, "tcgen05 output rows must be 16-byte aligned");
  auto kern = k_tc_gemm2<NSRC, EPI, ACT, PREACT>;
  int rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  const int64_t tiles = ceil_div(M, TC_BM) * B.n_tiles;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, kNumSMs);
  WD_LAUNCH(kern, grid, T2_THREADS, smem, s, p);
  return WD_OK;
}

// fourth-generation kernel for the gathered depth step (column tiles up to 304 wide, in-place transform, TMA residual)
template <int ACT, bool PREACT>
int tc_gemm4(cudaStream_t s, const AOperand& A, const TcPacked& B, const float* bias, const float* R, int ldr,
             float* C, int ldc, int64_t M, const Dropout& drop, const SignBits& sign) {
  T4Params p{};
  p.drop = drop;
  p.sign = sign;
  if (!tc_aligned16(A.X, A.ldx) || !tc_aligned16(A.X2, A.ldx2) || !tc_aligned16(R, ldr) || !tc_aligned16(C, ldc))
    return fail(WD_ERR_INVALID_ARG, "tcgen05 gathered step needs 16-byte aligned rows");
  if (B.BN > T4_MAX_BN || B.K2 != 0) return fail(WD_ERR_UNSUPPORTED, "tcgen05 gathered step: unsupported tile geometry");
  int rc = tma_make_2d(&p.mr, R, M, B.N, ldr, TC_BM, TMA_SW128, /*round_tf32=*/false);
  if (rc != WD_OK) return rc;
  p.X0 = A.X; p.X1 = A.X2; p.i0 = A.src; p.i1 = A.rev; p.ld0 = A.ldx; p.ld1 = A.ldx2; p.K = B.K1; p.KC = B.KC;
  p.Bp = B.img; p.bias = bias; p.C = C; p.ldc = ldc; p.M = M;
  p.N = B.N; p.BN = B.BN; p.n_tiles = B.n_tiles;
  p.S = A.S_out; p.lds = A.ldx2;
  static const bool x1_tma_off = [] { const char* e = getenv("WD_T4_X1_TMA"); return e && atoi(e) == 0; }();   // A/B runs
  p.x1_tma = (A.rev_adjacent && A.rev != nullptr && !x1_tma_off) ? 1 : 0;
  if (p.x1_tma) {
    rc = tma_make_2d(&p.mx1, A.X2, M, B.K1, A.ldx2, T4_X1_ROWS, TMA_SW128, /*round_tf32=*/false);
    if (rc != WD_OK) return rc;
  }
  const int smem = t4_smem_bytes(B.BN);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 tile does not fit shared memory");
  auto kern = k_tc_gemm4<ACT, PREACT>;
  rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  const int64_t tiles = ceil_div(M, TC_BM) * B.n_tiles;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, kNumSMs);
  WD_LAUNCH(kern, grid, T4_THREADS, smem, s, p);
  return WD_OK;
}

inline bool tc_use_gemm3() {   // WD_TC_GEMM=2 keeps plain operands on the second-generation (cp.async) kernel
  static bool v = [] { const char* e = getenv("WD_TC_GEMM"); return !(e && atoi(e) == 2); }();
  return v;
}

// third-generation kernel for plain / concatenated A operands: TMA tensor copies, no transform pass
template <int AMODE, int EPI, int ACT>
int tc_gemm3(cudaStream_t s, const AOperand& A, const TcPacked& B, const float* bias, float* C, int ldc, int64_t M,
             const Dropout& drop, const SignBits& sign) {
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
  auto kern = k_tc_gemm3<EPI, ACT>;
  rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  const int64_t tiles = ceil_div(M, TC_BM) * B.n_tiles;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, kNumSMs);
  WD_LAUNCH(kern, grid, T3_THREADS, smem, s, p);
  return WD_OK;
}

template <int AMODE, int EPI, int ACT, bool PREACT>
struct TcGemm3Dispatch {   // the TMA kernel exists for plain operands and the store / activation epilogues only
  static int run(cudaStream_t s, const AOperand& A, const TcPacked& B, const float* bias, float* C, int ldc, int64_t M,
                 const Dropout& drop, const SignBits& sign) {
    return tc_gemm3<AMODE, EPI, ACT>(s, A, B, bias, C, ldc, M, drop, sign);
  }
  static constexpr bool ok = true;
};
template <int EPI, int ACT, bool PREACT>
struct TcGemm3Dispatch<A_GATHER_DIFF, EPI, ACT, PREACT> {
  static int run(cudaStream_t, const AOperand&, const TcPacked&, const float*, float*, int, int64_t, const Dropout&, const SignBits&) { return WD_ERR_UNSUPPORTED; }
  static constexpr bool ok = false;
};
template <int AMODE, int ACT, bool PREACT>
struct TcGemm3Dispatch<AMODE, E_RES_ACT, ACT, PREACT> {
  static int run(cudaStream_t, const AOperand&, const TcPacked&, const float*, float*, int, int64_t, const Dropout&, const SignBits&) { return WD_ERR_UNSUPPORTED; }
  static constexpr bool ok = false;
};
template <int ACT, bool PREACT>
struct TcGemm3Dispatch<A_GATHER_DIFF, E_RES_ACT, ACT, PREACT> {
  static int run(cudaStream_t, const AOperand&, const TcPacked&, const float*, float*, int, int64_t, const Dropout&, const SignBits&) { return WD_ERR_UNSUPPORTED; }
  static constexpr bool ok = false;
};

// C[M, ldc] = epi( A[M, K1+K2] * B^T ) with B packed
template <int AMODE, int EPI, int ACT, bool PREACT>
int tc_gemm(cudaStream_t s, const AOperand& A, const TcPacked& B, const float* bias, const float* R, int ldr,
            float* C, int ldc, int64_t M, const Dropout& drop = Dropout{}, const SignBits& sign = SignBits{}) {
  if (!tc_use_v1()) {
    if (AMODE == A_GATHER_DIFF && EPI == E_RES_ACT && B.BN > T2_MAX_BN) return tc_gemm4<ACT, PREACT>(s, A, B, bias, R, ldr, C, ldc, M, drop, sign);
    using D3 = TcGemm3Dispatch<AMODE, EPI, ACT, PREACT>;
    if (D3::ok && tc_use_gemm3() && B.BN <= T2_MAX_BN && tc_aligned16(A.X, A.ldx) && tc_aligned16(C, ldc) &&
        (AMODE != A_CONCAT || tc_aligned16(A.X2, A.ldx2)))
      return D3::run(s, A, B, bias, C, ldc, M, drop, sign);
    return tc_gemm2<AMODE, EPI, ACT, PREACT>(s, A, B, bias, R, ldr, C, ldc, M, drop, sign);
  }
  if (drop.thresh != 0u) return fail(WD_ERR_UNSUPPORTED, "WD_TC_V1 kernels do not implement dropout");
  if (sign.p != nullptr) return fail(WD_ERR_UNSUPPORTED, "WD_TC_V1 kernels do not write sign bits");
  TcGemmParams p{};
  p.A = A;
  p.A.K1 = B.K1;
  p.K2 = B.K2;
  p.vec0 = tc_aligned16(A.X, A.ldx);
  p.vec1 = A.X2 ? tc_aligned16(A.X2, A.ldx2) : 0;
  if (AMODE == A_GATHER_DIFF && !(p.vec0 && p.vec1))
    return fail(WD_ERR_INVALID_ARG, "tcgen05 gather operand needs 16-byte aligned rows");
  p.Bp = B.img; p.bias = bias; p.R = R; p.ldr = ldr; p.C = C; p.ldc = ldc; p.M = M;
  p.N = B.N; p.BN = B.BN; p.n_tiles = B.n_tiles; p.KC1 = B.KC1; p.KC = B.KC;
  p.stages = tc_num_stages(B.BN);
  if (p.stages < 2) return fail(WD_ERR_UNSUPPORTED, "tcgen05 tile does not fit shared memory");
  const int smem = tc_smem_bytes(B.BN, p.stages);
  auto kern = k_tc_gemm<AMODE, EPI, ACT, PREACT>;
  int rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  const int64_t tiles = ceil_div(M, TC_BM) * B.n_tiles;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, kNumSMs);
  WD_LAUNCH(kern, grid, TC_THREADS, smem, s, p);
  return WD_OK;
}

// n2 tiles of a dW contraction: each segment split into tiles of <= TC_MAX_BN columns
inline int tc_dw_tiles(int w0, int w1, TcDwTile* out) {
  const int max_bn = tc_dw_version() == 1 ? tc_max_bn() : std::min(tc_max_bn(), D2_MAX_BN);
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
  const int64_t ctas = ceil_div(n1, tc_dw_version() >= 3 ? D3_MROWS : TC_BM) * std::max(1, n2_tiles);
  int64_t z = std::max<int64_t>(1, kNumSMs / ctas);
  z = std::min<int64_t>(z, std::max<int64_t>(1, rows / 128));
  z = std::min<int64_t>(z, kNumSMs);
  const int64_t rpz = round_up(ceil_div(rows, z), TC_BK);
  return (int)ceil_div(rows, rpz);
}

// second-generation dW kernel (raw row ring + transposing transform warps)
template <int YMODE, int ACT, bool PREACT>
int tc_dw2(cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1, float* P, int64_t R, int N1,
           int Z) {
  constexpr int NSRC = YMODE == A_GATHER_DIFF ? 2 : 1;
  D2Params p{};
  if (!tc_aligned16(X, ldx)) return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: X rows must be 16-byte aligned");
  p.X = X; p.ldx = ldx; p.N1 = N1; p.P = P; p.ldp = w0 + w1; p.R = R;
  p.Y0[0] = Y.X; p.ldy[0] = Y.ldx; p.align16[0] = tc_aligned16(Y.X, Y.ldx) ? 1 : 0;
  if (YMODE == A_GATHER_DIFF) {
    p.Y1 = Y.X2; p.ldy1 = Y.ldx2; p.i0 = Y.src; p.i1 = Y.rev;
    if (!p.align16[0] || !tc_aligned16(Y.X2, Y.ldx2))
      return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: gather operand needs 16-byte aligned rows");
  } else if (YMODE == A_CONCAT) {
    p.Y0[1] = Y.X2; p.ldy[1] = Y.ldx2; p.align16[1] = tc_aligned16(Y.X2, Y.ldx2) ? 1 : 0;
  }
  p.n2_tiles = tc_dw_tiles(w0, w1, p.t);
  if (p.n2_tiles <= 0) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW: too many column tiles");
  p.n1_tiles = (int)ceil_div(N1, TC_BM);
  p.rows_per_z = round_up(ceil_div(R, Z), TC_BK);
  if (ceil_div(R, p.rows_per_z) != Z) return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: inconsistent row chunking");
  p.maxBN = 16;
  for (int i = 0; i < p.n2_tiles; ++i) p.maxBN = std::max(p.maxBN, p.t[i].BN);
  p.raw_stages = D2_MAX_RAW_STAGES;
  while (p.raw_stages > 1 && d2_smem_bytes(p.maxBN, NSRC, p.raw_stages) > TC_SMEM_LIMIT) --p.raw_stages;
  const int smem = d2_smem_bytes(p.maxBN, NSRC, p.raw_stages);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW tile does not fit shared memory");
  auto kern = k_tc_dw2<NSRC, ACT, PREACT>;
  int rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  dim3 grid((unsigned)(p.n1_tiles * p.n2_tiles), (unsigned)Z);
  WD_LAUNCH(kern, grid, D2_THREADS, smem, s, p);
  return WD_OK;
}

// third-generation dW kernel (up to 384 X columns per CTA, 16-row chunks, SWIZZLE_64B operand stages)
template <int YMODE, int ACT, bool PREACT>
int tc_dw3(cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1, float* P, int64_t R, int N1,
           int Z) {
  constexpr int NSRC = YMODE == A_GATHER_DIFF ? 2 : 1;
  D3Params p{};
  if (!tc_aligned16(X, ldx)) return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: X rows must be 16-byte aligned");
  p.X = X; p.ldx = ldx; p.N1 = N1; p.P = P; p.ldp = w0 + w1; p.R = R;
  p.Y0[0] = Y.X; p.ldy[0] = Y.ldx; p.align16[0] = tc_aligned16(Y.X, Y.ldx) ? 1 : 0;
  if (YMODE == A_GATHER_DIFF) {
    p.Y1 = Y.X2; p.ldy1 = Y.ldx2; p.i0 = Y.src; p.i1 = Y.rev;
    if (!p.align16[0] || !tc_aligned16(Y.X2, Y.ldx2))
      return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: gather operand needs 16-byte aligned rows");
  } else if (YMODE == A_CONCAT) {
    p.Y0[1] = Y.X2; p.ldy[1] = Y.ldx2; p.align16[1] = tc_aligned16(Y.X2, Y.ldx2) ? 1 : 0;
  }
  p.n2_tiles = tc_dw_tiles(w0, w1, p.t);
  if (p.n2_tiles <= 0) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW: too many column tiles");
  p.n1_groups = (int)ceil_div(N1, D3_MROWS);
  p.xw = (int)round_up(std::min(N1, D3_MROWS), 32);
  p.rows_per_z = round_up(ceil_div(R, Z), TC_BK);
  if (ceil_div(R, p.rows_per_z) != Z) return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: inconsistent row chunking");
  p.maxBN = 16;
  for (int i = 0; i < p.n2_tiles; ++i) p.maxBN = std::max(p.maxBN, p.t[i].BN);
  p.raw_stages = D3_MAX_RAW_STAGES;
  while (p.raw_stages > 1 && d3_smem_bytes(p.xw, p.maxBN, NSRC, p.raw_stages) > TC_SMEM_LIMIT) --p.raw_stages;
  const int smem = d3_smem_bytes(p.xw, p.maxBN, NSRC, p.raw_stages);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW tile does not fit shared memory");
  auto kern = k_tc_dw3<NSRC, ACT, PREACT>;
  int rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  dim3 grid((unsigned)(p.n1_groups * p.n2_tiles), (unsigned)Z);
  WD_LAUNCH(kern, grid, D3_THREADS, smem, s, p);
  return WD_OK;
}

// fourth-generation dW kernel (MN-major operand stages written directly by the loaders, in-place TF32 rounding)
template <int YMODE, int ACT, bool PREACT>
int tc_dw4(cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1, float* P, int64_t R, int N1,
           int Z) {
  constexpr int NSRC = YMODE == A_GATHER_DIFF ? 2 : 1;
  D4Params p{};
  if (!tc_aligned16(X, ldx)) return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: X rows must be 16-byte aligned");
  p.X = X; p.ldx = ldx; p.N1 = N1; p.P = P; p.ldp = w0 + w1; p.R = R;
  p.Y0[0] = Y.X; p.ldy[0] = Y.ldx; p.align16[0] = tc_aligned16(Y.X, Y.ldx) ? 1 : 0;
  if (YMODE == A_GATHER_DIFF) {
    p.Y1 = Y.X2; p.ldy1 = Y.ldx2; p.i0 = Y.src; p.i1 = Y.rev;
    if (!p.align16[0] || !tc_aligned16(Y.X2, Y.ldx2))
      return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: gather operand needs 16-byte aligned rows");
  } else if (YMODE == A_CONCAT) {
    p.Y0[1] = Y.X2; p.ldy[1] = Y.ldx2; p.align16[1] = tc_aligned16(Y.X2, Y.ldx2) ? 1 : 0;
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
  p.stages = D4_MAX_STAGES;
  while (p.stages > 2 && d4_smem_bytes(p.XB, p.YB, NSRC, p.stages) > TC_SMEM_LIMIT) --p.stages;
  const int smem = d4_smem_bytes(p.XB, p.YB, NSRC, p.stages);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW tile does not fit shared memory");
  auto kern = k_tc_dw4<NSRC, ACT, PREACT>;
  int rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  dim3 grid((unsigned)(p.n1_groups * p.n2_tiles), (unsigned)Z);
  WD_LAUNCH(kern, grid, D4_THREADS, smem, s, p);
  return WD_OK;
}

// fifth-generation dW kernel: TMA tensor copies (rounding to tf32 in flight) for X and a plain Y; needs 16-byte
// aligned rows everywhere, otherwise the caller falls back to generation 4
template <int YMODE, int ACT, bool PREACT>
int tc_dw5(cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1, float* P, int64_t R, int N1,
           int Z) {
  constexpr int NSRC = YMODE == A_GATHER_DIFF ? 2 : 1;
  constexpr int BK = d5_bk(NSRC);
  D5Params p{};
  p.N1 = N1; p.P = P; p.ldp = w0 + w1; p.R = R;
  int rc = tma_make_2d(&p.mx, X, R, N1, ldx, BK, TMA_SW128_ATOM32);
  if (rc != WD_OK) return rc;
  if (YMODE == A_GATHER_DIFF) {
    p.Y0 = Y.X; p.ldy0 = Y.ldx; p.Y1 = Y.X2; p.ldy1 = Y.ldx2; p.i0 = Y.src; p.i1 = Y.rev;
  } else {
    rc = tma_make_2d(&p.my[0], Y.X, R, w0, Y.ldx, BK, TMA_SW128_ATOM32);
    if (rc != WD_OK) return rc;
    if (YMODE == A_CONCAT) {
      rc = tma_make_2d(&p.my[1], Y.X2, R, w1, Y.ldx2, BK, TMA_SW128_ATOM32);
      if (rc != WD_OK) return rc;
    }
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
  while (p.stages > 2 && d5_smem_bytes(p.XB, p.YB, NSRC, p.stages) > TC_SMEM_LIMIT) --p.stages;
  const int smem = d5_smem_bytes(p.XB, p.YB, NSRC, p.stages);
  if (smem > TC_SMEM_LIMIT) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW tile does not fit shared memory");
  auto kern = k_tc_dw5<NSRC, ACT, PREACT>;
  rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  dim3 grid((unsigned)(p.n1_groups * p.n2_tiles), (unsigned)Z);
  WD_LAUNCH(kern, grid, d5_threads(NSRC), smem, s, p);
  return WD_OK;
}

// P[0..Z) <- partial X^T Y over Z row chunks (Z from tc_dw_z)
template <int YMODE, int ACT, bool PREACT>
int tc_dw(cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1, float* P, int64_t R, int N1,
          int Z) {
  if (tc_dw_version() == 5) {
    const bool aligned = tc_aligned16(X, ldx) && tc_aligned16(Y.X, Y.ldx) && (YMODE == A_PLAIN || tc_aligned16(Y.X2, Y.ldx2));
    if (aligned) return tc_dw5<YMODE, ACT, PREACT>(s, X, ldx, Y, w0, w1, P, R, N1, Z);
  }
  if (tc_dw_version() >= 4) return tc_dw4<YMODE, ACT, PREACT>(s, X, ldx, Y, w0, w1, P, R, N1, Z);
  if (tc_dw_version() == 3) return tc_dw3<YMODE, ACT, PREACT>(s, X, ldx, Y, w0, w1, P, R, N1, Z);
  if (tc_dw_version() == 2) return tc_dw2<YMODE, ACT, PREACT>(s, X, ldx, Y, w0, w1, P, R, N1, Z);
  TcDwParams p{};
  p.X = X; p.ldx = ldx; p.N1 = N1; p.Y = Y; p.P = P; p.ldp = w0 + w1; p.R = R;
  p.n2_tiles = tc_dw_tiles(w0, w1, p.t);
  if (p.n2_tiles <= 0) return fail(WD_ERR_UNSUPPORTED, "tcgen05 dW: too many column tiles");
  p.n1_tiles = (int)ceil_div(N1, TC_BM);
  p.rows_per_z = round_up(ceil_div(R, Z), TC_BK);
  if (ceil_div(R, p.rows_per_z) != Z) return fail(WD_ERR_INVALID_ARG, "tcgen05 dW: inconsistent row chunking");
  p.maxBN = 16;
  for (int i = 0; i < p.n2_tiles; ++i) p.maxBN = std::max(p.maxBN, p.t[i].BN);
  p.stages = tc_num_stages(p.maxBN);
  const int smem = tc_smem_bytes(p.maxBN, p.stages);
  auto kern = k_tc_dw<YMODE, ACT, PREACT>;
  int rc = tc_set_smem(kern, smem);
  if (rc != WD_OK) return rc;
  dim3 grid((unsigned)(p.n1_tiles * p.n2_tiles), (unsigned)Z);
  WD_LAUNCH(kern, grid, TC_THREADS, smem, s, p);
  return WD_OK;
}

}  // namespace wd
