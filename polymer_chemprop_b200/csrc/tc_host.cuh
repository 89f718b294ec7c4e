// Host-side launch helpers of the tcgen05 contraction kernels (tc_gemm.cuh / tc_dw.cuh).
#pragma once
#include "tc_dw.cuh"

namespace wd {

// A packed weight operand: geometry + device image (k_tc_pack_b layout).
struct TcPacked {
  int N, K1, K2, BN, n_tiles, KC1, KC;
  float* img;
  int64_t floats() const { return (int64_t)n_tiles * KC * BN * 32; }
};

inline TcPacked tc_packed_geometry(int N, int K1, int K2) {
  TcTiling t = tc_tiling(N);
  TcPacked g{N, K1, K2, t.BN, t.n_tiles, (int)ceil_div(K1, TC_BK), 0, nullptr};
  g.KC = g.KC1 + (int)ceil_div(K2, TC_BK);
  return g;
}

inline int tc_pack(cudaStream_t s, const float* W, int ldb, int bmode, const TcPacked& g) {
  const int64_t total = g.floats();
  WD_LAUNCH(k_tc_pack_b, (unsigned)ceil_div(total, 256), 256, 0, s, W, ldb, bmode, g.N, g.K1, g.K2, g.BN, g.n_tiles,
            g.KC1, g.KC, g.img);
  return WD_OK;
}

inline bool tc_aligned16(const void* p, int ld) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && (ld & 3) == 0; }

template <typename Kernel>
inline int tc_set_smem(Kernel k, int bytes) {
  WD_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return WD_OK;
}

// C[M, ldc] = epi( A[M, K1+K2] * B^T ) with B packed
template <int AMODE, int EPI, int ACT, bool PREACT>
int tc_gemm(cudaStream_t s, const AOperand& A, const TcPacked& B, const float* bias, const float* R, int ldr,
            float* C, int ldc, int64_t M) {
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
  int n = 0;
  const int widths[2] = {w0, w1};
  int out0 = 0;
  for (int seg = 0; seg < 2; ++seg) {
    const int w = widths[seg];
    if (w <= 0) continue;
    TcTiling t = tc_tiling(w);
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
  const int64_t ctas = ceil_div(n1, TC_BM) * std::max(1, n2_tiles);
  int64_t z = std::max<int64_t>(1, kNumSMs / ctas);
  z = std::min<int64_t>(z, std::max<int64_t>(1, rows / 128));
  z = std::min<int64_t>(z, 64);
  const int64_t rpz = round_up(ceil_div(rows, z), TC_BK);
  return (int)ceil_div(rows, rpz);
}

// P[0..Z) <- partial X^T Y over Z row chunks (Z from tc_dw_z)
template <int YMODE, int ACT, bool PREACT>
int tc_dw(cudaStream_t s, const float* X, int ldx, const AOperand& Y, int w0, int w1, float* P, int64_t R, int N1,
          int Z) {
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
