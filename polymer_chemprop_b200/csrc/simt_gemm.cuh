// fp32 CUDA-core contraction kernels (math_mode FP32: the <=1e-5 parity path).
// Register-blocked smem-tiled FFMA; A operands are virtual (operands.cuh), epilogues fused.
#pragma once
#include "operands.cuh"

namespace wd {

// ---------------------------------------------------------------------------------------------
// C[M, ldc] (first N columns, the rest zero-filled) = epi( A[M,K] * B[N,K]^T )
// tile 128 x 64 x 16, 256 threads, 8 x 4 outputs per thread.
// ---------------------------------------------------------------------------------------------
constexpr int SG_BM = 128, SG_BN = 64, SG_BK = 16, SG_THREADS = 256;

template <int AMODE, int BMODE, int EPI, int ACT, bool PREACT>
__global__ void __launch_bounds__(SG_THREADS)
k_simt_gemm(AOperand A, const float* __restrict__ W, int ldb, const float* __restrict__ bias,
            const float* __restrict__ R, int ldr, float* __restrict__ C, int ldc, int64_t M, int N,
            int K, Dropout drop) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  __shared__ float As[SG_BK][SG_BM + 4];
  __shared__ float Bs[SG_BK][SG_BN + 4];
  const int tid = threadIdx.x;
  const int64_t row0 = (int64_t)blockIdx.x * SG_BM;
  const int col0 = blockIdx.y * SG_BN;
  const int ty = tid >> 4, tx = tid & 15;  // 16 x 16 thread grid: 8 rows x 4 cols each

  // A-load mapping: thread loads rows (tid>>4) + 16*i, column k = tid & 15
  const int lk = tid & 15;
  int sr[8], rr[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int64_t r = row0 + (tid >> 4) + 16 * i;
    sr[i] = rr[i] = 0;
    if (AMODE == A_GATHER_DIFF && r < M) {
      sr[i] = __ldg(A.src + r);
      rr[i] = __ldg(A.rev + r);
    }
  }
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += SG_BK) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      int rl = (tid >> 4) + 16 * i;
      int64_t r = row0 + rl;
      int k = k0 + lk;
      float v = 0.f;
      if (r < M && k < K) v = a_elem<AMODE, ACT, PREACT>(A, r, k, sr[i], rr[i]);
      As[lk][rl] = v;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float v = 0.f;
      if (BMODE == B_NK) {  // k fastest in memory
        int nl = (tid >> 4) + 16 * i, n = col0 + nl, k = k0 + lk;
        if (n < N && k < K) v = __ldg(W + (int64_t)n * ldb + k);
        Bs[lk][nl] = v;
      } else {              // n fastest in memory
        int nl = tid & 63, kl = (tid >> 6) + 4 * i, n = col0 + nl, k = k0 + kl;
        if (n < N && k < K) v = __ldg(W + (int64_t)k * ldb + n);
        Bs[kl][nl] = v;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SG_BK; ++kk) {
      float a[8], b[4];
      float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 8]);
      float4 a1 = *reinterpret_cast<const float4*>(&As[kk][ty * 8 + 4]);
      float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w;
      a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  const int n0 = col0 + tx * 4;
  if (n0 >= ldc) return;
  float bv[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) bv[j] = (bias != nullptr && n0 + j < N) ? __ldg(bias + n0 + j) : 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int64_t r = row0 + ty * 8 + i;
    if (r >= M) continue;
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v = acc[i][j] + bv[j];
      if (EPI == E_RES_ACT) v += (n0 + j < N) ? __ldg(R + r * ldr + n0 + j) : 0.f;
      if (EPI != E_STORE) v = act_fwd<ACT>(v);
      o[j] = (n0 + j < N) ? v : 0.f;   // pad columns [N, ldc) are kept at zero
    }
    float4 o4 = make_float4(o[0], o[1], o[2], o[3]);
    if (EPI != E_STORE && drop.thresh != 0u) o4 = dropout4(o4, drop, r * ldc + n0);
    st4(C + r * ldc + n0, o4);  // ldc % 4 == 0, n0 % 4 == 0
  }
}

// ---------------------------------------------------------------------------------------------
// Weight-gradient contraction over the row (bond / atom) dimension, deterministic split:
//   P[z][n1, n2] = sum_{r in chunk z} X[r, n1] * Y(r, n2)        (Y virtual, operands.cuh)
// followed by k_reduce_partials (fixed z order).  tile 64 x 64, 16 rows per step.
// ---------------------------------------------------------------------------------------------
constexpr int DW_BM = 64, DW_BN = 64, DW_BK = 16, DW_THREADS = 256;

template <int YMODE, int ACT, bool PREACT>
__global__ void __launch_bounds__(DW_THREADS)
k_simt_dw(const float* __restrict__ X, int ldx, AOperand Y, float* __restrict__ P, int64_t R,
          int N1, int N2, int64_t rows_per_chunk) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  __shared__ float Xs[DW_BK][DW_BM + 4];
  __shared__ float Ys[DW_BK][DW_BN + 4];
  const int tid = threadIdx.x;
  const int n1_0 = blockIdx.y * DW_BM, n2_0 = blockIdx.x * DW_BN;
  const int64_t r_begin = (int64_t)blockIdx.z * rows_per_chunk;
  const int64_t r_end = min(R, r_begin + rows_per_chunk);
  const int ty = tid >> 4, tx = tid & 15;  // 4 x 4 outputs per thread
  const int lc = tid & 63, lr = tid >> 6;  // load: column lc, rows lr + 4*i
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int64_t r0 = r_begin; r0 < r_end; r0 += DW_BK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int rl = lr + 4 * i;
      int64_t r = r0 + rl;
      float xv = 0.f, yv = 0.f;
      if (r < r_end) {
        if (n1_0 + lc < N1) xv = __ldg(X + r * ldx + n1_0 + lc);
        if (n2_0 + lc < N2) {
          int s = 0, rv = 0;
          if (YMODE == A_GATHER_DIFF) { s = __ldg(Y.src + r); rv = __ldg(Y.rev + r); }
          yv = a_elem<YMODE, ACT, PREACT>(Y, r, n2_0 + lc, s, rv);
        }
      }
      Xs[rl][lc] = xv;
      Ys[rl][lc] = yv;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < DW_BK; ++kk) {
      float4 a = *reinterpret_cast<const float4*>(&Xs[kk][ty * 4]);
      float4 b = *reinterpret_cast<const float4*>(&Ys[kk][tx * 4]);
      float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* Pz = P + (int64_t)blockIdx.z * N1 * N2;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int n1 = n1_0 + ty * 4 + i;
    if (n1 >= N1) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n2 = n2_0 + tx * 4 + j;
      if (n2 < N2) Pz[(int64_t)n1 * N2 + n2] = acc[i][j];
    }
  }
}

// column sums over a row chunk: P[z][n] = sum_{r in chunk z} X[r, n]   (bias gradients)
static __global__ void k_colsum_partial(const float* __restrict__ X, int ldx, float* __restrict__ P,
                                 int64_t R, int N, int64_t rows_per_chunk) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  int64_t r_begin = (int64_t)blockIdx.y * rows_per_chunk, r_end = min(R, r_begin + rows_per_chunk);
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;    // fixed association order: deterministic
  int64_t r = r_begin;
  for (; r + 4 <= r_end; r += 4) {
    s0 += __ldg(X + (r + 0) * ldx + n); s1 += __ldg(X + (r + 1) * ldx + n);
    s2 += __ldg(X + (r + 2) * ldx + n); s3 += __ldg(X + (r + 3) * ldx + n);
  }
  for (; r < r_end; ++r) s0 += __ldg(X + r * ldx + n);
  P[(int64_t)blockIdx.y * N + n] = (s0 + s1) + (s2 + s3);
}
// the same sums with four columns per thread (16-byte loads, eight rows in flight); needs N % 4 == 0 and 16-byte aligned
// rows.  Per column the additions happen in the same order as above (rows r, r + 4, ... into accumulator r % 4), so
// the bits are the same.
static __global__ void k_colsum_partial4(const float* __restrict__ X, int ldx, float* __restrict__ P,
                                  int64_t R, int N, int64_t rows_per_chunk) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  const int n = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (n >= N) return;
  const int64_t r_begin = (int64_t)blockIdx.y * rows_per_chunk, r_end = min(R, r_begin + rows_per_chunk);
  float4 s[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) s[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  int64_t r = r_begin;
  for (; r + 8 <= r_end; r += 8) {
    float4 v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = ldg4(X + (r + k) * ldx + n);
#pragma unroll
    for (int k = 0; k < 8; ++k) s[k & 3] = make_float4(s[k & 3].x + v[k].x, s[k & 3].y + v[k].y, s[k & 3].z + v[k].z, s[k & 3].w + v[k].w);
  }
  for (; r + 4 <= r_end; r += 4) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4 v = ldg4(X + (r + k) * ldx + n);
      s[k] = make_float4(s[k].x + v.x, s[k].y + v.y, s[k].z + v.z, s[k].w + v.w);
    }
  }
  for (; r < r_end; ++r) {
    const float4 v = ldg4(X + r * ldx + n);
    s[0] = make_float4(s[0].x + v.x, s[0].y + v.y, s[0].z + v.z, s[0].w + v.w);
  }
  float* o = P + (int64_t)blockIdx.y * N + n;
  o[0] = (s[0].x + s[1].x) + (s[2].x + s[3].x);
  o[1] = (s[0].y + s[1].y) + (s[2].y + s[3].y);
  o[2] = (s[0].z + s[1].z) + (s[2].z + s[3].z);
  o[3] = (s[0].w + s[1].w) + (s[2].w + s[3].w);
}

// out[i] = sum_z P[z][i].  The Z partials are split into RP_GROUPS contiguous ranges summed concurrently
// (ascending z inside a range, four independent accumulators), then the range sums are added in ascending order:
// a fixed association order, hence bitwise reproducible; the split keeps 8x more loads in flight than one
// thread walking all Z partials (Z is ~74..148 at batch 4096).
constexpr int RP_COLS = 32, RP_GROUPS = 8, RP_MAX_JOBS = 8;
// One launch reduces several partial buffers (a backward pass has up to six: dW_o, db_o, dW_h, db_h, dW_i, db_i):
// every reduction is tiny, so separate launches were latency-bound -- 5 x 22 us at batch 4096, a fifth of the step
// at batch 50.  Job j owns blocks [block0, block0 + ceil(n / RP_COLS)).
struct ReduceJob {
  const float* P;
  float* out;
  int64_t n;
  int Z;
  unsigned block0;
};
struct ReduceJobs {
  ReduceJob j[RP_MAX_JOBS];
  int count;
};
// V = 4: every thread owns four consecutive outputs and reads the partials as 16-byte words (all jobs need n % 4 == 0
// and 16-byte aligned partials); the order of additions per output is the same as for V = 1, hence the same bits.
template <int V>
static __global__ void __launch_bounds__(RP_COLS * RP_GROUPS) k_reduce_partials(const ReduceJobs jobs) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  __shared__ float part[RP_GROUPS][RP_COLS * V];
  int ji = 0;
#pragma unroll
  for (int k = 1; k < RP_MAX_JOBS; ++k)
    if (k < jobs.count && blockIdx.x >= jobs.j[k].block0) ji = k;
  const float* __restrict__ P = jobs.j[ji].P;
  float* __restrict__ out = jobs.j[ji].out;
  const int64_t n = jobs.j[ji].n;
  const int Z = jobs.j[ji].Z;
  const int tx = threadIdx.x & (RP_COLS - 1), ty = threadIdx.x / RP_COLS;
  const int64_t i = ((int64_t)(blockIdx.x - jobs.j[ji].block0) * RP_COLS + tx) * V;
  const int z0 = (int)((int64_t)Z * ty / RP_GROUPS), z1 = (int)((int64_t)Z * (ty + 1) / RP_GROUPS);
  float s0[V], s1[V], s2[V], s3[V];
#pragma unroll
  for (int v = 0; v < V; ++v) s0[v] = s1[v] = s2[v] = s3[v] = 0.f;
  auto add = [&](float (&acc)[V], int z) {
    if (V == 4) {
      const float4 x = *reinterpret_cast<const float4*>(P + (int64_t)z * n + i);
      acc[0] += x.x; acc[1 % V] += x.y; acc[2 % V] += x.z; acc[3 % V] += x.w;
    } else {
      acc[0] += P[(int64_t)z * n + i];
    }
  };
  if (i < n) {
    int z = z0;
    for (; z + 4 <= z1; z += 4) { add(s0, z); add(s1, z + 1); add(s2, z + 2); add(s3, z + 3); }
    for (; z < z1; ++z) add(s0, z);
  }
#pragma unroll
  for (int v = 0; v < V; ++v) part[ty][tx * V + v] = (s0[v] + s1[v]) + (s2[v] + s3[v]);
  __syncthreads();
  if (ty == 0 && i < n) {
#pragma unroll
    for (int v = 0; v < V; ++v) {
      float t = part[0][tx * V + v];
#pragma unroll
      for (int g = 1; g < RP_GROUPS; ++g) t += part[g][tx * V + v];
      out[i + v] = t;
    }
  }
}

}  // namespace wd
