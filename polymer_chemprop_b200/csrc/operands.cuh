// Operand descriptors shared by the SIMT (fp32) and tcgen05 (tf32) contraction kernels.
//
// Every contraction of the encoder is  C[r, n] = epi( sum_k A(r, k) * B(n, k) )  with a
// "virtual" A operand that is never materialised in HBM:
//   A_PLAIN        A(r,k) = X[r*ldx + k]
//   A_CONCAT       A(r,k) = k < K1 ? X[r*ldx + k] : X2[r*ldx2 + k-K1]   ([f_atoms, a_message], mpn.py:132)
//   A_GATHER_DIFF  A(r,k) = Agg[src[r]*ldx + k] - U(rev[r], k)           (mpn.py:119-120)
//                  with U(x,k) = PREACT ? act(X2[x*ldx2+k]) : X2[x*ldx2+k]  (message = act(input), :97)
#pragma once
#include "common.cuh"

namespace wd {

enum { A_PLAIN = 0, A_CONCAT = 1, A_GATHER_DIFF = 2 };
enum { B_NK = 0, B_KN = 1 };                       // B(n,k) = W[n*ldb+k]  |  W[k*ldb+n]
enum { E_STORE = 0, E_ACT = 1, E_RES_ACT = 2 };    // acc(+bias) | act(acc+bias) | act(R + acc + bias)

struct AOperand {
  const float* X;      // PLAIN/CONCAT: first source; GATHER_DIFF: Agg [N_a, ldx]
  const float* X2;     // CONCAT: second source; GATHER_DIFF: previous messages [N_b, ldx2]
  const int32_t* src;  // GATHER_DIFF: b2a
  const int32_t* rev;  // GATHER_DIFF: b2revb
  int ldx, ldx2, K1;
  int rev_adjacent;    // GATHER_DIFF: rev[r] == ((r-1)^1)+1 for every r >= 1 (WdGraph.rev_adjacent): X2 rows can come by TMA
  float* S_out;        // GATHER_DIFF on the tensor-core path: the operand rows X[src] - act?(X2[rev]) are also
                       // written here ([M, ldx2], fp32) so that the weight-gradient pass can read them as a plain matrix
};

template <int AMODE, int ACT, bool PREACT>
__device__ __forceinline__ float a_elem(const AOperand& A, int64_t r, int k, int sr, int rr) {
  if (AMODE == A_PLAIN) return __ldg(A.X + r * A.ldx + k);
  if (AMODE == A_CONCAT)
    return k < A.K1 ? __ldg(A.X + r * A.ldx + k) : __ldg(A.X2 + r * A.ldx2 + (k - A.K1));
  float u = __ldg(A.X2 + (int64_t)rr * A.ldx2 + k);
  if (PREACT) u = act_fwd<ACT>(u);
  return __ldg(A.X + (int64_t)sr * A.ldx + k) - u;
}

}  // namespace wd
