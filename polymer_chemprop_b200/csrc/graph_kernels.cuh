// HBM-bound graph kernels of the encoder: CSR aggregate / transpose-CSR scatter, the backward
// message-gradient kernel, undirected symmetrisation and the weighted per-molecule readout.
// All use one thread per (row, float4 column): adjacent lanes read adjacent 16-byte words of the
// same row, so every gathered row is fetched as full 128-byte lines.  No atomics anywhere.
#pragma once
#include "common.cuh"

namespace wd {

// out[a, :] = sum_{j in [ptr[a], ptr[a+1])} w[idx[j]] * U(idx[j], :)
//   forward : idx = in_idx,  w = w_bonds   (a_message, mpn.py:112-118 / :127-131)
//   backward: idx = out_idx, w = nullptr   (dA[a] = sum over out-edges of dS, transpose of a_message[b2a])
template <int ACT, bool PREACT>
__global__ void k_aggregate(const float* __restrict__ U, int ld, const float* __restrict__ w,
                            const int32_t* __restrict__ ptr, const int32_t* __restrict__ idx,
                            float* __restrict__ out, int64_t n_rows, int nc4, int rows_per_block) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  // block-local run of rows, element index split by a float multiply (see k_message_grad)
  const int64_t row_base = (int64_t)blockIdx.x * rows_per_block;
  const int n_el = (int)min((int64_t)rows_per_block, n_rows - row_base) * nc4;
  const float inv = 1.f / (float)nc4;
  for (int e = threadIdx.x; e < n_el; e += blockDim.x) {
  const int rl = (int)(((float)e + 0.5f) * inv);
  const int64_t a = row_base + rl;
  const int c = (e - rl * nc4) * 4;
  int j0 = __ldg(ptr + a), j1 = __ldg(ptr + a + 1);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  // four rows in flight per thread (lists hold ~3 entries): indices first, then the independent row loads;
  // the accumulation order is the list order (the reference's order), missing slots add an exact 0
  for (int j = j0; j < j1; j += 4) {
    int b[4];
    float wv[4];
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) b[u] = (j + u < j1) ? __ldg(idx + j + u) : -1;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      v[u] = b[u] >= 0 ? ldg4(U + (int64_t)b[u] * ld + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      wv[u] = b[u] >= 0 ? (w ? __ldg(w + b[u]) : 1.f) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (PREACT) v[u] = f4_act<ACT>(v[u]);
      if (b[u] >= 0) acc = f4_fma(wv[u], v[u], acc);
    }
  }
  st4(out + a * ld + c, acc);
  }
}

// U[b] = (X(b) + X(rev b)) / 2      (mpn.py:101-102)
template <int ACT, bool PREACT>
__global__ void k_symmetrize(const float* __restrict__ X, int ld, const int32_t* __restrict__ rev,
                             float* __restrict__ U, int64_t n_rows, int nc4) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows * nc4) return;
  int64_t b = i / nc4;
  int c = (int)(i - b * nc4) * 4;
  float4 x = ldg4(X + b * ld + c), y = ldg4(X + (int64_t)__ldg(rev + b) * ld + c);
  if (PREACT) { x = f4_act<ACT>(x); y = f4_act<ACT>(y); }
  st4(U + b * ld + c, make_float4((x.x + y.x) / 2.f, (x.y + y.y) / 2.f, (x.z + y.z) / 2.f, (x.w + y.w) / 2.f));
}

// Backward of one message-passing step's inputs:
//   g(b)  = w_b * dA[dst b] - dS[rev b]              (dS absent for the last step: HAS_DS = false)
//   UNDIRECTED && HAS_DS:  g(b) = (g(b) + g(rev b)) / 2
//   dZ[b] = g(b) * act'(Y[b])   with Y = M_t (PREACT: Y = act(I))
//   dI[b] = (dI_in ? dI_in[b] (+ dI_in2[b]) : 0) + dZ[b]     (WRITE_DI; dI_in may alias dI or be an earlier step's dZ;
//                                                            dI_in2: a second earlier dZ, see backward_impl)
// WRITE_DZ = false for the last step (nothing reads dZ_0), WRITE_DI = false for the first one (dI == dZ then).
constexpr int MG_ROWS = 64;   // most rows per block of k_message_grad (fewer when the batch is small: see message_grad())

// SIGNS (ReLU only): act'(Y) comes from the sign bits the forward epilogues left (common.cuh SignBits) instead of
// from Y itself -- one 4-byte word per 32 columns is read in place of 128 bytes, and a dropped element has bit 0.
template <int ACT, bool PREACT, bool HAS_DS, bool UNDIRECTED, bool WRITE_DZ, bool WRITE_DI, bool SIGNS>
__global__ void k_message_grad(const float* __restrict__ dA, const float* __restrict__ dS,
                               const float* __restrict__ Y, int ld, const float* __restrict__ w,
                               const int32_t* __restrict__ dst, const int32_t* __restrict__ rev,
                               float* __restrict__ dZ, const float* dI_in, const float* dI_in2, float* dI, int64_t n_rows, int nc4,
                               Dropout drop, SignBits sb, int rows_per_block) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  // A block walks rows_per_block (<= MG_ROWS) consecutive rows as one flat run of (row, column quad) elements, 256 at a time: adjacent
  // lanes touch adjacent 16-byte words exactly as a flat grid would, but the element -> (row, column) split is a
  // float multiply on a block-local index instead of a 64-bit division.  ncu had this kernel ISSUE-bound (issue
  // slots 78 % busy at 3.1 TB/s for the first step): with the sign bits it moves as little as 20 bytes per element.
  const int64_t row_base = (int64_t)blockIdx.x * rows_per_block;
  const int n_el = (int)min((int64_t)rows_per_block, n_rows - row_base) * nc4;
  const float inv = 1.f / (float)nc4;
  for (int e = threadIdx.x; e < n_el; e += blockDim.x) {
  const int rl = (int)(((float)e + 0.5f) * inv);          // exact: e < MG_ROWS * nc4 < 2^22, (e + 0.5) / nc4 is >= 0.5 / nc4 away from an integer
  const int64_t b = row_base + rl;
  const int c = (e - rl * nc4) * 4;
  float wb = __ldg(w + b);
  float4 da = ldg4(dA + (int64_t)__ldg(dst + b) * ld + c);
  float4 g = make_float4(wb * da.x, wb * da.y, wb * da.z, wb * da.w);
  if (HAS_DS) {
    int r = __ldg(rev + b);
    float4 ds = ldg4(dS + (int64_t)r * ld + c);
    g.x -= ds.x; g.y -= ds.y; g.z -= ds.z; g.w -= ds.w;
    if (UNDIRECTED) {
      float wr = __ldg(w + r);
      float4 dar = ldg4(dA + (int64_t)__ldg(dst + r) * ld + c);
      float4 dsb = ldg4(dS + b * ld + c);
      g.x = (g.x + (wr * dar.x - dsb.x)) / 2.f; g.y = (g.y + (wr * dar.y - dsb.y)) / 2.f;
      g.z = (g.z + (wr * dar.z - dsb.z)) / 2.f; g.w = (g.w + (wr * dar.w - dsb.w)) / 2.f;
    }
  }
  float4 dy;                                   // d (stored activation) / d (pre-activation)
  if (SIGNS) {
    static_assert(!SIGNS || ACT == WD_ACT_RELU, "sign bits carry the derivative of ReLU only");
    const uint32_t bits = sign_bits4(sb, b, c);
    const float on = (!PREACT && drop.thresh != 0u) ? drop.scale : 1.f;     // kept elements were scaled by 1 / (1 - p)
    dy = make_float4((bits & 1u) ? on : 0.f, (bits & 0x100u) ? on : 0.f, (bits & 0x10000u) ? on : 0.f,
                     (bits & 0x1000000u) ? on : 0.f);
  } else {
    float4 y = ldg4(Y + b * ld + c);
    if (PREACT) y = f4_act<ACT>(y);
    if (!PREACT && drop.thresh != 0u) dy = act_bwd_dropped4<ACT>(y, drop, b * ld + c);   // Y = dropout(act(z)), mpn.py:124
    else dy = make_float4(act_bwd_from_out<ACT>(y.x), act_bwd_from_out<ACT>(y.y), act_bwd_from_out<ACT>(y.z),
                          act_bwd_from_out<ACT>(y.w));
  }
  float4 dz = make_float4(g.x * dy.x, g.y * dy.y, g.z * dy.z, g.w * dy.w);
  if (WRITE_DZ) st4(dZ + b * ld + c, dz);
  if (WRITE_DI) {
    if (dI_in != nullptr) {
      float4 di = *reinterpret_cast<const float4*>(dI_in + b * ld + c);
      if (dI_in2 != nullptr) {                 // (dZ_a + dZ_b) first: the same association as the running sum it replaces
        const float4 d2 = *reinterpret_cast<const float4*>(dI_in2 + b * ld + c);
        di = make_float4(di.x + d2.x, di.y + d2.y, di.z + d2.z, di.w + d2.w);
      }
      dz = make_float4(di.x + dz.x, di.y + dz.y, di.z + dz.z, di.w + dz.w);
    }
    st4(dI + b * ld + c, dz);
  }
  }
}

// mscale[g] = dop[g] / den[g], den = sum w_atoms | 1 | aggregation_norm  (mpn.py:158-167);
// 0 for empty molecules (cached_zero_vector branch, :148-149).
__global__ void k_mol_scale(const float* __restrict__ w_atoms, const int32_t* __restrict__ a_start,
                            const int32_t* __restrict__ a_size, const float* __restrict__ dop,
                            int aggregation, float aggregation_norm, float* __restrict__ mscale,
                            int64_t n_mols) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_mols) return;
  int s = a_start[g], n = a_size[g];
  if (n == 0) { mscale[g] = 0.f; return; }
  float den = 1.f;
  if (aggregation == WD_AGG_MEAN) {
    den = 0.f;
    for (int i = 0; i < n; ++i) den += w_atoms[s + i];
  } else if (aggregation == WD_AGG_NORM) {
    den = aggregation_norm;
  }
  mscale[g] = dop[g] / den;
}

// out[g, :] = mscale[g] * sum_{a in scope g} w_atoms[a] * H[a, :]      (mpn.py:146-171)
__global__ void k_readout(const float* __restrict__ H, int ld, const float* __restrict__ w_atoms,
                          const int32_t* __restrict__ a_start, const int32_t* __restrict__ a_size,
                          const float* __restrict__ mscale, float* __restrict__ out, int h,
                          int64_t n_mols) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_mols * h) return;
  int64_t g = i / h;
  int c = (int)(i - g * h);
  int s = a_start[g], n = a_size[g];
  float acc = 0.f;
  for (int k = 0; k < n; ++k) acc = fmaf(__ldg(w_atoms + s + k), __ldg(H + (int64_t)(s + k) * ld + c), acc);
  out[g * h + c] = mscale[g] * acc;
}

// dZo[a, :] = w_atoms[a] * mscale[mol a] * d_out[mol a, :] * act'(H[a, :])   (0 for atoms in no scope)
template <int ACT, bool VEC4>
__global__ void k_readout_grad(const float* __restrict__ d_out, int h, const float* __restrict__ H,
                               int ld, const float* __restrict__ w_atoms,
                               const int32_t* __restrict__ mol_of, const float* __restrict__ mscale,
                               float* __restrict__ dZo, int64_t n_atoms, Dropout drop, int rows_per_block) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  const int nc4 = ld / 4;
  // block-local run of rows, element index split by a float multiply (see k_message_grad)
  const int64_t row_base = (int64_t)blockIdx.x * rows_per_block;
  const int n_el = (int)min((int64_t)rows_per_block, n_atoms - row_base) * nc4;
  const float inv = 1.f / (float)nc4;
  for (int e = threadIdx.x; e < n_el; e += blockDim.x) {
  const int rl = (int)(((float)e + 0.5f) * inv);
  const int64_t a = row_base + rl;
  const int c = (e - rl * nc4) * 4;
  int g = __ldg(mol_of + a);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (g >= 0 && c < h) {
    const float s = __ldg(w_atoms + a) * __ldg(mscale + g);
    const float4 y = *reinterpret_cast<const float4*>(H + a * ld + c);
    const float* dp = d_out + (int64_t)g * h + c;
    float4 d;
    if (VEC4) {               // h % 4 == 0 and d_out 16-byte aligned
      d = ldg4(dp);
    } else {
      d.x = __ldg(dp);
      d.y = c + 1 < h ? __ldg(dp + 1) : 0.f;
      d.z = c + 2 < h ? __ldg(dp + 2) : 0.f;
      d.w = c + 3 < h ? __ldg(dp + 3) : 0.f;
    }
    float4 dy;                                 // H = dropout(act(z)), mpn.py:133-134
    if (drop.thresh != 0u) dy = act_bwd_dropped4<ACT>(y, drop, a * ld + c);
    else dy = make_float4(act_bwd_from_out<ACT>(y.x), act_bwd_from_out<ACT>(y.y), act_bwd_from_out<ACT>(y.z),
                          act_bwd_from_out<ACT>(y.w));
    v.x = s * d.x * dy.x;
    v.y = c + 1 < h ? s * d.y * dy.y : 0.f;
    v.z = c + 2 < h ? s * d.z * dy.z : 0.f;
    v.w = c + 3 < h ? s * d.w * dy.w : 0.f;
  }
  st4(dZo + a * ld + c, v);
  }
}

}  // namespace wd
