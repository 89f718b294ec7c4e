// Shared helpers for the wD-MPNN sm_100a kernels: error plumbing, launch accounting,
// activations (chemprop/nn_utils.py:70-99) and small vector utilities.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <utility>

#include "../../include/wdmpnn.h"

namespace wd {

extern thread_local char g_err[512];
extern std::atomic<int64_t> g_launches;
extern std::atomic<int64_t> g_graph_launches, g_graph_builds;

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define WD_CUDA_OK(expr)                                                                   \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess)                                                                 \
      return wd::fail(WD_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                      __FILE__, __LINE__);                                                 \
  } while (0)

// Optional per-category CUDA-event timing of the library's own launches (wd_profile_*).
extern bool g_prof_on;
extern thread_local int g_prof_cat;
void prof_begin(cudaStream_t s);
void prof_end(cudaStream_t s);
struct ProfScope {  // sets the category of the launches issued while it is alive
  int prev;
  explicit ProfScope(int cat) : prev(g_prof_cat) { g_prof_cat = cat; }
  ~ProfScope() { g_prof_cat = prev; }
};

// Programmatic dependent launch (PDL).  Every kernel of this library starts with pdl_prologue() and is launched
// with programmaticStreamSerializationAllowed: the next kernel in the stream is scheduled while its predecessor
// drains (the implicit trigger when the predecessor's CTAs exit) and blocks in `griddepcontrol.wait` until the
// predecessor has completed and flushed its memory -- so no kernel touches memory before its predecessor is done;
// only launch latency and CTA ramp-up overlap with the predecessor's tail.  Measured on B200 (profiles/
// r1_j_pdl_variants.txt): 2.72 -> 2.68 ms/step at batch 4096, 0.540 -> 0.473 at 512, 0.290 -> 0.268 at 50.
// An EARLY trigger (`griddepcontrol.launch_dependents` at the top of every kernel) was measured too and is
// slower at large batches (2.90 ms): the dependents' CTAs become resident long before they can run.
// The wait is a no-op for launches without the attribute (WD_PDL=0, or profiling with events between launches).
__device__ __forceinline__ void pdl_prologue() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

extern int g_pdl;   // -1 = read WD_PDL from the environment on first use (default on)
inline bool pdl_enabled() {
  if (g_pdl < 0) { const char* e = getenv("WD_PDL"); g_pdl = (e && atoi(e) == 0) ? 0 : 1; }
  return g_pdl != 0;
}

// Launches that do not carry the programmatic-serialisation attribute: the attribute describes ONE primary kernel per
// secondary.  A kernel with two predecessors -- everything on the side branch of a captured pass (its own predecessor
// there + the main-stream kernel it was forked after), and the first main-stream kernel after a join -- is therefore
// launched with ordinary (full) dependencies.  No measurable cost at batch 50 (0.191 vs 0.192 ms per step);
// scripts/side_branch_stress.py passes either way (WD_SIDE_PLAIN=0 keeps the attribute, for A/B runs).
struct PdlGuard {
  cudaStream_t plain_stream = nullptr;   // every launch on this stream is a plain one
  bool plain_once = false;               // the next launch (on any stream) is a plain one
};
inline PdlGuard& pdl_guard() { thread_local PdlGuard g; return g; }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                 Args&&... args) {
  PdlGuard& pg = pdl_guard();
  static const bool guard_on = [] { const char* e = getenv("WD_SIDE_PLAIN"); return !(e && atoi(e) == 0); }();
  const bool plain = guard_on && (pg.plain_once || (pg.plain_stream != nullptr && s == pg.plain_stream));
  if (s != pg.plain_stream) pg.plain_once = false;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = (pdl_enabled() && !g_prof_on && !plain) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

// Every kernel launch goes through this so wd_launch_count() is exact.
#define WD_LAUNCH(kernel, grid, block, smem, stream, ...)                                   \
  do {                                                                                      \
    if (wd::g_prof_on) wd::prof_begin(stream);                                              \
    cudaError_t _e = wd::launch_kernel(kernel, (grid), (block), (smem), (stream), __VA_ARGS__); \
    if (wd::g_prof_on) wd::prof_end(stream);                                                \
    wd::g_launches.fetch_add(1, std::memory_order_relaxed);                                 \
    if (_e == cudaSuccess) _e = cudaGetLastError();                                         \
    if (_e != cudaSuccess)                                                                  \
      return wd::fail(WD_ERR_CUDA, "launch of %s failed: %s (%s:%d)", #kernel,              \
                      cudaGetErrorString(_e), __FILE__, __LINE__);                          \
  } while (0)

constexpr int kNumSMs = 148;  // B200

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ inline int64_t round_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }

// ---- activations ---------------------------------------------------------------------------
constexpr float kSeluAlpha = 1.6732632423543772848170429916717f;
constexpr float kSeluScale = 1.0507009873554804934193349852946f;

template <int ACT>
__device__ __forceinline__ float act_fwd(float x) {
  if (ACT == WD_ACT_RELU) return fmaxf(x, 0.f);
  if (ACT == WD_ACT_LEAKY_RELU) return x > 0.f ? x : 0.1f * x;
  if (ACT == WD_ACT_TANH) return tanhf(x);
  if (ACT == WD_ACT_SELU) return kSeluScale * (x > 0.f ? x : kSeluAlpha * expm1f(x));
  return x > 0.f ? x : expm1f(x);  // ELU(alpha=1)
}

// derivative expressed through the activation OUTPUT y (so pre-activations are never stored)
template <int ACT>
__device__ __forceinline__ float act_bwd_from_out(float y) {
  if (ACT == WD_ACT_RELU) return y > 0.f ? 1.f : 0.f;
  if (ACT == WD_ACT_LEAKY_RELU) return y > 0.f ? 1.f : 0.1f;
  if (ACT == WD_ACT_TANH) return 1.f - y * y;
  if (ACT == WD_ACT_SELU) return y > 0.f ? kSeluScale : y + kSeluScale * kSeluAlpha;
  return y > 0.f ? 1.f : y + 1.f;
}

// ---- dropout (nn.Dropout, mpn.py:38; applied at :124 and :134) --------------------------------
// The keep/drop decision of element e (= row * ld + col of the [rows, ld] activation) is word (e & 3) of
// Philox4x32-10 with counter (e >> 2, salt, 0) and key = seed: stateless, so the backward regenerates the
// mask instead of storing it, and the CPU oracle can reproduce it exactly (oracle/mpn_oracle.py).
struct Dropout {
  uint32_t thresh;            // drop iff word < thresh (= p * 2^32); 0 = dropout off (p == 0 / eval mode)
  float scale;                // 1 / (1 - p)
  uint32_t seed_lo, seed_hi;  // key
  uint32_t salt;              // site: depth step t = 1 .. d-1, d = atom hiddens
};

__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                       uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// multipliers (0 or scale) of the four elements e .. e+3 (e % 4 == 0)
__device__ __forceinline__ float4 dropout_keep4(const Dropout& d, int64_t e) {
  uint32_t w[4];
  const uint64_t q = (uint64_t)e >> 2;
  philox4x32_10((uint32_t)q, (uint32_t)(q >> 32), d.salt, 0u, d.seed_lo, d.seed_hi, w);
  return make_float4(w[0] < d.thresh ? 0.f : d.scale, w[1] < d.thresh ? 0.f : d.scale,
                     w[2] < d.thresh ? 0.f : d.scale, w[3] < d.thresh ? 0.f : d.scale);
}
__device__ __forceinline__ float4 dropout4(float4 x, const Dropout& d, int64_t e) {
  const float4 k = dropout_keep4(d, e);
  return make_float4(x.x * k.x, x.y * k.y, x.z * k.z, x.w * k.w);
}
// d act(z)/dz * keep * scale from the STORED (post-dropout) activation y = keep * scale * act(z)
template <int ACT>
__device__ __forceinline__ float4 act_bwd_dropped4(float4 y, const Dropout& d, int64_t e) {
  const float4 k = dropout_keep4(d, e);
  const float inv = 1.f / d.scale;
  return make_float4(act_bwd_from_out<ACT>(y.x * inv) * k.x, act_bwd_from_out<ACT>(y.y * inv) * k.y,
                     act_bwd_from_out<ACT>(y.z * inv) * k.z, act_bwd_from_out<ACT>(y.w * inv) * k.w);
}

// Sign bits of a stored activation, written by the tensor-core epilogues and read by k_message_grad in place of
// the activation itself when the derivative only needs (y > 0) (ReLU): 1 bit instead of 32 per element.
// A "sign tile" is the column range one epilogue call covers (a column tile of the plain kernels, an accumulator
// half of the gathered kernel; at most 160 columns = 5 words): column c belongs to column tile c / bn, half
// (c % bn) / half of it.  Tile j stores its rows back to back: word (j * n_rows + row) * wpt + k covers the
// tile's columns [32k, 32k + 32), and bit 8 * u + q of it is column 32k + 4q + u (q = 0..7, u = 0..3) -- the order
// in which the epilogue's write-out lanes hold the columns (tc_gemm2.cuh: t2_epilogue), so a word is four ballots.
struct SignBits {
  uint32_t* p;                // nullptr: not written / not used
  int64_t n_rows;
  int bn, half, tpn, wpt;     // column tile width, sign tile width, sign tiles per column tile, words per sign tile row
};
__device__ __forceinline__ uint32_t sign_bits4(const SignBits& m, int64_t row, int c) {   // bits of columns c .. c+3 (c % 4 == 0)
  int nt = 0, cc = c;                    // column tile / half by subtraction: one or two steps, cheaper than dividing
  while (cc >= m.bn) { cc -= m.bn; ++nt; }
  const int hb = cc >= m.half ? 1 : 0, c2 = cc - hb * m.half;      // at most two halves per column tile
  const uint32_t w = __ldg(m.p + ((int64_t)(nt * m.tpn + hb) * m.n_rows + row) * m.wpt + (c2 >> 5));
  return w >> ((c2 & 31) >> 2);          // bit 0 / 8 / 16 / 24 = column c / c+1 / c+2 / c+3
}
// Epilogue side: the 8 lanes (pc = lane & 7) of row group rq = lane >> 3 hold 32 columns of one row; lane pc keeps
// the words of 32-column chunk pc for its 8 rows (sw[i]: row rq + 4 i) until the sign tile is complete.
__device__ __forceinline__ void sign_collect(uint32_t& sw, float4 x, int rq, int pc, int chunk) {
  const uint32_t b0 = __ballot_sync(0xffffffffu, x.x > 0.f), b1 = __ballot_sync(0xffffffffu, x.y > 0.f),
                 b2 = __ballot_sync(0xffffffffu, x.z > 0.f), b3 = __ballot_sync(0xffffffffu, x.w > 0.f);
  const int sh = 8 * rq;
  const uint32_t w = ((b0 >> sh) & 0xffu) | (((b1 >> sh) & 0xffu) << 8) | (((b2 >> sh) & 0xffu) << 16) |
                     (((b3 >> sh) & 0xffu) << 24);
  if (pc == chunk) sw = w;
}
__device__ __forceinline__ void sign_flush(const SignBits& m, const uint32_t (&sw)[8], int tile, int64_t row_w0, int64_t M,
                                           int rq, int pc, int n_chunks) {
  if (pc >= n_chunks) return;
  uint32_t* base = m.p + ((int64_t)tile * m.n_rows + row_w0 + rq) * m.wpt + pc;   // row_w0: first row of this warp
#pragma unroll
  for (int i = 0; i < 8; ++i)
    if (row_w0 + rq + 4 * i < M) base[(int64_t)(4 * i) * m.wpt] = sw[i];
}

#define WD_DISPATCH_ACT(act, ...)                                          \
  switch (act) {                                                           \
    case WD_ACT_RELU: { constexpr int ACT = WD_ACT_RELU; __VA_ARGS__; } break;             \
    case WD_ACT_LEAKY_RELU: { constexpr int ACT = WD_ACT_LEAKY_RELU; __VA_ARGS__; } break; \
    case WD_ACT_TANH: { constexpr int ACT = WD_ACT_TANH; __VA_ARGS__; } break;             \
    case WD_ACT_SELU: { constexpr int ACT = WD_ACT_SELU; __VA_ARGS__; } break;             \
    default: { constexpr int ACT = WD_ACT_ELU; __VA_ARGS__; } break;                       \
  }

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 f4_fma(float s, float4 a, float4 acc) {
  acc.x = fmaf(s, a.x, acc.x); acc.y = fmaf(s, a.y, acc.y);
  acc.z = fmaf(s, a.z, acc.z); acc.w = fmaf(s, a.w, acc.w);
  return acc;
}
template <int ACT>
__device__ __forceinline__ float4 f4_act(float4 v) {
  return make_float4(act_fwd<ACT>(v.x), act_fwd<ACT>(v.y), act_fwd<ACT>(v.z), act_fwd<ACT>(v.w));
}

}  // namespace wd
