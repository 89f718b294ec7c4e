// Shared helpers for the wD-MPNN sm_100a kernels: error plumbing, launch accounting,
// activations (chemprop/nn_utils.py:70-99) and small vector utilities.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/wdmpnn.h"

namespace wd {

extern thread_local char g_err[512];
extern std::atomic<int64_t> g_launches;

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

// Every kernel launch goes through this so wd_launch_count() is exact.
#define WD_LAUNCH(kernel, grid, block, smem, stream, ...)                                   \
  do {                                                                                      \
    if (wd::g_prof_on) wd::prof_begin(stream);                                              \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                             \
    if (wd::g_prof_on) wd::prof_end(stream);                                                \
    wd::g_launches.fetch_add(1, std::memory_order_relaxed);                                 \
    cudaError_t _e = cudaGetLastError();                                                    \
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
