// Fused optimiser step over one flat fp32 parameter buffer.
//
// The reference steps torch.optim.Adam / AdamW over model.parameters() (chemprop/utils.py:295-310) after an
// optional nn.utils.clip_grad_norm_ (chemprop/train/train.py:80-82): with torch's for-each implementation that is
// ~10 small kernels per step plus the norm reduction; here parameters, gradients and both moments are flat buffers
// and the whole update is ONE pass (two more tiny launches when clipping: a fixed-order two-stage sum of squares).
// HBM-bound byte work: 16 bytes read + 12 written per parameter, float4 per thread.
#include <cmath>
#include <mutex>

#include "common.cuh"

namespace wd {
namespace {

constexpr int kNormBlocks = 296;      // 2 per SM; partial j sums elements j, j + 296 * 256, ... in a fixed order
constexpr int kNormThreads = 256;

__global__ void k_sumsq_partial(const float* __restrict__ g, int64_t n, float* __restrict__ partial) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  __shared__ float sm[kNormThreads];
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float x = __ldg(g + i);
    acc = fmaf(x, x, acc);
  }
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int s = kNormThreads / 2; s > 0; s >>= 1) {       // fixed tree: bitwise repeatable
    if ((int)threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sm[0];
}

// norm[0] = sqrt(sum of the partials), in ascending order
__global__ void k_sumsq_final(const float* __restrict__ partial, int n_partial, float* __restrict__ norm,
                              float* __restrict__ norm_out) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    float acc = 0.f;
    for (int i = 0; i < n_partial; ++i) acc += partial[i];
    const float r = sqrtf(acc);
    norm[0] = r;
    if (norm_out) norm_out[0] = r;
  }
}

struct AdamK {     // fp32 roundings of scalars formed in double on the host, as torch forms them
  float beta2, omb1, omb2, eps, wd, decay, step_size, bc2_sqrt, max_norm;   // omb = 1 - beta, decay = 1 - lr * wd
  int decoupled;
};

__device__ __forceinline__ void adam1(float& p, float g, float& m, float& v, const AdamK& k, float scale) {
  g *= scale;
  if (k.decoupled) p *= k.decay;
  else if (k.wd != 0.f) g = fmaf(k.wd, p, g);
  m = fmaf(k.omb1, g - m, m);
  v = fmaf(k.beta2, v, k.omb2 * g * g);
  const float denom = sqrtf(v) / k.bc2_sqrt + k.eps;
  p -= k.step_size * (m / denom);
}

__global__ void k_adam_flat(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, int64_t n, AdamK k, const float* __restrict__ norm) {
  pdl_prologue();   // programmatic dependent launch: see common.cuh
  float scale = 1.f;
  if (k.max_norm > 0.f) scale = fminf(1.f, k.max_norm / (__ldg(norm) + 1e-6f));     // clip_grad_norm_'s clamp
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  if (i + 4 <= n) {
    float4 pp = *reinterpret_cast<float4*>(p + i), mm = *reinterpret_cast<float4*>(m + i),
           vv = *reinterpret_cast<float4*>(v + i);
    const float4 gg = ldg4(g + i);
    adam1(pp.x, gg.x, mm.x, vv.x, k, scale); adam1(pp.y, gg.y, mm.y, vv.y, k, scale);
    adam1(pp.z, gg.z, mm.z, vv.z, k, scale); adam1(pp.w, gg.w, mm.w, vv.w, k, scale);
    st4(p + i, pp); st4(m + i, mm); st4(v + i, vv);
  } else {
    for (int64_t j = i; j < n; ++j) adam1(p[j], g[j], m[j], v[j], k, scale);
  }
}

}  // namespace
}  // namespace wd

extern "C" int64_t wd_adam_scratch_floats(void) { return wd::kNormBlocks + 4; }

extern "C" int wd_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                            const WdAdam* cfg, float* scratch, float* grad_norm_out, void* stream_) {
  using namespace wd;
  cudaStream_t s = (cudaStream_t)stream_;
  if (!params || !grads || !exp_avg || !exp_avg_sq || !cfg || n < 1)
    return fail(WD_ERR_INVALID_ARG, "wd_adam_step: NULL pointer or empty buffer");
  if (((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) | reinterpret_cast<uintptr_t>(exp_avg) |
        reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15) != 0)
    return fail(WD_ERR_INVALID_ARG, "wd_adam_step: the flat buffers must be 16-byte aligned");
  const WdAdam& c = *cfg;
  if (c.step < 1 || !(c.beta1 >= 0.0 && c.beta1 < 1.0) || !(c.beta2 >= 0.0 && c.beta2 < 1.0) || !(c.eps >= 0.0) ||
      !(c.lr >= 0.0) || !(c.weight_decay >= 0.0))
    return fail(WD_ERR_INVALID_ARG, "wd_adam_step: step must be >= 1, betas in [0, 1), lr / eps / weight_decay >= 0");
  const bool clip = c.max_grad_norm > 0.0;
  if (clip && !scratch) return fail(WD_ERR_INVALID_ARG, "wd_adam_step: clipping needs the scratch buffer");
  // bias corrections as torch computes them: in double on the host, then used as fp32 scalars
  const double bc1 = 1.0 - std::pow(c.beta1, (double)c.step);
  const double bc2 = 1.0 - std::pow(c.beta2, (double)c.step);
  AdamK k{(float)c.beta2, (float)(1.0 - c.beta1), (float)(1.0 - c.beta2), (float)c.eps, (float)c.weight_decay,
          (float)(1.0 - c.lr * c.weight_decay), (float)(c.lr / bc1), (float)std::sqrt(bc2),
          clip ? (float)c.max_grad_norm : 0.f, c.decoupled ? 1 : 0};
  if (clip) {
    WD_LAUNCH(k_sumsq_partial, kNormBlocks, kNormThreads, 0, s, grads, n, scratch + 4);
    WD_LAUNCH(k_sumsq_final, 1, 32, 0, s, (const float*)(scratch + 4), (int)kNormBlocks, scratch, grad_norm_out);
  }
  WD_LAUNCH(k_adam_flat, (unsigned)ceil_div(ceil_div(n, 4), 256), 256, 0, s, params, grads, exp_avg, exp_avg_sq, n, k,
            (const float*)scratch);
  return WD_OK;
}

// ---- wd_adam_step_static: the same step for callers that cannot allocate (no counterpart in the reference)
// --------------------------------------------------------------------------------------
// The clipping scratch (the norm slot + kNormBlocks partial sums of squares) lives in this file-scope device array
// instead of a caller buffer: 2 KB that every GPU which loads the library holds until the process exits.  In HOST code
// the array's name is only a host-side stand-in, not a device address, so its device address is looked up once per
// device -- with that device current -- by cudaGetSymbolAddress and cached; the kernels get that pointer.
// Every call on a device shares the array and it keeps the last call's contents.  Nothing in it has to start at zero:
// each call first overwrites all kNormBlocks partials (k_sumsq_partial's grid is fixed), then the norm slot
// (k_sumsq_final), then reads the norm (k_adam_flat), all stream-ordered.  Calls that OVERLAP on the GPU (different
// streams without ordering between them, or several host threads) race on the array; serialise them.
namespace wd {
namespace {

constexpr int kStaticScratchFloats = 512;
static_assert(kNormBlocks + 4 <= kStaticScratchFloats, "static scratch too small for the norm partials");
static __device__ __align__(256) float g_adam_static_scratch[kStaticScratchFloats];

constexpr int kMaxDevices = 64;

int static_scratch(float** out) {
  static std::mutex mu;
  static float* ptr[kMaxDevices] = {};
  int dev = 0;
  WD_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(WD_ERR_UNSUPPORTED, "wd_adam_step_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptr[dev] == nullptr) {     // once per device, with that device current
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_adam_static_scratch);
    if (e != cudaSuccess || p == nullptr)
      return fail(WD_ERR_CUDA, "wd_adam_step_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
    ptr[dev] = reinterpret_cast<float*>(p);
  }
  *out = ptr[dev];
  return WD_OK;
}

}  // namespace
}  // namespace wd

extern "C" int64_t wd_adam_static_max_params(void) { return (int64_t)WD_ADAM_STATIC_MAX_PARAMS; }

extern "C" int wd_adam_step_static(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                                   const WdAdam* cfg, float* grad_norm_out, void* stream) {
  using namespace wd;
  if (n > (int64_t)WD_ADAM_STATIC_MAX_PARAMS)
    return fail(WD_ERR_INVALID_ARG, "wd_adam_step_static: n = %lld exceeds the maximum of %lld parameters of the static-scratch "
                "variant; use wd_adam_step with a caller-provided scratch buffer", (long long)n, (long long)WD_ADAM_STATIC_MAX_PARAMS);
  float* scratch = nullptr;
  const int rc = static_scratch(&scratch);
  if (rc != WD_OK) return rc;
  return wd_adam_step(params, grads, exp_avg, exp_avg_sq, n, cfg, scratch, grad_norm_out, stream);
}
