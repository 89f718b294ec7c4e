// wd_profile_*: CUDA-event timing of the library's own kernel launches, by category.
#include <mutex>
#include <vector>

#include "common.cuh"

namespace wd {

bool g_prof_on = false;
int g_pdl = -1;
thread_local int g_prof_cat = WD_PROF_MISC;

namespace {
struct Rec { int cat; cudaEvent_t a, b; };
std::mutex g_mu;
std::vector<Rec> g_recs;
std::vector<cudaEvent_t> g_pool;
thread_local cudaEvent_t t_start = nullptr;

cudaEvent_t take_event() {
  if (!g_pool.empty()) { cudaEvent_t e = g_pool.back(); g_pool.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
}  // namespace

void prof_begin(cudaStream_t s) {
  std::lock_guard<std::mutex> lk(g_mu);
  t_start = take_event();
  cudaEventRecord(t_start, s);
}

void prof_end(cudaStream_t s) {
  std::lock_guard<std::mutex> lk(g_mu);
  cudaEvent_t e = take_event();
  cudaEventRecord(e, s);
  g_recs.push_back(Rec{g_prof_cat, t_start, e});
  t_start = nullptr;
}

}  // namespace wd

extern "C" void wd_profile_enable(int on) { wd::g_prof_on = on != 0; }

extern "C" void wd_profile_reset(void) {
  std::lock_guard<std::mutex> lk(wd::g_mu);
  for (auto& r : wd::g_recs) { wd::g_pool.push_back(r.a); wd::g_pool.push_back(r.b); }
  wd::g_recs.clear();
}

extern "C" int wd_profile_read(int category, double* total_ms, int64_t* launches) {
  using namespace wd;
  if (category < 0 || category >= WD_PROF_NUM || !total_ms || !launches)
    return fail(WD_ERR_INVALID_ARG, "wd_profile_read: bad category or NULL output");
  std::lock_guard<std::mutex> lk(g_mu);
  double ms = 0.0;
  int64_t n = 0;
  for (auto& r : g_recs) {
    if (r.cat != category) continue;
    WD_CUDA_OK(cudaEventSynchronize(r.b));
    float t = 0.f;
    WD_CUDA_OK(cudaEventElapsedTime(&t, r.a, r.b));
    ms += t;
    ++n;
  }
  *total_ms = ms;
  *launches = n;
  return WD_OK;
}

// ABI version of include/wdmpnn.h (3: wd_expand_features, wd_encoder_infer, wd_store_assemble, WdConfig.dropout*, WdGraph.rev_adjacent;
// 4: wd_adam_step)
extern "C" int wd_abi_version(void) { return 7; }
