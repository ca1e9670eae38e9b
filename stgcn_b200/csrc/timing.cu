// Optional per-kernel-class device timing: cudaEvent pairs recorded on the launching stream
// around the library's own launches.  bench.py uses it to get the live duration (and the
// algorithmic work) of the dominant kernel inside the timed region; off by default, and
// never active during CUDA-graph capture.
#include <mutex>
#include <vector>

#include "kernels.cuh"

namespace stgcn {

static const char* kClassNames[KC_COUNT] = {"umma_conv_gemm", "simt_conv_gemm", "wgrad", "agg_fwd", "agg_bwd", "agg_dA",
                                            "bn_stats", "elementwise", "pack_misc", "head", "umma_wgrad", "umma_spatial"};
struct Rec { cudaEvent_t a, b; int cls; double work; };
static std::mutex g_tmu;
static bool g_timing = false;
static std::vector<Rec> g_recs;
static std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_pool;
static double g_ms[KC_COUNT], g_work[KC_COUNT];
static uint64_t g_n[KC_COUNT];

static void drain_locked() {
  for (auto& r : g_recs) {
    float ms = 0.f;
    if (cudaEventSynchronize(r.b) == cudaSuccess && cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
      g_ms[r.cls] += ms; g_work[r.cls] += r.work; g_n[r.cls] += 1;
    }
    g_pool.emplace_back(r.a, r.b);
  }
  g_recs.clear();
}

KernelTimer::KernelTimer(int cls, double work, cudaStream_t st) : idx_(-1), st_(st) {
  if (!g_timing) return;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return;
  std::lock_guard<std::mutex> lk(g_tmu);
  if (g_recs.size() >= 8192) drain_locked();
  Rec r; r.cls = cls; r.work = work;
  if (!g_pool.empty()) { r.a = g_pool.back().first; r.b = g_pool.back().second; g_pool.pop_back(); }
  else { cudaEventCreate(&r.a); cudaEventCreate(&r.b); }
  cudaEventRecord(r.a, st);
  g_recs.push_back(r);
  idx_ = (int)g_recs.size() - 1;
}
KernelTimer::~KernelTimer() {
  if (idx_ < 0) return;
  std::lock_guard<std::mutex> lk(g_tmu);
  if (idx_ < (int)g_recs.size()) cudaEventRecord(g_recs[idx_].b, st_);
}

}  // namespace stgcn

using namespace stgcn;
extern "C" {
void stgcn_timing_enable(int32_t on) { std::lock_guard<std::mutex> lk(g_tmu); g_timing = on != 0; }
void stgcn_timing_reset(void) {
  std::lock_guard<std::mutex> lk(g_tmu);
  drain_locked();
  for (int i = 0; i < KC_COUNT; ++i) { g_ms[i] = 0; g_work[i] = 0; g_n[i] = 0; }
}
int32_t stgcn_timing_classes(void) { return KC_COUNT; }
const char* stgcn_timing_class_name(int32_t cls) { return cls >= 0 && cls < KC_COUNT ? kClassNames[cls] : ""; }
int32_t stgcn_timing_read(int32_t cls, double* ms, double* work, uint64_t* launches) {
  if (cls < 0 || cls >= KC_COUNT) return STGCN_ERR_INVALID;
  std::lock_guard<std::mutex> lk(g_tmu);
  drain_locked();
  if (ms) *ms = g_ms[cls];
  if (work) *work = g_work[cls];
  if (launches) *launches = g_n[cls];
  return STGCN_OK;
}
}
