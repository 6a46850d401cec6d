// Launch counter and optional per-category CUDA-event timing (bench.py's roofline leg).
// Off by default; when enabled, categories are timed with event pairs on the launching stream.
#include <atomic>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace aether {

std::atomic<unsigned long long> g_launch_count{0};

namespace {
struct Span { cudaEvent_t a, b; int cat; };
std::mutex g_mu;
std::vector<Span> g_spans;
bool g_on = false;
constexpr size_t MAX_SPANS = 1 << 15;
}  // namespace

ProfScope::ProfScope(int cat, cudaStream_t st) : st_(st), idx_(-1) {
  if (!g_on) return;
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_spans.size() >= MAX_SPANS) return;
  Span s;
  s.cat = cat;
  if (cudaEventCreate(&s.a) != cudaSuccess) return;
  if (cudaEventCreate(&s.b) != cudaSuccess) { cudaEventDestroy(s.a); return; }
  cudaEventRecord(s.a, st);
  g_spans.push_back(s);
  idx_ = (long)g_spans.size() - 1;
}

ProfScope::~ProfScope() {
  if (idx_ < 0) return;
  std::lock_guard<std::mutex> lk(g_mu);
  if ((size_t)idx_ < g_spans.size()) cudaEventRecord(g_spans[idx_].b, st_);
}

}  // namespace aether

using namespace aether;

extern "C" uint64_t aether_launch_count(void) { return g_launch_count.load(std::memory_order_relaxed); }

extern "C" int aether_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_mu);
  g_on = on != 0;
  return 0;
}

// Synchronises the device, sums elapsed milliseconds and span counts per category, clears the spans.
extern "C" int aether_profile_collect(double* ms_by_cat, int64_t* spans_by_cat, int32_t n_cat) {
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return (int)e;
  std::lock_guard<std::mutex> lk(g_mu);
  for (int i = 0; i < n_cat; ++i) { ms_by_cat[i] = 0.0; spans_by_cat[i] = 0; }
  for (auto& s : g_spans) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess && s.cat >= 0 && s.cat < n_cat) {
      ms_by_cat[s.cat] += ms;
      spans_by_cat[s.cat] += 1;
    }
    cudaEventDestroy(s.a);
    cudaEventDestroy(s.b);
  }
  g_spans.clear();
  return 0;
}
