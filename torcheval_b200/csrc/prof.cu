#include "prof.cuh"

#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace te {
namespace {

struct Span {
  const char* name;
  cudaEvent_t a, b;
};
std::mutex g_mu;
bool g_on = false;
std::vector<Span> g_spans;

}  // namespace

ProfScope::ProfScope(const char* name, cudaStream_t st) : slot_(-1), st_(st) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_on) return;
  Span s{name, nullptr, nullptr};
  if (cudaEventCreate(&s.a) != cudaSuccess || cudaEventCreate(&s.b) != cudaSuccess) return;
  cudaEventRecord(s.a, st);
  slot_ = (int)g_spans.size();
  g_spans.push_back(s);
}

ProfScope::~ProfScope() {
  if (slot_ < 0) return;
  std::lock_guard<std::mutex> lk(g_mu);
  if (slot_ < (int)g_spans.size()) cudaEventRecord(g_spans[slot_].b, st_);
}

}  // namespace te

using namespace te;

extern "C" int te_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (auto& s : g_spans) {
    cudaEventDestroy(s.a);
    cudaEventDestroy(s.b);
  }
  g_spans.clear();
  g_on = on != 0;
  return TE_OK;
}

extern "C" int te_profile_read(char* names, size_t names_bytes, double* total_ms, int64_t* launches,
                               int32_t max_entries) {
  std::lock_guard<std::mutex> lk(g_mu);
  std::vector<std::string> uniq;
  std::vector<double> ms;
  std::vector<int64_t> cnt;
  for (auto& s : g_spans) {
    if (cudaEventSynchronize(s.b) != cudaSuccess) continue;
    float t = 0.f;
    if (cudaEventElapsedTime(&t, s.a, s.b) != cudaSuccess) continue;
    size_t i = 0;
    for (; i < uniq.size(); ++i)
      if (uniq[i] == s.name) break;
    if (i == uniq.size()) {
      uniq.emplace_back(s.name);
      ms.push_back(0.0);
      cnt.push_back(0);
    }
    ms[i] += t;
    cnt[i] += 1;
  }
  int n = 0;
  size_t off = 0;
  if (names && names_bytes) names[0] = 0;
  for (size_t i = 0; i < uniq.size() && n < max_entries; ++i) {
    const size_t need = uniq[i].size() + 1;
    if (!names || off + need + 1 > names_bytes) break;
    memcpy(names + off, uniq[i].c_str(), need - 1);
    names[off + need - 1] = '\n';
    names[off + need] = 0;
    off += need;
    total_ms[n] = ms[i];
    launches[n] = cnt[i];
    ++n;
  }
  return n;
}
