// common.cu — library-wide bookkeeping for libplaincv_b200.so.
#include "pcv_internal.h"
#include <atomic>
#include <cstdlib>

namespace pcv {
static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
static std::atomic<int> g_pdl{-1};
bool pdl_enabled() {
  int v = g_pdl.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("PCV_PDL");
    v = (e && e[0] == '1') ? 1 : 0;  // off by default: measured no gain (46.43 vs 46.37 ms per step at C3)
    g_pdl.store(v, std::memory_order_relaxed);
  }
  return v != 0;
}
}  // namespace pcv

extern "C" int pcv_version(void) { return 100; }
extern "C" long long pcv_launch_count(void) { return pcv::g_launches.load(std::memory_order_relaxed); }
extern "C" int pcv_set_pdl(int on) {
  pcv::g_pdl.store(on ? 1 : 0, std::memory_order_relaxed);
  return PCV_OK;
}
