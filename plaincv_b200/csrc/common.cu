// common.cu — library-wide bookkeeping for libplaincv_b200.so.
#include "pcv_internal.h"
#include <atomic>

namespace pcv {
static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace pcv

extern "C" int pcv_version(void) { return 100; }
extern "C" long long pcv_launch_count(void) { return pcv::g_launches.load(std::memory_order_relaxed); }
