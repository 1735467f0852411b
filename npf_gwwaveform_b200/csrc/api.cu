// Library-level entry points: version, last error, launch counter.
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace npfk {
static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace npfk

extern "C" {
int npfk_version(void) { return 100; }
const char* npfk_last_error(void) { return npfk::g_err; }
int64_t npfk_launch_count(void) { return npfk::g_launches.load(std::memory_order_relaxed); }
}
