// Error plumbing and bookkeeping behind the C ABI (include/soda_b200.h).
#include <stdarg.h>

#include "common.cuh"

namespace soda {

static thread_local char t_error[512] = "";
std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_error, sizeof(t_error), fmt, ap);
  va_end(ap);
}

}  // namespace soda

extern "C" int soda_abi_version(void) { return SODA_ABI_VERSION; }
extern "C" const char* soda_last_error(void) { return soda::t_error; }
extern "C" int64_t soda_launch_count(void) { return soda::g_launches.load(); }
