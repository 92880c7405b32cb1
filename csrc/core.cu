// Library-wide state: error text, launch counter, device queries.
#include <stdarg.h>

#include "common.cuh"

namespace dcnn {
static thread_local char g_err[1024] = "";
unsigned long long g_launch_count = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }
}  // namespace dcnn

DCNN_API const char* dcnn_last_error() { return dcnn::last_error(); }
DCNN_API unsigned long long dcnn_launch_count() { return dcnn::g_launch_count; }
DCNN_API void dcnn_reset_launch_count() { dcnn::g_launch_count = 0; }
DCNN_API const char* dcnn_version() { return "dcnn_b200 0.1.0 (sm_100a)"; }
DCNN_API int dcnn_device_sms() { return dcnn::num_sms(); }
