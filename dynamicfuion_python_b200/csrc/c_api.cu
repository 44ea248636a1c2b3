// Error channel and version of the C ABI.
#include "host_util.h"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {
static thread_local char g_last_error[512] = "";
void set_last_error(const char* fmt, ...) {
	va_list args;
	va_start(args, fmt);
	vsnprintf(g_last_error, sizeof(g_last_error), fmt, args);
	va_end(args);
}
}  // namespace nnrt_b200

extern "C" const char* nnrt_b200_last_error(void) { return nnrt_b200::g_last_error; }
extern "C" int nnrt_b200_version(void) { return 100; }
