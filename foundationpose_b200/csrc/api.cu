// Error reporting and bookkeeping shared by all C-ABI entry points.
#include "common.cuh"
#include <string.h>

static thread_local char g_err[512] = "";
static thread_local long long g_launches = 0;

void fpb_set_error(const char* fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void fpb_count_launch(int n) { g_launches += n; }

FPB_EXPORT const char* fpb_last_error(void) { return g_err; }
FPB_EXPORT int fpb_version(void) { return 100; }
FPB_EXPORT long long fpb_launch_count(int reset)
{
  long long v = g_launches;
  if (reset) g_launches = 0;
  return v;
}
