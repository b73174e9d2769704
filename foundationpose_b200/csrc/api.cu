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

// FoundationPose.register() prologue (src/estimater.py:159-205) as ONE host call and without a host round trip:
// erode -> bilateral -> mask statistics -> translation guess -> N hypotheses -> xyz map.
extern "C" {
int fpb_erode_depth_f32(const float*, float*, int, int, int, float, float, float, void*);
int fpb_bilateral_depth_f32(const float*, float*, int, int, int, float, float, float, void*);
int fpb_depth_to_xyz(const float*, const double*, float, float*, int, int, int, void*);
int fpb_mask_depth_stats(const float*, const uint8_t*, int, int, int32_t*, float*, void*);
int fpb_hypotheses_from_stats(const int32_t*, const float*, const double*, const float*, int, int, float*, double*, void*);
}
FPB_EXPORT int fpb_register_prologue(const float* depth_in, const uint8_t* mask, int H, int W, const double* K9_h,
                                     const double* Kinv9_h, float* depth_tmp, float* depth_out, float* xyz_out,
                                     const float* rot_grid, int N, int min_valid, float* poses_out, int32_t* stats_i32,
                                     float* stats_f32, double* center_out, void* stream)
{
  FPB_CHECK_ARG(depth_in && mask && K9_h && Kinv9_h && depth_tmp && depth_out && xyz_out && rot_grid && poses_out && stats_i32 &&
                stats_f32 && H > 0 && W > 0 && N > 0);
  int rc;
  // defaults of erode_depth / bilateral_filter_depth as called by register(): src/estimater.py:171-172
  if ((rc = fpb_erode_depth_f32(depth_in, depth_tmp, H, W, 2, 0.001f, 0.8f, 100.0f, stream))) return rc;
  if ((rc = fpb_bilateral_depth_f32(depth_tmp, depth_out, H, W, 2, 100.0f, 2.0f, 100000.0f, stream))) return rc;
  if ((rc = fpb_mask_depth_stats(depth_out, mask, H, W, stats_i32, stats_f32, stream))) return rc;
  if ((rc = fpb_hypotheses_from_stats(stats_i32, stats_f32, Kinv9_h, rot_grid, N, min_valid, poses_out, center_out, stream))) return rc;
  return fpb_depth_to_xyz(depth_out, K9_h, 3.0e38f, xyz_out, H, W, 1, stream);
}
