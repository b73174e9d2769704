// Pose-error metrics over a batch of hypotheses:
//   add_err    src/Utils.py:232-240   mean_i || (R_p x_i + t_p) - (R_g x_i + t_g) ||
//   adds_err   src/Utils.py:242-253   mean_i min_j || gt_i - pred_j ||     (cKDTree query in the reference)
// One CTA per pose; the nearest-neighbour search of ADD-S is an exhaustive scan over shared-memory tiles of the
// transformed prediction points (P^2 fp32 distance evaluations per pose: ~1e8 for 10k points, far below a kd-tree
// build on the host, and the batch dimension fills the SMs).  Sums are reduced in a fixed order: deterministic.
// Compiled with -fmad=false like the other geometry units.
#include "common.cuh"

namespace {

constexpr int MT = 256;

__device__ __forceinline__ void xform(const float* T, float x, float y, float z, float& ox, float& oy, float& oz)
{
  ox = (T[0] * x + T[1] * y) + (T[2] * z + T[3]);
  oy = (T[4] * x + T[5] * y) + (T[6] * z + T[7]);
  oz = (T[8] * x + T[9] * y) + (T[10] * z + T[11]);
}

__device__ double block_sum(double v, double* sh)
{
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < MT / 32; w++) t += sh[w];
  return t;
}

__global__ void __launch_bounds__(MT) add_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int gt_stride,
                                                 const float* __restrict__ pts, int P, float* __restrict__ out)
{
  __shared__ float Tp[12], Tg[12];
  __shared__ double sh[MT / 32];
  const int b = blockIdx.x;
  if (threadIdx.x < 12) { Tp[threadIdx.x] = pred[(size_t)b * 16 + threadIdx.x]; Tg[threadIdx.x] = gt[(size_t)b * gt_stride + threadIdx.x]; }
  __syncthreads();
  double acc = 0.0;
  for (int i = threadIdx.x; i < P; i += MT) {
    const float x = pts[3 * i], y = pts[3 * i + 1], z = pts[3 * i + 2];
    float ax, ay, az, bx, by, bz;
    xform(Tp, x, y, z, ax, ay, az);
    xform(Tg, x, y, z, bx, by, bz);
    const float dx = ax - bx, dy = ay - by, dz = az - bz;
    acc += (double)sqrtf((dx * dx + dy * dy) + dz * dz);
  }
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) out[b] = (float)(t / (double)P);
}

__global__ void __launch_bounds__(MT) adds_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int gt_stride,
                                                  const float* __restrict__ pts, int P, float* __restrict__ out)
{
  __shared__ float Tp[12], Tg[12];
  __shared__ float4 tile[MT];
  __shared__ double sh[MT / 32];
  const int b = blockIdx.x;
  if (threadIdx.x < 12) { Tp[threadIdx.x] = pred[(size_t)b * 16 + threadIdx.x]; Tg[threadIdx.x] = gt[(size_t)b * gt_stride + threadIdx.x]; }
  __syncthreads();
  double acc = 0.0;
  for (int i0 = 0; i0 < P; i0 += MT) {                      // MT ground-truth points per pass, one per thread
    const int i = i0 + threadIdx.x;
    float gx = 0.0f, gy = 0.0f, gz = 0.0f;
    if (i < P) xform(Tg, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2], gx, gy, gz);
    float best = 3.0e38f;
    for (int j0 = 0; j0 < P; j0 += MT) {
      const int j = j0 + threadIdx.x;
      __syncthreads();
      if (j < P) {
        float4 q;
        xform(Tp, pts[3 * j], pts[3 * j + 1], pts[3 * j + 2], q.x, q.y, q.z);
        q.w = 0.0f;
        tile[threadIdx.x] = q;
      }
      __syncthreads();
      const int n = min(MT, P - j0);
#pragma unroll 8
      for (int k = 0; k < n; k++) {
        const float4 q = tile[k];
        const float dx = gx - q.x, dy = gy - q.y, dz = gz - q.z;
        best = fminf(best, (dx * dx + dy * dy) + dz * dz);
      }
    }
    if (i < P) acc += (double)sqrtf(best);
  }
  __syncthreads();
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) out[b] = (float)(t / (double)P);
}

}  // namespace

FPB_EXPORT int fpb_add_err(const float* pred, const float* gt, int gt_batched, const float* pts, int P, int B, float* out,
                           void* stream)
{
  FPB_CHECK_ARG(pred && gt && pts && out && P > 0 && B > 0);
  add_kernel<<<B, MT, 0, (cudaStream_t)stream>>>(pred, gt, gt_batched ? 16 : 0, pts, P, out);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

FPB_EXPORT int fpb_adds_err(const float* pred, const float* gt, int gt_batched, const float* pts, int P, int B, float* out,
                            void* stream)
{
  FPB_CHECK_ARG(pred && gt && pts && out && P > 0 && B > 0);
  adds_kernel<<<B, MT, 0, (cudaStream_t)stream>>>(pred, gt, gt_batched ? 16 : 0, pts, P, out);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}
