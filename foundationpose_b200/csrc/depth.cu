// Depth pre-processing stencils, depth->xyz and guess_translation statistics.
// Compiled with -fmad=false: every fp32 op rounds on its own so results are
// bit-identical to the reference's Warp kernels / eager arithmetic.
//   erode_depth_kernel             src/Utils.py:359-384
//   bilateral_filter_depth_kernel  src/Utils.py:304-343
//   depth2xyzmap / _batch          src/Utils.py:399-438
//   guess_translation              src/estimater.py:137-156
#include "common.cuh"
#include <cooperative_groups.h>
#include <stdlib.h>
namespace cg = cooperative_groups;

namespace {

constexpr int TX = 32, TY = 16, MAXR = 4;

// Stage a (TX+2r) x (TY+2r) tile; out-of-image taps are flagged with NaN-free sentinel -1
// and skipped by coordinates (the reference skips them, it does not treat them as invalid).
template <int MODE>  // 0 erode, 1 bilateral
__global__ void __launch_bounds__(TX* TY) stencil_kernel(const float* __restrict__ depth, float* __restrict__ out,
                                                         int H, int W, int radius, float p0, float p1, float p2,
                                                         float zfar)
{
  __shared__ float tile[(TY + 2 * MAXR)][(TX + 2 * MAXR) + 1];
  const int bx = blockIdx.x * TX, by = blockIdx.y * TY;
  const int tw = TX + 2 * radius, th = TY + 2 * radius;
  for (int i = threadIdx.y * TX + threadIdx.x; i < tw * th; i += TX * TY) {
    int ly = i / tw, lx = i - ly * tw;
    int gy = by + ly - radius, gx = bx + lx - radius;
    tile[ly][lx] = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? __ldg(depth + (size_t)gy * W + gx) : 0.0f;
  }
  __syncthreads();
  const int w = bx + threadIdx.x, h = by + threadIdx.y;
  if (w >= W || h >= H) return;
  const int lx0 = threadIdx.x + radius, ly0 = threadIdx.y + radius;
  const float dc = tile[ly0][lx0];
  if (MODE == 0) {
    const float diff_thres = p0, ratio_thres = p1;
    float bad = 0.0f, total = 0.0f;
    for (int du = -radius; du <= radius; du++) {
      int u = w + du;
      if (u < 0 || u >= W) continue;
      for (int dv = -radius; dv <= radius; dv++) {
        int v = h + dv;
        if (v < 0 || v >= H) continue;
        float cur = tile[ly0 + dv][lx0 + du];
        total += 1.0f;
        if (cur < 0.001f || cur >= zfar || fabsf(cur - dc) > diff_thres) bad += 1.0f;
      }
    }
    out[(size_t)h * W + w] = (bad / total > ratio_thres) ? 0.0f : dc;
  } else {
    const float sigmaD = p0, sigmaR = p1;
    float res = 0.0f, mean = 0.0f;
    int nvalid = 0;
    for (int du = -radius; du <= radius; du++) {
      int u = w + du;
      if (u < 0 || u >= W) continue;
      for (int dv = -radius; dv <= radius; dv++) {
        int v = h + dv;
        if (v < 0 || v >= H) continue;
        float cur = tile[ly0 + dv][lx0 + du];
        if (cur >= 0.001f && cur < zfar) { nvalid++; mean += cur; }
      }
    }
    if (nvalid > 0) {
      mean /= (float)nvalid;
      const float denD = 2.0f * sigmaD * sigmaD, denR = 2.0f * sigmaR * sigmaR;
      float sw = 0.0f, s = 0.0f;
      for (int du = -radius; du <= radius; du++) {
        int u = w + du;
        if (u < 0 || u >= W) continue;
        for (int dv = -radius; dv <= radius; dv++) {
          int v = h + dv;
          if (v < 0 || v >= H) continue;
          float cur = tile[ly0 + dv][lx0 + du];
          if (cur >= 0.001f && cur < zfar && fabsf(cur - mean) < 0.01f) {
            float a = -(float)(du * du + dv * dv) / denD;
            float b = (dc - cur) * (dc - cur) / denR;
            float wgt = expf(a - b);
            sw += wgt;
            s += wgt * cur;
          }
        }
      }
      if (sw > 0.0f) res = s / sw;
    }
    out[(size_t)h * W + w] = res;
  }
}

__global__ void depth2xyz_kernel(const float* __restrict__ depth, float* __restrict__ xyz, int H, int W, double fx,
                                 double fy, double cx, double cy, float zfar, int math_f64)
{
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= H * W) return;
  int v = idx / W, u = idx - v * W;
  float z = depth[idx];
  float x = 0.0f, y = 0.0f, zz = 0.0f;
  bool invalid = math_f64 ? (z < 0.001f) : (z < 0.001f || z > zfar);
  if (!invalid) {
    if (math_f64) {
      x = (float)(((double)u - cx) * (double)z / fx);
      y = (float)(((double)v - cy) * (double)z / fy);
    } else {
      x = ((float)u - (float)cx) * z / (float)fx;
      y = ((float)v - (float)cy) * z / (float)fy;
    }
    zz = z;
  }
  xyz[(size_t)idx * 3 + 0] = x;
  xyz[(size_t)idx * 3 + 1] = y;
  xyz[(size_t)idx * 3 + 2] = zz;
}

// ---- guess_translation statistics: one CTA, exact order statistics by 8-bit radix select ----
constexpr int ST = 1024;

__device__ inline bool px_valid(const float* depth, const uint8_t* mask, int i, float& d)
{
  d = depth[i];
  return mask[i] != 0 && d >= 0.001f;
}

// exact rank-th smallest valid depth inside the mask's bounding box [u0, u0+bw) x [v0, v0+bh) (4 x 8-bit radix passes)
__device__ uint32_t radix_select(const float* depth, const uint8_t* mask, int W, int u0, int v0, int bw, int bh, uint32_t rank,
                                 uint32_t* hist, uint32_t* sh)
{
  uint32_t prefix = 0, pmask = 0;
  const int n = bw * bh;
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int i = threadIdx.x; i < 256; i += ST) hist[i] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += ST) {
      const int r = i / bw;
      float d;
      if (px_valid(depth, mask, (v0 + r) * W + u0 + (i - r * bw), d)) {
        uint32_t b = __float_as_uint(d);
        if ((b & pmask) == prefix) atomicAdd(&hist[(b >> shift) & 255u], 1u);
      }
    }
    __syncthreads();
    if (threadIdx.x < 32) {                       // warp 0: first bin whose inclusive prefix count exceeds the rank
      const int lane = threadIdx.x;
      uint32_t c[8], tot = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) { c[k] = hist[lane * 8 + k]; tot += c[k]; }
      uint32_t incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      const uint32_t excl = incl - tot;
      const bool mine = excl <= rank && rank < incl;          // exactly one lane when rank < total count
      if (mine) {
        uint32_t acc = excl;
        int k = 0;
        for (; k < 8; k++) {
          if (acc + c[k] > rank) break;
          acc += c[k];
        }
        sh[0] = (uint32_t)(lane * 8 + k);
        sh[1] = rank - acc;
      }
    }
    __syncthreads();
    prefix |= sh[0] << shift;
    pmask |= 255u << shift;
    rank = sh[1];
    __syncthreads();
  }
  return prefix;
}

// One block: (1) bounding box + counts over the whole mask, 16 mask bytes per load (most of the frame is outside the
// mask); (2) the two middle order statistics of the valid depths by radix selection restricted to the bounding box.
__global__ void __launch_bounds__(ST) mask_stats_kernel(const float* __restrict__ depth, const uint8_t* __restrict__ mask,
                                                        int H, int W, int32_t* out_i, float* out_f)
{
  __shared__ uint32_t hist[256];
  __shared__ uint32_t sh[2];
  __shared__ int s_umin, s_umax, s_vmin, s_vmax, s_nmask, s_nvalid;
  const int n = H * W;
  if (threadIdx.x == 0) { s_umin = W; s_umax = -1; s_vmin = H; s_vmax = -1; s_nmask = 0; s_nvalid = 0; }
  __syncthreads();
  int umin = W, umax = -1, vmin = H, vmax = -1, nm = 0, nv = 0;
  auto visit = [&](int i) {
    const int v = i / W, u = i - v * W;
    umin = min(umin, u); umax = max(umax, u); vmin = min(vmin, v); vmax = max(vmax, v);
    nm++;
    if (depth[i] >= 0.001f) nv++;
  };
  const int head = min(n, (int)((16 - ((uintptr_t)mask & 15)) & 15));       // bytes before the first 16-byte boundary
  const int nvec = (n - head) / 16;
  for (int i = threadIdx.x; i < head; i += ST)
    if (mask[i]) visit(i);
  const uint4* mv = reinterpret_cast<const uint4*>(mask + head);
  for (int q = threadIdx.x; q < nvec; q += ST) {
    const uint4 m = __ldg(mv + q);
    if ((m.x | m.y | m.z | m.w) == 0u) continue;
    const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
    for (int k = 0; k < 16; k++)
      if ((w[k >> 2] >> ((k & 3) * 8)) & 255u) visit(head + q * 16 + k);
  }
  for (int i = head + nvec * 16 + threadIdx.x; i < n; i += ST)
    if (mask[i]) visit(i);
  atomicMin(&s_umin, umin); atomicMax(&s_umax, umax); atomicMin(&s_vmin, vmin); atomicMax(&s_vmax, vmax);
  atomicAdd(&s_nmask, nm); atomicAdd(&s_nvalid, nv);
  __syncthreads();
  const int nvalid = s_nvalid;
  float m0 = 0.0f, m1 = 0.0f;
  if (nvalid > 0) {
    const int u0 = s_umin, v0 = s_vmin, bw = s_umax - s_umin + 1, bh = s_vmax - s_vmin + 1;
    uint32_t a = radix_select(depth, mask, W, u0, v0, bw, bh, (uint32_t)((nvalid - 1) / 2), hist, sh);
    uint32_t b = (nvalid & 1) ? a : radix_select(depth, mask, W, u0, v0, bw, bh, (uint32_t)(nvalid / 2), hist, sh);
    m0 = __uint_as_float(a);
    m1 = __uint_as_float(b);
  }
  if (threadIdx.x == 0) {
    out_i[0] = s_umin; out_i[1] = s_umax; out_i[2] = s_vmin; out_i[3] = s_vmax;
    out_i[4] = s_nmask; out_i[5] = nvalid;
    out_f[0] = m0; out_f[1] = m1;
  }
}

// ---- the same statistics on a thread-block cluster of 8 CTAs (8192 threads) ----
// The one-CTA kernel above took 79 us at 640 x 480 -- rank-redundant fixed cost of every register(), 1.2 % of the 8-GPU
// step.  Here the mask scan and every radix pass are spread over the cluster: each CTA counts its slice into its own
// shared-memory histogram, the non-empty bins are added to CTA 0's histogram through distributed shared memory, CTA 0
// picks the digit and the other CTAs read it back from its shared memory.  Integer counts: the result is identical.
constexpr int MS_CL = 8;

__device__ uint32_t radix_select_cluster(cg::cluster_group& cluster, const float* depth, const uint8_t* mask, int W, int u0,
                                         int v0, int bw, int bh, uint32_t rank, uint32_t* hist, uint32_t* hist0, uint32_t* sh,
                                         const uint32_t* sh0)
{
  const unsigned crank = cluster.block_rank();
  uint32_t prefix = 0, pmask = 0;
  const int n = bw * bh;
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int i = threadIdx.x; i < 256; i += ST) hist[i] = 0;
    cluster.sync();                                            // CTA 0's histogram is clear (and sh of the last pass consumed)
    __shared__ uint32_t local[256];
    for (int i = threadIdx.x; i < 256; i += ST) local[i] = 0;
    __syncthreads();
    for (int i = crank * ST + threadIdx.x; i < n; i += MS_CL * ST) {
      const int r = i / bw;
      float d;
      if (px_valid(depth, mask, (v0 + r) * W + u0 + (i - r * bw), d)) {
        uint32_t b = __float_as_uint(d);
        if ((b & pmask) == prefix) atomicAdd(&local[(b >> shift) & 255u], 1u);
      }
    }
    __syncthreads();
    if (threadIdx.x < 256 && local[threadIdx.x]) atomicAdd(hist0 + threadIdx.x, local[threadIdx.x]);   // DSMEM (CTA 0: its own)
    cluster.sync();
    if (crank == 0 && threadIdx.x < 32) {                       // warp 0 of CTA 0: first bin whose inclusive prefix count exceeds the rank
      const int lane = threadIdx.x;
      uint32_t c[8], tot = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) { c[k] = hist[lane * 8 + k]; tot += c[k]; }
      uint32_t incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      const uint32_t excl = incl - tot;
      if (excl <= rank && rank < incl) {
        uint32_t a = excl;
        int k = 0;
        for (; k < 8; k++) {
          if (a + c[k] > rank) break;
          a += c[k];
        }
        sh[0] = (uint32_t)(lane * 8 + k);
        sh[1] = rank - a;
      }
    }
    cluster.sync();
    prefix |= sh0[0] << shift;
    pmask |= 255u << shift;
    rank = sh0[1];
  }
  cluster.sync();                                              // nobody still reads CTA 0's sh when the next select starts
  return prefix;
}

__global__ void __cluster_dims__(MS_CL, 1, 1) __launch_bounds__(ST)
mask_stats_cluster_kernel(const float* __restrict__ depth, const uint8_t* __restrict__ mask, int H, int W, int32_t* out_i,
                          float* out_f)
{
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ uint32_t hist[256];
  __shared__ uint32_t sh[2];
  __shared__ int s_box[6];                                     // umin, umax, vmin, vmax, nmask, nvalid
  const unsigned crank = cluster.block_rank();
  uint32_t* hist0 = cluster.map_shared_rank(hist, 0);
  const uint32_t* sh0 = cluster.map_shared_rank(sh, 0);
  int* box0 = cluster.map_shared_rank(s_box, 0);
  const int n = H * W;
  if (threadIdx.x == 0) { s_box[0] = W; s_box[1] = -1; s_box[2] = H; s_box[3] = -1; s_box[4] = 0; s_box[5] = 0; }
  cluster.sync();
  int umin = W, umax = -1, vmin = H, vmax = -1, nm = 0, nv = 0;
  auto visit = [&](int i) {
    const int v = i / W, u = i - v * W;
    umin = min(umin, u); umax = max(umax, u); vmin = min(vmin, v); vmax = max(vmax, v);
    nm++;
    if (depth[i] >= 0.001f) nv++;
  };
  const int gt = crank * ST + threadIdx.x, gs = MS_CL * ST;
  const int head = min(n, (int)((16 - ((uintptr_t)mask & 15)) & 15));       // bytes before the first 16-byte boundary
  const int nvec = (n - head) / 16;
  for (int i = gt; i < head; i += gs)
    if (mask[i]) visit(i);
  const uint4* mv = reinterpret_cast<const uint4*>(mask + head);
  for (int q = gt; q < nvec; q += gs) {
    const uint4 m = __ldg(mv + q);
    if ((m.x | m.y | m.z | m.w) == 0u) continue;
    const uint32_t w[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
    for (int k = 0; k < 16; k++)
      if ((w[k >> 2] >> ((k & 3) * 8)) & 255u) visit(head + q * 16 + k);
  }
  for (int i = head + nvec * 16 + gt; i < n; i += gs)
    if (mask[i]) visit(i);
  // warp-level reduction first, then one set of DSMEM atomics per warp
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    umin = min(umin, __shfl_xor_sync(0xffffffffu, umin, o)); umax = max(umax, __shfl_xor_sync(0xffffffffu, umax, o));
    vmin = min(vmin, __shfl_xor_sync(0xffffffffu, vmin, o)); vmax = max(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    nm += __shfl_xor_sync(0xffffffffu, nm, o); nv += __shfl_xor_sync(0xffffffffu, nv, o);
  }
  if ((threadIdx.x & 31) == 0 && nm > 0) {
    atomicMin(box0 + 0, umin); atomicMax(box0 + 1, umax); atomicMin(box0 + 2, vmin); atomicMax(box0 + 3, vmax);
    atomicAdd(box0 + 4, nm); atomicAdd(box0 + 5, nv);
  }
  cluster.sync();
  const int b_umin = box0[0], b_umax = box0[1], b_vmin = box0[2], b_vmax = box0[3], nmask = box0[4], nvalid = box0[5];
  float m0 = 0.0f, m1 = 0.0f;
  if (nvalid > 0) {                                            // cluster-uniform
    const int bw = b_umax - b_umin + 1, bh = b_vmax - b_vmin + 1;
    uint32_t a = radix_select_cluster(cluster, depth, mask, W, b_umin, b_vmin, bw, bh, (uint32_t)((nvalid - 1) / 2), hist, hist0, sh, sh0);
    uint32_t b = (nvalid & 1) ? a
                              : radix_select_cluster(cluster, depth, mask, W, b_umin, b_vmin, bw, bh, (uint32_t)(nvalid / 2), hist, hist0, sh, sh0);
    m0 = __uint_as_float(a);
    m1 = __uint_as_float(b);
  }
  if (crank == 0 && threadIdx.x == 0) {
    out_i[0] = b_umin; out_i[1] = b_umax; out_i[2] = b_vmin; out_i[3] = b_vmax;
    out_i[4] = nmask; out_i[5] = nvalid;
    out_f[0] = m0; out_f[1] = m1;
  }
  cluster.sync();                                              // CTA 0's shared memory stays alive until every CTA has read it
}

}  // namespace

FPB_EXPORT int fpb_erode_depth_f32(const float* in, float* out, int H, int W, int radius, float diff_thres,
                                   float ratio_thres, float zfar, void* stream)
{
  FPB_CHECK_ARG(in && out && H > 0 && W > 0 && radius >= 0 && radius <= MAXR && in != out);
  dim3 grid(fpb_cdiv(W, TX), fpb_cdiv(H, TY)), block(TX, TY);
  stencil_kernel<0><<<grid, block, 0, (cudaStream_t)stream>>>(in, out, H, W, radius, diff_thres, ratio_thres, 0.f, zfar);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

FPB_EXPORT int fpb_bilateral_depth_f32(const float* in, float* out, int H, int W, int radius, float zfar,
                                       float sigmaD, float sigmaR, void* stream)
{
  FPB_CHECK_ARG(in && out && H > 0 && W > 0 && radius >= 0 && radius <= MAXR && in != out);
  dim3 grid(fpb_cdiv(W, TX), fpb_cdiv(H, TY)), block(TX, TY);
  stencil_kernel<1><<<grid, block, 0, (cudaStream_t)stream>>>(in, out, H, W, radius, sigmaD, sigmaR, 0.f, zfar);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

FPB_EXPORT int fpb_depth_to_xyz(const float* depth, const double* K9_h, float zfar, float* xyz_hw3, int H, int W,
                                int math_f64, void* stream)
{
  FPB_CHECK_ARG(depth && K9_h && xyz_hw3 && H > 0 && W > 0);
  depth2xyz_kernel<<<fpb_cdiv((long long)H * W, 256), 256, 0, (cudaStream_t)stream>>>(
      depth, xyz_hw3, H, W, K9_h[0], K9_h[4], K9_h[2], K9_h[5], zfar, math_f64);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

FPB_EXPORT int fpb_mask_depth_stats(const float* depth, const uint8_t* mask, int H, int W, int32_t* out_i32,
                                    float* out_f32, void* stream)
{
  FPB_CHECK_ARG(depth && mask && out_i32 && out_f32 && H > 0 && W > 0);
  static int one_cta = -1;        // FPB_MASK_STATS_1CTA=1: the one-block kernel (kept as the reference for the cluster version)
  if (one_cta < 0) { const char* e = getenv("FPB_MASK_STATS_1CTA"); one_cta = e ? atoi(e) : 0; }
  if (one_cta) mask_stats_kernel<<<1, ST, 0, (cudaStream_t)stream>>>(depth, mask, H, W, out_i32, out_f32);
  else mask_stats_cluster_kernel<<<MS_CL, ST, 0, (cudaStream_t)stream>>>(depth, mask, H, W, out_i32, out_f32);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}
