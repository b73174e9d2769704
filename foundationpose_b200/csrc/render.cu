// Batched mesh rasteriser with fused shading (network input A) and the fused
// crop / perspective-warp of the observed frame (network input B).
// Compiled with -fmad=false: arithmetic is bit-identical to oracle/fp_oracle.c.
//   nvdiffrast_render                 src/Utils.py:133-219 (+ :752-802 projection)
//   make_crop_data_batch (refiner)    learning/training/predict_pose_refine.py:26-89
//   make_crop_data_batch (scorer)     learning/training/predict_score.py:57-114
//   transform_batch                   learning/datasets/h5_dataset.py:79-127, 137-179
//
// Design: all N hypotheses in one launch per stage.
//   1. vertex kernel   (N x V threads)  clip-space xyzw + camera xyz + diffuse term + the 1/16-px snapped window
//                                       position of every vertex (the two divisions happen once per vertex, not once
//                                       per face corner)                            -> L2-resident scratch
//   2. raster kernel   (N x F threads)  integer coverage on the snapped positions; 64-bit (z/w | face) atomicMin
//                                       z-buffer; the float clip coordinates are only fetched by triangles that cover
//                                       a pixel centre; big triangles are re-distributed over the warp
//   3. shade kernel    (N x S x S)      barycentrics -> xyz / colour|texture / lighting -> normalised
//                                       8-channel NHWC pixel, one 128-bit store (bf16) per pixel
#include "common.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

struct fpb_mesh {
  int V, F, Vt, Ht, Wt;
  float* pos;       // [V,3]
  float* vnormals;  // [V,3]
  int32_t* faces;   // [F,3]
  float* vcolor;    // [V,3] or null
  float4* vcolor4;  // [V] rgb0 (one 128-bit load per corner in the shader) or null
  float* uv;        // [Vt,2] or null
  int32_t* uv_idx;  // [F,3] or null
  float* tex;       // [Ht,Wt,3] or null
};

namespace {

__device__ __forceinline__ uint32_t f2ord(float f)
{
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__device__ __forceinline__ void bary_zw(const float4 c0, const float4 c1, const float4 c2, float fx, float fy, float& b0,
                                        float& b1, float& zw)
{
  float p0x = c0.x - fx * c0.w, p0y = c0.y - fy * c0.w;
  float p1x = c1.x - fx * c1.w, p1y = c1.y - fy * c1.w;
  float p2x = c2.x - fx * c2.w, p2y = c2.y - fy * c2.w;
  float a0 = p1x * p2y - p1y * p2x;
  float a1 = p2x * p0y - p2y * p0x;
  float a2 = p0x * p1y - p0y * p1x;
  float iw = 1.0f / ((a0 + a1) + a2);
  float z = (c0.z * a0 + c1.z * a1) + c2.z * a2;
  float w = (c0.w * a0 + c1.w * a1) + c2.w * a2;
  b0 = fminf(fmaxf(a0 * iw, 0.0f), 1.0f);
  b1 = fminf(fmaxf(a1 * iw, 0.0f), 1.0f);
  zw = fminf(fmaxf(z / w, -1.0f), 1.0f);
}

// Coverage + depth of a triangle that has a vertex behind the camera plane (w <= 0) or beyond the fixed-point range:
// nvdiffrast clips such triangles against the frustum; the equivalent test without clipping is the 2-D homogeneous one
// (Olano & Greer): with a_i the edge functions of the clip-space corners seen from the pixel, the pixel is inside iff
// all a_i share the sign of their sum, the interpolated w is positive (in front of the camera) and z/w lies inside
// the near / far planes.  Float arithmetic (no snapping, no top-left rule): exact ties on shared edges are not resolved.
__device__ __forceinline__ bool homog_cover(const float4 c0, const float4 c1, const float4 c2, float fx, float fy, float& zw)
{
  float p0x = c0.x - fx * c0.w, p0y = c0.y - fy * c0.w;
  float p1x = c1.x - fx * c1.w, p1y = c1.y - fy * c1.w;
  float p2x = c2.x - fx * c2.w, p2y = c2.y - fy * c2.w;
  float a0 = p1x * p2y - p1y * p2x;
  float a1 = p2x * p0y - p2y * p0x;
  float a2 = p0x * p1y - p0y * p1x;
  float sum = (a0 + a1) + a2;
  const bool pos = a0 >= 0.0f && a1 >= 0.0f && a2 >= 0.0f && sum > 0.0f;
  const bool neg = a0 <= 0.0f && a1 <= 0.0f && a2 <= 0.0f && sum < 0.0f;
  if (!(pos || neg)) return false;
  float z = (c0.z * a0 + c1.z * a1) + c2.z * a2;
  float w = (c0.w * a0 + c1.w * a1) + c2.w * a2;
  if (!((w > 0.0f) == (sum > 0.0f)) || w == 0.0f) return false;      // interpolated w <= 0: behind the camera
  zw = z / w;
  return zw >= -1.0f && zw <= 1.0f;                                     // near / far clip
}

constexpr int SNAP_INVALID = INT32_MIN;   // vertex behind the camera (w <= 0) or beyond the fixed-point range

// Clip-space position of one vertex (P * glcam * pose, then the crop zoom): the ONE instruction sequence used by the
// vertex pass and, in lean mode, by the rasteriser and the shader when they recompute a corner (bit-identical values).
__device__ __forceinline__ float4 vertex_clip(float x, float y, float z, const float* M, const float* R)
{
  float c[4];
#pragma unroll
  for (int j = 0; j < 4; j++) c[j] = ((M[j * 4] * x + M[j * 4 + 1] * y) + M[j * 4 + 2] * z) + M[j * 4 + 3];
  float4 cl;
  cl.x = c[0] * R[0] + c[3] * R[2];
  cl.y = c[1] * R[1] + c[3] * R[3];
  cl.z = c[2];
  cl.w = c[3];
  return cl;
}

// camera-frame position + diffuse term (headlight: -normal_z clamped) of one vertex
__device__ __forceinline__ float4 vertex_camd(float x, float y, float z, float nx, float ny, float nz, const float* T)
{
  float cam[3], cn[3];
#pragma unroll
  for (int j = 0; j < 3; j++) {
    cam[j] = ((T[j * 4] * x + T[j * 4 + 1] * y) + T[j * 4 + 2] * z) + T[j * 4 + 3];
    cn[j] = (T[j * 4] * nx + T[j * 4 + 1] * ny) + T[j * 4 + 2] * nz;
  }
  float nrm = sqrtf((cn[0] * cn[0] + cn[1] * cn[1]) + cn[2] * cn[2]);
  float d = -(cn[2] / fmaxf(nrm, 1e-12f));
  d = fminf(fmaxf(d, 0.0f), 1.0f);
  return make_float4(cam[0], cam[1], cam[2], d);
}

// Where the rasteriser / shader get the float clip coordinates of a corner from: the array written by the vertex pass,
// or (lean mode: large meshes, where that array is 16 B x V x N of HBM writes for a few thousand visible corners)
// recomputed from the mesh position with the per-hypothesis matrices.
struct ClipSrc {
  const float4* cl;      // [V] of this hypothesis, or null in lean mode
  const float* pos;      // [V,3]
  const float* M;        // 16 floats (params + 13)
  const float* R;        // 4 floats (params + 29)
  __device__ __forceinline__ float4 get(int i) const
  {
    if (cl) return cl[i];
    return vertex_clip(pos[i * 3], pos[i * 3 + 1], pos[i * 3 + 2], M, R);
  }
};

// LEAN: only the snapped window positions are written (8 B per vertex instead of 40)
template <bool LEAN>
__global__ void __launch_bounds__(256) vertex_kernel(const float* __restrict__ pos, const float* __restrict__ vnormals, int V,
                                                     const float* __restrict__ poses, const float* __restrict__ params,
                                                     int S, float4* __restrict__ clip, float4* __restrict__ camd,
                                                     int2* __restrict__ snap)
{
  const int n = blockIdx.y;
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  __shared__ float sT[16], sM[16], sR[4];
  if (threadIdx.x < 16) {
    sT[threadIdx.x] = poses[(size_t)n * 16 + threadIdx.x];
    sM[threadIdx.x] = params[(size_t)n * FPB_CROP_STRIDE + 13 + threadIdx.x];
  }
  if (threadIdx.x < 4) sR[threadIdx.x] = params[(size_t)n * FPB_CROP_STRIDE + 29 + threadIdx.x];
  __syncthreads();
  if (v >= V) return;
  const float x = pos[v * 3], y = pos[v * 3 + 1], z = pos[v * 3 + 2];
  const float4 cl = vertex_clip(x, y, z, sM, sR);
  if (!LEAN) {
    clip[(size_t)n * V + v] = cl;
    camd[(size_t)n * V + v] = vertex_camd(x, y, z, vnormals[v * 3], vnormals[v * 3 + 1], vnormals[v * 3 + 2], sT);
  }
  // window position in 1/16 pixel units (8*S sub-pixels per clip unit), snapped once per vertex
  int2 sn = make_int2(SNAP_INVALID, 0);
  if (cl.w > 0.0f) {
    const float sub = (float)(8 * S);
    const float sx = (cl.x / cl.w) * sub + sub;
    const float sy = (cl.y / cl.w) * sub + sub;
    if (fabsf(sx) < 1073741824.0f && fabsf(sy) < 1073741824.0f) sn = make_int2((int)rintf(sx), (int)rintf(sy));
  }
  snap[(size_t)n * V + v] = sn;
}

struct TriSetup {
  int X[3], Y[3];     // |X|, |Y| < 2^30
  int px0, px1, py0, py1;
  int ok;
  int small;          // every |X|, |Y| <= 16000 (within ~1000 pixels of the crop): |E| <= 2 * 32000 * 32016 < 2^31
};

__device__ __forceinline__ int floordiv16(int a) { return a >> 4; }
__device__ __forceinline__ int ceildiv16(int a) { return -((-a) >> 4); }

__device__ __forceinline__ void tri_setup(const int2 s0, const int2 s1, const int2 s2, int S, TriSetup& t)
{
  t.ok = 0;
  t.small = 0;
  if (s0.x == SNAP_INVALID || s1.x == SNAP_INVALID || s2.x == SNAP_INVALID) return;
  t.X[0] = s0.x; t.Y[0] = s0.y; t.X[1] = s1.x; t.Y[1] = s1.y; t.X[2] = s2.x; t.Y[2] = s2.y;
  const long long area = (long long)(t.X[1] - t.X[0]) * (t.Y[2] - t.Y[0]) - (long long)(t.Y[1] - t.Y[0]) * (t.X[2] - t.X[0]);
  if (area == 0) return;
  if (area < 0) {  // orient for the coverage test only
    int tx = t.X[1], ty = t.Y[1];
    t.X[1] = t.X[2]; t.Y[1] = t.Y[2]; t.X[2] = tx; t.Y[2] = ty;
  }
  const int xmin = min(t.X[0], min(t.X[1], t.X[2])), xmax = max(t.X[0], max(t.X[1], t.X[2]));
  const int ymin = min(t.Y[0], min(t.Y[1], t.Y[2])), ymax = max(t.Y[0], max(t.Y[1], t.Y[2]));
  const int px0 = max(0, ceildiv16(xmin - 8)), px1 = min(S - 1, floordiv16(xmax - 8));
  const int py0 = max(0, ceildiv16(ymin - 8)), py1 = min(S - 1, floordiv16(ymax - 8));
  if (px0 > px1 || py0 > py1) return;
  t.px0 = px0; t.px1 = px1; t.py0 = py0; t.py1 = py1;
  t.ok = 1;
  t.small = xmin >= -16000 && xmax <= 16000 && ymin >= -16000 && ymax <= 16000;
}

__device__ __forceinline__ bool covers(const TriSetup& t, int px, int py)
{
  const long long cx = 16LL * px + 8, cy = 16LL * py + 8;
#pragma unroll
  for (int e = 0; e < 3; e++) {
    const int a = e, b = (e + 1) % 3;
    long long dx = (long long)t.X[b] - t.X[a], dy = (long long)t.Y[b] - t.Y[a];
    long long E = dx * (cy - t.Y[a]) - dy * (cx - t.X[a]);
    if (E < 0) return false;
    if (E == 0 && !(dy > 0 || (dy == 0 && dx < 0))) return false;
  }
  return true;
}

// the same test in 32-bit arithmetic for triangles whose vertices lie within ~1000 pixels of the crop (exact, hence
// identical results): |dx|, |dy|, |cx - X|, |cy - Y| < 2^15 -> |E| < 2^31
struct Tri32 { int X[3], Y[3]; };
__device__ __forceinline__ bool covers32(const Tri32& t, int px, int py)
{
  const int cx = 16 * px + 8, cy = 16 * py + 8;
#pragma unroll
  for (int e = 0; e < 3; e++) {
    const int a = e, b = (e + 1) % 3;
    const int dx = t.X[b] - t.X[a], dy = t.Y[b] - t.Y[a];
    const int E = dx * (cy - t.Y[a]) - dy * (cx - t.X[a]);
    if (E < 0) return false;
    if (E == 0 && !(dy > 0 || (dy == 0 && dx < 0))) return false;
  }
  return true;
}

__device__ __forceinline__ void shade_fragment(const float4 c0, const float4 c1, const float4 c2, int px, int py, int S,
                                               uint32_t face, unsigned long long* zb)
{
  const float xs = 2.0f / (float)S, xo = 1.0f / (float)S - 1.0f;
  float fx = (float)px * xs + xo, fy = (float)py * xs + xo;
  float b0, b1, zw;
  bary_zw(c0, c1, c2, fx, fy, b0, b1, zw);
  if (zw != zw) return;
  unsigned long long key = ((unsigned long long)f2ord(zw) << 32) | face;
  atomicMin(zb + (size_t)py * S + px, key);
}

__device__ __forceinline__ void raster_pixel(const TriSetup& t, const float4 c0, const float4 c1, const float4 c2, int px,
                                             int py, int S, uint32_t face, unsigned long long* zb)
{
  if (!covers(t, px, py)) return;
  const float xs = 2.0f / (float)S, xo = 1.0f / (float)S - 1.0f;
  float fx = (float)px * xs + xo, fy = (float)py * xs + xo;
  float b0, b1, zw;
  bary_zw(c0, c1, c2, fx, fy, b0, b1, zw);
  if (zw != zw) return;
  unsigned long long key = ((unsigned long long)f2ord(zw) << 32) | face;
  atomicMin(zb + (size_t)py * S + px, key);
}

constexpr int SMALL_TRI_PIXELS = 16;
// Triangles whose pixel bounding box exceeds SMALL_TRI_PIXELS (low-poly / CAD-like meshes: the 12-face cube of BASELINE
// cfg 1 covers ~10^4 pixels per face) are not rasterised here: their face index goes to a per-hypothesis queue and the
// shading kernel resolves them TILE-PARALLEL -- the raster kernel queues the finished set-up record (snapped corners, pixel box,
// clip coordinates), every 32 x 8 pixel block of shade_kernel copies the records to shared memory, skips those whose
// box misses the tile, and each pixel thread z-tests the rest in registers
// against the z-buffer value the small-triangle path left (same integer edge functions, same (z/w | face) key: bit-
// identical images, no atomics).  The first version swept each big triangle's box with ONE warp, serially: 686 us per
// launch for 12 hypotheses x 12 faces.  BIG_CAP bounds the queue; overflow falls back to that warp sweep.
constexpr int BIG_CAP = 256;
constexpr int BIG_CHUNK = 64;
// one queued big triangle: the raster kernel's set-up + clip coordinates, 96 bytes (six 16-byte words)
struct __align__(16) BigTri {
  int X[3], Y[3];
  int px0, px1, py0, py1;
  int small;
  uint32_t face;
  float4 c0, c1, c2;
};
static_assert(sizeof(BigTri) == 96, "BigTri record");

// One triangle per lane; the whole warp must call this together (the large-footprint path is warp-cooperative).
// valid: this lane holds face f with corner indices i0..i2 and their snapped positions s0..s2.
__device__ __forceinline__ void raster_triangle(bool valid, int f, int i0, int i1, int i2, const int2 s0, const int2 s1,
                                                const int2 s2, const ClipSrc cl, int S,
                                                unsigned long long* __restrict__ zb, uint32_t* __restrict__ bigcnt,
                                                BigTri* __restrict__ bigq)
{
  TriSetup t;
  t.ok = 0;
  if (valid) tri_setup(s0, s1, s2, S, t);
  if (valid && (s0.x == SNAP_INVALID || s1.x == SNAP_INVALID || s2.x == SNAP_INVALID)) {
    // a corner is behind the camera plane or outside the fixed-point range: if any corner is in front, the triangle may
    // still cross the crop -> queue it for the homogeneous (clip-free) test over the whole crop (small = 2)
    const float4 c0 = cl.get(i0), c1 = cl.get(i1), c2 = cl.get(i2);
    if (c0.w > 0.0f || c1.w > 0.0f || c2.w > 0.0f) {
      const uint32_t slot = atomicAdd(bigcnt, 1u) + 1u;
      if (slot < (uint32_t)BIG_CAP) {                 // (a full queue drops it: it has no integer set-up for the sweep)
        BigTri r;
#pragma unroll
        for (int e = 0; e < 3; e++) { r.X[e] = 0; r.Y[e] = 0; }
        r.px0 = 0; r.px1 = S - 1; r.py0 = 0; r.py1 = S - 1; r.small = 2; r.face = (uint32_t)f;
        r.c0 = c0; r.c1 = c1; r.c2 = c2;
        const uint4* src = reinterpret_cast<const uint4*>(&r);
        uint4* dst = reinterpret_cast<uint4*>(bigq + slot);
#pragma unroll
        for (int e = 0; e < 6; e++) dst[e] = src[e];
      }
    }
  }
  int big = 0;
  if (t.ok) {
    int npx = (t.px1 - t.px0 + 1) * (t.py1 - t.py0 + 1);
    if (npx <= SMALL_TRI_PIXELS && t.small) {
      // incremental 32-bit edge functions (exact integers, so identical to covers32): E_e at the first pixel centre,
      // biased by the top-left rule so that "covered" is E_e >= 0 for all three edges, i.e. the sign bit of their OR
      int E[3], ex[3], ey[3];
#pragma unroll
      for (int e = 0; e < 3; e++) {
        const int a = e, b = (e + 1) % 3;
        const int dx = t.X[b] - t.X[a], dy = t.Y[b] - t.Y[a];
        const int cx = 16 * t.px0 + 8, cy = 16 * t.py0 + 8;
        const int tl = (dy > 0 || (dy == 0 && dx < 0)) ? 0 : 1;     // edges that do not own their boundary need E > 0
        E[e] = dx * (cy - t.Y[a]) - dy * (cx - t.X[a]) - tl;
        ex[e] = -16 * dy;                                            // one pixel to the right
        ey[e] = 16 * dx;                                             // one pixel down
      }
      // pass 1 (cheap, lanes in lockstep): coverage bits of the <= 16 pixels of the bounding box, row-major
      const int bw = t.px1 - t.px0 + 1;
      uint32_t mask = 0, bit = 1;
      for (int py = t.py0; py <= t.py1; py++) {
        int r0 = E[0], r1 = E[1], r2 = E[2];
        for (int px = t.px0; px <= t.px1; px++) {
          if ((r0 | r1 | r2) >= 0) mask |= bit;
          bit <<= 1;
          r0 += ex[0]; r1 += ex[1]; r2 += ex[2];
        }
        E[0] += ey[0]; E[1] += ey[1]; E[2] += ey[2];
      }
      // pass 2: one fragment per iteration and lane, so the trip count is the largest number of COVERED pixels in the
      // warp (1-3 for pixel-sized triangles) instead of the largest bounding box; the float clip coordinates are
      // fetched only by triangles that cover a pixel centre
      if (mask) {
        const float4 c0 = cl.get(i0), c1 = cl.get(i1), c2 = cl.get(i2);
        const uint32_t mdiv = (256u + (uint32_t)bw - 1u) / (uint32_t)bw;     // b / bw == (b * mdiv) >> 8 for b < 16, bw <= 16
        do {
          const uint32_t b = (uint32_t)__ffs(mask) - 1u;
          mask &= mask - 1u;
          const uint32_t dy = (b * mdiv) >> 8, dx = b - dy * (uint32_t)bw;
          shade_fragment(c0, c1, c2, t.px0 + (int)dx, t.py0 + (int)dy, S, (uint32_t)f, zb);
        } while (mask);
      }
    } else {
      // large footprint, or coordinates beyond the 32-bit fast path: queue it for the tile-parallel resolve in the shading
      // kernel (the counter starts at 0xffffffff: one memset initialises it together with the z-buffer)
      const uint32_t slot = atomicAdd(bigcnt, 1u) + 1u;
      if (slot < (uint32_t)BIG_CAP) {
        BigTri r;
#pragma unroll
        for (int e = 0; e < 3; e++) { r.X[e] = t.X[e]; r.Y[e] = t.Y[e]; }
        r.px0 = t.px0; r.px1 = t.px1; r.py0 = t.py0; r.py1 = t.py1; r.small = t.small; r.face = (uint32_t)f;
        r.c0 = cl.get(i0); r.c1 = cl.get(i1); r.c2 = cl.get(i2);
        const uint4* src = reinterpret_cast<const uint4*>(&r);
        uint4* dst = reinterpret_cast<uint4*>(bigq + slot);
#pragma unroll
        for (int e = 0; e < 6; e++) dst[e] = src[e];
      } else {
        big = 1;     // queue full: warp-cooperative 64-bit sweep below
      }
    }
  }
  // triangles with a large footprint: the whole warp sweeps the bounding box together
  unsigned m = __ballot_sync(0xffffffffu, big);
  const int lane = threadIdx.x & 31;
  while (m) {
    const int src = __ffs(m) - 1;
    m &= m - 1;
    TriSetup s;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      s.X[k] = __shfl_sync(0xffffffffu, t.X[k], src);
      s.Y[k] = __shfl_sync(0xffffffffu, t.Y[k], src);
    }
    s.px0 = __shfl_sync(0xffffffffu, t.px0, src); s.px1 = __shfl_sync(0xffffffffu, t.px1, src);
    s.py0 = __shfl_sync(0xffffffffu, t.py0, src); s.py1 = __shfl_sync(0xffffffffu, t.py1, src);
    const float4 d0 = cl.get(__shfl_sync(0xffffffffu, i0, src));    // same address in every lane: one broadcast load
    const float4 d1 = cl.get(__shfl_sync(0xffffffffu, i1, src));
    const float4 d2 = cl.get(__shfl_sync(0xffffffffu, i2, src));
    const uint32_t face = (uint32_t)__shfl_sync(0xffffffffu, f, src);     // lanes hold arbitrary faces (compacted queue)
    const int bw = s.px1 - s.px0 + 1, tot = bw * (s.py1 - s.py0 + 1);
    for (int i = lane; i < tot; i += 32) {
      int py = s.py0 + i / bw, px = s.px0 + i % bw;
      raster_pixel(s, d0, d1, d2, px, py, S, face, zb);
    }
  }
}

// Two phases per block of RASTER_FACES = 1024 consecutive faces (4 per thread):
//   1. cull (all lanes busy): all corner indices are fetched first, then all snapped positions (12 independent gathers
//      in flight per thread), then the integer set-up decides whether the triangle's pixel-centre bounding box inside
//      the crop is non-empty.  Survivors are appended to a shared-memory queue (warp-aggregated atomics).
//   2. rasterise (dense): the queue is walked with one surviving triangle per lane.
// ncu on cfg 4 (1008 hypotheses x 500 k faces, ~40 triangles per pixel) showed the one-phase kernel issue-bound at
// 16.6 active threads per instruction: 45 % of its instructions ran in the coverage / fragment code with 1-3 active
// lanes, because a warp executes that path as soon as ONE of its 32 triangles survives the cull.  (Staging the positions
// in shared memory by bulk async copy -- whole array, per-block vertex range, with or without overlap with the index
// loads -- and a persistent variant with next-chunk prefetch all measured equal or slower: latency and L2 sector
// traffic are not what bounds this kernel.)
constexpr int RASTER_TPT = 4;
constexpr int RASTER_FACES = 256 * RASTER_TPT;

__global__ void __launch_bounds__(256, 3) raster_cull_kernel(const int32_t* __restrict__ faces, int F, int V,
                                                          const float4* __restrict__ clip, const int2* __restrict__ snap,
                                                          const float* __restrict__ pos, const float* __restrict__ params,
                                                          int S, unsigned long long* __restrict__ zbuf,
                                                          uint32_t* __restrict__ bigcnt, BigTri* __restrict__ bigq)
{
  __shared__ int q_f[RASTER_FACES];
  __shared__ int q_i[3][RASTER_FACES];
  __shared__ int2 q_s[3][RASTER_FACES];
  __shared__ int q_count;
  const int n = blockIdx.y;
  const int f0 = blockIdx.x * RASTER_FACES + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const int2* sn = snap + (size_t)n * V;
  if (threadIdx.x == 0) q_count = 0;
  int idx[RASTER_TPT][3];
#pragma unroll
  for (int k = 0; k < RASTER_TPT; k++) {
    const int f = f0 + k * 256;
#pragma unroll
    for (int c = 0; c < 3; c++) idx[k][c] = f < F ? faces[f * 3 + c] : 0;
  }
  int2 sp[RASTER_TPT][3];
#pragma unroll
  for (int k = 0; k < RASTER_TPT; k++)
#pragma unroll
    for (int c = 0; c < 3; c++) sp[k][c] = sn[idx[k][c]];
  __syncthreads();                                   // q_count = 0 is visible
#pragma unroll
  for (int k = 0; k < RASTER_TPT; k++) {
    const int f = f0 + k * 256;
    TriSetup t;
    t.ok = 0;
    if (f < F) tri_setup(sp[k][0], sp[k][1], sp[k][2], S, t);
    // survivors: a non-empty pixel box, or a corner behind the camera plane (handled by raster_triangle's clip-free path)
    const bool keep = t.ok || (f < F && (sp[k][0].x == SNAP_INVALID || sp[k][1].x == SNAP_INVALID || sp[k][2].x == SNAP_INVALID));
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (m) {
      int base = 0;
      if (lane == 0) base = atomicAdd(&q_count, __popc(m));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (keep) {
        const int e = base + __popc(m & ((1u << lane) - 1u));
        q_f[e] = f;
#pragma unroll
        for (int c = 0; c < 3; c++) { q_i[c][e] = idx[k][c]; q_s[c][e] = sp[k][c]; }
      }
    }
  }
  __syncthreads();
  const int count = q_count;
  const ClipSrc cl = {clip ? clip + (size_t)n * V : nullptr, pos, params + (size_t)n * FPB_CROP_STRIDE + 13,
                      params + (size_t)n * FPB_CROP_STRIDE + 29};
  unsigned long long* zb = zbuf + (size_t)n * S * S;
  for (int base = 0; base < count; base += 256) {    // warp-uniform trip count
    const int e = base + (int)threadIdx.x;
    const bool valid = e < count;
    const int ee = valid ? e : 0;
    raster_triangle(valid, q_f[ee], q_i[0][ee], q_i[1][ee], q_i[2][ee], q_s[0][ee], q_s[1][ee], q_s[2][ee], cl, S, zb,
                    bigcnt + n, bigq + (size_t)n * BIG_CAP);
  }
}

// One-phase variant for meshes with few triangles per pixel (most triangles survive the cull, so the queue only adds
// overhead: cfg 2, 20 k faces on ~15 k covered pixels, 234 vs 245 us per pass): same loads, triangles rasterised in place.
__global__ void __launch_bounds__(256, 3) raster_kernel(const int32_t* __restrict__ faces, int F, int V,
                                                     const float4* __restrict__ clip, const int2* __restrict__ snap,
                                                     const float* __restrict__ pos, const float* __restrict__ params,
                                                     int S, unsigned long long* __restrict__ zbuf,
                                                     uint32_t* __restrict__ bigcnt, BigTri* __restrict__ bigq)
{
  const int n = blockIdx.y;
  const int f0 = blockIdx.x * RASTER_FACES + threadIdx.x;
  const int2* sn = snap + (size_t)n * V;
  int idx[RASTER_TPT][3];
#pragma unroll
  for (int k = 0; k < RASTER_TPT; k++) {
    const int f = f0 + k * 256;
#pragma unroll
    for (int c = 0; c < 3; c++) idx[k][c] = f < F ? faces[f * 3 + c] : 0;
  }
  int2 sp[RASTER_TPT][3];
#pragma unroll
  for (int k = 0; k < RASTER_TPT; k++)
#pragma unroll
    for (int c = 0; c < 3; c++) sp[k][c] = sn[idx[k][c]];
#pragma unroll
  for (int k = 0; k < RASTER_TPT; k++) {
    const int f = f0 + k * 256;
    const ClipSrc cl = {clip ? clip + (size_t)n * V : nullptr, pos, params + (size_t)n * FPB_CROP_STRIDE + 13,
                        params + (size_t)n * FPB_CROP_STRIDE + 29};
    raster_triangle(f < F, f, idx[k][0], idx[k][1], idx[k][2], sp[k][0], sp[k][1], sp[k][2], cl, S,
                    zbuf + (size_t)n * S * S, bigcnt + n, bigq + (size_t)n * BIG_CAP);
  }
}

// transform_batch's `rgb.cuda().float() / 255.0` (h5_dataset.py:123-124): on CUDA tensors ATen evaluates a division by
// a Python scalar as a multiplication with the fp32 reciprocal ("a * reciprocal(b)", BinaryDivTrueKernel.cu), and the
// reference only runs on CUDA, so x * (1.0f / 255.0f) IS the reference value (it differs from the correctly rounded
// quotient by one ulp on about a quarter of the inputs).  make_mesh_tensors' `/ 255.0` is the same torch op.
__device__ __forceinline__ float div255(float a) { return a * (1.0f / 255.0f); }

__device__ __forceinline__ void xyz_norm(const float xyz[3], const float* T, float inv_radius, float zthr, float o[3])
{
  const bool inv = xyz[2] < zthr;
  const float t[3] = {T[3], T[7], T[11]};
#pragma unroll
  for (int k = 0; k < 3; k++) {
    float x = (xyz[k] - t[k]) * inv_radius;
    o[k] = (inv || fabsf(x) >= 2.0f) ? 0.0f : x;
  }
}

// DT: element type of the network input buffer -- 0 fp32, 1 bf16, 2 fp16 (FPB_PREC_*)
template <int DT>
__device__ __forceinline__ void store_px8(void* base, size_t pix, const float v[6])
{
  if (DT == 1) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]),
                   c = __floats2bfloat162_rn(v[4], v[5]);
    uint4 q;
    q.x = *reinterpret_cast<uint32_t*>(&a); q.y = *reinterpret_cast<uint32_t*>(&b);
    q.z = *reinterpret_cast<uint32_t*>(&c); q.w = 0u;
    reinterpret_cast<uint4*>(base)[pix] = q;
  } else if (DT == 2) {
    __half2 a = __floats2half2_rn(v[0], v[1]), b = __floats2half2_rn(v[2], v[3]), c = __floats2half2_rn(v[4], v[5]);
    uint4 q;
    q.x = *reinterpret_cast<uint32_t*>(&a); q.y = *reinterpret_cast<uint32_t*>(&b);
    q.z = *reinterpret_cast<uint32_t*>(&c); q.w = 0u;
    reinterpret_cast<uint4*>(base)[pix] = q;
  } else {
    float4* p = reinterpret_cast<float4*>(base) + pix * 2;
    p[0] = make_float4(v[0], v[1], v[2], v[3]);
    p[1] = make_float4(v[4], v[5], 0.0f, 0.0f);
  }
}

// Tile-parallel resolve of the queued big triangles of one hypothesis for the 32 x 8 pixel block of the caller: returns
// min(key, keys of the queued triangles covering raster pixel (px, py)).  Block-cooperative (every thread of the block
// must call it; `inb` = this thread owns a pixel inside the crop).  rx0..ry1: the block's raster-space pixel rectangle.
__device__ __noinline__ unsigned long long resolve_big_triangles(unsigned long long key, bool inb, int px, int py, int rx0,
                                                                 int rx1, int ry0, int ry1, int S, int nbig,
                                                                 const BigTri* __restrict__ bigq, BigTri* sh)
{
  const int tid = threadIdx.y * 32 + threadIdx.x;
  const float xs = 2.0f / (float)S, xo = 1.0f / (float)S - 1.0f;
  const float fx = (float)px * xs + xo, fy = (float)py * xs + xo;
  for (int base = 0; base < nbig; base += BIG_CHUNK) {
    const int cnt = min(BIG_CHUNK, nbig - base);
    __syncthreads();                                   // the previous chunk has been consumed
    for (int w = tid; w < cnt * 6; w += 256)
      reinterpret_cast<uint4*>(sh)[w] = __ldg(reinterpret_cast<const uint4*>(bigq + base) + w);
    __syncthreads();
    if (inb) {
      for (int k = 0; k < cnt; k++) {
        const BigTri& b = sh[k];
        if (b.px1 < rx0 || b.px0 > rx1 || b.py1 < ry0 || b.py0 > ry1) continue;     // misses this tile (block-uniform)
        if (px < b.px0 || px > b.px1 || py < b.py0 || py > b.py1) continue;
        if (b.small == 2) {                              // clip-free homogeneous test (corner behind the camera plane)
          float zh;
          if (homog_cover(b.c0, b.c1, b.c2, fx, fy, zh)) {
            const unsigned long long kk = ((unsigned long long)f2ord(zh) << 32) | b.face;
            key = kk < key ? kk : key;
          }
          continue;
        }
        bool in;
        if (b.small) {
          Tri32 q;
#pragma unroll
          for (int e = 0; e < 3; e++) { q.X[e] = b.X[e]; q.Y[e] = b.Y[e]; }
          in = covers32(q, px, py);
        } else {
          TriSetup t;
#pragma unroll
          for (int e = 0; e < 3; e++) { t.X[e] = b.X[e]; t.Y[e] = b.Y[e]; }
          in = covers(t, px, py);
        }
        if (!in) continue;
        float b0, b1, zw;
        bary_zw(b.c0, b.c1, b.c2, fx, fy, b0, b1, zw);
        if (zw != zw) continue;
        const unsigned long long kk = ((unsigned long long)f2ord(zw) << 32) | b.face;
        key = kk < key ? kk : key;
      }
    }
  }
  return key;
}

// use_light / w_ambient / w_diffuse of nvdiffrast_render (src/Utils.py:200-211, default head light) and the optional
// normal_map output (:192-197): interpolated camera-frame vertex normals, F.normalize'd, background 0
struct ShadeOpts {
  int use_light; float w_ambient, w_diffuse; float* out_normal;
  const uint32_t* bigcnt; const BigTri* bigq;      // queue of big triangles (written by the raster kernels)
};

// FULL: the off-hot-path arguments of nvdiffrast_render (normal_map output, non-default light) are compiled in; the hot
// path instantiates FULL = false so that they do not cost it registers (62 vs 40 per thread: half the occupancy of a
// latency-bound kernel)
template <int BF16, bool LEAN, bool FULL>
__global__ void __launch_bounds__(256, (FULL || LEAN) ? 1 : 6) shade_kernel(fpb_mesh mesh, const float4* __restrict__ clip,
                                                    const float4* __restrict__ camd,
                                                    const unsigned long long* __restrict__ zbuf,
                                                    const float* __restrict__ poses, const float* __restrict__ params,
                                                    int N, int S, float inv_radius,
                                                    float zthr, void* __restrict__ outA, int out_pitch, int out_xoff,
                                                    float* __restrict__ out_color,
                                                    float* __restrict__ out_xyz, int32_t* __restrict__ out_tri,
                                                    ShadeOpts opt)
{
  // block = 32 x 8 pixels of one crop: grid (ceil(S/32), ceil(S/8), N)
  const int n = blockIdx.z;
  const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y;
  const bool inb = j < S && i < S;
  const int py = S - 1 - i, px = j;  // row flip (src/Utils.py:197,216-218)
  unsigned long long key = inb ? zbuf[(size_t)n * S * S + (size_t)py * S + px] : ~0ull;
  // big triangles queued by the raster kernel (none for dense meshes: one load, no barrier)
  const uint32_t stored = opt.bigcnt[n] + 1u;          // counter starts at 0xffffffff
  if (stored != 0u) {
    __shared__ BigTri sh_big[BIG_CHUNK];
    const int nbig = (int)(stored < (uint32_t)BIG_CAP ? stored : (uint32_t)BIG_CAP);
    const int rx0 = blockIdx.x * 32, rx1 = rx0 + 31, ry1 = S - 1 - (int)blockIdx.y * 8, ry0 = ry1 - 7;
    key = resolve_big_triangles(key, inb, px, py, rx0, rx1, ry0, ry1, S, nbig, opt.bigq + (size_t)n * BIG_CAP, sh_big);
  }
  if (!inb) return;
  const size_t idx = ((size_t)n * S + i) * S + j;
  float col[3] = {0.f, 0.f, 0.f}, xyz[3] = {0.f, 0.f, 0.f};
  int tri = 0;
  if (key == ~0ull && zthr > 0.0f) {      // background: colour 0, xyz 0 -> z below the validity threshold -> the whole pixel is zero
    if (out_color) { out_color[idx * 3] = 0.f; out_color[idx * 3 + 1] = 0.f; out_color[idx * 3 + 2] = 0.f; }
    if (out_xyz) { out_xyz[idx * 3] = 0.f; out_xyz[idx * 3 + 1] = 0.f; out_xyz[idx * 3 + 2] = 0.f; }
    if (out_tri) out_tri[idx] = 0;
    if (FULL && opt.out_normal) { opt.out_normal[idx * 3] = 0.f; opt.out_normal[idx * 3 + 1] = 0.f; opt.out_normal[idx * 3 + 2] = 0.f; }
    const float z6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    store_px8<BF16>(outA, ((size_t)n * S + i) * out_pitch + out_xoff + j, z6);
    return;
  }
  if (key != ~0ull) {
    const int f = (int)(uint32_t)(key & 0xffffffffu);
    const int i0 = mesh.faces[f * 3], i1 = mesh.faces[f * 3 + 1], i2 = mesh.faces[f * 3 + 2];
    const float xs = 2.0f / (float)S, xo = 1.0f / (float)S - 1.0f;
    float fx = (float)px * xs + xo, fy = (float)py * xs + xo;
    float b0, b1, zw;
    float4 q0, q1, q2;
    if (!LEAN) {
      const float4* cl = clip + (size_t)n * mesh.V;
      const float4* cd = camd + (size_t)n * mesh.V;
      bary_zw(cl[i0], cl[i1], cl[i2], fx, fy, b0, b1, zw);
      q0 = cd[i0]; q1 = cd[i1]; q2 = cd[i2];
    } else {
      // lean mode: the three corners of the visible triangle are recomputed (same instruction sequence as the vertex pass)
      const float* T = poses + (size_t)n * 16;
      const float* M = params + (size_t)n * FPB_CROP_STRIDE + 13;
      const float* R = params + (size_t)n * FPB_CROP_STRIDE + 29;
      const int iv[3] = {i0, i1, i2};
      float4 cc[3], qq[3];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const float x = mesh.pos[iv[k] * 3], y = mesh.pos[iv[k] * 3 + 1], z = mesh.pos[iv[k] * 3 + 2];
        cc[k] = vertex_clip(x, y, z, M, R);
        qq[k] = vertex_camd(x, y, z, mesh.vnormals[iv[k] * 3], mesh.vnormals[iv[k] * 3 + 1], mesh.vnormals[iv[k] * 3 + 2], T);
      }
      bary_zw(cc[0], cc[1], cc[2], fx, fy, b0, b1, zw);
      q0 = qq[0]; q1 = qq[1]; q2 = qq[2];
    }
    const float b2 = (1.0f - b0) - b1;
    xyz[0] = (b0 * q0.x + b1 * q1.x) + b2 * q2.x;
    xyz[1] = (b0 * q0.y + b1 * q1.y) + b2 * q2.y;
    xyz[2] = (b0 * q0.z + b1 * q1.z) + b2 * q2.z;
    const float dm = (b0 * q0.w + b1 * q1.w) + b2 * q2.w;
    float c[3];
    if (mesh.tex) {
      const int t0 = mesh.uv_idx[f * 3], t1 = mesh.uv_idx[f * 3 + 1], t2 = mesh.uv_idx[f * 3 + 2];
      float tu = (b0 * mesh.uv[t0 * 2] + b1 * mesh.uv[t1 * 2]) + b2 * mesh.uv[t2 * 2];
      float tv = (b0 * mesh.uv[t0 * 2 + 1] + b1 * mesh.uv[t1 * 2 + 1]) + b2 * mesh.uv[t2 * 2 + 1];
      float u = tu * (float)mesh.Wt - 0.5f, v = tv * (float)mesh.Ht - 0.5f;
      float uf = floorf(u), vf = floorf(v);
      float du = u - uf, dv = v - vf;
      int iu0 = (int)uf, iv0 = (int)vf, iu1 = iu0 + 1, iv1 = iv0 + 1;
      const int Wt = mesh.Wt, Ht = mesh.Ht;
      iu0 = ((iu0 % Wt) + Wt) % Wt; iu1 = ((iu1 % Wt) + Wt) % Wt;
      iv0 = ((iv0 % Ht) + Ht) % Ht; iv1 = ((iv1 % Ht) + Ht) % Ht;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        float a00 = mesh.tex[(iv0 * Wt + iu0) * 3 + k], a10 = mesh.tex[(iv0 * Wt + iu1) * 3 + k];
        float a01 = mesh.tex[(iv1 * Wt + iu0) * 3 + k], a11 = mesh.tex[(iv1 * Wt + iu1) * 3 + k];
        float top = a00 + du * (a10 - a00), bot = a01 + du * (a11 - a01);
        c[k] = top + dv * (bot - top);
      }
    } else {
      const float4 k0 = mesh.vcolor4[i0], k1 = mesh.vcolor4[i1], k2 = mesh.vcolor4[i2];
      c[0] = (b0 * k0.x + b1 * k1.x) + b2 * k2.x;
      c[1] = (b0 * k0.y + b1 * k1.y) + b2 * k2.y;
      c[2] = (b0 * k0.z + b1 * k1.z) + b2 * k2.z;
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
      float cv = FULL ? (opt.use_light ? c[k] * opt.w_ambient + (dm * c[k]) * opt.w_diffuse : c[k]) : c[k] * 0.8f + (dm * c[k]) * 0.5f;
      col[k] = fminf(fmaxf(cv, 0.0f), 1.0f);
    }
    tri = f + 1;
    if (FULL && opt.out_normal) {     // off the hot path: the rotated corner normals are recomputed here
      const float* T = poses + (size_t)n * 16;
      const int iv[3] = {i0, i1, i2};
      float cn[3][3];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const float nx = mesh.vnormals[iv[k] * 3], ny = mesh.vnormals[iv[k] * 3 + 1], nz = mesh.vnormals[iv[k] * 3 + 2];
#pragma unroll
        for (int jj = 0; jj < 3; jj++) cn[k][jj] = (T[jj * 4] * nx + T[jj * 4 + 1] * ny) + T[jj * 4 + 2] * nz;
      }
      float nm[3];
#pragma unroll
      for (int k = 0; k < 3; k++) nm[k] = (b0 * cn[0][k] + b1 * cn[1][k]) + b2 * cn[2][k];
      const float dn = fmaxf(sqrtf((nm[0] * nm[0] + nm[1] * nm[1]) + nm[2] * nm[2]), 1e-12f);
#pragma unroll
      for (int k = 0; k < 3; k++) opt.out_normal[idx * 3 + k] = nm[k] / dn;
    }
  } else if (FULL && opt.out_normal) {
    opt.out_normal[idx * 3] = 0.f; opt.out_normal[idx * 3 + 1] = 0.f; opt.out_normal[idx * 3 + 2] = 0.f;
  }
  if (out_color) { out_color[idx * 3] = col[0]; out_color[idx * 3 + 1] = col[1]; out_color[idx * 3 + 2] = col[2]; }
  if (out_xyz) { out_xyz[idx * 3] = xyz[0]; out_xyz[idx * 3 + 1] = xyz[1]; out_xyz[idx * 3 + 2] = xyz[2]; }
  if (out_tri) out_tri[idx] = tri;
  float v6[6];
#pragma unroll
  for (int k = 0; k < 3; k++) v6[k] = div255(col[k] * 255.0f);
  xyz_norm(xyz, poses + (size_t)n * 16, inv_radius, zthr, v6 + 3);
  store_px8<BF16>(outA, ((size_t)n * S + i) * out_pitch + out_xoff + j, v6);
}

__device__ __forceinline__ int nearest_src(float p, float scale, int size)
{
  float r = nearbyintf(p * scale - 0.5f);
  if (!(r >= 0.0f && r <= (float)(size - 1))) return -1;
  return (int)r;
}

struct K9f { float k[9]; };

template <int BF16, bool U8>
__global__ void __launch_bounds__(256) warp_kernel(const void* __restrict__ rgb_, const float* __restrict__ xyz_map,
                                                   const float* __restrict__ depth, int H, int W,
                                                   const float* __restrict__ poses, const float* __restrict__ params, K9f K,
                                                   int N, int S, float inv_radius, int mode, float wsc, float hsc,
                                                   float ssc, void* __restrict__ outB, int out_pitch, int out_xoff)
{
  // block = 32 x 8 pixels of one crop: grid (ceil(S/32), ceil(S/8), N); wsc = W/(W-1), hsc = H/(H-1), ssc = S/(S-1)
  const int n = blockIdx.z;
  const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y;
  if (j >= S || i >= S) return;
  const float* pr = params + (size_t)n * FPB_CROP_STRIDE;
  const float i00 = pr[33], i02 = pr[34], i11 = pr[35], i12 = pr[36];
  const float zthr = mode ? 0.1f : 0.001f;
  const float px = (float)j * i00 + i02, py = (float)i * i11 + i12;
  float v6[6];
  {
    const float ix = px * wsc - 0.5f, iy = py * hsc - 0.5f;
    const float x0 = floorf(ix), y0 = floorf(iy), x1 = x0 + 1.0f, y1 = y0 + 1.0f;
    const float wnw = (x1 - ix) * (y1 - iy), wne = (ix - x0) * (y1 - iy);
    const float wsw = (x1 - ix) * (iy - y0), wse = (ix - x0) * (iy - y0);
    const bool inx0 = x0 >= 0.0f && x0 <= (float)(W - 1), inx1 = x1 >= 0.0f && x1 <= (float)(W - 1);
    const bool iny0 = y0 >= 0.0f && y0 <= (float)(H - 1), iny1 = y1 >= 0.0f && y1 <= (float)(H - 1);
    // out-of-range taps are skipped (zero padding); their clamped offsets are never dereferenced
    const int xi0 = inx0 ? (int)x0 : 0, xi1 = inx1 ? (int)x1 : 0, yi0 = iny0 ? (int)y0 : 0, yi1 = iny1 ? (int)y1 : 0;
    const int o00 = (yi0 * W + xi0) * 3, o01 = (yi0 * W + xi1) * 3, o10 = (yi1 * W + xi0) * 3, o11 = (yi1 * W + xi1) * 3;
    auto tap = [&](int o) -> float {
      // uint8 -> float on the FMA pipe: 2^23 + b is exact, so (2^23 | b) - 2^23 == (float)b
      if (U8) return __uint_as_float(0x4B000000u | (uint32_t)reinterpret_cast<const uint8_t*>(rgb_)[o]) - 8388608.0f;
      return reinterpret_cast<const float*>(rgb_)[o];
    };
    const bool p00 = inx0 && iny0, p01 = inx1 && iny0, p10 = inx0 && iny1, p11 = inx1 && iny1;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      float acc = 0.0f;
      if (p00) acc += tap(o00 + k) * wnw;
      if (p01) acc += tap(o01 + k) * wne;
      if (p10) acc += tap(o10 + k) * wsw;
      if (p11) acc += tap(o11 + k) * wse;
      v6[k] = div255(acc);
    }
  }
  float xyz[3] = {0.f, 0.f, 0.f};
  const int u = nearest_src(px, wsc, W), v = nearest_src(py, hsc, H);
  if (u >= 0 && v >= 0) {
    if (mode == 0) {
      const float* s = xyz_map + ((size_t)v * W + u) * 3;
      xyz[0] = s[0]; xyz[1] = s[1]; xyz[2] = s[2];
    } else {
      const float qx = pr[0] * (float)u + pr[2], qy = pr[4] * (float)v + pr[5];
      const int j2 = nearest_src(qx, ssc, S), i2 = nearest_src(qy, ssc, S);
      if (j2 >= 0 && i2 >= 0) {
        const float px1 = (float)j2 * i00 + i02, py1 = (float)i2 * i11 + i12;
        const int u1 = nearest_src(px1, wsc, W), v1 = nearest_src(py1, hsc, H);
        if (u1 >= 0 && v1 >= 0) {
          const float z = depth[(size_t)v1 * W + u1];
          if (!(z < 0.001f)) {
            xyz[0] = ((float)u - K.k[2]) * z / K.k[0];
            xyz[1] = ((float)v - K.k[5]) * z / K.k[4];
            xyz[2] = z;
          }
        }
      }
    }
  }
  xyz_norm(xyz, poses + (size_t)n * 16, inv_radius, zthr, v6 + 3);
  store_px8<BF16>(outB, ((size_t)n * S + i) * out_pitch + out_xoff + j, v6);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
namespace {
__global__ void pack_color_kernel(const float* __restrict__ c3, int V, float4* __restrict__ c4)
{
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v < V) c4[v] = make_float4(c3[v * 3], c3[v * 3 + 1], c3[v * 3 + 2], 0.0f);
}
}  // namespace

template <typename T>
static int dev_copy(T** dst, const T* src, size_t count, cudaStream_t st)
{
  *dst = nullptr;
  if (!src || count == 0) return FPB_OK;
  FPB_CUDA(cudaMalloc((void**)dst, count * sizeof(T)));
  FPB_CUDA(cudaMemcpyAsync(*dst, src, count * sizeof(T), cudaMemcpyDeviceToDevice, st));
  return FPB_OK;
}

FPB_EXPORT int fpb_mesh_create(fpb_mesh** out, const float* pos, const float* vnormals, int V, const int32_t* faces, int F,
                               const float* vcolor, const float* uv, int Vt, const int32_t* uv_idx, const float* tex,
                               int Ht, int Wt, void* stream)
{
  FPB_CHECK_ARG(out && pos && vnormals && faces && V > 0 && F > 0);
  FPB_CHECK_ARG((vcolor != nullptr) != (tex != nullptr));
  FPB_CHECK_ARG(!tex || (uv && uv_idx && Vt > 0 && Ht > 0 && Wt > 0));
  cudaStream_t st = (cudaStream_t)stream;
  fpb_mesh* m = new fpb_mesh();
  m->V = V; m->F = F; m->Vt = Vt; m->Ht = Ht; m->Wt = Wt;
  int rc = FPB_OK;
  if ((rc = dev_copy(&m->pos, pos, (size_t)V * 3, st))) goto fail;
  if ((rc = dev_copy(&m->vnormals, vnormals, (size_t)V * 3, st))) goto fail;
  if ((rc = dev_copy(&m->faces, faces, (size_t)F * 3, st))) goto fail;
  if ((rc = dev_copy(&m->vcolor, vcolor, (size_t)V * 3, st))) goto fail;
  if (vcolor) {
    if (cudaMalloc((void**)&m->vcolor4, (size_t)V * sizeof(float4)) != cudaSuccess) {
      fpb_set_error("fpb_mesh_create: cudaMalloc of the packed vertex colours failed");
      rc = FPB_ERR_CUDA;
      goto fail;
    }
    pack_color_kernel<<<fpb_cdiv(V, 256), 256, 0, st>>>(m->vcolor, V, m->vcolor4);
  }
  if ((rc = dev_copy(&m->uv, uv, tex ? (size_t)Vt * 2 : 0, st))) goto fail;
  if ((rc = dev_copy(&m->uv_idx, uv_idx, tex ? (size_t)F * 3 : 0, st))) goto fail;
  if ((rc = dev_copy(&m->tex, tex, tex ? (size_t)Ht * Wt * 3 : 0, st))) goto fail;
  FPB_CUDA(cudaStreamSynchronize(st));
  *out = m;
  return FPB_OK;
fail:
  fpb_mesh_destroy(m);
  return rc;
}

FPB_EXPORT int fpb_mesh_destroy(fpb_mesh* m)
{
  if (!m) return FPB_OK;
  cudaFree(m->pos); cudaFree(m->vnormals); cudaFree(m->faces); cudaFree(m->vcolor); cudaFree(m->vcolor4);
  cudaFree(m->uv); cudaFree(m->uv_idx); cudaFree(m->tex);
  delete m;
  return FPB_OK;
}

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

FPB_EXPORT size_t fpb_render_ws_bytes(int N, int V, int S)
{
  if (N > 65535) N = 65535;      // larger batches are rendered in chunks on the same workspace
  return 2 * align256((size_t)N * V * sizeof(float4)) + align256((size_t)N * V * sizeof(int2)) +
         align256((size_t)N * S * S * sizeof(unsigned long long) + (size_t)N * sizeof(uint32_t)) +
         align256((size_t)N * BIG_CAP * sizeof(BigTri));
}

static int render_chunk(const fpb_mesh* mesh, const float* poses, const float* params, int N, int S, float mesh_diameter,
                        float zthr, void* ws, void* out_A, int a_is_bf16, int out_pitch, int out_xoff, float* out_color,
                        float* out_xyz, int32_t* out_tri, ShadeOpts opt, cudaStream_t st)
{
  const int V = mesh->V, F = mesh->F;
  char* w = (char*)ws;
  float4* clip = (float4*)w; w += align256((size_t)N * V * sizeof(float4));
  float4* camd = (float4*)w; w += align256((size_t)N * V * sizeof(float4));
  int2* snap = (int2*)w; w += align256((size_t)N * V * sizeof(int2));
  unsigned long long* zbuf = (unsigned long long*)w;
  uint32_t* bigcnt = (uint32_t*)(zbuf + (size_t)N * S * S);     // directly behind the z-buffer: one memset initialises both
  w += align256((size_t)N * S * S * sizeof(unsigned long long) + (size_t)N * sizeof(uint32_t));
  BigTri* bigq = (BigTri*)w;
  FPB_CUDA(cudaMemsetAsync(zbuf, 0xff, (size_t)N * S * S * sizeof(unsigned long long) + (size_t)N * sizeof(uint32_t), st));
  opt.bigcnt = bigcnt; opt.bigq = bigq;
  // lean mode: for large meshes the float clip / camera arrays (32 B x V x N, written once and read for a few thousand
  // visible corners) are not materialised; rasteriser and shader recompute the corners they need.  FPB_RENDER_LEAN:
  // 1 forces it (tests), 0 disables it; default: when those arrays would exceed 512 MB
  static int lean_env = -2;
  if (lean_env == -2) { const char* e = getenv("FPB_RENDER_LEAN"); lean_env = e ? atoi(e) : -1; }
  const bool lean = lean_env >= 0 ? lean_env != 0 : (size_t)N * V * 32 > ((size_t)512 << 20);
  if (lean) {
    vertex_kernel<true><<<dim3(fpb_cdiv(V, 256), N), 256, 0, st>>>(mesh->pos, mesh->vnormals, V, poses, params, S, clip, camd, snap);
    clip = nullptr;
    camd = nullptr;
  } else {
    vertex_kernel<false><<<dim3(fpb_cdiv(V, 256), N), 256, 0, st>>>(mesh->pos, mesh->vnormals, V, poses, params, S, clip, camd, snap);
  }
  FPB_LAUNCH_CHECK();
  // dense meshes (several triangles per crop pixel): cull first, rasterise the survivors densely
  static int cull_min = -1;
  if (cull_min < 0) { const char* e = getenv("FPB_RASTER_CULL_MIN"); cull_min = e ? atoi(e) : 4; }
  if ((long long)F >= (long long)cull_min * S * S)
    raster_cull_kernel<<<dim3(fpb_cdiv(F, RASTER_FACES), N), 256, 0, st>>>(mesh->faces, F, V, clip, snap, mesh->pos, params, S, zbuf, bigcnt, bigq);
  else
    raster_kernel<<<dim3(fpb_cdiv(F, RASTER_FACES), N), 256, 0, st>>>(mesh->faces, F, V, clip, snap, mesh->pos, params, S, zbuf, bigcnt, bigq);
  FPB_LAUNCH_CHECK();
  const float radius = mesh_diameter / 2.0f, inv_radius = 1.0f / radius;
  const dim3 blocks(fpb_cdiv(S, 32), fpb_cdiv(S, 8), N), threads(32, 8);
  const bool full = opt.out_normal || !opt.use_light || opt.w_ambient != 0.8f || opt.w_diffuse != 0.5f;
#define LAUNCH_SHADE(BF, LN)                                                                                          \
  do {                                                                                                                \
    if (full) shade_kernel<BF, LN, true><<<blocks, threads, 0, st>>>(*mesh, clip, camd, zbuf, poses, params, N, S, inv_radius, zthr, \
                                                                     out_A, out_pitch, out_xoff, out_color, out_xyz, out_tri, opt); \
    else shade_kernel<BF, LN, false><<<blocks, threads, 0, st>>>(*mesh, clip, camd, zbuf, poses, params, N, S, inv_radius, zthr, \
                                                                 out_A, out_pitch, out_xoff, out_color, out_xyz, out_tri, opt); \
  } while (0)
  if (a_is_bf16 == 1)      { if (lean) LAUNCH_SHADE(1, true); else LAUNCH_SHADE(1, false); }
  else if (a_is_bf16 == 2) { if (lean) LAUNCH_SHADE(2, true); else LAUNCH_SHADE(2, false); }
  else                     { if (lean) LAUNCH_SHADE(0, true); else LAUNCH_SHADE(0, false); }
#undef LAUNCH_SHADE
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

// grid.z carries the hypothesis index (<= 65535 per launch): larger batches run in chunks on the same workspace
// (the reference batches its renders in chunks of 512, predict_pose_refine.py:50-54, for any N)
constexpr int MAX_N_PER_LAUNCH = 65535;

FPB_EXPORT int fpb_render_crops_ex(const fpb_mesh* mesh, const float* poses, const float* params, int N, int S,
                                   float mesh_diameter, float zthr, void* ws, void* out_A, int a_is_bf16, int out_pitch,
                                   int out_xoff, float* out_color, float* out_xyz, int32_t* out_tri, int use_light,
                                   float w_ambient, float w_diffuse, float* out_normal, void* stream)
{
  FPB_CHECK_ARG(mesh && poses && params && ws && out_A && N > 0 && S > 1 && S <= 4096);
  FPB_CHECK_ARG(out_xoff >= 0 && out_pitch >= S + out_xoff && a_is_bf16 >= 0 && a_is_bf16 <= 2);
  const size_t esz = a_is_bf16 ? 2 : 4;
  for (int n0 = 0; n0 < N; n0 += MAX_N_PER_LAUNCH) {
    const int n = (N - n0) < MAX_N_PER_LAUNCH ? (N - n0) : MAX_N_PER_LAUNCH;
    const size_t px = (size_t)n0 * S * S;
    ShadeOpts opt = {use_light, w_ambient, w_diffuse, out_normal ? out_normal + px * 3 : nullptr, nullptr, nullptr};
    const int rc = render_chunk(mesh, poses + (size_t)n0 * 16, params + (size_t)n0 * FPB_CROP_STRIDE, n, S, mesh_diameter, zthr,
                                ws, (char*)out_A + (size_t)n0 * S * out_pitch * 8 * esz, a_is_bf16, out_pitch, out_xoff,
                                out_color ? out_color + px * 3 : nullptr, out_xyz ? out_xyz + px * 3 : nullptr,
                                out_tri ? out_tri + px : nullptr, opt, (cudaStream_t)stream);
    if (rc) return rc;
  }
  return FPB_OK;
}

FPB_EXPORT int fpb_render_crops(const fpb_mesh* mesh, const float* poses, const float* params, int N, int S,
                                float mesh_diameter, float zthr, void* ws, void* out_A, int a_is_bf16, int out_pitch,
                                int out_xoff, float* out_color, float* out_xyz, int32_t* out_tri, void* stream)
{
  return fpb_render_crops_ex(mesh, poses, params, N, S, mesh_diameter, zthr, ws, out_A, a_is_bf16, out_pitch, out_xoff,
                             out_color, out_xyz, out_tri, 1, 0.8f, 0.5f, nullptr, stream);
}

FPB_EXPORT int fpb_warp_crops(const void* rgb, int rgb_is_u8, const float* xyz_hw3, const float* depth_hw, int H, int W,
                              const float* poses, const float* params, const float* K9_h, int N, int S,
                              float mesh_diameter, int mode, void* out_B, int b_is_bf16, int out_pitch, int out_xoff,
                              void* stream)
{
  FPB_CHECK_ARG(rgb && poses && params && K9_h && out_B && N > 0 && S > 1 && H > 1 && W > 1);
  FPB_CHECK_ARG(out_xoff >= 0 && out_pitch >= S + out_xoff);
  FPB_CHECK_ARG(mode == 0 ? xyz_hw3 != nullptr : depth_hw != nullptr);
  cudaStream_t st = (cudaStream_t)stream;
  K9f K;
  for (int i = 0; i < 9; i++) K.k[i] = K9_h[i];
  const float radius = mesh_diameter / 2.0f, inv_radius = 1.0f / radius;
  const float wsc = (float)W / (float)(W - 1), hsc = (float)H / (float)(H - 1), ssc = (float)S / (float)(S - 1);
  const size_t esz = b_is_bf16 ? 2 : 4;
  for (int n0 = 0; n0 < N; n0 += MAX_N_PER_LAUNCH) {      // grid.z limit: chunks of <= 65535 hypotheses
    const int n = (N - n0) < MAX_N_PER_LAUNCH ? (N - n0) : MAX_N_PER_LAUNCH;
    const dim3 blocks(fpb_cdiv(S, 32), fpb_cdiv(S, 8), n), threads(32, 8);
    const float* po = poses + (size_t)n0 * 16;
    const float* pa = params + (size_t)n0 * FPB_CROP_STRIDE;
    void* ob = (char*)out_B + (size_t)n0 * S * out_pitch * 8 * esz;
#define LAUNCH_WARP(BF, U8)                                                                                           \
  warp_kernel<BF, U8><<<blocks, threads, 0, st>>>(rgb, xyz_hw3, depth_hw, H, W, po, pa, K, n, S, inv_radius, mode,   \
                                                  wsc, hsc, ssc, ob, out_pitch, out_xoff)
    if (b_is_bf16 == 1)      { if (rgb_is_u8) LAUNCH_WARP(1, true); else LAUNCH_WARP(1, false); }
    else if (b_is_bf16 == 2) { if (rgb_is_u8) LAUNCH_WARP(2, true); else LAUNCH_WARP(2, false); }
    else                     { if (rgb_is_u8) LAUNCH_WARP(0, true); else LAUNCH_WARP(0, false); }
#undef LAUNCH_WARP
    FPB_LAUNCH_CHECK();
  }
  return FPB_OK;
}
