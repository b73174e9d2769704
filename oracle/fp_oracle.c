/*
 * fp_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * Plain-C restatement of the per-pixel arithmetic on FoundationPose's
 * render-and-compare path.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library.
 *
 * PARITY UNPINNED: the reference ships no tests or golden vectors, and the
 * third-party code it calls on this path (nvdiffrast, kornia, warp-lang) is
 * not vendored, so every function below is a restatement of the published
 * algorithm anchored on the reference call sites that are cited per function.
 *
 * Build (see oracle/Makefile):
 *   gcc -O2 -fPIC -shared -fopenmp -ffp-contract=off -fno-fast-math ...
 * -ffp-contract=off matters: every fp32 op rounds separately, like the
 * reference's eager torch / Warp kernels, and like the CUDA side which is
 * compiled with -fmad=false for these stages.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------
 * erode_depth_kernel            reference: src/Utils.py:359-384
 * loop order: u (column) outer, v (row) inner, as in the Warp kernel.
 * ---------------------------------------------------------------------- */
ORC_API void orc_erode_depth(const float* depth, float* out, int H, int W, int radius,
                             float depth_diff_thres, float ratio_thres, float zfar)
{
#pragma omp parallel for schedule(static)
  for (int h = 0; h < H; h++) {
    for (int w = 0; w < W; w++) {
      float d_ori = depth[h * W + w];
      float bad_cnt = 0.0f, total = 0.0f;
      for (int u = w - radius; u <= w + radius; u++) {
        if (u < 0 || u >= W) continue;
        for (int v = h - radius; v <= h + radius; v++) {
          if (v < 0 || v >= H) continue;
          float cur = depth[v * W + u];
          total += 1.0f;
          if (cur < 0.001f || cur >= zfar || fabsf(cur - d_ori) > depth_diff_thres) bad_cnt += 1.0f;
        }
      }
      out[h * W + w] = (bad_cnt / total > ratio_thres) ? 0.0f : d_ori;
    }
  }
}

/* ------------------------------------------------------------------------
 * bilateral_filter_depth_kernel  reference: src/Utils.py:304-343
 * ---------------------------------------------------------------------- */
ORC_API void orc_bilateral_depth(const float* depth, float* out, int H, int W, int radius,
                                 float zfar, float sigmaD, float sigmaR)
{
#pragma omp parallel for schedule(static)
  for (int h = 0; h < H; h++) {
    for (int w = 0; w < W; w++) {
      float res = 0.0f;
      float mean_depth = 0.0f;
      int num_valid = 0;
      for (int u = w - radius; u <= w + radius; u++) {
        if (u < 0 || u >= W) continue;
        for (int v = h - radius; v <= h + radius; v++) {
          if (v < 0 || v >= H) continue;
          float cur = depth[v * W + u];
          if (cur >= 0.001f && cur < zfar) { num_valid++; mean_depth += cur; }
        }
      }
      if (num_valid > 0) {
        mean_depth /= (float)num_valid;
        float dc = depth[h * W + w];
        float sum_weight = 0.0f, sum = 0.0f;
        const float denD = 2.0f * sigmaD * sigmaD;
        const float denR = 2.0f * sigmaR * sigmaR;
        for (int u = w - radius; u <= w + radius; u++) {
          if (u < 0 || u >= W) continue;
          for (int v = h - radius; v <= h + radius; v++) {
            if (v < 0 || v >= H) continue;
            float cur = depth[v * W + u];
            if (cur >= 0.001f && cur < zfar && fabsf(cur - mean_depth) < 0.01f) {
              float a = -(float)((u - w) * (u - w) + (h - v) * (h - v)) / denD;
              float b = (dc - cur) * (dc - cur) / denR;
              float weight = expf(a - b);
              sum_weight += weight;
              sum += weight * cur;
            }
          }
        }
        if (sum_weight > 0.0f) res = sum / sum_weight;
      }
      out[h * W + w] = res;
    }
  }
}

/* ------------------------------------------------------------------------
 * depth2xyzmap (numpy flavour, float64 math)   reference: src/Utils.py:399-417
 * depth2xyzmap_batch (torch flavour, fp32)     reference: src/Utils.py:420-438
 * K9: row-major 3x3 in double.
 * ---------------------------------------------------------------------- */
ORC_API void orc_depth2xyz_f64(const float* depth, const double* K9, float* xyz, int H, int W)
{
  const double fx = K9[0], cx = K9[2], fy = K9[4], cy = K9[5];
#pragma omp parallel for schedule(static)
  for (int v = 0; v < H; v++)
    for (int u = 0; u < W; u++) {
      float z = depth[v * W + u];
      float* o = xyz + (size_t)(v * W + u) * 3;
      if (z < 0.001f) { o[0] = o[1] = o[2] = 0.0f; continue; }
      o[0] = (float)(((double)u - cx) * (double)z / fx);
      o[1] = (float)(((double)v - cy) * (double)z / fy);
      o[2] = z;
    }
}

ORC_API void orc_depth2xyz_f32(const float* depth, const float* K9, float zfar, float* xyz, int H, int W)
{
  const float fx = K9[0], cx = K9[2], fy = K9[4], cy = K9[5];
#pragma omp parallel for schedule(static)
  for (int v = 0; v < H; v++)
    for (int u = 0; u < W; u++) {
      float z = depth[v * W + u];
      float* o = xyz + (size_t)(v * W + u) * 3;
      if (z < 0.001f || z > zfar) { o[0] = o[1] = o[2] = 0.0f; continue; }
      o[0] = ((float)u - cx) * z / fx;
      o[1] = ((float)v - cy) * z / fy;
      o[2] = z;
    }
}

/* ------------------------------------------------------------------------
 * Crop-window parameters per hypothesis.
 *   compute_crop_window_tf_batch(method='box_3d')   src/Utils.py:577-621
 *   bbox2d_ori = tf_to_crops^-1 . {(0,0),(S-1,S-1)}  predict_pose_refine.py:44-45,
 *                                                    predict_score.py:74-75
 *   bbox zoom + projection composition               src/Utils.py:159-181
 * Closed-form inverse of the affine (scale+shift) crop transform replaces
 * torch.linalg.inv (LU, library) -- documented in DESIGN.md.
 * All fp32, one rounding per op, sums left to right.
 *   out_tf   [N,9]  tf_to_crop (row-major 3x3)
 *   out_bbox [N,4]  umin,vmin,umax,vmax
 *   out_mtx  [N,16] P @ (glcam_in_cvcam @ pose)
 *   out_rtf  [N,4]  tf00, tf11, tf30, tf31 of the render zoom
 *   out_inv  [N,4]  i00, i02, i11, i12 of tf_to_crop^-1
 * ---------------------------------------------------------------------- */
ORC_API void orc_crop_params(const float* poses, int N, const float* K9, float radius, int S, int H, int W,
                             const float* P16, float* out_tf, float* out_bbox, float* out_mtx,
                             float* out_rtf, float* out_inv)
{
  for (int n = 0; n < N; n++) {
    const float* T = poses + n * 16;
    const float tx = T[3], ty = T[7], tz = T[11];
    const float offs[5][3] = {{0, 0, 0}, {radius, 0, 0}, {-radius, 0, 0}, {0, radius, 0}, {0, -radius, 0}};
    float us[5], vs[5];
    for (int k = 0; k < 5; k++) {
      float X = tx + offs[k][0], Y = ty + offs[k][1], Z = tz + offs[k][2];
      float px = (K9[0] * X + K9[1] * Y) + K9[2] * Z;
      float py = (K9[3] * X + K9[4] * Y) + K9[5] * Z;
      float pz = (K9[6] * X + K9[7] * Y) + K9[8] * Z;
      us[k] = px / pz;
      vs[k] = py / pz;
    }
    float rad = 0.0f;
    for (int k = 0; k < 5; k++) {
      float du = fabsf(us[k] - us[0]), dv = fabsf(vs[k] - vs[0]);
      if (du > rad) rad = du;
      if (dv > rad) rad = dv;
    }
    float left = rintf(us[0] - rad), right = rintf(us[0] + rad);
    float top = rintf(vs[0] - rad), bottom = rintf(vs[0] + rad);
    float sx = (float)S / (right - left), sy = (float)S / (bottom - top);
    float ttx = sx * (-left), tty = sy * (-top);
    float* tf = out_tf + n * 9;
    tf[0] = sx; tf[1] = 0; tf[2] = ttx; tf[3] = 0; tf[4] = sy; tf[5] = tty; tf[6] = 0; tf[7] = 0; tf[8] = 1;
    float i00 = 1.0f / sx, i02 = -(ttx / sx), i11 = 1.0f / sy, i12 = -(tty / sy);
    out_inv[n * 4 + 0] = i00; out_inv[n * 4 + 1] = i02; out_inv[n * 4 + 2] = i11; out_inv[n * 4 + 3] = i12;
    float umin = i02, vmin = i12;
    float umax = i00 * (float)(S - 1) + i02, vmax = i11 * (float)(S - 1) + i12;
    float* bb = out_bbox + n * 4;
    bb[0] = umin; bb[1] = vmin; bb[2] = umax; bb[3] = vmax;
    float l = umin, t = (float)H - vmin, r = umax, b = (float)H - vmax;
    float* rt = out_rtf + n * 4;
    rt[0] = (float)W / (r - l);
    rt[1] = (float)H / (t - b);
    rt[2] = (((float)W - r) - l) / (r - l);
    rt[3] = (((float)H - t) - b) / (t - b);
    float G[16];
    for (int j = 0; j < 4; j++) { G[j] = T[j]; G[4 + j] = -T[4 + j]; G[8 + j] = -T[8 + j]; G[12 + j] = T[12 + j]; }
    float* M = out_mtx + n * 16;
    for (int i = 0; i < 4; i++)
      for (int j = 0; j < 4; j++)
        M[i * 4 + j] = ((P16[i * 4 + 0] * G[j] + P16[i * 4 + 1] * G[4 + j]) + P16[i * 4 + 2] * G[8 + j]) + P16[i * 4 + 3] * G[12 + j];
  }
}

/* ordered-uint mapping of a float so that unsigned compare == float compare */
static inline uint32_t f2ord(float f) { uint32_t u; memcpy(&u, &f, 4); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }

static inline int64_t floordiv16(int64_t a) { return a >> 4; }           /* arithmetic shift = floor */
static inline int64_t ceildiv16(int64_t a) { return -((-a) >> 4); }

/* pixel-centre barycentrics + z/w, nvdiffrast forward-shader formula */
static inline void bary_zw(const float* c0, const float* c1, const float* c2, float fx, float fy,
                           float* b0, float* b1, float* zw)
{
  float p0x = c0[0] - fx * c0[3], p0y = c0[1] - fy * c0[3];
  float p1x = c1[0] - fx * c1[3], p1y = c1[1] - fy * c1[3];
  float p2x = c2[0] - fx * c2[3], p2y = c2[1] - fy * c2[3];
  float a0 = p1x * p2y - p1y * p2x;
  float a1 = p2x * p0y - p2y * p0x;
  float a2 = p0x * p1y - p0y * p1x;
  float iw = 1.0f / ((a0 + a1) + a2);
  float z = (c0[2] * a0 + c1[2] * a1) + c2[2] * a2;
  float w = (c0[3] * a0 + c1[3] * a1) + c2[3] * a2;
  float B0 = a0 * iw, B1 = a1 * iw, ZW = z / w;
  *b0 = fminf(fmaxf(B0, 0.0f), 1.0f);
  *b1 = fminf(fmaxf(B1, 0.0f), 1.0f);
  *zw = fminf(fmaxf(ZW, -1.0f), 1.0f);
}

/* coverage + z/w of a triangle with a corner behind the camera plane: all edge functions share the sign of their sum,
 * interpolated w > 0, z/w inside the near / far planes (float arithmetic, no tie rule) */
static inline int homog_cover(const float* c0, const float* c1, const float* c2, float fx, float fy, float* zw)
{
  float p0x = c0[0] - fx * c0[3], p0y = c0[1] - fy * c0[3];
  float p1x = c1[0] - fx * c1[3], p1y = c1[1] - fy * c1[3];
  float p2x = c2[0] - fx * c2[3], p2y = c2[1] - fy * c2[3];
  float a0 = p1x * p2y - p1y * p2x;
  float a1 = p2x * p0y - p2y * p0x;
  float a2 = p0x * p1y - p0y * p1x;
  float sum = (a0 + a1) + a2;
  int pos = a0 >= 0.0f && a1 >= 0.0f && a2 >= 0.0f && sum > 0.0f;
  int neg = a0 <= 0.0f && a1 <= 0.0f && a2 <= 0.0f && sum < 0.0f;
  if (!(pos || neg)) return 0;
  float z = (c0[2] * a0 + c1[2] * a1) + c2[2] * a2;
  float w = (c0[3] * a0 + c1[3] * a1) + c2[3] * a2;
  if (!((w > 0.0f) == (sum > 0.0f)) || w == 0.0f) return 0;
  *zw = z / w;
  return *zw >= -1.0f && *zw <= 1.0f;
}

/* ------------------------------------------------------------------------
 * Batched mesh render at crop resolution: restatement of nvdiffrast_render
 * (src/Utils.py:133-219) = dr.rasterize + dr.interpolate (+ dr.texture) +
 * lighting + row flip.  Coverage: vertices snapped to 1/16 px, integer edge
 * functions, antisymmetric tie rule, no back-face culling; visibility: nearest
 * clamped z/w, ties to the lower face index (deterministic).
 *   pos,vnormals [V,3]; vcolor [V,3] or NULL; faces [F,3];
 *   uv [Vt,2] (v already flipped), uv_idx [F,3], tex [Ht,Wt,3] when textured.
 * Outputs (image row order): color [N,S,S,3], xyz [N,S,S,3], tri [N,S,S] (id+1).
 * ---------------------------------------------------------------------- */
ORC_API void orc_render(const float* poses, const float* mtx, const float* rtf, int N, int S,
                        const float* pos, const float* vnormals, const float* vcolor, int V,
                        const int32_t* faces, int F,
                        const float* uv, const int32_t* uv_idx, const float* tex, int Ht, int Wt,
                        float* out_color, float* out_xyz, int32_t* out_tri,
                        int use_light, float w_ambient, float w_diffuse, float* out_normal)
{
  /* use_light / w_ambient / w_diffuse: src/Utils.py:200-211 with the default head light (light_dir = (0,0,1));
   * out_normal (may be NULL): the normal_map output, src/Utils.py:192-197 (interpolated camera-frame vertex normals,
   * F.normalize'd, background 0). */
  const float xs = 2.0f / (float)S, xo = 1.0f / (float)S - 1.0f;
  const float sub = (float)(8 * S);
#pragma omp parallel for schedule(dynamic, 1)
  for (int n = 0; n < N; n++) {
    const float* T = poses + n * 16;
    const float* M = mtx + n * 16;
    const float* rt = rtf + n * 4;
    float* clip = (float*)malloc(sizeof(float) * 4 * V);
    float* cam = (float*)malloc(sizeof(float) * 3 * V);
    float* dif = (float*)malloc(sizeof(float) * V);
    float* cnv = (float*)malloc(sizeof(float) * 3 * V);
    uint64_t* zb = (uint64_t*)malloc(sizeof(uint64_t) * S * S);
    for (int i = 0; i < S * S; i++) zb[i] = ~0ull;
    for (int v = 0; v < V; v++) {
      float x = pos[v * 3], y = pos[v * 3 + 1], z = pos[v * 3 + 2];
      float c[4];
      for (int j = 0; j < 4; j++) c[j] = ((M[j * 4] * x + M[j * 4 + 1] * y) + M[j * 4 + 2] * z) + M[j * 4 + 3];
      clip[v * 4 + 0] = c[0] * rt[0] + c[3] * rt[2];
      clip[v * 4 + 1] = c[1] * rt[1] + c[3] * rt[3];
      clip[v * 4 + 2] = c[2];
      clip[v * 4 + 3] = c[3];
      for (int j = 0; j < 3; j++) cam[v * 3 + j] = ((T[j * 4] * x + T[j * 4 + 1] * y) + T[j * 4 + 2] * z) + T[j * 4 + 3];
      float nx = vnormals[v * 3], ny = vnormals[v * 3 + 1], nz = vnormals[v * 3 + 2];
      float cn[3];
      for (int j = 0; j < 3; j++) cn[j] = (T[j * 4] * nx + T[j * 4 + 1] * ny) + T[j * 4 + 2] * nz;
      for (int j = 0; j < 3; j++) cnv[v * 3 + j] = cn[j];
      float nrm = sqrtf((cn[0] * cn[0] + cn[1] * cn[1]) + cn[2] * cn[2]);
      float den = fmaxf(nrm, 1e-12f);
      float d = -(cn[2] / den);
      dif[v] = fminf(fmaxf(d, 0.0f), 1.0f);
    }
    for (int f = 0; f < F; f++) {
      const int i0 = faces[f * 3], i1 = faces[f * 3 + 1], i2 = faces[f * 3 + 2];
      const float *c0 = clip + i0 * 4, *c1 = clip + i1 * 4, *c2 = clip + i2 * 4;
      float sxf[3], syf[3];
      const float* cc[3] = {c0, c1, c2};
      int bad = !(c0[3] > 0.0f && c1[3] > 0.0f && c2[3] > 0.0f);
      int64_t X[3], Y[3];
      for (int k = 0; k < 3 && !bad; k++) {
        sxf[k] = (cc[k][0] / cc[k][3]) * sub + sub;
        syf[k] = (cc[k][1] / cc[k][3]) * sub + sub;
        if (!(fabsf(sxf[k]) < 1073741824.0f) || !(fabsf(syf[k]) < 1073741824.0f)) bad = 1;
        X[k] = (int64_t)rintf(sxf[k]);
        Y[k] = (int64_t)rintf(syf[k]);
      }
      if (bad) {
        /* a corner behind the camera plane (w <= 0) or beyond the fixed-point range: nvdiffrast clips the triangle
         * against the frustum; the equivalent clip-free test is the 2-D homogeneous one over the whole crop */
        if (!(c0[3] > 0.0f || c1[3] > 0.0f || c2[3] > 0.0f)) continue;
        for (int py = 0; py < S; py++)
          for (int px = 0; px < S; px++) {
            float fx = (float)px * xs + xo, fy = (float)py * xs + xo;
            float zw;
            if (!homog_cover(c0, c1, c2, fx, fy, &zw)) continue;
            uint64_t key = ((uint64_t)f2ord(zw) << 32) | (uint32_t)f;
            if (key < zb[py * S + px]) zb[py * S + px] = key;
          }
        continue;
      }
      int64_t area = (X[1] - X[0]) * (Y[2] - Y[0]) - (Y[1] - Y[0]) * (X[2] - X[0]);
      if (area == 0) continue;
      int a = 0, b = 1, c = 2;
      if (area < 0) { b = 2; c = 1; }
      int64_t xmin = X[0], xmax = X[0], ymin = Y[0], ymax = Y[0];
      for (int k = 1; k < 3; k++) {
        if (X[k] < xmin) xmin = X[k]; if (X[k] > xmax) xmax = X[k];
        if (Y[k] < ymin) ymin = Y[k]; if (Y[k] > ymax) ymax = Y[k];
      }
      int64_t px0 = ceildiv16(xmin - 8), px1 = floordiv16(xmax - 8);
      int64_t py0 = ceildiv16(ymin - 8), py1 = floordiv16(ymax - 8);
      if (px0 < 0) px0 = 0; if (py0 < 0) py0 = 0;
      if (px1 > S - 1) px1 = S - 1; if (py1 > S - 1) py1 = S - 1;
      const int ea[3] = {a, b, c}, eb[3] = {b, c, a};
      for (int64_t py = py0; py <= py1; py++)
        for (int64_t px = px0; px <= px1; px++) {
          int64_t cxp = 16 * px + 8, cyp = 16 * py + 8;
          int inside = 1;
          for (int e = 0; e < 3 && inside; e++) {
            int64_t dx = X[eb[e]] - X[ea[e]], dy = Y[eb[e]] - Y[ea[e]];
            int64_t E = dx * (cyp - Y[ea[e]]) - dy * (cxp - X[ea[e]]);
            if (E < 0) inside = 0;
            else if (E == 0 && !(dy > 0 || (dy == 0 && dx < 0))) inside = 0;
          }
          if (!inside) continue;
          float fx = (float)px * xs + xo, fy = (float)py * xs + xo;
          float b0, b1, zw;
          bary_zw(c0, c1, c2, fx, fy, &b0, &b1, &zw);
          if (zw != zw) continue;
          uint64_t key = ((uint64_t)f2ord(zw) << 32) | (uint32_t)f;
          if (key < zb[py * S + px]) zb[py * S + px] = key;
        }
    }
    for (int i = 0; i < S; i++)
      for (int j = 0; j < S; j++) {
        const int py = S - 1 - i, px = j;
        const size_t o = ((size_t)n * S + i) * S + j;
        uint64_t key = zb[py * S + px];
        if (key == ~0ull) {
          for (int k = 0; k < 3; k++) { out_color[o * 3 + k] = 0.0f; out_xyz[o * 3 + k] = 0.0f; }
          if (out_normal) for (int k = 0; k < 3; k++) out_normal[o * 3 + k] = 0.0f;
          out_tri[o] = 0;
          continue;
        }
        const int f = (int)(uint32_t)(key & 0xffffffffu);
        const int i0 = faces[f * 3], i1 = faces[f * 3 + 1], i2 = faces[f * 3 + 2];
        float fx = (float)px * xs + xo, fy = (float)py * xs + xo;
        float b0, b1, zw;
        bary_zw(clip + i0 * 4, clip + i1 * 4, clip + i2 * 4, fx, fy, &b0, &b1, &zw);
        float b2 = (1.0f - b0) - b1;
        float col[3];
        for (int k = 0; k < 3; k++)
          out_xyz[o * 3 + k] = (b0 * cam[i0 * 3 + k] + b1 * cam[i1 * 3 + k]) + b2 * cam[i2 * 3 + k];
        if (tex) {
          const int t0 = uv_idx[f * 3], t1 = uv_idx[f * 3 + 1], t2 = uv_idx[f * 3 + 2];
          float tu = (b0 * uv[t0 * 2] + b1 * uv[t1 * 2]) + b2 * uv[t2 * 2];
          float tv = (b0 * uv[t0 * 2 + 1] + b1 * uv[t1 * 2 + 1]) + b2 * uv[t2 * 2 + 1];
          float u = tu * (float)Wt - 0.5f, v = tv * (float)Ht - 0.5f;
          float uf = floorf(u), vf = floorf(v);
          float du = u - uf, dv = v - vf;
          int iu0 = (int)uf, iv0 = (int)vf;
          int iu1 = iu0 + 1, iv1 = iv0 + 1;
          iu0 = ((iu0 % Wt) + Wt) % Wt; iu1 = ((iu1 % Wt) + Wt) % Wt;
          iv0 = ((iv0 % Ht) + Ht) % Ht; iv1 = ((iv1 % Ht) + Ht) % Ht;
          for (int k = 0; k < 3; k++) {
            float a00 = tex[(iv0 * Wt + iu0) * 3 + k], a10 = tex[(iv0 * Wt + iu1) * 3 + k];
            float a01 = tex[(iv1 * Wt + iu0) * 3 + k], a11 = tex[(iv1 * Wt + iu1) * 3 + k];
            float top = a00 + du * (a10 - a00), bot = a01 + du * (a11 - a01);
            col[k] = top + dv * (bot - top);
          }
        } else {
          for (int k = 0; k < 3; k++)
            col[k] = (b0 * vcolor[i0 * 3 + k] + b1 * vcolor[i1 * 3 + k]) + b2 * vcolor[i2 * 3 + k];
        }
        float dm = (b0 * dif[i0] + b1 * dif[i1]) + b2 * dif[i2];
        for (int k = 0; k < 3; k++) {
          float cv = use_light ? col[k] * w_ambient + (dm * col[k]) * w_diffuse : col[k];
          out_color[o * 3 + k] = fminf(fmaxf(cv, 0.0f), 1.0f);
        }
        if (out_normal) {
          float nm[3];
          for (int k = 0; k < 3; k++) nm[k] = (b0 * cnv[i0 * 3 + k] + b1 * cnv[i1 * 3 + k]) + b2 * cnv[i2 * 3 + k];
          float nn = sqrtf((nm[0] * nm[0] + nm[1] * nm[1]) + nm[2] * nm[2]);
          float dn = fmaxf(nn, 1e-12f);
          for (int k = 0; k < 3; k++) out_normal[o * 3 + k] = nm[k] / dn;
        }
        out_tri[o] = f + 1;
      }
    free(clip); free(cam); free(dif); free(cnv); free(zb);
  }
}

/* ------------------------------------------------------------------------
 * Network input A from a render: predict_pose_refine.py:56-58 (rgb*255),
 * h5_dataset.py:79-127 / 137-179 (rgb/255; xyz -= t; xyz *= 1/radius;
 * per-channel zeroing of invalid | |xyz|>=2).  zthr = 0.001 (refiner) or
 * 0.1 (scorer).  Output NHWC fp32 [N,S,S,6] = (r,g,b,x,y,z).
 * ---------------------------------------------------------------------- */
static inline void xyz_norm(const float* xyz, const float* T, float inv_radius, float zthr, float* o)
{
  int inv = xyz[2] < zthr;
  const float t[3] = {T[3], T[7], T[11]};
  for (int k = 0; k < 3; k++) {
    float x = (xyz[k] - t[k]) * inv_radius;
    o[k] = (inv || fabsf(x) >= 2.0f) ? 0.0f : x;
  }
}

ORC_API void orc_make_A(const float* color, const float* xyz, const float* poses, int N, int S,
                        float mesh_diameter, float zthr, float* A)
{
  const float radius = mesh_diameter / 2.0f;
  const float inv_radius = 1.0f / radius;
  /* transform_batch's `rgb.cuda().float()/255.0` (h5_dataset.py:123-124): on a CUDA tensor ATen divides by a Python
   * scalar as a multiplication with the fp32 reciprocal (BinaryDivTrueKernel.cu, "a * reciprocal(b)"), and the
   * reference only runs on CUDA -- so the faithful value is x * (1.0f/255.0f), not the correctly rounded quotient */
  const float inv255 = 1.0f / 255.0f;
#pragma omp parallel for schedule(static)
  for (int n = 0; n < N; n++)
    for (int p = 0; p < S * S; p++) {
      size_t o = (size_t)n * S * S + p;
      for (int k = 0; k < 3; k++) A[o * 6 + k] = (color[o * 3 + k] * 255.0f) * inv255;
      xyz_norm(xyz + o * 3, poses + n * 16, inv_radius, zthr, A + o * 6 + 3);
    }
}

/* ------------------------------------------------------------------------
 * Network input B: kornia.warp_perspective(align_corners=False) of the
 * observed frame into each crop (predict_pose_refine.py:63,72;
 * predict_score.py:89-90; h5_dataset.py:137-170), composed with ATen's
 * grid_sample rules (bilinear: nw,ne,sw,se accumulation, zero padding;
 * nearest: nearbyint).  Source location of crop pixel (j,i):
 *     p = tf^-1 (j,i,1);  ix = p_x * W/(W-1) - 0.5;  iy = p_y * H/(H-1) - 0.5
 *   mode 0 (refiner): xyz = nearest(xyz_map),           zthr 0.001
 *   mode 1 (scorer) : depth crop -> full -> xyz (fp32) -> crop, three nearest
 *                     look-ups as pure index math,       zthr 0.1
 * rgb is [H,W,3] fp32 in 0..255; xyz_map [H,W,3]; depth [H,W].
 * ---------------------------------------------------------------------- */
static inline int nearest_src(float p, float scale, int size)
{
  float ix = p * scale - 0.5f;
  float r = nearbyintf(ix);
  if (!(r >= 0.0f && r <= (float)(size - 1))) return -1;
  return (int)r;
}

ORC_API void orc_make_B(const float* rgb, const float* xyz_map, const float* depth, int H, int W,
                        const float* poses, const float* tf9, const float* inv4, const float* K9,
                        int N, int S, float mesh_diameter, int mode, float* B)
{
  const float radius = mesh_diameter / 2.0f;
  const float inv_radius = 1.0f / radius;
  const float wsc = (float)W / (float)(W - 1), hsc = (float)H / (float)(H - 1);
  const float ssc = (float)S / (float)(S - 1);
  const float zthr = mode ? 0.1f : 0.001f;
  const float inv255 = 1.0f / 255.0f;
#pragma omp parallel for schedule(static)
  for (int n = 0; n < N; n++) {
    const float* iv = inv4 + n * 4;
    const float* tf = tf9 + n * 9;
    for (int i = 0; i < S; i++)
      for (int j = 0; j < S; j++) {
        float* o = B + (((size_t)n * S + i) * S + j) * 6;
        float px = (float)j * iv[0] + iv[1];
        float py = (float)i * iv[2] + iv[3];
        /* bilinear rgb */
        float ix = px * wsc - 0.5f, iy = py * hsc - 0.5f;
        float x0 = floorf(ix), y0 = floorf(iy), x1 = x0 + 1.0f, y1 = y0 + 1.0f;
        float wnw = (x1 - ix) * (y1 - iy), wne = (ix - x0) * (y1 - iy);
        float wsw = (x1 - ix) * (iy - y0), wse = (ix - x0) * (iy - y0);
        int inx0 = x0 >= 0.0f && x0 <= (float)(W - 1), inx1 = x1 >= 0.0f && x1 <= (float)(W - 1);
        int iny0 = y0 >= 0.0f && y0 <= (float)(H - 1), iny1 = y1 >= 0.0f && y1 <= (float)(H - 1);
        for (int k = 0; k < 3; k++) {
          float acc = 0.0f;
          if (inx0 && iny0) acc += rgb[((size_t)(int)y0 * W + (int)x0) * 3 + k] * wnw;
          if (inx1 && iny0) acc += rgb[((size_t)(int)y0 * W + (int)x1) * 3 + k] * wne;
          if (inx0 && iny1) acc += rgb[((size_t)(int)y1 * W + (int)x0) * 3 + k] * wsw;
          if (inx1 && iny1) acc += rgb[((size_t)(int)y1 * W + (int)x1) * 3 + k] * wse;
          o[k] = acc * inv255;      /* see orc_make_A */
        }
        /* nearest xyz */
        float xyz[3] = {0.0f, 0.0f, 0.0f};
        int u = nearest_src(px, wsc, W), v = nearest_src(py, hsc, H);
        if (u >= 0 && v >= 0) {
          if (mode == 0) {
            const float* s = xyz_map + ((size_t)v * W + u) * 3;
            xyz[0] = s[0]; xyz[1] = s[1]; xyz[2] = s[2];
          } else {
            /* stage 2: full pixel (u,v) reads crop pixel (j2,i2) */
            float qx = tf[0] * (float)u + tf[2], qy = tf[4] * (float)v + tf[5];
            int j2 = nearest_src(qx, ssc, S), i2 = nearest_src(qy, ssc, S);
            if (j2 >= 0 && i2 >= 0) {
              /* stage 1: crop pixel (j2,i2) reads full pixel (u1,v1) */
              float px1 = (float)j2 * iv[0] + iv[1], py1 = (float)i2 * iv[2] + iv[3];
              int u1 = nearest_src(px1, wsc, W), v1 = nearest_src(py1, hsc, H);
              if (u1 >= 0 && v1 >= 0) {
                float z = depth[(size_t)v1 * W + u1];
                if (!(z < 0.001f)) {
                  xyz[0] = ((float)u - K9[2]) * z / K9[0];
                  xyz[1] = ((float)v - K9[5]) * z / K9[4];
                  xyz[2] = z;
                }
              }
            }
          }
        }
        (void)zthr;
        xyz_norm(xyz, poses + n * 16, inv_radius, zthr, o + 3);
      }
  }
}

/* ------------------------------------------------------------------------
 * cluster_poses            reference: mycpp/src/app/pybind_api.cpp:24-68,
 * rotationGeodesicDistance mycpp/src/Utils.cpp:21-26.  fp32 like Eigen's
 * Matrix4f; products row.column accumulated left to right.
 * Returns the number kept; keep_idx receives the kept input indices.
 * ---------------------------------------------------------------------- */
ORC_API int orc_cluster_poses(float angle_diff_deg, float dist_diff, const float* poses, int N,
                              const float* sym, int nsym, int32_t* keep_idx)
{
  const float radian_thres = (float)((double)angle_diff_deg / 180.0 * M_PI);
  int nk = 0;
  if (N <= 0) return 0;
  keep_idx[nk++] = 0;
  for (int i = 1; i < N; i++) {
    const float* cur = poses + i * 16;
    int isnew = 1;
    for (int c = 0; c < nk && isnew; c++) {
      const float* cl = poses + keep_idx[c] * 16;
      float dx = cl[3] - cur[3], dy = cl[7] - cur[7], dz = cl[11] - cur[11];
      float nrm = sqrtf((dx * dx + dy * dy) + dz * dz);
      if (nrm >= dist_diff) continue;
      for (int s = 0; s < nsym; s++) {
        const float* tf = sym + s * 16;
        float R1[9];
        for (int r = 0; r < 3; r++)
          for (int cc = 0; cc < 3; cc++)
            R1[r * 3 + cc] = ((cur[r * 4 + 0] * tf[0 * 4 + cc] + cur[r * 4 + 1] * tf[1 * 4 + cc]) +
                              cur[r * 4 + 2] * tf[2 * 4 + cc]) + cur[r * 4 + 3] * tf[3 * 4 + cc];
        float tr = 0.0f;
        for (int r = 0; r < 3; r++)
          tr += (R1[r * 3 + 0] * cl[r * 4 + 0] + R1[r * 3 + 1] * cl[r * 4 + 1]) + R1[r * 3 + 2] * cl[r * 4 + 2];
        float cosv = (float)(((double)tr - 1.0) / 2.0);
        cosv = fmaxf(fminf(cosv, 1.0f), -1.0f);
        if (acosf(cosv) < radian_thres) { isnew = 0; break; }
      }
    }
    if (isnew) keep_idx[nk++] = i;
  }
  return nk;
}
