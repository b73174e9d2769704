// Hypothesis-side small kernels: cluster_poses, crop windows, pose update.
// Compiled with -fmad=false (see depth.cu).
//   cluster_poses                     mycpp/src/app/pybind_api.cpp:24-68, mycpp/src/Utils.cpp:21-26
//   compute_crop_window_tf_batch      src/Utils.py:577-621
//   bbox2d_ori / render zoom / mtx    predict_pose_refine.py:44-45, src/Utils.py:159-181
//   pose update                       predict_pose_refine.py:195-231, src/Utils.py:848-855
#include "common.cuh"

namespace {

struct K9 { float k[9]; };
struct P16 { float p[16]; };

// Greedy, order-dependent de-duplication: one CTA; pose i is tested against all kept poses in
// parallel, the append is serial, so the kept set is identical to the sequential C++ loop.
__global__ void __launch_bounds__(256) cluster_kernel(const float* __restrict__ poses, int N, const float* __restrict__ sym,
                                                      int S, float radian_thres, float dist_diff,
                                                      int32_t* __restrict__ keep, int32_t* __restrict__ n_keep)
{
  __shared__ int s_nk;
  if (threadIdx.x == 0) { s_nk = 1; keep[0] = 0; }
  __syncthreads();
  for (int i = 1; i < N; i++) {
    const float* cur = poses + (size_t)i * 16;
    const int nk = s_nk;
    int dup = 0;
    for (int c = threadIdx.x; c < nk && !dup; c += blockDim.x) {
      const float* cl = poses + (size_t)keep[c] * 16;
      float dx = cl[3] - cur[3], dy = cl[7] - cur[7], dz = cl[11] - cur[11];
      float nrm = sqrtf((dx * dx + dy * dy) + dz * dz);
      if (nrm >= dist_diff) continue;
      for (int s = 0; s < S; s++) {
        const float* tf = sym + (size_t)s * 16;
        float tr = 0.0f;
        for (int r = 0; r < 3; r++) {
          float R1[3];
          for (int cc = 0; cc < 3; cc++)
            R1[cc] = ((cur[r * 4 + 0] * tf[0 * 4 + cc] + cur[r * 4 + 1] * tf[1 * 4 + cc]) + cur[r * 4 + 2] * tf[2 * 4 + cc]) +
                     cur[r * 4 + 3] * tf[3 * 4 + cc];
          tr += (R1[0] * cl[r * 4 + 0] + R1[1] * cl[r * 4 + 1]) + R1[2] * cl[r * 4 + 2];
        }
        float cosv = (float)(((double)tr - 1.0) / 2.0);
        cosv = fmaxf(fminf(cosv, 1.0f), -1.0f);
        if (acosf(cosv) < radian_thres) { dup = 1; break; }
      }
    }
    int any = __syncthreads_or(dup);
    if (threadIdx.x == 0 && !any) { keep[nk] = i; s_nk = nk + 1; }
    __syncthreads();
  }
  if (threadIdx.x == 0) n_keep[0] = s_nk;
}

__global__ void crop_params_kernel(const float* __restrict__ poses, int N, K9 K, float radius, int S, int H, int W, P16 P,
                                   float* __restrict__ params)
{
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const float* T = poses + (size_t)n * 16;
  float* o = params + (size_t)n * FPB_CROP_STRIDE;
  const float tx = T[3], ty = T[7], tz = T[11];
  float u0 = 0.f, v0 = 0.f, rad = 0.0f;
  for (int k = 0; k < 5; k++) {
    float ox = (k == 1) ? radius : (k == 2 ? -radius : 0.0f);
    float oy = (k == 3) ? radius : (k == 4 ? -radius : 0.0f);
    float X = tx + ox, Y = ty + oy, Z = tz + 0.0f;
    float px = (K.k[0] * X + K.k[1] * Y) + K.k[2] * Z;
    float py = (K.k[3] * X + K.k[4] * Y) + K.k[5] * Z;
    float pz = (K.k[6] * X + K.k[7] * Y) + K.k[8] * Z;
    float u = px / pz, v = py / pz;
    if (k == 0) { u0 = u; v0 = v; }
    float du = fabsf(u - u0), dv = fabsf(v - v0);
    if (du > rad) rad = du;
    if (dv > rad) rad = dv;
  }
  float left = rintf(u0 - rad), right = rintf(u0 + rad), top = rintf(v0 - rad), bottom = rintf(v0 + rad);
  float sx = (float)S / (right - left), sy = (float)S / (bottom - top);
  float ttx = sx * (-left), tty = sy * (-top);
  o[0] = sx; o[1] = 0; o[2] = ttx; o[3] = 0; o[4] = sy; o[5] = tty; o[6] = 0; o[7] = 0; o[8] = 1;
  float i00 = 1.0f / sx, i02 = -(ttx / sx), i11 = 1.0f / sy, i12 = -(tty / sy);
  float umin = i02, vmin = i12, umax = i00 * (float)(S - 1) + i02, vmax = i11 * (float)(S - 1) + i12;
  o[9] = umin; o[10] = vmin; o[11] = umax; o[12] = vmax;
  float l = umin, t = (float)H - vmin, r = umax, b = (float)H - vmax;
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) {
      float g0 = T[j], g1 = -T[4 + j], g2 = -T[8 + j], g3 = T[12 + j];
      o[13 + i * 4 + j] = ((P.p[i * 4 + 0] * g0 + P.p[i * 4 + 1] * g1) + P.p[i * 4 + 2] * g2) + P.p[i * 4 + 3] * g3;
    }
  o[29] = (float)W / (r - l);
  o[30] = (float)H / (t - b);
  o[31] = (((float)W - r) - l) / (r - l);
  o[32] = (((float)H - t) - b) / (t - b);
  o[33] = i00; o[34] = i02; o[35] = i11; o[36] = i12;
  o[37] = 0; o[38] = 0; o[39] = 0;
}

struct F3 { float v[3]; };

struct PoseUpdateArgs {
  float radius, rot_normalizer;
  F3 tn;
  int normalize_xyz, trans_rep, rot_rep, S;
  float K[9];
};

__global__ void pose_update_kernel(const float* __restrict__ poses, const float* __restrict__ trans,
                                   const float* __restrict__ rot, int N, PoseUpdateArgs a,
                                   const float* __restrict__ crop_params, float* __restrict__ out,
                                   float* __restrict__ out_td, float* __restrict__ out_rd)
{
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const float* T = poses + (size_t)n * 16;
  float td[3];
  if (a.trans_rep == FPB_TRANS_DEEPIM) {
    // predict_pose_refine.py:201-215: the network predicts the shift of the projected object centre inside the
    // crop (in units of the crop size) and the depth ratio; back-project through the inverse window and inverse K
    const float* pr = crop_params + (size_t)n * FPB_CROP_STRIDE;
    const float* K = a.K;
    float uvw[3];
    for (int i = 0; i < 3; i++) uvw[i] = (K[i * 3] * T[3] + K[i * 3 + 1] * T[7]) + K[i * 3 + 2] * T[11];
    const float u = uvw[0] / uvw[2], v = uvw[1] / uvw[2];
    const float uc = (pr[0] * u + pr[1] * v) + pr[2], vc = (pr[3] * u + pr[4] * v) + pr[5];      // tf_to_crop @ (u, v, 1)
    const float z_pred = trans[n * 3 + 2] * T[11];
    const float up = uc + trans[n * 3] * (float)a.S, vp = vc + trans[n * 3 + 1] * (float)a.S;
    // inverse window (scale + shift): params[33..36] = (1/s_x, -t_x/s_x, 1/s_y, -t_y/s_y)
    const float uf = pr[33] * up + pr[34], vf = pr[35] * vp + pr[36];
    // inverse of the upper-triangular K applied to (uf, vf, 1)
    const float yn = (vf - K[5]) / K[4];
    const float xn = ((uf - K[2]) - K[1] * yn) / K[0];
    td[0] = xn * z_pred - T[3]; td[1] = yn * z_pred - T[7]; td[2] = z_pred - T[11];
    if (a.normalize_xyz)
      for (int k = 0; k < 3; k++) td[k] *= a.radius;         // applied to every representation (:226-227)
  } else {
    for (int k = 0; k < 3; k++) {
      float t = trans[n * 3 + k];
      td[k] = a.normalize_xyz ? t * a.radius : tanhf(t) * a.tn.v[k];
    }
  }
  float R[9];
  if (a.rot_rep == FPB_ROT_6D) {
    // pytorch3d rotation_6d_to_matrix: rows b1, b2, b3 (Gram-Schmidt; F.normalize eps 1e-12)
    const float* d6 = rot + (size_t)n * 6;
    float b1[3], b2[3], b3[3];
    float n1 = sqrtf((d6[0] * d6[0] + d6[1] * d6[1]) + d6[2] * d6[2]);
    n1 = fmaxf(n1, 1e-12f);
    for (int k = 0; k < 3; k++) b1[k] = d6[k] / n1;
    const float dot = (b1[0] * d6[3] + b1[1] * d6[4]) + b1[2] * d6[5];
    for (int k = 0; k < 3; k++) b2[k] = d6[3 + k] - dot * b1[k];
    float n2 = sqrtf((b2[0] * b2[0] + b2[1] * b2[1]) + b2[2] * b2[2]);
    n2 = fmaxf(n2, 1e-12f);
    for (int k = 0; k < 3; k++) b2[k] = b2[k] / n2;
    b3[0] = b1[1] * b2[2] - b1[2] * b2[1];
    b3[1] = b1[2] * b2[0] - b1[0] * b2[2];
    b3[2] = b1[0] * b2[1] - b1[1] * b2[0];
    for (int k = 0; k < 3; k++) { R[k] = b1[k]; R[3 + k] = b2[k]; R[6 + k] = b3[k]; }
  } else {
    float r[3];
    for (int k = 0; k < 3; k++) r[k] = tanhf(rot[n * 3 + k]) * a.rot_normalizer;
    // so3_exp_map (pytorch3d, eps 1e-4)
    float nrm = (r[0] * r[0] + r[1] * r[1]) + r[2] * r[2];
    float theta = sqrtf(fmaxf(nrm, 1e-4f));
    float f1 = sinf(theta) / theta;
    float f2 = (1.0f - cosf(theta)) / (theta * theta);
    float Km[9] = {0.f, -r[2], r[1], r[2], 0.f, -r[0], -r[1], r[0], 0.f};
    float K2[9];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) K2[i * 3 + j] = (Km[i * 3] * Km[j] + Km[i * 3 + 1] * Km[3 + j]) + Km[i * 3 + 2] * Km[6 + j];
    for (int i = 0; i < 9; i++) R[i] = (f1 * Km[i] + f2 * K2[i]) + ((i % 4 == 0) ? 1.0f : 0.0f);
  }
  float* o = out + (size_t)n * 16;
  // rot_delta = R^T (the .permute(0,2,1) of predict_pose_refine.py:222,224); B[:3,:3] = rot_delta @ A[:3,:3]
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) o[i * 4 + j] = (R[0 * 3 + i] * T[0 * 4 + j] + R[1 * 3 + i] * T[1 * 4 + j]) + R[2 * 3 + i] * T[2 * 4 + j];
  o[3] = T[3] + td[0]; o[7] = T[7] + td[1]; o[11] = T[11] + td[2];
  o[12] = 0.f; o[13] = 0.f; o[14] = 0.f; o[15] = 1.f;
  // trans_delta / rot_mat_delta as the reference keeps them in last_trans_update / last_rot_update (:238-239)
  if (out_td) { out_td[n * 3] = td[0]; out_td[n * 3 + 1] = td[1]; out_td[n * 3 + 2] = td[2]; }
  if (out_rd)
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) out_rd[n * 9 + i * 3 + j] = R[j * 3 + i];
}


// ---- hypotheses from the mask statistics, on the device (src/estimater.py:127-156 without the host round trip) ----
struct Mat3d { double v[9]; };
__global__ void hypo_from_stats_kernel(const int32_t* __restrict__ si, const float* __restrict__ sf, Mat3d Kinv,
                                       const float* __restrict__ rot_grid, int N, int min_valid, float* __restrict__ poses,
                                       double* __restrict__ center_out)
{
  // same operation order as _center_from_stats (float64 bbox centre, float32 median of the two middle depths)
  double c[3] = {0.0, 0.0, 0.0};
  const bool ok = si[4] > 0 && si[5] > 0;
  if (ok) {
    const double uc = (double)(si[0] + si[1]) / 2.0, vc = (double)(si[2] + si[3]) / 2.0;
    const float zc = (sf[0] + sf[1]) / 2.0f;
    for (int i = 0; i < 3; i++) c[i] = (Kinv.v[i * 3 + 0] * uc + Kinv.v[i * 3 + 1] * vc + Kinv.v[i * 3 + 2]) * (double)zc;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && center_out) { center_out[0] = c[0]; center_out[1] = c[1]; center_out[2] = c[2]; }
  // fewer than min_valid valid pixels: register() returns early on the host; the pipeline that is already enqueued
  // runs on a harmless dummy translation instead of a degenerate one
  float t[3] = {(float)c[0], (float)c[1], (float)c[2]};
  if (si[5] < min_valid) { t[0] = 0.0f; t[1] = 0.0f; t[2] = 1.0f; }
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const float* g = rot_grid + (size_t)n * 16;
  float* o = poses + (size_t)n * 16;
  for (int i = 0; i < 16; i++) o[i] = g[i];
  o[3] = t[0]; o[7] = t[1]; o[11] = t[2];
}

}  // namespace

FPB_EXPORT int fpb_cluster_poses_f32(const float* poses, int N, const float* sym, int S, float angle_deg, float dist,
                                     int32_t* out_keep_idx, int32_t* out_n_keep, void* stream)
{
  FPB_CHECK_ARG(poses && sym && out_keep_idx && out_n_keep && N > 0 && S > 0);
  const float radian_thres = (float)((double)angle_deg / 180.0 * 3.14159265358979323846);
  cluster_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(poses, N, sym, S, radian_thres, dist, out_keep_idx, out_n_keep);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

FPB_EXPORT int fpb_crop_params(const float* poses, int N, const float* K9_h, float radius, int S, int H, int W,
                               const float* P16_h, float* params, void* stream)
{
  FPB_CHECK_ARG(poses && K9_h && P16_h && params && N > 0 && S > 1 && H > 1 && W > 1);
  K9 K; P16 P;
  for (int i = 0; i < 9; i++) K.k[i] = K9_h[i];
  for (int i = 0; i < 16; i++) P.p[i] = P16_h[i];
  crop_params_kernel<<<fpb_cdiv(N, 128), 128, 0, (cudaStream_t)stream>>>(poses, N, K, radius, S, H, W, P, params);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

FPB_EXPORT int fpb_pose_update_deltas(const float* poses, const float* trans, const float* rot, int N, float mesh_diameter,
                                      float rot_normalizer, const float* tn3_h, int normalize_xyz, int trans_rep,
                                      int rot_rep, const float* crop_params, const float* K9_h, int S, float* out_poses,
                                      float* out_trans_delta, float* out_rot_delta, void* stream)
{
  FPB_CHECK_ARG(poses && trans && rot && out_poses && N > 0);
  FPB_CHECK_ARG(trans_rep == FPB_TRANS_TRACKNET || trans_rep == FPB_TRANS_DEEPIM);
  FPB_CHECK_ARG(rot_rep == FPB_ROT_AXIS_ANGLE || rot_rep == FPB_ROT_6D);
  FPB_CHECK_ARG(trans_rep == FPB_TRANS_DEEPIM ? (crop_params && K9_h && S > 0) : (normalize_xyz || tn3_h));
  PoseUpdateArgs a;
  a.radius = mesh_diameter / 2.0f;
  a.rot_normalizer = rot_normalizer;
  for (int i = 0; i < 3; i++) a.tn.v[i] = tn3_h ? tn3_h[i] : 0.f;
  a.normalize_xyz = normalize_xyz; a.trans_rep = trans_rep; a.rot_rep = rot_rep; a.S = S;
  for (int i = 0; i < 9; i++) a.K[i] = K9_h ? K9_h[i] : 0.f;
  pose_update_kernel<<<fpb_cdiv(N, 128), 128, 0, (cudaStream_t)stream>>>(poses, trans, rot, N, a, crop_params, out_poses,
                                                                          out_trans_delta, out_rot_delta);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

FPB_EXPORT int fpb_pose_update_ex(const float* poses, const float* trans, const float* rot, int N, float mesh_diameter,
                                  float rot_normalizer, const float* tn3_h, int normalize_xyz, int trans_rep, int rot_rep,
                                  const float* crop_params, const float* K9_h, int S, float* out_poses, void* stream)
{
  return fpb_pose_update_deltas(poses, trans, rot, N, mesh_diameter, rot_normalizer, tn3_h, normalize_xyz, trans_rep, rot_rep,
                                crop_params, K9_h, S, out_poses, nullptr, nullptr, stream);
}

FPB_EXPORT int fpb_pose_update(const float* poses, const float* trans, const float* rot, int N, float mesh_diameter,
                               float rot_normalizer, const float* tn3_h, int normalize_xyz, float* out_poses, void* stream)
{
  return fpb_pose_update_ex(poses, trans, rot, N, mesh_diameter, rot_normalizer, tn3_h, normalize_xyz, FPB_TRANS_TRACKNET,
                            FPB_ROT_AXIS_ANGLE, nullptr, nullptr, 0, out_poses, stream);
}

FPB_EXPORT int fpb_hypotheses_from_stats(const int32_t* stats_i32, const float* stats_f32, const double* Kinv9_h,
                                         const float* rot_grid, int N, int min_valid, float* poses_out, double* center_out,
                                         void* stream)
{
  FPB_CHECK_ARG(stats_i32 && stats_f32 && Kinv9_h && rot_grid && poses_out && N > 0);
  Mat3d Ki;
  for (int i = 0; i < 9; i++) Ki.v[i] = Kinv9_h[i];
  hypo_from_stats_kernel<<<fpb_cdiv(N, 128), 128, 0, (cudaStream_t)stream>>>(stats_i32, stats_f32, Ki, rot_grid, N, min_valid,
                                                                              poses_out, center_out);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

// ---- ranking tail of register(): src/estimater.py:221-236 (argsort(descending) of the scores, gather of poses and
//      scores, best pose @ tf_to_centered_mesh) in one block: bitonic sort of (score, index) keys in shared memory;
//      equal scores keep their input order (index ascending) ----
constexpr int RANK_MAX = 4096;
__global__ void __launch_bounds__(1024) rank_poses_kernel(const float* __restrict__ scores, const float* __restrict__ poses,
                                                           const float* __restrict__ tf, int N, int P, long long* __restrict__ out_ids,
                                                           float* __restrict__ out_scores, float* __restrict__ out_poses,
                                                           float* __restrict__ out_best)
{
  extern __shared__ unsigned long long keys[];      // P = next power of two >= N
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    unsigned long long k = ~0ull;                    // padding sorts last
    if (i < N) {
      uint32_t u = __float_as_uint(scores[i]);
      u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);          // ascending unsigned order == ascending float order
      k = ((unsigned long long)(~u) << 32) | (uint32_t)i;       // descending score, then ascending index (NaN-free scores)
    }
    keys[i] = k;
  }
  __syncthreads();
  for (int size = 2; size <= P; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = threadIdx.x; i < P; i += blockDim.x) {
        const int j = i ^ stride;
        if (j > i) {
          const bool up = (i & size) == 0;
          const unsigned long long a = keys[i], b = keys[j];
          if ((a > b) == up) { keys[i] = b; keys[j] = a; }
        }
      }
      __syncthreads();
    }
  for (int r = threadIdx.x; r < N; r += blockDim.x) {
    const int src = (int)(uint32_t)(keys[r] & 0xffffffffu);
    out_ids[r] = src;
    out_scores[r] = scores[src];
    for (int e = 0; e < 16; e++) out_poses[(size_t)r * 16 + e] = poses[(size_t)src * 16 + e];
  }
  if (threadIdx.x < 16 && out_best) {                // best = poses[ids[0]] @ tf  (fp32, k ascending like a GEMM row)
    const int src = (int)(uint32_t)(keys[0] & 0xffffffffu);
    const int i = threadIdx.x >> 2, j = threadIdx.x & 3;
    const float* Tp = poses + (size_t)src * 16;
    float s = 0.0f;
    for (int k = 0; k < 4; k++) s = fmaf(Tp[i * 4 + k], tf[k * 4 + j], s);
    out_best[threadIdx.x] = s;
  }
}

FPB_EXPORT int fpb_rank_poses(const float* scores, const float* poses, const float* tf16, int N, long long* out_ids,
                              float* out_scores, float* out_poses, float* out_best16, void* stream)
{
  FPB_CHECK_ARG(scores && poses && tf16 && out_ids && out_scores && out_poses && N > 0);
  if (N > RANK_MAX) { fpb_set_error("fpb_rank_poses: at most %d hypotheses per call", RANK_MAX); return FPB_ERR_UNSUPPORTED; }
  int P = 1;
  while (P < N) P <<= 1;
  rank_poses_kernel<<<1, 1024, (size_t)P * sizeof(unsigned long long), (cudaStream_t)stream>>>(scores, poses, tf16, N, P, out_ids, out_scores,
                                                                                            out_poses, out_best16);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

// ---- small multi-destination store: the refined poses of this rank's hypotheses into every rank's pose table
//      (peer-mapped symmetric memory), the pose half of the sharded register()'s exchange ----
namespace {
struct Dst8 { float* p[FPB_MAX_PEERS]; int n; };
__global__ void store_multi_kernel(const float* __restrict__ src, long long count, Dst8 dst)
{
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const float v = src[i];
  for (int d = 0; d < dst.n; d++) dst.p[d][i] = v;
}
}  // namespace

FPB_EXPORT int fpb_store_multi_f32(const float* src, long long count, float* const* dst_ptrs_h, int n_dst, void* stream)
{
  FPB_CHECK_ARG(src && count > 0 && dst_ptrs_h && n_dst >= 1 && n_dst <= FPB_MAX_PEERS);
  Dst8 d;
  d.n = n_dst;
  for (int i = 0; i < n_dst; i++) {
    FPB_CHECK_ARG(dst_ptrs_h[i]);
    d.p[i] = dst_ptrs_h[i];
  }
  store_multi_kernel<<<fpb_cdiv(count, 256), 256, 0, (cudaStream_t)stream>>>(src, count, d);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}
