/*
 * fpb200.h -- C ABI of libfpb200.so: the B200-native (sm_100a) kernels behind
 * FoundationPose's render-and-compare hypothesis path.
 *
 * Conventions:
 *   - every pointer is a DEVICE pointer unless its name ends in _h (host);
 *   - the caller owns every buffer, including workspaces; handles own only their
 *     packed weights;  no allocation, no implicit sync inside compute calls;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it;
 *   - return 0 on success, <0 on error; text via fpb_last_error() (thread-local);
 *   - layouts: images [H,W] / [H,W,3] row-major; poses [N,4,4] row-major fp32;
 *     network inputs NHWC with 8 channels (r,g,b,x,y,z,0,0).
 * Each entry point cites the reference interface it replaces (paths relative to the
 * reference repository root).
 */
#ifndef FPB200_H
#define FPB200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define FPB_OK 0
#define FPB_ERR_ARG (-1)
#define FPB_ERR_CUDA (-2)
#define FPB_ERR_UNSUPPORTED (-3)

/* precision modes of the network path */
#define FPB_PREC_FP32 0   /* SIMT fp32 parity mode */
#define FPB_PREC_BF16 1   /* tcgen05 bf16 operands, fp32 accumulation */
#define FPB_PREC_FP16 2   /* tcgen05 IEEE fp16 operands, fp32 accumulation: the dtype of the reference's autocast path
                             (predict_pose_refine.py:190, predict_score.py:193); same instruction, same speed, 3 more
                             mantissa bits than bf16 */
/* element type of 16-bit image / activation buffers handed to fpb_render_crops / fpb_warp_crops (a_is_bf16 /
 * b_is_bf16 arguments): 0 fp32, 1 bf16, 2 fp16 (= the FPB_PREC_* value of the network that will read them) */

const char* fpb_last_error(void);
int fpb_version(void);

/* ---- depth pre-processing: src/Utils.py:359-395 (erode_depth), :304-356
 *      (bilateral_filter_depth), :399-417 (depth2xyzmap, math_f64=1),
 *      :420-438 (depth2xyzmap_batch, math_f64=0) ------------------------- */
int fpb_erode_depth_f32(const float* in, float* out, int H, int W, int radius, float diff_thres,
                        float ratio_thres, float zfar, void* stream);
int fpb_bilateral_depth_f32(const float* in, float* out, int H, int W, int radius, float zfar,
                            float sigmaD, float sigmaR, void* stream);
int fpb_depth_to_xyz(const float* depth, const double* K9_h, float zfar, float* xyz_hw3, int H, int W,
                     int math_f64, void* stream);

/* ---- guess_translation statistics: src/estimater.py:137-156.
 *      out_i32[6] = umin, umax, vmin, vmax, n_mask, n_valid;
 *      out_f32[2] = the two middle order statistics of depth[mask & depth>=0.001]
 *      (equal when n_valid is odd).  mask is uint8 [H,W] (non-zero = set). ---- */
int fpb_mask_depth_stats(const float* depth, const uint8_t* mask, int H, int W, int32_t* out_i32,
                         float* out_f32, void* stream);

/* ---- hypotheses on the device: generate_random_pose_hypo + guess_translation, src/estimater.py:127-156.
 *      poses_out[N,4,4] = rot_grid with translation inv(K) @ [uc, vc, 1] * median depth (float64 math as on the
 *      host); center_out[3] (device float64, may be NULL) = that translation, zeros for an empty mask / no valid
 *      depth.  With fewer than min_valid valid pixels the poses get the dummy translation (0,0,1): register() returns
 *      early in that case (src/estimater.py:176-180) and the caller discards whatever was enqueued. ---- */
int fpb_hypotheses_from_stats(const int32_t* stats_i32, const float* stats_f32, const double* Kinv9_h,
                              const float* rot_grid, int N, int min_valid, float* poses_out, double* center_out,
                              void* stream);

/* ---- register() prologue in one call, no host synchronisation: src/estimater.py:159-205
 *      (erode_depth -> bilateral_filter_depth -> guess_translation -> hypotheses -> depth2xyzmap).
 *      depth_tmp / depth_out [H,W], xyz_out [H,W,3]; stats as fpb_mask_depth_stats. ---- */
int fpb_register_prologue(const float* depth_in, const uint8_t* mask, int H, int W, const double* K9_h,
                          const double* Kinv9_h, float* depth_tmp, float* depth_out, float* xyz_out,
                          const float* rot_grid, int N, int min_valid, float* poses_out, int32_t* stats_i32,
                          float* stats_f32, double* center_out, void* stream);

/* ---- cluster_poses: mycpp/src/app/pybind_api.cpp:24-68.
 *      out_keep_idx[N] receives kept input indices, out_n_keep[1] their count. ---- */
int fpb_cluster_poses_f32(const float* poses, int N, const float* sym, int S, float angle_deg, float dist,
                          int32_t* out_keep_idx, int32_t* out_n_keep, void* stream);

/* ---- pose-error metrics over a batch of B poses (row-major 4x4 float), model points [P,3]:
 *      add_err src/Utils.py:232-240, adds_err src/Utils.py:242-253 (exhaustive nearest neighbour instead of cKDTree).
 *      gt is one pose (gt_batched = 0) or B poses; out[B] on the device. ---- */
int fpb_add_err(const float* pred, const float* gt, int gt_batched, const float* pts, int P, int B, float* out,
                void* stream);
int fpb_adds_err(const float* pred, const float* gt, int gt_batched, const float* pts, int P, int B, float* out,
                 void* stream);

/* ---- crop windows: compute_crop_window_tf_batch(method='box_3d') src/Utils.py:577-621,
 *      bbox2d_ori predict_pose_refine.py:44-45, render matrices src/Utils.py:159-181.
 *      params layout per hypothesis (FPB_CROP_STRIDE floats):
 *        [0:9) tf_to_crop  [9:13) bbox umin,vmin,umax,vmax  [13:29) P@glcam@pose
 *        [29:33) render zoom tf00,tf11,tf30,tf31  [33:37) inverse i00,i02,i11,i12 ---- */
#define FPB_CROP_STRIDE 40
int fpb_crop_params(const float* poses, int N, const float* K9_h, float radius, int S, int H, int W,
                    const float* P16_h, float* params, void* stream);

/* ---- mesh handle: make_mesh_tensors src/Utils.py:104-130 (device arrays are COPIED) ---- */
typedef struct fpb_mesh fpb_mesh;
int fpb_mesh_create(fpb_mesh** out, const float* pos, const float* vnormals, int V, const int32_t* faces,
                    int F, const float* vcolor, const float* uv, int Vt, const int32_t* uv_idx,
                    const float* tex, int Ht, int Wt, void* stream);
int fpb_mesh_destroy(fpb_mesh* m);

/* ---- batched render at crop resolution + fused shading into network input A:
 *      nvdiffrast_render src/Utils.py:133-219 + rgb*255 (predict_pose_refine.py:56) +
 *      transform_batch (h5_dataset.py:79-127 / 137-179).
 *      ws must hold fpb_render_ws_bytes(N,V,S) bytes.  out_A: [N,S,out_pitch,8] bf16 or fp32
 *      (a_is_bf16); crop column j is written at pixel column out_xoff + j (the networks take
 *      out_pitch = FPB_IN_PITCH, out_xoff = FPB_IN_XOFF; pad pixels are never written).  Optional fp32 debug outputs (may be NULL): color [N,S,S,3],
 *      xyz [N,S,S,3] (camera frame, un-normalised), tri [N,S,S] int32 (face id + 1). ---- */
size_t fpb_render_ws_bytes(int N, int V, int S);
int fpb_render_crops(const fpb_mesh* mesh, const float* poses, const float* params, int N, int S,
                     float mesh_diameter, float zthr, void* ws, void* out_A, int a_is_bf16, int out_pitch,
                     int out_xoff, float* out_color, float* out_xyz, int32_t* out_tri, void* stream);
/* the remaining arguments of nvdiffrast_render (src/Utils.py:133): use_light (0: unlit colour), w_ambient / w_diffuse
 * (:211; fpb_render_crops uses the hot path's 1 / 0.8 / 0.5 with the default head light light_dir = (0,0,1)) and the
 * normal_map output (:192-197, get_normal): out_normal [N,S,S,3] fp32 or NULL.  N is unbounded: batches above 65535
 * hypotheses run in chunks on the same workspace (fpb_render_ws_bytes clamps N accordingly; so does fpb_warp_crops). */
int fpb_render_crops_ex(const fpb_mesh* mesh, const float* poses, const float* params, int N, int S,
                        float mesh_diameter, float zthr, void* ws, void* out_A, int a_is_bf16, int out_pitch,
                        int out_xoff, float* out_color, float* out_xyz, int32_t* out_tri, int use_light,
                        float w_ambient, float w_diffuse, float* out_normal, void* stream);

/* ---- observed-side crops into network input B: kornia warp_perspective call sites
 *      predict_pose_refine.py:63,72; predict_score.py:89-90; h5_dataset.py:137-170.
 *      mode 0 refiner (nearest xyz_map), mode 1 scorer (depth crop->full->xyz->crop).
 *      rgb: uint8 [H,W,3] (rgb_is_u8) or fp32 [H,W,3] in 0..255. ---- */
int fpb_warp_crops(const void* rgb, int rgb_is_u8, const float* xyz_hw3, const float* depth_hw, int H, int W,
                   const float* poses, const float* params, const float* K9_h, int N, int S,
                   float mesh_diameter, int mode, void* out_B, int b_is_bf16, int out_pitch, int out_xoff,
                   void* stream);

/* ---- pose update: predict_pose_refine.py:195-231, egocentric_delta_pose_to_pose
 *      src/Utils.py:848-855, pytorch3d so3_exp_map.  tn3_h used when !normalize_xyz. ---- */
int fpb_pose_update(const float* poses, const float* trans, const float* rot, int N, float mesh_diameter,
                    float rot_normalizer, const float* tn3_h, int normalize_xyz, float* out_poses,
                    void* stream);
/* ---- the config-gated representations of the same update (predict_pose_refine.py:201-224):
 *      trans_rep FPB_TRANS_DEEPIM  : trans = (du, dv) in crop units + z ratio; needs the crop parameters of the
 *                                    iteration (fpb_crop_params output), K and the crop size S (= input_resize[0]);
 *      rot_rep   FPB_ROT_6D        : rot is [N,6], pytorch3d rotation_6d_to_matrix (Gram-Schmidt), transposed.
 *      FPB_TRANS_TRACKNET + FPB_ROT_AXIS_ANGLE is fpb_pose_update. ---- */
#define FPB_TRANS_TRACKNET 0
#define FPB_TRANS_DEEPIM 1
#define FPB_ROT_AXIS_ANGLE 0
#define FPB_ROT_6D 1
int fpb_pose_update_ex(const float* poses, const float* trans, const float* rot, int N, float mesh_diameter,
                       float rot_normalizer, const float* tn3_h, int normalize_xyz, int trans_rep, int rot_rep,
                       const float* crop_params, const float* K9_h, int S, float* out_poses, void* stream);
/* same, additionally storing what the reference keeps as PoseRefinePredictor.last_trans_update / last_rot_update
 * (predict_pose_refine.py:238-239): trans_delta [N,3] (after the normaliser / radius scaling) and rot_mat_delta
 * [N,3,3] (the transposed exponential map); either pointer may be NULL. */
int fpb_pose_update_deltas(const float* poses, const float* trans, const float* rot, int N, float mesh_diameter,
                           float rot_normalizer, const float* tn3_h, int normalize_xyz, int trans_rep, int rot_rep,
                           const float* crop_params, const float* K9_h, int S, float* out_poses,
                           float* out_trans_delta, float* out_rot_delta, void* stream);

/* ---- ranking tail of register(): src/estimater.py:221-236 -- ids = argsort(scores, descending) (equal scores keep
 *      their input order), out_scores = scores[ids], out_poses = poses[ids], out_best16 = poses[ids[0]] @ tf16
 *      (tf16 = get_tf_to_centered_mesh(), device; out_best16 may be NULL).  N <= 4096 (one block, bitonic sort). ---- */
int fpb_rank_poses(const float* scores, const float* poses, const float* tf16, int N, long long* out_ids,
                   float* out_scores, float* out_poses, float* out_best16, void* stream);

/* ---- networks: learning/models/refine_network.py:26-93, score_network.py:27-90.
 *      Weights arrive as a flat host blob produced by foundationpose_b200.nets.pack_* (BN folded). ---- */
typedef struct fpb_net fpb_net;
#define FPB_NET_REFINE 0
#define FPB_NET_SCORE 1
#define FPB_NET_REFINE_6D 2   /* RefineNet built with rot_rep='6d' (refine_network.py:61-64): out_rot is [N,6] */
int fpb_net_create(fpb_net** out, int kind, int precision, const void* blob_h, size_t bytes, void* stream);
int fpb_net_destroy(fpb_net* net);
size_t fpb_net_ws_bytes(int kind, int precision, int N);
size_t fpb_net_blob_floats(int kind);              /* expected blob length (host-only helper) */
size_t fpb_score_cross_ws_bytes(int L, int bs);
/* AB: [2N,S,FPB_IN_PITCH,8] (first N = rendered A, last N = observed B), S = 160: rows are padded to
 * FPB_IN_PITCH pixels, image pixels at columns [FPB_IN_XOFF, FPB_IN_XOFF+S), pad pixels must be zero. */
#define FPB_IN_PITCH 168
#define FPB_IN_XOFF 4
/* fpb_refine_forward is stream-ordered on `stream` for the caller; internally the rot head runs on a private stream of
 * the net beside the trans head (forked from / joined to `stream` by events, legal under stream capture).  One net (and
 * workspace) per concurrently used stream / host thread. */
int fpb_refine_forward(const fpb_net* net, const void* AB, int N, void* ws, float* out_trans, float* out_rot,
                       void* stream);
/* same network, AB = [N+1,S,FPB_IN_PITCH,8]: N rendered crops + ONE observed crop shared by all N hypotheses.  In the
 * first refine iteration of register() every hypothesis has the guessed translation (src/estimater.py:127-134), hence
 * the same box_3d crop window (src/Utils.py:577-621), hence an identical observed crop B: the reference encodes it N
 * times (refine_network.py:78-80), here encodeA runs on N+1 images and B's features are broadcast.  Results are
 * bit-identical to fpb_refine_forward on the replicated input. */
int fpb_refine_forward_shared_b(const fpb_net* net, const void* AB, int N, void* ws, float* out_trans, float* out_rot,
                                void* stream);
int fpb_score_features(const fpb_net* net, const void* AB, int N, void* ws, float* out_feat_n512, void* stream);
/* ---- multi-GPU (new; the reference has no distributed code): hypotheses sharded over the GPUs of one box.
 *      ScoreNet's cross-hypothesis attention (score_network.py:86) needs all N feature rows on every rank, so the
 *      pooling kernel that produces a rank's [N_r,512] rows stores them straight into every rank's feature table
 *      (peer-mapped symmetric memory, stores travel over NVLink / NVSwitch): the all-gather is fused into the
 *      producing kernel.  out_feat_ptrs_h[i] = destination i already offset to this rank's first row (host array
 *      of n_dst <= FPB_MAX_PEERS device pointers; entry 0 is normally the local table).  fpb_store_multi_f32 does
 *      the same for the [N_r,16] refined poses.  The caller orders the exchange with a device barrier. ---- */
#define FPB_MAX_PEERS 8
int fpb_score_features_multi(const fpb_net* net, const void* AB, int N, void* ws, float* const* out_feat_ptrs_h,
                             int n_dst, void* stream);
int fpb_store_multi_f32(const float* src, long long count, float* const* dst_ptrs_h, int n_dst, void* stream);
int fpb_score_cross(const fpb_net* net, const float* feat, int L, int bs, void* ws, float* out_logits,
                    void* stream);

/* ---- fpb_score_cross without any allocation or caller workspace.  NO COUNTERPART IN THE REFERENCE: added
 *      for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 *      loops, harnesses that forbid it).  Same kernels and results as fpb_score_cross(net, feat, L, 1, ws, ...); the
 *      scratch is a file-scope __device__ array of 5,243,904 bytes (= fpb_score_cross_ws_bytes(256, 1)) that every GPU
 *      loading the library holds until exit.  Limit: L <= 256 hypotheses, one group (bs = 1); larger L returns
 *      FPB_ERR_UNSUPPORTED with a message.  Calls no cudaMalloc / cudaFree / allocating library routine; the first
 *      call on a device resolves the array's device address (cudaGetSymbolAddress).  Every call shares the array:
 *      it is cleared (stream-ordered) at the start of each call, and calls that overlap on the GPU -- different
 *      streams, or different threads -- RACE on it; serialise them. ---- */
int fpb_score_cross_static(const fpb_net* net, const float* feat, int L, float* out_logits, void* stream);

/* unit-test entry: one implicit GEMM C = epi(im2col(A) . W^T); A is NHWC [*,Hin,Win,Cin], W is [N][K], K = KH*KW*Cin.
 * precision/mode 0: SIMT fp32; 1: tcgen05 bf16 with cp.async producers; 2 / 3: persistent tcgen05 + TMA (im2col
 * tensor maps) kernel with cta_group 1 / 2.  _ex adds the row-padded "wide pixel" addressing of the 7x7 stem. */
int fpb_debug_igemm(int precision, const void* A, const void* W, void* C, int M, int N, int K, int Hin, int Win,
                    int Cin, int Ho, int Wo, int KH, int KW, int stride, int pad, const float* scale,
                    const float* shift, const void* residual, int relu, int out_bf16, void* stream);

int fpb_debug_igemm_ex(int mode, const void* A, const void* W, void* C, int M, int N, int K, int Hin, int Win, int Cin,
                       int Ho, int Wo, int KH, int KW, int stride, int pad_h, int pad_w, int row_pitch, int pix_pitch,
                       const float* scale, const float* shift, const void* residual, int relu, int out_bf16,
                       void* stream);

/* unit-test entry: 3x3 stride-1 convolution in the zero-bordered "halo" activation layout (positions p =
 * img*(H+1)*(W+1) + (y+1)*(W+1) + x; see csrc/net_kernels.cuh) and the dense<->position output row remapping. */
int fpb_debug_conv3x3_halo(int cta_group, const void* A, const void* W, void* C, int nimg, int H, int Wd, int Cin, int Cout,
                           int halo, int row_remap, const float* scale, const float* shift, const void* residual,
                           int relu, void* stream);

/* unit-test entry: fused bf16 multi-head self-attention over groups x 400 tokens (4 heads x 128):
 * qk [R][1024] = Q | K projections, vt [512][vt_pitch] = V^T (row h*128+d, column = token row), bv fp32 [512] V bias,
 * ctx [R][512] out;  R = groups*400.  Replaces nn.MultiheadAttention's core (refine_network.py:56-70). */
int fpb_debug_attention(const void* qk, const void* vt, long long vt_pitch, const float* bv, void* ctx, int groups,
                        void* stream);

/* unit-test entry: row-tile kernels of the encoder-layer tail (512-wide rows, csrc/net_ffn.cu).
 * mode 0: out[M][512] (16-bit) = LayerNorm(A . W2^T + b2 + residual) * gamma + beta   (out_proj + norm1,
 *         refine_network.py:56-70 / nn.TransformerEncoderLayer);
 * mode 1: out[M/tokens][512] (fp32) = mean over each group's tokens of LayerNorm(A + relu(A . W1^T + b1) . W2^T + b2)
 *         * gamma + beta   (linear1 + ReLU + linear2 + norm2 + `.mean(dim=1)`, refine_network.py:88-93);
 *         scratch: (M/16) x 512 floats, tokens % 16 == 0. */
int fpb_debug_ffn(int mode, const void* A, const void* W1, const float* b1, const void* W2, const float* b2,
                  const void* residual, const float* gamma, const float* beta, void* out, float* scratch, long long M,
                  int tokens, void* stream);

/* dev tool: clock64 stamps of the epilogue phases recorded by the last row-tile kernel launched with FPB_FFN_DBG=2048 */
int fpb_debug_ffn_stamps(long long* out_host, int n);

/* operand type (FPB_PREC_BF16 default, or FPB_PREC_FP16) the fpb_debug_* unit-test entries run in */
int fpb_debug_set_precision(int precision);

/* number of kernels this library launched on the calling thread since the last reset */
long long fpb_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif
