// Launch-level interface between net.cu (orchestration) and the kernel translation units
// net_ref.cu (SIMT fp32 parity kernels + small ops) and net_tc.cu (tcgen05 bf16 implicit GEMM).
#pragma once
#include "common.cuh"
#include <string.h>

// One implicit-GEMM problem:  C[m][n] = epi( sum_k A(m,k) * W[n][k] )
//   A(m,k): m -> (img, oy, ox) over Ho x Wo outputs; k -> (ky, kx, c) with c fastest;
//           element = in[((img*Hin + oy*stride-pad+ky)*row_pitch + ox*stride-pad_w+kx)*pix_pitch + c]  (0 outside
//           0 <= iy < Hin, 0 <= ix < Win); row_pitch (pixels per image row, 0 -> Win) and a separate, possibly
//           negative pad_w let the 7x7 stem read "wide pixels" (8 taps x 8 channels = 64 contiguous elements)
//           out of a row-padded image.
//   plain GEMM: Hin=Win=Ho=Wo=1, KH=KW=1, Cin=K, pix_pitch = lda.
//   W: [Ncols][K] row-major (K contiguous).
// Batched: blockIdx.z = b;  X_off = (b / b_inner) * X_outer + (b % b_inner) * X_inner   (elements)
// Epilogue: v = acc*alpha; v = v*scale[n]+shift[n] (scale may be null -> v += shift[n], shift may be null);
//           v += residual[m*ldr + n]; relu; v += pe[(m % pe_rows)*pe_ld + n];
//           out[(m % m_split) * ldc + (m / m_split) * split_col_off + n]
struct GemmProblem {
  const void* A;      // fp32 (ref) or bf16 (tc)
  const void* W;      // fp32 (ref) or bf16 (tc)
  void* C;            // fp32 or bf16 (out_bf16)
  int M, N, K;
  int Hin, Win, Cin, Ho, Wo, KH, KW, stride, pad, pad_w, row_pitch, pix_pitch;
  int ldw;            // row pitch of W in elements (>= K)
  int batch, b_inner;
  long long A_outer, A_inner, W_outer, W_inner, C_outer, C_inner;
  float alpha;
  const float* scale;
  const float* shift;
  const void* residual;  // same dtype as A-side activations (fp32 ref / bf16 tc)
  int ldr;
  int relu;
  const float* pe;
  int pe_rows, pe_ld;
  int ldc, m_split, split_col_off;
  int out_bf16;
  int n_valid;        // columns >= n_valid are not stored (N padding)
  // "halo" 3x3 stride-1 convolution over a zero-bordered activation (tc2 kernel only).  Positions p = img*IS +
  // (y+1)*Wp + x with Wp = img_w+1, IS = (img_h+1)*Wp: every image row is followed by one zero pixel and every image
  // is preceded by one zero row, so filter tap (ky,kx) of 128 consecutive positions is the same 128 rows of the
  // flattened [positions][C] matrix shifted by (ky-1)*Wp + (kx-1): ONE TMA box per 64-channel chunk serves all 9
  // taps.  M counts positions; outputs at border positions are written as zeros.
  int halo, img_h, img_w;
  int c_transposed;   // tc3 kernel only: out[n * ldc + m] (bf16), i.e. the product is stored transposed
  const float* row_shift;   // optional per-ROW additive term (after scale/shift; tc2 kernel only): v += row_shift[m]
  int row_remap;      // output rows: 0 as computed, 1 dense pixel index -> position, 2 position -> dense (borders skipped)
  // optional scratch for split-K (tc2 kernel only): with few output tiles (single-hypothesis batches: track_one) the K
  // loop is divided over ksplit CTAs per tile, fp32 partial tiles land here ([ksplit][M][N]) and a second kernel sums
  // them in a fixed order and applies the epilogue
  void* splitk_ws;
  long long splitk_ws_bytes;
};

static inline GemmProblem gemm_problem_default()
{
  GemmProblem p;
  memset(&p, 0, sizeof(p));
  p.Hin = p.Win = p.Ho = p.Wo = p.KH = p.KW = p.stride = 1;
  p.batch = 1; p.b_inner = 1;
  p.alpha = 1.0f;
  p.m_split = 0x7fffffff;
  return p;
}

struct MultiDst { float* p[FPB_MAX_PEERS]; int n; };

// Everything below exists twice: the network translation units are compiled once with bf16 operands (namespace bf16)
// and once with fp16 operands (namespace f16), see act_type.cuh.  "bf16" in the flag names below (out_bf16, x_bf16,
// is_bf16) means "the 16-bit activation type of this namespace" as opposed to fp32.
#define FPB_NET_KERNEL_DECLS \
/* 1 unless FPB_PDL=0: launch the tcgen05 kernels with programmatic stream serialization (see tc_ptx.cuh) */ \
int fpb_pdl_enabled(); \
 \
int ref_gemm(const GemmProblem& p, cudaStream_t st); /* fp32 SIMT */ \
int tc_gemm(const GemmProblem& p, cudaStream_t st); /* bf16 tcgen05, fp32 accumulate (cp.async producers) */ \
int tc2_gemm(const GemmProblem& p, cudaStream_t st); /* bf16 tcgen05, persistent TMA kernel; FPB_ERR_UNSUPPORTED -> use tc_gemm */ \
int tc3_gemm(const GemmProblem& p, cudaStream_t st, int force); /* bf16 tcgen05, operand-swapped (Cout 64 / 128); FPB_ERR_UNSUPPORTED -> tc2_gemm */ \
/* dedicated 7x7 stride-2 stem kernel (net_stem.cu): row-padded input [nimg,160,168,8] bf16 -> [nimg*6400,64] bf16 */ \
int stem_pack_weights(const void* w_packed, int is_bf16, void* w_stacked, cudaStream_t st); /* [64][7][8][8] -> [9][128][64] bf16 */ \
int stem_conv(const void* in, int nimg, const void* w_stacked, const float* scale, const float* shift, int relu, void* out, \
              cudaStream_t st, int force); \
void tc2_force_cta_group(int cg); /* tests: 0 = default, 1 / 2 = force */ \
 \
int fpb_tiled_map_act(const void* ptr, long long rows, long long cols, long long pitch, int box_rows, void* out_map, int* num_sms); \
int fpb_im2col_map_act(const GemmProblem& p, int nimg, void* out_map); /* 128 output pixels x 64 channels per request */ \
/* fused multi-head self-attention (bf16, tcgen05): qk [R][1024] (Q | K, head h at +128h), vt [512][vt_pitch] */ \
/* (V transposed: row = h*128+d, column = token row), bv [512] fp32 V bias -> ctx [R][512];  R = groups * 400 */ \
int tc_attention(const void* qk, const void* vt, long long vt_pitch, const float* bv, void* ctx, int groups, cudaStream_t st); \
 \
/* small ops (fp32 math; io dtype flags) */ \
int softmax_rows(const float* in, void* out, int out_bf16, long long rows, int cols, int ld_in, int ld_out, \
                 cudaStream_t st); /* pad columns [cols, ld_out) written as 0 */ \
int layernorm_rows(const void* x, int x_bf16, const float* gamma, const float* beta, void* out, int out_bf16, \
                   long long rows, int cols, float eps, cudaStream_t st); \
int token_mean(const void* x, int x_bf16, float* out, int groups, int tokens, int cols, cudaStream_t st); \
/* the same rows stored to up to FPB_MAX_PEERS destinations (local and peer-mapped buffers: the all-gather of the sharded */ \
/* scorer fused into the pooling kernel, see fpb_score_features_multi) */ \
int token_mean_multi(const void* x, int x_bf16, const MultiDst& out, int groups, int tokens, int cols, cudaStream_t st); \
int layernorm_token_mean_512(const void* x, const float* gamma, const float* beta, float* out, int groups, int tokens, float eps, \
                             cudaStream_t st); /* bf16 rows of 512: mean over tokens of LayerNorm(x), fused */ \
int small_linear(const float* x, const float* W, const float* b, float* out, int rows, int K, int N, \
                 cudaStream_t st); /* out[r][n] = x[r].W[n] + b[n] */ \
int transpose_v(const void* qkv, void* vt, int is_bf16, int groups, int tokens, int heads, int dh, int ld_qkv, int v_off, \
                int ld_vt, cudaStream_t st); /* vt[(g*heads+h)][d][t] */ \
int f32_to_act(const float* in, void* out, long long n, cudaStream_t st); /* fp32 -> act_t (bf16 / fp16) */ \
/* row-tile kernels of the encoder-layer tail (net_ffn.cu); FPB_ERR_UNSUPPORTED (FPB_FFN=0, misaligned) -> unfused path */ \
/* y = LayerNorm(A . W^T + b + residual) * gamma + beta on [M][512] 16-bit rows */ \
int proj_ln_512(const void* A, const void* W, const float* b, const void* residual, const float* gamma, const float* beta, \
                float eps, void* out, long long M, cudaStream_t st); \
/* pooled[g] = mean_t LayerNorm(x + relu(x . W1^T + b1) . W2^T + b2) * gamma + beta; gs: (groups*tokens/16) x 512 floats scratch */ \
/* Wf != null: the head's output Linear on the pooled row in the same pooling launch: out[g][n] = pooled[g] . Wf[n] + bf[n] */ \
int ffn_ln_token_mean_512(const void* x, const void* W1, const float* b1, const void* W2, const float* b2, const float* gamma, \
                          const float* beta, float eps, float* gs, float* pooled, int groups, int tokens, const float* Wf, \
                          const float* bf, float* out, int n_out, cudaStream_t st); \
int ffn_debug_stamps(long long* out_host, int n); /* dev tool: clock64 stamps of the last FPB_FFN_DBG=2048 launch */ \

namespace bf16 { FPB_NET_KERNEL_DECLS }
namespace f16 { FPB_NET_KERNEL_DECLS }

// function table of one namespace, so that net.cu can pick the operand type at run time (fpb_net::precision)
struct NetKernels {
  decltype(&bf16::ref_gemm) ref_gemm;
  decltype(&bf16::tc_gemm) tc_gemm;
  decltype(&bf16::tc2_gemm) tc2_gemm;
  decltype(&bf16::tc3_gemm) tc3_gemm;
  decltype(&bf16::stem_pack_weights) stem_pack_weights;
  decltype(&bf16::stem_conv) stem_conv;
  decltype(&bf16::tc2_force_cta_group) tc2_force_cta_group;
  decltype(&bf16::tc_attention) tc_attention;
  decltype(&bf16::softmax_rows) softmax_rows;
  decltype(&bf16::layernorm_rows) layernorm_rows;
  decltype(&bf16::token_mean) token_mean;
  decltype(&bf16::token_mean_multi) token_mean_multi;
  decltype(&bf16::layernorm_token_mean_512) layernorm_token_mean_512;
  decltype(&bf16::small_linear) small_linear;
  decltype(&bf16::transpose_v) transpose_v;
  decltype(&bf16::f32_to_act) f32_to_act;
  decltype(&bf16::proj_ln_512) proj_ln_512;
  decltype(&bf16::ffn_ln_token_mean_512) ffn_ln_token_mean_512;
  decltype(&bf16::ffn_debug_stamps) ffn_debug_stamps;
};
#define FPB_NET_KERNEL_TABLE(ns)                                                                                        \
  {ns::ref_gemm, ns::tc_gemm, ns::tc2_gemm, ns::tc3_gemm, ns::stem_pack_weights, ns::stem_conv, ns::tc2_force_cta_group, \
   ns::tc_attention, ns::softmax_rows, ns::layernorm_rows, ns::token_mean, ns::token_mean_multi,                        \
   ns::layernorm_token_mean_512, ns::small_linear, ns::transpose_v, ns::f32_to_act, ns::proj_ln_512,              \
   ns::ffn_ln_token_mean_512, ns::ffn_debug_stamps}
