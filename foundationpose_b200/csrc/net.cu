// RefineNet / ScoreNetMultiPair forward orchestration over the GEMM kernels (net_ref.cu / net_tc.cu).
//   RefineNet.forward           learning/models/refine_network.py:73-93
//   ScoreNetMultiPair.forward   learning/models/score_network.py:60-90
// Weights arrive as one flat fp32 blob (BatchNorm folded into per-channel scale/shift by
// foundationpose_b200/nets.py); in bf16 mode a bf16 shadow of the blob feeds the tcgen05 kernel.
#include "net_kernels.cuh"
#include <stdlib.h>
#include <mutex>
#include <vector>

namespace {

struct ConvSpec { int Cin, Cout, KH, KW, stride, pad, Hin, Ho; };
// The 7x7 stride-2 stem reads the network input as [imgs,160,IN_PITCH,8] (6 channels padded to 8, IN_XOFF zero
// pixels left of each row, >= 4 right) so that the 8 taps x 8 channels of one kernel row are 64 CONTIGUOUS
// elements: it runs as a 7x1 convolution over 161 overlapping "wide pixels" of 64 channels (pad_w = -1).
constexpr int IN_S = 160, IN_PITCH = 168, IN_XOFF = 4, IN_C = 8;
const ConvSpec kConvs[15] = {
    {64, 64, 7, 1, 2, 3, 160, 80},   {64, 128, 3, 3, 2, 1, 80, 40},   {128, 128, 3, 3, 1, 1, 40, 40},
    {128, 128, 3, 3, 1, 1, 40, 40},  {128, 128, 3, 3, 1, 1, 40, 40},  {128, 128, 3, 3, 1, 1, 40, 40},
    {256, 256, 3, 3, 1, 1, 40, 40},  {256, 256, 3, 3, 1, 1, 40, 40},  {256, 256, 3, 3, 1, 1, 40, 40},
    {256, 256, 3, 3, 1, 1, 40, 40},  {256, 512, 3, 3, 2, 1, 40, 20},  {512, 512, 3, 3, 1, 1, 20, 20},
    {512, 512, 3, 3, 1, 1, 20, 20},  {512, 512, 3, 3, 1, 1, 20, 20},  {512, 512, 3, 3, 1, 1, 20, 20}};

constexpr int D = 512, HEADS = 4, DH = 128, TOK = 400, TOKP = 448;

struct Lin { size_t w, b; int N, K; };            // offsets (in floats) into the blob
struct Conv { size_t w, scale, shift; };
struct Mha { Lin in, out; };
struct EncLayer { Mha att; Lin l1, l2; size_t n1g, n1b, n2g, n2b; Lin fin; };

}  // namespace

struct fpb_net {
  int kind, precision;
  int rot_dim;        // refine: 3 (axis_angle) or 6 (rot_rep='6d', refine_network.py:61-64)
  size_t n_floats;
  float* w32;
  uint16_t* w16;           // 16-bit shadow of the blob in the operand type of `precision` (bf16 or fp16)
  uint16_t* w_stem;        // stem weights stacked for net_stem.cu ([9][128][64], 16-bit modes only)
  cudaStream_t st2;        // refine, 16-bit modes: the rot head runs on this stream beside the trans head (fork / join by events)
  cudaEvent_t ev_fork, ev_join;
  // 16-bit modes: the V^T projection of a self-attention block runs beside its Q|K projection (both only read the tokens);
  // chain 0 = caller's stream (+ the ScoreNet head), chain 1 = st2
  cudaStream_t st_v[2];
  cudaEvent_t ev_vf[2], ev_vj[2];
  Conv conv[15];
  size_t pe;
  EncLayer head[2];   // refine: trans, rot
  Mha att, att_cross; // score
  Lin lin;            // score
};

namespace {

struct Cursor {
  size_t off = 0;
  size_t take(size_t n) { size_t o = off; off += (n + 7) & ~(size_t)7; return o; }   // keep 8-element alignment
};

Lin take_lin(Cursor& c, int N, int K) { Lin l; l.N = N; l.K = K; l.w = c.take((size_t)N * K); l.b = c.take(N); return l; }
Mha take_mha(Cursor& c) { Mha m; m.in = take_lin(c, 3 * D, D); m.out = take_lin(c, D, D); return m; }

size_t layout(fpb_net* n)
{
  Cursor c;
  for (int i = 0; i < 15; i++) {
    const ConvSpec& s = kConvs[i];
    n->conv[i].w = c.take((size_t)s.Cout * s.KH * s.KW * s.Cin);   // stem: 64 x 7 x (8 taps x 8 ch)
    n->conv[i].scale = c.take(s.Cout);
    n->conv[i].shift = c.take(s.Cout);
  }
  n->pe = c.take((size_t)TOK * D);
  if (n->kind == FPB_NET_REFINE) {
    for (int h = 0; h < 2; h++) {
      EncLayer& e = n->head[h];
      e.att = take_mha(c);
      e.l1 = take_lin(c, D, D);
      e.l2 = take_lin(c, D, D);
      e.n1g = c.take(D); e.n1b = c.take(D); e.n2g = c.take(D); e.n2b = c.take(D);
      e.fin = take_lin(c, h == 1 ? n->rot_dim : 3, D);
    }
  } else {
    n->att = take_mha(c);
    n->att_cross = take_mha(c);
    n->lin = take_lin(c, 1, D);
  }
  return c.off;
}

int net_chunk()
{
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("FPB_NET_CHUNK");
    v = e ? atoi(e) : 256;
    if (v < 1) v = 1;
  }
  return v;
}

size_t al(size_t x) { return (x + 255) & ~(size_t)255; }

// workspace carve-up for a chunk of n hypotheses; esz = activation element size
struct Ws {
  char *c0, *c1, *c2, *c3, *ab0, *ab1, *ab2, *d0, *d1, *d2, *qkv, *P, *vt, *ctx, *y1, *y, *ff;
  float *scores, *pooled;
  char* splitk;       // split-K scratch of the persistent GEMM kernel (small batches only)
  char* splitk2;      // the same for the head that runs on the second stream
  size_t total;
};
constexpr size_t SPLITK_BYTES = 16u << 20;

Ws carve(char* base, int n, size_t esz)
{
  Ws w;
  size_t o = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + o : nullptr; o += al(bytes); return p; };
  w.c0 = take((size_t)2 * n * 6400 * 64 * esz);
  w.c1 = take((size_t)2 * n * 1681 * 128 * esz);      // 41 x 41 zero-bordered positions per image (halo layout)
  w.c2 = take((size_t)2 * n * 1681 * 128 * esz);
  w.c3 = take((size_t)2 * n * 1681 * 128 * esz);
  w.ab0 = take((size_t)n * 1600 * 256 * esz);
  w.ab1 = take((size_t)n * 1600 * 256 * esz);
  w.ab2 = take((size_t)n * 1600 * 256 * esz);
  w.d0 = take((size_t)n * TOK * D * esz);
  w.d1 = take((size_t)n * TOK * D * esz);
  w.d2 = take((size_t)n * TOK * D * esz);
  w.qkv = take((size_t)n * TOK * 3 * D * esz);
  w.P = take((size_t)n * HEADS * TOK * TOKP * esz);
  w.vt = take((size_t)n * HEADS * DH * TOKP * esz);
  w.ctx = take((size_t)n * TOK * D * esz);
  w.y1 = take((size_t)n * TOK * D * esz);
  w.y = take((size_t)n * TOK * D * esz);
  w.ff = take((size_t)n * TOK * D * esz);
  w.scores = (float*)take((size_t)n * HEADS * TOK * TOKP * sizeof(float));
  w.pooled = (float*)take((size_t)n * D * sizeof(float));
  w.splitk = take(SPLITK_BYTES);
  w.splitk2 = take(SPLITK_BYTES);
  w.total = o;
  return w;
}

const NetKernels kKernelsBf16 = FPB_NET_KERNEL_TABLE(bf16), kKernelsF16 = FPB_NET_KERNEL_TABLE(f16);
inline const NetKernels* kernels_of(int precision) { return precision == FPB_PREC_FP16 ? &kKernelsF16 : &kKernelsBf16; }
inline bool is16(int precision) { return precision == FPB_PREC_BF16 || precision == FPB_PREC_FP16; }

struct Ctx {
  const fpb_net* net;
  cudaStream_t st;
  bool bf;                    // 16-bit tensor-core mode (bf16 or fp16 operands); false = fp32 SIMT parity mode
  size_t esz;
  const NetKernels* k;        // kernels of the operand type (fp32 mode: either table, the SIMT kernels are the same)
  void* splitk = nullptr;     // set per chunk from the workspace (Ws::splitk)
  int chain = 0;              // which of the net's V^T side streams this context may fork onto
  const void* W(size_t off) const { return bf ? (const void*)(net->w16 + off) : (const void*)(net->w32 + off); }
  const float* F(size_t off) const { return net->w32 + off; }
  int gemm(const GemmProblem& p_in) const
  {
    if (!bf) return k->ref_gemm(p_in, st);
    GemmProblem p = p_in;
    p.splitk_ws = splitk;
    p.splitk_ws_bytes = splitk ? (long long)SPLITK_BYTES : 0;
    int rc = k->tc3_gemm(p, st, 0);                          // narrow layers (Cout 64 / 128): operand-swapped kernel
    if (rc == FPB_ERR_UNSUPPORTED) rc = k->tc2_gemm(p, st);  // persistent TMA kernel
    return rc == FPB_ERR_UNSUPPORTED ? k->tc_gemm(p, st) : rc;  // batched / odd shapes: cp.async kernel
  }
};

#define RET(x) do { int rc__ = (x); if (rc__) return rc__; } while (0)

bool head_streams()
{
  static int v = -1;
  if (v < 0) { const char* e = getenv("FPB_HEAD_STREAMS"); v = e ? atoi(e) : 1; }
  return v != 0;
}

bool use_halo()
{
  static int v = -1;
  if (v < 0) { const char* e = getenv("FPB_HALO"); v = e ? atoi(e) : 0; }   // measured slower than im2col at one tap per stage: off
  return v != 0;
}

// in_pos / out_pos: activation in the zero-bordered position layout (GemmProblem::halo) instead of dense NHWC
int run_conv(const Ctx& c, int li, const void* in, int nimg, void* out, const void* residual, int relu, bool pe,
             int m_split, int split_off, int ldc, int in_pos = 0, int out_pos = 0)
{
  const ConvSpec& s = kConvs[li];
  GemmProblem p = gemm_problem_default();
  p.A = in; p.W = c.W(c.net->conv[li].w); p.C = out;
  p.M = nimg * s.Ho * s.Ho; p.N = s.Cout; p.K = s.KH * s.KW * s.Cin;
  if (in_pos || out_pos) {
    p.img_h = p.img_w = s.Ho;
    const int IS = (s.Ho + 1) * (s.Ho + 1);
    if (in_pos) { p.halo = 1; p.M = nimg * IS; p.row_remap = out_pos ? 0 : 2; }
    else {
      p.row_remap = 1;    // dense rows -> positions: the borders are never written, clear them first
      FPB_CUDA(cudaMemsetAsync(out, 0, (size_t)nimg * IS * s.Cout * c.esz, c.st));
    }
  }
  p.Hin = p.Win = s.Hin; p.Cin = s.Cin; p.Ho = p.Wo = s.Ho; p.KH = s.KH; p.KW = s.KW; p.stride = s.stride;
  p.pad = p.pad_w = s.pad;
  p.pix_pitch = s.Cin; p.ldw = p.K;
  if (li == 0) { p.Win = IN_PITCH - 7; p.row_pitch = IN_PITCH; p.pix_pitch = IN_C; p.pad_w = 3 - IN_XOFF; }
  p.scale = c.F(c.net->conv[li].scale); p.shift = c.F(c.net->conv[li].shift);
  p.residual = residual; p.ldr = s.Cout; p.relu = relu;
  if (pe) { p.pe = c.F(c.net->pe); p.pe_rows = TOK; p.pe_ld = D; }
  p.ldc = ldc > 0 ? ldc : s.Cout;
  if (m_split > 0) { p.m_split = m_split; p.split_col_off = split_off; }
  p.out_bf16 = c.bf;
  return c.gemm(p);
}

int run_linear(const Ctx& c, const Lin& l, const void* in, long long rows, void* out, const void* residual, int relu,
               int out_bf16)
{
  GemmProblem p = gemm_problem_default();
  p.A = in; p.W = c.W(l.w); p.C = out;
  p.M = (int)rows; p.N = l.N; p.K = l.K; p.Cin = l.K; p.pix_pitch = l.K; p.ldw = l.K;
  p.shift = c.F(l.b);
  p.residual = residual; p.ldr = l.N; p.relu = relu; p.ldc = l.N; p.out_bf16 = out_bf16;
  return c.gemm(p);
}

// shared-B mode: the observed-side features of hypothesis 0 (channels [128,256) of its 1600 rows, `seg` bytes per row at
// byte offset `off` inside a row of `pitch` bytes) are copied to hypotheses 1..n-1
__global__ void __launch_bounds__(256) broadcast_b_kernel(char* __restrict__ ab, int n, int rows, int pitch, int off, int seg)
{
  const int per_row = seg / 16;
  const long long total = (long long)(n - 1) * rows * per_row;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(i % per_row);
    const long long t = i / per_row;
    const int r = (int)(t % rows);
    const int h = (int)(t / rows) + 1;
    const uint4 v = *reinterpret_cast<const uint4*>(ab + (size_t)r * pitch + off + q * 16);
    *reinterpret_cast<uint4*>(ab + ((size_t)h * rows + r) * pitch + off + q * 16) = v;
  }
}

// both encoders: AB [2n,160,160,8] -> tokens (+pe) in w.d0 [n*400, 512].
// shared_b: AB is [n+1,...] = n rendered crops + ONE observed crop shared by all n hypotheses (first refine iteration of
// register(): every hypothesis has the same translation, hence the same crop window, hence the same observed crop):
// encodeA runs on n+1 images and the observed-side features are broadcast -- identical results, ~half of encodeA saved.
int run_encoders(const Ctx& c, const void* AB, int n, Ws& w, bool shared_b = false)
{
  const int nA = shared_b ? n + 1 : 2 * n;
  int rc0 = FPB_ERR_UNSUPPORTED;
  if (c.bf && c.net->w_stem)      // dedicated stem kernel: input rows loaded once, two output rows stacked in the MMA M dimension
    rc0 = c.k->stem_conv(AB, nA, c.net->w_stem, c.F(c.net->conv[0].scale), c.F(c.net->conv[0].shift), 1, w.c0, c.st, 0);
  if (rc0 == FPB_ERR_UNSUPPORTED) rc0 = run_conv(c, 0, AB, nA, w.c0, nullptr, 1, false, 0, 0, 0);
  RET(rc0);
  // optional (FPB_HALO=1): 3x3 stride-1 blocks in the zero-bordered position layout (one TMA box per 64 channels feeds all
  // 9 taps); kept for the record with its tests -- per-stage pipeline overhead makes it slower than the im2col path today
  const int hp = (c.bf && use_halo()) ? 1 : 0;
  RET(run_conv(c, 1, w.c0, nA, w.c1, nullptr, 1, false, 0, 0, 0, 0, hp));
  RET(run_conv(c, 2, w.c1, nA, w.c2, nullptr, 1, false, 0, 0, 0, hp, hp));
  RET(run_conv(c, 3, w.c2, nA, w.c3, w.c1, 1, false, 0, 0, 0, hp, hp));
  RET(run_conv(c, 4, w.c3, nA, w.c2, nullptr, 1, false, 0, 0, 0, hp, hp));
  // last encodeA conv writes cat((a,b),1): rows of the first n images -> channels [0,128), last n -> [128,256)
  RET(run_conv(c, 5, w.c2, nA, w.ab0, w.c3, 1, false, n * 1600, 128, 256, hp, 0));
  if (shared_b && n > 1) {      // the single B image landed in hypothesis 0's rows: copy to the others
    const int pitch = 256 * (int)c.esz, seg = 128 * (int)c.esz;
    broadcast_b_kernel<<<148 * 8, 256, 0, c.st>>>((char*)w.ab0, n, 1600, pitch, seg, seg);
    FPB_LAUNCH_CHECK();
  }
  RET(run_conv(c, 6, w.ab0, n, w.ab1, nullptr, 1, false, 0, 0, 0));
  RET(run_conv(c, 7, w.ab1, n, w.ab2, w.ab0, 1, false, 0, 0, 0));
  RET(run_conv(c, 8, w.ab2, n, w.ab1, nullptr, 1, false, 0, 0, 0));
  RET(run_conv(c, 9, w.ab1, n, w.ab0, w.ab2, 1, false, 0, 0, 0));
  RET(run_conv(c, 10, w.ab0, n, w.d0, nullptr, 1, false, 0, 0, 0));
  RET(run_conv(c, 11, w.d0, n, w.d1, nullptr, 1, false, 0, 0, 0));
  RET(run_conv(c, 12, w.d1, n, w.d2, w.d0, 1, false, 0, 0, 0));
  RET(run_conv(c, 13, w.d2, n, w.d1, nullptr, 1, false, 0, 0, 0));
  RET(run_conv(c, 14, w.d1, n, w.d0, w.d2, 1, true, 0, 0, 0));   // + positional embedding -> tokens in d0
  return FPB_OK;
}

// multi-head self-attention core on packed qkv [g*tokens, 3*D] -> ctx [g*tokens, D]
int run_attention_core(const Ctx& c, const void* qkv, int g, int tokens, int tokp, float* scores, void* P, void* vt,
                       void* ctx, bool bf)
{
  const size_t esz = bf ? 2 : 4;
  RET(c.k->transpose_v(qkv, vt, bf, g, tokens, HEADS, DH, 3 * D, 2 * D, tokp, c.st));
  GemmProblem p = gemm_problem_default();
  p.A = qkv; p.W = (const char*)qkv + (size_t)D * esz; p.C = scores;
  p.M = tokens; p.N = tokens; p.K = DH; p.Cin = DH; p.pix_pitch = 3 * D; p.ldw = 3 * D;
  p.batch = g * HEADS; p.b_inner = HEADS;
  p.A_outer = (long long)tokens * 3 * D; p.A_inner = DH; p.W_outer = p.A_outer; p.W_inner = DH;
  p.C_outer = (long long)HEADS * tokens * tokp; p.C_inner = (long long)tokens * tokp; p.ldc = tokp;
  p.alpha = 0.08838834764831845f;   // 1/sqrt(128)
  p.n_valid = tokens; p.out_bf16 = 0;
  RET(bf ? c.k->tc_gemm(p, c.st) : c.k->ref_gemm(p, c.st));
  RET(c.k->softmax_rows(scores, P, bf, (long long)g * HEADS * tokens, tokens, tokp, tokp, c.st));
  GemmProblem q = gemm_problem_default();
  q.A = P; q.W = vt; q.C = ctx;
  q.M = tokens; q.N = DH; q.K = tokp; q.Cin = tokp; q.pix_pitch = tokp; q.ldw = tokp;
  q.batch = g * HEADS; q.b_inner = HEADS;
  q.A_outer = (long long)HEADS * tokens * tokp; q.A_inner = (long long)tokens * tokp;
  q.W_outer = (long long)HEADS * DH * tokp; q.W_inner = (long long)DH * tokp;
  q.C_outer = (long long)tokens * D; q.C_inner = DH; q.ldc = D; q.out_bf16 = bf;
  RET(bf ? c.k->tc_gemm(q, c.st) : c.k->ref_gemm(q, c.st));
  return FPB_OK;
}

// self-attention block input x [n*400, 512] -> ctx [n*400, 512] (before out_proj)
int run_self_attention(const Ctx& c, const Mha& m, const void* x, int n, Ws& w)
{
  const long long R = (long long)n * TOK;
  if (!c.bf) {
    RET(run_linear(c, m.in, x, R, w.qkv, nullptr, 0, 0));
    return run_attention_core(c, w.qkv, n, TOK, TOKP, w.scores, w.P, w.vt, w.ctx, false);
  }
  // bf16: Q|K projection [R,1024]; V is produced already transposed, V^T = Wv . X^T  ([512, R], tokens contiguous),
  // by swapping the GEMM operands, so the fused attention kernel can TMA it as a K-major B operand; the V bias is a
  // per-row term of that transposed product (row_shift).
  uint16_t* qk = (uint16_t*)w.qkv;
  uint16_t* vt = qk + R * 2 * D;
  Lin lqk = m.in; lqk.N = 2 * D;
  // V^T on a side stream beside Q|K (FPB_HEAD_STREAMS): for one hypothesis the two projections are two 15 us latencies
  cudaStream_t sv = (c.net->st_v[c.chain] && head_streams()) ? c.net->st_v[c.chain] : nullptr;
  if (sv) {
    FPB_CUDA(cudaEventRecord(c.net->ev_vf[c.chain], c.st));
    FPB_CUDA(cudaStreamWaitEvent(sv, c.net->ev_vf[c.chain], 0));
  }
  RET(run_linear(c, lqk, x, R, qk, nullptr, 0, 1));
  {
    // preferred: V = X . Wv^T + bv through the operand-swapped kernel, which stores its (channel-major) accumulator
    // layout directly: vt[d][token]
    GemmProblem t = gemm_problem_default();
    t.A = x; t.W = c.W(m.in.w + (size_t)2 * D * D); t.C = vt;
    t.M = (int)R; t.N = D; t.K = D; t.Cin = D; t.pix_pitch = D; t.ldw = D; t.ldc = (int)R; t.out_bf16 = 1;
    t.shift = c.F(m.in.b + 2 * D); t.c_transposed = 1;
    const int rc3 = c.k->tc3_gemm(t, sv ? sv : c.st, 0);
    if (sv) {                                  // join (also when the kernel did not take the shape: nothing was enqueued then)
      FPB_CUDA(cudaEventRecord(c.net->ev_vj[c.chain], sv));
      FPB_CUDA(cudaStreamWaitEvent(c.st, c.net->ev_vj[c.chain], 0));
    }
    if (rc3 != FPB_ERR_UNSUPPORTED) {
      RET(rc3);
      return c.k->tc_attention(qk, vt, R, nullptr, w.ctx, n, c.st);
    }
  }
  GemmProblem p = gemm_problem_default();
  p.A = c.W(m.in.w + (size_t)2 * D * D); p.W = x; p.C = vt;
  p.M = D; p.N = (int)R; p.K = D; p.Cin = D; p.pix_pitch = D; p.ldw = D; p.ldc = (int)R; p.out_bf16 = 1;
  p.row_shift = c.F(m.in.b + 2 * D);
  RET(c.gemm(p));
  return c.k->tc_attention(qk, vt, R, nullptr, w.ctx, n, c.st);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
FPB_EXPORT int fpb_net_create(fpb_net** out, int kind, int precision, const void* blob_h, size_t bytes, void* stream)
{
  FPB_CHECK_ARG(out && blob_h && (kind == FPB_NET_REFINE || kind == FPB_NET_SCORE || kind == FPB_NET_REFINE_6D));
  FPB_CHECK_ARG(precision == FPB_PREC_FP32 || precision == FPB_PREC_BF16 || precision == FPB_PREC_FP16);
  cudaStream_t st = (cudaStream_t)stream;
  fpb_net* n = new fpb_net();
  memset(n, 0, sizeof(*n));
  n->kind = kind == FPB_NET_REFINE_6D ? FPB_NET_REFINE : kind; n->precision = precision;
  n->rot_dim = kind == FPB_NET_REFINE_6D ? 6 : 3;
  n->n_floats = layout(n);
  if (bytes != n->n_floats * sizeof(float)) {
    fpb_set_error("fpb_net_create: blob has %zu bytes, expected %zu", bytes, n->n_floats * sizeof(float));
    delete n;
    return FPB_ERR_ARG;
  }
  FPB_CUDA(cudaMalloc((void**)&n->w32, bytes));
  FPB_CUDA(cudaMemcpyAsync(n->w32, blob_h, bytes, cudaMemcpyHostToDevice, st));
  if (is16(precision)) {
    const NetKernels* k = kernels_of(precision);
    FPB_CUDA(cudaMalloc((void**)&n->w16, n->n_floats * sizeof(uint16_t)));
    int rc = k->f32_to_act(n->w32, n->w16, (long long)n->n_floats, st);
    if (rc) return rc;
    FPB_CUDA(cudaMalloc((void**)&n->w_stem, (size_t)9 * 128 * 64 * sizeof(uint16_t)));
    rc = k->stem_pack_weights(n->w32 + n->conv[0].w, 0, n->w_stem, st);
    if (rc) return rc;
  }
  if (n->kind == FPB_NET_REFINE && is16(precision)) {
    FPB_CUDA(cudaStreamCreateWithFlags(&n->st2, cudaStreamNonBlocking));
    FPB_CUDA(cudaEventCreateWithFlags(&n->ev_fork, cudaEventDisableTiming));
    FPB_CUDA(cudaEventCreateWithFlags(&n->ev_join, cudaEventDisableTiming));
  }
  if (is16(precision)) {
    for (int i = 0; i < (n->kind == FPB_NET_REFINE ? 2 : 1); i++) {
      FPB_CUDA(cudaStreamCreateWithFlags(&n->st_v[i], cudaStreamNonBlocking));
      FPB_CUDA(cudaEventCreateWithFlags(&n->ev_vf[i], cudaEventDisableTiming));
      FPB_CUDA(cudaEventCreateWithFlags(&n->ev_vj[i], cudaEventDisableTiming));
    }
  }
  FPB_CUDA(cudaStreamSynchronize(st));
  *out = n;
  return FPB_OK;
}

FPB_EXPORT int fpb_net_destroy(fpb_net* n)
{
  if (!n) return FPB_OK;
  cudaFree(n->w32);
  cudaFree(n->w16);
  cudaFree(n->w_stem);
  if (n->st2) { cudaStreamDestroy(n->st2); cudaEventDestroy(n->ev_fork); cudaEventDestroy(n->ev_join); }
  for (int i = 0; i < 2; i++)
    if (n->st_v[i]) { cudaStreamDestroy(n->st_v[i]); cudaEventDestroy(n->ev_vf[i]); cudaEventDestroy(n->ev_vj[i]); }
  delete n;
  return FPB_OK;
}

FPB_EXPORT size_t fpb_net_ws_bytes(int kind, int precision, int N)
{
  (void)kind;
  const int n = N < net_chunk() ? N : net_chunk();
  return carve(nullptr, n, is16(precision) ? 2 : 4).total + 1024;
}

FPB_EXPORT size_t fpb_net_blob_floats(int kind)
{
  fpb_net tmp;
  memset(&tmp, 0, sizeof(tmp));
  tmp.kind = kind == FPB_NET_REFINE_6D ? FPB_NET_REFINE : kind;
  tmp.rot_dim = kind == FPB_NET_REFINE_6D ? 6 : 3;
  return layout(&tmp);
}

static int encoder_layer(const Ctx& c, const EncLayer& e, int n, Ws& w, float* out3)
{
  const long long R = (long long)n * TOK;
  RET(run_self_attention(c, e.att, w.d0, n, w));
  if (c.bf && D == 512) {
    // row-tile kernels (net_ffn.cu): out_proj + residual + LayerNorm1 in one launch, then linear1 + ReLU + linear2 +
    // residual + LayerNorm2 + token sums in one launch with the 512-wide intermediate kept in shared memory
    // (w.scores is idle in the 16-bit modes: scratch for the per-16-token column sums)
    int rc = c.k->proj_ln_512(w.ctx, c.W(e.att.out.w), c.F(e.att.out.b), w.d0, c.F(e.n1g), c.F(e.n1b), 1e-5f, w.y, R, c.st);
    if (rc == FPB_OK) {
      rc = c.k->ffn_ln_token_mean_512(w.y, c.W(e.l1.w), c.F(e.l1.b), c.W(e.l2.w), c.F(e.l2.b), c.F(e.n2g), c.F(e.n2b), 1e-5f,
                                      w.scores, w.pooled, n, TOK, c.F(e.fin.w), c.F(e.fin.b), out3, e.fin.N, c.st);
      if (rc == FPB_OK) return FPB_OK;
    }
    if (rc != FPB_ERR_UNSUPPORTED) return rc;
  }
  RET(run_linear(c, e.att.out, w.ctx, R, w.y1, w.d0, 0, c.bf));                       // x + SA(x)
  RET(c.k->layernorm_rows(w.y1, c.bf, c.F(e.n1g), c.F(e.n1b), w.y, c.bf, R, D, 1e-5f, c.st));
  RET(run_linear(c, e.l1, w.y, R, w.ff, nullptr, 1, c.bf));
  RET(run_linear(c, e.l2, w.ff, R, w.y1, w.y, 0, c.bf));                               // y + FF(y)
  if (c.bf && D == 512) {                                                              // LayerNorm + token mean in one pass
    RET(c.k->layernorm_token_mean_512(w.y1, c.F(e.n2g), c.F(e.n2b), w.pooled, n, TOK, 1e-5f, c.st));
  } else {
    RET(c.k->layernorm_rows(w.y1, c.bf, c.F(e.n2g), c.F(e.n2b), w.ff, c.bf, R, D, 1e-5f, c.st));
    RET(c.k->token_mean(w.ff, c.bf, w.pooled, n, TOK, D, c.st));                            // mean commutes with Linear
  }
  RET(c.k->small_linear(w.pooled, c.F(e.fin.w), c.F(e.fin.b), out3, n, D, e.fin.N, c.st));
  return FPB_OK;
}

static int refine_forward_impl(const fpb_net* net, const void* AB, int N, void* ws, float* out_trans, float* out_rot,
                               void* stream, bool shared_b)
{
  FPB_CHECK_ARG(net && net->kind == FPB_NET_REFINE && AB && ws && out_trans && out_rot && N > 0);
  Ctx c{net, (cudaStream_t)stream, is16(net->precision), (size_t)(is16(net->precision) ? 2 : 4), kernels_of(net->precision)};
  const int chunk = net_chunk();
  char* base = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  const size_t img = (size_t)IN_S * IN_PITCH * IN_C * c.esz;
  // AB holds all N rendered crops first, then all N observed crops; a chunk needs them adjacent -> gather
  for (int b0 = 0; b0 < N; b0 += chunk) {
    const int n = (N - b0) < chunk ? (N - b0) : chunk;
    Ws w = carve(base, n, c.esz);
    c.splitk = N == 1 ? w.splitk : nullptr;   // split-K only for single-hypothesis calls (track_one): every larger batch keeps
                                              // per-hypothesis results independent of the batch it is evaluated in
    const void* in = AB;
    if (!(b0 == 0 && n == N)) {
      // reuse c3/c2 region? keep it simple: stage the chunk's [A;B] pairs at the start of c1 (dead until conv1 runs)
      char* stage = w.c1;
      static_assert(2 * 1600 * 128 >= 2 * IN_S * IN_PITCH * IN_C / 2, "staging fits");
      FPB_CUDA(cudaMemcpyAsync(stage, (const char*)AB + (size_t)b0 * img, (size_t)n * img, cudaMemcpyDeviceToDevice, c.st));
      if (shared_b)      // the one observed crop sits behind the N rendered ones
        FPB_CUDA(cudaMemcpyAsync(stage + (size_t)n * img, (const char*)AB + (size_t)N * img, img, cudaMemcpyDeviceToDevice, c.st));
      else
        FPB_CUDA(cudaMemcpyAsync(stage + (size_t)n * img, (const char*)AB + (size_t)(N + b0) * img, (size_t)n * img,
                                 cudaMemcpyDeviceToDevice, c.st));
      in = stage;
    }
    RET(run_encoders(c, in, n, w, shared_b));
    if (c.bf && net->st2 && head_streams()) {
      // The two pose heads are independent chains over the same tokens: the rot head runs on a second stream (fork / join
      // by events, also under stream capture), so that its kernels fill the SMs the trans head's partial last waves leave
      // idle -- and, for a single hypothesis (track_one: 4-16 CTAs per kernel), the two chains simply run side by side.
      // Its buffers alias encoder buffers that are dead once the tokens (w.d0) exist; results do not depend on the overlap.
      Ws w2 = w;
      w2.qkv = w.c0; w2.ctx = w.d1; w2.y = w.d2; w2.y1 = w.ab1; w2.ff = w.ab2;
      w2.scores = (float*)w.c1; w2.pooled = (float*)w.c2;
      Ctx c2 = c;
      c2.st = net->st2;
      c2.chain = 1;
      c2.splitk = c.splitk ? w.splitk2 : nullptr;
      FPB_CUDA(cudaEventRecord(net->ev_fork, c.st));
      FPB_CUDA(cudaStreamWaitEvent(net->st2, net->ev_fork, 0));
      const int rc0 = encoder_layer(c, net->head[0], n, w, out_trans + (size_t)b0 * 3);
      const int rc1 = rc0 ? rc0 : encoder_layer(c2, net->head[1], n, w2, out_rot + (size_t)b0 * net->rot_dim);
      FPB_CUDA(cudaEventRecord(net->ev_join, net->st2));          // always join, also on an error: the fork is pending
      FPB_CUDA(cudaStreamWaitEvent(c.st, net->ev_join, 0));
      RET(rc1);
    } else {
      RET(encoder_layer(c, net->head[0], n, w, out_trans + (size_t)b0 * 3));
      RET(encoder_layer(c, net->head[1], n, w, out_rot + (size_t)b0 * net->rot_dim));
    }
  }
  return FPB_OK;
}

FPB_EXPORT int fpb_refine_forward(const fpb_net* net, const void* AB, int N, void* ws, float* out_trans, float* out_rot,
                                  void* stream)
{
  return refine_forward_impl(net, AB, N, ws, out_trans, out_rot, stream, false);
}

FPB_EXPORT int fpb_refine_forward_shared_b(const fpb_net* net, const void* AB, int N, void* ws, float* out_trans, float* out_rot,
                                           void* stream)
{
  return refine_forward_impl(net, AB, N, ws, out_trans, out_rot, stream, true);
}

static int score_features_impl(const fpb_net* net, const void* AB, int N, void* ws, const MultiDst& dst, void* stream)
{
  FPB_CHECK_ARG(net && net->kind == FPB_NET_SCORE && AB && ws && N > 0);
  Ctx c{net, (cudaStream_t)stream, is16(net->precision), (size_t)(is16(net->precision) ? 2 : 4), kernels_of(net->precision)};
  const int chunk = net_chunk();
  char* base = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  const size_t img = (size_t)IN_S * IN_PITCH * IN_C * c.esz;
  for (int b0 = 0; b0 < N; b0 += chunk) {
    const int n = (N - b0) < chunk ? (N - b0) : chunk;
    Ws w = carve(base, n, c.esz);
    c.splitk = N == 1 ? w.splitk : nullptr;   // split-K only for single-hypothesis calls (track_one): every larger batch keeps
                                              // per-hypothesis results independent of the batch it is evaluated in
    const void* in = AB;
    if (!(b0 == 0 && n == N)) {
      char* stage = w.c1;
      FPB_CUDA(cudaMemcpyAsync(stage, (const char*)AB + (size_t)b0 * img, (size_t)n * img, cudaMemcpyDeviceToDevice, c.st));
      FPB_CUDA(cudaMemcpyAsync(stage + (size_t)n * img, (const char*)AB + (size_t)(N + b0) * img, (size_t)n * img,
                               cudaMemcpyDeviceToDevice, c.st));
      in = stage;
    }
    RET(run_encoders(c, in, n, w));
    const long long R = (long long)n * TOK;
    RET(run_self_attention(c, net->att, w.d0, n, w));
    RET(run_linear(c, net->att.out, w.ctx, R, w.y1, nullptr, 0, c.bf));
    MultiDst d = dst;
    for (int i = 0; i < d.n; i++) d.p[i] += (size_t)b0 * D;
    RET(c.k->token_mean_multi(w.y1, c.bf, d, n, TOK, D, c.st));
  }
  return FPB_OK;
}

FPB_EXPORT int fpb_score_features(const fpb_net* net, const void* AB, int N, void* ws, float* out_feat_n512, void* stream)
{
  FPB_CHECK_ARG(out_feat_n512);
  MultiDst d;
  d.n = 1;
  d.p[0] = out_feat_n512;
  return score_features_impl(net, AB, N, ws, d, stream);
}

FPB_EXPORT int fpb_score_features_multi(const fpb_net* net, const void* AB, int N, void* ws, float* const* out_feat_ptrs_h,
                                        int n_dst, void* stream)
{
  FPB_CHECK_ARG(out_feat_ptrs_h && n_dst >= 1 && n_dst <= FPB_MAX_PEERS);
  MultiDst d;
  d.n = n_dst;
  for (int i = 0; i < n_dst; i++) {
    FPB_CHECK_ARG(out_feat_ptrs_h[i]);
    d.p[i] = out_feat_ptrs_h[i];
  }
  return score_features_impl(net, AB, N, ws, d, stream);
}


// cross-hypothesis attention + Linear(512,1): always fp32 SIMT (0.33 GMAC at L=252)
FPB_EXPORT int fpb_score_cross(const fpb_net* net, const float* feat, int L, int bs, void* ws, float* out_logits,
                               void* stream)
{
  FPB_CHECK_ARG(net && net->kind == FPB_NET_SCORE && feat && ws && out_logits && L > 0 && bs > 0);
  cudaStream_t st = (cudaStream_t)stream;
  Ctx c{net, st, false, 4, kernels_of(net->precision)};
  const int Lp = (L + 3) & ~3;
  const long long R = (long long)bs * L;
  char* base = (char*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  size_t o = 0;
  auto take = [&](size_t bytes) { char* p = base + o; o += al(bytes); return p; };
  float* qkv = (float*)take((size_t)R * 3 * D * 4);
  float* scores = (float*)take((size_t)bs * HEADS * L * Lp * 4);
  float* P = (float*)take((size_t)bs * HEADS * L * Lp * 4);
  float* vt = (float*)take((size_t)bs * HEADS * DH * Lp * 4);
  float* ctx = (float*)take((size_t)R * D * 4);
  float* y = (float*)take((size_t)R * D * 4);
  RET(run_linear(c, net->att_cross.in, feat, R, qkv, nullptr, 0, 0));
  RET(run_attention_core(c, qkv, bs, L, Lp, scores, P, vt, ctx, false));
  RET(run_linear(c, net->att_cross.out, ctx, R, y, nullptr, 0, 0));
  RET(c.k->small_linear(y, c.F(net->lin.w), c.F(net->lin.b), out_logits, (int)R, D, 1, st));
  return FPB_OK;
}

FPB_EXPORT size_t fpb_score_cross_ws_bytes(int L, int bs)
{
  const size_t Lp = (L + 3) & ~3, R = (size_t)bs * L;
  return al(R * 3 * D * 4) + 2 * al((size_t)bs * HEADS * L * Lp * 4) + al((size_t)bs * HEADS * DH * Lp * 4) + 2 * al(R * D * 4) + 1024;
}

// ---------------------------------------------------------------------------------------------
// fpb_score_cross_static: fpb_score_cross for callers that cannot allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The scratch of the cross-hypothesis attention (qkv, scores, probabilities, V^T, context, output rows: 5.0 MB at the
// maximum L = 256, bs = 1) lives in a file-scope __device__ array that every GPU loading the library holds until exit.
// In host code the array's NAME is only a host-side stand-in: the device address is resolved per device with
// cudaGetSymbolAddress (with that device current) and that pointer is what the kernels get.
// ---------------------------------------------------------------------------------------------
constexpr int SCORE_CROSS_STATIC_MAX_L = 256;
constexpr size_t SCORE_CROSS_STATIC_BYTES = 5243904;      // fpb_score_cross_ws_bytes(256, 1)
static __device__ __align__(256) unsigned char g_score_cross_scratch[SCORE_CROSS_STATIC_BYTES];
constexpr int SCORE_CROSS_MAX_DEVICES = 64;
static void* g_score_cross_scratch_ptr[SCORE_CROSS_MAX_DEVICES];     // device address of the array, per device ordinal
static std::mutex g_score_cross_mutex;

FPB_EXPORT int fpb_score_cross_static(const fpb_net* net, const float* feat, int L, float* out_logits, void* stream)
{
  FPB_CHECK_ARG(net && net->kind == FPB_NET_SCORE && feat && out_logits && L > 0);
  if (L > SCORE_CROSS_STATIC_MAX_L || fpb_score_cross_ws_bytes(L, 1) > SCORE_CROSS_STATIC_BYTES) {
    fpb_set_error("fpb_score_cross_static: L = %d exceeds the static scratch (at most %d hypotheses per call); use fpb_score_cross "
                  "with a caller-owned workspace", L, SCORE_CROSS_STATIC_MAX_L);
    return FPB_ERR_UNSUPPORTED;
  }
  int dev = -1;
  FPB_CUDA(cudaGetDevice(&dev));
  FPB_CHECK_ARG(dev >= 0 && dev < SCORE_CROSS_MAX_DEVICES);
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_score_cross_mutex);
    if (!g_score_cross_scratch_ptr[dev]) {           // set-up, once per device, with that device current
      void* p = nullptr;
      FPB_CUDA(cudaGetSymbolAddress(&p, g_score_cross_scratch));
      g_score_cross_scratch_ptr[dev] = p;
    }
    scratch = g_score_cross_scratch_ptr[dev];
  }
  // the array keeps the previous call's contents: clear the part this call uses (nothing in the chain reads a scratch
  // element before writing it, the memset keeps that true if the chain changes); stream-ordered, no allocation
  FPB_CUDA(cudaMemsetAsync(scratch, 0, fpb_score_cross_ws_bytes(L, 1), (cudaStream_t)stream));
  return fpb_score_cross(net, feat, L, 1, scratch, out_logits, stream);
}

// operand type the debug / unit-test entries below run in: FPB_PREC_BF16 (default) or FPB_PREC_FP16
static int g_debug_precision = FPB_PREC_BF16;
FPB_EXPORT int fpb_debug_set_precision(int precision)
{
  FPB_CHECK_ARG(precision == FPB_PREC_BF16 || precision == FPB_PREC_FP16);
  g_debug_precision = precision;
  return FPB_OK;
}
#define DK (kernels_of(g_debug_precision))

// Debug / unit-test entry: one implicit GEMM through one kernel family (see GemmProblem).
//   mode 0: fp32 operands (SIMT); 1: bf16 tcgen05 with cp.async producers; 2 / 3: bf16 tcgen05 persistent TMA
//   kernel with cta_group 1 / 2; 4: operand-swapped kernel (64 / 128 output channels, or plain GEMMs with K <= 512 and
//   N % 128 == 0 through its weight-resident mode; out_bf16 = 2 stores the product transposed).  out is fp32 unless out_bf16.  row_pitch / pix_pitch / pad_w: 0 / 0 / pad_h for
//   dense NHWC; the stem uses (168, 8, -1) with Cin = 64 "wide pixels".
FPB_EXPORT int fpb_debug_igemm_ex(int mode, const void* A, const void* W, void* C, int M, int N, int K, int Hin, int Win,
                                  int Cin, int Ho, int Wo, int KH, int KW, int stride, int pad_h, int pad_w, int row_pitch,
                                  int pix_pitch, const float* scale, const float* shift, const void* residual, int relu,
                                  int out_bf16, void* stream)
{
  GemmProblem p = gemm_problem_default();
  p.A = A; p.W = W; p.C = C; p.M = M; p.N = N; p.K = K;
  p.Hin = Hin; p.Win = Win; p.Cin = Cin; p.Ho = Ho; p.Wo = Wo; p.KH = KH; p.KW = KW; p.stride = stride;
  p.pad = pad_h; p.pad_w = pad_w; p.row_pitch = row_pitch; p.pix_pitch = pix_pitch > 0 ? pix_pitch : Cin;
  p.ldw = K; p.scale = scale; p.shift = shift; p.residual = residual; p.ldr = N; p.relu = relu;
  p.ldc = N; p.out_bf16 = out_bf16 ? 1 : 0;
  if (out_bf16 == 2) { p.c_transposed = 1; p.ldc = M; }      // bf16 output stored transposed: C[n][m] (mode 4 only)
  cudaStream_t st = (cudaStream_t)stream;
  switch (mode) {
    case 0: return DK->ref_gemm(p, st);
    case 1: return DK->tc_gemm(p, st);
    case 2: DK->tc2_force_cta_group(1); break;
    case 3: DK->tc2_force_cta_group(2); break;
    case 4: {
      const int rc3 = DK->tc3_gemm(p, st, 1);
      if (rc3 == FPB_ERR_UNSUPPORTED) fpb_set_error("fpb_debug_igemm: shape not supported by the operand-swapped kernel");
      return rc3;
    }
    case 5: {                                      // dedicated stem kernel; W = packed stem weights [64][7][8][8] bf16
      if (!(Hin == 160 && row_pitch == 168 && pix_pitch == 8 && K == 448 && N == 64 && Ho == 80 && M % 6400 == 0 && out_bf16 == 1 && !residual)) {
        fpb_set_error("fpb_debug_igemm: mode 5 is the 7x7 stem on the row-padded input only");
        return FPB_ERR_ARG;
      }
      static void* ws = nullptr;                   // debug entry only: never freed, re-packed on every call
      if (!ws) FPB_CUDA(cudaMalloc(&ws, (size_t)9 * 128 * 64 * 2));
      int rc5 = DK->stem_pack_weights(W, 1, ws, st);
      if (!rc5) rc5 = DK->stem_conv(A, M / 6400, ws, scale, shift, relu, C, st, 1);
      return rc5;
    }
    default: fpb_set_error("fpb_debug_igemm: unknown mode %d", mode); return FPB_ERR_ARG;
  }
  const int rc = DK->tc2_gemm(p, st);
  DK->tc2_force_cta_group(0);
  if (rc == FPB_ERR_UNSUPPORTED) fpb_set_error("fpb_debug_igemm: shape not supported by the TMA kernel");
  return rc;
}

// unit-test entry of the halo 3x3 convolution / row remapping modes of the persistent TMA kernel:
//   A: zero-bordered activation [nimg*(H+1)*(W+1) positions][Cin] (halo=1) or dense NHWC (halo=0, remap=1, 3x3 s1 p1);
//   row_remap 0: output at positions, 1: dense rows -> positions (halo must be 0), 2: positions -> dense rows.
FPB_EXPORT int fpb_debug_conv3x3_halo(int cta_group, const void* A, const void* W, void* C, int nimg, int H, int Wd, int Cin,
                                      int Cout, int halo, int row_remap, const float* scale, const float* shift,
                                      const void* residual, int relu, void* stream)
{
  GemmProblem p = gemm_problem_default();
  p.A = A; p.W = W; p.C = C; p.N = Cout; p.K = 9 * Cin; p.Cin = Cin; p.pix_pitch = Cin; p.ldw = p.K;
  p.KH = p.KW = 3; p.stride = 1; p.pad = p.pad_w = 1; p.Hin = p.Ho = H; p.Win = p.Wo = Wd;
  p.img_h = H; p.img_w = Wd; p.halo = halo; p.row_remap = row_remap;
  p.M = halo ? nimg * (H + 1) * (Wd + 1) : nimg * H * Wd;
  p.scale = scale; p.shift = shift; p.residual = residual; p.ldr = Cout; p.relu = relu; p.ldc = Cout; p.out_bf16 = 1;
  if (cta_group == 3) {                              // operand-swapped kernel (net_tc3.cu)
    const int rc3 = DK->tc3_gemm(p, (cudaStream_t)stream, 1);
    if (rc3 == FPB_ERR_UNSUPPORTED) fpb_set_error("fpb_debug_conv3x3_halo: shape not supported by the operand-swapped kernel");
    return rc3;
  }
  DK->tc2_force_cta_group(cta_group);
  const int rc = DK->tc2_gemm(p, (cudaStream_t)stream);
  DK->tc2_force_cta_group(0);
  if (rc == FPB_ERR_UNSUPPORTED) fpb_set_error("fpb_debug_conv3x3_halo: shape not supported");
  return rc;
}

// unit-test entry of the row-tile encoder-layer kernels (net_ffn.cu).  mode 0: out = LN(A.W2^T + b2 + residual)*gamma+beta
// ([M][512] 16-bit); mode 1: out = pooled [M/tokens][512] fp32 of the fused FFN + LayerNorm + token mean (residual = A),
// scratch >= (M/16)*512 floats.
FPB_EXPORT int fpb_debug_ffn(int mode, const void* A, const void* W1, const float* b1, const void* W2, const float* b2,
                             const void* residual, const float* gamma, const float* beta, void* out, float* scratch,
                             long long M, int tokens, void* stream)
{
  FPB_CHECK_ARG(A && W2 && b2 && gamma && beta && out && M > 0);
  int rc;
  if (mode == 0) {
    FPB_CHECK_ARG(residual);
    rc = DK->proj_ln_512(A, W2, b2, residual, gamma, beta, 1e-5f, out, M, (cudaStream_t)stream);
  } else {
    FPB_CHECK_ARG(W1 && b1 && scratch && tokens > 0 && M % tokens == 0);
    rc = DK->ffn_ln_token_mean_512(A, W1, b1, W2, b2, gamma, beta, 1e-5f, scratch, (float*)out, (int)(M / tokens), tokens,
                                   nullptr, nullptr, nullptr, 0, (cudaStream_t)stream);
  }
  if (rc == FPB_ERR_UNSUPPORTED) fpb_set_error("fpb_debug_ffn: disabled (FPB_FFN=0) or unsupported alignment / shape");
  return rc;
}

// dev tool: clock64 stamps recorded by the last row-tile kernel launched with FPB_FFN_DBG=2048 (tools/ffn_bench.py --stamps)
FPB_EXPORT int fpb_debug_ffn_stamps(long long* out_host, int n)
{
  FPB_CHECK_ARG(out_host && n > 0);
  return DK->ffn_debug_stamps(out_host, n);
}

// unit-test entry of the fused self-attention kernel (net_attn.cu)
FPB_EXPORT int fpb_debug_attention(const void* qk, const void* vt, long long vt_pitch, const float* bv, void* ctx, int groups,
                                   void* stream)
{
  return DK->tc_attention(qk, vt, vt_pitch, bv, ctx, groups, (cudaStream_t)stream);
}

FPB_EXPORT int fpb_debug_igemm(int precision, const void* A, const void* W, void* C, int M, int N, int K, int Hin, int Win,
                               int Cin, int Ho, int Wo, int KH, int KW, int stride, int pad, const float* scale,
                               const float* shift, const void* residual, int relu, int out_bf16, void* stream)
{
  return fpb_debug_igemm_ex(precision, A, W, C, M, N, K, Hin, Win, Cin, Ho, Wo, KH, KW, stride, pad, pad, 0, 0, scale, shift,
                            residual, relu, out_bf16, stream);
}
