// SIMT fp32 kernels: the parity-mode implicit GEMM (conv / linear / batched attention products) and the
// small row-wise ops used by both precision modes (softmax, LayerNorm, token mean, tiny linear, V transpose).
//   ConvBNReLU / ResnetBasicBlock   learning/models/network_modules.py:37-50, 73-111
//   TransformerEncoderLayer / MHA   learning/models/refine_network.py:56-70, score_network.py:53-54
#include "net_kernels.cuh"
#include "act_type.cuh"
#include <stdlib.h>

namespace FPB_ACT_NS {

namespace {

constexpr int BM = 64, BN = 64, BK = 16;

// 256 threads, 64x64 tile, 4x4 outputs per thread, fp32 FMA.
__global__ void __launch_bounds__(256) ref_gemm_kernel(GemmProblem p)
{
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int b = blockIdx.z;
  const float* A = (const float*)p.A + (long long)(b / p.b_inner) * p.A_outer + (long long)(b % p.b_inner) * p.A_inner;
  const float* W = (const float*)p.W + (long long)(b / p.b_inner) * p.W_outer + (long long)(b % p.b_inner) * p.W_inner;
  const long long c_off = (long long)(b / p.b_inner) * p.C_outer + (long long)(b % p.b_inner) * p.C_inner;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int tid = threadIdx.x;
  const int lr = tid >> 2, lk = (tid & 3) * 4;   // loader: row 0..63, k offset 0,4,8,12
  // decode the A row handled by this loader thread
  const int m = m0 + lr;
  int img = 0, oy = 0, ox = 0;
  const bool mrow_ok = m < p.M;
  if (mrow_ok) {
    const int hw = p.Ho * p.Wo;
    img = m / hw;
    const int r = m - img * hw;
    oy = r / p.Wo; ox = r - oy * p.Wo;
  }
  const int wn = n0 + lr;
  const bool wrow_ok = wn < p.N;
  const int ty = tid >> 4, tx = tid & 15;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j] = 0.0f;

  // operand tile of K step k0 into registers (zero outside the image / the matrix)
  auto fetch = [&](int k0, float4& av, float4& wv) {
    const int k = k0 + lk;
    av = make_float4(0.f, 0.f, 0.f, 0.f);
    wv = av;
    if (mrow_ok && k < p.K) {
      const int tap = k / p.Cin, c = k - tap * p.Cin;
      const int ky = tap / p.KW, kx = tap - ky * p.KW;
      const int iy = oy * p.stride - p.pad + ky, ix = ox * p.stride - p.pad_w + kx;
      if (iy >= 0 && iy < p.Hin && ix >= 0 && ix < p.Win)
        av = *reinterpret_cast<const float4*>(A + ((long long)(img * p.Hin + iy) * (p.row_pitch > 0 ? p.row_pitch : p.Win) + ix) * p.pix_pitch + c);
    }
    if (wrow_ok && k < p.K) wv = *reinterpret_cast<const float4*>(W + (long long)wn * p.ldw + k);
  };
  // the loads of step k0 + BK are issued before the FMAs of step k0 (the loop was bound by one global round trip per step)
  float4 av, wv;
  fetch(0, av, wv);
  for (int k0 = 0; k0 < p.K; k0 += BK) {
    As[lk + 0][lr] = av.x; As[lk + 1][lr] = av.y; As[lk + 2][lr] = av.z; As[lk + 3][lr] = av.w;
    Bs[lk + 0][lr] = wv.x; Bs[lk + 1][lr] = wv.y; Bs[lk + 2][lr] = wv.z; Bs[lk + 3][lr] = wv.w;
    __syncthreads();
    if (k0 + BK < p.K) fetch(k0 + BK, av, wv);
#pragma unroll
    for (int kk = 0; kk < BK; kk++) {
      float a[4], w[4];
#pragma unroll
      for (int i = 0; i < 4; i++) { a[i] = As[kk][ty * 4 + i]; w[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    }
    __syncthreads();
  }
  const int nst = p.n_valid > 0 ? p.n_valid : p.N;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int mm = m0 + ty * 4 + i;
    if (mm >= p.M) continue;
    const long long orow = (long long)(mm % p.m_split) * p.ldc + (long long)(mm / p.m_split) * p.split_col_off + c_off;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int nn = n0 + tx * 4 + j;
      if (nn >= nst) continue;
      float v = acc[i][j] * p.alpha;
      if (p.scale) v = v * p.scale[nn] + (p.shift ? p.shift[nn] : 0.0f);
      else if (p.shift) v += p.shift[nn];
      if (p.residual) v += ((const float*)p.residual)[(long long)mm * p.ldr + nn];
      if (p.relu) v = fmaxf(v, 0.0f);
      if (p.pe) v += p.pe[(long long)(mm % p.pe_rows) * p.pe_ld + nn];
      if (p.out_bf16) ((act_t*)p.C)[orow + nn] = f2act(v);
      else ((float*)p.C)[orow + nn] = v;
    }
  }
}

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// one warp per row
__global__ void __launch_bounds__(256) softmax_kernel(const float* __restrict__ in, void* __restrict__ out, int out_bf16,
                                                      long long rows, int cols, int ld_in, int ld_out)
{
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* x = in + row * ld_in;
  float mx = -INFINITY;
  for (int c = lane; c < cols; c += 32) mx = fmaxf(mx, x[c]);
  mx = warp_max(mx);
  float s = 0.0f;
  for (int c = lane; c < cols; c += 32) s += __expf(x[c] - mx);
  s = warp_sum(s);
  const float inv = 1.0f / s;
  for (int c = lane; c < ld_out; c += 32) {
    float v = c < cols ? __expf(x[c] - mx) * inv : 0.0f;
    if (out_bf16) ((act_t*)out)[row * ld_out + c] = f2act(v);
    else ((float*)out)[row * ld_out + c] = v;
  }
}

template <bool XBF>
__global__ void __launch_bounds__(256) layernorm_kernel(const void* __restrict__ x_, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, void* __restrict__ out,
                                                        int out_bf16, long long rows, int cols, float eps)
{
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  auto ld = [&](int c) -> float {
    if (XBF) return act2f(((const act_t*)x_)[row * cols + c]);
    return ((const float*)x_)[row * cols + c];
  };
  float s = 0.0f;
  for (int c = lane; c < cols; c += 32) s += ld(c);
  const float mean = warp_sum(s) / (float)cols;
  float v = 0.0f;
  for (int c = lane; c < cols; c += 32) { float d = ld(c) - mean; v += d * d; }
  const float rstd = rsqrtf(warp_sum(v) / (float)cols + eps);
  for (int c = lane; c < cols; c += 32) {
    float y = (ld(c) - mean) * rstd * gamma[c] + beta[c];
    if (out_bf16) ((act_t*)out)[row * cols + c] = f2act(y);
    else ((float*)out)[row * cols + c] = y;
  }
}


// ---- bf16 fast paths for 512-wide rows (the transformer heads): one warp per row, the row lives in registers
//      (lane l holds columns 8l..8l+7 and 256+8l..256+8l+7: two 16-byte loads, 512 contiguous bytes per warp each)
struct Ln512 {
  float v[16];
  __device__ __forceinline__ void load(const act_t* row, int lane)
  {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(row) + lane), b = __ldg(reinterpret_cast<const uint4*>(row) + 32 + lane);
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 8; i++) { v[2 * i] = actbits2f((unsigned short)(w[i] & 0xffffu)); v[2 * i + 1] = actbits2f((unsigned short)(w[i] >> 16)); }
  }
  // (x - mean) * rstd * gamma + beta, in place
  __device__ __forceinline__ void normalize(const float (&g)[16], const float (&b)[16], float eps)
  {
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 16; i++) s += v[i];
    const float mean = warp_sum(s) * (1.0f / 512.0f);
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < 16; i++) { const float d = v[i] - mean; q += d * d; }
    const float rstd = rsqrtf(warp_sum(q) * (1.0f / 512.0f) + eps);
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = (v[i] - mean) * rstd * g[i] + b[i];
  }
};
__device__ __forceinline__ void ln512_params(const float* gamma, const float* beta, int lane, float (&g)[16], float (&b)[16])
{
#pragma unroll
  for (int h = 0; h < 2; h++)
#pragma unroll
    for (int i = 0; i < 8; i++) { g[h * 8 + i] = __ldg(gamma + h * 256 + lane * 8 + i); b[h * 8 + i] = __ldg(beta + h * 256 + lane * 8 + i); }
}

__global__ void __launch_bounds__(256) ln512_kernel(const act_t* __restrict__ x, const float* __restrict__ gamma,
                                                    const float* __restrict__ beta, act_t* __restrict__ out,
                                                    long long rows, float eps)
{
  const int lane = threadIdx.x & 31;
  float g[16], b[16];
  ln512_params(gamma, beta, lane, g, b);
  const long long nw = (long long)gridDim.x * 8;
  auto store = [&](const Ln512& r, long long row) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const act2_t h = f2act2(r.v[2 * i], r.v[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    uint4* o = reinterpret_cast<uint4*>(out + row * 512);
    o[lane] = make_uint4(w[0], w[1], w[2], w[3]);
    o[32 + lane] = make_uint4(w[4], w[5], w[6], w[7]);
  };
  for (long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); row < rows; row += 2 * nw) {   // two rows in flight per warp
    Ln512 r0, r1;
    const bool two = row + nw < rows;
    r0.load(x + row * 512, lane);
    if (two) r1.load(x + (row + nw) * 512, lane);
    r0.normalize(g, b, eps);
    if (two) r1.normalize(g, b, eps);
    store(r0, row);
    if (two) store(r1, row + nw);
  }
}

// LayerNorm followed by the mean over a group's tokens (the normalised rows are never written): one block per group.
// The summation order is EXACTLY that of the 8-CTA cluster kernel below (token t belongs to "rank" (t / 8) % 8 and warp
// t % 8; per rank the 8 warp partials are added in warp order, the 8 rank rows in rank order), so that a hypothesis gets
// bit-identical head outputs whether it is evaluated in a 252-hypothesis batch (this kernel) or in a 32-hypothesis
// shard of it (cluster kernel): sharded and unsharded register() must agree exactly.
__global__ void __launch_bounds__(256) ln512_token_mean_kernel(const act_t* __restrict__ x, const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, float* __restrict__ out,
                                                               int tokens, float eps)
{
  __shared__ float part[8][512];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float g[16], b[16];
  ln512_params(gamma, beta, lane, g, b);
  const act_t* xg = x + (long long)blockIdx.x * tokens * 512;
  float total[2] = {0.0f, 0.0f};                    // columns threadIdx.x and threadIdx.x + 256
  for (int rank = 0; rank < 8; rank++) {
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = 0.0f;
    // two rows in flight per warp (independent load -> reduce -> normalise chains); accumulated in token order
    for (int t = rank * 8 + warp; t < tokens; t += 128) {
      Ln512 r0, r1;
      const bool two = t + 64 < tokens;
      r0.load(xg + (long long)t * 512, lane);
      if (two) r1.load(xg + (long long)(t + 64) * 512, lane);
      r0.normalize(g, b, eps);
      if (two) r1.normalize(g, b, eps);
#pragma unroll
      for (int i = 0; i < 16; i++) acc[i] += r0.v[i];
      if (two) {
#pragma unroll
        for (int i = 0; i < 16; i++) acc[i] += r1.v[i];
      }
    }
#pragma unroll
    for (int h = 0; h < 2; h++)
#pragma unroll
      for (int i = 0; i < 8; i++) part[warp][h * 256 + lane * 8 + i] = acc[h * 8 + i];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const int c = threadIdx.x + k * 256;
      float s = 0.0f;
#pragma unroll
      for (int w = 0; w < 8; w++) s += part[w][c];
      total[k] += s;
    }
    __syncthreads();
  }
  out[(long long)blockIdx.x * 512 + threadIdx.x] = total[0] / (float)tokens;
  out[(long long)blockIdx.x * 512 + threadIdx.x + 256] = total[1] / (float)tokens;
}

// Small batches (track_one: one group; 8-GPU shards: 32 groups): a cluster of 8 CTAs per group, each CTA takes an
// eighth of the tokens and the leader sums the eight partial rows through distributed shared memory in rank order
// (deterministic).  One block per group would leave the device idle for 400 dependent row reductions (44 us at N = 1).
constexpr int LNM_CLUSTER = 8;
__global__ void __cluster_dims__(LNM_CLUSTER, 1, 1) __launch_bounds__(256)
ln512_token_mean_cluster_kernel(const act_t* __restrict__ x, const float* __restrict__ gamma,
                                const float* __restrict__ beta, float* __restrict__ out, int tokens, float eps)
{
  __shared__ float part[8][512];
  __shared__ float row[512];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int group = blockIdx.x / LNM_CLUSTER;
  float g[16], b[16], acc[16];
  ln512_params(gamma, beta, lane, g, b);
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = 0.0f;
  const act_t* xg = x + (long long)group * tokens * 512;
  for (int t = (int)rank * 8 + warp; t < tokens; t += 16 * LNM_CLUSTER) {      // two rows in flight, token order kept
    Ln512 r0, r1;
    const bool two = t + 8 * LNM_CLUSTER < tokens;
    r0.load(xg + (long long)t * 512, lane);
    if (two) r1.load(xg + (long long)(t + 8 * LNM_CLUSTER) * 512, lane);
    r0.normalize(g, b, eps);
    if (two) r1.normalize(g, b, eps);
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] += r0.v[i];
    if (two) {
#pragma unroll
      for (int i = 0; i < 16; i++) acc[i] += r1.v[i];
    }
  }
#pragma unroll
  for (int h = 0; h < 2; h++)
#pragma unroll
    for (int i = 0; i < 8; i++) part[warp][h * 256 + lane * 8 + i] = acc[h * 8 + i];
  __syncthreads();
  for (int c = threadIdx.x; c < 512; c += 256) {
    float s = 0.0f;
#pragma unroll
    for (int w = 0; w < 8; w++) s += part[w][c];
    row[c] = s;
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (rank == 0) {
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(row);
    for (int c = threadIdx.x; c < 512; c += 256) {
      float s = 0.0f;
#pragma unroll
      for (uint32_t r = 0; r < LNM_CLUSTER; r++) {
        uint32_t remote;
        float v;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(base + (uint32_t)c * 4u), "r"(r));
        asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
        s += v;
      }
      out[(long long)group * 512 + c] = s / (float)tokens;
    }
  }
  // peers must not exit while the leader still reads their shared memory
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// out[g][c] = mean_t x[g][t][c]; one block per (g, 64-column slab), 256 threads = 4 token lanes x 64 cols
template <bool XBF>
__global__ void __launch_bounds__(256) token_mean_kernel(const void* __restrict__ x_, const MultiDst out, int tokens, int cols)
{
  __shared__ float part[4][64];
  const int g = blockIdx.x, c = blockIdx.y * 64 + (threadIdx.x & 63), tl = threadIdx.x >> 6;
  float s = 0.0f;
  if (c < cols)
    for (int t = tl; t < tokens; t += 4) {
      const long long i = ((long long)g * tokens + t) * cols + c;
      s += XBF ? act2f(((const act_t*)x_)[i]) : ((const float*)x_)[i];
    }
  part[tl][threadIdx.x & 63] = s;
  __syncthreads();
  if (tl == 0 && c < cols) {
    const float v = (((part[0][threadIdx.x] + part[1][threadIdx.x]) + part[2][threadIdx.x]) + part[3][threadIdx.x]) / (float)tokens;
    for (int d = 0; d < out.n; d++) out.p[d][(long long)g * cols + c] = v;      // peer destinations: stores over NVLink
  }
}

// one warp per output element
__global__ void __launch_bounds__(256) small_linear_kernel(const float* __restrict__ x, const float* __restrict__ W,
                                                           const float* __restrict__ b, float* __restrict__ out, int rows,
                                                           int K, int N)
{
  const long long o = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (o >= (long long)rows * N) return;
  const int r = (int)(o / N), n = (int)(o % N);
  float s = 0.0f;
  for (int k = lane; k < K; k += 32) s = fmaf(x[(long long)r * K + k], W[(long long)n * K + k], s);
  s = warp_sum(s);
  if (lane == 0) out[o] = s + (b ? b[n] : 0.0f);
}

// vt[(g*heads+h)][d][t] = qkv[(g*tokens+t)][v_off + h*dh + d]; pad columns [tokens, ld_vt) are zeroed
template <typename T>
__global__ void __launch_bounds__(256) transpose_v_kernel(const T* __restrict__ qkv, T* __restrict__ vt, int tokens,
                                                          int heads, int dh, int ld_qkv, int v_off, int ld_vt)
{
  __shared__ T tile[32][33];
  const int gh = blockIdx.z, g = gh / heads, h = gh % heads;
  const int t0 = blockIdx.x * 32, d0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  for (int i = ty; i < 32; i += 8) {
    const int t = t0 + i, d = d0 + tx;
    tile[i][tx] = (t < tokens && d < dh) ? qkv[((long long)g * tokens + t) * ld_qkv + v_off + h * dh + d]
                                          : T(0.0f);
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int d = d0 + i, t = t0 + tx;
    if (d < dh && t < ld_vt) vt[((long long)gh * dh + d) * ld_vt + t] = tile[tx][i];
  }
}

__global__ void f32_to_act_kernel(const float* __restrict__ in, act_t* __restrict__ out, long long n)
{
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = f2act(in[i]);
}

// Plain fp32 GEMMs with few rows (ScoreNet's cross-hypothesis attention: 252 x 512 x 1536 and smaller, score_network.py:86):
// the 64 x 64 x 16 kernel above puts 4 x 24 blocks on 148 SMs and walks 32 dependent global round trips.  32 x 64 tiles
// with 32-deep K steps double the blocks and halve the trips; every output is still one sequential fmaf chain over k,
// i.e. bit-identical to ref_gemm_kernel.  (160 -> ~40 us for the whole att_cross block at L = 252.)
constexpr int SM_ = 32, SN_ = 64, SK_ = 32;
__global__ void __launch_bounds__(256) small_gemm_kernel(GemmProblem p)
{
  __shared__ float As[SK_][SM_ + 4];
  __shared__ float Bs[SK_][SN_ + 4];
  const int b = blockIdx.z;
  const float* A = (const float*)p.A + (long long)(b / p.b_inner) * p.A_outer + (long long)(b % p.b_inner) * p.A_inner;
  const float* W = (const float*)p.W + (long long)(b / p.b_inner) * p.W_outer + (long long)(b % p.b_inner) * p.W_inner;
  const long long c_off = (long long)(b / p.b_inner) * p.C_outer + (long long)(b % p.b_inner) * p.C_inner;
  const int m0 = blockIdx.x * SM_, n0 = blockIdx.y * SN_;
  const int tid = threadIdx.x;
  const int ar = tid >> 3, ak = (tid & 7) * 4;          // A loader: row 0..31, k offset 0..28
  const int wr = tid >> 3, wk = (tid & 7) * 4;          // W loader: rows wr and wr + 32
  const int ty = tid >> 4, tx = tid & 15;               // outputs: rows ty*2 + {0,1}, columns tx*4 + {0..3}
  float acc[2][4];
#pragma unroll
  for (int i = 0; i < 2; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j] = 0.0f;
  auto fetch = [&](int k0, float4& av, float4& w0, float4& w1) {
    av = make_float4(0.f, 0.f, 0.f, 0.f); w0 = av; w1 = av;
    if (m0 + ar < p.M && k0 + ak < p.K) av = *reinterpret_cast<const float4*>(A + (long long)(m0 + ar) * p.pix_pitch + k0 + ak);
    if (k0 + wk < p.K) {
      if (n0 + wr < p.N) w0 = *reinterpret_cast<const float4*>(W + (long long)(n0 + wr) * p.ldw + k0 + wk);
      if (n0 + wr + 32 < p.N) w1 = *reinterpret_cast<const float4*>(W + (long long)(n0 + wr + 32) * p.ldw + k0 + wk);
    }
  };
  float4 av, w0, w1;
  fetch(0, av, w0, w1);
  for (int k0 = 0; k0 < p.K; k0 += SK_) {
    As[ak + 0][ar] = av.x; As[ak + 1][ar] = av.y; As[ak + 2][ar] = av.z; As[ak + 3][ar] = av.w;
    Bs[wk + 0][wr] = w0.x; Bs[wk + 1][wr] = w0.y; Bs[wk + 2][wr] = w0.z; Bs[wk + 3][wr] = w0.w;
    Bs[wk + 0][wr + 32] = w1.x; Bs[wk + 1][wr + 32] = w1.y; Bs[wk + 2][wr + 32] = w1.z; Bs[wk + 3][wr + 32] = w1.w;
    __syncthreads();
    if (k0 + SK_ < p.K) fetch(k0 + SK_, av, w0, w1);
#pragma unroll
    for (int kk = 0; kk < SK_; kk++) {
      const float a0 = As[kk][ty * 2], a1 = As[kk][ty * 2 + 1];
      const float4 w = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      acc[0][0] = fmaf(a0, w.x, acc[0][0]); acc[0][1] = fmaf(a0, w.y, acc[0][1]);
      acc[0][2] = fmaf(a0, w.z, acc[0][2]); acc[0][3] = fmaf(a0, w.w, acc[0][3]);
      acc[1][0] = fmaf(a1, w.x, acc[1][0]); acc[1][1] = fmaf(a1, w.y, acc[1][1]);
      acc[1][2] = fmaf(a1, w.z, acc[1][2]); acc[1][3] = fmaf(a1, w.w, acc[1][3]);
    }
    __syncthreads();
  }
  const int nst = p.n_valid > 0 ? p.n_valid : p.N;
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const int mm = m0 + ty * 2 + i;
    if (mm >= p.M) continue;
    const long long orow = (long long)mm * p.ldc + c_off;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int nn = n0 + tx * 4 + j;
      if (nn >= nst) continue;
      float v = acc[i][j] * p.alpha;
      if (p.scale) v = v * p.scale[nn] + (p.shift ? p.shift[nn] : 0.0f);
      else if (p.shift) v += p.shift[nn];
      if (p.residual) v += ((const float*)p.residual)[(long long)mm * p.ldr + nn];
      if (p.relu) v = fmaxf(v, 0.0f);
      ((float*)p.C)[orow + nn] = v;
    }
  }
}

}  // namespace

int ref_gemm(const GemmProblem& p, cudaStream_t st)
{
  FPB_CHECK_ARG(p.A && p.W && p.C && p.M > 0 && p.N > 0 && p.K > 0 && p.K % 4 == 0 && p.Cin % 4 == 0 && !p.row_shift && !p.c_transposed);
  const bool plain = p.KH == 1 && p.KW == 1 && p.Hin == 1 && p.Win == 1 && p.Ho == 1 && p.Wo == 1 && p.stride == 1 && p.pad == 0 &&
                     p.pad_w == 0 && p.Cin == p.K && !p.pe && !p.out_bf16 && p.m_split >= p.M && p.pix_pitch % 4 == 0 && p.ldw % 4 == 0;
  if (plain && p.M <= 2048) {
    dim3 g(fpb_cdiv(p.M, SM_), fpb_cdiv(p.N, SN_), p.batch);
    small_gemm_kernel<<<g, 256, 0, st>>>(p);
    FPB_LAUNCH_CHECK();
    return FPB_OK;
  }
  dim3 grid(fpb_cdiv(p.M, BM), fpb_cdiv(p.N, BN), p.batch);
  ref_gemm_kernel<<<grid, 256, 0, st>>>(p);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

int softmax_rows(const float* in, void* out, int out_bf16, long long rows, int cols, int ld_in, int ld_out, cudaStream_t st)
{
  softmax_kernel<<<fpb_cdiv(rows, 8), 256, 0, st>>>(in, out, out_bf16, rows, cols, ld_in, ld_out);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

int layernorm_rows(const void* x, int x_bf16, const float* gamma, const float* beta, void* out, int out_bf16,
                   long long rows, int cols, float eps, cudaStream_t st)
{
  if (x_bf16 && out_bf16 && cols == 512 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 15) == 0) {
    const long long blocks = fpb_cdiv(rows, 8);
    ln512_kernel<<<(unsigned)(blocks < 148 * 8 ? blocks : 148 * 8), 256, 0, st>>>((const act_t*)x, gamma, beta, (act_t*)out, rows, eps);
  } else if (x_bf16) layernorm_kernel<true><<<fpb_cdiv(rows, 8), 256, 0, st>>>(x, gamma, beta, out, out_bf16, rows, cols, eps);
  else layernorm_kernel<false><<<fpb_cdiv(rows, 8), 256, 0, st>>>(x, gamma, beta, out, out_bf16, rows, cols, eps);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

// out[g][c] = mean_t LayerNorm(x[g][t][:])[c]  (bf16 rows of 512 columns; the normalised rows stay in registers, fp32)
int layernorm_token_mean_512(const void* x, const float* gamma, const float* beta, float* out, int groups, int tokens, float eps,
                             cudaStream_t st)
{
  if ((uintptr_t)x & 15) { fpb_set_error("layernorm_token_mean_512: unaligned input"); return FPB_ERR_ARG; }
  static int cluster_max = -1;
  if (cluster_max < 0) { const char* e = getenv("FPB_LNM_CLUSTER_MAX"); cluster_max = e ? atoi(e) : 74; }
  if (groups <= cluster_max)      // default 74: <= 592 CTAs, two waves of clusters at most
    ln512_token_mean_cluster_kernel<<<groups * LNM_CLUSTER, 256, 0, st>>>((const act_t*)x, gamma, beta, out, tokens, eps);
  else
    ln512_token_mean_kernel<<<groups, 256, 0, st>>>((const act_t*)x, gamma, beta, out, tokens, eps);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

int token_mean(const void* x, int x_bf16, float* out, int groups, int tokens, int cols, cudaStream_t st)
{
  MultiDst d;
  d.n = 1;
  d.p[0] = out;
  return token_mean_multi(x, x_bf16, d, groups, tokens, cols, st);
}

int token_mean_multi(const void* x, int x_bf16, const MultiDst& out, int groups, int tokens, int cols, cudaStream_t st)
{
  dim3 grid(groups, fpb_cdiv(cols, 64));
  if (x_bf16) token_mean_kernel<true><<<grid, 256, 0, st>>>(x, out, tokens, cols);
  else token_mean_kernel<false><<<grid, 256, 0, st>>>(x, out, tokens, cols);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

int small_linear(const float* x, const float* W, const float* b, float* out, int rows, int K, int N, cudaStream_t st)
{
  small_linear_kernel<<<fpb_cdiv((long long)rows * N, 8), 256, 0, st>>>(x, W, b, out, rows, K, N);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

int transpose_v(const void* qkv, void* vt, int is_bf16, int groups, int tokens, int heads, int dh, int ld_qkv, int v_off,
                int ld_vt, cudaStream_t st)
{
  dim3 grid(fpb_cdiv(ld_vt, 32), fpb_cdiv(dh, 32), groups * heads);
  if (is_bf16)
    transpose_v_kernel<act_t><<<grid, 256, 0, st>>>((const act_t*)qkv, (act_t*)vt, tokens, heads,
                                                             dh, ld_qkv, v_off, ld_vt);
  else
    transpose_v_kernel<float><<<grid, 256, 0, st>>>((const float*)qkv, (float*)vt, tokens, heads, dh, ld_qkv, v_off, ld_vt);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

int f32_to_act(const float* in, void* out, long long n, cudaStream_t st)
{
  f32_to_act_kernel<<<fpb_cdiv(n, 256), 256, 0, st>>>(in, (act_t*)out, n);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

}  // namespace FPB_ACT_NS
