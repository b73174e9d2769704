// tcgen05 / TMEM bf16 implicit-GEMM for sm_100a: every conv and linear layer of RefineNet / ScoreNet and the
// batched attention products (QK^T, PV) run through this one kernel.
//   ConvBNReLU / ResnetBasicBlock   learning/models/network_modules.py:37-50, 73-111
//   encoders + heads                learning/models/refine_network.py:38-70, score_network.py:36-57
//
// CTA = 9 warps:  warps 0-3 epilogue (TMEM lanes 32w..32w+31), warps 4-7 operand producers, warp 8 TMEM
// allocator + single-thread MMA issuer.  Per 64-wide K chunk the producers gather the im2col rows of A
// (128 rows x 128 B) and BN rows of W with 16-byte cp.async into the 128B-swizzled K-major layout UMMA reads,
// a ring of STAGES slots is handed over through mbarriers (full: 128 producer arrivals after
// cp.async.wait_group + fence.proxy.async; empty: tcgen05.commit), fp32 accumulators live in TMEM
// (128 lanes x BN columns) and the epilogue applies folded BN / bias, residual, ReLU, positional embedding and
// writes bf16 (or fp32) rows with 16-byte stores.
#include "net_kernels.cuh"
#include "act_type.cuh"

namespace FPB_ACT_NS {

namespace {

constexpr int TM = 128;                 // UMMA M
constexpr int KC = 64;                  // bf16 elements per K chunk = 128 bytes = one swizzle row
constexpr int NUM_THREADS = 288;
constexpr int A_STAGE_BYTES = TM * 128;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra.uni WAIT_DONE;\n\t"
      "bra.uni WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes)
{
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// start>>4 [0,14) | LBO>>4 [16,30) | SBO>>4 [32,46) | version=1 [46,48) | layout=SWIZZLE_128B(2) [61,64)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr)
{
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;                  // LBO (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;        // SBO: 8 rows x 128 B
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// kind::f16 instruction descriptor: D=f32 [4,6)=1, A=bf16 [7,10)=1, B=bf16 [10,13)=1, K-major both,
// N>>3 at [17,23), M>>4 at [24,29)
__host__ __device__ constexpr uint32_t make_idesc(int n)
{
  return (1u << 4) | (ACT_IDESC_FMT << 7) | (ACT_IDESC_FMT << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum)
{
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar)
{
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32])
{
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct SmemCtl {
  uint64_t full[8];
  uint64_t empty[8];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(NUM_THREADS, 1) tc_gemm_kernel(GemmProblem p)
{
  constexpr int B_STAGE_BYTES = BN * 128;
  constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  constexpr int LAG = STAGES - 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  SmemCtl* ctl = (SmemCtl*)(smem + STAGES * STAGE_BYTES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z;
  const long long a_off = (long long)(b / p.b_inner) * p.A_outer + (long long)(b % p.b_inner) * p.A_inner;
  const long long w_off = (long long)(b / p.b_inner) * p.W_outer + (long long)(b % p.b_inner) * p.W_inner;
  const long long c_off = (long long)(b / p.b_inner) * p.C_outer + (long long)(b % p.b_inner) * p.C_inner;
  const int m0 = blockIdx.x * TM, n0 = blockIdx.y * BN;
  const int nk = p.K / KC;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) { mbar_init(&ctl->full[s], 128); mbar_init(&ctl->empty[s], 1); }
    mbar_init(&ctl->tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ctl->tmem_base)),
                 "r"((uint32_t)BN)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp >= 4 && warp < 8) {
    // ------------------------------------------------------------------ producers
    const int pt = threadIdx.x - 128;            // 0..127 : A row of the tile
    const act_t* A = (const act_t*)p.A + a_off;
    const act_t* W = (const act_t*)p.W + w_off;
    const int m = m0 + pt;
    const bool m_ok = m < p.M;
    int img = 0, oy = 0, ox = 0;
    if (m_ok) {
      const int hw = p.Ho * p.Wo;
      img = m / hw;
      const int r = m - img * hw;
      oy = r / p.Wo; ox = r - oy * p.Wo;
    }
    const int iy0 = oy * p.stride - p.pad, ix0 = ox * p.stride - p.pad_w;
    const int rp = p.row_pitch > 0 ? p.row_pitch : p.Win;
    const long long img_base = (long long)img * p.Hin * rp;
    const uint32_t sw = (uint32_t)(pt & 7);
    const bool cin64 = (p.Cin % KC) == 0;
    for (int it = 0; it < nk + LAG; it++) {
      if (it < nk) {
        const int s = it % STAGES;
        mbar_wait(&ctl->empty[s], ((it / STAGES) & 1) ^ 1);
        const uint32_t sa = smem_u32(smem + s * STAGE_BYTES) + pt * 128;
        const int k0 = it * KC;
        if (cin64) {
          const int tap = k0 / p.Cin, c = k0 - tap * p.Cin;
          const int ky = tap / p.KW, kx = tap - ky * p.KW;
          const int iy = iy0 + ky, ix = ix0 + kx;
          const bool ok = m_ok && iy >= 0 && iy < p.Hin && ix >= 0 && ix < p.Win;
          const act_t* src = ok ? A + (img_base + (long long)iy * rp + ix) * p.pix_pitch + c : A;
#pragma unroll
          for (int j = 0; j < 8; j++) cp_async16(sa + ((j ^ sw) << 4), src + (ok ? j * 8 : 0), ok ? 16u : 0u);
        } else {
#pragma unroll
          for (int j = 0; j < 8; j++) {
            const int k = k0 + j * 8;
            const int tap = k / p.Cin, c = k - tap * p.Cin;
            const int ky = tap / p.KW, kx = tap - ky * p.KW;
            const int iy = iy0 + ky, ix = ix0 + kx;
            const bool ok = m_ok && iy >= 0 && iy < p.Hin && ix >= 0 && ix < p.Win;
            const act_t* src = ok ? A + (img_base + (long long)iy * rp + ix) * p.pix_pitch + c : A;
            cp_async16(sa + ((j ^ sw) << 4), src, ok ? 16u : 0u);
          }
        }
        const uint32_t sb = smem_u32(smem + s * STAGE_BYTES + A_STAGE_BYTES);
#pragma unroll
        for (int rr = 0; rr < (BN + 127) / 128; rr++) {
          const int row = pt + rr * 128;
          if (row < BN) {
            const int n = n0 + row;
            const bool ok = n < p.N;
            const act_t* src = ok ? W + (long long)n * p.ldw + k0 : W;
            const uint32_t dst = sb + row * 128;
            const uint32_t swb = (uint32_t)(row & 7);
#pragma unroll
            for (int j = 0; j < 8; j++) cp_async16(dst + ((j ^ swb) << 4), src + (ok ? j * 8 : 0), ok ? 16u : 0u);
          }
        }
      }
      cp_async_commit();
      if (it >= LAG) {
        cp_async_wait<LAG>();
        fence_proxy_async();
        mbar_arrive(&ctl->full[(it - LAG) % STAGES]);
      }
    }
  } else if (warp == 8) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BN);
      for (int it = 0; it < nk; it++) {
        const int s = it % STAGES;
        mbar_wait(&ctl->full[s], (it / STAGES) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * STAGE_BYTES);
        const uint32_t sb = sa + A_STAGE_BYTES;
        const uint64_t da = make_desc(sa), db = make_desc(sb);
#pragma unroll
        for (int k = 0; k < KC / 16; k++)   // +32 bytes per UMMA_K=16 inside the swizzle atom (>>4 => +2)
          umma_bf16(tmem_base, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (it | k) ? 1u : 0u);
        umma_commit(&ctl->empty[s]);
      }
      umma_commit(&ctl->tmem_full);
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 0-3)
    mbar_wait(&ctl->tmem_full, 0);
    tc_fence_after();
    const int row = warp * 32 + lane;
    const int m = m0 + row;
    const bool m_ok = m < p.M;
    const int nst = p.n_valid > 0 ? p.n_valid : p.N;
    const long long orow = m_ok ? (long long)(m % p.m_split) * p.ldc + (long long)(m / p.m_split) * p.split_col_off + c_off : 0;
    const float* pe_row = (p.pe && m_ok) ? p.pe + (long long)(m % p.pe_rows) * p.pe_ld : nullptr;
    const act_t* res_row = (p.residual && m_ok) ? (const act_t*)p.residual + (long long)m * p.ldr : nullptr;
#pragma unroll 1
    for (int ch = 0; ch < BN / 32; ch++) {
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(ch * 32), r);
      const int nb = n0 + ch * 32;
      if (!m_ok || nb >= nst) continue;
      float v[32];
#pragma unroll
      for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]) * p.alpha;
      const bool full = nb + 32 <= nst;
      if (p.scale) {
#pragma unroll
        for (int i = 0; i < 32; i++)
          if (full || nb + i < nst) v[i] = v[i] * __ldg(p.scale + nb + i) + (p.shift ? __ldg(p.shift + nb + i) : 0.0f);
      } else if (p.shift) {
#pragma unroll
        for (int i = 0; i < 32; i++)
          if (full || nb + i < nst) v[i] += __ldg(p.shift + nb + i);
      }
      if (res_row) {
        if (full) {
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const uint4 rv = *reinterpret_cast<const uint4*>(res_row + nb + q * 8);
            const act2_t* h = reinterpret_cast<const act2_t*>(&rv);
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const float2 f = act22f2(h[e]);
              v[q * 8 + e * 2] += f.x; v[q * 8 + e * 2 + 1] += f.y;
            }
          }
        } else {
          for (int i = 0; i < 32 && nb + i < nst; i++) v[i] += act2f(res_row[nb + i]);
        }
      }
      if (p.relu) {
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] = fmaxf(v[i], 0.0f);
      }
      if (pe_row) {
#pragma unroll
        for (int i = 0; i < 32; i++)
          if (full || nb + i < nst) v[i] += __ldg(pe_row + nb + i);
      }
      if (p.out_bf16) {
        act_t* o = (act_t*)p.C + orow + nb;
        if (full) {
#pragma unroll
          for (int q = 0; q < 4; q++) {
            uint4 pk;
            act2_t h0 = f2act2(v[q * 8 + 0], v[q * 8 + 1]), h1 = f2act2(v[q * 8 + 2], v[q * 8 + 3]);
            act2_t h2 = f2act2(v[q * 8 + 4], v[q * 8 + 5]), h3 = f2act2(v[q * 8 + 6], v[q * 8 + 7]);
            pk.x = *reinterpret_cast<uint32_t*>(&h0); pk.y = *reinterpret_cast<uint32_t*>(&h1);
            pk.z = *reinterpret_cast<uint32_t*>(&h2); pk.w = *reinterpret_cast<uint32_t*>(&h3);
            *reinterpret_cast<uint4*>(o + q * 8) = pk;
          }
        } else {
          for (int i = 0; i < 32 && nb + i < nst; i++) o[i] = f2act(v[i]);
        }
      } else {
        float* o = (float*)p.C + orow + nb;
        if (full) {
#pragma unroll
          for (int q = 0; q < 8; q++)
            *reinterpret_cast<float4*>(o + q * 4) = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
        } else {
          for (int i = 0; i < 32 && nb + i < nst; i++) o[i] = v[i];
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN) : "memory");
  }
}

template <int BN, int STAGES>
int launch_tc(const GemmProblem& p, cudaStream_t st)
{
  constexpr int smem = STAGES * (A_STAGE_BYTES + BN * 128) + 1024 + (int)sizeof(SmemCtl);
  static bool configured = false;
  if (!configured) {
    FPB_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(fpb_cdiv(p.M, TM), fpb_cdiv(p.n_valid > 0 ? p.n_valid : p.N, BN), p.batch);
  tc_gemm_kernel<BN, STAGES><<<grid, NUM_THREADS, smem, st>>>(p);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

}  // namespace

int tc_gemm(const GemmProblem& p, cudaStream_t st)
{
  FPB_CHECK_ARG(p.A && p.W && p.C && p.M > 0 && p.N > 0 && p.K > 0 && !p.row_shift && !p.halo && !p.row_remap && !p.c_transposed);
  FPB_CHECK_ARG(p.K % KC == 0 && p.Cin % 8 == 0 && p.pix_pitch % 8 == 0 && p.ldw % 8 == 0 && p.ldc % 8 == 0);
  FPB_CHECK_ARG(p.split_col_off % 8 == 0 && (p.residual == nullptr || p.ldr % 8 == 0));
  FPB_CHECK_ARG((p.A_outer | p.A_inner | p.W_outer | p.W_inner | p.C_outer | p.C_inner) % 8 == 0);
  const int ncols = p.n_valid > 0 ? p.n_valid : p.N;
  if (ncols % 256 == 0 || ncols > 384) return launch_tc<256, 4>(p, st);
  if (ncols > 64) return launch_tc<128, 4>(p, st);
  return launch_tc<64, 4>(p, st);
}

}  // namespace FPB_ACT_NS
