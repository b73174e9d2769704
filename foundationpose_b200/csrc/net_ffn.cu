// Row-tile tcgen05 kernels for the 512-wide tail of a transformer encoder layer (pose heads of RefineNet):
//   learning/models/refine_network.py:56-70 (nn.TransformerEncoderLayer: out_proj + residual + LayerNorm1,
//   linear1 + ReLU + linear2 + residual + LayerNorm2) followed by `.mean(dim=1)` over the 400 tokens (:88-93).
//
// The K <= 512 linears used to run one by one on the weight-resident operand-swapped kernel (a row's 512 channels spread
// over 4 CTAs) with LayerNorm / LayerNorm + token-mean as separate HBM passes.  LayerNorm is a ROW operation, so here a
// CTA pair (tcgen05 cta_group::2) owns 256 full rows: the accumulator of a 128 x 512 output tile fills the CTA's whole
// TMEM (512 fp32 columns, lane = row), and the thread that owns a row computes its statistics straight from TMEM.
//   ffn_kernel<false>  ("proj+LN"):   y = LN(A . W^T + b + residual) -> 16-bit rows            (out_proj + LayerNorm1)
//   ffn_kernel<true>   ("FFN+LN+sum"): h = relu(A . W1^T + b1) stays in shared memory as the A operand of the second
//                      product (written by the epilogue threads in the 128B-swizzled K-major layout TMA would produce),
//                      v = h . W2^T + b2 + A, n = (v - mean) * rstd, and only the per-16-row column sums of n leave the
//                      chip: 16 divides the 400 tokens of a hypothesis and the 128-row tiles, so a group never straddles
//                      a hypothesis and the summation order of a hypothesis does not depend on the batch (or shard) it
//                      is evaluated in.  pool_groups_kernel adds a hypothesis' 25 group rows in order and applies
//                      gamma / beta / (1 / tokens) (the affine map commutes with the mean).
// Roles per CTA (10 warps): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (leader CTA), warps 2-9 =
// epilogue (two warps per TMEM lane quarter, alternating 64-column groups).  TMEM holds ONE accumulator, so the MMA and
// epilogue phases of a tile alternate; the weight stream (W1 / W2 through a 4-deep ring, half of every 256-row weight
// block per CTA of the pair) keeps prefetching underneath.
#include "net_kernels.cuh"
#include "tc_ptx.cuh"
#include <cuda.h>
#include <stdlib.h>

namespace FPB_ACT_NS {

namespace {

constexpr int FD = 512;                    // row width = K = N of every product here
constexpr int KC = 64;                     // K elements per stage (128 bytes: one swizzle atom row)
constexpr int NKC = FD / KC;
constexpr int TMR = 128;                   // rows per CTA
constexpr int CHUNK = TMR * 128;           // one [128 rows x 64 K] operand block, 16 KB
constexpr int NTHREADS = 320;

struct FfnParams {
  int M, m_tiles;                          // rows, tiles of 256 rows
  const act_t* residual;                   // [M][512]
  const float* b1;                         // FFN: bias of the first product
  const float* b2;                         // bias of the (last) product
  const float* gamma;                      // proj+LN only (the FFN kernel leaves gamma / beta to pool_groups_kernel)
  const float* beta;
  float eps;
  act_t* out;                              // proj+LN: [M][512]
  float* gs;                               // FFN: [M/16][512] column sums of the normalised rows of each 16-row group
  int dbg;
};

struct FfnCtl {
  uint64_t full_a[4], empty_a[4], full_b[4], empty_b[4];
  uint64_t acc_full, acc_free, stg_free;
  uint32_t tmem_base;
};

template <bool FFN>
struct FCfg {
  static constexpr int NA = FFN ? 2 : 4;                     // must divide NKC (slot of K chunk kc = kc % NA)
  static constexpr int NB = 4;
  static constexpr int OFF_H = 0;
  static constexpr int OFF_B = FFN ? NKC * CHUNK : 0;
  static constexpr int OFF_A = OFF_B + NB * CHUNK;
  // 8 warp-private 4 KB staging tiles.  FFN: they overlay the A ring (32 KB), which is idle from the end of the first
  // product until the next tile (the producer waits for stg_free before it refills the ring).
  static constexpr int OFF_STG = FFN ? OFF_A : OFF_A + NA * CHUNK;
  static constexpr int OFF_XCH = OFF_STG + 8 * 4096;         // proj+LN: 8 x 256 B row-statistics exchange
  static constexpr int DATA = FFN ? OFF_A + NA * CHUNK : OFF_XCH + 8 * 256;
  static constexpr int SMEM = DATA + 1024 + (int)sizeof(FfnCtl);
  static_assert(NKC % NA == 0, "ring slot of a K chunk must not depend on the tile");
  static_assert(!FFN || NA * CHUNK == 8 * 4096, "staging overlays the A ring exactly");
  static_assert(SMEM <= 232448, "shared memory budget");
};

__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity)
{
  const uint32_t addr = smem_u32(bar);
  uint32_t ok = 0;
  long long t0 = 0;
  for (uint32_t spin = 0; !ok; spin++) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    if (!ok && (spin & 1023u) == 1023u) {
      const long long now = clock64();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 8000000000LL) __trap();
    }
  }
}
// release at cluster scope: orders this thread's earlier shared-memory writes (the h tile) before the arrival
__device__ __forceinline__ void mbar_arrive_cluster_release(uint64_t* bar, uint32_t cta)
{
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_local(uint64_t* bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float* v)
{
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]), "f"(v[8]), "f"(v[9]), "f"(v[10]),
      "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]), "f"(v[16]), "f"(v[17]), "f"(v[18]), "f"(v[19]), "f"(v[20]),
      "f"(v[21]), "f"(v[22]), "f"(v[23]), "f"(v[24]), "f"(v[25]), "f"(v[26]), "f"(v[27]), "f"(v[28]), "f"(v[29]), "f"(v[30]),
      "f"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// v[i] += bias[n0 + i], 64 columns, warp-uniform 16-byte loads
__device__ __forceinline__ void add_bias64(float (&v)[64], const float* __restrict__ bias, int n0)
{
  const float4* b4 = reinterpret_cast<const float4*>(bias + n0);
#pragma unroll
  for (int c = 0; c < 16; c++) {
    const float4 b = __ldg(b4 + c);
    v[c * 4 + 0] += b.x; v[c * 4 + 1] += b.y; v[c * 4 + 2] += b.z; v[c * 4 + 3] += b.w;
  }
}

template <bool FFN>
__global__ void __launch_bounds__(NTHREADS, 1)
ffn_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapW1,
           const __grid_constant__ CUtensorMap mapW2, const FfnParams q)
{
  using C = FCfg<FFN>;
  constexpr int NA = C::NA, NB = C::NB;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  FfnCtl* ctl = (FfnCtl*)(smem + C::DATA);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  pdl_launch_dependents();

  if (threadIdx.x == 0) {
    for (int s = 0; s < 4; s++) {
      mbar_init(&ctl->full_a[s], 1); mbar_init(&ctl->empty_a[s], 1);
      mbar_init(&ctl->full_b[s], 1); mbar_init(&ctl->empty_b[s], 1);
    }
    mbar_init(&ctl->acc_full, 1);
    mbar_init(&ctl->acc_free, 16);                 // 8 epilogue warps x 2 CTAs, on the leader's barrier
    mbar_init(&ctl->stg_free, 8);                  // 8 epilogue warps of this CTA
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ctl->tmem_base)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapW1) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapW2) : "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;
  pdl_wait();

  const uint32_t sH = smem_u32(smem + C::OFF_H), sB = smem_u32(smem + C::OFF_B), sA = smem_u32(smem + C::OFF_A);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (one lane, both CTAs)
    if (lane == 0) {
      const uint32_t full_a0 = mapa_u32(smem_u32(&ctl->full_a[0]), 0), full_b0 = mapa_u32(smem_u32(&ctl->full_b[0]), 0);
      int sa = 0, sb = 0;
      uint32_t pa = 1, pb = 1;                               // producer waits with inverted parity
      uint32_t it = 0;
      for (int t = cluster_id; t < q.m_tiles; t += n_clusters, it++) {
        const int m0 = (t * 2 + (int)rank) * TMR;
        if (FFN && it > 0) mbar_wait(&ctl->stg_free, (it - 1) & 1u);   // the staging tiles that overlay the A ring are idle again
        for (int kc = 0; kc < NKC; kc++) {
          mbar_wait(&ctl->empty_a[sa], pa);
          if (rank == 0) mbar_arrive_expect_tx(&ctl->full_a[sa], 2u * CHUNK);
          tma_load_2d<2>(sA + sa * CHUNK, &mapA, full_a0 + 8u * sa, kc * KC, m0);
          if (++sa == NA) { sa = 0; pa ^= 1u; }
#pragma unroll 1
          for (int nh = 0; nh < 2; nh++) {
            mbar_wait(&ctl->empty_b[sb], pb);
            if (rank == 0) mbar_arrive_expect_tx(&ctl->full_b[sb], 2u * CHUNK);
            tma_load_2d<2>(sB + sb * CHUNK, FFN ? &mapW1 : &mapW2, full_b0 + 8u * sb, kc * KC, nh * 256 + (int)rank * TMR);
            if (++sb == NB) { sb = 0; pb ^= 1u; }
          }
        }
        if (FFN) {
          for (int kc = 0; kc < NKC; kc++) {
#pragma unroll 1
            for (int nh = 0; nh < 2; nh++) {
              mbar_wait(&ctl->empty_b[sb], pb);
              if (rank == 0) mbar_arrive_expect_tx(&ctl->full_b[sb], 2u * CHUNK);
              tma_load_2d<2>(sB + sb * CHUNK, &mapW2, full_b0 + 8u * sb, kc * KC, nh * 256 + (int)rank * TMR);
              if (++sb == NB) { sb = 0; pb ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA, one lane)
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc(256, 256);
      int sa = 0, sb = 0;
      uint32_t pa = 0, pb = 0, fph = 0;
      bool first = true;
      // one 256 x 512 x 512 product into the pair's accumulators: A from the ring (streamed) or from the resident h tile
      auto product = [&](bool a_ring) {
        if (!first) {                                        // accumulator drained (and, before the second product, h written)
          mbar_wait_cluster(&ctl->acc_free, fph);
          fph ^= 1u;
          tc_fence_after();
        }
        first = false;
        for (int kc = 0; kc < NKC; kc++) {
          uint64_t da;
          if (a_ring) {
            mbar_wait(&ctl->full_a[sa], pa);
            tc_fence_after();
            da = make_desc(sA + sa * CHUNK);
          } else {
            da = make_desc(sH + kc * CHUNK);
          }
#pragma unroll 1
          for (int nh = 0; nh < 2; nh++) {
            mbar_wait(&ctl->full_b[sb], pb);
            tc_fence_after();
            const uint64_t db = make_desc(sB + sb * CHUNK);
            const uint32_t d_tmem = tmem_base + (uint32_t)(nh * 256);
            if (!(q.dbg & 8)) {
#pragma unroll
              for (int k = 0; k < KC / 16; k++)
                umma_bf16<2>(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kc | k) ? 1u : 0u);
            }
            umma_commit<2>(&ctl->empty_b[sb]);
            if (++sb == NB) { sb = 0; pb ^= 1u; }
          }
          if (a_ring) {
            umma_commit<2>(&ctl->empty_a[sa]);
            if (++sa == NA) { sa = 0; pa ^= 1u; }
          }
        }
        umma_commit<2>(&ctl->acc_full);
      };
      for (int t = cluster_id; t < q.m_tiles; t += n_clusters) {
        product(true);
        if (FFN) product(false);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    const int quarter = warp & 3;                 // TMEM lanes 32*quarter .. +31 are accessible to this warp
    const int ehalf = (warp - 2) >> 2;            // which of the two warps sharing this lane quarter
    const int widx = warp - 2;
    const int row = quarter * 32 + lane;          // row inside the CTA's 128-row tile
    const int sub = lane >> 3, cj = lane & 7;     // coalesced role: row 4i + sub, 16-byte chunk cj
    const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t s_stg = smem_u32(smem + C::OFF_STG + widx * 4096);
    // row-statistics exchange between the two warps of a lane quarter: 256 B per warp
    float2* xch_mine = reinterpret_cast<float2*>(FFN ? smem + C::OFF_STG + widx * 4096 : smem + C::OFF_XCH + widx * 256);
    float2* xch_peer = reinterpret_cast<float2*>(FFN ? smem + C::OFF_STG + (widx ^ 4) * 4096 : smem + C::OFF_XCH + (widx ^ 4) * 256);
    uint32_t full_ph = 0;
    uint4 rr[8];
    auto load_res = [&](int mw_, int nb_) {       // residual rows mw_ + 4i + sub, columns nb_ + 8 cj .. + 7
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const int mr = mw_ + 4 * i + sub;
        rr[i] = mr < q.M ? __ldg(reinterpret_cast<const uint4*>(q.residual + (long long)mr * FD + nb_) + cj) : make_uint4(0u, 0u, 0u, 0u);
      }
    };
    for (int t = cluster_id; t < q.m_tiles; t += n_clusters) {
      const int m_w = (t * 2 + (int)rank) * TMR + quarter * 32;      // first row of this warp
      if (FFN) {
        // ---- first product done: h = relu(acc + b1) -> shared memory, K-major with the 128B swizzle (chunk g = K columns
        //      64g .. 64g+63: row r at r * 128, 16-byte piece j at (j ^ (r & 7)) * 16), i.e. exactly what TMA writes
        mbar_wait(&ctl->acc_full, full_ph);
        full_ph ^= 1u;
        tc_fence_after();
#pragma unroll 1
        for (int g = ehalf; g < NKC; g += 2) {
          uint32_t r0[32], r1[32];
          tmem_ld32_nowait(t_row + (uint32_t)(g * 64), r0);
          tmem_ld32_nowait(t_row + (uint32_t)(g * 64 + 32), r1);
          tmem_ld_wait();
          float v[64];
#pragma unroll
          for (int i = 0; i < 32; i++) { v[i] = __uint_as_float(r0[i]); v[32 + i] = __uint_as_float(r1[i]); }
          add_bias64(v, q.b1, g * 64);
          const act2_t zero2 = f2act2(0.0f, 0.0f);
          const uint32_t hb = sH + (uint32_t)(g * CHUNK);
#pragma unroll
          for (int j = 0; j < 8; j++) {
            act2_t h0 = __hmax2(f2act2(v[j * 8 + 0], v[j * 8 + 1]), zero2), h1 = __hmax2(f2act2(v[j * 8 + 2], v[j * 8 + 3]), zero2);
            act2_t h2 = __hmax2(f2act2(v[j * 8 + 4], v[j * 8 + 5]), zero2), h3 = __hmax2(f2act2(v[j * 8 + 6], v[j * 8 + 7]), zero2);
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(hb + stg_off(row, j)), "r"(*reinterpret_cast<uint32_t*>(&h0)),
                         "r"(*reinterpret_cast<uint32_t*>(&h1)), "r"(*reinterpret_cast<uint32_t*>(&h2)), "r"(*reinterpret_cast<uint32_t*>(&h3)) : "memory");
          }
        }
        // generic-proxy writes of h -> visible to the tensor core's async-proxy reads; accumulator reads done
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster_release(&ctl->acc_free, 0);
      }
      // ---- last product done: v = acc + b2 + residual; LayerNorm statistics per row; v parked in TMEM
      load_res(m_w, ehalf * 64);
      mbar_wait(&ctl->acc_full, full_ph);
      full_ph ^= 1u;
      tc_fence_after();
      float shift = 0.0f, s1 = 0.0f, s2 = 0.0f;
#pragma unroll 1
      for (int g = ehalf; g < NKC; g += 2) {
        uint32_t r0[32], r1[32];
        tmem_ld32_nowait(t_row + (uint32_t)(g * 64), r0);
        tmem_ld32_nowait(t_row + (uint32_t)(g * 64 + 32), r1);
        tmem_ld_wait();
        float v[64];
#pragma unroll
        for (int i = 0; i < 32; i++) { v[i] = __uint_as_float(r0[i]); v[32 + i] = __uint_as_float(r1[i]); }
        add_bias64(v, q.b2, g * 64);
        // residual: coalesced registers -> warp-private swizzled tile -> lane = row
#pragma unroll
        for (int i = 0; i < 8; i++)
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(s_stg + stg_off(4 * i + sub, cj)), "r"(rr[i].x), "r"(rr[i].y), "r"(rr[i].z), "r"(rr[i].w) : "memory");
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; j++) {
          uint4 rv;
          asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(rv.x), "=r"(rv.y), "=r"(rv.z), "=r"(rv.w) : "r"(s_stg + stg_off(lane, j)) : "memory");
          const act2_t* h = reinterpret_cast<const act2_t*>(&rv);
#pragma unroll
          for (int e = 0; e < 4; e++) {
            const float2 f = act22f2(h[e]);
            v[j * 8 + e * 2] += f.x; v[j * 8 + e * 2 + 1] += f.y;
          }
        }
        __syncwarp();
        if (g + 2 < NKC) load_res(m_w, (g + 2) * 64);
        if (g == ehalf) shift = v[0];             // shifted sums: no cancellation for rows with a large mean
#pragma unroll
        for (int i = 0; i < 64; i++) { const float d = v[i] - shift; s1 += d; s2 = fmaf(d, d, s2); }
        tmem_st32(t_row + (uint32_t)(g * 64), v);
        tmem_st32(t_row + (uint32_t)(g * 64 + 32), v + 32);
      }
      tmem_st_wait();
      // combine the two half rows (256 columns each): mean_w = shift + s1/256, M2_w = s2 - s1^2/256
      const float mean_w = shift + s1 * (1.0f / 256.0f), m2_w = s2 - s1 * s1 * (1.0f / 256.0f);
      xch_mine[lane] = make_float2(mean_w, m2_w);
      asm volatile("bar.sync %0, 64;" ::"r"(2 + quarter) : "memory");
      const float2 o = xch_peer[lane];
      const float mean = 0.5f * (mean_w + o.x);
      const float dm = mean_w - o.x;
      const float var = (m2_w + o.y + dm * dm * 128.0f) * (1.0f / 512.0f);
      const float rstd = rsqrtf(var + q.eps);
      if (FFN) {
        // staging (and the statistics in it) no longer needed: the producer may refill the A ring underneath it.  The
        // proxy fence orders this warp's generic-proxy accesses before the TMA (async-proxy) writes that follow.
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive_local(&ctl->stg_free);
      }
      // ---- second pass over TMEM
#pragma unroll 1
      for (int g = ehalf; g < NKC; g += 2) {
        uint32_t r0[32], r1[32];
        tmem_ld32_nowait(t_row + (uint32_t)(g * 64), r0);
        tmem_ld32_nowait(t_row + (uint32_t)(g * 64 + 32), r1);
        tmem_ld_wait();
        if (g + 2 >= NKC) {                        // last TMEM read of this warp: hand the accumulator back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(&ctl->acc_free, 0);
        }
        float v[64];
#pragma unroll
        for (int i = 0; i < 32; i++) { v[i] = (__uint_as_float(r0[i]) - mean) * rstd; v[32 + i] = (__uint_as_float(r1[i]) - mean) * rstd; }
        if (FFN) {
          // column sums over the 16 rows of each half warp: halving exchange, lane l (mod 16) ends with columns 4l .. 4l+3
          const unsigned FULL = 0xffffffffu;
          float a[32], b[16], c[8], d[4];
          const bool b3 = lane & 8, b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
#pragma unroll
          for (int i = 0; i < 32; i++) { const float keep = b3 ? v[32 + i] : v[i], send = b3 ? v[i] : v[32 + i]; a[i] = keep + __shfl_xor_sync(FULL, send, 8); }
#pragma unroll
          for (int i = 0; i < 16; i++) { const float keep = b2 ? a[16 + i] : a[i], send = b2 ? a[i] : a[16 + i]; b[i] = keep + __shfl_xor_sync(FULL, send, 4); }
#pragma unroll
          for (int i = 0; i < 8; i++) { const float keep = b1 ? b[8 + i] : b[i], send = b1 ? b[i] : b[8 + i]; c[i] = keep + __shfl_xor_sync(FULL, send, 2); }
#pragma unroll
          for (int i = 0; i < 4; i++) { const float keep = b0 ? c[4 + i] : c[i], send = b0 ? c[i] : c[4 + i]; d[i] = keep + __shfl_xor_sync(FULL, send, 1); }
          const int mg = m_w + (lane >> 4) * 16;   // first row of this half warp's group
          if (mg < q.M && !(q.dbg & 1))
            *reinterpret_cast<float4*>(q.gs + (long long)(mg >> 4) * FD + g * 64 + (lane & 15) * 4) = make_float4(d[0], d[1], d[2], d[3]);
        } else {
          const float4* g4 = reinterpret_cast<const float4*>(q.gamma + g * 64);
          const float4* b4 = reinterpret_cast<const float4*>(q.beta + g * 64);
#pragma unroll
          for (int c4 = 0; c4 < 16; c4++) {
            const float4 gm = __ldg(g4 + c4), bt = __ldg(b4 + c4);
            v[c4 * 4 + 0] = fmaf(v[c4 * 4 + 0], gm.x, bt.x); v[c4 * 4 + 1] = fmaf(v[c4 * 4 + 1], gm.y, bt.y);
            v[c4 * 4 + 2] = fmaf(v[c4 * 4 + 2], gm.z, bt.z); v[c4 * 4 + 3] = fmaf(v[c4 * 4 + 3], gm.w, bt.w);
          }
#pragma unroll
          for (int j = 0; j < 8; j++) {
            act2_t h0 = f2act2(v[j * 8 + 0], v[j * 8 + 1]), h1 = f2act2(v[j * 8 + 2], v[j * 8 + 3]);
            act2_t h2 = f2act2(v[j * 8 + 4], v[j * 8 + 5]), h3 = f2act2(v[j * 8 + 6], v[j * 8 + 7]);
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(s_stg + stg_off(lane, j)), "r"(*reinterpret_cast<uint32_t*>(&h0)),
                         "r"(*reinterpret_cast<uint32_t*>(&h1)), "r"(*reinterpret_cast<uint32_t*>(&h2)), "r"(*reinterpret_cast<uint32_t*>(&h3)) : "memory");
          }
          __syncwarp();
          char* ob = (char*)(q.out + (long long)(m_w + sub) * FD + g * 64 + cj * 8);
          const int rows_left = q.M - m_w - sub;   // row 4i + sub is stored iff 4i < rows_left
#pragma unroll
          for (int i = 0; i < 8; i++) {
            uint4 ov;
            asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(ov.x), "=r"(ov.y), "=r"(ov.z), "=r"(ov.w) : "r"(s_stg + stg_off(4 * i + sub, cj)) : "memory");
            if (4 * i < rows_left && !(q.dbg & 1)) *reinterpret_cast<uint4*>(ob + (size_t)i * (4 * FD * sizeof(act_t))) = ov;
          }
          __syncwarp();
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// pooled[h][c] = gamma[c] * (sum of the `per` group rows of hypothesis h, in order) / tokens + beta[c]
__global__ void __launch_bounds__(128) pool_groups_kernel(const float* __restrict__ gs, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float* __restrict__ out, int per,
                                                          float inv_tokens)
{
  const int c = threadIdx.x * 4;
  const float4* src = reinterpret_cast<const float4*>(gs + (long long)blockIdx.x * per * FD + c);
  float4 s = src[0];
  for (int j = 1; j < per; j++) {
    const float4 x = src[(long long)j * (FD / 4)];
    s.x += x.x; s.y += x.y; s.z += x.z; s.w += x.w;
  }
  const float4 gm = *reinterpret_cast<const float4*>(gamma + c), bt = *reinterpret_cast<const float4*>(beta + c);
  *reinterpret_cast<float4*>(out + (long long)blockIdx.x * FD + c) =
      make_float4(fmaf(s.x * inv_tokens, gm.x, bt.x), fmaf(s.y * inv_tokens, gm.y, bt.y), fmaf(s.z * inv_tokens, gm.z, bt.z),
                  fmaf(s.w * inv_tokens, gm.w, bt.w));
}

int ffn_enabled()
{
  static int v = -1;
  if (v < 0) { const char* e = getenv("FPB_FFN"); v = e ? atoi(e) : 1; }
  return v;
}

template <bool FFN>
int launch_ffn(const CUtensorMap& mA, const CUtensorMap& mW1, const CUtensorMap& mW2, const FfnParams& q, int num_sms, cudaStream_t st)
{
  using C = FCfg<FFN>;
  static bool configured = false;
  if (!configured) {
    FPB_CUDA(cudaFuncSetAttribute(ffn_kernel<FFN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    configured = true;
  }
  int clusters = num_sms / 2;
  if (clusters > q.m_tiles) clusters = q.m_tiles;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(clusters * 2, 1, 1);
  cfg.blockDim = dim3(NTHREADS, 1, 1);
  cfg.dynamicSmemBytes = C::SMEM;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = fpb_pdl_enabled() ? 2 : 1;
  FPB_CUDA(cudaLaunchKernelEx(&cfg, ffn_kernel<FFN>, mA, mW1, mW2, q));
  fpb_count_launch();
  return FPB_OK;
}

bool aligned16(const void* p) { return (((uintptr_t)p) & 15) == 0; }

}  // namespace

// y = LayerNorm(A . W^T + b + residual) * gamma + beta; A, residual, y: [M][512] 16-bit rows, W: [512][512] (K contiguous)
int proj_ln_512(const void* A, const void* W, const float* b, const void* residual, const float* gamma, const float* beta, float eps,
                void* out, long long M, cudaStream_t st)
{
  if (!ffn_enabled() || M <= 0 || M > 0x7fffff00LL || !aligned16(A) || !aligned16(W) || !aligned16(residual) || !aligned16(out) ||
      !aligned16(b) || !aligned16(gamma) || !aligned16(beta))
    return FPB_ERR_UNSUPPORTED;
  CUtensorMap mA, mW;
  int num_sms = 0;
  int rc = fpb_tiled_map_act(A, M, FD, FD, TMR, &mA, &num_sms);
  if (rc) return rc;
  rc = fpb_tiled_map_act(W, FD, FD, FD, TMR, &mW, nullptr);
  if (rc) return rc;
  FfnParams q;
  memset(&q, 0, sizeof(q));
  q.M = (int)M; q.m_tiles = fpb_cdiv(M, 2 * TMR);
  q.residual = (const act_t*)residual; q.b2 = b; q.gamma = gamma; q.beta = beta; q.eps = eps; q.out = (act_t*)out;
  { const char* e = getenv("FPB_FFN_DBG"); q.dbg = e ? atoi(e) : 0; }
  return launch_ffn<false>(mA, mW, mW, q, num_sms, st);
}

// pooled[g] = mean over the `tokens` rows of group g of LayerNorm(x + relu(x . W1^T + b1) . W2^T + b2) * gamma + beta
// x: [groups * tokens][512] 16-bit rows; gs: scratch of (groups * tokens / 16) x 512 floats; tokens % 16 == 0
int ffn_ln_token_mean_512(const void* x, const void* W1, const float* b1, const void* W2, const float* b2, const float* gamma,
                          const float* beta, float eps, float* gs, float* pooled, int groups, int tokens, cudaStream_t st)
{
  const long long M = (long long)groups * tokens;
  if (!ffn_enabled() || groups <= 0 || tokens <= 0 || tokens % 16 != 0 || M > 0x7fffff00LL || !aligned16(x) || !aligned16(W1) ||
      !aligned16(W2) || !aligned16(b1) || !aligned16(b2) || !aligned16(gamma) || !aligned16(beta) || !aligned16(gs) || !aligned16(pooled))
    return FPB_ERR_UNSUPPORTED;
  CUtensorMap mA, mW1, mW2;
  int num_sms = 0;
  int rc = fpb_tiled_map_act(x, M, FD, FD, TMR, &mA, &num_sms);
  if (rc) return rc;
  rc = fpb_tiled_map_act(W1, FD, FD, FD, TMR, &mW1, nullptr);
  if (rc) return rc;
  rc = fpb_tiled_map_act(W2, FD, FD, FD, TMR, &mW2, nullptr);
  if (rc) return rc;
  FfnParams q;
  memset(&q, 0, sizeof(q));
  q.M = (int)M; q.m_tiles = fpb_cdiv(M, 2 * TMR);
  q.residual = (const act_t*)x; q.b1 = b1; q.b2 = b2; q.eps = eps; q.gs = gs;
  { const char* e = getenv("FPB_FFN_DBG"); q.dbg = e ? atoi(e) : 0; }
  rc = launch_ffn<true>(mA, mW1, mW2, q, num_sms, st);
  if (rc) return rc;
  pool_groups_kernel<<<groups, 128, 0, st>>>(gs, gamma, beta, pooled, tokens / 16, 1.0f / (float)tokens);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

}  // namespace FPB_ACT_NS
