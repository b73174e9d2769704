// Fused multi-head self-attention over the 400 tokens of one hypothesis (d_model 512, 4 heads of 128) on tcgen05.
//   TransformerEncoderLayer.self_attn   learning/models/refine_network.py:56-70   (via nn.MultiheadAttention)
//   ScoreNetMultiPair.att               learning/models/score_network.py:53,73
// One CTA per (hypothesis, head); it walks the 4 query tiles of 128 rows:
//   S = Q K^T      two tcgen05.mma (N = 208 + 192) per 16-wide k step, fp32 scores in TMEM columns 0..399
//   P = exp2(..)   8 softmax warps (lane = query row, two warps per TMEM lane quarter taking alternate 64-key chunks):
//                  pass 1 row max (halves combined through smem), pass 2 un-normalised probabilities written as bf16
//                  into a 2-stage smem ring in the 128B-swizzled K-major layout UMMA reads
//   O = P V'       tcgen05.mma (N = 144) per 64-key chunk against V^T tiles (TMA ring) extended by a constant row of
//                  ones, so column 128 of the accumulator is the softmax denominator computed from the same bf16 P as
//                  the numerator (no per-element adds in the softmax warps).  O overlays TMEM columns 368..511, i.e.
//                  the score columns of the two LAST key chunks; chunks are therefore processed in reverse order and
//                  the first P.V MMA waits until both have been consumed.
//   epilogue       O / rowsum (+ V bias when not already folded into V^T: softmax rows sum to 1, so the bias commutes)
// No score / probability tensor ever reaches HBM (the unfused path wrote 0.7 GB of fp32 scores per layer call).
// Warp roles: 0 TMA producer, 1 TMEM allocator + MMA issuer, 2-9 softmax / epilogue.
// Measured: with 4 softmax warps (one per scheduler) every tcgen05.wait / mbarrier wait was exposed and the kernel
// ran at 272 us per call with all MMAs and exps disabled it still took 168 us; see DESIGN.md 4.3.
#include "net_kernels.cuh"
#include "act_type.cuh"
#include <cuda.h>
#include <stdlib.h>

namespace FPB_ACT_NS {

namespace {

constexpr int TOK = 400, DH = 128, HEADS = 4, DM = 512;
constexpr int NTHREADS = 320;
constexpr int K_CHUNK_BYTES = TOK * 128;          // one 64-wide d chunk of K: 400 rows x 128 B
constexpr int Q_CHUNK_BYTES = 128 * 128;
constexpr int T16 = 16384;
constexpr int N_KEY_CHUNKS = 7;                   // 6 x 64 + 16 keys
constexpr int N1 = 208, N2 = 192;                 // S = [N1 | N2] columns
constexpr int NO = 144;                           // P.V accumulator columns: 128 head dims + the ones row (+ pad to 16)
constexpr int O_COL = 512 - NO;                   // 368: O overlays the scores of key chunks 5 (partly) and 6
constexpr int V_STAGE = T16 + 2048;               // V^T tile + 16 constant rows (row 128 = ones)
constexpr int NV = 2;                             // ring depth (a third V^T stage measured no gain: the stream is not the bound)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
  const uint32_t addr = smem_u32(bar);
  uint32_t ok = 0;
  long long t0 = 0;
  for (uint32_t spin = 0; !ok; spin++) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
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
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr)      // K-major, SWIZZLE_128B (see net_tc.cu)
{
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t make_idesc(int n)
{
  return (1u << 4) | (ACT_IDESC_FMT << 7) | (ACT_IDESC_FMT << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum)
{
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar)
{
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int x, int y)
{
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y)
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
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, uint32_t* r)
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
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float tmem_ld1(uint32_t taddr)
{
  uint32_t r;
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  return __uint_as_float(r);
}
__device__ __forceinline__ float ex2(float x)
{
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[32])
{
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct AttnCtl {
  uint64_t k_full, q_full, q_empty, s_full, o_full, tmem_free;
  uint64_t slot_full[2], slot_empty[2];       // ring slot = V^T tile (TMA) + P tile (softmax warps)
  uint32_t tmem_base;
};

constexpr int SMEM_K = 2 * K_CHUNK_BYTES, SMEM_Q = 2 * Q_CHUNK_BYTES, SMEM_V = NV * V_STAGE, SMEM_P = 2 * T16;
constexpr int SMEM_MAX = 2 * 2 * 128 * 4;         // partial row maxima: [tile parity][warp set][row]
constexpr int SMEM_STAGE = 8 * 2048;              // context staging: 2 KB per softmax warp
constexpr int SMEM_TOTAL = SMEM_K + SMEM_Q + SMEM_V + SMEM_P + SMEM_MAX + SMEM_STAGE + 1024 + (int)sizeof(AttnCtl);
static_assert(SMEM_TOTAL <= 227 * 1024, "attention shared memory budget");

// DBG (FPB_ATTN_DBG, timing experiments only; results are wrong when != 0): 1 no row-max pass, 2 no exp2, 4 no P.V MMAs,
// 8 no ctx stores, 16 no S MMAs
template <int DBG>
__global__ void __launch_bounds__(NTHREADS, 1)
attn_kernel(const __grid_constant__ CUtensorMap mapQ, const __grid_constant__ CUtensorMap mapK,
            const __grid_constant__ CUtensorMap mapV, const float* __restrict__ bv, act_t* __restrict__ ctx,
            float c1)
{
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sK = smem;
  uint8_t* sQ = sK + SMEM_K;
  uint8_t* sV = sQ + SMEM_Q;
  uint8_t* sP = sV + SMEM_V;
  float* smax = (float*)(sP + SMEM_P);
  uint8_t* sStage = sP + SMEM_P + SMEM_MAX;
  AttnCtl* ctl = (AttnCtl*)(sStage + SMEM_STAGE);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = blockIdx.x / HEADS, h = blockIdx.x - g * HEADS;
  const int row0 = g * TOK;                                  // first token row of this hypothesis
  constexpr int N_MT = (TOK + 127) / 128;                    // 4 query tiles
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // programmatic dependent launch, see tc_ptx.cuh

  if (threadIdx.x == 0) {
    mbar_init(&ctl->k_full, 1); mbar_init(&ctl->q_full, 1); mbar_init(&ctl->q_empty, 1);
    mbar_init(&ctl->s_full, 1); mbar_init(&ctl->o_full, 1); mbar_init(&ctl->tmem_free, 8);
    for (int s = 0; s < 2; s++) { mbar_init(&ctl->slot_full[s], 129); mbar_init(&ctl->slot_empty[s], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = threadIdx.x; i < NV * 128; i += NTHREADS) {
    // the 16 constant rows behind each V^T stage: row 128 = ones (its dot product with P is the row sum), the rest 0
    const int s = i >> 7, o = (i & 127) * 16;
    const uint32_t w = o < 128 ? ACT_ONE_X2 : 0u;
    *reinterpret_cast<uint4*>(sV + s * V_STAGE + T16 + o) = make_uint4(w, w, w, w);
  }
  fence_proxy_async();
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ctl->tmem_base)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;
  asm volatile("griddepcontrol.wait;" ::: "memory");         // predecessor outputs (Q|K, V^T) complete from here on
  // DBG & 32: block 0 records clock64 stamps (warp 1 lane 0 -> slots 0..63, warp 2 lane 0 -> 64..127, warp 6 lane 0 -> 128..191)
  long long* stamps = reinterpret_cast<long long*>(ctx);
  int n_st = 0;
  auto stamp = [&](int base) {
    if ((DBG & 32) && blockIdx.x == 0 && lane == 0 && n_st < 64 && base < 1000) stamps[base + n_st++] = clock64();
  };

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (lane == 0) {
      mbar_arrive_expect_tx(&ctl->k_full, (uint32_t)SMEM_K);
      for (int kc = 0; kc < 2; kc++)
        for (int half = 0; half < 2; half++)
          tma_load_2d(smem_u32(sK + kc * K_CHUNK_BYTES + half * 200 * 128), &mapK, &ctl->k_full, DM + h * DH + kc * 64,
                      row0 + half * 200);
      uint32_t gi = 0;
      for (int mt = 0; mt < N_MT; mt++) {
        mbar_wait(&ctl->q_empty, (uint32_t)((mt & 1) ^ 1));
        mbar_arrive_expect_tx(&ctl->q_full, (uint32_t)SMEM_Q);
        for (int kc = 0; kc < 2; kc++)
          tma_load_2d(smem_u32(sQ + kc * Q_CHUNK_BYTES), &mapQ, &ctl->q_full, h * DH + kc * 64, row0 + mt * 128);
        for (int oi = 0; oi < N_KEY_CHUNKS; oi++, gi++) {       // key chunks in reverse order (see header)
          const uint32_t s = gi & 1u, use = gi >> 1;
          mbar_wait(&ctl->slot_empty[s], (use & 1u) ^ 1u);
          mbar_arrive_expect_tx(&ctl->slot_full[s], (uint32_t)T16);
          tma_load_2d(smem_u32(sV + s * V_STAGE), &mapV, &ctl->slot_full[s], row0 + (N_KEY_CHUNKS - 1 - oi) * 64, h * DH);
        }
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    // Measured with clock stamps (FPB_ATTN_DBG=32): this thread's per-chunk turnaround (two waits, two commits) was
    // ~1000 cycles and paced the whole CTA, so a ring slot couples the V^T tile and the P tile behind ONE full barrier
    // (128 softmax arrivals + the TMA transaction) and ONE empty barrier, and all descriptors are hoisted.
    if (lane == 0) {
      constexpr uint32_t id1 = make_idesc(N1), id2 = make_idesc(N2), idv = make_idesc(NO);
      const uint64_t dq0 = make_desc(smem_u32(sQ)), dq1 = make_desc(smem_u32(sQ + Q_CHUNK_BYTES));
      const uint64_t dka0 = make_desc(smem_u32(sK)), dka1 = make_desc(smem_u32(sK + K_CHUNK_BYTES));
      const uint64_t dkb0 = make_desc(smem_u32(sK + N1 * 128)), dkb1 = make_desc(smem_u32(sK + K_CHUNK_BYTES + N1 * 128));
      const uint64_t dps[2] = {make_desc(smem_u32(sP)), make_desc(smem_u32(sP + T16))};
      const uint64_t dvs[2] = {make_desc(smem_u32(sV)), make_desc(smem_u32(sV + V_STAGE))};
      // scores of keys 0..207 (TMEM columns 0..207) / keys 208..399 (columns 208..399)
      auto issue_s1 = [&]() {
        if (DBG & 16) return;
#pragma unroll
        for (int k = 0; k < 4; k++) umma(tmem_base, dq0 + (uint64_t)(k * 2), dka0 + (uint64_t)(k * 2), id1, k ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 4; k++) umma(tmem_base, dq1 + (uint64_t)(k * 2), dka1 + (uint64_t)(k * 2), id1, 1u);
      };
      auto issue_s2 = [&]() {
        if (DBG & 16) return;
#pragma unroll
        for (int k = 0; k < 4; k++) umma(tmem_base + N1, dq0 + (uint64_t)(k * 2), dkb0 + (uint64_t)(k * 2), id2, k ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 4; k++) umma(tmem_base + N1, dq1 + (uint64_t)(k * 2), dkb1 + (uint64_t)(k * 2), id2, 1u);
      };
      mbar_wait(&ctl->k_full, 0);
      mbar_wait(&ctl->q_full, 0);
      tc_fence_after();
      issue_s1();
      uint32_t gi = 0;
      for (int mt = 0; mt < N_MT; mt++) {
        // the first half of this tile's scores was issued early (below); the second half overlaps the previous tile's
        // O columns and has to wait for its epilogue
        mbar_wait(&ctl->tmem_free, (uint32_t)((mt & 1) ^ 1));
        tc_fence_after();
        issue_s2();
        stamp(0);                                             // S MMAs issued
        umma_commit(&ctl->s_full);
        umma_commit(&ctl->q_empty);
        // O overlays score columns 368..399 (key chunks 6 and 5 = the first two in processing order): both must
        // have been turned into P before the first P.V MMA
        mbar_wait(&ctl->slot_full[gi & 1u], (gi >> 1) & 1u);
        mbar_wait(&ctl->slot_full[(gi + 1) & 1u], ((gi + 1) >> 1) & 1u);
        tc_fence_after();
        if (!(DBG & 4)) umma(tmem_base + O_COL, dps[gi & 1u], dvs[gi & 1u], idv, 0u);   // 16-key tail chunk: one k step
        stamp(0);
        umma_commit(&ctl->slot_empty[gi & 1u]);
        gi++;
#pragma unroll 1
        for (int oi = 1; oi < N_KEY_CHUNKS; oi++, gi++) {
          const uint32_t s = gi & 1u;
          mbar_wait(&ctl->slot_full[s], (gi >> 1) & 1u);
          tc_fence_after();
          if (!(DBG & 4)) {
#pragma unroll
            for (int k = 0; k < 4; k++) umma(tmem_base + O_COL, dps[s] + (uint64_t)(k * 2), dvs[s] + (uint64_t)(k * 2), idv, 1u);
          }
          stamp(0);                                           // P.V MMAs of chunk oi issued
          umma_commit(&ctl->slot_empty[s]);
        }
        umma_commit(&ctl->o_full);
        if (mt + 1 < N_MT) {
          // every softmax warp has consumed this tile's scores (all slot_full waits passed): the next tile's first 208
          // score columns do not touch the O columns and can be computed under this tile's epilogue
          mbar_wait(&ctl->q_full, (uint32_t)((mt + 1) & 1));
          tc_fence_after();
          issue_s1();
        }
      }
    }
  } else {
    // ---------------------------------------------------------------- softmax + epilogue (warps 2..9)
    const int quarter = warp & 3;                 // TMEM lanes (query rows) 32*quarter .. +31
    const int set = (warp - 2) >> 2;              // set 0: key chunks 6,4,2,0; set 1: key chunks 5,3,1
    const int row = quarter * 32 + lane;
    const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16);
    for (int mt = 0; mt < N_MT; mt++) {
      stamp(warp == 2 ? 64 : (warp == 6 ? 128 : 10000000));
      mbar_wait(&ctl->s_full, (uint32_t)(mt & 1));
      tc_fence_after();
      stamp(warp == 2 ? 64 : (warp == 6 ? 128 : 10000000));  // scores ready
      // ---- pass 1: row max over this warp's key chunks, then combined with the partner warp's
      float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll 1
      for (int c = 1 - set; c < ((DBG & 1) ? 0 : N_KEY_CHUNKS - 1); c += 2) {
        uint32_t rb[64];
        tmem_ld32_nowait(t_row + (uint32_t)(c * 64), rb);
        tmem_ld32_nowait(t_row + (uint32_t)(c * 64 + 32), rb + 32);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 64; i++) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(rb[i]));
      }
      if (set == 0) {
        uint32_t r[32];
        tmem_ld16(t_row + (uint32_t)((N_KEY_CHUNKS - 1) * 64), r);
#pragma unroll
        for (int i = 0; i < 16; i++) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(r[i]));
      }
      float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      float* sm = smax + (mt & 1) * 256;
      sm[set * 128 + row] = mx;
      asm volatile("bar.sync %0, 64;" ::"r"(1 + quarter) : "memory");
      mx = fmaxf(mx, sm[(set ^ 1) * 128 + row]);
      const float off = -mx * c1;
      stamp(warp == 2 ? 64 : (warp == 6 ? 128 : 10000000));  // row max done
      // ---- pass 2: un-normalised probabilities of this warp's chunks -> bf16 P tiles in the smem ring
#pragma unroll 1
      for (int oi = set; oi < N_KEY_CHUNKS; oi += 2) {
        const uint32_t gi = (uint32_t)(mt * N_KEY_CHUNKS + oi), s = gi & 1u, use = gi >> 1;
        const int c = N_KEY_CHUNKS - 1 - oi;
        const bool tail = oi == 0;
        uint32_t r2[64];
        if (tail) {
          uint32_t r[32];
          tmem_ld16(t_row + (uint32_t)(c * 64), r);
#pragma unroll
          for (int i = 0; i < 16; i++) r2[i] = r[i];
        } else {
          tmem_ld32_nowait(t_row + (uint32_t)(c * 64), r2);
          tmem_ld32_nowait(t_row + (uint32_t)(c * 64 + 32), r2 + 32);
        }
        mbar_wait(&ctl->slot_empty[s], (use & 1u) ^ 1u);     // overlaps the TMEM load latency
        tmem_ld_wait();
        const uint32_t sp = smem_u32(sP + s * T16) + (uint32_t)(row * 128);
        const int n16 = tail ? 2 : 8;                          // 16-byte chunks (8 keys each) of this key chunk
#pragma unroll
        for (int jj = 0; jj < 8; jj++) {
          if (jj < n16) {
            uint32_t pk[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
              float p0 = fmaf(__uint_as_float(r2[jj * 8 + e * 2]), c1, off), p1 = fmaf(__uint_as_float(r2[jj * 8 + e * 2 + 1]), c1, off);
              if (!(DBG & 2)) { p0 = ex2(p0); p1 = ex2(p1); }
              act2_t hh = f2act2(p0, p1);
              pk[e] = *reinterpret_cast<uint32_t*>(&hh);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(sp + (uint32_t)((jj ^ (row & 7)) << 4)), "r"(pk[0]),
                         "r"(pk[1]), "r"(pk[2]), "r"(pk[3])
                         : "memory");
          }
        }
        tc_fence_before();
        fence_proxy_async();
        mbar_arrive(&ctl->slot_full[s]);
        stamp(warp == 2 ? 64 : (warp == 6 ? 128 : 10000000));  // P chunk written
      }
      // ---- epilogue: O / rowsum (+ b_v) -> ctx; set 0 writes head dims 0..63, set 1 dims 64..127
      mbar_wait(&ctl->o_full, (uint32_t)(mt & 1));
      tc_fence_after();
      stamp(warp == 2 ? 64 : (warp == 6 ? 128 : 10000000));  // O ready
      const float inv = 1.0f / tmem_ld1(t_row + (uint32_t)(O_COL + DH));
      uint32_t o[64];
      tmem_ld32_nowait(t_row + (uint32_t)(O_COL + set * 64), o);
      tmem_ld32_nowait(t_row + (uint32_t)(O_COL + set * 64 + 32), o + 32);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl->tmem_free);          // all TMEM reads of this warp for this tile are done
      // 32 rows x 32 head dims (64 B) at a time are transposed through this warp's private 2 KB staging tile so that 4
      // lanes store one contiguous 64-byte row segment (two full sectors)
      const uint32_t stg = smem_u32(sStage) + (uint32_t)((warp - 2) * 2048);
      const int srow = lane >> 2, sc = lane & 3;
#pragma unroll
      for (int half = 0; half < 2; half++) {
#pragma unroll
        for (int q8 = 0; q8 < 4; q8++) {
          uint32_t pw[4];
#pragma unroll
          for (int e = 0; e < 4; e++) {
            const int d = half * 32 + q8 * 8 + e * 2;
            float a = __uint_as_float(o[d]) * inv, b = __uint_as_float(o[d + 1]) * inv;
            if (bv) { a += __ldg(bv + h * DH + set * 64 + d); b += __ldg(bv + h * DH + set * 64 + d + 1); }
            act2_t hh = f2act2(a, b);
            pw[e] = *reinterpret_cast<uint32_t*>(&hh);
          }
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(stg + (uint32_t)(lane * 64 + ((q8 ^ ((lane >> 1) & 3)) << 4))), "r"(pw[0]),
                       "r"(pw[1]), "r"(pw[2]), "r"(pw[3]) : "memory");
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const int rr = 8 * i + srow;                        // row within this warp's 32 query rows
          uint4 ov;
          asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(ov.x), "=r"(ov.y), "=r"(ov.z), "=r"(ov.w)
                       : "r"(stg + (uint32_t)(rr * 64 + ((sc ^ ((rr >> 1) & 3)) << 4))) : "memory");
          const int tk = mt * 128 + quarter * 32 + rr;
          if (tk < TOK && !(DBG & (8 | 32)))
            *reinterpret_cast<uint4*>(ctx + (long long)(row0 + tk) * DM + h * DH + set * 64 + half * 32 + sc * 8) = ov;
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

}  // namespace

// bv may be null when the V bias is already part of vt (folded into the V^T projection, see net.cu)
int tc_attention(const void* qk, const void* vt, long long vt_pitch, const float* bv, void* ctx, int groups, cudaStream_t st)
{
  FPB_CHECK_ARG(qk && vt && ctx && groups > 0 && vt_pitch % 8 == 0);
  const long long R = (long long)groups * TOK;
  CUtensorMap mQ, mK, mV;
  int rc;
  if ((rc = fpb_tiled_map_act(qk, R, 2 * DM, 2 * DM, 128, &mQ, nullptr))) return rc;
  if ((rc = fpb_tiled_map_act(qk, R, 2 * DM, 2 * DM, 200, &mK, nullptr))) return rc;
  if ((rc = fpb_tiled_map_act(vt, DM, R, vt_pitch, 128, &mV, nullptr))) return rc;
  static int dbg = -1;
  if (dbg < 0) { const char* e = getenv("FPB_ATTN_DBG"); dbg = e ? atoi(e) : 0; }
  const float c1 = 0.08838834764831845f * 1.4426950408889634f;   // 1/sqrt(128) * log2(e)
#define ATTN_LAUNCH(D)                                                                                             \
  do {                                                                                                             \
    static bool configured = false;                                                                                \
    if (!configured) {                                                                                             \
      FPB_CUDA(cudaFuncSetAttribute(attn_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL));     \
      configured = true;                                                                                           \
    }                                                                                                              \
    cudaLaunchConfig_t cfg;                                                                                        \
    memset(&cfg, 0, sizeof(cfg));                                                                                  \
    cfg.gridDim = dim3(groups * HEADS, 1, 1);                                                                      \
    cfg.blockDim = dim3(NTHREADS, 1, 1);                                                                           \
    cfg.dynamicSmemBytes = SMEM_TOTAL;                                                                             \
    cfg.stream = st;                                                                                               \
    cudaLaunchAttribute attr[1];                                                                                   \
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                               \
    attr[0].val.programmaticStreamSerializationAllowed = 1;                                                        \
    cfg.attrs = attr;                                                                                              \
    cfg.numAttrs = fpb_pdl_enabled() ? 1 : 0;                                                                      \
    FPB_CUDA(cudaLaunchKernelEx(&cfg, attn_kernel<D>, mQ, mK, mV, bv, (act_t*)ctx, c1));                    \
  } while (0)
  switch (dbg) {
    case 0: ATTN_LAUNCH(0); break;
    case 1: ATTN_LAUNCH(1); break;
    case 2: ATTN_LAUNCH(2); break;
    case 4: ATTN_LAUNCH(4); break;
    case 16: ATTN_LAUNCH(16); break;
    case 20: ATTN_LAUNCH(20); break;
    case 23: ATTN_LAUNCH(23); break;
    case 31: ATTN_LAUNCH(31); break;
    case 32: ATTN_LAUNCH(32); break;
    default: fpb_set_error("FPB_ATTN_DBG=%d is not an instantiated experiment", dbg); return FPB_ERR_ARG;
  }
#undef ATTN_LAUNCH
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

}  // namespace FPB_ACT_NS
