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
// epilogue (two warps per TMEM lane quarter; each owns two 64-column groups in either accumulator half).  TMEM holds ONE
// 512-column accumulator, handled as two 256-column halves with their own full / free barriers: products with a streamed A
// operand fill half 0, then half 1 (A passes through the ring twice), so the epilogue work on half 0 runs under the MMAs
// of half 1 and the next tile's half 0 under the second pass of this tile's half 1.  Residual rows never go through
// registers + staging: every epilogue warp TMA-loads its own [32 rows x 64 columns] boxes into the R/H region (proj+LN:
// half 0 a tile ahead, half 1 when the first pass over half 0 is done; FFN: chunk by chunk as the second product releases
// the h tile), reads them lane = row, writes the output rows over them and TMA-stores the boxes.
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
constexpr int BOX = 32 * 128;              // one warp's [32 rows x 64 columns] piece of a chunk, 4 KB
constexpr int NTHREADS = 320;
constexpr int NSTG_MAX = 5;                // operand ring: stages of two 16 KB blocks (3 beside the h tile, 5 in proj+LN)

struct FfnParams {
  int M, m_tiles;                          // rows, tiles of 256 rows
  const float* b1;                         // FFN: bias of the first product
  const float* b2;                         // bias of the (last) product
  const float* gamma;                      // proj+LN only (the FFN kernel leaves gamma / beta to pool_groups_kernel)
  const float* beta;
  float eps;
  float* gs;                               // FFN: [M/16][512] column sums of the normalised rows of each 16-row group
  int dbg;
};

// FPB_FFN_DBG & 2048: the first epilogue warp of CTA 0 records clock64 stamps of its phases here (dev tool:
// fpb_debug_ffn_stamps copies them out; tools/ffn_bench.py --stamps)
__device__ long long g_ffn_stamps[512];

struct FfnCtl {
  uint64_t full[NSTG_MAX], empty[NSTG_MAX];
  uint64_t acc_full[2], acc_free[2];       // per accumulator half (TMEM columns 0-255 / 256-511)
  uint64_t h_free[NKC];                    // FFN: K chunk kc of the h tile has been consumed by the second product
  uint64_t r_full[8][2];                   // per epilogue warp and column half: its residual boxes have landed
  uint32_t tmem_base;
};

// Shared memory: [R/H region][ring: NSTG x 32 KB][ctl].  The R/H region holds, per chunk (64 columns) and row r, 128 bytes
// at chunk*CHUNK + r*128 with the 16-byte pieces XOR-swizzled by (r & 7) -- the layout TMA writes / reads with SWIZZLE_128B
// and tcgen05.mma reads as a K-major operand.
//   FFN:     8 chunks (128 KB): the h tile (A operand of the second product, written by the epilogue threads), then,
//            chunk by chunk as the second product releases it, the residual boxes.  3 ring stages fit beside it.
//   proj+LN: 4 chunks (64 KB) used in turn for the residual boxes of accumulator half 0, those of half 1 (loaded when the
//            first pass over half 0 has consumed the former), the output boxes of half 0 and the output boxes of half 1
//            (group g lives in chunk g & 3).  The 64 KB this saves make the ring 5 stages deep: the kernel is bound by
//            bytes in flight, so ring depth is throughput.
template <bool FFN>
struct FLayout {
  static constexpr int R_CHUNKS = FFN ? NKC : NKC / 2;
  static constexpr int NSTG = FFN ? 3 : 5;
  static constexpr int OFF_RING = R_CHUNKS * CHUNK;
  static constexpr int DATA_BYTES = OFF_RING + NSTG * 2 * CHUNK;
  static constexpr int SMEM = DATA_BYTES + 1024 + (int)sizeof(FfnCtl);
  static_assert(NSTG <= NSTG_MAX && SMEM <= 232448, "shared memory budget");
};

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
// TMA store of one [32 rows x 64 columns] box from shared memory (rows beyond the tensor are clipped)
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int x, int y)
{
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map), "r"(src), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// mapA: rows of the streamed A operand, box 64 x 128.  mapW1 / mapW2: weights [512][512], box 64 x 128.
// mapR: residual rows, box 64 x 32.  mapO: output rows (proj+LN), box 64 x 32.
template <bool FFN>
__global__ void __launch_bounds__(NTHREADS, 1)
ffn_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapW1,
           const __grid_constant__ CUtensorMap mapW2, const __grid_constant__ CUtensorMap mapR,
           const __grid_constant__ CUtensorMap mapO, const __grid_constant__ FfnParams q)
{
  using L = FLayout<FFN>;
  constexpr int NSTG = L::NSTG;
  constexpr int CMASK = L::R_CHUNKS - 1;        // chunk of group g in the R/H region: g & CMASK
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  FfnCtl* ctl = (FfnCtl*)(smem + L::DATA_BYTES);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);     // provably warp-uniform: constant-bank indexing
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  pdl_launch_dependents();

  if (threadIdx.x == 0) {
    for (int s = 0; s < NSTG; s++) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
    for (int b = 0; b < 2; b++) { mbar_init(&ctl->acc_full[b], 1); mbar_init(&ctl->acc_free[b], 16); }   // 8 warps x 2 CTAs
    for (int k = 0; k < NKC; k++) mbar_init(&ctl->h_free[k], 1);
    for (int w = 0; w < 8; w++) { mbar_init(&ctl->r_full[w][0], 1); mbar_init(&ctl->r_full[w][1], 1); }
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
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapR) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapO) : "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;
  pdl_wait();

  const uint32_t sH = smem_u32(smem), sRing = smem_u32(smem + L::OFF_RING);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (one lane, both CTAs)
    if (lane == 0) {
      const uint32_t full0 = mapa_u32(smem_u32(&ctl->full[0]), 0);     // the leader's barriers collect both CTAs' bytes
      int st = 0;
      uint32_t ph = 1;                                       // producer waits with inverted parity
      // one ring stage = two 16 KB blocks per CTA
      auto stage = [&](const CUtensorMap* m0, int x0, int y0, const CUtensorMap* m1, int x1, int y1) {
        mbar_wait(&ctl->empty[st], ph);
        if (rank == 0) mbar_arrive_expect_tx(&ctl->full[st], 4u * CHUNK);
        const uint32_t dst = sRing + st * (2 * CHUNK);
        tma_load_2d<2>(dst, m0, full0 + 8u * st, x0, y0);
        tma_load_2d<2>(dst + CHUNK, m1, full0 + 8u * st, x1, y1);
        if (++st == NSTG) { st = 0; ph ^= 1u; }
      };
      for (int t = cluster_id; t < q.m_tiles; t += n_clusters) {
        const int m0 = (t * 2 + (int)rank) * TMR;
        // product with a streamed A operand, one accumulator half after the other (A passes twice)
        for (int nh = 0; nh < 2; nh++)
          for (int kc = 0; kc < NKC; kc++)
            stage(&mapA, kc * KC, m0, FFN ? &mapW1 : &mapW2, kc * KC, nh * 256 + (int)rank * TMR);
        if (FFN)
          for (int kc = 0; kc < NKC; kc++)
            stage(&mapW2, kc * KC, (int)rank * TMR, &mapW2, kc * KC, 256 + (int)rank * TMR);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA, one lane)
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc(256, 256);
      int st = 0;
      uint32_t ph = 0, fph[2] = {0u, 0u};
      bool used[2] = {false, false};
      // accumulator half b drained by the epilogue (and, before the second product, the h columns of that half written)
      auto wait_free = [&](int b) {
        if (used[b]) {
          mbar_wait(&ctl->acc_free[b], fph[b]);
          fph[b] ^= 1u;
          tc_fence_after();
        }
        used[b] = true;
      };
      for (int t = cluster_id; t < q.m_tiles; t += n_clusters) {
        for (int nh = 0; nh < 2; nh++) {
          wait_free(nh);
          const uint32_t d_tmem = tmem_base + (uint32_t)(nh * 256);
          for (int kc = 0; kc < NKC; kc++) {
            mbar_wait(&ctl->full[st], ph);
            tc_fence_after();
            const uint32_t sa = sRing + st * (2 * CHUNK);
            const uint64_t da = make_desc(sa), db = make_desc(sa + CHUNK);
            if (!(q.dbg & 8)) {
#pragma unroll
              for (int k = 0; k < KC / 16; k++)
                umma_bf16<2>(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kc | k) ? 1u : 0u);
            }
            umma_commit<2>(&ctl->empty[st]);
            if (++st == NSTG) { st = 0; ph ^= 1u; }
          }
          umma_commit<2>(&ctl->acc_full[nh]);
        }
        if (FFN) {
          wait_free(0);
          wait_free(1);
          for (int kc = 0; kc < NKC; kc++) {
            mbar_wait(&ctl->full[st], ph);
            tc_fence_after();
            const uint32_t sb = sRing + st * (2 * CHUNK);
            const uint64_t da = make_desc(sH + kc * CHUNK), db0 = make_desc(sb), db1 = make_desc(sb + CHUNK);
            if (!(q.dbg & 8)) {
#pragma unroll
              for (int k = 0; k < KC / 16; k++)
                umma_bf16<2>(tmem_base, da + (uint64_t)(k * 2), db0 + (uint64_t)(k * 2), idesc, (kc | k) ? 1u : 0u);
#pragma unroll
              for (int k = 0; k < KC / 16; k++)
                umma_bf16<2>(tmem_base + 256u, da + (uint64_t)(k * 2), db1 + (uint64_t)(k * 2), idesc, (kc | k) ? 1u : 0u);
            }
            umma_commit<2>(&ctl->empty[st]);
            umma_commit<2>(&ctl->h_free[kc]);               // this K chunk of h may be overwritten (residual boxes)
            if (++st == NSTG) { st = 0; ph ^= 1u; }
          }
          umma_commit<2>(&ctl->acc_full[0]);
          umma_commit<2>(&ctl->acc_full[1]);
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    const int quarter = warp & 3;                 // TMEM lanes 32*quarter .. +31 are accessible to this warp
    const int ehalf = (warp - 2) >> 2;            // which of the two warps sharing this lane quarter
    const int widx = warp - 2;
    const int row = quarter * 32 + lane;          // row inside the CTA's 128-row tile
    const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16);
    // shared-memory address of 16-byte piece j of this thread's row in group g
    auto piece = [&](int g, int j) { return sH + (uint32_t)((g & CMASK) * CHUNK) + stg_off(row, j); };
    const uint32_t xch_mine = piece(5 + 2 * ehalf, 0), xch_peer = piece(5 + 2 * (ehalf ^ 1), 0);
    uint32_t full_ph[2] = {0u, 0u};
    uint32_t it = 0;
    auto wait_acc = [&](int b) {
      mbar_wait(&ctl->acc_full[b], full_ph[b]);
      full_ph[b] ^= 1u;
      tc_fence_after();
    };
    // residual boxes of column half hb for the tile whose first row (of this warp) is mw
    auto load_boxes = [&](int hb, int mw) {
      if (lane == 0) {
        mbar_arrive_expect_tx(&ctl->r_full[widx][hb], 2u * BOX);
#pragma unroll
        for (int i = 0; i < 2; i++) {
          const int g = hb * 4 + 2 * ehalf + i;
          tma_load_2d<1>(sH + (uint32_t)((g & CMASK) * CHUNK + quarter * BOX), &mapR, smem_u32(&ctl->r_full[widx][hb]), g * 64, mw);
        }
      }
    };
    if (!FFN && cluster_id < q.m_tiles)            // proj+LN: the residual boxes of half 0 are prefetched a tile ahead
      load_boxes(0, (cluster_id * 2 + (int)rank) * TMR + quarter * 32);
    // The epilogue works in slabs of 32 columns (one tcgen05.ld.x32): slab j = 0..3 of half hb is columns
    // 64 * g + 32 * (j & 1) of group g = 4 hb + 2 ehalf + (j >> 1).  The TMEM load of slab j+1 is issued before slab j is
    // processed, and the per-column parameters (bias / gamma / beta: warp-uniform 16-byte loads) and the residual pieces
    // are requested BEFORE the wait for slab j, so that none of these latencies is exposed per slab.
    uint32_t rb[2][32];
    int n_st = 0;
    auto stamp = [&]() {
      if ((q.dbg & 2048) && blockIdx.x == 0 && warp == 2 && lane == 0 && n_st < 512) g_ffn_stamps[n_st++] = clock64();
    };
    for (int t = cluster_id; t < q.m_tiles; t += n_clusters, it++) {
      const int m_w = (t * 2 + (int)rank) * TMR + quarter * 32;      // first row of this warp
      if (FFN) {
        // ---- first product: h = relu(acc + b1) -> shared memory (A operand of the second product), half by half
#pragma unroll 1
        for (int hb = 0; hb < 2; hb++) {
          const int g0 = hb * 4 + 2 * ehalf;
          stamp();
          wait_acc(hb);
          stamp();
          tmem_ld32_nowait(t_row + (uint32_t)(g0 * 64), rb[0]);
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const int g = g0 + (j >> 1), c0 = g * 64 + (j & 1) * 32;
            float4 bb[8];
#pragma unroll
            for (int k = 0; k < 8; k++) bb[k] = __ldg(reinterpret_cast<const float4*>(q.b1 + c0) + k);
            tmem_ld_wait();
            if (j < 3) tmem_ld32_nowait(t_row + (uint32_t)((g0 + ((j + 1) >> 1)) * 64 + ((j + 1) & 1) * 32), rb[(j + 1) & 1]);
            const uint32_t(&r)[32] = rb[j & 1];
            const act2_t zero2 = f2act2(0.0f, 0.0f);
#pragma unroll
            for (int p = 0; p < 4; p++) {
              const float4 ba = bb[2 * p], bc = bb[2 * p + 1];
              act2_t h0 = __hmax2(f2act2(__uint_as_float(r[p * 8 + 0]) + ba.x, __uint_as_float(r[p * 8 + 1]) + ba.y), zero2);
              act2_t h1 = __hmax2(f2act2(__uint_as_float(r[p * 8 + 2]) + ba.z, __uint_as_float(r[p * 8 + 3]) + ba.w), zero2);
              act2_t h2 = __hmax2(f2act2(__uint_as_float(r[p * 8 + 4]) + bc.x, __uint_as_float(r[p * 8 + 5]) + bc.y), zero2);
              act2_t h3 = __hmax2(f2act2(__uint_as_float(r[p * 8 + 6]) + bc.z, __uint_as_float(r[p * 8 + 7]) + bc.w), zero2);
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(piece(g, (j & 1) * 4 + p)), "r"(*reinterpret_cast<uint32_t*>(&h0)),
                           "r"(*reinterpret_cast<uint32_t*>(&h1)), "r"(*reinterpret_cast<uint32_t*>(&h2)), "r"(*reinterpret_cast<uint32_t*>(&h3)) : "memory");
            }
          }
          // generic-proxy writes of h -> visible to the tensor core's async-proxy reads; accumulator half drained.  The
          // hand-over to the MMA issuer (possibly in the other CTA of the pair) is: these st.shared -> fence.proxy.async
          // (completes only when they are performed) -> remote mbarrier arrive -> the issuer's wait -> tcgen05.mma, whose
          // operand reads happen in the SM that owns the shared memory.  A cluster-scope release / acquire pair on the
          // barrier is not needed for that chain and was measured to be expensive: the acquire compiles to CCTL.IVALL,
          // which wipes this SM's L1 twice per tile (every per-column parameter load of the epilogue then misses), the
          // release to MEMBAR.ALL.GPU on the critical path between the two products.
          fence_async_smem();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(&ctl->acc_free[hb], 0);
        }
        // ---- while the second product runs: as soon as it has consumed K chunk g of h, this warp's box of that chunk
        //      is refilled with the residual rows (TMA, warp-private)
        for (int hb = 0; hb < 2; hb++) {
          mbar_wait(&ctl->h_free[hb * 4 + 2 * ehalf], it & 1u);
          mbar_wait(&ctl->h_free[hb * 4 + 2 * ehalf + 1], it & 1u);
          load_boxes(hb, m_w);
        }
      }
      // ---- last product: v = acc + b2 + residual; LayerNorm statistics per row; v parked in TMEM
      float shift = 0.0f, s1 = 0.0f, s2 = 0.0f;
#pragma unroll 1
      for (int hb = 0; hb < 2; hb++) {
        const int g0 = hb * 4 + 2 * ehalf;
        stamp();
        wait_acc(hb);
        stamp();
        mbar_wait(&ctl->r_full[widx][hb], it & 1u);
        stamp();
        tmem_ld32_nowait(t_row + (uint32_t)(g0 * 64), rb[0]);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int g = g0 + (j >> 1), c0 = g * 64 + (j & 1) * 32;
          float4 bb[8];
          uint4 rv[4];
#pragma unroll
          for (int k = 0; k < 8; k++) bb[k] = __ldg(reinterpret_cast<const float4*>(q.b2 + c0) + k);
#pragma unroll
          for (int p = 0; p < 4; p++)           // residual: lane = row reads of the swizzled box, conflict-free
            asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(rv[p].x), "=r"(rv[p].y), "=r"(rv[p].z), "=r"(rv[p].w) : "r"(piece(g, (j & 1) * 4 + p)) : "memory");
          tmem_ld_wait();
          if (j < 3) tmem_ld32_nowait(t_row + (uint32_t)((g0 + ((j + 1) >> 1)) * 64 + ((j + 1) & 1) * 32), rb[(j + 1) & 1]);
          const uint32_t(&r)[32] = rb[j & 1];
          float v[32];
#pragma unroll
          for (int p = 0; p < 4; p++) {
            const float4 ba = bb[2 * p], bc = bb[2 * p + 1];
            const act2_t* h = reinterpret_cast<const act2_t*>(&rv[p]);
            const float2 f0 = act22f2(h[0]), f1 = act22f2(h[1]), f2 = act22f2(h[2]), f3 = act22f2(h[3]);
            v[p * 8 + 0] = (__uint_as_float(r[p * 8 + 0]) + ba.x) + f0.x; v[p * 8 + 1] = (__uint_as_float(r[p * 8 + 1]) + ba.y) + f0.y;
            v[p * 8 + 2] = (__uint_as_float(r[p * 8 + 2]) + ba.z) + f1.x; v[p * 8 + 3] = (__uint_as_float(r[p * 8 + 3]) + ba.w) + f1.y;
            v[p * 8 + 4] = (__uint_as_float(r[p * 8 + 4]) + bc.x) + f2.x; v[p * 8 + 5] = (__uint_as_float(r[p * 8 + 5]) + bc.y) + f2.y;
            v[p * 8 + 6] = (__uint_as_float(r[p * 8 + 6]) + bc.z) + f3.x; v[p * 8 + 7] = (__uint_as_float(r[p * 8 + 7]) + bc.w) + f3.y;
          }
          if (hb == 0 && j == 0) shift = v[0];    // shifted sums: no cancellation for rows with a large mean
          float t1 = 0.0f, t2 = 0.0f;            // two accumulator pairs shorten the dependent chains
#pragma unroll
          for (int k = 0; k < 32; k += 2) {
            const float d0 = v[k] - shift, d1 = v[k + 1] - shift;
            s1 += d0; s2 = fmaf(d0, d0, s2);
            t1 += d1; t2 = fmaf(d1, d1, t2);
          }
          s1 += t1; s2 += t2;
          tmem_st32(t_row + (uint32_t)c0, v);
        }
        if (!FFN && hb == 0) {                   // proj+LN: the boxes of half 0 are consumed, those of half 1 take their place
          __syncwarp();                          // (a whole accumulator half of MMAs before they are needed)
          load_boxes(1, m_w);
        }
      }
      tmem_st_wait();
      stamp();
      // combine the two warps' 256 columns each: mean_w = shift + s1/256, M2_w = s2 - s1^2/256.  The exchange slot is the
      // first piece of this thread's row in its last box (the residual there has been consumed).
      const float mean_w = shift + s1 * (1.0f / 256.0f), m2_w = s2 - s1 * s1 * (1.0f / 256.0f);
      asm volatile("st.shared.v2.f32 [%0], {%1,%2};" ::"r"(xch_mine), "f"(mean_w), "f"(m2_w) : "memory");
      asm volatile("bar.sync %0, 64;" ::"r"(2 + quarter) : "memory");
      float2 o;
      asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(o.x), "=f"(o.y) : "r"(xch_peer) : "memory");
      asm volatile("bar.sync %0, 64;" ::"r"(2 + quarter) : "memory");      // both have read before either overwrites its slot
      const float mean = 0.5f * (mean_w + o.x);
      const float dm = mean_w - o.x;
      const float var = (m2_w + o.y + dm * dm * 128.0f) * (1.0f / 512.0f);
      const float rstd = rsqrtf(var + q.eps);
      const float nmr = -mean * rstd;
      stamp();
      // ---- second pass over TMEM, half by half (the MMA issuer refills a half as soon as all warps have read it)
#pragma unroll 1
      for (int hb = 0; hb < 2; hb++) {
        const int g0 = hb * 4 + 2 * ehalf;
        if (hb) stamp();
        tmem_ld32_nowait(t_row + (uint32_t)(g0 * 64), rb[0]);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int g = g0 + (j >> 1), c0 = g * 64 + (j & 1) * 32;
          if (!FFN && hb == 1 && !(j & 1)) {
            // proj+LN: this box still holds an output box of half 0 until its TMA store has read it (stores complete in
            // order: at most the most recent one may still be pending)
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncwarp();
          }
          float4 gm[8], bt[8];
          if (!FFN) {
#pragma unroll
            for (int k = 0; k < 8; k++) { gm[k] = __ldg(reinterpret_cast<const float4*>(q.gamma + c0) + k); bt[k] = __ldg(reinterpret_cast<const float4*>(q.beta + c0) + k); }
          }
          tmem_ld_wait();
          if (j < 3) {
            tmem_ld32_nowait(t_row + (uint32_t)((g0 + ((j + 1) >> 1)) * 64 + ((j + 1) & 1) * 32), rb[(j + 1) & 1]);
          } else {                                 // last TMEM read of this half: hand it back
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&ctl->acc_free[hb], 0);
          }
          const uint32_t(&r)[32] = rb[j & 1];
          float v[32];
#pragma unroll
          for (int k = 0; k < 32; k++) v[k] = fmaf(__uint_as_float(r[k]), rstd, nmr);
          if (FFN) {
            // column sums over the 16 rows of each half warp: halving exchange, lane l (mod 16) ends with columns 2l, 2l+1
            const unsigned FULL = 0xffffffffu;
            float a[16], b[8], c[4], d[2];
            const bool b3 = lane & 8, b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
#pragma unroll
            for (int k = 0; k < 16; k++) { const float keep = b3 ? v[16 + k] : v[k], send = b3 ? v[k] : v[16 + k]; a[k] = keep + __shfl_xor_sync(FULL, send, 8); }
#pragma unroll
            for (int k = 0; k < 8; k++) { const float keep = b2 ? a[8 + k] : a[k], send = b2 ? a[k] : a[8 + k]; b[k] = keep + __shfl_xor_sync(FULL, send, 4); }
#pragma unroll
            for (int k = 0; k < 4; k++) { const float keep = b1 ? b[4 + k] : b[k], send = b1 ? b[k] : b[4 + k]; c[k] = keep + __shfl_xor_sync(FULL, send, 2); }
#pragma unroll
            for (int k = 0; k < 2; k++) { const float keep = b0 ? c[2 + k] : c[k], send = b0 ? c[k] : c[2 + k]; d[k] = keep + __shfl_xor_sync(FULL, send, 1); }
            const int mg = m_w + (lane >> 4) * 16;   // first row of this half warp's group
            if (mg < q.M && !(q.dbg & 1))
              *reinterpret_cast<float2*>(q.gs + (long long)(mg >> 4) * FD + c0 + (lane & 15) * 2) = make_float2(d[0], d[1]);
          } else {
            // the output rows replace the residual rows in this warp's box; one TMA store per box
#pragma unroll
            for (int p = 0; p < 4; p++) {
              const float4 ga = gm[2 * p], gc = gm[2 * p + 1], ba = bt[2 * p], bc = bt[2 * p + 1];
              act2_t h0 = f2act2(fmaf(v[p * 8 + 0], ga.x, ba.x), fmaf(v[p * 8 + 1], ga.y, ba.y));
              act2_t h1 = f2act2(fmaf(v[p * 8 + 2], ga.z, ba.z), fmaf(v[p * 8 + 3], ga.w, ba.w));
              act2_t h2 = f2act2(fmaf(v[p * 8 + 4], gc.x, bc.x), fmaf(v[p * 8 + 5], gc.y, bc.y));
              act2_t h3 = f2act2(fmaf(v[p * 8 + 6], gc.z, bc.z), fmaf(v[p * 8 + 7], gc.w, bc.w));
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(piece(g, (j & 1) * 4 + p)), "r"(*reinterpret_cast<uint32_t*>(&h0)),
                           "r"(*reinterpret_cast<uint32_t*>(&h1)), "r"(*reinterpret_cast<uint32_t*>(&h2)), "r"(*reinterpret_cast<uint32_t*>(&h3)) : "memory");
            }
            if (j & 1) {                           // box of group g complete
              fence_async_smem();
              __syncwarp();
              if (lane == 0 && !(q.dbg & 1)) {
                tma_store_2d(&mapO, sH + (uint32_t)((g & CMASK) * CHUNK + quarter * BOX), g * 64, m_w);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
              }
            }
          }
        }
      }
      stamp();
      if (!FFN) {
        // boxes free again once the stores have read them: prefetch the next tile's residual rows
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
        if (t + n_clusters < q.m_tiles) load_boxes(0, ((t + n_clusters) * 2 + (int)rank) * TMR + quarter * 32);
      }
    }
    if (!FFN && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// pooled[h][c] = gamma[c] * (sum of the `per` group rows of hypothesis h, in order) / tokens + beta[c], then the output
// Linear(512, n_out) of the head on the pooled row (mean over tokens commutes with it): out[h][n] = pooled[h] . Wf[n] + bf[n].
// One block per hypothesis: 128 threads x 4 columns for the pooling, then one warp per output (same lane-strided fmaf order
// + butterfly sum as small_linear_kernel, which this replaces together with the separate pooling launch).
__global__ void __launch_bounds__(128) pool_groups_kernel(const float* __restrict__ gs, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float* __restrict__ pooled, int per,
                                                          float inv_tokens, const float* __restrict__ Wf,
                                                          const float* __restrict__ bf, float* __restrict__ out, int n_out)
{
  __shared__ float row[FD];
  const int c = threadIdx.x * 4;
  const float4* src = reinterpret_cast<const float4*>(gs + (long long)blockIdx.x * per * FD + c);
  float4 s = src[0];
  for (int j = 1; j < per; j++) {
    const float4 x = src[(long long)j * (FD / 4)];
    s.x += x.x; s.y += x.y; s.z += x.z; s.w += x.w;
  }
  const float4 gm = *reinterpret_cast<const float4*>(gamma + c), bt = *reinterpret_cast<const float4*>(beta + c);
  const float4 p = make_float4(fmaf(s.x * inv_tokens, gm.x, bt.x), fmaf(s.y * inv_tokens, gm.y, bt.y), fmaf(s.z * inv_tokens, gm.z, bt.z),
                               fmaf(s.w * inv_tokens, gm.w, bt.w));
  *reinterpret_cast<float4*>(pooled + (long long)blockIdx.x * FD + c) = p;
  if (!Wf) return;
  *reinterpret_cast<float4*>(row + c) = p;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int n = warp; n < n_out; n += 4) {
    float a = 0.0f;
    for (int k = lane; k < FD; k += 32) a = fmaf(row[k], Wf[(long long)n * FD + k], a);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if (lane == 0) out[(long long)blockIdx.x * n_out + n] = a + (bf ? bf[n] : 0.0f);
  }
}

int ffn_dbg()
{
  static int v = -1;
  if (v < 0) { const char* e = getenv("FPB_FFN_DBG"); v = e ? atoi(e) : 0; }
  return v;
}

int ffn_enabled()
{
  static int v = -1;
  if (v < 0) { const char* e = getenv("FPB_FFN"); v = e ? atoi(e) : 1; }
  return v;
}

template <bool FFN>
int launch_ffn(const CUtensorMap& mA, const CUtensorMap& mW1, const CUtensorMap& mW2, const CUtensorMap& mR, const CUtensorMap& mO,
               const FfnParams& q, int num_sms, cudaStream_t st)
{
  static bool configured = false;
  if (!configured) {
    FPB_CUDA(cudaFuncSetAttribute(ffn_kernel<FFN>, cudaFuncAttributeMaxDynamicSharedMemorySize, FLayout<FFN>::SMEM));
    configured = true;
  }
  int clusters = num_sms / 2;
  if (clusters > q.m_tiles) clusters = q.m_tiles;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(clusters * 2, 1, 1);
  cfg.blockDim = dim3(NTHREADS, 1, 1);
  cfg.dynamicSmemBytes = FLayout<FFN>::SMEM;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = fpb_pdl_enabled() ? 2 : 1;
  FPB_CUDA(cudaLaunchKernelEx(&cfg, ffn_kernel<FFN>, mA, mW1, mW2, mR, mO, q));
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
  CUtensorMap mA, mW, mR, mO;
  int num_sms = 0;
  int rc = fpb_tiled_map_act(A, M, FD, FD, TMR, &mA, &num_sms);
  if (rc) return rc;
  rc = fpb_tiled_map_act(W, FD, FD, FD, TMR, &mW, nullptr);
  if (rc) return rc;
  rc = fpb_tiled_map_act(residual, M, FD, FD, 32, &mR, nullptr);
  if (rc) return rc;
  rc = fpb_tiled_map_act(out, M, FD, FD, 32, &mO, nullptr);
  if (rc) return rc;
  FfnParams q;
  memset(&q, 0, sizeof(q));
  q.M = (int)M; q.m_tiles = fpb_cdiv(M, 2 * TMR);
  q.b2 = b; q.gamma = gamma; q.beta = beta; q.eps = eps;
  q.dbg = ffn_dbg();
  return launch_ffn<false>(mA, mW, mW, mR, mO, q, num_sms, st);
}

// pooled[g] = mean over the `tokens` rows of group g of LayerNorm(x + relu(x . W1^T + b1) . W2^T + b2) * gamma + beta
// x: [groups * tokens][512] 16-bit rows; gs: scratch of (groups * tokens / 16) x 512 floats; tokens % 16 == 0.
// Wf != null: also out[g][n] = pooled[g] . Wf[n] + bf[n] for n < n_out (the head's output Linear) in the pooling launch.
int ffn_ln_token_mean_512(const void* x, const void* W1, const float* b1, const void* W2, const float* b2, const float* gamma,
                          const float* beta, float eps, float* gs, float* pooled, int groups, int tokens, const float* Wf,
                          const float* bf, float* out, int n_out, cudaStream_t st)
{
  const long long M = (long long)groups * tokens;
  if (!ffn_enabled() || groups <= 0 || tokens <= 0 || tokens % 16 != 0 || M > 0x7fffff00LL || !aligned16(x) || !aligned16(W1) ||
      !aligned16(W2) || !aligned16(b1) || !aligned16(b2) || !aligned16(gamma) || !aligned16(beta) || !aligned16(gs) || !aligned16(pooled))
    return FPB_ERR_UNSUPPORTED;
  CUtensorMap mA, mW1, mW2, mR;
  int num_sms = 0;
  int rc = fpb_tiled_map_act(x, M, FD, FD, TMR, &mA, &num_sms);
  if (rc) return rc;
  rc = fpb_tiled_map_act(W1, FD, FD, FD, TMR, &mW1, nullptr);
  if (rc) return rc;
  rc = fpb_tiled_map_act(W2, FD, FD, FD, TMR, &mW2, nullptr);
  if (rc) return rc;
  rc = fpb_tiled_map_act(x, M, FD, FD, 32, &mR, nullptr);
  if (rc) return rc;
  FfnParams q;
  memset(&q, 0, sizeof(q));
  q.M = (int)M; q.m_tiles = fpb_cdiv(M, 2 * TMR);
  q.b1 = b1; q.b2 = b2; q.eps = eps; q.gs = gs;
  q.dbg = ffn_dbg();
  rc = launch_ffn<true>(mA, mW1, mW2, mR, mR, q, num_sms, st);
  if (rc) return rc;
  pool_groups_kernel<<<groups, 128, 0, st>>>(gs, gamma, beta, pooled, tokens / 16, 1.0f / (float)tokens, Wf, bf, out, n_out);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

int ffn_debug_stamps(long long* out_host, int n)
{
  if (n > 512) n = 512;
  FPB_CUDA(cudaMemcpyFromSymbol(out_host, g_ffn_stamps, (size_t)n * sizeof(long long)));
  return FPB_OK;
}

}  // namespace FPB_ACT_NS
