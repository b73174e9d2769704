// Persistent, warp-specialised tcgen05 / TMEM / TMA bf16 implicit GEMM for sm_100a (second-generation kernel).
//   ConvBNReLU / ResnetBasicBlock   learning/models/network_modules.py:37-50, 73-111
//   encoders + heads                learning/models/refine_network.py:38-70, score_network.py:36-57
//
// One CTA (CG=1) or one CTA pair (CG=2, tcgen05 cta_group::2) per SM / TPC, looping over output tiles of
// (128*CG) x BN.  Roles per CTA (6 warps):
//   warp 0      TMA producer: A through an im2col tensor map (cp.async.bulk.tensor.4d...im2col: 128 output pixels x
//               64 channels of one filter tap per request, zero-filled padding) or a tiled 2-D map (plain GEMM);
//               W through a tiled 2-D map; both land in the 128B-swizzled K-major layout UMMA reads.
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer; accumulators are double-buffered in TMEM
//               (2 x BN columns) so the epilogue of tile i overlaps the main loop of tile i+1.
//   warps 2-5   epilogue: tcgen05.ld -> folded BN / bias, residual, ReLU, positional embedding -> bf16/fp32 stores.
// Pipelines: smem full/empty mbarriers (TMA <-> MMA), TMEM full/empty mbarriers (MMA <-> epilogue).
// In CG=2 mode each CTA loads its own 128 rows of A and half of the W rows; the leader CTA issues the MMAs for
// both SMs and tcgen05.commit multicasts the barrier arrivals to the pair.
#include "net_kernels.cuh"
#include "tc_ptx.cuh"
#include <cuda.h>
#include <stdlib.h>
#include <map>
#include <mutex>
#include <vector>

namespace FPB_ACT_NS {

namespace {

constexpr int TM = 128;
constexpr int KC = 64;
constexpr int NTHREADS = 320;                 // 1 TMA warp + 1 MMA warp + 8 epilogue warps
constexpr int A_STAGE_BYTES = TM * 128;


struct Tc2Params {
  GemmProblem p;
  int a_im2col;          // 1: A through the im2col map, 0: tiled 2-D map over [M][K]
  int lower_w, lower_h;  // im2col bounding-box lower corner (= -pad)
  int m_tiles, n_tiles;  // tiles of (TM*CG) x BN
  int halo, Wp, IS, hrows, na, nb, tps;   // halo mode: position pitch, image stride, A box rows, A / W ring depths, taps per W stage
  int fast_epi;          // 1: smem-transposed, fully coalesced bf16 epilogue (ncols % 64 == 0)
  int ksplit;            // split-K factor (1 = off): work item t -> output tile t / ksplit, K slice t % ksplit (nk % ksplit == 0)
  long long c_split_stride;   // elements between the fp32 partial outputs of consecutive K slices
  int dbg;               // FPB_TC2_DBG experiment bits: 1 no stores, 2 no scale/shift, 4 no residual, 8 no MMA
};

constexpr int A_HALO_BYTES = 27 * 1024;          // >= (128 + 2*Wp + 2) rows x 128 B for Wp <= 42
template <int STAGES>
struct SmemCtl {
  uint64_t full[8];
  uint64_t empty[8];
  uint64_t afull[4];
  uint64_t aempty[4];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

template <int BN, int CG>
struct Cfg {
  static constexpr int B_ROWS = BN / CG;                      // W rows loaded by each CTA
  static constexpr int B_STAGE_BYTES = B_ROWS * 128;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  // per epilogue warp: 4 KB output staging + 4 KB residual staging + 4 KB positional-embedding staging + scale/shift
  // 2 warps per TMEM lane quarter split the 64-column groups -- except for the CTA-pair kernel of the long-K layers
  // (BN = 256, CG = 2, K >= 2304): its main loop (>= 18 k cycles per tile) hides a 4-warp epilogue, and the 16 KB of
  // staging this frees buys a 6th pipeline stage.  The operand stream of these layers is latency x bandwidth bound
  // (bytes in flight / ~2.5 k cycles of loaded TMA latency), so ring depth is throughput (FPB_TC2_EPI8 = old layout).
#ifdef FPB_TC2_EPI8
  static constexpr int EPI_WARPS = BN >= 128 ? 8 : 4;
#else
  static constexpr int EPI_WARPS = (BN >= 128 && !(BN == 256 && CG == 2)) ? 8 : 4;
#endif
  // one 4 KB staging tile per epilogue warp (residual in, positional embedding in, result out share it) plus a
  // double-buffered scale/shift table shared by the CTA
  static constexpr int EPI_WARP_BYTES = 4096;
  static constexpr int EPI_BYTES = EPI_WARPS * EPI_WARP_BYTES + 2 * (2 * BN * 4);
  static constexpr int STAGES_RAW = (224 * 1024 - EPI_BYTES) / STAGE_BYTES;
  static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
  static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
  static constexpr int SMEM = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + (int)sizeof(SmemCtl<STAGES>);
};

template <int BN, int CG>
__global__ void __launch_bounds__(NTHREADS, 1)
tc2_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const Tc2Params q)
{
  using C = Cfg<BN, CG>;
  constexpr int STAGES = C::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* epi_smem = smem + STAGES * C::STAGE_BYTES;
  SmemCtl<STAGES>* ctl = (SmemCtl<STAGES>*)(epi_smem + C::EPI_BYTES);
  const GemmProblem& p = q.p;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
  const int cluster_id = blockIdx.x / CG, n_clusters = gridDim.x / CG;
  const int ksp = q.ksplit;
  const int total_tiles = q.m_tiles * q.n_tiles * ksp;       // work items: (output tile, K slice)
  const int nk = p.K / KC / ksp;                              // K stages per work item
  pdl_launch_dependents();

  if (threadIdx.x == 0) {
    for (int s = 0; s < 8; s++) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
    for (int s = 0; s < 4; s++) { mbar_init(&ctl->afull[s], 1); mbar_init(&ctl->aempty[s], 1); }
    for (int a = 0; a < 2; a++) { mbar_init(&ctl->tmem_full[a], 1); mbar_init(&ctl->tmem_empty[a], C::EPI_WARPS * CG); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if (CG == 1)
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ctl->tmem_base)),
                   "r"((uint32_t)C::TMEM_COLS)
                   : "memory");
    else
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ctl->tmem_base)),
                   "r"((uint32_t)C::TMEM_COLS)
                   : "memory");
    if (CG == 1) asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    else asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapB) : "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;
  pdl_wait();                                   // the predecessor's outputs (our inputs) are complete from here on
  // FPB_TC2_DBG & 2048: CTA 0 records clock64 stamps into the output buffer (slots: warp 0 -> 0.., warp 1 -> 64.., warp 2 -> 128..)
  int n_st = 0;
  auto stamp = [&](int base) {
    if ((q.dbg & 2048) && blockIdx.x == 0 && lane == 0 && n_st < 64) reinterpret_cast<long long*>(p.C)[base + n_st++] = clock64();
  };

  // Both single-thread roles keep ring positions as (stage, phase) pairs advanced incrementally: every integer division
  // or modulo by a run-time value in these loops showed up directly as pipeline time (measured: the halo loop went
  // from 890 to the im2col loop's 410 cycles per stage with no loads and no MMAs issued).
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (one lane, every CTA)
    if (lane == 0) {
      const uint32_t full0 = CG == 2 ? mapa_u32(smem_u32(&ctl->full[0]), 0) : smem_u32(&ctl->full[0]);     // leader's barriers
      const uint32_t afull0 = CG == 2 ? mapa_u32(smem_u32(&ctl->afull[0]), 0) : smem_u32(&ctl->afull[0]);
      const bool no_a = q.dbg & 128, no_b = q.dbg & 256;
      if (q.halo) {
        const int nchunk = p.Cin / KC;
        const uint32_t sA0 = smem_u32(smem), sB0 = sA0 + q.na * A_HALO_BYTES;
        const uint32_t a_tx = no_a ? 0u : (uint32_t)(CG * q.hrows * 128), b_tx = no_b ? 0u : (uint32_t)(CG * C::B_STAGE_BYTES);
        int sa = 0, sb = 0;
        uint32_t pa = 1, pb = 1;                               // producer waits with inverted parity
        for (int t = cluster_id; t < total_tiles; t += n_clusters) {
          const int mt = t / q.n_tiles, nt = t - mt * q.n_tiles;
          const int m0 = (mt * CG + (int)rank) * TM;
          const int nrow0 = nt * BN + (int)rank * C::B_ROWS;
          for (int c = 0; c < nchunk; c++) {
            mbar_wait(&ctl->aempty[sa], pa);
            if (rank == 0) mbar_arrive_expect_tx(&ctl->afull[sa], a_tx);
            if (!no_a) tma_load_2d<CG>(sA0 + sa * A_HALO_BYTES, &mapA, afull0 + 8u * sa, c * KC, m0 - q.Wp - 1);
            if (++sa == q.na) { sa = 0; pa ^= 1u; }
            int kcol = c * KC;
            for (int tap = 0; tap < 9; tap += q.tps) {           // q.tps filter taps (one kernel row when 3) share a stage
              mbar_wait(&ctl->empty[sb], pb);
              if (rank == 0) mbar_arrive_expect_tx(&ctl->full[sb], b_tx * q.tps);
              for (int u = 0; u < q.tps; u++, kcol += p.Cin)
                if (!no_b) tma_load_2d<CG>(sB0 + (sb * q.tps + u) * C::B_STAGE_BYTES, &mapB, full0 + 8u * sb, kcol, nrow0);
              if (++sb == q.nb) { sb = 0; pb ^= 1u; }
            }
          }
        }
      } else {
        const int hw = p.Ho * p.Wo;
        const uint32_t tx = (uint32_t)(CG * ((no_a ? 0 : A_STAGE_BYTES) + (no_b ? 0 : C::B_STAGE_BYTES)));
        const uint32_t s0 = smem_u32(smem);
        int s = 0;
        uint32_t ph = 1;
        for (int t = cluster_id; t < total_tiles; t += n_clusters) {
          const int tile = ksp == 1 ? t : t / ksp, ks = t - tile * ksp;
          const int mt = tile / q.n_tiles, nt = tile - mt * q.n_tiles;
          const int m0 = (mt * CG + (int)rank) * TM;
          const int nrow0 = nt * BN + (int)rank * C::B_ROWS;
          int cw = 0, ch = 0, img = 0;
          if (q.a_im2col) {
            img = m0 / hw;
            const int r = m0 - img * hw;
            const int oy = r / p.Wo, ox = r - oy * p.Wo;
            cw = ox * p.stride + q.lower_w;
            ch = oy * p.stride + q.lower_h;
          }
          int k0 = 0, c0 = 0, kx = 0, ky = 0;                  // K walks (ky, kx, c) with c fastest
          if (ks) {                                            // split-K: this item starts at K stage ks * nk
            k0 = ks * nk * KC;
            const int tap = k0 / p.Cin;
            c0 = k0 - tap * p.Cin;
            ky = tap / p.KW; kx = tap - ky * p.KW;
          }
          for (int kc = 0; kc < nk; kc++, k0 += KC) {
            mbar_wait(&ctl->empty[s], ph);
            const uint32_t sa = s0 + s * C::STAGE_BYTES;
            if (rank == 0) mbar_arrive_expect_tx(&ctl->full[s], tx);
            if (!no_a) {
              if (q.a_im2col) tma_load_im2col<CG>(sa, &mapA, full0 + 8u * s, c0, cw, ch, img, (uint16_t)kx, (uint16_t)ky);
              else tma_load_2d<CG>(sa, &mapA, full0 + 8u * s, k0, m0);
            }
            if (!no_b) tma_load_2d<CG>(sa + A_STAGE_BYTES, &mapB, full0 + 8u * s, k0, nrow0);
            c0 += KC;
            if (c0 == p.Cin) { c0 = 0; if (++kx == p.KW) { kx = 0; ky++; } }
            if (++s == STAGES) { s = 0; ph ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA, one lane)
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc(TM * CG, BN);
      const bool no_mma = q.dbg & 8;
      uint32_t acc = 0, acc_ph = 1;                            // waits on tmem_empty with inverted parity
      if (q.halo) {
        const int nchunk = p.Cin / KC;
        const uint32_t sA0 = smem_u32(smem), sB0 = sA0 + q.na * A_HALO_BYTES;
        const uint32_t row_step = (uint32_t)(q.Wp * 128);          // one image row of positions
        int sa = 0, sb = 0;
        uint32_t pa = 0, pb = 0;
        for (int t = cluster_id; t < total_tiles; t += n_clusters) {
          mbar_wait(&ctl->tmem_empty[acc], acc_ph);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * BN;
          for (int c = 0; c < nchunk; c++) {
            mbar_wait(&ctl->afull[sa], pa);
            // the 128B swizzle is a function of the absolute smem address (measured: the descriptor base_offset must
            // stay 0 for start rows that are not multiples of 8), so a filter tap is just a start-address shift
            uint32_t a_addr = sA0 + sa * A_HALO_BYTES;
            for (int ky = 0; ky < 3; ky++, a_addr += row_step) {
              for (int kx = 0; kx < 3; kx++) {
                const int u = q.tps == 3 ? kx : 0;
                if (u == 0) {
                  mbar_wait(&ctl->full[sb], pb);
                  tc_fence_after();
                }
                const uint64_t da = make_desc(a_addr + 128u * kx), db = make_desc(sB0 + (sb * q.tps + u) * C::B_STAGE_BYTES);
                if (!no_mma) {
#pragma unroll
                  for (int k = 0; k < KC / 16; k++)
                    umma_bf16<CG>(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (c | ky | kx | k) ? 1u : 0u);
                }
                if (u == q.tps - 1) {
                  umma_commit<CG>(&ctl->empty[sb]);
                  if (++sb == q.nb) { sb = 0; pb ^= 1u; }
                }
              }
            }
            umma_commit<CG>(&ctl->aempty[sa]);
            if (++sa == q.na) { sa = 0; pa ^= 1u; }
          }
          umma_commit<CG>(&ctl->tmem_full[acc]);
          if (++acc == 2) { acc = 0; acc_ph ^= 1u; }
        }
      } else {
        const uint32_t s0 = smem_u32(smem);
        int s = 0;
        uint32_t ph = 0;
        for (int t = cluster_id; t < total_tiles; t += n_clusters) {
          mbar_wait(&ctl->tmem_empty[acc], acc_ph);
          tc_fence_after();
          stamp(64);                                             // accumulator buffer free
          const uint32_t d_tmem = tmem_base + acc * BN;
          for (int kc = 0; kc < nk; kc++) {
            mbar_wait(&ctl->full[s], ph);
            tc_fence_after();
            const uint32_t sa = s0 + s * C::STAGE_BYTES;
            const uint64_t da = make_desc(sa), db = make_desc(sa + A_STAGE_BYTES);
            if (!no_mma) {
#pragma unroll
              for (int k = 0; k < KC / 16; k++)
                umma_bf16<CG>(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kc | k) ? 1u : 0u);
            }
            umma_commit<CG>(&ctl->empty[s]);
            if (++s == STAGES) { s = 0; ph ^= 1u; }
          }
          stamp(64);                                             // all MMAs of the tile issued
          umma_commit<CG>(&ctl->tmem_full[acc]);
          if (++acc == 2) { acc = 0; acc_ph ^= 1u; }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    const int quarter = warp & 3;                 // TMEM lanes 32*quarter .. +31 are accessible to this warp
    const int ehalf = (warp - 2) >> 2;            // which of the two warps sharing this lane quarter
    constexpr int GSTEP = C::EPI_WARPS / 4;       // 64-column groups are interleaved over the warps of a quarter
    if (warp - 2 < C::EPI_WARPS) {
    const int row = quarter * 32 + lane;
    const int nst = p.n_valid > 0 ? p.n_valid : p.N;
    uint32_t tc = 0;
    // residual tile of the NEXT 64-column group is fetched one group ahead (8 x 16 B per lane, coalesced)
    uint4 rr[8];
    const bool use_res = q.fast_epi && p.residual && !(q.dbg & 4);
    auto load_res = [&](int mw_, int nb_) {                  // mw_: first row of this warp, nb_: first column of the group
      if (nb_ >= nst) return;
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const int mr = mw_ + 4 * i + (lane >> 3);
        rr[i] = mr < p.M ? __ldg(reinterpret_cast<const uint4*>((const act_t*)p.residual + (long long)mr * p.ldr + nb_) + (lane & 7))
                         : make_uint4(0u, 0u, 0u, 0u);
      }
    };
    // first row of this warp / first column of tile `tile`
    auto tile_mw = [&](int tile, int& mw_, int& n0_) {
      const int mt_ = tile / q.n_tiles;
      mw_ = (mt_ * CG + (int)rank) * TM + (warp & 3) * 32;
      n0_ = (tile - mt_ * q.n_tiles) * BN;
    };
    if (use_res && cluster_id < total_tiles) {
      int mw0, n00;
      tile_mw(cluster_id, mw0, n00);
      load_res(mw0, n00 + ehalf * 64);
    }
    for (int t = cluster_id; t < total_tiles; t += n_clusters, tc++) {
      const int tile = ksp == 1 ? t : t / ksp, ks = t - tile * ksp;
      const int mt = tile / q.n_tiles, nt = tile - mt * q.n_tiles;
      const int m = (mt * CG + (int)rank) * TM + row;
      const int n0 = nt * BN;
      const uint32_t acc = tc & 1u;
      if (warp == 2) stamp(128);
      mbar_wait(&ctl->tmem_full[acc], (tc >> 1) & 1u);
      tc_fence_after();
      if (warp == 2) stamp(128);                               // accumulator ready
      if (q.fast_epi) {
        // ---- coalesced path: TMEM -> registers (lane = row) -> math -> bf16 -> warp-private swizzled smem tile ->
        //      8 lanes per row read it back and store 128 contiguous bytes per row; residual / positional embedding
        //      travel the same way in the opposite direction.
        uint8_t* ew = epi_smem + (warp - 2) * C::EPI_WARP_BYTES;
        const uint32_t s_out = smem_u32(ew), s_res = s_out, s_pe = s_out;   // one tile, used in turn (guarded by __syncwarp)
        float* ss = (float*)(epi_smem + C::EPI_WARPS * C::EPI_WARP_BYTES) + (tc & 1u) * (2 * BN);   // scale[BN], shift[BN] of this n-tile
        const int m_w = (mt * CG + (int)rank) * TM + quarter * 32;   // first row of this warp
        const int sub = lane >> 3, cj = lane & 7;               // coalesced role: row 4i+sub, 16-byte chunk cj
        int mw_next = 0, n0_next = 0;
        if (use_res && t + n_clusters < total_tiles) tile_mw(t + n_clusters, mw_next, n0_next);
        // a warp's 32 rows never straddle m_split (host guarantees m_split % 32 == 0 or m_split >= M)
        const int seg = p.m_split >= p.M ? 0 : m_w / p.m_split;
        act_t* out_w = (act_t*)p.C + (long long)(m_w - seg * p.m_split) * p.ldc + (long long)seg * p.split_col_off;
        const int pe_r0 = p.pe ? m_w % p.pe_rows : 0;
        bool interior = true;                                   // halo mode: border positions produce zeros
        if (q.halo) {
          const int r_ = (m_w + lane) % q.IS, yy = r_ / q.Wp;
          interior = yy >= 1 && r_ - yy * q.Wp < p.img_w;
        }
        // global address of output row mr (nullptr: not stored)
        auto row_ptr = [&](int mr) -> act_t* {
          if (mr >= p.M) return nullptr;
          if (p.row_remap == 0) return out_w + (long long)(mr - m_w) * p.ldc;
          const int hw = p.img_h * p.img_w;
          if (p.row_remap == 1) {                               // dense pixel index -> zero-bordered position
            const int img = mr / hw, r_ = mr - img * hw, y = r_ / p.img_w;
            return (act_t*)p.C + ((long long)img * q.IS + (long long)(y + 1) * q.Wp + (r_ - y * p.img_w)) * p.ldc;
          }
          const int img = mr / q.IS, r_ = mr - img * q.IS, yy = r_ / q.Wp, xx = r_ - yy * q.Wp;   // position -> dense
          if (yy < 1 || xx >= p.img_w) return nullptr;
          const int md = img * hw + (yy - 1) * p.img_w + xx;
          return (act_t*)p.C + (long long)(md % p.m_split) * p.ldc + (long long)(md / p.m_split) * p.split_col_off;
        };
        // the epilogue warps fill the table of this tile together; one named barrier per tile orders it (the table is
        // double buffered, and passing the barrier of tile i+1 implies every warp has finished tile i)
        for (int i = (warp - 2) * 32 + lane; i < BN; i += C::EPI_WARPS * 32) {
          const bool ok = n0 + i < nst;
          ss[i] = ((p.scale && ok && !(q.dbg & 2)) ? __ldg(p.scale + n0 + i) : 1.0f) * p.alpha;
          ss[BN + i] = (p.shift && ok && !(q.dbg & 2)) ? __ldg(p.shift + n0 + i) : 0.0f;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(C::EPI_WARPS * 32) : "memory");
        if (warp == 2) stamp(128);                             // scale/shift table ready
        const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BN;
#pragma unroll 1
        for (int g = ehalf; g < BN / 64; g += GSTEP) {
          const int nb = n0 + g * 64;
          const bool g_ok = nb < nst;                           // warp-uniform
          auto prefetch_next = [&]() {
            if (!use_res) return;
            if (g + GSTEP < BN / 64) load_res(m_w, nb + GSTEP * 64);
            else if (t + n_clusters < total_tiles) load_res(mw_next, n0_next + ehalf * 64);
          };
          uint32_t r0[32], r1[32];
          tmem_ld32_nowait(t_row + (uint32_t)(g * 64), r0);
          tmem_ld32_nowait(t_row + (uint32_t)(g * 64 + 32), r1);
          tmem_ld_wait();
          if (g + GSTEP >= BN / 64) {
            // release this accumulator buffer to the MMA issuer (leader CTA's barrier)
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (CG == 1) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&ctl->tmem_empty[acc])) : "memory");
              else mbar_arrive_cluster(&ctl->tmem_empty[acc], 0);
            }
          }
          if (!g_ok) { prefetch_next(); continue; }
          float v[64];
#pragma unroll
          for (int i = 0; i < 32; i++) { v[i] = __uint_as_float(r0[i]); v[32 + i] = __uint_as_float(r1[i]); }
          const uint32_t ss_a = smem_u32(ss) + (uint32_t)(g * 256);
#pragma unroll
          for (int c4 = 0; c4 < 16; c4++) {                   // alpha is folded into the scale table
            float4 sc, sh;
            asm("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(sc.x), "=f"(sc.y), "=f"(sc.z), "=f"(sc.w) : "r"(ss_a + c4 * 16));
            asm("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(sh.x), "=f"(sh.y), "=f"(sh.z), "=f"(sh.w) : "r"(ss_a + BN * 4 + c4 * 16));
            v[c4 * 4 + 0] = fmaf(v[c4 * 4 + 0], sc.x, sh.x); v[c4 * 4 + 1] = fmaf(v[c4 * 4 + 1], sc.y, sh.y);
            v[c4 * 4 + 2] = fmaf(v[c4 * 4 + 2], sc.z, sh.z); v[c4 * 4 + 3] = fmaf(v[c4 * 4 + 3], sc.w, sh.w);
          }
          if (p.row_shift) {
            const float rs = m < p.M ? __ldg(p.row_shift + m) : 0.0f;
#pragma unroll
            for (int i = 0; i < 64; i++) v[i] += rs;
          }
          if (use_res) {
#pragma unroll
            for (int i = 0; i < 8; i++)
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(s_res + stg_off(4 * i + sub, cj)), "r"(rr[i].x), "r"(rr[i].y), "r"(rr[i].z), "r"(rr[i].w) : "memory");
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; j++) {
              uint4 rv;
              asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(rv.x), "=r"(rv.y), "=r"(rv.z), "=r"(rv.w) : "r"(s_res + stg_off(lane, j)) : "memory");
              const act2_t* h = reinterpret_cast<const act2_t*>(&rv);
#pragma unroll
              for (int e = 0; e < 4; e++) {
                const float2 f = act22f2(h[e]);
                v[j * 8 + e * 2] += f.x; v[j * 8 + e * 2 + 1] += f.y;
              }
            }
          }
          const bool relu_packed = p.relu && !p.pe;             // max(bf16(x), 0) == bf16(max(x, 0))
          if (p.relu && p.pe) {
#pragma unroll
            for (int i = 0; i < 64; i++) v[i] = fmaxf(v[i], 0.0f);
          }
          if (p.pe) {
#pragma unroll
            for (int hlf = 0; hlf < 2; hlf++) {                 // 32 fp32 columns = 128 B per row per half
              __syncwarp();
#pragma unroll
              for (int i = 0; i < 8; i++) {
                int per = pe_r0 + 4 * i + sub;
                if (per >= p.pe_rows) per -= p.pe_rows;
                const float4 pv = __ldg(reinterpret_cast<const float4*>(p.pe + (long long)per * p.pe_ld + nb + hlf * 32) + cj);
                asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(s_pe + stg_off(4 * i + sub, cj)), "f"(pv.x), "f"(pv.y), "f"(pv.z), "f"(pv.w) : "memory");
              }
              __syncwarp();
#pragma unroll
              for (int j = 0; j < 8; j++) {
                float4 pv;
                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(pv.x), "=f"(pv.y), "=f"(pv.z), "=f"(pv.w) : "r"(s_pe + stg_off(lane, j)) : "memory");
                v[hlf * 32 + j * 4 + 0] += pv.x; v[hlf * 32 + j * 4 + 1] += pv.y;
                v[hlf * 32 + j * 4 + 2] += pv.z; v[hlf * 32 + j * 4 + 3] += pv.w;
              }
            }
          }
          if (!interior) {
#pragma unroll
            for (int i = 0; i < 64; i++) v[i] = 0.0f;
          }
          if ((q.dbg & 1) && v[0] != 123456.0f) { prefetch_next(); continue; }
          __syncwarp();                                         // staging tile: residual / pe reads done before it is reused
          const act2_t zero2 = f2act2(0.0f, 0.0f);
#pragma unroll
          for (int j = 0; j < 8; j++) {
            act2_t h0 = f2act2(v[j * 8 + 0], v[j * 8 + 1]), h1 = f2act2(v[j * 8 + 2], v[j * 8 + 3]);
            act2_t h2 = f2act2(v[j * 8 + 4], v[j * 8 + 5]), h3 = f2act2(v[j * 8 + 6], v[j * 8 + 7]);
            if (relu_packed) { h0 = __hmax2(h0, zero2); h1 = __hmax2(h1, zero2); h2 = __hmax2(h2, zero2); h3 = __hmax2(h3, zero2); }
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(s_out + stg_off(lane, j)), "r"(*reinterpret_cast<uint32_t*>(&h0)),
                         "r"(*reinterpret_cast<uint32_t*>(&h1)), "r"(*reinterpret_cast<uint32_t*>(&h2)), "r"(*reinterpret_cast<uint32_t*>(&h3)) : "memory");
          }
          __syncwarp();
          prefetch_next();
          if (p.row_remap == 0) {
            char* ob = (char*)(out_w + (long long)sub * p.ldc + nb + cj * 8);
            uint32_t pitch4 = (uint32_t)p.ldc * 8u;            // 4 rows, in bytes
            asm volatile("" : "+r"(pitch4));                  // one IMAD.WIDE per store instead of 8 live 64-bit offsets
            const int rows_left = p.M - m_w - sub;             // row 4i+sub is stored iff 4i < rows_left
#pragma unroll
            for (int i = 0; i < 8; i++) {
              uint4 ov;
              asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(ov.x), "=r"(ov.y), "=r"(ov.z), "=r"(ov.w) : "r"(s_out + stg_off(4 * i + sub, cj)) : "memory");
              if (4 * i < rows_left && !(q.dbg & 2048)) *reinterpret_cast<uint4*>(ob + (unsigned long long)i * pitch4) = ov;
            }
          } else {
#pragma unroll
            for (int i = 0; i < 8; i++) {
              const int mr = m_w + 4 * i + sub;
              uint4 ov;
              asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(ov.x), "=r"(ov.y), "=r"(ov.z), "=r"(ov.w) : "r"(s_out + stg_off(4 * i + sub, cj)) : "memory");
              act_t* op = row_ptr(mr);
              if (op && !(q.dbg & 2048)) *(reinterpret_cast<uint4*>(op + nb) + cj) = ov;
            }
          }
          __syncwarp();
          if (warp == 2) stamp(128);                           // 64-column group done
        }
        continue;
      }
      const bool m_ok = m < p.M;
      const long long orow = (m_ok ? (long long)(m % p.m_split) * p.ldc + (long long)(m / p.m_split) * p.split_col_off : 0) +
                             (long long)ks * q.c_split_stride;
      const float* pe_row = (p.pe && m_ok) ? p.pe + (long long)(m % p.pe_rows) * p.pe_ld : nullptr;
      const act_t* res_row = (p.residual && m_ok) ? (const act_t*)p.residual + (long long)m * p.ldr : nullptr;
#pragma unroll 1
      for (int chn = ehalf; chn < BN / 32; chn += GSTEP) {
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BN + (uint32_t)(chn * 32), r);
        const int nb = n0 + chn * 32;
        if (!m_ok || nb >= nst) continue;
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]) * p.alpha;
        const bool full = nb + 32 <= nst;
        if (p.scale && !(q.dbg & 2)) {
#pragma unroll
          for (int i = 0; i < 32; i++)
            if (full || nb + i < nst) v[i] = v[i] * __ldg(p.scale + nb + i) + (p.shift ? __ldg(p.shift + nb + i) : 0.0f);
        } else if (p.shift) {
#pragma unroll
          for (int i = 0; i < 32; i++)
            if (full || nb + i < nst) v[i] += __ldg(p.shift + nb + i);
        }
        if (p.row_shift) {
          const float rs = __ldg(p.row_shift + m);
#pragma unroll
          for (int i = 0; i < 32; i++) v[i] += rs;
        }
        if (res_row && !(q.dbg & 4)) {
          if (full) {
#pragma unroll
            for (int qq = 0; qq < 4; qq++) {
              const uint4 rv = *reinterpret_cast<const uint4*>(res_row + nb + qq * 8);
              const act2_t* h = reinterpret_cast<const act2_t*>(&rv);
#pragma unroll
              for (int e = 0; e < 4; e++) {
                const float2 f = act22f2(h[e]);
                v[qq * 8 + e * 2] += f.x; v[qq * 8 + e * 2 + 1] += f.y;
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
        if ((q.dbg & 1) && v[0] != 123456.0f) continue;
        if (p.out_bf16) {
          act_t* o = (act_t*)p.C + orow + nb;
          if (full) {
#pragma unroll
            for (int qq = 0; qq < 4; qq++) {
              uint4 pk;
              act2_t h0 = f2act2(v[qq * 8 + 0], v[qq * 8 + 1]), h1 = f2act2(v[qq * 8 + 2], v[qq * 8 + 3]);
              act2_t h2 = f2act2(v[qq * 8 + 4], v[qq * 8 + 5]), h3 = f2act2(v[qq * 8 + 6], v[qq * 8 + 7]);
              pk.x = *reinterpret_cast<uint32_t*>(&h0); pk.y = *reinterpret_cast<uint32_t*>(&h1);
              pk.z = *reinterpret_cast<uint32_t*>(&h2); pk.w = *reinterpret_cast<uint32_t*>(&h3);
              *reinterpret_cast<uint4*>(o + qq * 8) = pk;
            }
          } else {
            for (int i = 0; i < 32 && nb + i < nst; i++) o[i] = f2act(v[i]);
          }
        } else {
          float* o = (float*)p.C + orow + nb;
          if (full) {
#pragma unroll
            for (int qq = 0; qq < 8; qq++)
              *reinterpret_cast<float4*>(o + qq * 4) = make_float4(v[qq * 4], v[qq * 4 + 1], v[qq * 4 + 2], v[qq * 4 + 3]);
          } else {
            for (int i = 0; i < 32 && nb + i < nst; i++) o[i] = v[i];
          }
        }
      }
      // release this accumulator buffer to the MMA issuer (leader CTA's barrier)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG == 1) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&ctl->tmem_empty[acc])) : "memory");
        else mbar_arrive_cluster(&ctl->tmem_empty[acc], 0);
      }
    }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    if (CG == 1)
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS) : "memory");
    else
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// host side: tensor maps (driver entry points resolved at run time; no -lcuda link dependency)
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_tiled = nullptr;
EncodeIm2colFn g_im2col = nullptr;
int g_num_sms = 0;

int resolve_driver()
{
  static std::once_flag once;
  static int rc = FPB_OK;
  std::call_once(once, [] {
    cudaDriverEntryPointQueryResult qr;
    void* f = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess || !f) { rc = FPB_ERR_CUDA; return; }
    g_tiled = (EncodeTiledFn)f;
    f = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f, cudaEnableDefault, &qr) != cudaSuccess || !f) { rc = FPB_ERR_CUDA; return; }
    g_im2col = (EncodeIm2colFn)f;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
  });
  if (rc) fpb_set_error("cuTensorMapEncode* driver entry points are unavailable");
  return rc;
}

struct MapKey {
  const void* ptr;
  long long v[12];
  bool operator<(const MapKey& o) const { return memcmp(this, &o, sizeof(MapKey)) < 0; }
};
std::mutex g_map_mu;
std::map<MapKey, CUtensorMap> g_maps;

// tiled 2-D map over a row-major [rows][pitch] bf16 matrix; box = 64 columns x box_rows rows, 128B swizzle
int get_tiled_map(const void* ptr, long long rows, long long cols, long long pitch, int box_rows, CUtensorMap* out)
{
  MapKey k;
  memset(&k, 0, sizeof(k));
  k.ptr = ptr; k.v[0] = 1; k.v[1] = rows; k.v[2] = cols; k.v[3] = pitch; k.v[4] = box_rows;
  std::lock_guard<std::mutex> lk(g_map_mu);
  auto itf = g_maps.find(k);
  if (itf != g_maps.end()) { *out = itf->second; return FPB_OK; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
  cuuint32_t box[2] = {(cuuint32_t)KC, (cuuint32_t)box_rows};
  cuuint32_t es[2] = {1, 1};
  CUtensorMap m;
  CUresult r = g_tiled(&m, ACT_TMAP_DTYPE, 2, (void*)ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fpb_set_error("cuTensorMapEncodeTiled failed (%d): rows=%lld cols=%lld pitch=%lld box_rows=%d", (int)r, rows, cols, pitch, box_rows);
    return FPB_ERR_CUDA;
  }
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps[k] = m;
  *out = m;
  return FPB_OK;
}

// im2col map over an NHWC bf16 activation: 128 output pixels x 64 channels per request
int get_im2col_map(const GemmProblem& p, int nimg, CUtensorMap* out)
{
  const int row_pitch = p.row_pitch > 0 ? p.row_pitch : p.Win;
  MapKey k;
  memset(&k, 0, sizeof(k));
  k.ptr = p.A; k.v[0] = 2; k.v[1] = nimg; k.v[2] = p.Hin; k.v[3] = p.Win; k.v[4] = p.Cin; k.v[5] = p.pix_pitch;
  k.v[6] = p.KH; k.v[7] = p.KW; k.v[8] = p.stride; k.v[9] = p.pad; k.v[10] = p.pad_w; k.v[11] = ((long long)row_pitch << 20) | (p.Ho << 10) | p.Wo;
  std::lock_guard<std::mutex> lk(g_map_mu);
  auto itf = g_maps.find(k);
  if (itf != g_maps.end()) { *out = itf->second; return FPB_OK; }
  cuuint64_t dims[4] = {(cuuint64_t)p.Cin, (cuuint64_t)p.Win, (cuuint64_t)p.Hin, (cuuint64_t)nimg};
  cuuint64_t strides[3] = {(cuuint64_t)p.pix_pitch * 2, (cuuint64_t)row_pitch * p.pix_pitch * 2,
                           (cuuint64_t)p.Hin * row_pitch * p.pix_pitch * 2};
  // bounding box of the filter-window origin: [lower, dim - 1 + upper]; it must hold exactly Wo x Ho origins
  int lower[2] = {-p.pad_w, -p.pad};
  int upper[2] = {(p.Wo - 1) * p.stride - p.pad_w - (p.Win - 1), (p.Ho - 1) * p.stride - p.pad - (p.Hin - 1)};
  cuuint32_t es[4] = {1, (cuuint32_t)p.stride, (cuuint32_t)p.stride, 1};
  CUtensorMap m;
  CUresult r = g_im2col(&m, ACT_TMAP_DTYPE, 4, (void*)p.A, dims, strides, lower, upper, (cuuint32_t)KC,
                        (cuuint32_t)TM, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fpb_set_error("cuTensorMapEncodeIm2col failed (%d): nimg=%d H=%d W=%d C=%d pitch=%d k=%dx%d s=%d pad=%d,%d", (int)r, nimg,
                  p.Hin, p.Win, p.Cin, p.pix_pitch, p.KH, p.KW, p.stride, p.pad, p.pad_w);
    return FPB_ERR_CUDA;
  }
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps[k] = m;
  *out = m;
  return FPB_OK;
}

template <int BN, int CG>
int launch_tc2(const CUtensorMap& mA, const CUtensorMap& mB, const Tc2Params& q_in, cudaStream_t st)
{
  using C = Cfg<BN, CG>;
  Tc2Params q = q_in;
  if (q.halo) {                                    // split the pipeline smem into an A-halo ring and a W ring
    const int region = C::STAGES * C::STAGE_BYTES;
    q.na = 2;
    q.tps = C::B_STAGE_BYTES <= 8192 ? 3 : 1;                 // small W tiles: one stage = one kernel row (3 taps)
    q.nb = (region - q.na * A_HALO_BYTES) / (C::B_STAGE_BYTES * q.tps);
    if (q.nb > 8) q.nb = 8;
    if (q.nb < 2) { fpb_set_error("tc2 halo mode: not enough shared memory"); return FPB_ERR_ARG; }
  }
  static bool configured = false;
  if (!configured) {
    FPB_CUDA(cudaFuncSetAttribute(tc2_kernel<BN, CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    configured = true;
  }
  const int total = q.m_tiles * q.n_tiles * q.ksplit;
  int clusters = g_num_sms / CG;
  if (clusters > total) clusters = total;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(clusters * CG, 1, 1);
  cfg.blockDim = dim3(NTHREADS, 1, 1);
  cfg.dynamicSmemBytes = C::SMEM;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = fpb_pdl_enabled() ? 2 : 1;
  FPB_CUDA(cudaLaunchKernelEx(&cfg, tc2_kernel<BN, CG>, mA, mB, q));
  fpb_count_launch();
  return FPB_OK;
}

int g_force_cg = 0;

// Second half of a split-K GEMM: out = epilogue(sum over K slices of the fp32 partial tiles), slices added in index order
// (deterministic).  One thread per (row, 4 columns); same operation order as the kernel's epilogue.
__global__ void __launch_bounds__(256) splitk_reduce_kernel(GemmProblem p, const float* __restrict__ part, int ksplit,
                                                            long long stride, int ld, int nst)
{
  const int n4 = nst >> 2;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)p.M * n4) return;
  const int m = (int)(i / n4), n = (int)(i - (long long)m * n4) * 4;
  float4 a = *reinterpret_cast<const float4*>(part + (long long)m * ld + n);
  for (int k = 1; k < ksplit; k++) {
    const float4 b = *reinterpret_cast<const float4*>(part + k * stride + (long long)m * ld + n);
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
  }
  float v[4] = {a.x * p.alpha, a.y * p.alpha, a.z * p.alpha, a.w * p.alpha};
#pragma unroll
  for (int j = 0; j < 4; j++) {
    if (p.scale) v[j] = v[j] * p.scale[n + j] + (p.shift ? p.shift[n + j] : 0.0f);
    else if (p.shift) v[j] += p.shift[n + j];
    if (p.row_shift) v[j] += p.row_shift[m];
    if (p.residual) v[j] += act2f(((const act_t*)p.residual)[(long long)m * p.ldr + n + j]);
    if (p.relu) v[j] = fmaxf(v[j], 0.0f);
    if (p.pe) v[j] += p.pe[(long long)(m % p.pe_rows) * p.pe_ld + n + j];
  }
  const long long o = (long long)(m % p.m_split) * p.ldc + (long long)(m / p.m_split) * p.split_col_off + n;
  if (p.out_bf16) {
    act2_t h0 = f2act2(v[0], v[1]), h1 = f2act2(v[2], v[3]);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&h0); pk.y = *reinterpret_cast<uint32_t*>(&h1);
    *reinterpret_cast<uint2*>((act_t*)p.C + o) = pk;
  } else {
    *reinterpret_cast<float4*>((float*)p.C + o) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// split-K factor for a problem with `tiles` output tiles on `clusters` CTAs / CTA pairs: the largest divisor s of the
// nk K stages that keeps >= 4 stages per slice, fits the scratch buffer and does not exceed one wave
int choose_ksplit(const GemmProblem& p, int tiles, int clusters, int nk, int ncols)
{
  static int enabled = -1;
  if (enabled < 0) { const char* e = getenv("FPB_SPLITK"); enabled = e ? atoi(e) : 1; }
  if (!enabled || !p.splitk_ws || p.halo || p.row_remap || tiles * 2 > clusters || nk < 8) return 1;
  if (ncols % 4 || p.ldc % 4 || p.split_col_off % 4 || (p.residual && p.ldr % 2)) return 1;
  const long long per = (long long)p.M * ncols * 4;
  int smax = clusters / tiles;
  if (smax > nk / 4) smax = nk / 4;
  for (int s = smax; s >= 2; s--)
    if (nk % s == 0 && per * s <= p.splitk_ws_bytes) return s;
  return 1;
}


int tc2_cta_group()
{
  if (g_force_cg) return g_force_cg;
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("FPB_TC_CTA_GROUP");
    v = e ? atoi(e) : 2;
    if (v != 1 && v != 2) v = 2;
  }
  return v;
}

}  // namespace

void tc2_force_cta_group(int cg) { g_force_cg = cg; }

int fpb_pdl_enabled()
{
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("FPB_PDL");
    v = e ? (atoi(e) != 0) : 1;
  }
  return v;
}

// shared with net_attn.cu: cached tiled 2-D bf16 tensor map (box = 64 columns x box_rows rows, 128B swizzle)
int fpb_tiled_map_act(const void* ptr, long long rows, long long cols, long long pitch, int box_rows, void* out_map, int* num_sms)
{
  int rc = resolve_driver();
  if (rc) return rc;
  if (num_sms) *num_sms = g_num_sms;
  return get_tiled_map(ptr, rows, cols, pitch, box_rows, (CUtensorMap*)out_map);
}

int fpb_im2col_map_act(const GemmProblem& p, int nimg, void* out_map)
{
  int rc = resolve_driver();
  if (rc) return rc;
  return get_im2col_map(p, nimg, (CUtensorMap*)out_map);
}

// Returns FPB_ERR_UNSUPPORTED (without setting up anything) when the problem does not fit this kernel, so the
// caller can fall back to the first-generation cp.async kernel (tc_gemm).
int tc2_gemm(const GemmProblem& p, cudaStream_t st)
{
  if (p.batch != 1 || p.K % KC != 0 || p.Cin % KC != 0 || p.pix_pitch % 8 != 0 || p.ldw % 8 != 0) {
    if (p.halo || p.row_remap) { fpb_set_error("tc2_gemm: halo / row_remap problem has an unsupported shape"); return FPB_ERR_ARG; }
    return FPB_ERR_UNSUPPORTED;
  }
  if (((uintptr_t)p.A | (uintptr_t)p.W) & 15) return FPB_ERR_UNSUPPORTED;
  if (p.c_transposed) { fpb_set_error("tc2_gemm: transposed output is only implemented by the operand-swapped kernel"); return FPB_ERR_ARG; }
  const bool plain = !p.halo && p.KH == 1 && p.KW == 1 && p.stride == 1 && p.pad == 0 && p.pad_w == 0 && p.Ho == 1 && p.Wo == 1;
  if (!plain && (p.pad > 127 || p.KH > 127)) return FPB_ERR_UNSUPPORTED;
  int rc = resolve_driver();
  if (rc) return rc;
  const int ncols = p.n_valid > 0 ? p.n_valid : p.N;
  const int BN = (ncols % 256 == 0 || ncols > 384) ? 256 : (ncols > 64 ? 128 : 64);
  int CG = tc2_cta_group();
  if (BN / CG < 64) CG = 1;                        // keep >= 64 W rows per CTA
  if (!g_force_cg && !getenv("FPB_TC_CTA_GROUP") && (p.K < 2304 || BN < 256)) CG = 1;   // measured: pairs pay off only for long K
  Tc2Params q;
  q.p = p;
  q.a_im2col = plain ? 0 : 1;
  q.lower_w = -p.pad_w; q.lower_h = -p.pad;
  { const char* e = getenv("FPB_TC2_DBG"); q.dbg = e ? atoi(e) : 0; }
  q.fast_epi = (p.out_bf16 && ncols % 64 == 0 && !(q.dbg & 16) && (!p.residual || p.ldr % 8 == 0) && p.ldc % 8 == 0 &&
                p.split_col_off % 8 == 0 && (!p.pe || (p.pe_ld % 4 == 0 && p.pe_rows >= 32)) &&
                (p.m_split % 32 == 0 || p.m_split >= p.M)) ? 1 : 0;
  q.halo = p.halo ? 1 : 0;
  q.Wp = p.img_w + 1; q.IS = (p.img_h + 1) * q.Wp; q.hrows = TM + 2 * q.Wp + 2; q.na = q.nb = 0; q.tps = 1;
  if (p.halo || p.row_remap) {
    const bool ok = q.fast_epi && p.img_h > 0 && p.img_w > 0 && (!p.halo || (p.KH == 3 && p.KW == 3 && p.stride == 1 && q.hrows * 128 <= A_HALO_BYTES)) &&
                    (p.row_remap == 2 || p.m_split >= p.M) && !p.pe;
    if (!ok) { fpb_set_error("tc2_gemm: unsupported halo / row_remap configuration"); return FPB_ERR_ARG; }
    if (p.halo && !g_force_cg && !getenv("FPB_TC_CTA_GROUP") && BN >= 128) CG = 2;     // halves the weight stream, now the larger one
  }
  q.m_tiles = fpb_cdiv(p.M, TM * CG);
  q.n_tiles = fpb_cdiv(ncols, BN);
  q.ksplit = choose_ksplit(p, q.m_tiles * q.n_tiles, g_num_sms / CG, p.K / KC, ncols);
  q.c_split_stride = 0;
  if (q.ksplit > 1) {
    // first kernel: raw fp32 partial tiles [ksplit][M][ncols] into the scratch buffer, no epilogue math
    GemmProblem& s = q.p;
    s.C = p.splitk_ws; s.out_bf16 = 0; s.alpha = 1.0f; s.scale = nullptr; s.shift = nullptr; s.row_shift = nullptr;
    s.residual = nullptr; s.relu = 0; s.pe = nullptr; s.ldc = ncols; s.m_split = 0x7fffffff; s.split_col_off = 0;
    q.fast_epi = 0;
    q.c_split_stride = (long long)p.M * ncols;
  }
  CUtensorMap mA, mB;
  if (p.halo) {
    rc = get_tiled_map(p.A, p.M, p.Cin, p.pix_pitch, q.hrows, &mA);       // [positions][C], box = halo rows x 64 channels
  } else if (plain) {
    rc = get_tiled_map(p.A, p.M, p.K, p.pix_pitch, TM, &mA);
  } else {
    const int nimg = fpb_cdiv(p.M, p.Ho * p.Wo);
    rc = get_im2col_map(p, nimg, &mA);
  }
  if (rc) return rc;
  rc = get_tiled_map(p.W, p.N, p.K, p.ldw, BN / CG, &mB);
  if (rc) return rc;
  if (CG == 2) rc = BN == 256 ? launch_tc2<256, 2>(mA, mB, q, st) : launch_tc2<128, 2>(mA, mB, q, st);
  else if (BN == 256) rc = launch_tc2<256, 1>(mA, mB, q, st);
  else if (BN == 128) rc = launch_tc2<128, 1>(mA, mB, q, st);
  else rc = launch_tc2<64, 1>(mA, mB, q, st);
  if (rc || q.ksplit == 1) return rc;
  const long long work = (long long)p.M * (ncols / 4);
  splitk_reduce_kernel<<<fpb_cdiv(work, 256), 256, 0, st>>>(p, (const float*)p.splitk_ws, q.ksplit, q.c_split_stride, ncols, ncols);
  FPB_LAUNCH_CHECK();
  return FPB_OK;
}

}  // namespace FPB_ACT_NS
