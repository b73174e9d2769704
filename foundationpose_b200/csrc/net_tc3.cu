// Operand-swapped persistent tcgen05 / TMEM / TMA bf16 implicit GEMM for the narrow layers (Cout = 64 or 128):
//   stem 7x7, conv1, the 128-channel ResnetBasicBlocks    learning/models/refine_network.py:38-49, score_network.py:36-47
//                                                          learning/models/network_modules.py:37-50, 73-111
//
// Measured on B200 (tools/gemm_bench.py, FPB_TC2_DBG=384: no operand loads): a 128 x N x 16 tcgen05.mma with both
// operands in shared memory takes the same time for N = 128 as for N = 256, i.e. the N <= 128 layers run the tensor
// pipe at half rate (or a quarter, N = 64) when the pixels are the M operand.  This kernel therefore computes the
// TRANSPOSED product  D^T[channel][pixel] = W[channel][k] * X[pixel][k]^T:  the weights are the 128-row M operand
// (64-channel layers: the upper 64 rows of every weight tile stay zero in shared memory and are never loaded), 256
// pixels are the N operand, so every instruction is the full-rate 128 x 256 x 16 shape.  TMEM lanes hold channels,
// TMEM columns hold pixels; the epilogue transposes through a warp-private staging tile so that global stores and
// residual loads stay NHWC-coalesced (64 contiguous bytes per pixel per warp).
//
// Roles per CTA (one CTA per SM, persistent over tiles of 256 pixels):
//   warp 0     TMA producer  (im2col tensor map: 2 x 128 pixels x 64 channels per stage; tiled map for plain GEMMs;
//              "halo" mode: one [256 + 2*Wp + 2 positions] x 64-channel box serves the 9 taps of a 3x3 stride-1 conv,
//              see GemmProblem::halo, and only the weights stream per tap)
//   warp 1     TMEM allocator + single-thread MMA issuer, accumulators double buffered (2 x 256 columns)
//   warps 2-9  epilogue: tcgen05.ld (lane = channel) -> folded BN, residual, ReLU -> bf16 -> transposed stores
#include "net_kernels.cuh"
#include "tc_ptx.cuh"
#include <stdlib.h>

namespace FPB_ACT_NS {

namespace {

constexpr int TP = 256;                       // pixels per tile (MMA N)
constexpr int KC = 64;                        // K elements per pipeline stage (128 B rows)
constexpr int NT3 = 320;
constexpr int PIX_BYTES = TP * 128;           // 32 KB
constexpr int W_BYTES = 128 * 128;            // 16 KB (128 weight rows, the upper 64 zero for 64-channel layers)
constexpr int STAGE3 = PIX_BYTES + W_BYTES;
constexpr int NSTAGE3 = 4;
constexpr int RING_BYTES = NSTAGE3 * STAGE3;  // 192 KB
constexpr int HALO_BYTES = 44 * 1024;         // >= (256 + 2*Wp + 2) x 128 B for Wp <= 42
constexpr int NA3 = 2;                        // halo ring depth
constexpr int NB3 = (RING_BYTES - NA3 * HALO_BYTES) / W_BYTES;   // weight ring depth in halo mode (6)
constexpr int EPI3_WARP_BYTES = 4096;         // 2 KB result tile + 2 KB residual tile per epilogue warp
constexpr int EPI3_BYTES = 8 * EPI3_WARP_BYTES;

struct Ctl3 {
  uint64_t full[8];
  uint64_t empty[8];
  uint64_t afull[2];
  uint64_t aempty[2];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint64_t wfull;
  uint32_t tmem_base;
};
constexpr int SMEM3 = RING_BYTES + EPI3_BYTES + 1024 + (int)sizeof(Ctl3);
static_assert(SMEM3 <= 227 * 1024, "tc3 shared memory budget");
static_assert(NB3 >= 4 && NB3 <= 8, "tc3 halo weight ring");

struct Tc3Params {
  GemmProblem p;
  int mode;              // 0 plain tiled A, 1 im2col A, 2 halo, 3 plain A with the weight slab resident in smem
  int n_slabs, nstage;   // mode 3: 128-channel slabs of the output (one per CTA), pixel ring depth
  int lower_w, lower_h;  // im2col bounding-box lower corner
  int tiles;             // tiles of TP pixels
  int Wp, IS, hrows;     // halo mode: position pitch, image stride, A box rows
  int ch;                // 64 or 128 output channels
  int dbg;               // FPB_TC2_DBG bits: 1 no stores, 4 no residual, 8 no MMA, 128 no A loads, 256 no W loads
};

// CG = 2 (mode 3 only): a CTA pair shares every pixel tile -- each CTA keeps its own 128-channel weight slab resident,
// loads half of the 256 pixels (16 KB per stage) and the leader issues cta_group::2 MMAs (M = 2 x 128 channels,
// N = 256 pixels), so the activation stream per SM halves again and the ring is twice as deep.
template <int CG>
__global__ void __launch_bounds__(NT3, 1)
tc3_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapA2,
           const __grid_constant__ CUtensorMap mapW, const Tc3Params q)
{
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* epi_smem = smem + RING_BYTES;
  Ctl3* ctl = (Ctl3*)(epi_smem + EPI3_BYTES);
  const GemmProblem& p = q.p;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = p.K / KC;
  const bool halo = q.mode == 2;
  // tiles of this CTA: t_first, t_first + t_step, ...; mode 3 pins a CTA to one 128-channel slab of the output
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
  const int unit = (int)blockIdx.x / CG, n_units = (int)gridDim.x / CG;       // CTA (CG 1) or CTA pair (CG 2)
  const int n_groups = q.n_slabs / CG;                                          // slab groups: one per unit
  const int slab = q.mode == 3 ? (unit % n_groups) * CG + (int)rank : 0;
  const int t_first = q.mode == 3 ? unit / n_groups : unit;
  const int t_step = q.mode == 3 ? n_units / n_groups : n_units;
  pdl_launch_dependents();

  if (q.ch == 64) {
    // rows 64..127 of every weight slot are the zero half of the M operand; TMA only ever writes rows 0..63
    const int nslot = halo ? NB3 : NSTAGE3;
    for (int i = threadIdx.x; i < nslot * 512; i += NT3) {
      const int s = i >> 9, o = (i & 511) * 16;
      uint8_t* base = halo ? smem + NA3 * HALO_BYTES + s * W_BYTES : smem + s * STAGE3 + PIX_BYTES;
      *reinterpret_cast<uint4*>(base + 8192 + o) = make_uint4(0u, 0u, 0u, 0u);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < 8; s++) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
    for (int s = 0; s < 2; s++) { mbar_init(&ctl->afull[s], 1); mbar_init(&ctl->aempty[s], 1); }
    for (int a = 0; a < 2; a++) { mbar_init(&ctl->tmem_full[a], 1); mbar_init(&ctl->tmem_empty[a], 8 * CG); }
    mbar_init(&ctl->wfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if (CG == 1) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ctl->tmem_base)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ctl->tmem_base)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapW) : "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;
  pdl_wait();
  // FPB_TC2_DBG & 2048: CTA 0 records clock64 stamps into the output buffer (MMA thread -> slots 64.., warp 2 -> 128..)
  int n_st = 0;
  auto stamp = [&](int base) {
    if ((q.dbg & 2048) && blockIdx.x == 0 && lane == 0 && n_st < 64) reinterpret_cast<long long*>(p.C)[base + n_st++] = clock64();
  };
  const uint32_t w_tx = (q.dbg & 256) ? 0u : (uint32_t)(q.ch * 128);
  const bool no_a = q.dbg & 128, no_w = q.dbg & 256;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      if (halo) {
        const int nchunk = p.Cin / KC;
        const uint32_t sA0 = smem_u32(smem), sW0 = sA0 + NA3 * HALO_BYTES;
        const uint32_t a_tx = no_a ? 0u : (uint32_t)(q.hrows * 128);
        int sa = 0, sb = 0;
        uint32_t pa = 1, pb = 1;
        for (int t = t_first; t < q.tiles; t += t_step) {
          const int r0 = t * TP - q.Wp - 1;
          for (int c = 0; c < nchunk; c++) {
            mbar_wait(&ctl->aempty[sa], pa);
            mbar_arrive_expect_tx(&ctl->afull[sa], a_tx);
            if (!no_a) {
              tma_load_2d<1>(sA0 + sa * HALO_BYTES, &mapA, smem_u32(&ctl->afull[sa]), c * KC, r0);
              tma_load_2d<1>(sA0 + sa * HALO_BYTES + PIX_BYTES, &mapA2, smem_u32(&ctl->afull[sa]), c * KC, r0 + TP);
            }
            if (++sa == NA3) { sa = 0; pa ^= 1u; }
            int kcol = c * KC;
            for (int tap = 0; tap < 9; tap++, kcol += p.Cin) {
              mbar_wait(&ctl->empty[sb], pb);
              mbar_arrive_expect_tx(&ctl->full[sb], w_tx);
              if (!no_w) tma_load_2d<1>(sW0 + sb * W_BYTES, &mapW, smem_u32(&ctl->full[sb]), kcol, 0);
              if (++sb == NB3) { sb = 0; pb ^= 1u; }
            }
          }
        }
      } else if (q.mode == 3) {
        // weight slab [128 channels][K] once (nk tiles of 16 KB), then only the pixel stream: 32 KB / CG per stage.
        // CG 2: both CTAs signal the leader's barriers (the leader arms them with the pair's byte count)
        constexpr uint32_t PIXB = PIX_BYTES / CG;
        const uint32_t sW = smem_u32(smem), s0 = sW + nk * W_BYTES;
        const uint32_t wfull_a = CG == 2 ? mapa_u32(smem_u32(&ctl->wfull), 0) : smem_u32(&ctl->wfull);
        const uint32_t full_a = CG == 2 ? mapa_u32(smem_u32(&ctl->full[0]), 0) : smem_u32(&ctl->full[0]);
        if (rank == 0) mbar_arrive_expect_tx(&ctl->wfull, no_w ? 0u : (uint32_t)(CG * nk * W_BYTES));
        if (!no_w)
          for (int kc = 0; kc < nk; kc++) tma_load_2d<CG>(sW + kc * W_BYTES, &mapW, wfull_a, kc * KC, slab * 128);
        int s = 0;
        uint32_t ph = 1;
        for (int t = t_first; t < q.tiles; t += t_step) {
          const int m0 = t * TP + (int)rank * (TP / CG);
          for (int kc = 0; kc < nk; kc++) {
            mbar_wait(&ctl->empty[s], ph);
            if (rank == 0) mbar_arrive_expect_tx(&ctl->full[s], no_a ? 0u : (uint32_t)PIX_BYTES);
            if (!no_a) tma_load_2d<CG>(s0 + s * PIXB, &mapA, full_a + 8u * s, kc * KC, m0);
            if (++s == q.nstage) { s = 0; ph ^= 1u; }
          }
        }
      } else {
        const int hw = p.Ho * p.Wo;
        const uint32_t s0 = smem_u32(smem);
        int s = 0;
        uint32_t ph = 1;
        for (int t = t_first; t < q.tiles; t += t_step) {
          const int m0 = t * TP;
          const bool second = m0 + 128 < p.M;                  // a tile's upper 128 pixels may lie beyond M
          int cw[2] = {0, 0}, chh[2] = {0, 0}, img[2] = {0, 0};
          if (q.mode == 1) {
            for (int h = 0; h < 2; h++) {
              const int mm = m0 + 128 * h;
              img[h] = mm / hw;
              const int r = mm - img[h] * hw;
              const int oy = r / p.Wo, ox = r - oy * p.Wo;
              cw[h] = ox * p.stride + q.lower_w;
              chh[h] = oy * p.stride + q.lower_h;
            }
          }
          const uint32_t a_bytes = no_a ? 0u : (q.mode == 1 ? (second ? (uint32_t)PIX_BYTES : (uint32_t)PIX_BYTES / 2) : (uint32_t)PIX_BYTES);
          int k0 = 0, c0 = 0, kx = 0, ky = 0;
          for (int kc = 0; kc < nk; kc++, k0 += KC) {
            mbar_wait(&ctl->empty[s], ph);
            const uint32_t sp = s0 + s * STAGE3, bar = smem_u32(&ctl->full[s]);
            mbar_arrive_expect_tx(&ctl->full[s], a_bytes + w_tx);
            if (!no_a) {
              if (q.mode == 1) {
                tma_load_im2col<1>(sp, &mapA, bar, c0, cw[0], chh[0], img[0], (uint16_t)kx, (uint16_t)ky);
                if (second) tma_load_im2col<1>(sp + PIX_BYTES / 2, &mapA, bar, c0, cw[1], chh[1], img[1], (uint16_t)kx, (uint16_t)ky);
              } else {
                tma_load_2d<1>(sp, &mapA, bar, k0, m0);
              }
            }
            if (!no_w) tma_load_2d<1>(sp + PIX_BYTES, &mapW, bar, k0, 0);
            c0 += KC;
            if (c0 == p.Cin) { c0 = 0; if (++kx == p.KW) { kx = 0; ky++; } }
            if (++s == NSTAGE3) { s = 0; ph ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer: D[channel][pixel] += W . X^T
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(128, TP);
      const bool no_mma = q.dbg & 8;
      uint32_t acc = 0, acc_ph = 1;
      if (halo) {
        const int nchunk = p.Cin / KC;
        const uint32_t sA0 = smem_u32(smem), sW0 = sA0 + NA3 * HALO_BYTES;
        const uint32_t row_step = (uint32_t)(q.Wp * 128);
        int sa = 0, sb = 0;
        uint32_t pa = 0, pb = 0;
        for (int t = t_first; t < q.tiles; t += t_step) {
          mbar_wait(&ctl->tmem_empty[acc], acc_ph);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * TP;
          for (int c = 0; c < nchunk; c++) {
            mbar_wait(&ctl->afull[sa], pa);
            uint32_t x_addr = sA0 + sa * HALO_BYTES;            // a filter tap is a start-address shift (see net_tc2.cu)
            for (int ky = 0; ky < 3; ky++, x_addr += row_step) {
              for (int kx = 0; kx < 3; kx++) {
                mbar_wait(&ctl->full[sb], pb);
                tc_fence_after();
                const uint64_t dw = make_desc(sW0 + sb * W_BYTES), dx = make_desc(x_addr + 128u * kx);
                if (!no_mma) {
#pragma unroll
                  for (int k = 0; k < KC / 16; k++)
                    umma_bf16<1>(d_tmem, dw + (uint64_t)(k * 2), dx + (uint64_t)(k * 2), idesc, (c | ky | kx | k) ? 1u : 0u);
                }
                umma_commit<1>(&ctl->empty[sb]);
                if (++sb == NB3) { sb = 0; pb ^= 1u; }
              }
            }
            umma_commit<1>(&ctl->aempty[sa]);
            if (++sa == NA3) { sa = 0; pa ^= 1u; }
          }
          umma_commit<1>(&ctl->tmem_full[acc]);
          if (++acc == 2) { acc = 0; acc_ph ^= 1u; }
        }
      } else if (q.mode == 3) {
        if (rank == 0) {
          constexpr uint32_t PIXB = PIX_BYTES / CG;
          constexpr uint32_t idesc3 = make_idesc(128 * CG, TP);
          const uint32_t sW = smem_u32(smem), s0 = sW + nk * W_BYTES;
          const uint64_t dw0 = make_desc(sW);
          int s = 0;
          uint32_t ph = 0;
          mbar_wait(&ctl->wfull, 0);
          for (int t = t_first; t < q.tiles; t += t_step) {
            mbar_wait(&ctl->tmem_empty[acc], acc_ph);
            tc_fence_after();
            stamp(64);
            const uint32_t d_tmem = tmem_base + acc * TP;
            for (int kc = 0; kc < nk; kc++) {
              mbar_wait(&ctl->full[s], ph);
              tc_fence_after();
              const uint64_t dx = make_desc(s0 + s * PIXB), dw = dw0 + (uint64_t)(kc * (W_BYTES >> 4));
              if (!no_mma) {
#pragma unroll
                for (int k = 0; k < KC / 16; k++)
                  umma_bf16<CG>(d_tmem, dw + (uint64_t)(k * 2), dx + (uint64_t)(k * 2), idesc3, (kc | k) ? 1u : 0u);
              }
              umma_commit<CG>(&ctl->empty[s]);
              if (++s == q.nstage) { s = 0; ph ^= 1u; }
            }
            stamp(64);
            umma_commit<CG>(&ctl->tmem_full[acc]);
            if (++acc == 2) { acc = 0; acc_ph ^= 1u; }
          }
        }
      } else {
        const uint32_t s0 = smem_u32(smem);
        int s = 0;
        uint32_t ph = 0;
        for (int t = t_first; t < q.tiles; t += t_step) {
          mbar_wait(&ctl->tmem_empty[acc], acc_ph);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * TP;
          for (int kc = 0; kc < nk; kc++) {
            mbar_wait(&ctl->full[s], ph);
            tc_fence_after();
            const uint32_t sp = s0 + s * STAGE3;
            const uint64_t dx = make_desc(sp), dw = make_desc(sp + PIX_BYTES);
            if (!no_mma) {
#pragma unroll
              for (int k = 0; k < KC / 16; k++)
                umma_bf16<1>(d_tmem, dw + (uint64_t)(k * 2), dx + (uint64_t)(k * 2), idesc, (kc | k) ? 1u : 0u);
            }
            umma_commit<1>(&ctl->empty[s]);
            if (++s == NSTAGE3) { s = 0; ph ^= 1u; }
          }
          umma_commit<1>(&ctl->tmem_full[acc]);
          if (++acc == 2) { acc = 0; acc_ph ^= 1u; }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    const int quarter = warp & 3;                 // TMEM lanes (= channels) 32*quarter .. +31
    const int half = (warp - 2) >> 2;             // pixel columns 128*half .. +127 of the tile
    const int ch0 = slab * 128 + quarter * 32;     // first output channel of this warp
    const bool active = quarter * 32 < q.ch;
    const float sc = ((active && p.scale) ? __ldg(p.scale + ch0 + lane) : 1.0f) * p.alpha;
    const float sh = (active && p.shift) ? __ldg(p.shift + ch0 + lane) : 0.0f;
    const uint32_t s_out = smem_u32(epi_smem + (warp - 2) * EPI3_WARP_BYTES), s_res = s_out + 2048;
    const int spx = lane >> 2, scj = lane & 3;    // coalesced role: pixel 8i + spx, 16-byte chunk scj (8 channels)
    const bool use_res = active && p.residual && !(q.dbg & 4);
    const int hw = p.img_h * p.img_w;
    // global address of the 64 channels-bytes this lane stores for pixel row mr (nullptr: not stored)
    auto row_ptr = [&](int mr) -> act_t* {
      if (mr >= p.M) return nullptr;
      if (p.row_remap == 0) {
        if (p.m_split >= p.M) return (act_t*)p.C + (long long)mr * p.ldc;
        const int seg = mr / p.m_split;
        return (act_t*)p.C + (long long)(mr - seg * p.m_split) * p.ldc + (long long)seg * p.split_col_off;
      }
      if (p.row_remap == 1) {                               // dense pixel index -> zero-bordered position
        const int img = mr / hw, r_ = mr - img * hw, y = r_ / p.img_w;
        return (act_t*)p.C + ((long long)img * q.IS + (long long)(y + 1) * q.Wp + (r_ - y * p.img_w)) * p.ldc;
      }
      const int img = mr / q.IS, r_ = mr - img * q.IS, yy = r_ / q.Wp, xx = r_ - yy * q.Wp;   // position -> dense
      if (yy < 1 || xx >= p.img_w) return nullptr;
      const int md = img * hw + (yy - 1) * p.img_w + xx;
      const int seg = md / p.m_split;
      return (act_t*)p.C + (long long)(md - seg * p.m_split) * p.ldc + (long long)seg * p.split_col_off;
    };
    // Direct path (dense output rows): lane = channel, so one warp instruction moves the 32 channels x 2 B = 64
    // contiguous bytes (two full sectors) of one pixel; no transposition, no staging, ~4x fewer instructions than
    // the staged path below (measured: the staged epilogue of a 64-channel layer, 4 active warps, was the bound).
    const bool direct = p.row_remap == 0 && !halo;
    unsigned short rq[32];
    auto load_res_direct = [&](int mbase) {
      if (mbase >= p.M) return;
      const char* rp = (const char*)((const unsigned short*)p.residual + (long long)mbase * p.ldr + ch0 + lane);
      uint32_t pitch = (uint32_t)p.ldr * 2u;
      asm volatile("" : "+r"(pitch));                 // keep the 32 row offsets out of registers (one IMAD.WIDE per access)
      if (mbase + 32 <= p.M) {
#pragma unroll
        for (int j = 0; j < 32; j++) rq[j] = __ldg((const unsigned short*)(rp + (unsigned long long)j * pitch));
      } else {
#pragma unroll
        for (int j = 0; j < 32; j++) rq[j] = mbase + j < p.M ? __ldg((const unsigned short*)(rp + (unsigned long long)j * pitch)) : (unsigned short)0;
      }
    };
    uint4 rr[4];
    auto load_res = [&](int mbase) {
      if (direct) { load_res_direct(mbase); return; }
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int mr = mbase + 8 * i + spx;
        rr[i] = mr < p.M ? __ldg(reinterpret_cast<const uint4*>((const act_t*)p.residual + (long long)mr * p.ldr + ch0) + scj)
                         : make_uint4(0u, 0u, 0u, 0u);
      }
    };
    if (use_res && t_first < q.tiles) load_res(t_first * TP + half * 128);
    uint32_t tc = 0;
    for (int t = t_first; t < q.tiles; t += t_step, tc++) {
      const uint32_t acc = tc & 1u;
      if (warp == 2) stamp(128);
      mbar_wait(&ctl->tmem_full[acc], (tc >> 1) & 1u);
      tc_fence_after();
      if (warp == 2) stamp(128);
      const int mh = t * TP + half * 128;          // first pixel of this warp's columns
      const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * TP + (uint32_t)(half * 128);
#pragma unroll 1
      for (int g = 0; g < 4; g++) {
        const int mg = mh + g * 32;
        uint32_t r[32];
        if (active) tmem_ld32(t_addr + (uint32_t)(g * 32), r);
        if (warp == 2) stamp(128);                   // TMEM load of this group returned
        if (g == 3) {                               // all TMEM reads of this warp are done: release the accumulator
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (CG == 1) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&ctl->tmem_empty[acc])) : "memory");
            else mbar_arrive_cluster(&ctl->tmem_empty[acc], 0);
          }
        }
        if (!active || mg >= p.M) continue;         // warp-uniform
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] = fmaf(__uint_as_float(r[i]), sc, sh);
        if (direct) {
          if (use_res) {
#pragma unroll
            for (int j = 0; j < 32; j++) v[j] += actbits2f(rq[j]);
            if (g < 3) load_res_direct(mg + 32);
            else if (t + t_step < q.tiles) load_res_direct((t + t_step) * TP + half * 128);
          }
          if (p.relu) {
#pragma unroll
            for (int i = 0; i < 32; i++) v[i] = fmaxf(v[i], 0.0f);
          }
          if ((q.dbg & 1) && v[0] != 123456.0f) continue;
          if (p.c_transposed) {
            // out[channel][pixel]: the accumulator layout itself; 32 pixels of this lane's channel = 64 contiguous bytes
            act_t* ot = (act_t*)p.C + (long long)(ch0 + lane) * p.ldc + mg;
            if (mg + 32 <= p.M) {
#pragma unroll
              for (int j4 = 0; j4 < 4; j4++) {
                uint4 pk;
                uint32_t* pw = reinterpret_cast<uint32_t*>(&pk);
#pragma unroll
                for (int e = 0; e < 4; e++) {
                  act2_t hh = f2act2(v[j4 * 8 + e * 2], v[j4 * 8 + e * 2 + 1]);
                  pw[e] = *reinterpret_cast<uint32_t*>(&hh);
                }
                reinterpret_cast<uint4*>(ot)[j4] = pk;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; j++)
                if (mg + j < p.M) ot[j] = f2act(v[j]);
            }
            continue;
          }
          // a group of 32 pixels never straddles m_split (host: m_split % 32 == 0 or m_split >= M)
          const int seg = p.m_split >= p.M ? 0 : mg / p.m_split;
          // Sub-word stores pace the LSU (measured: 32 STG.U16 per group made the epilogue the bound of the short-K
          // layers), so neighbouring lanes swap one value per pixel pair and every lane stores a bf16x2: even lanes
          // write channels (c, c+1) of pixel j, odd lanes channels (c-1, c) of pixel j+1 -- 16 four-byte stores.
          const int odd = lane & 1;
          char* op = (char*)((act_t*)p.C + (long long)(mg - seg * p.m_split + odd) * p.ldc + (long long)seg * p.split_col_off + ch0 + (lane & ~1));
          uint32_t pitch2 = (uint32_t)p.ldc * 4u;            // two pixel rows, in bytes
          asm volatile("" : "+r"(pitch2));
          const int left = p.M - mg - odd;                   // pixel j + odd is stored iff j < left
#pragma unroll
          for (int jj = 0; jj < 16; jj++) {
            const float mine_keep = odd ? v[2 * jj + 1] : v[2 * jj];
            const float other = __shfl_xor_sync(0xffffffffu, odd ? v[2 * jj] : v[2 * jj + 1], 1);
            const act2_t hh = odd ? f2act2(other, mine_keep) : f2act2(mine_keep, other);
            if (2 * jj < left && !(q.dbg & 2048)) *(act2_t*)(op + (unsigned long long)jj * pitch2) = hh;
          }
          continue;
        }
        if (use_res) {
#pragma unroll
          for (int i = 0; i < 4; i++)
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(s_res + (uint32_t)((8 * i + spx) * 64 + scj * 16)), "r"(rr[i].x), "r"(rr[i].y), "r"(rr[i].z), "r"(rr[i].w) : "memory");
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 32; j++) {
            uint16_t hv;
            asm volatile("ld.shared.u16 %0, [%1];" : "=h"(hv) : "r"(s_res + (uint32_t)(j * 64 + lane * 2)) : "memory");
            v[j] += actbits2f(hv);
          }
          // prefetch the residual of the next 32 pixels (next group, or the first group of the next tile)
          if (g < 3) load_res(mg + 32);
          else if (t + t_step < q.tiles) load_res((t + t_step) * TP + half * 128);
        }
        if (p.relu) {
#pragma unroll
          for (int i = 0; i < 32; i++) v[i] = fmaxf(v[i], 0.0f);
        }
        if ((q.dbg & 1) && v[0] != 123456.0f) continue;
        if (!(q.dbg & 1024))
#pragma unroll
        for (int j = 0; j < 32; j++) {
          const act_t hb = f2act(v[j]);
          asm volatile("st.shared.u16 [%0], %1;" ::"r"(s_out + (uint32_t)(j * 64 + lane * 2)), "h"(*reinterpret_cast<const uint16_t*>(&hb)) : "memory");
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const int mr = mg + 8 * i + spx;
          uint4 ov;
          asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(ov.x), "=r"(ov.y), "=r"(ov.z), "=r"(ov.w) : "r"(s_out + (uint32_t)((8 * i + spx) * 64 + scj * 16)) : "memory");
          act_t* op = row_ptr(mr);
          if (halo && p.row_remap == 0 && op) {     // border positions of the halo layout are written as zeros
            const int r_ = mr % q.IS, yy = r_ / q.Wp;
            if (yy < 1 || r_ - yy * q.Wp >= p.img_w) ov = make_uint4(0u, 0u, 0u, 0u);
          }
          if (op && !(q.dbg & 512)) *(reinterpret_cast<uint4*>(op + ch0) + scj) = ov;
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    if (CG == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

int tc3_enabled()
{
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("FPB_TC3");
    v = e ? atoi(e) : 1;
  }
  return v;
}

}  // namespace

// Returns FPB_ERR_UNSUPPORTED when the problem is not a narrow (64 / 128 output channels) bf16 layer this kernel
// covers; the caller then uses tc2_gemm.
int tc3_gemm(const GemmProblem& p, cudaStream_t st, int force)
{
  if (!force && !tc3_enabled()) return FPB_ERR_UNSUPPORTED;
  const int ncols = p.n_valid > 0 ? p.n_valid : p.N;
  if (p.batch != 1 || p.K % KC != 0 || p.Cin % KC != 0 || p.pix_pitch % 8 != 0 || p.ldw % 8 != 0) return FPB_ERR_UNSUPPORTED;
  const bool plain0 = !p.halo && p.KH == 1 && p.KW == 1 && p.stride == 1 && p.pad == 0 && p.pad_w == 0 && p.Ho == 1 && p.Wo == 1;
  // wide plain GEMMs with a short K (the transformer linears, K = 512): one 128-channel weight slab stays resident
  // in shared memory per CTA and only the activations stream (measured: these layers are bound by the operand
  // stream from L2, ~57 B/clk/SM, and 2/3 of that stream were weights)
  static int wres_env = -1;
  if (wres_env < 0) { const char* e = getenv("FPB_TC3_WRES"); wres_env = e ? atoi(e) : 1; }
  const bool wres = (wres_env || force) && plain0 && ncols % 128 == 0 && ncols >= 256 && p.N == ncols && p.K <= 512 && !p.row_remap && p.m_split >= p.M;
  if (!wres && ((ncols != 64 && ncols != 128) || p.N != ncols)) return FPB_ERR_UNSUPPORTED;
  if (!p.out_bf16 || p.pe || p.row_shift) return FPB_ERR_UNSUPPORTED;
  if (p.c_transposed && (p.residual || p.halo || p.row_remap || p.m_split < p.M)) return FPB_ERR_UNSUPPORTED;
  if (p.ldc % 8 != 0 || p.split_col_off % 8 != 0 || (p.residual && p.ldr % 8 != 0)) return FPB_ERR_UNSUPPORTED;
  if (p.m_split < p.M && p.m_split % 32 != 0) return FPB_ERR_UNSUPPORTED;
  if (((uintptr_t)p.A | (uintptr_t)p.W | (uintptr_t)p.C | (uintptr_t)p.residual) & 15) return FPB_ERR_UNSUPPORTED;
  const bool plain = !p.halo && p.KH == 1 && p.KW == 1 && p.stride == 1 && p.pad == 0 && p.pad_w == 0 && p.Ho == 1 && p.Wo == 1;
  if (!plain && (p.pad > 127 || p.KH > 127)) return FPB_ERR_UNSUPPORTED;
  Tc3Params q;
  q.p = p;
  q.mode = p.halo ? 2 : (wres ? 3 : (plain ? 0 : 1));
  q.lower_w = -p.pad_w; q.lower_h = -p.pad;
  q.ch = wres ? 128 : ncols;
  q.n_slabs = wres ? ncols / 128 : 1;
  // pairs (cta_group::2) whenever the slabs pair up; FPB_TC3_CG=1 forces single CTAs
  static int cg_env = -1;
  if (cg_env < 0) { const char* e = getenv("FPB_TC3_CG"); cg_env = e ? atoi(e) : 2; }
  const int CGn = (wres && q.n_slabs % 2 == 0 && cg_env == 2) ? 2 : 1;
  q.nstage = wres ? (RING_BYTES + EPI3_BYTES - (p.K / KC) * W_BYTES) / (PIX_BYTES / CGn) : NSTAGE3;
  if (q.nstage > 8) q.nstage = 8;
  { const char* e = getenv("FPB_TC2_DBG"); q.dbg = e ? atoi(e) : 0; }
  q.Wp = p.img_w + 1; q.IS = (p.img_h + 1) * q.Wp; q.hrows = TP + 2 * q.Wp + 2;
  if (p.halo || p.row_remap) {
    const bool ok = p.img_h > 0 && p.img_w > 0 && (!p.halo || (p.KH == 3 && p.KW == 3 && p.stride == 1 && q.hrows * 128 <= HALO_BYTES && q.hrows - TP <= 256)) &&
                    (p.row_remap == 2 || p.m_split >= p.M);
    if (!ok) return FPB_ERR_UNSUPPORTED;
  }
  q.tiles = fpb_cdiv(p.M, TP);
  int num_sms = 0;
  CUtensorMap mA, mA2, mW;
  int rc;
  if (p.halo) {
    rc = fpb_tiled_map_act(p.A, p.M, p.Cin, p.pix_pitch, TP, &mA, &num_sms);
    if (!rc) rc = fpb_tiled_map_act(p.A, p.M, p.Cin, p.pix_pitch, q.hrows - TP, &mA2, &num_sms);
  } else if (plain) {
    rc = fpb_tiled_map_act(p.A, p.M, p.K, p.pix_pitch, TP / CGn, &mA, &num_sms);
    mA2 = mA;
  } else {
    rc = fpb_im2col_map_act(p, fpb_cdiv(p.M, p.Ho * p.Wo), &mA);
    mA2 = mA;
  }
  if (rc) return rc;
  rc = fpb_tiled_map_act(p.W, p.N, p.K, p.ldw, wres ? 128 : ncols, &mW, &num_sms);
  if (rc) return rc;
  static bool configured = false;
  if (!configured) {
    FPB_CUDA(cudaFuncSetAttribute(tc3_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM3));
    FPB_CUDA(cudaFuncSetAttribute(tc3_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM3));
    configured = true;
  }
  int grid = q.tiles < num_sms ? q.tiles : num_sms;
  if (wres) {                                       // CTAs (pairs) per slab (pair) x number of slab groups
    const int groups = q.n_slabs / CGn;
    int per = (num_sms / CGn) / groups;
    if (per > q.tiles) per = q.tiles;
    if (per < 1) return FPB_ERR_UNSUPPORTED;
    grid = per * groups * CGn;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(NT3, 1, 1);
  cfg.dynamicSmemBytes = SMEM3;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CGn; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = fpb_pdl_enabled() ? 2 : 1;
  if (CGn == 2) FPB_CUDA(cudaLaunchKernelEx(&cfg, tc3_kernel<2>, mA, mA2, mW, q));
  else FPB_CUDA(cudaLaunchKernelEx(&cfg, tc3_kernel<1>, mA, mA2, mW, q));
  fpb_count_launch();
  return FPB_OK;
}

}  // namespace FPB_ACT_NS
