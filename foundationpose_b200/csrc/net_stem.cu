// 7x7 stride-2 stem convolution (6 -> 64 channels, 160x160 -> 80x80) as a dedicated persistent tcgen05 kernel.
//   learning/models/refine_network.py:38-41, score_network.py:36-39 (encodeA[0] = ConvBNReLU(C_in, 64, kernel 7, stride 2))
//
// Why a kernel of its own: as an implicit GEMM over im2col rows the stem re-reads every input row 7 times (once per
// kernel row) in 128-byte windows that start every 32 bytes, with a 64-row M operand -- measured 411 us per 504-image
// pass, bound by the L2 -> SM operand stream (2.9 GB) and the half-empty MMA shape.  Here
//   * the input rows are fetched ONCE per tile as plain boxes of the row-padded network input [img][160][168][8]
//     (the render / warp kernels write that layout): seen through a 5-D tensor map
//     [16 elements][84 pixel pairs][4 row phases][40 rows][img], the rows with the same (row mod 4) form a "plane";
//   * a pixel pair (2 px x 8 ch = 16 bf16 = 32 B) is exactly one K step of a kind::f16 MMA, and a plane sits in shared
//     memory in the K-major SWIZZLE_32B layout (32-byte rows), so the operand "pixel pair ox + t of plane row r + s" for
//     ALL 256 pixels of a tile is just a start-address shift of ((s * 84 + t) * 32 bytes) in the B descriptor -- the
//     swizzle is a function of the absolute shared-memory address (net_tc2.cu, halo mode), no im2col copy exists;
//   * two output rows are stacked in the 128 M rows: lanes 0-63 = 64 channels of output row 2r, lanes 64-127 = the same
//     channels of output row 2r+1 (input rows 4r+ky'-3 with the weights of kernel row ky' resp. ky'-2), so one
//     128 x 256 x 16 instruction is fully used: 36 MMAs produce 2 x 256 pixels x 64 channels (im2col: 2 x 28 half-empty);
//   * the stacked weights [9 ky'][128][4 pairs x 16] (144 KB) stay resident in shared memory for the whole kernel.
// A tile = 3 "double rows" (6 output rows) of one image = 252 flat pixels (+4 ignored) with 84 pixels per row, 4 of
// them junk; 14 tiles per image.
#include "net_kernels.cuh"
#include "tc_ptx.cuh"
#include <map>
#include <mutex>
#include <stdlib.h>
#include <string.h>

namespace FPB_ACT_NS {

namespace {

constexpr int ST_NT = 320;                         // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue
constexpr int ST_IMG = 160, ST_PITCH = 168, ST_C = 8, ST_OUT = 80, ST_CH = 64;
constexpr int ST_PAIRS = ST_PITCH / 2;             // 84 pixel pairs per input row
constexpr int ST_PROWS = ST_IMG / 4;               // 40 rows per plane = 40 double output rows
constexpr int ST_TROWS = 3;                        // double rows per tile
constexpr int ST_TILES_PER_IMG = (ST_PROWS + ST_TROWS - 1) / ST_TROWS;   // 14
constexpr int ST_BOX_BYTES = 5 * ST_PAIRS * 32;    // largest box: 5 plane rows (plane 1), the others 4
constexpr int ST_SLOT = 13824;                     // >= (2*84 + 3 + 256) * 32 = 13664, multiple of 256
constexpr int ST_NSLOT = 6;
constexpr int ST_WCHUNK = 128 * 128;               // one kernel row ky': 128 stacked rows x 64 K elements
constexpr int ST_WBYTES = 9 * ST_WCHUNK;           // 147456
static_assert(ST_SLOT % 256 == 0 && ST_SLOT >= (2 * ST_PAIRS + 3 + 256) * 32 && ST_SLOT >= ST_BOX_BYTES, "stem slot size");

struct StemCtl {
  uint64_t full[ST_NSLOT];
  uint64_t empty[ST_NSLOT];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint64_t wfull;
  uint32_t tmem_base;
};
constexpr int ST_SMEM = 1024 + ST_WBYTES + ST_NSLOT * ST_SLOT + (int)sizeof(StemCtl);
static_assert(ST_SMEM <= 227 * 1024, "stem shared memory budget");

struct StemParams {
  const float* scale;      // [64] folded BatchNorm
  const float* shift;
  act_t* out;      // [nimg][80][80][64]
  int nimg, relu;
  int dbg;                 // FPB_STEM_DBG bits: 1 no stores, 8 no MMA, 128 no input loads (timing experiments)
};

// plane pl = input rows with (row mod 4) == pl; it serves the kernel rows ky' = ((pl + 3) & 3) + 4 i, i < plane_n(pl), whose
// operand starts i plane rows below the first row of the box; the box starts plane_first(pl) rows from the tile's first
// double row and has plane_n(pl) + 2 rows
__device__ __forceinline__ constexpr int plane_n(int pl) { return pl == 1 ? 3 : 2; }
__device__ __forceinline__ constexpr int plane_ky(int pl, int i) { return ((pl + 3) & 3) + 4 * i; }
__device__ __forceinline__ constexpr int plane_first(int pl) { return pl == 0 ? 0 : -1; }

// K-major SWIZZLE_32B descriptor: 32-byte rows, 8-row groups 256 B apart
__device__ __forceinline__ uint64_t make_desc_sw32(uint32_t saddr)
{
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(256 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)6 << 61;
  return d;
}

__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* map, uint32_t mbar, int c0, int c1, int c2, int c3, int c4)
{
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(dst),
      "l"(map), "r"(mbar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

__global__ void __launch_bounds__(ST_NT, 1)
stem_kernel(const __grid_constant__ CUtensorMap mapX4, const __grid_constant__ CUtensorMap mapX5, const __grid_constant__ CUtensorMap mapW,
            const StemParams q)
{
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* ring = smem + ST_WBYTES;
  StemCtl* ctl = (StemCtl*)(ring + ST_NSLOT * ST_SLOT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = q.nimg * ST_TILES_PER_IMG;
  pdl_launch_dependents();

  if (threadIdx.x == 0) {
    for (int s = 0; s < ST_NSLOT; s++) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
    for (int a = 0; a < 2; a++) { mbar_init(&ctl->tmem_full[a], 1); mbar_init(&ctl->tmem_empty[a], 8); }
    mbar_init(&ctl->wfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapX4) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapX5) : "memory");
    // the weights are constants of the network: fetch them while the predecessor kernel is still draining
    mbar_arrive_expect_tx(&ctl->wfull, (uint32_t)ST_WBYTES);
    for (int k = 0; k < 9; k++) tma_load_2d<1>(smem_u32(smem + k * ST_WCHUNK), &mapW, smem_u32(&ctl->wfull), 0, k * 128);
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ctl->tmem_base)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;
  pdl_wait();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer: one box per plane per tile
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 1;                                     // waits on empty with inverted parity
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int img = t / ST_TILES_PER_IMG, r0 = (t - img * ST_TILES_PER_IMG) * ST_TROWS;
#pragma unroll
        for (int pl = 0; pl < 4; pl++) {
          mbar_wait(&ctl->empty[s], ph);
          const uint32_t bytes = (uint32_t)((plane_n(pl) + 2) * ST_PAIRS * 32);
          mbar_arrive_expect_tx(&ctl->full[s], (q.dbg & 128) ? 0u : bytes);
          if (!(q.dbg & 128))
            tma_load_5d(smem_u32(ring + s * ST_SLOT), pl == 1 ? &mapX5 : &mapX4, smem_u32(&ctl->full[s]), 0, 0, pl, r0 + plane_first(pl), img);
          if (++s == ST_NSLOT) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(128, 256);
      const uint32_t sW = smem_u32(smem), sR = smem_u32(ring);
      int s = 0;
      uint32_t ph = 0, acc = 0, acc_ph = 1;
      mbar_wait(&ctl->wfull, 0);
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        mbar_wait(&ctl->tmem_empty[acc], acc_ph);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
#pragma unroll
        for (int pl = 0; pl < 4; pl++) {
          mbar_wait(&ctl->full[s], ph);
          tc_fence_after();
          const uint32_t sb = sR + s * ST_SLOT;
          if (!(q.dbg & 8)) {
#pragma unroll
            for (int i = 0; i < plane_n(pl); i++) {
              const uint64_t dw = make_desc(sW + plane_ky(pl, i) * ST_WCHUNK);
              const uint64_t db = make_desc_sw32(sb + (uint32_t)(i * ST_PAIRS * 32));
#pragma unroll
              for (int k = 0; k < 4; k++)
                umma_bf16<1>(d_tmem, dw + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (pl | i | k) ? 1u : 0u);
            }
          }
          umma_commit<1>(&ctl->empty[s]);
          if (++s == ST_NSLOT) { s = 0; ph ^= 1u; }
        }
        umma_commit<1>(&ctl->tmem_full[acc]);
        if (++acc == 2) { acc = 0; acc_ph ^= 1u; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9): lane = stacked channel
    const int quarter = warp & 3;                 // TMEM lanes 32*quarter .. +31
    const int half = (warp - 2) >> 2;             // pixel columns 128*half .. +127
    const int ch0 = (quarter & 1) * 32;           // first channel of this warp
    const int par = quarter >> 1;                 // 0: even output row of the double row, 1: odd
    const float sc = q.scale ? __ldg(q.scale + ch0 + lane) : 1.0f;
    const float sh = q.shift ? __ldg(q.shift + ch0 + lane) : 0.0f;
    const int odd = lane & 1;
    const float lo = q.relu ? 0.0f : -INFINITY;
    uint32_t tc = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, tc++) {
      const uint32_t acc = tc & 1u;
      const int img = t / ST_TILES_PER_IMG, r0 = (t - img * ST_TILES_PER_IMG) * ST_TROWS;
      mbar_wait(&ctl->tmem_full[acc], (tc >> 1) & 1u);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * 256 + (uint32_t)(half * 128);
      act_t* obase = q.out + (long long)img * (ST_OUT * ST_OUT * ST_CH) + ch0 + (lane & ~1);
#pragma unroll 1
      for (int g = 0; g < 4; g++) {
        uint32_t r[32];
        tmem_ld32(t_addr + (uint32_t)(g * 32), r);
        if (g == 3) {                               // all TMEM reads of this warp are done: release the accumulator
          tc_fence_before();
          __syncwarp();
          if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&ctl->tmem_empty[acc])) : "memory");
        }
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] = fmaxf(fmaf(__uint_as_float(r[i]), sc, sh), lo);
        // flat pixel n = 84 * (double row) + ox; neighbouring lanes swap one value per pixel pair so that every lane
        // stores a bf16x2: even lanes channels (c, c+1) of pixel 2jj, odd lanes channels (c-1, c) of pixel 2jj+1
        uint32_t hh[16];
#pragma unroll
        for (int jj = 0; jj < 16; jj++) {
          const float mine_keep = odd ? v[2 * jj + 1] : v[2 * jj];
          const float other = __shfl_xor_sync(0xffffffffu, odd ? v[2 * jj] : v[2 * jj + 1], 1);
          const act2_t h2 = odd ? f2act2(other, mine_keep) : f2act2(mine_keep, other);
          hh[jj] = *reinterpret_cast<const uint32_t*>(&h2);
        }
        // The 32 pixels of a group cross at most one row end.  Pixel 2jj + odd of this lane sits at ox = oxm + 2jj: it is
        // stored through p0 while ox < 80 (jj < ja), skipped for the 4 junk columns, and stored through p1 (next double
        // row: + 160 - 84 pixels) from ox >= 84 on (jj >= jb).  Branch-free: two predicated stores with immediate offsets
        // (measured: per-pixel index arithmetic + divergent branches made this epilogue the bound of the kernel).
        const int n0 = half * 128 + g * 32;
        const int rl0 = n0 / ST_PAIRS, oxm = n0 - rl0 * ST_PAIRS + odd;
        const bool va = r0 + rl0 < ST_PROWS && !(q.dbg & 1);
        const bool vb = rl0 + 1 < ST_TROWS && r0 + rl0 + 1 < ST_PROWS && !(q.dbg & 1);
        const int ja = va ? (ST_OUT - oxm + 1) >> 1 : 0;                   // first jj with ox >= 80 (may exceed 16)
        const int jb = vb ? (ST_PAIRS - oxm + 1) >> 1 : 16;                // first jj with ox >= 84
        char* p0 = (char*)(obase + ((2 * (r0 + rl0) + par) * ST_OUT + oxm) * ST_CH);
        char* p1 = p0 + (2 * ST_OUT - ST_PAIRS) * ST_CH * 2;
#pragma unroll
        for (int jj = 0; jj < 16; jj++) {
          asm volatile(
              "{\n\t.reg .pred pa, pb;\n\t"
              "setp.gt.s32 pa, %3, %5;\n\t"
              "setp.le.s32 pb, %4, %5;\n\t"
              "@pa st.global.b32 [%0], %2;\n\t"
              "@pb st.global.b32 [%1], %2;\n\t}" ::"l"(p0 + jj * (2 * ST_CH * 2)), "l"(p1 + jj * (2 * ST_CH * 2)), "r"(hh[jj]), "r"(ja), "r"(jb), "r"(jj)
              : "memory");
        }
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

// stacked weights [9 ky'][128 rows][4 pairs][2 px][8 ch] from the packed stem weights [64][7][8 taps][8 ch]
// (tap j of the packed layout = kernel column j; here pixel 2t+h of the window is kernel column 2t+h-1)
template <typename T>
__global__ void stem_pack_kernel(const T* __restrict__ w, act_t* __restrict__ out)
{
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 9 * 128 * 64) return;
  const int kyp = idx / 8192, m = (idx >> 6) & 127, col = idx & 63;
  const int co = m & 63, ky = m < 64 ? kyp : kyp - 2;
  const int kx = (col >> 3) - 1, c = col & 7;
  float v = 0.0f;
  if (ky >= 0 && ky < 7 && kx >= 0 && kx < 7) v = (float)w[((co * 7 + ky) * 8 + kx) * 8 + c];
  out[idx] = f2act(v);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
int g_sms = 0;
std::mutex g_mu;
std::map<std::pair<const void*, int>, CUtensorMap> g_xmaps;

int stem_enabled()
{
  static int v = -1;
  if (v < 0) { const char* e = getenv("FPB_STEM"); v = e ? atoi(e) : 1; }
  return v;
}

}  // namespace

int stem_pack_weights(const void* w_packed, int is_bf16, void* w_stacked, cudaStream_t st)
{
  const int n = 9 * 128 * 64;
  if (is_bf16) stem_pack_kernel<act_t><<<(n + 255) / 256, 256, 0, st>>>((const act_t*)w_packed, (act_t*)w_stacked);
  else stem_pack_kernel<float><<<(n + 255) / 256, 256, 0, st>>>((const float*)w_packed, (act_t*)w_stacked);
  FPB_CUDA(cudaGetLastError());
  return FPB_OK;
}

// in: row-padded network input [nimg][160][168][8] bf16; w_stacked: stem_pack_weights output; out: [nimg*6400][64] bf16.
// Returns FPB_ERR_UNSUPPORTED when the kernel is switched off (FPB_STEM=0) so the caller uses the im2col path.
int stem_conv(const void* in, int nimg, const void* w_stacked, const float* scale, const float* shift, int relu, void* out,
              cudaStream_t st, int force)
{
  if (!force && !stem_enabled()) return FPB_ERR_UNSUPPORTED;
  if (nimg < 1 || (((uintptr_t)in | (uintptr_t)w_stacked | (uintptr_t)out) & 15)) return FPB_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_encode) {
    cudaDriverEntryPointQueryResult qr;
    void* f = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess || !f) {
      fpb_set_error("cuTensorMapEncodeTiled driver entry point is unavailable");
      return FPB_ERR_CUDA;
    }
    g_encode = (EncodeTiledFn)f;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
    FPB_CUDA(cudaFuncSetAttribute(stem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM));
  }
  CUtensorMap mX[2], mW;                                   // boxes of 4 and 5 plane rows
  for (int bi = 0; bi < 2; bi++) {
    auto key = std::make_pair(in, nimg * 2 + bi);
    auto it = g_xmaps.find(key);
    if (it != g_xmaps.end()) { mX[bi] = it->second; continue; }
    const cuuint64_t row_b = (cuuint64_t)ST_PITCH * ST_C * 2;                  // 2688 B per input row
    cuuint64_t dims[5] = {16, (cuuint64_t)ST_PAIRS, 4, (cuuint64_t)ST_PROWS, (cuuint64_t)nimg};
    cuuint64_t strides[4] = {32, row_b, 4 * row_b, (cuuint64_t)ST_IMG * row_b};
    cuuint32_t box[5] = {16, (cuuint32_t)ST_PAIRS, 1, (cuuint32_t)(4 + bi), 1};
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = g_encode(&mX[bi], ACT_TMAP_DTYPE, 5, (void*)in, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { fpb_set_error("stem_conv: cuTensorMapEncodeTiled (input) failed (%d)", (int)r); return FPB_ERR_CUDA; }
    if (g_xmaps.size() > 256) g_xmaps.clear();
    g_xmaps[key] = mX[bi];
  }
  int sms = 0;
  int rc = fpb_tiled_map_act(w_stacked, 9 * 128, 64, 64, 128, &mW, &sms);
  if (rc) return rc;
  StemParams q;
  q.scale = scale; q.shift = shift; q.out = (act_t*)out; q.nimg = nimg; q.relu = relu;
  { const char* e = getenv("FPB_STEM_DBG"); q.dbg = e ? atoi(e) : 0; }
  const int tiles = nimg * ST_TILES_PER_IMG;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(tiles < g_sms ? tiles : g_sms, 1, 1);
  cfg.blockDim = dim3(ST_NT, 1, 1);
  cfg.dynamicSmemBytes = ST_SMEM;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (fpb_pdl_enabled() && !force) ? 1 : 0;   // force = debug entry: the weights were packed by the previous kernel
  FPB_CUDA(cudaLaunchKernelEx(&cfg, stem_kernel, mX[0], mX[1], mW, q));
  fpb_count_launch();
  return FPB_OK;
}

}  // namespace FPB_ACT_NS
