// Activation / weight element type of the tcgen05 network kernels.  Every network translation unit is compiled twice
// (foundationpose_b200/build.py): as-is with bf16 operands (namespace bf16) and with -DFPB_ACT_FP16 with IEEE fp16
// operands (namespace f16) -- the dtype the reference's autocast path runs in (predict_pose_refine.py:190,
// predict_score.py:193).  Both are 16-bit, so layouts, TMA boxes, swizzles and the tcgen05.mma kind::f16 instruction
// are identical; only the A/B format field of the instruction descriptor and the conversions differ.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#ifdef FPB_ACT_FP16
#define FPB_ACT_NS f16
typedef __half act_t;
typedef __half2 act2_t;
#define ACT_IDESC_FMT 0u                                   /* kind::f16 descriptor: A / B format F16 */
#define ACT_TMAP_DTYPE CU_TENSOR_MAP_DATA_TYPE_FLOAT16
__device__ __forceinline__ act2_t f2act2(float a, float b) { return __floats2half2_rn(a, b); }
__device__ __forceinline__ act_t f2act(float a) { return __float2half_rn(a); }
__device__ __forceinline__ float act2f(act_t a) { return __half2float(a); }
__device__ __forceinline__ float2 act22f2(act2_t a) { return __half22float2(a); }
__device__ __forceinline__ float actbits2f(unsigned short u) { return __half2float(__ushort_as_half(u)); }   /* raw 16-bit pattern -> float */
#define ACT_ONE_X2 0x3C003C00u                             /* two packed 1.0 */
#else
#define FPB_ACT_NS bf16
typedef __nv_bfloat16 act_t;
typedef __nv_bfloat162 act2_t;
#define ACT_IDESC_FMT 1u                                   /* kind::f16 descriptor: A / B format BF16 */
#define ACT_TMAP_DTYPE CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
__device__ __forceinline__ act2_t f2act2(float a, float b) { return __floats2bfloat162_rn(a, b); }
__device__ __forceinline__ act_t f2act(float a) { return __float2bfloat16(a); }
__device__ __forceinline__ float act2f(act_t a) { return __bfloat162float(a); }
__device__ __forceinline__ float2 act22f2(act2_t a) { return __bfloat1622float2(a); }
__device__ __forceinline__ float actbits2f(unsigned short u) { return __uint_as_float((uint32_t)u << 16); }
#define ACT_ONE_X2 0x3F803F80u
#endif
