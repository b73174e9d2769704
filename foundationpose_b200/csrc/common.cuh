// Shared helpers for libfpb200.so translation units.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/fpb200.h"

#define FPB_EXPORT extern "C" __attribute__((visibility("default")))

void fpb_set_error(const char* fmt, ...);
void fpb_count_launch(int n = 1);

#define FPB_CHECK_ARG(cond)                                                        \
  do {                                                                             \
    if (!(cond)) {                                                                 \
      fpb_set_error("%s:%d: bad argument: %s", __FILE__, __LINE__, #cond);         \
      return FPB_ERR_ARG;                                                          \
    }                                                                              \
  } while (0)

#define FPB_CUDA(call)                                                             \
  do {                                                                             \
    cudaError_t e__ = (call);                                                      \
    if (e__ != cudaSuccess) {                                                      \
      fpb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      return FPB_ERR_CUDA;                                                         \
    }                                                                              \
  } while (0)

#define FPB_LAUNCH_CHECK()                                                         \
  do {                                                                             \
    fpb_count_launch();                                                            \
    FPB_CUDA(cudaGetLastError());                                                  \
  } while (0)

static inline int fpb_cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }
