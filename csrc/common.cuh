// Shared helpers for the b200rnn CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../include/b200rnn.h"

namespace b200rnn {

// Thread-local last-error string, exposed through b200rnn_last_error().
void set_error(const char* fmt, ...);

#define B200_CHECK_CUDA(expr)                                                        \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      ::b200rnn::set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr,        \
                           cudaGetErrorString(_e));                                  \
      return B200RNN_ECUDA;                                                          \
    }                                                                                \
  } while (0)

#define B200_REQUIRE(cond, code, ...)                                                \
  do {                                                                               \
    if (!(cond)) {                                                                   \
      ::b200rnn::set_error(__VA_ARGS__);                                             \
      return (code);                                                                 \
    }                                                                                \
  } while (0)

#define B200_CHECK_LAUNCH() B200_CHECK_CUDA(cudaGetLastError())

static inline int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// ---- device math: fp32-accurate activations on the fast exp path ----
// sigmoid(x) = 1/(1+e^-x); tanh(x) = 2*sigmoid(2x)-1.  ex2.approx is ~2 ulp,
// rcp.approx ~1 ulp: absolute error ~1e-7, far inside the 1e-5 test tolerance.
__device__ __forceinline__ float sigmoid_f(float x) {
  return __fdividef(1.0f, 1.0f + __expf(-x));
}
__device__ __forceinline__ float tanh_f(float x) {
  // clamp avoids inf/inf for huge |x|
  float xc = fminf(fmaxf(x, -15.0f), 15.0f);
  float e = __expf(2.0f * xc);
  return __fdividef(e - 1.0f, e + 1.0f);
}

// Packed fp32 FMA (FFMA2 in SASS, new with sm_100): acc.{x,y} += a.{x,y} * b.{x,y} in ONE issue slot.  A plain
// three-register FFMA issues every other cycle per SM sub-partition on Blackwell, so a CUDA-core GEMV only
// reaches the fp32 peak through the packed form; operands are natural register pairs (adjacent K elements).
__device__ __forceinline__ void ffma2(float2& acc, const float2& a, const float2& b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;"
      : "+l"(reinterpret_cast<unsigned long long&>(acc))
      : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace b200rnn
