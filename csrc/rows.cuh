// Block-level, fixed-order reduction of the backward kernels' per-CTA gradient rows — the first phase of the
// fused update (core.cu) and of the in-kernel all-reduce (comm/collectives.cu), so that no separate
// reduce_rows launch sits between the backward pass and the optimizer (the c10d Reducer's flatten + the
// per-tensor Adam loop in the reference: src/motion/trainer/ddp.py:19, trainer/base.py:117).
//
// A block of kRowThreads = kRG x kC4 threads owns kC4 float4 columns (kRowElems consecutive elements).
// Thread (rg, c) adds rows rg, rg + kRG, rg + 2 kRG, ... of column c with 128-bit loads, four in flight;
// the kRG partial sums of a column are then added in the order rg = 0 .. kRG-1.  The order depends only on
// n_rows, so every rank computes bit-identical sums (README.md:9 "equal final parameters").
#pragma once
#include "common.cuh"

namespace b200rnn {

constexpr int kRG = 8;                       // row groups
constexpr int kC4 = 32;                      // float4 columns per block
constexpr int kRowThreads = kRG * kC4;       // 256
constexpr int kRowElems = 4 * kC4;           // 128 elements per block

__device__ __forceinline__ float4 f4_add(float4 a, const float4& b) {
  a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
  return a;
}

// Elements [e0, e0 + 4) of row r; elements >= n_valid read as zero.  VEC: rows 16-byte aligned at every e0.
template <bool VEC>
__device__ __forceinline__ float4 load_row4(const float* __restrict__ rows, long long r, long long row_stride,
                                            long long e0, long long n_valid) {
  const float* p = rows + r * row_stride + e0;
  if (VEC) {
    if (e0 + 3 < n_valid) return __ldcg(reinterpret_cast<const float4*>(p));
  }
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (e0 + 0 < n_valid) v.x = __ldcg(p + 0);
  if (e0 + 1 < n_valid) v.y = __ldcg(p + 1);
  if (e0 + 2 < n_valid) v.z = __ldcg(p + 2);
  if (e0 + 3 < n_valid) v.w = __ldcg(p + 3);
  return v;
}

// Sum over rows of this block's columns.  sm: kRowThreads float4.  On return (after the trailing
// __syncthreads) sm[c], c < kC4, holds the full sum of column c for every thread to read.
template <bool VEC>
__device__ __forceinline__ void block_reduce_rows(const float* __restrict__ rows, int n_rows, long long row_stride,
                                                  long long block_e0, long long n_valid, float4* sm) {
  const int c = threadIdx.x % kC4, rg = threadIdx.x / kC4;
  const long long e0 = block_e0 + 4 * c;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (e0 < n_valid) {
    int r = rg;
    for (; r + 3 * kRG < n_rows; r += 4 * kRG) {
      const float4 v0 = load_row4<VEC>(rows, r, row_stride, e0, n_valid);
      const float4 v1 = load_row4<VEC>(rows, r + kRG, row_stride, e0, n_valid);
      const float4 v2 = load_row4<VEC>(rows, r + 2 * kRG, row_stride, e0, n_valid);
      const float4 v3 = load_row4<VEC>(rows, r + 3 * kRG, row_stride, e0, n_valid);
      acc = f4_add(f4_add(f4_add(f4_add(acc, v0), v1), v2), v3);
    }
    for (; r < n_rows; r += kRG) acc = f4_add(acc, load_row4<VEC>(rows, r, row_stride, e0, n_valid));
  }
  sm[rg * kC4 + c] = acc;
  __syncthreads();
  if (rg == 0) {
#pragma unroll
    for (int g = 1; g < kRG; ++g) acc = f4_add(acc, sm[g * kC4 + c]);
    sm[c] = acc;
  }
  __syncthreads();
}

static inline bool rows_vectorizable(const void* rows, long long row_stride) {
  return (reinterpret_cast<uintptr_t>(rows) & 15) == 0 && (row_stride & 3) == 0;
}

}  // namespace b200rnn
