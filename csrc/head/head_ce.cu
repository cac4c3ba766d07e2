// K5, standalone form — classifier head + mean cross-entropy + correct count, forward AND backward in two
// launches, for models whose recurrent layers do not run in the small-hidden fused kernels (which carry
// their own fused head, csrc/rnn/lstm_small.inc).  Replaces the reference's
//   output = fc(out[:, -1, :]);  loss = CrossEntropyLoss()(output, labels);  correct = (argmax == labels).sum()
// and their autograd backward (src/motion/model.py:16, src/motion/trainer/base.py:112-116): no cuBLAS GEMM,
// no ATen log_softmax / nll_loss / argmax kernels and no host sync.
//
//   kernel 1 (one warp per sample): logits (lane c keeps class c), log-sum-exp, loss, argmax, dlogits,
//             dh[b,:] = dlogits[b,:] x W — everything that is per sample;
//   kernel 2 (one thread per weight): dW[c,k] = sum_b dlogits[b,c] h[b,k], db, and the fixed-order sum of the
//             per-block loss / correct partials — everything that reduces over the batch, deterministically.
// Every block of kernel 2 counts itself on the reducer's retire counter (see csrc/runtime/reducer.cpp), so the
// head's gradient bucket can be all-reduced while the recurrent layers are still in their backward pass.
#include <math.h>
#include "../common.cuh"

namespace b200rnn {
namespace {

constexpr int kWarpsPerBlock = 4;

__global__ void __launch_bounds__(kWarpsPerBlock * 32) head_ce_sample_kernel(
    const float* __restrict__ h, int64_t ld_h, const float* __restrict__ W, const float* __restrict__ bias,
    const int64_t* __restrict__ labels, const int64_t* __restrict__ idx, int B, int H, int C, float loss_scale,
    float* __restrict__ logits_out, float* __restrict__ dlogits, float* __restrict__ dh, int64_t ld_dh,
    float* __restrict__ part /* [grid][2] */, int* __restrict__ status) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * kWarpsPerBlock + warp;
  __shared__ float dl_s[kWarpsPerBlock][32];
  float loss = 0.f, corr = 0.f;
  if (b < B) {
    const float* hb = h + (int64_t)b * ld_h;
    float mine = -INFINITY;   // lane c < C ends up with logit c
    for (int c = 0; c < C; ++c) {
      float acc = 0.f;
      for (int k = lane; k < H; k += 32) acc = fmaf(W[(int64_t)c * H + k], hb[k], acc);
      acc = warp_sum(acc) + bias[c];
      if (lane == c) mine = acc;
    }
    if (logits_out != nullptr && lane < C) logits_out[(int64_t)b * C + lane] = mine;
    // max / first argmax over the C lanes
    float m = mine; int arg = lane < C ? lane : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, m, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
    }
    const float e = lane < C ? expf(mine - m) : 0.f;
    const float lse = m + logf(warp_sum(e));
    if (labels != nullptr) {
      int y = (int)labels[idx != nullptr ? idx[b] : (int64_t)b];
      if ((unsigned)y >= (unsigned)C) { if (lane == 0) atomicOr(status, 1); y = 0; }
      const float ly = __shfl_sync(0xffffffffu, mine, y);
      loss = lse - ly;
      corr = (arg == y) ? 1.f : 0.f;
      const float dl = lane < C ? (expf(mine - lse) - (lane == y ? 1.f : 0.f)) * loss_scale : 0.f;
      if (dlogits != nullptr && lane < C) dlogits[(int64_t)b * C + lane] = dl;
      if (dh != nullptr) {
        // dlogits of this sample through shared memory: the k loop below is ragged across lanes (H need not be a
        // multiple of 32), so it must not contain warp-synchronous shuffles
        dl_s[warp][lane] = dl;
        __syncwarp();
        float* o = dh + (int64_t)b * ld_dh;
        for (int k = lane; k < H; k += 32) {
          float acc = 0.f;
          for (int c = 0; c < C; ++c) acc = fmaf(dl_s[warp][c], W[(int64_t)c * H + k], acc);
          o[k] = acc;
        }
      }
    }
  }
  __shared__ float red[2][kWarpsPerBlock];
  if (lane == 0) { red[0][warp] = loss; red[1][warp] = corr; }
  __syncthreads();
  if (threadIdx.x == 0 && part != nullptr) {
    float l = 0.f, c = 0.f;
#pragma unroll
    for (int w = 0; w < kWarpsPerBlock; ++w) { l += red[0][w]; c += red[1][w]; }
    part[2 * blockIdx.x] = l; part[2 * blockIdx.x + 1] = c;
  }
}

__global__ void __launch_bounds__(256) head_ce_wgrad_kernel(
    const float* __restrict__ h, int64_t ld_h, const float* __restrict__ dlogits, int B, int H, int C,
    float* __restrict__ dW, float* __restrict__ db, const float* __restrict__ part, int n_part,
    float* __restrict__ metrics, unsigned int* __restrict__ retire) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (int64_t)C * H) {
    const int c = (int)(i / H), k = (int)(i % H);
    float acc = 0.f;
    for (int b = 0; b < B; ++b) acc = fmaf(dlogits[(int64_t)b * C + c], h[(int64_t)b * ld_h + k], acc);
    dW[i] = acc;
  }
  if (blockIdx.x == 0) {
    if (threadIdx.x < C) {
      float acc = 0.f;
      for (int b = 0; b < B; ++b) acc += dlogits[(int64_t)b * C + threadIdx.x];
      db[threadIdx.x] = acc;
    }
    if (threadIdx.x == 32 && metrics != nullptr) {
      float l = 0.f, c = 0.f;
      for (int p = 0; p < n_part; ++p) { l += part[2 * p]; c += part[2 * p + 1]; }
      metrics[0] = l; metrics[1] = c;
    }
  }
  if (retire != nullptr) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(retire) : "memory");
  }
}

__device__ int g_head_status = 0;

}  // namespace
}  // namespace b200rnn

using namespace b200rnn;

extern "C" int64_t b200rnn_head_ce_scratch_floats(int32_t batch, int32_t classes) {
  const int64_t blocks = (batch + kWarpsPerBlock - 1) / kWarpsPerBlock;
  return (int64_t)batch * classes + 2 * blocks;
}

extern "C" int b200rnn_head_ce(const float* h_last, int64_t ld_h, const float* fc_w, const float* fc_b,
                               const int64_t* labels, const int64_t* idx, int32_t batch, int32_t hidden,
                               int32_t classes, float loss_scale, float* d_h, int64_t ld_dh, float* d_w,
                               float* d_b, float* metrics, float* scratch, uint32_t* retire_counter,
                               int32_t* retire_count, void* stream) {
  B200_REQUIRE(h_last && fc_w && fc_b && labels && d_h && d_w && d_b && scratch && batch >= 1 && hidden >= 1,
               B200RNN_EINVAL, "head_ce: bad argument");
  B200_REQUIRE(classes >= 1 && classes <= 32, B200RNN_ETOOBIG, "head_ce: classes=%d outside [1, 32]", classes);
  int* status = nullptr;
  B200_CHECK_CUDA(cudaGetSymbolAddress((void**)&status, g_head_status));
  const int blocks1 = (batch + kWarpsPerBlock - 1) / kWarpsPerBlock;
  float* dlogits = scratch;
  float* part = scratch + (int64_t)batch * classes;
  cudaStream_t s = (cudaStream_t)stream;
  head_ce_sample_kernel<<<blocks1, kWarpsPerBlock * 32, 0, s>>>(h_last, ld_h, fc_w, fc_b, labels, idx, batch, hidden, classes,
                                                                loss_scale, nullptr, dlogits, d_h, ld_dh, part, status);
  B200_CHECK_LAUNCH();
  const int blocks2 = (int)(((int64_t)classes * hidden + 255) / 256);
  head_ce_wgrad_kernel<<<blocks2, 256, 0, s>>>(h_last, ld_h, dlogits, batch, hidden, classes, d_w, d_b, part, blocks1,
                                               metrics, retire_counter);
  B200_CHECK_LAUNCH();
  if (retire_count) *retire_count = blocks2;
  return B200RNN_OK;
}

extern "C" int b200rnn_head_status(int32_t* status) {
  B200_REQUIRE(status != nullptr, B200RNN_EINVAL, "head_status: null");
  int v = 0;
  B200_CHECK_CUDA(cudaMemcpyFromSymbol(&v, g_head_status, sizeof(int)));
  *status = v;
  return B200RNN_OK;
}
