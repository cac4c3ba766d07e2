// Error plumbing + the small memory-bound kernels: K11 dataset gather, K5 head
// (standalone form), deterministic row reduction, K6 fused optimizer, L2 flush.
#include "common.cuh"
#include "optim.cuh"
#include "rows.cuh"
#include <math.h>

namespace b200rnn {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

namespace {

// ---- K11 gather: one warp per row segment, 128-bit copies when aligned ----
__global__ void gather_kernel(const float* __restrict__ feats, const int64_t* __restrict__ labels,
                              const int64_t* __restrict__ idx, int batch, int row_elems,
                              float* __restrict__ x_out, int64_t* __restrict__ y_out) {
  // grid-stride over rows: a capped grid (b200rnn_gather_batch_capped) keeps a gather that streams from pinned
  // host memory on a side stream from occupying the SMs the training kernels run on
  for (int b = blockIdx.x; b < batch; b += gridDim.x) {
    const int64_t src = idx[b];
    const float* s = feats + src * (int64_t)row_elems;
    float* d = x_out + (int64_t)b * row_elems;
    if ((row_elems & 3) == 0) {
      const float4* s4 = reinterpret_cast<const float4*>(s);
      float4* d4 = reinterpret_cast<float4*>(d);
      const int n4 = row_elems / 4;
      int i = threadIdx.x;
      for (; i + 3 * (int)blockDim.x < n4; i += 4 * blockDim.x) {   // four independent 16-byte loads in flight
        const float4 v0 = __ldg(&s4[i]), v1 = __ldg(&s4[i + blockDim.x]);
        const float4 v2 = __ldg(&s4[i + 2 * blockDim.x]), v3 = __ldg(&s4[i + 3 * blockDim.x]);
        d4[i] = v0; d4[i + blockDim.x] = v1; d4[i + 2 * blockDim.x] = v2; d4[i + 3 * blockDim.x] = v3;
      }
      for (; i < n4; i += blockDim.x) d4[i] = __ldg(&s4[i]);
    } else {
      for (int i = threadIdx.x; i < row_elems; i += blockDim.x) d[i] = __ldg(&s[i]);
    }
    if (threadIdx.x == 0 && labels != nullptr && y_out != nullptr) y_out[b] = labels[src];
  }
}

// K11, time-major form: x_out[t][b][:] = feats[idx[b]][t][:] — the large-hidden path reads its input time-major
// (one GEMM over all timesteps), so the gather and the [B,T,I] -> [T,B,I] transpose are one pass.
__global__ void gather_tm_kernel(const float* __restrict__ feats, const int64_t* __restrict__ idx, int batch, int T, int I,
                                 float* __restrict__ x_out) {
  const int b = blockIdx.x;
  const int64_t src = idx != nullptr ? idx[b] : (int64_t)b;
  const float* s = feats + src * (int64_t)T * I;
  for (int e = threadIdx.x; e < T * I; e += blockDim.x) {
    const int t = e / I, i = e - t * I;
    x_out[((int64_t)t * batch + b) * I + i] = __ldg(&s[e]);
  }
}

// ---- deterministic row reduction ----
__global__ void reduce_rows_kernel(const float* __restrict__ rows, int n_rows, int64_t row_floats,
                                   float* __restrict__ out, int accumulate) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= row_floats) return;
  float acc = 0.0f;
  for (int r = 0; r < n_rows; ++r) acc += rows[(int64_t)r * row_floats + i];
  out[i] = accumulate ? out[i] + acc : acc;
}

// ---- K6 fused optimizer (W = 1 form; the multi-GPU form lives in comm/) ----
// One block = 128 consecutive elements: fixed-order sum of the gradient rows (rows.cuh), then the
// optimizer on 4 elements per thread of the first warp.
template <bool VEC>
__global__ void __launch_bounds__(kRowThreads) fused_update_kernel(
    const float* __restrict__ rows, int n_rows, int64_t n_params, float* __restrict__ params,
    float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq, OptScalars o,
    float* __restrict__ grad_out, float* __restrict__ metrics) {
  __shared__ float4 sm[kRowThreads];
  const int64_t row_floats = n_params + 2;
  const int64_t block_e0 = (int64_t)blockIdx.x * kRowElems;
  block_reduce_rows<VEC>(rows, n_rows, row_floats, block_e0, row_floats, sm);
  if (threadIdx.x >= kC4) return;
  const int64_t e0 = block_e0 + 4 * threadIdx.x;
  const float4 s4 = sm[threadIdx.x];
  const float acc[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int64_t i = e0 + e;
    if (i >= row_floats) break;
    if (i >= n_params) {
      if (metrics != nullptr) metrics[i - n_params] = acc[e];
      continue;
    }
    const float g = acc[e] * o.grad_scale;
    if (grad_out != nullptr) grad_out[i] = g;
    float p = params[i];
    float m = exp_avg ? exp_avg[i] : 0.0f;
    float v = exp_avg_sq ? exp_avg_sq[i] : 0.0f;
    apply_update(o, g, p, m, v);
    params[i] = p;
    if (exp_avg) exp_avg[i] = m;
    if (exp_avg_sq) exp_avg_sq[i] = v;
  }
}

// Single gradient row (e.g. one bucket of an arena whose gradient came from GEMMs): plain streaming update, no
// shared memory, so it co-resides with any compute kernel when it runs on a side stream.
__global__ void __launch_bounds__(256) update_vec_kernel(
    const float* __restrict__ g_in, int64_t n, float* __restrict__ params, float* __restrict__ exp_avg,
    float* __restrict__ exp_avg_sq, OptScalars o, float* __restrict__ grad_out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float g = g_in[i] * o.grad_scale;
    if (grad_out != nullptr) grad_out[i] = g;
    float p = params[i];
    float m = exp_avg ? exp_avg[i] : 0.0f;
    float v = exp_avg_sq ? exp_avg_sq[i] : 0.0f;
    apply_update(o, g, p, m, v);
    params[i] = p;
    if (exp_avg) exp_avg[i] = m;
    if (exp_avg_sq) exp_avg_sq[i] = v;
  }
}

// ---- K5 head, standalone: one warp per sample ----
__global__ void head_fwd_kernel(const float* __restrict__ h_last, int64_t ld_h,
                                const float* __restrict__ fc_w, const float* __restrict__ fc_b,
                                const int64_t* __restrict__ labels, int batch, int hidden, int classes,
                                float* __restrict__ logits, float* __restrict__ part /* [grid][2] */) {
  extern __shared__ float sm[];  // per-warp logits [warps][classes]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int warps = blockDim.x >> 5;
  float* lg = sm + warp * classes;
  float loss_acc = 0.f, corr_acc = 0.f;
  for (int b = blockIdx.x * warps + warp; b < batch; b += gridDim.x * warps) {
    const float* h = h_last + (int64_t)b * ld_h;
    for (int c = 0; c < classes; ++c) {
      float acc = 0.f;
      for (int k = lane; k < hidden; k += 32) acc = fmaf(fc_w[(int64_t)c * hidden + k], h[k], acc);
      acc = warp_sum(acc);
      if (lane == 0) {
        acc += fc_b[c];
        lg[c] = acc;
        if (logits) logits[(int64_t)b * classes + c] = acc;
      }
    }
    __syncwarp();
    if (lane == 0 && labels != nullptr) {
      float m = lg[0]; int arg = 0;
      for (int c = 1; c < classes; ++c) if (lg[c] > m) { m = lg[c]; arg = c; }
      float se = 0.f;
      for (int c = 0; c < classes; ++c) se += expf(lg[c] - m);
      const int y = (int)labels[b];
      loss_acc += (m + logf(se)) - lg[y];
      corr_acc += (arg == y) ? 1.f : 0.f;
    }
    __syncwarp();
  }
  // block reduce (lane 0 of each warp holds values)
  __shared__ float red[2][32];
  if (lane == 0) { red[0][warp] = loss_acc; red[1][warp] = corr_acc; }
  __syncthreads();
  if (threadIdx.x == 0 && part != nullptr) {
    float l = 0.f, c = 0.f;
    for (int w = 0; w < warps; ++w) { l += red[0][w]; c += red[1][w]; }
    part[2 * blockIdx.x] = l;
    part[2 * blockIdx.x + 1] = c;
  }
}

__global__ void head_finish_kernel(const float* __restrict__ part, int n, float* __restrict__ metrics) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    float l = 0.f, c = 0.f;
    for (int i = 0; i < n; ++i) { l += part[2 * i]; c += part[2 * i + 1]; }
    metrics[0] = l; metrics[1] = c;
  }
}

__global__ void l2_flush_kernel(float4* __restrict__ p, int64_t n4, float val) {
  const float4 v = make_float4(val, val, val, val);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}

}  // namespace
}  // namespace b200rnn

using namespace b200rnn;

extern "C" const char* b200rnn_last_error(void) { return g_err; }
extern "C" int b200rnn_version(void) { return 100; }

extern "C" int b200rnn_gather_batch_capped(const float* feats, const int64_t* labels, const int64_t* idx,
                                           int32_t batch, int32_t row_elems, float* x_out, int64_t* y_out,
                                           int32_t max_blocks, void* stream) {
  B200_REQUIRE(feats && idx && x_out && batch >= 0 && row_elems > 0, B200RNN_EINVAL, "gather_batch: bad argument");
  if (batch == 0) return B200RNN_OK;
  const int grid = (max_blocks > 0 && max_blocks < batch) ? max_blocks : batch;
  gather_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(feats, labels, idx, batch, row_elems, x_out, y_out);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_gather_batch(const float* feats, const int64_t* labels, const int64_t* idx,
                                    int32_t batch, int32_t row_elems, float* x_out, int64_t* y_out,
                                    void* stream) {
  return b200rnn_gather_batch_capped(feats, labels, idx, batch, row_elems, x_out, y_out, 0, stream);
}

extern "C" int b200rnn_gather_batch_tm(const float* feats, const int64_t* idx, int32_t batch, int32_t seq_len,
                                       int32_t features, float* x_tm_out, void* stream) {
  B200_REQUIRE(feats && x_tm_out && batch >= 0 && seq_len > 0 && features > 0, B200RNN_EINVAL, "gather_batch_tm: bad argument");
  if (batch == 0) return B200RNN_OK;
  gather_tm_kernel<<<batch, 256, 0, (cudaStream_t)stream>>>(feats, idx, batch, seq_len, features, x_tm_out);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_reduce_rows(const float* rows, int32_t n_rows, int64_t row_floats, float* out,
                                   int32_t accumulate, void* stream) {
  B200_REQUIRE(rows && out && n_rows >= 0 && row_floats > 0, B200RNN_EINVAL, "reduce_rows: bad argument");
  const int threads = 256;
  const int64_t blocks = (row_floats + threads - 1) / threads;
  reduce_rows_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(rows, n_rows, row_floats, out, accumulate);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_fused_update(const float* rows, int32_t n_rows, int64_t n_params, float* params,
                                    float* exp_avg, float* exp_avg_sq, const b200rnn_optim_desc* opt,
                                    float* grad_out, float* metrics, void* stream) {
  B200_REQUIRE(rows && params && opt && n_rows >= 1 && n_params > 0, B200RNN_EINVAL, "fused_update: bad argument");
  B200_REQUIRE(opt->kind == 0 || opt->kind == 1, B200RNN_EINVAL, "fused_update: unknown optimizer kind %d", opt->kind);
  B200_REQUIRE(opt->kind != 0 || (exp_avg && exp_avg_sq && opt->step >= 1), B200RNN_EINVAL,
               "fused_update: Adam needs exp_avg/exp_avg_sq and step >= 1");
  B200_REQUIRE(!(opt->kind == 1 && opt->momentum != 0.0f && !exp_avg), B200RNN_EINVAL,
               "fused_update: SGD momentum needs a momentum buffer (exp_avg)");
  const OptScalars o = make_opt_scalars(opt);
  if (n_rows == 1 && metrics == nullptr) {
    int64_t vb = (n_params + 255) / 256;
    if (vb > 8 * num_sms()) vb = 8 * num_sms();
    update_vec_kernel<<<(unsigned)vb, 256, 0, (cudaStream_t)stream>>>(rows, n_params, params, exp_avg, exp_avg_sq, o, grad_out);
    B200_CHECK_LAUNCH();
    return B200RNN_OK;
  }
  const int64_t blocks = (n_params + 2 + kRowElems - 1) / kRowElems;
  if (rows_vectorizable(rows, n_params + 2))
    fused_update_kernel<true><<<(unsigned)blocks, kRowThreads, 0, (cudaStream_t)stream>>>(
        rows, n_rows, n_params, params, exp_avg, exp_avg_sq, o, grad_out, metrics);
  else
    fused_update_kernel<false><<<(unsigned)blocks, kRowThreads, 0, (cudaStream_t)stream>>>(
        rows, n_rows, n_params, params, exp_avg, exp_avg_sq, o, grad_out, metrics);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

namespace {
constexpr int kHeadFwdThreads = 128, kHeadFwdMaxGrid = 64;
inline int head_fwd_grid(int32_t batch) {
  const int warps = kHeadFwdThreads / 32;
  const int grid = (batch + warps - 1) / warps;
  return grid > kHeadFwdMaxGrid ? kHeadFwdMaxGrid : (grid < 1 ? 1 : grid);
}
}  // namespace

extern "C" int64_t b200rnn_head_fwd_metrics_floats(int32_t batch) { return 2 * ((int64_t)head_fwd_grid(batch) + 1); }

extern "C" int b200rnn_head_fwd(const float* h_last, int64_t ld_h, const float* fc_w, const float* fc_b,
                                const int64_t* labels, int32_t batch, int32_t hidden, int32_t classes,
                                float* logits, float* metrics, void* stream) {
  B200_REQUIRE(h_last && fc_w && fc_b && batch >= 1 && hidden >= 1 && classes >= 1, B200RNN_EINVAL,
               "head_fwd: bad argument");
  B200_REQUIRE(classes <= 1024, B200RNN_ETOOBIG, "head_fwd: classes=%d exceeds 1024", classes);
  // metrics doubles as the partial buffer: b200rnn_head_fwd_metrics_floats(batch) floats when labels are given.
  const int threads = kHeadFwdThreads, warps = threads / 32;
  const int grid = head_fwd_grid(batch);
  float* part = (labels && metrics) ? metrics + 2 : nullptr;
  head_fwd_kernel<<<grid, threads, warps * classes * sizeof(float), (cudaStream_t)stream>>>(
      h_last, ld_h, fc_w, fc_b, labels, batch, hidden, classes, logits, part);
  B200_CHECK_LAUNCH();
  if (part) {
    head_finish_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(part, grid, metrics);
    B200_CHECK_LAUNCH();
  }
  return B200RNN_OK;
}

extern "C" int b200rnn_l2_flush(void* scratch, int64_t bytes, void* stream) {
  B200_REQUIRE(scratch && bytes >= 16, B200RNN_EINVAL, "l2_flush: bad argument");
  static unsigned counter = 0;
  l2_flush_kernel<<<num_sms() * 4, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<float4*>(scratch), bytes / 16, (float)(++counter));
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

// Non-zero once a bounded wait of the layer-wavefront kernels (csrc/rnn/lstm_small.inc) timed out.
extern "C" int b200rnn_wave_status_lstm_fwd(void);
extern "C" int b200rnn_wave_status_lstm_bwd(void);
extern "C" int b200rnn_wave_status_lstm_bwd_dx(void);
extern "C" int b200rnn_wave_status_gru_fwd(void);
extern "C" int b200rnn_wave_status_gru_bwd(void);
extern "C" int b200rnn_wave_status_gru_bwd_dx(void);
extern "C" int b200rnn_wave_status_lstm_fwd_h16(void);
extern "C" int b200rnn_wave_status_lstm_bwd_h16(void);
extern "C" int b200rnn_wave_status_lstm_bwd_dx_h16(void);
extern "C" int b200rnn_wave_status_gru_fwd_h16(void);
extern "C" int b200rnn_wave_status_gru_bwd_h16(void);
extern "C" int b200rnn_wave_status_gru_bwd_dx_h16(void);
extern "C" int b200rnn_wave_status(int32_t* status) {
  B200_REQUIRE(status != nullptr, B200RNN_EINVAL, "wave_status: null");
  *status = b200rnn_wave_status_lstm_fwd() | b200rnn_wave_status_lstm_bwd() | b200rnn_wave_status_lstm_bwd_dx() |
            b200rnn_wave_status_gru_fwd() | b200rnn_wave_status_gru_bwd() | b200rnn_wave_status_gru_bwd_dx() |
            b200rnn_wave_status_lstm_fwd_h16() | b200rnn_wave_status_lstm_bwd_h16() | b200rnn_wave_status_lstm_bwd_dx_h16() |
            b200rnn_wave_status_gru_fwd_h16() | b200rnn_wave_status_gru_bwd_h16() | b200rnn_wave_status_gru_bwd_dx_h16();
  return B200RNN_OK;
}
