// Large-hidden LSTM / GRU path (H >= 64, any input width): GEMM-shaped work on the tcgen05 GEMM
// (csrc/gemm/gemm_tc.cu), everything else fused into one pointwise kernel per timestep.
//
//   forward, per layer:   gx = X[T*B, in] x W_ih^T + (b_ih [+ b_hh])         one GEMM over ALL timesteps (K1)
//                         for t: g_t = gx_t (+)= h_{t-1} x W_hh^T            tcgen05 GEMM, accumulate into gx_t
//                                cell kernel: gates, c_t, h_t (fp32 + bf16)   fused pointwise
//   backward, per layer:  for t (reverse): cell-backward kernel -> dG_t; dh_{t-1} = dG_t x W_hh
//                         dX = dG x W_ih; dW_ih = dG^T x X; dW_hh = dG^T x H_prev; db = colsum(dG)   (K4)
//
// bf16 operands, fp32 accumulation / cell state (BASELINE.json config 4: "2-layer LSTM hidden=1024,
// seq_len=512, bf16").  The whole time loop is driven from this file (one C call per direction).
// Where the persistent kernels of rnn_persist.cu cover the shape (LSTM / GRU, H % 64 == 0, H / 8 <= #SMs,
// B <= 128) the time loop of a layer is ONE cooperative launch with W_hh sharded across the SMs'
// shared memory; otherwise it is the per-timestep formulation below (a GEMM + a cell kernel per step,
// weights re-read from L2 each step).
// Reference call sites: src/motion/model.py:9-17 (nn.LSTM fwd), trainer/base.py:116 (autograd bwd).
#include <cuda_bf16.h>
#include "../common.cuh"

extern "C" int b200rnn_gemm_bf16(const void*, const void*, void*, const float*, int32_t, int32_t, int32_t, int32_t,
                                 int32_t, int32_t, int32_t, int32_t, void*);
extern "C" int b200rnn_gemm_bf16_retire(const void*, const void*, void*, const float*, int32_t, int32_t, int32_t, int32_t,
                                        int32_t, int32_t, int32_t, int32_t, uint32_t*, int32_t*, void*);
extern "C" int b200rnn_transpose_cast_bf16(const float*, void*, int32_t, int32_t, int64_t, int64_t, void*);
// persistent time loops (rnn_persist.cu)
extern "C" int b200rnn_rnn_persist_supported(const b200rnn_model_desc*);
extern "C" int64_t b200rnn_rnn_persist_stream_bytes(int32_t);
extern "C" int b200rnn_lstm_persist_fwd(int32_t, int32_t, int32_t, float*, const void*, float*, float*, void*, void*, void*);
extern "C" int b200rnn_lstm_persist_bwd(int32_t, int32_t, int32_t, const float*, const float*, const float*, const float*,
                                        const float*, const void*, float*, void*, void*, void*, void*);
extern "C" int b200rnn_gru_persist_fwd(int32_t, int32_t, int32_t, float*, float*, const void*, const float*, float*, void*, void*, void*);
extern "C" int b200rnn_gru_persist_bwd(int32_t, int32_t, int32_t, const float*, const float*, const float*, const float*,
                                       const float*, const void*, float*, void*, float*, void*, void*, void*, void*);

namespace b200rnn {
namespace {

typedef __nv_bfloat16 bf16;

// ---- pointwise cell kernels: one thread per (b, u) ----
// LSTM forward.  g: [B,4H] pre-activations (i|f|g|o blocks) -> overwritten with activations.
__global__ void lstm_cell_fwd(float* __restrict__ g, const float* __restrict__ c_prev, float* __restrict__ c_out,
                              float* __restrict__ h_out, bf16* __restrict__ h_b, int B, int H, int ldhb) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int b = idx / H, u = idx % H;
  float* gr = g + (int64_t)b * 4 * H;
  const float ig = sigmoid_f(gr[u]), fg = sigmoid_f(gr[H + u]), gg = tanh_f(gr[2 * H + u]), og = sigmoid_f(gr[3 * H + u]);
  const float c = fmaf(fg, c_prev ? c_prev[idx] : 0.0f, ig * gg);
  const float h = og * tanh_f(c);
  gr[u] = ig; gr[H + u] = fg; gr[2 * H + u] = gg; gr[3 * H + u] = og;
  c_out[idx] = c; h_out[idx] = h;
  h_b[(int64_t)b * ldhb + u] = __float2bfloat16(h);
}
// LSTM backward.  dh_ext: gradient from above at this step (nullable), dh_rec: recurrent gradient
// (nullable), dc: carry (in/out).  Writes dG fp32 [B,4H] and its bf16 copy.
__global__ void lstm_cell_bwd(const float* __restrict__ g, const float* __restrict__ c, const float* __restrict__ c_prev,
                              const float* __restrict__ dh_ext, int64_t dhe_sb, const float* __restrict__ dh_rec,
                              float* __restrict__ dc, float* __restrict__ dG, bf16* __restrict__ dG_b, int B, int H) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int b = idx / H, u = idx % H;
  const float* gr = g + (int64_t)b * 4 * H;
  const float ig = gr[u], fg = gr[H + u], gg = gr[2 * H + u], og = gr[3 * H + u];
  float dh = 0.0f;
  if (dh_ext) dh += dh_ext[(int64_t)b * dhe_sb + u];
  if (dh_rec) dh += dh_rec[idx];
  const float tc = tanh_f(c[idx]);
  const float dct = fmaf(dh * og, 1.0f - tc * tc, dc[idx]);
  const float cp = c_prev ? c_prev[idx] : 0.0f;
  const float d0 = dct * gg * ig * (1.0f - ig), d1 = dct * cp * fg * (1.0f - fg);
  const float d2 = dct * ig * (1.0f - gg * gg), d3 = dh * tc * og * (1.0f - og);
  dc[idx] = dct * fg;
  float* o = dG + (int64_t)b * 4 * H;
  bf16* ob = dG_b + (int64_t)b * 4 * H;
  o[u] = d0; o[H + u] = d1; o[2 * H + u] = d2; o[3 * H + u] = d3;
  ob[u] = __float2bfloat16(d0); ob[H + u] = __float2bfloat16(d1);
  ob[2 * H + u] = __float2bfloat16(d2); ob[3 * H + u] = __float2bfloat16(d3);
}
// GRU forward (torch semantics: n = tanh(x_n + b_in + r * (h W_hn^T + b_hn)); h' = (1-z) n + z h).
// gx: [B,3H] = x-projection + b_ih; gh: [B,3H] = h-projection + b_hh.  Stashes r, z, n in gx and hn in gh.
__global__ void gru_cell_fwd(float* __restrict__ gx, float* __restrict__ gh, const float* __restrict__ h_prev,
                             float* __restrict__ h_out, bf16* __restrict__ h_b, int B, int H, int ldhb) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int b = idx / H, u = idx % H;
  float* x = gx + (int64_t)b * 3 * H;
  float* hh = gh + (int64_t)b * 3 * H;
  const float r = sigmoid_f(x[u] + hh[u]), z = sigmoid_f(x[H + u] + hh[H + u]);
  const float hn = hh[2 * H + u];
  const float n = tanh_f(x[2 * H + u] + r * hn);
  const float hp = h_prev ? h_prev[idx] : 0.0f;
  const float h = (1.0f - z) * n + z * hp;
  x[u] = r; x[H + u] = z; x[2 * H + u] = n;
  h_out[idx] = h;
  h_b[(int64_t)b * ldhb + u] = __float2bfloat16(h);
}
// GRU backward: produces dGx (grad wrt x-projection pre-activations) and dGh (wrt h-projection) and
// the direct part of dh_prev (z * dh); the GEMM part (dGh x W_hh) is added by the caller.
__global__ void gru_cell_bwd(const float* __restrict__ gx, const float* __restrict__ gh, const float* __restrict__ h_prev,
                             const float* __restrict__ dh_ext, int64_t dhe_sb, const float* __restrict__ dh_rec,
                             float* __restrict__ dGx, bf16* __restrict__ dGx_b, float* __restrict__ dGh,
                             bf16* __restrict__ dGh_b, float* __restrict__ dh_direct, int B, int H) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int b = idx / H, u = idx % H;
  const float* x = gx + (int64_t)b * 3 * H;
  const float r = x[u], z = x[H + u], n = x[2 * H + u];
  const float hn = gh[(int64_t)b * 3 * H + 2 * H + u];
  float dh = 0.0f;
  if (dh_ext) dh += dh_ext[(int64_t)b * dhe_sb + u];
  if (dh_rec) dh += dh_rec[idx];
  const float hp = h_prev ? h_prev[idx] : 0.0f;
  const float dn = dh * (1.0f - z) * (1.0f - n * n);
  const float dz = dh * (hp - n) * z * (1.0f - z);
  const float dr = dn * hn * r * (1.0f - r);
  const int64_t o = (int64_t)b * 3 * H;
  dGx[o + u] = dr; dGx[o + H + u] = dz; dGx[o + 2 * H + u] = dn;
  dGh[o + u] = dr; dGh[o + H + u] = dz; dGh[o + 2 * H + u] = dn * r;
  dGx_b[o + u] = __float2bfloat16(dr); dGx_b[o + H + u] = __float2bfloat16(dz); dGx_b[o + 2 * H + u] = __float2bfloat16(dn);
  dGh_b[o + u] = __float2bfloat16(dr); dGh_b[o + H + u] = __float2bfloat16(dz); dGh_b[o + 2 * H + u] = __float2bfloat16(dn * r);
  dh_direct[idx] = dh * z;
}

__global__ void add_inplace(float* __restrict__ a, const float* __restrict__ b, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) a[i] += b[i];
}
// out[c] = sum_r in[r][c], deterministic, in two passes: block (x, y) sums row slab y of 32 columns with
// 8 row lanes and a fixed-order shared-memory combine -> part[y][c]; colsum_final adds the slabs in order.
// (One thread per column walking all T*B rows took 2.15 ms per layer at T*B = 32768, 4H = 4096.)
__global__ void colsum_partial(const float* __restrict__ in, int64_t rows, int cols, int64_t slab, float* __restrict__ part) {
  __shared__ float sm[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * slab;
  const int64_t r1 = r0 + slab < rows ? r0 + slab : rows;
  float acc = 0.0f;
  if (c < cols)
    for (int64_t r = r0 + threadIdx.y; r < r1; r += 8) acc += in[r * cols + c];
  sm[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float t = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sm[i][threadIdx.x];
    part[(int64_t)blockIdx.y * cols + c] = t;
  }
}
// out2 (nullable): second destination of the same sums (LSTM: db_hh == db_ih).  retire (nullable): counted once
// per block after its stores, like the GEMM tiles (see gemm_tc.cu).
__global__ void colsum_final(const float* __restrict__ part, int slabs, int cols, float* __restrict__ out,
                             float* __restrict__ out2, unsigned int* __restrict__ retire) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < cols) {
    float acc = 0.0f;
    for (int s = 0; s < slabs; ++s) acc += part[(int64_t)s * cols + c];
    out[c] = acc;
    if (out2 != nullptr) out2[c] = acc;
  }
  if (retire != nullptr) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(retire) : "memory");
  }
}
// fp32 [rows, cols] -> bf16 [rows, ld_out] (zero padded columns)
__global__ void cast_pad_bf16(const float* __restrict__ in, int64_t rows, int cols, int64_t ld_in, bf16* __restrict__ out, int ld_out) {
  const int64_t n = rows * ld_out;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / ld_out;
    const int c = (int)(i % ld_out);
    out[i] = __float2bfloat16(c < cols ? in[r * ld_in + c] : 0.0f);
  }
}

// o[r][c] = b[c]
__global__ void bcast_rows(float* __restrict__ o, const float* __restrict__ b, int64_t n, int cols) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = b[i % cols];
}

inline int grid_for(int64_t n, int threads = 256, int cap = 2368) {
  int64_t g = (n + threads - 1) / threads;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}
inline int pad8(int x) { return (x + 7) & ~7; }
// tmp must hold colsum_slabs(rows) * cols floats
inline int colsum_slabs(int64_t rows) { int64_t s = rows / 8; return (int)(s < 1 ? 1 : (s > 64 ? 64 : s)); }
inline int colsum(const float* in, int64_t rows, int cols, float* out, float* tmp, cudaStream_t st,
                  float* out2 = nullptr, unsigned int* retire = nullptr) {
  const int slabs = colsum_slabs(rows);
  const int64_t slab = (rows + slabs - 1) / slabs;
  colsum_partial<<<dim3((cols + 31) / 32, slabs), dim3(32, 8), 0, st>>>(in, rows, cols, slab, tmp);
  colsum_final<<<(cols + 127) / 128, 128, 0, st>>>(tmp, slabs, cols, out, out2, retire);
  return (cols + 127) / 128;   // blocks that will count themselves on `retire`
}

#define RC(x) do { int _rc = (x); if (_rc != B200RNN_OK) return _rc; } while (0)

}  // namespace
}  // namespace b200rnn

using namespace b200rnn;

// Workspace layout is computed by these helpers so Python only allocates flat buffers.
//   per layer l (in_l = input width, G = gates): all sizes in bytes, 256-aligned
struct LargeLayout {
  int64_t xb, gates, gh, c, h, hb, total;   // xb: bf16 [TB, in_pad]; gates: fp32 [TB, G*H]; gh (GRU): fp32 [TB, 3H];
};                                          // c: fp32 [TB,H] (LSTM); h: fp32 [TB,H]; hb: bf16 [TB,H]

static LargeLayout layer_layout(int cell, int in_l, int H, int64_t TB) {
  auto al = [](int64_t x) { return (x + 255) & ~int64_t(255); };
  const int G = cell == B200RNN_CELL_LSTM ? 4 : 3;
  LargeLayout L;
  int64_t off = 0;
  L.xb = off; off = al(off + TB * pad8(in_l) * 2);
  L.gates = off; off = al(off + TB * G * H * 4);
  L.gh = off; off = al(off + (cell == B200RNN_CELL_GRU ? TB * 3 * H * 4 : 0));
  L.c = off; off = al(off + (cell == B200RNN_CELL_LSTM ? TB * H * 4 : 0));
  L.h = off; off = al(off + TB * H * 4);
  L.hb = off; off = al(off + TB * H * 2);
  L.total = off;
  return L;
}

extern "C" int64_t b200rnn_rnn_large_stash_bytes(const b200rnn_model_desc* d) {
  if (!d) return 0;
  int64_t tot = 0;
  for (int l = 0; l < d->num_layers; ++l)
    tot += layer_layout(d->cell, l == 0 ? d->input_size : d->hidden_size, d->hidden_size, (int64_t)d->seq_len * d->batch).total;
  return tot;
}

extern "C" int64_t b200rnn_rnn_large_h_offset(const b200rnn_model_desc* d, int32_t layer) {
  if (!d || layer < 0 || layer >= d->num_layers) return -1;
  int64_t o = 0;
  const int64_t TB = (int64_t)d->seq_len * d->batch;
  for (int l = 0; l < layer; ++l) o += layer_layout(d->cell, l == 0 ? d->input_size : d->hidden_size, d->hidden_size, TB).total;
  return o + layer_layout(d->cell, layer == 0 ? d->input_size : d->hidden_size, d->hidden_size, TB).h;
}
extern "C" int64_t b200rnn_rnn_large_c_offset(const b200rnn_model_desc* d, int32_t layer) {
  if (!d || layer < 0 || layer >= d->num_layers || d->cell != B200RNN_CELL_LSTM) return -1;
  int64_t o = 0;
  const int64_t TB = (int64_t)d->seq_len * d->batch;
  for (int l = 0; l < layer; ++l) o += layer_layout(d->cell, l == 0 ? d->input_size : d->hidden_size, d->hidden_size, TB).total;
  return o + layer_layout(d->cell, layer == 0 ? d->input_size : d->hidden_size, d->hidden_size, TB).c;
}

// Packed bf16 weights: per layer W_ih [G*H, pad8(in)], W_hh [G*H, H], W_ih^T [in, pad8(G*H)... ] etc. are
// produced on the fly from the fp32 master copy (cheap next to the recurrence).
extern "C" int64_t b200rnn_rnn_large_scratch_bytes(const b200rnn_model_desc* d) {
  if (!d) return 0;
  const int H = d->hidden_size, G = d->cell == B200RNN_CELL_LSTM ? 4 : 3;
  const int inmax = d->input_size > H ? d->input_size : H;
  const int64_t TB = (int64_t)d->seq_len * d->batch;
  auto al = [](int64_t x) { return (x + 255) & ~int64_t(255); };
  int64_t tot = 0;
  tot += al((int64_t)G * H * pad8(inmax) * 2);      // W_ih bf16
  tot += al((int64_t)G * H * H * 2);                // W_hh bf16
  tot += al((int64_t)pad8(inmax) * G * H * 2);      // W_ih^T bf16 [in_pad? , G*H]
  tot += al((int64_t)H * G * H * 2);                // W_hh^T bf16
  tot += al(TB * G * H * 4) + al(TB * G * H * 2);   // dG fp32 + bf16 (GRU: dGx)
  tot += al(TB * G * H * 4) + al(TB * G * H * 2);   // GRU: dGh fp32 + bf16
  tot += al((int64_t)G * H * TB * 2);               // dG^T bf16
  tot += al((int64_t)pad8(inmax) * TB * 2);         // X^T / Hprev^T bf16
  tot += al(TB * H * 4) * 2;                        // dh_above ping-pong [TB,H]
  tot += al((int64_t)d->batch * H * 4) * 4;         // dh_rec, dc, dh_direct, bias tmp
  tot += al((int64_t)G * H * 4);                    // fused bias
  tot += al((TB * pad8(inmax) > (int64_t)G * H ? TB * pad8(inmax) : (int64_t)G * H) * 4);   // dX fp32 / colsum partials
  tot += 256;                                       // grid-barrier words of the persistent kernels
  tot += al(b200rnn_rnn_persist_stream_bytes(H));   // their double-buffered operand tile image
  return tot;
}

namespace {
struct Carver {
  char* p;
  template <typename T> T* take(int64_t bytes) { T* r = reinterpret_cast<T*>(p); p += (bytes + 255) & ~int64_t(255); return r; }
};
struct Scratch {
  bf16 *wih_b, *whh_b, *wihT_b, *whhT_b;
  float* dG; bf16* dG_b; float* dG2; bf16* dG2_b;
  bf16 *dGT_b, *xT_b;
  float* dxa[2];
  float *dh0, *dh1, *dc, *tmp, *bias, *dX;
  void* sync;
  void* stream_img;
};
Scratch carve(const b200rnn_model_desc* d, void* scratch) {
  const int H = d->hidden_size, G = d->cell == B200RNN_CELL_LSTM ? 4 : 3;
  const int inmax = d->input_size > H ? d->input_size : H;
  const int64_t TB = (int64_t)d->seq_len * d->batch, Bh = (int64_t)d->batch * H;
  const int64_t TBp = (TB + 7) & ~int64_t(7);
  Carver c{(char*)scratch};
  Scratch s;
  s.wih_b = c.take<bf16>((int64_t)G * H * pad8(inmax) * 2);
  s.whh_b = c.take<bf16>((int64_t)G * H * H * 2);
  s.wihT_b = c.take<bf16>((int64_t)pad8(inmax) * G * H * 2);
  s.whhT_b = c.take<bf16>((int64_t)H * G * H * 2);
  s.dG = c.take<float>(TB * G * H * 4); s.dG_b = c.take<bf16>(TB * G * H * 2);
  s.dG2 = c.take<float>(TB * G * H * 4); s.dG2_b = c.take<bf16>(TB * G * H * 2);
  s.dGT_b = c.take<bf16>((int64_t)G * H * TBp * 2);
  s.xT_b = c.take<bf16>((int64_t)pad8(inmax) * TBp * 2);
  s.dxa[0] = c.take<float>(TB * H * 4); s.dxa[1] = c.take<float>(TB * H * 4);
  s.dh0 = c.take<float>(Bh * 4); s.dh1 = c.take<float>(Bh * 4); s.dc = c.take<float>(Bh * 4); s.tmp = c.take<float>(Bh * 4);
  s.bias = c.take<float>((int64_t)G * H * 4);
  s.dX = c.take<float>((TB * pad8(inmax) > (int64_t)G * H ? TB * pad8(inmax) : (int64_t)G * H) * 4);
  s.sync = c.take<char>(256);
  s.stream_img = c.take<char>(b200rnn_rnn_persist_stream_bytes(H));
  return s;
}
}  // namespace

extern "C" int b200rnn_rnn_large_fwd(const b200rnn_model_desc* d, const float* params, const float* x_tm /* [T,B,I] time-major */,
                                     void* stash, void* scratch, void* stream) {
  B200_REQUIRE(d && params && x_tm && stash && scratch, B200RNN_EINVAL, "rnn_large_fwd: null argument");
  const int H = d->hidden_size, B = d->batch, T = d->seq_len, L = d->num_layers;
  const int G = d->cell == B200RNN_CELL_LSTM ? 4 : 3;
  B200_REQUIRE(H % 8 == 0 && H >= 8, B200RNN_EUNSUPPORTED, "rnn_large: hidden size must be a multiple of 8 (H=%d)", H);
  const int64_t TB = (int64_t)T * B, Bh = (int64_t)B * H;
  cudaStream_t s = (cudaStream_t)stream;
  const Scratch sc = carve(d, scratch);
  char* st = (char*)stash;
  const float* inp = x_tm;
  int in_l = d->input_size;
  for (int l = 0; l < L; ++l) {
    const LargeLayout lay = layer_layout(d->cell, in_l, H, TB);
    bf16* xb = (bf16*)(st + lay.xb);
    float* gates = (float*)(st + lay.gates);
    float* gh = (float*)(st + lay.gh);
    float* c = (float*)(st + lay.c);
    float* h = (float*)(st + lay.h);
    bf16* hb = (bf16*)(st + lay.hb);
    const int inp8 = pad8(in_l);
    // operands -> bf16 (zero-padded K)
    cast_pad_bf16<<<grid_for(TB * inp8), 256, 0, s>>>(inp, TB, in_l, in_l, xb, inp8);
    cast_pad_bf16<<<grid_for((int64_t)G * H * inp8), 256, 0, s>>>(params + d->off_wih[l], (int64_t)G * H, in_l, in_l, sc.wih_b, inp8);
    cast_pad_bf16<<<grid_for((int64_t)G * H * H), 256, 0, s>>>(params + d->off_whh[l], (int64_t)G * H, H, H, sc.whh_b, H);
    B200_CHECK_LAUNCH();
    if (d->cell == B200RNN_CELL_LSTM) {
      // fused bias b_ih + b_hh enters through the epilogue of the all-timestep GEMM (K1)
      B200_CHECK_CUDA(cudaMemcpyAsync(sc.bias, params + d->off_bih[l], (size_t)G * H * 4, cudaMemcpyDeviceToDevice, s));
      add_inplace<<<grid_for((int64_t)G * H), 256, 0, s>>>(sc.bias, params + d->off_bhh[l], (int64_t)G * H);
      RC(b200rnn_gemm_bf16(xb, sc.wih_b, gates, sc.bias, (int32_t)TB, G * H, inp8, inp8, inp8, G * H, 0, 0, stream));
      if (b200rnn_rnn_persist_supported(d)) {
        RC(b200rnn_lstm_persist_fwd(T, B, H, gates, sc.whh_b, c, h, sc.stream_img, sc.sync, stream));
      } else
      for (int t = 0; t < T; ++t) {
        float* gt = gates + t * Bh * G;
        if (t > 0) RC(b200rnn_gemm_bf16(hb + (t - 1) * Bh, sc.whh_b, gt, nullptr, B, G * H, H, H, H, G * H, 1, 0, stream));
        lstm_cell_fwd<<<grid_for(Bh, 256, 1 << 20), 256, 0, s>>>(gt, t > 0 ? c + (t - 1) * Bh : nullptr, c + t * Bh,
                                                                  h + t * Bh, hb + t * Bh, B, H, H);
      }
    } else {
      RC(b200rnn_gemm_bf16(xb, sc.wih_b, gates, params + d->off_bih[l], (int32_t)TB, G * H, inp8, inp8, inp8, G * H, 0, 0, stream));
      if (b200rnn_rnn_persist_supported(d)) {
        RC(b200rnn_gru_persist_fwd(T, B, H, gates, gh, sc.whh_b, params + d->off_bhh[l], h, sc.stream_img, sc.sync, stream));
      } else
      for (int t = 0; t < T; ++t) {
        float* ght = gh + t * Bh * G;
        if (t > 0) RC(b200rnn_gemm_bf16(hb + (t - 1) * Bh, sc.whh_b, ght, params + d->off_bhh[l], B, G * H, H, H, H, G * H, 0, 0, stream));
        else bcast_rows<<<grid_for(Bh * G, 256, 1 << 20), 256, 0, s>>>(ght, params + d->off_bhh[l], Bh * G, G * H);  // h_{-1} = 0
        gru_cell_fwd<<<grid_for(Bh, 256, 1 << 20), 256, 0, s>>>(gates + t * Bh * G, ght, t > 0 ? h + (t - 1) * Bh : nullptr,
                                                                 h + t * Bh, hb + t * Bh, B, H, H);
      }
    }
    B200_CHECK_LAUNCH();
    inp = h;
    in_l = H;
    st += lay.total;
  }
  return B200RNN_OK;
}

// Backward through time of ONE layer (the body of b200rnn_rnn_large_bwd).  Gradients of this layer's four
// parameter tensors are complete when the call's kernels have run; every kernel that writes one of them (the two
// weight-gradient GEMMs tile by tile, the bias column sums block by block) counts itself on `retire` (nullable)
// as it finishes, and *retire_count receives how many such counts this call will add — the reducer's trigger
// for the layer's gradient bucket (csrc/runtime/reducer.cpp; c10d's per-bucket "grad ready" hooks fire at
// the same point, src/motion/trainer/ddp.py:19).
static int large_bwd_layer(const b200rnn_model_desc* d, const float* params, const float* x_tm, const void* stash,
                           const Scratch& sc, int l, const float* d_out_tm, const float* d_hn, const float* d_cn,
                           float* grad_out, float* dx_tm, unsigned int* retire, int32_t* retire_count, void* stream) {
  const int H = d->hidden_size, B = d->batch, T = d->seq_len, L = d->num_layers;
  const bool lstm = d->cell == B200RNN_CELL_LSTM;
  const int G = lstm ? 4 : 3;
  const int64_t TB = (int64_t)T * B, Bh = (int64_t)B * H;
  const int64_t TBp = (TB + 7) & ~int64_t(7);
  cudaStream_t s = (cudaStream_t)stream;
  int32_t counted = 0, n = 0;
  // layer offsets inside the stash
  int64_t loff[B200RNN_MAX_LAYERS];
  {
    int64_t o = 0;
    for (int k = 0; k < L; ++k) { loff[k] = o; o += layer_layout(d->cell, k == 0 ? d->input_size : H, H, TB).total; }
  }
  {
    const int in_l = l == 0 ? d->input_size : H;
    const LargeLayout lay = layer_layout(d->cell, in_l, H, TB);
    const char* st = (const char*)stash + loff[l];
    const float* gates = (const float*)(st + lay.gates);
    const float* gh = (const float*)(st + lay.gh);
    const float* c = (const float*)(st + lay.c);
    const float* h = (const float*)(st + lay.h);
    const float* xin = l == 0 ? x_tm : (const float*)((const char*)stash + loff[l - 1] + layer_layout(d->cell, l == 1 ? d->input_size : H, H, TB).h);
    const float* dh_ext = (l == L - 1) ? d_out_tm : sc.dxa[(l + 1) & 1];
    // W^T operands for the "x W" products
    RC(b200rnn_transpose_cast_bf16(params + d->off_whh[l], sc.whhT_b, G * H, H, H, G * H, stream));
    RC(b200rnn_transpose_cast_bf16(params + d->off_wih[l], sc.wihT_b, G * H, in_l, in_l, G * H, stream));
    // carries
    if (d_hn) B200_CHECK_CUDA(cudaMemcpyAsync(sc.dh0, d_hn + l * Bh, Bh * 4, cudaMemcpyDeviceToDevice, s));
    else B200_CHECK_CUDA(cudaMemsetAsync(sc.dh0, 0, Bh * 4, s));
    if (lstm) {
      if (d_cn) B200_CHECK_CUDA(cudaMemcpyAsync(sc.dc, d_cn + l * Bh, Bh * 4, cudaMemcpyDeviceToDevice, s));
      else B200_CHECK_CUDA(cudaMemsetAsync(sc.dc, 0, Bh * 4, s));
    }
    float* dh_cur = sc.dh0;
    float* dh_nxt = sc.dh1;
    if (lstm && b200rnn_rnn_persist_supported(d)) {
      RC(b200rnn_lstm_persist_bwd(T, B, H, gates, c, dh_ext, d_hn ? d_hn + l * Bh : nullptr, d_cn ? d_cn + l * Bh : nullptr,
                                  sc.whhT_b, sc.dG, sc.dG_b, sc.stream_img, sc.sync, stream));
    } else if (!lstm && b200rnn_rnn_persist_supported(d)) {
      RC(b200rnn_gru_persist_bwd(T, B, H, gates, gh, h, dh_ext, d_hn ? d_hn + l * Bh : nullptr, sc.whhT_b,
                                 sc.dG, sc.dG_b, sc.dG2, sc.dG2_b, sc.stream_img, sc.sync, stream));
    } else
    for (int t = T - 1; t >= 0; --t) {
      const float* dhe = dh_ext ? dh_ext + t * Bh : nullptr;
      if (lstm) {
        lstm_cell_bwd<<<grid_for(Bh, 256, 1 << 20), 256, 0, s>>>(
            gates + t * Bh * G, c + t * Bh, t > 0 ? c + (t - 1) * Bh : nullptr, dhe, H, dh_cur, sc.dc,
            sc.dG + t * Bh * G, sc.dG_b + t * Bh * G, B, H);
        if (t > 0) RC(b200rnn_gemm_bf16(sc.dG_b + t * Bh * G, sc.whhT_b, dh_nxt, nullptr, B, H, G * H, G * H, G * H, H, 0, 0, stream));
      } else {
        gru_cell_bwd<<<grid_for(Bh, 256, 1 << 20), 256, 0, s>>>(
            gates + t * Bh * G, gh + t * Bh * G, t > 0 ? h + (t - 1) * Bh : nullptr, dhe, H, dh_cur,
            sc.dG + t * Bh * G, sc.dG_b + t * Bh * G, sc.dG2 + t * Bh * G, sc.dG2_b + t * Bh * G, dh_nxt, B, H);
        if (t > 0) RC(b200rnn_gemm_bf16(sc.dG2_b + t * Bh * G, sc.whhT_b, dh_nxt, nullptr, B, H, G * H, G * H, G * H, H, 1, 0, stream));
      }
      float* tmp = dh_cur; dh_cur = dh_nxt; dh_nxt = tmp;
    }
    B200_CHECK_LAUNCH();
    // ---- bulk GEMMs over all timesteps (K4) ----
    // dX = dG x W_ih  -> gradient for the layer below (or dx)
    if (l > 0 || dx_tm != nullptr) {
      float* dX = l > 0 ? sc.dxa[l & 1] : dx_tm;
      RC(b200rnn_gemm_bf16(sc.dG_b, sc.wihT_b, dX, nullptr, (int32_t)TB, in_l, G * H, G * H, G * H, in_l, 0, 0, stream));
    }
    // dW_ih = dG^T x X ; dW_hh = dGh^T x H_prev
    B200_CHECK_CUDA(cudaMemsetAsync(sc.dGT_b, 0, (size_t)G * H * TBp * 2, s));
    RC(b200rnn_transpose_cast_bf16(sc.dG, sc.dGT_b, (int32_t)TB, G * H, G * H, TBp, stream));
    B200_CHECK_CUDA(cudaMemsetAsync(sc.xT_b, 0, (size_t)pad8(in_l > H ? in_l : H) * TBp * 2, s));
    RC(b200rnn_transpose_cast_bf16(xin, sc.xT_b, (int32_t)TB, in_l, in_l, TBp, stream));
    RC(b200rnn_gemm_bf16_retire(sc.dGT_b, sc.xT_b, grad_out + d->off_wih[l], nullptr, G * H, in_l, (int32_t)TBp, (int32_t)TBp,
                                (int32_t)TBp, in_l, 0, 0, retire, &n, stream));
    counted += n;
    if (!lstm) {
      B200_CHECK_CUDA(cudaMemsetAsync(sc.dGT_b, 0, (size_t)G * H * TBp * 2, s));
      RC(b200rnn_transpose_cast_bf16(sc.dG2, sc.dGT_b, (int32_t)TB, G * H, G * H, TBp, stream));
    }
    B200_CHECK_CUDA(cudaMemsetAsync(sc.xT_b, 0, (size_t)pad8(in_l > H ? in_l : H) * TBp * 2, s));
    if (T > 1) RC(b200rnn_transpose_cast_bf16(h, sc.xT_b + B, (int32_t)(TB - B), H, H, TBp, stream));   // H_prev = h shifted by one step
    RC(b200rnn_gemm_bf16_retire(sc.dGT_b, sc.xT_b, grad_out + d->off_whh[l], nullptr, G * H, H, (int32_t)TBp, (int32_t)TBp,
                                (int32_t)TBp, H, 0, 0, retire, &n, stream));
    counted += n;
    // biases
    // sc.dX (TB x pad8(inmax) floats >= slabs x G*H) is free here and serves as the partial-sum buffer
    if (lstm) counted += colsum(sc.dG, TB, G * H, grad_out + d->off_bih[l], sc.dX, s, grad_out + d->off_bhh[l], retire);
    else {
      counted += colsum(sc.dG, TB, G * H, grad_out + d->off_bih[l], sc.dX, s, nullptr, retire);
      counted += colsum(sc.dG2, TB, G * H, grad_out + d->off_bhh[l], sc.dX, s, nullptr, retire);
    }
    B200_CHECK_LAUNCH();
  }
  if (retire_count) *retire_count = counted;
  return B200RNN_OK;
}

// Backward through time.  d_out_tm: gradient of the top layer's output sequence, time-major [T,B,H]
// (nullable); d_hn / d_cn: [L,B,H] (nullable).  grad_out: flat gradient in parameter layout
// (d->n_params floats, fully overwritten for the RNN parameters).  dx_tm: [T,B,I] (nullable).
extern "C" int b200rnn_rnn_large_bwd(const b200rnn_model_desc* d, const float* params, const float* x_tm, const void* stash,
                                     void* scratch, const float* d_out_tm, const float* d_hn, const float* d_cn,
                                     float* grad_out, float* dx_tm, void* stream) {
  B200_REQUIRE(d && params && x_tm && stash && scratch && grad_out, B200RNN_EINVAL, "rnn_large_bwd: null argument");
  const Scratch sc = carve(d, scratch);
  for (int l = d->num_layers - 1; l >= 0; --l)
    RC(large_bwd_layer(d, params, x_tm, stash, sc, l, d_out_tm, d_hn, d_cn, grad_out, dx_tm, nullptr, nullptr, stream));
  return B200RNN_OK;
}

// One layer of the above (call for layer = L-1 down to 0 on the same stash / scratch: the gradient that flows
// to the layer below is kept in the scratch between calls).  See large_bwd_layer for retire / retire_count.
extern "C" int b200rnn_rnn_large_bwd_layer(const b200rnn_model_desc* d, const float* params, const float* x_tm,
                                           const void* stash, void* scratch, int32_t layer, const float* d_out_tm,
                                           const float* d_hn, const float* d_cn, float* grad_out, float* dx_tm,
                                           uint32_t* retire_counter, int32_t* retire_count, void* stream) {
  B200_REQUIRE(d && params && x_tm && stash && scratch && grad_out && layer >= 0 && layer < d->num_layers, B200RNN_EINVAL,
               "rnn_large_bwd_layer: bad argument");
  const Scratch sc = carve(d, scratch);
  return large_bwd_layer(d, params, x_tm, stash, sc, layer, d_out_tm, d_hn, d_cn, grad_out, dx_tm, retire_counter,
                         retire_count, stream);
}
