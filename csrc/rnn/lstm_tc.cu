// K2 (tensor-core form) — persistent LSTM forward for H = 32 on tcgen05 / TMEM.
//
// Same contract and stash layout as b200rnn_lstm_fwd (lstm_small.inc), different engine: the gate
// pre-activations of a whole batch tile are ONE small GEMM per timestep,
//
//        D[128 gate rows, 16 samples]  =  Wcat[128, K]  x  [x_t | h_{t-1}]^T[K, 16]
//
// issued by a single thread as tcgen05.mma (M=128, N=16, kind::tf32) with the accumulator in
// TMEM; Wcat = [W_ih | W_hh] stays resident in shared memory for all T steps and both operands
// use the no-swizzle K-major core-matrix layout, written by ordinary stores (no TMA needed: the
// B operand is produced by the CTA itself every step).  The gate rows are permuted so that TMEM
// lane 32w+l holds gate (l & 3) of hidden unit 8w + (l >> 2): after tcgen05.ld each thread
// applies its gate's non-linearity to its 16 samples, a 4-lane register transpose gives every
// lane all four gates of 4 samples, and the cell update, the next-step operand (h_t, written
// back into the B tile) and the training stash are produced in place.
//
// fp32 fidelity: operands are split a = a_hi + a_lo (a_hi = top 19 bits, i.e. exactly what the
// tf32 datapath reads) and D = a_hi*b_hi + a_lo*b_hi + a_hi*b_lo is accumulated in fp32 — the
// "3xTF32" scheme; dropped terms are O(2^-20) relative, the same order as fp32 summation error.
//
// Replaces the FFMA gate GEMV of lstm_small.inc (FFMA issue bound: ~83 M warp-FFMAs per forward)
// with ~18-24 MMA instructions per step.  Reference call site: src/motion/model.py:14.
#include <stdlib.h>
#include "../gemm/tc_common.cuh"

namespace b200rnn {
namespace {

using namespace tc;

constexpr int H = 32;
constexpr int NT = 128;          // threads = gate rows = TMEM lanes
constexpr int NS = 16;           // sample columns of the MMA (N)
constexpr int KMAX = 64;         // max operand width (stacked layers: H + H)
constexpr int B_BYTES = NS * KMAX * 4;    // 4 KiB per split part per buffer
constexpr int TC_SMEM = 4 * B_BYTES + 64 + 128;

struct TcFwdArgs {
  const float* x;
  const float* params;
  float* h_seq;
  float* c_seq;
  float* g_seq;
  int64_t off_wih[B200RNN_MAX_LAYERS], off_whh[B200RNN_MAX_LAYERS];
  int64_t off_bih[B200RNN_MAX_LAYERS], off_bhh[B200RNN_MAX_LAYERS];
  int I, T, B, L, BT, training;
  int* error;
};

__device__ int g_tc_error = 0;
__device__ long long g_tc_prof[2][8];   // per-phase clock64 stamps of one step (profiling aid)
#define TC_STAMP(i) do { if (stamp) g_tc_prof[tid == 0 ? 0 : 1][i] = clock64(); } while (0)

// byte offset of element (row, k) in a K-major no-swizzle operand tile whose rows are kpad wide
__device__ __forceinline__ uint32_t op_off(int row, int k, int kpad) {
  return (uint32_t)((row >> 3) * (kpad * 32) + (k >> 2) * 128 + (row & 7) * 16 + (k & 3) * 4);
}
__device__ __forceinline__ void split_store(uint8_t* hi_base, uint8_t* lo_base, uint32_t off, float v) {
  const float hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
  *reinterpret_cast<float*>(hi_base + off) = hi;
  *reinterpret_cast<float*>(lo_base + off) = v - hi;
}

// 4-lane transpose: on entry lane q (= lane & 3) holds a[16] = its gate for samples 0..15;
// on return o[g][s] = gate g of sample 4q + s.
__device__ __forceinline__ void gate_transpose(const float (&a)[16], int q, float (&o)[4][4]) {
  const bool hi = (q & 2) != 0, lo = (q & 1) != 0;
  float k0[2][4], k1[2][4];   // [local block][sample]: own gate (k0) and gate q^2 (k1)
#pragma unroll
  for (int bl = 0; bl < 2; ++bl)
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const float keep = hi ? a[8 + 4 * bl + s] : a[4 * bl + s];
      const float send = hi ? a[4 * bl + s] : a[8 + 4 * bl + s];
      k0[bl][s] = keep;
      k1[bl][s] = __shfl_xor_sync(0xffffffffu, send, 2);
    }
  float r0[4], r1[4], m0[4], m1[4];   // mine: gates q, q^2; received: gates q^1, q^3
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    m0[s] = lo ? k0[1][s] : k0[0][s];
    m1[s] = lo ? k1[1][s] : k1[0][s];
    const float s0 = lo ? k0[0][s] : k0[1][s];
    const float s1 = lo ? k1[0][s] : k1[1][s];
    r0[s] = __shfl_xor_sync(0xffffffffu, s0, 1);
    r1[s] = __shfl_xor_sync(0xffffffffu, s1, 1);
  }
#pragma unroll
  for (int g = 0; g < 4; ++g)
#pragma unroll
    for (int s = 0; s < 4; ++s)
    {
      // branch-free 4-way select on the lane-dependent index g ^ q (0: m0, 2: m1, 1: r0, 3: r1);
      // a nested ternary on q compiles to divergent branches (measured: ~2000 cycles per step)
      const int x = g ^ q;
      const float t0 = (x & 2) ? m1[s] : m0[s];
      const float t1 = (x & 2) ? r1[s] : r0[s];
      o[g][s] = (x & 1) ? t1 : t0;
    }
}

// TMEM column map (256 columns allocated per CTA; two CTAs fill the SM's 512):
//   [0,64)   W_hi  — resident A operand, row m = TMEM lane m, one tf32 element per column
//   [64,128) W_lo
//   [128,144) D0 = W_hi x v_hi   [144,160) D1 = W_lo x v_hi   [160,176) D2 = W_hi x v_lo
// Three independent accumulators keep the three 3xTF32 product chains from serialising on one D.
constexpr uint32_t COL_ALO = 64, COL_D0 = 128, COL_D1 = 144, COL_D2 = 160;
constexpr int TMEM_COLS = 256;

template <int KPAD, int NSPLIT, int NB>
__device__ __forceinline__ void tc_fwd_layer(
    const float* inp, int64_t in_sb, int64_t in_st, int in_dim,
    const float* __restrict__ w_ih, const float* __restrict__ w_hh,
    const float* __restrict__ b_ih, const float* __restrict__ b_hh,
    float* h_seq, float* c_seq, float* g_seq,
    int T, int B, int b0, int BTc, int training,
    uint8_t* Bop, uint64_t* mbar, uint32_t tmem_base, uint32_t& phase, bool& ok) {
  constexpr int INP = KPAD - H;
  constexpr int NV = 4 * NB;                       // sample slots this instantiation processes
  constexpr int NPF = (NV * INP + NT - 1) / NT;
  constexpr uint32_t IDESC = make_idesc(2, 128, NS);
  constexpr uint32_t SBO = KPAD * 32;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q = lane & 3;                    // my gate (as a TMEM lane) / my sample block (after the transpose)
  const int u = warp * 8 + (lane >> 2);      // my hidden unit
  const int r_pt = q * H + u;                // my gate row in torch's [i|f|g|o] layout
  const uint32_t lane_base = tmem_base + (uint32_t(warp * 32) << 16);

  // ---- resident A operand: my row of [W_ih | W_hh] -> TMEM, split hi/lo ----
#pragma unroll 1
  for (int c = 0; c < KPAD / 16; ++c) {
    float hi[16], lo[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int k = c * 16 + j;
      float v;
      if (k < INP) v = (k < in_dim) ? w_ih[(int64_t)r_pt * in_dim + k] : 0.0f;
      else v = w_hh[(int64_t)r_pt * H + (k - INP)];
      hi[j] = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
      lo[j] = v - hi[j];
    }
    tmem_st16(lane_base + c * 16, hi);
    tmem_st16(lane_base + COL_ALO + c * 16, lo);
  }
  tmem_st_wait();
  const float bias = b_ih[r_pt] + b_hh[r_pt];
  // ---- B operand buffers: zero everything (rows >= BTc and padded K stay zero forever) ----
  for (int i = tid; i < 4 * B_BYTES / 4; i += NT) reinterpret_cast<float*>(Bop)[i] = 0.0f;
  __syncthreads();

  float pf[NPF];
  auto load_inputs = [&](int t) {
#pragma unroll
    for (int i = 0; i < NPF; ++i) {
      const int idx = tid + i * NT;
      const int b = idx / INP, k = idx % INP;
      float v = 0.0f;
      if (idx < NV * INP && b < BTc && k < in_dim) v = inp[(int64_t)(b0 + b) * in_sb + (int64_t)t * in_st + k];
      pf[i] = v;
    }
  };
  auto store_inputs = [&](int buf) {
    uint8_t* hi = Bop + buf * 2 * B_BYTES;
#pragma unroll
    for (int i = 0; i < NPF; ++i) {
      const int idx = tid + i * NT;
      if (idx < NV * INP) split_store(hi, hi + B_BYTES, op_off(idx / INP, idx % INP, KPAD), pf[i]);
    }
  };
  load_inputs(0);
  store_inputs(0);
  float c_reg[4] = {0.f, 0.f, 0.f, 0.f};
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();

  for (int t = 0; t < T; ++t) {
    const int cur = t & 1, nxt = cur ^ 1;
    const bool stamp = (t == 8) && (blockIdx.x == 0) && (tid == 0 || tid == 127) && (INP == H);
    TC_STAMP(0);
    if (tid == 0) {
      tc_fence_after_sync();
      const uint32_t b_hi = smem_u32(Bop + cur * 2 * B_BYTES), b_lo = b_hi + B_BYTES;
#pragma unroll
      for (int ks = 0; ks < KPAD / 8; ++ks) {
        const uint64_t dbh = make_smem_desc(b_hi + ks * 256, 128, SBO, 0);
        mma_tf32_ts(tmem_base + COL_D0, tmem_base + ks * 8, dbh, IDESC, ks > 0 ? 1u : 0u);
        if (NSPLIT == 3) {
          const uint64_t dbl = make_smem_desc(b_lo + ks * 256, 128, SBO, 0);
          mma_tf32_ts(tmem_base + COL_D1, tmem_base + COL_ALO + ks * 8, dbh, IDESC, ks > 0 ? 1u : 0u);
          mma_tf32_ts(tmem_base + COL_D2, tmem_base + ks * 8, dbl, IDESC, ks > 0 ? 1u : 0u);
        }
      }
      mma_commit(mbar);
    }
    TC_STAMP(1);
    if (t + 1 < T) load_inputs(t + 1);        // global loads overlap the MMA
    if (ok && !mbar_wait(mbar, phase)) { ok = false; atomicExch(&g_tc_error, 1); }
    phase ^= 1;
    tc_fence_after_sync();
    TC_STAMP(2);
    float v[16];
    {
      uint32_t r0[16], r1[16], r2[16];
      tmem_ld16_nowait(lane_base + COL_D0, r0);
      if (NSPLIT == 3) { tmem_ld16_nowait(lane_base + COL_D1, r1); tmem_ld16_nowait(lane_base + COL_D2, r2); }
      tmem_ld_wait();
#pragma unroll
      for (int b = 0; b < 16; ++b) {
        float x = __uint_as_float(r0[b]);
        if (NSPLIT == 3) x += __uint_as_float(r1[b]) + __uint_as_float(r2[b]);
        v[b] = x;
      }
    }
    // my gate's non-linearity on the NV live sample slots, branch-free (sigmoid for i,f,o;
    // tanh = 2*sigmoid(2x)-1 for g) so the NV chains interleave
    TC_STAMP(3);
    const float sc = (q == 2) ? 2.0f : 1.0f;
    const float mul = (q == 2) ? 2.0f : 1.0f, add = (q == 2) ? -1.0f : 0.0f;
#pragma unroll
    for (int b = 0; b < NV; ++b) v[b] = fmaf(mul, sigmoid_f(sc * (v[b] + bias)), add);
    float g4[4][4];
    gate_transpose(v, q, g4);
    uint8_t* bn_hi = Bop + nxt * 2 * B_BYTES;
    TC_STAMP(4);
    if (q < NB) {
      float hh[4], cc[4];
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        cc[s] = fmaf(g4[1][s], c_reg[s], g4[0][s] * g4[2][s]);
        hh[s] = g4[3][s] * tanh_f(cc[s]);
        c_reg[s] = cc[s];
      }
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const int b = 4 * q + s;
        if (b < BTc) {
          split_store(bn_hi, bn_hi + B_BYTES, op_off(b, INP + u, KPAD), hh[s]);
          const int64_t e = ((int64_t)t * B + (b0 + b)) * H + u;
          h_seq[e] = hh[s];
          if (training) {
            c_seq[e] = cc[s];
            reinterpret_cast<float4*>(g_seq)[e] = make_float4(g4[0][s], g4[1][s], g4[2][s], g4[3][s]);
          }
        }
      }
    }
    TC_STAMP(5);
    if (t + 1 < T) store_inputs(nxt);
    fence_proxy_async_smem();     // generic-proxy operand writes -> visible to the tensor core
    tc_fence_before_sync();       // my tcgen05.ld completed before the next MMA overwrites D
    TC_STAMP(6);
    __syncthreads();
    TC_STAMP(7);
  }
}

template <int KPAD0, int NSPLIT, int NB>
__global__ void __launch_bounds__(NT) lstm_fwd_tc_kernel(const TcFwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint8_t* Bop = smem;                               // [buf][hi|lo]
  uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + 4 * B_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);

  const int b0 = blockIdx.x * a.BT;
  const int BTc = min(a.BT, a.B - b0);
  if (BTc <= 0) return;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(mbar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  uint32_t phase = 0;
  bool ok = true;
  const int64_t lsz = (int64_t)a.T * a.B * H;
  tc_fwd_layer<KPAD0, NSPLIT, NB>(a.x, (int64_t)a.T * a.I, a.I, a.I,
                      a.params + a.off_wih[0], a.params + a.off_whh[0], a.params + a.off_bih[0], a.params + a.off_bhh[0],
                      a.h_seq, a.c_seq, a.g_seq, a.T, a.B, b0, BTc, a.training,
                      Bop, mbar, tmem_base, phase, ok);
  for (int l = 1; l < a.L; ++l) {
    __syncthreads();
    tc_fwd_layer<2 * H, NSPLIT, NB>(a.h_seq + (l - 1) * lsz, H, (int64_t)a.B * H, H,
                        a.params + a.off_wih[l], a.params + a.off_whh[l], a.params + a.off_bih[l], a.params + a.off_bhh[l],
                        a.h_seq + l * lsz, a.c_seq + l * lsz, a.g_seq + 4 * l * lsz, a.T, a.B, b0, BTc, a.training,
                        Bop, mbar, tmem_base, phase, ok);
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<TMEM_COLS>(tmem_base);
}

int tc_choose_bt(int B) {
  int bt = (B + 2 * num_sms() - 1) / (2 * num_sms());   // two CTAs per SM: one's MMA hides the other's epilogue
  if (bt < 1) bt = 1;
  if (bt > NS) bt = NS;
  return bt;
}

}  // namespace
}  // namespace b200rnn

using namespace b200rnn;

extern "C" int b200rnn_lstm_fwd_tc(const b200rnn_model_desc* d, const float* params, const float* x, float* stash,
                                   int32_t training, void* stream) {
  B200_REQUIRE(d && params && x && stash, B200RNN_EINVAL, "lstm_fwd_tc: null argument");
  B200_REQUIRE(d->cell == B200RNN_CELL_LSTM && b200rnn_rnn_small_supported(d), B200RNN_EUNSUPPORTED,
               "lstm_fwd_tc: shape (cell=%d I=%d H=%d L=%d) not covered", d->cell, d->input_size, d->hidden_size, d->num_layers);
  TcFwdArgs a;
  a.x = x; a.params = params;
  const int64_t lsz = (int64_t)d->num_layers * d->seq_len * d->batch * d->hidden_size;
  a.h_seq = stash; a.c_seq = stash + lsz; a.g_seq = stash + 2 * lsz;
  for (int l = 0; l < B200RNN_MAX_LAYERS; ++l) {
    a.off_wih[l] = d->off_wih[l]; a.off_whh[l] = d->off_whh[l];
    a.off_bih[l] = d->off_bih[l]; a.off_bhh[l] = d->off_bhh[l];
  }
  a.I = d->input_size; a.T = d->seq_len; a.B = d->batch; a.L = d->num_layers;
  a.BT = tc_choose_bt(d->batch); a.training = training; a.error = nullptr;
  const int n_cta = (a.B + a.BT - 1) / a.BT;
  cudaStream_t s = (cudaStream_t)stream;
  static int split = 0;
  if (split == 0) {
    // B200RNN_TC_SPLIT=1 selects single-pass TF32 (10-bit mantissa operands); default 3 = 3xTF32 (fp32 fidelity)
    const char* e = getenv("B200RNN_TC_SPLIT");
    split = (e && e[0] == '1') ? 1 : 3;
  }
  const bool narrow = d->input_size <= 16;
  const int nb = (a.BT + 3) / 4;
#define B200_TC_LAUNCH(KP, SP, NBV)                                                                    \
  do {                                                                                                 \
    B200_CHECK_CUDA(cudaFuncSetAttribute(lstm_fwd_tc_kernel<KP, SP, NBV>,                              \
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM));      \
    lstm_fwd_tc_kernel<KP, SP, NBV><<<n_cta, NT, TC_SMEM, s>>>(a);                                     \
  } while (0)
#define B200_TC_NB(KP, SP)                                                                             \
  switch (nb) {                                                                                        \
    case 1: B200_TC_LAUNCH(KP, SP, 1); break;                                                          \
    case 2: B200_TC_LAUNCH(KP, SP, 2); break;                                                          \
    case 3: B200_TC_LAUNCH(KP, SP, 3); break;                                                          \
    default: B200_TC_LAUNCH(KP, SP, 4); break;                                                         \
  }
  if (split == 3) { if (narrow) { B200_TC_NB(48, 3) } else { B200_TC_NB(64, 3) } }
  else            { if (narrow) { B200_TC_NB(48, 1) } else { B200_TC_NB(64, 1) } }
#undef B200_TC_NB
#undef B200_TC_LAUNCH
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_tc_status(int32_t* status) {
  B200_REQUIRE(status != nullptr, B200RNN_EINVAL, "tc_status: null");
  int v = 0;
  B200_CHECK_CUDA(cudaMemcpyFromSymbol(&v, g_tc_error, sizeof(int)));
  *status = v;
  return B200RNN_OK;
}

extern "C" int b200rnn_tc_prof(long long* out16) {
  B200_REQUIRE(out16 != nullptr, B200RNN_EINVAL, "tc_prof: null");
  B200_CHECK_CUDA(cudaMemcpyFromSymbol(out16, g_tc_prof, sizeof(long long) * 16));
  return B200RNN_OK;
}
