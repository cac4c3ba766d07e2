// K2/K3 — fused persistent LSTM forward / backward-through-time, small-hidden family.
//
// One launch runs the WHOLE sequence for ALL layers: each CTA owns a tile of
// batch rows and walks t = 0..T-1 (forward) or T-1..0 (backward) with the
// recurrent + input weights held in REGISTERS for the entire sequence, the
// per-step operand vector [x_t | h_{t-1}] staged in shared memory, and the gate
// GEMV + sigmoid/tanh + cell update fused.  Backward additionally fuses the
// classifier head + cross-entropy + correct-count, keeps dW accumulators in
// registers across all T steps and emits one partial gradient row per CTA
// (summed in fixed order downstream => deterministic, rank-independent results).
//
// Replaces ATen's per-timestep lstm (src/motion/model.py:9-17) and its autograd
// backward (src/motion/trainer/base.py:116) for the reference's model geometry
// (H=32, I=9, L=2, T=128).  fp32 FFMA math: the reference is fp32 and at H=32
// the step is bound by the serial dependency chain, not by FLOPs.
//
// Thread mapping (H = hidden, KQ lanes per hidden unit, NT = H*KQ threads):
//   "A" mapping  tid = u*KQ + kq : thread owns hidden unit u, the 4 gate rows of
//                u, and the K-slice [kq*KPT, (kq+1)*KPT) of the operand vector.
//   "B" mapping  tid = k*JQ + jq : (backward only) thread owns output column k of
//                dG x [W_ih | W_hh] and the gate-row slice jq.
#include "../common.cuh"

namespace b200rnn {
namespace {

constexpr int kMaxBT = 16;

struct FwdArgs {
  const float* x;
  const float* params;
  float* h_seq;
  float* c_seq;
  float* g_seq;
  int64_t off_wih[B200RNN_MAX_LAYERS], off_whh[B200RNN_MAX_LAYERS];
  int64_t off_bih[B200RNN_MAX_LAYERS], off_bhh[B200RNN_MAX_LAYERS];
  int I, T, B, L, BT, training;
};

struct BwdArgs {
  const float* x;
  const float* params;
  const float* h_seq;
  const float* c_seq;
  const float* g_seq;
  const int64_t* labels;  // nullable
  const float* d_out;     // nullable, [B,T,H]
  const float* d_hn;      // nullable, [L][B][H]
  const float* d_cn;      // nullable, [L][B][H]
  float* rows;            // [nCTA][row_floats]
  float* dxbuf;           // [2][T][B][H]
  int64_t off_wih[B200RNN_MAX_LAYERS], off_whh[B200RNN_MAX_LAYERS];
  int64_t off_bih[B200RNN_MAX_LAYERS], off_bhh[B200RNN_MAX_LAYERS];
  int64_t off_fcw, off_fcb, n_params, row_floats;
  float loss_scale;
  int I, T, B, L, BT, C;
};

// Column of operand element k inside a staged row: each thread-slice of KPT columns is padded by 4
// floats when KPT is a multiple of 8, so the KQ lanes of a hidden unit (which read the same row at
// different slices) hit different banks with their 128-bit loads.
template <int KPT>
__device__ __forceinline__ constexpr int vcol(int k) { return (KPT % 8 == 0) ? (k / KPT) * (KPT + 4) + (k % KPT) : k; }
// Position of gate-gradient element j' (= unit*4 + gate) inside a dG row: the JQ slices read by the
// B mapping are padded by 4 floats each; rows are DG_LD = 4H + 24 floats apart (== 24 mod 32 banks)
template <int JR>
__device__ __forceinline__ constexpr int dgcol(int jp) { return (jp / JR) * (JR + 4) + (jp % JR); }

template <int KQ>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = KQ / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// s[si][q]: partial gate sums (q = gate) of sample si held by each of the KQ lanes of a hidden
// unit.  On return lane kq holds in p[q] the full sums of sample kq.
template <int KQ>
__device__ __forceinline__ void group_transpose_reduce(float (&s)[KQ][4], int kq, float (&p)[4]) {
  static_assert(KQ == 4, "written for 4 lanes per hidden unit");
  const bool hi = (kq & 2) != 0, lo = (kq & 1) != 0;
  float k0[4], k1[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float send0 = hi ? s[0][q] : s[2][q];
    const float send1 = hi ? s[1][q] : s[3][q];
    k0[q] = (hi ? s[2][q] : s[0][q]) + __shfl_xor_sync(0xffffffffu, send0, 2);
    k1[q] = (hi ? s[3][q] : s[1][q]) + __shfl_xor_sync(0xffffffffu, send1, 2);
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float send = lo ? k0[q] : k1[q];
    p[q] = (lo ? k1[q] : k0[q]) + __shfl_xor_sync(0xffffffffu, send, 1);
  }
}

// ---------------------------------------------------------------------------
// Forward, one layer.  vs = [2][BTMAX][VS_LD] operand staging (double buffered).
// ---------------------------------------------------------------------------
template <int H, int KQ, int KPT, int NG, int LS>
__device__ __forceinline__ void fwd_layer(
    const float* inp /* may be h_seq of the layer below, written earlier by this CTA: no __restrict__ */,
    int64_t in_sb, int64_t in_st, int in_dim,
    const float* __restrict__ w_ih, const float* __restrict__ w_hh,
    const float* __restrict__ b_ih, const float* __restrict__ b_hh,
    float* h_seq, float* c_seq, float* g_seq,
    int T, int B, int b0, int BTc, int training, float* vs) {
  constexpr int NT = H * KQ;
  constexpr int KPAD = KQ * KPT;
  constexpr int INP = KPAD - H;           // padded input width of this layer
  constexpr int BTMAX = NG * KQ;
  constexpr int VS_LD = 2 * H + 4 * KQ + 4;   // row stride of vs (max KPAD = 2H plus slice padding)
  constexpr int NPF = (BTMAX * INP + NT - 1) / NT;
  static_assert(INP > 0 && INP % 4 == 0 && KPT % 4 == 0, "bad K split");

  const int tid = threadIdx.x;
  const int u = tid / KQ;
  const int kq = tid % KQ;

  // --- weights -> registers (resident for the whole sequence) ---
  float w[4][KPT];
  float bias[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const int r = g * H + u;
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
      const int k = kq * KPT + j;
      float v;
      if (k < INP) v = (k < in_dim) ? w_ih[(int64_t)r * in_dim + k] : 0.0f;
      else v = w_hh[(int64_t)r * H + (k - INP)];
      w[g][j] = v;
    }
    bias[g] = b_ih[r] + b_hh[r];
  }

  float pf[NPF];
  auto load_inputs = [&](int t) {
#pragma unroll
    for (int i = 0; i < NPF; ++i) {
      const int idx = tid + i * NT;
      const int b = idx / INP, k = idx % INP;
      float v = 0.0f;
      if (idx < BTMAX * INP && b < BTc && k < in_dim)
        v = inp[(int64_t)(b0 + b) * in_sb + (int64_t)t * in_st + k];
      pf[i] = v;
    }
  };
  auto store_inputs = [&](int buf) {
#pragma unroll
    for (int i = 0; i < NPF; ++i) {
      const int idx = tid + i * NT;
      const int b = idx / INP, k = idx % INP;
      if (idx < BTMAX * INP) vs[(buf * BTMAX + b) * VS_LD + vcol<KPT>(k)] = pf[i];
    }
  };

  load_inputs(0);
  // all rows of both buffers start finite (rows >= BTc are computed but never used)
  __syncthreads();
  for (int idx = tid; idx < 2 * BTMAX * VS_LD; idx += NT) vs[idx] = 0.0f;
  __syncthreads();
  store_inputs(0);
  float c_reg[NG];
#pragma unroll
  for (int g = 0; g < NG; ++g) c_reg[g] = 0.0f;
  __syncthreads();

  for (int t = 0; t < T; ++t) {
    const int cur = t & 1, nxt = cur ^ 1;
    if (t + 1 < T) load_inputs(t + 1);
#pragma unroll
    for (int g = 0; g < NG; ++g) {
      // branch-free: the KQ samples of a group form 4*KQ independent FMA chains (ILP);
      // rows past the tile end hold zeros and their results are simply not used.
      float s[KQ][4];
#pragma unroll
      for (int si = 0; si < KQ; ++si) {
        const int b = g * KQ + si;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        const float4* vp = reinterpret_cast<const float4*>(&vs[(cur * BTMAX + b) * VS_LD + vcol<KPT>(kq * KPT)]);
        const bool live = (g < NG - 1) || (si < LS);   // compile-time: the tile holds (NG-1)*KQ + LS samples
#pragma unroll
        for (int j4 = 0; j4 < (live ? KPT / 4 : 0); ++j4) {
          const float4 v = vp[j4];
          a0 = fmaf(w[0][4 * j4 + 0], v.x, a0); a1 = fmaf(w[1][4 * j4 + 0], v.x, a1);
          a2 = fmaf(w[2][4 * j4 + 0], v.x, a2); a3 = fmaf(w[3][4 * j4 + 0], v.x, a3);
          a0 = fmaf(w[0][4 * j4 + 1], v.y, a0); a1 = fmaf(w[1][4 * j4 + 1], v.y, a1);
          a2 = fmaf(w[2][4 * j4 + 1], v.y, a2); a3 = fmaf(w[3][4 * j4 + 1], v.y, a3);
          a0 = fmaf(w[0][4 * j4 + 2], v.z, a0); a1 = fmaf(w[1][4 * j4 + 2], v.z, a1);
          a2 = fmaf(w[2][4 * j4 + 2], v.z, a2); a3 = fmaf(w[3][4 * j4 + 2], v.z, a3);
          a0 = fmaf(w[0][4 * j4 + 3], v.w, a0); a1 = fmaf(w[1][4 * j4 + 3], v.w, a1);
          a2 = fmaf(w[2][4 * j4 + 3], v.w, a2); a3 = fmaf(w[3][4 * j4 + 3], v.w, a3);
        }
        s[si][0] = a0; s[si][1] = a1; s[si][2] = a2; s[si][3] = a3;
      }
      // transpose-reduce over the KQ lanes of this hidden unit: lane kq ends with the 4 gate
      // sums of sample g*KQ + kq (12 shuffles instead of 32 for a full butterfly)
      float p[4];
      group_transpose_reduce<KQ>(s, kq, p);
      const int b = g * KQ + kq;
      if (b < BTc) {
        const float ig = sigmoid_f(p[0] + bias[0]);
        const float fg = sigmoid_f(p[1] + bias[1]);
        const float gg = tanh_f(p[2] + bias[2]);
        const float og = sigmoid_f(p[3] + bias[3]);
        const float c = fmaf(fg, c_reg[g], ig * gg);
        const float h = og * tanh_f(c);
        c_reg[g] = c;
        vs[(nxt * BTMAX + b) * VS_LD + vcol<KPT>(INP + u)] = h;
        const int64_t e = ((int64_t)t * B + (b0 + b)) * H + u;
        h_seq[e] = h;
        if (training) {
          c_seq[e] = c;
          reinterpret_cast<float4*>(g_seq)[e] = make_float4(ig, fg, gg, og);
        }
      }
    }
    if (t + 1 < T) store_inputs(nxt);
    __syncthreads();
  }
}

template <int H, int KQ, int KPT0, int NG, int LS>
__global__ void __launch_bounds__(H* KQ) lstm_fwd_kernel(const FwdArgs a) {
  constexpr int KPT1 = 2 * H / KQ;
  constexpr int BTMAX = NG * KQ;
  constexpr int VS_LD = 2 * H + 4 * KQ + 4;
  __shared__ __align__(16) float vs[2 * BTMAX * VS_LD];
  const int b0 = blockIdx.x * a.BT;
  const int BTc = min(a.BT, a.B - b0);
  if (BTc <= 0) return;
  const int64_t lsz = (int64_t)a.T * a.B * H;
  fwd_layer<H, KQ, KPT0, NG, LS>(a.x, (int64_t)a.T * a.I, a.I, a.I,
                             a.params + a.off_wih[0], a.params + a.off_whh[0],
                             a.params + a.off_bih[0], a.params + a.off_bhh[0],
                             a.h_seq, a.c_seq, a.g_seq, a.T, a.B, b0, BTc, a.training, vs);
  for (int l = 1; l < a.L; ++l) {
    __syncthreads();  // layer l-1's h_seq tile (written by this CTA) is now visible
    fwd_layer<H, KQ, KPT1, NG, LS>(a.h_seq + (l - 1) * lsz, H, (int64_t)a.B * H, H,
                               a.params + a.off_wih[l], a.params + a.off_whh[l],
                               a.params + a.off_bih[l], a.params + a.off_bhh[l],
                               a.h_seq + l * lsz, a.c_seq + l * lsz, a.g_seq + 4 * l * lsz,
                               a.T, a.B, b0, BTc, a.training, vs);
  }
}

// Per-phase clock64 stamps of one backward step (profiling aid, read with b200rnn_bwd_prof()).
__device__ long long g_bwd_prof[2][8];
#define BWD_STAMP(role, i) do { if (stamp) g_bwd_prof[role][i] = clock64(); } while (0)

// ---------------------------------------------------------------------------
// Backward, one layer.  Warp-specialised: the CTA has 2*NT threads.
//   role 0 ("A", threads 0..NT-1, mapping tid = u*KQ + kq): gate gradients (phase A) and the
//           dW += dG^T [x_t | h_{t-1}] accumulation in registers (phase B2);
//   role 1 ("B", threads NT..2NT-1, mapping tid = k*JQ + jq): dv = dG x [W_ih | W_hh] with the
//           W^T slice in registers (phase B1: dh for step t-1 — the serial chain — and dx for the
//           layer below), plus the prefetch/staging of the next step's operands.
// B1 and B2 run concurrently on different warps; two __syncthreads per step.
// ---------------------------------------------------------------------------
template <int H, int KQ, int KPT, int NG, int BTU, bool HAS_DX>
__device__ __forceinline__ void bwd_layer(
    const float* __restrict__ inp, int64_t in_sb, int64_t in_st, int in_dim,
    const float* __restrict__ w_ih, const float* __restrict__ w_hh,
    const float* __restrict__ h_seq, const float* __restrict__ c_seq, const float* __restrict__ g_seq,
    const float* dabove /* may be dxbuf written earlier by this CTA: no __restrict__ */,
    int64_t da_sb, int64_t da_st,
    const float* __restrict__ d_cn, float* dx_out,
    float* row, int64_t off_wih, int64_t off_whh, int64_t off_bih, int64_t off_bhh,
    int T, int B, int b0, int BTc,
    float* vs, float* dG_s, float* dhrec_s) {
  constexpr int NT = H * KQ;
  constexpr int KPAD = KQ * KPT;
  constexpr int INP = KPAD - H;
  constexpr int BTMAX = NG * KQ;
  constexpr int VS_LD = 2 * H + 4 * KQ + 4;
  constexpr int DG_LD = 4 * H + 24;            // dG row stride (floats)
  constexpr int G4 = 4 * H;                    // gate rows
  constexpr int NOUT = HAS_DX ? KPAD : H;      // columns of dG x [W_ih|W_hh] we need
  constexpr int KOFF = HAS_DX ? 0 : INP;
  constexpr int JQ = NT / NOUT;                // gate-row split of the B mapping
  constexpr int JR = G4 / JQ;                  // gate rows per B-thread
  constexpr int NPFX = (BTMAX * INP + NT - 1) / NT;
  constexpr int NPFH = (BTMAX * H + NT - 1) / NT;
  static_assert(NT % NOUT == 0 && JQ >= 1 && (JQ & (JQ - 1)) == 0 && JR % 4 == 0, "bad B mapping");
  static_assert(BTU <= BTMAX && BTU > BTMAX - KQ, "BTU must fall in the last sample group");
  static_assert(!HAS_DX || INP == H, "dx only for stacked layers (input width == H)");

  // all rows of the staging buffers start at zero: the sample loops below are compile-time (BTMAX
  // iterations, full ILP) and rows past the tile end must contribute nothing
  for (int idx = threadIdx.x; idx < 2 * BTMAX * VS_LD; idx += 2 * NT) vs[idx] = 0.0f;
  for (int idx = threadIdx.x; idx < BTMAX * DG_LD; idx += 2 * NT) dG_s[idx] = 0.0f;
  __syncthreads();

  const bool roleA = threadIdx.x < NT;
  const int tid = roleA ? threadIdx.x : threadIdx.x - NT;   // role-local id
  const int u = tid / KQ, kq = tid % KQ;       // A mapping
  const int kB = tid / JQ, jq = tid % JQ;      // B mapping

  if (roleA) {
    // =========================== role A ===========================
    float dW[4][KPT];
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
      for (int j = 0; j < KPT; ++j) dW[g][j] = 0.0f;
    float dbacc[4] = {0.f, 0.f, 0.f, 0.f};
    float dc_reg[NG], c_t[NG], pf_cprev[NG], pf_da[NG];
    float4 pf_g[NG];

    auto load_step = [&](int t) {  // stash values phase A of step t needs
#pragma unroll
      for (int g = 0; g < NG; ++g) {
        const int b = g * KQ + kq;
        float4 gv = make_float4(0.f, 0.f, 0.f, 0.f);
        float cp = 0.f, da = 0.f;
        if (b < BTc) {
          const int64_t e = ((int64_t)t * B + (b0 + b)) * H + u;
          gv = reinterpret_cast<const float4*>(g_seq)[e];
          if (t > 0) cp = c_seq[e - (int64_t)B * H];
          if (dabove != nullptr) da = dabove[(int64_t)(b0 + b) * da_sb + (int64_t)t * da_st + u];
        }
        pf_g[g] = gv; pf_cprev[g] = cp; pf_da[g] = da;
      }
    };
#pragma unroll
    for (int g = 0; g < NG; ++g) {
      const int b = g * KQ + kq;
      float ct = 0.f, dc = 0.f;
      if (b < BTc) {
        const int64_t e = ((int64_t)(T - 1) * B + (b0 + b)) * H + u;
        ct = c_seq[e];
        if (d_cn != nullptr) dc = d_cn[(int64_t)(b0 + b) * H + u];
      }
      c_t[g] = ct; dc_reg[g] = dc;
    }
    load_step(T - 1);
    __syncthreads();  // (S0) operands of step T-1 staged by role B; caller's dhrec_s init visible

    for (int t = T - 1; t >= 0; --t) {
      const int cur = t & 1;
      const bool stamp = HAS_DX && t == T / 2 && blockIdx.x == 0 && threadIdx.x == 0;
      BWD_STAMP(0, 0);
      float4 gv[NG]; float cprev[NG], da[NG];
#pragma unroll
      for (int g = 0; g < NG; ++g) { gv[g] = pf_g[g]; cprev[g] = pf_cprev[g]; da[g] = pf_da[g]; }
      if (t > 0) load_step(t - 1);
      // ---- phase A: gate gradients for my (sample, unit) slots, branch-free so the NG slots'
      // MUFU / FMA chains interleave (slots past the tile end see zero inputs and produce zeros) ----
#pragma unroll
      for (int g = 0; g < NG; ++g) {
        const int b = g * KQ + kq;
        const float ig = gv[g].x, fg = gv[g].y, gg = gv[g].z, og = gv[g].w;
        const float dh = dhrec_s[b * H + u] + da[g];
        const float tc = tanh_f(c_t[g]);
        const float dct = fmaf(dh * og, 1.0f - tc * tc, dc_reg[g]);
        const float dpi = dct * gg * ig * (1.0f - ig);
        const float dpf = dct * cprev[g] * fg * (1.0f - fg);
        const float dpg = dct * ig * (1.0f - gg * gg);
        const float dpo = dh * tc * og * (1.0f - og);
        dc_reg[g] = dct * fg;
        c_t[g] = cprev[g];
        *reinterpret_cast<float4*>(&dG_s[b * DG_LD + dgcol<JR>(u * 4)]) = make_float4(dpi, dpf, dpg, dpo);
        dbacc[0] += dpi; dbacc[1] += dpf; dbacc[2] += dpg; dbacc[3] += dpo;
      }
      BWD_STAMP(0, 1);
      __syncthreads();  // (S1) dG_s complete
      BWD_STAMP(0, 2);
      // ---- phase B2: dW += dG^T [x_t | h_{t-1}] ----
#pragma unroll
      for (int b = 0; b < BTU; ++b) {
        const float4 d = *reinterpret_cast<const float4*>(&dG_s[b * DG_LD + dgcol<JR>(u * 4)]);
        const float4* vp = reinterpret_cast<const float4*>(&vs[(cur * BTMAX + b) * VS_LD + vcol<KPT>(kq * KPT)]);
#pragma unroll
        for (int j4 = 0; j4 < KPT / 4; ++j4) {
          const float4 v = vp[j4];
          dW[0][4 * j4 + 0] = fmaf(d.x, v.x, dW[0][4 * j4 + 0]); dW[0][4 * j4 + 1] = fmaf(d.x, v.y, dW[0][4 * j4 + 1]);
          dW[0][4 * j4 + 2] = fmaf(d.x, v.z, dW[0][4 * j4 + 2]); dW[0][4 * j4 + 3] = fmaf(d.x, v.w, dW[0][4 * j4 + 3]);
          dW[1][4 * j4 + 0] = fmaf(d.y, v.x, dW[1][4 * j4 + 0]); dW[1][4 * j4 + 1] = fmaf(d.y, v.y, dW[1][4 * j4 + 1]);
          dW[1][4 * j4 + 2] = fmaf(d.y, v.z, dW[1][4 * j4 + 2]); dW[1][4 * j4 + 3] = fmaf(d.y, v.w, dW[1][4 * j4 + 3]);
          dW[2][4 * j4 + 0] = fmaf(d.z, v.x, dW[2][4 * j4 + 0]); dW[2][4 * j4 + 1] = fmaf(d.z, v.y, dW[2][4 * j4 + 1]);
          dW[2][4 * j4 + 2] = fmaf(d.z, v.z, dW[2][4 * j4 + 2]); dW[2][4 * j4 + 3] = fmaf(d.z, v.w, dW[2][4 * j4 + 3]);
          dW[3][4 * j4 + 0] = fmaf(d.w, v.x, dW[3][4 * j4 + 0]); dW[3][4 * j4 + 1] = fmaf(d.w, v.y, dW[3][4 * j4 + 1]);
          dW[3][4 * j4 + 2] = fmaf(d.w, v.z, dW[3][4 * j4 + 2]); dW[3][4 * j4 + 3] = fmaf(d.w, v.w, dW[3][4 * j4 + 3]);
        }
      }
      BWD_STAMP(0, 3);
      __syncthreads();  // (S2) dhrec_s / next operands ready (role B); dG_s, vs[cur] free
      BWD_STAMP(0, 4);
    }
    // ---- emit this CTA's partial gradients (torch parameter layout) ----
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int r = g * H + u;
#pragma unroll
      for (int j = 0; j < KPT; ++j) {
        const int k = kq * KPT + j;
        if (k < INP) { if (k < in_dim) row[off_wih + (int64_t)r * in_dim + k] = dW[g][j]; }
        else row[off_whh + (int64_t)r * H + (k - INP)] = dW[g][j];
      }
      const float db = group_sum<KQ>(dbacc[g]);
      if (kq == 0) { row[off_bih + r] = db; row[off_bhh + r] = db; }
    }
  } else {
    // =========================== role B ===========================
    // W^T slice: wt[jj] = Wcat[row(j')][KOFF + kB], j' = jq*JR + jj = u'*4 + g'
    float wt[JR];
#pragma unroll
    for (int jj = 0; jj < JR; ++jj) {
      const int jp = jq * JR + jj;
      const int up = jp >> 2, gp = jp & 3;
      const int r = gp * H + up;
      const int kk = KOFF + kB;
      float v;
      if (kk < INP) v = (kk < in_dim) ? w_ih[(int64_t)r * in_dim + kk] : 0.0f;
      else v = w_hh[(int64_t)r * H + (kk - INP)];
      wt[jj] = v;
    }
    float pfx[NPFX], pfh[NPFH];
    auto load_operands = [&](int t) {  // [x_t | h_{t-1}] of my tile
#pragma unroll
      for (int i = 0; i < NPFX; ++i) {
        const int idx = tid + i * NT;
        const int b = idx / INP, k = idx % INP;
        float v = 0.0f;
        if (idx < BTMAX * INP && b < BTc && k < in_dim)
          v = inp[(int64_t)(b0 + b) * in_sb + (int64_t)t * in_st + k];
        pfx[i] = v;
      }
#pragma unroll
      for (int i = 0; i < NPFH; ++i) {
        const int idx = tid + i * NT;
        const int b = idx / H, k = idx % H;
        float v = 0.0f;
        if (idx < BTMAX * H && b < BTc && t > 0)
          v = h_seq[((int64_t)(t - 1) * B + (b0 + b)) * H + k];
        pfh[i] = v;
      }
    };
    auto store_operands = [&](int buf) {
#pragma unroll
      for (int i = 0; i < NPFX; ++i) {
        const int idx = tid + i * NT;
        if (idx < BTMAX * INP) vs[(buf * BTMAX + idx / INP) * VS_LD + vcol<KPT>(idx % INP)] = pfx[i];
      }
#pragma unroll
      for (int i = 0; i < NPFH; ++i) {
        const int idx = tid + i * NT;
        if (idx < BTMAX * H) vs[(buf * BTMAX + idx / H) * VS_LD + vcol<KPT>(INP + (idx % H))] = pfh[i];
      }
    };
    load_operands(T - 1);
    store_operands((T - 1) & 1);
    __syncthreads();  // (S0)

    for (int t = T - 1; t >= 0; --t) {
      const int nxt = (t & 1) ^ 1;
      const bool stamp = HAS_DX && t == T / 2 && blockIdx.x == 0 && threadIdx.x == NT;
      BWD_STAMP(1, 0);
      if (t > 0) load_operands(t - 1);   // global loads in flight across the barrier
      BWD_STAMP(1, 1);
      __syncthreads();  // (S1) dG_s complete
      BWD_STAMP(1, 2);
      // ---- phase B1: dv = dG x [W_ih | W_hh]  (dx for the layer below, dh for step t-1) ----
      // all dot products first (BTMAX x 4 independent FMA chains, loads hoistable), then the
      // lane reductions back to back, then the stores
      float res[BTU];
#pragma unroll
      for (int b = 0; b < BTU; ++b) {
        const float4* dp = reinterpret_cast<const float4*>(&dG_s[b * DG_LD + jq * (JR + 4)]);
        float acc0 = 0.0f, acc1 = 0.0f, acc2 = 0.0f, acc3 = 0.0f;
#pragma unroll
        for (int j4 = 0; j4 < JR / 4; ++j4) {
          const float4 d = dp[j4];
          acc0 = fmaf(wt[4 * j4 + 0], d.x, acc0); acc1 = fmaf(wt[4 * j4 + 1], d.y, acc1);
          acc2 = fmaf(wt[4 * j4 + 2], d.z, acc2); acc3 = fmaf(wt[4 * j4 + 3], d.w, acc3);
        }
        res[b] = (acc0 + acc1) + (acc2 + acc3);
      }
#pragma unroll
      for (int b = 0; b < BTU; ++b) res[b] = group_sum<JQ>(res[b]);
      if (jq == 0) {
        const int kk = KOFF + kB;
#pragma unroll
        for (int b = 0; b < BTU; ++b) {
          if (b < BTc) {
            if (HAS_DX && kk < INP) dx_out[((int64_t)t * B + (b0 + b)) * H + kk] = res[b];
            else dhrec_s[b * H + (kk - INP)] = res[b];
          }
        }
      }
      BWD_STAMP(1, 3);
      if (t > 0) store_operands(nxt);
      BWD_STAMP(1, 4);
      __syncthreads();  // (S2)
      BWD_STAMP(1, 5);
    }
  }
}

template <int H, int KQ, int KPT0, int NG, int BTU>
__global__ void __launch_bounds__(2 * H * KQ, 1) lstm_bwd_kernel(const BwdArgs a) {
  constexpr int NT2 = 2 * H * KQ;   // all threads (both roles)
  constexpr int KPT1 = 2 * H / KQ;
  constexpr int BTMAX = NG * KQ;
  constexpr int VS_LD = 2 * H + 4 * KQ + 4;
  constexpr int DG_LD = 4 * H + 24;
  constexpr int CMAX = 32;
  __shared__ __align__(16) float vs[2 * BTMAX * VS_LD];
  __shared__ __align__(16) float dG_s[BTMAX * DG_LD];
  __shared__ __align__(16) float dhrec_s[BTMAX * H];
  __shared__ float lg_s[BTMAX * CMAX];   // logits, then dlogits

  const int tid = threadIdx.x;
  const int b0 = blockIdx.x * a.BT;
  const int BTc = min(a.BT, a.B - b0);
  if (BTc <= 0) return;
  float* row = a.rows + (int64_t)blockIdx.x * a.row_floats;
  const int64_t lsz = (int64_t)a.T * a.B * H;
  const int L = a.L;
  // rows past the tile end must hold finite values: phase A is branch-free
  for (int idx = tid; idx < BTMAX * H; idx += NT2) dhrec_s[idx] = 0.0f;
  __syncthreads();

  // ---- top of the network: fused head + CE, or external d_hn ----
  const bool fused_head = (a.labels != nullptr && a.C > 0);
  if (fused_head) {
    const int C = a.C;
    const float* fcw = a.params + a.off_fcw;
    const float* fcb = a.params + a.off_fcb;
    float* hT = dG_s;  // reuse: [BTMAX][H]
    for (int idx = tid; idx < BTc * H; idx += NT2)
      hT[idx] = a.h_seq[(L - 1) * lsz + ((int64_t)(a.T - 1) * a.B + (b0 + idx / H)) * H + (idx % H)];
    __syncthreads();
    for (int idx = tid; idx < BTc * C; idx += NT2) {
      const int b = idx / C, c = idx % C;
      float acc = fcb[c];
#pragma unroll 8
      for (int k = 0; k < H; ++k) acc = fmaf(fcw[c * H + k], hT[b * H + k], acc);
      lg_s[b * CMAX + c] = acc;
    }
    __syncthreads();
    float loss_b = 0.f, corr_b = 0.f;
    if (tid < BTc) {
      const int b = tid;
      const int y = (int)a.labels[b0 + b];
      float m = lg_s[b * CMAX];
      int arg = 0;
      for (int c = 1; c < C; ++c) { const float v = lg_s[b * CMAX + c]; if (v > m) { m = v; arg = c; } }
      float se = 0.f;
      for (int c = 0; c < C; ++c) se += expf(lg_s[b * CMAX + c] - m);
      const float lse = m + logf(se);
      loss_b = lse - lg_s[b * CMAX + y];
      corr_b = (arg == y) ? 1.f : 0.f;
      for (int c = 0; c < C; ++c) {
        const float p = expf(lg_s[b * CMAX + c] - lse);
        lg_s[b * CMAX + c] = (p - (c == y ? 1.f : 0.f)) * a.loss_scale;
      }
    }
    if (tid < 32) {  // BTMAX <= 16 < 32: all live values sit in warp 0
      loss_b = warp_sum(loss_b);
      corr_b = warp_sum(corr_b);
      if (tid == 0) { row[a.n_params] = loss_b; row[a.n_params + 1] = corr_b; }
    }
    __syncthreads();
    for (int idx = tid; idx < BTc * H; idx += NT2) {
      const int b = idx / H, k = idx % H;
      float acc = 0.f;
      for (int c = 0; c < C; ++c) acc = fmaf(lg_s[b * CMAX + c], fcw[c * H + k], acc);
      if (a.d_hn != nullptr) acc += a.d_hn[((int64_t)(L - 1) * a.B + (b0 + b)) * H + k];
      dhrec_s[idx] = acc;
    }
    for (int idx = tid; idx < C * H; idx += NT2) {
      const int c = idx / H, k = idx % H;
      float acc = 0.f;
      for (int b = 0; b < BTc; ++b) acc = fmaf(lg_s[b * CMAX + c], hT[b * H + k], acc);
      row[a.off_fcw + idx] = acc;
    }
    for (int c = tid; c < C; c += NT2) {
      float acc = 0.f;
      for (int b = 0; b < BTc; ++b) acc += lg_s[b * CMAX + c];
      row[a.off_fcb + c] = acc;
    }
    __syncthreads();  // hT (aliases dG_s) fully consumed before phase A overwrites it
  } else {
    for (int idx = tid; idx < BTc * H; idx += NT2)
      dhrec_s[idx] = (a.d_hn != nullptr) ? a.d_hn[((int64_t)(L - 1) * a.B + (b0 + idx / H)) * H + (idx % H)] : 0.f;
    if (a.C > 0) {
      for (int idx = tid; idx < a.C * H; idx += NT2) row[a.off_fcw + idx] = 0.f;
      for (int c = tid; c < a.C; c += NT2) row[a.off_fcb + c] = 0.f;
    }
    if (tid == 0) { row[a.n_params] = 0.f; row[a.n_params + 1] = 0.f; }
  }

  for (int l = L - 1; l >= 1; --l) {
    const float* dabove; int64_t da_sb, da_st;
    if (l == L - 1) { dabove = a.d_out; da_sb = (int64_t)a.T * H; da_st = H; }
    else { dabove = a.dxbuf + ((l + 1) & 1) * lsz; da_sb = H; da_st = (int64_t)a.B * H; }
    bwd_layer<H, KQ, KPT1, NG, BTU, true>(
        a.h_seq + (l - 1) * lsz, H, (int64_t)a.B * H, H,
        a.params + a.off_wih[l], a.params + a.off_whh[l],
        a.h_seq + l * lsz, a.c_seq + l * lsz, a.g_seq + 4 * l * lsz,
        dabove, da_sb, da_st,
        a.d_cn ? a.d_cn + (int64_t)l * a.B * H : nullptr,
        a.dxbuf + (l & 1) * lsz,
        row, a.off_wih[l], a.off_whh[l], a.off_bih[l], a.off_bhh[l],
        a.T, a.B, b0, BTc, vs, dG_s, dhrec_s);
    __syncthreads();  // dx tile of layer l visible; smem free for reuse
    for (int idx = tid; idx < BTc * H; idx += NT2)
      dhrec_s[idx] = (a.d_hn != nullptr) ? a.d_hn[((int64_t)(l - 1) * a.B + (b0 + idx / H)) * H + (idx % H)] : 0.f;
  }
  {
    const float* dabove; int64_t da_sb, da_st;
    if (L == 1) { dabove = a.d_out; da_sb = (int64_t)a.T * H; da_st = H; }
    else { dabove = a.dxbuf + (1 & 1) * lsz; da_sb = H; da_st = (int64_t)a.B * H; }
    bwd_layer<H, KQ, KPT0, NG, BTU, false>(
        a.x, (int64_t)a.T * a.I, a.I, a.I,
        a.params + a.off_wih[0], a.params + a.off_whh[0],
        a.h_seq, a.c_seq, a.g_seq,
        dabove, da_sb, da_st,
        a.d_cn, nullptr,
        row, a.off_wih[0], a.off_whh[0], a.off_bih[0], a.off_bhh[0],
        a.T, a.B, b0, BTc, vs, dG_s, dhrec_s);
  }
}

// ---------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------
constexpr int kH = 32, kKQ = 4;

// Rows per CTA (any size 1..16; kernels are instantiated per (groups, live samples in the last group));
// pick the tile that minimises the per-SM load when CTAs spread evenly with at most
// `slots` co-resident CTAs per SM (register-limited), smaller tile on ties (more warps per SM).
int choose_bt_fwd(int B) {
  // Measured (tools/kernel_bench.py): a forward step costs about one unit per SAMPLE GROUP resident on
  // the SM (reduce + activation + barrier overhead) plus ~6% of that per live sample, so partially
  // filled groups are only worth it when the whole tile is a single group.
  const int sms = num_sms(), slots = 4;
  const int cand[7] = {1, 2, 3, 4, 8, 12, 16};
  int best = 4;
  float best_cost = 1e30f;
  for (int i = 0; i < 7; ++i) {
    const int bt = cand[i];
    const int n = (B + bt - 1) / bt;
    const int per_sm = (n + sms - 1) / sms;
    const int waves = (per_sm + slots - 1) / slots;
    const int resident = waves == 1 ? per_sm : waves * slots;
    const float cost = resident * ((bt + kKQ - 1) / kKQ + 0.06f * bt);
    if (cost < best_cost) { best_cost = cost; best = bt; }
  }
  return best;
}
// Backward: the sample loops are compile-time over the tile (multiples of KQ), one CTA of 8 warps
// per SM (its register budget): minimise the per-SM load like the forward does.
int choose_bt(int B) {
  const int sms = num_sms();
  int best = 4, best_load = 1 << 30;
  const int cand[10] = {1, 2, 3, 4, 5, 6, 8, 10, 12, 16};   // instantiated tile sizes
  for (int i = 0; i < 10; ++i) {
    const int bt = cand[i];
    const int n = (B + bt - 1) / bt;
    const int load = ((n + sms - 1) / sms) * bt;
    if (load <= best_load) { best_load = load; best = bt; }   // ties: larger tile = a single wave
  }
  return best;
}

bool small_supported(const b200rnn_model_desc* d) {
  return d->cell == B200RNN_CELL_LSTM && d->hidden_size == kH && d->input_size >= 1 &&
         d->input_size <= 32 && d->num_layers >= 1 && d->num_layers <= B200RNN_MAX_LAYERS &&
         d->num_classes >= 0 && d->num_classes <= 32 && d->seq_len >= 1 && d->batch >= 1;
}

template <int KPT0, int NG, int LS>
int launch_fwd(const FwdArgs& a, int n_cta, cudaStream_t s) {
  lstm_fwd_kernel<kH, kKQ, KPT0, NG, LS><<<n_cta, kH * kKQ, 0, s>>>(a);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}
template <int KPT0, int NG, int BTU = 4 * NG>
int launch_bwd(const BwdArgs& a, int n_cta, cudaStream_t s) {
  lstm_bwd_kernel<kH, kKQ, KPT0, NG, BTU><<<n_cta, 2 * kH * kKQ, 0, s>>>(a);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}
// tile size -> instantiation (sample groups NG = ceil(BT / 4), compile-time sample count BT)
int dispatch_bwd(const BwdArgs& a, int n_cta, bool narrow, cudaStream_t s) {
#define B200_BWD_CASE(BT, NG) case BT: return narrow ? launch_bwd<12, NG, BT>(a, n_cta, s) : launch_bwd<16, NG, BT>(a, n_cta, s);
  switch (a.BT) {
    B200_BWD_CASE(1, 1) B200_BWD_CASE(2, 1) B200_BWD_CASE(3, 1) B200_BWD_CASE(4, 1) B200_BWD_CASE(5, 2)
    B200_BWD_CASE(6, 2) B200_BWD_CASE(8, 2) B200_BWD_CASE(10, 3) B200_BWD_CASE(12, 3) B200_BWD_CASE(16, 4)
  }
#undef B200_BWD_CASE
  set_error("lstm_bwd: no kernel instantiated for a tile of %d rows", a.BT);
  return B200RNN_EINVAL;
}

}  // namespace
}  // namespace b200rnn

using namespace b200rnn;

extern "C" int b200rnn_rnn_small_supported(const b200rnn_model_desc* d) {
  return (d != nullptr && small_supported(d)) ? 1 : 0;
}

extern "C" int b200rnn_rnn_small_tiling(const b200rnn_model_desc* d, int32_t* rows_per_cta, int32_t* n_cta) {
  B200_REQUIRE(d != nullptr && d->batch >= 1, B200RNN_EINVAL, "tiling: bad desc");
  const int bt = choose_bt(d->batch);
  if (rows_per_cta) *rows_per_cta = bt;
  if (n_cta) *n_cta = (d->batch + bt - 1) / bt;
  return B200RNN_OK;
}

extern "C" int64_t b200rnn_rnn_stash_floats(const b200rnn_model_desc* d) {
  if (!d) return 0;
  return (int64_t)d->num_layers * d->seq_len * d->batch * d->hidden_size * 6;
}

extern "C" int64_t b200rnn_grad_row_floats(const b200rnn_model_desc* d) {
  return d ? d->n_params + 2 : 0;
}

extern "C" int64_t b200rnn_rnn_bwd_scratch_floats(const b200rnn_model_desc* d) {
  if (!d) return 0;
  const int bt = choose_bt(d->batch);
  const int64_t n_cta = (d->batch + bt - 1) / bt;
  return n_cta * (d->n_params + 2) + 2ll * d->seq_len * d->batch * d->hidden_size;
}

extern "C" int b200rnn_lstm_fwd(const b200rnn_model_desc* d, const float* params, const float* x,
                                float* stash, int32_t training, void* stream) {
  B200_REQUIRE(d && params && x && stash, B200RNN_EINVAL, "lstm_fwd: null argument");
  B200_REQUIRE(small_supported(d), B200RNN_EUNSUPPORTED,
               "lstm_fwd: shape (cell=%d I=%d H=%d L=%d C=%d) not covered by the small-hidden family",
               d->cell, d->input_size, d->hidden_size, d->num_layers, d->num_classes);
  FwdArgs a;
  a.x = x; a.params = params;
  const int64_t lsz = (int64_t)d->num_layers * d->seq_len * d->batch * d->hidden_size;
  a.h_seq = stash; a.c_seq = stash + lsz; a.g_seq = stash + 2 * lsz;
  for (int l = 0; l < B200RNN_MAX_LAYERS; ++l) {
    a.off_wih[l] = d->off_wih[l]; a.off_whh[l] = d->off_whh[l];
    a.off_bih[l] = d->off_bih[l]; a.off_bhh[l] = d->off_bhh[l];
  }
  a.I = d->input_size; a.T = d->seq_len; a.B = d->batch; a.L = d->num_layers;
  a.BT = choose_bt_fwd(d->batch); a.training = training;
  const int n_cta = (a.B + a.BT - 1) / a.BT;
  const int ng = (a.BT + kKQ - 1) / kKQ;
  cudaStream_t s = (cudaStream_t)stream;
  const bool narrow = d->input_size <= 16;
  const int ls = a.BT - (ng - 1) * kKQ;
#define B200_FWD_LS(NG, LS) case LS: return narrow ? launch_fwd<12, NG, LS>(a, n_cta, s) : launch_fwd<16, NG, LS>(a, n_cta, s);
#define B200_DISPATCH_FWD(NG)                                                     \
  case NG: switch (ls) { B200_FWD_LS(NG, 1) B200_FWD_LS(NG, 2) B200_FWD_LS(NG, 3) B200_FWD_LS(NG, 4) } break;
  switch (ng) {
    B200_DISPATCH_FWD(1) B200_DISPATCH_FWD(2) B200_DISPATCH_FWD(3) B200_DISPATCH_FWD(4)
  }
#undef B200_FWD_LS
#undef B200_DISPATCH_FWD
  set_error("lstm_fwd: internal tiling error (ng=%d)", ng);
  return B200RNN_EINVAL;
}

extern "C" int b200rnn_lstm_bwd(const b200rnn_model_desc* d, const float* params, const float* x,
                                const float* stash, const int64_t* labels, float loss_scale,
                                const float* d_out, const float* d_hn, const float* d_cn,
                                float* scratch, void* stream) {
  B200_REQUIRE(d && params && x && stash && scratch, B200RNN_EINVAL, "lstm_bwd: null argument");
  B200_REQUIRE(small_supported(d), B200RNN_EUNSUPPORTED,
               "lstm_bwd: shape (cell=%d I=%d H=%d L=%d C=%d) not covered by the small-hidden family",
               d->cell, d->input_size, d->hidden_size, d->num_layers, d->num_classes);
  BwdArgs a;
  a.x = x; a.params = params;
  const int64_t lsz = (int64_t)d->num_layers * d->seq_len * d->batch * d->hidden_size;
  a.h_seq = stash; a.c_seq = stash + lsz; a.g_seq = stash + 2 * lsz;
  a.labels = labels; a.d_out = d_out; a.d_hn = d_hn; a.d_cn = d_cn;
  for (int l = 0; l < B200RNN_MAX_LAYERS; ++l) {
    a.off_wih[l] = d->off_wih[l]; a.off_whh[l] = d->off_whh[l];
    a.off_bih[l] = d->off_bih[l]; a.off_bhh[l] = d->off_bhh[l];
  }
  a.off_fcw = d->off_fcw; a.off_fcb = d->off_fcb; a.n_params = d->n_params;
  a.row_floats = d->n_params + 2;
  a.loss_scale = loss_scale;
  a.I = d->input_size; a.T = d->seq_len; a.B = d->batch; a.L = d->num_layers;
  a.C = d->num_classes;
  a.BT = choose_bt(d->batch);
  const int n_cta = (a.B + a.BT - 1) / a.BT;
  a.rows = scratch;
  a.dxbuf = scratch + (int64_t)n_cta * a.row_floats;
  const int ng = (a.BT + kKQ - 1) / kKQ;
  cudaStream_t s = (cudaStream_t)stream;
  const bool narrow = d->input_size <= 16;
  return dispatch_bwd(a, n_cta, narrow, s);
}


// ---------------------------------------------------------------------------
// Allocation-free backward (see include/b200rnn.h: b200rnn_lstm_bwd_static).
//
// Some hosts of this library cannot give the backward pass a scratch buffer: a plain C/C++
// program that links the C ABI without torch's caching allocator, a service that handed the
// whole GPU to its own memory pool at start-up (so cudaMalloc has nothing left and fails),
// or a latency-sensitive loop that must not call cudaMalloc/cudaFree because both serialise
// the device.  For them the per-CTA gradient rows and the inter-layer dx buffers live in the
// fixed array below instead of caller memory.  Cost: every GPU that loads the library keeps
// B200RNN_STATIC_SCRATCH_FLOATS * 4 bytes (16 MiB) reserved until the process exits.
// ---------------------------------------------------------------------------
static __device__ __align__(256) float g_bwd_static_scratch[B200RNN_STATIC_SCRATCH_FLOATS];

namespace b200rnn {
namespace {

constexpr int kMaxDevices = 64;
float* g_static_ptr[kMaxDevices] = {nullptr};  // device address of g_bwd_static_scratch, per device

// In host code `g_bwd_static_scratch` is only a handle for the runtime, not a device address;
// the real address differs per device and has to be asked for with that device current.
int static_scratch_for_current_device(float** out) {
  int dev = -1;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kMaxDevices, B200RNN_EINVAL, "lstm_bwd_static: device index %d out of range", dev);
  if (g_static_ptr[dev] == nullptr) {
    float* p = nullptr;
    B200_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_bwd_static_scratch));
    B200_REQUIRE(p != nullptr, B200RNN_ECUDA, "lstm_bwd_static: cudaGetSymbolAddress returned NULL");
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return B200RNN_OK;
}

}  // namespace
}  // namespace b200rnn

extern "C" int b200rnn_lstm_bwd_static(const b200rnn_model_desc* d, const float* params, const float* x,
                                       const float* stash, const int64_t* labels, float loss_scale,
                                       const float* d_out, const float* d_hn, const float* d_cn,
                                       float* grad_out, void* stream) {
  B200_REQUIRE(d && params && x && stash && grad_out, B200RNN_EINVAL, "lstm_bwd_static: null argument");
  B200_REQUIRE(small_supported(d), B200RNN_EUNSUPPORTED,
               "lstm_bwd_static: shape (cell=%d I=%d H=%d L=%d C=%d) not covered by the small-hidden family",
               d->cell, d->input_size, d->hidden_size, d->num_layers, d->num_classes);
  const int64_t need = b200rnn_rnn_bwd_scratch_floats(d);
  B200_REQUIRE(need <= (int64_t)B200RNN_STATIC_SCRATCH_FLOATS, B200RNN_ETOOBIG,
               "lstm_bwd_static: this shape needs %lld scratch floats but the built-in scratch holds %lld "
               "(B200RNN_STATIC_SCRATCH_FLOATS); use b200rnn_lstm_bwd with a caller-provided buffer",
               (long long)need, (long long)B200RNN_STATIC_SCRATCH_FLOATS);
  float* scratch = nullptr;
  int rc = static_scratch_for_current_device(&scratch);
  if (rc != B200RNN_OK) return rc;

  // Same launch as b200rnn_lstm_bwd, pointed at the built-in scratch.  Nothing in it has to start
  // at zero: every gradient-row slot and every dx element is written by the kernel before it is read.
  BwdArgs a;
  a.x = x; a.params = params;
  const int64_t lsz = (int64_t)d->num_layers * d->seq_len * d->batch * d->hidden_size;
  a.h_seq = stash; a.c_seq = stash + lsz; a.g_seq = stash + 2 * lsz;
  a.labels = labels; a.d_out = d_out; a.d_hn = d_hn; a.d_cn = d_cn;
  for (int l = 0; l < B200RNN_MAX_LAYERS; ++l) {
    a.off_wih[l] = d->off_wih[l]; a.off_whh[l] = d->off_whh[l];
    a.off_bih[l] = d->off_bih[l]; a.off_bhh[l] = d->off_bhh[l];
  }
  a.off_fcw = d->off_fcw; a.off_fcb = d->off_fcb; a.n_params = d->n_params;
  a.row_floats = d->n_params + 2;
  a.loss_scale = loss_scale;
  a.I = d->input_size; a.T = d->seq_len; a.B = d->batch; a.L = d->num_layers;
  a.C = d->num_classes;
  a.BT = choose_bt(d->batch);
  const int n_cta = (a.B + a.BT - 1) / a.BT;
  a.rows = scratch;
  a.dxbuf = scratch + (int64_t)n_cta * a.row_floats;
  const int ng = (a.BT + kKQ - 1) / kKQ;
  cudaStream_t s = (cudaStream_t)stream;
  const bool narrow = d->input_size <= 16;
  rc = dispatch_bwd(a, n_cta, narrow, s);
  if (rc != B200RNN_OK) return rc;
  // fixed-order sum of the rows into the caller's gradient vector (a kernel launch, no allocation)
  return b200rnn_reduce_rows(scratch, n_cta, a.row_floats, grad_out, 0, stream);
}

extern "C" int b200rnn_bwd_prof(long long* out16) {
  B200_REQUIRE(out16 != nullptr, B200RNN_EINVAL, "bwd_prof: null");
  B200_CHECK_CUDA(cudaMemcpyFromSymbol(out16, g_bwd_prof, sizeof(long long) * 16));
  return B200RNN_OK;
}
