// Persistent recurrent kernels for the large-hidden LSTM path: ONE launch walks all T timesteps of a
// layer, with the recurrent weights sharded across the SMs' shared memory for the whole sequence.
//
// The per-timestep formulation (rnn_large.cu) re-reads W_hh (8 MiB bf16 at H = 1024) from L2 every
// step and pays two launches per step.  Here CTA j owns hidden units [8j, 8j+8):
//
//   forward   its 32 rows of W_hh (4 gates x 8 units, all K = H columns) stay in shared memory; per
//             step it streams h_{t-1} [B, H] (bf16, written by all CTAs the step before) through a
//             TMA ring, accumulates  h_{t-1} x W_shard^T  -> D[B, 32] in TMEM with tcgen05.mma, and
//             its epilogue warps add the precomputed input projection, apply the gates, keep c_t in
//             registers, and publish h_t (bf16 for the next step, fp32 + c_t + gate activations for
//             the backward stash);
//   backward  its 8 rows of W_hh^T (K = 4H columns) stay in shared memory; per step it streams
//             dG_{t+1} [B, 4H] (bf16), accumulates  dG_{t+1} x W_hh[:, units]  -> D[B, 8] = dh_rec,
//             and the epilogue does the cell backward for its units (dc carried in registers),
//             emitting dG_t fp32 (for the bulk dW / dX GEMMs afterwards) and bf16 (next step).
//
// The streamed operand never takes the tensor-map path: the epilogue that produces it writes it straight
// into the 128B-swizzled K-major tile image tcgen05 wants ([K chunk][row][64] with the 16-byte units
// XOR-ed by row % 8), double-buffered by step parity, so the consumers fetch it with plain 1-D bulk copies
// of 32 KiB issued from four warps (see the measurements quoted at the producers below).
//
// Steps are separated by a grid-wide barrier on one global counter (red.release / ld.acquire at
// gpu scope; the kernel is launched cooperatively so all CTAs are co-resident).  Only the bf16
// operand the next step consumes is stored before the arrive; the fp32 stash stores follow it, off
// the critical path, and the next step's pointwise inputs are prefetched before the wait.
// Every wait is bounded: a timeout raises an abort flag that drains all roles of all CTAs and is
// reported through b200rnn_persist_status().
//
// Replaces the time loop of ATen's lstm (src/motion/model.py:14) and of its autograd backward
// (src/motion/trainer/base.py:116) for BASELINE.json's scaled configuration (hidden 1024, T 512).
#include <cuda_bf16.h>
#include <cstdlib>
#include "../gemm/tc_common.cuh"

namespace b200rnn {
namespace {

using namespace tc;
typedef __nv_bfloat16 bf16;

constexpr int PK = 64;            // K elements per ring stage (one 128-byte swizzle row)
constexpr int MAX_STAGES = 32;    // ring depth is chosen at launch: as many stages as shared memory holds (<= nk)
constexpr int UC = 8;             // hidden units per CTA
constexpr long long kSpinBudget = 600000000LL;   // ~0.3 s at 2 GHz

__device__ int g_persist_error = 0;
// clock64 stamps of one mid-sequence step on CTA 0 (profiling aid, b200rnn_persist_prof()): [dir][role*8 + i]
__device__ long long g_persist_prof[2][24];
#define PSTAMP(i) do { if (stamp) g_persist_prof[BWD ? 1 : 0][i] = clock64(); } while (0)

struct PersistSync {
  unsigned int arrive;   // monotonic: +1 per CTA per completed step
  unsigned int abort;    // set by any CTA whose wait timed out
};

__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
// ---- thread-block-cluster helpers (2-D backward: partial sums exchanged through distributed shared memory) ----
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float x, float y, float z, float w) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote_release(uint32_t addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_acq_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// Bounded waits that also drain quickly once anybody has aborted.
__device__ __forceinline__ bool wait_mbar(uint64_t* bar, uint32_t parity, volatile int* s_abort, PersistSync* sync, int code) {
  if (mbar_try_wait(bar, parity)) return true;
  const long long t0 = clock64();
  for (;;) {
    if (mbar_try_wait(bar, parity)) return true;
    if (*s_abort) return false;
    if (clock64() - t0 > kSpinBudget) {
      *s_abort = 1;
      atomicExch(&sync->abort, 1u);
      atomicCAS(&g_persist_error, 0, code);
      return false;
    }
  }
}
__device__ __forceinline__ bool wait_mbar_cluster(uint64_t* bar, uint32_t parity, volatile int* s_abort, PersistSync* sync, int code) {
  if (mbar_try_wait_acq_cluster(bar, parity)) return true;
  const long long t0 = clock64();
  for (;;) {
    if (mbar_try_wait_acq_cluster(bar, parity)) return true;
    if (*s_abort) return false;
    if (clock64() - t0 > kSpinBudget) {
      *s_abort = 1;
      atomicExch(&sync->abort, 1u);
      atomicCAS(&g_persist_error, 0, code);
      return false;
    }
  }
}
__device__ __forceinline__ bool wait_grid(PersistSync* sync, unsigned target, volatile int* s_abort, int code) {
  const long long t0 = clock64();
  for (;;) {
    if (ld_acquire_gpu(&sync->arrive) >= target) return true;
    if (*s_abort) return false;
    if (*(volatile unsigned*)&sync->abort) { *s_abort = 1; return false; }
    if (clock64() - t0 > kSpinBudget) {
      *s_abort = 1;
      atomicExch(&sync->abort, 1u);
      atomicCAS(&g_persist_error, 0, code);
      return false;
    }
  }
}

// Address of the 16-byte unit holding elements [k, k+8) of row `row` inside the swizzled tile image:
// chunk k/64 is a [BM][64] bf16 tile whose 16-byte units are XOR-ed with row % 8 (the 128B swizzle).
__device__ __forceinline__ bf16* stream_unit(bf16* stream, int buf, int nk, int bm, int row, int k) {
  const int chunk = k >> 6, j = (k & 63) >> 3;
  return stream + (((size_t)buf * nk + chunk) * bm + row) * PK + ((j ^ (row & 7)) << 3);
}

// GRU notes (torch semantics n = tanh(x_n + r * (W_hn h + b_hn)), gate order r, z, n): `gates` is [T,B,3H]
// (x-projection + b_ih in, activations r, z, n out), `gh` [T,B,3H] keeps hn = W_hn h + b_hn at [2H, 3H) for the
// backward pass, the forward accumulator has a duplicate of gate n in the fourth 8-column group (N stays 32),
// and the backward emits two gate-gradient sets: dGx = (dr, dz, dn) for the input-side GEMMs and
// dGh = (dr, dz, dn * r) for the recurrent ones (dGh is what the next step streams).
struct PersistArgs {
  float* gates;          // fwd: [T,B,G*H] input projection + bias in, gate activations out; bwd: activations in
  float* gh;             // GRU: [T,B,3H]
  const float* bhh;      // GRU fwd: b_hh [3H]
  float* dG2;            // GRU bwd: dGh fp32 [T,B,3H]
  bf16* dG2_b;           // GRU bwd: dGh bf16 [T,B,3H]
  float* c;              // [T,B,H]  fwd: out; bwd: in
  float* h;              // [T,B,H]  fwd: out
  bf16* stream;          // [2][K/64][BM][64] swizzled tile image of the streamed operand, by step parity
  const float* dh_ext;   // bwd: [T,B,H] gradient from above (nullable)
  const float* d_hn;     // bwd: [B,H] (nullable)
  const float* d_cn;     // bwd: [B,H] (nullable)
  float* dG;             // bwd: [T,B,4H] out
  bf16* dG_b;            // bwd: [T,B,4H] out (row-major copy for the bulk GEMMs)
  PersistSync* sync;
  int B, H, T, stages;
};

// MODE 0: forward.  MODE 1: backward, CTA = 8 output units x all K = 4H (every CTA streams the whole dG).
// MODE 2: backward, 2-D: a cluster of 4 CTAs owns 32 output units; CTA `ks` of the cluster multiplies the
//         K quarter [ks*H, (ks+1)*H) (= gate ks of dG) into partial sums for all 32 units, the four partials
//         are exchanged through distributed shared memory, and each CTA finishes 8 of the units.  Every CTA
//         then streams a quarter of dG (what the forward streams of h) instead of all of it.
template <int BM, int MODE> struct PersistCfg {
  static constexpr bool BWD = MODE != 0;
  static constexpr bool TWOD = MODE == 2;
  // MMA N: fwd 4 gates x 8 units; 2-D bwd 32 units; 1-D bwd 8 units (M = 128 needs N % 16 == 0: the 8 rows
  // are loaded twice and accumulator columns 8-15 are ignored)
  static constexpr int N = MODE == 1 ? (BM == 128 ? 2 * UC : UC) : 4 * UC;
  static constexpr int XBUF = TWOD ? 4 * BM * UC * 4 : 0;   // bytes: [4 source CTAs][BM rows][8 units] fp32
  static constexpr int W_CHUNK = N * PK * 2;          // bytes of the resident weight shard per K chunk
  static constexpr int A_TILE = BM * PK * 2;          // bytes of the streamed operand per K chunk
  // one accumulator per MMA-issuing warp (TMEM columns [i*ACC_STRIDE, +N)); the epilogue adds them up
  static constexpr int ACC_STRIDE = N < 16 ? 16 : N;
  static constexpr int TM_COLS = 128;
  static constexpr int NACC = 4;
  static constexpr int CG = 32768 / A_TILE;           // K chunks per TMA instruction / ring stage (32 KiB)
  static constexpr int STAGE = CG * A_TILE;
};
constexpr int NPROD = 4;          // issuing warps: 0, 2, 3 and 8
constexpr int NISS = 4;           // MMA-issuing warps: 1, 9, 10, 11 (one per K-step of a 64-wide chunk)
constexpr int NTHREADS = 384;     // warps 4-7: epilogue (TMEM lane quarters 0-3)
static_assert(NISS == PK / 16, "issuer i owns K-step i of every chunk");

template <int BM, int MODE, int CELL>
__global__ void __launch_bounds__(NTHREADS, 1)
rnn_persist_kernel(const __grid_constant__ CUtensorMap tma_w, const PersistArgs a) {
  using Cfg = PersistCfg<BM, MODE>;
  constexpr int N = Cfg::N;
  constexpr bool BWD = Cfg::BWD, TWOD = Cfg::TWOD;
  constexpr bool GRU = CELL == B200RNN_CELL_GRU;
  constexpr int G = GRU ? 3 : 4;                        // gates
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int H = a.H, B = a.B, T = a.T;
  const int KTOT = MODE == 1 ? G * H : (MODE == 2 ? G * H / 4 : H);   // K extent of this CTA's product
  const int nk = KTOT / PK;
  const int nk_img = (BWD ? G * H : H) / PK;           // chunks of the streamed operand's tile image
  const int ks = TWOD ? (int)(blockIdx.x & 3u) : 0;    // 2-D: rank in the cluster = K quarter
  const int ngi = (int)(blockIdx.x >> 2);              // 2-D: group of 32 output units
  const int ng = (nk + Cfg::CG - 1) / Cfg::CG;         // chunk groups per step
  uint8_t* sW = smem;                                  // [nk][N rows][128 B] swizzled atoms
  uint8_t* sA = sW + (size_t)nk * Cfg::W_CHUNK;        // ring of [CG][BM rows][128 B] stages
  const int PSTAGES = a.stages;
  float* xbuf = reinterpret_cast<float*>(sA + PSTAGES * Cfg::STAGE);   // 2-D: partials received from the cluster
  uint64_t* full = reinterpret_cast<uint64_t*>(sA + PSTAGES * Cfg::STAGE + Cfg::XBUF);
  uint64_t* empty = full + PSTAGES;
  uint64_t* tmem_full = empty + PSTAGES;
  uint64_t* wbar = tmem_full + 1;
  uint64_t* xbar = wbar + 1;                           // 2-D: one arrival per peer CTA and step
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(xbar + 1);
  volatile int* s_abort = reinterpret_cast<volatile int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int u0 = blockIdx.x * UC;
  const unsigned ncta = gridDim.x;
  const int gshift = (int)((blockIdx.x * 5u) % (unsigned)ng);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tma_w);
    *s_abort = 0;
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < PSTAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], NISS); }
    mbar_init(tmem_full, NISS);
    mbar_init(wbar, 1);
    mbar_init(xbar, 3);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::TM_COLS>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if ((warp == 0 || warp == 2 || warp == 3 || warp == 8) && lane == 0) {
    // ===== copy producers (4 warps) =====
    // Measured on B200 (tools/microbench/tma_*.cu): one thread's bulk copies complete ~650-1000
    // cycles apart whatever their size (2-32 KiB), copies from different warps proceed in parallel, and a
    // tensor-map copy of 128-byte rows costs ~10 cycles per row on top.  Four warps issuing contiguous 32 KiB
    // 1-D copies move 128 KiB into an SM in ~1400 cycles with all 128 CTAs reading at once; one thread
    // issuing 8 KiB tensor tiles needed ~11000.
    const int pidx = warp == 0 ? 0 : (warp == 8 ? 3 : warp - 1);
    if (pidx == 0) {
      // resident weight shard, once
      mbar_expect_tx(wbar, (uint32_t)(nk * Cfg::W_CHUNK));
      for (int kb = 0; kb < nk; ++kb) {
        if (TWOD) {
#pragma unroll
          for (int r = 0; r < 4; ++r)                                   // W_hh^T rows = my group's 32 units, my K quarter
            tma_load_2d(sW + (size_t)kb * Cfg::W_CHUNK + r * (UC * PK * 2), &tma_w, wbar, ks * KTOT + kb * PK, ngi * 32 + r * UC);
        } else if (BWD) {
#pragma unroll
          for (int r = 0; r < N / UC; ++r)                                                  // W_hh^T rows = my units
            tma_load_2d(sW + (size_t)kb * Cfg::W_CHUNK + r * (UC * PK * 2), &tma_w, wbar, kb * PK, u0);
        } else {
#pragma unroll
          for (int g = 0; g < 4; ++g)                                   // W_hh rows g*H + my units (GRU: gate n twice)
            tma_load_2d(sW + (size_t)kb * Cfg::W_CHUNK + g * (UC * PK * 2), &tma_w, wbar, kb * PK, (g < G ? g : G - 1) * H + u0);
        }
      }
    }
    int it = 0;
    for (int s_i = 1; s_i < T; ++s_i) {
      // step s_i consumes what step s_i - 1 wrote (forward: h_{t-1}; backward: dG_{t+1}) into buffer (s_i-1)&1
      const uint8_t* src = reinterpret_cast<const uint8_t*>(a.stream) +
                           ((size_t)((s_i - 1) & 1) * nk_img + (size_t)ks * nk) * Cfg::A_TILE;
      const bool stamp = blockIdx.x == 0 && s_i == T / 2 && pidx == 0;
      PSTAMP(0);
      const bool go = wait_grid(a.sync, (unsigned)s_i * ncta, s_abort, 10);
      PSTAMP(1);
      fence_proxy_async_all();   // peers' generic-proxy stores (acquired above) -> my async-proxy (TMA) reads
      for (int g = 0; g < ng; ++g, ++it) {
        if (g % NPROD != pidx) continue;
        const int s = it % PSTAGES;
        const uint32_t ph = (it / PSTAGES) & 1;
        if (!go || !wait_mbar(&empty[s], ph ^ 1, s_abort, a.sync, 11)) continue;
        // every CTA walks the chunk groups from a different starting point (all CTAs read the same operand)
        const int kc0 = ((g + gshift) % ng) * Cfg::CG;
        const uint32_t bytes = (uint32_t)((nk - kc0 < Cfg::CG ? nk - kc0 : Cfg::CG) * Cfg::A_TILE);
        mbar_expect_tx(&full[s], bytes);
        bulk_load_1d(sA + s * Cfg::STAGE, src + (size_t)kc0 * Cfg::A_TILE, bytes, &full[s]);
      }
      PSTAMP(2);
    }
  } else if ((warp == 1 || warp >= 9) && lane == 0) {
    // ===== MMA issuers (4 warps) =====
    // Measured (tools/microbench/mma_issue_probe.cu): one thread issues a tcgen05.mma every ~129 cycles
    // whatever its shape (N 8..256), while issuers in different warps proceed in parallel until the
    // shared-memory read of A saturates (~33 cycles per M=64 MMA with four issuers, ~40 at M=128).  At
    // N <= 32 a single issuer would leave the tensor core idle 3/4 of the time, so issuer i takes K-step i
    // of every 64-wide chunk and accumulates into its own TMEM columns; the epilogue adds the four partials.
    const int iidx = warp == 1 ? 0 : warp - 8;
    constexpr uint32_t idesc = make_idesc(1, BM, N);
    const uint32_t tacc = tmem_base + (uint32_t)(iidx * Cfg::ACC_STRIDE);
    bool okw = wait_mbar(wbar, 0, s_abort, a.sync, 20);
    tc_fence_after_sync();
    int it = 0;
    for (int s_i = 1; s_i < T; ++s_i) {
      const bool stamp = blockIdx.x == 0 && s_i == T / 2 && iidx == 0;
      uint32_t accum = 0;
      for (int g = 0; g < ng; ++g, ++it) {
        const int s = it % PSTAGES;
        const uint32_t ph = (it / PSTAGES) & 1;
        if (!okw || !wait_mbar(&full[s], ph, s_abort, a.sync, 21)) continue;
        if (g == 0) PSTAMP(8);
        tc_fence_after_sync();
        const int kc0 = ((g + gshift) % ng) * Cfg::CG;
#pragma unroll
        for (int c = 0; c < Cfg::CG; ++c) {
          if (kc0 + c < nk) {   // the last group may hold fewer chunks
            const uint32_t a0 = smem_u32(sA + s * Cfg::STAGE + c * Cfg::A_TILE);
            const uint32_t b0 = smem_u32(sW + (size_t)(kc0 + c) * Cfg::W_CHUNK);
            // 128B-swizzled K-major tiles: 8-row groups 1024 B apart; K-step i of the chunk = +32 B * i
            const uint64_t da = make_smem_desc(a0 + iidx * 32, 16, 1024, 2);
            const uint64_t db = make_smem_desc(b0 + iidx * 32, 16, 1024, 2);
            mma_bf16(tacc, da, db, idesc, accum);
            accum = 1;
          }
        }
        mma_commit(&empty[s]);
      }
      mma_commit(tmem_full);
      PSTAMP(9);
    }
  } else if (warp >= 4 && warp < 8) {
    // ===== epilogue: one thread per batch row, UC hidden units each =====
    const int q = warp & 3;
    // TMEM lane -> accumulator row: M = 128 uses all 32 lanes of each quarter, M = 64 the lower 16
    const int row = (BM == 128) ? q * 32 + lane : q * 16 + lane;
    const bool live = (BM == 128 || lane < 16) && row < B;
    const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16);
    float carry[UC];   // fwd: c_{t-1}; bwd: dc flowing into step t
#pragma unroll
    for (int i = 0; i < UC; ++i) carry[i] = 0.0f;
    float bh[3][UC];   // GRU forward: b_hh of my units
    if (GRU && !BWD) {
#pragma unroll
      for (int g = 0; g < 3; ++g)
#pragma unroll
        for (int i = 0; i < UC; ++i) bh[g][i] = a.bhh[g * H + u0 + i];
    }
    if (!GRU && BWD && live && a.d_cn != nullptr) {
#pragma unroll
      for (int i = 0; i < UC; ++i) carry[i] = a.d_cn[(int64_t)row * H + u0 + i];
    }
    for (int s_i = 0; s_i < T; ++s_i) {
      const int t = BWD ? (T - 1 - s_i) : s_i;
      const int64_t rb = (int64_t)t * B + row;
      const bool stamp = blockIdx.x == 0 && s_i == T / 2 && threadIdx.x == 128;
      PSTAMP(16);
      // ---- pointwise inputs of this step: independent of the GEMM, fetched before the wait ----
      float4 gv[4][2];
      float4 aux[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};     // bwd: dh_ext / d_hn
      float4 cv[2], cpv[2];
      if (live) {
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const float4* gp = reinterpret_cast<const float4*>(a.gates + rb * G * H + (int64_t)g * H + u0);
          gv[g][0] = gp[0]; gv[g][1] = gp[1];
        }
        if (GRU) gv[3][0] = gv[3][1] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (BWD) {
          if (GRU) {   // gv[3] = hn; cpv = h_{t-1}; cv unused
            const float4* hp = reinterpret_cast<const float4*>(a.gh + rb * 3 * H + 2 * (int64_t)H + u0);
            gv[3][0] = hp[0]; gv[3][1] = hp[1];
            cv[0] = cv[1] = make_float4(0.f, 0.f, 0.f, 0.f);
          } else {
            const float4* cp = reinterpret_cast<const float4*>(a.c + rb * H + u0);
            cv[0] = cp[0]; cv[1] = cp[1];
          }
          if (t > 0) {
            const float4* pp = reinterpret_cast<const float4*>((GRU ? a.h : a.c) + (rb - B) * H + u0);
            cpv[0] = pp[0]; cpv[1] = pp[1];
          } else {
            cpv[0] = cpv[1] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
          if (a.dh_ext != nullptr) {
            const float4* ep = reinterpret_cast<const float4*>(a.dh_ext + rb * H + u0);
            aux[0] = ep[0]; aux[1] = ep[1];
          }
          if (s_i == 0 && a.d_hn != nullptr) {
            const float4* ep = reinterpret_cast<const float4*>(a.d_hn + (int64_t)row * H + u0);
            const float4 e0 = ep[0], e1 = ep[1];
            aux[0].x += e0.x; aux[0].y += e0.y; aux[0].z += e0.z; aux[0].w += e0.w;
            aux[1].x += e1.x; aux[1].y += e1.y; aux[1].z += e1.z; aux[1].w += e1.w;
          }
        }
      }
      // ---- recurrent product of this step ----
      float acc[N];
#pragma unroll
      for (int i = 0; i < N; ++i) acc[i] = 0.0f;
      if (s_i > 0) {
        const bool ok = __all_sync(0xffffffffu, wait_mbar(tmem_full, (uint32_t)(s_i - 1) & 1, s_abort, a.sync, 30));
        tc_fence_after_sync();
        PSTAMP(17);
        if (ok) {
          // sum the issuers' partial accumulators (each issuer ran at least one MMA: K-step i of chunk 0)
#pragma unroll
          for (int ai = 0; ai < Cfg::NACC; ++ai) {
            uint32_t r0[16];
            tmem_ld16_nowait(taddr + ai * Cfg::ACC_STRIDE, r0);
            if (N > 16) {
              uint32_t r1[16];
              tmem_ld16_nowait(taddr + ai * Cfg::ACC_STRIDE + 16, r1);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) { acc[i] += __uint_as_float(r0[i]); acc[(16 + i) % N] += __uint_as_float(r1[i]); }
            } else {
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < N; ++i) acc[i] += __uint_as_float(r0[i]);
            }
          }
        }
      }
      // ---- 2-D backward: acc holds my K quarter's partial sums for the cluster's 32 units.  Send the 8-unit
      // slices the peers finish into their xbuf[my rank], keep mine, and add what the peers sent me. ----
      float dhrec[UC];
#pragma unroll
      for (int i = 0; i < UC; ++i) dhrec[i] = BWD ? acc[i % N] : 0.0f;
      if (TWOD && s_i > 0) {
        const uint32_t xb_local = smem_u32(xbuf + ((size_t)ks * BM + row) * UC);
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          if (p == ks) {
#pragma unroll
            for (int i = 0; i < UC; ++i) dhrec[i] = acc[(p * UC + i) % N];
          } else if (live) {
            const uint32_t ra = map_to_cta(xb_local, (uint32_t)p);
            st_cluster_v4(ra, acc[(p * UC + 0) % N], acc[(p * UC + 1) % N], acc[(p * UC + 2) % N], acc[(p * UC + 3) % N]);
            st_cluster_v4(ra + 16, acc[(p * UC + 4) % N], acc[(p * UC + 5) % N], acc[(p * UC + 6) % N], acc[(p * UC + 7) % N]);
          }
        }
        epi_bar_sync();   // the CTA's slices are stored; one release-arrive per peer covers them all
        if (threadIdx.x == 128) {
          const uint32_t xbar_local = smem_u32(xbar);
#pragma unroll
          for (int p = 0; p < 4; ++p)
            if (p != ks) mbar_arrive_remote_release(map_to_cta(xbar_local, (uint32_t)p));
        }
        const bool okx = __all_sync(0xffffffffu, wait_mbar_cluster(xbar, (uint32_t)(s_i - 1) & 1, s_abort, a.sync, 31));
        if (okx && live) {
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            if (p != ks) {
              const float4* xp = reinterpret_cast<const float4*>(xbuf + ((size_t)p * BM + row) * UC);
              const float4 x0 = xp[0], x1 = xp[1];
              dhrec[0] += x0.x; dhrec[1] += x0.y; dhrec[2] += x0.z; dhrec[3] += x0.w;
              dhrec[4] += x1.x; dhrec[5] += x1.y; dhrec[6] += x1.z; dhrec[7] += x1.w;
            }
          }
        }
      }
      tc_fence_before_sync();   // my TMEM reads are done before the arrive lets the next step's MMAs overwrite D
      const float gin[4][UC] = {
          {gv[0][0].x, gv[0][0].y, gv[0][0].z, gv[0][0].w, gv[0][1].x, gv[0][1].y, gv[0][1].z, gv[0][1].w},
          {gv[1][0].x, gv[1][0].y, gv[1][0].z, gv[1][0].w, gv[1][1].x, gv[1][1].y, gv[1][1].z, gv[1][1].w},
          {gv[2][0].x, gv[2][0].y, gv[2][0].z, gv[2][0].w, gv[2][1].x, gv[2][1].y, gv[2][1].z, gv[2][1].w},
          {gv[3][0].x, gv[3][0].y, gv[3][0].z, gv[3][0].w, gv[3][1].x, gv[3][1].y, gv[3][1].z, gv[3][1].w}};
      if (!BWD) {
        float go[4][UC], hv[UC];
        if (live) {
#pragma unroll
          for (int i = 0; i < UC; ++i) {
            if constexpr (!GRU) {
              const float ig = sigmoid_f(gin[0][i] + acc[(0 * UC + i) % N]);
              const float fg = sigmoid_f(gin[1][i] + acc[(1 * UC + i) % N]);
              const float gg = tanh_f(gin[2][i] + acc[(2 * UC + i) % N]);
              const float og = sigmoid_f(gin[3][i] + acc[(3 * UC + i) % N]);
              const float cc = fmaf(fg, carry[i], ig * gg);
              carry[i] = cc;
              hv[i] = og * tanh_f(cc);
              go[0][i] = ig; go[1][i] = fg; go[2][i] = gg; go[3][i] = og;
            } else {
              // carry = h_{t-1}; go[3] = hn, stashed for the backward pass
              const float rg = sigmoid_f(gin[0][i] + acc[(0 * UC + i) % N] + bh[0][i]);
              const float zg = sigmoid_f(gin[1][i] + acc[(1 * UC + i) % N] + bh[1][i]);
              const float hn = acc[(2 * UC + i) % N] + bh[2][i];
              const float ng = tanh_f(fmaf(rg, hn, gin[2][i]));
              hv[i] = fmaf(zg, carry[i] - ng, ng);
              carry[i] = hv[i];
              go[0][i] = rg; go[1][i] = zg; go[2][i] = ng; go[3][i] = hn;
            }
          }
          // the operand of the next step first ...
          __nv_bfloat162 p0 = __floats2bfloat162_rn(hv[0], hv[1]), p1 = __floats2bfloat162_rn(hv[2], hv[3]);
          __nv_bfloat162 p2 = __floats2bfloat162_rn(hv[4], hv[5]), p3 = __floats2bfloat162_rn(hv[6], hv[7]);
          uint4 pk;
          pk.x = *reinterpret_cast<uint32_t*>(&p0); pk.y = *reinterpret_cast<uint32_t*>(&p1);
          pk.z = *reinterpret_cast<uint32_t*>(&p2); pk.w = *reinterpret_cast<uint32_t*>(&p3);
          *reinterpret_cast<uint4*>(stream_unit(a.stream, s_i & 1, nk_img, BM, row, u0)) = pk;
        }
        PSTAMP(18);
        epi_bar_sync();
        PSTAMP(19);
        // first epilogue thread publishes this CTA's step: the release is cumulative over the stores the
        // CTA barrier above ordered before it
        if (threadIdx.x == 128) red_release_gpu(&a.sync->arrive, 1u);
        PSTAMP(20);
        // ... then the stash for the backward pass, off the critical path
        if (live) {
#pragma unroll
          for (int g = 0; g < G; ++g) {
            float4* gp = reinterpret_cast<float4*>(a.gates + rb * G * H + (int64_t)g * H + u0);
            gp[0] = make_float4(go[g][0], go[g][1], go[g][2], go[g][3]);
            gp[1] = make_float4(go[g][4], go[g][5], go[g][6], go[g][7]);
          }
          // LSTM: c_t; GRU: hn into gh[:, 2H + u]
          float4* cp = GRU ? reinterpret_cast<float4*>(a.gh + rb * 3 * H + 2 * (int64_t)H + u0)
                           : reinterpret_cast<float4*>(a.c + rb * H + u0);
          if constexpr (GRU) {
            cp[0] = make_float4(go[3][0], go[3][1], go[3][2], go[3][3]);
            cp[1] = make_float4(go[3][4], go[3][5], go[3][6], go[3][7]);
          } else {
            cp[0] = make_float4(carry[0], carry[1], carry[2], carry[3]);
            cp[1] = make_float4(carry[4], carry[5], carry[6], carry[7]);
          }
          float4* hp = reinterpret_cast<float4*>(a.h + rb * H + u0);
          hp[0] = make_float4(hv[0], hv[1], hv[2], hv[3]);
          hp[1] = make_float4(hv[4], hv[5], hv[6], hv[7]);
        }
      } else {
        float dg[4][UC];
        uint4 dgp[4];
        if (live) {
          const float cvv[UC] = {cv[0].x, cv[0].y, cv[0].z, cv[0].w, cv[1].x, cv[1].y, cv[1].z, cv[1].w};
          const float cpp[UC] = {cpv[0].x, cpv[0].y, cpv[0].z, cpv[0].w, cpv[1].x, cpv[1].y, cpv[1].z, cpv[1].w};
          const float dhe[UC] = {aux[0].x, aux[0].y, aux[0].z, aux[0].w, aux[1].x, aux[1].y, aux[1].z, aux[1].w};
#pragma unroll
          for (int i = 0; i < UC; ++i) {
            if constexpr (!GRU) {
              const float ig = gin[0][i], fg = gin[1][i], gg = gin[2][i], og = gin[3][i];
              const float dh = dhe[i] + dhrec[i];
              const float tc = tanh_f(cvv[i]);
              const float dct = fmaf(dh * og, 1.0f - tc * tc, carry[i]);
              dg[0][i] = dct * gg * ig * (1.0f - ig);
              dg[1][i] = dct * cpp[i] * fg * (1.0f - fg);
              dg[2][i] = dct * ig * (1.0f - gg * gg);
              dg[3][i] = dh * tc * og * (1.0f - og);
              carry[i] = dct * fg;
            } else {
              // carry = dh_t * z_t flowing straight into h_{t-1}; dg[0..2] = dGh (dr, dz, dn * r), dg[3] = dn
              const float rg = gin[0][i], zg = gin[1][i], ng = gin[2][i], hn = gin[3][i];
              const float dh = dhe[i] + dhrec[i] + carry[i];
              const float dn = dh * (1.0f - zg) * (1.0f - ng * ng);
              dg[0][i] = dn * hn * rg * (1.0f - rg);
              dg[1][i] = dh * (cpp[i] - ng) * zg * (1.0f - zg);
              dg[2][i] = dn * rg;
              dg[3][i] = dn;
              carry[i] = dh * zg;
              (void)cvv;
            }
          }
#pragma unroll
          for (int g = 0; g < G; ++g) {
            __nv_bfloat162 p0 = __floats2bfloat162_rn(dg[g][0], dg[g][1]), p1 = __floats2bfloat162_rn(dg[g][2], dg[g][3]);
            __nv_bfloat162 p2 = __floats2bfloat162_rn(dg[g][4], dg[g][5]), p3 = __floats2bfloat162_rn(dg[g][6], dg[g][7]);
            uint4 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&p0); pk.y = *reinterpret_cast<uint32_t*>(&p1);
            pk.z = *reinterpret_cast<uint32_t*>(&p2); pk.w = *reinterpret_cast<uint32_t*>(&p3);
            *reinterpret_cast<uint4*>(stream_unit(a.stream, s_i & 1, nk_img, BM, row, g * H + u0)) = pk;
            dgp[g] = pk;
          }
        }
        PSTAMP(18);
        epi_bar_sync();
        PSTAMP(19);
        if (threadIdx.x == 128) red_release_gpu(&a.sync->arrive, 1u);
        PSTAMP(20);
        if (live) {
#pragma unroll
          for (int g = 0; g < G; ++g) {
            // LSTM: dG; GRU: dGh (what was streamed) into dG2 and dGx = (dr, dz, dn) into dG
            float* d32 = GRU ? a.dG2 : a.dG;
            bf16* d16 = GRU ? a.dG2_b : a.dG_b;
            float4* gp = reinterpret_cast<float4*>(d32 + rb * G * H + (int64_t)g * H + u0);
            gp[0] = make_float4(dg[g][0], dg[g][1], dg[g][2], dg[g][3]);
            gp[1] = make_float4(dg[g][4], dg[g][5], dg[g][6], dg[g][7]);
            *reinterpret_cast<uint4*>(d16 + rb * G * H + (int64_t)g * H + u0) = dgp[g];
            if constexpr (GRU) {
              const int gs = g == 2 ? 3 : g;   // dGx differs from dGh only in the n gate (dn instead of dn * r)
              float4* xp = reinterpret_cast<float4*>(a.dG + rb * G * H + (int64_t)g * H + u0);
              xp[0] = make_float4(dg[gs][0], dg[gs][1], dg[gs][2], dg[gs][3]);
              xp[1] = make_float4(dg[gs][4], dg[gs][5], dg[gs][6], dg[gs][7]);
              __nv_bfloat162 p0 = __floats2bfloat162_rn(dg[gs][0], dg[gs][1]), p1 = __floats2bfloat162_rn(dg[gs][2], dg[gs][3]);
              __nv_bfloat162 p2 = __floats2bfloat162_rn(dg[gs][4], dg[gs][5]), p3 = __floats2bfloat162_rn(dg[gs][6], dg[gs][7]);
              uint4 pk;
              pk.x = *reinterpret_cast<uint32_t*>(&p0); pk.y = *reinterpret_cast<uint32_t*>(&p1);
              pk.z = *reinterpret_cast<uint32_t*>(&p2); pk.w = *reinterpret_cast<uint32_t*>(&p3);
              *reinterpret_cast<uint4*>(a.dG_b + rb * G * H + (int64_t)g * H + u0) = pk;
            }
          }
        }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) { __syncwarp(); tmem_dealloc<Cfg::TM_COLS>(tmem_base); }
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiled persist_encode() {
  static EncodeTiled fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess) fn = (EncodeTiled)p;
  }
  return fn;
}
// bf16 [rows][K] weight matrix: box = one K chunk x UC rows (one 8-row swizzle atom)
int make_map_weight(CUtensorMap* map, const void* ptr, int rows, int K) {
  EncodeTiled enc = persist_encode();
  B200_REQUIRE(enc != nullptr, B200RNN_ECUDA, "rnn_persist: cuTensorMapEncodeTiled entry point not available");
  const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
  const cuuint32_t box[2] = {(cuuint32_t)PK, (cuuint32_t)UC};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, B200RNN_ECUDA, "rnn_persist: cuTensorMapEncodeTiled(weights) failed with %d", (int)r);
  return B200RNN_OK;
}

constexpr size_t kSmemLimit = 227 * 1024;
constexpr size_t kSmemFixed = 2 * MAX_STAGES * 8 + 64 + 1024;   // barriers + slots + alignment slack
constexpr int kMinStages = 2;

inline int persist_nk(int mode, int H, int G) { return (mode == 1 ? G * H : (mode == 2 ? G * H / 4 : H)) / PK; }

// ring depth: the whole streamed operand of a step if it fits next to the weight shard, else what does
template <int BM, int MODE>
int persist_stages(int H, int G) {
  using Cfg = PersistCfg<BM, MODE>;
  const int nk = persist_nk(MODE, H, G);
  const int ng = (nk + Cfg::CG - 1) / Cfg::CG;
  const size_t w = (size_t)nk * Cfg::W_CHUNK + Cfg::XBUF;
  if (w + kSmemFixed + (size_t)kMinStages * Cfg::STAGE > kSmemLimit) return 0;
  int st = (int)((kSmemLimit - w - kSmemFixed) / Cfg::STAGE);
  if (st > ng && ng >= kMinStages) st = ng;
  if (st > MAX_STAGES) st = MAX_STAGES;
  return st;
}
template <int BM, int MODE>
size_t persist_smem(int H, int G) {
  using Cfg = PersistCfg<BM, MODE>;
  return (size_t)persist_nk(MODE, H, G) * Cfg::W_CHUNK + (size_t)persist_stages<BM, MODE>(H, G) * Cfg::STAGE + Cfg::XBUF + kSmemFixed;
}

// Returns B200RNN_EUNSUPPORTED (without launching) when the grid cannot be made co-resident, so that the
// caller can fall back to another decomposition.
template <int BM, int MODE, int CELL>
int launch_persist(const CUtensorMap& mw, PersistArgs a, cudaStream_t s) {
  constexpr int G = CELL == B200RNN_CELL_GRU ? 3 : 4;
  B200_REQUIRE(MODE != 2 || (G * a.H) % (4 * PK) == 0, B200RNN_EUNSUPPORTED,
               "rnn_persist: K = %d does not split into four 64-aligned quarters", G * a.H);
  a.stages = persist_stages<BM, MODE>(a.H, G);
  B200_REQUIRE(a.stages >= kMinStages, B200RNN_EUNSUPPORTED, "rnn_persist: weight shard of H=%d leaves no room for the operand ring", a.H);
  const size_t smem = persist_smem<BM, MODE>(a.H, G);
  auto kern = rnn_persist_kernel<BM, MODE, CELL>;
  B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = a.H / UC;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(NTHREADS); cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  if (MODE == 2) {
    // clusters of 4 CTAs (the K quarters of one 32-unit group), all clusters co-resident
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = 4; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
    cfg.numAttrs = 2;
    int n_clusters = 0;
    const cudaError_t e = cudaOccupancyMaxActiveClusters(&n_clusters, (void*)kern, &cfg);
    if (e != cudaSuccess) { cudaGetLastError(); n_clusters = 0; }
    B200_REQUIRE(n_clusters * 4 >= grid, B200RNN_EUNSUPPORTED,
                 "rnn_persist: %d clusters of 4 CTAs cannot be co-resident (%d fit)", grid / 4, n_clusters);
  } else {
    int per_sm = 0;
    B200_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NTHREADS, smem));
    B200_REQUIRE(per_sm * num_sms() >= grid, B200RNN_EUNSUPPORTED,
                 "rnn_persist: %d CTAs cannot be co-resident (%d per SM x %d SMs)", grid, per_sm, num_sms());
  }
  B200_CHECK_CUDA(cudaMemsetAsync(a.sync, 0, sizeof(PersistSync), s));
  void* args[2] = {(void*)&mw, (void*)&a};
  const cudaError_t le = cudaLaunchKernelExC(&cfg, (void*)kern, args);
  if (le != cudaSuccess && MODE == 2) {   // e.g. a driver that refuses cooperative + cluster launches
    cudaGetLastError();
    set_error("rnn_persist: clustered cooperative launch refused (%s)", cudaGetErrorString(le));
    return B200RNN_EUNSUPPORTED;
  }
  B200_CHECK_CUDA(le);
  return B200RNN_OK;
}

template <int CELL>
int run_fwd(const CUtensorMap& mw, const PersistArgs& a, cudaStream_t s) {
  return a.B <= 64 ? launch_persist<64, 0, CELL>(mw, a, s) : launch_persist<128, 0, CELL>(mw, a, s);
}
// 2-D decomposition first (a quarter of the operand traffic per SM); the 1-D one when K does not split, its
// clusters do not fit, or B200RNN_BWD2D=0
template <int CELL>
int run_bwd(const CUtensorMap& mw, const PersistArgs& a, cudaStream_t s) {
  static const bool twod = [] { const char* e = std::getenv("B200RNN_BWD2D"); return !(e && e[0] == '0'); }();
  if (twod) {
    const int rc = a.B <= 64 ? launch_persist<64, 2, CELL>(mw, a, s) : launch_persist<128, 2, CELL>(mw, a, s);
    if (rc != B200RNN_EUNSUPPORTED) return rc;
  }
  return a.B <= 64 ? launch_persist<64, 1, CELL>(mw, a, s) : launch_persist<128, 1, CELL>(mw, a, s);
}

}  // namespace
}  // namespace b200rnn

using namespace b200rnn;

// Shapes the persistent kernels cover: LSTM / GRU, H a multiple of 64 with H/8 CTAs co-resident (one per
// SM), batch <= 128, and a weight shard + ring that fits shared memory.  B200RNN_PERSIST=0 disables.
extern "C" int b200rnn_rnn_persist_supported(const b200rnn_model_desc* d) {
  if (d == nullptr || (d->cell != B200RNN_CELL_LSTM && d->cell != B200RNN_CELL_GRU)) return 0;
  static const bool enabled = [] { const char* e = std::getenv("B200RNN_PERSIST"); return !(e && e[0] == '0'); }();
  if (!enabled) return 0;
  const int H = d->hidden_size, B = d->batch, G = d->cell == B200RNN_CELL_LSTM ? 4 : 3;
  if (H % 64 != 0 || H < 64 || B < 1 || B > 128 || d->seq_len < 1) return 0;
  if (H / UC > num_sms()) return 0;
  const int st = B <= 64 ? (persist_stages<64, 1>(H, G) < persist_stages<64, 0>(H, G) ? persist_stages<64, 1>(H, G) : persist_stages<64, 0>(H, G))
                         : (persist_stages<128, 1>(H, G) < persist_stages<128, 0>(H, G) ? persist_stages<128, 1>(H, G) : persist_stages<128, 0>(H, G));
  return st >= kMinStages ? 1 : 0;
}

// Bytes of the double-buffered tile image of the streamed operand (backward: K = 4H, up to 128 rows).
extern "C" int64_t b200rnn_rnn_persist_stream_bytes(int32_t H) { return 2ll * (4 * H / PK) * 128 * PK * 2; }

// Forward time loop of one LSTM layer.  gates: [T,B,4H] fp32 holding x-projection + biases (overwritten with
// the gate activations); whh_b: bf16 [4H,H]; outputs c, h (fp32), both [T,B,H]; stream: scratch of
// b200rnn_rnn_persist_stream_bytes(H); sync: 256 bytes.
extern "C" int b200rnn_lstm_persist_fwd(int32_t T, int32_t B, int32_t H, float* gates, const void* whh_b,
                                        float* c, float* h, void* stream_buf, void* sync, void* stream) {
  B200_REQUIRE(gates && whh_b && c && h && stream_buf && sync, B200RNN_EINVAL, "lstm_persist_fwd: null argument");
  CUtensorMap mw;
  int rc = make_map_weight(&mw, whh_b, 4 * H, H);
  if (rc) return rc;
  PersistArgs a{};
  a.gates = gates; a.c = c; a.h = h; a.stream = (bf16*)stream_buf; a.sync = (PersistSync*)sync;
  a.B = B; a.H = H; a.T = T;
  return run_fwd<B200RNN_CELL_LSTM>(mw, a, (cudaStream_t)stream);
}

// Backward time loop of one LSTM layer.  gates / c: the forward stash; dh_ext: [T,B,H] gradient from above
// (nullable); d_hn / d_cn: [B,H] (nullable); whhT_b: bf16 [H,4H] (W_hh transposed); outputs dG fp32 and
// dG_b bf16, both [T,B,4H] row-major.
extern "C" int b200rnn_lstm_persist_bwd(int32_t T, int32_t B, int32_t H, const float* gates, const float* c,
                                        const float* dh_ext, const float* d_hn, const float* d_cn,
                                        const void* whhT_b, float* dG, void* dG_b, void* stream_buf, void* sync,
                                        void* stream) {
  B200_REQUIRE(gates && c && whhT_b && dG && dG_b && stream_buf && sync, B200RNN_EINVAL, "lstm_persist_bwd: null argument");
  CUtensorMap mw;
  int rc = make_map_weight(&mw, whhT_b, H, 4 * H);
  if (rc) return rc;
  PersistArgs a{};
  a.gates = const_cast<float*>(gates); a.c = const_cast<float*>(c);
  a.dh_ext = dh_ext; a.d_hn = d_hn; a.d_cn = d_cn; a.dG = dG; a.dG_b = (bf16*)dG_b;
  a.stream = (bf16*)stream_buf; a.sync = (PersistSync*)sync;
  a.B = B; a.H = H; a.T = T;
  return run_bwd<B200RNN_CELL_LSTM>(mw, a, (cudaStream_t)stream);
}

// GRU forward time loop of one layer.  gates: [T,B,3H] x-projection + b_ih in, (r, z, n) out; gh: [T,B,3H],
// receives hn at [2H, 3H); whh_b: bf16 [3H,H]; bhh: fp32 [3H]; h: [T,B,H] out.
extern "C" int b200rnn_gru_persist_fwd(int32_t T, int32_t B, int32_t H, float* gates, float* gh, const void* whh_b,
                                       const float* bhh, float* h, void* stream_buf, void* sync, void* stream) {
  B200_REQUIRE(gates && gh && whh_b && bhh && h && stream_buf && sync, B200RNN_EINVAL, "gru_persist_fwd: null argument");
  CUtensorMap mw;
  int rc = make_map_weight(&mw, whh_b, 3 * H, H);
  if (rc) return rc;
  PersistArgs a{};
  a.gates = gates; a.gh = gh; a.bhh = bhh; a.h = h; a.stream = (bf16*)stream_buf; a.sync = (PersistSync*)sync;
  a.B = B; a.H = H; a.T = T;
  return run_fwd<B200RNN_CELL_GRU>(mw, a, (cudaStream_t)stream);
}

// GRU backward time loop of one layer.  gates / gh / h: the forward stash; whhT_b: bf16 [H,3H]; outputs
// dGx = (dr, dz, dn) in dG / dG_b and dGh = (dr, dz, dn * r) in dG2 / dG2_b, all [T,B,3H] row-major.
extern "C" int b200rnn_gru_persist_bwd(int32_t T, int32_t B, int32_t H, const float* gates, const float* gh, const float* h,
                                       const float* dh_ext, const float* d_hn, const void* whhT_b, float* dG, void* dG_b,
                                       float* dG2, void* dG2_b, void* stream_buf, void* sync, void* stream) {
  B200_REQUIRE(gates && gh && h && whhT_b && dG && dG_b && dG2 && dG2_b && stream_buf && sync, B200RNN_EINVAL,
               "gru_persist_bwd: null argument");
  CUtensorMap mw;
  int rc = make_map_weight(&mw, whhT_b, H, 3 * H);
  if (rc) return rc;
  PersistArgs a{};
  a.gates = const_cast<float*>(gates); a.gh = const_cast<float*>(gh); a.h = const_cast<float*>(h);
  a.dh_ext = dh_ext; a.d_hn = d_hn; a.dG = dG; a.dG_b = (bf16*)dG_b; a.dG2 = dG2; a.dG2_b = (bf16*)dG2_b;
  a.stream = (bf16*)stream_buf; a.sync = (PersistSync*)sync;
  a.B = B; a.H = H; a.T = T;
  return run_bwd<B200RNN_CELL_GRU>(mw, a, (cudaStream_t)stream);
}

extern "C" int b200rnn_persist_status(int32_t* status) {
  B200_REQUIRE(status != nullptr, B200RNN_EINVAL, "persist_status: null");
  int v = 0;
  B200_CHECK_CUDA(cudaMemcpyFromSymbol(&v, g_persist_error, sizeof(int)));
  *status = v;
  return B200RNN_OK;
}

extern "C" int b200rnn_persist_prof(long long* out48) {
  B200_REQUIRE(out48 != nullptr, B200RNN_EINVAL, "persist_prof: null");
  B200_CHECK_CUDA(cudaMemcpyFromSymbol(out48, g_persist_prof, sizeof(long long) * 48));
  return B200RNN_OK;
}
