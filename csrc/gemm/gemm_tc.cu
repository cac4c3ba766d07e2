// K1 / K4 — bf16 GEMM on the 5th-generation tensor cores: TMA-fed, tcgen05.mma, TMEM accumulators.
//
//   C[M,N] (fp32 or bf16) (+)= A[M,K] (bf16, row-major) x B[N,K]^T (bf16, row-major) + bias[N]
//
// Both operands K-major, which is how every LSTM/GRU GEMM reads its weights without transposes:
// the input projection over ALL timesteps at once  X[T*B, in] x W_ih[4H, in]^T + b   (K1; replaces
// the per-timestep `linear` of ATen's lstm, src/motion/model.py:14), the per-step recurrent
// projection h[B,H] x W_hh[4H,H]^T for large hidden sizes, and the weight-gradient GEMMs
// dW = dG^T x [x | h] over T*B (K4; replaces autograd's addmm backward, trainer/base.py:116).
//
// Structure (one 128x128 output tile per CTA, 256 threads, warp-specialised):
//   warp 0   : TMA producer — cp.async.bulk.tensor 128B-swizzled A/B tiles (BLOCK_K = 64 bf16) into a
//              4-stage shared-memory ring, completion on `full` mbarriers;
//   warp 1   : MMA issuer — one elected thread issues 4 x tcgen05.mma (M128 N128 K16) per stage into a
//              128-column fp32 TMEM accumulator, tcgen05.commit releases the stage (`empty`) and,
//              after the last k-block, signals the epilogue (`tmem_full`);
//   warp 2   : TMEM allocator / deallocator;
//   warps 4-7: epilogue — tcgen05.ld (32 lanes x 16 columns per instruction), + bias, optional
//              accumulate, fp32 or bf16 store.
// Out-of-range rows / columns / K are zero-filled by TMA, so M, N, K need no padding (K % 8 == 0
// for the 16-byte global stride rule).  Every mbarrier wait is bounded; a timeout sets an error
// word that b200rnn_gemm_status() reports.
#include <cuda_bf16.h>
#include <cstdlib>
#include "tc_common.cuh"

namespace b200rnn {
namespace {

using namespace tc;

constexpr int BM = 128, BK = 64, STAGES = 4;
constexpr int TILE_BYTES = BM * BK * 2;           // 16 KiB for the A tile per stage
// BN = 256 for large problems: one thread issues a tcgen05.mma every ~129 cycles whatever its shape
// (profiles/r1_microbench_tma_mma.md), so only a 128x256 MMA (128 cycles of tensor time) keeps the tensor core
// busy from a single issuer; BN = 128 when N or the grid is too small for that; BN = 32 for skinny problems
// (M <= 128: the per-timestep recurrent projection h[B,H] x W_hh^T) so that N / 32 CTAs cover the SMs.
template <int BN> constexpr int smem_bytes() { return STAGES * (TILE_BYTES + BN * BK * 2) + 256 + 1024; }

__device__ int g_gemm_error = 0;

template <int BN>
__global__ void __launch_bounds__(256, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                 void* __restrict__ Cout, const float* __restrict__ bias, int M, int N, int K, int ldc,
                 int accumulate, int c_bf16, unsigned int* __restrict__ retire) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  constexpr int B_TILE = BN * BK * 2;
  constexpr int TM_COLS = BN < 32 ? 32 : BN;
  uint8_t* sB = smem + STAGES * TILE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (TILE_BYTES + B_TILE));
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = (K + BK - 1) / BK;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tma_a); tma_prefetch_desc(&tma_b); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<TM_COLS>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer =====
    for (int kb = 0; kb < nk; ++kb) {
      const int s = kb % STAGES;
      const uint32_t ph = (kb / STAGES) & 1;
      if (!mbar_wait(&empty[s], ph ^ 1)) { atomicExch(&g_gemm_error, 1); break; }
      mbar_expect_tx(&full[s], TILE_BYTES + B_TILE);
      tma_load_2d(sA + s * TILE_BYTES, &tma_a, &full[s], kb * BK, m0);
      tma_load_2d(sB + s * B_TILE, &tma_b, &full[s], kb * BK, n0);
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = make_idesc(1, BM, BN);
    bool ok = true;
    for (int kb = 0; kb < nk && ok; ++kb) {
      const int s = kb % STAGES;
      const uint32_t ph = (kb / STAGES) & 1;
      if (!mbar_wait(&full[s], ph)) { atomicExch(&g_gemm_error, 2); ok = false; break; }
      tc_fence_after_sync();
      const uint32_t a0 = smem_u32(sA + s * TILE_BYTES), b0 = smem_u32(sB + s * B_TILE);
#pragma unroll
      for (int k = 0; k < BK / 16; ++k) {
        // 128B-swizzled K-major tile: 8-row groups 1024 B apart; a K step of 16 bf16 = +32 B
        const uint64_t da = make_smem_desc(a0 + k * 32, 16, 1024, 2);
        const uint64_t db = make_smem_desc(b0 + k * 32, 16, 1024, 2);
        mma_bf16(tmem_base, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
      }
      mma_commit(&empty[s]);
    }
    mma_commit(tmem_full);
  } else if (warp >= 4) {
    // ===== epilogue =====
    const bool ok = mbar_wait(tmem_full, 0);
    if (!ok && lane == 0) atomicExch(&g_gemm_error, 3);
    tc_fence_after_sync();
    const int q = warp & 3;
    float* Cf = reinterpret_cast<float*>(Cout);
    __nv_bfloat16* Cb = reinterpret_cast<__nv_bfloat16*>(Cout);
    // TMEM gives each lane one output ROW; writing it out directly would scatter every store instruction
    // over 32 rows (measured: 512 MB of fp32 output at ~400 GB/s).  Stage the warp's 32 x BN block through
    // shared memory (the operand ring is idle once tmem_full has fired) and write rows out coalesced.
    constexpr int LDS = BN + 1;   // odd stride: conflict-free for both the row-wise writes and the column-wise reads
    float* stage = reinterpret_cast<float*>(smem) + q * 32 * LDS;
#pragma unroll 1
    for (int c = 0; c < BN / 16; ++c) {
      float v[16];
      tmem_ld16(tmem_base + (uint32_t(q * 32) << 16) + c * 16, v);
#pragma unroll
      for (int j = 0; j < 16; ++j) stage[lane * LDS + c * 16 + j] = v[j];
    }
    __syncwarp();
    if (ok) {
#pragma unroll 1
      for (int r = 0; r < 32; ++r) {
        const int grow = m0 + q * 32 + r;
        if (grow >= M) break;
#pragma unroll
        for (int k = 0; k < BN / 32; ++k) {
          const int col = n0 + k * 32 + lane;
          if (col < N) {
            float val = stage[r * LDS + k * 32 + lane] + (bias != nullptr ? bias[col] : 0.0f);
            const int64_t e = (int64_t)grow * ldc + col;
            if (c_bf16) {
              if (accumulate) val += __bfloat162float(Cb[e]);
              Cb[e] = __float2bfloat16(val);
            } else {
              if (accumulate) val += Cf[e];
              Cf[e] = val;
            }
          }
        }
      }
    }
  }
  if (retire != nullptr) __threadfence();   // this tile's stores are visible device-wide before it is counted
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc<TM_COLS>(tmem_base);
  // Tile retired: one release-add per CTA on the caller's counter.  A consumer kernel on another stream
  // (the gradient bucket's all-reduce) spins on it and starts the moment the LAST tile of the weight-gradient
  // GEMMs is out — the device-side "grad ready" trigger of the reducer (csrc/runtime/reducer.cpp).
  if (retire != nullptr && threadIdx.x == 0)
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(retire) : "memory");
}


// ---------------------------------------------------------------------------
// Persistent form for the bulk GEMMs (input projection over all timesteps, dX, dW): one CTA per SM walks a
// static list of 128x256 output tiles.  Two 256-column fp32 accumulators fill the SM's 512 TMEM columns, so
// the epilogue of tile i (TMEM -> registers -> 128B-swizzled shared memory -> TMA store) runs while the MMA
// warp is already accumulating tile i+1; the operand ring (4 stages x 48 KiB) never drains between tiles.
//   warp 0   : TMA producer (A 128x64, B 256x64 bf16 boxes, 128B swizzle)
//   warp 1   : MMA issuer (4 x tcgen05.mma M128 N256 K16 per stage), commits `empty[s]` per stage and
//              `acc_full[a]` per tile; waits `acc_empty[a]` before reusing accumulator a
//   warp 2   : TMEM allocator
//   warps 4-7: epilogue, one TMEM lane quarter (32 output rows) each.  Per 128-byte column chunk (32 fp32 or
//              64 bf16 columns): tcgen05.ld, + bias, swizzled st.shared, then ONE cp.async.bulk.tensor store
//              (cp.reduce...add when accumulating) issued by an elected lane — `UTMASTG` / `UTMAREDG` in SASS.  Each warp owns two
//              4 KiB staging slabs, so chunk c+1 is being assembled while chunk c drains.
// Tiles are ordered M-fastest: CTAs that run concurrently share the B (weight) panel in L2.
// ---------------------------------------------------------------------------
namespace persist {
constexpr int BN = 256, NACC = 2;
constexpr int B_TILE = BN * BK * 2;                               // 32 KiB
constexpr int RING = STAGES * (TILE_BYTES + B_TILE);              // 192 KiB
constexpr int SLAB = 32 * 128;                                    // 32 rows x 128 bytes
constexpr int STAGING = 4 * 2 * SLAB;                             // 4 warps x 2 slabs = 32 KiB
constexpr int SMEM = RING + STAGING + 256 + 1024;
}  // namespace persist

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1, bool add) {
  if (add)
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
  else
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

template <bool C_BF16>
__global__ void __launch_bounds__(256, 1)
gemm_bf16_persist_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         const __grid_constant__ CUtensorMap tma_c, const float* __restrict__ bias, int M, int N, int K,
                         int accumulate, unsigned int* __restrict__ retire) {
  using namespace persist;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * TILE_BYTES;
  uint8_t* stage_base = smem + RING;                               // 1024-byte aligned (RING is a multiple of 1024)
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + RING + STAGING);
  uint64_t* empty = full + STAGES;
  uint64_t* acc_full = empty + STAGES;
  uint64_t* acc_empty = acc_full + NACC;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + NACC);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = (K + BK - 1) / BK;
  const int tiles_m = (M + BM - 1) / BM, tiles_n = (N + BN - 1) / BN;
  const int n_tiles = tiles_m * tiles_n;

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tma_a); tma_prefetch_desc(&tma_b); tma_prefetch_desc(&tma_c); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < NACC; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer =====
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int m0 = (tile % tiles_m) * BM, n0 = (tile / tiles_m) * BN;
      for (int kb = 0; kb < nk; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        if (!mbar_wait(&empty[s], ph ^ 1)) { atomicExch(&g_gemm_error, 1); return; }
        mbar_expect_tx(&full[s], TILE_BYTES + B_TILE);
        tma_load_2d(sA + s * TILE_BYTES, &tma_a, &full[s], kb * BK, m0);
        tma_load_2d(sB + s * B_TILE, &tma_b, &full[s], kb * BK, n0);
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = make_idesc(1, BM, BN);
    int it = 0, t_local = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++t_local) {
      const int a = t_local & 1;
      const uint32_t aph = (t_local >> 1) & 1;
      if (!mbar_wait(&acc_empty[a], aph ^ 1)) { atomicExch(&g_gemm_error, 4); return; }   // epilogue drained this accumulator
      tc_fence_after_sync();
      const uint32_t d_tmem = tmem_base + a * BN;
      for (int kb = 0; kb < nk; ++kb, ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        if (!mbar_wait(&full[s], ph)) { atomicExch(&g_gemm_error, 2); return; }
        tc_fence_after_sync();
        const uint32_t a0 = smem_u32(sA + s * TILE_BYTES), b0 = smem_u32(sB + s * B_TILE);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t da = make_smem_desc(a0 + k * 32, 16, 1024, 2);
          const uint64_t db = make_smem_desc(b0 + k * 32, 16, 1024, 2);
          mma_bf16(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        mma_commit(&empty[s]);
      }
      mma_commit(&acc_full[a]);
    }
  } else if (warp >= 4) {
    // ===== epilogue =====
    const int q = warp & 3;
    constexpr int CHUNK = C_BF16 ? 64 : 32;           // columns per 128-byte chunk
    uint8_t* slab0 = stage_base + q * 2 * SLAB;
    int t_local = 0, chunk_it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++t_local) {
      const int m0 = (tile % tiles_m) * BM, n0 = (tile / tiles_m) * BN;
      const int a = t_local & 1;
      const uint32_t aph = (t_local >> 1) & 1;
      const bool ok = mbar_wait(&acc_full[a], aph);
      if (!ok && lane == 0) atomicExch(&g_gemm_error, 3);
      tc_fence_after_sync();
      const uint32_t t_addr = tmem_base + a * BN + (uint32_t(q * 32) << 16);
      const bool row_live = m0 + q * 32 < M;          // whole slab out of range: nothing to store
#pragma unroll 1
      for (int c = 0; c < BN / CHUNK; ++c, ++chunk_it) {
        if (n0 + c * CHUNK >= N) break;
        uint8_t* slab = slab0 + (chunk_it & 1) * SLAB;
        // the store that last used this slab (two chunks ago) must have read it
        if (lane == 0) bulk_wait_read<1>();
        __syncwarp();
        uint32_t r[C_BF16 ? 64 : 32];
        {
          uint32_t (&r0)[16] = *reinterpret_cast<uint32_t (*)[16]>(&r[0]);
          uint32_t (&r1)[16] = *reinterpret_cast<uint32_t (*)[16]>(&r[16]);
          tmem_ld16_nowait(t_addr + c * CHUNK, r0);
          tmem_ld16_nowait(t_addr + c * CHUNK + 16, r1);
          if constexpr (C_BF16) {
            uint32_t (&r2)[16] = *reinterpret_cast<uint32_t (*)[16]>(&r[32]);
            uint32_t (&r3)[16] = *reinterpret_cast<uint32_t (*)[16]>(&r[48]);
            tmem_ld16_nowait(t_addr + c * CHUNK + 32, r2);
            tmem_ld16_nowait(t_addr + c * CHUNK + 48, r3);
          }
          tmem_ld_wait();
        }
        const int colbase = n0 + c * CHUNK;
        // lane = output row; 16-byte pieces go to the 128B-swizzled position the C tensor map expects
        const uint32_t row_off = smem_u32(slab) + lane * 128;
        if constexpr (!C_BF16) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 v;
            v.x = __uint_as_float(r[4 * j + 0]); v.y = __uint_as_float(r[4 * j + 1]);
            v.z = __uint_as_float(r[4 * j + 2]); v.w = __uint_as_float(r[4 * j + 3]);
            if (bias != nullptr) {
              const int col = colbase + 4 * j;
              if (col + 3 < N) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(bias + col));
                v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
              } else {
                if (col + 0 < N) v.x += __ldg(bias + col + 0);
                if (col + 1 < N) v.y += __ldg(bias + col + 1);
                if (col + 2 < N) v.z += __ldg(bias + col + 2);
              }
            }
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(row_off + ((j ^ (lane & 7)) << 4)),
                         "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint32_t pk[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int col = colbase + 8 * j + 2 * e;
              float lo = __uint_as_float(r[8 * j + 2 * e]), hi = __uint_as_float(r[8 * j + 2 * e + 1]);
              if (bias != nullptr) {
                if (col < N) lo += __ldg(bias + col);
                if (col + 1 < N) hi += __ldg(bias + col + 1);
              }
              const __nv_bfloat162 h2 = __floats2bfloat162_rn(lo, hi);
              pk[e] = *reinterpret_cast<const uint32_t*>(&h2);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_off + ((j ^ (lane & 7)) << 4)),
                         "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (ok && row_live) tma_store_2d(&tma_c, slab, colbase, m0 + q * 32, accumulate != 0);
          bulk_commit();
        }
      }
      // this warp's quarter of the accumulator is in registers / shared memory: hand it back to the MMA warp
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[a]);
      if (retire != nullptr) {   // tile-retire count (see the non-persistent kernel): once per (tile, warp)
        if (lane == 0) {
          bulk_wait_all();
          asm volatile("fence.proxy.async;" ::: "memory");
          __threadfence();
          asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(retire) : "memory");
        }
        __syncwarp();
      }
    }
    if (lane == 0) bulk_wait_all();
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem_base);
}

// ---------------------------------------------------------------------------
// CTA-pair form of the persistent kernel (tcgen05 `cta_group::2`): two CTAs of a cluster — two SMs of one TPC —
// compute one 256 x 256 output tile.  Each CTA stages ITS 128 rows of A and ITS 128 columns of B (32 KiB per stage
// instead of 48, so the ring is 6 deep), the leader CTA's single MMA thread issues M = 256 instructions that drive
// both SMs' tensor cores (twice the tensor work per issued instruction: a lone issuer no longer caps the pipe), and
// every CTA drains its own 128 x 256 half of the accumulator from its own TMEM.  Cross-CTA signalling:
//   * both producers' TMA loads complete on the LEADER's `full[s]` (cp.async.bulk.tensor ... cta_group::2 with the
//     leader's barrier address): the MMA thread waits once for all four boxes of a stage;
//   * tcgen05.commit ... multicast::cluster arrives on `empty[s]` / `acc_full[a]` of BOTH CTAs;
//   * the peer's epilogue warps arrive remotely on the leader's `acc_empty[a]`.
// All operands and the tensor maps cover this GPU's memory only.
// ---------------------------------------------------------------------------
namespace pair {
constexpr int BN = 256, NACC = 2, STG = 6;
constexpr int A_TILE = BM * BK * 2;                               // 16 KiB: my 128 rows of A
constexpr int B_TILE = (BN / 2) * BK * 2;                         // 16 KiB: my 128 columns of B
constexpr int RING = STG * (A_TILE + B_TILE);                     // 192 KiB
constexpr int SMEM = RING + persist::STAGING + 256 + 1024;
}  // namespace pair

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (an address in my own shared memory) inside CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void mma_bf16_pair(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit_pair(uint64_t* bar) {   // arrives on `bar` of both CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
// bounded wait that also gives up as soon as any role has reported a pipeline error
__device__ __forceinline__ bool mbar_wait_abortable(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return true;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 2000000000LL || *reinterpret_cast<volatile int*>(&g_gemm_error) != 0) return false;
  }
  return true;
}

template <bool C_BF16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(256, 1)
gemm_bf16_pair_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                      const __grid_constant__ CUtensorMap tma_c, const float* __restrict__ bias, int M, int N, int K,
                      int accumulate, unsigned int* __restrict__ retire) {
  using namespace pair;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STG * A_TILE;
  uint8_t* stage_base = smem + RING;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + RING + persist::STAGING);
  uint64_t* empty = full + STG;
  uint64_t* acc_full = empty + STG;
  uint64_t* acc_empty = acc_full + NACC;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + NACC);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();          // 0 = leader (issues the MMAs)
  const int pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int nk = (K + BK - 1) / BK;
  const int tiles_m = (M + 2 * BM - 1) / (2 * BM), tiles_n = (N + BN - 1) / BN;
  const int n_tiles = tiles_m * tiles_n;

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tma_a); tma_prefetch_desc(&tma_b); tma_prefetch_desc(&tma_c); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STG; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < NACC; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 8); }   // 4 warps x 2 CTAs
    fence_mbar_init();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before_sync();
  __syncthreads();
  __syncwarp();
  cluster_sync_all();     // both CTAs' barriers exist before anybody signals across the pair
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer (both CTAs): my half of each operand, completion on the LEADER's full[s] =====
    int it = 0;
    for (int tile = pair_id; tile < n_tiles; tile += n_pairs) {
      const int m0 = (tile % tiles_m) * 2 * BM + (int)rank * BM, n0 = (tile / tiles_m) * BN + (int)rank * (BN / 2);
      for (int kb = 0; kb < nk; ++kb, ++it) {
        const int s = it % STG;
        const uint32_t ph = (it / STG) & 1;
        if (!mbar_wait_abortable(&empty[s], ph ^ 1)) { atomicExch(&g_gemm_error, 1); break; }
        const uint32_t lead_full = mapa_u32(smem_u32(&full[s]), 0);
        if (rank == 0) mbar_expect_tx(&full[s], 2 * (A_TILE + B_TILE));   // four boxes land per stage
        tma_load_2d_pair(sA + s * A_TILE, &tma_a, lead_full, kb * BK, m0);
        tma_load_2d_pair(sB + s * B_TILE, &tma_b, lead_full, kb * BK, n0);
      }
    }
  } else if (warp == 1 && lane == 0 && rank == 0) {
    // ===== MMA issuer (leader CTA only) =====
    constexpr uint32_t idesc = make_idesc(1, 2 * BM, BN);
    int it = 0, t_local = 0;
    bool ok = true;
    for (int tile = pair_id; tile < n_tiles && ok; tile += n_pairs, ++t_local) {
      const int a = t_local & 1;
      const uint32_t aph = (t_local >> 1) & 1;
      if (!mbar_wait_abortable(&acc_empty[a], aph ^ 1)) { atomicExch(&g_gemm_error, 4); break; }
      tc_fence_after_sync();
      const uint32_t d_tmem = tmem_base + a * BN;
      for (int kb = 0; kb < nk; ++kb, ++it) {
        const int s = it % STG;
        const uint32_t ph = (it / STG) & 1;
        if (!mbar_wait_abortable(&full[s], ph)) { atomicExch(&g_gemm_error, 2); ok = false; break; }
        tc_fence_after_sync();
        const uint32_t a0 = smem_u32(sA + s * A_TILE), b0 = smem_u32(sB + s * B_TILE);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t da = make_smem_desc(a0 + k * 32, 16, 1024, 2);
          const uint64_t db = make_smem_desc(b0 + k * 32, 16, 1024, 2);
          mma_bf16_pair(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        mma_commit_pair(&empty[s]);
      }
      if (ok) mma_commit_pair(&acc_full[a]);
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs): my 128 rows of the tile =====
    const int q = warp & 3;
    constexpr int CHUNK = C_BF16 ? 64 : 32;
    uint8_t* slab0 = stage_base + q * 2 * persist::SLAB;
    int t_local = 0, chunk_it = 0;
    for (int tile = pair_id; tile < n_tiles; tile += n_pairs, ++t_local) {
      const int m0 = (tile % tiles_m) * 2 * BM + (int)rank * BM, n0 = (tile / tiles_m) * BN;
      const int a = t_local & 1;
      const uint32_t aph = (t_local >> 1) & 1;
      const bool ok = mbar_wait_abortable(&acc_full[a], aph);
      if (!ok && lane == 0) atomicExch(&g_gemm_error, 3);
      tc_fence_after_sync();
      const uint32_t t_addr = tmem_base + a * BN + (uint32_t(q * 32) << 16);
      const bool row_live = m0 + q * 32 < M;
#pragma unroll 1
      for (int c = 0; c < BN / CHUNK; ++c, ++chunk_it) {
        if (n0 + c * CHUNK >= N) break;
        uint8_t* slab = slab0 + (chunk_it & 1) * persist::SLAB;
        if (lane == 0) bulk_wait_read<1>();
        __syncwarp();
        uint32_t r[C_BF16 ? 64 : 32];
        {
          uint32_t (&r0)[16] = *reinterpret_cast<uint32_t (*)[16]>(&r[0]);
          uint32_t (&r1)[16] = *reinterpret_cast<uint32_t (*)[16]>(&r[16]);
          tmem_ld16_nowait(t_addr + c * CHUNK, r0);
          tmem_ld16_nowait(t_addr + c * CHUNK + 16, r1);
          if constexpr (C_BF16) {
            uint32_t (&r2)[16] = *reinterpret_cast<uint32_t (*)[16]>(&r[32]);
            uint32_t (&r3)[16] = *reinterpret_cast<uint32_t (*)[16]>(&r[48]);
            tmem_ld16_nowait(t_addr + c * CHUNK + 32, r2);
            tmem_ld16_nowait(t_addr + c * CHUNK + 48, r3);
          }
          tmem_ld_wait();
        }
        const int colbase = n0 + c * CHUNK;
        const uint32_t row_off = smem_u32(slab) + lane * 128;
        if constexpr (!C_BF16) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 v;
            v.x = __uint_as_float(r[4 * j + 0]); v.y = __uint_as_float(r[4 * j + 1]);
            v.z = __uint_as_float(r[4 * j + 2]); v.w = __uint_as_float(r[4 * j + 3]);
            if (bias != nullptr) {
              const int col = colbase + 4 * j;
              if (col + 3 < N) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(bias + col));
                v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
              } else {
                if (col + 0 < N) v.x += __ldg(bias + col + 0);
                if (col + 1 < N) v.y += __ldg(bias + col + 1);
                if (col + 2 < N) v.z += __ldg(bias + col + 2);
              }
            }
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(row_off + ((j ^ (lane & 7)) << 4)),
                         "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint32_t pk[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int col = colbase + 8 * j + 2 * e;
              float lo = __uint_as_float(r[8 * j + 2 * e]), hi = __uint_as_float(r[8 * j + 2 * e + 1]);
              if (bias != nullptr) {
                if (col < N) lo += __ldg(bias + col);
                if (col + 1 < N) hi += __ldg(bias + col + 1);
              }
              const __nv_bfloat162 h2 = __floats2bfloat162_rn(lo, hi);
              pk[e] = *reinterpret_cast<const uint32_t*>(&h2);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_off + ((j ^ (lane & 7)) << 4)),
                         "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (ok && row_live) tma_store_2d(&tma_c, slab, colbase, m0 + q * 32, accumulate != 0);
          bulk_commit();
        }
      }
      // my quarter of my CTA's accumulator half is out of TMEM: tell the leader's MMA thread
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&acc_empty[a]), 0));
      if (retire != nullptr) {
        if (lane == 0) {
          bulk_wait_all();
          asm volatile("fence.proxy.async;" ::: "memory");
          __threadfence();
          asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(retire) : "memory");
        }
        __syncwarp();
      }
    }
    if (lane == 0) bulk_wait_all();
  }
  tc_fence_before_sync();
  __syncthreads();
  __syncwarp();
  cluster_sync_all();     // nobody leaves (or frees TMEM / shared memory) while the pair still works on a tile
  if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
}

// fp32 -> bf16 cast, optionally transposing [R, C] -> [C, R] (operands of the dW GEMMs)
__global__ void cast_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = __float2bfloat16(in[i]);
}
__global__ void transpose_cast_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, int R, int C,
                                           int64_t ld_in, int64_t ld_out) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (r < R && c < C) ? in[(int64_t)r * ld_in + c] : 0.0f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (c < C && r < R) out[(int64_t)c * ld_out + r] = __float2bfloat16(tile[threadIdx.x][i]);
  }
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiled get_encode() {
  static EncodeTiled fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess) fn = (EncodeTiled)p;
  }
  return fn;
}

int make_map(CUtensorMap* map, const void* ptr, int rows, int k, int ld, int box_rows, const char* what) {
  EncodeTiled enc = get_encode();
  B200_REQUIRE(enc != nullptr, B200RNN_ECUDA, "gemm: cuTensorMapEncodeTiled entry point not available");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld % 8) == 0, B200RNN_EINVAL,
               "gemm: operand %s must be 16-byte aligned with a leading dimension multiple of 8 (ld=%d)", what, ld);
  const cuuint64_t dims[2] = {(cuuint64_t)k, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, B200RNN_ECUDA, "gemm: cuTensorMapEncodeTiled(%s) failed with %d", what, (int)r);
  return B200RNN_OK;
}

// C tile map for the persistent kernel's TMA stores: box = 128 bytes of columns x 32 rows, 128B swizzle.
int make_c_map(CUtensorMap* map, void* ptr, int rows, int cols, int ldc, bool bf16_out) {
  EncodeTiled enc = get_encode();
  B200_REQUIRE(enc != nullptr, B200RNN_ECUDA, "gemm: cuTensorMapEncodeTiled entry point not available");
  const int es = bf16_out ? 2 : 4;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ldc * es};
  const cuuint32_t box[2] = {(cuuint32_t)(128 / es), 32u};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, bf16_out ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ptr, dims,
                         strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, B200RNN_ECUDA, "gemm: cuTensorMapEncodeTiled(C) failed with %d", (int)r);
  return B200RNN_OK;
}

int pair_mode() {   // B200RNN_GEMM_PAIR=0 keeps the bulk GEMMs on the single-CTA persistent kernel
  static const int v = [] { const char* e = std::getenv("B200RNN_GEMM_PAIR"); return (e && e[0] == '0') ? 0 : 1; }();
  return v;
}
bool persist_enabled() {   // B200RNN_GEMM_PERSIST=0 keeps every GEMM on the one-tile-per-CTA kernel
  static const bool on = [] { const char* e = std::getenv("B200RNN_GEMM_PERSIST"); return !(e && e[0] == '0'); }();
  return on;
}

}  // namespace
}  // namespace b200rnn

using namespace b200rnn;

extern "C" int b200rnn_gemm_bf16_retire(const void* A, const void* B, void* C, const float* bias, int32_t M, int32_t N,
                                        int32_t K, int32_t lda, int32_t ldb, int32_t ldc, int32_t accumulate,
                                        int32_t c_is_bf16, uint32_t* retire_counter, int32_t* retire_count, void* stream);

extern "C" int b200rnn_gemm_bf16(const void* A, const void* B, void* C, const float* bias, int32_t M, int32_t N,
                                 int32_t K, int32_t lda, int32_t ldb, int32_t ldc, int32_t accumulate,
                                 int32_t c_is_bf16, void* stream) {
  return b200rnn_gemm_bf16_retire(A, B, C, bias, M, N, K, lda, ldb, ldc, accumulate, c_is_bf16, nullptr, nullptr, stream);
}

extern "C" int b200rnn_gemm_bf16_retire(const void* A, const void* B, void* C, const float* bias, int32_t M, int32_t N,
                                        int32_t K, int32_t lda, int32_t ldb, int32_t ldc, int32_t accumulate,
                                        int32_t c_is_bf16, uint32_t* retire_counter, int32_t* retire_count, void* stream) {
  B200_REQUIRE(A && B && C && M > 0 && N > 0 && K > 0, B200RNN_EINVAL, "gemm_bf16: bad argument");
  const bool skinny = (M <= BM) && ((N + 127) / 128 < num_sms() / 2) && N >= 32;
  const int64_t ctas256 = (int64_t)((N + 255) / 256) * ((M + BM - 1) / BM);
  const bool wide = !skinny && N >= 256 && ctas256 >= num_sms() / 2;
  const int bn = skinny ? 32 : (wide ? 256 : 128);
  CUtensorMap ma, mb;
  // bulk shapes with a TMA-storable C (16-byte aligned rows): the persistent kernel
  const int es = c_is_bf16 ? 2 : 4;
  if (wide && persist_enabled() && (reinterpret_cast<uintptr_t>(C) & 15) == 0 && ((int64_t)ldc * es) % 16 == 0) {
    CUtensorMap mc;
    int rc2 = make_map(&ma, A, M, K, lda, BM, "A");
    if (rc2) return rc2;
    rc2 = make_map(&mb, B, N, K, ldb, persist::BN, "B");
    if (rc2) return rc2;
    rc2 = make_c_map(&mc, C, M, N, ldc, c_is_bf16 != 0);
    if (rc2) return rc2;
    static bool pattr = false;
    if (!pattr) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_persist_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, persist::SMEM));
      B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_persist_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, persist::SMEM));
      pattr = true;
    }
    if (pair_mode() && M >= 2 * BM) {
      // CTA-pair kernel: B box = my 128 columns
      rc2 = make_map(&mb, B, N, K, ldb, pair::BN / 2, "B");
      if (rc2) return rc2;
      static bool pair_attr = false;
      if (!pair_attr) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, pair::SMEM));
        B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, pair::SMEM));
        pair_attr = true;
      }
      const int tiles2 = ((M + 2 * BM - 1) / (2 * BM)) * ((N + pair::BN - 1) / pair::BN);
      const int pairs = tiles2 < num_sms() / 2 ? tiles2 : num_sms() / 2;
      if (retire_count) *retire_count = 8 * tiles2;   // 4 epilogue warps x 2 CTAs per tile
      if (c_is_bf16) gemm_bf16_pair_kernel<true><<<2 * pairs, 256, pair::SMEM, (cudaStream_t)stream>>>(ma, mb, mc, bias, M, N, K, accumulate, retire_counter);
      else gemm_bf16_pair_kernel<false><<<2 * pairs, 256, pair::SMEM, (cudaStream_t)stream>>>(ma, mb, mc, bias, M, N, K, accumulate, retire_counter);
      B200_CHECK_LAUNCH();
      return B200RNN_OK;
    }
    const int tiles = ((M + BM - 1) / BM) * ((N + persist::BN - 1) / persist::BN);
    const int grid_p = tiles < num_sms() ? tiles : num_sms();
    if (retire_count) *retire_count = 4 * tiles;   // every epilogue warp counts each of its tile quarters
    if (c_is_bf16) gemm_bf16_persist_kernel<true><<<grid_p, 256, persist::SMEM, (cudaStream_t)stream>>>(ma, mb, mc, bias, M, N, K, accumulate, retire_counter);
    else gemm_bf16_persist_kernel<false><<<grid_p, 256, persist::SMEM, (cudaStream_t)stream>>>(ma, mb, mc, bias, M, N, K, accumulate, retire_counter);
    B200_CHECK_LAUNCH();
    return B200RNN_OK;
  }
  int rc = make_map(&ma, A, M, K, lda, BM, "A");
  if (rc) return rc;
  rc = make_map(&mb, B, N, K, ldb, bn, "B");
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes<128>()));
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes<32>()));
    B200_CHECK_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes<256>()));
    attr_set = true;
  }
  dim3 grid((N + bn - 1) / bn, (M + BM - 1) / BM);
  if (retire_count) *retire_count = (int32_t)(grid.x * grid.y);   // increments this launch adds to retire_counter
  unsigned int* rc_ptr = retire_counter;
  if (skinny) gemm_bf16_kernel<32><<<grid, 256, smem_bytes<32>(), (cudaStream_t)stream>>>(ma, mb, C, bias, M, N, K, ldc, accumulate, c_is_bf16, rc_ptr);
  else if (wide) gemm_bf16_kernel<256><<<grid, 256, smem_bytes<256>(), (cudaStream_t)stream>>>(ma, mb, C, bias, M, N, K, ldc, accumulate, c_is_bf16, rc_ptr);
  else gemm_bf16_kernel<128><<<grid, 256, smem_bytes<128>(), (cudaStream_t)stream>>>(ma, mb, C, bias, M, N, K, ldc, accumulate, c_is_bf16, rc_ptr);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_gemm_status(int32_t* status) {
  B200_REQUIRE(status != nullptr, B200RNN_EINVAL, "gemm_status: null");
  int v = 0;
  B200_CHECK_CUDA(cudaMemcpyFromSymbol(&v, g_gemm_error, sizeof(int)));
  *status = v;
  return B200RNN_OK;
}

extern "C" int b200rnn_cast_bf16(const float* in, void* out, int64_t n, void* stream) {
  B200_REQUIRE(in && out && n > 0, B200RNN_EINVAL, "cast_bf16: bad argument");
  int64_t blocks = (n + 255) / 256;
  if (blocks > 4096) blocks = 4096;
  cast_bf16_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(in, (__nv_bfloat16*)out, n);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_transpose_cast_bf16(const float* in, void* out, int32_t rows, int32_t cols, int64_t ld_in,
                                           int64_t ld_out, void* stream) {
  B200_REQUIRE(in && out && rows > 0 && cols > 0, B200RNN_EINVAL, "transpose_cast_bf16: bad argument");
  dim3 grid((cols + 31) / 32, (rows + 31) / 32), block(32, 8);
  transpose_cast_bf16_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(in, (__nv_bfloat16*)out, rows, cols, ld_in, ld_out);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}
