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
    const int row = m0 + q * 32 + lane;
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
