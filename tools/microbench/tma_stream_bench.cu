// Microbenchmark behind the persistent-RNN design: how fast can every SM stream the SAME
// [rows x K] bf16 operand out of L2 each "step"?  128 CTAs (one per SM), one producer thread each,
// a ring of DEPTH stages, REPS passes over the region.  Variants:
//   0  TMA tensor tiles {64 cols x ROWS} 128B-swizzled, all CTAs read the same region
//   1  same tiles, every CTA reads a private region
//   2  cp.async.bulk (1-D, non-tensor) of contiguous 8 KiB pieces of the shared region
//   3  as 0, but the region is rewritten by the grid between passes (like h_t in the real kernel)
//   4  as 0, waiting with mbarrier.test_wait (pure spin) instead of try_wait
//   5  as 0, all loads of a pass issued before the first wait is irrelevant here (depth covers it); 32-row tiles
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o build/tma_stream_bench tools/microbench/tma_stream_bench.cu -lcuda
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint64_t* b, uint32_t ph) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(ph) : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_test(uint64_t* b, uint32_t ph) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(ph) : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_wait(uint64_t* b, uint32_t ph, bool spin) {
  const long long t0 = clock64();
  if (spin) { while (!mbar_test(b, ph)) if (clock64() - t0 > 400000000LL) return false; }
  else { while (!mbar_try(b, ph)) if (clock64() - t0 > 400000000LL) return false; }
  return true;
}
__device__ __forceinline__ void tma_3d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
               ::"r"(smem_u32(dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void bulk_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct Args { const __nv_bfloat16* base; __nv_bfloat16* wbase; unsigned* sync; long long* cycles; int rows, K, reps, depth, variant; };

__global__ void __launch_bounds__(128, 1) bench(const __grid_constant__ CUtensorMap map, const Args a) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  const int tile = a.rows * 128;
  uint64_t* full = (uint64_t*)(smem + a.depth * tile);
  const int nk = a.K / 64;
  if (threadIdx.x == 0) {
    for (int i = 0; i < a.depth; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  long long t0 = 0, total = 0;
  for (int r = 0; r < a.reps; ++r) {
    if (a.variant == 3) {
      // every CTA rewrites its 8 columns of the region, then a grid barrier (as the real kernel does)
      if (threadIdx.x < a.rows) {
        uint4 v = make_uint4(r, r + 1, r + 2, r + 3);
        *(uint4*)(a.wbase + (size_t)threadIdx.x * a.K + blockIdx.x * 8) = v;
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(a.sync), "r"(1u) : "memory");
        unsigned v;
        const long long w0 = clock64();
        do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.sync) : "memory"); } while (v < (unsigned)(r + 1) * gridDim.x && clock64() - w0 < 400000000LL);
        asm volatile("fence.proxy.async;" ::: "memory");
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      t0 = clock64();
      const int z = (a.variant == 1) ? blockIdx.x : 0;
      int issued = 0, done = 0;
      while (done < nk) {
        while (issued < nk && issued - done < a.depth) {
          const int s = issued % a.depth;
          const int kc = (issued + blockIdx.x * 5) % nk;
          mbar_expect_tx(&full[s], tile);
          if (a.variant == 2) bulk_1d(smem + s * tile, (const uint8_t*)a.base + (size_t)kc * tile, tile, &full[s]);
          else tma_3d(smem + s * tile, &map, &full[s], kc * 64, 0, z);
          ++issued;
        }
        const int s = done % a.depth;
        const uint32_t ph = ((r * nk + done) / a.depth) & 1;
        if (!mbar_wait(&full[s], ph, a.variant == 4)) { a.cycles[blockIdx.x] = -1; return; }
        ++done;
      }
      total += clock64() - t0;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) a.cycles[blockIdx.x] = total;
}

int main(int argc, char** argv) {
  const int rows = argc > 1 ? atoi(argv[1]) : 64, K = argc > 2 ? atoi(argv[2]) : 1024, reps = 200, ncta = 128;
  typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                          const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
  Enc enc = (Enc)p;
  __nv_bfloat16* buf; unsigned* sync; long long* cyc;
  const size_t region = (size_t)rows * K * 2;
  CK(cudaMalloc(&buf, region * ncta)); CK(cudaMemset(buf, 0, region * ncta));
  CK(cudaMalloc(&sync, 256)); CK(cudaMalloc(&cyc, ncta * 8));
  CUtensorMap map;
  const cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)rows, (cuuint64_t)ncta};
  const cuuint64_t strides[2] = {(cuuint64_t)K * 2, (cuuint64_t)region};
  const cuuint32_t box[3] = {64, (cuuint32_t)rows, 1}, es[3] = {1, 1, 1};
  if (enc(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, buf, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); return 1; }
  const int depths[3] = {4, 8, 16};
  for (int variant = 0; variant < 5; ++variant)
    for (int di = 0; di < 3; ++di) {
      const int depth = depths[di];
      const size_t smem = (size_t)depth * rows * 128 + 1024 + 256 + 100 * 1024;   // +100 KiB: force one CTA per SM
      if (smem > 227 * 1024) continue;
      CK(cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      Args a{buf, buf, sync, cyc, rows, K, reps, depth, variant};
      CK(cudaMemset(sync, 0, 256));
      void* args[2] = {&map, &a};
      CK(cudaLaunchCooperativeKernel((void*)bench, dim3(ncta), dim3(128), args, smem, 0));
      CK(cudaDeviceSynchronize());
      std::vector<long long> h(ncta);
      CK(cudaMemcpy(h.data(), cyc, ncta * 8, cudaMemcpyDeviceToHost));
      long long mx = 0, sum = 0; bool bad = false;
      for (long long v : h) { if (v < 0) bad = true; if (v > mx) mx = v; sum += v; }
      const double per_pass = (double)sum / ncta / reps;
      printf("variant %d depth %2d rows %d K %d: %s avg %.0f cyc/pass (max CTA %.0f) -> %.1f B/clk/SM, %.0f B/clk chip\n", variant, depth, rows, K,
             bad ? "TIMEOUT" : "ok", per_pass, (double)mx / reps, region / per_pass, region / per_pass * ncta);
    }
  return 0;
}
