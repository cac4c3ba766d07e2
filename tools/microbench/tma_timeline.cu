// Timeline of 8 back-to-back bulk copies issued by one thread: issue begin/end and completion time of each.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ bool mbar_test(uint64_t* b, uint32_t ph) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(ph) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bulk_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
constexpr int NL = 8;
// mode 0: one thread issues all, then the SAME thread polls.  mode 1: one thread issues, warp 1 lane 0 polls.
// mode 2: one thread issues with a separate expect_tx pass first (all expects, then all copies).
__global__ void __launch_bounds__(64, 1) timeline(const uint8_t* base, int bytes, int mode, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = (uint64_t*)(smem + NL * bytes);
  __shared__ long long t_is[NL], t_ie[NL], t_c[NL], t00;
  if (threadIdx.x == 0) {
    for (int i = 0; i < NL; ++i) mbar_init(&bars[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  for (int rep = 0; rep < 3; ++rep) {   // last repetition is the one reported (warm)
    if (threadIdx.x == 0) {
      t00 = clock64();
      if (mode == 2) for (int i = 0; i < NL; ++i) mbar_expect_tx(&bars[i], bytes);
      for (int i = 0; i < NL; ++i) {
        t_is[i] = clock64();
        if (mode != 2) mbar_expect_tx(&bars[i], bytes);
        bulk_1d(smem + i * bytes, base + (size_t)(rep * NL + i) * bytes, bytes, &bars[i]);
        t_ie[i] = clock64();
      }
    }
    const bool poller = (mode == 1) ? (threadIdx.x == 32) : (threadIdx.x == 0);
    if (poller) {
      for (int i = 0; i < NL; ++i) {
        while (!mbar_test(&bars[i], rep & 1)) {}
        t_c[i] = clock64();
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0)
    for (int i = 0; i < NL; ++i) { out[3 * i] = t_is[i] - t00; out[3 * i + 1] = t_ie[i] - t00; out[3 * i + 2] = t_c[i] - t00; }
}
// Ring of depth D reused for NC copies by one thread (issue -> wait oldest -> reissue), as a pipelined producer does.
constexpr int NC = 24;
__global__ void __launch_bounds__(64, 1) ring_timeline(const uint8_t* base, int bytes, int depth, int use_try, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = (uint64_t*)(smem + depth * bytes);
  __shared__ long long t_i[NC], t_c[NC];
  if (threadIdx.x == 0) {
    for (int i = 0; i < depth; ++i) mbar_init(&bars[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long t00 = clock64();
    int issued = 0, done = 0;
    while (done < NC) {
      while (issued < NC && issued - done < depth) {
        const int s = issued % depth;
        mbar_expect_tx(&bars[s], bytes);
        bulk_1d(smem + s * bytes, base + (size_t)issued * bytes, bytes, &bars[s]);
        t_i[issued] = clock64() - t00;
        ++issued;
      }
      const uint32_t ph = (done / depth) & 1;
      if (use_try) {
        uint32_t ok = 0;
        while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(&bars[done % depth])), "r"(ph) : "memory");
      } else {
        while (!mbar_test(&bars[done % depth], ph)) {}
      }
      t_c[done] = clock64() - t00;
      ++done;
    }
    for (int i = 0; i < NC; ++i) { out[2 * i] = t_i[i]; out[2 * i + 1] = t_c[i]; }
  }
}
int main() {
  {
    uint8_t* buf; long long* out;
    CK(cudaMalloc(&buf, 8 << 20)); CK(cudaMemset(buf, 1, 8 << 20)); CK(cudaMalloc(&out, 2 * NC * 8));
    for (int use_try = 0; use_try < 2; ++use_try)
      for (int depth : {4, 8}) {
        const int bytes = 8192;
        const size_t smem = (size_t)depth * bytes + 2048;
        CK(cudaFuncSetAttribute(ring_timeline, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        for (int w = 0; w < 2; ++w) ring_timeline<<<1, 64, smem>>>(buf, bytes, depth, use_try, out);
        CK(cudaDeviceSynchronize());
        long long h[2 * NC];
        CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
        printf("ring depth %d try_wait %d:", depth, use_try);
        for (int i = 0; i < NC; ++i) printf(" [i%lld c%lld]", h[2 * i], h[2 * i + 1]);
        printf("\n");
      }
  }
  uint8_t* buf; long long* out;
  CK(cudaMalloc(&buf, 8 << 20)); CK(cudaMemset(buf, 1, 8 << 20)); CK(cudaMalloc(&out, 3 * NL * 8));
  for (int bytes : {2048, 8192, 16384})
    for (int mode = 0; mode < 3; ++mode) {
      const size_t smem = (size_t)NL * bytes + 2048;
      CK(cudaFuncSetAttribute(timeline, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      timeline<<<1, 64, smem>>>(buf, bytes, mode, out);
      CK(cudaDeviceSynchronize());
      long long h[3 * NL];
      CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
      printf("bytes %5d mode %d:", bytes, mode);
      for (int i = 0; i < NL; ++i) printf("  [%lld-%lld c%lld]", h[3 * i], h[3 * i + 1], h[3 * i + 2]);
      printf("\n");
    }
  return 0;
}
