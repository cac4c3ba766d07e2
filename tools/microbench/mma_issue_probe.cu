// Probe: cost per tcgen05.mma at the small shapes the persistent RNN kernels use.
// One CTA; operands are zeroed shared memory (128B-swizzled K-major tiles); cycles measured around
// NMMA back-to-back MMAs (issue time) and until their commit arrives (completion time).
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I csrc -o build/mma_issue_probe tools/microbench/mma_issue_probe.cu
#include <cstdio>
#include <cstdlib>
#include "gemm/tc_common.cuh"
using namespace b200rnn::tc;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

constexpr int NMMA = 256;
// nissue warps (0..nissue-1) each issue NMMA/nissue MMAs; accumulators rotate over `nacc` column groups
template <int M, int N>
__global__ void __launch_bounds__(288, 1) probe(int nissue, int nacc, int same_addr, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                   // 4 chunks of [M][64] bf16
  uint8_t* sB = smem + 4 * M * 128;     // 2 chunks of [<=256][64]
  uint64_t* bars = (uint64_t*)(sB + 2 * 256 * 128);
  uint32_t* slot = (uint32_t*)(bars + 8);
  for (int i = threadIdx.x; i < (4 * M * 128 + 2 * 256 * 128) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1); fence_mbar_init(); }
  if (warp == 8) tmem_alloc<512>(slot);
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = *slot;
  constexpr uint32_t idesc = make_idesc(1, M, N);
  constexpr int stride = N < 32 ? 32 : N;
  long long t0 = 0, t1 = 0, t2 = 0;
  if (warp < nissue && lane == 0) {
    const int per = NMMA / nissue;
    t0 = clock64();
    for (int i = 0; i < per; ++i) {
      const int chunk = same_addr ? 0 : (i >> 2) & 3, k = same_addr ? 0 : i & 3;
      const uint64_t da = make_smem_desc(smem_u32(sA + chunk * M * 128) + k * 32, 16, 1024, 2);
      const uint64_t db = make_smem_desc(smem_u32(sB + (chunk & 1) * 256 * 128) + k * 32, 16, 1024, 2);
      const int acc = (warp * (nacc / nissue > 0 ? nacc / nissue : 1) + i % (nacc / nissue > 0 ? nacc / nissue : 1));
      mma_bf16(tm + (uint32_t)((acc * stride) % 512), da, db, idesc, i > 0 ? 1u : 0u);
    }
    t1 = clock64();
    mma_commit(&bars[warp]);
    mbar_wait(&bars[warp], 0);
    t2 = clock64();
    out[warp * 2] = t1 - t0;
    out[warp * 2 + 1] = t2 - t0;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 8) tmem_dealloc<512>(tm);
}

template <int M, int N>
void run(int nissue, int nacc, int same_addr, long long* dout) {
  const size_t smem = 4 * M * 128 + 2 * 256 * 128 + 2048;
  CK(cudaFuncSetAttribute(probe<M, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  for (int w = 0; w < 2; ++w) probe<M, N><<<1, 288, smem>>>(nissue, nacc, same_addr, dout);
  CK(cudaDeviceSynchronize());
  long long h[16];
  CK(cudaMemcpy(h, dout, sizeof(h), cudaMemcpyDeviceToHost));
  const int per = NMMA / nissue;
  printf("M %3d N %3d issuers %d accumulators %d same_addr %d: issue %.1f cyc/MMA, complete %.1f cyc/MMA (per issuer; %d MMAs each)\n",
         M, N, nissue, nacc, same_addr, (double)h[0] / per, (double)h[1] / per, per);
}

int main() {
  long long* dout;
  CK(cudaMalloc(&dout, 128));
  run<64, 32>(1, 1, 0, dout);
  run<64, 32>(2, 2, 0, dout);
  run<64, 32>(4, 4, 0, dout);
  run<64, 32>(8, 8, 0, dout);
  run<64, 8>(4, 4, 0, dout);
  run<64, 8>(8, 8, 0, dout);
  run<128, 32>(4, 4, 0, dout);
  run<128, 32>(8, 8, 0, dout);
  run<128, 16>(8, 8, 0, dout);
  run<128, 256>(2, 2, 0, dout);
  return 0;
}
