// Probe: cost of issuing one bulk copy, latency of one isolated copy, and throughput with 1 vs 4 issuing
// warps (each warp's lane 0 owns its own ring).  One CTA per SM, `ncta` CTAs, plain (non-cooperative) launch.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
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
struct Out { long long issue_cyc, single_lat, pass_cyc; };
// nwarps issuing warps; each issues `per` copies of `bytes` with ring depth `depth`
__global__ void __launch_bounds__(128, 1) probe(const uint8_t* base0, size_t region, int bytes, int per, int depth, int nwarps, int reps, Out* out, int priv) {
  const uint8_t* base = base0 + (priv ? (size_t)blockIdx.x * 131072 : 0);
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint64_t* bars = (uint64_t*)(smem + (size_t)nwarps * depth * bytes);
  if (threadIdx.x == 0) {
    for (int i = 0; i < nwarps * depth + 1; ++i) mbar_init(&bars[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  long long issue = 0, lat = 0;
  if (threadIdx.x == 0) {   // isolated copy: issue cost and latency
    uint64_t* b = &bars[nwarps * depth];
    mbar_expect_tx(b, bytes);
    const long long t0 = clock64();
    bulk_1d(smem, base, bytes, b);
    const long long t1 = clock64();
    while (!mbar_test(b, 0)) {}
    lat = clock64() - t0; issue = t1 - t0;
  }
  __syncthreads();
  const long long p0 = clock64();
  if (warp < nwarps && lane == 0) {
    uint8_t* ring = smem + (size_t)warp * depth * bytes;
    uint64_t* full = bars + warp * depth;
    const uint8_t* src = base + ((size_t)warp * per * bytes) % 131072;
    for (int r = 0; r < reps; ++r) {
      int issued = 0, done = 0;
      while (done < per) {
        while (issued < per && issued - done < depth) {
          const int s = issued % depth;
          mbar_expect_tx(&full[s], bytes);
          bulk_1d(ring + s * bytes, src + (size_t)issued * bytes, bytes, &full[s]);
          ++issued;
        }
        const uint32_t ph = ((r * per + done) / depth) & 1;
        while (!mbar_test(&full[done % depth], ph)) {}
        ++done;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) { out[blockIdx.x].issue_cyc = issue; out[blockIdx.x].single_lat = lat; out[blockIdx.x].pass_cyc = (clock64() - p0) / reps; }
}
int main() {
  const size_t region = 64 << 20;
  uint8_t* buf; Out* out;
  CK(cudaMalloc(&buf, region)); CK(cudaMemset(buf, 1, region)); CK(cudaMalloc(&out, 148 * sizeof(Out)));
  const int reps = 100;
  for (int priv = 0; priv < 2; ++priv)
  for (int ncta : {1, 32, 128})
    for (int bytes : {8192, 32768})
      for (int nwarps : {1, 4}) {
        const int depth = 1;
        const int per = 131072 / bytes / nwarps;   // 128 KiB per CTA per pass
        if (per < 1) continue;
        const size_t ring = (size_t)nwarps * depth * bytes;
        const size_t smem = ring + 2048 + (ring < 120 * 1024 ? 100 * 1024 : 0);
        CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        probe<<<ncta, 128, smem>>>(buf, region, bytes, per, depth, nwarps, reps, out, priv);
        CK(cudaDeviceSynchronize());
        std::vector<Out> h(ncta);
        CK(cudaMemcpy(h.data(), out, ncta * sizeof(Out), cudaMemcpyDeviceToHost));
        double pc = 0; for (auto& o : h) pc += o.pass_cyc; pc /= ncta;
        printf("%s ncta %3d bytes %5d warps %d depth %d: pass(128 KiB) %.0f cyc -> %.1f B/clk/SM, %.0f B/clk chip\n",
               priv ? "private" : "shared ", ncta, bytes, nwarps, depth, pc, 131072.0 / pc, 131072.0 / pc * ncta);
      }
  return 0;
}
