// N7 — the gradient reducer: bucket plan + device-side "bucket ready" triggers.
//
// c10d's Reducer, which the reference uses through DistributedDataParallel(model)
// (src/motion/trainer/ddp.py:19), does two things: it assigns parameters to buckets in REVERSE
// registration order (gradients become ready back to front; 1 MiB first bucket, 25 MiB cap), and it
// hooks every gradient so that a bucket's all-reduce starts DURING loss.backward(), the moment the
// bucket's last gradient exists.  Here a bucket is a contiguous slice of the flat gradient arena
// (so flatten / unflatten cost nothing) and "the last gradient exists" is observed on the device:
// every kernel that writes part of a bucket's gradient (the weight-gradient GEMMs tile by tile, the
// bias sums and the head's weight gradient block by block) adds one count to the bucket's counter as
// it retires (red.release.gpu), and a one-thread wait kernel at the head of the bucket's all-reduce on
// the communication stream spins on that counter (ld.acquire.gpu).  No host event, no stream
// dependency: the collective of layer l starts while the backward recurrence of layer l-1 is running
// on the compute stream.  Counters only grow (target += expected counts per step), so nothing is
// reset between steps.  The wait is bounded (status word) and stamps %globaltimer when it fires —
// together with b200rnn_reducer_stamp this yields the overlap timeline without a profiler.
#include <vector>
#include "../common.cuh"

namespace b200rnn {
namespace {

constexpr int kStampsPerBucket = 4;   // 0: trigger seen by the wait kernel; 1..3: user marks (b200rnn_reducer_stamp)

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void reducer_wait_kernel(const unsigned int* __restrict__ counter, unsigned int target, long long budget,
                                    unsigned long long* __restrict__ stamp, int* __restrict__ status, int bucket) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const long long t0 = clock64();
  for (;;) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
    if ((int)(v - target) >= 0) break;
    if (clock64() - t0 > budget) { atomicOr(status, 1 << (bucket & 30)); break; }
    __nanosleep(200);
  }
  *stamp = globaltimer_ns();
}

__global__ void reducer_stamp_kernel(unsigned long long* __restrict__ stamp) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *stamp = globaltimer_ns();
}

}  // namespace
}  // namespace b200rnn

using namespace b200rnn;

struct b200rnn_reducer {
  int n_buckets = 0;
  std::vector<int64_t> lo, hi;
  std::vector<unsigned int> target;        // counts the producers have promised so far, per bucket
  unsigned int* counters = nullptr;        // device [n_buckets]
  unsigned long long* stamps = nullptr;    // device [n_buckets][kStampsPerBucket]
  int* status = nullptr;                   // device
  long long budget = 120000000000LL;       // ~60 s of SM clocks
};

extern "C" int b200rnn_bucket_plan(const int64_t* sizes_bytes, int32_t n, int64_t first_cap, int64_t cap,
                                   int32_t* bucket_of /* [n] */, int32_t* n_buckets) {
  B200_REQUIRE(sizes_bytes && bucket_of && n_buckets && n >= 0 && first_cap > 0 && cap > 0, B200RNN_EINVAL,
               "bucket_plan: bad argument");
  int32_t cur = 0;
  int64_t cur_bytes = 0, limit = first_cap;
  bool open = false;
  for (int32_t i = n - 1; i >= 0; --i) {
    if (open && cur_bytes + sizes_bytes[i] > limit) { ++cur; cur_bytes = 0; limit = cap; }
    bucket_of[i] = cur;
    cur_bytes += sizes_bytes[i];
    open = true;
  }
  *n_buckets = n > 0 ? cur + 1 : 0;
  return B200RNN_OK;
}

extern "C" int b200rnn_reducer_create(int32_t n_buckets, const int64_t* lo, const int64_t* hi, b200rnn_reducer** out) {
  B200_REQUIRE(out && lo && hi && n_buckets >= 1 && n_buckets <= 30, B200RNN_EINVAL, "reducer_create: bad argument");
  for (int i = 0; i < n_buckets; ++i)
    B200_REQUIRE(lo[i] >= 0 && hi[i] > lo[i], B200RNN_EINVAL, "reducer_create: bucket %d has an empty range", i);
  b200rnn_reducer* r = new b200rnn_reducer();
  r->n_buckets = n_buckets;
  r->lo.assign(lo, lo + n_buckets);
  r->hi.assign(hi, hi + n_buckets);
  r->target.assign(n_buckets, 0u);
  const size_t bytes = n_buckets * (sizeof(unsigned int) + kStampsPerBucket * sizeof(unsigned long long)) + 64;
  char* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e == cudaSuccess) e = cudaMemset(p, 0, bytes);
  if (e != cudaSuccess) {
    if (p) cudaFree(p);
    delete r;
    set_error("reducer_create: %s", cudaGetErrorString(e));
    return B200RNN_ECUDA;
  }
  r->stamps = reinterpret_cast<unsigned long long*>(p);
  r->counters = reinterpret_cast<unsigned int*>(p + n_buckets * kStampsPerBucket * sizeof(unsigned long long));
  r->status = reinterpret_cast<int*>(r->counters + n_buckets);
  *out = r;
  return B200RNN_OK;
}

extern "C" int b200rnn_reducer_destroy(b200rnn_reducer* r) {
  if (r == nullptr) return B200RNN_OK;
  if (r->stamps) cudaFree(r->stamps);
  delete r;
  return B200RNN_OK;
}

extern "C" int b200rnn_reducer_range(const b200rnn_reducer* r, int32_t bucket, int64_t* lo, int64_t* hi) {
  B200_REQUIRE(r && bucket >= 0 && bucket < r->n_buckets && lo && hi, B200RNN_EINVAL, "reducer_range: bad argument");
  *lo = r->lo[bucket]; *hi = r->hi[bucket];
  return B200RNN_OK;
}

extern "C" uint32_t* b200rnn_reducer_counter(b200rnn_reducer* r, int32_t bucket) {
  if (r == nullptr || bucket < 0 || bucket >= r->n_buckets) return nullptr;
  return r->counters + bucket;
}

extern "C" int b200rnn_reducer_expect(b200rnn_reducer* r, int32_t bucket, int32_t counts) {
  B200_REQUIRE(r && bucket >= 0 && bucket < r->n_buckets && counts >= 0, B200RNN_EINVAL, "reducer_expect: bad argument");
  r->target[bucket] += (unsigned int)counts;
  return B200RNN_OK;
}

extern "C" int b200rnn_reducer_wait(b200rnn_reducer* r, int32_t bucket, void* stream) {
  B200_REQUIRE(r && bucket >= 0 && bucket < r->n_buckets, B200RNN_EINVAL, "reducer_wait: bad argument");
  reducer_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(r->counters + bucket, r->target[bucket], r->budget,
                                                          r->stamps + bucket * kStampsPerBucket, r->status, bucket);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_reducer_stamp(b200rnn_reducer* r, int32_t bucket, int32_t which, void* stream) {
  B200_REQUIRE(r && bucket >= 0 && bucket < r->n_buckets && which >= 1 && which < kStampsPerBucket, B200RNN_EINVAL,
               "reducer_stamp: bad argument");
  reducer_stamp_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(r->stamps + bucket * kStampsPerBucket + which);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_reducer_read(b200rnn_reducer* r, uint64_t* stamps_out /* [n_buckets][4] ns */, int32_t* status_out) {
  B200_REQUIRE(r != nullptr, B200RNN_EINVAL, "reducer_read: null");
  B200_CHECK_CUDA(cudaDeviceSynchronize());
  if (stamps_out)
    B200_CHECK_CUDA(cudaMemcpy(stamps_out, r->stamps, (size_t)r->n_buckets * kStampsPerBucket * sizeof(unsigned long long),
                               cudaMemcpyDeviceToHost));
  if (status_out) B200_CHECK_CUDA(cudaMemcpy(status_out, r->status, sizeof(int), cudaMemcpyDeviceToHost));
  return B200RNN_OK;
}
