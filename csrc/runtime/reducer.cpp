// N7 — bucket planning of the gradient reducer (host C++).
// c10d's Reducer, which the reference uses through DistributedDataParallel(model)
// (src/motion/trainer/ddp.py:19), assigns parameters to buckets in REVERSE registration order
// (gradients become ready back to front) with a small first bucket (1 MiB) so communication
// starts early and a 25 MiB cap afterwards.  Same policy here; a bucket is a contiguous slice
// of the flat gradient arena, so "flatten / unflatten" costs nothing.
#include "../common.cuh"

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
