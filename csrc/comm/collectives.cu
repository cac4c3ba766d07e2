// K7-K9 — collectives issued from INSIDE kernels over NVLink / NVSwitch peer memory.
//
// Every rank owns one symmetric region (flags | inbox | heap); every region is mapped into
// every rank (CUDA IPC, csrc/runtime/symm.cpp), so a kernel on rank r can store straight into
// rank p's memory and raise a flag there.  No host involvement, no NCCL: the reference's
// equivalents are c10d Reducer -> MPI_Allreduce / MPI_Bcast / MPI_Send/Recv on CPU tensors
// (src/motion/trainer/ddp.py:18-19, src/example/example_distributed.py:11-14) and Horovod's
// fusion buffer + MPI allreduce (trainer/horovod.py:35).
//
// Protocol (all flags are monotonically increasing sequence numbers, never reset):
//   writer : data stores to the peer -> __threadfence_system -> bar.sync -> st.release.sys flag
//   reader : ld.acquire.sys spin on its OWN memory until flag >= seq -> bar.sync -> ld.cg data
// Inboxes are double-buffered by the parity of seq; a slot is only overwritten two calls
// later, which the peers' own progress already orders after the last read.
// Every spin has a clock64 watchdog that records the missing source in a status word instead
// of hanging the GPU.  Sums are taken in rank order 0..W-1 on every rank => bit-identical
// results everywhere (the reference's "equal final parameters" criterion, README.md:9).
#include "../common.cuh"
#include "../optim.cuh"
#include "../rows.cuh"
#include <cstdlib>

namespace b200rnn {
namespace {

constexpr int kThreads = 256;
constexpr int kElems = 4;                       // elements per thread in the two-shot kernel
constexpr int kBlockElems = kThreads * kElems;  // 1024 parameters per two-shot block
// Watchdog of every in-kernel wait, in SM clocks (~2 GHz).  Ranks may legitimately be seconds apart (rank 0
// validating / checkpointing while the others already wait in the next step's all-reduce, a first-step lazy
// module load), so the default is generous: 60 s.  B200RNN_COMM_TIMEOUT_MS overrides it for the process,
// ctx->timeout_cycles per communicator.  A wait that does time out sets a status bit (b200rnn_comm_status);
// the values computed by that call are garbage and the trainers abort at the next check.
constexpr long long kDefaultTimeout = 120000000000LL;
long long default_timeout_cycles() {
  static const long long v = [] {
    const char* e = std::getenv("B200RNN_COMM_TIMEOUT_MS");
    if (e != nullptr && e[0] != 0) {
      const long long ms = std::atoll(e);
      if (ms > 0) return ms * 2000000LL;
    }
    return kDefaultTimeout;
  }();
  return v;
}

struct Ctx {  // device-side view
  int rank, world, max_blocks, delay_ns, loss_ppm, loss_rto_ns;
  long long inbox_floats, timeout, ll_floats;
  char* base[B200RNN_MAX_PEERS];
  long long off_flags_ar, off_flags_pr, off_flags_sig, off_flags_bar, off_status, off_inbox, off_heap, off_ll;
};

Ctx to_dev(const b200rnn_comm_ctx* c) {
  Ctx d;
  d.rank = c->rank; d.world = c->world; d.max_blocks = c->max_blocks; d.delay_ns = c->delay_ns;
  d.loss_ppm = c->loss_ppm; d.loss_rto_ns = c->loss_rto_ns > 0 ? c->loss_rto_ns : 200000;
  d.inbox_floats = c->inbox_floats;
  d.timeout = c->timeout_cycles > 0 ? c->timeout_cycles : default_timeout_cycles();
  d.ll_floats = c->ll_floats; d.off_ll = c->off_ll;
  for (int i = 0; i < B200RNN_MAX_PEERS; ++i) d.base[i] = (char*)c->base[i];
  d.off_flags_ar = c->off_flags_ar; d.off_flags_pr = c->off_flags_pr; d.off_flags_sig = c->off_flags_sig;
  d.off_flags_bar = c->off_flags_bar; d.off_status = c->off_status; d.off_inbox = c->off_inbox;
  d.off_heap = c->off_heap;
  return d;
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t* flags_ar(const Ctx& c, int owner, int src, int blk) {
  return reinterpret_cast<uint32_t*>(c.base[owner] + c.off_flags_ar) + (long long)src * c.max_blocks + blk;
}
__device__ __forceinline__ uint32_t* flags_pr(const Ctx& c, int owner, int blk) {
  return reinterpret_cast<uint32_t*>(c.base[owner] + c.off_flags_pr) + blk;
}
__device__ __forceinline__ float* inbox(const Ctx& c, int owner, uint32_t seq, int src) {
  return reinterpret_cast<float*>(c.base[owner] + c.off_inbox) + ((long long)(seq & 1) * c.world + src) * c.inbox_floats;
}
// Network fault injection in front of a publication (the reference's `tc netem delay` / `tc netem loss` rules,
// fabfile.py:146-152, at NVLink scale).  delay_ns lengthens every message.  loss_ppm "drops" a block's message with
// that probability (a hash of seq, rank and block: the same on every run, different per message); a dropped message
// costs one retransmission timeout loss_rto_ns and may be dropped again, up to 8 times in a row - what a lossy link
// does to a reliable transport.  Nothing is ever really lost: the stores go out after the penalty.
__device__ __forceinline__ uint32_t mix32(uint32_t h) {
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}
__device__ __forceinline__ void spin_ns(long long ns) {
  const long long t0 = clock64();
  const long long cycles = ns * 2;  // ~2 GHz SM clock
  while (clock64() - t0 < cycles) { }
}
__device__ __forceinline__ void delay(const Ctx& c, uint32_t seq) {
  if (c.delay_ns > 0) spin_ns(c.delay_ns);
  if (c.loss_ppm > 0) {
    uint32_t h = mix32(seq * 0x9E3779B1u ^ mix32((uint32_t)c.rank * 0x01000193u + blockIdx.x + 0x5BD1E995u));
    for (int tries = 0; tries < 8 && (int)(h % 1000000u) < c.loss_ppm; ++tries) {
      spin_ns(c.loss_rto_ns);
      h = mix32(h + 0x9E3779B9u);
    }
  }
}
// Spin until *f >= seq (wrap-safe); on timeout set bit `src` in this rank's status word.
__device__ __forceinline__ void wait_flag(const Ctx& c, const uint32_t* f, uint32_t seq, int src) {
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys(f) - seq) < 0) {
    if (clock64() - t0 > c.timeout) {
      atomicOr(reinterpret_cast<unsigned int*>(c.base[c.rank] + c.off_status), 1u << src);
      return;
    }
  }
}

// ---------------------------------------------------------------------------
// Element access helpers: 4 consecutive elements per thread, 128-bit where alignment allows.
// ---------------------------------------------------------------------------
struct F4 { float v[4]; };
__device__ __forceinline__ F4 reduce_rows4(const float* __restrict__ rows, int n_rows, long long row_stride, long long i, long long n_valid) {
  F4 r; r.v[0] = r.v[1] = r.v[2] = r.v[3] = 0.0f;
  if (n_rows == 1 && ((reinterpret_cast<uintptr_t>(rows + i) & 15) == 0) && i + 3 < n_valid) {
    const float4 t = *reinterpret_cast<const float4*>(rows + i);
    r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w;
    return r;
  }
  for (int k = 0; k < n_rows; ++k) {
    const float* p = rows + (long long)k * row_stride + i;
#pragma unroll
    for (int e = 0; e < 4; ++e) if (i + e < n_valid) r.v[e] += p[e];
  }
  return r;
}
__device__ __forceinline__ void store4(float* dst, const F4& x, long long i, long long n_valid) {
  if (i + 3 < n_valid && (reinterpret_cast<uintptr_t>(dst + i) & 15) == 0) {
    *reinterpret_cast<float4*>(dst + i) = make_float4(x.v[0], x.v[1], x.v[2], x.v[3]);
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e) if (i + e < n_valid) dst[i + e] = x.v[e];
  }
}
__device__ __forceinline__ F4 load4_cg(const float* src, long long i, long long n_valid) {
  F4 r;
  if (i + 3 < n_valid && (reinterpret_cast<uintptr_t>(src + i) & 15) == 0) {
    const float4 t = __ldcg(reinterpret_cast<const float4*>(src + i));
    r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w;
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e) r.v[e] = (i + e < n_valid) ? __ldcg(src + i + e) : 0.0f;
  }
  return r;
}

// ---------------------------------------------------------------------------
// One-shot: every rank pushes its whole (locally reduced) vector into every peer's inbox,
// then sums the W inbox slices it received.  1 NVLink one-way latency; traffic W*n per rank:
// the right trade for the 56.6 KB gradient of the reference's model.
// Block b handles elements [b*1024, (b+1)*1024) (4 per thread, 128-bit peer stores); flags are per
// (source, block), so blocks are independent and no grid-wide synchronisation exists.
// ---------------------------------------------------------------------------
template <bool UPDATE>
__global__ void __launch_bounds__(kThreads) oneshot_kernel(
    Ctx c, uint32_t seq, const float* __restrict__ rows, long long n, long long row_stride,
    float* buf /* UPDATE: params, else in/out vector */, float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq,
    OptScalars o, float scale, float* __restrict__ metrics) {
  const long long i = ((long long)blockIdx.x * kThreads + threadIdx.x) * kElems;
  F4 g;
  if (UPDATE) {
    g = reduce_rows4(rows, 1, row_stride, i, n + 2);
    if (metrics != nullptr) {
#pragma unroll
      for (int e = 0; e < kElems; ++e) if (i + e >= n && i + e < n + 2) metrics[i + e - n] = g.v[e];   // stay local
    }
  } else {
    g = reduce_rows4(buf, 1, 0, i, n);
  }
  if (i < n) {
    for (int p = 0; p < c.world; ++p) store4(inbox(c, p, seq, c.rank), g, i, n);   // peer stores (incl. self)
  }
  __syncthreads();
  // ONE system fence (issued by the W publishing lanes together, after the barrier made every thread's peer
  // stores visible to them), then the W flags go out in parallel as relaxed stores.
  if (threadIdx.x < c.world) {
    if (threadIdx.x == 0) delay(c, seq);
    __syncwarp(__activemask());
    asm volatile("fence.acq_rel.sys;" ::: "memory");
    st_relaxed_sys(flags_ar(c, threadIdx.x, c.rank, blockIdx.x), seq);
    wait_flag(c, flags_ar(c, c.rank, threadIdx.x, blockIdx.x), seq, threadIdx.x);
  }
  __syncthreads();
  if (i >= n) return;
  F4 sum; sum.v[0] = sum.v[1] = sum.v[2] = sum.v[3] = 0.0f;
  for (int s = 0; s < c.world; ++s) {   // fixed rank order => identical bits on every rank
    const F4 t = load4_cg(inbox(c, c.rank, seq, s), i, n);
#pragma unroll
    for (int e = 0; e < 4; ++e) sum.v[e] += t.v[e];
  }
#pragma unroll
  for (int e = 0; e < kElems; ++e) {
    const long long j = i + e;
    if (j >= n) break;
    if (UPDATE) {
      float p = buf[j];
      float m = exp_avg ? exp_avg[j] : 0.0f;
      float v = exp_avg_sq ? exp_avg_sq[j] : 0.0f;
      apply_update(o, sum.v[e] * o.grad_scale, p, m, v);
      buf[j] = p;
      if (exp_avg) exp_avg[j] = m;
      if (exp_avg_sq) exp_avg_sq[j] = v;
    } else {
      buf[j] = sum.v[e] * scale;
    }
  }
}

// ---------------------------------------------------------------------------
// One-shot, low-latency form ("LL") for the small vectors an LSTM produces (the reference's whole gradient
// is 56.6 KB): every 8-byte word that crosses NVLink carries 4 bytes of data and the 4-byte sequence
// number of the call, so the data is its own arrival flag — no fence, no separate flag store, one NVLink
// one-way latency in total.  A block (rows.cuh geometry: 256 threads, 128 elements) first reduces the
// backward kernel's per-CTA gradient rows in fixed order, then its 8 warps push the 32 result columns to
// 8 peers at a time (two 16-byte stores {d0,seq,d1,seq} {d2,seq,d3,seq} per column), then warp s polls the
// words that source s wrote into this rank's LL inbox, and the first warp adds the W contributions in
// rank order and runs the optimizer.  Inboxes are double-buffered by the parity of seq like the plain
// inboxes; a stale word can never match because seq only grows.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint4* ll_slot(const Ctx& c, int owner, uint32_t seq, int src) {
  return reinterpret_cast<uint4*>(c.base[owner] + c.off_ll) + ((long long)(seq & 1) * c.world + src) * (c.ll_floats / 2);
}
__device__ __forceinline__ void st_ll(uint4* p, float a, float b, uint32_t seq) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(__float_as_uint(a)), "r"(seq),
               "r"(__float_as_uint(b)), "r"(seq) : "memory");
}
__device__ __forceinline__ uint4 ld_ll(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

template <bool UPDATE, bool VEC>
__global__ void __launch_bounds__(kRowThreads) oneshot_ll_kernel(
    Ctx c, uint32_t seq, const float* __restrict__ rows, int n_rows, long long n, long long row_stride,
    float* buf /* UPDATE: params, else in/out vector */, float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq,
    OptScalars o, float scale, float* __restrict__ metrics) {
  __shared__ float4 sm[(B200RNN_MAX_PEERS > kRG ? B200RNN_MAX_PEERS : kRG) * kC4];
  const long long block_e0 = (long long)blockIdx.x * kRowElems;
  const int col = threadIdx.x % kC4, grp = threadIdx.x / kC4;
  const long long e0 = block_e0 + 4 * col;            // my column's first element
  const long long col4 = e0 / 4;                      // float4 index of my column in the vector
  // ---- phase 0: my rank's vector for these columns ----
  if (UPDATE) block_reduce_rows<VEC>(rows, n_rows, row_stride, block_e0, n + 2, sm);
  else block_reduce_rows<VEC>(buf, 1, 0, block_e0, n, sm);
  const float4 mine = sm[col];
  if (UPDATE && grp == 0 && metrics != nullptr) {      // row tail (loss sum, correct count) stays local
    const float mv[4] = {mine.x, mine.y, mine.z, mine.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) if (e0 + e >= n && e0 + e < n + 2) metrics[e0 + e - n] = mv[e];
  }
  // ---- phase 1: push (data, seq) words into every peer's LL inbox: warp `grp` serves peers grp, grp+8 ----
  if (e0 < n) {
    delay(c, seq);
    for (int p = grp; p < c.world; p += kRG) {
      uint4* dst = ll_slot(c, p, seq, c.rank) + 2 * col4;
      st_ll(dst, mine.x, mine.y, seq);
      st_ll(dst + 1, mine.z, mine.w, seq);
    }
  }
  __syncthreads();   // sm is about to be reused
  // ---- phase 2: warp `grp` collects the contributions of sources grp, grp+8 ----
  if (e0 < n) {
    for (int src = grp; src < c.world; src += kRG) {
      const uint4* slot = ll_slot(c, c.rank, seq, src) + 2 * col4;
      uint4 a = ld_ll(slot), b = ld_ll(slot + 1);
      if (a.y != seq || a.w != seq || b.y != seq || b.w != seq) {
        const long long t0 = clock64();
        do {
          a = ld_ll(slot); b = ld_ll(slot + 1);
          if (clock64() - t0 > c.timeout) {
            atomicOr(reinterpret_cast<unsigned int*>(c.base[c.rank] + c.off_status), 1u << src);
            break;
          }
        } while (a.y != seq || a.w != seq || b.y != seq || b.w != seq);
      }
      sm[src * kC4 + col] = make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x), __uint_as_float(b.z));
    }
  }
  __syncthreads();
  if (grp != 0 || e0 >= n) return;
  // ---- phase 3: sum in rank order (identical bits on every rank), optimizer ----
  float4 sum = sm[col];
  for (int src = 1; src < c.world; ++src) sum = f4_add(sum, sm[src * kC4 + col]);
  const float sv[4] = {sum.x, sum.y, sum.z, sum.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const long long j = e0 + e;
    if (j >= n) break;
    if (UPDATE) {
      float p = buf[j];
      float m = exp_avg ? exp_avg[j] : 0.0f;
      float v = exp_avg_sq ? exp_avg_sq[j] : 0.0f;
      apply_update(o, sv[e] * o.grad_scale, p, m, v);
      buf[j] = p;
      if (exp_avg) exp_avg[j] = m;
      if (exp_avg_sq) exp_avg_sq[j] = v;
    } else {
      buf[j] = sv[e] * scale;
    }
  }
}

// Grid-wide barrier on a monotonically increasing local counter (all CTAs co-resident: grid = #SMs).
__device__ __forceinline__ void grid_sync(const Ctx& c, unsigned int* counter, unsigned int target) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    atomicAdd(counter, 1u);
    const long long t0 = clock64();
    while ((int)(*reinterpret_cast<volatile unsigned int*>(counter) - target) < 0) {
      if (clock64() - t0 > c.timeout) {   // CTAs not co-resident / lost: report instead of hanging
        atomicOr(reinterpret_cast<unsigned int*>(c.base[c.rank] + c.off_status), 1u << 30);
        break;
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------
// Two-shot, bulk form: reduce-scatter the gradient to its owning rank (shard p -> rank p), run the
// optimizer on the owner's shard, all-gather the UPDATED PARAMETERS by peer stores into every rank's
// (symmetric) arena.  Traffic 2(W-1)/W * n per rank.  The whole grid (one CTA per SM) streams each
// phase with 128-bit accesses; phases are separated by a local grid barrier and ONE flag per
// (source, destination) pair — 2W flags per call instead of one per 1 KiB block.
// ---------------------------------------------------------------------------
template <bool UPDATE>
__global__ void __launch_bounds__(kThreads) twoshot_kernel(
    Ctx c, uint32_t seq, unsigned int sync_target, const float* __restrict__ rows, int n_rows, long long n,
    long long row_stride, float* buf, long long buf_heap_off, float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq,
    OptScalars o, float scale, float* __restrict__ metrics, long long shard) {
  unsigned int* counter = reinterpret_cast<unsigned int*>(c.base[c.rank] + c.off_flags_bar) + 2 * B200RNN_MAX_PEERS + 1;
  const long long tstride = (long long)gridDim.x * kThreads * kElems;
  const long long t0 = ((long long)blockIdx.x * kThreads + threadIdx.x) * kElems;
  if (UPDATE && blockIdx.x == 0 && threadIdx.x < 2 && metrics != nullptr) {
    float g = 0.0f;
    for (int r = 0; r < n_rows; ++r) g += rows[(long long)r * row_stride + n + threadIdx.x];
    metrics[threadIdx.x] = g;
  }
  // ---- phase 1: push shard p of my (row-reduced) vector into owner p's inbox ----
  for (int dp = 1; dp <= c.world; ++dp) {
    const int p = (c.rank + dp) % c.world;          // rotate so the W ranks hit W different owners
    const long long lo = (long long)p * shard;
    const long long hi = lo + shard < n ? lo + shard : n;
    float* dst = inbox(c, p, seq, c.rank);
    constexpr int U = 4;   // independent 128-bit loads in flight per thread: the loop is latency-bound otherwise
    for (long long i0 = lo + t0; i0 < hi; i0 += U * tstride) {
      F4 g[U];
#pragma unroll
      for (int k = 0; k < U; ++k) {
        const long long i = i0 + k * tstride;
        if (i < hi) g[k] = UPDATE ? reduce_rows4(rows, n_rows, row_stride, i, hi) : reduce_rows4(buf, 1, 0, i, hi);
      }
#pragma unroll
      for (int k = 0; k < U; ++k) {
        const long long i = i0 + k * tstride;
        if (i < hi) store4(dst - lo, g[k], i, hi);
      }
    }
  }
  grid_sync(c, counter, sync_target - gridDim.x);
  if (threadIdx.x == 0 && blockIdx.x < c.world) {
    delay(c, seq);
    st_release_sys(flags_ar(c, blockIdx.x, c.rank, 0), seq);
  }
  // ---- phase 2: owner reduces its shard, applies the optimizer, broadcasts the new parameters ----
  if (threadIdx.x < c.world) wait_flag(c, flags_ar(c, c.rank, threadIdx.x, 0), seq, threadIdx.x);
  __syncthreads();
  {
    const long long lo = (long long)c.rank * shard;
    const long long hi = lo + shard < n ? lo + shard : n;
    constexpr int U2 = 2;   // two elements per thread and iteration: 2 x W independent inbox loads in flight
    for (long long i0 = lo + t0; i0 < hi; i0 += U2 * tstride) {
      F4 sum[U2];
#pragma unroll
      for (int k = 0; k < U2; ++k) sum[k].v[0] = sum[k].v[1] = sum[k].v[2] = sum[k].v[3] = 0.0f;
#pragma unroll 4
      for (int s = 0; s < c.world; ++s) {   // fixed rank order => identical bits on every rank
#pragma unroll
        for (int k = 0; k < U2; ++k) {
          const long long i = i0 + k * tstride;
          if (i < hi) {
            const F4 t = load4_cg(inbox(c, c.rank, seq, s) - lo, i, hi);
#pragma unroll
            for (int e = 0; e < 4; ++e) sum[k].v[e] += t.v[e];
          }
        }
      }
#pragma unroll
      for (int k = 0; k < U2; ++k) {
        const long long i = i0 + k * tstride;
        if (i >= hi) break;
        F4 out;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const long long j = i + e;
          float r = sum[k].v[e] * scale;
          if (UPDATE && j < hi) {
            float p = buf[j];
            float m = exp_avg ? exp_avg[j] : 0.0f;
            float v = exp_avg_sq ? exp_avg_sq[j] : 0.0f;
            apply_update(o, sum[k].v[e] * o.grad_scale, p, m, v);
            if (exp_avg) exp_avg[j] = m;
            if (exp_avg_sq) exp_avg_sq[j] = v;
            r = p;
          }
          out.v[e] = r;
        }
        for (int dp = 0; dp < c.world; ++dp) {   // all-gather by peer stores (self first)
          const int p = (c.rank + dp) % c.world;
          store4(reinterpret_cast<float*>(c.base[p] + c.off_heap + buf_heap_off), out, i, hi);
        }
      }
    }
  }
  grid_sync(c, counter, sync_target);
  if (threadIdx.x == 0 && blockIdx.x < c.world) st_release_sys(flags_pr(c, blockIdx.x, c.rank), seq);
  // ---- phase 3: do not retire before every owner's shard has landed here ----
  if (blockIdx.x == 0 && threadIdx.x < c.world) wait_flag(c, flags_pr(c, c.rank, threadIdx.x), seq, threadIdx.x);
}

// ---------------------------------------------------------------------------
// NVLS: the reduction happens INSIDE the NVSwitch.  Every rank writes its (row-reduced) gradient into
// the same heap slot of its own region; after a cross-rank barrier the owner of shard p issues
// multimem.ld_reduce on the MULTICAST address of that slot (the switch returns the sum over all ranks
// in one load), runs the optimizer on its shard and multimem.st's the updated parameters, which the
// switch replicates into every rank's arena.  Per-GPU NVLink traffic ~ n instead of 2(W-1)/W * n * ...
// and no inboxes.  Needs a region created by b200rnn_symm_vmm_create with multicast.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc_ptr) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc_ptr) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st(float* mc_ptr, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(mc_ptr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// all-to-all flag barrier between the ranks (block 0 signals, every block's first W threads wait)
__device__ __forceinline__ void rank_barrier(const Ctx& c, uint32_t seq, bool use_pr) {
  if (blockIdx.x == 0 && threadIdx.x < c.world) {
    __threadfence_system();
    uint32_t* f = use_pr ? flags_pr(c, threadIdx.x, c.rank) : flags_ar(c, threadIdx.x, c.rank, 0);
    st_release_sys(f, seq);
  }
  if (threadIdx.x < c.world) {
    const uint32_t* f = use_pr ? flags_pr(c, c.rank, threadIdx.x) : flags_ar(c, c.rank, threadIdx.x, 0);
    wait_flag(c, f, seq, threadIdx.x);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kThreads) nvls_update_kernel(
    Ctx c, char* mc_base, uint32_t seq, unsigned int sync_target, const float* __restrict__ rows, int n_rows, long long n,
    long long row_stride, float* buf, long long buf_heap_off, long long grad_heap_off, float* __restrict__ exp_avg,
    float* __restrict__ exp_avg_sq, OptScalars o, float* __restrict__ metrics, long long shard) {
  unsigned int* counter = reinterpret_cast<unsigned int*>(c.base[c.rank] + c.off_flags_bar) + 2 * B200RNN_MAX_PEERS + 1;
  const long long tstride = (long long)gridDim.x * kThreads * kElems;
  const long long t0 = ((long long)blockIdx.x * kThreads + threadIdx.x) * kElems;
  const long long n4 = (n + 3) & ~3LL;
  if (blockIdx.x == 0 && threadIdx.x < 2 && metrics != nullptr) {
    float g = 0.0f;
    for (int r = 0; r < n_rows; ++r) g += rows[(long long)r * row_stride + n + threadIdx.x];
    metrics[threadIdx.x] = g;
  }
  // ---- phase 0: my row-reduced gradient -> my copy of the shared slot (zero padded to n4) ----
  float* slot = reinterpret_cast<float*>(c.base[c.rank] + c.off_heap + grad_heap_off);
  for (long long i = t0; i < n4; i += tstride) {
    const F4 g = reduce_rows4(rows, n_rows, row_stride, i, n);
    *reinterpret_cast<float4*>(slot + i) = make_float4(g.v[0], g.v[1], g.v[2], g.v[3]);
  }
  asm volatile("fence.proxy.alias;" ::: "memory");
  grid_sync(c, counter, sync_target - gridDim.x);
  if (blockIdx.x == 0 && threadIdx.x == 0) delay(c, seq);
  rank_barrier(c, seq, false);
  asm volatile("fence.proxy.alias;" ::: "memory");
  // ---- phase 1: my shard: in-switch reduction, optimizer, in-switch broadcast of the parameters ----
  {
    const long long lo = (long long)c.rank * shard;
    const long long hi = lo + shard < n4 ? lo + shard : n4;
    const float* mc_grad = reinterpret_cast<const float*>(mc_base + c.off_heap + grad_heap_off);
    float* mc_par = reinterpret_cast<float*>(mc_base + c.off_heap + buf_heap_off);
    constexpr int U = 4;   // in-switch loads in flight per thread
    for (long long i0 = lo + t0; i0 < hi; i0 += U * tstride) {
      float4 s4[U];
#pragma unroll
      for (int k = 0; k < U; ++k) {
        const long long i = i0 + k * tstride;
        if (i < hi) s4[k] = multimem_ld_reduce_add(mc_grad + i);
      }
#pragma unroll
      for (int k = 0; k < U; ++k) {
        const long long i = i0 + k * tstride;
        if (i >= hi) break;
        const float sv[4] = {s4[k].x, s4[k].y, s4[k].z, s4[k].w};
        float out[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const long long j = i + e;
          float p = 0.0f;
          if (j < n) {
            p = buf[j];
            float m = exp_avg ? exp_avg[j] : 0.0f;
            float v = exp_avg_sq ? exp_avg_sq[j] : 0.0f;
            apply_update(o, sv[e] * o.grad_scale, p, m, v);
            if (exp_avg) exp_avg[j] = m;
            if (exp_avg_sq) exp_avg_sq[j] = v;
          }
          out[e] = p;
        }
        multimem_st(mc_par + i, make_float4(out[0], out[1], out[2], out[3]));
      }
    }
  }
  asm volatile("fence.proxy.alias;" ::: "memory");
  grid_sync(c, counter, sync_target);
  rank_barrier(c, seq, true);
  asm volatile("fence.proxy.alias;" ::: "memory");
}

// K8 broadcast on the symmetric heap: root copies its bytes into every peer's heap, then flags.
__global__ void __launch_bounds__(kThreads) broadcast_kernel(Ctx c, uint32_t seq, long long heap_off, long long n16,
                                                              long long tail_bytes, int root) {
  uint32_t* bar_local = reinterpret_cast<uint32_t*>(c.base[c.rank] + c.off_flags_bar);
  if (c.rank == root) {
    const int4* src = reinterpret_cast<const int4*>(c.base[root] + c.off_heap + heap_off);
    for (int p = 0; p < c.world; ++p) {
      if (p == root) continue;
      int4* dst = reinterpret_cast<int4*>(c.base[p] + c.off_heap + heap_off);
      for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n16; i += (long long)gridDim.x * kThreads)
        dst[i] = src[i];
      if (blockIdx.x == 0 && threadIdx.x < tail_bytes) {
        const char* s8 = reinterpret_cast<const char*>(src + n16);
        reinterpret_cast<char*>(dst + n16)[threadIdx.x] = s8[threadIdx.x];
      }
    }
    __threadfence_system();
    __syncthreads();
    // count finished CTAs in the root's own slot; the last one publishes to the peers
    __shared__ unsigned int last;
    if (threadIdx.x == 0) {
      unsigned int* cnt = reinterpret_cast<unsigned int*>(bar_local + 2 * B200RNN_MAX_PEERS);
      const unsigned int done = atomicAdd(cnt, 1u) + 1;
      last = (done == gridDim.x) ? 1u : 0u;
      if (last) *cnt = 0;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
      __threadfence_system();
      delay(c, seq);
      for (int p = 0; p < c.world; ++p)
        if (p != root) st_release_sys(reinterpret_cast<uint32_t*>(c.base[p] + c.off_flags_bar) + B200RNN_MAX_PEERS + root, seq);
    }
  } else if (blockIdx.x == 0 && threadIdx.x == 0) {
    wait_flag(c, bar_local + B200RNN_MAX_PEERS + root, seq, root);
  }
}

// K9 put + signal / wait.
__global__ void __launch_bounds__(kThreads) put_kernel(Ctx c, uint32_t seq, int dst, const char* __restrict__ src,
                                                        long long heap_off, long long bytes) {
  char* d = c.base[dst] + c.off_heap + heap_off;
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(d)) & 15) == 0) {
    const long long n16 = bytes / 16;
    for (long long i = threadIdx.x; i < n16; i += kThreads) reinterpret_cast<int4*>(d)[i] = reinterpret_cast<const int4*>(src)[i];
    for (long long i = n16 * 16 + threadIdx.x; i < bytes; i += kThreads) d[i] = src[i];
  } else {
    for (long long i = threadIdx.x; i < bytes; i += kThreads) d[i] = src[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    delay(c, seq);
    st_release_sys(reinterpret_cast<uint32_t*>(c.base[dst] + c.off_flags_sig) + c.rank, seq);
  }
}
__global__ void wait_kernel(Ctx c, uint32_t seq, int src) {
  if (threadIdx.x == 0) wait_flag(c, reinterpret_cast<uint32_t*>(c.base[c.rank] + c.off_flags_sig) + src, seq, src);
}

__global__ void barrier_kernel(Ctx c, uint32_t seq) {
  if (threadIdx.x < c.world) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(c.base[threadIdx.x] + c.off_flags_bar) + c.rank, seq);
    wait_flag(c, reinterpret_cast<uint32_t*>(c.base[c.rank] + c.off_flags_bar) + threadIdx.x, seq, threadIdx.x);
  }
}

// K10 — one-sided parameter server.  The server rank only HOSTS the shared parameter vector in its
// symmetric heap; workers pull it with peer loads and apply their optimizer step to it with
// red.global.add over NVLink (per-worker Adam state stays on the worker, as every worker of the
// reference owns its own remote optimizer instance).  No barrier between workers: asynchronous,
// Hogwild-style, like src/motion/param_server (worker.py:44-59) — minus its four RPC round trips.
__global__ void __launch_bounds__(kThreads) ps_push_kernel(
    Ctx c, int server, long long heap_off, const float* __restrict__ rows, int n_rows, long long n, long long row_stride,
    float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq, OptScalars o, float* __restrict__ metrics) {
  const long long i = (long long)blockIdx.x * kThreads + threadIdx.x;
  if (i >= n + 2) return;
  float g = 0.0f;
  for (int r = 0; r < n_rows; ++r) g += rows[(long long)r * row_stride + i];
  if (i >= n) { if (metrics != nullptr) metrics[i - n] = g; return; }
  g *= o.grad_scale;
  float* sp = reinterpret_cast<float*>(c.base[server] + c.off_heap + heap_off) + i;
  // The DELTA of this step goes to the server: apply_update on p = 0 yields it directly; with weight decay the
  // update depends on the parameter itself, so the server's current value is read over NVLink first (another
  // worker may change it before the add lands: asynchronous by design, like the reference's parameter server).
  const float p0 = (o.wd != 0.0f) ? *reinterpret_cast<const volatile float*>(sp) : 0.0f;
  float p = p0;
  float m = exp_avg ? exp_avg[i] : 0.0f;
  float v = exp_avg_sq ? exp_avg_sq[i] : 0.0f;
  apply_update(o, g, p, m, v);
  if (exp_avg) exp_avg[i] = m;
  if (exp_avg_sq) exp_avg_sq[i] = v;
  atomicAdd(sp, p - p0);                // RED.ADD.F32 to peer memory
}
__global__ void __launch_bounds__(kThreads) ps_pull_kernel(Ctx c, int server, long long heap_off, float* __restrict__ dst, long long n) {
  const volatile float* sp = reinterpret_cast<const volatile float*>(c.base[server] + c.off_heap + heap_off);
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += (long long)gridDim.x * kThreads) dst[i] = sp[i];
}

// Host-side count of two-shot launches per symmetric region (its grid-barrier counter lives in the
// region and only ever increases).  b200rnn_comm_layout-time callers reset it via b200rnn_comm_forget.
struct SyncEntry { const void* base; unsigned int calls; };
SyncEntry g_sync[64];
int g_sync_n = 0;
unsigned int next_sync_call(const void* base) {
  for (int i = 0; i < g_sync_n; ++i) if (g_sync[i].base == base) return ++g_sync[i].calls;
  if (g_sync_n < 64) { g_sync[g_sync_n].base = base; g_sync[g_sync_n].calls = 1; ++g_sync_n; return 1; }
  g_sync[0].base = base; g_sync[0].calls = 1;
  return 1;
}

int check_ctx(const b200rnn_comm_ctx* c, const char* who) {
  B200_REQUIRE(c != nullptr, B200RNN_EINVAL, "%s: null context", who);
  B200_REQUIRE(c->world >= 1 && c->world <= B200RNN_MAX_PEERS && c->rank >= 0 && c->rank < c->world, B200RNN_EINVAL,
               "%s: bad rank/world %d/%d", who, c->rank, c->world);
  for (int i = 0; i < c->world; ++i)
    B200_REQUIRE(c->base[i] != nullptr, B200RNN_EINVAL, "%s: region of rank %d is not mapped", who, i);
  return B200RNN_OK;
}

}  // namespace
}  // namespace b200rnn

using namespace b200rnn;

extern "C" int64_t b200rnn_comm_layout(b200rnn_comm_ctx* ctx) {
  if (!ctx || ctx->world < 1 || ctx->world > B200RNN_MAX_PEERS || ctx->max_blocks < 1 || ctx->inbox_floats < 1) return -1;
  auto align = [](int64_t x) { return (x + 255) & ~int64_t(255); };
  int64_t off = 0;
  ctx->off_flags_ar = off;  off = align(off + (int64_t)ctx->world * ctx->max_blocks * 4);
  ctx->off_flags_pr = off;  off = align(off + (int64_t)ctx->max_blocks * 4);
  ctx->off_flags_sig = off; off = align(off + B200RNN_MAX_PEERS * 4);
  ctx->off_flags_bar = off; off = align(off + (2 * B200RNN_MAX_PEERS + 4) * 4);   // barrier | bcast | counter
  ctx->off_status = off;    off = align(off + 64);
  ctx->off_inbox = off;     off = align(off + 2 * (int64_t)ctx->world * ctx->inbox_floats * 4);
  if (ctx->ll_floats <= 0) ctx->ll_floats = 64 * 1024;     // LL one-shot: vectors up to 256 KB
  ctx->ll_floats = (ctx->ll_floats + 3) & ~int64_t(3);
  ctx->off_ll = off;        off = align(off + 2 * (int64_t)ctx->world * ctx->ll_floats * 8);   // (data, seq) words
  ctx->off_heap = off;      off = align(off + ctx->heap_bytes);
  return off;
}

static int launch_allreduce(const b200rnn_comm_ctx* ctx, uint32_t seq, const float* rows, int n_rows, int64_t n,
                            float* buf, int64_t heap_off, float* m, float* v, const b200rnn_optim_desc* opt,
                            float scale, float* metrics, int algo, bool update, cudaStream_t s) {
  const Ctx c = to_dev(ctx);
  OptScalars o{};
  if (update) o = make_opt_scalars(opt);
  const int64_t row_stride = n + 2;
  const bool ll_fits = ((n + 3) & ~int64_t(3)) <= ctx->ll_floats;
  if (algo == B200RNN_ALGO_AUTO) {
    // Fallback rule for callers without a measured table (the Python layer passes an explicit choice from
    // profiles/allreduce_table.json): LL while the vector fits its inbox and W * bytes stays small (its
    // traffic is 2 W n), flag-based one-shot up to ~128 KiB of W * bytes, two-shot above.
    if (ctx->world == 1 || (ll_fits && n * 4 * ctx->world <= 2 * 1024 * 1024)) algo = B200RNN_ALGO_LL;
    else algo = (n * 4 * ctx->world <= 128 * 1024 || heap_off < 0) ? B200RNN_ALGO_ONESHOT : B200RNN_ALGO_TWOSHOT;
  }
  if (algo == B200RNN_ALGO_LL) {
    B200_REQUIRE(ll_fits, B200RNN_ETOOBIG, "comm LL one-shot: n=%lld exceeds the LL inbox capacity %lld",
                 (long long)n, (long long)ctx->ll_floats);
    const int64_t blocks = ((update ? n + 2 : n) + kRowElems - 1) / kRowElems;
    const bool vec = update ? rows_vectorizable(rows, row_stride) : ((reinterpret_cast<uintptr_t>(buf) & 15) == 0);
#define B200_LL(UPD, VEC) oneshot_ll_kernel<UPD, VEC><<<(unsigned)blocks, kRowThreads, 0, s>>>( \
        c, seq, rows, n_rows, n, row_stride, buf, m, v, o, scale, metrics)
    if (update) { if (vec) B200_LL(true, true); else B200_LL(true, false); }
    else { if (vec) B200_LL(false, true); else B200_LL(false, false); }
#undef B200_LL
  } else if (algo == B200RNN_ALGO_ONESHOT) {
    B200_REQUIRE(!update || n_rows == 1, B200RNN_EINVAL,
                 "comm one-shot: reduce the %d gradient rows first (b200rnn_reduce_rows) or use the LL form", n_rows);
    const int64_t one_blocks = ((update ? n + 2 : n) + kBlockElems - 1) / kBlockElems;
    B200_REQUIRE(one_blocks <= ctx->max_blocks, B200RNN_ETOOBIG, "comm one-shot: %lld blocks exceed max_blocks=%d",
                 (long long)one_blocks, ctx->max_blocks);
    B200_REQUIRE(n <= ctx->inbox_floats, B200RNN_ETOOBIG, "comm one-shot: n=%lld exceeds inbox capacity %lld",
                 (long long)n, (long long)ctx->inbox_floats);
    if (update) oneshot_kernel<true><<<(unsigned)one_blocks, kThreads, 0, s>>>(c, seq, rows, n, row_stride, buf, m, v, o, scale, metrics);
    else oneshot_kernel<false><<<(unsigned)one_blocks, kThreads, 0, s>>>(c, seq, rows, n, row_stride, buf, m, v, o, scale, metrics);
  } else {
    int64_t shard = (n + ctx->world - 1) / ctx->world;
    shard = (shard + 3) & ~int64_t(3);                       // 16-byte aligned shard starts
    B200_REQUIRE(shard <= ctx->inbox_floats, B200RNN_ETOOBIG,
                 "comm two-shot: shard of %lld floats exceeds inbox capacity %lld", (long long)shard,
                 (long long)ctx->inbox_floats);
    B200_REQUIRE(heap_off >= 0 && heap_off + n * 4 <= ctx->heap_bytes && (heap_off & 15) == 0, B200RNN_EINVAL,
                 "comm two-shot: the vector must live in the symmetric heap, 16-byte aligned (heap_off=%lld)", (long long)heap_off);
    // local grid barrier bookkeeping: fixed grid (one CTA per SM), monotonically increasing target
    const int grid = num_sms();
    B200_REQUIRE(ctx->world <= grid, B200RNN_EINVAL, "comm two-shot: world larger than the grid");
    const unsigned int target = next_sync_call(ctx->base[ctx->rank]) * 2u * (unsigned)grid;
    if (update) twoshot_kernel<true><<<grid, kThreads, 0, s>>>(c, seq, target, rows, n_rows, n, row_stride, buf, heap_off, m, v, o, scale, metrics, shard);
    else twoshot_kernel<false><<<grid, kThreads, 0, s>>>(c, seq, target, rows, n_rows, n, row_stride, buf, heap_off, m, v, o, scale, metrics, shard);
  }
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_comm_allreduce_update(const b200rnn_comm_ctx* ctx, uint32_t seq, const float* rows,
                                             int32_t n_rows, int64_t n_params, float* params,
                                             int64_t params_heap_off, float* exp_avg, float* exp_avg_sq,
                                             const b200rnn_optim_desc* opt, float* metrics, int32_t algo, void* stream) {
  int rc = check_ctx(ctx, "comm_allreduce_update");
  if (rc) return rc;
  B200_REQUIRE(rows && params && opt && n_rows >= 1 && n_params > 0, B200RNN_EINVAL, "comm_allreduce_update: bad argument");
  B200_REQUIRE(opt->kind != 0 || (exp_avg && exp_avg_sq && opt->step >= 1), B200RNN_EINVAL,
               "comm_allreduce_update: Adam needs exp_avg/exp_avg_sq and step >= 1");
  return launch_allreduce(ctx, seq, rows, n_rows, n_params, params, params_heap_off, exp_avg, exp_avg_sq, opt, 1.0f,
                          metrics, algo, true, (cudaStream_t)stream);
}

extern "C" int b200rnn_comm_allreduce(const b200rnn_comm_ctx* ctx, uint32_t seq, float* buf, int64_t n, float scale,
                                      void* stream) {
  int rc = check_ctx(ctx, "comm_allreduce");
  if (rc) return rc;
  B200_REQUIRE(buf && n > 0, B200RNN_EINVAL, "comm_allreduce: bad argument");
  const bool ll = ((n + 3) & ~int64_t(3)) <= ctx->ll_floats && n * 4 * ctx->world <= 2 * 1024 * 1024;
  return launch_allreduce(ctx, seq, nullptr, 1, n, buf, -1, nullptr, nullptr, nullptr, scale, nullptr,
                          ll ? B200RNN_ALGO_LL : B200RNN_ALGO_ONESHOT, false, (cudaStream_t)stream);
}
extern "C" int b200rnn_comm_allreduce_algo(const b200rnn_comm_ctx* ctx, uint32_t seq, float* buf, int64_t n, float scale,
                                           int32_t algo, void* stream) {
  int rc = check_ctx(ctx, "comm_allreduce_algo");
  if (rc) return rc;
  B200_REQUIRE(buf && n > 0 && (algo == B200RNN_ALGO_LL || algo == B200RNN_ALGO_ONESHOT), B200RNN_EINVAL,
               "comm_allreduce_algo: bad argument (only the one-shot forms work on a vector outside the heap)");
  return launch_allreduce(ctx, seq, nullptr, 1, n, buf, -1, nullptr, nullptr, nullptr, scale, nullptr, algo, false,
                          (cudaStream_t)stream);
}

extern "C" int b200rnn_comm_broadcast(const b200rnn_comm_ctx* ctx, uint32_t seq, int64_t heap_off, int64_t bytes,
                                      int32_t root, void* stream) {
  int rc = check_ctx(ctx, "comm_broadcast");
  if (rc) return rc;
  B200_REQUIRE(heap_off >= 0 && (heap_off & 15) == 0 && bytes > 0 && heap_off + bytes <= ctx->heap_bytes && root >= 0 &&
               root < ctx->world, B200RNN_EINVAL, "comm_broadcast: bad range/root");
  if (ctx->world == 1) return B200RNN_OK;
  const Ctx c = to_dev(ctx);
  const int64_t n16 = bytes / 16;
  int grid = 1;
  if (ctx->rank == root) { grid = (int)((n16 + kThreads - 1) / kThreads); if (grid < 1) grid = 1; if (grid > 64) grid = 64; }
  broadcast_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(c, seq, heap_off, n16, bytes - n16 * 16, root);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_comm_put(const b200rnn_comm_ctx* ctx, uint32_t seq, int32_t dst, const void* src,
                                int64_t heap_off, int64_t bytes, void* stream) {
  int rc = check_ctx(ctx, "comm_put");
  if (rc) return rc;
  B200_REQUIRE(src && dst >= 0 && dst < ctx->world && heap_off >= 0 && bytes > 0 && heap_off + bytes <= ctx->heap_bytes,
               B200RNN_EINVAL, "comm_put: bad argument");
  put_kernel<<<1, kThreads, 0, (cudaStream_t)stream>>>(to_dev(ctx), seq, dst, (const char*)src, heap_off, bytes);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_comm_wait(const b200rnn_comm_ctx* ctx, uint32_t seq, int32_t src, void* stream) {
  int rc = check_ctx(ctx, "comm_wait");
  if (rc) return rc;
  B200_REQUIRE(src >= 0 && src < ctx->world, B200RNN_EINVAL, "comm_wait: bad source rank");
  wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(to_dev(ctx), seq, src);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_comm_barrier(const b200rnn_comm_ctx* ctx, uint32_t seq, void* stream) {
  int rc = check_ctx(ctx, "comm_barrier");
  if (rc) return rc;
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(to_dev(ctx), seq);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_comm_status(const b200rnn_comm_ctx* ctx, int32_t* status, void* stream) {
  int rc = check_ctx(ctx, "comm_status");
  if (rc) return rc;
  B200_REQUIRE(status != nullptr, B200RNN_EINVAL, "comm_status: null output");
  B200_CHECK_CUDA(cudaMemcpyAsync(status, (char*)ctx->base[ctx->rank] + ctx->off_status, 4, cudaMemcpyDeviceToHost,
                                  (cudaStream_t)stream));
  B200_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return B200RNN_OK;
}

extern "C" int b200rnn_comm_ps_push(const b200rnn_comm_ctx* ctx, int32_t server, int64_t heap_off, const float* rows,
                                    int32_t n_rows, int64_t n_params, float* exp_avg, float* exp_avg_sq,
                                    const b200rnn_optim_desc* opt, float* metrics, void* stream) {
  int rc = check_ctx(ctx, "comm_ps_push");
  if (rc) return rc;
  B200_REQUIRE(rows && opt && n_rows >= 1 && n_params > 0 && server >= 0 && server < ctx->world && heap_off >= 0 &&
               heap_off + n_params * 4 <= ctx->heap_bytes, B200RNN_EINVAL, "comm_ps_push: bad argument");
  B200_REQUIRE(opt->kind != 0 || (exp_avg && exp_avg_sq && opt->step >= 1), B200RNN_EINVAL,
               "comm_ps_push: Adam needs exp_avg/exp_avg_sq and step >= 1");
  const OptScalars o = make_opt_scalars(opt);
  const int64_t blocks = (n_params + 2 + kThreads - 1) / kThreads;
  ps_push_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(to_dev(ctx), server, heap_off, rows, n_rows, n_params,
                                                                          n_params + 2, exp_avg, exp_avg_sq, o, metrics);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_comm_ps_pull(const b200rnn_comm_ctx* ctx, int32_t server, int64_t heap_off, float* dst,
                                    int64_t n_params, void* stream) {
  int rc = check_ctx(ctx, "comm_ps_pull");
  if (rc) return rc;
  B200_REQUIRE(dst && n_params > 0 && server >= 0 && server < ctx->world && heap_off >= 0 &&
               heap_off + n_params * 4 <= ctx->heap_bytes, B200RNN_EINVAL, "comm_ps_pull: bad argument");
  int64_t blocks = (n_params + kThreads - 1) / kThreads;
  if (blocks > 4 * num_sms()) blocks = 4 * num_sms();
  ps_pull_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(to_dev(ctx), server, heap_off, dst, n_params);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}

extern "C" int b200rnn_comm_forget(const void* region_base) {
  for (int i = 0; i < g_sync_n; ++i)
    if (g_sync[i].base == region_base) { g_sync[i] = g_sync[g_sync_n - 1]; --g_sync_n; break; }
  return B200RNN_OK;
}

extern "C" int b200rnn_comm_allreduce_update_nvls(const b200rnn_comm_ctx* ctx, void* mc_base, uint32_t seq, const float* rows,
                                                  int32_t n_rows, int64_t n_params, float* params, int64_t params_heap_off,
                                                  int64_t grad_heap_off, float* exp_avg, float* exp_avg_sq,
                                                  const b200rnn_optim_desc* opt, float* metrics, void* stream) {
  int rc = check_ctx(ctx, "comm_allreduce_update_nvls");
  if (rc) return rc;
  B200_REQUIRE(mc_base && rows && params && opt && n_rows >= 1 && n_params > 0, B200RNN_EINVAL, "comm_allreduce_update_nvls: bad argument");
  const int64_t n4 = (n_params + 3) & ~int64_t(3);
  B200_REQUIRE(params_heap_off >= 0 && grad_heap_off >= 0 && (params_heap_off & 15) == 0 && (grad_heap_off & 15) == 0 &&
               params_heap_off + n4 * 4 <= ctx->heap_bytes && grad_heap_off + n4 * 4 <= ctx->heap_bytes, B200RNN_EINVAL,
               "comm_allreduce_update_nvls: parameter arena and gradient slot must live 16-byte aligned in the symmetric heap");
  B200_REQUIRE(opt->kind != 0 || (exp_avg && exp_avg_sq && opt->step >= 1), B200RNN_EINVAL,
               "comm_allreduce_update_nvls: Adam needs exp_avg/exp_avg_sq and step >= 1");
  const OptScalars o = make_opt_scalars(opt);
  int64_t shard = (n4 + ctx->world - 1) / ctx->world;
  shard = (shard + 3) & ~int64_t(3);
  int grid = num_sms();
  const int64_t need = (n4 / kElems + kThreads - 1) / kThreads;
  if (need < grid) grid = (int)(need < ctx->world ? ctx->world : need);   // small vectors: fewer CTAs, cheaper grid barrier
  // the grid-barrier counter target depends on the grid: keep it fixed per region by always counting num_sms
  grid = num_sms();
  const unsigned int target = next_sync_call(ctx->base[ctx->rank]) * 2u * (unsigned)grid;
  nvls_update_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(to_dev(ctx), (char*)mc_base, seq, target, rows, n_rows, n_params,
                                                                  n_params + 2, params, params_heap_off, grad_heap_off, exp_avg,
                                                                  exp_avg_sq, o, metrics, shard);
  B200_CHECK_LAUNCH();
  return B200RNN_OK;
}
