/*
 * b200rnn — C ABI of the B200-native (sm_100a) RNN training kernels.
 *
 * Every entry point takes raw device pointers plus a CUDA stream (passed as
 * void* so this header needs no CUDA include) and returns 0 on success or a
 * negative B200RNN_E* code; b200rnn_last_error() gives the message.
 * None of the functions in this header allocate device memory unless their
 * comment says so: the caller (the Python package, through torch's caching
 * allocator) owns all buffers and sizes them with the *_floats() queries.
 *
 * Parity citations are paths into the reference tree (jkhlr/pytorch-distributed-rnn).
 */
#ifndef B200RNN_H_
#define B200RNN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200RNN_MAX_LAYERS 8
#define B200RNN_MAX_PEERS 16

enum {
  B200RNN_OK = 0,
  B200RNN_EINVAL = -1,      /* bad argument / unsupported shape */
  B200RNN_ECUDA = -2,       /* a CUDA call failed */
  B200RNN_EUNSUPPORTED = -3,/* shape not covered by this kernel family */
  B200RNN_ETOOBIG = -4,     /* request exceeds a fixed capacity */
  B200RNN_ESYS = -5         /* OS / socket error (runtime bootstrap) */
};

enum { B200RNN_CELL_LSTM = 0, B200RNN_CELL_GRU = 1 };

/* Model geometry + where each parameter lives inside the flat fp32 arena.
 * Parameter order and shapes are torch.nn.LSTM's (src/motion/model.py:9-10):
 * weight_ih_l{k} [G*H, in], weight_hh_l{k} [G*H, H], bias_ih_l{k} [G*H],
 * bias_hh_l{k} [G*H] (G = 4 gates i,f,g,o for LSTM; 3 gates r,z,n for GRU),
 * then the optional classifier head fc.weight [C, H], fc.bias [C]. */
typedef struct {
  int32_t cell;          /* B200RNN_CELL_* */
  int32_t input_size;    /* I */
  int32_t hidden_size;   /* H */
  int32_t num_layers;    /* L <= B200RNN_MAX_LAYERS */
  int32_t batch;         /* B (per rank) */
  int32_t seq_len;       /* T */
  int32_t num_classes;   /* C, 0 = no head */
  int32_t reserved;
  int64_t off_wih[B200RNN_MAX_LAYERS];
  int64_t off_whh[B200RNN_MAX_LAYERS];
  int64_t off_bih[B200RNN_MAX_LAYERS];
  int64_t off_bhh[B200RNN_MAX_LAYERS];
  int64_t off_fcw;
  int64_t off_fcb;
  int64_t n_params;      /* arena length in floats (gradient rows carry 2 extra slots) */
} b200rnn_model_desc;

const char* b200rnn_last_error(void);
int b200rnn_version(void);

/* ---- K11: device-resident dataset gather (replaces DataLoader collate,
 * src/motion/trainer/base.py:56-60,107).  x_out[b] = feats[idx[b]] (row_elems
 * floats per row), y_out[b] = labels[idx[b]]. */
int b200rnn_gather_batch(const float* feats, const int64_t* labels,
                         const int64_t* idx, int32_t batch, int32_t row_elems,
                         float* x_out, int64_t* y_out, void* stream);
/* Same with at most max_blocks CTAs (0 = one per row): for a gather that streams its rows out of PINNED HOST
 * memory on a side stream (feats / labels may be any device-accessible addresses) without taking the SMs away
 * from the kernels it overlaps with. */
int b200rnn_gather_batch_capped(const float* feats, const int64_t* labels,
                                const int64_t* idx, int32_t batch, int32_t row_elems,
                                float* x_out, int64_t* y_out, int32_t max_blocks, void* stream);
/* Time-major form for the large-hidden path: x_tm_out[t][b][:] = feats[idx[b]][t][:] (idx nullable = identity),
 * i.e. gather and [B,T,I] -> [T,B,I] transpose in one pass. */
int b200rnn_gather_batch_tm(const float* feats, const int64_t* idx, int32_t batch, int32_t seq_len,
                            int32_t features, float* x_tm_out, void* stream);

/* ---- K2/K3: fused persistent recurrent kernels, small-hidden family
 * (whole sequence, all layers, one launch; recurrent weights register-resident).
 * Replaces ATen lstm fwd/bwd (src/motion/model.py:14, trainer/base.py:116). */

/* 1 if (cell, I, H, L) is covered by the fused small-hidden family. */
int b200rnn_rnn_small_supported(const b200rnn_model_desc* d);
/* Batch rows per CTA and CTA count the fused kernels will use for d->batch. */
int b200rnn_rnn_small_tiling(const b200rnn_model_desc* d, int32_t* rows_per_cta, int32_t* n_cta);
/* Activation stash written by forward (training=1) and read by backward. */
int64_t b200rnn_rnn_stash_floats(const b200rnn_model_desc* d);
/* Backward scratch: per-CTA gradient partial rows + inter-layer dx buffers. */
int64_t b200rnn_rnn_bwd_scratch_floats(const b200rnn_model_desc* d);
/* Length of one gradient row: n_params + 2 (loss sum, correct count). */
int64_t b200rnn_grad_row_floats(const b200rnn_model_desc* d);

/* x [B,T,I] fp32 contiguous. stash layout (floats):
 *   h_seq [L][T][B][H] | c_seq [L][T][B][H] | gates [L][T][B][H][4]
 * (c_seq and gates only written when training != 0).  The top layer's
 * h_seq[L-1] is the module output in time-major order. */
int b200rnn_lstm_fwd(const b200rnn_model_desc* d, const float* params,
                     const float* x, float* stash, int32_t training, void* stream);

/* Tensor-core form of b200rnn_lstm_fwd (same arguments, same stash): the per-step gate
 * pre-activations of a batch tile are one tcgen05.mma GEMM (M=128 gate rows, N=16 samples,
 * kind::tf32, 3xTF32 split for fp32 fidelity) with [W_ih|W_hh] resident in shared memory and the
 * accumulator in TMEM (csrc/rnn/lstm_tc.cu). */
int b200rnn_lstm_fwd_tc(const b200rnn_model_desc* d, const float* params,
                        const float* x, float* stash, int32_t training, void* stream);
/* 0 = no tensor-core recurrent kernel has hit its mbarrier watchdog since load. */
int b200rnn_tc_status(int32_t* status);

/* Backward through time for all layers.  Gradient sources (all optional):
 *   labels != NULL && d->num_classes > 0: the classifier head + mean
 *       cross-entropy + correct-count are evaluated in-kernel on
 *       h_seq[L-1][T-1] (src/motion/model.py:16, trainer/base.py:112-114),
 *       loss scaled by loss_scale (normally 1/B);
 *   d_out  [B,T,H] gradient of the top layer's full output sequence;
 *   d_hn / d_cn [L][B][H] gradients of the final hidden / cell states.
 * Results: scratch holds n_cta partial gradient rows (row r = CTA r's
 * contribution, b200rnn_grad_row_floats() long, torch parameter order, then
 * [n_params] = sum of per-sample losses, [n_params+1] = correct count);
 * reduce them with b200rnn_reduce_rows() or b200rnn_fused_update(). */
int b200rnn_lstm_bwd(const b200rnn_model_desc* d, const float* params,
                     const float* x, const float* stash, const int64_t* labels,
                     float loss_scale, const float* d_out, const float* d_hn,
                     const float* d_cn, float* scratch, void* stream);

/* GRU members of the same fused family (hidden_size 32 or 16, d->cell == B200RNN_CELL_GRU;
 * torch gate order r, z, n; SURVEY P22 — the reference itself only builds nn.LSTM).  Same
 * tiling, stash layout (h_seq | unused | per-step (r, z, n, W_hn h + b_hn)), scratch layout and
 * fused head + cross-entropy as the LSTM entry points; a GRU has no cell state, so no d_cn. */
int b200rnn_gru_fwd(const b200rnn_model_desc* d, const float* params,
                    const float* x, float* stash, int32_t training, void* stream);
int b200rnn_gru_bwd(const b200rnn_model_desc* d, const float* params,
                    const float* x, const float* stash, const int64_t* labels,
                    float loss_scale, const float* d_out, const float* d_hn,
                    float* scratch, void* stream);

/* Backward variants that additionally write the gradient of the network input: dx is
 * [T][B][hidden_size] (time-major, the input zero-padded to hidden_size columns; columns
 * >= input_size come out as zeros).  Same scratch and results as the plain entry points.
 * (No counterpart in the reference, which never differentiates w.r.t. its input; needed for
 * a drop-in recurrent layer that sits on top of other trainable layers.) */
int b200rnn_lstm_bwd_dx(const b200rnn_model_desc* d, const float* params,
                        const float* x, const float* stash, const int64_t* labels,
                        float loss_scale, const float* d_out, const float* d_hn,
                        const float* d_cn, float* scratch, float* dx, void* stream);
int b200rnn_gru_bwd_dx(const b200rnn_model_desc* d, const float* params,
                       const float* x, const float* stash, const int64_t* labels,
                       float loss_scale, const float* d_out, const float* d_hn,
                       float* scratch, float* dx, void* stream);

/* Training-loop forms of the fused forward / backward (LSTM and GRU): the same kernels with two extras.
 *  - idx (nullable, [B] int64): batch row b is row idx[b] of x and of labels, i.e. x / labels are the
 *    whole device-resident dataset and idx is the sampler's slice for this step — the K11 gather folded
 *    into the kernels' own loaders (DataLoader + DistributedSampler, src/motion/trainer/base.py:56-60,107).
 *  - flags: hand-over of the layer-wavefront sentinels between consecutive calls on the same buffers,
 *    which removes the fill launch in front of a wavefront kernel.
 *      B200RNN_STEP_ARMED  forward: h_seq of the layers below the top already holds the sentinel
 *                          everywhere; backward: so do the inter-layer dx buffers in `scratch`.
 *      B200RNN_STEP_REARM  backward only: write the sentinel back over every h_seq element of the lower
 *                          layers and every dx element once consumed, so that the NEXT forward / backward
 *                          on the same stash / scratch may be called with B200RNN_STEP_ARMED.  The stash is
 *                          therefore not const here, and h_seq of the lower layers is gone afterwards.
 *    A buffer is armed only by such a backward; any other call that writes it (a forward, a backward
 *    without REARM) disarms it.  With flags == 0 these are b200rnn_lstm_fwd / b200rnn_lstm_bwd plus idx. */
#define B200RNN_STEP_ARMED 1
#define B200RNN_STEP_REARM 2
int b200rnn_lstm_fwd_ex(const b200rnn_model_desc* d, const float* params, const float* x, const int64_t* idx,
                        float* stash, int32_t training, int32_t flags, void* stream);
int b200rnn_lstm_bwd_ex(const b200rnn_model_desc* d, const float* params, const float* x, const int64_t* idx,
                        float* stash, const int64_t* labels, float loss_scale, const float* d_out,
                        const float* d_hn, const float* d_cn, float* scratch, int32_t flags, void* stream);
int b200rnn_gru_fwd_ex(const b200rnn_model_desc* d, const float* params, const float* x, const int64_t* idx,
                       float* stash, int32_t training, int32_t flags, void* stream);
int b200rnn_gru_bwd_ex(const b200rnn_model_desc* d, const float* params, const float* x, const int64_t* idx,
                       float* stash, const int64_t* labels, float loss_scale, const float* d_out,
                       const float* d_hn, float* scratch, int32_t flags, void* stream);

/* Capacity of the scratch array compiled into the library for b200rnn_lstm_bwd_static. */
#define B200RNN_STATIC_SCRATCH_FLOATS (4 * 1024 * 1024)

/* Backward for callers that cannot supply scratch memory.  (No counterpart in the reference:
 * it never manages device memory.)  Meant for hosts of this C ABI that have no allocator to
 * ask: a plain C/C++ program without torch, a process whose own pool already owns the whole
 * GPU so that cudaMalloc fails, or a loop that cannot afford the device-wide stall of
 * cudaMalloc/cudaFree.  Same arguments and math as b200rnn_lstm_bwd followed by
 * b200rnn_reduce_rows, except that the per-CTA gradient rows and the inter-layer dx buffers
 * sit in a B200RNN_STATIC_SCRATCH_FLOATS-float (16 MiB) array inside the library, and the
 * result is written to grad_out[n_params + 2] (gradients in torch parameter order, then the
 * sum of per-sample losses and the correct count).  The call makes no cudaMalloc/cudaFree and
 * calls no allocating library routine; inputs and grad_out remain the caller's.
 *  - Every GPU that loads the library holds those 16 MiB until the process exits.
 *  - Shapes with b200rnn_rnn_bwd_scratch_floats(d) > B200RNN_STATIC_SCRATCH_FLOATS are
 *    refused with B200RNN_ETOOBIG (message in b200rnn_last_error()); use b200rnn_lstm_bwd.
 *  - The first call on each device looks up the array's device address
 *    (cudaGetSymbolAddress with that device current) and caches it; make that first call
 *    outside any stream capture.  Later calls only launch kernels.  The device that owns
 *    `stream` must be the current device (checked; B200RNN_EINVAL otherwise).
 *  - All calls on a device share the one array and it keeps the previous call's contents.
 *    Nothing in it needs clearing (each element is written before it is read within a call),
 *    but two calls that overlap on the GPU (different streams, no dependency) race on it:
 *    keep calls on one stream or order them with events. */
int b200rnn_lstm_bwd_static(const b200rnn_model_desc* d, const float* params,
                            const float* x, const float* stash, const int64_t* labels,
                            float loss_scale, const float* d_out, const float* d_hn,
                            const float* d_cn, float* grad_out, void* stream);

/* out[i] (+)= sum_r rows[r][i] in fixed row order (deterministic). */
int b200rnn_reduce_rows(const float* rows, int32_t n_rows, int64_t row_floats,
                        float* out, int32_t accumulate, void* stream);

/* ---- K5: classifier head, standalone form (evaluation and autograd path).
 * h_last [B,H] (row stride ld_h floats), logits out [B,C] (nullable),
 * metrics out: [0] = sum of per-sample CE losses, [1] = correct count.  When labels are given, `metrics` also
 * carries the per-block partial sums behind those two values (combined in fixed order by a second launch): it
 * must then hold b200rnn_head_fwd_metrics_floats(batch) floats (at most 130), not 2. */
int64_t b200rnn_head_fwd_metrics_floats(int32_t batch);
int b200rnn_head_fwd(const float* h_last, int64_t ld_h, const float* fc_w, const float* fc_b,
                     const int64_t* labels, int32_t batch, int32_t hidden, int32_t classes,
                     float* logits, float* metrics, void* stream);

/* K5 with its backward: head + mean cross-entropy + correct count + all their gradients in two launches
 * (csrc/head/head_ce.cu) — what models outside the small-hidden fused family use instead of nn.Linear +
 * CrossEntropyLoss + argmax and their autograd backward (src/motion/model.py:16, trainer/base.py:112-116).
 * labels[idx ? idx[b] : b] (idx nullable); loss scaled by loss_scale (normally 1/B) for the gradients.
 * Outputs: d_h [B,H] (row stride ld_dh) = gradient of h_last, d_w [C,H], d_b [C], metrics[2] = (sum of
 * per-sample losses, correct count).  scratch: b200rnn_head_ce_scratch_floats(B, C) floats.  Every block of
 * the weight-gradient kernel adds one count to retire_counter (nullable) when done; *retire_count (nullable)
 * receives how many.  b200rnn_head_status: bit 0 = a label outside [0, C) was met (treated as class 0). */
int64_t b200rnn_head_ce_scratch_floats(int32_t batch, int32_t classes);
int b200rnn_head_ce(const float* h_last, int64_t ld_h, const float* fc_w, const float* fc_b,
                    const int64_t* labels, const int64_t* idx, int32_t batch, int32_t hidden,
                    int32_t classes, float loss_scale, float* d_h, int64_t ld_dh, float* d_w,
                    float* d_b, float* metrics, float* scratch, uint32_t* retire_counter,
                    int32_t* retire_count, void* stream);
int b200rnn_head_status(int32_t* status);

/* ---- K6: fused optimizers on the flat arena (src/motion/trainer/base.py:42-43,117;
 * src/example/example_ddp.py:54,82). */
typedef struct {
  int32_t kind;        /* 0 = Adam, 1 = SGD */
  int32_t step;        /* 1-based step number (Adam bias correction) */
  float lr, beta1, beta2, eps, weight_decay, momentum;
  float grad_scale;    /* multiplies the summed gradient (1/world for DDP mean) */
  int32_t reserved;
} b200rnn_optim_desc;

/* grad = grad_scale * sum_r rows[r][i]; apply optimizer in place on
 * params/exp_avg/exp_avg_sq; copy row tail (loss sum, correct) to metrics[2]
 * (nullable); optionally also store the reduced gradient to grad_out (nullable). */
int b200rnn_fused_update(const float* rows, int32_t n_rows, int64_t n_params,
                         float* params, float* exp_avg, float* exp_avg_sq,
                         const b200rnn_optim_desc* opt, float* grad_out,
                         float* metrics, void* stream);

/* ---- K1/K4: bf16 GEMM on tcgen05 tensor cores (TMA-fed, TMEM accumulators).
 * C[M,N] (+)= A[M,K] x B[N,K]^T + bias[N]; A, B bf16 row-major (K contiguous, lda/ldb multiples
 * of 8, 16-byte aligned), C fp32 or bf16 (c_is_bf16), accumulate != 0 adds to the existing C.
 * Used for the input projection over all timesteps (src/motion/model.py:14), the per-step
 * recurrent projection at large hidden sizes and the weight-gradient GEMMs (trainer/base.py:116). */
int b200rnn_gemm_bf16(const void* A, const void* B, void* C, const float* bias, int32_t M, int32_t N,
                      int32_t K, int32_t lda, int32_t ldb, int32_t ldc, int32_t accumulate,
                      int32_t c_is_bf16, void* stream);
/* Same GEMM whose CTAs each add one count (red.release.gpu) to *retire_counter once their output tile is
 * stored; *retire_count receives the number of tiles.  The reducer's wait kernel (b200rnn_reducer_wait) on
 * another stream starts the gradient bucket's all-reduce when the last tile of the weight-gradient GEMMs has
 * retired (K4 -> K7 trigger; c10d fires its bucket hooks at the same point, src/motion/trainer/ddp.py:19). */
int b200rnn_gemm_bf16_retire(const void* A, const void* B, void* C, const float* bias, int32_t M, int32_t N,
                             int32_t K, int32_t lda, int32_t ldb, int32_t ldc, int32_t accumulate,
                             int32_t c_is_bf16, uint32_t* retire_counter, int32_t* retire_count, void* stream);
/* 0 = no tensor-core pipeline timeout has been observed since the library was loaded. */
int b200rnn_gemm_status(int32_t* status);
/* fp32 -> bf16 cast; and cast + transpose [rows, cols] -> [cols, rows]. */
int b200rnn_cast_bf16(const float* in, void* out, int64_t n, void* stream);
int b200rnn_transpose_cast_bf16(const float* in, void* out, int32_t rows, int32_t cols, int64_t ld_in,
                                int64_t ld_out, void* stream);

/* ---- Large-hidden LSTM / GRU path (H % 8 == 0; csrc/rnn/rnn_large.cu): input projection over all
 * timesteps, per-step recurrent projection and all weight-gradient products run on the tcgen05 GEMM
 * (bf16 operands, fp32 accumulate / state), gates + cell update fused in one pointwise kernel per
 * step; the time loop is driven natively.  x_tm is TIME-major [T,B,I].  Buffers are the caller's:
 * size them with the *_bytes queries.  stash: per layer  bf16 X | gates | (GRU) h-projection |
 * (LSTM) c | h fp32 | h bf16;  the top layer's h ([T,B,H] fp32) is the module output. */
int64_t b200rnn_rnn_large_stash_bytes(const b200rnn_model_desc* d);
int64_t b200rnn_rnn_large_scratch_bytes(const b200rnn_model_desc* d);
/* Byte offset of layer l's fp32 h sequence [T,B,H] (and, LSTM, c sequence) inside the stash. */
int64_t b200rnn_rnn_large_h_offset(const b200rnn_model_desc* d, int32_t layer);
int64_t b200rnn_rnn_large_c_offset(const b200rnn_model_desc* d, int32_t layer);
int b200rnn_rnn_large_fwd(const b200rnn_model_desc* d, const float* params, const float* x_tm,
                          void* stash, void* scratch, void* stream);
/* d_out_tm [T,B,H] / d_hn, d_cn [L,B,H] nullable; grad_out: d->n_params floats in parameter layout;
 * dx_tm [T,B,I] nullable. */
int b200rnn_rnn_large_bwd(const b200rnn_model_desc* d, const float* params, const float* x_tm,
                          const void* stash, void* scratch, const float* d_out_tm, const float* d_hn,
                          const float* d_cn, float* grad_out, float* dx_tm, void* stream);

/* One layer of b200rnn_rnn_large_bwd: call for layer = L-1 down to 0 on the same stash / scratch (the gradient
 * flowing to the layer below stays in the scratch between calls).  When the call's kernels have run, the
 * gradients of this layer's four parameter tensors are complete; every kernel that writes one of them counts
 * itself on retire_counter (nullable) and *retire_count (nullable) receives how many counts will be added. */
int b200rnn_rnn_large_bwd_layer(const b200rnn_model_desc* d, const float* params, const float* x_tm,
                                const void* stash, void* scratch, int32_t layer, const float* d_out_tm,
                                const float* d_hn, const float* d_cn, float* grad_out, float* dx_tm,
                                uint32_t* retire_counter, int32_t* retire_count, void* stream);

/* Persistent time loops of the large-hidden LSTM (csrc/rnn/rnn_persist.cu): one cooperative launch
 * per layer and direction, W_hh sharded across the SMs' shared memory, h_t / dG_t exchanged through
 * L2 with a grid barrier per step.  b200rnn_rnn_large_fwd/bwd use them automatically when
 * b200rnn_rnn_persist_supported(d) (LSTM / GRU, hidden % 64 == 0, hidden / 8 <= #SMs, batch <= 128;
 * B200RNN_PERSIST=0 disables).  `sync` is 256 bytes of device memory, zeroed by the call.
 * b200rnn_persist_status() returns non-zero after a bounded wait inside the kernels timed out. */
int b200rnn_rnn_persist_supported(const b200rnn_model_desc* d);
int64_t b200rnn_rnn_persist_stream_bytes(int32_t H);   /* size of `stream_img` below */
int b200rnn_lstm_persist_fwd(int32_t T, int32_t B, int32_t H, float* gates, const void* whh_bf16,
                             float* c, float* h, void* stream_img, void* sync, void* stream);
int b200rnn_lstm_persist_bwd(int32_t T, int32_t B, int32_t H, const float* gates, const float* c,
                             const float* dh_ext, const float* d_hn, const float* d_cn,
                             const void* whhT_bf16, float* dG, void* dG_bf16, void* stream_img, void* sync,
                             void* stream);
/* GRU members: gates [T,B,3H] (x-projection + b_ih in, r / z / n out), gh [T,B,3H] (receives
 * W_hn h + b_hn at [2H,3H)), bhh fp32 [3H]; the backward writes dGx = (dr, dz, dn) to dG / dG_bf16 and
 * dGh = (dr, dz, dn*r) to dG2 / dG2_bf16. */
int b200rnn_gru_persist_fwd(int32_t T, int32_t B, int32_t H, float* gates, float* gh, const void* whh_bf16,
                            const float* bhh, float* h, void* stream_img, void* sync, void* stream);
int b200rnn_gru_persist_bwd(int32_t T, int32_t B, int32_t H, const float* gates, const float* gh, const float* h,
                            const float* dh_ext, const float* d_hn, const void* whhT_bf16, float* dG,
                            void* dG_bf16, float* dG2, void* dG2_bf16, void* stream_img, void* sync, void* stream);
int b200rnn_persist_status(int32_t* status);

/* Status word of the small-hidden family, sticky since load.  Bit 0: a bounded wait of the
 * layer-wavefront mode timed out (a consumer CTA never saw its producer's values; results of that
 * call are garbage; B200RNN_WAVE=0 keeps the layers of a batch tile in one CTA, in sequence).
 * Bit 1: the fused head met a label outside [0, num_classes) (it was treated as class 0). */
int b200rnn_wave_status(int32_t* status);
/* Whether the small-hidden forward / backward of this shape is launched as a layer wavefront (each then
 * issues one extra small kernel, the sentinel pre-fill, before the main one). */
int b200rnn_rnn_small_fwd_wave(const b200rnn_model_desc* d);
int b200rnn_rnn_small_bwd_wave(const b200rnn_model_desc* d);

/* ---- L2 flush helper for benchmarking (writes `bytes` of scratch). */
int b200rnn_l2_flush(void* scratch, int64_t bytes, void* stream);

/* ======================================================================================
 * Runtime bootstrap + NVLink symmetric memory + in-kernel collectives (csrc/runtime, csrc/comm).
 * Replaces what the reference gets from mpirun + c10d ProcessGroupMPI + the Horovod core
 * (src/motion/trainer/ddp.py:18-19, trainer/horovod.py:17,35,41, src/example/example_distributed.py).
 * ====================================================================================== */

/* ---- B1 rendezvous: star over TCP (rank 0 listens).  Host only, no CUDA. ---- */
typedef struct b200rnn_boot b200rnn_boot;
/* rank 0: bind + listen on *port (0 = ephemeral; the chosen port is written back). */
int b200rnn_boot_listen(int32_t* port, int32_t* listen_fd);
/* All ranks.  rank 0 passes the fd from b200rnn_boot_listen; others pass -1 and connect to addr:port. */
int b200rnn_boot_init(int32_t rank, int32_t world, const char* addr, int32_t port, int32_t listen_fd,
                      int32_t timeout_ms, b200rnn_boot** out);
int b200rnn_boot_allgather(b200rnn_boot* b, const void* send, int32_t bytes, void* recv /* world*bytes */);
int b200rnn_boot_barrier(b200rnn_boot* b);
void b200rnn_boot_destroy(b200rnn_boot* b);

/* ---- B1 topology: what the GPUs of this node can do to each other. ---- */
typedef struct {
  int32_t n_devices;
  int32_t sm_count[B200RNN_MAX_PEERS];
  int32_t multicast_supported[B200RNN_MAX_PEERS];          /* NVLS (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED) */
  int32_t can_access[B200RNN_MAX_PEERS][B200RNN_MAX_PEERS]; /* cudaDeviceCanAccessPeer */
  int32_t perf_rank[B200RNN_MAX_PEERS][B200RNN_MAX_PEERS];  /* cudaDevP2PAttrPerformanceRank (0 = best) */
  int32_t native_atomics[B200RNN_MAX_PEERS][B200RNN_MAX_PEERS];
  char name[B200RNN_MAX_PEERS][64];
} b200rnn_topology;
int b200rnn_topology_probe(b200rnn_topology* out);

/* ---- Symmetric memory: one region per rank, every region mapped into every rank.
 * b200rnn_symm_create ALLOCATES device memory (cudaMalloc + CUDA IPC export/import) and
 * enables peer access; collective over the bootstrap. ---- */
typedef struct b200rnn_symm b200rnn_symm;
int b200rnn_symm_create(b200rnn_boot* b, int32_t device, int64_t bytes, b200rnn_symm** out);
int b200rnn_symm_pointers(b200rnn_symm* s, void** out /* [world]; out[rank] is the local region */);
int b200rnn_symm_destroy(b200rnn_symm* s);

/* Symmetric memory from the CUDA virtual-memory API (cuMemCreate, POSIX-fd exchange over Unix sockets)
 * plus, when want_multicast != 0 and the node supports it, an NVSwitch multicast mapping of the same
 * regions (csrc/runtime/symm_vmm.cpp).  ALLOCATES device memory; collective over the bootstrap; `tag`
 * must be unique per node and job.  mc_out is NULL when multicast could not be set up. */
typedef struct b200rnn_symm_vmm b200rnn_symm_vmm;
int b200rnn_symm_vmm_create(b200rnn_boot* b, int32_t rank, int32_t world, int32_t device, int64_t bytes,
                            int32_t tag, int32_t want_multicast, b200rnn_symm_vmm** out);
int b200rnn_symm_vmm_pointers(b200rnn_symm_vmm* s, void** out /* [world] */, void** mc_out);
int b200rnn_symm_vmm_destroy(b200rnn_symm_vmm* s);

/* ---- N7 reducer (csrc/runtime/reducer.cpp).
 * Bucket assignment in reverse parameter order (c10d policy: 1 MiB first bucket, then `cap`):
 * bucket_of[i] = bucket index of parameter i (bucket 0 holds the LAST parameters, whose gradients are
 * ready first). */
int b200rnn_bucket_plan(const int64_t* sizes_bytes, int32_t n, int64_t first_cap, int64_t cap,
                        int32_t* bucket_of, int32_t* n_buckets);
/* Device-side "bucket ready" triggers.  A bucket is the arena slice [lo, hi) (floats).  Producers — the
 * kernels that write the bucket's gradient — count themselves on b200rnn_reducer_counter(r, bucket) as they
 * retire; the host promises how many counts a step adds (b200rnn_reducer_expect) and enqueues
 * b200rnn_reducer_wait on the communication stream in front of the bucket's all-reduce: a one-thread kernel
 * that spins until the promised counts have arrived (bounded; b200rnn_reducer_read reports a timeout as
 * bit `bucket` of the status word) and stamps %globaltimer.  b200rnn_reducer_stamp enqueues a kernel that
 * writes %globaltimer into mark `which` (1..3) of the bucket — e.g. "collective done", "next layer's
 * backward begins / ends" — so the overlap can be read back as a timeline (stamps_out[bucket][0..3], ns).
 * b200rnn_reducer_create ALLOCATES a few hundred bytes of device memory. */
typedef struct b200rnn_reducer b200rnn_reducer;
int b200rnn_reducer_create(int32_t n_buckets, const int64_t* lo, const int64_t* hi, b200rnn_reducer** out);
int b200rnn_reducer_destroy(b200rnn_reducer* r);
int b200rnn_reducer_range(const b200rnn_reducer* r, int32_t bucket, int64_t* lo, int64_t* hi);
uint32_t* b200rnn_reducer_counter(b200rnn_reducer* r, int32_t bucket);
int b200rnn_reducer_expect(b200rnn_reducer* r, int32_t bucket, int32_t counts);
int b200rnn_reducer_wait(b200rnn_reducer* r, int32_t bucket, void* stream);
int b200rnn_reducer_stamp(b200rnn_reducer* r, int32_t bucket, int32_t which, void* stream);
int b200rnn_reducer_read(b200rnn_reducer* r, uint64_t* stamps_out, int32_t* status_out);

/* ---- Communicator context handed (by value) to every comm kernel. ---- */
typedef struct {
  int32_t rank, world;
  int32_t max_blocks;       /* flag slots per source rank (blocks of the all-reduce) */
  int32_t delay_ns;         /* injected delay before publishing flags: the analogue of the
                               reference's tc-netem experiment (fabfile.py:130-191) */
  int64_t inbox_floats;     /* capacity per (parity, source) inbox */
  int64_t timeout_cycles;   /* watchdog for flag waits (0 = default) */
  int64_t heap_bytes;
  int64_t ll_floats;        /* capacity (floats, per (parity, source)) of the low-latency one-shot inbox;
                               0 = default (64 Ki floats) */
  int32_t loss_ppm;         /* injected message loss, parts per million: a block's publication is "dropped" with this
                               probability and then costs loss_rto_ns per retransmission (tc netem loss analogue) */
  int32_t loss_rto_ns;      /* retransmission timeout charged per dropped message (0 = 200 000 ns) */
  /* byte offsets inside a region (identical on every rank) */
  int64_t off_flags_ar, off_flags_pr, off_flags_sig, off_flags_bar, off_status, off_inbox, off_heap, off_ll;
  void* base[B200RNN_MAX_PEERS];
} b200rnn_comm_ctx;

/* Fills the off_* fields from world / max_blocks / inbox_floats / heap_bytes; returns region bytes. */
int64_t b200rnn_comm_layout(b200rnn_comm_ctx* ctx);

enum { B200RNN_ALGO_AUTO = 0, B200RNN_ALGO_ONESHOT = 1, B200RNN_ALGO_TWOSHOT = 2, B200RNN_ALGO_NVLS = 3,
       B200RNN_ALGO_LL = 4 /* one-shot, (data, seq) words: no fence, no flags; vectors <= ll_floats */ };

/* K7: ONE kernel = fixed-order reduction of the n_rows local gradient rows + all-reduce over
 * NVLink peer memory (one-shot: every rank pushes its vector into every peer's inbox; two-shot:
 * reduce-scatter to the owning rank, optimizer on the owner's shard, all-gather of the updated
 * PARAMETERS by peer stores) + gradient averaging (opt->grad_scale) + Adam/SGD step.  Ranks sum
 * in rank order 0..W-1, so all ranks end bit-identical.  For two-shot `params` must live in the
 * symmetric heap at heap offset params_heap_off.  metrics[2] (nullable) gets this rank's
 * (loss sum, correct count) from the row tail.  seq must increase by 1 per call on all ranks. */
int b200rnn_comm_allreduce_update(const b200rnn_comm_ctx* ctx, uint32_t seq, const float* rows,
                                  int32_t n_rows, int64_t n_params, float* params,
                                  int64_t params_heap_off, float* exp_avg, float* exp_avg_sq,
                                  const b200rnn_optim_desc* opt, float* metrics, int32_t algo,
                                  void* stream);
/* NVLS form of K7: region from b200rnn_symm_vmm_create with multicast; the parameter arena
 * (params_heap_off) and a gradient slot of the same padded size (grad_heap_off) live in the heap.
 * The switch performs the reduction (multimem.ld_reduce) and the parameter broadcast (multimem.st). */
int b200rnn_comm_allreduce_update_nvls(const b200rnn_comm_ctx* ctx, void* mc_base, uint32_t seq, const float* rows,
                                       int32_t n_rows, int64_t n_params, float* params, int64_t params_heap_off,
                                       int64_t grad_heap_off, float* exp_avg, float* exp_avg_sq,
                                       const b200rnn_optim_desc* opt, float* metrics, void* stream);
/* Plain in-place all-reduce (sum * scale) of buf[n] (n <= inbox_floats), same kernels without
 * the optimizer epilogue.  Shares the seq counter with b200rnn_comm_allreduce_update. */
int b200rnn_comm_allreduce(const b200rnn_comm_ctx* ctx, uint32_t seq, float* buf, int64_t n,
                           float scale, void* stream);
/* Same with an explicit one-shot form (B200RNN_ALGO_LL or B200RNN_ALGO_ONESHOT): what the latency sweep times. */
int b200rnn_comm_allreduce_algo(const b200rnn_comm_ctx* ctx, uint32_t seq, float* buf, int64_t n,
                                float scale, int32_t algo, void* stream);
/* K8: broadcast `bytes` at heap offset `heap_off` from root's heap into every rank's heap. */
int b200rnn_comm_broadcast(const b200rnn_comm_ctx* ctx, uint32_t seq, int64_t heap_off, int64_t bytes,
                           int32_t root, void* stream);
/* K9: one-sided put of src[bytes] into dst's heap at heap_off, then signal; wait for a signal. */
int b200rnn_comm_put(const b200rnn_comm_ctx* ctx, uint32_t seq, int32_t dst, const void* src,
                     int64_t heap_off, int64_t bytes, void* stream);
int b200rnn_comm_wait(const b200rnn_comm_ctx* ctx, uint32_t seq, int32_t src, void* stream);
/* K10: one-sided parameter server (the src/motion/param_server package).  The server rank hosts the shared
 * parameters at heap offset heap_off of ITS symmetric heap and does nothing else.  ps_pull copies them
 * into dst with peer loads; ps_push reduces the worker's gradient rows, runs the worker's OWN
 * optimizer state (exp_avg / exp_avg_sq live on the worker) and adds the resulting parameter delta to
 * the server's vector with red.global.add over NVLink.  Asynchronous: no ordering between workers. */
int b200rnn_comm_ps_push(const b200rnn_comm_ctx* ctx, int32_t server, int64_t heap_off, const float* rows,
                         int32_t n_rows, int64_t n_params, float* exp_avg, float* exp_avg_sq,
                         const b200rnn_optim_desc* opt, float* metrics, void* stream);
int b200rnn_comm_ps_pull(const b200rnn_comm_ctx* ctx, int32_t server, int64_t heap_off, float* dst,
                         int64_t n_params, void* stream);
/* Tell the library that a (new) symmetric region starts at region_base: resets its host-side
 * bookkeeping for that address.  Call once after b200rnn_symm_create. */
int b200rnn_comm_forget(const void* region_base);
/* Device-side barrier across all ranks. */
int b200rnn_comm_barrier(const b200rnn_comm_ctx* ctx, uint32_t seq, void* stream);
/* Reads this rank's watchdog status word (synchronises the stream): 0 = ok, else a bit mask of
 * (1 << source rank) whose flag never arrived. */
int b200rnn_comm_status(const b200rnn_comm_ctx* ctx, int32_t* status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200RNN_H_ */
