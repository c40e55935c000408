/*
 * pdt.h -- C ABI of the B200-native data-parallel training step.
 *
 * The reference (Yun-960/Pytorch-Distributed-Template) is pure Python and has no FFI; the
 * functions below are what a binding for its hot path -- Trainer._train_epoch's inner loop,
 * trainer/trainer.py:45-72 -- would call.  Each declaration cites the reference lines whose
 * work it replaces.  The Python side (pytorch-distributed-template_b200/pdt/native.py) binds
 * them with ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *  - every function returns 0 on success or a negative PDT_E* code; pdt_last_error() gives
 *    the message for the calling thread.  Nothing throws across the boundary.
 *  - all pointers are DEVICE pointers owned by the caller (16-byte aligned, contiguous);
 *    no ownership is transferred and, unless a comment says so, nothing is allocated.
 *  - `stream` is a cudaStream_t passed as void*; all work is asynchronous on it.
 *  - n is the number of samples on this GPU.  Internal activations use the "lane" layout:
 *    samples are grouped in tiles of 32 and stored [tile][feature][32], so buffers named
 *    *_lane must hold pdt_num_tiles(n) * features * 32 elements.  One buffer differs: p1_lane
 *    (conv1's pooled output, 1,440 features = [10][12][12]) is stored [tile][feature/4][32][4] --
 *    four consecutive features (columns) of one sample are one 16-byte group; element (f, lane)
 *    of a tile is at ((f / 4) * 32 + lane) * 4 + f % 4.  Same size, same tiles.
 *  - the 21,840 MnistModel parameters are addressed through one flat fp32 vector in
 *    state_dict order: conv1.weight, conv1.bias, conv2.weight, conv2.bias, fc1.weight,
 *    fc1.bias, fc2.weight, fc2.bias (offsets PDT_OFF_*).
 */
#ifndef PDT_H
#define PDT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDT_OK 0
#define PDT_EINVAL (-1)
#define PDT_ECUDA (-2)
#define PDT_ENCCL (-3)

#define PDT_TILE 32
#define PDT_NUM_PARAMS 21840
#define PDT_OFF_W1 0
#define PDT_OFF_B1 250
#define PDT_OFF_W2 260
#define PDT_OFF_B2 5260
#define PDT_OFF_FW1 5280
#define PDT_OFF_FB1 21280
#define PDT_OFF_FW2 21330
#define PDT_OFF_FB2 21830
#define PDT_P1_FEATS 1440 /* 10 x 12 x 12 */
#define PDT_P2_FEATS 320  /* 20 x 4 x 4  */
#define PDT_H_FEATS 50
#define PDT_IDX1_BYTES 360 /* per sample: 10 ch x 36 blocks, 4 x 2-bit argmax each */
#define PDT_IDX2_BYTES 320 /* per sample: one byte (2-bit argmax code) per pooled conv2 output, same index as p2 */
#define PDT_TAIL 4         /* floats after the gradients in the flat buffer */

/* loss/metric record written by pdt_nll_metrics_fwd (8 x 4 bytes) */
typedef struct {
    float loss_mean;  /* loss_sum / n_valid                                  */
    float loss_sum;   /* -sum logp[n, t_n] over valid rows                   */
    int32_t top1;     /* rows whose argmax == target                         */
    int32_t topk;     /* rows whose target is among the k largest            */
    int32_t n_valid;  /* rows with target != ignore_index and 0 <= t < C     */
    int32_t n_rows;
    int32_t pad[2];
} pdt_stats;

/* hyper-parameters live on the device so a captured CUDA graph sees new values */
typedef struct {
    double lr, momentum, dampening, weight_decay; /* doubles, as torch keeps them in Python */
    int32_t nesterov;
    int32_t step; /* number of updates done so far; advanced by the kernel */
    uint32_t ticket; /* must start at 0 */
    int32_t pad;
} pdt_sgd_hyper; /* 48 bytes; the first 40 are the part a host re-uploads when lr changes */

typedef struct {
    double lr, beta1, beta2, eps, weight_decay;
    int32_t amsgrad;
    int32_t step; /* number of updates done so far; advanced by the kernel */
    uint32_t ticket; /* must start at 0 */
    int32_t pad;
} pdt_adam_hyper; /* 56 bytes; the first 44 are the host-owned part */

const char* pdt_last_error(void);
int pdt_version(void);
int64_t pdt_num_tiles(int64_t n);
/* bytes of scratch the backward pass needs for n samples (per-tile partial gradients) */
int64_t pdt_mnist_scratch_bytes(int64_t n);
/* bytes of scratch pdt_nll_metrics_fwd needs for n rows */
int64_t pdt_nll_scratch_bytes(int64_t n);

/* ---- forward: model/model.py:15-22 ------------------------------------------------- */

/* model.py:16  relu(max_pool2d(conv1(x), 2)).  x [n,1,28,28] row-major; w [10,1,5,5]; b [10].
 * Writes p1_lane [T][360][32][4] (see "lane" layout above) and the 2-bit pool argmax codes
 * idx1_lane [T][360][32] (byte (c*36 + by*6 + bx) = the four codes of the 2x2 block (by,bx) of pooled
 * positions, 2 bits each, row-major). */
int pdt_conv1_pool_relu_fwd(const float* x, const float* w, const float* b, float* p1_lane,
                            uint8_t* idx1_lane, int64_t n, void* stream);

/* The *_u8 variants take the batch as RAW BYTES, x [n,1,28,28] uint8 (what an MNIST file holds), and apply the
 * reference loader's ToTensor + Normalize((mean,), (std,)) -- data_loader/data_loaders.py:13-16: (v / 255 - mean) / std,
 * two fp32 operations -- while loading each pixel, so the host ships 1 byte per pixel instead of 4
 * (trainer/trainer.py:46) and the kernels see bit-identical values.  Otherwise identical to their namesakes. */
int pdt_conv1_pool_relu_fwd_u8(const uint8_t* x, float mean, float std, const float* w, const float* b,
                               float* p1_lane, uint8_t* idx1_lane, int64_t n, void* stream);

/* model.py:17  relu(max_pool2d(conv2_drop(conv2(p1)), 2)).  keep2d [n,20] uint8 keep mask of
 * Dropout2d (NULL = eval mode, no dropout); kept channels are multiplied by `scale` (2.0 for
 * p=0.5).  Writes p2_lane [T][320][32] (feature order = view(-1,320)) and idx2_lane. */
int pdt_conv2_drop_pool_relu_fwd(const float* p1_lane, const float* w, const float* b,
                                 const uint8_t* keep2d, float scale, float* p2_lane,
                                 uint8_t* idx2_lane, int64_t n, void* stream);

/* The same call on the tensor cores (tcgen05.mma, accumulators in tensor memory): conv2 as an implicit GEMM of
 * 128 (2 samples x 64 positions) x 32 x 8 tiles in split-TF32 arithmetic (three TF32 passes over hi/lo parts of both
 * operands, fp32 accumulation: equal to the fp32 result within its own rounding noise).  Same arguments, layouts,
 * outputs and pooling / dropout rules as pdt_conv2_drop_pool_relu_fwd (model.py:17). */
int pdt_conv2_drop_pool_relu_fwd_tc(const float* p1_lane, const float* w, const float* b,
                                    const uint8_t* keep2d, float scale, float* p2_lane,
                                    uint8_t* idx2_lane, int64_t n, void* stream);

/* model.py:18-22  log_softmax(fc2(dropout(relu(fc1(p2))))).  keep1d [n,50] uint8 (NULL = eval).
 * Writes hd_lane [T][50][32] (post-dropout hidden) and logp [n,10] row-major. */
int pdt_head_fwd(const float* p2_lane, const float* fc1w, const float* fc1b, const float* fc2w,
                 const float* fc2b, const uint8_t* keep1d, float scale, float* hd_lane,
                 float* logp, int64_t n, void* stream);

/* pdt_head_fwd with the criterion folded into its tail (trainer/trainer.py:49-50: output = model(data);
 * loss = criterion(output, target)): the CTA that forms a tile's log-probabilities also takes that tile's
 * nll_loss terms and top-1 / top-k counts (same rules as pdt_nll_metrics_fwd), and the last tile to finish
 * folds all tiles in a fixed order into `out`.  classes = 10, logp is still written in full.
 * scratch: pdt_head_nll_scratch_bytes(n) bytes, zero-initialised once by the caller.
 * ring (may be NULL): as pdt_nll_metrics_fwd_log. */
int64_t pdt_head_nll_scratch_bytes(int64_t n);
int pdt_head_fwd_nll(const float* p2_lane, const float* fc1w, const float* fc1b, const float* fc2w,
                     const float* fc2b, const uint8_t* keep1d, float scale, float* hd_lane, float* logp,
                     int64_t n, const int64_t* target, int topk, int64_t ignore_index, pdt_stats* out,
                     void* scratch, float* ring, int32_t* ring_count, int ring_capacity, void* stream);

/* The middle of a TRAINING step as one launch (trainer/trainer.py:49-57 for the default model and criterion):
 * pdt_conv2_drop_pool_relu_fwd + pdt_head_fwd_nll + pdt_head_bwd_nll.  The four CTAs that compute conv2 for a
 * tile form a thread-block cluster; fc1 is split over their features (one exchange through distributed shared
 * memory), and every output of the three calls is produced: p2_lane, idx2_lane, hd_lane, logp, `out`,
 * dq2_lane and the fc1/fc2 partial gradients in `scratch`.  The loss gradient is taken for grad_loss = 1:
 * pass the real one to pdt_grad_reduce / pdt_adam_step_reduce as grad_scale_dev.  n <= PDT_TRAIN_FUSED_MAX_ROWS.
 * nll_scratch: pdt_head_nll_scratch_bytes(n) bytes, zero-initialised once. */
int pdt_conv2_head_train(const float* p1_lane, const float* w, const float* b, const uint8_t* keep2d,
                         float scale2d, float* p2_lane, uint8_t* idx2_lane, const float* fc1w,
                         const float* fc1b, const float* fc2w, const float* fc2b, const uint8_t* keep1d,
                         float scale1d, float* hd_lane, float* logp, const int64_t* target, int topk,
                         int64_t ignore_index, pdt_stats* out, void* nll_scratch, float* ring,
                         int32_t* ring_count, int ring_capacity, float* dq2_lane, void* scratch,
                         int64_t n, void* stream);

/* pdt_conv2_head_train with conv2 on the tensor cores (the arithmetic of pdt_conv2_drop_pool_relu_fwd_tc): CTA q of a
 * tile's 4-CTA cluster computes conv2 for samples 8q..8q+7 and the CTAs exchange 80-feature x 8-sample blocks through
 * distributed shared memory before the head phases.  Same arguments, outputs and limits. */
int pdt_conv2_head_train_tc(const float* p1_lane, const float* w, const float* b, const uint8_t* keep2d,
                            float scale2d, float* p2_lane, uint8_t* idx2_lane, const float* fc1w,
                            const float* fc1b, const float* fc2w, const float* fc2b, const uint8_t* keep1d,
                            float scale1d, float* hd_lane, float* logp, const int64_t* target, int topk,
                            int64_t ignore_index, pdt_stats* out, void* nll_scratch, float* ring,
                            int32_t* ring_count, int ring_capacity, float* dq2_lane, void* scratch,
                            int64_t n, void* stream);

/* Dropout keep masks (model.py:11,17,20) from a counter-based generator: keep2d [n,20] and
 * keep1d [n,50], each unit kept with probability 1-p.  rng_state is three uint64 on the device
 * {seed, offset, 0}; the kernel advances offset so a replayed graph draws fresh masks. */
int pdt_dropout_masks(uint8_t* keep2d, uint8_t* keep1d, int64_t n, float p, uint64_t* rng_state,
                      void* stream);

/* pdt_conv1_pool_relu_fwd and pdt_dropout_masks as ONE launch (training forward): the mask CTAs ride at the
 * end of conv1's grid.  Same outputs and the same mask stream as the two separate calls. */
int pdt_conv1_fwd_masks(const float* x, const float* w, const float* b, float* p1_lane, uint8_t* idx1_lane,
                        uint8_t* keep2d, uint8_t* keep1d, float p, uint64_t* rng_state, int64_t n, void* stream);

int pdt_conv1_fwd_masks_u8(const uint8_t* x, float mean, float std, const float* w, const float* b, float* p1_lane,
                           uint8_t* idx1_lane, uint8_t* keep2d, uint8_t* keep1d, float p, uint64_t* rng_state, int64_t n,
                           void* stream);

/* ---- loss + metrics: model/loss.py:4-5, model/metric.py:4-20 ------------------------- */

/* nll_loss (mean over rows with target != ignore_index) and the integer correct-counts of
 * accuracy / top_k_acc in one pass.  logp [n,classes] row-major, target int64 [n].
 * Rank convention under ties: class j is ahead of the target t when o_j > o_t, or
 * o_j == o_t and j < t (equals torch.argmax's first-index rule at k=1).
 * scratch: pdt_nll_scratch_bytes(n) bytes, zero-initialised once by the caller. */
int pdt_nll_metrics_fwd(const float* logp, const int64_t* target, int64_t n, int classes, int topk,
                        int64_t ignore_index, pdt_stats* out, void* scratch, void* stream);

/* pdt_nll_metrics_fwd that also appends the mean loss to a device ring, ring[(*ring_count)++ % capacity]
 * (what pdt_ring_push does as a separate launch): the trainer's per-step loss log at no extra launch. */
int pdt_nll_metrics_fwd_log(const float* logp, const int64_t* target, int64_t n, int classes, int topk,
                            int64_t ignore_index, pdt_stats* out, void* scratch, float* ring,
                            int32_t* ring_count, int ring_capacity, void* stream);

#define PDT_TRAIN_FUSED_MAX_ROWS 8192 /* largest n pdt_conv2_head_train accepts */

/* Same result as pdt_nll_metrics_fwd, for callers that are not allowed to allocate device memory.
 * This entry point has no counterpart in the reference.  Some hosts hand every byte of the GPU to
 * their own pool at start-up (a serving process, another framework's arena) and forbid
 * cudaMalloc afterwards; others sit in a latency-critical loop where cudaMalloc's device-wide
 * synchronisation is not acceptable.  Such a caller has nowhere to get the scratch buffer the
 * regular entry point wants, so this one brings its own: a 2 KB array built into the library
 * image (each GPU that loads the library holds it until the process exits).  It neither allocates
 * nor frees anything and calls no allocating routine.
 *   - limit: n <= 16,384 rows per call (PDT_NLL_STATIC_MAX_ROWS); larger n returns PDT_EINVAL with
 *     a message naming the limit -- use pdt_nll_metrics_fwd with your own scratch instead.
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress,
 *     with that device current) and caches it; make that first call outside any stream capture,
 *     and from one host thread at a time (later calls only read the cache).  The cached address
 *     does not survive cudaDeviceReset.
 *   - all calls on a device share the one array.  Each call clears the part that must start at
 *     zero, on `stream`, before its kernel; two calls that overlap on the GPU (different streams,
 *     no ordering) race on it and give undefined results.  Calls ordered on one stream are fine. */
#define PDT_NLL_STATIC_MAX_ROWS 16384
int pdt_nll_metrics_fwd_static(const float* logp, const int64_t* target, int64_t n, int classes,
                               int topk, int64_t ignore_index, pdt_stats* out, void* stream);

/* backward of nll_loss: dlogp[n,c] = -(grad_loss / n_valid) at c == target, else 0.
 * grad_loss may be NULL (treated as 1.0). */
int pdt_nll_bwd(const int64_t* target, const pdt_stats* stats, const float* grad_loss, float* dlogp,
                int64_t n, int classes, int64_t ignore_index, void* stream);

/* ---- backward: autograd of model.py:15-22 (trainer/trainer.py:57) -------------------- */

/* log_softmax, fc2, dropout, relu, fc1 backward.  Writes dq2_lane [T][320][32] = gradient
 * reaching conv2's pooled output (already masked by relu and scaled by the dropout factor),
 * and per-tile partial gradients of fc1/fc2 into `scratch`. */
int pdt_head_bwd(const float* dlogp, const float* logp, const float* hd_lane, const float* p2_lane,
                 const float* fc1w, const float* fc2w, float scale1d, float scale2d,
                 float* dq2_lane, void* scratch, int64_t n, void* stream);

/* pdt_nll_bwd + pdt_head_bwd as one launch: the gradient of nll_loss w.r.t. the log-probabilities is formed
 * in registers from target / stats->n_valid / grad_loss (NULL = 1.0) instead of being read from a dlogp
 * buffer.  Valid when the loss is the only consumer of the model output (the trainer's step). */
int pdt_head_bwd_nll(const int64_t* target, const pdt_stats* stats, const float* grad_loss, int64_t ignore_index,
                     const float* logp, const float* hd_lane, const float* p2_lane, const float* fc1w,
                     const float* fc2w, float scale1d, float scale2d, float* dq2_lane, void* scratch,
                     int64_t n, void* stream);

/* conv2 data gradient: dq1_lane [T][1440][32] = dL/dp1 masked by p1 > 0. */
int pdt_conv2_bwd_data(const float* dq2_lane, const uint8_t* idx2_lane, const float* w,
                       const float* p1_lane, float* dq1_lane, int64_t n, void* stream);

/* pdt_conv2_bwd_data + pdt_conv1_bwd_weight as ONE launch: the CTA that forms a quadrant of the conv2 data
 * gradient goes on to correlate it with the input image (x [n,1,28,28]) for conv1's weight/bias partials, so
 * dq1_lane is never written to memory.  Writes conv1's partials into `scratch` (same slots as
 * pdt_conv1_bwd_weight; summation grouped by quadrant instead of by row band). */
int pdt_conv2_bwd_data_conv1_bwd_weight(const float* dq2_lane, const uint8_t* idx2_lane, const float* w,
                                        const float* p1_lane, const float* x, const uint8_t* idx1_lane,
                                        void* scratch, int64_t n, void* stream);

int pdt_conv2_bwd_data_conv1_bwd_weight_u8(const float* dq2_lane, const uint8_t* idx2_lane, const float* w,
                                           const float* p1_lane, const uint8_t* x, float mean, float std,
                                           const uint8_t* idx1_lane, void* scratch, int64_t n, void* stream);

/* conv2 weight/bias gradient partials into scratch. */
int pdt_conv2_bwd_weight(const float* dq2_lane, const uint8_t* idx2_lane, const float* p1_lane,
                         void* scratch, int64_t n, void* stream);

/* conv1 weight/bias gradient partials into scratch (x needs no gradient, trainer.py:46). */
int pdt_conv1_bwd_weight(const float* dq1_lane, const uint8_t* idx1_lane, const float* x,
                         void* scratch, int64_t n, void* stream);

int pdt_conv1_bwd_weight_u8(const float* dq1_lane, const uint8_t* idx1_lane, const uint8_t* x, float mean, float std,
                            void* scratch, int64_t n, void* stream);

/* Sum the per-tile partials in a fixed order into the flat gradient vector:
 *   grad[i] = beta * grad[i] + scale * sum_slots partial[slot][i]      (i < 21840)
 * and, when stats != NULL, write the tail grad[21840..21843] = {loss_mean * scale, top1, topk,
 * n_valid} so that ONE all-reduce carries gradients, BaseTrainer.reduce_loss
 * (base/base_trainer.py:165-174) and the metric counts.  scale = 1/world_size reproduces DDP's
 * averaging (train.py:46-52).  grad_scale_dev (may be NULL = 1): one float on the device that also multiplies
 * the sums -- the gradient of the loss that backward() was called with, when the partials were formed for 1. */
int pdt_grad_reduce(const void* scratch, int64_t n, float beta, float scale, const pdt_stats* stats,
                    float* flat_grad, const float* grad_scale_dev, void* stream);

/* ---- flat buffer + collective: DDP bucket (train.py:46-52), reduce_loss ---------------- */

/* dst[i] = src[i] * scale, fp32 -> {fp32 (dst_dtype 0), bf16 (1)}; in place allowed for fp32. */
int pdt_flat_scale_cast(const float* src, void* dst, int64_t count, float scale, int dst_dtype,
                        void* stream);

/* NCCL communicator over NVLink/NVSwitch.  pdt_comm_unique_id fills 128 bytes on one rank; the
 * caller ships them to every rank (e.g. through torch.distributed) and all call pdt_comm_init.
 * These allocate NCCL's own resources; pdt_comm_destroy releases them. */
int pdt_comm_unique_id(void* id128);
int pdt_comm_init(const void* id128, int rank, int world, void** comm);
int pdt_comm_destroy(void* comm);
/* in-place sum all-reduce of `count` fp32 values (gradients + tail). */
int pdt_allreduce_flat(void* comm, float* buf, int64_t count, void* stream);
/* NCCL user-buffer registration for a long-lived flat buffer (DDP's bucket, train.py:46-52, is re-used every step):
 * pdt_comm_mem_alloc returns memory from ncclMemAlloc (this DOES allocate; pdt_comm_mem_free releases it),
 * pdt_comm_register / pdt_comm_deregister wrap ncclCommRegister / ncclCommDeregister.  All-reduces on a registered
 * buffer go through pdt_allreduce_flat unchanged. */
int pdt_comm_mem_alloc(int64_t bytes, void** dev_ptr);
int pdt_comm_mem_free(void* dev_ptr);
int pdt_comm_register(void* comm, void* buf, int64_t bytes, void** handle);
int pdt_comm_deregister(void* comm, void* handle);

/* One-shot all-reduce over NVLink peer memory for SMALL buffers (the 87 KB MNIST gradient: latency-bound
 * under NCCL).  Replaces the same reference lines as pdt_allreduce_flat (DDP's bucket all-reduce,
 * train.py:46-52, and reduce_loss, base_trainer.py:165-174); SURVEY.md 8(f) N4.
 *   pdt_p2p_alloc   cudaMalloc's this rank's exchange block (pdt_p2p_bytes(capacity) bytes, zeroed) and
 *                   returns its 64-byte CUDA IPC handle; the caller ships the handles to every rank.
 *   pdt_p2p_open    maps a peer's block (cudaIpcOpenMemHandle); the rank's own entry is its dev_ptr.
 *   pdt_p2p_allreduce  in-place fp32 sum of buf[0..count) over the W ranks, one kernel: copy-in, flag
 *                   exchange with system-scope release/acquire, rank-ordered sum (bit-identical on all
 *                   ranks).  peer_bases: HOST array of W mapped base pointers.  count <= capacity, buf
 *                   16-byte aligned and padded to a multiple of 4 floats, W <= 8.  Every rank must call
 *                   it the same number of times.  A wait for a peer is bounded (pdt_p2p_set_timeout_ms, default
 *                   30,000 ms, or PDT_P2P_TIMEOUT_MS in the environment): on time-out the rank raises an error word
 *                   on EVERY rank and the call leaves buf untouched; every later call is a no-op as well until the
 *                   host has seen the error (pdt_p2p_error != 0 -> stop training; DDP/NCCL would block instead,
 *                   train.py:46-52). */
int64_t pdt_p2p_bytes(int64_t capacity_floats);
int pdt_p2p_alloc(int64_t capacity_floats, void** dev_ptr, void* ipc_handle64);
int pdt_p2p_open(const void* ipc_handle64, void** peer_ptr);
int pdt_p2p_close(void* peer_ptr);
int pdt_p2p_free(void* dev_ptr);
int pdt_p2p_allreduce(void* const* peer_bases, int rank, int world, float* buf, int64_t count,
                      int64_t capacity_floats, void* stream);
int pdt_p2p_set_timeout_ms(int64_t ms);

/* The data-parallel step's tail as ONE kernel over peer memory: pdt_grad_reduce (per-tile partials -> flat
 * gradient, x scale, x *grad_scale_dev) + pdt_p2p_allreduce (sum over the ranks) + pdt_adam_step_flat, for the
 * 21,840 MnistModel parameters and the 4-float tail.  flat_grad receives the summed gradient and tail, as after
 * the three separate calls.  Same exchange allocation and protocol as pdt_p2p_allreduce (calls may be mixed).
 * On a time-out (see pdt_p2p_allreduce) nothing is written: flat_grad, the weights, the Adam state and
 * hyper->step keep their values. */
int pdt_p2p_reduce_adam(void* const* peer_bases, int rank, int world, const void* scratch, int64_t n,
                        float beta, float scale, const pdt_stats* stats, const float* grad_scale_dev,
                        float* flat_grad, float* param, float* exp_avg, float* exp_avg_sq,
                        float* max_exp_avg_sq, pdt_adam_hyper* hyper, int64_t capacity_floats,
                        void* stream);
int pdt_p2p_error(const void* own_base, int* error_out);
/* the same word copied asynchronously on `stream` into pinned host memory: a training loop can look at it one
 * log interval later without draining the stream (trainer/trainer.py:60-62 is where the reference synchronises). */
int pdt_p2p_error_async(const void* own_base, int* pinned_out, void* stream);

/* ---- optimizer: optimizer.step() trainer/trainer.py:58; torch/optim/{sgd,adam}.py ------ */

/* One fused launch over the whole flat parameter vector.  `grad_scale` multiplies the gradient
 * first (1/world_size when the all-reduce summed unscaled gradients, else 1).  State vectors
 * have `count` elements; momentum_buf may be NULL when momentum == 0; vmax may be NULL when
 * amsgrad == 0.  Hyper-parameters are read from device memory; hyper->step is advanced. */
int pdt_sgd_step_flat(float* param, const float* grad, float* momentum_buf, int64_t count,
                      float grad_scale, pdt_sgd_hyper* hyper, void* stream);
int pdt_adam_step_flat(float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                       float* max_exp_avg_sq, int64_t count, float grad_scale,
                       pdt_adam_hyper* hyper, void* stream);

/* pdt_grad_reduce followed by pdt_adam_step_flat as ONE launch, for steps with no collective in between
 * (single GPU): sums the per-tile partials of n samples in the same fixed order, writes the flat gradient
 * (+ tail when stats != NULL) and applies Adam to the 21,840 MnistModel parameters. */
int pdt_adam_step_reduce(const void* scratch, int64_t n, float beta, float scale, const pdt_stats* stats,
                         float* flat_grad, float* param, float* exp_avg, float* exp_avg_sq,
                         float* max_exp_avg_sq, pdt_adam_hyper* hyper, const float* grad_scale_dev, void* stream);

/* ---- host-loop plumbing: trainer/trainer.py:46 (input copy) and :62 (loss read-back) -------- */

/* cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, stream): the per-step input copy from pinned host
 * memory without a framework dispatch in between (the host loop, not the GPU, bounds small steps). */
int pdt_memcpy_async(void* dst, const void* src, int64_t bytes, void* stream);

/* ring[(*counter)++ % capacity] = *value, on the device: a captured step logs its loss without a
 * per-step host copy; the trainer reads the ring at its log steps. */
int pdt_ring_push(float* ring, int32_t* counter, const float* value, int capacity, void* stream);

/* ---- measurement helper ------------------------------------------------------------- */

/* FP32 FMA micro-benchmark used to measure the CUDA-core roofline on the box: every thread
 * runs `iters` x 64 FMAs as 8 independent dependent-chains.  mode 0: scalar FFMA; mode 1: the packed
 * fma.rn.f32x2 of sm_100 (FFMA2, two FMAs per lane per issue slot); mode 2: FFMA2 interleaved with
 * shared-memory loads.  out[0] receives a checksum so the work cannot be optimised away. */
int pdt_fma_peak(float* out, int blocks, int threads, int iters, int mode, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PDT_H */
