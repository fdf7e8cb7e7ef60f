/* b200ds.h -- C ABI of libb200ds.so: the B200 (sm_100a) implementation of the one
 * data-parallel hot path of dashabalashova/distributed-gpu-test, the training step of
 * train.py (forward, BCE-with-logits loss, backward, gradient all-reduce, Adam).
 *
 * The reference has no FFI of its own: train.py reaches its arithmetic through the
 * Python APIs of torch and DeepSpeed.  Each entry point below therefore cites the
 * reference CALL SITE (file:line into the reference repo) whose work it replaces; the
 * Python shim `b200ds` (distributed-gpu-test_b200/b200ds) mirrors the DeepSpeed names
 * train.py uses and binds these symbols with ctypes.
 *
 * Conventions
 *   - every pointer named *_dev / documented "device" is a raw CUDA device pointer
 *     (tensor.data_ptr()); the caller (torch) owns the memory and keeps it alive until
 *     the stream work that uses it has finished.  Kernels that need scratch take it from the
 *     caller (see the *_scratch_bytes queries); the per-kernel entry points allocate nothing.
 *     A context owns its side stream, events and NCCL communicator; two calls make it own more,
 *     each on FIRST use only: b200ds_ctx_enable_p2p (the peer-mapped exchange buffer, a flag row,
 *     a pinned mailbox) and b200ds_train_step_host_prefetch (a copy stream and event, a capture
 *     stream, one device word and one pinned word, and the CUDA graphs of the step) -- all freed
 *     by b200ds_ctx_destroy.
 *   - `stream` is a cudaStream_t passed as uintptr_t (0 = legacy default stream).  All
 *     work is stream-ordered; no entry point synchronises the host unless it says so.
 *   - return value: 0 on success, otherwise one of B200DS_ERR_*; b200ds_last_error()
 *     returns a thread-local message.  Nothing throws across this boundary.
 *   - one caller thread per context.
 *   - flat parameter layout (fp32), the order of SimpleNet.parameters() (train.py:29-33):
 *       [ W1 (H*F, row-major [H,F]) | b1 (H) | W2 (H) | b2 (1) ],  P = H*F + 2H + 1.
 *     Gradients, Adam exp_avg and exp_avg_sq use the same layout.
 */
#ifndef B200DS_H_
#define B200DS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200DS_API __attribute__((visibility("default")))
#else
#define B200DS_API
#endif

#define B200DS_OK 0
#define B200DS_ERR_ARG 1      /* bad argument / unsupported shape */
#define B200DS_ERR_CUDA 2     /* a CUDA runtime call failed */
#define B200DS_ERR_NCCL 3     /* an NCCL call failed */
#define B200DS_ERR_STATE 4    /* call made in the wrong state (e.g. no communicator) */

#define B200DS_NCCL_ID_BYTES 128

typedef struct b200ds_ctx b200ds_ctx;

/* ---- library ------------------------------------------------------------------- */
B200DS_API const char* b200ds_last_error(void);
B200DS_API int b200ds_version(void);                       /* 10000*major + 100*minor + patch */
B200DS_API int b200ds_device_sm_count(int device, int* sm_count);

/* ---- context: replaces deepspeed.init_distributed() + set_device (train.py:59-61) --
 * rank/world come from the launcher's RANK/WORLD_SIZE (train.py:56-57).  world == 1
 * needs no id.  For world > 1 rank 0 calls b200ds_nccl_unique_id and publishes the 128
 * bytes (the shim uses the torch TCP store at MASTER_ADDR:MASTER_PORT); every rank then
 * calls b200ds_ctx_create with the same bytes.  Collective: all ranks must call. */
B200DS_API int b200ds_nccl_unique_id(void* id128);
B200DS_API int b200ds_ctx_create(int device, int rank, int world, const void* nccl_id128, b200ds_ctx** out);
B200DS_API int b200ds_ctx_destroy(b200ds_ctx* ctx);
B200DS_API int b200ds_ctx_info(const b200ds_ctx* ctx, int* device, int* rank, int* world);

/* ---- synthetic data: replaces make_synthetic_data (train.py:19-24) for the
 * generated-batch configs.  Philox4x32-10 + Box-Muller with the element mapping of
 * torch.randn(numel, device='cuda') under (seed, offset): bit-identical to it on the
 * same GPU model.  *offset_advance (host, may be NULL) receives how far the generator
 * offset moves, as torch's CUDA generator would. */
B200DS_API int b200ds_philox_normal(float* x_dev, int64_t numel, uint64_t seed, uint64_t offset,
                         uint64_t* offset_advance, uintptr_t stream);
/* Same stream, CUDA-graph friendly: the k-th replay (k = *draw_counter_dev, a device uint64
 * the caller's step increments, e.g. the step counter of b200ds_bce_loss_grad) produces what
 * the (k+1)-th consecutive torch.randn(numel) call after (seed, offset0) would. */
B200DS_API int b200ds_philox_normal_counted(float* x_dev, int64_t numel, uint64_t seed, uint64_t offset0,
                                            const uint64_t* draw_counter_dev, uintptr_t stream);
/* y[b] = 1 if sum_{j < F/2} x[b,j] > 0 else 0   (train.py:22-23) */
B200DS_API int b200ds_label_rule(const float* x_dev, float* y_dev, int64_t batch, int n_features,
                      uintptr_t stream);
/* Batch gather from the device-resident dataset: replaces TensorDataset indexing +
 * default_collate + .to(device) (train.py:65,74,104-106).  idx_dev: int64 [batch]. */
B200DS_API int b200ds_gather_batch(const float* x_all_dev, const float* y_all_dev, const int64_t* idx_dev,
                        int64_t batch, int n_features, float* x_out_dev, float* y_out_dev,
                        uintptr_t stream);

/* ---- fp32 step, one kernel per engine call ---------------------------------------- */
/* model_engine(batch_x) (train.py:107 -> SimpleNet.forward, train.py:35-36):
 * logits[b] = relu(x[b,:] W1^T + b1) . W2 + b2.  x: [batch,F] fp32, logits: [batch]. */
B200DS_API int b200ds_mlp_forward_f32(const float* params_dev, const float* x_dev, int64_t batch,
                           int n_features, int hidden, float* logits_dev, uintptr_t stream);

/* criterion(logits, y) (train.py:96,112) fused with the first node of
 * model_engine.backward (train.py:113): loss = mean_b[(1-y) z - logsigmoid(z)],
 * dlogits[b] = (sigmoid(z) - y)/batch.  loss_dev: 1 float.  dlogits_dev may be NULL.
 * step_counter_dev (may be NULL): uint64 incremented by one (Adam's t, read by
 * b200ds_adam_step_counted).  scratch: b200ds_bce_scratch_bytes(batch) bytes, device. */
B200DS_API size_t b200ds_bce_scratch_bytes(int64_t batch);
B200DS_API int b200ds_bce_loss_grad(const float* logits_dev, const float* y_dev, int64_t batch,
                         float* loss_dev, float* dlogits_dev, uint64_t* step_counter_dev,
                         void* scratch_dev, size_t scratch_bytes, uintptr_t stream);

/* The same pass without any scratch argument.  No counterpart in the reference: it exists for
 * hosts of this ABI that cannot allocate device memory themselves -- a launcher that only
 * holds the buffers its framework gave it, or a step loop under a caching allocator that
 * already owns the GPU, where a stray cudaMalloc may fail or stall the step.  The scratch is a
 * 1280-byte array built into the library (present on every GPU that loads it); nothing is
 * allocated or freed.  Limit: batch <= 4194304 rows, larger batches are rejected
 * (B200DS_ERR_ARG) -- use b200ds_bce_loss_grad for those.  The first call on a device looks the
 * array's address up and should be made outside stream capture.  All calls share the array:
 * calls whose GPU work may overlap must be ordered by the caller. */
B200DS_API int b200ds_bce_loss_grad_noalloc(const float* logits_dev, const float* y_dev, int64_t batch,
                                            float* loss_dev, float* dlogits_dev,
                                            uint64_t* step_counter_dev, uintptr_t stream);

/* model_engine.backward(loss) up to the all-reduce (train.py:113): recomputes the hidden
 * activation, writes dW1, db1, dW2, db2 into grads_dev (flat layout, overwritten).
 * No dx is formed (the input does not require grad).  F <= 256 on this path. */
B200DS_API size_t b200ds_mlp_backward_f32_scratch_bytes(int64_t batch, int n_features, int hidden);
B200DS_API int b200ds_mlp_backward_f32(const float* params_dev, const float* x_dev, const float* dlogits_dev,
                            int64_t batch, int n_features, int hidden, float* grads_dev,
                            void* scratch_dev, size_t scratch_bytes, uintptr_t stream);

/* ---- gradient exchange: the all-reduce inside model_engine.backward (train.py:113,
 * DeepSpeed stage 0, ds_config train.py:78-83).  SUM over ranks of bucket_dev[0..n) in
 * place with ncclAllReduce on the context's side stream: the side stream waits for the
 * work already enqueued on `stream`, and `stream` then waits for the all-reduce, so the
 * caller's ordering is preserved without a host sync.  The 1/world scale is NOT applied
 * here; pass grad_scale = 1/world to b200ds_adam_step.  world == 1: no-op. */
B200DS_API int b200ds_allreduce_bucket(b200ds_ctx* ctx, float* bucket_dev, int64_t n, uintptr_t stream);
/* The same exchange in two halves, so that work which does not need the gradients (the next
 * batch's generation, say) can be enqueued on `stream` between them and run beside the
 * all-reduce: _begin orders the side stream after `stream` and starts the all-reduce, _end
 * makes `stream` wait for it.  One exchange in flight per context. */
B200DS_API int b200ds_allreduce_bucket_begin(b200ds_ctx* ctx, float* bucket_dev, int64_t n, uintptr_t stream);
B200DS_API int b200ds_allreduce_bucket_end(b200ds_ctx* ctx, uintptr_t stream);
/* ZeRO-1-style optimizer step (BASELINE config 5; an extension -- the reference runs stage 0,
 * train.py:82): reduce-scatter of the bucket, Adam on this rank's slice [rank*n/W, (rank+1)*n/W)
 * with slice-sized exp_avg / exp_avg_sq, all-gather of the parameters; replaces
 * b200ds_allreduce_bucket + b200ds_adam_step.  n % (4*world) == 0.  `step` as in b200ds_adam_step. */
B200DS_API int b200ds_zero1_step(b200ds_ctx* ctx, float* params_dev, float* grads_dev, float* exp_avg_shard_dev,
                                 float* exp_avg_sq_shard_dev, int64_t n, double lr, double beta1, double beta2,
                                 double eps, int64_t step, uintptr_t stream);
/* One NCCL collective on a raw buffer, ordered like b200ds_allreduce_bucket (side stream, event-chained), for the
 * fabric sweep tool (BASELINE config 5 "allreduce bus GB/s sweep vs bucket size"; the reference's own purpose is
 * an acceptance check of NCCL over the fabric, README.md:3).  op: 0 all-reduce(sum) in place, 1 reduce-scatter (the
 * rank's shard of the sum lands at its own offset rank*n/W), 2 all-gather (from that offset into the whole buffer).
 * dtype: 0 fp32, 1 bf16.  n elements; n % world == 0 for ops 1 and 2. */
B200DS_API int b200ds_collective(b200ds_ctx* ctx, int op, int dtype, void* buf_dev, int64_t n, uintptr_t stream);
/* Parameter broadcast from rank 0 done by deepspeed.initialize (train.py:87-93). */
B200DS_API int b200ds_broadcast_params(b200ds_ctx* ctx, float* params_dev, int64_t n, uintptr_t stream);
/* dist.barrier() at teardown (train.py:134): host-blocking. */
B200DS_API int b200ds_barrier(b200ds_ctx* ctx);

/* ---- optimizer: model_engine.step() (train.py:114) = torch.optim.Adam.step with the
 * defaults of train.py:85 (no weight decay, no amsgrad), then gradients cleared.
 *   g = grad*grad_scale; m += (g-m)(1-b1); v = b2 v + (1-b2) g g;
 *   p -= (lr/(1-b1^t)) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
 * `step` is t AFTER the increment (1 for the first update).  n must cover the flat
 * buffers; buffers must be 16-byte aligned (128-bit accesses).  zero_grad != 0 clears
 * grads_dev in the same pass. */
B200DS_API int b200ds_adam_step(float* params_dev, float* grads_dev, float* exp_avg_dev, float* exp_avg_sq_dev,
                     int64_t n, double lr, double beta1, double beta2, double eps, int64_t step,
                     float grad_scale, int zero_grad, uintptr_t stream);
/* Same, but t is read from a device counter (graph-replayable; see b200ds_bce_loss_grad). */
B200DS_API int b200ds_adam_step_counted(float* params_dev, float* grads_dev, float* exp_avg_dev,
                             float* exp_avg_sq_dev, int64_t n, double lr, double beta1, double beta2,
                             double eps, const uint64_t* step_counter_dev, float grad_scale,
                             int zero_grad, uintptr_t stream);

/* ---- fused tiny-shape step (train.py:104-117 in one launch) ------------------------
 * For the reference's own shape class (F <= 32, H <= 256, per-step batch <= 32): one CTA
 * keeps the 1409 parameters and their Adam state in registers and runs `n_steps`
 * consecutive training steps: gather rows order_dev[s*batch .. ) of the resident
 * dataset, forward, loss, backward, and -- when apply_adam != 0 (world == 1) -- the Adam
 * update.  With apply_adam == 0, n_steps must be 1 and the gradients are written to
 * grads_dev for b200ds_allreduce_bucket + b200ds_adam_step.
 *   order_dev : int64 [n_order] sample indices of this rank for the epoch
 *   first     : index into order_dev of the first sample of the first step
 *   losses_dev: float [n_steps], the per-step loss (what loss.item() returns, train.py:115)
 *   step0     : Adam t before the first of these steps.
 * The ragged last batch (train.py:74, drop_last=False) is handled (mean over what is left). */
B200DS_API int b200ds_tiny_train_steps_f32(float* params_dev, float* grads_dev, float* exp_avg_dev,
                                float* exp_avg_sq_dev, int n_features, int hidden,
                                const float* x_all_dev, const float* y_all_dev,
                                const int64_t* order_dev, int64_t n_order, int64_t first,
                                int batch, int n_steps, int apply_adam, double lr, double beta1,
                                double beta2, double eps, int64_t step0, float* losses_dev,
                                uintptr_t stream);

/* ---- bf16 tensor-core path (tcgen05.mma / TMEM / TMA) for the widened model ----------
 * BASELINE config 4 (H = 4096..8192, 64K samples per GPU).  The reference has no bf16
 * path (train.py:81 disables fp16); semantics defined by the build: GEMM operands rounded
 * to bf16, fp32 accumulation, fp32 master weights / Adam state / gradient bucket
 * (SURVEY H5).  Shape class: F + 2 <= 32, H % 128 == 0, batch % 128 == 0
 * (b200ds_tc_supported); anything else belongs to the fp32 entry points above.
 *
 * Operand packs (bf16, tiled for 1-D TMA, see csrc/tc_common.cuh):
 *   xq : [batch, 32]  = x[b,0..F-1], 1, 1, 0...        b200ds_tc_xq_bytes(batch)
 *   wq : [H, 32]      = W1[n,0..F-1], hi(b1[n]), lo(b1[n]), 0...  followed by the per-slice
 *        vectors v = w1q^T w2 the forward uses           b200ds_tc_wq_bytes(H)
 * x is packed once per batch (after generation / H2D), the weights once per optimizer step. */
B200DS_API int b200ds_tc_supported(int64_t batch, int n_features, int hidden);
B200DS_API size_t b200ds_tc_xq_bytes(int64_t batch);
B200DS_API size_t b200ds_tc_wq_bytes(int hidden);
B200DS_API int b200ds_pack_x_bf16(const float* x_dev, int64_t batch, int n_features, void* xq_dev,
                                  uintptr_t stream);
/* Same with a batch padded up to the shape class: x holds `rows` real rows, the pack gets `batch`
 * rows (batch % 8 == 0), rows [rows, batch) with zero features.  The engine uses it to run the
 * reference's own tiny batches (train.py:74, 4 rows, ragged tail) through the bf16 kernels; the
 * padded rows must be given dlogits = 0 in the backward. */
B200DS_API int b200ds_pack_x_bf16_rows(const float* x_dev, int64_t rows, int64_t batch, int n_features,
                                       void* xq_dev, uintptr_t stream);
/* Generated-batch configs: draw the batch and write it straight into the pack, with its
 * labels -- xq[b,k] = bf16(x[b,k]) where x = torch.randn(batch*F, device='cuda') under
 * (seed, offset) viewed as [batch,F] (same stream and mapping as b200ds_philox_normal), and
 * y[b] by the rule of train.py:22-23 on the fp32 values.  Replaces b200ds_philox_normal +
 * b200ds_label_rule + b200ds_pack_x_bf16 by one launch; fp32 x is never materialised.
 * The _counted form takes the draw index from a device counter (CUDA-graph replay). */
B200DS_API int b200ds_gen_batch_bf16(void* xq_dev, float* y_dev, int64_t batch, int n_features,
                                     uint64_t seed, uint64_t offset, uint64_t* offset_advance,
                                     uintptr_t stream);
B200DS_API int b200ds_gen_batch_bf16_counted(void* xq_dev, float* y_dev, int64_t batch, int n_features,
                                             uint64_t seed, uint64_t offset0,
                                             const uint64_t* draw_counter_dev, uintptr_t stream);
B200DS_API int b200ds_pack_params_bf16(const float* params_dev, int n_features, int hidden, void* wq_dev,
                                       uintptr_t stream);
/* model_engine(batch_x) (train.py:107): logits[b] = relu(xq[b,:] . w1q[n,:]) . W2 + b2, the
 * hidden activation stays in TMEM/registers.  W2 and b2 are read in fp32 from params_dev.
 * H > 2048 is split into slices of hidden units handled by different CTAs; their parts of a logit
 * are summed in slice order through `scratch` (b200ds_mlp_forward_tc_scratch_bytes; 0 bytes and a
 * NULL pointer for H <= 2048; zeroed once by the caller, left zeroed by every call), so the
 * result is bit-reproducible from run to run for every H. */
B200DS_API size_t b200ds_mlp_forward_tc_scratch_bytes(int64_t batch, int n_features, int hidden);
B200DS_API int b200ds_mlp_forward_tc(const void* xq_dev, const void* wq_dev, const float* params_dev,
                                     int64_t batch, int n_features, int hidden, float* logits_dev,
                                     void* scratch_dev, size_t scratch_bytes, uintptr_t stream);
/* model_engine.backward(loss) up to the all-reduce (train.py:113): recomputes the
 * activation mask on the tensor cores, accumulates S[n,k] = sum_b 1[z1[b,n]>0] dz[b] xq[b,k]
 * in TMEM, then dW1 = W2[n] S[n,:F], db1 = W2[n] S[n,F], dW2[n] = sum_k w1q[n,k] S[n,k],
 * db2 = sum_b dz[b]; grads_dev (flat layout) is overwritten.  dlogits 16-byte aligned. */
B200DS_API size_t b200ds_mlp_backward_tc_scratch_bytes(int64_t batch, int n_features, int hidden);
B200DS_API int b200ds_mlp_backward_tc(const void* xq_dev, const void* wq_dev, const float* params_dev,
                                      const float* dlogits_dev, int64_t batch, int n_features, int hidden,
                                      float* grads_dev, void* scratch_dev, size_t scratch_bytes,
                                      uintptr_t stream);

/* criterion(logits, y) and model_engine.backward(loss) up to the all-reduce (train.py:112-113) as
 * ONE pass: four extra warps of the backward kernel form dlogits = (sigmoid(z) - y) / B from the
 * logits and labels (the arithmetic of b200ds_bce_loss_grad: the W1 / b1 / W2 gradients are
 * bit-identical to that call followed by b200ds_mlp_backward_tc; db2 and the loss differ by
 * summation order only) and sum the loss terms on the side; dlogits is never written.
 * loss_dev receives the batch mean; step_counter_dev (NULL allowed) is incremented as
 * b200ds_bce_loss_grad does.  Scratch: b200ds_mlp_backward_tc_scratch_bytes.  logits and labels
 * 16-byte aligned. */
B200DS_API int b200ds_mlp_loss_backward_tc(const void* xq_dev, const void* wq_dev, const float* params_dev,
                                           const float* logits_dev, const float* y_dev, int64_t batch,
                                           int n_features, int hidden, float* loss_dev, float* grads_dev,
                                           uint64_t* step_counter_dev, void* scratch_dev, size_t scratch_bytes,
                                           uintptr_t stream);

/* ---- gradient exchange + optimizer step over NVLink peer memory ---------------------
 * model_engine.backward's all-reduce and model_engine.step()'s Adam (train.py:113-114) as ONE
 * kernel: every rank publishes its gradient bucket in a buffer its peers have mapped (CUDA
 * IPC), waits on arrival flags, reads all W buckets through NVLink, sums them in rank order,
 * applies 1/W and Adam.  No collective-library call on the step; parameters stay bit-equal
 * across ranks (fixed summation order).  grads_dev ends up holding the SUM, as after
 * b200ds_allreduce_bucket.  b200ds_ctx_enable_p2p is collective (every rank of the context
 * calls it once, with the same max_elems), allocates the exchange buffer (8 * max_elems bytes)
 * and exchanges the handles through the context's communicator; at most B200DS_P2P_MAX_WORLD
 * ranks, all on one NVLink domain.  n: multiple of 4, <= max_elems.  step / step_counter_dev
 * as in b200ds_adam_step / b200ds_adam_step_counted.  The kernel spins until every rank has
 * launched it: every rank must call it the same number of times, in the same order. */
#define B200DS_P2P_MAX_WORLD 8
B200DS_API int b200ds_ctx_enable_p2p(b200ds_ctx* ctx, int64_t max_elems);
B200DS_API int b200ds_allreduce_adam_p2p(b200ds_ctx* ctx, float* params_dev, float* grads_dev,
                                         float* exp_avg_dev, float* exp_avg_sq_dev, int64_t n, double lr,
                                         double beta1, double beta2, double eps, int64_t step,
                                         const uint64_t* step_counter_dev, uintptr_t stream);

/* ---- everything between the backward kernel and the next forward as one launch -------
 * b200ds_mlp_loss_backward_partials_tc = b200ds_mlp_loss_backward_tc without its finish pass:
 * the split partials stay in `scratch` (step_counter_dev is advanced by the kernel).
 * b200ds_update_tc then does, in ONE kernel: the finish pass (gradients from the partials, loss
 * mean into loss_dev if non-NULL), the gradient exchange over peer memory for W > 1 (protocol of
 * b200ds_allreduce_adam_p2p; needs b200ds_ctx_enable_p2p(ctx, >= b200ds_update_tc_exchange_elems(H))),
 * Adam (t = step, or *step_counter_dev), and the bf16 operand pack of the updated W1 / b1 into
 * wq_dev (what b200ds_pack_params_bf16 would produce).  grads_dev receives the (summed) gradients.
 * For one rank the result is bit-identical to b200ds_mlp_loss_backward_tc + b200ds_adam_step +
 * b200ds_pack_params_bf16.  Every rank must make the same sequence of calls. */
B200DS_API int b200ds_mlp_loss_backward_partials_tc(const void* xq_dev, const void* wq_dev, const float* logits_dev,
                                                    const float* y_dev, int64_t batch, int n_features, int hidden,
                                                    uint64_t* step_counter_dev, void* scratch_dev,
                                                    size_t scratch_bytes, uintptr_t stream);
B200DS_API int64_t b200ds_update_tc_exchange_elems(int hidden);
B200DS_API int b200ds_update_tc(b200ds_ctx* ctx, const void* scratch_dev, size_t scratch_bytes, int64_t batch,
                                int n_features, int hidden, float* params_dev, float* grads_dev,
                                float* exp_avg_dev, float* exp_avg_sq_dev, void* wq_dev, double lr, double beta1,
                                double beta2, double eps, int64_t step, const uint64_t* step_counter_dev,
                                float* loss_dev, uintptr_t stream);
/* The same update in two halves, same arguments, possibly on different streams (the caller orders `_end` after
 * `_begin` with an event): `_begin` = finish pass + publication of this rank's gradients in every rank's
 * exchange buffer (for one rank: the whole update); `_end` = wait for every rank's publication (a one-warp
 * kernel, so the SMs stay free for work enqueued beside it, e.g. the next batch's generation), rank-order sum,
 * Adam, operand pack (for one rank: nothing).  b200ds_update_tc = both, back to back. */
B200DS_API int b200ds_update_tc_begin(b200ds_ctx* ctx, const void* scratch_dev, size_t scratch_bytes, int64_t batch,
                                int n_features, int hidden, float* params_dev, float* grads_dev,
                                float* exp_avg_dev, float* exp_avg_sq_dev, void* wq_dev, double lr, double beta1,
                                double beta2, double eps, int64_t step, const uint64_t* step_counter_dev,
                                float* loss_dev, uintptr_t stream);
B200DS_API int b200ds_update_tc_end(b200ds_ctx* ctx, const void* scratch_dev, size_t scratch_bytes, int64_t batch,
                                int n_features, int hidden, float* params_dev, float* grads_dev,
                                float* exp_avg_dev, float* exp_avg_sq_dev, void* wq_dev, double lr, double beta1,
                                double beta2, double eps, int64_t step, const uint64_t* step_counter_dev,
                                float* loss_dev, uintptr_t stream);

/* ---- wide inputs: the widened model with n_features in the hundreds or thousands ("F = H", BASELINE config 4) --
 * SimpleNet takes any n_features (train.py:26-33); from F = 256 on layer 1 is a real GEMM and runs as K-loop
 * tcgen05.mma kernels fed by tensor-map TMA (csrc/mlp_fh.cu).  Shape class: F % 256 == 0, H % 128 == 0,
 * batch % 256 == 0 (b200ds_fh_supported).  Semantics: GEMM operands rounded to bf16, fp32 accumulation, fp32
 * master weights; the hidden activation is kept as bf16 between forward and backward (no recompute).
 * Operands are bf16 arrays in "layout L" (8 x 8 core matrices, see mlp_fh.cu): b200ds_fh_operand_bytes(rows, cols)
 * bytes each;  x: [batch, F], W1: [H, F], hT: [H, batch] (written by the forward, consumed by the backward).
 * b200ds_fh_pack_bf16 converts a row-major fp32 matrix (rows % 8 == 0, cols % 64 == 0). */
B200DS_API int b200ds_fh_supported(int64_t batch, int n_features, int hidden);
/* the same class narrowed to what the BACKWARD takes as well (batch % 2048 == 0): b200ds_train_step_host(_prefetch)
 * routes a step to these kernels only then; any other wide shape takes the fp32 kernels. */
B200DS_API int b200ds_fh_train_supported(int64_t batch, int n_features, int hidden);
B200DS_API size_t b200ds_fh_operand_bytes(int64_t rows, int64_t cols);
B200DS_API int b200ds_fh_pack_bf16(const float* src_dev, int64_t rows, int64_t cols, void* dst_dev, uintptr_t stream);
/* model_engine(batch_x) (train.py:107) up to the per-tile parts of the logits: hT[n, b] = bf16(relu(x_q[b,:] . W1_q[n,:] + b1[n]))
 * and, in `scratch` (b200ds_fh_forward_scratch_bytes), zpart[n / 128][b] = sum over the 128 units of the tile of W2[n] hT[n, b];
 * b200ds_fh_logits adds the parts up in tile order (+ b2) -- deterministic, no atomics. */
B200DS_API size_t b200ds_fh_forward_scratch_bytes(int64_t batch, int n_features, int hidden);
B200DS_API int b200ds_fh_forward(const void* xq_dev, const void* w1q_dev, const float* params_dev, int64_t batch,
                                 int n_features, int hidden, void* hT_dev, void* scratch_dev, size_t scratch_bytes,
                                 uintptr_t stream);

/* make_synthetic_data (train.py:19-24) for generated batches of wide inputs: xq (layout L) = bf16 of
 * torch.randn(batch*F, device='cuda') under (seed, offset) viewed as [batch, F], y[b] = 1[sum_{k < F/2} x[b,k] > 0]
 * on the fp32 values (32-column partial sums by shuffle, added in column order: deterministic).  Needs F to divide
 * the generator's thread count (b200ds_fh_gen_supported; powers of two in [64, 8192] for large batches);
 * scratch: b200ds_fh_gen_scratch_bytes.  The _counted form takes the draw index from a device counter. */
B200DS_API int b200ds_fh_gen_supported(int64_t batch, int n_features);
B200DS_API size_t b200ds_fh_gen_scratch_bytes(int64_t batch, int n_features);
B200DS_API int b200ds_fh_gen_batch(void* xq_dev, float* y_dev, int64_t batch, int n_features, uint64_t seed,
                                   uint64_t offset, uint64_t* offset_advance, void* scratch_dev, size_t scratch_bytes,
                                   uintptr_t stream);
B200DS_API int b200ds_fh_gen_batch_counted(void* xq_dev, float* y_dev, int64_t batch, int n_features, uint64_t seed,
                                           uint64_t offset0, const uint64_t* draw_counter_dev, void* scratch_dev,
                                           size_t scratch_bytes, uintptr_t stream);
B200DS_API int b200ds_fh_logits(const void* fwd_scratch_dev, const float* params_dev, int64_t batch, int n_features,
                                int hidden, float* logits_dev, uintptr_t stream);
/* model_engine.backward(loss) up to the all-reduce (train.py:113) for wide inputs, from dlogits (b200ds_bce_loss_grad):
 * relu's backward IN PLACE on hT (it becomes the A operand 1[h>0] * bf16(dlogits) of the weight-gradient GEMM -- hT
 * is consumed), dW2 = sum_b dz h, db1 = W2 * sum_b 1[h>0] dz, db2 = sum_b dz, dW1[n,:] = W2[n] * sum_b A[n,b] x_q[b,:]
 * (tcgen05 GEMM over the batch, two halves of the batch per output tile meeting through a two-addend
 * red.global.add: deterministic).  grads_dev (flat layout) is overwritten.  batch % 2048 == 0. */
B200DS_API size_t b200ds_fh_backward_scratch_bytes(int64_t batch, int n_features, int hidden);
B200DS_API int b200ds_fh_backward(const void* xq_dev, void* hT_dev, const float* dlogits_dev, const float* params_dev,
                                  int64_t batch, int n_features, int hidden, float* grads_dev, void* scratch_dev,
                                  size_t scratch_bytes, uintptr_t stream);

/* ---- end-to-end step with HOST buffers (what bench.py's `e2e` times) ---------------
 * One call = train.py:104-115 for one batch: H2D of x,y from (pinned) host memory,
 * forward, loss, backward, all-reduce, Adam (t = step0 + 1), D2H of the loss, and a
 * stream synchronise (this is the one entry point that blocks the host, as loss.item()
 * does at train.py:115).  Shapes inside the tiny class (F <= 32, H <= 256, batch <= 32)
 * take the fused one-launch kernel; shapes of the tensor-core class take the bf16 tcgen05
 * kernels when `xq` and `wq` are supplied (non-NULL); everything else the per-stage fp32
 * kernels.  The caller supplies every device buffer in `bufs` (all borrowed):
 *   scratch: b200ds_train_step_scratch_bytes(...) bytes, zeroed once by the caller;
 *   xq / wq: b200ds_tc_xq_bytes(batch) / b200ds_tc_wq_bytes(H) bytes, or NULL for fp32; for shapes of the
 *            wide-input class (b200ds_fh_supported) b200ds_fh_operand_bytes(batch, F) / (H, F) bytes.
 * With more than one rank the exchange is ncclAllReduce + the Adam kernel, or the single
 * peer-memory kernel when b200ds_ctx_enable_p2p has been called with room for the bucket. */
typedef struct b200ds_step_bufs {
  float* params;  float* grads;  float* exp_avg;  float* exp_avg_sq;   /* flat, length padded to 4 */
  float* x_stage; float* y_stage;          /* [batch*F], [batch] device staging */
  float* logits;  float* dlogits;          /* [batch] */
  float* loss;                             /* 1 float */
  void* scratch;  size_t scratch_bytes;
  void* xq;  void* wq;                     /* bf16 operand packs, or NULL */
} b200ds_step_bufs;
B200DS_API size_t b200ds_train_step_scratch_bytes(int64_t batch, int n_features, int hidden);
B200DS_API int b200ds_train_step_host(b200ds_ctx* ctx, const b200ds_step_bufs* bufs,
                           const float* x_host, const float* y_host, int64_t batch,
                           int n_features, int hidden, double lr, double beta1, double beta2,
                           double eps, int64_t step0, float* loss_host, uintptr_t stream);
/* The same step with the NEXT batch's host->device copy queued on the context's own copy
 * stream right behind the launch of this batch's kernels -- what DataLoader workers + pin_memory
 * give the reference loop (train.py:36-44,104): the copy of batch k+1 hides behind the kernels of
 * batch k.  x_stage / y_stage must hold TWO batches (2*batch*F and 2*batch floats).  The call whose
 * (x_host, y_host) are the pointers given as (x_next_host, y_next_host) by the previous call
 * uses the staged copy and copies nothing; any other pointers are copied as in
 * b200ds_train_step_host.  A prefetched host buffer must stay unchanged until the call that
 * consumes it returns.  NULL next pointers = no prefetch.  The announced batch is copied with THIS
 * call's batch and n_features, so it must have the same shape (do not announce a ragged last
 * batch: pass NULL); a consuming call whose shape differs from the staged one ignores the staged
 * copy and copies its own inputs.  Result per call: identical to
 * b200ds_train_step_host on the same batch.  The device part of the step is captured once per
 * (buffers, staging half, shape, hyper-parameters) as a CUDA graph on a stream of the context and
 * replayed on the caller's stream (one launch per step instead of ~8); the graphs are destroyed
 * by b200ds_ctx_destroy.  The Adam step number lives in a device word of the context that the step's
 * loss kernel increments; the host rewrites it (after a stream synchronise) only when step0 is not the
 * previous call's step0 + 1.  With bf16 packs, a supported shape and one rank the gradients take the
 * b200ds_update_tc route (finish + Adam + pack in one kernel); with several ranks finish, then
 * b200ds_allreduce_adam_p2p when b200ds_ctx_enable_p2p succeeded, else ncclAllReduce and the Adam
 * kernel.  bufs->wq is rewritten by the call either way, the caller never packs. */
B200DS_API int b200ds_train_step_host_prefetch(b200ds_ctx* ctx, const b200ds_step_bufs* bufs,
                           const float* x_host, const float* y_host,
                           const float* x_next_host, const float* y_next_host, int64_t batch,
                           int n_features, int hidden, double lr, double beta1, double beta2,
                           double eps, int64_t step0, float* loss_host, uintptr_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DS_H_ */
