// End-to-end training step over HOST buffers: the whole of train.py:104-115 for one batch
// behind one C call (H2D copies, forward, loss, backward, all-reduce, Adam, loss D2H,
// sync).  This is the call bench.py times for its `e2e` figure.
#include "ctx.cuh"
#include <string.h>

using namespace b200ds;

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static bool tiny_class(int64_t batch, int F, int H) { return F <= 32 && H <= 256 && batch <= 32; }

extern "C" {

size_t b200ds_train_step_scratch_bytes(int64_t batch, int n_features, int hidden) {
  size_t f32 = b200ds_mlp_backward_f32_scratch_bytes(batch, n_features, hidden);
  size_t tcb = b200ds_mlp_backward_tc_scratch_bytes(batch, n_features, hidden);
  if (b200ds_fh_train_supported(batch, n_features, hidden))      // wide inputs: [ loss | forward parts | backward parts | hT ]
    return align_up(b200ds_bce_scratch_bytes(batch), 256) + align_up(b200ds_fh_forward_scratch_bytes(batch, n_features, hidden), 256) +
           align_up(b200ds_fh_backward_scratch_bytes(batch, n_features, hidden), 256) + align_up(b200ds_fh_operand_bytes(hidden, batch), 256);
  // [ loss scratch | backward scratch (fp32 or tensor-core) | forward slice parts + tile tickets (tensor-core, H > 2048) ]
  return align_up(b200ds_bce_scratch_bytes(batch), 256) + align_up(f32 > tcb ? f32 : tcb, 256) +
         align_up(b200ds_mlp_forward_tc_scratch_bytes(batch, n_features, hidden), 256);
}

// the device part of the step on staged inputs (x_dev, y_dev), loss left in b->loss
static int run_staged_step(b200ds_ctx* ctx, const b200ds_step_bufs* b, const float* x_dev, const float* y_dev,
                           int64_t batch, int n_features, int hidden, double lr, double beta1, double beta2,
                           double eps, int64_t step0, uint64_t* step_counter /* graph replay: device word holding step0 when the step starts;
                           the loss kernel of the step makes it step0 + 1, the optimizer reads it.  NULL: eager, step0 + 1 by value */,
                           uintptr_t stream) {
  const Layout L{n_features, hidden};
  const int64_t P = L.count();
  const int64_t Ppad = (P + 3) / 4 * 4;
  int rc;
  const float scale = ctx->world > 1 ? (float)(1.0 / ctx->world) : 1.f;
  if (tiny_class(batch, n_features, hidden)) {
    const int fused_adam = ctx->world == 1;
    rc = b200ds_tiny_train_steps_f32(b->params, b->grads, b->exp_avg, b->exp_avg_sq, n_features, hidden,
                                     x_dev, y_dev, ctx->iota, batch, 0, (int)batch, 1, fused_adam,
                                     lr, beta1, beta2, eps, step0, b->loss, stream);
    if (rc) return rc;
    if (!fused_adam) {
      if ((rc = b200ds_allreduce_bucket(ctx, b->grads, P, stream))) return rc;
      if ((rc = b200ds_adam_step(b->params, b->grads, b->exp_avg, b->exp_avg_sq, Ppad, lr, beta1, beta2, eps,
                                 step0 + 1, scale, 0, stream))) return rc;
    }
    return B200DS_OK;
  }
  B200_REQUIRE(b->logits && b->dlogits, "train_step_host: logits/dlogits buffers required");
  size_t need = b200ds_train_step_scratch_bytes(batch, n_features, hidden);
  B200_REQUIRE(b->scratch && b->scratch_bytes >= need, "train_step_host: scratch too small (%zu < %zu)", b->scratch_bytes, need);
  char* sc = reinterpret_cast<char*>(b->scratch);
  size_t bce_bytes = align_up(b200ds_bce_scratch_bytes(batch), 256);
  if (b->xq && b->wq && b200ds_fh_train_supported(batch, n_features, hidden)) {
    // wide inputs (F = H): pack, layer-1 GEMM, logits, loss, backward (relu, dW2/db1/db2, weight-gradient GEMM)
    const size_t fwd_b = align_up(b200ds_fh_forward_scratch_bytes(batch, n_features, hidden), 256);
    const size_t bwd_b = align_up(b200ds_fh_backward_scratch_bytes(batch, n_features, hidden), 256);
    char* fwd_sc = sc + bce_bytes; char* bwd_sc = fwd_sc + fwd_b; char* hT = bwd_sc + bwd_b;
    if ((rc = b200ds_fh_pack_bf16(x_dev, batch, n_features, b->xq, stream))) return rc;
    if ((rc = b200ds_fh_pack_bf16(b->params, hidden, n_features, b->wq, stream))) return rc;
    if ((rc = b200ds_fh_forward(b->xq, b->wq, b->params, batch, n_features, hidden, hT, fwd_sc, fwd_b, stream))) return rc;
    if ((rc = b200ds_fh_logits(fwd_sc, b->params, batch, n_features, hidden, b->logits, stream))) return rc;
    if ((rc = b200ds_bce_loss_grad(b->logits, y_dev, batch, b->loss, b->dlogits, step_counter, sc, bce_bytes, stream))) return rc;
    if ((rc = b200ds_fh_backward(b->xq, hT, b->dlogits, b->params, batch, n_features, hidden, b->grads, bwd_sc, bwd_b, stream))) return rc;
    if ((rc = b200ds_allreduce_bucket(ctx, b->grads, P, stream))) return rc;
    if (step_counter)
      return b200ds_adam_step_counted(b->params, b->grads, b->exp_avg, b->exp_avg_sq, Ppad, lr, beta1, beta2, eps, step_counter, scale, 0, stream);
    return b200ds_adam_step(b->params, b->grads, b->exp_avg, b->exp_avg_sq, Ppad, lr, beta1, beta2, eps, step0 + 1, scale, 0, stream);
  }
  const bool use_tc = b->xq && b->wq && b200ds_tc_supported(batch, n_features, hidden);
  if (use_tc) {
    if ((rc = b200ds_pack_x_bf16(x_dev, batch, n_features, b->xq, stream))) return rc;
    if ((rc = b200ds_pack_params_bf16(b->params, n_features, hidden, b->wq, stream))) return rc;
    const size_t fwd_bytes = b200ds_mlp_forward_tc_scratch_bytes(batch, n_features, hidden);
    if ((rc = b200ds_mlp_forward_tc(b->xq, b->wq, b->params, batch, n_features, hidden, b->logits,
                                    fwd_bytes ? sc + need - align_up(fwd_bytes, 256) : nullptr, fwd_bytes, stream))) return rc;
  } else {
    if ((rc = b200ds_mlp_forward_f32(b->params, x_dev, batch, n_features, hidden, b->logits, stream))) return rc;
  }
  if (use_tc) {       // criterion + backward in one pass (dlogits is not materialised on this path)
    char* bwd_sc = sc + bce_bytes;
    const size_t bwd_bytes = need - bce_bytes - align_up(b200ds_mlp_forward_tc_scratch_bytes(batch, n_features, hidden), 256);
    if (ctx->world == 1) {
      // the split partials stay in scratch; finish pass + Adam + the operand pack are one kernel instead of three.  (With
      // several ranks the same kernel family does the exchange over peer memory in three launches; in this single-stream
      // step, with nothing to run beside them, that measured slower than finish + the one-launch all-reduce + Adam kernel
      // below: 144.7 against 135.5 us per call at 2 GPUs.)
      if ((rc = b200ds_mlp_loss_backward_partials_tc(b->xq, b->wq, b->logits, y_dev, batch, n_features, hidden, step_counter,
                                                     bwd_sc, bwd_bytes, stream))) return rc;
      return b200ds_update_tc(ctx, bwd_sc, bwd_bytes, batch, n_features, hidden, b->params, b->grads, b->exp_avg, b->exp_avg_sq, b->wq,
                              lr, beta1, beta2, eps, step0 + 1, step_counter, b->loss, stream);
    }
    if ((rc = b200ds_mlp_loss_backward_tc(b->xq, b->wq, b->params, b->logits, y_dev, batch, n_features, hidden, b->loss,
                                          b->grads, step_counter, bwd_sc, bwd_bytes, stream))) return rc;
  } else {
    if ((rc = b200ds_bce_loss_grad(b->logits, y_dev, batch, b->loss, b->dlogits, step_counter, sc, bce_bytes, stream))) return rc;
    if ((rc = b200ds_mlp_backward_f32(b->params, x_dev, b->dlogits, batch, n_features, hidden, b->grads,
                                      sc + bce_bytes, need - bce_bytes - align_up(b200ds_mlp_forward_tc_scratch_bytes(batch, n_features, hidden), 256), stream))) return rc;
  }
  if (ctx->world > 1 && ctx->p2p_cap >= Ppad)      // peer-memory exchange enabled on this context: one kernel for both
    return b200ds_allreduce_adam_p2p(ctx, b->params, b->grads, b->exp_avg, b->exp_avg_sq, Ppad, lr, beta1, beta2, eps,
                                     step0 + 1, step_counter, stream);
  if ((rc = b200ds_allreduce_bucket(ctx, b->grads, P, stream))) return rc;
  if (step_counter)
    return b200ds_adam_step_counted(b->params, b->grads, b->exp_avg, b->exp_avg_sq, Ppad, lr, beta1, beta2, eps,
                                    step_counter, scale, 0, stream);
  return b200ds_adam_step(b->params, b->grads, b->exp_avg, b->exp_avg_sq, Ppad, lr, beta1, beta2, eps,
                          step0 + 1, scale, 0, stream);
}

// The same device work as a CUDA graph: captured once per (buffers, staging half, shape, hyper-parameters) on
// the context's capture stream, replayed on the caller's stream.  The eager call costs ~8 kernel launches of
// host time per step; the replay one.  The Adam step number is the only thing that changes between replays: it lives
// in a device word that the step's loss kernel increments; the host writes it only when a call's step0 is not the
// previous call's step0 + 1 (first call, restart from a checkpoint) -- no copy in front of the graph otherwise.
static int run_staged_step_graph(b200ds_ctx* ctx, const b200ds_step_bufs* b, int slot, const float* x_dev, const float* y_dev,
                                 int64_t batch, int n_features, int hidden, double lr, double beta1, double beta2,
                                 double eps, int64_t step0, cudaStream_t s) {
  if (!ctx->cap) {
    B200_CUDA(cudaStreamCreateWithFlags(&ctx->cap, cudaStreamNonBlocking));
    B200_CUDA(cudaMalloc(&ctx->step_word, sizeof(unsigned long long)));
    B200_CUDA(cudaHostAlloc(&ctx->step_pinned, sizeof(unsigned long long), cudaHostAllocDefault));
  }
  unsigned char key[192];
  size_t n = 0;
  memset(key, 0, sizeof(key));
  auto put = [&](const void* p, size_t len) { memcpy(key + n, p, len); n += len; };
  static_assert(sizeof(b200ds_step_bufs) + 2 * sizeof(void*) + sizeof(int64_t) + 2 * sizeof(int) + 4 * sizeof(double) <= sizeof(key), "key");
  put(b, sizeof(*b)); put(&x_dev, sizeof(x_dev)); put(&y_dev, sizeof(y_dev)); put(&batch, sizeof(batch));
  put(&n_features, sizeof(int)); put(&hidden, sizeof(int)); put(&lr, sizeof(double)); put(&beta1, sizeof(double));
  put(&beta2, sizeof(double)); put(&eps, sizeof(double));
  b200ds_ctx::StepGraph& g = ctx->sgraph[slot & 1];
  if (!g.exec || g.key_len != n || memcmp(g.key, key, n) != 0) {
    if (g.exec) { B200_CUDA(cudaGraphExecDestroy(g.exec)); g.exec = nullptr; }
    cudaGraph_t graph = nullptr;
    B200_CUDA(cudaStreamBeginCapture(ctx->cap, cudaStreamCaptureModeThreadLocal));
    int rc = run_staged_step(ctx, b, x_dev, y_dev, batch, n_features, hidden, lr, beta1, beta2, eps, step0, reinterpret_cast<uint64_t*>(ctx->step_word),
                             (uintptr_t)ctx->cap);
    cudaError_t e = cudaStreamEndCapture(ctx->cap, &graph);
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    B200_CUDA(e);
    e = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    B200_CUDA(e);
    memcpy(g.key, key, n);
    g.key_len = n;
  }
  if (!ctx->step_word_known || ctx->step_word_value != step0) {
    // step_pinned may still be the source of an earlier copy that has not run: one word per write would be needed to
    // make this asynchronous; out-of-sequence steps are rare, so wait for the stream instead
    B200_CUDA(cudaStreamSynchronize(s));
    *ctx->step_pinned = (unsigned long long)step0;
    B200_CUDA(cudaMemcpyAsync(ctx->step_word, ctx->step_pinned, sizeof(unsigned long long), cudaMemcpyHostToDevice, s));
  }
  cudaError_t le = cudaGraphLaunch(g.exec, s);
  ctx->step_word_known = le == cudaSuccess;
  ctx->step_word_value = step0 + 1;        // what the word holds once this replay's loss kernel has run
  B200_CUDA(le);
  return B200DS_OK;
}

int b200ds_train_step_host(b200ds_ctx* ctx, const b200ds_step_bufs* b,
                           const float* x_host, const float* y_host, int64_t batch,
                           int n_features, int hidden, double lr, double beta1, double beta2,
                           double eps, int64_t step0, float* loss_host, uintptr_t stream) {
  B200_REQUIRE(ctx && b && x_host && y_host && loss_host, "train_step_host: null pointer");
  B200_REQUIRE(batch > 0 && n_features >= 1 && hidden >= 1, "train_step_host: bad shape");
  B200_REQUIRE(b->params && b->grads && b->exp_avg && b->exp_avg_sq && b->x_stage && b->y_stage && b->loss,
               "train_step_host: missing device buffer");
  cudaStream_t s = as_stream(stream);
  if (ctx->copy) B200_CUDA(cudaStreamSynchronize(ctx->copy));   // a prefetch nobody consumed may still be landing
  ctx->pf_x = ctx->pf_y = nullptr;          // this call overwrites the first half of the staging buffers
  B200_CUDA(cudaMemcpyAsync(b->x_stage, x_host, sizeof(float) * batch * n_features, cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaMemcpyAsync(b->y_stage, y_host, sizeof(float) * batch, cudaMemcpyHostToDevice, s));
  int rc = run_staged_step(ctx, b, b->x_stage, b->y_stage, batch, n_features, hidden, lr, beta1, beta2, eps, step0, nullptr, stream);
  if (rc) return rc;
  B200_CUDA(cudaMemcpyAsync(loss_host, b->loss, sizeof(float), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  return B200DS_OK;
}

int b200ds_train_step_host_prefetch(b200ds_ctx* ctx, const b200ds_step_bufs* b,
                                    const float* x_host, const float* y_host,
                                    const float* x_next_host, const float* y_next_host, int64_t batch,
                                    int n_features, int hidden, double lr, double beta1, double beta2,
                                    double eps, int64_t step0, float* loss_host, uintptr_t stream) {
  B200_REQUIRE(ctx && b && x_host && y_host && loss_host, "train_step_host_prefetch: null pointer");
  B200_REQUIRE((x_next_host == nullptr) == (y_next_host == nullptr), "train_step_host_prefetch: give both next pointers or neither");
  B200_REQUIRE(batch > 0 && n_features >= 1 && hidden >= 1, "train_step_host_prefetch: bad shape");
  B200_REQUIRE(b->params && b->grads && b->exp_avg && b->exp_avg_sq && b->x_stage && b->y_stage && b->loss,
               "train_step_host_prefetch: missing device buffer");
  cudaStream_t s = as_stream(stream);
  if (!ctx->copy) {                          // first use: the copy stream and its event
    B200_CUDA(cudaStreamCreateWithFlags(&ctx->copy, cudaStreamNonBlocking));
    B200_CUDA(cudaEventCreateWithFlags(&ctx->ev_copied, cudaEventDisableTiming));
  }
  const size_t xb = sizeof(float) * (size_t)batch * n_features, yb = sizeof(float) * (size_t)batch;
  int slot;
  if (ctx->pf_x == x_host && ctx->pf_y == y_host && ctx->pf_batch == batch && ctx->pf_F == n_features) {
    slot = ctx->pf_slot;                     // staged by the previous call: wait for that copy, not for a new one
    B200_CUDA(cudaStreamWaitEvent(s, ctx->ev_copied, 0));
  } else {
    slot = 0;
    B200_CUDA(cudaStreamSynchronize(ctx->copy));      // an unused prefetch may still be writing the staging buffers
    B200_CUDA(cudaMemcpyAsync(b->x_stage, x_host, xb, cudaMemcpyHostToDevice, s));
    B200_CUDA(cudaMemcpyAsync(b->y_stage, y_host, yb, cudaMemcpyHostToDevice, s));
  }
  float* xs = b->x_stage + (size_t)slot * batch * n_features;
  float* ys = b->y_stage + (size_t)slot * batch;
  ctx->pf_x = ctx->pf_y = nullptr;
  // the staged kernels as one graph replay (the one-launch tiny class stays eager)
  int rc = tiny_class(batch, n_features, hidden)
               ? run_staged_step(ctx, b, xs, ys, batch, n_features, hidden, lr, beta1, beta2, eps, step0, nullptr, stream)
               : run_staged_step_graph(ctx, b, slot, xs, ys, batch, n_features, hidden, lr, beta1, beta2, eps, step0, s);
  if (rc) return rc;
  B200_CUDA(cudaMemcpyAsync(loss_host, b->loss, sizeof(float), cudaMemcpyDeviceToHost, s));
  // The next batch's copy is queued AFTER this step's kernels: the caller waits for the loss below, so every host call
  // in front of the graph launch is GPU idle time; behind it, it costs nothing.
  if (x_next_host) {                         // the other half is idle: the previous call's work on it has been synchronised
    float* xn = b->x_stage + (size_t)(slot ^ 1) * batch * n_features;
    float* yn = b->y_stage + (size_t)(slot ^ 1) * batch;
    B200_CUDA(cudaMemcpyAsync(xn, x_next_host, xb, cudaMemcpyHostToDevice, ctx->copy));
    B200_CUDA(cudaMemcpyAsync(yn, y_next_host, yb, cudaMemcpyHostToDevice, ctx->copy));
    B200_CUDA(cudaEventRecord(ctx->ev_copied, ctx->copy));
    ctx->pf_x = x_next_host; ctx->pf_y = y_next_host; ctx->pf_slot = slot ^ 1;
    ctx->pf_batch = batch; ctx->pf_F = n_features;      // the staged copy has THIS call's shape; a call with another shape re-copies
  }
  B200_CUDA(cudaStreamSynchronize(s));
  return B200DS_OK;
}

}  // extern "C"
