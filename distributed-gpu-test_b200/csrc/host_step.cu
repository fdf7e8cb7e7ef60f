// End-to-end training step over HOST buffers: the whole of train.py:104-115 for one batch
// behind one C call (H2D copies, forward, loss, backward, all-reduce, Adam, loss D2H,
// sync).  This is the call bench.py times for its `e2e` figure.
#include "ctx.cuh"

using namespace b200ds;

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static bool tiny_class(int64_t batch, int F, int H) { return F <= 32 && H <= 256 && batch <= 32; }

extern "C" {

size_t b200ds_train_step_scratch_bytes(int64_t batch, int n_features, int hidden) {
  size_t f32 = b200ds_mlp_backward_f32_scratch_bytes(batch, n_features, hidden);
  size_t tcb = b200ds_mlp_backward_tc_scratch_bytes(batch, n_features, hidden);
  return align_up(b200ds_bce_scratch_bytes(batch), 256) + align_up(f32 > tcb ? f32 : tcb, 256);
}

// the device part of the step on staged inputs (x_dev, y_dev), loss left in b->loss
static int run_staged_step(b200ds_ctx* ctx, const b200ds_step_bufs* b, const float* x_dev, const float* y_dev,
                           int64_t batch, int n_features, int hidden, double lr, double beta1, double beta2,
                           double eps, int64_t step0, uintptr_t stream) {
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
  const bool use_tc = b->xq && b->wq && b200ds_tc_supported(batch, n_features, hidden);
  if (use_tc) {
    if ((rc = b200ds_pack_x_bf16(x_dev, batch, n_features, b->xq, stream))) return rc;
    if ((rc = b200ds_pack_params_bf16(b->params, n_features, hidden, b->wq, stream))) return rc;
    if ((rc = b200ds_mlp_forward_tc(b->xq, b->wq, b->params, batch, n_features, hidden, b->logits, stream))) return rc;
  } else {
    if ((rc = b200ds_mlp_forward_f32(b->params, x_dev, batch, n_features, hidden, b->logits, stream))) return rc;
  }
  if (use_tc) {       // criterion + backward in one pass (dlogits is not materialised on this path)
    if ((rc = b200ds_mlp_loss_backward_tc(b->xq, b->wq, b->params, b->logits, y_dev, batch, n_features, hidden, b->loss,
                                          b->grads, nullptr, sc + bce_bytes, b->scratch_bytes - bce_bytes, stream))) return rc;
  } else {
    if ((rc = b200ds_bce_loss_grad(b->logits, y_dev, batch, b->loss, b->dlogits, nullptr, sc, bce_bytes, stream))) return rc;
    if ((rc = b200ds_mlp_backward_f32(b->params, x_dev, b->dlogits, batch, n_features, hidden, b->grads,
                                      sc + bce_bytes, b->scratch_bytes - bce_bytes, stream))) return rc;
  }
  if (ctx->world > 1 && ctx->p2p_cap >= Ppad)      // peer-memory exchange enabled on this context: one kernel for both
    return b200ds_allreduce_adam_p2p(ctx, b->params, b->grads, b->exp_avg, b->exp_avg_sq, Ppad, lr, beta1, beta2, eps,
                                     step0 + 1, nullptr, stream);
  if ((rc = b200ds_allreduce_bucket(ctx, b->grads, P, stream))) return rc;
  return b200ds_adam_step(b->params, b->grads, b->exp_avg, b->exp_avg_sq, Ppad, lr, beta1, beta2, eps,
                          step0 + 1, scale, 0, stream);
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
  int rc = run_staged_step(ctx, b, b->x_stage, b->y_stage, batch, n_features, hidden, lr, beta1, beta2, eps, step0, stream);
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
  if (ctx->pf_x == x_host && ctx->pf_y == y_host) {
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
  if (x_next_host) {                         // the other half is idle: the previous call's work on it has been synchronised
    float* xn = b->x_stage + (size_t)(slot ^ 1) * batch * n_features;
    float* yn = b->y_stage + (size_t)(slot ^ 1) * batch;
    B200_CUDA(cudaMemcpyAsync(xn, x_next_host, xb, cudaMemcpyHostToDevice, ctx->copy));
    B200_CUDA(cudaMemcpyAsync(yn, y_next_host, yb, cudaMemcpyHostToDevice, ctx->copy));
    B200_CUDA(cudaEventRecord(ctx->ev_copied, ctx->copy));
    ctx->pf_x = x_next_host; ctx->pf_y = y_next_host; ctx->pf_slot = slot ^ 1;
  }
  int rc = run_staged_step(ctx, b, xs, ys, batch, n_features, hidden, lr, beta1, beta2, eps, step0, stream);
  if (rc) return rc;
  B200_CUDA(cudaMemcpyAsync(loss_host, b->loss, sizeof(float), cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  return B200DS_OK;
}

}  // extern "C"
