// Gradient exchange + optimizer step as ONE kernel over NVLink peer memory: what
// model_engine.backward's all-reduce (train.py:113) and model_engine.step()'s Adam
// (train.py:114) do in the reference, without a collective library call on the step.
//
// Every rank owns an exchange buffer of 2 x cap floats and a row of arrival flags, both
// cudaMalloc'ed by b200ds_ctx_enable_p2p and mapped into every peer with CUDA IPC (the handles
// travel through one ncclAllGather at set-up).  Per step, epoch e (a device word, so the launch
// replays from a CUDA graph):
//   1. each CTA copies its slice of the local gradients into half (e & 1) of the local buffer;
//      the last CTA to finish fences system-wide and PUSHES e into flags[rank] of every peer;
//   2. each CTA waits until its local flags[0..W-1] all hold e (local polling only);
//   3. each CTA reads its slice from every rank's buffer -- peers through NVLink loads -- sums in
//      rank order (identical on every rank: parameters stay bit-equal across ranks, which the
//      ring order of a library all-reduce does not promise), applies 1/W and Adam, and writes
//      the summed gradient back for callers that look at it.
// Two halves are enough: a rank rewrites half (e & 1) at epoch e + 2, after it has passed the
// flag wait of epoch e + 1, which every peer only signals after its epoch-e kernel -- and with it
// all reads of that half -- has completed.
#include "adam.cuh"
#include "ctx.cuh"
#include "p2p.cuh"
#include <string.h>
#include <stdlib.h>

namespace b200ds {

constexpr int kP2PThreads = 256;

template <bool kCounted>
__global__ void __launch_bounds__(kP2PThreads)
p2p_allreduce_adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                          int64_t n /* multiple of 4 */, AdamScalars s, double lr, double beta1, double beta2, int64_t step,
                          const unsigned long long* __restrict__ step_counter, P2PDev d) {
  __shared__ float sh[2];
  __shared__ bool last;
  const unsigned long long e = *d.epoch + 1;             // written back by the last CTA to leave
  const int64_t half = (int64_t)(e & 1) * d.cap;
  const int64_t nvec = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  // ---- 1. publish the local gradients
  {
    float4* dst = reinterpret_cast<float4*>(d.xchg[d.rank] + half);
    const float4* src = reinterpret_cast<const float4*>(g);
    for (int64_t i = t0; i < nvec; i += stride) dst[i] = src[i];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();                                     // device scope here; the last CTA's system fence below is cumulative
    last = atomicAdd(&d.tickets[0], 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last && threadIdx.x < d.world) {                   // all of this rank's slices are in place
    __threadfence_system();
    st_release_sys(&d.flags[threadIdx.x][d.rank], e);
    if (threadIdx.x == 0) d.tickets[0] = 0u;
  }
  if (threadIdx.x == 32) {                               // the optimizer's scalars, after the publication: two double-precision
    double t = kCounted ? (double)(*step_counter) : (double)step;      // pow() hide behind the NVLink round trip instead of delaying it
    double bc1 = 1.0 - pow(beta1, t);
    double bc2 = 1.0 - pow(beta2, t);
    sh[0] = (float)(-(lr / bc1));
    sh[1] = (float)sqrt(bc2);
  }
  // ---- 2. wait for every rank's publication (bounded in wall time, see p2p_wait_flag)
  if (threadIdx.x < d.world) p2p_wait_flag(&d.flags[d.rank][threadIdx.x], e, (int)threadIdx.x, d);
  __syncthreads();
  // ---- 3. sum in rank order, Adam
  const float neg_step_size = sh[0], bc2_sqrt = sh[1];
  float4* p4 = reinterpret_cast<float4*>(p);
  float4* g4 = reinterpret_cast<float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  for (int64_t i = t0; i < nvec; i += stride) {
    float4 pv = p4[i], mv = m4[i], vv = v4[i];
    float4 acc[B200DS_P2P_MAX_WORLD];
#pragma unroll
    for (int w = 0; w < B200DS_P2P_MAX_WORLD; ++w)
      if (w < d.world) acc[w] = ld_peer(reinterpret_cast<const float4*>(d.xchg[w] + half) + i);
    float4 gs = acc[0];
#pragma unroll
    for (int w = 1; w < B200DS_P2P_MAX_WORLD; ++w)
      if (w < d.world) { gs.x += acc[w].x; gs.y += acc[w].y; gs.z += acc[w].z; gs.w += acc[w].w; }
    g4[i] = gs;
    adam_one(pv.x, gs.x, mv.x, vv.x, s, neg_step_size, bc2_sqrt);
    adam_one(pv.y, gs.y, mv.y, vv.y, s, neg_step_size, bc2_sqrt);
    adam_one(pv.z, gs.z, mv.z, vv.z, s, neg_step_size, bc2_sqrt);
    adam_one(pv.w, gs.w, mv.w, vv.w, s, neg_step_size, bc2_sqrt);
    p4[i] = pv; m4[i] = mv; v4[i] = vv;
  }
  // ---- 4. the last CTA to leave closes the epoch
  __syncthreads();
  if (threadIdx.x == 0) {
    if (atomicAdd(&d.tickets[1], 1u) == gridDim.x - 1) {
      d.tickets[1] = 0u;
      *d.epoch = e;
      __threadfence();
    }
  }
}

}  // namespace b200ds

using namespace b200ds;

struct P2PHandles {
  cudaIpcMemHandle_t xchg, flags;
};

extern "C" {

// Undo whatever part of the set-up exists: close the peers' mappings, free the local buffers.
static void p2p_teardown(b200ds_ctx* ctx) {
  for (int w = 0; w < ctx->world && w < B200DS_P2P_MAX_WORLD; ++w) {
    if (w == ctx->rank) continue;
    if (ctx->p2p_xchg[w]) cudaIpcCloseMemHandle(ctx->p2p_xchg[w]);
    if (ctx->p2p_flags[w]) cudaIpcCloseMemHandle(ctx->p2p_flags[w]);
    ctx->p2p_xchg[w] = nullptr; ctx->p2p_flags[w] = nullptr;
  }
  if (ctx->p2p_xchg[ctx->rank]) cudaFree(ctx->p2p_xchg[ctx->rank]);
  if (ctx->p2p_flags[ctx->rank]) cudaFree(ctx->p2p_flags[ctx->rank]);
  ctx->p2p_xchg[ctx->rank] = nullptr; ctx->p2p_flags[ctx->rank] = nullptr;
  if (ctx->p2p_mailbox) cudaFreeHost(ctx->p2p_mailbox);
  ctx->p2p_mailbox = ctx->p2p_mailbox_dev = nullptr;
  ctx->p2p_own = false;
  ctx->p2p_cap = 0;
  cudaGetLastError();
}

// The local half of the set-up; on failure the caller still takes part in the vote below.
static int p2p_local_setup(b200ds_ctx* ctx, int64_t cap, void** words) {
  B200_CUDA(cudaMalloc(&ctx->p2p_xchg[ctx->rank], sizeof(float) * 2 * (size_t)cap * (size_t)(1 + ctx->world)));   // pull region + one push slot per rank, two halves each
  B200_CUDA(cudaMalloc(words, 256));      // flag row (W x u64) | epoch (u64) | tickets (2 x u32)
  ctx->p2p_flags[ctx->rank] = reinterpret_cast<unsigned long long*>(*words);
  B200_CUDA(cudaMemset(*words, 0, 256));
  B200_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&ctx->p2p_mailbox), 2 * sizeof(unsigned long long), cudaHostAllocMapped));
  ctx->p2p_mailbox[0] = ctx->p2p_mailbox[1] = 0ull;
  B200_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&ctx->p2p_mailbox_dev), ctx->p2p_mailbox, 0));
  return B200DS_OK;
}

int b200ds_ctx_enable_p2p(b200ds_ctx* ctx, int64_t max_elems) {
  B200_REQUIRE(ctx != nullptr, "ctx_enable_p2p: null context");
  B200_REQUIRE(max_elems > 0, "ctx_enable_p2p: max_elems must be positive");
  B200_REQUIRE(ctx->world <= B200DS_P2P_MAX_WORLD, "ctx_enable_p2p: at most %d ranks (one NVLink domain), got %d",
               B200DS_P2P_MAX_WORLD, ctx->world);
  B200_REQUIRE(ctx->p2p_cap == 0, "ctx_enable_p2p: already enabled");
  B200_CUDA(cudaSetDevice(ctx->device));
  const int W = ctx->world;
  const int64_t cap = (max_elems + 3) / 4 * 4;
  {
    const char* ms = getenv("B200DS_P2P_TIMEOUT_MS");
    long long v = ms ? atoll(ms) : 0;
    ctx->p2p_timeout_ns = (unsigned long long)(v > 0 ? v : 300000) * 1000000ull;
  }
  void* words = nullptr;
  ctx->p2p_own = true;
  int my_rc = p2p_local_setup(ctx, cap, &words);
  char my_err[256];
  snprintf(my_err, sizeof(my_err), "%s", my_rc ? b200ds_last_error() : "");
  if (W > 1) {
    // The outcome is COLLECTIVE: a rank that cannot map a peer must not leave the others enabled (their next
    // exchange would wait for a publication that never comes).  Every rank therefore walks the same sequence of
    // communicator calls whatever happens locally, and the last one sums a failure flag over all ranks.
    const NcclApi* api = nccl_api();
    if (!api) { p2p_teardown(ctx); return B200DS_ERR_NCCL; }
    if (!ctx->comm) { p2p_teardown(ctx); set_error("ctx_enable_p2p: context has no communicator"); return B200DS_ERR_STATE; }
    static_assert(sizeof(P2PHandles) == 128, "two IPC handles");
    P2PHandles mine;
    memset(&mine, 0, sizeof(mine));
    if (!my_rc) {
      cudaError_t e = cudaIpcGetMemHandle(&mine.xchg, ctx->p2p_xchg[ctx->rank]);
      if (e == cudaSuccess) e = cudaIpcGetMemHandle(&mine.flags, words);
      if (e != cudaSuccess) { my_rc = cuda_fail(e, "cudaIpcGetMemHandle", __FILE__, __LINE__); snprintf(my_err, sizeof(my_err), "%s", b200ds_last_error()); }
    }
    unsigned char* gather = nullptr;
    P2PHandles all[B200DS_P2P_MAX_WORLD];
    cudaError_t ge = cudaMalloc(&gather, sizeof(P2PHandles) * (size_t)W);
    if (ge != cudaSuccess) { p2p_teardown(ctx); return cuda_fail(ge, "cudaMalloc(handle gather)", __FILE__, __LINE__); }   // nothing collective can run without it
    cudaMemcpy(gather + sizeof(P2PHandles) * (size_t)ctx->rank, &mine, sizeof(mine), cudaMemcpyHostToDevice);
    ncclResult_t r = api->AllGather(gather + sizeof(P2PHandles) * (size_t)ctx->rank, gather, sizeof(P2PHandles), ncclChar, ctx->comm,
                                    ctx->side);
    cudaError_t se = cudaStreamSynchronize(ctx->side);
    if (r == ncclSuccess && se == cudaSuccess) se = cudaMemcpy(all, gather, sizeof(P2PHandles) * (size_t)W, cudaMemcpyDeviceToHost);
    cudaFree(gather);
    if (r != ncclSuccess || se != cudaSuccess) {
      p2p_teardown(ctx);
      if (r != ncclSuccess) { set_error("ctx_enable_p2p: ncclAllGather failed (%s)", api->GetErrorString(r)); return B200DS_ERR_NCCL; }
      return cuda_fail(se, "handle exchange", __FILE__, __LINE__);
    }
    for (int w = 0; w < W && my_rc == B200DS_OK; ++w) {
      if (w == ctx->rank) continue;
      void* px = nullptr; void* pf = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&px, all[w].xchg, cudaIpcMemLazyEnablePeerAccess);
      if (e == cudaSuccess) {
        ctx->p2p_xchg[w] = reinterpret_cast<float*>(px);
        e = cudaIpcOpenMemHandle(&pf, all[w].flags, cudaIpcMemLazyEnablePeerAccess);
      }
      if (e == cudaSuccess) {
        ctx->p2p_flags[w] = reinterpret_cast<unsigned long long*>(pf);
      } else {
        cudaGetLastError();
        snprintf(my_err, sizeof(my_err), "ctx_enable_p2p: cannot map rank %d's exchange buffer: %s", w, cudaGetErrorString(e));
        my_rc = B200DS_ERR_CUDA;
      }
    }
    // the vote; it is also the barrier after which a rank may publish (every mapping is open, every word zeroed)
    float vote = my_rc ? 1.f : 0.f, total = 1.f;
    cudaError_t ve = cudaMemcpy(ctx->barrier_word + 1, &vote, sizeof(float), cudaMemcpyHostToDevice);
    r = api->AllReduce(ctx->barrier_word + 1, ctx->barrier_word + 1, 1, ncclFloat32, ncclSum, ctx->comm, ctx->side);
    if (ve == cudaSuccess) ve = cudaStreamSynchronize(ctx->side);
    if (ve == cudaSuccess && r == ncclSuccess) ve = cudaMemcpy(&total, ctx->barrier_word + 1, sizeof(float), cudaMemcpyDeviceToHost);
    if (r != ncclSuccess || ve != cudaSuccess || total != 0.f) {
      p2p_teardown(ctx);
      if (my_rc) { set_error("%s", my_err); return my_rc; }
      if (r != ncclSuccess) { set_error("ctx_enable_p2p: ncclAllReduce failed (%s)", api->GetErrorString(r)); return B200DS_ERR_NCCL; }
      if (ve != cudaSuccess) return cuda_fail(ve, "set-up vote", __FILE__, __LINE__);
      set_error("ctx_enable_p2p: %d rank(s) could not map their peers; the peer-memory exchange stays off on every rank", (int)total);
      return B200DS_ERR_STATE;
    }
  } else if (my_rc) {
    p2p_teardown(ctx);
    set_error("%s", my_err);
    return my_rc;
  }
  ctx->p2p_cap = cap;
  return B200DS_OK;
}

// A wait that timed out in an earlier launch left a note in the mailbox (and a trapped context).
int p2p_check_mailbox(const b200ds_ctx* ctx, const char* who) {
  if (ctx->p2p_mailbox && ctx->p2p_mailbox[0] != 0ull) {
    set_error("%s: an earlier exchange on rank %d gave up waiting for rank %d at epoch %llu (B200DS_P2P_TIMEOUT_MS = %llu)", who,
              ctx->rank, (int)(ctx->p2p_mailbox[0] - 1ull), ctx->p2p_mailbox[1], ctx->p2p_timeout_ns / 1000000ull);
    return B200DS_ERR_STATE;
  }
  return B200DS_OK;
}

int b200ds_allreduce_adam_p2p(b200ds_ctx* ctx, float* params_dev, float* grads_dev, float* exp_avg_dev,
                              float* exp_avg_sq_dev, int64_t n, double lr, double beta1, double beta2, double eps,
                              int64_t step, const uint64_t* step_counter_dev, uintptr_t stream) {
  B200_REQUIRE(ctx != nullptr, "allreduce_adam_p2p: null context");
  B200_REQUIRE(ctx->p2p_cap > 0, "allreduce_adam_p2p: call b200ds_ctx_enable_p2p first");
  B200_REQUIRE(n > 0 && n % 4 == 0 && n <= ctx->p2p_cap, "allreduce_adam_p2p: n (%lld) must be a positive multiple of 4, at most %lld",
               (long long)n, (long long)ctx->p2p_cap);
  B200_REQUIRE(params_dev && grads_dev && exp_avg_dev && exp_avg_sq_dev, "allreduce_adam_p2p: null pointer");
  B200_REQUIRE(((((uintptr_t)params_dev) | ((uintptr_t)grads_dev) | ((uintptr_t)exp_avg_dev) | ((uintptr_t)exp_avg_sq_dev)) & 15) == 0,
               "allreduce_adam_p2p: buffers must be 16-byte aligned");
  B200_REQUIRE(step_counter_dev != nullptr || step >= 1, "allreduce_adam_p2p: step must be >= 1 (t after the increment)");
  if (int rc = p2p_check_mailbox(ctx, "allreduce_adam_p2p")) return rc;
  const P2PDev d = make_p2p_dev(ctx);
  const AdamScalars s = make_adam_scalars(beta1, beta2, eps, ctx->world > 1 ? (float)(1.0 / ctx->world) : 1.f);
  // every CTA spins in phase 2, so all of them must be resident at once: at most one CTA per SM
  int64_t want = ((n >> 2) + kP2PThreads - 1) / kP2PThreads;
  int grid = (int)(want < 1 ? 1 : (want > sm_count() ? sm_count() : want));
  if (step_counter_dev)
    p2p_allreduce_adam_kernel<true><<<grid, kP2PThreads, 0, as_stream(stream)>>>(
        params_dev, grads_dev, exp_avg_dev, exp_avg_sq_dev, n, s, lr, beta1, beta2, 0,
        reinterpret_cast<const unsigned long long*>(step_counter_dev), d);
  else
    p2p_allreduce_adam_kernel<false><<<grid, kP2PThreads, 0, as_stream(stream)>>>(
        params_dev, grads_dev, exp_avg_dev, exp_avg_sq_dev, n, s, lr, beta1, beta2, step, nullptr, d);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

}  // extern "C"
