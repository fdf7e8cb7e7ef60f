// Context life cycle and the gradient exchange of model_engine.backward (train.py:113):
// one flat fp32 bucket, ncclAllReduce(SUM) over NVLink 5 / NVSwitch on a side stream,
// event-chained to the caller's stream (no host sync).  DeepSpeed stage 0 semantics
// (ds_config train.py:78-83): sum, then x 1/world -- the scale is folded into the Adam
// kernel's gradient read.
#include "ctx.cuh"
#include <dlfcn.h>
#include <string.h>

namespace b200ds {

static NcclApi g_api;
static int g_api_state = 0;   // 0 = untried, 1 = ok, -1 = failed

const NcclApi* nccl_api() {
  if (g_api_state == 1) return &g_api;
  if (g_api_state == -1) return nullptr;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    set_error("cannot load libnccl.so.2: %s", dlerror());
    g_api_state = -1;
    return nullptr;
  }
#define B200_SYM(field, name)                                              \
  *(void**)(&g_api.field) = dlsym(h, name);                                \
  if (!g_api.field) { set_error("libnccl: missing symbol %s", name); g_api_state = -1; return nullptr; }
  B200_SYM(GetUniqueId, "ncclGetUniqueId")
  B200_SYM(CommInitRank, "ncclCommInitRank")
  B200_SYM(CommDestroy, "ncclCommDestroy")
  B200_SYM(AllReduce, "ncclAllReduce")
  B200_SYM(Broadcast, "ncclBroadcast")
  B200_SYM(ReduceScatter, "ncclReduceScatter")
  B200_SYM(AllGather, "ncclAllGather")
  B200_SYM(GetErrorString, "ncclGetErrorString")
#undef B200_SYM
  g_api_state = 1;
  return &g_api;
}

static int nccl_fail(const NcclApi* api, ncclResult_t r, const char* what) {
  set_error("NCCL error %d (%s) in %s", (int)r, api->GetErrorString(r), what);
  return B200DS_ERR_NCCL;
}

#define B200_NCCL(api, expr)                                              \
  do {                                                                    \
    ncclResult_t _r = (expr);                                             \
    if (_r != ncclSuccess) return nccl_fail(api, _r, #expr);              \
  } while (0)

__global__ void iota_kernel(int64_t* p, int n) {
  int i = threadIdx.x;
  if (i < n) p[i] = i;
}

}  // namespace b200ds

using namespace b200ds;

extern "C" {

int b200ds_nccl_unique_id(void* id128) {
  B200_REQUIRE(id128 != nullptr, "nccl_unique_id: null output");
  static_assert(sizeof(ncclUniqueId) == B200DS_NCCL_ID_BYTES, "ncclUniqueId size");
  const NcclApi* api = nccl_api();
  if (!api) return B200DS_ERR_NCCL;
  ncclUniqueId id;
  B200_NCCL(api, api->GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return B200DS_OK;
}

int b200ds_ctx_create(int device, int rank, int world, const void* nccl_id128, b200ds_ctx** out) {
  B200_REQUIRE(out != nullptr, "ctx_create: null output");
  *out = nullptr;
  B200_REQUIRE(world >= 1 && rank >= 0 && rank < world, "ctx_create: bad rank/world %d/%d", rank, world);
  B200_REQUIRE(world == 1 || nccl_id128 != nullptr, "ctx_create: world > 1 needs the NCCL unique id");
  int ndev = 0;
  B200_CUDA(cudaGetDeviceCount(&ndev));
  B200_REQUIRE(device >= 0 && device < ndev, "ctx_create: device %d not in [0,%d)", device, ndev);
  B200_CUDA(cudaSetDevice(device));
  int major = 0, minor = 0;
  B200_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  B200_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  B200_REQUIRE(major == 10, "ctx_create: this library is built for sm_100a only (device is sm_%d%d)", major, minor);
  b200ds_ctx* c = new b200ds_ctx();
  c->device = device; c->rank = rank; c->world = world;
  cudaError_t e;
  if ((e = cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaEventCreateWithFlags(&c->ev_ready, cudaEventDisableTiming)) != cudaSuccess ||
      (e = cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming)) != cudaSuccess ||
      (e = cudaMalloc(&c->iota, 32 * sizeof(int64_t))) != cudaSuccess ||
      (e = cudaMalloc(&c->barrier_word, 256)) != cudaSuccess) {
    b200ds_ctx_destroy(c);
    return cuda_fail(e, "context resources", __FILE__, __LINE__);
  }
  iota_kernel<<<1, 32, 0, c->side>>>(c->iota, 32);
  if ((e = cudaMemsetAsync(c->barrier_word, 0, 256, c->side)) != cudaSuccess ||
      (e = cudaStreamSynchronize(c->side)) != cudaSuccess) {
    b200ds_ctx_destroy(c);
    return cuda_fail(e, "context init", __FILE__, __LINE__);
  }
  if (world > 1) {
    const NcclApi* api = nccl_api();
    if (!api) { b200ds_ctx_destroy(c); return B200DS_ERR_NCCL; }
    ncclUniqueId id;
    memcpy(&id, nccl_id128, sizeof(id));
    ncclResult_t r = api->CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) {
      c->comm = nullptr;
      b200ds_ctx_destroy(c);
      return nccl_fail(api, r, "ncclCommInitRank");
    }
  }
  *out = c;
  return B200DS_OK;
}

int b200ds_ctx_destroy(b200ds_ctx* c) {
  if (!c) return B200DS_OK;
  cudaSetDevice(c->device);
  if (c->side) cudaStreamSynchronize(c->side);
  for (auto& g : c->sgraph)          // before the communicator: NCCL keeps it alive while a graph refers to it
    if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
  if (c->cap) cudaStreamDestroy(c->cap);
  if (c->step_word) cudaFree(c->step_word);
  if (c->step_pinned) cudaFreeHost(c->step_pinned);
  if (c->copy) { cudaStreamSynchronize(c->copy); cudaStreamDestroy(c->copy); }
  if (c->p2p_own) {
    cudaDeviceSynchronize();
    for (int w = 0; w < c->world && w < B200DS_P2P_MAX_WORLD; ++w) {
      if (w == c->rank) continue;
      if (c->p2p_xchg[w]) cudaIpcCloseMemHandle(c->p2p_xchg[w]);
      if (c->p2p_flags[w]) cudaIpcCloseMemHandle(c->p2p_flags[w]);
    }
    if (c->p2p_xchg[c->rank]) cudaFree(c->p2p_xchg[c->rank]);
    if (c->p2p_flags[c->rank]) cudaFree(c->p2p_flags[c->rank]);
    if (c->p2p_mailbox) cudaFreeHost(c->p2p_mailbox);
  }
  if (c->ev_copied) cudaEventDestroy(c->ev_copied);
  if (c->comm) { const NcclApi* api = nccl_api(); if (api) api->CommDestroy(c->comm); }
  if (c->ev_ready) cudaEventDestroy(c->ev_ready);
  if (c->ev_done) cudaEventDestroy(c->ev_done);
  if (c->side) cudaStreamDestroy(c->side);
  if (c->iota) cudaFree(c->iota);
  if (c->barrier_word) cudaFree(c->barrier_word);
  delete c;
  return B200DS_OK;
}

int b200ds_ctx_info(const b200ds_ctx* ctx, int* device, int* rank, int* world) {
  B200_REQUIRE(ctx != nullptr, "ctx_info: null context");
  if (device) *device = ctx->device;
  if (rank) *rank = ctx->rank;
  if (world) *world = ctx->world;
  return B200DS_OK;
}

int b200ds_allreduce_bucket_begin(b200ds_ctx* ctx, float* bucket_dev, int64_t n, uintptr_t stream) {
  B200_REQUIRE(ctx != nullptr, "allreduce_bucket: null context");
  B200_REQUIRE(n >= 0, "allreduce_bucket: negative n");
  if (ctx->world == 1 || n == 0) return B200DS_OK;
  B200_REQUIRE(bucket_dev != nullptr, "allreduce_bucket: null bucket");
  if (!ctx->comm) { set_error("allreduce_bucket: context has no communicator"); return B200DS_ERR_STATE; }
  const NcclApi* api = nccl_api();
  if (!api) return B200DS_ERR_NCCL;
  cudaStream_t s = as_stream(stream);
  B200_CUDA(cudaEventRecord(ctx->ev_ready, s));
  B200_CUDA(cudaStreamWaitEvent(ctx->side, ctx->ev_ready, 0));
  B200_NCCL(api, api->AllReduce(bucket_dev, bucket_dev, (size_t)n, ncclFloat32, ncclSum, ctx->comm, ctx->side));
  B200_CUDA(cudaEventRecord(ctx->ev_done, ctx->side));
  return B200DS_OK;
}

int b200ds_allreduce_bucket_end(b200ds_ctx* ctx, uintptr_t stream) {
  B200_REQUIRE(ctx != nullptr, "allreduce_bucket_end: null context");
  if (ctx->world == 1) return B200DS_OK;
  B200_CUDA(cudaStreamWaitEvent(as_stream(stream), ctx->ev_done, 0));
  return B200DS_OK;
}

int b200ds_allreduce_bucket(b200ds_ctx* ctx, float* bucket_dev, int64_t n, uintptr_t stream) {
  int rc = b200ds_allreduce_bucket_begin(ctx, bucket_dev, n, stream);
  if (rc) return rc;
  if (n == 0) return B200DS_OK;
  return b200ds_allreduce_bucket_end(ctx, stream);
}

// ZeRO-1-style step (BASELINE config 5; the reference itself runs stage 0, train.py:82):
// reduce-scatter the gradient bucket, Adam on this rank's 1/W slice only (so exp_avg /
// exp_avg_sq are slice-sized), all-gather the updated parameters.  Same arithmetic per element
// as stage 0 up to the summation order inside NCCL.
int b200ds_zero1_step(b200ds_ctx* ctx, float* params_dev, float* grads_dev, float* exp_avg_shard_dev,
                      float* exp_avg_sq_shard_dev, int64_t n, double lr, double beta1, double beta2, double eps,
                      int64_t step, uintptr_t stream) {
  B200_REQUIRE(ctx != nullptr, "zero1_step: null context");
  B200_REQUIRE(params_dev && grads_dev && exp_avg_shard_dev && exp_avg_sq_shard_dev, "zero1_step: null pointer");
  const int W = ctx->world;
  B200_REQUIRE(n > 0 && n % (4 * (int64_t)W) == 0, "zero1_step: n (%lld) must be a positive multiple of 4*world (%d)",
               (long long)n, 4 * W);
  const int64_t shard = n / W, off = shard * ctx->rank;
  if (W == 1)
    return b200ds_adam_step(params_dev, grads_dev, exp_avg_shard_dev, exp_avg_sq_shard_dev, n, lr, beta1, beta2, eps, step, 1.f, 0, stream);
  if (!ctx->comm) { set_error("zero1_step: context has no communicator"); return B200DS_ERR_STATE; }
  const NcclApi* api = nccl_api();
  if (!api) return B200DS_ERR_NCCL;
  cudaStream_t s = as_stream(stream);
  // exchange on the side stream, ordered after the caller's work; the Adam slice runs there too so the
  // three phases need no round trip through the caller's stream
  B200_CUDA(cudaEventRecord(ctx->ev_ready, s));
  B200_CUDA(cudaStreamWaitEvent(ctx->side, ctx->ev_ready, 0));
  B200_NCCL(api, api->ReduceScatter(grads_dev, grads_dev + off, (size_t)shard, ncclFloat32, ncclSum, ctx->comm, ctx->side));
  int rc = b200ds_adam_step(params_dev + off, grads_dev + off, exp_avg_shard_dev, exp_avg_sq_shard_dev, shard, lr, beta1, beta2,
                            eps, step, (float)(1.0 / W), 0, (uintptr_t)ctx->side);
  if (rc) return rc;
  B200_NCCL(api, api->AllGather(params_dev + off, params_dev, (size_t)shard, ncclFloat32, ctx->comm, ctx->side));
  B200_CUDA(cudaEventRecord(ctx->ev_done, ctx->side));
  B200_CUDA(cudaStreamWaitEvent(s, ctx->ev_done, 0));
  return B200DS_OK;
}

// One collective of the step's repertoire on a raw buffer, for the fabric sweep (tools/allreduce_sweep.py; BASELINE
// config 5, SURVEY 8f-4 -- and the reference's own purpose, README.md:3: an acceptance check of NCCL over the fabric).
int b200ds_collective(b200ds_ctx* ctx, int op, int dtype, void* buf_dev, int64_t n, uintptr_t stream) {
  B200_REQUIRE(ctx != nullptr, "collective: null context");
  B200_REQUIRE(op >= 0 && op <= 2, "collective: op must be 0 (all-reduce), 1 (reduce-scatter) or 2 (all-gather)");
  B200_REQUIRE(dtype == 0 || dtype == 1, "collective: dtype must be 0 (fp32) or 1 (bf16)");
  B200_REQUIRE(n >= 0, "collective: negative n");
  if (ctx->world == 1 || n == 0) return B200DS_OK;
  B200_REQUIRE(buf_dev != nullptr, "collective: null buffer");
  B200_REQUIRE(op == 0 || n % ctx->world == 0, "collective: n (%lld) must be a multiple of the world size for reduce-scatter / all-gather", (long long)n);
  if (!ctx->comm) { set_error("collective: context has no communicator"); return B200DS_ERR_STATE; }
  const NcclApi* api = nccl_api();
  if (!api) return B200DS_ERR_NCCL;
  const ncclDataType_t dt = dtype == 0 ? ncclFloat32 : ncclBfloat16;
  const size_t esz = dtype == 0 ? 4 : 2;
  const int64_t shard = n / ctx->world;
  char* base = reinterpret_cast<char*>(buf_dev);
  char* mine = base + (size_t)shard * ctx->rank * esz;
  cudaStream_t s = as_stream(stream);
  B200_CUDA(cudaEventRecord(ctx->ev_ready, s));
  B200_CUDA(cudaStreamWaitEvent(ctx->side, ctx->ev_ready, 0));
  if (op == 0) B200_NCCL(api, api->AllReduce(base, base, (size_t)n, dt, ncclSum, ctx->comm, ctx->side));
  else if (op == 1) B200_NCCL(api, api->ReduceScatter(base, mine, (size_t)shard, dt, ncclSum, ctx->comm, ctx->side));
  else B200_NCCL(api, api->AllGather(mine, base, (size_t)shard, dt, ctx->comm, ctx->side));
  B200_CUDA(cudaEventRecord(ctx->ev_done, ctx->side));
  B200_CUDA(cudaStreamWaitEvent(s, ctx->ev_done, 0));
  return B200DS_OK;
}

int b200ds_broadcast_params(b200ds_ctx* ctx, float* params_dev, int64_t n, uintptr_t stream) {
  B200_REQUIRE(ctx != nullptr, "broadcast_params: null context");
  if (ctx->world == 1 || n <= 0) return B200DS_OK;
  B200_REQUIRE(params_dev != nullptr, "broadcast_params: null buffer");
  if (!ctx->comm) { set_error("broadcast_params: context has no communicator"); return B200DS_ERR_STATE; }
  const NcclApi* api = nccl_api();
  if (!api) return B200DS_ERR_NCCL;
  B200_NCCL(api, api->Broadcast(params_dev, params_dev, (size_t)n, ncclFloat32, 0, ctx->comm, as_stream(stream)));
  return B200DS_OK;
}

int b200ds_barrier(b200ds_ctx* ctx) {
  B200_REQUIRE(ctx != nullptr, "barrier: null context");
  B200_CUDA(cudaSetDevice(ctx->device));
  if (ctx->world > 1) {
    if (!ctx->comm) { set_error("barrier: context has no communicator"); return B200DS_ERR_STATE; }
    const NcclApi* api = nccl_api();
    if (!api) return B200DS_ERR_NCCL;
    B200_NCCL(api, api->AllReduce(ctx->barrier_word, ctx->barrier_word, 1, ncclFloat32, ncclSum, ctx->comm, ctx->side));
  }
  B200_CUDA(cudaStreamSynchronize(ctx->side));
  return B200DS_OK;
}

}  // extern "C"
