// Fused BCE-with-logits loss + d(loss)/d(logits): replaces nn.BCEWithLogitsLoss()
// (train.py:96,112; ~5 ATen launches) and the first autograd node of
// model_engine.backward (train.py:113; 5 more).  One pass over z and y: 8 B read + 4 B
// written per sample, warp-shuffle reductions, one atomic ticket per CTA so the mean is
// finished by the last CTA in a fixed order (deterministic, single launch).
#include "common.cuh"

namespace b200ds {

constexpr int kLossThreads = 256;
constexpr int kLossMaxBlocks = 1184;   // 8 CTAs x 148 SMs

struct LossScratch {          // caller-provided, zeroed once; the kernel leaves ticket == 0
  unsigned int ticket;
  unsigned int pad[3];
  float partial[kLossMaxBlocks];
};

__global__ void __launch_bounds__(kLossThreads)
bce_loss_grad_kernel(const float* __restrict__ z, const float* __restrict__ y, int64_t batch,
                     float* __restrict__ loss, float* __restrict__ dz,
                     unsigned long long* __restrict__ step_counter, LossScratch* __restrict__ scr) {
  __shared__ float warp_part[kLossThreads / 32];
  __shared__ bool is_last;
  const float fb = (float)batch;
  float acc = 0.f;
  // 128-bit main body, scalar tail
  const int64_t nvec = batch >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool vec_ok = ((((uintptr_t)z) | ((uintptr_t)y) | ((uintptr_t)dz)) & 15) == 0;
  if (vec_ok) {
    for (int64_t i = t0; i < nvec; i += stride) {
      float4 zv = __ldg(reinterpret_cast<const float4*>(z) + i);
      float4 yv = __ldg(reinterpret_cast<const float4*>(y) + i);
      acc += bce_term(zv.x, yv.x);
      acc += bce_term(zv.y, yv.y);
      acc += bce_term(zv.z, yv.z);
      acc += bce_term(zv.w, yv.w);
      if (dz) {
        float4 g = make_float4(bce_grad(zv.x, yv.x, fb), bce_grad(zv.y, yv.y, fb),
                               bce_grad(zv.z, yv.z, fb), bce_grad(zv.w, yv.w, fb));
        reinterpret_cast<float4*>(dz)[i] = g;
      }
    }
    for (int64_t i = (nvec << 2) + t0; i < batch; i += stride) {
      float zi = z[i], yi = y[i];
      acc += bce_term(zi, yi);
      if (dz) dz[i] = bce_grad(zi, yi, fb);
    }
  } else {
    for (int64_t i = t0; i < batch; i += stride) {
      float zi = z[i], yi = y[i];
      acc += bce_term(zi, yi);
      if (dz) dz[i] = bce_grad(zi, yi, fb);
    }
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < kLossThreads / 32 ? warp_part[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) {
      if (gridDim.x == 1) {
        *loss = v / fb;
        if (step_counter) *step_counter += 1ull;
        is_last = false;
      } else {
        scr->partial[blockIdx.x] = v;
        __threadfence();
        unsigned t = atomicAdd(&scr->ticket, 1u);
        is_last = (t == gridDim.x - 1);
      }
    }
  }
  __syncthreads();
  if (is_last) {   // the CTA that arrived last sums the partials in block order, in double
    __threadfence();
    double s = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) s += (double)__ldcg(&scr->partial[i]);
    __shared__ double dpart[kLossThreads / 32];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) dpart[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double tot = 0.0;
      for (int w = 0; w < kLossThreads / 32; ++w) tot += dpart[w];
      *loss = (float)(tot / (double)batch);
      if (step_counter) *step_counter += 1ull;
      scr->ticket = 0u;
    }
  }
}

// ---- scratch for the allocation-free entry point -------------------------------------
// Lives in the library image: every GPU that loads libb200ds.so carries these 1280 bytes
// until the process exits.  256 per-CTA partial sums + the arrival ticket; the prefix of the
// layout is the same as LossScratch, so the kernel above runs on it unchanged as long as the
// grid has at most kStaticLossMaxBlocks CTAs.
constexpr int kStaticLossMaxBlocks = 256;
constexpr int64_t kStaticLossMaxBatch = 4194304;     // 2^22 rows: 4 rows per thread per pass, 16 passes at most
constexpr size_t kStaticLossBytes = 1280;
static_assert(16 + kStaticLossMaxBlocks * sizeof(float) <= kStaticLossBytes, "static loss scratch too small");
static __device__ __align__(256) unsigned char g_static_loss_scratch[kStaticLossBytes];

}  // namespace b200ds

using namespace b200ds;

extern "C" {

size_t b200ds_bce_scratch_bytes(int64_t) { return sizeof(LossScratch); }

int b200ds_bce_loss_grad(const float* logits_dev, const float* y_dev, int64_t batch,
                         float* loss_dev, float* dlogits_dev, uint64_t* step_counter_dev,
                         void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  B200_REQUIRE(batch > 0, "bce_loss_grad: batch must be positive (got %lld)", (long long)batch);
  B200_REQUIRE(logits_dev && y_dev && loss_dev, "bce_loss_grad: null pointer");
  int64_t want = (batch / 4 + kLossThreads - 1) / kLossThreads;
  int grid = (int)(want < 1 ? 1 : (want > kLossMaxBlocks ? kLossMaxBlocks : want));
  if (grid > 1)
    B200_REQUIRE(scratch_dev && scratch_bytes >= sizeof(LossScratch),
                 "bce_loss_grad: scratch of %zu bytes required (zeroed once)", sizeof(LossScratch));
  bce_loss_grad_kernel<<<grid, kLossThreads, 0, as_stream(stream)>>>(
      logits_dev, y_dev, batch, loss_dev, dlogits_dev,
      reinterpret_cast<unsigned long long*>(step_counter_dev), reinterpret_cast<LossScratch*>(scratch_dev));
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

// The same loss + gradient pass for callers that have no way to hand over device scratch.
// Two kinds of caller of this library are in that position: a launcher or plugin host written
// against the C ABI alone (it holds the tensors the framework gave it and has no allocator
// of its own to ask for more), and a step loop running under a framework whose caching
// allocator already owns the GPU's memory, where a raw cudaMalloc from a plugin can fail or
// stall the device mid-step.  Nothing is allocated or freed here, on any path.
//
// The scratch is the file-scope device array above.  Its device address is looked up once per
// device, by the first call made while that device is current (cudaGetSymbolAddress); make
// that first call outside of any stream capture.  Because the array is shared by every call
// and keeps whatever the previous call left in it, the arrival ticket is cleared on `stream`
// before each launch; two calls whose GPU work overlaps (different streams, no ordering
// between them) would race on it and must be serialised by the caller.
int b200ds_bce_loss_grad_noalloc(const float* logits_dev, const float* y_dev, int64_t batch,
                                 float* loss_dev, float* dlogits_dev, uint64_t* step_counter_dev,
                                 uintptr_t stream) {
  B200_REQUIRE(batch > 0, "bce_loss_grad_noalloc: batch must be positive (got %lld)", (long long)batch);
  B200_REQUIRE(batch <= kStaticLossMaxBatch,
               "bce_loss_grad_noalloc: batch %lld exceeds the built-in scratch limit of %lld rows; "
               "use b200ds_bce_loss_grad with caller scratch", (long long)batch, (long long)kStaticLossMaxBatch);
  B200_REQUIRE(logits_dev && y_dev && loss_dev, "bce_loss_grad_noalloc: null pointer");
  static void* scratch_of_device[64];
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < 64, "bce_loss_grad_noalloc: device index %d out of range", dev);
  if (scratch_of_device[dev] == nullptr) {
    // the array's name is only a host-side handle; the kernel needs the address on THIS device
    void* p = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&p, g_static_loss_scratch));
    scratch_of_device[dev] = p;
  }
  LossScratch* scr = reinterpret_cast<LossScratch*>(scratch_of_device[dev]);
  cudaStream_t st = as_stream(stream);
  int64_t want = (batch / 4 + kLossThreads - 1) / kLossThreads;
  int grid = (int)(want < 1 ? 1 : (want > kStaticLossMaxBlocks ? kStaticLossMaxBlocks : want));
  B200_CUDA(cudaMemsetAsync(scr, 0, 16, st));      // ticket must start at zero whatever ran before
  bce_loss_grad_kernel<<<grid, kLossThreads, 0, st>>>(logits_dev, y_dev, batch, loss_dev, dlogits_dev,
                                                      reinterpret_cast<unsigned long long*>(step_counter_dev), scr);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

}  // extern "C"
