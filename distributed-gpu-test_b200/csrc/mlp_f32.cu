// fp32 CUDA-core path of the model step, shape-generic (F <= 256 in backward): one kernel
// for model_engine(batch_x) (train.py:107 -> SimpleNet.forward train.py:35-36, 3 library
// launches in the reference) and one (+ a small finish) for the autograd part of
// model_engine.backward (train.py:113, ~11 launches).  This is the exact-precision path
// used by BASELINE configs 1-2 and as the on-device comparison for the bf16 tensor-core
// path; the hidden activation is recomputed in backward rather than stored.
#include "common.cuh"

namespace b200ds {

// ---------------------------------------------------------------- forward
// One warp per sample row; lanes stride over hidden units; warp-shuffle reduce for layer 2.
__global__ void __launch_bounds__(256)
mlp_forward_f32_kernel(const float* __restrict__ params, const float* __restrict__ x, int64_t batch,
                       Layout L, float* __restrict__ logits) {
  extern __shared__ float xs[];                 // [warps][F]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const float* w1 = params + L.off_w1();
  const float* b1 = params + L.off_b1();
  const float* w2 = params + L.off_w2();
  const float b2 = params[L.off_b2()];
  float* xrow = xs + warp * L.F;
  for (int64_t b = (int64_t)blockIdx.x * nwarp + warp; b < batch; b += (int64_t)gridDim.x * nwarp) {
    for (int f = lane; f < L.F; f += 32) xrow[f] = x[b * L.F + f];
    __syncwarp();
    float acc = 0.f;
    for (int n = lane; n < L.H; n += 32) {
      const float* wr = w1 + (int64_t)n * L.F;
      float z1 = b1[n];
      for (int f = 0; f < L.F; ++f) z1 = fmaf(xrow[f], __ldg(wr + f), z1);
      acc = fmaf(fmaxf(z1, 0.f), w2[n], acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) logits[b] = acc + b2;
    __syncwarp();
  }
}

// ---------------------------------------------------------------- backward
// grid = (hidden chunks of 64, row splits).  Per 32-row tile:
//   phase A: dz1[r][n] = dz[r] w2[n] 1[z1>0] and hd[r][n] = dz[r] relu(z1) into smem
//   phase B: thread-owned accumulators acc[n][c], c in [0,F+2): c<F -> dW1[n,c], F -> db1[n], F+1 -> dW2[n]
// Partials per row split go to scratch; backward_finish sums them in split order.
constexpr int kBwdChunk = 64;
constexpr int kBwdRows = 32;
constexpr int kBwdThreads = 256;

__global__ void __launch_bounds__(kBwdThreads)
mlp_backward_f32_kernel(const float* __restrict__ params, const float* __restrict__ x,
                        const float* __restrict__ dz, int64_t batch, Layout L,
                        float* __restrict__ partial /* [splits][H][F+2] */) {
  extern __shared__ float sm[];
  const int F = L.F, C = F + 2;
  float* w1s = sm;                               // [64][F+1]  (padded rows)
  float* xs = w1s + kBwdChunk * (F + 1);         // [32][F]
  float* g = xs + kBwdRows * F;                  // [32][64]
  float* hd = g + kBwdRows * kBwdChunk;          // [32][64]
  float* acc = hd + kBwdRows * kBwdChunk;        // [64][C]
  float* dzs = acc + kBwdChunk * C;              // [32]
  const int n0 = blockIdx.x * kBwdChunk;
  const int nn = min(kBwdChunk, L.H - n0);
  const float* w1 = params + L.off_w1();
  const float* b1 = params + L.off_b1();
  const float* w2 = params + L.off_w2();
  for (int i = threadIdx.x; i < kBwdChunk * F; i += kBwdThreads) {
    int n = i / F, f = i - n * F;
    w1s[n * (F + 1) + f] = n < nn ? w1[(int64_t)(n0 + n) * F + f] : 0.f;
  }
  for (int i = threadIdx.x; i < kBwdChunk * C; i += kBwdThreads) acc[i] = 0.f;
  const int64_t rows_per_split = ((batch + gridDim.y - 1) / gridDim.y + kBwdRows - 1) / kBwdRows * kBwdRows;
  const int64_t r_begin = (int64_t)blockIdx.y * rows_per_split;
  const int64_t r_end = min(batch, r_begin + rows_per_split);
  for (int64_t r0 = r_begin; r0 < r_end; r0 += kBwdRows) {
    const int nr = (int)min((int64_t)kBwdRows, r_end - r0);
    __syncthreads();
    for (int i = threadIdx.x; i < kBwdRows * F; i += kBwdThreads) {
      int r = i / F;
      xs[i] = r < nr ? x[r0 * F + i] : 0.f;
    }
    if (threadIdx.x < kBwdRows) dzs[threadIdx.x] = threadIdx.x < nr ? dz[r0 + threadIdx.x] : 0.f;
    __syncthreads();
    for (int i = threadIdx.x; i < kBwdRows * kBwdChunk; i += kBwdThreads) {
      int r = i / kBwdChunk, n = i - r * kBwdChunk;
      float gv = 0.f, hv = 0.f;
      if (n < nn) {
        float z1 = b1[n0 + n];
        const float* wr = w1s + n * (F + 1);
        const float* xr = xs + r * F;
        for (int f = 0; f < F; ++f) z1 = fmaf(xr[f], wr[f], z1);
        float d = dzs[r];
        gv = z1 > 0.f ? d * w2[n0 + n] : 0.f;
        hv = d * fmaxf(z1, 0.f);
      }
      g[i] = gv;
      hd[i] = hv;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < kBwdChunk * C; j += kBwdThreads) {
      int c = j / kBwdChunk, n = j - c * kBwdChunk;   // a warp shares c: xs reads broadcast
      float a = 0.f;
      if (c < F) {
        for (int r = 0; r < kBwdRows; ++r) a = fmaf(g[r * kBwdChunk + n], xs[r * F + c], a);
      } else if (c == F) {
        for (int r = 0; r < kBwdRows; ++r) a += g[r * kBwdChunk + n];
      } else {
        for (int r = 0; r < kBwdRows; ++r) a += hd[r * kBwdChunk + n];
      }
      acc[n * C + c] += a;
    }
  }
  __syncthreads();
  float* out = partial + ((int64_t)blockIdx.y * L.H + n0) * C;
  for (int i = threadIdx.x; i < nn * C; i += kBwdThreads) out[i] = acc[i];
}

// grads[...] = sum over splits, scattered into the flat layout; db2 = sum_b dz[b].
__global__ void __launch_bounds__(256)
mlp_backward_f32_finish(const float* __restrict__ partial, int splits, const float* __restrict__ dz,
                        int64_t batch, Layout L, float* __restrict__ grads) {
  const int C = L.F + 2;
  const int64_t total = (int64_t)L.H * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float a = 0.f;
    for (int s = 0; s < splits; ++s) a += partial[(int64_t)s * total + i];
    int n = (int)(i / C), c = (int)(i - (int64_t)n * C);
    if (c < L.F) grads[L.off_w1() + (int64_t)n * L.F + c] = a;
    else if (c == L.F) grads[L.off_b1() + n] = a;
    else grads[L.off_w2() + n] = a;
  }
  if (blockIdx.x == 0) {   // db2: block-wide deterministic sum
    __shared__ float wp[8];
    float a = 0.f;
    for (int64_t i = threadIdx.x; i < batch; i += blockDim.x) a += dz[i];
    a = warp_sum(a);
    if ((threadIdx.x & 31) == 0) wp[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int w = 0; w < 8; ++w) t += wp[w];
      grads[L.off_b2()] = t;
    }
  }
}

static int bwd_splits(int64_t batch, int hidden) {
  int chunks = (hidden + kBwdChunk - 1) / kBwdChunk;
  int64_t tiles = (batch + kBwdRows - 1) / kBwdRows;
  int64_t want = (2 * (int64_t)sm_count() + chunks - 1) / chunks;
  int64_t s = want < tiles ? want : tiles;
  return (int)(s < 1 ? 1 : s);
}

static size_t bwd_smem(int F) {
  return sizeof(float) * ((size_t)kBwdChunk * (F + 1) + (size_t)kBwdRows * F + 2 * (size_t)kBwdRows * kBwdChunk +
                          (size_t)kBwdChunk * (F + 2) + kBwdRows);
}

}  // namespace b200ds

using namespace b200ds;

extern "C" {

int b200ds_mlp_forward_f32(const float* params_dev, const float* x_dev, int64_t batch,
                           int n_features, int hidden, float* logits_dev, uintptr_t stream) {
  B200_REQUIRE(batch >= 0 && n_features >= 1 && hidden >= 1, "mlp_forward_f32: bad shape B=%lld F=%d H=%d",
               (long long)batch, n_features, hidden);
  if (batch == 0) return B200DS_OK;
  B200_REQUIRE(params_dev && x_dev && logits_dev, "mlp_forward_f32: null pointer");
  B200_REQUIRE(n_features <= 1024, "mlp_forward_f32: F <= 1024 on the fp32 path (got %d)", n_features);
  Layout L{n_features, hidden};
  const int warps = 8;
  int64_t want = (batch + warps - 1) / warps;
  int64_t cap = (int64_t)sm_count() * 8;
  int grid = (int)(want > cap ? cap : want);
  size_t smem = sizeof(float) * warps * n_features;
  mlp_forward_f32_kernel<<<grid, warps * 32, smem, as_stream(stream)>>>(params_dev, x_dev, batch, L, logits_dev);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

size_t b200ds_mlp_backward_f32_scratch_bytes(int64_t batch, int n_features, int hidden) {
  if (batch <= 0 || n_features < 1 || hidden < 1) return 0;
  return sizeof(float) * (size_t)bwd_splits(batch, hidden) * hidden * (n_features + 2);
}

int b200ds_mlp_backward_f32(const float* params_dev, const float* x_dev, const float* dlogits_dev,
                            int64_t batch, int n_features, int hidden, float* grads_dev,
                            void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  B200_REQUIRE(batch > 0 && n_features >= 1 && hidden >= 1, "mlp_backward_f32: bad shape B=%lld F=%d H=%d",
               (long long)batch, n_features, hidden);
  B200_REQUIRE(n_features <= 256, "mlp_backward_f32: F <= 256 on the fp32 path (got %d); use the bf16 path", n_features);
  B200_REQUIRE(params_dev && x_dev && dlogits_dev && grads_dev, "mlp_backward_f32: null pointer");
  size_t need = b200ds_mlp_backward_f32_scratch_bytes(batch, n_features, hidden);
  B200_REQUIRE(scratch_dev && scratch_bytes >= need, "mlp_backward_f32: scratch too small (%zu < %zu)", scratch_bytes, need);
  Layout L{n_features, hidden};
  int splits = bwd_splits(batch, hidden);
  dim3 grid((hidden + kBwdChunk - 1) / kBwdChunk, splits);
  size_t smem = bwd_smem(n_features);
  static bool attr_set[64];
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (smem > 48 * 1024 && !attr_set[dev & 63]) {
    B200_CUDA(cudaFuncSetAttribute(mlp_backward_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set[dev & 63] = true;
  }
  mlp_backward_f32_kernel<<<grid, kBwdThreads, smem, as_stream(stream)>>>(
      params_dev, x_dev, dlogits_dev, batch, L, reinterpret_cast<float*>(scratch_dev));
  B200_LAUNCH_CHECK();
  int64_t total = (int64_t)hidden * (n_features + 2);
  int fgrid = (int)((total + 255) / 256);
  if (fgrid > sm_count() * 4) fgrid = sm_count() * 4;
  mlp_backward_f32_finish<<<fgrid, 256, 0, as_stream(stream)>>>(reinterpret_cast<const float*>(scratch_dev), splits,
                                                                dlogits_dev, batch, L, grads_dev);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

}  // extern "C"
