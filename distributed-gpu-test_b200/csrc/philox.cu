// On-device synthetic data: Philox4x32-10 + Box-Muller with the element mapping of
// torch.randn(device='cuda'), the label rule of train.py:22-23, and the batch gather that
// replaces TensorDataset/DataLoader collation + .to(device) (train.py:65-74,104-106).
//
// The integer stream follows curand_philox4x32_x.h:149-195 (10 rounds, Weyl key bump) and
// curand_init(seed, subsequence = global thread, offset); the float transform follows
// _curand_box_muller (curand_normal.h:70-88).  The launch geometry follows ATen's
// calc_execution_policy / distribution_elementwise_grid_stride_kernel
// (ATen/native/cuda/DistributionTemplates.h:50-92) because it decides which counter each
// output element sees.  Do NOT build this file with --use_fast_math or -fmad=false: the
// bit pattern of logf and the FMA contraction below are part of the contract.
#include "common.cuh"
#include "philox.cuh"

namespace b200ds {

// One thread = one Philox subsequence, exactly as ATen launches it.
__global__ void __launch_bounds__(256) philox_normal_kernel(float* __restrict__ out, int64_t numel,
                                                            uint2 key, uint64_t ctr0, int calls,
                                                            const unsigned long long* __restrict__ step_counter) {
  // graph-replayable form: the draw index lives in device memory and every draw advances the
  // Philox counter by `calls`, exactly like consecutive torch.randn calls on one generator
  if (step_counter) ctr0 += (uint64_t)(*step_counter) * (uint64_t)calls;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (int c = 0; c < calls; ++c) {
    uint64_t ctr = ctr0 + (uint64_t)c;
    uint4 r = philox4x32_10(make_uint4((unsigned)ctr, (unsigned)(ctr >> 32), (unsigned)tid, (unsigned)((uint64_t)tid >> 32)), key);
    float2 p = box_muller(r.x, r.y), q = box_muller(r.z, r.w);
    float v[4] = {p.x, p.y, q.x, q.y};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      int64_t li = tid + nthreads * (4 * (int64_t)c + k);
      if (li < numel) out[li] = normal_transform(v[k]);
    }
  }
}

__global__ void label_rule_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t batch, int F) {
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  const float* row = x + b * F;
  float acc = 0.f;
  for (int j = 0; j < F / 2; ++j) acc += row[j];
  y[b] = acc > 0.f ? 1.f : 0.f;
}

// x_out[i,:] = x_all[idx[i],:], y_out[i] = y_all[idx[i]]; one warp per row.
__global__ void gather_batch_kernel(const float* __restrict__ x_all, const float* __restrict__ y_all,
                                    const int64_t* __restrict__ idx, int64_t batch, int F,
                                    float* __restrict__ x_out, float* __restrict__ y_out) {
  int lane = threadIdx.x & 31;
  int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= batch) return;
  int64_t src = idx[row];
  for (int j = lane; j < F; j += 32) x_out[row * F + j] = x_all[src * F + j];
  if (lane == 0) y_out[row] = y_all[src];
}

void philox_plan(int64_t numel, int* grid, int* calls) {
  int64_t blocks = (numel + 255) / 256;
  int64_t cap = (int64_t)sm_count() * (2048 / 256);   // maxThreadsPerMultiProcessor / block
  int g = (int)(blocks < cap ? blocks : cap);
  *grid = g;
  *calls = (int)((numel - 1) / ((int64_t)g * 256 * 4) + 1);
}

}  // namespace b200ds

using namespace b200ds;

extern "C" {

int b200ds_philox_normal(float* x_dev, int64_t numel, uint64_t seed, uint64_t offset,
                         uint64_t* offset_advance, uintptr_t stream) {
  B200_REQUIRE(numel >= 0, "philox_normal: negative numel");
  B200_REQUIRE(offset % 4 == 0, "philox_normal: offset must be a multiple of 4 (torch keeps it so)");
  if (offset_advance) *offset_advance = 0;
  if (numel == 0) return B200DS_OK;
  B200_REQUIRE(x_dev != nullptr, "philox_normal: null output");
  int grid, calls;
  philox_plan(numel, &grid, &calls);
  philox_normal_kernel<<<grid, 256, 0, as_stream(stream)>>>(
      x_dev, numel, make_uint2((unsigned)seed, (unsigned)(seed >> 32)), offset / 4, calls, nullptr);
  B200_LAUNCH_CHECK();
  if (offset_advance) *offset_advance = (uint64_t)calls * 4;
  return B200DS_OK;
}

int b200ds_philox_normal_counted(float* x_dev, int64_t numel, uint64_t seed, uint64_t offset0,
                                 const uint64_t* draw_counter_dev, uintptr_t stream) {
  B200_REQUIRE(numel > 0 && x_dev != nullptr, "philox_normal_counted: empty or null output");
  B200_REQUIRE(offset0 % 4 == 0, "philox_normal_counted: offset0 must be a multiple of 4");
  B200_REQUIRE(draw_counter_dev != nullptr, "philox_normal_counted: null draw counter");
  int grid, calls;
  philox_plan(numel, &grid, &calls);
  philox_normal_kernel<<<grid, 256, 0, as_stream(stream)>>>(
      x_dev, numel, make_uint2((unsigned)seed, (unsigned)(seed >> 32)), offset0 / 4, calls,
      reinterpret_cast<const unsigned long long*>(draw_counter_dev));
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

int b200ds_label_rule(const float* x_dev, float* y_dev, int64_t batch, int n_features, uintptr_t stream) {
  B200_REQUIRE(batch >= 0 && n_features >= 1, "label_rule: bad shape");
  if (batch == 0) return B200DS_OK;
  B200_REQUIRE(x_dev && y_dev, "label_rule: null pointer");
  label_rule_kernel<<<(unsigned)((batch + 255) / 256), 256, 0, as_stream(stream)>>>(x_dev, y_dev, batch, n_features);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

int b200ds_gather_batch(const float* x_all_dev, const float* y_all_dev, const int64_t* idx_dev,
                        int64_t batch, int n_features, float* x_out_dev, float* y_out_dev,
                        uintptr_t stream) {
  B200_REQUIRE(batch >= 0 && n_features >= 1, "gather_batch: bad shape");
  if (batch == 0) return B200DS_OK;
  B200_REQUIRE(x_all_dev && y_all_dev && idx_dev && x_out_dev && y_out_dev, "gather_batch: null pointer");
  int64_t threads = batch * 32;
  gather_batch_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, as_stream(stream)>>>(
      x_all_dev, y_all_dev, idx_dev, batch, n_features, x_out_dev, y_out_dev);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

}  // extern "C"
