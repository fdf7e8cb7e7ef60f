// tc_probe6: tcgen05.st throughput (registers -> TMEM), one CTA.
//   warps 0..W-1 each store x16 (32 lanes x 16 columns x 4 B = 2 KB) NST times between waits.
#include "../distributed-gpu-test_b200/csrc/tc_common.cuh"
#include <cstdio>
#include <cstdlib>
using namespace b200ds::tc;

__global__ void __launch_bounds__(256, 1) k(long long* out, int nwarps, int per_wait, int use_ld) {
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  uint32_t p[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) p[i] = tid * 16 + i;
  uint32_t v[32];
  const uint32_t base = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 256;
  __syncthreads();
  long long t0 = clock64();
  const int iters = 512;
  if (warp < nwarps) {
    for (int it = 0; it < iters; ++it) {
      if (use_ld) {
        for (int j = 0; j < per_wait; ++j) tmem_ld32(base + ((it * per_wait + j) & 7) * 32, v);
        tmem_wait_ld();
        p[0] ^= v[5];
      } else {
        for (int j = 0; j < per_wait; ++j) tmem_st16(base + ((it * per_wait + j) & 15) * 16, p);
        tmem_wait_st();
      }
    }
  }
  long long t1 = clock64();
  if (tid == 0) { out[0] = t1 - t0; out[1] = p[0]; }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

int main() {
  long long* d; cudaMalloc(&d, 64);
  for (int use_ld = 0; use_ld < 2; ++use_ld)
    for (int nw = 4; nw <= 8; nw += 4)
      for (int pw = 1; pw <= 4; pw *= 2) {
        for (int rep = 0; rep < 2; ++rep) k<<<1, 256>>>(d, nw, pw, use_ld);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
        long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        const double bytes = 512.0 * pw * nw * (use_ld ? 4096 : 2048);
        printf("%s warps %d, %d per wait: %lld cycles, %.1f B/clk, %.1f cycles per wait round\n", use_ld ? "ld.x32" : "st.x16", nw, pw,
               h[0], bytes / h[0], h[0] / 512.0);
      }
  return 0;
}
