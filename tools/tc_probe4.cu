// tc_probe4: does CUDA-core / TMEM-load activity on the same SM slow the tcgen05.mma stream?
// warp 9 issues 224 MMAs (M128 N128 K16, ring of 4 accumulators, no waits); warps 0-7 run, in a
// loop until the MMAs are done, one of: nothing | FFMA only | tcgen05.ld only | both.
#include "../distributed-gpu-test_b200/csrc/tc_common.cuh"
#include <cstdio>
#include <cstdlib>
using namespace b200ds::tc;

__global__ void __launch_bounds__(320, 1) k(long long* out, int mode, float* sink, int wmask) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  __shared__ volatile int done;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* w1s = smem;
  unsigned char* xs = smem + 16 * 8192;
  for (int i = tid; i < (17 * 8192) / 4; i += 320) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == 9) tmem_alloc<512>(&slot);
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); done = 0; }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 9) {
    const uint32_t idesc = idesc_bf16(128, 128, 0, 0);
    const uint64_t xdesc0 = smem_desc(smem_u32(xs), 128, 512);
    const uint64_t wdesc0 = smem_desc(smem_u32(w1s), 128, 512);
    long long t0 = clock64();
    for (int q = 0; q < 112; ++q) {
      const uint64_t wd = wdesc0 + (uint64_t)(((q & 15) * 8192) >> 4);
      const uint32_t dcol = tmem + (q & 3) * 128;
      if (elect_one()) { mma_ss(dcol, wd, xdesc0, idesc, 0); mma_ss(dcol, wd + 16, xdesc0 + 16, idesc, 1); }
      __syncwarp();
    }
    if (elect_one()) mma_commit(&bar);
    __syncwarp();
    mbar_wait(&bar, 0);
    tc_fence_after();
    if (lane == 0) { out[0] = clock64() - t0; done = 1; }
  } else if (warp < 8 && ((wmask >> warp) & 1)) {
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    float acc[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) acc[i] = 0.f;
    uint32_t va[32], vb[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) { va[i] = i + tid; vb[i] = i * 3 + tid; }
    float w = 1.0f + tid;
    long long iters = 0;
    while (!done) {
      if (mode & 2) {
        tmem_ld32(tmem + lane_base + (warp >> 2) * 64, va);
        tmem_ld32(tmem + lane_base + (warp >> 2) * 64 + 32, vb);
        tmem_wait_ld();
      }
      if (mode & 1) {
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[i] = fmaf(fabsf(__uint_as_float(va[i])), w, acc[i]);
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[32 + i] = fmaf(fabsf(__uint_as_float(vb[i])), w, acc[32 + i]);
      }
      ++iters;
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 64; ++i) s += acc[i];
    if (s == 123.456f) sink[0] = s;
    if (tid == 0) out[1] = iters;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<512>(tmem);
}

int main() {
  long long* d; float* sink; cudaMalloc(&d, 64); cudaMalloc(&sink, 64);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 17 * 8192 + 1024);
  const char* names[4] = {"epilogue warps idle-spin", "FFMA only", "tcgen05.ld only", "tcgen05.ld + FFMA"};
  for (int rep = 0; rep < 2; ++rep)
    for (int m = 0; m < 4; ++m) {
      k<<<(getenv("GRID") ? atoi(getenv("GRID")) : 1), 320, 17 * 8192 + 1024>>>(d, m, sink, getenv("WMASK") ? (int)strtol(getenv("WMASK"), 0, 16) : 0xff);
      if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
      long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      if (rep) printf("%-26s: 112 chunks in %lld cycles = %.1f / chunk; epilogue loop iterations %lld (%.1f cycles each)\n", names[m], h[0], h[0] / 112.0, h[1], h[1] ? (double)h[0] / h[1] : 0.0);
    }
  return 0;
}
