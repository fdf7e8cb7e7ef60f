// tc_probe2: MMA stream throughput of the forward kernel's exact operand pattern, one CTA:
// 128 KB W1 slice (16 tiles) + x tile in dynamic smem, 2 MMAs (K=32) per 128x128 chunk,
// ring of 4 accumulators, no epilogue.  Prints cycles per chunk for several variants.
#include "../distributed-gpu-test_b200/csrc/tc_common.cuh"
#include <cstdio>
using namespace b200ds::tc;

__global__ void __launch_bounds__(128, 1) k(long long* out, int variant, int nchunks) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  unsigned char* w1s = smem;                // 16 x 8 KB
  unsigned char* xs = smem + 16 * 8192;     // 8 KB
  for (int i = tid; i < (17 * 8192) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == 0) tmem_alloc<512>(&slot);
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  long long t0 = 0;
  if (warp == 1) {
    const uint32_t idesc = idesc_bf16(128, 128, 0, 0);
    const uint64_t xdesc0 = smem_desc(smem_u32(xs), 128, 512);
    const uint64_t wdesc0 = smem_desc(smem_u32(w1s), 128, 512);
    t0 = clock64();
    for (int q = 0; q < nchunks; ++q) {
      const int c = (variant & 1) ? 0 : (q & 15);           // variant bit0: always the same W1 tile
      const int r = (variant & 2) ? 0 : (q & 3);            // variant bit1: always the same accumulator
      const uint64_t wd = wdesc0 + (uint64_t)((c * 8192) >> 4);
      const uint32_t dcol = tmem + r * 128;
      if (elect_one()) {
        if (variant & 4) {   // x as A, w as B
          mma_ss(dcol, xdesc0, wd, idesc, 0);
          mma_ss(dcol, xdesc0 + 16, wd + 16, idesc, 1);
        } else {
          mma_ss(dcol, wd, xdesc0, idesc, 0);
          mma_ss(dcol, wd + 16, xdesc0 + 16, idesc, 1);
        }
      }
      __syncwarp();
    }
    long long t1 = clock64();
    if (elect_one()) mma_commit(&bar);
    __syncwarp();
    if (tid == 32) out[1] = t1 - t0;
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  if (tid == 32) out[0] = clock64() - t0;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

int main() {
  long long* d; cudaMalloc(&d, 64);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 17 * 8192 + 1024);
  const char* names[8] = {"16 W1 tiles, ring 4, A=w", "same W1 tile, ring 4, A=w", "16 tiles, same acc, A=w", "same tile, same acc, A=w",
                          "16 W1 tiles, ring 4, A=x", "same W1 tile, ring 4, A=x", "16 tiles, same acc, A=x", "same tile, same acc, A=x"};
  for (int rep = 0; rep < 2; ++rep)
    for (int v = 0; v < 8; ++v) {
      k<<<1, 128, 17 * 8192 + 1024>>>(d, v, 112);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
      long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      if (rep) printf("%-28s: 112 chunks (224 MMAs) in %lld cycles = %.1f / chunk (issue loop alone %.1f / chunk)\n", names[v], h[0], h[0] / 112.0, h[1] / 112.0);
    }
  return 0;
}
