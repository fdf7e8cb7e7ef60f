// tc_probe3: handshake latencies of the MMA <-> epilogue loop, one CTA, empty pipe.
//   L1: MMA issue (2 x M128 N128 K16) + tcgen05.commit  ->  another warp's mbarrier wait returns
//   L2: that warp's mbarrier.arrive -> the issuing warp's wait returns
#include "../distributed-gpu-test_b200/csrc/tc_common.cuh"
#include <cstdio>
using namespace b200ds::tc;

__global__ void __launch_bounds__(128, 1) k(long long* out, int n_mma) {
  __shared__ __align__(1024) unsigned char xs[8192];
  __shared__ __align__(1024) unsigned char ws[8192];
  __shared__ __align__(8) uint64_t full, freeb;
  __shared__ uint32_t slot;
  __shared__ long long stamp[64][4];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 2048; i += 128) { reinterpret_cast<uint32_t*>(xs)[i] = 0x3c003c00u; reinterpret_cast<uint32_t*>(ws)[i] = 0x3c003c00u; }
  if (warp == 0) tmem_alloc<512>(&slot);
  if (tid == 0) { mbar_init(&full, 1); mbar_init(&freeb, 1); mbar_fence_init(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  const uint32_t idesc = idesc_bf16(128, 128, 0, 0);
  const uint64_t xd = smem_desc(smem_u32(xs), 128, 512), wd = smem_desc(smem_u32(ws), 128, 512);
  const long long base = clock64();
  if (warp == 1) {
    uint32_t ph = 0;
    for (int it = 0; it < 32; ++it) {
      long long t0 = clock64();
      if (elect_one()) {
        for (int m = 0; m < n_mma; ++m) mma_ss(tmem, wd, xd, idesc, m > 0);
        mma_commit(&full);
      }
      __syncwarp();
      long long t1 = clock64();
      mbar_wait(&freeb, ph);
      long long t2 = clock64();
      if (lane == 0) { stamp[it][0] = t0 - base; stamp[it][1] = t1 - base; stamp[it][3] = t2 - base; }
      ph ^= 1;
    }
  } else if (warp == 2) {
    uint32_t ph = 0;
    for (int it = 0; it < 32; ++it) {
      mbar_wait(&full, ph);
      long long t = clock64();
      tc_fence_after();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&freeb); stamp[it][2] = t - base; }
      ph ^= 1;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) for (int i = 0; i < 32; ++i) for (int j = 0; j < 4; ++j) out[i * 4 + j] = stamp[i][j];
  if (warp == 0) tmem_dealloc<512>(tmem);
}

int main() {
  long long* d; cudaMalloc(&d, 128 * 8);
  for (int n = 0; n <= 4; n += 2) {
    k<<<1, 128>>>(d, n); k<<<1, 128>>>(d, n);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
    long long h[128]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double a = 0, b = 0, c = 0;
    for (int i = 8; i < 32; ++i) { a += h[i * 4 + 1] - h[i * 4 + 0]; b += h[i * 4 + 2] - h[i * 4 + 0]; c += h[i * 4 + 3] - h[i * 4 + 2]; }
    printf("%d MMAs per round: issue+commit %.0f cycles; issue start -> consumer woke %.0f; consumer woke -> issuer woke (arrive round trip) %.0f; full loop %.0f\n",
           n, a / 24, b / 24, c / 24, (double)(h[31 * 4] - h[8 * 4]) / 23);
  }
  return 0;
}
