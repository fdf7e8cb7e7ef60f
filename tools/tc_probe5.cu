// tc_probe5: what does the per-chunk handshake cost the tcgen05.mma stream?  One CTA, chunks of
// 2 x (M128 N128 K16) into a ring of 4 accumulators.
//   mode 0: one issuer, no per-chunk commit (lean stream)
//   mode 1: one issuer, tcgen05.commit per chunk to a ring of mbarriers nobody waits on
//   mode 2: one issuer + full/free handshake with 4 consumer warps that only wait and arrive
//   mode 3: as 2, two issuers on alternate chunks
//   mode 4: as 3, consumers also tcgen05.ld their 128 columns before they release
//   mode 5: as 4, one commit per TWO chunks (ring of 2 x 256 columns)
#include "../distributed-gpu-test_b200/csrc/tc_common.cuh"
#include <cstdio>
#include <cstdlib>
using namespace b200ds::tc;
constexpr int NQ = 256;

__global__ void __launch_bounds__(320, 1) k(long long* out, int mode) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t full[4], freeb[4], endbar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* w1s = smem;
  unsigned char* xs = smem + 16 * 8192;
  for (int i = tid; i < (17 * 8192) / 4; i += 320) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == 9) tmem_alloc<512>(&slot);
  const int n_cons = 4;
  if (tid == 0) {
    for (int i = 0; i < 4; ++i) { mbar_init(&full[i], 1); mbar_init(&freeb[i], n_cons); }
    mbar_init(&endbar, mode == 3 || mode == 4 || mode == 5 ? 2 : 1);
    mbar_fence_init();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  const int group = mode == 5 ? 2 : 1;           // chunks per handshake
  const int ring = 4 / group;
  const int n_issuers = (mode >= 3) ? 2 : 1;
  if (warp == 9 || (warp == 8 && n_issuers == 2)) {
    const int me = warp == 9 ? 0 : 1;
    const uint32_t idesc = idesc_bf16(128, 128, 0, 0);
    const uint64_t xdesc0 = smem_desc(smem_u32(xs), 128, 512);
    const uint64_t wdesc0 = smem_desc(smem_u32(w1s), 128, 512);
    long long t0 = clock64();
    for (int h = 0; h < NQ / group; ++h) {       // handshake unit
      if (n_issuers == 2 && (h & 1) != me) continue;
      const int r = h % ring; const uint32_t ph = (uint32_t)(h / ring) & 1u;
      if (mode >= 2) { mbar_wait(&freeb[r], ph ^ 1); tc_fence_after(); }
      const int q = h * group;
      const uint64_t wd = wdesc0 + (uint64_t)(((q & 15) * 8192) >> 4);
      const uint32_t dcol = tmem + (q & 3) * 128;
      if (group == 1) {
        if (mode >= 1) {
          if (elect_one()) { mma_ss(dcol, wd, xdesc0, idesc, 0); mma_ss(dcol, wd + 16, xdesc0 + 16, idesc, 1); mma_commit(&full[r]); }
        } else {
          if (elect_one()) { mma_ss(dcol, wd, xdesc0, idesc, 0); mma_ss(dcol, wd + 16, xdesc0 + 16, idesc, 1); }
        }
      } else {
        if (elect_one()) {
          mma_ss(dcol, wd, xdesc0, idesc, 0); mma_ss(dcol, wd + 16, xdesc0 + 16, idesc, 1);
          mma_ss(dcol + 128, wd + 512, xdesc0, idesc, 0); mma_ss(dcol + 128, wd + 512 + 16, xdesc0 + 16, idesc, 1);
          mma_commit(&full[r]);
        }
      }
      __syncwarp();
    }
    if (elect_one()) mma_commit(&endbar);
    __syncwarp();
    mbar_wait(&endbar, 0);
    tc_fence_after();
    if (lane == 0 && me == 0) out[0] = clock64() - t0;
  } else if (warp < 4 && mode >= 2) {
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    uint32_t va[32], keep = 0;
    for (int h = 0; h < NQ / group; ++h) {
      const int r = h % ring; const uint32_t ph = (uint32_t)(h / ring) & 1u;
      mbar_wait(&full[r], ph);
      tc_fence_after();
      if (mode >= 4) {
        for (int c = 0; c < 4 * group; ++c) { tmem_ld32(tmem + lane_base + r * 128 * group + c * 32, va); tmem_wait_ld(); keep ^= va[3]; }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&freeb[r]);
    }
    if (keep == 0x12345u) out[2] = keep;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<512>(tmem);
}

int main() {
  long long* d; cudaMalloc(&d, 64);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 17 * 8192 + 1024);
  const char* names[6] = {"lean stream, 1 issuer", "+ commit per chunk", "+ full/free handshake, idle consumers", "  same, 2 issuers",
                          "  2 issuers, consumers tcgen05.ld", "  same, one handshake per 2 chunks"};
  for (int rep = 0; rep < 2; ++rep)
    for (int m = 0; m < 6; ++m) {
      k<<<(getenv("GRID") ? atoi(getenv("GRID")) : 1), 320, 17 * 8192 + 1024>>>(d, m);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("mode %d error %s\n", m, cudaGetErrorString(e)); return 1; }
      long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      if (rep) printf("%-44s: %d chunks in %lld cycles = %.1f / chunk\n", names[m], NQ, h[0], h[0] / (double)NQ);
    }
  return 0;
}
