// tc_probe9: what does ONE tcgen05.mma (kind::f16, M = 128, K = 16) cost as a function of N, with A from shared memory (SS)
// and with A from TMEM (TS)?  One CTA, one issuing thread, a long stream into alternating accumulators, one commit at
// the end.  The backward's second GEMM is M128 N32 K16 x 8 per chunk with A from TMEM: if an instruction costs ~64 cycles
// whatever N <= 128 (the A operand, 128 x 16, streams at a fixed rate), that GEMM alone is ~512 cycles per chunk.
#include "../distributed-gpu-test_b200/csrc/tc_common.cuh"
#include <cstdio>
#include <cstdlib>
using namespace b200ds::tc;
constexpr int NI = 2048;

__global__ void __launch_bounds__(64, 1) k(long long* out, int N, int ts, int b_mn, int alt) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t endbar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < (48 * 1024) / 4; i += 64) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == 0) tmem_alloc<512>(&slot);
  if (tid == 0) { mbar_init(&endbar, 1); mbar_fence_init(); }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 1) {
    const uint32_t idesc = idesc_bf16(128, N, 0, b_mn);
    const uint64_t ad = smem_desc(smem_u32(smem), 128, 512);
    const uint64_t bd = b_mn ? smem_desc(smem_u32(smem + 16384), 512, 128) : smem_desc(smem_u32(smem + 16384), 128, 512);
    long long t0 = clock64();
    for (int i = 0; i < NI; i += 8) {                                         // eight instructions per elected block, as the kernels issue them
      const uint32_t dcol = tmem + ((i >> 3) & 1) * 256;
      if (elect_one()) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const uint32_t dc = alt ? tmem + (u & 1) * 256 : dcol;               // alt: consecutive instructions accumulate into different tiles
          if (ts) mma_ts(dc, tmem + 448 + u * 8, bd, idesc, 1u);              // A: 8 TMEM columns (16 bf16) per instruction
          else mma_ss(dc, ad + (uint64_t)((u & 1) * 16), bd, idesc, 1u);
        }
      }
      __syncwarp();
    }
    if (elect_one()) mma_commit(&endbar);
    __syncwarp();
    mbar_wait(&endbar, 0);
    tc_fence_after();
    if (lane == 0) out[0] = clock64() - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

int main() {
  long long* d; cudaMalloc(&d, 64);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 49 * 1024);
  printf("tcgen05.mma kind::f16 M128 K16: cycles per instruction (stream of %d, one issuing thread)\n", NI);
  for (int alt = 0; alt < 2; ++alt)
  for (int ts = 0; ts < 2; ++ts)
    for (int N : {16, 32, 64, 128, 256}) {
      long long h = 0;
      for (int rep = 0; rep < 2; ++rep) {
        k<<<1, 64, 49 * 1024>>>(d, N, ts, ts, alt);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("N %d ts %d: error %s\n", N, ts, cudaGetErrorString(e)); return 1; }
        cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      }
      printf("  %s A from %-6s N = %3d : %7.1f cycles   (4096 MAC/cycle would be %5.1f)\n", alt ? "alternating accumulators," : "one accumulator (chained),", ts ? "TMEM" : "smem", N, h / (double)NI, 128.0 * N * 16 / 4096);
    }
  return 0;
}
