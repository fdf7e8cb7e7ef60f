// tc_probe8: do the tensor core and the epilogue's tcgen05.ld compete inside TMEM?  One CTA:
//   an MMA issuer streams chunks of 2 x (M128 N128 K16) (the forward's layer-1 chunk, K = 32) into columns 256..511,
//   free-running (one commit at the very end), while the epilogue loop of tc_probe7 (tcgen05.ld + one FFMA per element)
//   reads columns 0..255.  Reported: cycles per 128 x 128 chunk of each side, alone and together.
#include "../distributed-gpu-test_b200/csrc/tc_common.cuh"
#include <cstdio>
#include <cstdlib>
using namespace b200ds::tc;

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}
constexpr int NCH = 384;

// EW epilogue warps (8: 64 columns each, 32-column pieces; 16: 32 columns each, 16-column pieces), 1 load in flight.
// warp EW = MMA issuer.  run_mma / run_epi select the sides.  mma_n: N of the MMA (128, or 256 = two chunks per instruction pair)
template <int EW>
__global__ void __launch_bounds__(EW * 32 + 32, 1) k(long long* out, float* sink, int run_mma, int run_epi, int accum) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t endbar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < (2 * 8192) / 4; i += EW * 32 + 32) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (warp == EW) tmem_alloc<512>(&slot);
  if (tid == 0) { mbar_init(&endbar, 1); mbar_fence_init(); }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == EW) {
    if (run_mma) {
      const uint32_t idesc = idesc_bf16(128, 128, 0, 0);
      const uint64_t xd = smem_desc(smem_u32(smem + 8192), 128, 512);
      const uint64_t wd = smem_desc(smem_u32(smem), 128, 512);
      long long t0 = clock64();
      for (int q = 0; q < NCH; ++q) {
        const uint32_t dcol = tmem + 256 + (q & 1) * 128;
        if (elect_one()) { mma_ss(dcol, wd, xd, idesc, 0); mma_ss(dcol, wd + 16, xd + 16, idesc, accum); }
        __syncwarp();
      }
      if (elect_one()) mma_commit(&endbar);
      __syncwarp();
      mbar_wait(&endbar, 0);
      tc_fence_after();
      if (lane == 0) out[32] = clock64() - t0;
    }
  } else if (run_epi) {
    constexpr int COLS = EW == 8 ? 64 : 32, PIECE = EW == 8 ? 32 : 16, PPC = COLS / PIECE;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const int cgrp = warp >> 2;
    float acc[COLS];
#pragma unroll
    for (int i = 0; i < COLS; ++i) acc[i] = 0.f;
    uint32_t buf[2][PIECE];
    float w = 0.001f * lane;
    auto issue = [&](int p, uint32_t (&dst)[PIECE]) {
      const int chunk = p / PPC, pc = p % PPC;
      const uint32_t a = tmem + lane_base + (uint32_t)((chunk & 1) * 128 + cgrp * COLS + pc * PIECE);
      if constexpr (PIECE == 32) tmem_ld32(a, dst); else tmem_ld16(a, dst);
    };
    const int NP = NCH * PPC;
    long long t0 = clock64();
    issue(0, buf[0]);
    for (int p0 = 0; p0 < NP; p0 += 2) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int p = p0 + u;
        tmem_wait_ld();
        if (p + 1 < NP) issue(p + 1, buf[(u + 1) & 1]);
        const int pc = u % PPC;
#pragma unroll
        for (int i = 0; i < PIECE; ++i) acc[pc * PIECE + i] = fmaf(fabsf(__uint_as_float(buf[u][i])), w, acc[pc * PIECE + i]);
      }
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < COLS; ++i) s += acc[i];
    if (lane == 0) out[warp] = t1 - t0;
    sink[tid] = s;
  }
  tc_fence_before(); __syncthreads();
  if (warp == EW) tmem_dealloc<512>(tmem);
}

int main() {
  long long* d; cudaMalloc(&d, 64 * 8); float* sink; cudaMalloc(&sink, 4096);
  long long h[64];
  cudaFuncSetAttribute(k<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 8192 + 1024);
  cudaFuncSetAttribute(k<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 8192 + 1024);
  auto run = [&](const char* name, int ew, int mma, int epi, int accum) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaMemset(d, 0, 64 * 8);
      if (ew == 8) k<8><<<1, 8 * 32 + 32, 2 * 8192 + 1024>>>(d, sink, mma, epi, accum);
      else k<16><<<1, 16 * 32 + 32, 2 * 8192 + 1024>>>(d, sink, mma, epi, accum);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s: error %s\n", name, cudaGetErrorString(e)); exit(1); }
    }
    cudaMemcpy(h, d, 64 * 8, cudaMemcpyDeviceToHost);
    long long m = 0; for (int i = 0; i < ew; ++i) if (h[i] > m) m = h[i];
    printf("%-64s  MMA %7.1f   epilogue %7.1f   cycles / chunk\n", name, mma ? h[32] / (double)NCH : 0.0, epi ? m / (double)NCH : 0.0);
  };
  run("MMA alone (K = 32: 2 instructions per chunk, 2nd accumulates)", 8, 1, 0, 1);
  run("MMA alone, 2nd instruction overwrites too (no TMEM read-back)", 8, 1, 0, 0);
  run("epilogue alone, 8 warps x 64 columns", 8, 0, 1, 1);
  run("both, 8 epilogue warps", 8, 1, 1, 1);
  run("both, 8 epilogue warps, MMAs without accumulate", 8, 1, 1, 0);
  run("epilogue alone, 16 warps x 32 columns (16-col pieces)", 16, 0, 1, 1);
  run("both, 16 epilogue warps", 16, 1, 1, 1);
  return 0;
}
