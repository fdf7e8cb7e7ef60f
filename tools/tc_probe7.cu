// tc_probe7: what can the CUDA-core side of the forward epilogue sustain, in isolation?
//   A  tcgen05.ld 32x32b.x32 alone: back-to-back throughput and ld -> wait::ld latency, for 1 / 4 / 8 / 16 warps
//   B  FFMA with an |.| source (acc += |v| * w) alone, 3-register form, per SM sub-partition; and fma.rn.f32x2
//   C  the epilogue loop of mlp_fwd_tc_kernel without the MMA side: per 128 x 128 fp32 chunk every element is read
//      from TMEM once and goes through one FFMA; variants: 8 warps x 64 columns or 16 warps x 32 columns,
//      1 or 2 or 3 loads in flight per warp, 32- or 16-column pieces.
// cycles are per 128 x 128 chunk for the whole CTA (the MMA floor of that chunk at K = 32 is 128 cycles).
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

constexpr int NCH = 384;       // chunks per measurement

// mode A: ld only.  nw warps; each warp owns lanes 32*(w%4); `dep` = wait after every load
__global__ void __launch_bounds__(512, 1) kA(long long* out, int nw, int dep) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = slot;
  if (warp < nw) {
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    uint32_t v[32], keep = 0;
    long long t0 = clock64();
    for (int i = 0; i < NCH * 4; ++i) {        // NCH*4 loads of 32 columns per warp
      tmem_ld32(tmem + lane_base + ((i * 32) & 511), v);
      if (dep) { tmem_wait_ld(); keep ^= v[5]; }
    }
    tmem_wait_ld();
    keep ^= v[7];
    long long t1 = clock64();
    if (lane == 0) out[warp] = t1 - t0;
    if (keep == 0x1234567u) out[63] = keep;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

// mode B: FFMA |v| * w + acc, 64 accumulators, values from registers.  pk = 1: fma.rn.f32x2 (no abs)
__global__ void __launch_bounds__(512, 1) kB(long long* out, float* sink, int nw, int pk) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp >= nw) return;
  float acc[64], v[32];
  for (int i = 0; i < 64; ++i) acc[i] = 0.f;
  for (int i = 0; i < 32; ++i) v[i] = (float)(lane + i) - 16.5f;
  float w = 0.001f * lane;
  long long t0 = clock64();
  if (!pk) {
    for (int it = 0; it < 1024; ++it) {
#pragma unroll
      for (int i = 0; i < 32; ++i) acc[i] = fmaf(fabsf(v[i]), w, acc[i]);
#pragma unroll
      for (int i = 0; i < 32; ++i) acc[32 + i] = fmaf(fabsf(v[i]), w, acc[32 + i]);
      w += 1e-6f;
    }
  } else {
    for (int it = 0; it < 1024; ++it) {
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        asm volatile("{.reg .b64 a, b, c; mov.b64 a, {%2, %3}; mov.b64 b, {%4, %4}; mov.b64 c, {%0, %1}; fma.rn.f32x2 c, a, b, c; mov.b64 {%0, %1}, c;}"
                     : "+f"(acc[i]), "+f"(acc[i + 1]) : "f"(v[i]), "f"(v[i + 1]), "f"(w));
      }
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        asm volatile("{.reg .b64 a, b, c; mov.b64 a, {%2, %3}; mov.b64 b, {%4, %4}; mov.b64 c, {%0, %1}; fma.rn.f32x2 c, a, b, c; mov.b64 {%0, %1}, c;}"
                     : "+f"(acc[32 + i]), "+f"(acc[33 + i]) : "f"(v[i]), "f"(v[i + 1]), "f"(w));
      }
      w += 1e-6f;
    }
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < 64; ++i) s += acc[i];
  if (lane == 0) out[warp] = t1 - t0;
  sink[threadIdx.x] = s;
}

// mode C: epilogue loop.  COLS columns per warp per chunk (64: 8 warps, 32: 16 warps), PIECE columns per load,
// DEPTH loads in flight.
constexpr int gcd_(int a, int b) { return b == 0 ? a : gcd_(b, a % b); }
template <int COLS, int PIECE, int DEPTH>
__global__ void __launch_bounds__(4 * (128 / COLS) * 32, 1) kC(long long* out, float* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = slot;
  {
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const int cgrp = warp >> 2;                     // which COLS-wide column group of each chunk
    float acc[COLS];
#pragma unroll
    for (int i = 0; i < COLS; ++i) acc[i] = 0.f;
    constexpr int PPC = COLS / PIECE;               // pieces per chunk per warp
    constexpr int RING = DEPTH + 1;
    constexpr int U = PPC * RING / gcd_(PPC, RING); // unroll: piece-in-chunk and buffer index both static
    uint32_t buf[RING][PIECE];
    float w = 0.001f * lane;
    auto issue = [&](int p, uint32_t (&dst)[PIECE]) {   // piece p of the stream
      const int chunk = p / PPC, pc = p % PPC;
      const uint32_t a = tmem + lane_base + (uint32_t)(((chunk & 3) * 128 + cgrp * COLS + pc * PIECE) & 511);
      if constexpr (PIECE == 32) tmem_ld32(a, dst); else tmem_ld16(a, dst);
    };
    const int NP = NCH * PPC;
    long long t0 = clock64();
#pragma unroll
    for (int d = 0; d < DEPTH; ++d) issue(d, buf[d]);
    for (int p0 = 0; p0 < NP; p0 += U) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int p = p0 + u;
        tmem_wait_ld();                              // everything outstanding (the oldest is piece p)
        if (p + DEPTH < NP) issue(p + DEPTH, buf[(u + DEPTH) % RING]);
        constexpr int dummy = 0; (void)dummy;
        const int pc = u % PPC;                      // static after unrolling
#pragma unroll
        for (int i = 0; i < PIECE; ++i)
          acc[pc * PIECE + i] = fmaf(fabsf(__uint_as_float(buf[u % RING][i])), w, acc[pc * PIECE + i]);
      }
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < COLS; ++i) s += acc[i];
    if (lane == 0) out[warp] = t1 - t0;
    sink[threadIdx.x] = s;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem);
}

static long long maxof(long long* h, int n) { long long m = 0; for (int i = 0; i < n; ++i) if (h[i] > m) m = h[i]; return m; }

int main() {
  long long* d; cudaMalloc(&d, 64 * 8); float* sink; cudaMalloc(&sink, 4096);
  long long h[64];
  auto run = [&](const char* name, int nw, double per, auto launch) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaMemset(d, 0, 64 * 8);
      launch();
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s: error %s\n", name, cudaGetErrorString(e)); exit(1); }
    }
    cudaMemcpy(h, d, 64 * 8, cudaMemcpyDeviceToHost);
    printf("%-72s %8.1f\n", name, maxof(h, nw) / per);
  };
  printf("A  tcgen05.ld x32 (4 KB per warp-load): cycles per load, slowest warp\n");
  for (int nw : {1, 4, 8, 16}) {
    char nm[96]; snprintf(nm, 96, "   %2d warps, back to back (throughput)", nw);
    run(nm, nw, NCH * 4.0, [&] { kA<<<1, 512>>>(d, nw, 0); });
    snprintf(nm, 96, "   %2d warps, wait::ld after each (latency)", nw);
    run(nm, nw, NCH * 4.0, [&] { kA<<<1, 512>>>(d, nw, 1); });
  }
  printf("B  64 FFMA per iteration: cycles per warp-FFMA, slowest warp\n");
  for (int nw : {4, 8, 16}) {
    char nm[96]; snprintf(nm, 96, "   %2d warps, fma |v|*w+acc", nw);
    run(nm, nw, 1024 * 64.0, [&] { kB<<<1, 512>>>(d, sink, nw, 0); });
    snprintf(nm, 96, "   %2d warps, fma.rn.f32x2 (32 per iteration, no abs)", nw);
    run(nm, nw, 1024 * 32.0, [&] { kB<<<1, 512>>>(d, sink, nw, 1); });
  }
  printf("C  epilogue loop (ld + one FFMA per element): cycles per 128x128 chunk, whole CTA\n");
  run("    8 warps x 64 cols, 32-col pieces, 1 load in flight (today)", 8, (double)NCH, [&] { kC<64, 32, 1><<<1, 256>>>(d, sink); });
  run("    8 warps x 64 cols, 32-col pieces, 2 loads in flight", 8, (double)NCH, [&] { kC<64, 32, 2><<<1, 256>>>(d, sink); });
  run("    8 warps x 64 cols, 16-col pieces, 3 loads in flight", 8, (double)NCH, [&] { kC<64, 16, 3><<<1, 256>>>(d, sink); });
  run("   16 warps x 32 cols, 32-col pieces, 1 load in flight", 16, (double)NCH, [&] { kC<32, 32, 1><<<1, 512>>>(d, sink); });
  run("   16 warps x 32 cols, 16-col pieces, 1 load in flight", 16, (double)NCH, [&] { kC<32, 16, 1><<<1, 512>>>(d, sink); });
  run("   16 warps x 32 cols, 16-col pieces, 2 loads in flight", 16, (double)NCH, [&] { kC<32, 16, 2><<<1, 512>>>(d, sink); });
  run("   16 warps x 32 cols, 16-col pieces, 3 loads in flight", 16, (double)NCH, [&] { kC<32, 16, 3><<<1, 512>>>(d, sink); });
  return 0;
}
