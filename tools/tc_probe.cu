// tc_probe: validates, on a real B200, every tcgen05/TMEM/TMA building block the
// tensor-core kernels rely on, one stage at a time, against a host reference:
//   1. TMA bulk copy of pre-tiled operands + SS MMA with no-swizzle K-major descriptors
//   2. in-place fp32 -> relu bf16 conversion in TMEM, TS MMA (A from TMEM) against a plain
//      bf16 vector read through an overlapping-core-matrix descriptor (layer 2, N = 16)
//   3. the transposed SS MMA used by the backward
//   4. mask conversion + TS MMA against the SAME x tile re-read as an MN-major operand
//   5. tcgen05.ld throughput (cycles per 32x32 column block per warp)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tc_probe tools/tc_probe.cu
#include "../distributed-gpu-test_b200/csrc/tc_common.cuh"
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

using namespace b200ds::tc;

constexpr int KP = 32;
constexpr uint32_t TINY2 = 0x33803380u;   // bf16x2 (2^-24, 2^-24)

struct ProbeOut {
  float d1[128 * 128];
  float z[128];
  float z16[128 * 16];
  float d1t[128 * 128];
  float s[128 * 32];
  long long ld_cycles[4];
  long long mma_cycles[8];
  long long mma2[16];
};

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __nv_bfloat16* __restrict__ xq, const __nv_bfloat16* __restrict__ wq,
             const __nv_bfloat16* __restrict__ w2v, int w2_mode, ProbeOut* out) {
  __shared__ __align__(128) unsigned char xs[128 * KP * 2];
  __shared__ __align__(128) unsigned char ws[128 * KP * 2];
  __shared__ __align__(128) unsigned char w2s[4096 + 1024];
  __shared__ __align__(8) uint64_t bar_tma, bar_mma;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;

  if (warp == 0) tmem_alloc<512>(&tmem_slot);
  if (tid == 0) {
    mbar_init(&bar_tma, 1);
    mbar_init(&bar_mma, 1);
    mbar_fence_init();
  }
  for (int i = tid; i < (4096 + 1024) / 4; i += 128) reinterpret_cast<uint32_t*>(w2s)[i] = 0;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (w2_mode == 1) {   // explicit K-major [16 rows][128 k], row 0 = w2: SBO = 16*128, LBO = 128
    if (tid < 128) {
      int k = tid;
      reinterpret_cast<__nv_bfloat16*>(w2s)[(k / 8) * 64 + (k % 8)] = w2v[k];
    }
    fence_proxy_async_smem();
  }
  __syncthreads();
  if (tid == 0) {
    uint32_t bytes = 8192 + 8192 + (w2_mode == 0 ? 256 : 0);
    mbar_arrive_expect_tx(&bar_tma, bytes);
    tma_load_1d(xs, xq, 8192, &bar_tma);
    tma_load_1d(ws, wq, 8192, &bar_tma);
    if (w2_mode == 0) tma_load_1d(w2s, w2v, 256, &bar_tma);
  }
  mbar_wait(&bar_tma, 0);

  const uint32_t SBO_K = KP * 16, LBO_K = 128;
  uint32_t mma_phase = 0;
  // ---- 1. D1[b,n] = X . W^T   (A = x tile, B = w tile, both K-major), cols [0,128)
  if (tid == 0) {
    tc_fence_after();
    for (int k = 0; k < KP / 16; ++k)
      mma_ss(tmem + 0, smem_desc(smem_u32(xs) + k * 2 * LBO_K, LBO_K, SBO_K),
             smem_desc(smem_u32(ws) + k * 2 * LBO_K, LBO_K, SBO_K), idesc_bf16(128, 128, 0, 0), k > 0);
    mma_commit(&bar_mma);
  }
  mbar_wait(&bar_mma, mma_phase); mma_phase ^= 1;
  tc_fence_after();
  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  const int row = warp * 32 + (tid & 31);
  long long t0 = clock64();
  for (int j = 0; j < 4; ++j) {
    uint32_t r[32];
    tmem_ld32(tmem + lane_base + j * 32, r);
    tmem_wait_ld();
    for (int c = 0; c < 32; ++c) out->d1[row * 128 + j * 32 + c] = __uint_as_float(r[c]);
    uint32_t p[16];
    for (int c = 0; c < 16; ++c) p[c] = pack_relu_bf16x2(__uint_as_float(r[2 * c]), __uint_as_float(r[2 * c + 1]));
    tmem_st16(tmem + lane_base + j * 16, p);
  }
  tmem_wait_st();
  tc_fence_before();
  __syncthreads();
  // ---- 2. z = relu(D1) . w2 : TS MMA, N = 16, cols [384,400)
  if (tid == 0) {
    tc_fence_after();
    for (int s = 0; s < 8; ++s) {
      uint64_t db = (w2_mode == 0) ? smem_desc(smem_u32(w2s) + s * 32, 16, 16)
                                   : smem_desc(smem_u32(w2s) + s * 2 * 128, 128, 16 * 128);
      mma_ts(tmem + 384, tmem + s * 8, db, idesc_bf16(128, 16, 0, 0), s > 0);
    }
    mma_commit(&bar_mma);
  }
  mbar_wait(&bar_mma, mma_phase); mma_phase ^= 1;
  tc_fence_after();
  {
    uint32_t r[32];
    tmem_ld32(tmem + lane_base + 384, r);
    tmem_wait_ld();
    out->z[row] = __uint_as_float(r[0]);
    for (int c = 0; c < 16; ++c) out->z16[row * 16 + c] = __uint_as_float(r[c]);
  }
  tc_fence_before();
  __syncthreads();
  // ---- 3. D1T[n,b] = W . X^T, cols [128,256)
  if (tid == 0) {
    tc_fence_after();
    for (int k = 0; k < KP / 16; ++k)
      mma_ss(tmem + 128, smem_desc(smem_u32(ws) + k * 2 * LBO_K, LBO_K, SBO_K),
             smem_desc(smem_u32(xs) + k * 2 * LBO_K, LBO_K, SBO_K), idesc_bf16(128, 128, 0, 0), k > 0);
    mma_commit(&bar_mma);
  }
  mbar_wait(&bar_mma, mma_phase); mma_phase ^= 1;
  tc_fence_after();
  for (int j = 0; j < 4; ++j) {
    uint32_t r[32];
    tmem_ld32(tmem + lane_base + 128 + j * 32, r);
    tmem_wait_ld();
    for (int c = 0; c < 32; ++c) out->d1t[row * 128 + j * 32 + c] = __uint_as_float(r[c]);
    uint32_t p[16];
    for (int c = 0; c < 16; ++c)
      p[c] = min_bf16x2(pack_relu_bf16x2(__uint_as_float(r[2 * c]), __uint_as_float(r[2 * c + 1])), TINY2);
    tmem_st16(tmem + lane_base + 128 + j * 16, p);
  }
  tmem_wait_st();
  tc_fence_before();
  __syncthreads();
  // ---- 4. S[n,f] = sum_b mask[n,b] x[b,f] : TS MMA, B = the x tile read MN-major, N = 32, cols [416,448)
  if (tid == 0) {
    tc_fence_after();
    for (int s = 0; s < 8; ++s)
      mma_ts(tmem + 416, tmem + 128 + s * 8, smem_desc(smem_u32(xs) + s * 2 * SBO_K, /*LBO: K groups*/ SBO_K, /*SBO: N groups*/ LBO_K),
             idesc_bf16(128, 32, 0, 1), s > 0);
    mma_commit(&bar_mma);
  }
  mbar_wait(&bar_mma, mma_phase); mma_phase ^= 1;
  tc_fence_after();
  {
    uint32_t r[32];
    tmem_ld32(tmem + lane_base + 416, r);
    tmem_wait_ld();
    for (int c = 0; c < 32; ++c) out->s[row * 32 + c] = __uint_as_float(r[c]) * 16777216.f;
  }
  // ---- 5. tcgen05.ld throughput: 64 back-to-back x32 loads per warp
  __syncthreads();
  long long t1 = clock64();
  uint32_t acc = 0;
  for (int it = 0; it < 16; ++it) {
    for (int j = 0; j < 4; ++j) {
      uint32_t r[32];
      tmem_ld32(tmem + lane_base + j * 32, r);
      tmem_wait_ld();
      acc ^= r[it & 31];
    }
  }
  long long t2 = clock64();
  if ((tid & 31) == 0) out->ld_cycles[warp] = (t2 - t1) + (acc == 0x12345678 ? 1 : 0) + 0 * (t1 - t0);
  // ---- 5b. two x32 loads in flight per warp
  __syncthreads();
  t1 = clock64();
  for (int it = 0; it < 16; ++it) {
    for (int j = 0; j < 4; j += 2) {
      uint32_t r[32], q[32];
      tmem_ld32(tmem + lane_base + j * 32, r);
      tmem_ld32(tmem + lane_base + j * 32 + 32, q);
      tmem_wait_ld();
      acc ^= r[it & 31] ^ q[it & 31];
    }
  }
  t2 = clock64();
  if ((tid & 31) == 0 && warp == 0) out->mma_cycles[7] = (t2 - t1) + (acc == 0x12345678 ? 1 : 0);
  tc_fence_before();
  __syncthreads();
  // ---- 6. MMA instruction cost: 64 back-to-back instructions of each kind, issue -> completion
  for (int kind = 0; kind < 6; ++kind) {
    long long ta = 0;
    if (tid == 0) {
      tc_fence_after();
      ta = clock64();
      for (int i = 0; i < 64; ++i) {
        int k = i & 1, s8 = i & 7;
        if (kind == 0)        // SS M128 N128 K16 (forward layer 1, 128-wide chunk)
          mma_ss(tmem + 0, smem_desc(smem_u32(xs) + k * 2 * LBO_K, LBO_K, SBO_K), smem_desc(smem_u32(ws) + k * 2 * LBO_K, LBO_K, SBO_K), idesc_bf16(128, 128, 0, 0), 1);
        else if (kind == 1)   // SS M128 N64
          mma_ss(tmem + 0, smem_desc(smem_u32(xs) + k * 2 * LBO_K, LBO_K, SBO_K), smem_desc(smem_u32(ws) + k * 2 * LBO_K, LBO_K, SBO_K), idesc_bf16(128, 64, 0, 0), 1);
        else if (kind == 2)   // TS M128 N16 (layer 2)
          mma_ts(tmem + 384, tmem + 128 + s8 * 8, smem_desc(smem_u32(w2s) + s8 * 32, 16, 16), idesc_bf16(128, 16, 0, 0), 1);
        else if (kind == 3)   // TS M128 N32 MN-major B (backward S accumulation)
          mma_ts(tmem + 416, tmem + 128 + s8 * 8, smem_desc(smem_u32(xs) + s8 * 2 * SBO_K, SBO_K, LBO_K), idesc_bf16(128, 32, 0, 1), 1);
        else if (kind == 4)   // SS M128 N256
          mma_ss(tmem + 0, smem_desc(smem_u32(xs) + k * 2 * LBO_K, LBO_K, SBO_K), smem_desc(smem_u32(ws) + k * 2 * LBO_K, LBO_K, 0 /* all row groups alias group 0: stays inside ws */), idesc_bf16(128, 256, 0, 0), 1);
        else                  // SS M128 N32
          mma_ss(tmem + 0, smem_desc(smem_u32(xs) + k * 2 * LBO_K, LBO_K, SBO_K), smem_desc(smem_u32(ws) + k * 2 * LBO_K, LBO_K, SBO_K), idesc_bf16(128, 32, 0, 0), 1);
      }
      mma_commit(&bar_mma);
    }
    mbar_wait(&bar_mma, mma_phase); mma_phase ^= 1;
    tc_fence_after();
    if (tid == 0) out->mma_cycles[kind] = clock64() - ta;
    tc_fence_before();
    __syncthreads();
  }
  // ---- 8. lean issue loop: descriptors precomputed, unrolled x8, invariant operands
  for (int kind = 0; kind < 4; ++kind) {
    long long ta = 0, tb = 0;
    if (tid == 0) {
      tc_fence_after();
      const uint64_t da0 = smem_desc(smem_u32(xs), LBO_K, SBO_K), da1 = smem_desc(smem_u32(xs) + 2 * LBO_K, LBO_K, SBO_K);
      const uint64_t db0 = smem_desc(smem_u32(ws), LBO_K, SBO_K), db1 = smem_desc(smem_u32(ws) + 2 * LBO_K, LBO_K, SBO_K);
      const uint64_t dw = smem_desc(smem_u32(w2s), 16, 16);
      const uint64_t dg = smem_desc(smem_u32(xs), SBO_K, LBO_K);
      ta = clock64();
#pragma unroll 1
      for (int i = 0; i < 8; ++i) {
        if (kind == 0) {
#pragma unroll
          for (int u = 0; u < 4; ++u) { mma_ss(tmem + u * 128, da0, db0, idesc_bf16(128, 128, 0, 0), 1); mma_ss(tmem + u * 128, da1, db1, idesc_bf16(128, 128, 0, 0), 1); }
        } else if (kind == 1) {
#pragma unroll
          for (int u = 0; u < 4; ++u) { mma_ss(tmem + u * 64, da0, db0, idesc_bf16(128, 64, 0, 0), 1); mma_ss(tmem + u * 64, da1, db1, idesc_bf16(128, 64, 0, 0), 1); }
        } else if (kind == 2) {
#pragma unroll
          for (int u = 0; u < 8; ++u) mma_ts(tmem + 384, tmem + 128 + u * 8, dw + (uint64_t)(u * 2), idesc_bf16(128, 16, 0, 0), 1);
        } else {
#pragma unroll
          for (int u = 0; u < 8; ++u) mma_ts(tmem + 416, tmem + 128 + u * 8, dg + (uint64_t)(u * 64), idesc_bf16(128, 32, 0, 1), 1);
        }
      }
      tb = clock64();
      mma_commit(&bar_mma);
    }
    mbar_wait(&bar_mma, mma_phase); mma_phase ^= 1;
    tc_fence_after();
    if (tid == 0) { out->mma2[10 + kind] = clock64() - ta; out->mma_cycles[6] = tb - ta; if (kind == 2) out->ld_cycles[1] = tb - ta; if (kind == 0) out->ld_cycles[2] = tb - ta; if (kind == 3) out->ld_cycles[3] = tb - ta; }
    tc_fence_before();
    __syncthreads();
  }
  // ---- 9. two warps issuing concurrently (lane 0 of warps 0 and 1), 64 lean TS N32 each
  {
    __shared__ __align__(8) uint64_t bar2;
    if (tid == 0) { mbar_init(&bar2, 2); mbar_fence_init(); }
    __syncthreads();
    long long ta = clock64();
    if (tid == 0 || tid == 32) {
      tc_fence_after();
      const uint64_t dg = smem_desc(smem_u32(xs), SBO_K, LBO_K);
      const uint32_t dcol = tid == 0 ? 416u : 448u;
#pragma unroll 1
      for (int i = 0; i < 8; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) mma_ts(tmem + dcol, tmem + 128 + u * 8, dg + (uint64_t)(u * 64), idesc_bf16(128, 32, 0, 1), 1);
      }
      mma_commit(&bar2);
    }
    mbar_wait(&bar2, 0);
    tc_fence_after();
    if (tid == 0) out->mma2[14] = clock64() - ta;
    tc_fence_before();
    __syncthreads();
  }
  // ---- 7. same, but round-robin over 4 independent accumulator tiles; also issue-only time
  for (int kind = 0; kind < 5; ++kind) {
    long long ta = 0, tb = 0;
    if (tid == 0) {
      tc_fence_after();
      ta = clock64();
      for (int i = 0; i < 64; ++i) {
        int k = i & 1, s8 = i & 7, d = i & 3;
        if (kind == 0)        // SS M128 N128, 4 accumulators at cols 0,128,256,384
          mma_ss(tmem + d * 128, smem_desc(smem_u32(xs) + k * 2 * LBO_K, LBO_K, SBO_K), smem_desc(smem_u32(ws) + k * 2 * LBO_K, LBO_K, SBO_K), idesc_bf16(128, 128, 0, 0), 1);
        else if (kind == 1)   // SS M128 N64
          mma_ss(tmem + d * 64, smem_desc(smem_u32(xs) + k * 2 * LBO_K, LBO_K, SBO_K), smem_desc(smem_u32(ws) + k * 2 * LBO_K, LBO_K, SBO_K), idesc_bf16(128, 64, 0, 0), 1);
        else if (kind == 2)   // TS M128 N16
          mma_ts(tmem + 384 + d * 16, tmem + 128 + s8 * 8, smem_desc(smem_u32(w2s) + s8 * 32, 16, 16), idesc_bf16(128, 16, 0, 0), 1);
        else if (kind == 3)   // TS M128 N32 MN-major
          mma_ts(tmem + 384 + d * 32, tmem + 128 + s8 * 8, smem_desc(smem_u32(xs) + s8 * 2 * SBO_K, SBO_K, LBO_K), idesc_bf16(128, 32, 0, 1), 1);
        else                  // SS M64 N128 (half-height)
          mma_ss(tmem + d * 128, smem_desc(smem_u32(xs) + k * 2 * LBO_K, LBO_K, SBO_K), smem_desc(smem_u32(ws) + k * 2 * LBO_K, LBO_K, SBO_K), idesc_bf16(64, 128, 0, 0), 1);
      }
      tb = clock64();
      mma_commit(&bar_mma);
    }
    mbar_wait(&bar_mma, mma_phase); mma_phase ^= 1;
    tc_fence_after();
    if (tid == 0) { out->mma2[2 * kind] = clock64() - ta; out->mma2[2 * kind + 1] = tb - ta; }
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) tmem_dealloc<512>(tmem);
}

static float bf(float v) { return __bfloat162float(__float2bfloat16(v)); }

int main() {
  std::vector<float> x(128 * KP), w(128 * KP), w2(128);
  srand(1);
  auto rnd = []() { return (float)rand() / RAND_MAX * 2.f - 1.f; };
  for (int r = 0; r < 128; ++r)
    for (int k = 0; k < KP; ++k) {
      x[r * KP + k] = k < 20 ? bf(rnd()) : (k < 22 ? 1.f : 0.f);
      w[r * KP + k] = k < 22 ? bf(rnd() * 0.3f) : 0.f;
    }
  for (int k = 0; k < 128; ++k) w2[k] = bf(rnd() * 0.2f);
  auto tile = [](const std::vector<float>& m) {
    std::vector<__nv_bfloat16> t(128 * KP);
    for (int r = 0; r < 128; ++r)
      for (int k = 0; k < KP; ++k)
        t[((r / 8) * (KP * 16) + (k / 8) * 128 + (r % 8) * 16 + (k % 8) * 2) / 2] = __float2bfloat16(m[r * KP + k]);
    return t;
  };
  auto xt = tile(x), wt = tile(w);
  std::vector<__nv_bfloat16> w2b(128);
  for (int k = 0; k < 128; ++k) w2b[k] = __float2bfloat16(w2[k]);
  __nv_bfloat16 *dx, *dw, *dw2;
  ProbeOut* dout;
  cudaMalloc(&dx, 8192); cudaMalloc(&dw, 8192); cudaMalloc(&dw2, 256); cudaMalloc(&dout, sizeof(ProbeOut));
  cudaMemcpy(dx, xt.data(), 8192, cudaMemcpyHostToDevice);
  cudaMemcpy(dw, wt.data(), 8192, cudaMemcpyHostToDevice);
  cudaMemcpy(dw2, w2b.data(), 256, cudaMemcpyHostToDevice);
  // host reference
  std::vector<double> d1(128 * 128), z(128, 0.0), s(128 * 32, 0.0);
  for (int b = 0; b < 128; ++b)
    for (int n = 0; n < 128; ++n) {
      double a = 0;
      for (int k = 0; k < KP; ++k) a += (double)x[b * KP + k] * w[n * KP + k];
      d1[b * 128 + n] = a;
    }
  for (int b = 0; b < 128; ++b)
    for (int n = 0; n < 128; ++n) z[b] += (double)bf((float)std::fmax(d1[b * 128 + n], 0.0)) * w2[n];
  for (int n = 0; n < 128; ++n)
    for (int b = 0; b < 128; ++b)
      if (d1[b * 128 + n] > 0)
        for (int f = 0; f < 32; ++f) s[n * 32 + f] += x[b * KP + f];
  int bad = 0;
  for (int mode = 0; mode < 2; ++mode) {
    cudaMemset(dout, 0xFF, sizeof(ProbeOut));
    probe_kernel<<<1, 128>>>(dx, dw, dw2, mode, dout);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("mode %d: CUDA error %s\n", mode, cudaGetErrorString(e)); return 2; }
    std::vector<ProbeOut> hov(1);
    ProbeOut& ho = hov[0];
    cudaMemcpy(&ho, dout, sizeof(ProbeOut), cudaMemcpyDeviceToHost);
    double e1 = 0, e1t = 0, ez = 0, es = 0, zmax = 0, smax = 0;
    for (int b = 0; b < 128; ++b)
      for (int n = 0; n < 128; ++n) {
        e1 = std::fmax(e1, std::fabs(ho.d1[b * 128 + n] - d1[b * 128 + n]));
        e1t = std::fmax(e1t, std::fabs(ho.d1t[n * 128 + b] - d1[b * 128 + n]));
      }
    for (int b = 0; b < 128; ++b) { ez = std::fmax(ez, std::fabs(ho.z[b] - z[b])); zmax = std::fmax(zmax, std::fabs(z[b])); }
    for (int i = 0; i < 128 * 32; ++i) { es = std::fmax(es, std::fabs(ho.s[i] - s[i])); smax = std::fmax(smax, std::fabs(s[i])); }
    printf("mode %d (w2 %s): |D1 err| %.3g  |D1T err| %.3g  |z err| %.3g (max |z| %.3g)  |S err| %.3g (max |S| %.3g)\n",
           mode, mode == 0 ? "overlapping-LBO vector" : "explicit K-major tile", e1, e1t, ez, zmax, es, smax);
    printf("   z[0..3] gpu %.5f %.5f %.5f %.5f  ref %.5f %.5f %.5f %.5f | z16 row0: %.4f %.4f %.4f\n", ho.z[0], ho.z[1], ho.z[2], ho.z[3],
           z[0], z[1], z[2], z[3], ho.z16[0], ho.z16[1], ho.z16[2]);
    printf("   S[0,0..3] gpu %.4f %.4f %.4f %.4f ref %.4f %.4f %.4f %.4f ; S[5,20..21] gpu %.3f %.3f ref %.3f %.3f\n", ho.s[0], ho.s[1], ho.s[2], ho.s[3],
           s[0], s[1], s[2], s[3], ho.s[5 * 32 + 20], ho.s[5 * 32 + 21], s[5 * 32 + 20], s[5 * 32 + 21]);
    printf("   tcgen05.ld: %lld cycles for 64 x (32 lanes x 32 cols) per warp, 4 warps concurrent -> %.1f B/cycle/SM\n",
           ho.ld_cycles[0], 4.0 * 64 * 4096 / (double)ho.ld_cycles[0]);
    printf("   tcgen05.ld, 2 in flight: %lld cycles for the same -> %.1f B/cycle/SM\n", ho.mma_cycles[7], 4.0 * 64 * 4096 / (double)ho.mma_cycles[7]);
    const char* kn[6] = {"SS M128 N128 K16", "SS M128 N64 K16", "TS M128 N16 K16", "TS M128 N32 K16 (MN-major B)", "SS M128 N256 K16", "SS M128 N32 K16"};
    for (int k = 0; k < 6; ++k) printf("   64 x %-30s: %lld cycles -> %.1f cycles/instr\n", kn[k], ho.mma_cycles[k], ho.mma_cycles[k] / 64.0);
    const char* kn2[5] = {"SS M128 N128 (4 accumulators)", "SS M128 N64 (4 acc)", "TS M128 N16 (4 acc)", "TS M128 N32 MN (4 acc)", "SS M64 N128 (4 acc)"};
    for (int k = 0; k < 5; ++k) printf("   64 x %-30s: %lld cycles total (%.1f/instr), issue loop alone %lld (%.1f/instr)\n", kn2[k], ho.mma2[2 * k], ho.mma2[2 * k] / 64.0, ho.mma2[2 * k + 1], ho.mma2[2 * k + 1] / 64.0);
    const char* kn3[4] = {"lean SS M128 N128", "lean SS M128 N64", "lean TS M128 N16", "lean TS M128 N32 MN"};
    for (int k = 0; k < 4; ++k) printf("   64 x %-30s: %lld cycles total (%.1f/instr)\n", kn3[k], ho.mma2[10 + k], ho.mma2[10 + k] / 64.0);
    printf("   lean issue-only: N128 %lld, N16 %lld, N32MN %lld cycles for 64\n", ho.ld_cycles[2], ho.ld_cycles[1], ho.ld_cycles[3]);
    printf("   2 warps x 64 lean TS N32 concurrently: %lld cycles (one warp alone: %lld)\n", ho.mma2[14], ho.mma2[13]);
    if (e1 > 1e-3 || e1t > 1e-3) bad |= 1;
    if (ez > 2e-2 * std::fmax(1.0, zmax)) bad |= (mode == 0 ? 2 : 4);
    if (es > 1e-2 * std::fmax(1.0, smax)) bad |= 8;
  }
  printf("probe result bits: %d (0 = all good; 1 SS, 2 TS+overlap trick, 4 TS explicit, 8 MN-major reuse)\n", bad);
  return 0;
}
