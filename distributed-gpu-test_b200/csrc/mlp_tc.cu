// bf16 tensor-core path of the model step for the widened model (BASELINE config 4):
// tcgen05.mma with fp32 accumulators in TMEM, operands fed by TMA bulk copies, and the
// rest of each layer fused around the MMA so that the hidden activation [B,H] never
// reaches HBM (materialising it makes the step HBM-bound, 24x slower).
//
// Replaces, for large shapes, model_engine(batch_x) (train.py:107 -> train.py:35-36) and the
// autograd part of model_engine.backward(loss) (train.py:113).
//
// Operands (see tc_common.cuh for the byte layouts):
//   xq  [B,  KP=32] bf16, tiled: x[b,0..F-1], 1, 1, 0...       (two ones columns)
//   w1q [H,  KP=32] bf16, tiled: W1[n,0..F-1], hi(b1[n]), lo(b1[n]), 0...
//   so that  z1[b,n] = sum_k xq[b,k] w1q[n,k] = x.W1[n,:] + b1[n]  comes straight out of the MMA
//   (b1 is carried as a bf16 hi/lo pair: ~16 mantissa bits, for free inside the K padding).
//
// forward   per (128-row b tile, 128-wide n chunk):  D1 = xq_tile . w1q_chunk^T   (SS MMA, K=32)
//           epilogue on CUDA cores straight from TMEM:  relu(a) = (a + |a|)/2, so
//             z[b] = b2 + 1/2 sum_n w2[n] z1[b,n] + 1/2 sum_n w2[n] |z1[b,n]|
//           the first sum is linear in x (= xq[b,:] . v, v = w1q^T w2, 32 numbers per slice),
//           the second is ONE FFMA with an |.| source modifier per accumulator element.
// backward  per (128 n block, 128-row b tile):  D1T = w1q_block . xq_tile^T   (SS MMA, recompute)
//           epilogue: A2[n,b] = 1[z1>0] * dz[b] as bf16, written back IN PLACE into TMEM,
//           S[n,k] += A2 . xq_tile   (TS MMA, A from TMEM, B = the same x tile read MN-major)
//           then  dW1[n,f] = w2[n] S[n,f],  db1[n] = w2[n] S[n,F],
//                 dW2[n]  = sum_k w1q[n,k] S[n,k]  (= sum_b dz[b] h[b,n]),  db2 = sum_b dz[b].
#include "common.cuh"
#include "tc_common.cuh"

namespace b200ds {
using namespace tc;

constexpr int KP = 32;                    // padded K of layer 1 (F + 2 <= 32)
constexpr int kTileRows = 128;            // rows per operand tile (b tile / n chunk)
constexpr uint32_t kTileBytes = kTileRows * KP * 2;      // 8192
constexpr uint32_t LBO_K = 128, SBO_K = KP * 16;          // K-major reading of a tile
constexpr int kMaxSliceRows = 2048;       // W1 rows resident per forward CTA (128 KB)
constexpr int kMaxSlices = 8;

__device__ __forceinline__ size_t tiled_byte(int64_t row, int k) {
  return (size_t)(row >> 3) * SBO_K + (size_t)(k >> 3) * LBO_K + (size_t)(row & 7) * 16 + (size_t)(k & 7) * 2;
}

// ------------------------------------------------------------------ packing
// thread -> (row group, k chunk, row in group): consecutive threads write consecutive 16 B
__global__ void __launch_bounds__(256)
pack_x_kernel(const float* __restrict__ x, int64_t batch, int F, uint4* __restrict__ xq) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= batch * (KP / 8)) return;
  int r8 = (int)(t & 7), c = (int)((t >> 3) & 3);
  int64_t row = ((t >> 5) << 3) + r8;
  const float* xr = x + row * F;
  __nv_bfloat16 v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int k = c * 8 + j;
    float f = k < F ? xr[k] : (k < F + 2 ? 1.f : 0.f);
    v[j] = __float2bfloat16(f);
  }
  xq[t] = *reinterpret_cast<uint4*>(v);
}

__global__ void __launch_bounds__(256)
pack_w1_kernel(const float* __restrict__ params, Layout L, uint4* __restrict__ w1q) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (int64_t)L.H * (KP / 8)) return;
  int r8 = (int)(t & 7), c = (int)((t >> 3) & 3);
  int n = (int)(((t >> 5) << 3) + r8);
  const float* wr = params + L.off_w1() + (int64_t)n * L.F;
  float b1 = params[L.off_b1() + n];
  __nv_bfloat16 hi = __float2bfloat16(b1);
  __nv_bfloat16 lo = __float2bfloat16(b1 - __bfloat162float(hi));
  __nv_bfloat16 v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int k = c * 8 + j;
    v[j] = k < L.F ? __float2bfloat16(wr[k]) : (k == L.F ? hi : (k == L.F + 1 ? lo : __float2bfloat16(0.f)));
  }
  w1q[t] = *reinterpret_cast<uint4*>(v);
}

// v[s][k] = sum_{n in slice s} w2[n] * w1q[n,k]   (deterministic order), one CTA per slice
__global__ void __launch_bounds__(1024)
lin_vector_kernel(const float* __restrict__ params, Layout L, const unsigned char* __restrict__ w1q, int slice_rows,
                  float* __restrict__ v) {
  __shared__ float part[32][33];
  const int k = threadIdx.x & 31, j = threadIdx.x >> 5;
  const int n0 = blockIdx.x * slice_rows;
  const float* w2 = params + L.off_w2();
  float a = 0.f;
  for (int n = n0 + j; n < n0 + slice_rows; n += 32)
    a = fmaf(w2[n], __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(w1q + tiled_byte(n, k))), a);
  part[j][k] = a;
  __syncthreads();
  if (j == 0) {
    float s = 0.f;
    for (int i = 0; i < 32; ++i) s += part[i][k];
    v[blockIdx.x * 32 + k] = s;
  }
}

// ------------------------------------------------------------------ forward
constexpr int kFwdXStages = 3;
constexpr int kFwdRing = 4;               // D1 accumulators in TMEM: 4 x 128 columns
constexpr int kFwdThreads = 32 * 10;      // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue (two groups)

struct FwdSmem {                           // fixed-size tail after the W1 slice / w2 slice / x stages
  float v[32];
  float zs[2][128];
  uint64_t w_bar, x_full[kFwdXStages], x_free[kFwdXStages], d1_full[kFwdRing], d1_free[kFwdRing];
  uint32_t tmem_slot;
};

__global__ void __launch_bounds__(kFwdThreads, 1)
mlp_fwd_tc_kernel(const unsigned char* __restrict__ xq, const unsigned char* __restrict__ w1q,
                  const float* __restrict__ vlin, const float* __restrict__ w2, const float* __restrict__ b2p,
                  float* __restrict__ logits, int n_btiles, int slice_rows, int n_slices, int ctas_per_slice) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slice = blockIdx.x % n_slices, cta_in_slice = blockIdx.x / n_slices;
  const int NC = slice_rows / kTileRows;                 // chunks per b tile
  unsigned char* w1s = smem;                             // [slice_rows][KP] bf16 tiled
  float* w2s = reinterpret_cast<float*>(w1s + (size_t)slice_rows * KP * 2);       // [slice_rows] fp32
  unsigned char* xs = reinterpret_cast<unsigned char*>(w2s + slice_rows);         // [stages][8192]
  FwdSmem* sh = reinterpret_cast<FwdSmem*>(xs + kFwdXStages * kTileBytes);

  if (threadIdx.x == 0) {
    mbar_init(&sh->w_bar, 1);
    for (int i = 0; i < kFwdXStages; ++i) { mbar_init(&sh->x_full[i], 1); mbar_init(&sh->x_free[i], 1); }
    for (int i = 0; i < kFwdRing; ++i) { mbar_init(&sh->d1_full[i], 1); mbar_init(&sh->d1_free[i], 4); }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<512>(&sh->tmem_slot);
  for (int i = threadIdx.x; i < slice_rows; i += blockDim.x) w2s[i] = w2[slice * slice_rows + i];
  if (threadIdx.x < 32) sh->v[threadIdx.x] = vlin[slice * 32 + threadIdx.x];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      const unsigned char* wsrc = w1q + (size_t)slice * slice_rows * KP * 2;
      mbar_arrive_expect_tx(&sh->w_bar, (uint32_t)slice_rows * KP * 2);
      for (int c = 0; c < NC; ++c) tma_load_1d(w1s + (size_t)c * kTileBytes, wsrc + (size_t)c * kTileBytes, kTileBytes, &sh->w_bar);
      int s = 0; uint32_t ph = 0;
      for (int t = cta_in_slice; t < n_btiles; t += ctas_per_slice) {
        mbar_wait(&sh->x_free[s], ph ^ 1);
        mbar_arrive_expect_tx(&sh->x_full[s], kTileBytes);
        tma_load_1d(xs + s * kTileBytes, xq + (size_t)t * kTileBytes, kTileBytes, &sh->x_full[s]);
        if (++s == kFwdXStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: D1[ring] = x_tile . w1_chunk^T, K = 32 =====
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(128, 128, 0, 0);
      const uint64_t xdesc0 = smem_desc(smem_u32(xs), LBO_K, SBO_K);
      const uint64_t wdesc0 = smem_desc(smem_u32(w1s), LBO_K, SBO_K);
      mbar_wait(&sh->w_bar, 0);
      int s = 0; uint32_t xph = 0; int r = 0; uint32_t rph = 0;
      for (int t = cta_in_slice; t < n_btiles; t += ctas_per_slice) {
        mbar_wait(&sh->x_full[s], xph);
        const uint64_t xd = xdesc0 + (uint64_t)((s * kTileBytes) >> 4);
        for (int c = 0; c < NC; ++c) {
          mbar_wait(&sh->d1_free[r], rph ^ 1);
          tc_fence_after();
          const uint64_t wd = wdesc0 + (uint64_t)((c * kTileBytes) >> 4);
          mma_ss(tmem + r * 128, xd, wd, idesc, 0);
          mma_ss(tmem + r * 128, xd + (2 * LBO_K >> 4), wd + (2 * LBO_K >> 4), idesc, 1);
          mma_commit(&sh->d1_full[r]);
          if (++r == kFwdRing) { r = 0; rph ^= 1; }
        }
        mma_commit(&sh->x_free[s]);
        if (++s == kFwdXStages) { s = 0; xph ^= 1; }
      }
    }
  } else {
    // ===== epilogue: two groups of 4 warps; group g takes the chunks with c % 2 == g =====
    const int ew = warp - 2, grp = ew >> 2, quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const float b2 = slice == 0 ? b2p[0] : 0.f;
    int s = 0; uint32_t xph = 0; int q = 0; int tcount = 0;
    for (int t = cta_in_slice; t < n_btiles; t += ctas_per_slice, ++tcount) {
      float lin = 0.f;
      if (grp == 0) {       // linear half of relu: xq[row,:] . v, from the x tile in smem
        mbar_wait(&sh->x_full[s], xph);
        const unsigned char* xr = xs + s * kTileBytes + (size_t)(row >> 3) * SBO_K + (size_t)(row & 7) * 16;
#pragma unroll
        for (int c = 0; c < KP / 8; ++c) {
          uint4 qv = *reinterpret_cast<const uint4*>(xr + c * LBO_K);
          const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&qv);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float2 f = __bfloat1622float2(h2[j]);
            lin = fmaf(f.x, sh->v[c * 8 + 2 * j], lin);
            lin = fmaf(f.y, sh->v[c * 8 + 2 * j + 1], lin);
          }
        }
      }
      if (++s == kFwdXStages) { s = 0; xph ^= 1; }
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
      for (int c = 0; c < NC; ++c, ++q) {
        if ((q & 1) != grp) continue;
        const int r = q % kFwdRing;
        const uint32_t rph = (uint32_t)(q / kFwdRing) & 1u;
        mbar_wait(&sh->d1_full[r], rph);
        tc_fence_after();
        const float* wc = w2s + c * kTileRows;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t v[32];
          tmem_ld32(tmem + lane_base + r * 128 + j * 32, v);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            float4 w4 = *reinterpret_cast<const float4*>(wc + j * 32 + i);
            a0 = fmaf(fabsf(__uint_as_float(v[i + 0])), w4.x, a0);
            a1 = fmaf(fabsf(__uint_as_float(v[i + 1])), w4.y, a1);
            a2 = fmaf(fabsf(__uint_as_float(v[i + 2])), w4.z, a2);
            a3 = fmaf(fabsf(__uint_as_float(v[i + 3])), w4.w, a3);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh->d1_free[r]);
      }
      const float absum = (a0 + a1) + (a2 + a3);
      float* zs = sh->zs[tcount & 1];
      if (grp == 1) zs[row] = absum;
      asm volatile("bar.sync 1, 256;" ::: "memory");       // the 8 epilogue warps only
      if (grp == 0) {
        float z = 0.5f * (lin + (absum + zs[row])) + b2;
        float* dst = logits + (size_t)t * kTileRows + row;
        if (n_slices == 1) *dst = z;
        else atomicAdd(dst, z);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------ backward
constexpr int kBwdXStages = 4;
constexpr int kBwdRing = 3;               // D1T accumulators: 3 x 128 columns (cols 0..383)
constexpr int kBwdNB = 4;                 // n blocks per CTA: S accumulators 4 x 32 columns (cols 384..511)
constexpr int kBwdThreads = 32 * 12;      // 0 TMA, 1 MMA1, 2 MMA2, 3 dz stager, 4-11 epilogue (two groups)

struct BwdSmem {
  uint32_t dzs[kBwdXStages][64];          // dz of the tile as bf16x2 pairs
  uint64_t w_bar, x_full[kBwdXStages], x_free[kBwdXStages], dz_full[kBwdXStages];
  uint64_t d1_full[kBwdRing], d1_free[kBwdRing], a_ready[kBwdRing], s_full;
  uint32_t tmem_slot;
};

__global__ void __launch_bounds__(kBwdThreads, 1)
mlp_bwd_tc_kernel(const unsigned char* __restrict__ xq, const unsigned char* __restrict__ w1q,
                  const float* __restrict__ dz, float* __restrict__ partial /* [splits][H][32] */,
                  int n_btiles, int hidden, int nb, int n_groups, int n_splits) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int group = blockIdx.x % n_groups, split = blockIdx.x / n_groups;
  unsigned char* w1s = smem;                                           // [nb][8192]
  unsigned char* xs = w1s + (size_t)nb * kTileBytes;                   // [stages][8192]
  BwdSmem* sh = reinterpret_cast<BwdSmem*>(xs + kBwdXStages * kTileBytes);
  // b tiles of this split: contiguous range
  const int per = (n_btiles + n_splits - 1) / n_splits;
  const int t_begin = split * per, t_end = min(n_btiles, t_begin + per);

  if (threadIdx.x == 0) {
    mbar_init(&sh->w_bar, 1);
    for (int i = 0; i < kBwdXStages; ++i) { mbar_init(&sh->x_full[i], 1); mbar_init(&sh->x_free[i], 1); mbar_init(&sh->dz_full[i], 1); }
    for (int i = 0; i < kBwdRing; ++i) { mbar_init(&sh->d1_full[i], 1); mbar_init(&sh->d1_free[i], 1); mbar_init(&sh->a_ready[i], 4); }
    mbar_init(&sh->s_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<512>(&sh->tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_slot;
  const uint32_t s_col0 = kBwdRing * 128;

  if (warp == 0) {
    if (lane == 0) {
      const unsigned char* wsrc = w1q + (size_t)group * nb * kTileBytes;
      mbar_arrive_expect_tx(&sh->w_bar, (uint32_t)nb * kTileBytes);
      for (int i = 0; i < nb; ++i) tma_load_1d(w1s + (size_t)i * kTileBytes, wsrc + (size_t)i * kTileBytes, kTileBytes, &sh->w_bar);
      int s = 0; uint32_t ph = 0;
      for (int t = t_begin; t < t_end; ++t) {
        mbar_wait(&sh->x_free[s], ph ^ 1);
        mbar_arrive_expect_tx(&sh->x_full[s], kTileBytes);
        tma_load_1d(xs + s * kTileBytes, xq + (size_t)t * kTileBytes, kTileBytes, &sh->x_full[s]);
        if (++s == kBwdXStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 3) {
    // ===== dz stager: dz[b0..b0+127] -> bf16x2 pairs in smem =====
    int s = 0; uint32_t ph = 0;
    for (int t = t_begin; t < t_end; ++t) {
      mbar_wait(&sh->x_free[s], ph ^ 1);
      float4 d = *reinterpret_cast<const float4*>(dz + (size_t)t * kTileRows + lane * 4);
      __nv_bfloat162 p0 = __floats2bfloat162_rn(d.x, d.y), p1 = __floats2bfloat162_rn(d.z, d.w);
      sh->dzs[s][lane * 2 + 0] = *reinterpret_cast<uint32_t*>(&p0);
      sh->dzs[s][lane * 2 + 1] = *reinterpret_cast<uint32_t*>(&p1);
      __syncwarp();
      if (lane == 0) mbar_arrive(&sh->dz_full[s]);
      if (++s == kBwdXStages) { s = 0; ph ^= 1; }
    }
  } else if (warp == 1) {
    // ===== MMA1 issuer: D1T[ring] = w1_block . x_tile^T (lanes = n, columns = b) =====
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(128, 128, 0, 0);
      const uint64_t xdesc0 = smem_desc(smem_u32(xs), LBO_K, SBO_K);
      const uint64_t wdesc0 = smem_desc(smem_u32(w1s), LBO_K, SBO_K);
      mbar_wait(&sh->w_bar, 0);
      int s = 0; uint32_t xph = 0; int r = 0; uint32_t rph = 0;
      for (int t = t_begin; t < t_end; ++t) {
        mbar_wait(&sh->x_full[s], xph);
        const uint64_t xd = xdesc0 + (uint64_t)((s * kTileBytes) >> 4);
        for (int i = 0; i < nb; ++i) {
          mbar_wait(&sh->d1_free[r], rph ^ 1);
          tc_fence_after();
          const uint64_t wd = wdesc0 + (uint64_t)((i * kTileBytes) >> 4);
          mma_ss(tmem + r * 128, wd, xd, idesc, 0);
          mma_ss(tmem + r * 128, wd + (2 * LBO_K >> 4), xd + (2 * LBO_K >> 4), idesc, 1);
          mma_commit(&sh->d1_full[r]);
          if (++r == kBwdRing) { r = 0; rph ^= 1; }
        }
        if (++s == kBwdXStages) { s = 0; xph ^= 1; }
      }
    }
  } else if (warp == 2) {
    // ===== MMA2 issuer: S[i] += A2[ring] (TMEM, bf16) . x_tile (MN-major view), K = 128 rows of b =====
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(128, 32, 0, 1);
      const uint64_t gdesc0 = smem_desc(smem_u32(xs), /*LBO = K groups*/ SBO_K, /*SBO = N groups*/ LBO_K);
      int s = 0; int r = 0; uint32_t rph = 0;
      for (int t = t_begin; t < t_end; ++t) {
        const uint64_t gd = gdesc0 + (uint64_t)((s * kTileBytes) >> 4);
        for (int i = 0; i < nb; ++i) {
          mbar_wait(&sh->a_ready[r], rph);
          tc_fence_after();
#pragma unroll
          for (int u = 0; u < 8; ++u)
            mma_ts(tmem + s_col0 + i * 32, tmem + r * 128 + u * 8, gd + (uint64_t)((u * 2 * SBO_K) >> 4), idesc,
                   (t > t_begin || u > 0) ? 1u : 0u);
          mma_commit(&sh->d1_free[r]);
          if (++r == kBwdRing) { r = 0; rph ^= 1; }
        }
        mma_commit(&sh->x_free[s]);
        if (++s == kBwdXStages) s = 0;
      }
      mma_commit(&sh->s_full);
    }
  } else {
    // ===== epilogue: mask * dz -> bf16, in place; chunk q goes to group q % 2 =====
    const int ew = warp - 4, grp = ew >> 2, quad = warp & 3;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    int q = 0;
    int s = 0; uint32_t xph = 0;
    for (int t = t_begin; t < t_end; ++t) {
      bool dz_waited = false;
      for (int i = 0; i < nb; ++i, ++q) {
        if ((q & 1) != grp) continue;
        const int r = q % kBwdRing;
        const uint32_t rph = (uint32_t)(q / kBwdRing) & 1u;
        if (!dz_waited) { mbar_wait(&sh->dz_full[s], xph); dz_waited = true; }
        mbar_wait(&sh->d1_full[r], rph);
        tc_fence_after();
        const uint4* dzp = reinterpret_cast<const uint4*>(sh->dzs[s]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t v[32];
          tmem_ld32(tmem + lane_base + r * 128 + j * 32, v);
          tmem_wait_ld();
          uint32_t p[16];
#pragma unroll
          for (int c4 = 0; c4 < 4; ++c4) {
            uint4 d4 = dzp[j * 4 + c4];
            const uint32_t dd[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int c = c4 * 4 + e;
              uint32_t h = pack_relu_bf16x2(__uint_as_float(v[2 * c]), __uint_as_float(v[2 * c + 1]));
              uint32_t m;
              asm("set.gt.u32.bf16x2 %0, %1, %2;" : "=r"(m) : "r"(h), "r"(0u));
              p[c] = m & dd[e];
            }
          }
          tmem_st16(tmem + lane_base + r * 128 + j * 16, p);
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh->a_ready[r]);
      }
      if (++s == kBwdXStages) { s = 0; xph ^= 1; }
    }
    // ---- drain: S[i] -> partial[split][n][0..31]
    if (grp == 0) {
      mbar_wait(&sh->s_full, 0);
      tc_fence_after();
      const bool any = t_end > t_begin;
      for (int i = 0; i < nb; ++i) {
        uint32_t v[32];
        tmem_ld32(tmem + lane_base + s_col0 + i * 32, v);
        tmem_wait_ld();
        const int n = (group * nb + i) * kTileRows + quad * 32 + lane;
        float4* dst = reinterpret_cast<float4*>(partial + ((size_t)split * hidden + n) * 32);
#pragma unroll
        for (int c = 0; c < 8; ++c)
          dst[c] = any ? make_float4(__uint_as_float(v[4 * c]), __uint_as_float(v[4 * c + 1]), __uint_as_float(v[4 * c + 2]),
                                     __uint_as_float(v[4 * c + 3]))
                       : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem);
}

// grads from S: one warp per hidden unit n, lane = k
__global__ void __launch_bounds__(256)
mlp_bwd_tc_finish(const float* __restrict__ partial, int n_splits, const float* __restrict__ params,
                  const float* __restrict__ dz, int64_t batch, Layout L, float* __restrict__ grads) {
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  for (int n = blockIdx.x * wpb + (threadIdx.x >> 5); n < L.H; n += gridDim.x * wpb) {
    float s = 0.f;
    for (int sp = 0; sp < n_splits; ++sp) s += partial[((size_t)sp * L.H + n) * 32 + lane];
    const float w2 = params[L.off_w2() + n];
    float wq;
    if (lane < L.F) {
      wq = __bfloat162float(__float2bfloat16(params[L.off_w1() + (int64_t)n * L.F + lane]));
      grads[L.off_w1() + (int64_t)n * L.F + lane] = w2 * s;
    } else if (lane == L.F || lane == L.F + 1) {
      float b1 = params[L.off_b1() + n];
      float hi = __bfloat162float(__float2bfloat16(b1));
      wq = lane == L.F ? hi : __bfloat162float(__float2bfloat16(b1 - hi));
      if (lane == L.F) grads[L.off_b1() + n] = w2 * s;
    } else {
      wq = 0.f;
    }
    float d2 = warp_sum(wq * s);
    if (lane == 0) grads[L.off_w2() + n] = d2;
  }
  if (blockIdx.x == 0) {
    __shared__ float wp[8];
    float a = 0.f;
    for (int64_t i = threadIdx.x; i < batch; i += blockDim.x) a += dz[i];
    a = warp_sum(a);
    if (lane == 0) wp[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += wp[w];
      grads[L.off_b2()] = t;
    }
  }
}

// ------------------------------------------------------------------ host side
struct FwdPlan { int slices, slice_rows, ctas_per_slice, n_btiles; size_t smem; };
struct BwdPlan { int nb, groups, splits, n_btiles; size_t smem; };

static bool tc_shape_ok(int64_t batch, int F, int H) {
  return F >= 1 && F + 2 <= KP && H >= 128 && H % 128 == 0 && batch >= 128 && batch % 128 == 0;
}

static bool fwd_plan(int64_t batch, int H, FwdPlan* p) {
  int slices = (H + kMaxSliceRows - 1) / kMaxSliceRows;
  while (slices <= kMaxSlices && (H % slices != 0 || (H / slices) % 128 != 0)) ++slices;
  if (slices > kMaxSlices) return false;
  p->slices = slices;
  p->slice_rows = H / slices;
  p->n_btiles = (int)(batch / kTileRows);
  int cps = sm_count() / slices;
  if (cps < 1) cps = 1;
  if (cps > p->n_btiles) cps = p->n_btiles;
  p->ctas_per_slice = cps;
  p->smem = (size_t)p->slice_rows * KP * 2 + (size_t)p->slice_rows * 4 + kFwdXStages * kTileBytes + sizeof(FwdSmem) + 1024;
  return true;
}

static bool bwd_plan(int64_t batch, int H, BwdPlan* p) {
  int nblocks = H / 128;
  int nb = kBwdNB;
  while (nb > 1 && nblocks % nb != 0) --nb;
  p->nb = nb;
  p->groups = nblocks / nb;
  p->n_btiles = (int)(batch / kTileRows);
  int splits = sm_count() / p->groups;
  if (splits < 1) splits = 1;
  if (splits > p->n_btiles) splits = p->n_btiles;
  p->splits = splits;
  p->smem = (size_t)nb * kTileBytes + kBwdXStages * kTileBytes + sizeof(BwdSmem) + 1024;
  return true;
}

}  // namespace b200ds

using namespace b200ds;

extern "C" {

int b200ds_tc_supported(int64_t batch, int n_features, int hidden) {
  FwdPlan fp;
  return tc_shape_ok(batch, n_features, hidden) && fwd_plan(batch, hidden, &fp) ? 1 : 0;
}

size_t b200ds_tc_xq_bytes(int64_t batch) { return batch > 0 ? (size_t)batch * KP * 2 : 0; }

size_t b200ds_tc_wq_bytes(int hidden) { return hidden > 0 ? (size_t)hidden * KP * 2 + kMaxSlices * 32 * sizeof(float) : 0; }

int b200ds_pack_x_bf16(const float* x_dev, int64_t batch, int n_features, void* xq_dev, uintptr_t stream) {
  B200_REQUIRE(batch > 0 && batch % 8 == 0, "pack_x_bf16: batch must be a positive multiple of 8 (got %lld)", (long long)batch);
  B200_REQUIRE(n_features >= 1 && n_features + 2 <= KP, "pack_x_bf16: F must be in [1,%d] (got %d)", KP - 2, n_features);
  B200_REQUIRE(x_dev && xq_dev, "pack_x_bf16: null pointer");
  int64_t threads = batch * (KP / 8);
  pack_x_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, as_stream(stream)>>>(x_dev, batch, n_features, reinterpret_cast<uint4*>(xq_dev));
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

int b200ds_pack_params_bf16(const float* params_dev, int n_features, int hidden, void* wq_dev, uintptr_t stream) {
  B200_REQUIRE(n_features >= 1 && n_features + 2 <= KP, "pack_params_bf16: F must be in [1,%d] (got %d)", KP - 2, n_features);
  B200_REQUIRE(hidden >= 128 && hidden % 128 == 0, "pack_params_bf16: H must be a multiple of 128 (got %d)", hidden);
  B200_REQUIRE(params_dev && wq_dev, "pack_params_bf16: null pointer");
  FwdPlan fp;
  B200_REQUIRE(fwd_plan(128, hidden, &fp), "pack_params_bf16: H = %d cannot be sliced into <= %d resident slices", hidden, kMaxSlices);
  Layout L{n_features, hidden};
  int64_t threads = (int64_t)hidden * (KP / 8);
  pack_w1_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, as_stream(stream)>>>(params_dev, L, reinterpret_cast<uint4*>(wq_dev));
  B200_LAUNCH_CHECK();
  unsigned char* w1q = reinterpret_cast<unsigned char*>(wq_dev);
  float* v = reinterpret_cast<float*>(w1q + (size_t)hidden * KP * 2);
  lin_vector_kernel<<<fp.slices, 1024, 0, as_stream(stream)>>>(params_dev, L, w1q, fp.slice_rows, v);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

int b200ds_mlp_forward_tc(const void* xq_dev, const void* wq_dev, const float* params_dev, int64_t batch,
                          int n_features, int hidden, float* logits_dev, uintptr_t stream) {
  B200_REQUIRE(tc_shape_ok(batch, n_features, hidden),
               "mlp_forward_tc: needs F <= %d, H %% 128 == 0, batch %% 128 == 0 (got B=%lld F=%d H=%d)", KP - 2,
               (long long)batch, n_features, hidden);
  B200_REQUIRE(xq_dev && wq_dev && params_dev && logits_dev, "mlp_forward_tc: null pointer");
  FwdPlan fp;
  B200_REQUIRE(fwd_plan(batch, hidden, &fp), "mlp_forward_tc: H = %d cannot be sliced", hidden);
  Layout L{n_features, hidden};
  cudaStream_t st = as_stream(stream);
  static bool attr_set[64];
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B200_CUDA(cudaFuncSetAttribute(mlp_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  if (fp.slices > 1) B200_CUDA(cudaMemsetAsync(logits_dev, 0, sizeof(float) * batch, st));
  const unsigned char* w1q = reinterpret_cast<const unsigned char*>(wq_dev);
  const float* v = reinterpret_cast<const float*>(w1q + (size_t)hidden * KP * 2);
  mlp_fwd_tc_kernel<<<fp.slices * fp.ctas_per_slice, kFwdThreads, fp.smem, st>>>(
      reinterpret_cast<const unsigned char*>(xq_dev), w1q, v, params_dev + L.off_w2(), params_dev + L.off_b2(), logits_dev,
      fp.n_btiles, fp.slice_rows, fp.slices, fp.ctas_per_slice);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

size_t b200ds_mlp_backward_tc_scratch_bytes(int64_t batch, int n_features, int hidden) {
  if (!tc_shape_ok(batch, n_features, hidden)) return 0;
  BwdPlan bp;
  bwd_plan(batch, hidden, &bp);
  return (size_t)bp.splits * hidden * 32 * sizeof(float);
}

int b200ds_mlp_backward_tc(const void* xq_dev, const void* wq_dev, const float* params_dev,
                           const float* dlogits_dev, int64_t batch, int n_features, int hidden,
                           float* grads_dev, void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  B200_REQUIRE(tc_shape_ok(batch, n_features, hidden),
               "mlp_backward_tc: needs F <= %d, H %% 128 == 0, batch %% 128 == 0 (got B=%lld F=%d H=%d)", KP - 2,
               (long long)batch, n_features, hidden);
  B200_REQUIRE(xq_dev && wq_dev && params_dev && dlogits_dev && grads_dev, "mlp_backward_tc: null pointer");
  B200_REQUIRE((((uintptr_t)dlogits_dev) & 15) == 0, "mlp_backward_tc: dlogits must be 16-byte aligned");
  BwdPlan bp;
  bwd_plan(batch, hidden, &bp);
  size_t need = (size_t)bp.splits * hidden * 32 * sizeof(float);
  B200_REQUIRE(scratch_dev && scratch_bytes >= need, "mlp_backward_tc: scratch too small (%zu < %zu)", scratch_bytes, need);
  Layout L{n_features, hidden};
  cudaStream_t st = as_stream(stream);
  static bool attr_set[64];
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B200_CUDA(cudaFuncSetAttribute(mlp_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const unsigned char* w1q = reinterpret_cast<const unsigned char*>(wq_dev);
  mlp_bwd_tc_kernel<<<bp.groups * bp.splits, kBwdThreads, bp.smem, st>>>(
      reinterpret_cast<const unsigned char*>(xq_dev), w1q, dlogits_dev, reinterpret_cast<float*>(scratch_dev), bp.n_btiles,
      hidden, bp.nb, bp.groups, bp.splits);
  B200_LAUNCH_CHECK();
  int fgrid = (hidden + 7) / 8;
  if (fgrid > sm_count() * 8) fgrid = sm_count() * 8;
  mlp_bwd_tc_finish<<<fgrid, 256, 0, st>>>(reinterpret_cast<const float*>(scratch_dev), bp.splits, params_dev, dlogits_dev,
                                           batch, L, grads_dev);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

}  // extern "C"
