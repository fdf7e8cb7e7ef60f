// The widened model with WIDE INPUTS (BASELINE config 4, extension "F = H": n_features 64 .. 8192): here layer 1 is a
// real GEMM -- B x H x F with F in the thousands -- and the tensor cores can be kept busy: K-loop tcgen05.mma kernels
// with fp32 accumulators in TMEM, operands brought in by tensor-map TMA (cp.async.bulk.tensor, SASS UTMALDG) through a
// ring of shared-memory stages, the rest of each layer fused around the GEMM.
//
// Replaces, for F >= 64: model_engine(batch_x) (train.py:107 -> train.py:35-36; SimpleNet takes any n_features,
// train.py:26-33) and the autograd part of model_engine.backward(loss) (train.py:113).
//
// Numerics (same as the F = 20 path): GEMM operands rounded to bf16, fp32 accumulation, fp32 master weights.
// With F in the thousands recomputing layer 1 in the backward would double the dominant cost, so the forward keeps the
// hidden activation: h = bf16(relu(x_q W1_q^T + b1)), 2 bytes per (sample, unit), written once and read once; layer 2
// and dW2 use that rounded h.
//
// Operand layout "L" of a [R, K] bf16 matrix (x: R = batch, K = F; W1: R = H, K = F; h^T: R = H, K = batch):
//     byte(r, k) = (r/8) * (K*16) + (k/8) * 128 + (r%8) * 16 + (k%8) * 2
// i.e. 8 x 8 core matrices of 128 contiguous bytes, k-fastest inside a group of 8 rows.  Any (8i rows) x (64j columns)
// tile is a set of 1 KB runs, and a tensor map over the array seen as uint32 [R/8][K/64][256] lands it in shared memory
// as the UMMA canonical no-swizzle layout -- K-major (core matrices 128 B apart along K, 1 KB apart along rows) for the
// forward's operands and the weight gradient's h^T, and, with a box of [8 row groups][4 k blocks][256], MN-major for x
// as the B operand of the weight gradient (a core matrix is its own transpose under the two readings: x is stored
// once and serves both GEMMs).  All tensor maps cover THIS GPU's memory only.
#include "common.cuh"
#include "tc_common.cuh"
#include <cuda.h>
#include <cudaTypedefs.h>
#include "philox.cuh"

namespace b200ds {
using namespace tc;

__host__ __device__ __forceinline__ size_t fh_byte(int64_t r, int64_t k, int64_t K) {
  return (size_t)(r >> 3) * (size_t)(K * 16) + (size_t)(k >> 3) * 128 + (size_t)(r & 7) * 16 + (size_t)(k & 7) * 2;
}

constexpr int kFhBK = 64;                  // K per pipeline stage
constexpr int kFhTM = 128;                 // M of a tile (TMEM lanes)
constexpr int kFhTN = 256;                 // N of a tile (TMEM columns of one accumulator)
constexpr int kFhStages = 4;
constexpr uint32_t kFhABytes = kFhTM * kFhBK * 2;     // 16 KB
constexpr uint32_t kFhBBytes = kFhTN * kFhBK * 2;     // 32 KB
constexpr int kFhThreads = 32 * 6;         // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue (TMEM lane quadrant = warp % 4)

// ---------------------------------------------------------------- TMA with a tensor map (3-D)
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
               ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tma_prefetch_map(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// ---------------------------------------------------------------- fp32 row-major -> layout L (bf16)
// thread t -> the t-th 16-byte chunk of the output: (row in group, k group, row group), so stores are contiguous
__global__ void __launch_bounds__(256)
fh_pack_kernel(const float* __restrict__ src, int64_t rows, int64_t K, uint4* __restrict__ dst) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= rows * (K / 8)) return;
  const int r8 = (int)(t & 7);
  const int64_t kg = (t >> 3) % (K / 8), rg = t / K;
  const float4* s = reinterpret_cast<const float4*>(src + (rg * 8 + r8) * K + kg * 8);
  const float4 a = __ldg(s), b = __ldg(s + 1);
  __nv_bfloat162 q[4] = {__floats2bfloat162_rn(a.x, a.y), __floats2bfloat162_rn(a.z, a.w), __floats2bfloat162_rn(b.x, b.y),
                         __floats2bfloat162_rn(b.z, b.w)};
  dst[t] = *reinterpret_cast<uint4*>(q);
}

// ---------------------------------------------------------------- forward GEMM
// D[n, b] = sum_k W1q[n, k] xq[b, k]   (A = W1 tile 128 x 64, B = x tile 256 x 64, both K-major)
// epilogue per element: h = bf16(relu(D + b1[n])) -> hT[n, b];  zpart[n tile][b] = sum_{n in tile} w2[n] h[n, b]
struct FhFwdSmem {
  float red[2][4][kFhTN];                  // [tile parity][lane quadrant][column]: per-warp column sums
  uint64_t full[kFhStages], empty[kFhStages], acc_full[2], acc_empty[2];
  uint32_t tmem_slot;
};

template <int N, int OFF>
__device__ __forceinline__ void fh_lane_fold(float (&a)[32], int lane) {
  const bool up = (lane & OFF) != 0;
#pragma unroll
  for (int i = 0; i < N / 2; ++i) {
    const float keep = up ? a[i + N / 2] : a[i];
    const float send = up ? a[i] : a[i + N / 2];
    a[i] = keep + __shfl_xor_sync(0xffffffffu, send, OFF);
  }
}

__global__ void __launch_bounds__(kFhThreads, 1)
fh_fwd_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x,
              const float* __restrict__ b1, const float* __restrict__ w2, unsigned char* __restrict__ hT,
              float* __restrict__ zpart /* [H/128][batch] */, int64_t batch, int F, int H) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* a_s = smem;                                           // [stages][16 KB]
  unsigned char* b_s = smem + kFhStages * kFhABytes;                   // [stages][32 KB]
  FhFwdSmem* sh = reinterpret_cast<FhFwdSmem*>(b_s + kFhStages * kFhBBytes);
  const int n_nt = H / kFhTM, n_bt = (int)(batch / kFhTN), n_kb = F / kFhBK;
  const int n_tiles = n_nt * n_bt;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kFhStages; ++i) { mbar_init(&sh->full[i], 1); mbar_init(&sh->empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sh->acc_full[i], 1); mbar_init(&sh->acc_empty[i], 4); }
    mbar_fence_init();
    tma_prefetch_map(&map_w);
    tma_prefetch_map(&map_x);
  }
  if (warp == 1) tmem_alloc<512>(&sh->tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int nt = tile % n_nt, bt = tile / n_nt;
        for (int kb = 0; kb < n_kb; ++kb) {
          mbar_wait(&sh->empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&sh->full[s], kFhABytes + kFhBBytes);
          tma_load_3d(a_s + (size_t)s * kFhABytes, &map_w, 0, kb, nt * (kFhTM / 8), &sh->full[s]);
          tma_load_3d(b_s + (size_t)s * kFhBBytes, &map_x, 0, kb, bt * (kFhTN / 8), &sh->full[s]);
          if (++s == kFhStages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    const uint32_t idesc = idesc_bf16(kFhTM, kFhTN, 0, 0);
    const uint64_t adesc0 = smem_desc(smem_u32(a_s), 128, 1024);
    const uint64_t bdesc0 = smem_desc(smem_u32(b_s), 128, 1024);
    int s = 0; uint32_t ph = 0; int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int ab = it & 1;
      mbar_wait(&sh->acc_empty[ab], ((uint32_t)(it >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t dcol = tmem + ab * kFhTN;
      for (int kb = 0; kb < n_kb; ++kb) {
        mbar_wait(&sh->full[s], ph);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t ad = adesc0 + (uint64_t)((s * kFhABytes) >> 4), bd = bdesc0 + (uint64_t)((s * kFhBBytes) >> 4);
#pragma unroll
          for (int k = 0; k < kFhBK / 16; ++k)
            mma_ss(dcol, ad + (uint64_t)(k * 16), bd + (uint64_t)(k * 16), idesc, (kb | k) ? 1u : 0u);   // 2 core matrices = 256 B per K step
          mma_commit(&sh->empty[s]);
        }
        __syncwarp();
        if (++s == kFhStages) { s = 0; ph ^= 1; }
      }
      if (elect_one()) mma_commit(&sh->acc_full[ab]);
      __syncwarp();
    }
  } else {
    // ===== epilogue =====
    const int quad = warp & 3;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const int et = threadIdx.x - 64;                                    // 0..127 among the epilogue threads
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int nt = tile % n_nt, bt = tile / n_nt, ab = it & 1;
      const int n = nt * kFhTM + quad * 32 + lane;
      const float b1n = __ldg(b1 + n), w2n = __ldg(w2 + n);
      mbar_wait(&sh->acc_full[ab], (uint32_t)(it >> 1) & 1u);
      tc_fence_after();
      float* red = sh->red[it & 1][quad];
      for (int c = 0; c < kFhTN / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem + lane_base + ab * kFhTN + c * 32, v);
        tmem_wait_ld();
        if (c == kFhTN / 32 - 1) {                                      // the accumulator is in registers: hand it back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sh->acc_empty[ab]);
        }
        float p[32];
        uint4* dst = reinterpret_cast<uint4*>(hT + fh_byte(n, (int64_t)bt * kFhTN + c * 32, batch));
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t q[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int i = j * 8 + e * 2;
            q[e] = pack_relu_bf16x2(__uint_as_float(v[i]) + b1n, __uint_as_float(v[i + 1]) + b1n);
            p[i] = w2n * __uint_as_float(q[e] << 16);
            p[i + 1] = w2n * __uint_as_float(q[e] & 0xffff0000u);
          }
          dst[j * 8] = make_uint4(q[0], q[1], q[2], q[3]);              // next 8 columns: 128 bytes on (one core matrix)
        }
        fh_lane_fold<32, 16>(p, lane);
        fh_lane_fold<16, 8>(p, lane);
        fh_lane_fold<8, 4>(p, lane);
        fh_lane_fold<4, 2>(p, lane);
        fh_lane_fold<2, 1>(p, lane);
        red[c * 32 + lane] = p[0];                                      // lane l ends with the sum of column l of the piece
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      {
        const float* r0 = sh->red[it & 1][0];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int col = et + 128 * j;
          zpart[(size_t)nt * batch + (size_t)bt * kFhTN + col] =
              (r0[col] + r0[kFhTN + col]) + (r0[2 * kFhTN + col] + r0[3 * kFhTN + col]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem);
}

// ---------------------------------------------------------------- generated batch, straight into layout L
// Element li = row*F + k of torch.randn(B*F, device='cuda') under (seed, offset) is output j % 4 of Philox call
// (counter offset/4 + j/4, subsequence sub), li = sub + nthreads*j (philox.cu, ATen's mapping).  When F divides nthreads
// (powers of two up to 8192: nthreads = 303104 = 2^13 * 37 for large batches) a subsequence stays in one COLUMN,
// k = sub % F, and walks down the rows r0 + (nthreads/F) * j, r0 = sub / F.  A CTA owns 8 consecutive r0 x 32
// consecutive k (warp = one r0): per Philox call every thread gets four values for four rows, eight lanes pack their
// eight neighbouring columns into one 16-byte store, and the CTA's eight warps fill whole 32-byte sectors of the core
// matrices (their eight rows are consecutive).  All four outputs of every Philox call are used; fp32 x is never stored.
// Labels (train.py:22-23) need sum_{k < F/2} x[row, k]: a warp's 32 columns are summed by shuffles and stored as
// ypart[row][k / 32]; fh_label_kernel adds a row's parts up in column order (deterministic) and thresholds.
__device__ __forceinline__ void fh_normals4(uint64_t ctr, uint64_t sub, uint2 key, float (&out)[4]) {
  const uint4 r = philox4x32_10(make_uint4((unsigned)ctr, (unsigned)(ctr >> 32), (unsigned)sub, (unsigned)(sub >> 32)), key);
  const float2 p = box_muller(r.x, r.y), q = box_muller(r.z, r.w);
  out[0] = normal_transform(p.x); out[1] = normal_transform(p.y);
  out[2] = normal_transform(q.x); out[3] = normal_transform(q.y);
}

__global__ void __launch_bounds__(256)
fh_gen_kernel(unsigned char* __restrict__ xq, float* __restrict__ ypart, int64_t batch, int F, uint2 key, uint64_t ctr0,
              uint32_t nthreads, int calls, int calls_per_cta, const unsigned long long* __restrict__ draw_counter) {
  if (draw_counter) ctr0 += (uint64_t)(*draw_counter) * (uint64_t)calls;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rstride = (int)(nthreads / (uint32_t)F);                    // rows between consecutive outputs of a subsequence
  const int r0 = blockIdx.y * 8 + warp;
  if (r0 >= rstride) return;
  const int k = blockIdx.x * 32 + lane;
  const uint64_t sub = (uint64_t)r0 * (uint64_t)F + (uint64_t)k;
  const int half_blocks = F / 64;                                       // 32-column blocks inside the first half of a row
  const int c_begin = blockIdx.z * calls_per_cta, c_end = min(calls, c_begin + calls_per_cta);
  for (int c = c_begin; c < c_end; ++c) {
    const int64_t row_c = (int64_t)r0 + (int64_t)rstride * 4 * c;       // row of output 0
    if (row_c >= batch) break;
    float v[4];
    fh_normals4(ctr0 + (uint64_t)c, sub, key, v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t row = row_c + (int64_t)rstride * i;
      if (row >= batch) break;                                          // warp-uniform
      // eight lanes' bf16 values -> one 16-byte store
      const uint32_t h16 = (uint32_t)__bfloat16_as_ushort(__float2bfloat16(v[i]));
      const uint32_t pair = h16 | (__shfl_down_sync(0xffffffffu, h16, 1) << 16);
      const uint32_t pair2 = __shfl_down_sync(0xffffffffu, pair, 2);
      const uint32_t q2 = __shfl_down_sync(0xffffffffu, pair, 4), q3 = __shfl_down_sync(0xffffffffu, pair2, 4);
      if ((lane & 7) == 0) *reinterpret_cast<uint4*>(xq + fh_byte(row, k, F)) = make_uint4(pair, pair2, q2, q3);
      if ((int)blockIdx.x < half_blocks) {                              // this block of 32 columns lies in the first half of the row
        float a = v[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0) ypart[(size_t)row * half_blocks + blockIdx.x] = a;
      }
    }
  }
}

__global__ void __launch_bounds__(256)
fh_label_kernel(const float* __restrict__ ypart, int64_t batch, int half_blocks, float* __restrict__ y) {
  const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= batch) return;
  const float* p = ypart + (size_t)row * half_blocks;
  float a = 0.f;
  for (int j = 0; j < half_blocks; ++j) a += p[j];
  y[row] = a > 0.f ? 1.f : 0.f;
}

// ---------------------------------------------------------------- logits from the forward's parts
__global__ void __launch_bounds__(256)
fh_logits_kernel(const float* __restrict__ zpart, int n_parts, int64_t batch, const float* __restrict__ b2, float* __restrict__ logits) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  float z = zpart[b];
  for (int p = 1; p < n_parts; ++p) z += zpart[(size_t)p * batch + b];      // tile order: the same sum on every run
  logits[b] = z + b2[0];
}

// ---------------------------------------------------------------- backward of relu, in place on hT
// hT[n, b] (the activation, bf16) becomes A2T[n, b] = 1[h > 0] * bf16(dz[b]): the A operand of the weight-gradient
// GEMM (dW1[n, :] = W2[n] * sum_b A2T[n, b] x[b, :]).  On the way the same pass forms, per split of the batch,
//   part[split][0][n] = sum_b dz[b] h[n, b]        (-> dW2[n])
//   part[split][1][n] = sum_b 1[h > 0] dz[b]       (-> db1[n] = W2[n] * that)
// One block = one group of 8 rows (units) x one split of the batch; thread = (row in group, k group), 16-byte accesses,
// a warp covers 512 contiguous bytes.  Reductions in a fixed order.
constexpr int kFhRbSplits = 8;
__global__ void __launch_bounds__(256)
fh_relu_bwd_kernel(unsigned char* __restrict__ hT, const float* __restrict__ dz, int64_t batch, int H, float* __restrict__ part) {
  __shared__ float red[2][8][8];                                        // [which][warp][row in group]
  const int t = threadIdx.x, r8 = t & 7, kg0 = t >> 3, warp = t >> 5, lane = t & 31;
  const int rg = blockIdx.x, split = blockIdx.y;
  const int64_t kgs = batch / 8, per = kgs / kFhRbSplits;               // k groups (of 8 samples) in total / per split
  uint4* row = reinterpret_cast<uint4*>(hT + (size_t)rg * (size_t)batch * 16) + r8;
  float aw = 0.f, ab = 0.f;
  for (int64_t kg = split * per + kg0; kg < (split + 1) * per; kg += 32) {
    uint4 hv = row[kg * 8];
    const float4 d0 = __ldg(reinterpret_cast<const float4*>(dz + kg * 8)), d1 = __ldg(reinterpret_cast<const float4*>(dz + kg * 8) + 1);
    const float dd[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
    uint32_t hw[4] = {hv.x, hv.y, hv.z, hv.w}, ow[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float h0 = __uint_as_float(hw[e] << 16), h1 = __uint_as_float(hw[e] & 0xffff0000u);
      const float g0 = dd[2 * e], g1 = dd[2 * e + 1];
      aw = fmaf(g0, h0, aw); aw = fmaf(g1, h1, aw);
      const float m0 = h0 > 0.f ? g0 : 0.f, m1 = h1 > 0.f ? g1 : 0.f;
      ab += m0; ab += m1;
      __nv_bfloat162 q = __floats2bfloat162_rn(m0, m1);
      ow[e] = *reinterpret_cast<uint32_t*>(&q);
    }
    row[kg * 8] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
  }
  aw += __shfl_xor_sync(0xffffffffu, aw, 8);  ab += __shfl_xor_sync(0xffffffffu, ab, 8);
  aw += __shfl_xor_sync(0xffffffffu, aw, 16); ab += __shfl_xor_sync(0xffffffffu, ab, 16);
  if (lane < 8) { red[0][warp][lane] = aw; red[1][warp][lane] = ab; }
  __syncthreads();
  if (t < 16) {
    const int which = t >> 3, r = t & 7;
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += red[which][w][r];
    part[((size_t)split * 2 + which) * H + rg * 8 + r] = a;
  }
}

// dW2, db1, db2 from the parts; one thread per hidden unit, block 0 also sums dz (fixed order)
__global__ void __launch_bounds__(256)
fh_finish_kernel(const float* __restrict__ part, const float* __restrict__ dz, int64_t batch, const float* __restrict__ params, Layout L,
                 float* __restrict__ grads) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < L.H) {
    float gw = 0.f, gb = 0.f;
    for (int s = 0; s < kFhRbSplits; ++s) { gw += part[((size_t)s * 2 + 0) * L.H + n]; gb += part[((size_t)s * 2 + 1) * L.H + n]; }
    grads[L.off_w2() + n] = gw;
    grads[L.off_b1() + n] = params[L.off_w2() + n] * gb;
  }
  if (blockIdx.x == 0) {
    __shared__ float red[256];
    float a = 0.f;
    for (int64_t b = threadIdx.x; b < batch; b += 256) a += dz[b];
    red[threadIdx.x] = a;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
      __syncthreads();
    }
    if (threadIdx.x == 0) grads[L.off_b2()] = red[0];
  }
}

// ---------------------------------------------------------------- weight-gradient GEMM
// S[n, k] = sum_b A2T[n, b] x[b, k]   (A = A2T tile 128 x 64 K-major, B = x tile [64 b][256 k] MN-major), dW1 = W2[n] S.
// A work unit is an output tile (n tile, k tile) x one of kFhWgSplits halves of the batch; the halves meet in the
// gradient buffer through red.global.add on zeros -- two addends, so the sum does not depend on their order.
constexpr int kFhWgSplits = 2;
struct FhWgSmem {
  uint64_t full[kFhStages], empty[kFhStages], acc_full[2], acc_empty[2];
  uint32_t tmem_slot;
};

__global__ void __launch_bounds__(kFhThreads, 1)
fh_wgrad_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_x,
                const float* __restrict__ w2, float* __restrict__ gw1 /* [H, F] fp32, zeroed */, int64_t batch, int F, int H) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* a_s = smem;
  unsigned char* b_s = smem + kFhStages * kFhABytes;
  FhWgSmem* sh = reinterpret_cast<FhWgSmem*>(b_s + kFhStages * kFhBBytes);
  const int n_nt = H / kFhTM, n_kt = F / kFhTN;
  const int n_units = n_nt * n_kt * kFhWgSplits;
  const int n_bb = (int)(batch / kFhBK) / kFhWgSplits;                   // 64-sample blocks per unit

  if (threadIdx.x == 0) {
    for (int i = 0; i < kFhStages; ++i) { mbar_init(&sh->full[i], 1); mbar_init(&sh->empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sh->acc_full[i], 1); mbar_init(&sh->acc_empty[i], 4); }
    mbar_fence_init();
    tma_prefetch_map(&map_a);
    tma_prefetch_map(&map_x);
  }
  if (warp == 1) tmem_alloc<512>(&sh->tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
        const int nt = u % n_nt, kt = (u / n_nt) % n_kt, sp = u / (n_nt * n_kt);
        for (int i = 0; i < n_bb; ++i) {
          const int bb = sp * n_bb + i;
          mbar_wait(&sh->empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&sh->full[s], kFhABytes + kFhBBytes);
          tma_load_3d(a_s + (size_t)s * kFhABytes, &map_a, 0, bb, nt * (kFhTM / 8), &sh->full[s]);          // 128 units x 64 samples
          tma_load_3d(b_s + (size_t)s * kFhBBytes, &map_x, 0, kt * (kFhTN / 64), bb * (kFhBK / 8), &sh->full[s]);   // 64 samples x 256 features
          if (++s == kFhStages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = idesc_bf16(kFhTM, kFhTN, 0, 1);
    const uint64_t adesc0 = smem_desc(smem_u32(a_s), 128, 1024);                       // K-major: 128 B along K, 1 KB along rows
    const uint64_t bdesc0 = smem_desc(smem_u32(b_s), /*along K (samples)*/ 4096, /*along N (features)*/ 128);   // MN-major
    int s = 0; uint32_t ph = 0; int it = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
      const int ab = it & 1;
      mbar_wait(&sh->acc_empty[ab], ((uint32_t)(it >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t dcol = tmem + ab * kFhTN;
      for (int i = 0; i < n_bb; ++i) {
        mbar_wait(&sh->full[s], ph);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t ad = adesc0 + (uint64_t)((s * kFhABytes) >> 4), bd = bdesc0 + (uint64_t)((s * kFhBBytes) >> 4);
#pragma unroll
          for (int k = 0; k < kFhBK / 16; ++k)
            mma_ss(dcol, ad + (uint64_t)(k * 16), bd + (uint64_t)(k * 512), idesc, (i | k) ? 1u : 0u);     // A: +256 B, B: +2 row groups of 4 KB
          mma_commit(&sh->empty[s]);
        }
        __syncwarp();
        if (++s == kFhStages) { s = 0; ph ^= 1; }
      }
      if (elect_one()) mma_commit(&sh->acc_full[ab]);
      __syncwarp();
    }
  } else {
    const int quad = warp & 3;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    int it = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++it) {
      const int nt = u % n_nt, kt = (u / n_nt) % n_kt, ab = it & 1;
      const int n = nt * kFhTM + quad * 32 + lane;
      const float w2n = __ldg(w2 + n);
      float* dst = gw1 + (size_t)n * F + (size_t)kt * kFhTN;
      mbar_wait(&sh->acc_full[ab], (uint32_t)(it >> 1) & 1u);
      tc_fence_after();
      for (int c = 0; c < kFhTN / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem + lane_base + ab * kFhTN + c * 32, v);
        tmem_wait_ld();
        if (c == kFhTN / 32 - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sh->acc_empty[ab]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + c * 32 + 4 * j), "f"(w2n * __uint_as_float(v[4 * j])),
                       "f"(w2n * __uint_as_float(v[4 * j + 1])), "f"(w2n * __uint_as_float(v[4 * j + 2])),
                       "f"(w2n * __uint_as_float(v[4 * j + 3]))
                       : "memory");
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem);
}

// ---------------------------------------------------------------- host side
static PFN_cuTensorMapEncodeTiled g_encode = nullptr;

static int fh_get_encode() {
  if (g_encode) return B200DS_OK;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  B200_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  if (!fn || qres != cudaDriverEntryPointSuccess) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return B200DS_ERR_CUDA; }
  g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(fn);
  return B200DS_OK;
}

// tensor map over a layout-L array [R, K] bf16 seen as uint32 [R/8][K/64][256]; box = [row_groups][k_blocks][256]
static int fh_make_map(CUtensorMap* map, const void* base, int64_t R, int64_t K, int row_groups, int k_blocks) {
  if (int rc = fh_get_encode()) return rc;
  const cuuint64_t dims[3] = {256, (cuuint64_t)(K / 64), (cuuint64_t)(R / 8)};
  const cuuint64_t strides[2] = {1024, (cuuint64_t)K * 16};
  const cuuint32_t box[3] = {256, (cuuint32_t)k_blocks, (cuuint32_t)row_groups};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (CUresult %d) for a [%lld, %lld] operand", (int)r, (long long)R, (long long)K); return B200DS_ERR_CUDA; }
  return B200DS_OK;
}

static bool fh_shape_ok(int64_t batch, int F, int H) {
  return F >= 64 && F % 64 == 0 && F % 256 == 0 && H >= 128 && H % 128 == 0 && batch >= 256 && batch % 256 == 0;
}

}  // namespace b200ds

using namespace b200ds;

extern "C" {

int b200ds_fh_supported(int64_t batch, int n_features, int hidden) { return fh_shape_ok(batch, n_features, hidden) ? 1 : 0; }

// forward AND backward: the backward also cuts the batch into kFhRbSplits x 256 rows per CTA (relu pass) and two halves (weight gradient)
static bool fh_train_ok(int64_t batch, int n_features, int hidden) {
  return fh_shape_ok(batch, n_features, hidden) && (batch / 8) % (kFhRbSplits * 32) == 0 && (batch / kFhBK) % kFhWgSplits == 0;
}
int b200ds_fh_train_supported(int64_t batch, int n_features, int hidden) { return fh_train_ok(batch, n_features, hidden) ? 1 : 0; }

size_t b200ds_fh_operand_bytes(int64_t rows, int64_t cols) { return rows > 0 && cols > 0 ? (size_t)rows * (size_t)cols * 2 : 0; }

int b200ds_fh_pack_bf16(const float* src_dev, int64_t rows, int64_t cols, void* dst_dev, uintptr_t stream) {
  B200_REQUIRE(rows > 0 && rows % 8 == 0 && cols > 0 && cols % 64 == 0, "fh_pack_bf16: rows %% 8 == 0 and cols %% 64 == 0 (got %lld x %lld)",
               (long long)rows, (long long)cols);
  B200_REQUIRE(src_dev && dst_dev, "fh_pack_bf16: null pointer");
  B200_REQUIRE((((uintptr_t)src_dev) & 15) == 0 && (((uintptr_t)dst_dev) & 15) == 0, "fh_pack_bf16: buffers must be 16-byte aligned");
  const int64_t chunks = rows * (cols / 8);
  fh_pack_kernel<<<(unsigned)((chunks + 255) / 256), 256, 0, as_stream(stream)>>>(src_dev, rows, cols, reinterpret_cast<uint4*>(dst_dev));
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

size_t b200ds_fh_forward_scratch_bytes(int64_t batch, int n_features, int hidden) {
  return fh_shape_ok(batch, n_features, hidden) ? (size_t)(hidden / kFhTM) * (size_t)batch * sizeof(float) : 0;
}

int b200ds_fh_forward(const void* xq_dev, const void* w1q_dev, const float* params_dev, int64_t batch, int n_features, int hidden,
                      void* hT_dev, void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  B200_REQUIRE(fh_shape_ok(batch, n_features, hidden), "fh_forward: needs F %% 256 == 0, H %% 128 == 0, batch %% 256 == 0 (got B=%lld F=%d H=%d)",
               (long long)batch, n_features, hidden);
  B200_REQUIRE(xq_dev && w1q_dev && params_dev && hT_dev, "fh_forward: null pointer");
  const size_t need = b200ds_fh_forward_scratch_bytes(batch, n_features, hidden);
  B200_REQUIRE(scratch_dev && scratch_bytes >= need, "fh_forward: scratch too small (%zu < %zu)", scratch_bytes, need);
  CUtensorMap map_w, map_x;
  if (int rc = fh_make_map(&map_w, w1q_dev, hidden, n_features, kFhTM / 8, 1)) return rc;
  if (int rc = fh_make_map(&map_x, xq_dev, batch, n_features, kFhTN / 8, 1)) return rc;
  const Layout L{n_features, hidden};
  const size_t smem = (size_t)kFhStages * (kFhABytes + kFhBBytes) + sizeof(FhFwdSmem) + 1024;
  static bool attr_set[64];
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B200_CUDA(cudaFuncSetAttribute(fh_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const int n_tiles = (hidden / kFhTM) * (int)(batch / kFhTN);
  const int grid = n_tiles < sm_count() ? n_tiles : sm_count();
  fh_fwd_kernel<<<grid, kFhThreads, smem, as_stream(stream)>>>(map_w, map_x, params_dev + L.off_b1(), params_dev + L.off_w2(),
                                                               reinterpret_cast<unsigned char*>(hT_dev),
                                                               reinterpret_cast<float*>(scratch_dev), batch, n_features, hidden);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

size_t b200ds_fh_gen_scratch_bytes(int64_t batch, int n_features) {
  return batch > 0 && n_features >= 64 ? (size_t)batch * (size_t)(n_features / 64) * sizeof(float) : 0;
}

int b200ds_fh_gen_supported(int64_t batch, int n_features) {
  if (batch <= 0 || batch % 8 != 0 || n_features < 64 || n_features % 64 != 0) return 0;
  int grid, calls;
  philox_plan(batch * n_features, &grid, &calls);
  return ((int64_t)grid * 256) % n_features == 0 ? 1 : 0;
}

static int fh_launch_gen(void* xq_dev, float* y_dev, int64_t batch, int n_features, uint64_t seed, uint64_t offset,
                         const uint64_t* draw_counter_dev, uint64_t* offset_advance, void* scratch_dev, size_t scratch_bytes,
                         uintptr_t stream) {
  B200_REQUIRE(b200ds_fh_gen_supported(batch, n_features),
               "fh_gen_batch: n_features must divide the generator's thread count (a power of two in [64, 8192] for large batches) and "
               "batch %% 8 == 0 (got B=%lld F=%d); otherwise use b200ds_philox_normal + b200ds_fh_pack_bf16", (long long)batch, n_features);
  B200_REQUIRE(offset % 4 == 0, "fh_gen_batch: offset must be a multiple of 4");
  B200_REQUIRE(xq_dev && y_dev, "fh_gen_batch: null pointer");
  const size_t need = b200ds_fh_gen_scratch_bytes(batch, n_features);
  B200_REQUIRE(scratch_dev && scratch_bytes >= need, "fh_gen_batch: scratch too small (%zu < %zu)", scratch_bytes, need);
  int grid, calls;
  philox_plan(batch * n_features, &grid, &calls);
  const uint32_t nthreads = (uint32_t)grid * 256u;
  const int rstride = (int)(nthreads / (uint32_t)n_features);
  // enough CTAs for ~8 waves of the machine: the call range is cut into pieces
  const int ctas_xy = (n_features / 32) * ((rstride + 7) / 8);
  int pieces = (sm_count() * 8 * 8 + ctas_xy - 1) / ctas_xy;
  if (pieces < 1) pieces = 1;
  if (pieces > calls) pieces = calls;
  const int per = (calls + pieces - 1) / pieces;
  pieces = (calls + per - 1) / per;
  cudaStream_t st = as_stream(stream);
  fh_gen_kernel<<<dim3(n_features / 32, (rstride + 7) / 8, pieces), 256, 0, st>>>(
      reinterpret_cast<unsigned char*>(xq_dev), reinterpret_cast<float*>(scratch_dev), batch, n_features,
      make_uint2((unsigned)seed, (unsigned)(seed >> 32)), offset / 4, nthreads, calls, per,
      reinterpret_cast<const unsigned long long*>(draw_counter_dev));
  B200_LAUNCH_CHECK();
  fh_label_kernel<<<(unsigned)((batch + 255) / 256), 256, 0, st>>>(reinterpret_cast<const float*>(scratch_dev), batch, n_features / 64, y_dev);
  B200_LAUNCH_CHECK();
  if (offset_advance) *offset_advance = (uint64_t)calls * 4;
  return B200DS_OK;
}

int b200ds_fh_gen_batch(void* xq_dev, float* y_dev, int64_t batch, int n_features, uint64_t seed, uint64_t offset,
                        uint64_t* offset_advance, void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  return fh_launch_gen(xq_dev, y_dev, batch, n_features, seed, offset, nullptr, offset_advance, scratch_dev, scratch_bytes, stream);
}

int b200ds_fh_gen_batch_counted(void* xq_dev, float* y_dev, int64_t batch, int n_features, uint64_t seed, uint64_t offset0,
                                const uint64_t* draw_counter_dev, void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  B200_REQUIRE(draw_counter_dev != nullptr, "fh_gen_batch_counted: null draw counter");
  return fh_launch_gen(xq_dev, y_dev, batch, n_features, seed, offset0, draw_counter_dev, nullptr, scratch_dev, scratch_bytes, stream);
}

int b200ds_fh_logits(const void* fwd_scratch_dev, const float* params_dev, int64_t batch, int n_features, int hidden,
                     float* logits_dev, uintptr_t stream) {
  B200_REQUIRE(fh_shape_ok(batch, n_features, hidden), "fh_logits: shape outside the wide-input class");
  B200_REQUIRE(fwd_scratch_dev && params_dev && logits_dev, "fh_logits: null pointer");
  const Layout L{n_features, hidden};
  fh_logits_kernel<<<(unsigned)((batch + 255) / 256), 256, 0, as_stream(stream)>>>(reinterpret_cast<const float*>(fwd_scratch_dev),
                                                                                   hidden / kFhTM, batch, params_dev + L.off_b2(), logits_dev);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

size_t b200ds_fh_backward_scratch_bytes(int64_t batch, int n_features, int hidden) {
  return fh_shape_ok(batch, n_features, hidden) ? (size_t)kFhRbSplits * 2 * (size_t)hidden * sizeof(float) : 0;
}

int b200ds_fh_backward(const void* xq_dev, void* hT_dev, const float* dlogits_dev, const float* params_dev, int64_t batch,
                       int n_features, int hidden, float* grads_dev, void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  B200_REQUIRE(fh_shape_ok(batch, n_features, hidden), "fh_backward: needs F %% 256 == 0, H %% 128 == 0, batch %% 256 == 0 (got B=%lld F=%d H=%d)",
               (long long)batch, n_features, hidden);
  B200_REQUIRE(fh_train_ok(batch, n_features, hidden), "fh_backward: batch must be a multiple of %d (b200ds_fh_train_supported)", 8 * kFhRbSplits * 32);
  B200_REQUIRE(xq_dev && hT_dev && dlogits_dev && params_dev && grads_dev, "fh_backward: null pointer");
  B200_REQUIRE((((uintptr_t)dlogits_dev) & 15) == 0 && (((uintptr_t)grads_dev) & 15) == 0, "fh_backward: dlogits / grads must be 16-byte aligned");
  const size_t need = b200ds_fh_backward_scratch_bytes(batch, n_features, hidden);
  B200_REQUIRE(scratch_dev && scratch_bytes >= need, "fh_backward: scratch too small (%zu < %zu)", scratch_bytes, need);
  const Layout L{n_features, hidden};
  cudaStream_t st = as_stream(stream);
  float* part = reinterpret_cast<float*>(scratch_dev);
  // 1. relu backward in place (hT -> A2T) + the parts of dW2 / db1;  2. their finish + db2
  fh_relu_bwd_kernel<<<dim3(hidden / 8, kFhRbSplits), 256, 0, st>>>(reinterpret_cast<unsigned char*>(hT_dev), dlogits_dev, batch, hidden, part);
  B200_LAUNCH_CHECK();
  fh_finish_kernel<<<(hidden + 255) / 256, 256, 0, st>>>(part, dlogits_dev, batch, params_dev, L, grads_dev);
  B200_LAUNCH_CHECK();
  // 3. dW1: the two batch halves of every output tile add into zeros
  B200_CUDA(cudaMemsetAsync(grads_dev + L.off_w1(), 0, sizeof(float) * (size_t)hidden * n_features, st));
  CUtensorMap map_a, map_x;
  if (int rc = fh_make_map(&map_a, hT_dev, hidden, batch, kFhTM / 8, 1)) return rc;
  if (int rc = fh_make_map(&map_x, xq_dev, batch, n_features, kFhBK / 8, kFhTN / 64)) return rc;
  const size_t smem = (size_t)kFhStages * (kFhABytes + kFhBBytes) + sizeof(FhWgSmem) + 1024;
  static bool attr_set[64];
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B200_CUDA(cudaFuncSetAttribute(fh_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const int n_units = (hidden / kFhTM) * (n_features / kFhTN) * kFhWgSplits;
  const int grid = n_units < sm_count() ? n_units : sm_count();
  fh_wgrad_kernel<<<grid, kFhThreads, smem, st>>>(map_a, map_x, params_dev + L.off_w2(), grads_dev + L.off_w1(), batch, n_features, hidden);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

}  // extern "C"
