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
//           epilogue: A2[n,b] = 1[z1 not negative] * dz[b] as bf16 (sign-bit test, 2 instructions per
//           pair; it differs from relu'(z1) only for z1 == +0.0 exactly), written back IN PLACE into TMEM,
//           S[n,k] += A2 . xq_tile   (TS MMA, A from TMEM, B = the same x tile read MN-major)
//           then  dW1[n,f] = w2[n] S[n,f],  db1[n] = w2[n] S[n,F],
//                 dW2[n]  = sum_k w1q[n,k] S[n,k]  (= sum_b dz[b] h[b,n]),  db2 = sum_b dz[b].
#include "common.cuh"
#include "tc_common.cuh"
#include "adam.cuh"
#include "p2p.cuh"
#include <type_traits>
#include "philox.cuh"
#include <stdlib.h>

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
// rows [rows, batch) of the pack are padding: zero features (the ones columns stay, so such a row sees z1 = b1)
__global__ void __launch_bounds__(256)
pack_x_kernel(const float* __restrict__ x, int64_t rows, int64_t batch, int F, uint4* __restrict__ xq) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= batch * (KP / 8)) return;
  int r8 = (int)(t & 7), c = (int)((t >> 3) & 3);
  int64_t row = ((t >> 5) << 3) + r8;
  const float* xr = x + row * F;
  const bool real = row < rows;
  __nv_bfloat16 v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int k = c * 8 + j;
    float f = k < F ? (real ? xr[k] : 0.f) : (k < F + 2 ? 1.f : 0.f);
    v[j] = __float2bfloat16(f);
  }
  xq[t] = *reinterpret_cast<uint4*>(v);
}

// Generated batch straight into the tensor-core pack (BASELINE config 4: the batch is drawn on the device every
// step).  Element (row, k) is element li = row*F + k of torch.randn(B*F, device='cuda') under (seed, offset): Philox
// counter offset/4 + j/4, subsequence sub, output j % 4, where li = sub + nthreads * j -- the mapping of
// philox_normal_kernel, i.e. of ATen's distribution kernels.  The kernel uses torch's own thread geometry: thread `sub`
// evaluates Philox4x32-10 once per call index c and uses ALL FOUR outputs -- the elements sub + nthreads*(4c + i),
// i = 0..3, four places of the batch ~nthreads/F rows apart -- converts each to bf16 and stores it at its place in the
// tiled pack.  (The first version gave every ELEMENT a thread and a Philox evaluation of its own, keeping one output of
// four: 3.3 times the generator work for the benchmark shape.)
// A row's 64-byte pack and its label (train.py:22-23: sum of columns 0 .. F/2-1 in index order, > 0) need the row's F
// neighbouring elements: the block keeps its fp32 values in shared memory, 256 consecutive elements per (c, i), plus a
// halo of F - 1 elements past the end that a few threads draw in addition (one more evaluation per call index serves
// all four i).  A row belongs to the block that holds its column 0: that block writes the whole row -- one 16-byte
// store per k chunk, the thread -> address pattern of pack_x_kernel -- and its label.  fp32 x never reaches global
// memory, and nothing is stored in pieces smaller than 16 bytes.
constexpr int kGenMaxCalls = 4;            // Philox calls per thread kept in flight (more calls -> outer loop)
constexpr int kGenHaloMax = 32;            // F - 1 <= 29

template <typename idx_t>
__device__ __forceinline__ void gen_divmod(idx_t li, uint32_t d, uint32_t inv, idx_t* q, uint32_t* r) {
  if (sizeof(idx_t) == 4) {                // li / d by multiply-high and one correction (64-bit div/mod would dominate)
    uint32_t qd = __umulhi((uint32_t)li, inv);
    uint32_t rm = (uint32_t)li - qd * d;
    const bool fix = rm >= d;
    qd += fix ? 1u : 0u;
    rm -= fix ? d : 0u;
    *q = (idx_t)qd; *r = rm;
  } else {
    *q = li / (idx_t)d;
    *r = (uint32_t)(li - *q * (idx_t)d);
  }
}

__device__ __forceinline__ void gen_normals4(uint64_t ctr, uint64_t sub, uint2 key, float (&out)[4]) {
  const uint4 r = philox4x32_10(make_uint4((unsigned)ctr, (unsigned)(ctr >> 32), (unsigned)sub, (unsigned)(sub >> 32)), key);
  const float2 p = box_muller(r.x, r.y), q = box_muller(r.z, r.w);
  out[0] = normal_transform(p.x); out[1] = normal_transform(p.y);
  out[2] = normal_transform(q.x); out[3] = normal_transform(q.y);
}

template <typename idx_t>
__global__ void __launch_bounds__(256, 8)      // torch's geometry is 8 blocks of 256 per SM: all of them resident (one wave) needs <= 32 registers
gen_batch_kernel(unsigned char* __restrict__ xq, float* __restrict__ y, int64_t batch, int F, uint2 key, uint64_t ctr0,
                 uint32_t nthreads, int calls, uint32_t inv_F /* floor(2^32 / F) */, uint32_t inv_nthreads,
                 const unsigned long long* __restrict__ draw_counter) {
  __shared__ float xs[kGenMaxCalls * 4][256 + kGenHaloMax];
  if (draw_counter) ctr0 += (uint64_t)(*draw_counter) * (uint64_t)calls;
  const int t = threadIdx.x;
  const idx_t n = (idx_t)batch * (idx_t)F;
  const int half = F >> 1, halo = F - 1;
  // a CTA walks torch's blocks vb = blockIdx.x, blockIdx.x + gridDim.x, ...: the launch may hold fewer CTAs than torch's
  // grid so that the generator leaves room on every SM for the kernels it runs beside (the gradient exchange)
  for (uint32_t vb = blockIdx.x; vb < nthreads / 256u; vb += gridDim.x)
  for (int c0 = 0; c0 < calls; c0 += kGenMaxCalls) {
    const uint32_t sub = vb * 256u + t;
    const int nc = min(kGenMaxCalls, calls - c0);
    // ---- this thread's elements: up to kGenMaxCalls independent Philox chains side by side
    float val[kGenMaxCalls][4];
#pragma unroll
    for (int cc = 0; cc < kGenMaxCalls; ++cc) {
      if (cc < nc && (idx_t)sub + (idx_t)nthreads * (idx_t)(4 * (c0 + cc)) < n)       // any of the four inside the batch
        gen_normals4(ctr0 + (uint64_t)(c0 + cc), sub, key, val[cc]);
      else
        val[cc][0] = val[cc][1] = val[cc][2] = val[cc][3] = 0.f;
    }
#pragma unroll
    for (int cc = 0; cc < kGenMaxCalls; ++cc) {
      if (cc >= nc) continue;
#pragma unroll
      for (int i = 0; i < 4; ++i) xs[cc * 4 + i][t] = val[cc][i];
    }
    // ---- halo: the `halo` elements after the block's 256, per (c, i)
    if (t < halo) {
      const uint32_t sub_h = vb * 256u + 256u + t;
      if (sub_h < nthreads) {              // same call index, next subsequences: one evaluation serves i = 0..3
        for (int cc = 0; cc < nc; ++cc) {
          float hv[4] = {0.f, 0.f, 0.f, 0.f};
          if ((idx_t)sub_h + (idx_t)nthreads * (idx_t)(4 * (c0 + cc)) < n) gen_normals4(ctr0 + (uint64_t)(c0 + cc), sub_h, key, hv);
#pragma unroll
          for (int i = 0; i < 4; ++i) xs[cc * 4 + i][256 + t] = hv[i];
        }
      } else {                             // last block: the element after (nthreads - 1, j) is (0, j + 1)
        for (int cc = 0; cc < nc; ++cc)
          for (int i = 0; i < 4; ++i) {
            const idx_t li = (idx_t)sub_h + (idx_t)nthreads * (idx_t)(4 * (c0 + cc) + i);
            float v1 = 0.f;
            if (li < n) {
              idx_t j; uint32_t sb;
              gen_divmod<idx_t>(li, nthreads, inv_nthreads, &j, &sb);
              float hv[4];
              gen_normals4(ctr0 + (uint64_t)(j >> 2), sb, key, hv);
              v1 = hv[(int)(j & 3)];
            }
            xs[cc * 4 + i][256 + t] = v1;
          }
      }
    }
    __syncthreads();
    // ---- rows whose column 0 lies among this block's 256 elements of some (c, i): at most 256 / F + 1 per (c, i);
    // a work item is (combo, row-in-combo, k chunk)
    const int rows_per = 256 / F + 1;
    for (int w = t; w < nc * 4 * rows_per * (KP / 8); w += 256) {
      const int c8 = w & (KP / 8 - 1), wr = w >> 2;                    // KP / 8 == 4
      const int combo = wr / rows_per, ri = wr - combo * rows_per;
      const idx_t base = (idx_t)vb * 256 + (idx_t)nthreads * (idx_t)(4 * c0 + combo);
      if (base >= n) continue;
      idx_t row0; uint32_t k0;
      gen_divmod<idx_t>(base, (uint32_t)F, inv_F, &row0, &k0);
      const idx_t row = row0 + (k0 ? 1 : 0) + (idx_t)ri;              // first row starting at or after `base`, then the next ones
      const idx_t start = row * (idx_t)F;                             // li of its column 0
      if (start >= base + 256 || row >= (idx_t)batch) continue;
      const float* xr = &xs[combo][(int)(start - base)];
      __nv_bfloat16 q[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = c8 * 8 + j;
        q[j] = __float2bfloat16(k < F ? xr[k] : (k < F + 2 ? 1.f : 0.f));
      }
      *reinterpret_cast<uint4*>(xq + tiled_byte((int64_t)row, c8 * 8)) = *reinterpret_cast<uint4*>(q);
      if (c8 == 0) {                                                   // label: sum of columns 0 .. F/2-1 in index order
        float acc = 0.f;
        for (int k = 0; k < half; ++k) acc += xr[k];
        y[row] = acc > 0.f ? 1.f : 0.f;
      }
    }
    __syncthreads();
  }
}

constexpr int kPackRows = 16;             // rows of W1 per block of the pack (pack_w1_kernel: 16 rows x 4 k chunks; the update kernel: 16 warps)

// Also emits, per block of 64 rows, vpart[block][k] = sum_{n in block} w2[n] * w1q[n,k]: the forward
// sums the blocks of its slice in a fixed order to get v = w1q^T w2 (the linear half of relu).
__global__ void __launch_bounds__(4 * kPackRows)
pack_w1_kernel(const float* __restrict__ params, Layout L, uint4* __restrict__ w1q, float* __restrict__ vpart) {
  __shared__ float prod[kPackRows][33];
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;     // H % 64 == 0: no partial blocks
  const int r8 = (int)(t & 7), c = (int)((t >> 3) & 3);
  const int n = (int)(((t >> 5) << 3) + r8);
  const float* wr = params + L.off_w1() + (int64_t)n * L.F;
  const float b1 = params[L.off_b1() + n];
  const float w2 = params[L.off_w2() + n];
  const __nv_bfloat16 hi = __float2bfloat16(b1);
  const __nv_bfloat16 lo = __float2bfloat16(b1 - __bfloat162float(hi));
  __nv_bfloat16 v[8];
  const int rl = n - blockIdx.x * kPackRows;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int k = c * 8 + j;
    v[j] = k < L.F ? __float2bfloat16(wr[k]) : (k == L.F ? hi : (k == L.F + 1 ? lo : __float2bfloat16(0.f)));
    prod[rl][k] = w2 * __bfloat162float(v[j]);
  }
  w1q[t] = *reinterpret_cast<uint4*>(v);
  __syncthreads();
  if (threadIdx.x < 32) {
    float a = 0.f;
    for (int r = 0; r < kPackRows; ++r) a += prod[r][threadIdx.x];
    vpart[(size_t)blockIdx.x * 32 + threadIdx.x] = a;
  }
}

// ------------------------------------------------------------------ forward
// D1T[n, b] = w1_chunk . x_tile^T puts the hidden unit n on the TMEM lane, so the layer-2 weight
// w2[n] is ONE register per thread per chunk (no shared-memory broadcast traffic next to the
// UMMA operand reads, which already use the full 128 B/clk), and each thread keeps 64 running
// sums acc[b] += w2[n] |z1[n,b]| across all chunks of the tile.  The sum over n (lanes, then
// the 4 lane quadrants) happens once per tile.
#ifdef B200DS_TRACE      // tuning builds only (tools/build_variant.sh trace "-DB200DS_TRACE", tools/tc_trace.py): clock64 stamps of CTA 0
__device__ long long g_trace[24 * 128 * 4];      // [role][chunk or group][event]
#define B200_TRACE(role, q, ev) do { if (blockIdx.x == 0 && lane == 0 && (q) < 128) g_trace[(((role) * 128) + (q)) * 4 + (ev)] = clock64(); } while (0)
#else
#define B200_TRACE(role, q, ev) do { } while (0)
#endif
constexpr int kFwdXStages = 3;
constexpr int kFwdRing = 4;               // D1T accumulators in TMEM: 4 x 128 columns
constexpr int kFwdEpiWarps = 8;           // warp (quad, half): 32 lanes (n) x 64 columns (b)
constexpr int kFwdThreads = 32 * 12;      // warps 0-7 epilogue, warp 8 TMA, warps 9 / 10 MMA (even / odd chunks), warp 11 slice sum.
                                          // The issuing warps sit at the highest warp ids: the SM's arbiter favours the
                                          // highest ready warp, and an MMA issuer that waits behind 64 FFMAs stalls the pipe.

struct FwdSmem {                           // fixed-size tail after the W1 slice / w2 slice / x stages
  float v[32];
  float red[2][2][4][64];                  // [tile parity][column half][lane quadrant][column]: per-warp column sums
  uint64_t w_bar, x_full[kFwdXStages], x_free[kFwdXStages], d1_full[kFwdRing], d1_free[kFwdRing];
  uint32_t tmem_slot;
  uint64_t z_full[2], z_free[2];           // red[parity] handed to the finishing warp / given back
  uint64_t v_ready;                        // v[] (formed by the epilogue warps in the prologue) may be read by the finishing warp
};

// sum over the 32 lanes of 64 per-lane values; lane l ends with the totals of columns
// col(l), col(l)+1 where col(l) = 2 * bitreverse5(l)... precisely: bit4->+32, bit3->+16, bit2->+8, bit1->+4, bit0->+2
template <int N, int OFF>
__device__ __forceinline__ void lane_fold(float (&a)[64], int lane) {
  const bool up = (lane & OFF) != 0;
#pragma unroll
  for (int i = 0; i < N / 2; ++i) {
    const float keep = up ? a[i + N / 2] : a[i];
    const float send = up ? a[i] : a[i + N / 2];
    a[i] = keep + __shfl_xor_sync(0xffffffffu, send, OFF);
  }
}

// G = chunks per full/free handshake (one tcgen05.commit and one mbarrier round per G chunks of
// 128 columns; the accumulator ring is 4 / G groups deep).
template <int G>
__global__ void __launch_bounds__(kFwdThreads, 1)
mlp_fwd_tc_kernel(const unsigned char* __restrict__ xq, const unsigned char* __restrict__ w1q,
                  const float* __restrict__ vlin, const float* __restrict__ w2, const float* __restrict__ b2p,
                  float* __restrict__ logits, float* __restrict__ zpart /* [n_slices][batch], n_slices > 1 */,
                  unsigned int* __restrict__ tickets /* [n_btiles], zero between launches */,
                  int n_btiles, int slice_rows, int n_slices, int ctas_per_slice) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slice = blockIdx.x % n_slices, cta_in_slice = blockIdx.x / n_slices;
  const int NC = slice_rows / kTileRows;                 // chunks per b tile
  unsigned char* w1s = smem;                             // [slice_rows][KP] bf16 tiled
  float* w2s = reinterpret_cast<float*>(w1s + (size_t)slice_rows * KP * 2);       // [slice_rows] fp32
  unsigned char* xs = reinterpret_cast<unsigned char*>(w2s + slice_rows);         // [stages][8192]
  FwdSmem* sh = reinterpret_cast<FwdSmem*>(xs + kFwdXStages * kTileBytes);
  constexpr int RING = kFwdRing / G;                     // accumulator groups in TMEM
  const int NG = NC / G;                                 // groups per b tile
  const uint32_t n_issuers = NG < 2 ? 1u : 2u;

  if (threadIdx.x == 0) {
    mbar_init(&sh->w_bar, 1);
    // an x stage has two kinds of readers: the MMAs (released by the issuers' tcgen05.commit) and the finishing
    // warp, which reads the tile's rows for the linear term -- both must be done with it before the producer may
    // refill the stage (with NG <= RING the MMAs alone run a whole ring ahead)
    for (int i = 0; i < kFwdXStages; ++i) { mbar_init(&sh->x_full[i], 1); mbar_init(&sh->x_free[i], n_issuers + 1); }
    for (int i = 0; i < kFwdRing; ++i) { mbar_init(&sh->d1_full[i], 1); mbar_init(&sh->d1_free[i], kFwdEpiWarps); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sh->z_full[i], kFwdEpiWarps); mbar_init(&sh->z_free[i], 1); }
    mbar_init(&sh->v_ready, 1);
    mbar_fence_init();
    // start the 128 KB W1 slice right away: it overlaps the rest of the prologue
    const unsigned char* wsrc = w1q + (size_t)slice * slice_rows * KP * 2;
    mbar_arrive_expect_tx(&sh->w_bar, (uint32_t)slice_rows * KP * 2);
    for (int c = 0; c < NC; ++c) tma_load_1d(w1s + (size_t)c * kTileBytes, wsrc + (size_t)c * kTileBytes, kTileBytes, &sh->w_bar);
  }
  if (warp == 9) tmem_alloc<512>(&sh->tmem_slot);
  tc_fence_before();
  __syncthreads();                         // barriers and TMEM are up: the producer and the MMA issuers start at once
  tc_fence_after();
  if (warp < kFwdEpiWarps) {
    // only the epilogue warps need w2 and v: they fetch them among themselves (named barrier 1, the one
    // their tile ends use) while the first MMAs already run
    for (int i = threadIdx.x; i < slice_rows; i += 32 * kFwdEpiWarps) w2s[i] = w2[slice * slice_rows + i];
    {    // v = sum over the 64-row blocks of this slice: 8 warps take 1/8 of the blocks each, then a fixed-order sum
      const int nblk = slice_rows / kPackRows;
      const float* vp = vlin + (size_t)slice * nblk * 32 + lane;
      float a = 0.f;
      for (int i = warp; i < nblk; i += 8) a += vp[(size_t)i * 32];
      (&sh->red[0][0][0][0])[warp * 32 + lane] = a;      // red[] is free until the first tile ends
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (threadIdx.x < 32) {
      float a = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) a += (&sh->red[0][0][0][0])[w * 32 + threadIdx.x];
      sh->v[threadIdx.x] = a;
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (threadIdx.x == 0) mbar_arrive(&sh->v_ready);
  }
  if (sh->tmem_slot != 0) __trap();        // all 512 columns are ours: the allocation can only start at 0,
  constexpr uint32_t tmem = 0;             // and a literal keeps the MMA issue path free of vector -> uniform moves

  if (warp == 8) {
    // ===== TMA producer =====
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (int t = cta_in_slice; t < n_btiles; t += ctas_per_slice) {
        mbar_wait(&sh->x_free[s], ph ^ 1);
        mbar_arrive_expect_tx(&sh->x_full[s], kTileBytes);
        tma_load_1d(xs + s * kTileBytes, xq + (size_t)t * kTileBytes, kTileBytes, &sh->x_full[s]);
        if (++s == kFwdXStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 9 || warp == 10) {
    // ===== MMA issuers (group parity): D1T[ring] = w1_chunk . x_tile^T, K = 32.  The whole warp
    // walks the loop (uniform control flow); one elected lane issues.  Two issuing warps because
    // one warp's wait + issue + commit round is longer than the MMAs of a group execute.  The body
    // is instantiated per issuer so that every operand of the issue path is provably warp-uniform
    // (the issuer shares its scheduler with two FFMA-heavy warps: each instruction it needs costs
    // several issue slots). =====
    auto issue = [&](auto me_tag) {
      constexpr int me = decltype(me_tag)::value;
      if (me >= NG) return;
      const uint32_t idesc = idesc_bf16(128, 128, 0, 0);
      const uint64_t xdesc0 = smem_desc(smem_u32(xs), LBO_K, SBO_K);
      const uint64_t wdesc0 = smem_desc(smem_u32(w1s), LBO_K, SBO_K);
      mbar_wait(&sh->w_bar, 0);
      int s = 0; uint32_t xph = 0; int gq0 = 0;
      for (int t = cta_in_slice; t < n_btiles; t += ctas_per_slice, gq0 += NG) {
        mbar_wait(&sh->x_full[s], xph);
        const uint64_t xd = xdesc0 + (uint64_t)((s * kTileBytes) >> 4);
        for (int g = me; g < NG; g += 2) {
          const int gq = gq0 + g;
          const int r = gq % RING;
          const uint32_t rph = (uint32_t)(gq / RING) & 1u;
          const uint64_t wd = wdesc0 + (uint64_t)((g * G * kTileBytes) >> 4);
          const uint32_t dcol = tmem + r * (128 * G);
          B200_TRACE(me, gq, 0);
          mbar_wait(&sh->d1_free[r], rph ^ 1);
          tc_fence_after();
          B200_TRACE(me, gq, 1);
          if (elect_one()) {
#pragma unroll
            for (int j = 0; j < G; ++j) {
              mma_ss(dcol + j * 128, wd + (uint64_t)((j * kTileBytes) >> 4), xd, idesc, 0);
              mma_ss(dcol + j * 128, wd + (uint64_t)((j * kTileBytes + 2 * LBO_K) >> 4), xd + (2 * LBO_K >> 4), idesc, 1);
            }
            mma_commit(&sh->d1_full[r]);
          }
          __syncwarp();
          B200_TRACE(me, gq, 2);
        }
        if (elect_one()) mma_commit(&sh->x_free[s]);
        __syncwarp();
        if (++s == kFwdXStages) { s = 0; xph ^= 1; }
      }
    };
    if (warp == 9) issue(std::integral_constant<int, 0>{});
    else issue(std::integral_constant<int, 1>{});
  } else if (warp == 11) {
    // ===== finisher: per tile, rows lane + 32 j.  It forms the linear half of relu itself, lin = xq[row,:] . v, from the x
    // tile; takes the eight epilogue warps' column sums (red[parity], one mbarrier hand-over: the epilogue warps do not
    // meet at a barrier at the end of a tile, each goes on to the next tile when its own fold is done); adds the four
    // lane quadrants in a fixed order; z = (lin + sum) / 2 [+ b2].  With several slices of hidden units (H > 2048, in
    // different CTAs) the CTA's part goes to global memory and whoever draws the LAST of the tile's n_slices tickets
    // adds the parts up in slice order: bit-reproducible whichever slice was faster, nothing to zero beforehand.  All
    // of it -- fences, L2 round trips of the ticket and of the parts -- runs beside the next tile's arithmetic. =====
    const float b2 = b2p[0];
    const size_t nrows = (size_t)n_btiles * kTileRows;
    int s = 0; uint32_t xph = 0; int it = 0;
    mbar_wait(&sh->v_ready, 0);
    for (int t = cta_in_slice; t < n_btiles; t += ctas_per_slice, ++it) {
      float lin[4] = {0.f, 0.f, 0.f, 0.f};
      mbar_wait(&sh->x_full[s], xph);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int row = lane + 32 * j;
        const unsigned char* xr = xs + s * kTileBytes + (size_t)(row >> 3) * SBO_K + (size_t)(row & 7) * 16;
#pragma unroll
        for (int c = 0; c < KP / 8; ++c) {
          uint4 qv = *reinterpret_cast<const uint4*>(xr + c * LBO_K);
          const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&qv);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float2 f = __bfloat1622float2(h2[e]);
            lin[j] = fmaf(f.x, sh->v[c * 8 + 2 * e], lin[j]);
            lin[j] = fmaf(f.y, sh->v[c * 8 + 2 * e + 1], lin[j]);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&sh->x_free[s]);            // the rows of the stage are in registers
      if (++s == kFwdXStages) { s = 0; xph ^= 1; }
      const int b = it & 1;
      mbar_wait(&sh->z_full[b], (uint32_t)(it >> 1) & 1u);
      float z[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int row = lane + 32 * j, hb = row >> 6, cb = row & 63;
        const float ab = (sh->red[b][hb][0][cb] + sh->red[b][hb][1][cb]) + (sh->red[b][hb][2][cb] + sh->red[b][hb][3][cb]);
        z[j] = 0.5f * (lin[j] + ab);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&sh->z_free[b]);
      const size_t g0 = (size_t)t * kTileRows + lane;
      if (n_slices == 1) {
#pragma unroll
        for (int j = 0; j < 4; ++j) logits[g0 + 32 * j] = z[j] + b2;
        continue;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) __stcg(zpart + (size_t)slice * nrows + g0 + 32 * j, z[j]);
      __threadfence();
      __syncwarp();
      unsigned tk = 0;
      if (lane == 0) tk = atomicAdd(&tickets[t], 1u);
      tk = __shfl_sync(0xffffffffu, tk, 0);
      if (tk == (unsigned)(n_slices - 1)) {
        __threadfence();
        float zs[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) zs[j] = __ldcg(zpart + g0 + 32 * j);
        for (int sl = 1; sl < n_slices; ++sl) {
#pragma unroll
          for (int j = 0; j < 4; ++j) zs[j] += __ldcg(zpart + (size_t)sl * nrows + g0 + 32 * j);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) logits[g0 + 32 * j] = zs[j] + b2;
        if (lane == 0) tickets[t] = 0u;
      }
    }
  } else {
    // ===== epilogue: 8 warps on every chunk =====
    const int half = warp >> 2, quad = warp & 3;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    int it = 0, gq = 0;
    for (int t = cta_in_slice; t < n_btiles; t += ctas_per_slice) {
      float acc[64];
#pragma unroll
      for (int i = 0; i < 64; ++i) acc[i] = 0.f;
      // Software pipeline across chunks: the TMEM load of one 32-column piece is in flight while
      // the FFMAs of the previous piece issue (va: columns 0-31 of the chunk, vb: 32-63).
      uint32_t va[32], vb[32];
      {
        const int r = gq % RING;
        mbar_wait(&sh->d1_full[r], (uint32_t)(gq / RING) & 1u);
        tc_fence_after();
        tmem_ld32(tmem + lane_base + r * (128 * G) + half * 64, va);
      }
      for (int g = 0; g < NG; ++g, ++gq) {
        const int r = gq % RING;
        B200_TRACE(2 + warp, gq, 0);
#pragma unroll
        for (int j = 0; j < G; ++j) {
          const float w2n = w2s[(g * G + j) * kTileRows + quad * 32 + lane];
          const uint32_t cbase = tmem + lane_base + r * (128 * G) + j * 128 + half * 64;
          tmem_wait_ld();                                              // va (piece 0 of this chunk) has landed
          tmem_ld32(cbase + 32, vb);                                   // piece 1
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[i] = fmaf(fabsf(__uint_as_float(va[i])), w2n, acc[i]);
          tmem_wait_ld();                                              // vb has landed
          if (j + 1 < G) {
            tmem_ld32(cbase + 128, va);                                // piece 0 of the next chunk of the group
          } else {
            tc_fence_before();                                         // the group is fully read: release it
            __syncwarp();
            if (lane == 0) mbar_arrive(&sh->d1_free[r]);
            B200_TRACE(2 + warp, gq, 1);
            if (g + 1 < NG) {                                          // piece 0 of the next group
              const int r1 = (gq + 1) % RING;
              mbar_wait(&sh->d1_full[r1], (uint32_t)((gq + 1) / RING) & 1u);
              tc_fence_after();
              B200_TRACE(2 + warp, gq, 2);
              tmem_ld32(tmem + lane_base + r1 * (128 * G) + half * 64, va);
            }
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[32 + i] = fmaf(fabsf(__uint_as_float(vb[i])), w2n, acc[32 + i]);
        }
        B200_TRACE(2 + warp, gq, 3);
      }
      // ---- sum over n: 32 lanes by a transposing butterfly, then the 4 quadrants through smem
      lane_fold<64, 16>(acc, lane);
      lane_fold<32, 8>(acc, lane);
      lane_fold<16, 4>(acc, lane);
      lane_fold<8, 2>(acc, lane);
      lane_fold<4, 1>(acc, lane);
      const int col = ((lane & 16) ? 32 : 0) + ((lane & 8) ? 16 : 0) + ((lane & 4) ? 8 : 0) + ((lane & 2) ? 4 : 0) + ((lane & 1) ? 2 : 0);
      {                                                    // this warp's column sums -> the finishing warp
        const int b = it & 1;
        mbar_wait(&sh->z_free[b], ((uint32_t)(it >> 1) & 1u) ^ 1u);
        *reinterpret_cast<float2*>(&sh->red[b][half][quad][col]) = make_float2(acc[0], acc[1]);
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh->z_full[b]);
      }
      ++it;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------ backward
constexpr int kBwdXStages = 12;
constexpr int kBwdRing = 3;               // D1T accumulators: 3 x 128 columns (cols 0..383)
constexpr int kBwdNB = 4;                 // n blocks per CTA: S accumulators 4 x 32 columns (cols 384..511)
constexpr int kBwdThreads = 32 * 12;      // 0-7 epilogue (two groups), 8 TMA + dz stager, 9 MMA1, 10-11 MMA2 (n block parity);
                                          // issuers at the highest warp ids, see the forward kernel
constexpr int kBwdLossThreads = 32 * 16;  // fused loss: warps 12-15 form dz (one row of the tile per thread) instead of warp 8

struct BwdSmem {
  uint32_t dzs[kBwdXStages][64];          // dz of the tile as bf16x2 pairs
  uint64_t w_bar, x_full[kBwdXStages], x_free[kBwdXStages], dz_full[kBwdXStages];
  uint64_t d1_full[kBwdRing], d1_free[kBwdRing], a_ready[kBwdRing], s_full;
  uint32_t tmem_slot;
  float lred[4][2];                       // fused loss: per-warp sums of dz and of the loss terms
};

// LOSS: the loss kernel is folded in -- `dz` holds the LOGITS; four extra warps form
// dz[b] = (sigmoid(z) - y) / B for the tile (one row per thread, the arithmetic of
// bce_loss_grad_kernel: the W1 / b1 / W2 gradients are bit-identical to the two-kernel path), and
// in the CTAs of group 0 they also sum the loss terms into loss_part[split]; the finish pass adds
// those up.  dlogits never reaches HBM.
template <bool LOSS>
__global__ void __launch_bounds__(kBwdLossThreads, 1)
mlp_bwd_tc_kernel(const unsigned char* __restrict__ xq, const unsigned char* __restrict__ w1q,
                  const float* __restrict__ dz, const float* __restrict__ ylab, float fbatch,
                  float* __restrict__ partial /* [splits][H][32] */, float* __restrict__ db2_part /* [splits] */,
                  float* __restrict__ loss_part /* [splits] */, unsigned long long* __restrict__ step_counter,
                  int n_btiles, int hidden, int nb, int n_groups, int n_splits) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int group = blockIdx.x % n_groups, split = blockIdx.x / n_groups;
  if (step_counter && blockIdx.x == 0 && threadIdx.x == 0) *step_counter += 1ull;   // nothing in this kernel reads it
  unsigned char* w1s = smem;                                           // [nb][8192]
  unsigned char* xs = w1s + (size_t)nb * kTileBytes;                   // [stages][8192]
  BwdSmem* sh = reinterpret_cast<BwdSmem*>(xs + kBwdXStages * kTileBytes);
  // b tiles of this split: contiguous range
  const int per = (n_btiles + n_splits - 1) / n_splits;
  const int t_begin = split * per, t_end = min(n_btiles, t_begin + per);

  if (threadIdx.x == 0) {
    mbar_init(&sh->w_bar, 1);
    const uint32_t n_issuers = nb < 2 ? 1u : 2u;     // MMA2 issuing threads that commit per tile
    for (int i = 0; i < kBwdXStages; ++i) { mbar_init(&sh->x_full[i], 1); mbar_init(&sh->x_free[i], n_issuers); mbar_init(&sh->dz_full[i], LOSS ? 4 : 1); }
    for (int i = 0; i < kBwdRing; ++i) { mbar_init(&sh->d1_full[i], 1); mbar_init(&sh->d1_free[i], 1); mbar_init(&sh->a_ready[i], 4); }
    mbar_init(&sh->s_full, n_issuers);
    mbar_fence_init();
  }
  if (warp == 9) tmem_alloc<512>(&sh->tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sh->tmem_slot;
  const uint32_t s_col0 = kBwdRing * 128;

  if (warp == 8) {
    // ===== TMA producer (lane 0) + dz stager (all lanes): dz[b0..b0+127] -> bf16x2 pairs in smem =====
    if (lane == 0) {
      const unsigned char* wsrc = w1q + (size_t)group * nb * kTileBytes;
      mbar_arrive_expect_tx(&sh->w_bar, (uint32_t)nb * kTileBytes);
      for (int i = 0; i < nb; ++i) tma_load_1d(w1s + (size_t)i * kTileBytes, wsrc + (size_t)i * kTileBytes, kTileBytes, &sh->w_bar);
    }
    int s = 0; uint32_t ph = 0;
    float dsum = 0.f;
    // dz is read three tiles ahead: a global load in this serial loop would otherwise put its
    // whole latency into every tile
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    auto dz_tile = [&](int t) { return t < t_end ? __ldg(reinterpret_cast<const float4*>(dz + (size_t)t * kTileRows) + lane) : zero4; };
    float4 d0 = dz_tile(t_begin), d1 = dz_tile(t_begin + 1), d2 = dz_tile(t_begin + 2);
    for (int t = t_begin; t < t_end; ++t) {
      mbar_wait(&sh->x_free[s], ph ^ 1);
      if (lane == 0) {
        mbar_arrive_expect_tx(&sh->x_full[s], kTileBytes);
        tma_load_1d(xs + s * kTileBytes, xq + (size_t)t * kTileBytes, kTileBytes, &sh->x_full[s]);
      }
      if (!LOSS) {
        const float4 d = d0;
        d0 = d1; d1 = d2; d2 = dz_tile(t + 3);
        dsum += (d.x + d.y) + (d.z + d.w);
        __nv_bfloat162 p0 = __floats2bfloat162_rn(d.x, d.y), p1 = __floats2bfloat162_rn(d.z, d.w);
        sh->dzs[s][lane * 2 + 0] = *reinterpret_cast<uint32_t*>(&p0);
        sh->dzs[s][lane * 2 + 1] = *reinterpret_cast<uint32_t*>(&p1);
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh->dz_full[s]);
      }
      if (++s == kBwdXStages) { s = 0; ph ^= 1; }
    }
    if (!LOSS) {
      dsum = warp_sum(dsum);       // this split's share of db2 = sum_b dz[b]
      if (group == 0 && lane == 0) db2_part[split] = dsum;
    }
  } else if (warp >= 12) {
    // ===== fused loss (LOSS only; the launch has no such warps otherwise): thread = row of the tile =====
    const int r = (warp - 12) * 32 + lane;
    auto ld = [&](const float* src, int t) { return t < t_end ? __ldg(src + (size_t)t * kTileRows + r) : 0.f; };
    // logits and labels are read three tiles ahead (see the dz loads of the plain path)
    float z0 = ld(dz, t_begin), z1 = ld(dz, t_begin + 1), z2 = ld(dz, t_begin + 2);
    float y0 = ld(ylab, t_begin), y1 = ld(ylab, t_begin + 1), y2 = ld(ylab, t_begin + 2);
    float dsum = 0.f, lsum = 0.f;
    int s = 0; uint32_t ph = 0;
    for (int t = t_begin; t < t_end; ++t) {
      const float z = z0, y = y0;
      z0 = z1; z1 = z2; z2 = ld(dz, t + 3);
      y0 = y1; y1 = y2; y2 = ld(ylab, t + 3);
      const float d = bce_grad(z, y, fbatch);
      dsum += d;
      if (group == 0) lsum += bce_term(z, y);
      mbar_wait(&sh->x_free[s], ph ^ 1);                  // the stage's previous tile is done with dzs[s]
      reinterpret_cast<__nv_bfloat16*>(sh->dzs[s])[r] = __float2bfloat16(d);
      __syncwarp();
      if (lane == 0) mbar_arrive(&sh->dz_full[s]);
      if (++s == kBwdXStages) { s = 0; ph ^= 1; }
    }
    dsum = warp_sum(dsum);
    lsum = warp_sum(lsum);
    if (lane == 0) { sh->lred[warp - 12][0] = dsum; sh->lred[warp - 12][1] = lsum; }
    asm volatile("bar.sync 2, 128;" ::: "memory");
    if (group == 0 && warp == 12 && lane == 0) {           // fixed order over the four warps
      db2_part[split] = (sh->lred[0][0] + sh->lred[1][0]) + (sh->lred[2][0] + sh->lred[3][0]);
      loss_part[split] = (sh->lred[0][1] + sh->lred[1][1]) + (sh->lred[2][1] + sh->lred[3][1]);
    }
  } else if (warp == 9) {
    // ===== MMA1 issuer: D1T[ring] = w1_block . x_tile^T (lanes = n, columns = b) =====
    {
      const uint32_t idesc = idesc_bf16(128, 128, 0, 0);
      const uint64_t xdesc0 = smem_desc(smem_u32(xs), LBO_K, SBO_K);
      const uint64_t wdesc0 = smem_desc(smem_u32(w1s), LBO_K, SBO_K);
      mbar_wait(&sh->w_bar, 0);
      int s = 0; uint32_t xph = 0; int r = 0; uint32_t rph = 0;
      for (int t = t_begin; t < t_end; ++t) {
        mbar_wait(&sh->x_full[s], xph);
        const uint64_t xd = xdesc0 + (uint64_t)((s * kTileBytes) >> 4);
        for (int i = 0; i < nb; ++i) {
          const uint64_t wd = wdesc0 + (uint64_t)((i * kTileBytes) >> 4);
          const uint32_t dcol = tmem + r * 128;
          B200_TRACE(10, (t - t_begin) * nb + i, 0);
          mbar_wait(&sh->d1_free[r], rph ^ 1);
          tc_fence_after();
          B200_TRACE(10, (t - t_begin) * nb + i, 1);
          if (elect_one()) {
            mma_ss(dcol, wd, xd, idesc, 0);
            mma_ss(dcol, wd + (2 * LBO_K >> 4), xd + (2 * LBO_K >> 4), idesc, 1);
            mma_commit(&sh->d1_full[r]);
          }
          __syncwarp();
          B200_TRACE(10, (t - t_begin) * nb + i, 2);
          if (++r == kBwdRing) { r = 0; rph ^= 1; }
        }
        if (++s == kBwdXStages) { s = 0; xph ^= 1; }
      }
    }
  } else if (warp == 10 || warp == 11) {
    // ===== MMA2 issuers: S[i] += A2[ring] (TMEM, bf16) . x_tile (MN-major view), K = 128 rows of b.
    // n block i belongs to issuer i % 2 for the whole kernel, so the accumulation order into
    // S[i] is that thread's program order; two issuing threads lift the per-thread issue rate.
    const int me = warp - 10;
    if (me < nb) {
      const uint32_t idesc = idesc_bf16(128, 32, 0, 1);
      const uint64_t gdesc0 = smem_desc(smem_u32(xs), /*LBO = K groups*/ SBO_K, /*SBO = N groups*/ LBO_K);
      int s = 0; int q = 0;
      for (int t = t_begin; t < t_end; ++t) {
        const uint64_t gd = gdesc0 + (uint64_t)((s * kTileBytes) >> 4);
        const uint32_t first = t > t_begin ? 1u : 0u;
        for (int i = 0; i < nb; ++i, ++q) {
          if ((i & 1) != me) continue;
          const int r = q % kBwdRing;
          const uint32_t rph = (uint32_t)(q / kBwdRing) & 1u;
          const uint32_t scol = tmem + s_col0 + i * 32, acol = tmem + r * 128;
          B200_TRACE(11, q, 0);
          mbar_wait(&sh->a_ready[r], rph);
          tc_fence_after();
          B200_TRACE(11, q, 1);
          if (elect_one()) {
            mma_ts(scol, acol, gd, idesc, first);
#pragma unroll
            for (int u = 1; u < 8; ++u)
              mma_ts(scol, acol + u * 8, gd + (uint64_t)((u * 2 * SBO_K) >> 4), idesc, 1u);
            mma_commit(&sh->d1_free[r]);
          }
          __syncwarp();
          B200_TRACE(11, q, 2);
        }
        if (elect_one()) mma_commit(&sh->x_free[s]);
        __syncwarp();
        if (++s == kBwdXStages) s = 0;
      }
      if (elect_one()) mma_commit(&sh->s_full);
      __syncwarp();
    }
  } else {
    // ===== epilogue: mask * dz -> bf16, in place; chunk q goes to group q % 2 =====
    const int grp = warp >> 2, quad = warp & 3;
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
        B200_TRACE(12 + quad, q, 0);
        mbar_wait(&sh->d1_full[r], rph);
        tc_fence_after();
        B200_TRACE(12 + quad, q, 1);
        const uint4* dzp = reinterpret_cast<const uint4*>(sh->dzs[s]);
        const uint32_t tbase = tmem + lane_base + r * 128;
        uint32_t va[32], vb[32], vc[32];
        uint4 d4[4];
        // A tcgen05.wait::ld round costs ~200 cycles and waits for every outstanding load: the four
        // 32-column pieces go through three register buffers so that only two rounds are exposed,
        // the second with its loads already 64 / 128 instructions old.
        tmem_ld32(tbase, va);
        tmem_ld32(tbase + 32, vb);
#define B200_BWD_CONVERT(J, CUR)                                                                   \
        {                                                                                          \
          _Pragma("unroll") for (int c4 = 0; c4 < 4; ++c4) d4[c4] = dzp[(J) * 4 + c4];            \
          uint32_t p[16];                                                                          \
          _Pragma("unroll") for (int c4 = 0; c4 < 4; ++c4) {                                      \
            const uint32_t dd[4] = {d4[c4].x, d4[c4].y, d4[c4].z, d4[c4].w};                       \
            _Pragma("unroll") for (int e = 0; e < 4; ++e) {                                       \
              const int c = c4 * 4 + e;                                                            \
              uint32_t neg;   /* 0xFFFF in a half whose z1 has the sign bit set (byte-permute with sign replication) */ \
              asm("prmt.b32 %0, %1, %2, 0xFFBB;" : "=r"(neg) : "r"(CUR[2 * c]), "r"(CUR[2 * c + 1]));      \
              p[c] = dd[e] & ~neg;                                                                 \
            }                                                                                      \
          }                                                                                        \
          tmem_st16(tbase + (J) * 16, p);                                                          \
        }
        tmem_wait_ld();                          // pieces 0 and 1
        tmem_ld32(tbase + 64, vc);
        B200_BWD_CONVERT(0, va)
        tmem_ld32(tbase + 96, va);
        B200_BWD_CONVERT(1, vb)
        tmem_wait_ld();                          // pieces 2 and 3
        B200_BWD_CONVERT(2, vc)
        B200_BWD_CONVERT(3, va)
#undef B200_BWD_CONVERT
        B200_TRACE(12 + quad, q, 2);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh->a_ready[r]);
        B200_TRACE(12 + quad, q, 3);
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
  if (warp == 9) tmem_dealloc<512>(tmem);
}

// grads from S: one warp per hidden unit n, lane = k; split partials summed in split order
__global__ void __launch_bounds__(256)
mlp_bwd_tc_finish(const float* __restrict__ partial, const float* __restrict__ db2_part, int n_splits,
                  const float* __restrict__ params, Layout L, float* __restrict__ grads,
                  const float* __restrict__ loss_part, double inv_batch, float* __restrict__ loss_out,
                  unsigned long long* __restrict__ step_counter) {
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  const size_t stride = (size_t)L.H * 32;
  for (int n = blockIdx.x * wpb + (threadIdx.x >> 5); n < L.H; n += gridDim.x * wpb) {
    const float* p = partial + (size_t)n * 32 + lane;
    float s = 0.f;
    int sp = 0;
    for (; sp + 8 <= n_splits; sp += 8) {           // 8 independent loads in flight, summed in order
      float v0 = p[(size_t)(sp + 0) * stride], v1 = p[(size_t)(sp + 1) * stride], v2 = p[(size_t)(sp + 2) * stride],
            v3 = p[(size_t)(sp + 3) * stride], v4 = p[(size_t)(sp + 4) * stride], v5 = p[(size_t)(sp + 5) * stride],
            v6 = p[(size_t)(sp + 6) * stride], v7 = p[(size_t)(sp + 7) * stride];
      s = ((((((((s + v0) + v1) + v2) + v3) + v4) + v5) + v6) + v7);
    }
    for (; sp < n_splits; ++sp) s += p[(size_t)sp * stride];
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
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    float t = 0.f;
    for (int sp = 0; sp < n_splits; ++sp) t += db2_part[sp];
    grads[L.off_b2()] = t;
    if (loss_part) {             // fused loss: mean over the batch, split partials summed in split order in double
      double ls = 0.0;
      for (int sp = 0; sp < n_splits; ++sp) ls += (double)loss_part[sp];
      *loss_out = (float)(ls * inv_batch);
      if (step_counter) *step_counter += 1ull;
    }
  }
}

// ------------------------------------------------------------------ finish + exchange + Adam + operand pack
// Everything between the backward kernel and the next forward: the split partials of S are summed
// (mlp_bwd_tc_finish), the gradients of the W ranks are exchanged through peer memory and summed in rank order
// (skipped for one rank), Adam is applied (adam.cu) and the new W1 / b1 are packed to bf16 with the partials of
// v = w1q^T w2 (pack_w1_kernel).  One warp per hidden unit n, lane = k, the mapping of mlp_bwd_tc_finish: lane k < F
// owns W1[n,k], lane F owns b1[n], lane F + 1 owns W2[n]; b2 belongs to lane 31 of the first warp of CTA 0.  A CTA is
// kPackRows units, H / kPackRows CTAs.  For one rank the result is bit-identical to the three kernels it replaces.
//   kPhase 0  one rank: all of it in one launch;
//   kPhase 1  several ranks, first launch: finish pass, then the gradients are PUSHED: every rank stores its own into
//             slot `rank` of every rank's exchange buffer (its own included), layout [unit][32] + db2, then -- last CTA
//             of the grid -- the epoch into its flag at every peer;
//   (gate)    a one-warp kernel waits for every rank's flag (they sit in local memory): while the ranks wait for the
//             slowest of them nothing but this warp is resident, the SMs belong to whatever runs beside the exchange;
//   kPhase 2  last launch: add the W slots up in rank order (local reads only: the data crossed NVLink one way,
//             once), Adam, pack.
// Separate launches because nothing here waits for a CTA of its own grid: no CTA has to be resident for another one to
// make progress, so the grids can be as large as the work wants.
template <bool kCounted, int kPhase>
__global__ void __launch_bounds__(32 * kPackRows)
mlp_update_tc_kernel(const float* __restrict__ partial, const float* __restrict__ db2_part, const float* __restrict__ loss_part,
                     int n_splits, double inv_batch, float* __restrict__ params, float* __restrict__ grads,
                     float* __restrict__ m, float* __restrict__ v, Layout L, unsigned char* __restrict__ w1q, float* __restrict__ vpart,
                     AdamScalars s, double lr, double beta1, double beta2, int64_t step,
                     const unsigned long long* __restrict__ step_counter, float* __restrict__ loss_out, P2PDev d) {
  __shared__ float prod[kPackRows][33];
  __shared__ float sh[2];
  __shared__ bool last;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int n = blockIdx.x * kPackRows + wib;                          // H % kPackRows == 0: no partial blocks
  const int F = L.F;
  const bool boss = blockIdx.x == 0 && threadIdx.x == 31;              // looks after b2 (besides its own lane's parameter, if it has one)
  const int64_t xi = (int64_t)n * 32 + lane, xb2 = (int64_t)L.H * 32;  // this thread's place in an exchange slot
  float g = 0.f, gb2 = 0.f;
  if (kPhase != 1 && threadIdx.x == 0) {                               // the optimizer's scalars
    double tt = kCounted ? (double)(*step_counter) : (double)step;
    double bc1 = 1.0 - pow(beta1, tt);
    double bc2 = 1.0 - pow(beta2, tt);
    sh[0] = (float)(-(lr / bc1));
    sh[1] = (float)sqrt(bc2);
  }
  if (kPhase != 2) {
    // ---- the finish pass: split partials of S[n, lane] summed in split order, gradients from S
    const float* p = partial + (size_t)n * 32 + lane;
    const size_t stride = (size_t)L.H * 32;
    float sacc = 0.f;
    int sp = 0;
    for (; sp + 8 <= n_splits; sp += 8) {           // 8 independent loads in flight, summed in order
      float v0 = __ldcg(p + (size_t)(sp + 0) * stride), v1 = __ldcg(p + (size_t)(sp + 1) * stride), v2 = __ldcg(p + (size_t)(sp + 2) * stride),
            v3 = __ldcg(p + (size_t)(sp + 3) * stride), v4 = __ldcg(p + (size_t)(sp + 4) * stride), v5 = __ldcg(p + (size_t)(sp + 5) * stride),
            v6 = __ldcg(p + (size_t)(sp + 6) * stride), v7 = __ldcg(p + (size_t)(sp + 7) * stride);
      sacc = ((((((((sacc + v0) + v1) + v2) + v3) + v4) + v5) + v6) + v7);
    }
    for (; sp < n_splits; ++sp) sacc += __ldcg(p + (size_t)sp * stride);
    const float w2 = params[L.off_w2() + n];
    float wq = 0.f;
    if (lane < F) {
      wq = __bfloat162float(__float2bfloat16(params[L.off_w1() + (int64_t)n * F + lane]));
      g = w2 * sacc;
    } else if (lane == F || lane == F + 1) {
      const float b1 = params[L.off_b1() + n];
      const float hi = __bfloat162float(__float2bfloat16(b1));
      wq = lane == F ? hi : __bfloat162float(__float2bfloat16(b1 - hi));
      if (lane == F) g = w2 * sacc;
    }
    const float d2 = warp_sum(wq * sacc);
    if (lane == F + 1) g = d2;
    if (boss) {
      for (int q = 0; q < n_splits; ++q) gb2 += db2_part[q];
      if (loss_part) {           // fused loss: mean over the batch, split partials summed in split order in double
        double ls = 0.0;
        for (int q = 0; q < n_splits; ++q) ls += (double)loss_part[q];
        *loss_out = (float)(ls * inv_batch);
      }
    }
  }
  if (kPhase == 1) {
    // ---- publish: slot `rank` of every rank's buffer, then (last CTA) the epoch into every peer's flag row
    const unsigned long long e = *d.epoch + 1;
    const int64_t half = (int64_t)(e & 1) * d.world * d.cap + (int64_t)d.rank * d.cap;
    for (int w = 0; w < d.world; ++w) {
      d.push[w][half + xi] = g;
      if (boss) d.push[w][half + xb2] = gb2;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();                            // this CTA's stores (ordered before by the barrier) are out
      last = atomicAdd(&d.tickets[0], 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x < d.world) {
      __threadfence_system();
      st_release_sys(&d.flags[threadIdx.x][d.rank], e);
      if (threadIdx.x == 0) d.tickets[0] = 0u;
    }
    return;
  }
  if (kPhase == 2) {
    const unsigned long long e = *d.epoch + 1;             // the gate kernel before this launch has seen every rank's flag at e
    const float* mine = d.push[d.rank] + (int64_t)(e & 1) * d.world * d.cap;
    g = ld_peer1(mine + xi);
    if (boss) gb2 = ld_peer1(mine + xb2);
    for (int w = 1; w < d.world; ++w) {
      g += ld_peer1(mine + (int64_t)w * d.cap + xi);
      if (boss) gb2 += ld_peer1(mine + (int64_t)w * d.cap + xb2);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      if (atomicAdd(&d.tickets[1], 1u) == gridDim.x - 1) {   // the last CTA to have read its slots closes the epoch
        d.tickets[1] = 0u;
        *d.epoch = e;
        __threadfence();
      }
    }
  } else {
    __syncthreads();                                     // sh[] written
  }
  // ---- Adam on the parameter this thread owns; the (summed) gradient goes back to the flat bucket
  const float neg_step_size = sh[0], bc2_sqrt = sh[1];
  int64_t idx = -1;
  if (lane < F) idx = L.off_w1() + (int64_t)n * F + lane;
  else if (lane == F) idx = L.off_b1() + n;
  else if (lane == F + 1) idx = L.off_w2() + n;
  float pw = 0.f;
  if (idx >= 0) {
    pw = params[idx];
    float mm = m[idx], vv = v[idx];
    grads[idx] = g;
    adam_one(pw, g, mm, vv, s, neg_step_size, bc2_sqrt);
    params[idx] = pw; m[idx] = mm; v[idx] = vv;
  }
  if (boss) {
    const int64_t ib = L.off_b2();
    float pp = params[ib], mm = m[ib], vv = v[ib];
    grads[ib] = gb2;
    adam_one(pp, gb2, mm, vv, s, neg_step_size, bc2_sqrt);
    params[ib] = pp; m[ib] = mm; v[ib] = vv;
  }
  // ---- bf16 pack of the new W1 / b1 and the partials of v (pack_w1_kernel)
  const float b1_new = __shfl_sync(0xffffffffu, pw, F);
  const float w2_new = __shfl_sync(0xffffffffu, pw, F + 1);
  const __nv_bfloat16 hi = __float2bfloat16(b1_new);
  const __nv_bfloat16 lo = __float2bfloat16(b1_new - __bfloat162float(hi));
  const __nv_bfloat16 q = lane < F ? __float2bfloat16(pw) : (lane == F ? hi : (lane == F + 1 ? lo : __float2bfloat16(0.f)));
  prod[wib][lane] = w2_new * __bfloat162float(q);
  // eight lanes' bf16 values -> one 16-byte store (the k chunk c = lane / 8 of row n)
  uint32_t h16 = (uint32_t)__bfloat16_as_ushort(q);
  uint32_t pair = h16 | (__shfl_down_sync(0xffffffffu, h16, 1) << 16);        // valid on even lanes
  uint32_t pair2 = __shfl_down_sync(0xffffffffu, pair, 2);                     // lanes % 4 == 0: {pair, pair2} = 4 values
  uint32_t q2 = __shfl_down_sync(0xffffffffu, pair, 4), q3 = __shfl_down_sync(0xffffffffu, pair2, 4);
  if ((lane & 7) == 0)
    *reinterpret_cast<uint4*>(w1q + tiled_byte(n, lane)) = make_uint4(pair, pair2, q2, q3);
  __syncthreads();
  if (threadIdx.x < 32) {
    float a = 0.f;
    for (int r = 0; r < kPackRows; ++r) a += prod[r][threadIdx.x];
    vpart[(size_t)blockIdx.x * 32 + threadIdx.x] = a;
  }
}

__global__ void __launch_bounds__(32)
p2p_gate_kernel(P2PDev d) {
  const unsigned long long e = *d.epoch + 1;
  if (threadIdx.x < d.world) p2p_wait_flag(&d.flags[d.rank][threadIdx.x], e, (int)threadIdx.x, d);
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

#ifdef B200DS_TRACE
B200DS_API int b200ds_debug_trace(long long* out_host, size_t n, int clear) {      // tuning builds only; not part of the ABI
  B200_CUDA(cudaDeviceSynchronize());
  if (out_host) B200_CUDA(cudaMemcpyFromSymbol(out_host, g_trace, n * sizeof(long long)));
  if (clear) { void* p = nullptr; B200_CUDA(cudaGetSymbolAddress(&p, g_trace)); B200_CUDA(cudaMemset(p, 0, sizeof(long long) * 24 * 128 * 4)); }
  return B200DS_OK;
}
#endif

int b200ds_tc_supported(int64_t batch, int n_features, int hidden) {
  FwdPlan fp;
  return tc_shape_ok(batch, n_features, hidden) && fwd_plan(batch, hidden, &fp) ? 1 : 0;
}

size_t b200ds_tc_xq_bytes(int64_t batch) { return batch > 0 ? (size_t)batch * KP * 2 : 0; }

size_t b200ds_tc_wq_bytes(int hidden) { return hidden > 0 ? (size_t)hidden * KP * 2 + (size_t)(hidden / kPackRows) * 32 * sizeof(float) : 0; }

int b200ds_pack_x_bf16_rows(const float* x_dev, int64_t rows, int64_t batch, int n_features, void* xq_dev, uintptr_t stream) {
  B200_REQUIRE(batch > 0 && batch % 8 == 0, "pack_x_bf16: batch must be a positive multiple of 8 (got %lld)", (long long)batch);
  B200_REQUIRE(rows >= 0 && rows <= batch, "pack_x_bf16: rows (%lld) must be in [0, batch = %lld]", (long long)rows, (long long)batch);
  B200_REQUIRE(n_features >= 1 && n_features + 2 <= KP, "pack_x_bf16: F must be in [1,%d] (got %d)", KP - 2, n_features);
  B200_REQUIRE(x_dev && xq_dev, "pack_x_bf16: null pointer");
  int64_t threads = batch * (KP / 8);
  pack_x_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, as_stream(stream)>>>(x_dev, rows, batch, n_features,
                                                                                  reinterpret_cast<uint4*>(xq_dev));
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

int b200ds_pack_x_bf16(const float* x_dev, int64_t batch, int n_features, void* xq_dev, uintptr_t stream) {
  return b200ds_pack_x_bf16_rows(x_dev, batch, batch, n_features, xq_dev, stream);
}

static int launch_gen_batch(void* xq_dev, float* y_dev, int64_t batch, int n_features, uint64_t seed, uint64_t offset,
                            const uint64_t* draw_counter_dev, uint64_t* offset_advance, uintptr_t stream) {
  B200_REQUIRE(batch > 0 && batch % 8 == 0, "gen_batch_bf16: batch must be a positive multiple of 8 (got %lld)", (long long)batch);
  B200_REQUIRE(n_features >= 1 && n_features + 2 <= KP, "gen_batch_bf16: F must be in [1,%d] (got %d)", KP - 2, n_features);
  B200_REQUIRE(offset % 4 == 0, "gen_batch_bf16: offset must be a multiple of 4");
  B200_REQUIRE(xq_dev && y_dev, "gen_batch_bf16: null pointer");
  int grid, calls;
  philox_plan(batch * n_features, &grid, &calls);         // torch's own geometry: the kernel IS that grid
  const uint2 key = make_uint2((unsigned)seed, (unsigned)(seed >> 32));
  const bool small = batch * n_features < (int64_t)1 << 31;
  const uint32_t nthreads = (uint32_t)grid * 256u;
  auto inv32 = [](uint64_t d) { uint64_t q = ((uint64_t)1 << 32) / d; return (uint32_t)(q > 0xFFFFFFFFull ? 0xFFFFFFFFull : q); };   // d = 1: 2^32 - 1
  const uint32_t inv_F = inv32((uint64_t)n_features), inv_nt = inv32((uint64_t)nthreads);
  static const int ctas_per_sm = getenv("B200DS_GEN_CTAS_PER_SM") ? atoi(getenv("B200DS_GEN_CTAS_PER_SM")) : 8;   // tuning: 8 = torch's own grid
  const int lgrid = grid < sm_count() * ctas_per_sm ? grid : sm_count() * ctas_per_sm;
  if (small)
    gen_batch_kernel<uint32_t><<<lgrid, 256, 0, as_stream(stream)>>>(
        reinterpret_cast<unsigned char*>(xq_dev), y_dev, batch, n_features, key, offset / 4, nthreads, calls, inv_F, inv_nt,
        reinterpret_cast<const unsigned long long*>(draw_counter_dev));
  else
    gen_batch_kernel<uint64_t><<<lgrid, 256, 0, as_stream(stream)>>>(
        reinterpret_cast<unsigned char*>(xq_dev), y_dev, batch, n_features, key, offset / 4, nthreads, calls, inv_F, inv_nt,
        reinterpret_cast<const unsigned long long*>(draw_counter_dev));
  B200_LAUNCH_CHECK();
  if (offset_advance) *offset_advance = (uint64_t)calls * 4;
  return B200DS_OK;
}

int b200ds_gen_batch_bf16(void* xq_dev, float* y_dev, int64_t batch, int n_features, uint64_t seed, uint64_t offset,
                          uint64_t* offset_advance, uintptr_t stream) {
  return launch_gen_batch(xq_dev, y_dev, batch, n_features, seed, offset, nullptr, offset_advance, stream);
}

int b200ds_gen_batch_bf16_counted(void* xq_dev, float* y_dev, int64_t batch, int n_features, uint64_t seed,
                                  uint64_t offset0, const uint64_t* draw_counter_dev, uintptr_t stream) {
  B200_REQUIRE(draw_counter_dev != nullptr, "gen_batch_bf16_counted: null draw counter");
  return launch_gen_batch(xq_dev, y_dev, batch, n_features, seed, offset0, draw_counter_dev, nullptr, stream);
}

int b200ds_pack_params_bf16(const float* params_dev, int n_features, int hidden, void* wq_dev, uintptr_t stream) {
  B200_REQUIRE(n_features >= 1 && n_features + 2 <= KP, "pack_params_bf16: F must be in [1,%d] (got %d)", KP - 2, n_features);
  B200_REQUIRE(hidden >= 128 && hidden % 128 == 0, "pack_params_bf16: H must be a multiple of 128 (got %d)", hidden);
  B200_REQUIRE(params_dev && wq_dev, "pack_params_bf16: null pointer");
  FwdPlan fp;
  B200_REQUIRE(fwd_plan(128, hidden, &fp), "pack_params_bf16: H = %d cannot be sliced into <= %d resident slices", hidden, kMaxSlices);
  Layout L{n_features, hidden};
  unsigned char* w1q = reinterpret_cast<unsigned char*>(wq_dev);
  float* vpart = reinterpret_cast<float*>(w1q + (size_t)hidden * KP * 2);
  pack_w1_kernel<<<hidden / kPackRows, 4 * kPackRows, 0, as_stream(stream)>>>(params_dev, L, reinterpret_cast<uint4*>(wq_dev), vpart);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

size_t b200ds_mlp_forward_tc_scratch_bytes(int64_t batch, int n_features, int hidden) {
  FwdPlan fp;
  if (!tc_shape_ok(batch, n_features, hidden) || !fwd_plan(batch, hidden, &fp) || fp.slices == 1) return 0;
  return (size_t)fp.slices * (size_t)batch * sizeof(float) + (size_t)fp.n_btiles * sizeof(unsigned int);
}

int b200ds_mlp_forward_tc(const void* xq_dev, const void* wq_dev, const float* params_dev, int64_t batch,
                          int n_features, int hidden, float* logits_dev, void* scratch_dev, size_t scratch_bytes,
                          uintptr_t stream) {
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
    B200_CUDA(cudaFuncSetAttribute(mlp_fwd_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    B200_CUDA(cudaFuncSetAttribute(mlp_fwd_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const size_t need = b200ds_mlp_forward_tc_scratch_bytes(batch, n_features, hidden);
  B200_REQUIRE(need == 0 || (scratch_dev && scratch_bytes >= need), "mlp_forward_tc: scratch too small (%zu < %zu)", scratch_bytes, need);
  float* zpart = reinterpret_cast<float*>(scratch_dev);
  unsigned int* tickets = need ? reinterpret_cast<unsigned int*>(zpart + (size_t)fp.slices * batch) : nullptr;
  const unsigned char* w1q = reinterpret_cast<const unsigned char*>(wq_dev);
  const float* v = reinterpret_cast<const float*>(w1q + (size_t)hidden * KP * 2);
  auto kern = (fp.slice_rows / kTileRows) % 2 == 0 ? mlp_fwd_tc_kernel<2> : mlp_fwd_tc_kernel<1>;
  kern<<<fp.slices * fp.ctas_per_slice, kFwdThreads, fp.smem, st>>>(
      reinterpret_cast<const unsigned char*>(xq_dev), w1q, v, params_dev + L.off_w2(), params_dev + L.off_b2(), logits_dev,
      zpart, tickets, fp.n_btiles, fp.slice_rows, fp.slices, fp.ctas_per_slice);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

size_t b200ds_mlp_backward_tc_scratch_bytes(int64_t batch, int n_features, int hidden) {
  if (!tc_shape_ok(batch, n_features, hidden)) return 0;
  BwdPlan bp;
  bwd_plan(batch, hidden, &bp);
  return (size_t)bp.splits * hidden * 32 * sizeof(float) + 2 * (size_t)bp.splits * sizeof(float);
}

static int launch_backward_tc(const char* who, const void* xq_dev, const void* wq_dev, const float* params_dev,
                              const float* dz_or_logits_dev, const float* y_dev, float* loss_dev, uint64_t* step_counter_dev,
                              int64_t batch, int n_features, int hidden, float* grads_dev, void* scratch_dev,
                              size_t scratch_bytes, bool finish, uintptr_t stream) {
  const bool fused = y_dev != nullptr;
  B200_REQUIRE(tc_shape_ok(batch, n_features, hidden),
               "%s: needs F <= %d, H %% 128 == 0, batch %% 128 == 0 (got B=%lld F=%d H=%d)", who, KP - 2,
               (long long)batch, n_features, hidden);
  B200_REQUIRE(xq_dev && wq_dev && dz_or_logits_dev, "%s: null pointer", who);
  B200_REQUIRE(!finish || (params_dev && grads_dev), "%s: null pointer", who);
  B200_REQUIRE((((uintptr_t)dz_or_logits_dev) & 15) == 0 && (((uintptr_t)y_dev) & 15) == 0,
               "%s: logits / dlogits / labels must be 16-byte aligned", who);
  BwdPlan bp;
  bwd_plan(batch, hidden, &bp);
  size_t need = b200ds_mlp_backward_tc_scratch_bytes(batch, n_features, hidden);
  B200_REQUIRE(scratch_dev && scratch_bytes >= need, "%s: scratch too small (%zu < %zu)", who, scratch_bytes, need);
  Layout L{n_features, hidden};
  cudaStream_t st = as_stream(stream);
  static bool attr_set[64];
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B200_CUDA(cudaFuncSetAttribute(mlp_bwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    B200_CUDA(cudaFuncSetAttribute(mlp_bwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const unsigned char* w1q = reinterpret_cast<const unsigned char*>(wq_dev);
  float* db2_part = reinterpret_cast<float*>(scratch_dev) + (size_t)bp.splits * hidden * 32;
  float* loss_part = db2_part + bp.splits;
  // without the finish pass the step counter is advanced by the backward kernel itself
  unsigned long long* ctr_in_kernel = finish ? nullptr : reinterpret_cast<unsigned long long*>(step_counter_dev);
  if (fused)
    mlp_bwd_tc_kernel<true><<<bp.groups * bp.splits, kBwdLossThreads, bp.smem, st>>>(
        reinterpret_cast<const unsigned char*>(xq_dev), w1q, dz_or_logits_dev, y_dev, (float)batch,
        reinterpret_cast<float*>(scratch_dev), db2_part, loss_part, ctr_in_kernel, bp.n_btiles, hidden, bp.nb, bp.groups, bp.splits);
  else
    mlp_bwd_tc_kernel<false><<<bp.groups * bp.splits, kBwdThreads, bp.smem, st>>>(
        reinterpret_cast<const unsigned char*>(xq_dev), w1q, dz_or_logits_dev, nullptr, (float)batch,
        reinterpret_cast<float*>(scratch_dev), db2_part, loss_part, ctr_in_kernel, bp.n_btiles, hidden, bp.nb, bp.groups, bp.splits);
  B200_LAUNCH_CHECK();
  if (!finish) return B200DS_OK;
  int fgrid = (hidden + 7) / 8;
  if (fgrid > sm_count() * 8) fgrid = sm_count() * 8;
  mlp_bwd_tc_finish<<<fgrid, 256, 0, st>>>(reinterpret_cast<const float*>(scratch_dev), db2_part, bp.splits, params_dev, L,
                                           grads_dev, fused ? loss_part : nullptr, 1.0 / (double)batch, loss_dev,
                                           reinterpret_cast<unsigned long long*>(step_counter_dev));
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

int b200ds_mlp_backward_tc(const void* xq_dev, const void* wq_dev, const float* params_dev,
                           const float* dlogits_dev, int64_t batch, int n_features, int hidden,
                           float* grads_dev, void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  return launch_backward_tc("mlp_backward_tc", xq_dev, wq_dev, params_dev, dlogits_dev, nullptr, nullptr, nullptr, batch,
                            n_features, hidden, grads_dev, scratch_dev, scratch_bytes, true, stream);
}

int b200ds_mlp_loss_backward_tc(const void* xq_dev, const void* wq_dev, const float* params_dev,
                                const float* logits_dev, const float* y_dev, int64_t batch, int n_features, int hidden,
                                float* loss_dev, float* grads_dev, uint64_t* step_counter_dev, void* scratch_dev,
                                size_t scratch_bytes, uintptr_t stream) {
  B200_REQUIRE(y_dev && loss_dev, "mlp_loss_backward_tc: null pointer");
  return launch_backward_tc("mlp_loss_backward_tc", xq_dev, wq_dev, params_dev, logits_dev, y_dev, loss_dev, step_counter_dev,
                            batch, n_features, hidden, grads_dev, scratch_dev, scratch_bytes, true, stream);
}

int b200ds_mlp_loss_backward_partials_tc(const void* xq_dev, const void* wq_dev, const float* logits_dev, const float* y_dev,
                                         int64_t batch, int n_features, int hidden, uint64_t* step_counter_dev,
                                         void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  B200_REQUIRE(y_dev != nullptr, "mlp_loss_backward_partials_tc: null pointer");
  return launch_backward_tc("mlp_loss_backward_partials_tc", xq_dev, wq_dev, nullptr, logits_dev, y_dev, nullptr, step_counter_dev,
                            batch, n_features, hidden, nullptr, scratch_dev, scratch_bytes, false, stream);
}

int64_t b200ds_update_tc_exchange_elems(int hidden) { return (int64_t)hidden * 32 + 4; }

// which = 1: finish + publish (or the whole update for one rank), 2: gate + collect + Adam + pack, 3: both
static int launch_update_tc(int which, const char* who, b200ds_ctx* ctx, const void* scratch_dev, size_t scratch_bytes, int64_t batch,
                            int n_features, int hidden, float* params_dev, float* grads_dev, float* exp_avg_dev,
                            float* exp_avg_sq_dev, void* wq_dev, double lr, double beta1, double beta2, double eps, int64_t step,
                            const uint64_t* step_counter_dev, float* loss_dev, uintptr_t stream) {
  B200_REQUIRE(ctx != nullptr, "%s: null context", who);
  B200_REQUIRE(tc_shape_ok(batch, n_features, hidden),
               "%s: needs F <= %d, H %% 128 == 0, batch %% 128 == 0 (got B=%lld F=%d H=%d)", who, KP - 2,
               (long long)batch, n_features, hidden);
  B200_REQUIRE(params_dev && grads_dev && exp_avg_dev && exp_avg_sq_dev && wq_dev, "%s: null pointer", who);
  B200_REQUIRE(step_counter_dev != nullptr || step >= 1, "%s: step must be >= 1 (t after the increment)", who);
  BwdPlan bp;
  bwd_plan(batch, hidden, &bp);
  size_t need = b200ds_mlp_backward_tc_scratch_bytes(batch, n_features, hidden);
  B200_REQUIRE(scratch_dev && scratch_bytes >= need, "%s: scratch too small (%zu < %zu)", who, scratch_bytes, need);
  const int W = ctx->world;
  B200_REQUIRE(W == 1 || ctx->p2p_cap >= b200ds_update_tc_exchange_elems(hidden),
               "%s: %d ranks need b200ds_ctx_enable_p2p(ctx, >= %lld) first", who, W,
               (long long)b200ds_update_tc_exchange_elems(hidden));
  if (W > 1) { if (int rc = p2p_check_mailbox(ctx, who)) return rc; }
  Layout L{n_features, hidden};
  const float* partial = reinterpret_cast<const float*>(scratch_dev);
  const float* db2_part = partial + (size_t)bp.splits * hidden * 32;
  const float* loss_part = loss_dev ? db2_part + bp.splits : nullptr;
  unsigned char* w1q = reinterpret_cast<unsigned char*>(wq_dev);
  float* vpart = reinterpret_cast<float*>(w1q + (size_t)hidden * KP * 2);
  const AdamScalars s = make_adam_scalars(beta1, beta2, eps, W > 1 ? (float)(1.0 / W) : 1.f);
  const P2PDev d = make_p2p_dev(ctx);
  const int grid = hidden / kPackRows, threads = 32 * kPackRows;
  const unsigned long long* ctr = reinterpret_cast<const unsigned long long*>(step_counter_dev);
  cudaStream_t st = as_stream(stream);
#define B200_UPD(COUNTED, PHASE)                                                                                        \
  mlp_update_tc_kernel<COUNTED, PHASE><<<grid, threads, 0, st>>>(partial, db2_part, loss_part, bp.splits, 1.0 / (double)batch, \
      params_dev, grads_dev, exp_avg_dev, exp_avg_sq_dev, L, w1q, vpart, s, lr, beta1, beta2, COUNTED ? 0 : step, COUNTED ? ctr : nullptr, loss_dev, d)
  if (W == 1) {
    if (which & 1) { if (ctr) B200_UPD(true, 0); else B200_UPD(false, 0); }
  } else {
    if (which & 1) { if (ctr) B200_UPD(true, 1); else B200_UPD(false, 1); B200_LAUNCH_CHECK(); }
    if (which & 2) {
      p2p_gate_kernel<<<1, 32, 0, st>>>(d);
      B200_LAUNCH_CHECK();
      if (ctr) B200_UPD(true, 2); else B200_UPD(false, 2);
    }
  }
#undef B200_UPD
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

int b200ds_update_tc(b200ds_ctx* ctx, const void* scratch_dev, size_t scratch_bytes, int64_t batch, int n_features, int hidden,
                     float* params_dev, float* grads_dev, float* exp_avg_dev, float* exp_avg_sq_dev, void* wq_dev,
                     double lr, double beta1, double beta2, double eps, int64_t step, const uint64_t* step_counter_dev,
                     float* loss_dev, uintptr_t stream) {
  return launch_update_tc(3, "update_tc", ctx, scratch_dev, scratch_bytes, batch, n_features, hidden, params_dev, grads_dev, exp_avg_dev,
                          exp_avg_sq_dev, wq_dev, lr, beta1, beta2, eps, step, step_counter_dev, loss_dev, stream);
}

int b200ds_update_tc_begin(b200ds_ctx* ctx, const void* scratch_dev, size_t scratch_bytes, int64_t batch, int n_features, int hidden,
                           float* params_dev, float* grads_dev, float* exp_avg_dev, float* exp_avg_sq_dev, void* wq_dev,
                           double lr, double beta1, double beta2, double eps, int64_t step, const uint64_t* step_counter_dev,
                           float* loss_dev, uintptr_t stream) {
  return launch_update_tc(1, "update_tc_begin", ctx, scratch_dev, scratch_bytes, batch, n_features, hidden, params_dev, grads_dev,
                          exp_avg_dev, exp_avg_sq_dev, wq_dev, lr, beta1, beta2, eps, step, step_counter_dev, loss_dev, stream);
}

int b200ds_update_tc_end(b200ds_ctx* ctx, const void* scratch_dev, size_t scratch_bytes, int64_t batch, int n_features, int hidden,
                         float* params_dev, float* grads_dev, float* exp_avg_dev, float* exp_avg_sq_dev, void* wq_dev,
                         double lr, double beta1, double beta2, double eps, int64_t step, const uint64_t* step_counter_dev,
                         float* loss_dev, uintptr_t stream) {
  return launch_update_tc(2, "update_tc_end", ctx, scratch_dev, scratch_bytes, batch, n_features, hidden, params_dev, grads_dev,
                          exp_avg_dev, exp_avg_sq_dev, wq_dev, lr, beta1, beta2, eps, step, step_counter_dev, loss_dev, stream);
}

}  // extern "C"
