"""CPU oracle for the on-device Philox batch generator -- TEST INFRASTRUCTURE only
(same import rule as oracle/reference_step.py).

The reference draws its data on the CPU generator (train.py:20-21).  The build's
generated-batch configs (BASELINE config 4) instead define the batch as
`torch.randn(B, F, device='cuda')` under a given (seed, offset), i.e. the integer stream
and the element mapping of ATen's CUDA normal kernel:

  * Philox4x32-10, key = seed, counter = (offset/4 + call, subsequence = global thread id)
      -- curand_philox4x32_x.h:149-195 and curand_init/skipahead in curand_kernel.h
  * block 256, grid = min(SMs * (maxThreadsPerSM/256), ceil(n/256)), unroll 4,
    element li = tid + nthreads*(4*call + lane4), offset advance ((n-1)/(256*grid*4)+1)*4
      -- ATen/native/cuda/DistributionTemplates.h:50-62,72-92
  * Box-Muller: u = x*2^-32 + 2^-33, v = y*2^-32*2pi + 2^-33*2pi,
    (sin v, cos v) * sqrt(-2 ln u)          -- curand_normal.h:70-88

The uint32 stream and the element mapping are integer work and are exact here; the
float32 normals go through numpy's log/sin/cos instead of CUDA's logf/__sincosf, so
they agree with the device only to a few ulp.  The bit-exact check of the floats is the
`-m gpu` test against torch.randn(device='cuda') itself.
"""
from __future__ import annotations

import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = np.uint32(0x9E3779B9)
_W1 = np.uint32(0xBB67AE85)
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr: np.ndarray, key: tuple[int, int]) -> np.ndarray:
    """ctr: uint32 [..., 4] (x,y,z,w); key: (lo, hi).  Returns uint32 [..., 4]."""
    c = [ctr[..., i].astype(np.uint64) for i in range(4)]
    k0 = np.uint32(key[0] & 0xFFFFFFFF)
    k1 = np.uint32(key[1] & 0xFFFFFFFF)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = _M0 * c[0]
            p1 = _M1 * c[2]
            hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
            hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
            c = [hi1 ^ c[1] ^ np.uint64(k0), lo1, hi0 ^ c[3] ^ np.uint64(k1), lo0]
            k0 = np.uint32((int(k0) + int(_W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(_W1)) & 0xFFFFFFFF)
    return np.stack([v.astype(np.uint32) for v in c], axis=-1)


def launch_plan(numel: int, sm_count: int, max_threads_per_sm: int = 2048):
    """(grid, nthreads, calls_per_thread, offset_advance) of ATen's calc_execution_policy."""
    block = 256
    grid = min(sm_count * (max_threads_per_sm // block), (numel + block - 1) // block)
    nthreads = grid * block
    calls = (numel - 1) // (nthreads * 4) + 1
    return grid, nthreads, calls, calls * 4


def randn_u32(numel: int, seed: int, offset: int, sm_count: int) -> np.ndarray:
    """uint32 [numel, 3]: for every output element the two Philox words its Box-Muller
    consumes and which output it takes (0 = sin branch, 1 = cos branch)."""
    assert offset % 4 == 0
    _, nthreads, _, _ = launch_plan(numel, sm_count)
    li = np.arange(numel, dtype=np.int64)
    tid = li % nthreads
    t = li // nthreads
    call, lane4 = t // 4, t % 4
    c64 = offset // 4 + call
    ctr = np.stack([(c64 & 0xFFFFFFFF), (c64 >> 32), (tid & 0xFFFFFFFF), (tid >> 32)], axis=-1).astype(np.uint32)
    out = philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))
    pair = lane4 // 2
    a = np.where(pair == 0, out[:, 0], out[:, 2])
    b = np.where(pair == 0, out[:, 1], out[:, 3])
    return np.stack([a, b, (lane4 % 2).astype(np.uint32)], axis=-1)


def box_muller(a: np.ndarray, b: np.ndarray, branch: np.ndarray) -> np.ndarray:
    """float32 restatement of _curand_box_muller; libm instead of CUDA intrinsics."""
    inv = np.float32(2.3283064e-10)
    inv2pi = np.float32(np.float32(2.3283064e-10) * np.float32(6.2831855))
    u = a.astype(np.float32) * inv + np.float32(inv / np.float32(2))
    v = b.astype(np.float32) * inv2pi + np.float32(inv2pi / np.float32(2))
    s = np.sqrt(np.float32(-2.0) * np.log(u.astype(np.float32))).astype(np.float32)
    r = np.where(branch == 0, np.sin(v), np.cos(v)).astype(np.float32)
    # ATen's normal_() ends with rand * std + mean (std=1, mean=0): turns the -0.0 of u == 1 into +0.0
    return ((r * s).astype(np.float32) * np.float32(1.0) + np.float32(0.0)).astype(np.float32)


def randn(numel: int, seed: int, offset: int, sm_count: int) -> np.ndarray:
    w = randn_u32(numel, seed, offset, sm_count)
    return box_muller(w[:, 0], w[:, 1], w[:, 2])


def labels(x: np.ndarray) -> np.ndarray:
    """train.py:22-23 on a generated batch, sequential fp32 sum over the first F//2 columns."""
    f = x.shape[1]
    acc = np.zeros(x.shape[0], dtype=np.float32)
    for j in range(f // 2):
        acc = (acc + x[:, j]).astype(np.float32)
    return (acc > 0).astype(np.float32)
