#!/usr/bin/env python3
"""Gradient-bucket collective sweep (BASELINE config 5 / SURVEY 8f-4): latency, algorithm and bus bandwidth against
bucket size, 4 KiB .. 256 MiB in x4 steps (plus the benchmark's own buckets), for
  * ncclAllReduce, fp32 and bf16                       (b200ds_collective op 0: the stage-0 exchange of train.py:113)
  * ncclReduceScatter + ncclAllGather, fp32 and bf16   (ops 1, 2: the ZeRO-1-style exchange of b200ds_zero1_step)
  * the library's own one-launch peer-memory all-reduce + Adam kernel (b200ds_allreduce_adam_p2p), up to 16 MiB --
    it is built for latency-bound buckets; its row includes the Adam arithmetic
Launch with torchrun, one rank per GPU:

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/allreduce_sweep.py [--quick]

busbw: all-reduce algbw * 2 (W-1)/W; reduce-scatter / all-gather algbw * (W-1)/W each (bytes = the full bucket), to be
read against 900 GB/s per direction per GPU (NVLink 5 nominal; measured 8-rank all-reduce reference on this pool:
725 GB/s at 1 GiB).  Prints one JSON document on rank 0.  It is also what the reference is for (README.md:3: a check
that NCCL works over the fabric): a run that completes with sane bandwidths is the acceptance signal."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch
import torch.distributed as dist
from b200ds import _lib, comm

OPS = {"allreduce": 0, "reduce_scatter": 1, "all_gather": 2}
DTYPES = {"f32": (0, torch.float32, 4), "bf16": (1, torch.bfloat16, 2)}


def sweep_sizes(quick):
    sizes = [1 << k for k in range(12, 29, 2)]                              # 4 KiB .. 256 MiB
    sizes += [4 * (20 * 4096 + 2 * 4096 + 4), 4 * (20 * 8192 + 2 * 8192 + 4), 4 * (4096 * 4096 + 2 * 4096 + 4)]   # wide4096 / wide8192 / wideFH4096 buckets
    if quick:
        sizes = [1 << 12, 1 << 18, 1 << 24, 4 * (20 * 4096 + 2 * 4096 + 4)]
    return sorted(set(sizes))


def main():
    quick = "--quick" in sys.argv
    lib = _lib.load()
    le = comm.init_distributed()
    dev = torch.device("cuda", le.local_rank)
    torch.cuda.set_device(dev)
    ctx = C.c_void_p()

    def make_id():
        buf = C.create_string_buffer(128)
        _lib.check(lib.b200ds_nccl_unique_id(buf))
        return buf.raw
    nid = comm.exchange_nccl_id(make_id, le)
    idbuf = C.create_string_buffer(nid, 128) if nid else None
    _lib.check(lib.b200ds_ctx_create(le.local_rank, le.rank, le.world_size, C.cast(idbuf, C.c_void_p) if idbuf else None, C.byref(ctx)))
    st = torch.cuda.current_stream().cuda_stream
    W = le.world_size
    p2p_max = 16 << 20
    p2p_ok = W > 1 and lib.b200ds_ctx_enable_p2p(ctx, p2p_max // 4) == 0
    if W > 1:
        okf = torch.tensor([1.0 if p2p_ok else 0.0], device=dev)
        dist.all_reduce(okf, op=dist.ReduceOp.MIN)
        p2p_ok = float(okf) > 0

    def timed(fn, nbytes):
        iters = 200 if nbytes <= (1 << 22) else 20
        for _ in range(5):
            fn()
        if W > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=dev)
        if W > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t) * 1e3                                               # us, slowest rank

    rows = []
    for nbytes in sweep_sizes(quick):
        for dname, (dcode, tdt, esz) in DTYPES.items():
            n = nbytes // esz // (4 * W) * (4 * W)
            if n == 0:
                continue
            buf = torch.ones(n, dtype=tdt, device=dev)
            for oname, ocode in OPS.items():
                us = timed(lambda: _lib.check(lib.b200ds_collective(ctx, ocode, dcode, buf.data_ptr(), n, st)), nbytes)
                algbw = n * esz / (us * 1e-6) / 1e9
                factor = 2.0 * (W - 1) / W if oname == "allreduce" else (W - 1) / W
                rows.append(dict(op=oname, dtype=dname, bytes=n * esz, us=round(us, 2), algbw_gbs=round(algbw, 2), busbw_gbs=round(algbw * factor, 2)))
                buf.fill_(1)
            del buf
        if p2p_ok and nbytes <= p2p_max:
            n = nbytes // 4 // 4 * 4
            p, g, m, v = (torch.zeros(n, device=dev) for _ in range(4))
            g.fill_(1e-3)
            us = timed(lambda: _lib.check(lib.b200ds_allreduce_adam_p2p(ctx, p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), n,
                                                                        1e-3, 0.9, 0.999, 1e-8, 1, None, st)), nbytes)
            algbw = n * 4 / (us * 1e-6) / 1e9
            rows.append(dict(op="p2p_allreduce_adam_kernel", dtype="f32", bytes=n * 4, us=round(us, 2), algbw_gbs=round(algbw, 2),
                             busbw_gbs=round(algbw * 2 * (W - 1) / W, 2)))
            del p, g, m, v
    if le.rank == 0:
        print(json.dumps(dict(tool="allreduce_sweep", world=W, nvlink_nominal_gbs_per_direction=900, rows=rows)))
    _lib.check(lib.b200ds_ctx_destroy(ctx))
    if W > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
