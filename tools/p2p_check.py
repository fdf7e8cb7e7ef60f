#!/usr/bin/env python3
"""Multi-GPU check of b200ds_allreduce_adam_p2p (one kernel: peer-memory all-reduce + Adam)
against b200ds_allreduce_bucket + b200ds_adam_step (NCCL, then Adam), under torchrun:

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/p2p_check.py

Same per-rank gradients into both paths for several steps; the parameters must agree to fp32
round-off (the summation order differs), the peer-memory path's replicas must be bit-equal
across ranks, and the bucket must hold the sum afterwards.  Prints the time per call of both.
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch
import torch.distributed as dist
from b200ds import _lib, comm


def main():
    lib = _lib.load()
    le = comm.read_launch_env()
    if le.world_size > 1:
        comm.init_distributed()
    dev = torch.device("cuda", le.local_rank)
    torch.cuda.set_device(dev)
    W, rank = le.world_size, le.rank

    def nccl_id():
        buf = C.create_string_buffer(128)
        _lib.check(lib.b200ds_nccl_unique_id(buf))
        return buf.raw

    nid = comm.exchange_nccl_id(nccl_id, le)
    idbuf = C.create_string_buffer(nid, 128) if nid else None
    ctx = C.c_void_p()
    _lib.check(lib.b200ds_ctx_create(le.local_rank, rank, W, C.cast(idbuf, C.c_void_p) if idbuf else None, C.byref(ctx)))
    st = torch.cuda.current_stream().cuda_stream
    args = (1e-3, 0.9, 0.999, 1e-8)
    ok = True
    n_max = 1 << 20
    _lib.check(lib.b200ds_ctx_enable_p2p(ctx, n_max))
    for n in (4, 90116, n_max):
        torch.manual_seed(7)                                  # same initial parameters on every rank
        pa = torch.randn(n, device=dev); pb = pa.clone()
        ma, va, mb, vb = (torch.zeros(n, device=dev) for _ in range(4))
        gen = torch.Generator(device=dev); gen.manual_seed(100 + rank)
        worst = 0.0
        for step in range(1, 8):
            g = torch.randn(n, device=dev, generator=gen)
            ga, gb = g.clone(), g.clone()
            _lib.check(lib.b200ds_allreduce_bucket(ctx, ga.data_ptr(), n, st))
            _lib.check(lib.b200ds_adam_step(pa.data_ptr(), ga.data_ptr(), ma.data_ptr(), va.data_ptr(), n, *args, step, 1.0 / W, 0, st))
            _lib.check(lib.b200ds_allreduce_adam_p2p(ctx, pb.data_ptr(), gb.data_ptr(), mb.data_ptr(), vb.data_ptr(), n, *args,
                                                     step, None, st))
            torch.cuda.synchronize()
            worst = max(worst, float((ga - gb).abs().max() / ga.abs().max()))
        perr = float((pa - pb).abs().max() / pa.abs().max())
        same = True
        if W > 1:
            gathered = [torch.zeros_like(pb) for _ in range(W)]
            dist.all_gather(gathered, pb)
            same = all(torch.equal(gathered[0], t) for t in gathered)
        good = perr <= 1e-5 and worst <= 1e-5 and same
        ok = ok and good
        # time per call (eager launches, CUDA events, after warm-up; every rank runs the same count)
        reps = 200
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        if W > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for i in range(reps):
            lib.b200ds_allreduce_bucket(ctx, ga.data_ptr(), n, st)
            lib.b200ds_adam_step(pa.data_ptr(), ga.data_ptr(), ma.data_ptr(), va.data_ptr(), n, *args, 8 + i, 1.0 / W, 0, st)
        e1.record()
        for i in range(reps):
            lib.b200ds_allreduce_adam_p2p(ctx, pb.data_ptr(), gb.data_ptr(), mb.data_ptr(), vb.data_ptr(), n, *args, 8 + i, None, st)
        e2.record()
        torch.cuda.synchronize()
        if rank == 0:
            print(f"p2p_check n={n} world={W}: bucket sum rel diff {worst:.1e}, param rel diff {perr:.1e}, replicas bit-equal {same}"
                  f" -> {'OK' if good else 'FAIL'};  NCCL all-reduce + Adam {1e3 * e0.elapsed_time(e1) / reps:.1f} us/call,"
                  f" peer-memory kernel {1e3 * e1.elapsed_time(e2) / reps:.1f} us/call", flush=True)
    if W > 1:
        dist.barrier()
    _lib.check(lib.b200ds_ctx_destroy(ctx))
    if W > 1:
        dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
