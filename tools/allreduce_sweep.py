#!/usr/bin/env python3
"""Gradient-bucket all-reduce sweep (BASELINE config 5 / SURVEY 8f-4): latency, algorithm and bus
bandwidth of b200ds_allreduce_bucket (NCCL over NVLink on the context's side stream) against
bucket size.  Launch with torchrun, one rank per GPU:

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/allreduce_sweep.py

busbw = algbw * 2 (W-1)/W, to be read against 900 GB/s per direction per GPU (NVLink 5 nominal;
measured 8-rank all-reduce reference on this pool: 725 GB/s at 1 GiB)."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch
import torch.distributed as dist
from b200ds import _lib, comm


def main():
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
    sizes = [1 << k for k in range(12, 29, 2)] + [4 * (20 * 4096 + 2 * 4096 + 1)]      # 4 KiB .. 256 MiB, and the H=4096 bucket
    rows = []
    for nbytes in sorted(sizes):
        n = nbytes // 4
        buf = torch.ones(n, dtype=torch.float32, device=dev)
        iters = 200 if nbytes <= (1 << 22) else 20
        for _ in range(10):
            _lib.check(lib.b200ds_allreduce_bucket(ctx, buf.data_ptr(), n, st))
        if W > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            _lib.check(lib.b200ds_allreduce_bucket(ctx, buf.data_ptr(), n, st))
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=dev)
        if W > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        us = float(t) * 1e3
        # the same exchange replayed from a CUDA graph (how bench.py runs the step)
        us_graph = None
        if os.environ.get("SWEEP_GRAPH") and nbytes <= (1 << 22):
            g = torch.cuda.CUDAGraph()
            cap = torch.cuda.Stream(device=dev)
            cap.wait_stream(torch.cuda.current_stream())
            with torch.cuda.graph(g, stream=cap):
                buf.mul_(1.0)
                _lib.check(lib.b200ds_allreduce_bucket(ctx, buf.data_ptr(), n, torch.cuda.current_stream().cuda_stream))
                buf.mul_(0.5)
            torch.cuda.current_stream().wait_stream(cap)
            for _ in range(20):
                g.replay()
            if W > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(iters):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            us_graph = round(e0.elapsed_time(e1) / iters * 1e3, 2)
            g.reset()
            del g
        algbw = nbytes / (us * 1e-6) / 1e9
        rows.append(dict(bytes=nbytes, us=round(us, 2), algbw_gbs=round(algbw, 2), busbw_gbs=round(algbw * 2 * (W - 1) / W, 2), graph_us_with_two_tiny_kernels=us_graph))
    if le.rank == 0:
        print(json.dumps(dict(tool="allreduce_sweep", world=W, dtype="f32", rows=rows)))
    _lib.check(lib.b200ds_ctx_destroy(ctx))
    if W > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
