#!/usr/bin/env python3
"""Multi-GPU check of the ZeRO-1-style step (BASELINE config 5): under torchrun, train the
train.py model for a few steps with stage 0 and with stage 1 from the same initial weights and
the same rank-sharded data; the parameters must agree to fp32 round-off on every rank.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/zero1_check.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch
import torch.distributed as dist
import b200ds


def run(stage, steps, F, H, B, rank, world, dev):
    torch.manual_seed(42)
    net = torch.nn.Sequential(torch.nn.Linear(F, H), torch.nn.ReLU(), torch.nn.Linear(H, 1))
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)
    cfg = {"train_batch_size": B * world, "gradient_accumulation_steps": 1, "fp16": {"enabled": False},
           "zero_optimization": {"stage": stage}}
    engine, _, _, _ = b200ds.initialize(model=net, optimizer=opt, config=cfg)
    crit = b200ds.BCEWithLogitsLoss()
    g = torch.Generator().manual_seed(1234 + rank)
    losses = []
    for _ in range(steps):
        x = torch.randn(B, F, generator=g)
        y = (x[:, : F // 2].sum(dim=1, keepdim=True) > 0).float()
        loss = crit(engine(x.to(dev)), y.to(dev))
        engine.backward(loss)
        engine.step()
        losses.append(loss.item())
    return engine.flat[: engine.P].clone(), losses


def main():
    le = b200ds.init_distributed()
    dev = torch.device("cuda", le.local_rank)
    torch.cuda.set_device(dev)
    ok = True
    for (F, H, B) in ((20, 64, 4), (20, 512, 256)):
        p0, l0 = run(0, 8, F, H, B, le.rank, le.world_size, dev)
        p1, l1 = run(1, 8, F, H, B, le.rank, le.world_size, dev)
        err = float((p0 - p1).abs().max() / p0.abs().max())
        lerr = max(abs(a - b) for a, b in zip(l0, l1))
        gathered = [torch.zeros_like(p1) for _ in range(le.world_size)]
        dist.all_gather(gathered, p1)
        same = all(torch.equal(gathered[0], t) for t in gathered)
        good = err <= 1e-5 and lerr <= 1e-5 and same
        ok = ok and good
        if le.rank == 0:
            print(f"zero1_check F={F} H={H} B={B} world={le.world_size}: max rel param diff vs stage 0 {err:.2e}, "
                  f"max loss diff {lerr:.2e}, replicas identical {same} -> {'OK' if good else 'FAIL'}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
