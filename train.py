#!/usr/bin/env python3
"""Drop-in for the reference's train.py on B200: same CLI, same launcher environment,
same engine-style calls, same stdout -- with the step executed by libb200ds.so.

    python train.py [--train-batch-size N] [--epochs N] [--lr F] [--seed N] [--save-dir S]
    torchrun --nproc-per-node 8 train.py ...          (RANK/LOCAL_RANK/WORLD_SIZE/MASTER_*)
    srun ... python train.py ...                      (SLURM_PROCID/SLURM_LOCALID/SLURM_NTASKS)

Kept from the reference (file:line into it):
  * flags and their defaults, unknown flags ignored                      train.py:42-49
  * seed -> dataset (2000 x 20) -> model (20-64-1) -> Adam(lr), this order  train.py:54,63,76,85
  * per-rank batch of 4, DistributedSampler(shuffle=True) when world > 1    train.py:65-74
  * ds_config (batch 64, GAS 1, fp16 off, ZeRO stage 0)                     train.py:78-83
  * initialize -> engine; per step engine(x), criterion, backward, step, loss.item()
                                                                            train.py:87-117
  * "Epoch {e:2d} finished. Avg loss: {v:.4f}" / "Training completed." from LOCAL_RANK 0,
    rank-local mean of the per-step losses                                  train.py:119-128
  * a checkpoint directory per epoch on every rank                          train.py:123-125
  * best-effort barrier + destroy_process_group at exit                     train.py:133-140
Changed on purpose: world size 1 runs instead of raising at sampler.set_epoch (train.py:101);
batches come from the device-resident copy of the dataset instead of DataLoader + .to(device).
Extension flags (all default off, the reference ignores unknown flags the same way):
  --fused-steps   run each epoch through the fused multi-step kernel (identical numbers)
  --no-checkpoint skip the per-epoch checkpoint
  --bf16          ds_config["bf16"] = {"enabled": True}: the tcgen05 kernels on bf16 operands (BASELINE config 3)
  --zero-stage N  ds_config["zero_optimization"]["stage"] = N (0 or 1; BASELINE config 5)
"""
import argparse
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(_ROOT, "distributed-gpu-test_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch.utils.data import TensorDataset  # noqa: E402
from torch.utils.data.distributed import DistributedSampler  # noqa: E402

import b200ds as deepspeed  # noqa: E402
from b200ds.accelerator import get_accelerator  # noqa: E402

N_SAMPLES, N_FEATURES, HIDDEN, RANK_BATCH = 2000, 20, 64, 4     # train.py:63,74,76


def cli(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--train-batch-size", type=int, default=64)   # parsed, unused: as in the reference
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--save-dir", type=str, default="ds_checkpoints")
    ap.add_argument("--fused-steps", action="store_true")
    ap.add_argument("--no-checkpoint", action="store_true")
    ap.add_argument("--bf16", action="store_true")
    ap.add_argument("--zero-stage", type=int, default=0)
    known, _unknown = ap.parse_known_args(argv)
    return known


def synthetic_dataset(n, f, seed):
    """The reference's generator calls in the reference's order: bit-identical data."""
    torch.manual_seed(seed)
    feats = torch.randn(n, f)
    labels = (feats[:, : f // 2].sum(dim=1) > 0).float().unsqueeze(1)
    return feats, labels


def build_model(f, h):
    class SimpleNet(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.net = torch.nn.Sequential(torch.nn.Linear(f, h), torch.nn.ReLU(), torch.nn.Linear(h, 1))

        def forward(self, inp):
            return self.net(inp)

    return SimpleNet()


def run(opts):
    torch.manual_seed(opts.seed)
    local_rank = int(os.environ.get("LOCAL_RANK", os.environ.get("SLURM_LOCALID", 0)))

    deepspeed.init_distributed()
    get_accelerator().set_device(local_rank)

    feats, labels = synthetic_dataset(N_SAMPLES, N_FEATURES, opts.seed)
    dataset = TensorDataset(feats, labels)
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    sampler = DistributedSampler(dataset, shuffle=True) if multi else None

    model = build_model(feats.shape[1], HIDDEN)
    ds_config = {
        "train_batch_size": 64,
        "gradient_accumulation_steps": 1,
        "fp16": {"enabled": False},
        "zero_optimization": {"stage": 0},
    }
    if opts.bf16:                  # BASELINE config 3: bf16 operands on the tensor cores (fp32 master weights)
        ds_config["bf16"] = {"enabled": True}
    if opts.zero_stage:            # BASELINE config 5: ZeRO-1-style sharded optimizer step
        ds_config["zero_optimization"] = {"stage": opts.zero_stage}
    optimizer = torch.optim.Adam(model.parameters(), lr=opts.lr)
    engine, optimizer, _, _ = deepspeed.initialize(args=ds_config, model=model, optimizer=optimizer,
                                                   model_parameters=model.parameters(), config=ds_config)
    device = engine.device
    loader = deepspeed.DeviceBatchLoader(feats, labels, batch_size=RANK_BATCH, device=device,
                                         sampler=sampler, shuffle=not multi)
    criterion = deepspeed.BCEWithLogitsLoss()

    for epoch in range(1, opts.epochs + 1):
        engine.train()
        if sampler is not None:
            sampler.set_epoch(epoch)
        running, steps = 0.0, 0
        if opts.fused_steps and engine.supports_fused_steps(RANK_BATCH):
            order = loader.begin_epoch()
            n_steps = len(loader)
            losses = engine.fused_steps(loader.x, loader.y, order, 0, RANK_BATCH, n_steps)
            for v in losses.cpu().tolist():       # one D2H per epoch; same fp64 running sum
                running += v
                steps += 1
        else:
            for bx, by in loader:
                logits = engine(bx)
                if logits.dim() == 1:
                    logits = logits.unsqueeze(1)
                loss = criterion(logits, by)
                engine.backward(loss)
                engine.step()
                running += loss.item()
                steps += 1
        mean_loss = running / max(1, steps)
        if local_rank == 0:
            print(f"Epoch {epoch:2d} finished. Avg loss: {mean_loss:.4f}", flush=True)
        if not opts.no_checkpoint:
            ckpt_dir = os.path.join(opts.save_dir, f"epoch{epoch}")
            os.makedirs(ckpt_dir, exist_ok=True)
            engine.save_checkpoint(ckpt_dir)

    if local_rank == 0:
        print("Training completed.", flush=True)


if __name__ == "__main__":
    run(cli())
    for teardown in (getattr(dist, "barrier", None), getattr(dist, "destroy_process_group", None)):
        try:
            if dist.is_initialized():
                teardown()
        except Exception:
            pass
