#!/usr/bin/env python3
"""Generate tests/golden/*.npz by RUNNING THE REFERENCE'S OWN CODE in this container.

The reference (/root/reference/train.py) imports `deepspeed`, which is not installed, so
a stub module is planted in sys.modules first; everything else is the reference's own
objects, unmodified: make_synthetic_data (train.py:19-24), SimpleNet (train.py:26-36),
parse_args defaults (train.py:42-49), plus the torch classes main() uses
(TensorDataset / DataLoader(batch_size=4) / DistributedSampler(shuffle=True),
nn.BCEWithLogitsLoss, torch.optim.Adam) driven in main()'s order (train.py:51-121).
The DeepSpeed engine calls (train.py:87-94,107,113,114) are replaced by what stage 0
does: module forward, loss.backward(), sum-then-scale gradient averaging over the
emulated ranks, optimizer.step(), zero_grad.  That last part is the unpinned bit.

/root/reference does not travel to the GPU box, hence committed fixtures.
Run:  python tests/golden/make_golden.py      (needs /root/reference, CPU only)
"""
import hashlib
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/train.py"


def load_reference():
    ds = types.ModuleType("deepspeed")
    acc = types.ModuleType("deepspeed.accelerator")
    acc.get_accelerator = lambda: None
    ds.accelerator = acc
    sys.modules["deepspeed"] = ds
    sys.modules["deepspeed.accelerator"] = acc
    spec = importlib.util.spec_from_file_location("ref_train", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def sha16(t: torch.Tensor) -> str:
    return hashlib.sha256(t.detach().contiguous().numpy().tobytes()).hexdigest()[:16]


def run_reference(ref, world: int, epochs: int, seed: int = 42, lr: float = 1e-3):
    """train.py main() order with `world` ranks emulated in one process."""
    from torch.utils.data import DataLoader, TensorDataset
    from torch.utils.data.distributed import DistributedSampler

    torch.manual_seed(seed)                                            # :54
    x, y = ref.make_synthetic_data(n_samples=2000, n_features=20, seed=seed)   # :63
    ds = TensorDataset(x, y)                                           # :65
    model = ref.SimpleNet(n_features=x.shape[1], hidden=64)            # :76
    init = [p.detach().clone() for p in model.parameters()]
    opt = torch.optim.Adam(model.parameters(), lr=lr)                  # :85
    crit = torch.nn.BCEWithLogitsLoss()                                # :96
    if world > 1:
        samplers = [DistributedSampler(ds, num_replicas=world, rank=r, shuffle=True) for r in range(world)]
        loaders = [DataLoader(ds, batch_size=4, sampler=s, shuffle=False) for s in samplers]   # :74
    else:
        samplers = [None]
        loaders = [DataLoader(ds, batch_size=4, sampler=None, shuffle=True)]
    step_losses = [[] for _ in range(world)]
    epoch_means = [[] for _ in range(world)]
    first_epoch_indices = None
    for epoch in range(1, epochs + 1):                                 # :99
        model.train()
        rng0 = torch.get_rng_state()
        iters = []
        for r in range(world):
            if samplers[r] is not None:
                samplers[r].set_epoch(epoch)                           # :101
            if world > 1:
                torch.set_rng_state(rng0)      # every rank is its own process with its own global RNG
            iters.append(iter(loaders[r]))
        sums = [0.0] * world
        steps = 0
        while True:
            try:
                bs = [next(it) for it in iters]
            except StopIteration:
                break
            grads = None
            for r, (bx, by) in enumerate(bs):
                logits = model(bx)                                     # :107
                loss = crit(logits, by)                                # :112
                opt.zero_grad(set_to_none=True)
                loss.backward()                                        # :113
                flat = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
                grads = flat.clone() if grads is None else grads + flat
                sums[r] += loss.item()                                 # :115
                step_losses[r].append(loss.item())
            if world > 1:
                grads = grads * (1.0 / world)
            o = 0
            for p in model.parameters():
                p.grad = grads[o:o + p.numel()].reshape(p.shape).clone()
                o += p.numel()
            opt.step()                                                 # :114
            steps += 1
        for r in range(world):
            epoch_means[r].append(sums[r] / max(1, steps))             # :119
    final = [p.detach().clone() for p in model.parameters()]
    return dict(x=x, y=y, init=init, final=final, step_losses=step_losses, epoch_means=epoch_means)


def main():
    torch.set_num_threads(1)
    ref = load_reference()
    sys.argv = ["train.py"]
    a = ref.parse_args()
    out = {
        "cli_defaults": np.array([a.train_batch_size, a.epochs, a.seed], dtype=np.int64),
        "cli_lr": np.array([a.lr], dtype=np.float64),
        "torch_version": np.array([torch.__version__]),
        "cpu_capability": np.array([torch.backends.cpu.get_cpu_capability()]),
    }
    for world, epochs in ((1, 5), (2, 5), (4, 5), (8, 5)):
        r = run_reference(ref, world, epochs)
        tag = f"w{world}"
        if world == 1:
            out["x"] = r["x"].numpy()
            out["y"] = r["y"].numpy()
            for i, name in enumerate(("w1", "b1", "w2", "b2")):
                out[f"init_{name}"] = r["init"][i].numpy()
            out["sha_x"] = np.array([sha16(r["x"])])
            out["sha_y"] = np.array([sha16(r["y"])])
            out["sha_init"] = np.array([sha16(t) for t in r["init"]])
        for i, name in enumerate(("w1", "b1", "w2", "b2")):
            out[f"{tag}_final_{name}"] = r["final"][i].numpy()
        out[f"{tag}_step_losses"] = np.array(r["step_losses"], dtype=np.float64)   # [rank, step]
        out[f"{tag}_epoch_means"] = np.array(r["epoch_means"], dtype=np.float64)   # [rank, epoch]
        print(tag, "epoch means rank0:", " ".join(f"{v:.4f}" for v in r["epoch_means"][0]),
              "steps", len(r["step_losses"][0]))
    np.savez_compressed(os.path.join(HERE, "train_golden.npz"), **out)
    print("x[0,:4] =", [float(v) for v in out["x"][0, :4]], "sum y =", float(out["y"].sum()))
    print("sha x/y:", out["sha_x"][0], out["sha_y"][0], "params:", list(out["sha_init"]))


if __name__ == "__main__":
    main()
