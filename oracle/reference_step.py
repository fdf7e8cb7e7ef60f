"""CPU oracle for the train.py hot path -- TEST INFRASTRUCTURE, not product code.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product path (distributed-gpu-test_b200/) never
does; it fails loudly when the CUDA extension is missing.

What it restates (file:line into /root/reference):
  * make_dataset        <- train.py:19-24   (seed, randn, label rule)
  * init_params         <- train.py:26-33,76 (two nn.Linear, default init, creation order)
  * rank_indices        <- train.py:65-74,101,104 (DistributedSampler / shuffle=True protocol)
  * forward             <- train.py:35-36
  * bce_with_logits     <- train.py:96,112  (nn.BCEWithLogitsLoss, mean reduction)
  * backward            <- train.py:113     (autograd of the above; no dx)
  * average_gradients   <- train.py:113     (engine.backward: stage-0 flat-bucket all-reduce)
  * adam_update         <- train.py:85,114  (torch.optim.Adam defaults, then grads cleared)
  * run_training        <- train.py:51-128  (the loop, the fp64 loss accumulator, the log line)

The arithmetic of the reference lives in two un-vendored third parties:
  torch      (unpinned, Dockerfile:11; this image has 2.11.0+cu128)
  deepspeed  (unpinned, Dockerfile:12; NOT installed here or on the GPU box)
The torch part is pinned: tests/test_oracle.py checks every function below against the
reference's own make_synthetic_data / SimpleNet (imported from /root/reference/train.py
with a stub `deepspeed` module by tests/golden/make_golden.py) and against torch's own
autograd / BCEWithLogitsLoss / optim.Adam.  The deepspeed part (sum-then-scale gradient
averaging, rank-0 parameter broadcast) is restated from its documented stage-0 semantics
and is PARITY UNPINNED: the reference holds no test or golden vector for it and the
package cannot be run here.

Deviations from the reference that the oracle makes on purpose:
  * world size 1 does not crash (train.py:101 calls set_epoch on None);
  * the device is a parameter (train.py:60-61 hard-code cuda).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch

ADAM_BETA1 = 0.9      # torch.optim.Adam defaults, train.py:85 passes only lr
ADAM_BETA2 = 0.999
ADAM_EPS = 1e-8


# --------------------------------------------------------------------------- data
def make_dataset(n_samples: int, n_features: int, seed: int):
    """train.py:19-24.  x ~ N(0,1) from the CPU generator, y = 1[sum of first half > 0]."""
    torch.manual_seed(seed)
    x = torch.randn(n_samples, n_features)
    half = x[:, : n_features // 2].sum(dim=1)
    y = (half > 0).float().unsqueeze(1)
    return x, y


def label_rule(x: torch.Tensor) -> torch.Tensor:
    """The label rule alone (train.py:22-23) for batches generated elsewhere."""
    f = x.shape[1]
    return (x[:, : f // 2].sum(dim=1) > 0).to(torch.float32).unsqueeze(1)


def init_params(n_features: int, hidden: int):
    """train.py:27-33.  Must be called right after make_dataset so that the global CPU
    generator is at the same stream position as in train.py:63,76."""
    l1 = torch.nn.Linear(n_features, hidden)
    l2 = torch.nn.Linear(hidden, 1)
    return {
        "w1": l1.weight.detach().clone(),   # [H, F]
        "b1": l1.bias.detach().clone(),     # [H]
        "w2": l2.weight.detach().clone(),   # [1, H]
        "b2": l2.bias.detach().clone(),     # [1]
    }


PARAM_ORDER = ("w1", "b1", "w2", "b2")      # == model.parameters() order of SimpleNet


def flatten(params) -> torch.Tensor:
    return torch.cat([params[k].reshape(-1) for k in PARAM_ORDER])


def unflatten(flat: torch.Tensor, n_features: int, hidden: int):
    f, h = n_features, hidden
    o = 0
    out = {}
    for k, shape in (("w1", (h, f)), ("b1", (h,)), ("w2", (1, h)), ("b2", (1,))):
        n = math.prod(shape)
        out[k] = flat[o:o + n].reshape(shape).clone()
        o += n
    return out


# --------------------------------------------------------------------------- sampler
def rank_indices(n: int, world: int, rank: int, epoch: int) -> list[int]:
    """Sample order for one epoch on one rank (train.py:65-74,101).

    world > 1: DistributedSampler(shuffle=True, seed=0, drop_last=False):
      randperm(n) from a private generator seeded `0 + epoch`, padded by wrap-around to
      world*ceil(n/world), then idx[rank::world]  (torch/utils/data/distributed.py:107-134).
    world == 1: DataLoader(shuffle=True) -> RandomSampler, whose per-epoch seed is drawn
      from the GLOBAL CPU generator (torch/utils/data/sampler.py:163); `epoch` is unused.
    """
    # DataLoader.__iter__ draws `_base_seed` from the global generator every epoch before
    # any index is produced (torch/utils/data/dataloader.py:706-710), for every world size.
    torch.empty((), dtype=torch.int64).random_()
    if world > 1:
        g = torch.Generator()
        g.manual_seed(0 + epoch)
        idx = torch.randperm(n, generator=g).tolist()
        per_rank = math.ceil(n / world)
        total = per_rank * world
        pad = total - len(idx)
        if pad > 0:
            if pad <= len(idx):
                idx += idx[:pad]
            else:
                idx += (idx * math.ceil(pad / len(idx)))[:pad]
        return idx[rank:total:world]
    seed = int(torch.empty((), dtype=torch.int64).random_().item())
    g = torch.Generator()
    g.manual_seed(seed)
    return torch.randperm(n, generator=g).tolist()


def batches(indices: list[int], batch_size: int):
    """DataLoader(batch_size=4, drop_last=False): consecutive chunks, ragged tail kept."""
    for i in range(0, len(indices), batch_size):
        yield indices[i:i + batch_size]


# --------------------------------------------------------------------------- step math
def forward(p, x: torch.Tensor):
    """train.py:35-36.  Returns (logits [B,1], hidden activation [B,H])."""
    z1 = torch.addmm(p["b1"], x, p["w1"].t())
    h = torch.clamp_min(z1, 0)
    z = torch.addmm(p["b2"], h, p["w2"].t())
    return z, h


def bce_with_logits(z: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """nn.BCEWithLogitsLoss(reduction='mean') written out: l_i = (1-y_i) z_i - logsigmoid(z_i)
    (ATen/native/Loss.cpp binary_cross_entropy_with_logits), then the mean."""
    li = (1 - y) * z - torch.nn.functional.logsigmoid(z)
    return li.mean()


def backward(p, x, y, z, h):
    """Autograd of forward+loss, written out (train.py:113 up to the all-reduce).
    dz = (sigmoid(z) - y) / B; no gradient w.r.t. x is formed (input does not require grad)."""
    b = z.shape[0]
    dz = (torch.sigmoid(z) - y) / b                    # [B,1]
    g_w2 = dz.t() @ h                                   # [1,H]
    g_b2 = dz.sum(dim=0)                                # [1]
    dh = dz @ p["w2"]                                   # [B,H]
    dz1 = dh * (h > 0).to(dh.dtype)                     # [B,H]
    g_w1 = dz1.t() @ x                                  # [H,F]
    g_b1 = dz1.sum(dim=0)                               # [H]
    return {"w1": g_w1, "b1": g_b1, "w2": g_w2, "b2": g_b2}


def average_gradients(per_rank_flat: list[torch.Tensor]) -> torch.Tensor:
    """DeepSpeed stage-0 gradient averaging over one flat fp32 bucket: all-reduce SUM in
    rank order, then multiply by 1/W.  [DS-UNVERIFIED order; see module docstring]"""
    w = len(per_rank_flat)
    acc = per_rank_flat[0].clone()
    for g in per_rank_flat[1:]:
        acc += g
    if w > 1:
        acc *= (1.0 / w)
    return acc


@dataclass
class AdamState:
    step: int = 0
    m: torch.Tensor | None = None
    v: torch.Tensor | None = None


def adam_update(flat_p: torch.Tensor, flat_g: torch.Tensor, st: AdamState, lr: float,
                beta1: float = ADAM_BETA1, beta2: float = ADAM_BETA2, eps: float = ADAM_EPS):
    """torch.optim.Adam single-tensor update (torch/optim/adam.py _single_tensor_adam),
    weight_decay=0, amsgrad=False, maximize=False; bias corrections in Python double."""
    if st.m is None:
        st.m = torch.zeros_like(flat_p)
        st.v = torch.zeros_like(flat_p)
    st.step += 1
    st.m.lerp_(flat_g, 1 - beta1)
    st.v.mul_(beta2).addcmul_(flat_g, flat_g, value=1 - beta2)
    bc1 = 1 - beta1 ** st.step
    bc2 = 1 - beta2 ** st.step
    step_size = lr / bc1
    bc2_sqrt = bc2 ** 0.5
    denom = (st.v.sqrt() / bc2_sqrt).add_(eps)
    flat_p.addcdiv_(st.m, denom, value=-step_size)
    return flat_p


# --------------------------------------------------------------------------- the loop
@dataclass
class RunResult:
    step_losses: list[list[float]] = field(default_factory=list)   # [rank][step]
    epoch_means: list[list[float]] = field(default_factory=list)   # [rank][epoch]
    log_lines: list[str] = field(default_factory=list)             # rank 0's stdout
    params: dict | None = None
    init_params: dict | None = None
    x: torch.Tensor | None = None
    y: torch.Tensor | None = None


def run_training(world: int = 1, epochs: int = 5, lr: float = 1e-3, seed: int = 42,
                 n_samples: int = 2000, n_features: int = 20, hidden: int = 64,
                 batch_size: int = 4, max_steps: int | None = None,
                 dtype: torch.dtype = torch.float32) -> RunResult:
    """train.py:51-128 with all `world` ranks emulated in this one process (identical
    arithmetic to W processes + all-reduce: every rank seeds the same way, train.py:54,63).

    `dtype=torch.bfloat16` is the restatement of the build's bf16 extension (BASELINE
    config 3; the reference has no bf16 path): GEMM operands rounded to bf16, fp32
    accumulation, fp32 master weights / Adam state / gradient bucket.
    """
    torch.manual_seed(seed)                                   # train.py:54
    x, y = make_dataset(n_samples, n_features, seed)          # train.py:63
    p = init_params(n_features, hidden)                       # train.py:76
    res = RunResult(x=x, y=y, init_params={k: v.clone() for k, v in p.items()})
    flat = flatten(p)
    st = AdamState()
    res.step_losses = [[] for _ in range(world)]
    res.epoch_means = [[] for _ in range(world)]
    done = 0
    for epoch in range(1, epochs + 1):                        # train.py:99 (1-based)
        if world > 1:
            # one process per rank in the reference: each rank's global generator is its own,
            # so emulate rank r's draw without letting it disturb rank 0's stream
            st0 = torch.get_rng_state()
            order = []
            for r in range(world):
                torch.set_rng_state(st0)
                order.append(rank_indices(n_samples, world, r, epoch))
        else:
            order = [rank_indices(n_samples, 1, 0, epoch)]
        per_rank_batches = [list(batches(o, batch_size)) for o in order]
        nsteps = len(per_rank_batches[0])
        sums = [0.0] * world                                  # Python doubles, train.py:102
        steps = 0
        for s in range(nsteps):
            if max_steps is not None and done >= max_steps:
                break
            p = unflatten(flat, n_features, hidden)
            grads = []
            for r in range(world):
                bi = torch.tensor(per_rank_batches[r][s])
                bx, by = x[bi], y[bi]
                if dtype == torch.bfloat16:
                    loss, g = _bf16_step(p, bx, by)
                else:
                    z, h = forward(p, bx)
                    loss = bce_with_logits(z, by)
                    g = backward(p, bx, by, z, h)
                grads.append(flatten(g))
                lv = float(loss)                              # loss.item(), train.py:115
                sums[r] += lv
                res.step_losses[r].append(lv)
            adam_update(flat, average_gradients(grads), st, lr)
            steps += 1
            done += 1
        for r in range(world):
            res.epoch_means[r].append(sums[r] / max(1, steps))   # train.py:119
        res.log_lines.append(f"Epoch {epoch:2d} finished. Avg loss: {res.epoch_means[0][-1]:.4f}")
    res.log_lines.append("Training completed.")
    res.params = unflatten(flat, n_features, hidden)
    return res


def _bf16_round(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32)


def _bf16_step(p, bx, by):
    """bf16-operand / fp32-accumulate restatement used by the bf16 configs (extension)."""
    xq = _bf16_round(bx)
    w1q = _bf16_round(p["w1"])
    z1 = xq @ w1q.t() + p["b1"]
    h = torch.clamp_min(z1, 0)
    z = h @ p["w2"].t() + p["b2"]
    loss = bce_with_logits(z, by)
    b = z.shape[0]
    dz = (torch.sigmoid(z) - by) / b
    g_w2 = dz.t() @ h
    g_b2 = dz.sum(dim=0)
    dz1 = (dz @ p["w2"]) * (h > 0).to(torch.float32)
    g_w1 = dz1.t() @ xq
    g_b1 = dz1.sum(dim=0)
    return loss, {"w1": g_w1, "b1": g_b1, "w2": g_w2, "b2": g_b2}


# --------------------------------------------------------------------------- torch-native cross-check
def torch_native_step(p, bx, by, opt_state: dict, lr: float, average_grads=None):
    """One step using the SAME torch objects train.py uses (nn.Linear/ReLU/Sequential,
    nn.BCEWithLogitsLoss, autograd, torch.optim.Adam) -- used by the tests to pin the
    written-out math above, and by bench.py --impl reference as the timed CPU path.
    `average_grads(list_of_param_grads)`, when given, runs between backward and the optimizer step:
    the stage-0 gradient averaging of engine.backward (train.py:113) for world > 1."""
    net = opt_state.get("net")
    if net is None:
        h, f = p["w1"].shape
        net = torch.nn.Sequential(torch.nn.Linear(f, h), torch.nn.ReLU(), torch.nn.Linear(h, 1))
        with torch.no_grad():
            net[0].weight.copy_(p["w1"]); net[0].bias.copy_(p["b1"])
            net[2].weight.copy_(p["w2"]); net[2].bias.copy_(p["b2"])
        opt_state["net"] = net
        opt_state["opt"] = torch.optim.Adam(net.parameters(), lr=lr)
        opt_state["crit"] = torch.nn.BCEWithLogitsLoss()
    opt = opt_state["opt"]
    logits = net(bx)
    loss = opt_state["crit"](logits, by)
    opt.zero_grad(set_to_none=True)
    loss.backward()
    if average_grads is not None:
        average_grads([q.grad for q in net.parameters()])
    opt.step()
    return loss.detach(), logits.detach()
