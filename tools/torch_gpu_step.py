#!/usr/bin/env python3
"""What stock PyTorch does with the benchmark's step on the same GPU (context for DESIGN 4.3, not a bench arm): the
widened model of train.py:26-36 (F = 20, H = 4096), 65536 samples drawn on the device each step with the label rule of
train.py:22-23, BCEWithLogitsLoss, backward, fused Adam -- eager and as one CUDA graph, fp32 and under bf16 autocast.
Self-contained (torch only; neither the library nor the oracle is imported).

    python tools/torch_gpu_step.py [--H 4096] [--F 20]"""
import argparse, json, time
import torch
import torch.nn as nn

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=65536); ap.add_argument("--F", type=int, default=20); ap.add_argument("--H", type=int, default=4096)
ap.add_argument("--steps", type=int, default=50)
a = ap.parse_args()
dev = torch.device("cuda", 0)
B, F, H = a.B, a.F, a.H
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def build(bf16):
    torch.manual_seed(42)
    net = nn.Sequential(nn.Linear(F, H), nn.ReLU(), nn.Linear(H, 1)).to(dev)
    opt = torch.optim.Adam(net.parameters(), lr=1e-3, fused=True, capturable=True)
    crit = nn.BCEWithLogitsLoss()
    loss_out = torch.zeros((), device=dev)

    def step():
        x = torch.randn(B, F, device=dev)
        y = (x[:, : F // 2].sum(dim=1, keepdim=True) > 0).float()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            z = net(x)
        loss = crit(z.float(), y)
        opt.zero_grad(set_to_none=False)
        loss.backward()
        opt.step()
        loss_out.copy_(loss.detach())
    return step, loss_out


def timed(fn, n):
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for i in range(n):
        flush.fill_(i & 0xFF)                       # the same L2 flush between steps as bench.py
        evs[i][0].record(); fn(); evs[i][1].record()
    torch.cuda.synchronize()
    return sum(e0.elapsed_time(e1) for e0, e1 in evs) / n * 1e3


res = {}
for bf16 in (False, True):
    step, loss_out = build(bf16)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(5):
            step()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    t0 = time.time()
    while time.time() - t0 < 1.0:                   # clocks up
        step()
    eager = timed(step, a.steps)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        step()
    for _ in range(200):
        g.replay()
    graph = timed(g.replay, a.steps)
    key = "bf16_autocast" if bf16 else "fp32"
    res[key] = dict(eager_us=round(eager, 1), graph_us=round(graph, 1), samples_per_s_graph=round(B / graph * 1e6), loss=float(loss_out))
print(json.dumps(dict(tool="torch_gpu_step", torch=torch.__version__, B=B, F=F, H=H, steps=a.steps, **res)))
