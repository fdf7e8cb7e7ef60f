"""Parity tests proper: every CUDA entry point is called through the C ABI (ctypes) on a
B200 and compared with the CPU oracle on the same seeded inputs, with the committed golden
vectors of the reference run, and -- at full sizes -- through size-independent properties.

Tolerances (north_star): generated batches and initial weights bit-exact; fp32 loss and
parameters within 1e-5 relative."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import philox as ph
from oracle import reference_step as ora

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RTOL = 1e-5


def _st():
    return torch.cuda.current_stream().cuda_stream


def _check(lib, rc):
    from b200ds import _lib
    _lib.check(rc)


def _flat_dev(p, dev):
    flat = ora.flatten(p)
    P = flat.numel()
    Ppad = (P + 3) // 4 * 4
    out = torch.zeros(Ppad, dtype=torch.float32, device=dev)
    out[:P] = flat.to(dev)
    return out, P, Ppad


def _rel(a, b):
    return float((a - b).abs().max() / max(1e-12, float(b.abs().max())))


# ------------------------------------------------------------------ Philox / data
@pytest.mark.parametrize("numel", [1, 6, 80, 255, 256, 257, 40000, 65536 * 20, 303104 * 4 + 5])
def test_philox_bit_exact_vs_torch_cuda(lib, cuda_dev, numel):
    seed = 42
    g = torch.Generator(device=cuda_dev)
    g.manual_seed(seed)
    off = 0
    for call in range(3):      # consecutive draws: the offset bookkeeping must track torch's generator
        want = torch.randn(numel, device=cuda_dev, generator=g)
        got = torch.empty(numel, dtype=torch.float32, device=cuda_dev)
        adv = C.c_uint64(0)
        _check(lib, lib.b200ds_philox_normal(got.data_ptr(), numel, seed, off, C.byref(adv), _st()))
        torch.cuda.synchronize()
        assert torch.equal(got.view(torch.int32), want.view(torch.int32)), (numel, call)
        off += adv.value
        assert off == g.get_offset() if hasattr(g, "get_offset") else True


def test_philox_matches_numpy_oracle_stream(lib, cuda_dev):
    sm = C.c_int(0)
    _check(lib, lib.b200ds_device_sm_count(0, C.byref(sm)))
    n = 5000
    got = torch.empty(n, dtype=torch.float32, device=cuda_dev)
    _check(lib, lib.b200ds_philox_normal(got.data_ptr(), n, 1234567, 8, None, _st()))
    want = ph.randn(n, 1234567, 8, sm.value)
    assert np.abs(got.cpu().numpy() - want).max() < 1e-5       # libm vs CUDA intrinsics: a few ulp


def test_philox_empty_and_bad_offset(lib, cuda_dev):
    assert lib.b200ds_philox_normal(None, 0, 1, 0, None, _st()) == 0
    assert lib.b200ds_philox_normal(None, 4, 1, 2, None, _st()) == 1


def test_label_rule_and_gather(lib, cuda_dev, golden):
    x = torch.from_numpy(golden["x"]).to(cuda_dev)
    y = torch.empty(2000, dtype=torch.float32, device=cuda_dev)
    _check(lib, lib.b200ds_label_rule(x.data_ptr(), y.data_ptr(), 2000, 20, _st()))
    assert np.array_equal(y.cpu().numpy(), golden["y"].reshape(-1))          # bit-exact labels (1031 positives)
    idx = torch.tensor([1999, 0, 7, 7, 1234], dtype=torch.int64, device=cuda_dev)
    bx = torch.empty(5, 20, dtype=torch.float32, device=cuda_dev)
    by = torch.empty(5, dtype=torch.float32, device=cuda_dev)
    _check(lib, lib.b200ds_gather_batch(x.data_ptr(), y.data_ptr(), idx.data_ptr(), 5, 20, bx.data_ptr(), by.data_ptr(), _st()))
    assert torch.equal(bx, x[idx]) and torch.equal(by, y[idx])


# ------------------------------------------------------------------ per-stage fp32 kernels
SHAPES = [(4, 20, 64), (2, 20, 64), (1, 20, 64), (3, 1, 1), (33, 7, 65), (250, 20, 64), (1000, 32, 256), (513, 64, 100)]


@pytest.mark.parametrize("B,F,H", SHAPES)
def test_forward_loss_backward_vs_oracle(lib, cuda_dev, B, F, H):
    torch.manual_seed(B * 7919 + F * 31 + H)
    p = ora.init_params(F, H)
    x = torch.randn(B, F)
    y = (torch.rand(B, 1) > 0.5).float()
    z_ref, h_ref = ora.forward(p, x)
    loss_ref = ora.bce_with_logits(z_ref, y)
    g_ref = ora.flatten(ora.backward(p, x, y, z_ref, h_ref))
    params, P, Ppad = _flat_dev(p, cuda_dev)
    xd, yd = x.to(cuda_dev), y.reshape(-1).to(cuda_dev)
    z = torch.empty(B, dtype=torch.float32, device=cuda_dev)
    dz = torch.empty(B, dtype=torch.float32, device=cuda_dev)
    loss = torch.zeros(1, dtype=torch.float32, device=cuda_dev)
    grads = torch.full((Ppad,), float("nan"), dtype=torch.float32, device=cuda_dev)
    sc = torch.zeros(max(256, lib.b200ds_train_step_scratch_bytes(B, F, H)), dtype=torch.uint8, device=cuda_dev)
    _check(lib, lib.b200ds_mlp_forward_f32(params.data_ptr(), xd.data_ptr(), B, F, H, z.data_ptr(), _st()))
    _check(lib, lib.b200ds_bce_loss_grad(z.data_ptr(), yd.data_ptr(), B, loss.data_ptr(), dz.data_ptr(), None,
                                         sc.data_ptr(), sc.numel(), _st()))
    _check(lib, lib.b200ds_mlp_backward_f32(params.data_ptr(), xd.data_ptr(), dz.data_ptr(), B, F, H, grads.data_ptr(),
                                            sc.data_ptr(), sc.numel(), _st()))
    torch.cuda.synchronize()
    assert _rel(z.cpu(), z_ref.reshape(-1)) <= RTOL
    assert abs(float(loss) - float(loss_ref)) <= RTOL * abs(float(loss_ref))
    dz_ref = (torch.sigmoid(z_ref) - y) / B
    assert _rel(dz.cpu(), dz_ref.reshape(-1)) <= RTOL
    assert _rel(grads[:P].cpu(), g_ref) <= RTOL


def test_loss_kernel_large_batch_multi_cta(lib, cuda_dev):
    # > 1 CTA: exercises the ticket + last-CTA reduction, twice (the ticket must reset itself)
    B = 65536 + 3
    torch.manual_seed(5)
    z = torch.randn(B) * 4
    y = (torch.rand(B) > 0.5).float()
    want = float(ora.bce_with_logits(z.reshape(-1, 1).double(), y.reshape(-1, 1).double()))
    zd, yd = z.to(cuda_dev), y.to(cuda_dev)
    dz = torch.empty(B, dtype=torch.float32, device=cuda_dev)
    loss = torch.zeros(1, dtype=torch.float32, device=cuda_dev)
    ctr = torch.zeros(1, dtype=torch.int64, device=cuda_dev)
    sc = torch.zeros(lib.b200ds_bce_scratch_bytes(B), dtype=torch.uint8, device=cuda_dev)
    for rep in range(2):
        loss.zero_()
        _check(lib, lib.b200ds_bce_loss_grad(zd.data_ptr(), yd.data_ptr(), B, loss.data_ptr(), dz.data_ptr(),
                                             ctr.data_ptr(), sc.data_ptr(), sc.numel(), _st()))
        torch.cuda.synchronize()
        assert abs(float(loss) - want) <= RTOL * abs(want), rep
    assert int(ctr) == 2
    dz_ref = (torch.sigmoid(z) - y) / B
    assert _rel(dz.cpu(), dz_ref) <= RTOL
    # property: sum of dz == (sum sigmoid - sum y)/B  (a checksum of the gradient)
    assert abs(float(dz.double().sum()) - float((torch.sigmoid(z.double()) - y.double()).sum() / B)) <= 1e-6


def test_loss_extreme_logits(lib, cuda_dev):
    z = torch.tensor([-100.0, -30.0, -1e-8, 0.0, 1e-8, 30.0, 100.0, 88.0])
    y = torch.tensor([0.0, 1.0, 0.0, 1.0, 1.0, 0.0, 1.0, 0.0])
    want = float(ora.bce_with_logits(z.reshape(-1, 1), y.reshape(-1, 1)))
    loss = torch.zeros(1, dtype=torch.float32, device=cuda_dev)
    dz = torch.empty(8, dtype=torch.float32, device=cuda_dev)
    zd, yd = z.to(cuda_dev), y.to(cuda_dev)
    _check(lib, lib.b200ds_bce_loss_grad(zd.data_ptr(), yd.data_ptr(), 8, loss.data_ptr(),
                                         dz.data_ptr(), None, None, 0, _st()))
    torch.cuda.synchronize()
    assert torch.isfinite(dz).all() and abs(float(loss) - want) <= RTOL * abs(want)


@pytest.mark.parametrize("n", [1, 3, 4, 1409, 1412, 90113 + 3, 1 << 20])
def test_adam_vs_oracle(lib, cuda_dev, n):
    torch.manual_seed(n)
    npad = (n + 3) // 4 * 4
    p0 = torch.randn(npad)
    mine = p0.clone().to(cuda_dev)
    m = torch.zeros(npad, device=cuda_dev)
    v = torch.zeros(npad, device=cuda_dev)
    ref = p0.clone()
    st = ora.AdamState()
    ctr = torch.zeros(1, dtype=torch.int64, device=cuda_dev)
    for t in range(1, 8):
        g = torch.randn(npad) * (10.0 ** (t % 4 - 2))
        gd = g.clone().to(cuda_dev)
        ora.adam_update(ref, g * 0.5, st, 3e-3)            # grad_scale = 1/world = 0.5
        if t % 2:
            _check(lib, lib.b200ds_adam_step(mine.data_ptr(), gd.data_ptr(), m.data_ptr(), v.data_ptr(), npad,
                                             3e-3, 0.9, 0.999, 1e-8, t, 0.5, 1, _st()))
            ctr += 1
        else:
            ctr += 1
            _check(lib, lib.b200ds_adam_step_counted(mine.data_ptr(), gd.data_ptr(), m.data_ptr(), v.data_ptr(), npad,
                                                     3e-3, 0.9, 0.999, 1e-8, ctr.data_ptr(), 0.5, 0, _st()))
        torch.cuda.synchronize()
        if t % 2:
            assert float(gd.abs().max()) == 0.0                 # zero_grad folded into the pass
    assert _rel(mine.cpu(), ref) <= RTOL
    assert _rel(m.cpu(), st.m) <= RTOL and _rel(v.cpu(), st.v) <= RTOL


# ------------------------------------------------------------------ fused tiny step vs golden reference run
def _tiny_state(golden, dev):
    p = {k: torch.from_numpy(golden[f"init_{k}"]) for k in ora.PARAM_ORDER}
    params, P, Ppad = _flat_dev(p, dev)
    z = lambda: torch.zeros(Ppad, dtype=torch.float32, device=dev)      # noqa: E731
    return params, z(), z(), z(), P, Ppad


def test_fused_epoch_matches_reference_world1(lib, cuda_dev, golden):
    """500 steps of the reference's own W=1 run (golden step losses) in ONE launch."""
    x = torch.from_numpy(golden["x"]).to(cuda_dev)
    y = torch.from_numpy(golden["y"]).reshape(-1).to(cuda_dev)
    params, grads, m, v, P, Ppad = _tiny_state(golden, cuda_dev)
    if str(golden["cpu_capability"][0]) != torch.backends.cpu.get_cpu_capability():
        pytest.skip("host CPU RNG path differs from the golden host: sample order not reproducible here")
    # the order the reference's DataLoader(shuffle=True) produced: redo its RNG draws
    torch.manual_seed(42)
    ora.make_dataset(2000, 20, 42)
    ora.init_params(20, 64)
    losses_all = []
    step = 0
    for epoch in range(1, 6):
        order = torch.tensor(ora.rank_indices(2000, 1, 0, epoch), dtype=torch.int64, device=cuda_dev)
        losses = torch.zeros(500, dtype=torch.float32, device=cuda_dev)
        _check(lib, lib.b200ds_tiny_train_steps_f32(params.data_ptr(), grads.data_ptr(), m.data_ptr(), v.data_ptr(), 20, 64,
                                                    x.data_ptr(), y.data_ptr(), order.data_ptr(), 2000, 0, 4, 500, 1,
                                                    1e-3, 0.9, 0.999, 1e-8, step, losses.data_ptr(), _st()))
        step += 500
        losses_all.append(losses.cpu().double().numpy())
    got = np.concatenate(losses_all)
    want = golden["w1_step_losses"][0]
    assert np.abs(got - want).max() <= RTOL * np.abs(want).max()
    means = [f"{got[e * 500:(e + 1) * 500].sum() / 500:.4f}" for e in range(5)]
    assert means == [f"{v:.4f}" for v in golden["w1_epoch_means"][0]]       # the logged line, to 4 decimals
    for k, (o, n) in zip(ora.PARAM_ORDER, ((0, 1280), (1280, 64), (1344, 64), (1408, 1))):
        ref = torch.from_numpy(golden[f"w1_final_{k}"]).reshape(-1)
        assert _rel(params[o:o + n].cpu(), ref) <= 5e-5, k     # 2500 chained Adam steps; per-step parity is the 1e-5 gate


@pytest.mark.parametrize("B,F,H,steps", [(4, 20, 64, 7), (2, 20, 64, 3), (32, 32, 256, 4), (5, 3, 33, 6), (1, 20, 64, 5)])
def test_fused_steps_vs_oracle_shapes(lib, cuda_dev, B, F, H, steps):
    torch.manual_seed(B + 10 * F + 100 * H)
    n = 200
    x = torch.randn(n, F)
    y = (torch.rand(n, 1) > 0.5).float()
    p = ora.init_params(F, H)
    order = torch.randperm(n)[: steps * B - (1 if B > 1 else 0)]     # ragged last batch when B > 1
    flat = ora.flatten(p)
    st = ora.AdamState()
    want = []
    for s in range(steps):
        bi = order[s * B:(s + 1) * B]
        pp = ora.unflatten(flat, F, H)
        z, h = ora.forward(pp, x[bi])
        want.append(float(ora.bce_with_logits(z, y[bi])))
        ora.adam_update(flat, ora.flatten(ora.backward(pp, x[bi], y[bi], z, h)), st, 1e-2)
    params, P, Ppad = _flat_dev(p, cuda_dev)
    zz = lambda: torch.zeros(Ppad, dtype=torch.float32, device=cuda_dev)   # noqa: E731
    grads, m, v = zz(), zz(), zz()
    losses = torch.zeros(steps, dtype=torch.float32, device=cuda_dev)
    od = order.to(torch.int64).to(cuda_dev)
    xd, yd = x.to(cuda_dev), y.reshape(-1).to(cuda_dev)
    _check(lib, lib.b200ds_tiny_train_steps_f32(params.data_ptr(), grads.data_ptr(), m.data_ptr(), v.data_ptr(), F, H,
                                                xd.data_ptr(), yd.data_ptr(),
                                                od.data_ptr(), od.numel(), 0, B, steps, 1, 1e-2, 0.9, 0.999, 1e-8, 0,
                                                losses.data_ptr(), _st()))
    torch.cuda.synchronize()
    assert np.abs(losses.cpu().double().numpy() - np.array(want)).max() <= RTOL * max(want)
    assert _rel(params[:P].cpu(), flat) <= RTOL


def test_fused_step_gradient_only_mode_matches_staged_kernels(lib, cuda_dev, golden):
    """apply_adam = 0 (the multi-rank form): gradients equal the per-stage kernels' bucket."""
    x = torch.from_numpy(golden["x"]).to(cuda_dev)
    y = torch.from_numpy(golden["y"]).reshape(-1).to(cuda_dev)
    params, grads, m, v, P, Ppad = _tiny_state(golden, cuda_dev)
    order = torch.tensor([3, 1999, 500, 77], dtype=torch.int64, device=cuda_dev)
    loss = torch.zeros(1, dtype=torch.float32, device=cuda_dev)
    _check(lib, lib.b200ds_tiny_train_steps_f32(params.data_ptr(), grads.data_ptr(), None, None, 20, 64, x.data_ptr(),
                                                y.data_ptr(), order.data_ptr(), 4, 0, 4, 1, 0, 1e-3, 0.9, 0.999, 1e-8, 0,
                                                loss.data_ptr(), _st()))
    bx, by = x[order].contiguous(), y[order].contiguous()
    z = torch.empty(4, device=cuda_dev); dz = torch.empty(4, device=cuda_dev); l2 = torch.zeros(1, device=cuda_dev)
    g2 = torch.zeros(Ppad, device=cuda_dev)
    sc = torch.zeros(max(256, lib.b200ds_train_step_scratch_bytes(4, 20, 64)), dtype=torch.uint8, device=cuda_dev)
    _check(lib, lib.b200ds_mlp_forward_f32(params.data_ptr(), bx.data_ptr(), 4, 20, 64, z.data_ptr(), _st()))
    _check(lib, lib.b200ds_bce_loss_grad(z.data_ptr(), by.data_ptr(), 4, l2.data_ptr(), dz.data_ptr(), None, sc.data_ptr(), sc.numel(), _st()))
    _check(lib, lib.b200ds_mlp_backward_f32(params.data_ptr(), bx.data_ptr(), dz.data_ptr(), 4, 20, 64, g2.data_ptr(), sc.data_ptr(), sc.numel(), _st()))
    torch.cuda.synchronize()
    assert abs(float(loss) - float(l2)) <= RTOL * abs(float(l2))
    assert _rel(grads[:P], g2[:P]) <= RTOL


def test_host_buffer_step_both_paths(lib, cuda_dev):
    """b200ds_train_step_host: the tiny class takes the fused kernel, larger shapes the staged
    kernels; both must follow the oracle over consecutive steps."""
    from b200ds import _lib
    for (B, F, H) in ((4, 20, 64), (300, 20, 512)):
        torch.manual_seed(B)
        p = ora.init_params(F, H)
        params, P, Ppad = _flat_dev(p, cuda_dev)
        z = lambda n=Ppad: torch.zeros(n, dtype=torch.float32, device=cuda_dev)    # noqa: E731
        grads, m, v = z(), z(), z()
        xs, ys, lg, dl, ls = z(B * F), z(B), z(B), z(B), z(1)
        sc = torch.zeros(max(256, lib.b200ds_train_step_scratch_bytes(B, F, H)), dtype=torch.uint8, device=cuda_dev)
        bufs = _lib.StepBufs(params.data_ptr(), grads.data_ptr(), m.data_ptr(), v.data_ptr(), xs.data_ptr(), ys.data_ptr(),
                             lg.data_ptr(), dl.data_ptr(), ls.data_ptr(), sc.data_ptr(), sc.numel(), None, None)
        ctx = C.c_void_p()
        _check(lib, lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
        flat = ora.flatten(p)
        st = ora.AdamState()
        xh = torch.empty(B, F).pin_memory(); yh = torch.empty(B).pin_memory(); lh = torch.zeros(1).pin_memory()
        for step in range(4):
            xh.copy_(torch.randn(B, F)); yh.copy_((torch.rand(B) > 0.5).float())
            pp = ora.unflatten(flat, F, H)
            zr, hr = ora.forward(pp, xh)
            want = float(ora.bce_with_logits(zr, yh.reshape(-1, 1)))
            ora.adam_update(flat, ora.flatten(ora.backward(pp, xh, yh.reshape(-1, 1), zr, hr)), st, 1e-3)
            _check(lib, lib.b200ds_train_step_host(ctx, C.byref(bufs), xh.data_ptr(), yh.data_ptr(), B, F, H,
                                                   1e-3, 0.9, 0.999, 1e-8, step, lh.data_ptr(), _st()))
            assert abs(float(lh) - want) <= RTOL * abs(want), (B, step)
        assert _rel(params[:P].cpu(), flat) <= RTOL
        _check(lib, lib.b200ds_ctx_destroy(ctx))


# ------------------------------------------------------------------ full-size properties (BASELINE config 4 sizes)
def test_full_size_fp32_properties(lib, cuda_dev):
    """B = 65536, H = 4096, F = 20: too big for the CPU oracle in seconds, so check
    (a) against the same written-out math evaluated by torch on the GPU in fp32/fp64,
    (b) linearity of backward in dlogits, (c) db2 == sum(dz) and db1 == row sums checksum."""
    B, F, H = 65536, 20, 4096
    torch.manual_seed(1)
    p = ora.init_params(F, H)
    params, P, Ppad = _flat_dev(p, cuda_dev)
    x = torch.empty(B * F, dtype=torch.float32, device=cuda_dev)
    _check(lib, lib.b200ds_philox_normal(x.data_ptr(), B * F, 42, 0, None, _st()))
    x = x.view(B, F)
    y = torch.empty(B, dtype=torch.float32, device=cuda_dev)
    _check(lib, lib.b200ds_label_rule(x.data_ptr(), y.data_ptr(), B, F, _st()))
    assert 0.45 < float(y.mean()) < 0.55
    pd = {k: v.to(cuda_dev) for k, v in p.items()}
    z_ref, h_ref = ora.forward({k: v.double() for k, v in pd.items()}, x.double())
    z = torch.empty(B, dtype=torch.float32, device=cuda_dev)
    _check(lib, lib.b200ds_mlp_forward_f32(params.data_ptr(), x.data_ptr(), B, F, H, z.data_ptr(), _st()))
    assert _rel(z.double(), z_ref.reshape(-1)) <= RTOL
    dz = torch.empty(B, dtype=torch.float32, device=cuda_dev)
    loss = torch.zeros(1, dtype=torch.float32, device=cuda_dev)
    sc = torch.zeros(max(256, lib.b200ds_train_step_scratch_bytes(B, F, H)), dtype=torch.uint8, device=cuda_dev)
    _check(lib, lib.b200ds_bce_loss_grad(z.data_ptr(), y.data_ptr(), B, loss.data_ptr(), dz.data_ptr(), None, sc.data_ptr(), sc.numel(), _st()))
    g = torch.zeros(Ppad, dtype=torch.float32, device=cuda_dev)
    off = lib.b200ds_bce_scratch_bytes(B)
    off = (off + 255) // 256 * 256
    _check(lib, lib.b200ds_mlp_backward_f32(params.data_ptr(), x.data_ptr(), dz.data_ptr(), B, F, H, g.data_ptr(),
                                            sc.data_ptr() + off, sc.numel() - off, _st()))
    g_ref = ora.backward({k: v.double() for k, v in pd.items()}, x.double(), y.double().reshape(-1, 1), z_ref, h_ref)
    g_ref = ora.flatten(g_ref)
    assert _rel(g[:P].double(), g_ref) <= 5e-5          # 65536-term fp32 sums vs fp64
    # linearity: backward(2*dz) == 2*backward(dz) exactly (power of two)
    g2 = torch.zeros(Ppad, dtype=torch.float32, device=cuda_dev)
    dz2 = dz * 2
    _check(lib, lib.b200ds_mlp_backward_f32(params.data_ptr(), x.data_ptr(), dz2.data_ptr(), B, F, H, g2.data_ptr(),
                                            sc.data_ptr() + off, sc.numel() - off, _st()))
    assert torch.equal(g2[:P], 2 * g[:P])
    assert abs(float(g[P - 1]) - float(dz.double().sum())) <= 1e-6


# ------------------------------------------------------------------ the engine API and train.py itself
def test_engine_api_step_matches_oracle(cuda_dev):
    import b200ds
    torch.manual_seed(42)
    x, y = ora.make_dataset(2000, 20, 42)
    net = torch.nn.Sequential(torch.nn.Linear(20, 64), torch.nn.ReLU(), torch.nn.Linear(64, 1))
    p = {"w1": net[0].weight.detach().clone(), "b1": net[0].bias.detach().clone(),
         "w2": net[2].weight.detach().clone(), "b2": net[2].bias.detach().clone()}
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)
    cfg = {"train_batch_size": 64, "gradient_accumulation_steps": 1, "fp16": {"enabled": False}, "zero_optimization": {"stage": 0}}
    engine, opt2, a, b = b200ds.initialize(args=cfg, model=net, optimizer=opt, model_parameters=net.parameters(), config=cfg)
    assert opt2 is opt and a is None and b is None and engine.device == cuda_dev
    assert torch.equal(net[0].weight.detach().cpu(), p["w1"])                 # initial weights bit-exact after the move
    crit = b200ds.BCEWithLogitsLoss()
    torch_crit = torch.nn.BCEWithLogitsLoss()
    flat = ora.flatten(p)
    st = ora.AdamState()
    engine.train()
    for s in range(6):
        bi = torch.arange(s * 4, s * 4 + (4 if s != 3 else 2))
        bx, by = x[bi].to(cuda_dev), y[bi].to(cuda_dev)
        logits = engine(bx)
        assert logits.shape == (len(bi), 1)
        loss = crit(logits, by) if s % 2 == 0 else torch_crit(logits, by)    # ours, and the reference's torch criterion
        engine.backward(loss)
        engine.step()
        pp = ora.unflatten(flat, 20, 64)
        z, h = ora.forward(pp, x[bi])
        want = float(ora.bce_with_logits(z, y[bi]))
        ora.adam_update(flat, ora.flatten(ora.backward(pp, x[bi], y[bi], z, h)), st, 1e-3)
        assert abs(loss.item() - want) <= RTOL * abs(want), s
    assert _rel(engine.flat[: engine.P].cpu(), flat) <= RTOL
    assert _rel(net[2].weight.detach().cpu().reshape(-1), flat[1344:1408]) <= RTOL   # module params are views of the flat buffer


@pytest.mark.parametrize("extra", [[], ["--fused-steps"]])
def test_train_py_stdout_matches_reference(golden, tmp_path, extra):
    if str(golden["cpu_capability"][0]) != torch.backends.cpu.get_cpu_capability():
        pytest.skip("host CPU RNG path differs from the golden host")
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    r = subprocess.run([sys.executable, os.path.join(ROOT, "train.py"), "--save-dir", str(tmp_path / "ck"),
                        "--distributed-backend", "nccl", "--deepspeed"] + extra,
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("Epoch") or ln.startswith("Training")]
    want = [f"Epoch {e + 1:2d} finished. Avg loss: {v:.4f}" for e, v in enumerate(golden["w1_epoch_means"][0])]
    assert lines == want + ["Training completed."]
    assert os.path.exists(tmp_path / "ck" / "epoch5" / "latest")


def _engine_on_reference_model(cfg_extra=None, seed=42):
    import b200ds
    torch.manual_seed(seed)
    ora.make_dataset(2000, 20, seed)          # the reference draws the dataset before it builds the model (train.py:63,76)
    net = torch.nn.Sequential(torch.nn.Linear(20, 64), torch.nn.ReLU(), torch.nn.Linear(64, 1))
    p = {"w1": net[0].weight.detach().clone(), "b1": net[0].bias.detach().clone(),
         "w2": net[2].weight.detach().clone(), "b2": net[2].bias.detach().clone()}
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)
    cfg = {"train_batch_size": 64, "gradient_accumulation_steps": 1, "fp16": {"enabled": False}, "zero_optimization": {"stage": 0}}
    cfg.update(cfg_extra or {})
    engine, _, _, _ = b200ds.initialize(model=net, optimizer=opt, config=cfg)
    return engine, net, opt, p


def test_checkpoint_holds_the_trained_tensors_and_loads_back(cuda_dev, tmp_path):
    """engine.save_checkpoint (train.py:123-125) after N steps: the .pt file's module.* tensors are the engine's
    flat buffer and the oracle's parameters after the same N steps; load_checkpoint into a fresh engine, then
    three more steps on both: parameters, Adam moments and losses stay bit-equal."""
    import b200ds
    x, y = ora.make_dataset(2000, 20, 42)
    engine, net, opt, p = _engine_on_reference_model()
    crit = b200ds.BCEWithLogitsLoss()
    flat = ora.flatten(p)
    st = ora.AdamState()

    def steps(eng, lo, hi, track=False):
        out = []
        for s in range(lo, hi):
            bi = torch.arange(s * 4, s * 4 + 4)
            loss = crit(eng(x[bi].to(cuda_dev)), y[bi].to(cuda_dev))
            eng.backward(loss)
            eng.step()
            out.append(loss.item())
            if track:
                pp = ora.unflatten(flat, 20, 64)
                z, h = ora.forward(pp, x[bi])
                ora.adam_update(flat, ora.flatten(ora.backward(pp, x[bi], y[bi], z, h)), st, 1e-3)
        return out

    steps(engine, 0, 7, track=True)
    assert engine.save_checkpoint(str(tmp_path / "ck"), client_state={"epoch": 3})
    tag = open(tmp_path / "ck" / "latest").read().strip()
    assert tag == "global_step7"
    state = torch.load(tmp_path / "ck" / tag / "mp_rank_00_model_states.pt", map_location="cpu", weights_only=False)
    names = ["0.weight", "0.bias", "2.weight", "2.bias"]
    assert sorted(state["module"]) == sorted(names)
    saved = torch.cat([state["module"][n].reshape(-1) for n in names])
    assert torch.equal(saved, engine.flat[: engine.P].cpu())                       # the file holds what the device holds
    assert _rel(saved, flat) <= RTOL                                               # ... which is the oracle's state after 7 steps
    assert state["global_steps"] == 7 and state["optimizer"]["step"] == 7 and state["client_state"] == {"epoch": 3}
    assert float(opt.state_dict()["state"][0]["step"]) == 7.0                      # optimizer.state_dict() reports Adam's t
    m_saved = torch.cat([state["optimizer"]["exp_avg"][n].reshape(-1) for n in names])
    assert torch.equal(m_saved, engine.exp_avg[: engine.P].cpu())
    # a fresh engine (different init) restored from the file continues bit for bit
    engine2, net2, _, _ = _engine_on_reference_model(seed=7)
    assert not torch.equal(engine2.flat, engine.flat)
    path, client = engine2.load_checkpoint(str(tmp_path / "ck"))
    assert path.endswith(tag) and client == {"epoch": 3} and engine2.global_steps == 7
    assert torch.equal(engine2.flat, engine.flat) and torch.equal(engine2.exp_avg, engine.exp_avg)
    assert torch.equal(net2[0].weight.detach(), net[0].weight.detach())            # module views follow
    la, lb = steps(engine, 7, 10), steps(engine2, 7, 10)
    assert la == lb
    assert torch.equal(engine2.flat, engine.flat) and torch.equal(engine2.exp_avg_sq, engine.exp_avg_sq)
    assert engine2.load_checkpoint(str(tmp_path / "nothing_here")) == (None, None)


def test_engine_bf16_on_the_reference_model_runs_the_tensor_core_kernels(cuda_dev, golden):
    """BASELINE config 3 semantics on the reference's own shapes (H = 64, batches of 4, ragged tail): bf16.enabled
    pads the hidden layer and the batch up to the tensor-core shape class instead of silently running fp32.  The
    first 200 reference steps (golden W = 1 losses, fp32) are tracked within north_star's bf16 tolerance 2e-2 on
    the running mean, the padded hidden units stay exactly zero, and a ragged batch of 2 gives 2 logits."""
    import b200ds
    x, y = torch.from_numpy(golden["x"]), torch.from_numpy(golden["y"])
    engine, net, opt, p = _engine_on_reference_model({"bf16": {"enabled": True}})
    assert engine.H == 128 and engine.H_model == 64 and engine.P == 20 * 128 + 2 * 128 + 1
    assert torch.equal(net[0].weight.detach().cpu(), torch.from_numpy(golden["init_w1"]))
    crit = b200ds.BCEWithLogitsLoss()
    g = torch.Generator(); g.manual_seed(1)            # any fixed order: compare against the fp32 oracle on the same batches
    order = torch.randperm(2000, generator=g)
    flat = ora.flatten(p)
    st = ora.AdamState()
    run16 = run32 = 0.0
    for s in range(200):
        bi = order[s * 4: s * 4 + (2 if s == 17 else 4)]
        logits = engine(x[bi].to(cuda_dev))
        assert engine._tc and logits.shape == (len(bi), 1)
        loss = crit(logits, y[bi].to(cuda_dev))
        engine.backward(loss)
        engine.step()
        pp = ora.unflatten(flat, 20, 64)
        z, h = ora.forward(pp, x[bi])
        want = float(ora.bce_with_logits(z, y[bi]))
        ora.adam_update(flat, ora.flatten(ora.backward(pp, x[bi], y[bi], z, h)), st, 1e-3)
        run16 += loss.item(); run32 += want
        assert abs(loss.item() - want) <= 2e-2 * abs(want) + 2e-3, (s, loss.item(), want)
    assert abs(run16 - run32) <= 2e-2 * run32
    F, H, Hm = 20, engine.H, 64
    fl = engine.flat.cpu()
    assert float(fl[Hm * F: H * F].abs().max()) == 0.0 and float(fl[H * F + Hm: H * F + H].abs().max()) == 0.0
    assert float(fl[H * F + H + Hm: H * F + 2 * H].abs().max()) == 0.0            # padded W1 rows, b1, W2: still zero
    got = torch.cat([t.detach().reshape(-1).cpu() for t in (net[0].weight, net[0].bias, net[2].weight, net[2].bias)])
    move = float((flat - ora.flatten(p)).norm())
    assert float((got - flat).norm()) <= 2e-2 * float(flat.norm()), (float((got - flat).norm()), move)   # the parameters, within 2e-2
    assert float((got - flat).abs().max()) <= 0.1 * 200 * 1e-3                    # and no element off by more than a tenth of Adam's reach
    with pytest.raises(NotImplementedError):           # F + 2 > 32: no tensor-core path, and no silent fp32 substitute
        wide = torch.nn.Sequential(torch.nn.Linear(40, 64), torch.nn.ReLU(), torch.nn.Linear(64, 1))
        b200ds.initialize(model=wide, optimizer=torch.optim.Adam(wide.parameters()), config={"bf16": {"enabled": True}})


def test_counted_philox_replays_consecutive_draws(lib, cuda_dev):
    """b200ds_philox_normal_counted with draw counter k == the (k+1)-th consecutive
    torch.randn(n) of a generator seeded the same way (what a CUDA-graph replay relies on)."""
    n, seed = 65536 * 20, 7
    g = torch.Generator(device=cuda_dev)
    g.manual_seed(seed)
    ctr = torch.zeros(1, dtype=torch.int64, device=cuda_dev)
    out = torch.empty(n, dtype=torch.float32, device=cuda_dev)
    for k in range(4):
        want = torch.randn(n, device=cuda_dev, generator=g)
        ctr.fill_(k)
        _check(lib, lib.b200ds_philox_normal_counted(out.data_ptr(), n, seed, 0, ctr.data_ptr(), _st()))
        torch.cuda.synchronize()
        assert torch.equal(out.view(torch.int32), want.view(torch.int32)), k
    assert lib.b200ds_philox_normal_counted(out.data_ptr(), n, seed, 0, None, _st()) == 1


@pytest.mark.parametrize("n", [4, 1412, 90116, 400000])
def test_peer_memory_allreduce_adam_with_one_rank_is_adam(lib, cuda_dev, n):
    """b200ds_allreduce_adam_p2p at world == 1: the whole publish / wait / read-back protocol runs
    against this rank's own buffer and must leave exactly what b200ds_adam_step leaves, over
    several epochs (both halves of the exchange buffer), with the step as an argument and from
    the device counter; the gradient bucket keeps its (one-rank) sum."""
    ctx = C.c_void_p()
    _check(lib, lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
    args = (1e-3, 0.9, 0.999, 1e-8)
    torch.manual_seed(n)
    p1 = torch.randn(n, device=cuda_dev); p2 = p1.clone()
    m1 = torch.zeros(n, device=cuda_dev); v1 = torch.zeros(n, device=cuda_dev)
    m2 = torch.zeros(n, device=cuda_dev); v2 = torch.zeros(n, device=cuda_dev)
    g = torch.empty(n, device=cuda_dev)
    assert lib.b200ds_allreduce_adam_p2p(ctx, p1.data_ptr(), g.data_ptr(), m1.data_ptr(), v1.data_ptr(), n, *args, 1, None, _st()) == 1
    assert b"enable_p2p" in lib.b200ds_last_error()
    _check(lib, lib.b200ds_ctx_enable_p2p(ctx, n))
    assert lib.b200ds_ctx_enable_p2p(ctx, n) == 1                       # once per context
    ctr = torch.zeros(1, dtype=torch.int64, device=cuda_dev)
    for step in range(1, 6):
        g.copy_(torch.randn(n, device=cuda_dev)); g0 = g.clone()
        if step % 2:
            _check(lib, lib.b200ds_allreduce_adam_p2p(ctx, p1.data_ptr(), g.data_ptr(), m1.data_ptr(), v1.data_ptr(), n, *args,
                                                     step, None, _st()))
        else:
            ctr.fill_(step)
            _check(lib, lib.b200ds_allreduce_adam_p2p(ctx, p1.data_ptr(), g.data_ptr(), m1.data_ptr(), v1.data_ptr(), n, *args,
                                                     0, ctr.data_ptr(), _st()))
        _check(lib, lib.b200ds_adam_step(p2.data_ptr(), g0.data_ptr(), m2.data_ptr(), v2.data_ptr(), n, *args, step, 1.0, 0, _st()))
        torch.cuda.synchronize()
        assert torch.equal(g, g0), step
        assert torch.equal(p1, p2) and torch.equal(m1, m2) and torch.equal(v1, v2), step
    assert lib.b200ds_allreduce_adam_p2p(ctx, p1.data_ptr(), g.data_ptr(), m1.data_ptr(), v1.data_ptr(), n + 4, *args, 1, None, _st()) == 1
    _check(lib, lib.b200ds_ctx_destroy(ctx))


def test_single_rank_context_exchange_entry_points_are_no_ops(lib, cuda_dev):
    """world == 1: the all-reduce (whole and in two halves), the broadcast and the barrier return
    at once and leave the bucket alone; the ZeRO-1-style step degenerates to plain Adam."""
    ctx = C.c_void_p()
    _check(lib, lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
    dev, rank, world = C.c_int(-1), C.c_int(-1), C.c_int(-1)
    _check(lib, lib.b200ds_ctx_info(ctx, C.byref(dev), C.byref(rank), C.byref(world)))
    assert (dev.value, rank.value, world.value) == (0, 0, 1)
    n = 1412
    torch.manual_seed(0)
    g = torch.randn(n, device=cuda_dev)
    g0 = g.clone()
    _check(lib, lib.b200ds_allreduce_bucket(ctx, g.data_ptr(), n, _st()))
    _check(lib, lib.b200ds_allreduce_bucket_begin(ctx, g.data_ptr(), n, _st()))
    _check(lib, lib.b200ds_allreduce_bucket_end(ctx, _st()))
    _check(lib, lib.b200ds_broadcast_params(ctx, g.data_ptr(), n, _st()))
    _check(lib, lib.b200ds_barrier(ctx))
    assert torch.equal(g, g0)
    p1 = torch.randn(n, device=cuda_dev); p2 = p1.clone()
    m1 = torch.zeros(n, device=cuda_dev); v1 = torch.zeros(n, device=cuda_dev)
    m2 = torch.zeros(n, device=cuda_dev); v2 = torch.zeros(n, device=cuda_dev)
    _check(lib, lib.b200ds_zero1_step(ctx, p1.data_ptr(), g.data_ptr(), m1.data_ptr(), v1.data_ptr(), n, 1e-3, 0.9, 0.999, 1e-8, 1, _st()))
    _check(lib, lib.b200ds_adam_step(p2.data_ptr(), g.data_ptr(), m2.data_ptr(), v2.data_ptr(), n, 1e-3, 0.9, 0.999, 1e-8, 1, 1.0, 0, _st()))
    torch.cuda.synchronize()
    assert torch.equal(p1, p2) and torch.equal(m1, m2) and torch.equal(v1, v2)
    assert lib.b200ds_zero1_step(ctx, p1.data_ptr(), g.data_ptr(), m1.data_ptr(), v1.data_ptr(), 1409, 1e-3, 0.9, 0.999, 1e-8, 1, _st()) == 1
    assert lib.b200ds_ctx_create(0, 1, 1, None, C.byref(C.c_void_p())) == 1        # rank outside world
    assert lib.b200ds_ctx_create(0, 0, 2, None, C.byref(C.c_void_p())) == 1        # world 2 without an id
    _check(lib, lib.b200ds_ctx_destroy(ctx))


def test_engine_rejects_what_is_outside_the_path(cuda_dev):
    import b200ds
    net = torch.nn.Sequential(torch.nn.Linear(20, 64), torch.nn.ReLU(), torch.nn.Linear(64, 1))
    with pytest.raises(NotImplementedError):
        b200ds.initialize(model=net, optimizer=torch.optim.SGD(net.parameters(), lr=0.1), config={})
    with pytest.raises(NotImplementedError):
        b200ds.initialize(model=net, optimizer=torch.optim.Adam(net.parameters(), weight_decay=0.1), config={})
    with pytest.raises(NotImplementedError):
        b200ds.initialize(model=torch.nn.Linear(20, 1), optimizer=torch.optim.Adam(net.parameters()), config={})
    engine, _, _, _ = b200ds.initialize(model=net, optimizer=torch.optim.Adam(net.parameters()), config={})
    with pytest.raises(ValueError):
        engine(torch.zeros(4, 19, device=cuda_dev))
    with pytest.raises(ValueError):
        engine(torch.zeros(4, 20))                       # host tensor: there is no CPU path
    with pytest.raises(RuntimeError):
        engine.backward(torch.zeros((), device=cuda_dev))


def test_sweep_tool_and_collective_entry_point(lib, cuda_dev):
    """tools/allreduce_sweep.py runs end to end on one rank (every collective is a no-op there) and prints its table;
    b200ds_collective checks its arguments."""
    import json
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "allreduce_sweep.py"), "--quick"], capture_output=True, text=True,
                       env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    doc = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    ops = {(row["op"], row["dtype"]) for row in doc["rows"]}
    assert {("allreduce", "f32"), ("allreduce", "bf16"), ("reduce_scatter", "f32"), ("all_gather", "bf16")} <= ops
    ctx = C.c_void_p()
    _check(lib, lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
    buf = torch.ones(64, device=cuda_dev)
    assert lib.b200ds_collective(ctx, 0, 0, buf.data_ptr(), 64, _st()) == 0          # one rank: nothing to do
    assert lib.b200ds_collective(ctx, 3, 0, buf.data_ptr(), 64, _st()) == 1          # no such op
    assert lib.b200ds_collective(ctx, 0, 2, buf.data_ptr(), 64, _st()) == 1          # no such dtype
    _check(lib, lib.b200ds_ctx_destroy(ctx))


def test_engine_host_buffer_step_equals_the_engine_calls(cuda_dev):
    """engine.make_host_step (b200ds_train_step_host under the engine's own buffers: one call per step over pinned host
    tensors) against engine(x) / backward / step on the same batches: same losses and parameters (the reference model,
    fp32: both run the fused one-CTA kernel); and a bf16 engine refuses a batch its tensor-core kernels cannot cut
    instead of stepping in fp32."""
    import b200ds
    x, y = ora.make_dataset(2000, 20, 42)
    ea, _, _, _ = _engine_on_reference_model()
    eb, _, _, _ = _engine_on_reference_model()
    crit = b200ds.BCEWithLogitsLoss()
    host_step = eb.make_host_step(4)
    xh = torch.empty(4, 20).pin_memory(); yh = torch.empty(4).pin_memory(); lh = torch.zeros(1).pin_memory()
    for s in range(6):
        bi = torch.arange(s * 4, s * 4 + 4)
        loss = crit(ea(x[bi].to(cuda_dev)), y[bi].to(cuda_dev))
        ea.backward(loss)
        ea.step()
        xh.copy_(x[bi]); yh.copy_(y[bi].reshape(-1))
        host_step(xh, yh, lh)
        assert abs(float(lh) - loss.item()) <= 1e-6 * max(1.0, abs(loss.item())), s
    assert eb.global_steps == ea.global_steps == 6
    assert _rel(eb.flat[: eb.P].cpu(), ea.flat[: ea.P].cpu()) <= 1e-6
    ec, _, _, _ = _engine_on_reference_model({"bf16": {"enabled": True}})
    with pytest.raises(NotImplementedError):
        ec.make_host_step(4)
    assert callable(ec.make_host_step(128))
