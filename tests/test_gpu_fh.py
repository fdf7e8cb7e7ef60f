"""Parity of the wide-input path (F in the hundreds / thousands: K-loop tcgen05 GEMMs fed by tensor-map TMA,
csrc/mlp_fh.cu) through the C ABI, against an fp64 evaluation of the same bf16-operand formulas on the GPU.
SimpleNet(n_features, hidden) takes any n_features (train.py:26-36); train.py:107,113 are the calls replaced."""
import ctypes as C

import pytest
import torch

from oracle import reference_step as ora

pytestmark = pytest.mark.gpu


def _st():
    return torch.cuda.current_stream().cuda_stream


def _chk(rc):
    from b200ds import _lib
    _lib.check(rc)


def _bf(t):
    return t.to(torch.bfloat16).to(torch.float64)


def unpack_L(buf, R, K):
    """layout L bytes -> [R, K] bf16 tensor (8 x 8 core matrices, k-fastest inside a group of 8 rows)"""
    return buf.view(torch.bfloat16).view(R // 8, K // 8, 8, 8).permute(0, 2, 1, 3).reshape(R, K)


def _setup(lib, dev, B, F, H, seed):
    torch.manual_seed(seed)
    p = {k: v.to(dev) for k, v in ora.init_params(F, H).items()}
    p["b1"] = p["b1"] + 0.05 * torch.randn(H, device=dev)
    x = torch.randn(B, F, device=dev)
    flat = torch.cat([p[k].reshape(-1) for k in ora.PARAM_ORDER])
    P = flat.numel(); Ppad = (P + 3) // 4 * 4
    params = torch.zeros(Ppad, device=dev); params[:P] = flat
    xq = torch.empty(lib.b200ds_fh_operand_bytes(B, F), dtype=torch.uint8, device=dev)
    wq = torch.empty(lib.b200ds_fh_operand_bytes(H, F), dtype=torch.uint8, device=dev)
    _chk(lib.b200ds_fh_pack_bf16(x.data_ptr(), B, F, xq.data_ptr(), _st()))
    _chk(lib.b200ds_fh_pack_bf16(params.data_ptr(), H, F, wq.data_ptr(), _st()))
    return p, x, params, P, xq, wq


def test_fh_pack_layout(lib, cuda_dev):
    R, K = 64, 256
    src = torch.randn(R, K, device=cuda_dev)
    dst = torch.empty(lib.b200ds_fh_operand_bytes(R, K), dtype=torch.uint8, device=cuda_dev)
    _chk(lib.b200ds_fh_pack_bf16(src.data_ptr(), R, K, dst.data_ptr(), _st()))
    torch.cuda.synchronize()
    assert torch.equal(unpack_L(dst, R, K), src.to(torch.bfloat16))
    assert lib.b200ds_fh_pack_bf16(src.data_ptr(), 63, K, dst.data_ptr(), _st()) == 1


@pytest.mark.parametrize("B,F,H", [(256, 256, 128), (512, 256, 256), (1024, 1024, 1024), (8192, 512, 384), (4096, 4096, 4096), (2048, 8192, 8192)])
def test_fh_forward_vs_fp64(lib, cuda_dev, B, F, H):
    assert lib.b200ds_fh_supported(B, F, H) == 1
    p, x, params, P, xq, wq = _setup(lib, cuda_dev, B, F, H, B + F + H)
    hT = torch.full((lib.b200ds_fh_operand_bytes(H, B),), 0xFF, dtype=torch.uint8, device=cuda_dev)
    sc = torch.full((lib.b200ds_fh_forward_scratch_bytes(B, F, H) // 4,), float("nan"), device=cuda_dev)
    for rep in range(2):
        _chk(lib.b200ds_fh_forward(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, hT.data_ptr(), sc.data_ptr(),
                                   sc.numel() * 4, _st()))
        torch.cuda.synchronize()
        z1 = _bf(x) @ _bf(p["w1"]).t() + p["b1"].double()                       # [B, H]
        z1_abs = _bf(x).abs() @ _bf(p["w1"]).abs().t() + p["b1"].double().abs()
        h_ref = torch.clamp_min(z1, 0)
        h = unpack_L(hT, H, B).double().t()                                     # [B, H]
        # h is the bf16 rounding of an fp32 sum of F products: 2^-8 of the value (bf16 half-ulp, with room for a
        # rounding-boundary flip) plus the fp32 accumulation error, measured against the sum of |terms|
        bad = (h - h_ref).abs() > 2.0 ** -8 * h_ref.abs() + 1e-5 * z1_abs
        assert not bool(bad.any()), (rep, int(bad.sum()), float((h - h_ref).abs().max()))
        # the logit parts: exact function of the h the kernel stored (fp32 sums of 128 products)
        zp = sc.view(H // 128, B).double()
        w2 = p["w2"].double().reshape(-1)
        want = (h.view(B, H // 128, 128) * w2.view(1, H // 128, 128)).sum(dim=2).t()          # [H/128, B]
        scale = (h.view(B, H // 128, 128) * w2.abs().view(1, H // 128, 128)).sum(dim=2).t()
        bad = (zp - want).abs() > 1e-5 * scale + 1e-30
        assert not bool(bad.any()), (rep, int(bad.sum()), float((zp - want).abs().max()))


@pytest.mark.parametrize("B,F,H", [(2048, 256, 128), (4096, 512, 384), (2048, 1024, 1024), (8192, 4096, 2048), (2048, 8192, 8192)])
def test_fh_backward_vs_fp64(lib, cuda_dev, B, F, H):
    p, x, params, P, xq, wq = _setup(lib, cuda_dev, B, F, H, 3 * B + F + H)
    hT = torch.empty(lib.b200ds_fh_operand_bytes(H, B), dtype=torch.uint8, device=cuda_dev)
    fsc = torch.zeros(lib.b200ds_fh_forward_scratch_bytes(B, F, H), dtype=torch.uint8, device=cuda_dev)
    _chk(lib.b200ds_fh_forward(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, hT.data_ptr(), fsc.data_ptr(), fsc.numel(), _st()))
    logits = torch.empty(B, device=cuda_dev)
    _chk(lib.b200ds_fh_logits(fsc.data_ptr(), params.data_ptr(), B, F, H, logits.data_ptr(), _st()))
    torch.cuda.synchronize()
    h = unpack_L(hT, H, B).double().t().contiguous()                           # [B, H], what the forward kept
    w2 = p["w2"].double().reshape(-1)
    z_ref = h @ w2 + p["b2"].double()
    assert float((logits.double() - z_ref).abs().max()) <= 1e-5 * float((h @ w2.abs()).max() + 1)
    y = (x[:, : F // 2].sum(dim=1) > 0).float()
    dz = ((torch.sigmoid(logits) - y) / B).contiguous()
    Ppad = params.numel()
    bsc = torch.zeros(lib.b200ds_fh_backward_scratch_bytes(B, F, H), dtype=torch.uint8, device=cuda_dev)
    for rep in range(2):
        if rep:                                                                  # the backward consumed hT: run the forward again
            _chk(lib.b200ds_fh_forward(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, hT.data_ptr(), fsc.data_ptr(), fsc.numel(), _st()))
        grads = torch.full((Ppad,), float("nan"), device=cuda_dev)
        _chk(lib.b200ds_fh_backward(xq.data_ptr(), hT.data_ptr(), dz.data_ptr(), params.data_ptr(), B, F, H, grads.data_ptr(),
                                    bsc.data_ptr(), bsc.numel(), _st()))
        torch.cuda.synchronize()
        g = grads[:P].double()
        assert torch.isfinite(g).all()
        mask = (h > 0).double()
        dzd = dz.double()
        a2 = mask * _bf(dz).view(B, 1)                                           # what the tensor core multiplies
        assert torch.equal(unpack_L(hT, H, B).double().t(), a2)                  # hT now holds the A operand
        xq64 = _bf(x)
        g_w1 = (a2.t() @ xq64) * w2.view(H, 1)
        s_w1 = (a2.abs().t() @ xq64.abs()) * w2.abs().view(H, 1)
        g_b1 = (mask * dzd.view(B, 1)).sum(dim=0) * w2
        s_b1 = (mask * dzd.abs().view(B, 1)).sum(dim=0) * w2.abs()
        g_w2 = (h * dzd.view(B, 1)).sum(dim=0)
        s_w2 = (h * dzd.abs().view(B, 1)).sum(dim=0)
        ref = torch.cat([g_w1.reshape(-1), g_b1, g_w2, dzd.sum().reshape(1)])
        scale = torch.cat([s_w1.reshape(-1), s_b1, s_w2, dzd.abs().sum().reshape(1)])
        bad = (g - ref).abs() > 2e-5 * scale + 1e-30                             # fp32 sums of exactly these terms
        assert not bool(bad.any()), (rep, int(bad.sum()), int(bad.nonzero()[0]), float((g - ref).abs().max()))


@pytest.mark.parametrize("B,F", [(2048, 256), (4096, 1024), (8192, 4096), (512, 512), (65536, 4096), (4096, 8192)])
def test_fh_generator_is_torch_randn_bit_for_bit(lib, cuda_dev, B, F):
    """b200ds_fh_gen_batch == bf16 pack of torch.randn(B*F, device='cuda') under the same seed / offset, three consecutive
    draws (the third through the device draw counter); labels = sign of the fp64 row sums of the fp32 values wherever that
    sum is not within rounding of zero."""
    import ctypes as C
    assert lib.b200ds_fh_gen_supported(B, F) == 1
    seed = 1234 + B + F
    g = torch.Generator(device=cuda_dev); g.manual_seed(seed)
    xq = torch.empty(lib.b200ds_fh_operand_bytes(B, F), dtype=torch.uint8, device=cuda_dev)
    xq_ref = torch.empty_like(xq)
    y = torch.empty(B, device=cuda_dev)
    sc = torch.zeros(lib.b200ds_fh_gen_scratch_bytes(B, F), dtype=torch.uint8, device=cuda_dev)
    ctr = torch.zeros(1, dtype=torch.int64, device=cuda_dev)
    off = 0
    for draw in range(3):
        x = torch.randn(B * F, device=cuda_dev, generator=g)
        _chk(lib.b200ds_fh_pack_bf16(x.data_ptr(), B, F, xq_ref.data_ptr(), _st()))
        xq.zero_(); y.fill_(-1)
        adv = C.c_uint64(0)
        if draw < 2:
            _chk(lib.b200ds_fh_gen_batch(xq.data_ptr(), y.data_ptr(), B, F, seed, off, C.byref(adv), sc.data_ptr(), sc.numel(), _st()))
            off += adv.value
        else:
            ctr.fill_(draw)
            _chk(lib.b200ds_fh_gen_batch_counted(xq.data_ptr(), y.data_ptr(), B, F, seed, 0, ctr.data_ptr(), sc.data_ptr(), sc.numel(), _st()))
        torch.cuda.synchronize()
        assert torch.equal(xq, xq_ref), (B, F, draw)
        s64 = x.view(B, F)[:, : F // 2].double().sum(dim=1)
        clear = s64.abs() > 1e-3
        assert torch.equal(y[clear], (s64[clear] > 0).float()), (B, F, draw)
        assert bool(((y == 0) | (y == 1)).all())
    assert lib.b200ds_fh_gen_supported(2048, 768) == 0        # 768 does not divide the generator's 303104 threads


def test_fh_n_steps_track_fp32_autograd(lib, cuda_dev):
    """The wide-input step as bench.py runs it (generator -> GEMM forward -> logits -> loss -> backward -> Adam -> pack, device
    counters), 10 steps eagerly and 10 steps as a replayed CUDA graph: same bits; and against torch's own fp32 modules
    (nn.Linear / ReLU / BCEWithLogitsLoss / autograd / optim.Adam, the objects train.py builds) on the same batches: every
    step's loss within 2e-2; parameters after the 10 steps: each group (W1, b1, W2, b2) within 2e-2 of the fp32 group in
    norm (north_star's bf16 tolerance), no element further away than Adam's reach n * lr, and the step each group took
    within 1e-1 of fp32's.  (The step is the looser figure on purpose: bf16 layer-1 operands flip the relu of the
    ~0.3 % of (sample, unit) pairs whose pre-activation is within rounding of zero, which perturbs a W1 gradient -- a
    random-sign sum over the batch -- by ~sqrt(0.003) = 5 %, and Adam turns a gradient into a step of fixed size.)"""
    import ctypes as C

    import bench
    from b200ds import _lib
    wl = dict(B=4096, F=1024, H=512)
    B, F, H = wl["B"], wl["F"], wl["H"]
    n = 10

    def run(graph):
        ctx = C.c_void_p()
        _lib.check(lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
        ws = bench.WideInputStep(lib, ctx, cuda_dev, wl, 1, 0, "bf16")
        losses, ys = [], []
        ws.step(_st()); torch.cuda.synchronize(); losses.append(float(ws.loss)); ys.append(None)
        g = None
        if graph:
            g = torch.cuda.CUDAGraph()
            cap = torch.cuda.Stream(device=cuda_dev)
            cap.wait_stream(torch.cuda.current_stream())
            with torch.cuda.graph(g, stream=cap):
                ws.step(torch.cuda.current_stream().cuda_stream)
            torch.cuda.current_stream().wait_stream(cap)
        for k in range(1, n):
            ys.append(ws.y.clone())
            g.replay() if graph else ws.step(_st())
            torch.cuda.synchronize()
            losses.append(float(ws.loss))
        if g is not None:
            g.reset()
        out = (losses, ys, ws.params[: ws.P].clone(), ws.seed)
        _lib.check(lib.b200ds_ctx_destroy(ctx))
        return out

    l_e, ys, p_e, seed = run(False)
    l_g, _, p_g, _ = run(True)
    assert l_e == l_g and torch.equal(p_e, p_g)
    # torch fp32 on the same batches
    flat0 = bench.init_flat(F, H).to(cuda_dev)
    net = torch.nn.Sequential(torch.nn.Linear(F, H), torch.nn.ReLU(), torch.nn.Linear(H, 1)).to(cuda_dev)
    with torch.no_grad():
        net[0].weight.copy_(flat0[: H * F].view(H, F)); net[0].bias.copy_(flat0[H * F: H * F + H])
        net[2].weight.copy_(flat0[H * F + H: H * F + 2 * H].view(1, H)); net[2].bias.copy_(flat0[H * F + 2 * H:])
    opt = torch.optim.Adam(net.parameters(), lr=bench.LR)
    crit = torch.nn.BCEWithLogitsLoss()
    gen = torch.Generator(device=cuda_dev); gen.manual_seed(seed)
    torch.backends.cuda.matmul.allow_tf32 = False
    for k in range(n):
        x = torch.randn(B * F, device=cuda_dev, generator=gen).view(B, F)
        y = (x[:, : F // 2].double().sum(dim=1) > 0).float() if ys[k] is None else ys[k]
        loss = crit(net(x), y.view(B, 1))
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        assert abs(l_e[k] - float(loss.detach())) <= 2e-2 * abs(float(loss.detach())), (k, l_e[k], float(loss.detach()))
    ref = torch.cat([t.detach().reshape(-1) for t in (net[0].weight, net[0].bias, net[2].weight, net[2].bias)])
    d = (p_e.double() - ref.double()).abs()
    assert float(d.max()) <= n * bench.LR, float(d.max())
    for name, o, cnt in (("w1", 0, H * F), ("b1", H * F, H), ("w2", H * F + H, H), ("b2", H * F + 2 * H, 1)):
        a, r, z0 = p_e[o:o + cnt].double(), ref[o:o + cnt].double(), flat0[o:o + cnt].double()
        assert float((a - r).norm()) <= 2e-2 * float(r.norm()), (name, float((a - r).norm() / r.norm()))
        assert float((a - r).norm()) <= 1e-1 * float((r - z0).norm()), (name, float((a - r).norm() / (r - z0).norm()))


def test_engine_routes_wide_inputs_through_the_gemm_path(cuda_dev):
    """deepspeed-style calls on a model with n_features = 512: bf16.enabled sends engine(x) / backward / step through the
    wide-input kernels (batch of 300 padded to 2048 rows, hidden 192 to 256); six steps track the oracle's fp32 step on
    the same batches within north_star's bf16 tolerance, the padded hidden units stay zero."""
    import b200ds
    F, H, B, steps = 512, 192, 300, 6
    torch.manual_seed(21)
    net = torch.nn.Sequential(torch.nn.Linear(F, H), torch.nn.ReLU(), torch.nn.Linear(H, 1))
    p = {"w1": net[0].weight.detach().clone(), "b1": net[0].bias.detach().clone(),
         "w2": net[2].weight.detach().clone(), "b2": net[2].bias.detach().clone()}
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)
    engine, _, _, _ = b200ds.initialize(model=net, optimizer=opt, config={"bf16": {"enabled": True}, "zero_optimization": {"stage": 0}})
    assert engine._wide and engine.H == 256
    crit = b200ds.BCEWithLogitsLoss()
    flat = ora.flatten(p)
    st = ora.AdamState()
    for s in range(steps):
        x = torch.randn(B, F)
        y = ora.label_rule(x)
        logits = engine(x.to(cuda_dev))
        assert logits.shape == (B, 1)
        loss = crit(logits, y.to(cuda_dev))
        engine.backward(loss)
        engine.step()
        pp = ora.unflatten(flat, F, H)
        z, h = ora.forward(pp, x)
        want = float(ora.bce_with_logits(z, y))
        ora.adam_update(flat, ora.flatten(ora.backward(pp, x, y, z, h)), st, 1e-3)
        assert abs(loss.item() - want) <= 2e-2 * abs(want), (s, loss.item(), want)
    got = torch.cat([t.detach().reshape(-1).cpu() for t in (net[0].weight, net[0].bias, net[2].weight, net[2].bias)])
    assert float((got - flat).norm()) <= 2e-2 * float(flat.norm())
    fl = engine.flat.cpu()
    assert float(fl[H * F: 256 * F].abs().max()) == 0.0          # padded W1 rows


def test_host_step_takes_wide_shapes_the_backward_cannot_cut(lib, cuda_dev):
    """b200ds_fh_supported is the forward's shape class; the backward needs batch % 2048 == 0 on top
    (b200ds_fh_train_supported).  A host-buffer step on a wide shape outside that (B = 256) must run -- on the fp32
    kernels -- and not fail in the middle of the GEMM route; its loss is the oracle's fp32 loss."""
    from b200ds import _lib
    B, F, H = 256, 256, 128
    assert lib.b200ds_fh_supported(B, F, H) == 1 and lib.b200ds_fh_train_supported(B, F, H) == 0
    assert lib.b200ds_fh_train_supported(2048, F, H) == 1
    torch.manual_seed(21)
    p = ora.init_params(F, H)
    flat = ora.flatten(p)
    P = flat.numel(); Ppad = (P + 3) // 4 * 4
    z = lambda n: torch.zeros(n, dtype=torch.float32, device=cuda_dev)      # noqa: E731
    params = z(Ppad); params[:P] = flat.to(cuda_dev)
    grads, m, v, xs, ys, lg, dl, ls = z(Ppad), z(Ppad), z(Ppad), z(B * F), z(B), z(B), z(B), z(1)
    sc = torch.zeros(max(256, lib.b200ds_train_step_scratch_bytes(B, F, H)), dtype=torch.uint8, device=cuda_dev)
    xq = torch.empty(lib.b200ds_fh_operand_bytes(B, F), dtype=torch.uint8, device=cuda_dev)
    wq = torch.empty(lib.b200ds_fh_operand_bytes(H, F), dtype=torch.uint8, device=cuda_dev)
    bufs = _lib.StepBufs(params.data_ptr(), grads.data_ptr(), m.data_ptr(), v.data_ptr(), xs.data_ptr(), ys.data_ptr(),
                         lg.data_ptr(), dl.data_ptr(), ls.data_ptr(), sc.data_ptr(), sc.numel(), xq.data_ptr(), wq.data_ptr())
    ctx = C.c_void_p()
    _chk(lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
    xh = torch.randn(B, F).pin_memory(); yh = ora.label_rule(xh).reshape(-1).contiguous().pin_memory(); lh = torch.zeros(1).pin_memory()
    zref, _ = ora.forward(ora.unflatten(flat, F, H), xh.clone())
    lref = ora.bce_with_logits(zref, yh.reshape(-1, 1).clone())
    _chk(lib.b200ds_train_step_host(ctx, C.byref(bufs), xh.data_ptr(), yh.data_ptr(), B, F, H, 1e-3, 0.9, 0.999, 1e-8, 0,
                                    lh.data_ptr(), torch.cuda.current_stream().cuda_stream))
    assert abs(float(lh) - float(lref)) <= 1e-5 * max(1.0, abs(float(lref)))      # the fp32 kernels' tolerance, not bf16's
    _chk(lib.b200ds_ctx_destroy(ctx))
