"""Parity of the bf16 tensor-core path (tcgen05/TMEM/TMA kernels) through the C ABI.

Checker: the oracle's bf16 restatement (oracle/reference_step._bf16_step semantics: GEMM
operands rounded to bf16, fp32 accumulation) evaluated in fp64 by torch on the GPU for the
big shapes.  Tolerances: logits 1e-4 relative (same bf16 operands, only the summation order
and the hi/lo bias split differ); gradients 2e-2 (north_star's bf16 tolerance; dz is rounded
to bf16 once on its way into the tensor core)."""
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


_FWD_SC = {}


def _fwd_tc(lib, xq, wq, params, B, F, H, logits):
    """b200ds_mlp_forward_tc with its slice scratch (zeroed once per shape, left zeroed by the kernel)."""
    key = (B, F, H, logits.device.index)
    if key not in _FWD_SC:
        _FWD_SC[key] = torch.zeros(max(256, lib.b200ds_mlp_forward_tc_scratch_bytes(B, F, H)), dtype=torch.uint8, device=logits.device)
    sc = _FWD_SC[key]
    _chk(lib.b200ds_mlp_forward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, logits.data_ptr(), sc.data_ptr(),
                                   sc.numel(), _st()))


def _bf(t):
    return t.to(torch.bfloat16).to(torch.float64)


def _reference(p, x, y):
    """fp64 evaluation of the bf16-operand step (oracle._bf16_step, written for any device).
    Besides the values it returns, per logit and per gradient element, the sum of the ABSOLUTE values of the
    terms that element is a sum of: the yardstick of an element-wise error bound (a bf16 rounding of one factor
    moves every term by 2^-9 of its size, whatever the terms cancel to)."""
    xq, w1q = _bf(x), _bf(p["w1"])
    z1 = xq @ w1q.t() + p["b1"].double()
    h = torch.clamp_min(z1, 0)
    w2 = p["w2"].double()
    z = h @ w2.t() + p["b2"].double()
    z_abs = h @ w2.abs().t() + p["b2"].double().abs()
    B = x.shape[0]
    dz = (torch.sigmoid(z) - y.double()) / B
    g_w2 = dz.t() @ h
    g_b2 = dz.sum(dim=0)
    mask = (h > 0).double()
    dz1 = (dz @ w2) * mask
    g_w1 = dz1.t() @ xq
    g_b1 = dz1.sum(dim=0)
    a1 = dz1.abs()
    cat = lambda *t: torch.cat([v.reshape(-1) for v in t])      # noqa: E731
    g = cat(g_w1, g_b1, g_w2, g_b2)
    g_abs = cat(a1.t() @ xq.abs(), a1.sum(dim=0), dz.abs().t() @ h, dz.abs().sum(dim=0))
    return z.reshape(-1), dz.reshape(-1), g, z_abs.reshape(-1), g_abs


# Element-wise bounds (every element is held to its own bound; nothing is normalised by the largest entry):
#   logits     |z - ref| <= 1e-5 * (sum of |terms| of that logit): same bf16 operands as the reference, only the fp32
#              summation order, the hi/lo split of b1 (2^-17) and the |.|-form of relu differ
#   gradients  |g - ref| <= 2e-2 * |ref| + 2^-8 * (sum of |terms|): dz goes into the tensor core rounded to bf16
#              (2^-9 per term), north_star's 2e-2 on the value itself
def _assert_logits(z, z_ref, z_abs):
    bad = (z.double() - z_ref).abs() > 1e-5 * z_abs + 1e-30
    assert not bool(bad.any()), (int(bad.sum()), float((z.double() - z_ref).abs().max()))


def _assert_grads(g, g_ref, g_abs, what=""):
    err = (g.double() - g_ref).abs()
    bad = err > 2e-2 * g_ref.abs() + 2.0 ** -8 * g_abs + 1e-30
    assert not bool(bad.any()), (what, int(bad.sum()), int(bad.nonzero()[0]), float(err.max()))


SHAPES = [(128, 20, 128), (256, 20, 256), (1024, 20, 384), (4096, 30, 2048), (2048, 1, 512), (65536, 20, 4096),
          (8192, 20, 8192), (384, 7, 640)]


@pytest.mark.parametrize("B,F,H", SHAPES)
def test_tc_forward_backward_vs_bf16_oracle(lib, cuda_dev, B, F, H):
    assert lib.b200ds_tc_supported(B, F, H) == 1
    torch.manual_seed(B + 31 * F + 7 * H)
    p = {k: v.to(cuda_dev) for k, v in ora.init_params(F, H).items()}
    p["b1"] = p["b1"] + 0.05 * torch.randn(H, device=cuda_dev)          # make the bias matter
    x = torch.randn(B, F, device=cuda_dev)
    y = (x[:, : max(1, F // 2)].sum(dim=1, keepdim=True) > 0).float()
    flat = torch.cat([p[k].reshape(-1) for k in ora.PARAM_ORDER])
    P = flat.numel()
    Ppad = (P + 3) // 4 * 4
    params = torch.zeros(Ppad, device=cuda_dev); params[:P] = flat
    xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=cuda_dev)
    wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=cuda_dev)
    logits = torch.full((B,), float("nan"), device=cuda_dev)
    grads = torch.full((Ppad,), float("nan"), device=cuda_dev)
    sc = torch.zeros(max(256, lib.b200ds_mlp_backward_tc_scratch_bytes(B, F, H)), dtype=torch.uint8, device=cuda_dev)
    _chk(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), _st()))
    _chk(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), _st()))
    _fwd_tc(lib, xq, wq, params, B, F, H, logits)
    torch.cuda.synchronize()
    z_ref, dz_ref, g_ref, z_abs, g_abs = _reference(p, x, y)
    _assert_logits(logits, z_ref, z_abs)
    dz = dz_ref.float().contiguous()
    for rep in range(2):       # twice: the second call must not depend on state left by the first
        grads.fill_(float("nan"))
        _chk(lib.b200ds_mlp_backward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), dz.data_ptr(), B, F, H,
                                        grads.data_ptr(), sc.data_ptr(), sc.numel(), _st()))
        torch.cuda.synchronize()
        g = grads[:P].double()
        assert torch.isfinite(g).all()
        _assert_grads(g, g_ref, g_abs, rep)


@pytest.mark.parametrize("B,H", [(131072, 128), (131072, 256), (131072, 384), (131072, 512), (262144, 128), (98304, 1024)])
def test_tc_forward_many_tiles_per_cta_small_hidden(lib, cuda_dev, B, H):
    """With few chunks per tile (H <= 512) the MMAs of a CTA run several batch tiles ahead of its epilogue, and
    every CTA walks five or more tiles: an x stage must not be refilled before the epilogue warps have read their
    rows of it for the linear term.  Run three times: a race would not show up every time."""
    F = 20
    torch.manual_seed(B // 128 + H)
    p = {k: v.to(cuda_dev) for k, v in ora.init_params(F, H).items()}
    p["b1"] = p["b1"] + 0.05 * torch.randn(H, device=cuda_dev)
    x = torch.randn(B, F, device=cuda_dev)
    y = (x[:, : F // 2].sum(dim=1, keepdim=True) > 0).float()
    flat = torch.cat([p[k].reshape(-1) for k in ora.PARAM_ORDER])
    P = flat.numel(); Ppad = (P + 3) // 4 * 4
    params = torch.zeros(Ppad, device=cuda_dev); params[:P] = flat
    xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=cuda_dev)
    wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=cuda_dev)
    _chk(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), _st()))
    _chk(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), _st()))
    z_ref, _, _, z_abs, _ = _reference(p, x, y)
    for rep in range(3):
        logits = torch.full((B,), float("nan"), device=cuda_dev)
        _fwd_tc(lib, xq, wq, params, B, F, H, logits)
        torch.cuda.synchronize()
        _assert_logits(logits, z_ref, z_abs)


def test_tc_forward_is_bit_reproducible_with_many_slices(lib, cuda_dev):
    """H = 8192 is four slices of hidden units in different CTAs: their parts of a logit are summed in slice
    order, so two runs give the same bits (and the scratch is left ready for the next call)."""
    B, F, H = 16384, 20, 8192
    torch.manual_seed(9)
    flat = ora.flatten(ora.init_params(F, H)).to(cuda_dev)
    P = flat.numel(); Ppad = (P + 3) // 4 * 4
    params = torch.zeros(Ppad, device=cuda_dev); params[:P] = flat
    x = torch.randn(B, F, device=cuda_dev)
    xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=cuda_dev)
    wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=cuda_dev)
    _chk(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), _st()))
    _chk(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), _st()))
    assert lib.b200ds_mlp_forward_tc_scratch_bytes(B, F, H) > 0
    outs = []
    for rep in range(4):
        z = torch.full((B,), float("nan"), device=cuda_dev)
        _fwd_tc(lib, xq, wq, params, B, F, H, z)
        torch.cuda.synchronize()
        outs.append(z)
    for z in outs[1:]:
        assert torch.equal(outs[0], z)
    assert lib.b200ds_mlp_forward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, outs[0].data_ptr(), None, 0, _st()) == 1


def test_tc_rejects_unsupported_shapes(lib, cuda_dev):
    assert lib.b200ds_tc_supported(100, 20, 128) == 0        # batch not a multiple of 128
    assert lib.b200ds_tc_supported(128, 31, 128) == 0        # F + 2 > 32
    assert lib.b200ds_tc_supported(128, 20, 64) == 0         # H not a multiple of 128
    assert lib.b200ds_mlp_forward_tc(None, None, None, 100, 20, 128, None, None, 0, _st()) == 1
    assert lib.b200ds_mlp_backward_tc_scratch_bytes(100, 20, 128) == 0


def test_tc_matches_fp32_path_within_bf16_tolerance(lib, cuda_dev):
    """The two product paths against each other at full size: same loss to 2e-2."""
    B, F, H = 65536, 20, 4096
    torch.manual_seed(3)
    flat = ora.flatten(ora.init_params(F, H)).to(cuda_dev)
    P = flat.numel(); Ppad = (P + 3) // 4 * 4
    params = torch.zeros(Ppad, device=cuda_dev); params[:P] = flat
    x = torch.empty(B * F, device=cuda_dev)
    _chk(lib.b200ds_philox_normal(x.data_ptr(), B * F, 42, 0, None, _st()))
    y = torch.empty(B, device=cuda_dev)
    _chk(lib.b200ds_label_rule(x.data_ptr(), y.data_ptr(), B, F, _st()))
    xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=cuda_dev)
    wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=cuda_dev)
    z32 = torch.empty(B, device=cuda_dev); z16 = torch.empty(B, device=cuda_dev)
    _chk(lib.b200ds_mlp_forward_f32(params.data_ptr(), x.data_ptr(), B, F, H, z32.data_ptr(), _st()))
    _chk(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), _st()))
    _chk(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), _st()))
    _fwd_tc(lib, xq, wq, params, B, F, H, z16)
    l32 = torch.zeros(1, device=cuda_dev); l16 = torch.zeros(1, device=cuda_dev)
    sc = torch.zeros(lib.b200ds_bce_scratch_bytes(B), dtype=torch.uint8, device=cuda_dev)
    _chk(lib.b200ds_bce_loss_grad(z32.data_ptr(), y.data_ptr(), B, l32.data_ptr(), None, None, sc.data_ptr(), sc.numel(), _st()))
    _chk(lib.b200ds_bce_loss_grad(z16.data_ptr(), y.data_ptr(), B, l16.data_ptr(), None, None, sc.data_ptr(), sc.numel(), _st()))
    torch.cuda.synchronize()
    assert abs(float(l16) - float(l32)) <= 2e-2 * abs(float(l32))
    assert float((z16 - z32).abs().max()) <= 2e-2 * float(z32.abs().max())


def test_engine_bf16_config_trains_within_tolerance(cuda_dev):
    """bf16.enabled in the engine config routes engine(x)/backward through the tcgen05
    kernels; N steps must track the oracle's bf16 restatement within 2e-2 (north_star)."""
    import b200ds
    F, H, B, steps = 20, 256, 256, 12
    torch.manual_seed(11)
    net = torch.nn.Sequential(torch.nn.Linear(F, H), torch.nn.ReLU(), torch.nn.Linear(H, 1))
    p = {"w1": net[0].weight.detach().clone(), "b1": net[0].bias.detach().clone(),
         "w2": net[2].weight.detach().clone(), "b2": net[2].bias.detach().clone()}
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)
    cfg = {"train_batch_size": B, "gradient_accumulation_steps": 1, "bf16": {"enabled": True}, "zero_optimization": {"stage": 0}}
    engine, _, _, _ = b200ds.initialize(model=net, optimizer=opt, config=cfg)
    crit = b200ds.BCEWithLogitsLoss()
    flat = ora.flatten(p)
    st = ora.AdamState()
    for s in range(steps):
        x = torch.randn(B, F)
        y = ora.label_rule(x)
        loss = crit(engine(x.to(cuda_dev)), y.to(cuda_dev))
        engine.backward(loss)
        engine.step()
        assert engine._tc
        pp = ora.unflatten(flat, F, H)
        lref, g = ora._bf16_step(pp, x, y)
        ora.adam_update(flat, ora.flatten(g), st, 1e-3)
        assert abs(loss.item() - float(lref)) <= 2e-2 * abs(float(lref)), s
    got = engine.flat[: engine.P].cpu()
    assert float((got - flat).abs().max()) <= 2e-2 * float(flat.abs().max())


def test_host_buffer_step_tc_path(lib, cuda_dev):
    from b200ds import _lib
    B, F, H = 512, 20, 256
    torch.manual_seed(5)
    p = ora.init_params(F, H)
    flat = ora.flatten(p)
    P = flat.numel(); Ppad = (P + 3) // 4 * 4
    z = lambda n: torch.zeros(n, dtype=torch.float32, device=cuda_dev)      # noqa: E731
    params = z(Ppad); params[:P] = flat.to(cuda_dev)
    grads, m, v, xs, ys, lg, dl, ls = z(Ppad), z(Ppad), z(Ppad), z(B * F), z(B), z(B), z(B), z(1)
    sc = torch.zeros(max(256, lib.b200ds_train_step_scratch_bytes(B, F, H)), dtype=torch.uint8, device=cuda_dev)
    xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=cuda_dev)
    wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=cuda_dev)
    bufs = _lib.StepBufs(params.data_ptr(), grads.data_ptr(), m.data_ptr(), v.data_ptr(), xs.data_ptr(), ys.data_ptr(),
                         lg.data_ptr(), dl.data_ptr(), ls.data_ptr(), sc.data_ptr(), sc.numel(), xq.data_ptr(), wq.data_ptr())
    ctx = C.c_void_p()
    _chk(lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
    st = ora.AdamState()
    xh = torch.empty(B, F).pin_memory(); yh = torch.empty(B).pin_memory(); lh = torch.zeros(1).pin_memory()
    for step in range(3):
        xh.copy_(torch.randn(B, F)); yh.copy_(ora.label_rule(xh).reshape(-1))
        pp = ora.unflatten(flat, F, H)
        lref, g = ora._bf16_step(pp, xh.clone(), yh.reshape(-1, 1).clone())
        ora.adam_update(flat, ora.flatten(g), st, 1e-3)
        _chk(lib.b200ds_train_step_host(ctx, C.byref(bufs), xh.data_ptr(), yh.data_ptr(), B, F, H, 1e-3, 0.9, 0.999, 1e-8,
                                        step, lh.data_ptr(), _st()))
        assert abs(float(lh) - float(lref)) <= 2e-2 * abs(float(lref)), step
    assert float((params[:P].cpu() - flat).abs().max()) <= 2e-2 * float(flat.abs().max())
    _chk(lib.b200ds_ctx_destroy(ctx))


@pytest.mark.parametrize("B,F,H", [(128, 20, 128), (1024, 20, 384), (65536, 20, 4096), (384, 7, 640)])
def test_fused_loss_backward_equals_loss_then_backward(lib, cuda_dev, B, F, H):
    """b200ds_mlp_loss_backward_tc == b200ds_bce_loss_grad + b200ds_mlp_backward_tc: W1 / b1 / W2
    gradients bit for bit (the same dlogits arithmetic inside the kernel), db2 and the loss to fp32
    round-off (another summation order), step counter incremented once; twice in a row."""
    torch.manual_seed(B + H)
    p = ora.init_params(F, H)
    flat = ora.flatten(p)
    P = flat.numel(); Ppad = (P + 3) // 4 * 4
    params = torch.zeros(Ppad, device=cuda_dev); params[:P] = flat.to(cuda_dev)
    x = torch.randn(B, F, device=cuda_dev)
    y = (x[:, : max(1, F // 2)].sum(dim=1) > 0).float().contiguous()
    xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=cuda_dev)
    wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=cuda_dev)
    z = torch.empty(B, device=cuda_dev)
    _chk(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), _st()))
    _chk(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), _st()))
    _fwd_tc(lib, xq, wq, params, B, F, H, z)
    sc = torch.zeros(lib.b200ds_mlp_backward_tc_scratch_bytes(B, F, H), dtype=torch.uint8, device=cuda_dev)
    bsc = torch.zeros(lib.b200ds_bce_scratch_bytes(B), dtype=torch.uint8, device=cuda_dev)
    dz = torch.empty(B, device=cuda_dev)
    la, lb = torch.zeros(1, device=cuda_dev), torch.zeros(1, device=cuda_dev)
    ga, gb = torch.zeros(Ppad, device=cuda_dev), torch.full((Ppad,), float("nan"), device=cuda_dev)
    ca, cb = torch.zeros(1, dtype=torch.int64, device=cuda_dev), torch.zeros(1, dtype=torch.int64, device=cuda_dev)
    _chk(lib.b200ds_bce_loss_grad(z.data_ptr(), y.data_ptr(), B, la.data_ptr(), dz.data_ptr(), ca.data_ptr(), bsc.data_ptr(),
                                  bsc.numel(), _st()))
    _chk(lib.b200ds_mlp_backward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), dz.data_ptr(), B, F, H, ga.data_ptr(),
                                    sc.data_ptr(), sc.numel(), _st()))
    for rep in range(2):
        gb.fill_(float("nan"))
        _chk(lib.b200ds_mlp_loss_backward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), z.data_ptr(), y.data_ptr(), B, F, H,
                                             lb.data_ptr(), gb.data_ptr(), cb.data_ptr(), sc.data_ptr(), sc.numel(), _st()))
        torch.cuda.synchronize()
        assert torch.equal(ga[:P - 1], gb[:P - 1]), rep
        assert abs(float(ga[P - 1]) - float(gb[P - 1])) <= 1e-5 * max(1e-30, abs(float(ga[P - 1]))) + 1e-9, rep
        assert abs(float(la) - float(lb)) <= 1e-5 * abs(float(la)), rep
    assert int(ca) == 1 and int(cb) == 2


@pytest.mark.parametrize("B,F,H", [(128, 20, 128), (1024, 20, 384), (65536, 20, 4096), (384, 7, 640), (256, 30, 256), (256, 1, 128)])
def test_fused_update_equals_finish_adam_pack(lib, cuda_dev, B, F, H):
    """b200ds_mlp_loss_backward_partials_tc + b200ds_update_tc (one rank) against
    b200ds_mlp_loss_backward_tc + b200ds_adam_step + b200ds_pack_params_bf16 over three steps:
    parameters, Adam moments, gradients, the packed operand and the loss bit for bit."""
    torch.manual_seed(7 * B + H)
    flat = ora.flatten(ora.init_params(F, H))
    P = flat.numel(); Ppad = (P + 3) // 4 * 4
    z = lambda n: torch.zeros(n, dtype=torch.float32, device=cuda_dev)      # noqa: E731
    st = [dict(p=z(Ppad), g=z(Ppad), m=z(Ppad), v=z(Ppad), loss=z(1), ctr=torch.zeros(1, dtype=torch.int64, device=cuda_dev),
               wq=torch.zeros(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=cuda_dev),
               sc=torch.zeros(lib.b200ds_mlp_backward_tc_scratch_bytes(B, F, H), dtype=torch.uint8, device=cuda_dev)) for _ in range(2)]
    for s_ in st:
        s_["p"][:P] = flat.to(cuda_dev)
        _chk(lib.b200ds_pack_params_bf16(s_["p"].data_ptr(), F, H, s_["wq"].data_ptr(), _st()))
    ctx = C.c_void_p()
    _chk(lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
    xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=cuda_dev)
    lg = z(B)
    args = (1e-3, 0.9, 0.999, 1e-8)
    for step in range(1, 4):
        x = torch.randn(B, F, device=cuda_dev)
        y = (x[:, : max(1, F // 2)].sum(dim=1) > 0).float().contiguous()
        _chk(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), _st()))
        a, b = st
        _fwd_tc(lib, xq, a["wq"], a["p"], B, F, H, lg)
        _chk(lib.b200ds_mlp_loss_backward_tc(xq.data_ptr(), a["wq"].data_ptr(), a["p"].data_ptr(), lg.data_ptr(), y.data_ptr(), B, F, H,
                                             a["loss"].data_ptr(), a["g"].data_ptr(), a["ctr"].data_ptr(), a["sc"].data_ptr(),
                                             a["sc"].numel(), _st()))
        ga = a["g"].clone()
        _chk(lib.b200ds_adam_step(a["p"].data_ptr(), a["g"].data_ptr(), a["m"].data_ptr(), a["v"].data_ptr(), Ppad, *args, step, 1.0, 0, _st()))
        _chk(lib.b200ds_pack_params_bf16(a["p"].data_ptr(), F, H, a["wq"].data_ptr(), _st()))
        _fwd_tc(lib, xq, b["wq"], b["p"], B, F, H, lg)
        _chk(lib.b200ds_mlp_loss_backward_partials_tc(xq.data_ptr(), b["wq"].data_ptr(), lg.data_ptr(), y.data_ptr(), B, F, H,
                                                      b["ctr"].data_ptr(), b["sc"].data_ptr(), b["sc"].numel(), _st()))
        if step % 2:
            _chk(lib.b200ds_update_tc(ctx, b["sc"].data_ptr(), b["sc"].numel(), B, F, H, b["p"].data_ptr(), b["g"].data_ptr(),
                                      b["m"].data_ptr(), b["v"].data_ptr(), b["wq"].data_ptr(), *args, step, None,
                                      b["loss"].data_ptr(), _st()))
        else:
            _chk(lib.b200ds_update_tc(ctx, b["sc"].data_ptr(), b["sc"].numel(), B, F, H, b["p"].data_ptr(), b["g"].data_ptr(),
                                      b["m"].data_ptr(), b["v"].data_ptr(), b["wq"].data_ptr(), *args, 0, b["ctr"].data_ptr(),
                                      b["loss"].data_ptr(), _st()))
        torch.cuda.synchronize()
        assert torch.equal(ga[:P], b["g"][:P]), step
        for k in ("p", "m", "v", "wq", "loss", "ctr"):
            assert torch.equal(a[k], b[k]), (k, step)
    _chk(lib.b200ds_ctx_destroy(ctx))


def _host_step_state(lib, dev, B, F, H, stage_batches):
    from b200ds import _lib
    torch.manual_seed(11)
    flat = ora.flatten(ora.init_params(F, H))
    P = flat.numel(); Ppad = (P + 3) // 4 * 4
    z = lambda n: torch.zeros(n, dtype=torch.float32, device=dev)      # noqa: E731
    t = dict(params=z(Ppad), grads=z(Ppad), m=z(Ppad), v=z(Ppad), xs=z(stage_batches * B * F), ys=z(stage_batches * B),
             lg=z(B), dl=z(B), ls=z(1),
             sc=torch.zeros(max(256, lib.b200ds_train_step_scratch_bytes(B, F, H)), dtype=torch.uint8, device=dev),
             xq=torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=dev),
             wq=torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=dev))
    t["params"][:P] = flat.to(dev)
    bufs = _lib.StepBufs(t["params"].data_ptr(), t["grads"].data_ptr(), t["m"].data_ptr(), t["v"].data_ptr(),
                         t["xs"].data_ptr(), t["ys"].data_ptr(), t["lg"].data_ptr(), t["dl"].data_ptr(), t["ls"].data_ptr(),
                         t["sc"].data_ptr(), t["sc"].numel(), t["xq"].data_ptr(), t["wq"].data_ptr())
    return t, bufs, P


@pytest.mark.parametrize("B,F,H", [(512, 20, 256), (4, 20, 64)])
def test_prefetching_host_step_equals_the_plain_one(lib, cuda_dev, B, F, H):
    """b200ds_train_step_host_prefetch: same losses and parameters, bit for bit, as
    b200ds_train_step_host over the same batches -- with the next batch announced, with a
    batch that was NOT the announced one (prefetch discarded), and with no announcement."""
    torch.manual_seed(3)
    n = 6
    xh = [torch.randn(B, F).pin_memory() for _ in range(n)]
    yh = [ora.label_rule(x).reshape(-1).contiguous().pin_memory() for x in xh]
    lh = torch.zeros(1).pin_memory()
    order = [0, 1, 2, 4, 5, 3]               # step 3 consumes batch 4 although batch 3 was announced
    announce = [1, 2, 3, 5, None, None]      # step 4 announces nothing, step 5 arrives unannounced
    step0s = [0, 1, 2, 3, 40, 41]            # the Adam step number jumps (a restart from a checkpoint): the graph path keeps it in a
                                             # device word that its loss kernel increments and must notice a call that is out of sequence
    res = []
    for prefetch in (False, True):
        t, bufs, P = _host_step_state(lib, cuda_dev, B, F, H, 2 if prefetch else 1)
        ctx = C.c_void_p()
        _chk(lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
        losses = []
        for step, i, nxt in zip(step0s, order, announce):
            if prefetch:
                xn = xh[nxt].data_ptr() if nxt is not None else None
                yn = yh[nxt].data_ptr() if nxt is not None else None
                _chk(lib.b200ds_train_step_host_prefetch(ctx, C.byref(bufs), xh[i].data_ptr(), yh[i].data_ptr(), xn, yn, B, F, H,
                                                         1e-3, 0.9, 0.999, 1e-8, step, lh.data_ptr(), _st()))
            else:
                _chk(lib.b200ds_train_step_host(ctx, C.byref(bufs), xh[i].data_ptr(), yh[i].data_ptr(), B, F, H,
                                                1e-3, 0.9, 0.999, 1e-8, step, lh.data_ptr(), _st()))
            losses.append(float(lh))
        res.append((losses, t["params"][:P].cpu().clone()))
        _chk(lib.b200ds_ctx_destroy(ctx))
    assert res[0][0] == res[1][0]
    assert torch.equal(res[0][1], res[1][1])


@pytest.mark.parametrize("B,F", [(128, 20), (65536, 20), (1024, 30), (256, 1), (4096, 7)])
def test_fused_batch_generator_is_bit_identical_to_the_three_kernel_path(lib, cuda_dev, B, F):
    """b200ds_gen_batch_bf16 == philox_normal -> label_rule -> pack_x, byte for byte, also at a
    later offset and through the device draw counter."""
    seed = 42 + B
    x = torch.empty(B * F, device=cuda_dev)
    y_ref = torch.empty(B, device=cuda_dev)
    xq_ref = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=cuda_dev)
    xq = torch.empty_like(xq_ref)
    y = torch.empty(B, device=cuda_dev)
    adv = C.c_uint64(0)
    off = 0
    ctr = torch.zeros(1, dtype=torch.int64, device=cuda_dev)
    for draw in range(3):
        _chk(lib.b200ds_philox_normal(x.data_ptr(), B * F, seed, off, C.byref(adv), _st()))
        _chk(lib.b200ds_label_rule(x.data_ptr(), y_ref.data_ptr(), B, F, _st()))
        _chk(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq_ref.data_ptr(), _st()))
        xq.zero_(); y.fill_(-1)
        adv2 = C.c_uint64(0)
        if draw < 2:
            _chk(lib.b200ds_gen_batch_bf16(xq.data_ptr(), y.data_ptr(), B, F, seed, off, C.byref(adv2), _st()))
            assert adv2.value == adv.value
        else:
            ctr.fill_(draw)
            _chk(lib.b200ds_gen_batch_bf16_counted(xq.data_ptr(), y.data_ptr(), B, F, seed, 0, ctr.data_ptr(), _st()))
        torch.cuda.synchronize()
        assert torch.equal(xq, xq_ref), (B, F, draw)
        assert torch.equal(y, y_ref), (B, F, draw)
        off += adv.value
    # and the fp32 values underneath are torch's own CUDA generator stream
    g = torch.Generator(device=cuda_dev); g.manual_seed(seed)
    assert torch.equal(torch.randn(B * F, device=cuda_dev, generator=g), _first_draw(lib, cuda_dev, B * F, seed))


def _first_draw(lib, dev, n, seed):
    out = torch.empty(n, device=dev)
    _chk(lib.b200ds_philox_normal(out.data_ptr(), n, seed, 0, None, _st()))
    return out
