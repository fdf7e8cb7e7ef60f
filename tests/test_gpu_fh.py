"""Parity of the wide-input path (F in the hundreds / thousands: K-loop tcgen05 GEMMs fed by tensor-map TMA,
csrc/mlp_fh.cu) through the C ABI, against an fp64 evaluation of the same bf16-operand formulas on the GPU.
SimpleNet(n_features, hidden) takes any n_features (train.py:26-36); train.py:107,113 are the calls replaced."""
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


@pytest.mark.parametrize("B,F,H", [(256, 256, 128), (512, 256, 256), (1024, 1024, 1024), (8192, 512, 384), (4096, 4096, 4096)])
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


@pytest.mark.parametrize("B,F,H", [(2048, 256, 128), (4096, 512, 384), (2048, 1024, 1024), (8192, 4096, 2048)])
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
