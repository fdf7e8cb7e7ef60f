"""The loss + gradient entry point that takes no scratch argument.

Some hosts of the C ABI cannot give the library device scratch: a launcher that owns no
allocator (it only forwards the tensors it was handed), or a training loop whose framework
allocator already holds the GPU's memory, where a plugin-side cudaMalloc can fail or stall
the step.  b200ds_bce_loss_grad_noalloc serves them from a small array built into the
library.  Here it is called through ctypes, twice in a row on different inputs (the array is
shared and keeps the previous call's contents), and compared with the caller-scratch entry
point at the tolerance the other fp32 tests use.  Collected last on purpose.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _st():
    return torch.cuda.current_stream().cuda_stream


def _both(lib, dev, batch, seed):
    from b200ds import _lib
    torch.manual_seed(seed)
    z = (torch.randn(batch) * 3).to(dev)
    y = (torch.rand(batch) > 0.5).float().to(dev)
    sc = torch.zeros(lib.b200ds_bce_scratch_bytes(batch), dtype=torch.uint8, device=dev)
    loss_a = torch.zeros(1, device=dev); dz_a = torch.empty(batch, device=dev)
    loss_b = torch.full((1,), float("nan"), device=dev); dz_b = torch.full((batch,), float("nan"), device=dev)
    ctr = torch.zeros(1, dtype=torch.int64, device=dev)
    _lib.check(lib.b200ds_bce_loss_grad(z.data_ptr(), y.data_ptr(), batch, loss_a.data_ptr(), dz_a.data_ptr(), None,
                                        sc.data_ptr(), sc.numel(), _st()), "bce_loss_grad")
    _lib.check(lib.b200ds_bce_loss_grad_noalloc(z.data_ptr(), y.data_ptr(), batch, loss_b.data_ptr(), dz_b.data_ptr(),
                                                ctr.data_ptr(), _st()), "bce_loss_grad_noalloc")
    torch.cuda.synchronize()
    return float(loss_a), float(loss_b), dz_a, dz_b, int(ctr)


@pytest.mark.parametrize("batches", [(4, 3), (1000, 65536 + 3), (4194304, 2049)])
def test_noalloc_matches_caller_scratch_entry_point_twice(lib, cuda_dev, batches):
    for call, batch in enumerate(batches):          # two calls in a row, second one sees the first one's leftovers
        la, lb, dza, dzb, ctr = _both(lib, cuda_dev, batch, seed=batch + call)
        assert abs(la - lb) <= RTOL * abs(la), (batch, call, la, lb)
        assert float((dza - dzb).abs().max()) <= RTOL * float(dza.abs().max()), (batch, call)
        assert ctr == 1                              # the optional step counter ticks once per call


def test_noalloc_rejects_what_it_cannot_serve(lib, cuda_dev):
    z = torch.zeros(8, device=cuda_dev)
    out = torch.zeros(1, device=cuda_dev)
    assert lib.b200ds_bce_loss_grad_noalloc(z.data_ptr(), z.data_ptr(), 4194304 + 1, out.data_ptr(), None, None, _st()) == 1
    assert b"exceeds the built-in scratch limit" in lib.b200ds_last_error()
    assert lib.b200ds_bce_loss_grad_noalloc(z.data_ptr(), z.data_ptr(), 0, out.data_ptr(), None, None, _st()) == 1
    assert lib.b200ds_bce_loss_grad_noalloc(None, z.data_ptr(), 8, out.data_ptr(), None, None, _st()) == 1
