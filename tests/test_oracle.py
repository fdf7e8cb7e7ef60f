"""Pins the CPU oracle (oracle/) against (a) the golden vectors produced by running the
reference's own code (tests/golden/make_golden.py), (b) torch's own autograd / loss /
Adam / samplers -- the third-party code the reference's arithmetic lives in."""
import hashlib
import math

import numpy as np
import pytest
import torch

from oracle import philox as ph
from oracle import reference_step as ora

torch.set_num_threads(1)


def _sha16(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def _same_cpu_path(golden):
    return str(golden["cpu_capability"][0]) == torch.backends.cpu.get_cpu_capability()


def test_golden_known_answers(golden):
    # SURVEY 8c known-answer vectors (seed 42, AVX2 Box-Muller path)
    assert float(golden["x"][0, 0]) == 1.9269152879714966
    assert float(golden["x"][-1, -1]) == 1.2584049701690674
    assert float(golden["y"].sum()) == 1031.0
    assert str(golden["sha_x"][0]) == "50469934153700f7" == _sha16(golden["x"])
    assert str(golden["sha_y"][0]) == "e506d387ed85bb9b" == _sha16(golden["y"])
    assert list(golden["cli_defaults"]) == [64, 5, 42] and float(golden["cli_lr"][0]) == 1e-3


def test_dataset_and_init_bit_exact(golden):
    if not _same_cpu_path(golden):
        pytest.skip("host ATen CPU capability differs from the golden host (SURVEY H1): randn bits differ by design")
    torch.manual_seed(42)
    x, y = ora.make_dataset(2000, 20, 42)
    p = ora.init_params(20, 64)
    assert np.array_equal(x.numpy(), golden["x"])
    assert np.array_equal(y.numpy(), golden["y"])
    for k in ora.PARAM_ORDER:
        assert np.array_equal(p[k].numpy(), golden[f"init_{k}"]), k


@pytest.mark.parametrize("world", [1, 2, 8])
def test_training_run_matches_reference(golden, world):
    if not _same_cpu_path(golden):
        pytest.skip("needs the golden host's CPU RNG path")
    r = ora.run_training(world=world)
    sl = np.array(r.step_losses)
    gl = golden[f"w{world}_step_losses"]
    assert sl.shape == gl.shape
    assert np.abs(sl - gl).max() <= 1e-6
    assert [f"{v:.4f}" for v in r.epoch_means[0]] == [f"{v:.4f}" for v in golden[f"w{world}_epoch_means"][0]]
    for k in ora.PARAM_ORDER:
        assert np.abs(r.params[k].numpy() - golden[f"w{world}_final_{k}"]).max() <= 1e-6, k
    assert r.log_lines[-1] == "Training completed."
    assert r.log_lines[0].startswith("Epoch  1 finished. Avg loss: ")


def test_ragged_last_batch_world8():
    # 2000/8 = 250 per rank = 62 batches of 4 + one of 2 (train.py:74, drop_last=False)
    idx = ora.rank_indices(2000, 8, 3, 1)
    bs = list(ora.batches(idx, 4))
    assert len(idx) == 250 and len(bs) == 63 and len(bs[-1]) == 2


@pytest.mark.parametrize("world,rank", [(2, 0), (2, 1), (8, 5), (3, 2)])
def test_rank_indices_match_distributed_sampler(world, rank):
    from torch.utils.data import TensorDataset
    from torch.utils.data.distributed import DistributedSampler
    ds = TensorDataset(torch.zeros(2000, 1))
    s = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=True)
    for epoch in (1, 2, 5):
        s.set_epoch(epoch)
        assert list(iter(s)) == ora.rank_indices(2000, world, rank, epoch)


def test_world1_order_matches_dataloader():
    from torch.utils.data import DataLoader, TensorDataset
    ds = TensorDataset(torch.arange(2000))
    torch.manual_seed(7)
    got = [int(v) for (b,) in DataLoader(ds, batch_size=4, shuffle=True) for v in b]
    torch.manual_seed(7)
    assert got == ora.rank_indices(2000, 1, 0, 1)


@pytest.mark.parametrize("B,F,H", [(4, 20, 64), (2, 20, 64), (1, 3, 5), (33, 20, 128), (7, 32, 256)])
def test_written_out_math_matches_autograd(B, F, H):
    torch.manual_seed(B * 1000 + H)
    p = ora.init_params(F, H)
    x = torch.randn(B, F)
    y = (torch.rand(B, 1) > 0.5).float()
    z, h = ora.forward(p, x)
    loss = ora.bce_with_logits(z, y)
    g = ora.backward(p, x, y, z, h)
    q = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    zz = torch.nn.functional.linear(torch.relu(torch.nn.functional.linear(x, q["w1"], q["b1"])), q["w2"], q["b2"])
    ll = torch.nn.BCEWithLogitsLoss()(zz, y)
    ll.backward()
    assert torch.allclose(z, zz.detach(), rtol=1e-6, atol=1e-7)
    assert abs(float(loss) - float(ll.detach())) <= 1e-6 * max(1.0, abs(float(ll.detach())))
    for k in ora.PARAM_ORDER:
        assert torch.allclose(g[k], q[k].grad, rtol=1e-5, atol=1e-7), k


def test_bce_extreme_logits():
    z = torch.tensor([[-80.0], [-20.0], [0.0], [20.0], [80.0]])
    for yv in (0.0, 1.0):
        y = torch.full_like(z, yv)
        a = float(ora.bce_with_logits(z, y))
        b = float(torch.nn.BCEWithLogitsLoss()(z, y))
        assert math.isfinite(a) and abs(a - b) <= 1e-6 * max(1.0, abs(b))


def test_adam_matches_torch_optim():
    torch.manual_seed(0)
    p0 = torch.randn(1409)
    ref = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=1e-3)
    mine = p0.clone()
    st = ora.AdamState()
    for t in range(25):
        g = torch.randn(1409) * (10.0 ** (t % 5 - 3))
        ref.grad = g.clone()
        opt.step()
        ora.adam_update(mine, g.clone(), st, 1e-3)
        assert torch.equal(mine, ref.detach()), t


def test_average_gradients_sum_then_scale():
    gs = [torch.full((5,), float(i + 1)) for i in range(4)]
    assert torch.equal(ora.average_gradients(gs), torch.full((5,), 2.5))
    assert torch.equal(ora.average_gradients(gs[:1]), gs[0])


def test_bf16_restatement_within_tolerance():
    r32 = ora.run_training(world=1, epochs=1, max_steps=60)
    r16 = ora.run_training(world=1, epochs=1, max_steps=60, dtype=torch.bfloat16)
    a, b = np.array(r32.step_losses[0]), np.array(r16.step_losses[0])
    assert np.abs(a - b).max() <= 2e-2 * np.abs(a).max()       # north_star bf16 tolerance


# ------------------------------------------------------------------ Philox
def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32 10 rounds
    def h(c, k):
        return [int(v) for v in ph.philox4x32_10(np.array([c], dtype=np.uint32), k)[0]]
    assert h([0, 0, 0, 0], (0, 0)) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    assert h([0xFFFFFFFF] * 4, (0xFFFFFFFF, 0xFFFFFFFF)) == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    assert h([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], (0xA4093822, 0x299F31D0)) == \
        [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]


def test_philox_torch_cuda_known_values():
    # torch.manual_seed(42); torch.randn(6, device='cuda') -- the widely quoted first values
    got = ph.randn(6, 42, 0, 148)
    want = np.array([0.1940, 2.1614, -0.1721, 0.8491, -1.9244, 0.6530], dtype=np.float32)
    assert np.abs(got - want).max() < 5e-5


def test_philox_launch_plan():
    # 148 SMs x 8 CTAs x 256 threads; 65536x20 elements -> two Philox calls per thread
    grid, nthreads, calls, adv = ph.launch_plan(65536 * 20, 148)
    assert (grid, nthreads, calls, adv) == (1184, 303104, 2, 8)
    assert ph.launch_plan(80, 148) == (1, 256, 1, 4)
    assert ph.launch_plan(1, 148) == (1, 256, 1, 4)


def test_philox_offset_continuation():
    # element stream at offset 4 == what the second call of every thread would have produced
    n, sm = 256 * 8, 1
    _, nthreads, calls, adv = ph.launch_plan(n, sm)
    assert nthreads == 2048 and calls == 1 and adv == 4
    a = ph.randn_u32(2 * 4 * nthreads, 7, 0, sm)     # two calls per thread
    b = ph.randn_u32(4 * nthreads, 7, 4, sm)
    assert np.array_equal(a[4 * nthreads:], b)


def test_label_rule_matches_reference_rule():
    x = torch.randn(500, 20)
    y_ref = ora.label_rule(x).reshape(-1).numpy()
    y = ph.labels(x.numpy())
    margin = np.abs(x[:, :10].sum(dim=1).numpy()) > 1e-5
    assert np.array_equal(y[margin], y_ref[margin])


def test_philox_negative_zero_edge_case():
    """Found on the GPU against torch.randn: when the first Philox word is >= 0xFFFFFF80 the uniform
    rounds to exactly 1, Box-Muller gives sqrt(-2 log 1) = -0.0, and ATen's final `* std + mean`
    makes it +0.0.  Seed 7, third draw of 65536*20 values (offset 16), elements 50328 / 353432."""
    w = ph.randn_u32(65536 * 20, 7, 16, 148)
    assert int(w[50328, 0]) == 0xFFFFFFF2 and int(w[353432, 0]) == 0xFFFFFFF2
    v = ph.box_muller(w[[50328, 353432], 0], w[[50328, 353432], 1], w[[50328, 353432], 2])
    assert v.view(np.uint32).tolist() == [0, 0]          # +0.0, not 0x80000000
