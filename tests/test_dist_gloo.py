"""world_size-2 host-side checks on CPU with the gloo backend: launcher env -> process
group, NCCL-id exchange through the torch store, rank sharding of the epoch order, and
the sum-then-scale gradient averaging the bucket all-reduce implements."""
import os
import socket
import sys

import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
    os.environ.update({"RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(world),
                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
    import torch.distributed as dist
    from torch.utils.data import TensorDataset
    from torch.utils.data.distributed import DistributedSampler

    from b200ds import comm
    from oracle import reference_step as ora

    torch.set_num_threads(1)
    le = comm.init_distributed(dist_backend="gloo")
    assert (le.rank, le.world_size) == (rank, world) and dist.get_world_size() == world
    assert comm.init_distributed() is le                                    # idempotent
    nid = comm.exchange_nccl_id(lambda: bytes(range(128)), le)
    assert nid == bytes(range(128))
    nid2 = comm.exchange_nccl_id(lambda: bytes([7] * 128), le)              # a second engine gets a fresh key
    assert nid2 == bytes([7] * 128)

    # epoch order: the shim uses torch's DistributedSampler; the oracle restates it
    torch.manual_seed(42)
    x, y = ora.make_dataset(2000, 20, 42)
    p = ora.init_params(20, 64)
    s = DistributedSampler(TensorDataset(x, y), shuffle=True)
    s.set_epoch(1)
    mine = list(iter(s))
    assert mine == ora.rank_indices(2000, world, rank, 1)

    # one data-parallel step: local grads -> all_reduce(SUM) -> x 1/W -> Adam
    flat = ora.flatten(p)
    st = ora.AdamState()
    bi = torch.tensor(mine[:4])
    z, h = ora.forward(p, x[bi])
    g = ora.flatten(ora.backward(p, x[bi], y[bi], z, h))
    bucket = g.clone()
    dist.all_reduce(bucket, op=dist.ReduceOp.SUM)
    bucket *= 1.0 / world
    ora.adam_update(flat, bucket, st, 1e-3)
    gathered = [torch.zeros_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    assert all(torch.equal(gathered[0], t) for t in gathered)              # replicas stay identical
    if rank == 0:
        torch.save({"flat": flat, "loss": float(ora.bce_with_logits(z, y[bi]))}, out)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_step_matches_single_process_emulation(tmp_path):
    from oracle import reference_step as ora
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    ref = ora.run_training(world=2, epochs=1, max_steps=1)
    assert abs(got["loss"] - ref.step_losses[0][0]) <= 1e-7
    assert torch.allclose(got["flat"], ora.flatten(ref.params), rtol=0, atol=1e-7)
