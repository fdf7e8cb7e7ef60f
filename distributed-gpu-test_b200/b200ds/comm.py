"""Process-group bootstrap: the drop-in for `deepspeed.init_distributed()` (train.py:59).

Reads the launcher's environment exactly as the reference's stack does -- torchrun /
torch.distributed.launch set RANK, LOCAL_RANK, WORLD_SIZE, MASTER_ADDR, MASTER_PORT
(slurm_train.sbatch:18-23) -- and, as an addition, maps bare SLURM task variables
(SLURM_PROCID / SLURM_LOCALID / SLURM_NTASKS, slurm_train.sbatch:36) when the torchrun
ones are absent.  torch.distributed is initialised because train.py itself asks it for
the world size (train.py:68) and calls barrier/destroy at exit (train.py:134,138).
"""
from __future__ import annotations

import os
import re
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class LaunchEnv:
    rank: int
    local_rank: int
    world_size: int
    master_addr: str
    master_port: int
    source: str            # "torchrun" | "slurm" | "single"


def _first_host(nodelist: str) -> str:
    """First host of a SLURM nodelist such as 'gpu[03-06,09],login1' -> 'gpu03'."""
    m = re.match(r"^([^\[,]+)(?:\[([^\]]+)\])?", nodelist.strip())
    if not m:
        return "127.0.0.1"
    prefix, inner = m.group(1), m.group(2)
    if not inner:
        return prefix
    first = inner.split(",")[0].split("-")[0]
    return prefix + first


def read_launch_env(env=None) -> LaunchEnv:
    env = os.environ if env is None else env
    if "RANK" in env and "WORLD_SIZE" in env:
        return LaunchEnv(int(env["RANK"]), int(env.get("LOCAL_RANK", 0)), int(env["WORLD_SIZE"]),
                         env.get("MASTER_ADDR", "127.0.0.1"), int(env.get("MASTER_PORT", 29500)), "torchrun")
    if "SLURM_PROCID" in env and "SLURM_NTASKS" in env:
        addr = env.get("MASTER_ADDR") or _first_host(env.get("SLURM_JOB_NODELIST", env.get("SLURM_NODELIST", "127.0.0.1")))
        return LaunchEnv(int(env["SLURM_PROCID"]), int(env.get("SLURM_LOCALID", 0)), int(env["SLURM_NTASKS"]),
                         addr, int(env.get("MASTER_PORT", 12345)), "slurm")   # 12345: slurm_train.sbatch:15
    return LaunchEnv(0, int(env.get("LOCAL_RANK", 0)), int(env.get("WORLD_SIZE", 1)), "127.0.0.1",
                     int(env.get("MASTER_PORT", 29500)), "single")


_STATE = {"env": None}


def init_distributed(dist_backend: str | None = None, **_ignored) -> LaunchEnv:
    """Idempotent.  Backend: nccl when CUDA is present (as DeepSpeed picks on CUDA), gloo
    otherwise (CPU-only host logic tests)."""
    if _STATE["env"] is not None and (dist.is_initialized() or _STATE["env"].world_size == 1):
        return _STATE["env"]
    le = read_launch_env()
    if le.world_size > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", le.master_addr)
        os.environ.setdefault("MASTER_PORT", str(le.master_port))
        os.environ.setdefault("RANK", str(le.rank))
        os.environ.setdefault("WORLD_SIZE", str(le.world_size))
        os.environ.setdefault("LOCAL_RANK", str(le.local_rank))
        backend = dist_backend or ("nccl" if torch.cuda.is_available() else "gloo")
        kwargs = {}
        if backend == "nccl":
            torch.cuda.set_device(le.local_rank)
            kwargs["device_id"] = torch.device("cuda", le.local_rank)
        dist.init_process_group(backend=backend, rank=le.rank, world_size=le.world_size, **kwargs)
    _STATE["env"] = le
    return le


def launch_env() -> LaunchEnv:
    return _STATE["env"] if _STATE["env"] is not None else read_launch_env()


def exchange_nccl_id(make_id, le: LaunchEnv) -> bytes:
    """Rank 0 creates the 128-byte NCCL unique id; everyone reads it from the torch store."""
    if le.world_size == 1:
        return b""
    store = dist.distributed_c10d._get_default_store()
    key = "b200ds/nccl_id/%d" % _STATE.setdefault("id_seq", 0)
    _STATE["id_seq"] += 1
    if le.rank == 0:
        store.set(key, make_id())
    return bytes(store.get(key))
