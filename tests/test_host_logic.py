"""Host-side logic of the drop-in: launcher environment, config handling, CLI."""
import os
import subprocess
import sys

import pytest
import torch

from b200ds import comm
from b200ds import engine as eng

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_env_torchrun():
    le = comm.read_launch_env({"RANK": "5", "LOCAL_RANK": "1", "WORLD_SIZE": "8", "MASTER_ADDR": "10.0.0.1", "MASTER_PORT": "12345"})
    assert (le.rank, le.local_rank, le.world_size, le.master_addr, le.master_port, le.source) == (5, 1, 8, "10.0.0.1", 12345, "torchrun")


def test_env_slurm_fallback():
    le = comm.read_launch_env({"SLURM_PROCID": "3", "SLURM_LOCALID": "3", "SLURM_NTASKS": "16", "SLURM_JOB_NODELIST": "gpu[03-06,09]"})
    assert (le.rank, le.local_rank, le.world_size, le.master_addr, le.master_port, le.source) == (3, 3, 16, "gpu03", 12345, "slurm")
    assert comm._first_host("nodeA,nodeB") == "nodeA"
    assert comm._first_host("n[7,9-12]") == "n7"


def test_env_single_defaults():
    le = comm.read_launch_env({})
    assert (le.rank, le.local_rank, le.world_size, le.source) == (0, 0, 1, "single")


def test_config_parsing():
    ok = eng._parse_config({"train_batch_size": 64, "gradient_accumulation_steps": 1, "fp16": {"enabled": False},
                            "zero_optimization": {"stage": 0}})           # train.py:78-83
    assert ok["precision"] == "fp32" and ok["stage"] == 0
    assert eng._parse_config({"bf16": {"enabled": True}})["precision"] == "bf16"
    assert eng._parse_config({"zero_optimization": {"stage": 1}})["stage"] == 1      # ZeRO-1-style extension (config 5)
    for bad in ({"gradient_accumulation_steps": 2}, {"fp16": {"enabled": True}}, {"zero_optimization": {"stage": 3}}):
        with pytest.raises(NotImplementedError):
            eng._parse_config(bad)


def test_model_family_detection():
    net = torch.nn.Sequential(torch.nn.Linear(20, 64), torch.nn.ReLU(), torch.nn.Linear(64, 1))

    class Wrap(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.net = net

    l1, l2 = eng._find_mlp(Wrap())
    assert (l1.in_features, l1.out_features, l2.out_features) == (20, 64, 1)
    with pytest.raises(NotImplementedError):
        eng._find_mlp(torch.nn.Sequential(torch.nn.Linear(20, 64), torch.nn.Tanh(), torch.nn.Linear(64, 1)))
    with pytest.raises(NotImplementedError):
        eng._find_mlp(torch.nn.Linear(3, 1))


def test_cli_matches_reference_defaults(golden):
    sys.path.insert(0, ROOT)
    import importlib
    train = importlib.import_module("train")
    a = train.cli([])
    assert [a.train_batch_size, a.epochs, a.seed] == list(golden["cli_defaults"])
    assert a.lr == float(golden["cli_lr"][0]) and a.save_dir == "ds_checkpoints"
    # launcher-injected flags are swallowed (slurm_train.sbatch:27-28, torch.distributed.launch)
    b = train.cli(["--distributed-backend", "nccl", "--deepspeed", "--local-rank=3", "--epochs", "2"])
    assert b.epochs == 2


def test_train_py_fails_loudly_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "train.py"), "--epochs", "1"], capture_output=True, text=True,
                       env={k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")})
    assert r.returncode != 0 and "Epoch" not in r.stdout


def test_bench_reference_arm_lines(tmp_path):
    """`bench.py --impl reference`: one JSON line with the contract's keys; with --gpus 2 rank 0 spawns two gloo
    processes and the other torchrun ranks print nothing."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    for gpus in (1, 2):
        r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", str(gpus), "--steps", "3",
                            "--warmup", "1", "--workload", "tiny"], capture_output=True, text=True, env=env, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        line = json.loads(r.stdout.strip().splitlines()[-1])
        assert line["impl"] == "reference" and line["n_gpus"] == gpus and line["unit"] == "samples/s" and line["value"] > 0
        assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0
        assert line["config"]["parallelism"] == f"dp{gpus}" and line["steps"] == 3
    env2 = dict(env, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "2", "--warmup", "1",
                        "--workload", "tiny"], capture_output=True, text=True, env=env2, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_bench_reference_arm_under_torchrun():
    """The driver launches `bench.py --impl reference --gpus N` under torchrun for N > 1.  The launcher's environment
    (TORCHELASTIC_USE_AGENT_STORE above all: it turns a tcp:// rendezvous into a client of the agent's store) must not
    reach the gloo workers rank 0 spawns -- the run once hung exactly there.  One line, rc 0, in bounded time."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29541", os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "3",
                        "--warmup", "1", "--workload", "tiny"], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0 and "2 process(es)" in line["config"]["arm"]


def test_bench_traffic_lookup_follows_the_kernel_source(tmp_path, monkeypatch):
    """bench.py's roofline.traffic comes from profiles/ncu_traffic.json and only while the profiled kernel's source file
    still has the hash it was profiled with; a workload that was never profiled, another precision, or an edited source
    gives None (never a stale literal)."""
    import hashlib, json, os, sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    src = tmp_path / "csrc" / "k.cu"
    src.parent.mkdir()
    src.write_text("__global__ void k() {}\n")
    (tmp_path / "profiles").mkdir()
    ent = {"kernel": "k", "traffic_bytes": 1234, "dtype": "bf16", "source": "csrc/k.cu",
           "source_sha16": hashlib.sha256(src.read_bytes()).hexdigest()[:16], "report": "profiles/x.txt"}
    (tmp_path / "profiles" / "ncu_traffic.json").write_text(json.dumps({"wide4096": ent}))
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))
    assert bench.profiled_traffic("wide4096", "bf16") == 1234
    assert bench.profiled_traffic("wide4096", "fp32") is None
    assert bench.profiled_traffic("wide8192", "bf16") is None
    src.write_text("__global__ void k() { /* edited */ }\n")
    assert bench.profiled_traffic("wide4096", "bf16") is None
    # the committed table is current for the committed sources
    monkeypatch.undo()
    committed = json.load(open(os.path.join(bench.ROOT, "profiles", "ncu_traffic.json")))
    for wl, e in committed.items():
        assert bench.profiled_traffic(wl, "bf16") == e["traffic_bytes"], wl
