#!/usr/bin/env python3
"""train.py itself on W GPUs (torchrun), rank-0 stdout against the golden epoch means of the reference's own objects
(tests/golden/train_golden.npz, w{W}_epoch_means: what /root/reference/train.py's data / model / sampler code gives
with stage-0 gradient averaging, see tests/golden/make_golden.py).

    python tools/run_train_ranks.py [--ranks 2 4 8] [--out gpurun_out]

Per world size three runs: the reference configuration (fp32, stage 0): log lines equal to 4 decimals; `--bf16`
(BASELINE config 3: tensor-core kernels on the padded model): epoch means within 2e-2; `--zero-stage 1`
(config 5): equal to 4 decimals again (same arithmetic, sharded optimizer).  Writes the logs and prints one
verdict line per run; exit code 0 only if every run passes."""
import argparse
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, nargs="+", default=[2])
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out"))
    ap.add_argument("--modes", nargs="+", default=["fp32", "bf16", "zero1"])
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    gold = np.load(os.path.join(ROOT, "tests", "golden", "train_golden.npz"))
    import torch
    if str(gold["cpu_capability"][0]) != torch.backends.cpu.get_cpu_capability():
        print("SKIP: this host's CPU RNG path (%s) differs from the golden host's (%s): the dataset would differ"
              % (torch.backends.cpu.get_cpu_capability(), gold["cpu_capability"][0]))
        return 0
    ok_all = True
    port = 29600
    for W in a.ranks:
        want = [float(v) for v in gold[f"w{W}_epoch_means"][0]]
        for mode in a.modes:
            extra = {"fp32": [], "bf16": ["--bf16"], "zero1": ["--zero-stage", "1"]}[mode]
            port += 1
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={W}", "--master-addr", "127.0.0.1",
                   "--master-port", str(port), os.path.join(ROOT, "train.py"), "--no-checkpoint", "--distributed-backend", "nccl",
                   "--deepspeed"] + extra
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
            log = os.path.join(a.out, f"train_w{W}_{mode}.log")
            with open(log, "w") as f:
                f.write("$ " + " ".join(cmd[1:]) + "\n" + r.stdout + ("\n[stderr tail]\n" + r.stderr[-1500:] if r.returncode else ""))
            lines = [ln for ln in r.stdout.splitlines() if ln.startswith("Epoch")]
            got = [float(ln.rsplit(":", 1)[1]) for ln in lines]
            done = "Training completed." in r.stdout
            if mode == "bf16":
                good = len(got) == len(want) and all(abs(g - w) <= 2e-2 * abs(w) + 5e-5 for g, w in zip(got, want))
                crit = "within 2e-2 of the golden fp32 means"
            else:
                good = lines == [f"Epoch {e + 1:2d} finished. Avg loss: {v:.4f}" for e, v in enumerate(want)]
                crit = "log lines equal to 4 decimals"
            good = good and done and r.returncode == 0
            ok_all &= good
            print(f"train.py W={W} {mode:5s}: {'OK ' if good else 'FAIL'} ({crit})  got {' '.join(f'{g:.4f}' for g in got)}  "
                  f"golden {' '.join(f'{w:.4f}' for w in want)}  rc={r.returncode}  -> {os.path.relpath(log, ROOT)}", flush=True)
    return 0 if ok_all else 1


if __name__ == "__main__":
    sys.exit(main())
