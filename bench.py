#!/usr/bin/env python3
"""bench.py -- train samples/s of the train.py step hot path on N B200s (one process per
GPU), beside the reference's CPU path on the same box.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload wide4096|wide8192|wideFH4096|wideFH8192|wideFH1024|tiny]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's CPU implementation, same metric/config

One "step" = train.py:104-117 for one per-rank batch: on-device Philox batch generation
+ label rule, forward, BCE loss, backward, gradient all-reduce (N > 1), Adam.
Default workload = BASELINE.json configs[3] (widened model: F=20, H=4096, 65536
samples/GPU), the largest single-GPU configuration and the one the metric's roofline
fraction is defined on (the B=4 reference shape is launch-latency-bound, its
numbers ride along under "tiny").  Prints ONE JSON line on rank 0.

Environment switches, for tuning and A/B runs only (the driver's runs set none of them):
B200DS_BENCH_NCCL=1 (W > 1: ncclAllReduce + Adam kernel instead of the peer-memory update), B200DS_BENCH_SPLIT_UPDATE=1
(finish / exchange + Adam / pack as separate kernels), B200DS_BENCH_SEPARATE_LOSS=1 (loss kernel, then backward),
B200DS_BENCH_SETTLE=n (clock warm-up replays; profiling runs shorten it), B200DS_BENCH_NOFLUSH=1 (no L2 flush between
timed steps: invalid as a bench number), B200DS_BENCH_SKIP_E2E=1, B200DS_BENCH_STEP_TIMES=1 (per-step device times in the
line), B200DS_BENCH_TRACE=1 (progress on stderr).
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))

import torch  # noqa: E402

WORKLOADS = {
    "wide4096": dict(B=65536, F=20, H=4096),
    "wide8192": dict(B=65536, F=20, H=8192),
    "tiny": dict(B=4, F=20, H=64),
    # wide inputs ("F = H", SURVEY 8d-4: the extension where layer 1 is a real GEMM); a smaller one for quick runs
    "wideFH4096": dict(B=65536, F=4096, H=4096),
    "wideFH1024": dict(B=65536, F=1024, H=1024),
    "wideFH8192": dict(B=65536, F=8192, H=8192),      # 67 M parameters, 17.6 TFLOP per step, 268 MB gradient bucket
}
SEED = 42
LR = 1e-3


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="wide4096", choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default="auto", choices=["auto", "fp32", "bf16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-tiny", action="store_true")
    ap.add_argument("--ref-rank", type=int, default=-1)        # internal: a gloo worker of the multi-process CPU reference
    ap.add_argument("--ref-port", type=int, default=29611)
    ap.add_argument("--ref-threads", type=int, default=0)
    return ap.parse_known_args()[0]


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sus=d.get("bf16_tflops_sustained", d["bf16_tflops"]), src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sus=1400.0, src="fallback")   # B200_PROFILING.md fallback


def profiled_traffic(workload, precision):
    """DRAM bytes per launch of the dominant kernel (dram__bytes_read.sum + dram__bytes_write.sum) from the committed
    `ncu --set full` capture, profiles/ncu_traffic.json -- only for a workload that was profiled, and only while the
    kernel's source file is still the one that was profiled (sha-256 of the file); otherwise None."""
    import hashlib
    try:
        ent = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(workload)
        if not ent or ent.get("dtype") != ("bf16" if precision == "bf16" else "f32"):
            return None
        sha = hashlib.sha256(open(os.path.join(ROOT, ent["source"]), "rb").read()).hexdigest()[:16]
        return ent["traffic_bytes"] if sha == ent["source_sha16"] else None
    except Exception:       # noqa: BLE001
        return None


def trace(msg):
    if os.environ.get("B200DS_BENCH_TRACE"):
        sys.stderr.write(f"[bench rank {os.environ.get('RANK', 0)}] {msg}\n")
        sys.stderr.flush()


def init_flat(F, H):
    """Initial weights in the reference's creation order (train.py:54,76): seed, then the two
    nn.Linear default inits; flat layout [W1|b1|W2|b2]."""
    torch.manual_seed(SEED)
    l1, l2 = torch.nn.Linear(F, H), torch.nn.Linear(H, 1)
    return torch.cat([t.detach().reshape(-1) for t in (l1.weight, l1.bias, l2.weight, l2.bias)])


def label_rule(x):
    f = x.shape[1]
    return (x[:, : f // 2].sum(dim=1) > 0).float()


def step_flops(B, F, H):
    return B * (4 * F * H + 6 * H)           # algorithmic FLOPs per sample = 4FH + 6H


# ------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:      # noqa: BLE001
            self.nv = None
            self.err = repr(e)

    def sample(self):
        if self.nv is None:
            return
        nv = self.nv
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            bits = fn(self.h)
            names = {
                "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
            }
            for k, b in names.items():
                if bits & b:
                    self.reasons.add(k)
        except Exception:           # noqa: BLE001
            pass

    def start(self):
        def run():
            while not self._stop.wait(0.01):
                self.sample()
        self._thr = threading.Thread(target=run, daemon=True)
        self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------ reference arm (CPU)
def workload_config(name, wl, world):
    """The part of `config` both arms print identically (the driver compares the two lines)."""
    return {"workload": f"{name}: train.py step (train.py:104-117), F={wl['F']} H={wl['H']}, {wl['B']} samples per rank per step "
                        f"drawn from N(0,1) every step with the label rule of train.py:22-23, Adam, {world} rank(s)",
            "per_gpu_batch": wl["B"], "n_features": wl["F"], "hidden": wl["H"], "parallelism": f"dp{world}"}


def cpu_reference_run(wl, steps, warmup, budget_s=None, threads=None, rank=0, world=1):
    """The reference's own CPU path for this step (torch nn.Sequential + BCEWithLogitsLoss
    + autograd + optim.Adam, as train.py:26-36,85,96,107-114 use them) via the oracle's torch-native
    restatement, on FULL per-rank batches (wl["B"] samples per step, as the GPU arm).  world > 1: this is one of
    `world` gloo processes; gradients are summed over the ranks, then scaled by 1/world (DeepSpeed stage 0,
    train.py:113), through one flat bucket."""
    from oracle import reference_step as ora
    cores = len(os.sched_getaffinity(0))
    threads = threads or cores
    torch.set_num_threads(threads)
    B, F, H = wl["B"], wl["F"], wl["H"]
    torch.manual_seed(SEED)
    p = ora.init_params(F, H)
    state = {}
    torch.manual_seed(SEED + 1000003 * rank)        # a different batch stream per rank, as on the GPU arm
    avg = None
    if world > 1:
        import torch.distributed as dist

        def avg(grads):
            flat = torch.cat([g.reshape(-1) for g in grads])
            dist.all_reduce(flat, op=dist.ReduceOp.SUM)
            flat.mul_(1.0 / world)
            o = 0
            for g in grads:
                g.copy_(flat[o:o + g.numel()].view_as(g)); o += g.numel()

    def one():
        x = torch.randn(B, F)
        y = ora.label_rule(x)
        loss, _ = ora.torch_native_step(p, x, y, state, LR, average_grads=avg)
        return float(loss)

    for _ in range(max(1, warmup)):
        one()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        one()
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    return dict(value=world * B * done / dt, unit="samples/s", cores=threads * world, kind="port",
                sample=f"{done} full steps of {B} samples per rank (F={F}, H={H}), {world} process(es) x {threads} thread(s) of "
                       f"{cores} host cores{', gloo all-reduce of the flat gradient bucket' if world > 1 else ''}, torch "
                       f"{torch.__version__} CPU fp32, batch drawn with torch.randn on the host each step",
                ms_per_step=1e3 * dt / done, steps=done, Bs=B, threads=threads)


def reference_worker(a):
    """One gloo rank of the multi-process CPU reference (spawned by rank 0 of `--impl reference` when --gpus > 1)."""
    import torch.distributed as dist
    rank, world = a.ref_rank, a.gpus
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{a.ref_port}", rank=rank, world_size=world)
    r = cpu_reference_run(WORKLOADS[a.workload], a.steps, a.warmup, threads=a.ref_threads, rank=rank, world=world)
    t = torch.tensor([r["ms_per_step"]], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        r["ms_per_step"] = float(t)
        r["value"] = world * r["Bs"] * 1e3 / float(t)
        print("REFWORKER " + json.dumps(r), flush=True)
    dist.destroy_process_group()


# ------------------------------------------------------------------------------ our arm
class WideStep:
    """Device-resident step for the widened workloads, every stage a libb200ds.so kernel."""

    def __init__(self, lib, ctx, dev, wl, world, rank, precision):
        from b200ds import _lib
        self.lib, self.ctx, self._lib = lib, ctx, _lib
        B, F, H = wl["B"], wl["F"], wl["H"]
        self.B, self.F, self.H, self.world = B, F, H, world
        self.P = F * H + 2 * H + 1
        self.Ppad = (self.P + 3) // 4 * 4
        self.precision = precision
        flat = init_flat(F, H)
        z = lambda n: torch.zeros(n, dtype=torch.float32, device=dev)      # noqa: E731
        self.params = z(self.Ppad); self.params[: self.P] = flat.to(dev)
        self.grads, self.m, self.v = z(self.Ppad), z(self.Ppad), z(self.Ppad)
        self.x, self.y = z(B * F), z(B)
        self.logits, self.dz, self.loss = z(B), z(B), z(1)
        self.bce_sc = torch.zeros(lib.b200ds_bce_scratch_bytes(B), dtype=torch.uint8, device=dev)
        self.bwd_sc = torch.zeros(max(256, lib.b200ds_mlp_backward_f32_scratch_bytes(B, F, H)), dtype=torch.uint8, device=dev)
        self.counter = torch.zeros(1, dtype=torch.int64, device=dev)
        self.seed = SEED + 1000003 * rank           # distinct Philox key per rank
        self.offset = 0
        self.tc = precision == "bf16"
        self.primed = False
        if self.tc:
            assert lib.b200ds_tc_supported(B, F, H), "bf16 tensor-core path does not take this shape"
            self.xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=dev)
            self.wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=dev)
            self.bwd_sc = torch.zeros(max(256, lib.b200ds_mlp_backward_tc_scratch_bytes(B, F, H)), dtype=torch.uint8, device=dev)
            self.fwd_sc = torch.zeros(max(256, lib.b200ds_mlp_forward_tc_scratch_bytes(B, F, H)), dtype=torch.uint8, device=dev)
            self.dominant = dict(name="mlp_bwd_tc_kernel<LOSS> (+finish)", bound="tensor")
        # bf16: gen_batch, pack_w1, forward, loss+backward, finish, adam; fp32: philox, label, forward, loss, backward, finish, adam
        self.launches_per_step = 6 if precision == "bf16" else 7          # 4 with the fused update (set below)
        # W > 1: gradient exchange + Adam as one peer-memory kernel (NVLink loads of the peers' buckets,
        # fixed summation order); B200DS_BENCH_NCCL=1 selects ncclAllReduce + the Adam kernel instead
        self.p2p = world > 1 and not os.environ.get("B200DS_BENCH_NCCL")
        # finish pass + gradient exchange + Adam + operand pack as b200ds_update_tc (one launch at one rank, two with
        # several); B200DS_BENCH_SPLIT_UPDATE=1 selects the separate finish / exchange+Adam / pack kernels (A/B arm)
        self.fused_update = self.tc and not os.environ.get("B200DS_BENCH_SPLIT_UPDATE") and (world == 1 or self.p2p)
        if self.p2p:
            # collective; if any rank cannot map its peers (no peer access on this box) every rank takes the NCCL exchange
            import torch.distributed as dist
            rc = lib.b200ds_ctx_enable_p2p(ctx, max(self.Ppad, lib.b200ds_update_tc_exchange_elems(H) if self.tc else 0))
            okf = torch.tensor([1.0 if rc == 0 else 0.0], device=dev)
            dist.all_reduce(okf, op=dist.ReduceOp.MIN)
            if float(okf) < 1.0:
                if rank == 0:
                    print("bench: peer-memory exchange unavailable (%s); using ncclAllReduce + Adam"
                          % lib.b200ds_last_error().decode("utf-8", "replace"), file=sys.stderr, flush=True)
                self.p2p = False
                self.fused_update = False
        if self.tc and self.fused_update:          # the finish pass belongs to the update kernel on this arm: the bracket holds the backward alone
            self.dominant = dict(name="mlp_bwd_tc_kernel<LOSS>", bound="tensor")
        self.side = torch.cuda.Stream(device=dev, priority=-1)     # the exchange: its CTAs go first whenever an SM has room

    @property
    def launches(self):
        """kernels of this library launched per step"""
        if not self.tc:
            return 7                      # philox, label, forward, loss, backward, finish, adam
        if self.fused_update:
            return 4 if self.world == 1 else 6     # gen_batch, forward, loss+backward, update (W > 1: publish, gate, collect)
        return 6                                   # ... finish, adam (or exchange+adam), pack_w1

    def step(self, st, ev_pair=None):
        lib, chk = self.lib, self._lib.check
        B, F, H = self.B, self.F, self.H
        # the draw index is the device step counter (incremented by the loss kernel), so the same
        # launch sequence can be replayed from a CUDA graph and still get a fresh batch each step
        if self.tc:
            if not self.primed:      # batch 0 and the first weight pack; later ones are made at the end of the previous step
                self.gen_next(st)
                self.pack(st)
                self.primed = True
            chk(lib.b200ds_mlp_forward_tc(self.xq.data_ptr(), self.wq.data_ptr(), self.params.data_ptr(), B, F, H,
                                          self.logits.data_ptr(), self.fwd_sc.data_ptr(), self.fwd_sc.numel(), st))
        else:
            chk(lib.b200ds_philox_normal_counted(self.x.data_ptr(), B * F, self.seed, 0, self.counter.data_ptr(), st))
            chk(lib.b200ds_label_rule(self.x.data_ptr(), self.y.data_ptr(), B, F, st))
            chk(lib.b200ds_mlp_forward_f32(self.params.data_ptr(), self.x.data_ptr(), B, F, H, self.logits.data_ptr(), st))
        if ev_pair:
            ev_pair[0].record()
        if self.tc and self.fused_update:
            # criterion + backward in one pass (four warps of the backward kernel form dlogits from the logits and
            # labels); the split partials stay in scratch for the update kernel
            chk(lib.b200ds_mlp_loss_backward_partials_tc(self.xq.data_ptr(), self.wq.data_ptr(), self.logits.data_ptr(),
                                                         self.y.data_ptr(), B, F, H, self.counter.data_ptr(),
                                                         self.bwd_sc.data_ptr(), self.bwd_sc.numel(), st))
        elif self.tc and os.environ.get("B200DS_BENCH_SEPARATE_LOSS"):      # A/B arm: loss kernel, then backward
            chk(lib.b200ds_bce_loss_grad(self.logits.data_ptr(), self.y.data_ptr(), B, self.loss.data_ptr(), self.dz.data_ptr(),
                                         self.counter.data_ptr(), self.bce_sc.data_ptr(), self.bce_sc.numel(), st))
            chk(lib.b200ds_mlp_backward_tc(self.xq.data_ptr(), self.wq.data_ptr(), self.params.data_ptr(), self.dz.data_ptr(),
                                           B, F, H, self.grads.data_ptr(), self.bwd_sc.data_ptr(), self.bwd_sc.numel(), st))
        elif self.tc:
            chk(lib.b200ds_mlp_loss_backward_tc(self.xq.data_ptr(), self.wq.data_ptr(), self.params.data_ptr(),
                                                self.logits.data_ptr(), self.y.data_ptr(), B, F, H, self.loss.data_ptr(),
                                                self.grads.data_ptr(), self.counter.data_ptr(), self.bwd_sc.data_ptr(),
                                                self.bwd_sc.numel(), st))
        else:
            chk(lib.b200ds_bce_loss_grad(self.logits.data_ptr(), self.y.data_ptr(), B, self.loss.data_ptr(), self.dz.data_ptr(),
                                         self.counter.data_ptr(), self.bce_sc.data_ptr(), self.bce_sc.numel(), st))
            chk(lib.b200ds_mlp_backward_f32(self.params.data_ptr(), self.x.data_ptr(), self.dz.data_ptr(), B, F, H,
                                            self.grads.data_ptr(), self.bwd_sc.data_ptr(), self.bwd_sc.numel(), st))
        if ev_pair:
            ev_pair[1].record()
        # ---- exchange + Adam, and (bf16) the next step's inputs: the next batch does not need this step's
        # gradients, so it is generated on the main stream while the exchange, Adam and the new weight pack
        # run on a side stream; both are done before the next step starts
        if self.tc and self.fused_update:
            # finish pass + gradient exchange (W > 1, peer memory) + Adam + operand pack: one kernel.  With several
            # ranks it mostly waits on NVLink round trips, so it runs on a side stream beside the generator.
            def update(fn, stream_handle):
                chk(fn(self.ctx, self.bwd_sc.data_ptr(), self.bwd_sc.numel(), B, F, H, self.params.data_ptr(),
                       self.grads.data_ptr(), self.m.data_ptr(), self.v.data_ptr(), self.wq.data_ptr(),
                       LR, 0.9, 0.999, 1e-8, 0, self.counter.data_ptr(), self.loss.data_ptr(), stream_handle))
            if self.world == 1:        # (the update beside the generator on a second stream was measured: 99 us against 94)
                update(lib.b200ds_update_tc, st)
                self.gen_next(st)
                return
            # several ranks: finish + publication, the wait for the other ranks (one warp), the rank-order sum, Adam and the
            # pack go to the (high-priority) side stream, the next batch is drawn on the main stream meanwhile.  Measured at
            # 2 GPUs: 105.5 us per step; publication on the main stream and only the second half on the side stream 107;
            # the generator on the side stream and the whole update on the main stream 110.5.
            main = torch.cuda.current_stream()
            fork = torch.cuda.Event(); fork.record(main)
            self.side.wait_event(fork)
            update(lib.b200ds_update_tc, self.side.cuda_stream)
            join = torch.cuda.Event(); join.record(self.side)
            self.gen_next(st)
            main.wait_event(join)
            return
        if self.tc and self.p2p:
            # the exchange kernel mostly waits on NVLink round trips (at most one CTA per SM): it overlaps with
            # the generator.  (With one rank the same fork was measured and gains nothing: the generator fills
            # every thread slot, so Adam and the pack only start when it drains.)
            main = torch.cuda.current_stream()
            fork = torch.cuda.Event(); fork.record(main)
            self.side.wait_event(fork)
            sst = self.side.cuda_stream
            chk(lib.b200ds_allreduce_adam_p2p(self.ctx, self.params.data_ptr(), self.grads.data_ptr(), self.m.data_ptr(),
                                              self.v.data_ptr(), self.Ppad, LR, 0.9, 0.999, 1e-8, 0, self.counter.data_ptr(), sst))
            self.pack(sst)
            join = torch.cuda.Event(); join.record(self.side)
            self.gen_next(st)
            main.wait_event(join)
            return
        if self.p2p:        # fp32 kernels
            chk(lib.b200ds_allreduce_adam_p2p(self.ctx, self.params.data_ptr(), self.grads.data_ptr(), self.m.data_ptr(),
                                              self.v.data_ptr(), self.Ppad, LR, 0.9, 0.999, 1e-8, 0, self.counter.data_ptr(), st))
            return
        if self.tc:         # NCCL arm: the all-reduce on the context's side stream beside the generator
            chk(lib.b200ds_allreduce_bucket_begin(self.ctx, self.grads.data_ptr(), self.P, st))
            self.gen_next(st)
            chk(lib.b200ds_allreduce_bucket_end(self.ctx, st))
        else:
            chk(lib.b200ds_allreduce_bucket(self.ctx, self.grads.data_ptr(), self.P, st))
        chk(lib.b200ds_adam_step_counted(self.params.data_ptr(), self.grads.data_ptr(), self.m.data_ptr(), self.v.data_ptr(),
                                         self.Ppad, LR, 0.9, 0.999, 1e-8, self.counter.data_ptr(),
                                         1.0 / self.world, 0, st))
        if self.tc:
            self.pack(st)

    def pack(self, st):
        """bf16 operand pack of the current W1 / b1 (+ the partials of v = w1q^T w2) for the next forward."""
        self._lib.check(self.lib.b200ds_pack_params_bf16(self.params.data_ptr(), self.F, self.H, self.wq.data_ptr(), st))

    def gen_next(self, st):
        """Batch number *counter of the Philox stream (the loss kernel has already counted this
        step), drawn straight into the bf16 pack together with its labels."""
        self._lib.check(self.lib.b200ds_gen_batch_bf16_counted(self.xq.data_ptr(), self.y.data_ptr(), self.B, self.F,
                                                               self.seed, 0, self.counter.data_ptr(), st))

    dominant = dict(name="mlp_backward_f32_kernel (+finish)", bound="tensor")   # overridden per instance for bf16

    def dominant_flops(self):
        # algorithmic work of the backward: dW1 2FH, dh 2H, dW2 2H per sample (recompute not counted)
        return self.B * (2 * self.F * self.H + 4 * self.H)


class WideInputStep:
    """Device-resident step for the wide-input workloads (F = H): layer 1 as K-loop tcgen05 GEMMs (csrc/mlp_fh.cu).
    Same surface as WideStep."""

    tc = True
    fused_update = False
    p2p = False

    def __init__(self, lib, ctx, dev, wl, world, rank, precision):
        from b200ds import _lib
        assert precision == "bf16", "the wide-input path is the bf16 tensor-core path"
        self.lib, self.ctx, self._lib = lib, ctx, _lib
        B, F, H = wl["B"], wl["F"], wl["H"]
        assert lib.b200ds_fh_supported(B, F, H) and lib.b200ds_fh_gen_supported(B, F)
        self.B, self.F, self.H, self.world = B, F, H, world
        self.P = F * H + 2 * H + 1
        self.Ppad = (self.P + 3) // 4 * 4
        flat = init_flat(F, H)
        z = lambda n: torch.zeros(n, dtype=torch.float32, device=dev)      # noqa: E731
        u8 = lambda n: torch.zeros(max(256, n), dtype=torch.uint8, device=dev)      # noqa: E731
        self.params = z(self.Ppad); self.params[: self.P] = flat.to(dev)
        self.grads, self.m, self.v = z(self.Ppad), z(self.Ppad), z(self.Ppad)
        self.y, self.logits, self.dz, self.loss = z(B), z(B), z(B), z(1)
        self.xq = u8(lib.b200ds_fh_operand_bytes(B, F))
        self.wq = u8(lib.b200ds_fh_operand_bytes(H, F))
        self.hT = u8(lib.b200ds_fh_operand_bytes(H, B))
        self.gen_sc = u8(lib.b200ds_fh_gen_scratch_bytes(B, F))
        self.fwd_sc = u8(lib.b200ds_fh_forward_scratch_bytes(B, F, H))
        self.bwd_sc = u8(lib.b200ds_fh_backward_scratch_bytes(B, F, H))
        self.bce_sc = u8(lib.b200ds_bce_scratch_bytes(B))
        self.counter = torch.zeros(1, dtype=torch.int64, device=dev)
        self.seed = SEED + 1000003 * rank
        self.primed = False
        self.dominant = dict(name="fh_fwd_kernel (layer-1 GEMM + fused epilogue)", bound="tensor")

    launches = 11      # gen, labels, forward, logits, loss, relu backward, finish, (memset,) wgrad, adam, pack

    def dominant_flops(self):
        return 2 * self.B * self.F * self.H

    def gen_next(self, st):
        self._lib.check(self.lib.b200ds_fh_gen_batch_counted(self.xq.data_ptr(), self.y.data_ptr(), self.B, self.F, self.seed, 0,
                                                             self.counter.data_ptr(), self.gen_sc.data_ptr(), self.gen_sc.numel(), st))

    def pack(self, st):
        self._lib.check(self.lib.b200ds_fh_pack_bf16(self.params.data_ptr(), self.H, self.F, self.wq.data_ptr(), st))

    def step(self, st, ev_pair=None):
        lib, chk = self.lib, self._lib.check
        B, F, H = self.B, self.F, self.H
        if not self.primed:
            self.gen_next(st)
            self.pack(st)
            self.primed = True
        if ev_pair:
            ev_pair[0].record()
        chk(lib.b200ds_fh_forward(self.xq.data_ptr(), self.wq.data_ptr(), self.params.data_ptr(), B, F, H, self.hT.data_ptr(),
                                  self.fwd_sc.data_ptr(), self.fwd_sc.numel(), st))
        if ev_pair:
            ev_pair[1].record()
        chk(lib.b200ds_fh_logits(self.fwd_sc.data_ptr(), self.params.data_ptr(), B, F, H, self.logits.data_ptr(), st))
        chk(lib.b200ds_bce_loss_grad(self.logits.data_ptr(), self.y.data_ptr(), B, self.loss.data_ptr(), self.dz.data_ptr(),
                                     self.counter.data_ptr(), self.bce_sc.data_ptr(), self.bce_sc.numel(), st))
        chk(lib.b200ds_fh_backward(self.xq.data_ptr(), self.hT.data_ptr(), self.dz.data_ptr(), self.params.data_ptr(), B, F, H,
                                   self.grads.data_ptr(), self.bwd_sc.data_ptr(), self.bwd_sc.numel(), st))
        # the 4 P-byte gradient bucket crosses NVLink (NCCL, side stream) while the next batch is drawn: the generator is
        # pure arithmetic, the all-reduce pure data movement.  (Drawing the next batch into a second buffer BESIDE the
        # backward GEMM was measured too: 4.30 ms against 4.25 ms per step -- under the 1 kW power cap the generator's
        # arithmetic only takes clock from the tensor cores.)
        chk(lib.b200ds_allreduce_bucket_begin(self.ctx, self.grads.data_ptr(), self.P, st))
        self.gen_next(st)
        chk(lib.b200ds_allreduce_bucket_end(self.ctx, st))
        chk(lib.b200ds_adam_step_counted(self.params.data_ptr(), self.grads.data_ptr(), self.m.data_ptr(), self.v.data_ptr(),
                                         self.Ppad, LR, 0.9, 0.999, 1e-8, self.counter.data_ptr(), 1.0 / self.world, 0, st))
        self.pack(st)


def run_ours(a):
    import torch.distributed as dist

    from b200ds import _lib, comm
    lib = _lib.load()
    le = comm.read_launch_env()
    world, rank, local = le.world_size, le.rank, le.local_rank
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback); use --impl reference for the CPU arm"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        comm.init_distributed()
    wl = WORKLOADS[a.workload]
    precision = a.precision
    if precision == "auto":
        precision = "bf16" if (a.workload != "tiny" and (lib.b200ds_tc_supported(wl["B"], wl["F"], wl["H"]) or
                                                          lib.b200ds_fh_supported(wl["B"], wl["F"], wl["H"]))) else "fp32"
    ctx = C.c_void_p()
    nid = comm.exchange_nccl_id(lambda: _nccl_id(lib), le)
    idbuf = C.create_string_buffer(nid, 128) if nid else None
    _lib.check(lib.b200ds_ctx_create(local, rank, world, C.cast(idbuf, C.c_void_p) if idbuf else None, C.byref(ctx)))
    pk = peaks()
    st = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    out = {}
    if a.workload == "tiny":
        res = bench_tiny(lib, ctx, dev, world, rank, a, barrier)
        out.update(res)
    else:
        wide_inputs = a.workload.startswith("wideFH")
        ws = (WideInputStep if wide_inputs else WideStep)(lib, ctx, dev, wl, world, rank, precision)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
        W = max(3, a.warmup)
        for _ in range(W):
            ws.step(st)
        barrier()
        trace("eager warm-up done")
        # ---- the step as ONE CUDA graph (kernels + the side-stream all-reduce): no per-kernel launch cost
        graph = torch.cuda.CUDAGraph()
        cap = torch.cuda.Stream(device=dev)
        cap.wait_stream(torch.cuda.current_stream())
        with torch.cuda.graph(graph, stream=cap):
            ws.step(torch.cuda.current_stream().cuda_stream)
        torch.cuda.current_stream().wait_stream(cap)
        trace("graph captured")
        graph.replay()
        torch.cuda.synchronize()
        trace("first replay done")
        # untimed: let the SM clock settle under load.  A FIXED replay count: every replay holds a
        # collective, so all ranks must run the same number of them.
        clk = ClockSampler(local)
        if rank == 0:             # one sampler per job (NVML queries from 8 processes at once stall the launch threads);
            clk.start()           # it watches the identical load of the warm-up replays and the timed region
        settle = int(os.environ.get("B200DS_BENCH_SETTLE", 60 if wide_inputs else (4000 if wl["H"] <= 4096 else 2000)))   # profiling runs shorten it
        for _ in range(settle):
            graph.replay()
        barrier()
        trace("clock warm-up replays done")
        K = a.steps
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        align = torch.zeros(1, device=dev)
        wall0 = time.perf_counter()
        for i in range(K):
            if not os.environ.get("B200DS_BENCH_NOFLUSH"):
                flush.fill_(i & 0xFF)             # L2 flush between timed steps, outside the event bracket
            if i == 0:
                # The barrier above lines the HOSTS up; they then reach this loop tens of microseconds apart and submit the
                # first replay just in time, so the first step of an early rank would time its wait for the last rank's
                # submission (measured at 8 GPUs: 256-287 us against 111 for every later step).  A few more flush writes
                # (outside every event bracket, like the flush itself) keep the device busy while each host gets several
                # steps ahead, and one collective queued behind them lines the DEVICES up right in front of step 0.
                for _ in range(8):
                    flush.fill_(0)
                if world > 1:
                    dist.all_reduce(align)
            evs[i][0].record()
            graph.replay()
            evs[i][1].record()
        if rank == 0:
            clk.sample()          # GPU still busy with the queued steps; nothing NVML-related runs inside the loop
        barrier()
        wall = time.perf_counter() - wall0
        clk.stop()
        trace("timed region done")
        dev_ms = sum(e0.elapsed_time(e1) for e0, e1 in evs)
        if os.environ.get("B200DS_BENCH_STEP_TIMES"):      # tuning runs: this rank's per-step device times
            out["step_ms"] = [round(e0.elapsed_time(e1), 4) for e0, e1 in evs]
        # ---- dominant kernel: the same step graph captured once more with two event-record nodes around that kernel
        # (external events: they stay readable after a replay), replayed K2 times with the same L2 flush in front; the
        # kernel is timed where it runs in the timed region -- inside the graph, behind the forward, on its own stream
        K2 = max(3, min(K, 10))
        k_note = "kernel_ms: CUDA events recorded by nodes of the step graph around the kernel, %d replays right after the timed region" % K2
        try:
            kev = (torch.cuda.Event(enable_timing=True, external=True), torch.cuda.Event(enable_timing=True, external=True))
            graph2 = torch.cuda.CUDAGraph()
            cap.wait_stream(torch.cuda.current_stream())
            with torch.cuda.graph(graph2, stream=cap):
                ws.step(torch.cuda.current_stream().cuda_stream, kev)
            torch.cuda.current_stream().wait_stream(cap)
            k_ms = 0.0
            for i in range(K2):
                flush.fill_(i & 0xFF)
                graph2.replay()
                torch.cuda.synchronize()
                k_ms += kev[0].elapsed_time(kev[1])
            k_ms /= K2
            graph2.reset()
            del graph2
        except Exception as e:      # noqa: BLE001  (a torch without external events: the eager pass, which adds launch gaps)
            trace("in-graph kernel timing unavailable (%r); eager pass" % (e,))
            k_note = "kernel_ms: CUDA events around the kernel in an eager pass of the same step right after the timed region"
            kevs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K2)]
            for i in range(K2):
                flush.fill_(i & 0xFF)
                ws.step(st, kevs[i])
            torch.cuda.synchronize()
            k_ms = sum(e0.elapsed_time(e1) for e0, e1 in kevs) / K2
        t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t)
        ms_per_step = dev_ms / K
        value = world * wl["B"] * K / (dev_ms / 1e3)
        loss_last = float(ws.loss)
        achieved = ws.dominant_flops() / (k_ms / 1e3) / 1e12
        traffic = profiled_traffic(a.workload, precision)
        roof = dict(bound="tensor", achieved=achieved, peak=pk["tf_sus"], unit="TFLOP/s", frac=achieved / pk["tf_sus"],
                    traffic=traffic, kernel=ws.dominant["name"], kernel_ms=k_ms, peak_source=f"{pk['src']} bf16 sustained",
                    step_frac=step_flops(**wl) / (ms_per_step / 1e3) / 1e12 / pk["tf_sus"],
                    note=k_note)
        # ---- e2e: the C-ABI host-buffer call, H2D/D2H inside the timed region
        graph.reset()          # NCCL will not tear a communicator down while a captured graph still refers to it
        del graph
        trace("kernel pass done")
        if world > 1:
            out["xrank"] = cross_rank_checks(ws, st, dev, world)
            trace("cross-rank checks done")
            out["allreduce"] = bucket_allreduce_rate(lib, ctx, ws, st, dev, world)
        e2e = None if os.environ.get("B200DS_BENCH_SKIP_E2E") else bench_e2e(lib, ctx, dev, wl, world, a, barrier, precision)   # tuning runs only
        trace("e2e done")
        out.update(dict(value=value, ms_per_step=ms_per_step, roofline=roof, e2e=e2e, clocks=clk.summary(),
                        gpu_launches=K * ws.launches, wall_s=wall, loss_last=loss_last,
                        config={**workload_config(a.workload, wl, world),
                                "l2": "256 MiB device buffer rewritten between timed steps, outside the per-step CUDA-event brackets",
                                "exchange": (("finish + push over peer memory + Adam + pack (b200ds_update_tc, two launches)" if ws.fused_update else
                                              "peer-memory all-reduce + Adam in one kernel (b200ds_allreduce_adam_p2p)") if ws.p2p else
                                             "ncclAllReduce on a side stream + Adam kernel") if world > 1 else "none (1 rank)",
                                "timing": "each step = one CUDA-graph replay of the whole step (all kernels + the gradient exchange), "
                                          "bracketed by CUDA events on the launching stream; sum over steps, max over ranks"},
                        dtype="f32" if precision == "fp32" else "bf16"))
        if not a.no_tiny and rank == 0 and world == 1:
            try:
                out["tiny"] = bench_tiny(lib, ctx, dev, 1, 0, a, barrier, brief=True)
            except Exception as e:      # noqa: BLE001
                out["tiny"] = {"error": repr(e)}
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cb = cpu_reference_run(wl if a.workload != "tiny" else WORKLOADS["tiny"], steps=1000 if a.workload == "tiny" else 12,
                               warmup=1, budget_s=15.0)
        out["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
    elif rank == 0:
        out["cpu_baseline"] = None
    _lib.check(lib.b200ds_ctx_destroy(ctx))
    if rank == 0:
        line = dict(metric="train_samples_per_sec", value=out.pop("value"), unit="samples/s", n_gpus=world, steps=a.steps,
                    warmup=max(3, a.warmup), ms_per_step=out.pop("ms_per_step"), higher_is_better=True, scaling="weak",
                    vs_baseline=None, dtype=out.pop("dtype"), data="synthetic: on-device Philox N(0,1) batches with the label rule "
                    "of train.py:22-23; weights random-init with nn.Linear's default init in train.py's order",
                    config=out.pop("config"))
        line.update(out)
        print(json.dumps(line), flush=True)
    if world > 1:
        try:
            dist.barrier()
            dist.destroy_process_group()
        except Exception:       # noqa: BLE001
            pass


def bucket_allreduce_rate(lib, ctx, ws, st, dev, world):
    """The step's gradient bucket (4 P bytes) through ncclAllReduce alone, outside the timed region: bus GB/s against
    NVLink 5's 900 GB/s per direction per GPU (north_star: "allreduce bus GB/s is reported against NVLink peak")."""
    import torch.distributed as dist
    from b200ds import _lib
    buf = torch.ones(ws.Ppad, dtype=torch.float32, device=dev)
    n = ws.P
    for _ in range(3):
        _lib.check(lib.b200ds_allreduce_bucket(ctx, buf.data_ptr(), n, st))
    dist.barrier(); torch.cuda.synchronize()
    iters = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        _lib.check(lib.b200ds_allreduce_bucket(ctx, buf.data_ptr(), n, st))
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    us = float(t) * 1e3
    algbw = 4 * n / (us * 1e-6) / 1e9
    bus = algbw * 2 * (world - 1) / world
    return dict(bucket_bytes=4 * n, us=round(us, 2), algbw_gbs=round(algbw, 1), busbw_gbs=round(bus, 1), nvlink_nominal_gbs=900,
                frac_of_nominal=round(bus / 900, 3), note="ncclAllReduce of the flat fp32 gradient bucket, eager back-to-back calls")


def cross_rank_checks(ws, st, dev, world):
    """Outside the timed region, N > 1: (1) the replicas are bit-equal -- a 64-bit checksum of the whole parameter
    vector and a 1 KB sample, all-gathered and compared on every rank; (2) the exchange bench.py runs (peer memory)
    against the NCCL arm: 3 more steps from the same state with each, largest relative parameter difference
    (the peer kernel sums in rank order, NCCL in ring/tree order: round-off only).  train.py:113-114."""
    import torch.distributed as dist
    P = ws.P

    def fingerprint():
        torch.cuda.synchronize()
        bits = ws.params[:P].view(torch.int32).to(torch.int64)
        chk = torch.stack([bits.sum(), (bits * (torch.arange(P, device=dev) % 8191 + 1)).sum()])
        sample = ws.params[:256].view(torch.int32).to(torch.int64)
        mine = torch.cat([chk, sample])
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        return all(torch.equal(allr[0], a) for a in allr), int(chk[0])

    equal, chk = fingerprint()
    res = {"params_bit_equal_across_ranks": bool(equal), "params_checksum": chk}
    if ws.tc:
        keep = {k: getattr(ws, k).clone() for k in ("params", "grads", "m", "v", "xq", "y", "wq", "counter", "loss")}

        def restore():
            for k, t in keep.items():
                getattr(ws, k).copy_(t)
            torch.cuda.synchronize()
            dist.barrier()

        arm = (ws.p2p, ws.fused_update)
        for _ in range(3):
            ws.step(st)
        eq2, _ = fingerprint()
        ours = ws.params[:P].clone()
        restore()
        ws.p2p, ws.fused_update = False, False           # ncclAllReduce on the side stream + the Adam kernel + the pack kernel
        for _ in range(3):
            ws.step(st)
        torch.cuda.synchronize()
        ref = ws.params[:P]
        res.update({"extra_steps": 3, "bit_equal_after_extra_steps": bool(eq2),
                    "exchange_vs_nccl_max_rel_diff": float(((ours - ref).abs() / ref.abs().clamp_min(1e-6)).max()),
                    "exchange_vs_nccl_max_abs_diff": float((ours - ref).abs().max()),
                    "exchange": "b200ds_update_tc (push over peer memory)" if arm[1] else
                                ("b200ds_allreduce_adam_p2p" if arm[0] else "ncclAllReduce")})
        restore()
        ws.p2p, ws.fused_update = arm
    if not equal:
        raise RuntimeError("replicas differ across ranks after the timed region: %r" % (res,))
    return res


def _nccl_id(lib):
    from b200ds import _lib
    buf = C.create_string_buffer(128)
    _lib.check(lib.b200ds_nccl_unique_id(buf))
    return buf.raw


def bench_e2e(lib, ctx, dev, wl, world, a, barrier, precision):
    """Same step through b200ds_train_step_host: inputs in pinned host memory, loss read
    back to the host every step."""
    import torch.distributed as dist

    from b200ds import _lib
    B, F, H = wl["B"], wl["F"], wl["H"]
    P = F * H + 2 * H + 1
    Ppad = (P + 3) // 4 * 4
    flat = init_flat(F, H)
    z = lambda n: torch.zeros(n, dtype=torch.float32, device=dev)      # noqa: E731
    params = z(Ppad); params[:P] = flat.to(dev)
    grads, m, v, xs, ys, lg, dl, ls = z(Ppad), z(Ppad), z(Ppad), z(2 * B * F), z(2 * B), z(B), z(B), z(1)   # staging: two batches
    sc = torch.zeros(max(256, lib.b200ds_train_step_scratch_bytes(B, F, H)), dtype=torch.uint8, device=dev)
    tc = precision == "bf16"
    fh = tc and bool(lib.b200ds_fh_train_supported(B, F, H))
    xq = torch.empty(lib.b200ds_fh_operand_bytes(B, F) if fh else lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=dev) if tc else None
    wq = torch.empty(lib.b200ds_fh_operand_bytes(H, F) if fh else lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=dev) if tc else None
    bufs = _lib.StepBufs(params.data_ptr(), grads.data_ptr(), m.data_ptr(), v.data_ptr(), xs.data_ptr(), ys.data_ptr(),
                         lg.data_ptr(), dl.data_ptr(), ls.data_ptr(), sc.data_ptr(), sc.numel(),
                         xq.data_ptr() if tc else None, wq.data_ptr() if tc else None)
    nbuf = 2 if fh else 4                    # wide inputs: 1 GiB of pinned host memory per batch
    xh = [torch.randn(B, F).pin_memory() for _ in range(nbuf)]
    yh = [label_rule(x).contiguous().pin_memory() for x in xh]
    lh = torch.zeros(1).pin_memory()
    st = torch.cuda.current_stream().cuda_stream
    step = 0
    K = max(3, min(a.steps, 8 if fh else 20))
    def host_step(i):
        # batch i from pinned host memory; batch i+1 announced so that its copy runs behind this step's kernels
        j, jn = i % nbuf, (i + 1) % nbuf
        _lib.check(lib.b200ds_train_step_host_prefetch(ctx, C.byref(bufs), xh[j].data_ptr(), yh[j].data_ptr(),
                                                       xh[jn].data_ptr(), yh[jn].data_ptr(), B, F, H,
                                                       LR, 0.9, 0.999, 1e-8, step, lh.data_ptr(), st))

    for i in range(3):
        host_step(i)
        step += 1
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(3, 3 + K):
        host_step(i)
        step += 1
    e1.record()
    barrier()
    dt = max(time.perf_counter() - t0, e0.elapsed_time(e1) / 1e3)
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    # the same host->device copies alone, all ranks at once: what the box's host side can feed (the e2e step needs
    # 5.5 MB per rank per step; with 8 ranks the copies share the host's memory and PCIe root complexes)
    def copies(n):
        for i in range(n):
            xs[: B * F].copy_(xh[i % nbuf].view(-1), non_blocking=True)
            ys[:B].copy_(yh[i % nbuf].view(-1), non_blocking=True)
    copies(3)                                # the first copies of a loop pay torch's one-time set-up for pinned sources
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(3):                       # best of three passes: a single pass of 20 copies right after an idle period reads low
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        copies(K)
        c1.record()
        barrier()
        best = min(best, c0.elapsed_time(c1) / 1e3)
    tc_ = torch.tensor([best], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tc_, op=dist.ReduceOp.MAX)
    h2d = 4 * (B * F + B)
    return dict(value=world * B * K / float(t), unit="samples/s", h2d_bytes_per_step=h2d, d2h_bytes_per_step=4,
                h2d_gbs_per_gpu=h2d * K / float(t) / 1e9, h2d_only_gbs_per_gpu=h2d * K / float(tc_) / 1e9,
                h2d_only_samples_per_s=world * B * K / float(tc_),
                steps=K, api="b200ds_train_step_host_prefetch (C ABI, pinned host x/y in, host loss out, stream sync per step; "
                "the copy of batch k+1 is queued on a copy stream right behind the launch of batch k's kernels)",
                loss_last=float(lh))


def bench_tiny(lib, ctx, dev, world, rank, a, barrier, brief=False):
    """The reference's own shape (B=4/rank, F=20, H=64, N=2000): epochs of 500/W steps through
    the fused kernel(s) on the device-resident dataset; latency-bound, no roofline quoted."""
    import torch.distributed as dist

    from b200ds import _lib
    torch.manual_seed(SEED)
    x = torch.randn(2000, 20)
    y = label_rule(x)
    flat = init_flat(20, 64)
    P, Ppad = 1409, 1412
    z = lambda n: torch.zeros(n, dtype=torch.float32, device=dev)      # noqa: E731
    params = z(Ppad); params[:P] = flat.to(dev)
    grads, m, v = z(Ppad), z(Ppad), z(Ppad)
    xd, yd = x.to(dev), y.reshape(-1).to(dev)
    g = torch.Generator(); g.manual_seed(1)
    order = torch.randperm(2000, generator=g)[rank::world].to(torch.int64).to(dev)      # DistributedSampler-style shard
    n_steps = (order.numel() + 3) // 4
    losses = z(n_steps)
    st = torch.cuda.current_stream().cuda_stream
    step0 = [0]

    def epoch():
        if world == 1:
            _lib.check(lib.b200ds_tiny_train_steps_f32(params.data_ptr(), grads.data_ptr(), m.data_ptr(), v.data_ptr(), 20, 64,
                                                       xd.data_ptr(), yd.data_ptr(), order.data_ptr(), order.numel(), 0, 4,
                                                       n_steps, 1, LR, 0.9, 0.999, 1e-8, step0[0], losses.data_ptr(), st))
            step0[0] += n_steps
            return 1
        for s in range(n_steps):
            _lib.check(lib.b200ds_tiny_train_steps_f32(params.data_ptr(), grads.data_ptr(), None, None, 20, 64, xd.data_ptr(),
                                                       yd.data_ptr(), order.data_ptr(), order.numel(), 4 * s, 4, 1, 0,
                                                       LR, 0.9, 0.999, 1e-8, step0[0], losses.data_ptr() + 4 * s, st))
            _lib.check(lib.b200ds_allreduce_bucket(ctx, grads.data_ptr(), P, st))
            step0[0] += 1
            _lib.check(lib.b200ds_adam_step(params.data_ptr(), grads.data_ptr(), m.data_ptr(), v.data_ptr(), Ppad,
                                            LR, 0.9, 0.999, 1e-8, step0[0], 1.0 / world, 0, st))
        return 2 * n_steps
    for _ in range(3):
        epoch()
    barrier()
    reps = 5 if brief else max(5, a.steps // 8)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clk = ClockSampler(dev.index)
    e0.record()
    launches = 0
    for _ in range(reps):
        launches += epoch()
    e1.record()
    clk.sample()
    barrier()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t)
    steps_total = reps * n_steps
    samples = world * reps * 2000 if world > 1 else reps * 2000
    res = dict(samples_per_s=samples / (ms / 1e3), us_per_step=1e3 * ms / steps_total, steps=steps_total,
               launches=launches, epoch_mean_loss=float(losses.double().mean()),
               note="reference shape B=4/rank F=20 H=64 N=2000; fused one-CTA kernel, device-resident dataset; latency-bound")
    if brief:
        return res
    return dict(value=res["samples_per_s"], ms_per_step=res["us_per_step"] / 1e3, dtype="f32", gpu_launches=launches,
                clocks=clk.summary(), roofline=None, e2e=None, tiny=res,
                config={"workload": "tiny: train.py reference shape, F=20 H=64, 4 samples/rank/step, N=2000 resident dataset",
                        "parallelism": f"dp{world}", "l2": "not flushed: 168 kB working set, latency-bound by design"})


def run_reference(a):
    """`--impl reference`: the reference's CPU implementation of the step on the box's host cores, same workload,
    metric and unit as our arm, full-size steps.  N = 1: one process with every core.  N > 1:
    N gloo processes, OMP threads = cores // N each, spawned by rank 0 (the other torchrun ranks exit at once)."""
    import subprocess
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    world = max(1, a.gpus)
    wl = WORKLOADS[a.workload]
    cores = len(os.sched_getaffinity(0))
    extra = {}
    if world == 1:
        r = cpu_reference_run(wl, steps=a.steps, warmup=max(1, a.warmup))
    else:
        threads = max(1, cores // world)
        # the workers rendezvous among themselves (tcp://127.0.0.1:port): nothing of the launcher's environment may reach
        # them -- with TORCHELASTIC_USE_AGENT_STORE set, torch's tcp:// rendezvous turns every rank into a CLIENT of the
        # elastic agent's store and waits for a server that does not exist on that port
        drop = ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "GROUP_RANK", "ROLE_RANK",
                "GROUP_WORLD_SIZE", "ROLE_WORLD_SIZE", "ROLE_NAME")
        env = {k: v for k, v in os.environ.items() if k not in drop and not k.startswith("TORCHELASTIC_")}
        env["OMP_NUM_THREADS"] = str(threads)
        port = 29611 + (os.getpid() % 2000)
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--gpus", str(world),
                                   "--steps", str(a.steps), "--warmup", str(max(1, a.warmup)), "--workload", a.workload,
                                   "--ref-rank", str(i), "--ref-port", str(port), "--ref-threads", str(threads)],
                                  env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for i in range(world)]
        outs = [p.communicate(timeout=1500) for p in procs]
        if any(p.returncode for p in procs):
            raise RuntimeError("CPU reference worker failed: " + " | ".join(o[1][-400:] for o in outs))
        r = json.loads([ln for ln in outs[0][0].splitlines() if ln.startswith("REFWORKER ")][-1][len("REFWORKER "):])
    if a.workload != "tiny":
        # the figure the reference itself pins (slurm_train.sbatch:16, OMP_NUM_THREADS=1): one thread, a bounded sample
        one = cpu_reference_run(wl, steps=2, warmup=1, budget_s=20.0, threads=1)
        extra["one_thread"] = {"value": one["value"], "unit": "samples/s", "cores": 1, "sample": one["sample"]}
        torch.set_num_threads(cores)
    cfg = workload_config(a.workload, wl, world)
    cfg["arm"] = ("CPU reference path: %d process(es) x %d thread(s), gloo" % (world, r["threads"])) if world > 1 else \
                 ("CPU reference path: one process, %d threads" % r["threads"])
    line = dict(impl="reference", metric="train_samples_per_sec", value=r["value"], unit="samples/s", n_gpus=world,
                steps=r["steps"], warmup=max(1, a.warmup), ms_per_step=r["ms_per_step"], higher_is_better=True,
                scaling="weak", vs_baseline=None, dtype="f32",
                data="synthetic: host torch.randn batches with the label rule of train.py:22-23; nn.Linear default init",
                config=cfg,
                cpu_baseline={k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
                e2e=dict(value=r["value"], unit="samples/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                gpu_launches=0, **extra)
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference" and args.ref_rank >= 0:
        reference_worker(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
