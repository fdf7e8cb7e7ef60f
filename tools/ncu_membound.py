#!/usr/bin/env python3
"""The memory-bound kernels of the step at sizes where HBM GB/s means something (north_star: >= 80 % of the HBM
roofline): Adam at P = 16.8 M and 67.1 M parameters (28 B/param), the fused BCE loss + gradient at 4 194 304 rows
(12 B/row), the Philox normal generator at 64 Mi elements (4 B written per element), the bf16 operand pack of W1.
Run plain (prints CUDA-event GB/s) or under `ncu --set full -k regex:adam_kernel|bce_loss_grad_kernel|philox_normal_kernel|fh_pack_kernel`;
tools/ncu_summary.py turns the report into profiles/r02_ncu_membound.txt."""
import ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch
from b200ds import _lib
lib = _lib.load()
dev = torch.device("cuda", 0)
st = torch.cuda.current_stream().cuda_stream
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
reps = int(os.environ.get("REPS", 5))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(name, nbytes, fn):
    best = 1e9
    for i in range(reps):
        flush.fill_(i)                       # cold L2 (the operands are larger than L2 anyway)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    gbs = nbytes / best / 1e6
    print(f"{name:44s} {best * 1e3:9.1f} us  {gbs:8.1f} GB/s  {gbs / peak:5.2f} of {peak:.0f} (measured copy)", flush=True)


for n in (16785412, 67125252):              # P + pad for F = H = 4096 / 8192
    p, g, m, v = (torch.randn(n, device=dev) * 0.01 for _ in range(4))
    v.abs_()
    timed(f"adam_kernel n={n}", 28 * n, lambda: _lib.check(lib.b200ds_adam_step(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), n,
                                                                                  1e-3, 0.9, 0.999, 1e-8, 3, 1.0, 0, st)))
    del p, g, m, v
B = 4194304
z, y = torch.randn(B, device=dev), (torch.rand(B, device=dev) > 0.5).float()
dz, loss = torch.empty(B, device=dev), torch.zeros(1, device=dev)
sc = torch.zeros(lib.b200ds_bce_scratch_bytes(B), dtype=torch.uint8, device=dev)
timed(f"bce_loss_grad_kernel rows={B}", 12 * B, lambda: _lib.check(lib.b200ds_bce_loss_grad(z.data_ptr(), y.data_ptr(), B, loss.data_ptr(), dz.data_ptr(),
                                                                                             None, sc.data_ptr(), sc.numel(), st)))
B = 67108864                                 # more rows: the loss kernel's fixed costs amortised
z, y = torch.randn(B, device=dev), (torch.rand(B, device=dev) > 0.5).float()
dz = torch.empty(B, device=dev)
sc = torch.zeros(lib.b200ds_bce_scratch_bytes(B), dtype=torch.uint8, device=dev)
timed(f"bce_loss_grad_kernel rows={B}", 12 * B, lambda: _lib.check(lib.b200ds_bce_loss_grad(z.data_ptr(), y.data_ptr(), B, loss.data_ptr(), dz.data_ptr(),
                                                                                             None, sc.data_ptr(), sc.numel(), st)))
del z, y, dz
n = 64 << 20
x = torch.empty(n, device=dev)
timed(f"philox_normal_kernel n={n}", 4 * n, lambda: _lib.check(lib.b200ds_philox_normal(x.data_ptr(), n, 42, 0, None, st)))
R, K = 8192, 8192
w = torch.randn(R, K, device=dev)
wq = torch.empty(lib.b200ds_fh_operand_bytes(R, K), dtype=torch.uint8, device=dev)
timed(f"fh_pack_kernel {R}x{K}", 6 * R * K, lambda: _lib.check(lib.b200ds_fh_pack_bf16(w.data_ptr(), R, K, wq.data_ptr(), st)))
