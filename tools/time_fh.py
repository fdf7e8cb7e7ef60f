#!/usr/bin/env python3
"""Device time of the wide-input kernels (CUDA events around single launches after warm-up; these run for
milliseconds, launch cost does not matter)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch
from b200ds import _lib
lib = _lib.load()
B, F, H = int(os.environ.get("B", 65536)), int(os.environ.get("F", 4096)), int(os.environ.get("H", 4096))
dev = torch.device("cuda", 0)
torch.manual_seed(0)
P = F * H + 2 * H + 1; Ppad = (P + 3) // 4 * 4
params = torch.randn(Ppad, device=dev) * 0.02
x = torch.randn(B, F, device=dev)
st = torch.cuda.current_stream().cuda_stream
xq = torch.empty(lib.b200ds_fh_operand_bytes(B, F), dtype=torch.uint8, device=dev)
wq = torch.empty(lib.b200ds_fh_operand_bytes(H, F), dtype=torch.uint8, device=dev)
hT = torch.empty(lib.b200ds_fh_operand_bytes(H, B), dtype=torch.uint8, device=dev)
_lib.check(lib.b200ds_fh_pack_bf16(x.data_ptr(), B, F, xq.data_ptr(), st))
_lib.check(lib.b200ds_fh_pack_bf16(params.data_ptr(), H, F, wq.data_ptr(), st))
del x
fsc = torch.zeros(lib.b200ds_fh_forward_scratch_bytes(B, F, H), dtype=torch.uint8, device=dev)
bsc = torch.zeros(lib.b200ds_fh_backward_scratch_bytes(B, F, H), dtype=torch.uint8, device=dev)
dz = torch.randn(B, device=dev) * 1e-5
grads = torch.zeros(Ppad, device=dev)
def fwd(): _lib.check(lib.b200ds_fh_forward(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, hT.data_ptr(), fsc.data_ptr(), fsc.numel(), st))
def bwd(): _lib.check(lib.b200ds_fh_backward(xq.data_ptr(), hT.data_ptr(), dz.data_ptr(), params.data_ptr(), B, F, H, grads.data_ptr(), bsc.data_ptr(), bsc.numel(), st))
def pack(): _lib.check(lib.b200ds_fh_pack_bf16(params.data_ptr(), H, F, wq.data_ptr(), st))
flops = 2.0 * B * F * H
for name, fn, fl in (("fh_forward", fwd, flops), ("fh_backward (relu bwd + finish + memset + wgrad)", bwd, flops), ("pack W1", pack, 0)):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    n = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"B={B} F={F} H={H} {name:50s} {ms:8.3f} ms" + (f"  {fl / ms / 1e9:8.1f} TFLOP/s" if fl else ""), flush=True)
