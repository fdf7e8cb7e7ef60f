#!/usr/bin/env python3
"""Dump the per-chunk clock trace of CTA 0 of the forward kernel (needs a lib built with -DB200DS_TRACE)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch, numpy as np
from b200ds import _lib
lib = _lib.load()
B, F, H = int(os.environ.get("B", 303104)), 20, int(os.environ.get("H", 4096))
dev = torch.device("cuda", 0)
P = F * H + 2 * H + 1; Ppad = (P + 3) // 4 * 4
params = torch.randn(Ppad, device=dev) * 0.1
x = torch.randn(B, F, device=dev)
xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=dev)
wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=dev)
z = torch.empty(B, device=dev)
st = torch.cuda.current_stream().cuda_stream
_lib.check(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), st))
_lib.check(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), st))
for _ in range(200):
    _lib.check(lib.b200ds_mlp_forward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, z.data_ptr(), st))
torch.cuda.synchronize()
out = np.zeros(4 * 512, dtype=np.int64)
fn = lib.b200ds_debug_trace; fn.restype = C.c_int; fn.argtypes = [C.c_void_p]
assert fn(out.ctypes.data) == 0
t = out.reshape(4, 512); t0 = t[0, 0]
names = ["free_seen", "committed", "full_seen", "released"]
print("q   " + " ".join(f"{n:>10s}" for n in names) + "   full-commit  rel-full  period")
for q in range(int(os.environ.get("Q0", 64)), int(os.environ.get("Q1", 112))):
    r = t[:, q] - t0
    print(f"{q:3d} " + " ".join(f"{v:10d}" for v in r) + f"   {r[2]-r[1]:8d} {r[3]-r[2]:8d} {t[1,q]-t[1,q-1]:8d}")
